// Context: parameters, optimizer state, activations/workspace and the step orchestration.
#pragma once
#include <map>
#include <string>
#include <utility>
#include <vector>

#include "../../include/b200xf.h"
#include "internal.h"
#include "nvls.h"

namespace b200xf {

struct Dense { size_t w = 0, b = 0; int K = 0, N = 0, ld = 0; };  // kernel [K, N] (row pitch ld), bias [N]
struct LnP { size_t g = 0, b = 0; };
struct EncLayer { Dense qkv, o, f1, f2; LnP ln1, ln2; };
struct DecLayer { Dense qkv1, o1, q2, o2, f1, f2; LnP ln1, ln2, ln3; };
struct ParamView { std::string name; int64_t rows, cols, ld, offset; int ndim; };

struct EncAct { bf16 *qkv, *ctx, *z1, *o1, *h, *z2; float *lse, *stats; };
struct DecAct { bf16 *qkv, *ctx1, *z1, *o1, *q2, *ctx2, *z2, *o2, *z3, *h; float *lse1, *lse2, *stats; };

int comm_unique_id(void* out128);
void positional_encoding_host(float* out, int positions, int d);

struct Context {
  b200xf_hparams hp{};
  int device = 0;
  // All work runs on a stream the context owns (high priority, capturable); every entry point first makes it wait for
  // the caller's stream and afterwards makes the caller's stream wait for it, so the caller sees ordinary stream order.
  cudaStream_t stream = nullptr, user_stream = nullptr;
  cudaEvent_t ev_user = nullptr, ev_own = nullptr;
  int enter();
  int leave();

  // ---- parameters (flat) and their logical views
  size_t total_params = 0;
  int vt_ld = 0;
  Dense final_layer, kv_all;
  size_t enc_emb = 0, dec_emb = 0;
  std::vector<EncLayer> enc;
  std::vector<DecLayer> dec;
  std::map<std::string, ParamView> views;
  std::vector<ParamView> table;
  std::vector<std::pair<size_t, size_t>> sections;  // [begin, end) ranges in backward completion order
  float *params = nullptr, *grads = nullptr, *adam_m = nullptr, *adam_v = nullptr;
  bf16* shadow = nullptr;
  bool grads_clean = true;
  AdamHyper adam{};
  long long* iterations_dev = nullptr;
  long long iterations_host = 0;

  // ---- workspace
  char* arena = nullptr; size_t arena_size = 0, arena_used = 0;
  bf16 *enc_x = nullptr, *dec_x = nullptr, *kv = nullptr, *dkv = nullptr, *logits = nullptr;
  std::vector<EncAct> ea; std::vector<DecAct> da;
  // Weight-gradient GEMMs and bias column sums run on a second stream, off the dgrad / LayerNorm / attention critical
  // path.  Buffers they read are double buffered: before the compute stream rewrites a copy it waits for the side-stream
  // readers of that copy (normally long finished).
  struct SideBuf { bf16* buf[2] = {nullptr, nullptr}; cudaEvent_t done[2] = {nullptr, nullptr}; bool pending[2] = {false, false}; int cur = 0; };
  SideBuf sb_dz, sb_dzm, sb_dh, sb_dqkv, sb_dq2;
  cudaStream_t side_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool side_on = true, side_dirty = false;
  int sb_write(SideBuf& b, bf16** out);
  int side_launch_begin(cudaStream_t* s);
  int side_launch_end(const void* read_ptr);
  int join_side();
  // optional phase timing (B200XF_PHASE_TIMING=1): events after forward / loss / backward / update, averaged at destroy
  bool phase_timing = false; cudaEvent_t ev_phase[5] = {nullptr, nullptr, nullptr, nullptr, nullptr}; double phase_ms[4] = {0, 0, 0, 0}; long phase_n = 0;
  int phase_mark(int i);
  bool dropout_on = false; unsigned long long dropout_seed = 0x5EEDull;
  bf16 *dzm = nullptr, *dsub = nullptr;
  bf16 *dA = nullptr, *dB = nullptr, *dz = nullptr, *dctx = nullptr, *dq2 = nullptr, *dh = nullptr, *dqkv = nullptr;
  float *dq_acc = nullptr, *delta = nullptr, *row_loss = nullptr, *red_scratch = nullptr, *pe_in = nullptr, *pe_tgt = nullptr, *metrics = nullptr;
  int *row_correct = nullptr, *nonpad_count = nullptr;
  size_t red_capacity = 0, red_used = 0;
  std::vector<ReduceJob> red_jobs;
  long long *tok_inp = nullptr, *tok_target = nullptr, *tok_tgt_in = nullptr, *tok_labels = nullptr;
  char* emb_scratch = nullptr; size_t emb_scratch_bytes = 0;
  float* host_out = nullptr;

  // ---- data parallel
  void* comm = nullptr; int rank = 0, world = 1;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_ready = nullptr, ev_done = nullptr;
  // sharded update, pipelined across buckets: collectives on comm_stream, Adam / refresh on upd_stream
  cudaStream_t upd_stream = nullptr;
  cudaEvent_t ev_rs = nullptr, ev_adam[2] = {nullptr, nullptr}, ev_ag = nullptr;
  struct PendingGather { bool valid = false; size_t begin = 0, off = 0, sh = 0, n = 0; int slot = 0; } pending_gather;
  int adam_slot = 0;
  int flush_pending_gather();
  // B200XF_PIPELINED_UPDATE: 0 one chain per bucket on comm_stream, 1 every bucket pipelined, 2 only the trailing buckets.
  // Measured on 2 GPUs: 4.44 / 4.47 / 4.42 ms per step -- spreading the update over two streams in mid-backward disturbs the
  // GEMMs more than it hides; for the buckets that end up behind the last backward kernel it shortens the exposed chain.
  int pipelined_update = 2;
  bool upd_dirty = false;
  size_t bucket_bytes = size_t(32) << 20, bucket_begin = 0, next_section = 0;
  // With several ranks each bucket is reduce-scattered, updated on the local 1/world shard and all-gathered (Adam m / v
  // are then only maintained for the local shard of every bucket).  B200XF_SHARDED_UPDATE=0 keeps the replicated update.
  // 2 = also with a one-rank communicator (exercises the sharded / pipelined chain on a single GPU; tests).
  int sharded_update = 1;
  bool sharded_active() const { return comm && sharded_update && (world > 1 || sharded_update == 2); }
  // After a sharded step Adam's m / v are current only on this rank's shard of every bucket; gather_optimizer_state
  // (a collective: every rank calls it) completes them everywhere.
  bool opt_state_complete = true;
  // With a communicator a bucket must amortise a collective's latency; alone, buckets of >= 8 MB.
  size_t bucket_want_bytes() const { return bucket_bytes_set ? bucket_bytes : (comm ? bucket_bytes : (size_t(8) << 20)); }
  std::vector<std::pair<size_t, size_t>> bucket_ranges() const;
  int gather_optimizer_state();
  // the bf16 weight copy is only written by the optimizer, never by the kernel just before a forward / data-gradient
  // GEMM on the compute stream, so those GEMMs may fetch their first weight tiles ahead of the dependency wait
  // Bit mask of the loads issued ahead of the dependency wait (B200XF_WEIGHT_PREFETCH): 1 weight tiles in forward /
  // data-gradient GEMMs, 2 saved Q/K/V in the attention kernels (backward: all three; forward cross-attention: K, V).
  // Measured and dropped: LayerNorm-backward saved state and the weight-gradient GEMM's activation operand (both slower).
  int weights_static = 3;
  int comm_sms = 0;  // SMs left to the collectives' kernels during backward (B200XF_COMM_SMS); 0 = GEMMs use every SM
  int backward_gemm_cap() const;
  // B200XF_BWD_GEMM_CTAS: cap on the grids of the backward GEMMs, so that a data-gradient GEMM (compute stream) and a weight-
  // gradient GEMM (side stream) can hold SMs at the same time, each with several tiles per CTA (0 = every SM; measured: DESIGN 4b)
  int bwd_gemm_ctas = 0;
  bool bucket_bytes_set = false;  // B200XF_BUCKET_MB given
  bool bucket_join = true;  // a closing bucket makes the compute stream wait for the side stream (B200XF_BUCKET_JOIN)
  bool overlap_update = false;
  bool overlap_single = true;   // single GPU too: each bucket's Adam runs on the update stream under the rest of backward (B200XF_OVERLAP_UPDATE=0 disables)

  static size_t arena_bytes_for(const b200xf_hparams& hp, int vt_ld, size_t* red_capacity_out, size_t* emb_scratch_out);
  // ---- NVLS update path (nvls_update.cu): parameters and gradients in caller-provided symmetric memory
  bool external_state = false;           // params / grads were handed in (use_external_state): not ours to free
  int use_external_state(float* params_ext, float* grads_ext);
  NvlsComm nvls; bool nvls_on = false;
  unsigned* nvls_words = nullptr;        // [0] launch counter of the update kernels, [1] CTA counter
  int nvls_ctas = 16, nvls_tail_ctas = 64;  // CTAs of the update kernel under backward / for the buckets that close after it (B200XF_NVLS_CTAS, _TAIL_CTAS)
  int nvls_attach(void* grads_mc, void* params_mc, void* const* pads, void* const* grads_peers, void* const* params_peers, int p2p, int rank_, int world_);
  ~Context();
  int init(const b200xf_hparams& h, int dev, cudaStream_t s);
  void build_layout();
  size_t add_tensor(size_t& cursor, size_t elems);
  Dense add_dense(size_t& cursor, int K, int N, int ld);
  void add_view(const std::string& name, size_t offset, int64_t rows, int64_t cols, int64_t ld, int ndim);
  template <typename T> int arena_alloc(T** out, size_t count);

  int params_changed();
  int set_iterations(long long it);
  int get_iterations(long long* it);
  int comm_init(const void* id128, int rank, int world);
  int comm_destroy();
  int allreduce(float* ptr, size_t count, cudaStream_t s);

  int dense_fwd(const bf16* x, int M, const Dense& p, bf16* y, int relu, const bf16* resid, int drop_site = -1);
  DropArgs site_drop(int site) const;
  int dense_dgrad(const bf16* dy, int M, const Dense& p, bf16* dx, const bf16* resid, int resid_op);
  // bias_grad: the weight-gradient GEMM also sums dy's columns (the Dense's bias gradient) from the tiles it stages anyway
  int dense_wgrad(const bf16* x, const bf16* dy, int M, const Dense& p, bool bias_grad = false);
  // y = LayerNorm(z), z = resid + drop(x W + b): one GEMM whose epilogue normalises (B200XF_FUSE_LN=0: GEMM, then LayerNorm kernel)
  int dense_ln_fwd(const bf16* x, int M, const Dense& p, bf16* z, const bf16* resid, int drop_site, const LnP& ln, bf16* y, float* mean, float* rstd);
  // B200XF_FUSE_LN / B200XF_FUSE_COLSUM: 1 always, 0 never (default), -1 from the batch size (>= B200XF_FUSE_MIN_TOKENS tokens).
  // Both fused epilogues are correct (tests/test_kernels_gpu.py, tests/test_bench_path_gpu.py run them) and save launches and an
  // HBM pass, yet measured slower inside the step at every configuration (profiles/r02_fusion_ab.md): the LayerNorm epilogue
  // is fully exposed behind a one-tile main loop and gives up the persistent kernel's epilogue / main-loop overlap; the column
  // sums read the staged tiles from shared memory, whose bandwidth the tensor core's own operand reads already use up.
  int fuse_ln = 0, fuse_colsum = 0, fuse_min_tokens = 16384;
  bool fused(int mode, int tokens) const { return mode > 0 || (mode < 0 && tokens >= fuse_min_tokens); }
  int ln_bwd(const bf16* dy, const bf16* z, const float* mean, const float* rstd, const LnP& p, size_t dcol_off, int rows, int drop_site);
  int colsum(const bf16* x, int ld, int rows, int cols, size_t dst_off);
  float* red_alloc(size_t floats);
  int flush_reductions(cudaStream_t on = nullptr);  // the partials scratch is not reused within a step, so this may run late
  int last_B = 0, last_S = 0, last_T = 0; const long long *last_inp = nullptr, *last_tgt = nullptr;
  int attention_weights(int layer, int block, float* out);
  int check_shapes(int B, int S, int T);
  int forward(const long long* inp, const long long* tgt_in, int B, int S, int T, float* logits_f32, bool training);
  int forward_encoder(const long long* inp, int B, int S);   // encoder stack + the K|V projections of every decoder layer's cross-attention
  // Incremental greedy decoding (tester.py:27-42 re-runs the whole model for every token; here the encoder output and the
  // cross-attention K|V are computed once and each decoder layer keeps the K|V of the tokens decoded so far).
  int dec_B = 0, dec_S = 0, dec_Tmax = 0; const long long* dec_inp = nullptr;
  int decode_begin(const long long* inp, int B, int S, int Tmax);
  int decode_step(const long long* tokens_t, int t, float* logits_f32, long long* next_tokens);
  int backward(const long long* inp, const long long* tgt_in, int B, int S, int T);
  int section_done();
  int loss_and_grad(int Nt, int workers_count, int write_grad);
  int step_device(const long long* inp, const long long* target, int B, int S, int T1, int workers_count, int mode, float* out6);
  int step_body(const long long* inp, const long long* target, int B, int S, int T1, int workers_count, int mode);
  // A train step's ~330 launches and ~300 event operations cost about as much host time as the GPU needs to run them, so
  // once a (B, S, T+1, workers) shape has been seen a few times the whole step is captured into a CUDA graph (PDL edges,
  // side streams and NCCL included) and replayed.  B200XF_GRAPH=0 disables this.
  struct StepGraph { cudaGraphExec_t exec = nullptr; int seen = 0; int launches = 0; bool counted = false; };
  std::map<std::vector<int>, StepGraph> graphs;
  bool use_graph = true;
  void drop_graphs();
  int step_host(const long long* inp, const long long* target, int B, int S, int T1, int workers_count, float* out6_host);
};

}  // namespace b200xf
