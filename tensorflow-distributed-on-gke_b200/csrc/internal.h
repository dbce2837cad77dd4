// Internal (C++) interfaces shared by the kernels, the step orchestration and the C ABI.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

#include "ptx.cuh"

namespace b200xf {

typedef __nv_bfloat16 bf16;

// ---- error / bookkeeping (thread-local last error; every launcher returns 0 on success)
int set_error(const char* fmt, ...);
const char* last_error_string();
int num_sms();
void count_launch(int n = 1);
long long launches_total();

bool pdl_enabled();  // B200XF_PDL=0 switches programmatic dependent launch off
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE property of a kernel: raise it once per (kernel, device),
// with the launching device current.  Fixed-size table, no heap use.
int ensure_dynamic_smem(const void* kernel, int bytes);
template <typename K> inline int ensure_dynamic_smem(K* kernel, int bytes) { return ensure_dynamic_smem(reinterpret_cast<const void*>(kernel), bytes); }
int stream_priority(cudaStream_t stream);  // cached cudaStreamGetPriority

// Launch with the programmatic-stream-serialization attribute (see ptx.cuh: pdl_wait / pdl_trigger).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  // the stream's priority, stated on the launch itself so that it is certainly part of a captured graph's kernel node
  attr[1].id = cudaLaunchAttributePriority;
  attr[1].val.priority = stream_priority(stream);
  cfg.attrs = attr; cfg.numAttrs = 2;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// Same, as thread-block clusters of `cluster_x` consecutive CTAs (grid.x must be a multiple of it).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, int cluster_x,
                                    Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  attr[1].id = cudaLaunchAttributePriority;
  attr[1].val.priority = stream_priority(stream);
  attr[2].id = cudaLaunchAttributeClusterDimension;
  attr[2].val.clusterDim.x = cluster_x; attr[2].val.clusterDim.y = 1; attr[2].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = cluster_x > 1 ? 3 : 2;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

#define B200XF_CUDA(expr)                                                                          \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) return ::b200xf::set_error("%s: %s", #expr, cudaGetErrorString(_e));    \
  } while (0)
#define B200XF_TRY(expr)          \
  do {                            \
    int _rc = (expr);             \
    if (_rc) return _rc;          \
  } while (0)

// ---- GEMM  D[M,N] = alpha * A[M,K] . B[N,K]^T (+bias)(relu)(+resid | *relumask(resid))
struct GemmArgs {
  const void* a = nullptr; int lda = 0; int a_mn = 0;  // a_mn=0: stored [M][K]; 1: stored [K][M]
  const void* b = nullptr; int ldb = 0; int b_mn = 0;  // b_mn=0: stored [N][K]; 1: stored [K][N]
  void* d = nullptr; int ldd = 0; int d_fp32 = 0;
  int accumulate = 0;  // fp32 atomic accumulate into d (enables split-K)
  int split_k = 0;     // 0 = auto
  int M = 0, N = 0, K = 0;
  const float* bias = nullptr;
  int relu = 0;
  const void* resid = nullptr; int ldr = 0; int resid_op = 0;  // 1 add, 2 relu-mask
  float alpha = 1.f;
  int force_bn = 0;
  int no_tma_store = 0;  // debug: force the direct-store epilogue
  int max_ctas = 0;      // cap on the persistent grid (0 = every SM)
  int b_static = 0;      // B (weights) is not written by the kernel launched just before: prefetch it ahead of the dependency wait
  int pair = 0;          // CTA pairs (cta_group::2, 256-row tiles): 0 auto, 1 force when M >= 256, -1 never
  DropArgs drop;         // inverted dropout on (acc + bias) before the residual add (transformer_model.py:196,202,232,238,244)
  // Weight-gradient layout only (a_mn = b_mn = 1): also emit the column sums of B over the contraction dimension -- the bias
  // gradient sum_tokens dY[token][n] of the Dense whose weight gradient this is -- as GemmInfo::splits * GemmInfo::tiles_m
  // partial rows of N floats (summed later by reduce_partials).  B's tiles are already in shared memory.
  float* colsum_partials = nullptr;
  // LayerNorm fused into the epilogue (N = the normalised width, bf16 output, no accumulate): d receives
  // z = epilogue(acc) (bias, dropout, residual add), ln_y receives (z - mean) * rstd * gamma + beta with the statistics taken
  // over the bf16-rounded z, ln_mean / ln_rstd the per-row statistics.  The CTAs that hold one row block's N tiles form a
  // thread-block cluster and exchange their per-row partial statistics through distributed shared memory.
  const float* ln_gamma = nullptr; const float* ln_beta = nullptr;
  void* ln_y = nullptr; int ln_ldy = 0; float* ln_mean = nullptr; float* ln_rstd = nullptr; float ln_eps = 1e-6f;
};
struct GemmInfo { int splits = 1; int tiles_m = 0, tiles_n = 0; int bn = 0; int pair = 0; int fused_ln = 0; };
bool gemm_ln_fusable(int M, int N);  // whether a GEMM of width N can carry the fused LayerNorm epilogue
int gemm_bf16(const GemmArgs& a, cudaStream_t stream, GemmInfo* info = nullptr);
int gemm_plan(const GemmArgs& a, GemmInfo* info);  // the tiling gemm_bf16 would choose (e.g. the number of K splits), nothing launched
void gemm_clear_tensor_map_cache();
void gemm_set_pair_mode(int mode);  // CTA pairs: -1 heuristic (default), 0 never, 1 whenever M > 128 and the tile is >= 128 wide
// Cached 2-D TMA descriptor (128B swizzle) of a row-major matrix: kind 0 = bf16 with a 64-column box, 1 = fp32 with a
// 32-column box; `out_cu_tensor_map` points at a CUtensorMap.
int tensor_map_2d(const void* ptr, unsigned long long inner, unsigned long long outer, unsigned long long ld, unsigned box_outer,
                  void* out_cu_tensor_map, int kind);
// debug: per-CTA phase time stamps (8 x u64 per CTA, 148 CTAs per slot); successive launches use successive slots of
// the ring (stamps are baked into a captured graph, so a replayed step fills its slots again); null to disable
void gemm_set_trace(unsigned long long* dev_ptr, int slots = 1);

// ---- attention (fused, masks derived in-tile from token ids; transformer_model.py:73-109,350-363)
struct AttnArgs {
  const bf16* q = nullptr; int ldq = 0;   // [B*Sq, ldq], head h at column h*dk
  const bf16* k = nullptr; int ldk = 0;   // [B*Sk, ldk]
  const bf16* v = nullptr; int ldv = 0;
  bf16* o = nullptr; int ldo = 0;         // [B*Sq, ldo]
  float* lse = nullptr;                   // [B,H,Sq,2] {row max, log2(row sum)} of the scaled+masked scores, log2 domain
  const long long* key_tokens = nullptr;  // [B,Sk] int64; key j masked iff token == 0
  int causal = 0;                         // additionally mask j > i
  int B = 0, H = 0, Sq = 0, Sk = 0, dk = 0;
  // backward only
  const bf16* d_o = nullptr; int lddo = 0;
  bf16* dq = nullptr; int lddq = 0;
  bf16* dk_out = nullptr; int lddk = 0;
  bf16* dv_out = nullptr; int lddv = 0;
  float* delta = nullptr;                 // scratch [B,H,Sq]
  float* dq_acc = nullptr;                // optional scratch: fp32 [B*Sq, H*dk] dQ accumulator (enables the multi-tile tcgen05 backward)
  // bit 0: q, bit 1: k and v are NOT outputs of the kernel launched just before this one on the stream (e.g. saved
  // forward activations in the backward pass): the single-tile tcgen05 kernels fetch them ahead of the dependency wait
  int early = 0;
  int generic = 0;  // forward: take the general (mma.sync) kernel whatever the shape (incremental decoding: Sq = 1 with strided rows)
};
int attention_fwd(const AttnArgs& a, cudaStream_t stream);
int attention_bwd(const AttnArgs& a, cudaStream_t stream);
// softmax weights [B,H,Sq,Sk] fp32 recomputed from q, k and the saved row statistics (inference API)
int attention_probs(const AttnArgs& a, float* out, cudaStream_t stream);
// tcgen05 single-tile path (Sq, Sk <= 128, head size 64); B200XF_ATTENTION_TC=0 disables it
bool attention_tc_supported(const AttnArgs& a);
int attention_fwd_tc(const AttnArgs& a, cudaStream_t stream);
int attention_bwd_tc(const AttnArgs& a, cudaStream_t stream);
// tcgen05 multi-tile backward (any Sq, Sk; head size 64; needs AttnArgs::dq_acc)
bool attention_tc_multi_supported(const AttnArgs& a);
bool attention_tc_fwd_multi_supported(const AttnArgs& a);
int attention_fwd_tc_multi(const AttnArgs& a, cudaStream_t stream);
int attention_bwd_tc_multi(const AttnArgs& a, cudaStream_t stream);

// ---- LayerNorm (+ optional residual add) fwd / bwd (transformer_model.py:187-188,198,204)
int layernorm_fwd(const bf16* x, const bf16* resid, const float* gamma, const float* beta, bf16* y, bf16* z_out,
                  float* mean, float* rstd, int rows, int d, float eps, cudaStream_t stream);
// Column sums are produced as per-block partials and summed later by one batched launch (reduce_partials), in a
// fixed order: deterministic, and the many small reductions of a backward pass share one launch.
struct ReduceJob { const float* src; float* dst; int nparts; int cols; int stride; };  // dst[c] = sum_p src[p*stride + c]
int reduce_partials(const ReduceJob* jobs, int njobs, cudaStream_t stream);
// dz = LN'(dy); partials [ln_bwd_parts(rows, d)][3][d] = per-block sums of (dy*xhat, dy, dz) -> dgamma, dbeta, colsum(dz).
int ln_bwd_parts(int rows, int d);
size_t ln_bwd_scratch_floats(int rows, int d);
// dzm (optional, with drop.thresh != 0): dz * dropout mask of the sub-layer output that was added to the residual — the
// gradient the producing Dense sees; the third partial sum is then taken over dzm.
int layernorm_bwd_partial(const bf16* dy, const bf16* z, const float* mean, const float* rstd, const float* gamma, bf16* dz,
                          int rows, int d, float* partials, cudaStream_t stream, bf16* dzm = nullptr, DropArgs drop = DropArgs());
// convenience: partial + reduce; scratch holds ln_bwd_scratch_floats(rows, d) floats.  dcol may be null.
int layernorm_bwd(const bf16* dy, const bf16* z, const float* mean, const float* rstd, const float* gamma, bf16* dz,
                  float* dgamma, float* dbeta, float* dcol, int rows, int d, float* scratch, cudaStream_t stream);

// ---- embedding * sqrt(d) + positional encoding (transformer_model.py:273-275,305-307)
int embed_pe_fwd(const long long* tokens, const float* table, int vocab, const float* pe, bf16* out, int B, int S, int d,
                 cudaStream_t stream, DropArgs drop = DropArgs());
size_t embed_bwd_scratch_bytes(int n_tokens);
// The two phases of embed_bwd separately.  (The sort needs only the token ids; running it early on another stream, under the
// first backward GEMMs, was measured: 3.992 against 3.982 ms/step -- it stays in line.)
bool embed_sort_is_single_cta(int n_tokens, int vocab);   // one CTA (batches up to 4 096 tokens) instead of a device-wide radix sort
int embed_bwd_sort(const long long* tokens, int vocab, int n_tokens, void* scratch, size_t scratch_bytes, cudaStream_t stream);
int embed_bwd_sorted(const bf16* dx, float* dtable, int vocab, int n_tokens, int d, const void* scratch, cudaStream_t stream, DropArgs drop = DropArgs());
int embed_bwd(const long long* tokens, const bf16* dx, float* dtable, int vocab, int n_tokens, int d, void* scratch,
              size_t scratch_bytes, cudaStream_t stream, DropArgs drop = DropArgs());
int dropout_mask_fill(float* out, int rows, int cols, DropArgs drop, cudaStream_t stream);  // out = scale or 0 per element (tests)

// ---- targets: shift, count (__main__.py:106-107; transformer_model.py:12)
int prep_targets(const long long* target, long long* tgt_in, long long* labels, int* nonpad_count, int B, int T_plus_1,
                 cudaStream_t stream);

// ---- masked softmax cross-entropy over the vocab, fwd + bwd in place (transformer_model.py:7-26)
// logits bf16 [rows, ldl]; writes dlogits in place; row_loss/row_correct fp32/int per row; finalize reduces them.
int vocab_ce_fwd_bwd(bf16* logits, int ldl, const long long* labels, int rows, int V, const int* nonpad_count,
                     float inv_workers, float* row_loss, int* row_correct, int write_grad, cudaStream_t stream);
// loss / accuracy terms from fp32 logits (no rounding, no gradient): the reference's public loss functions on Transformer.call's output
int vocab_ce_f32(const float* logits, int ldl, const long long* labels, int rows, int V, float* row_loss, int* row_correct, cudaStream_t stream);
int ce_finalize(const float* row_loss, const int* row_correct, const int* nonpad_count, int rows, float inv_workers,
                float* out3 /* loss, acc_num, acc_den */, cudaStream_t stream);

// ---- Adam + LR schedule over a flat fp32 range (__main__.py:56-73,113-114)
struct AdamHyper { float beta1, beta2, eps, rsqrt_d_model, warmup_pow_m15; };  // warmup_pow_m15 = float(warmup ** -1.5)
int adam_step(float* w, float* g, float* m, float* v, bf16* w_bf16, size_t n, const long long* iterations_dev,
              AdamHyper h, int zero_grad, cudaStream_t stream);
int increment_counter(long long* iterations_dev, cudaStream_t stream);
// shadow = bf16(w), g = 0 over n elements (after the all-gather of a sharded update)
int refresh_shadow_zero_grad(const float* w, bf16* shadow, float* g, size_t n, cudaStream_t stream);

// ---- small helpers
// partials [colsum_parts(rows, cols)][cols] of the column sums of a bf16 matrix
int colsum_parts(int rows, int cols);
int colsum_partial(const bf16* x, int ld, int rows, int cols, float* partials, cudaStream_t stream);
int cast_f32_to_bf16(const float* in, bf16* out, size_t n, cudaStream_t stream);
int argmax_rows(const float* x, int ld, int rows, int cols, long long* out, cudaStream_t stream);  // first maximum wins
// out[r][j] = j < len(first_row + r) ? flat[offsets[first_row + r] + j] : 0 for r < rows, j < L (english_portugese_dataset.py:36-39);
// *too_long_dev (caller-zeroed) receives the longest length that did not fit L
int assemble_padded(const long long* flat, const long long* offsets, int first_row, int rows, int L, long long* out,
                    int* too_long_dev, cudaStream_t stream);
int cast_bf16_to_f32_2d(const bf16* in, int ldi, float* out, int ldo, int rows, int cols, cudaStream_t stream);

}  // namespace b200xf
