// extern "C" surface declared in include/b200xf.h.
#include <cmath>
#include <cstring>
#include <new>

#include "context.h"

using namespace b200xf;

struct b200xf_ctx { Context c; };

extern "C" {

const char* b200xf_last_error(void) { return last_error_string(); }
int b200xf_version(void) { return 100; }
long long b200xf_launch_count(void) { return launches_total(); }

int b200xf_ctx_create(const b200xf_hparams* hp, int device, void* cuda_stream, b200xf_ctx** out) {
  if (!hp || !out) return set_error("null argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return set_error("no CUDA device available: this library has no CPU fallback");
  b200xf_ctx* ctx = new (std::nothrow) b200xf_ctx();
  if (!ctx) return set_error("out of host memory");
  int rc = ctx->c.init(*hp, device, static_cast<cudaStream_t>(cuda_stream));
  if (rc) { delete ctx; return rc; }
  *out = ctx;
  return 0;
}
int b200xf_ctx_destroy(b200xf_ctx* ctx) { delete ctx; return 0; }

int b200xf_device_bytes(const b200xf_hparams* hp, int64_t* state_bytes, int64_t* workspace_bytes) {
  if (!hp || !state_bytes || !workspace_bytes) return set_error("null argument");
  Context c;                      // layout only: nothing is allocated
  c.hp = *hp;
  c.build_layout();
  *state_bytes = int64_t(c.total_params) * 18;   // fp32 weights, gradients, Adam m and v + the bf16 operand copy
  *workspace_bytes = int64_t(Context::arena_bytes_for(*hp, c.vt_ld, nullptr, nullptr));
  return 0;
}

int b200xf_param_count(b200xf_ctx* ctx) { return ctx ? int(ctx->c.table.size()) : 0; }
int b200xf_param_info(b200xf_ctx* ctx, int index, char* name_buf, size_t name_buf_len, int64_t* rows, int64_t* cols, int64_t* ld,
                      int64_t* offset, int32_t* ndim) {
  if (!ctx || index < 0 || index >= int(ctx->c.table.size())) return set_error("param index out of range");
  const ParamView& v = ctx->c.table[index];
  if (name_buf && name_buf_len) { strncpy(name_buf, v.name.c_str(), name_buf_len - 1); name_buf[name_buf_len - 1] = 0; }
  if (rows) *rows = v.rows; if (cols) *cols = v.cols; if (ld) *ld = v.ld; if (offset) *offset = v.offset; if (ndim) *ndim = v.ndim;
  return 0;
}
int64_t b200xf_param_total(b200xf_ctx* ctx) { return ctx ? int64_t(ctx->c.total_params) : 0; }
float* b200xf_param_ptr(b200xf_ctx* ctx) { return ctx ? ctx->c.params : nullptr; }
float* b200xf_grad_ptr(b200xf_ctx* ctx) { return ctx ? ctx->c.grads : nullptr; }
float* b200xf_adam_m_ptr(b200xf_ctx* ctx) { return ctx ? ctx->c.adam_m : nullptr; }
float* b200xf_adam_v_ptr(b200xf_ctx* ctx) { return ctx ? ctx->c.adam_v : nullptr; }
#define B200XF_SCOPED(ctx, expr)            \
  do {                                      \
    B200XF_TRY((ctx)->c.enter());           \
    const int _rc2 = (expr);                \
    if (_rc2) return _rc2;                  \
    return (ctx)->c.leave();                \
  } while (0)

int b200xf_params_changed(b200xf_ctx* ctx) {
  if (!ctx) return set_error("null ctx");
  B200XF_SCOPED(ctx, ctx->c.params_changed());
}
int b200xf_set_iterations(b200xf_ctx* ctx, int64_t it) {
  if (!ctx) return set_error("null ctx");
  B200XF_SCOPED(ctx, ctx->c.set_iterations(it));
}
int b200xf_get_iterations(b200xf_ctx* ctx, int64_t* it) {
  if (!ctx || !it) return set_error("null argument");
  long long v = 0;
  B200XF_TRY(ctx->c.enter());
  int rc = ctx->c.get_iterations(&v);
  *it = v;
  return rc ? rc : ctx->c.leave();
}

int b200xf_forward(b200xf_ctx* ctx, const int64_t* inp, const int64_t* tgt_in, int B, int S, int T, int training, float* logits_out) {
  if (!ctx || !inp || !tgt_in) return set_error("null argument");
  B200XF_SCOPED(ctx, ctx->c.forward(reinterpret_cast<const long long*>(inp), reinterpret_cast<const long long*>(tgt_in), B, S, T, logits_out, training != 0));
}
int b200xf_train_step_device(b200xf_ctx* ctx, const int64_t* inp, const int64_t* target, int B, int S, int T1, int workers, float* out6) {
  if (!ctx || !inp || !target) return set_error("null argument");
  B200XF_SCOPED(ctx, ctx->c.step_device(reinterpret_cast<const long long*>(inp), reinterpret_cast<const long long*>(target), B, S, T1, workers, 2, out6));
}
int b200xf_train_step_host(b200xf_ctx* ctx, const int64_t* inp, const int64_t* target, int B, int S, int T1, int workers, float* out6) {
  if (!ctx || !inp || !target || !out6) return set_error("null argument");
  B200XF_SCOPED(ctx, ctx->c.step_host(reinterpret_cast<const long long*>(inp), reinterpret_cast<const long long*>(target), B, S, T1, workers, out6));
}
int b200xf_eval_step_device(b200xf_ctx* ctx, const int64_t* inp, const int64_t* target, int B, int S, int T1, float* out6) {
  if (!ctx || !inp || !target) return set_error("null argument");
  B200XF_SCOPED(ctx, ctx->c.step_device(reinterpret_cast<const long long*>(inp), reinterpret_cast<const long long*>(target), B, S, T1, 1, 0, out6));
}
int b200xf_forward_backward_device(b200xf_ctx* ctx, const int64_t* inp, const int64_t* target, int B, int S, int T1, int workers, float* out6) {
  if (!ctx || !inp || !target) return set_error("null argument");
  B200XF_SCOPED(ctx, ctx->c.step_device(reinterpret_cast<const long long*>(inp), reinterpret_cast<const long long*>(target), B, S, T1, workers, 1, out6));
}
int b200xf_set_dropout_seed(b200xf_ctx* ctx, uint64_t seed) {
  if (!ctx) return set_error("null ctx");
  ctx->c.dropout_seed = seed;
  ctx->c.drop_graphs();  // captured steps carry the old key
  return 0;
}
int b200xf_dropout_mask(b200xf_ctx* ctx, int site, int rows, int cols, float* out) {
  if (!ctx || !out) return set_error("null argument");
  Context& c = ctx->c;
  const bool saved = c.dropout_on;
  c.dropout_on = c.hp.droput_rate > 0.f;
  const DropArgs d = c.site_drop(site);
  c.dropout_on = saved;
  if (!d.thresh) return set_error("dropout is disabled (rate 0)");
  B200XF_SCOPED(ctx, dropout_mask_fill(out, rows, cols, d, c.stream));
}
int b200xf_decode_begin(b200xf_ctx* ctx, const int64_t* inp, int B, int S, int max_target_length) {
  if (!ctx || !inp) return set_error("null argument");
  B200XF_SCOPED(ctx, ctx->c.decode_begin(reinterpret_cast<const long long*>(inp), B, S, max_target_length));
}
int b200xf_decode_step(b200xf_ctx* ctx, const int64_t* tokens_t, int t, float* logits_out, int64_t* next_tokens) {
  if (!ctx || !tokens_t) return set_error("null argument");
  B200XF_SCOPED(ctx, ctx->c.decode_step(reinterpret_cast<const long long*>(tokens_t), t, logits_out, reinterpret_cast<long long*>(next_tokens)));
}
int b200xf_attention_weights(b200xf_ctx* ctx, int decoder_layer, int block, float* out) {
  if (!ctx || !out) return set_error("null argument");
  B200XF_SCOPED(ctx, ctx->c.attention_weights(decoder_layer, block, out));
}
int b200xf_argmax_rows(const float* x, int ld, int rows, int cols, int64_t* out, void* stream) {
  if (!x || !out) return set_error("null argument");
  return argmax_rows(x, ld, rows, cols, reinterpret_cast<long long*>(out), static_cast<cudaStream_t>(stream));
}
int b200xf_assemble_padded(const int64_t* flat_tokens, const int64_t* offsets, int first_row, int rows, int length, int64_t* out,
                           int* too_long, void* stream) {
  if (!flat_tokens || !offsets || !out || !too_long) return set_error("null argument");
  return assemble_padded(reinterpret_cast<const long long*>(flat_tokens), reinterpret_cast<const long long*>(offsets), first_row, rows,
                         length, reinterpret_cast<long long*>(out), too_long, static_cast<cudaStream_t>(stream));
}
int b200xf_synchronize(b200xf_ctx* ctx) {
  if (!ctx) return set_error("null ctx");
  B200XF_CUDA(cudaStreamSynchronize(ctx->c.stream));
  B200XF_CUDA(cudaStreamSynchronize(ctx->c.comm_stream));
  B200XF_CUDA(cudaStreamSynchronize(ctx->c.upd_stream));
  B200XF_CUDA(cudaStreamSynchronize(ctx->c.side_stream));
  B200XF_CUDA(cudaStreamSynchronize(ctx->c.user_stream));
  return 0;
}

int b200xf_comm_unique_id(void* id) { return id ? comm_unique_id(id) : set_error("null argument"); }
int b200xf_comm_init(b200xf_ctx* ctx, const void* id, int rank, int world) {
  if (!ctx || !id) return set_error("null argument");
  return ctx->c.comm_init(id, rank, world);
}
int b200xf_comm_destroy(b200xf_ctx* ctx) { return ctx ? ctx->c.comm_destroy() : set_error("null ctx"); }
int b200xf_use_external_state(b200xf_ctx* ctx, float* params, float* grads) {
  if (!ctx) return set_error("null ctx");
  return ctx->c.use_external_state(params, grads);
}
int b200xf_nvls_attach(b200xf_ctx* ctx, void* grads_multicast, void* params_multicast, void* const* flag_pads, void* const* grads_peers,
                       void* const* params_peers, int point_to_point, int rank, int world) {
  if (!ctx) return set_error("null ctx");
  return ctx->c.nvls_attach(grads_multicast, params_multicast, flag_pads, grads_peers, params_peers, point_to_point, rank, world);
}
int b200xf_debug_nvls_update(b200xf_ctx* ctx, int64_t begin, int64_t count, int ctas, int refresh) {
  if (!ctx || !ctx->c.nvls_on) return set_error("debug_nvls_update: the NVLS update is not attached");
  Context& c = ctx->c;
  if (begin < 0 || count <= 0 || size_t(begin + count) > c.total_params || count % (int64_t(c.world) * 64) || begin % 64) return set_error("debug_nvls_update: bad range");
  const size_t sh = size_t(count) / c.world, off = size_t(begin) + size_t(c.rank) * sh;
  B200XF_TRY(c.enter());
  B200XF_TRY(nvls_reduce_adam_broadcast(c.nvls, c.params, c.adam_m, c.adam_v, off, sh, c.iterations_dev, c.adam, c.nvls_words, c.nvls_words + 1, ctas, c.stream));
  if (refresh) B200XF_TRY(refresh_shadow_zero_grad(c.params + begin, c.shadow + begin, c.grads + begin, size_t(count), c.stream));
  return c.leave();
}
int b200xf_nvls_error(b200xf_ctx* ctx) {
  if (!ctx || !ctx->c.nvls_on) return 0;
  unsigned v = 0;
  if (cudaMemcpy(&v, ctx->c.nvls.pad_local + kNvlsError, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  return int(v);
}
int b200xf_gather_optimizer_state(b200xf_ctx* ctx) {
  if (!ctx) return set_error("null ctx");
  B200XF_SCOPED(ctx, ctx->c.gather_optimizer_state());
}
int b200xf_optimizer_state_complete(b200xf_ctx* ctx) { return (ctx && ctx->c.opt_state_complete) ? 1 : 0; }
int b200xf_allreduce_bucket(b200xf_ctx* ctx, float* ptr, int64_t count) {
  if (!ctx || !ptr) return set_error("null argument");
  B200XF_TRY(ctx->c.enter());
  B200XF_CUDA(cudaEventRecord(ctx->c.ev_ready, ctx->c.stream));
  B200XF_CUDA(cudaStreamWaitEvent(ctx->c.comm_stream, ctx->c.ev_ready, 0));
  B200XF_TRY(ctx->c.allreduce(ptr, size_t(count), ctx->c.comm_stream));
  B200XF_CUDA(cudaEventRecord(ctx->c.ev_done, ctx->c.comm_stream));
  B200XF_CUDA(cudaStreamWaitEvent(ctx->c.stream, ctx->c.ev_done, 0));
  return ctx->c.leave();
}

// ---------------------------------------------------------------- per-kernel entry points
int b200xf_gemm_bf16(const void* a, int lda, int a_mn, const void* b, int ldb, int b_mn, void* d, int ldd, int d_fp32, int accumulate,
                     int M, int N, int K, const float* bias, int relu, const void* resid, int ldr, int resid_op, float alpha,
                     int force_bn, int split_k, void* stream) {
  GemmArgs g;
  g.a = a; g.lda = lda; g.a_mn = a_mn; g.b = b; g.ldb = ldb; g.b_mn = b_mn; g.d = d; g.ldd = ldd; g.d_fp32 = d_fp32;
  g.accumulate = accumulate; g.M = M; g.N = N; g.K = K; g.bias = bias; g.relu = relu; g.resid = resid; g.ldr = ldr;
  g.resid_op = resid_op; g.alpha = alpha; g.force_bn = force_bn; g.split_k = split_k;
  return gemm_bf16(g, static_cast<cudaStream_t>(stream));
}

int b200xf_dense_layernorm_fwd(const void* x, int ldx, const void* w, int ldw, const float* bias, const void* resid, int ldr, void* z, int ldz,
                               const float* gamma, const float* beta, void* y, int ldy, float* mean, float* rstd, int rows, int d_out, int d_in,
                               float eps, void* stream) {
  if (!x || !w || !z || !y || !gamma || !beta || !mean || !rstd) return set_error("dense_layernorm_fwd: null pointer");
  if (!gemm_ln_fusable(rows, d_out)) return set_error("dense_layernorm_fwd: width %d unsupported (multiple of 8, <= 2048)", d_out);
  GemmArgs g;
  g.a = x; g.lda = ldx; g.a_mn = 0; g.b = w; g.ldb = ldw; g.b_mn = 1; g.d = z; g.ldd = ldz;
  g.M = rows; g.N = d_out; g.K = d_in; g.bias = bias; g.resid = resid; g.ldr = ldr; g.resid_op = resid ? 1 : 0;
  g.ln_gamma = gamma; g.ln_beta = beta; g.ln_y = y; g.ln_ldy = ldy; g.ln_mean = mean; g.ln_rstd = rstd; g.ln_eps = eps;
  return gemm_bf16(g, static_cast<cudaStream_t>(stream));
}

int b200xf_dense_wgrad_bias(const void* x, int ldx, const void* dy, int lddy, float* dw, int lddw, float* dbias, float* scratch, size_t scratch_floats,
                            int tokens, int d_in, int d_out, void* stream) {
  if (!x || !dy || !dw || !dbias || !scratch) return set_error("dense_wgrad_bias: null pointer");
  GemmArgs g;
  g.a = x; g.lda = ldx; g.a_mn = 1; g.b = dy; g.ldb = lddy; g.b_mn = 1; g.d = dw; g.ldd = lddw; g.d_fp32 = 1; g.accumulate = 1;
  g.M = d_in; g.N = d_out; g.K = tokens;
  GemmInfo info;
  B200XF_TRY(gemm_plan(g, &info));
  const int rows = info.splits * info.tiles_m;
  if (size_t(rows) * d_out > scratch_floats) return set_error("dense_wgrad_bias: scratch of %zu floats needed", size_t(rows) * d_out);
  g.colsum_partials = scratch;
  B200XF_TRY(gemm_bf16(g, static_cast<cudaStream_t>(stream)));
  ReduceJob job{scratch, dbias, rows, d_out, d_out};
  return reduce_partials(&job, 1, static_cast<cudaStream_t>(stream));
}

int b200xf_debug_gemm_trace(void* device_u64_buffer) {
  gemm_set_trace(static_cast<unsigned long long*>(device_u64_buffer));
  return 0;
}

int b200xf_gemm_set_pair_mode(int mode) {
  if (mode < -1 || mode > 1) return set_error("gemm pair mode must be -1 (heuristic), 0 (never) or 1 (always)");
  gemm_set_pair_mode(mode);
  return 0;
}

int b200xf_debug_gemm_trace_ring(void* device_u64_buffer, int slots) {
  if (device_u64_buffer && slots <= 0) return set_error("gemm trace ring: slots must be positive");
  gemm_set_trace(static_cast<unsigned long long*>(device_u64_buffer), slots);
  return 0;
}

int b200xf_attention_fwd(const void* q, int ldq, const void* k, int ldk, const void* v, int ldv, void* o, int ldo, float* lse,
                         const int64_t* key_tokens, int causal, int B, int H, int Sq, int Sk, int dk, void* stream) {
  AttnArgs a;
  a.q = static_cast<const bf16*>(q); a.ldq = ldq; a.k = static_cast<const bf16*>(k); a.ldk = ldk; a.v = static_cast<const bf16*>(v); a.ldv = ldv;
  a.o = static_cast<bf16*>(o); a.ldo = ldo; a.lse = lse; a.key_tokens = reinterpret_cast<const long long*>(key_tokens); a.causal = causal;
  a.B = B; a.H = H; a.Sq = Sq; a.Sk = Sk; a.dk = dk;
  return attention_fwd(a, static_cast<cudaStream_t>(stream));
}
int b200xf_attention_bwd(const void* q, int ldq, const void* k, int ldk, const void* v, int ldv, const void* o, int ldo, const float* lse,
                         const int64_t* key_tokens, int causal, int B, int H, int Sq, int Sk, int dk, const void* d_o, int lddo,
                         void* dq, int lddq, void* dk_out, int lddk, void* dv_out, int lddv, float* delta_scratch, float* dq_acc_scratch,
                         void* stream) {
  AttnArgs a;
  a.q = static_cast<const bf16*>(q); a.ldq = ldq; a.k = static_cast<const bf16*>(k); a.ldk = ldk; a.v = static_cast<const bf16*>(v); a.ldv = ldv;
  a.o = const_cast<bf16*>(static_cast<const bf16*>(o)); a.ldo = ldo; a.lse = const_cast<float*>(lse);
  a.key_tokens = reinterpret_cast<const long long*>(key_tokens); a.causal = causal;
  a.B = B; a.H = H; a.Sq = Sq; a.Sk = Sk; a.dk = dk;
  a.d_o = static_cast<const bf16*>(d_o); a.lddo = lddo; a.dq = static_cast<bf16*>(dq); a.lddq = lddq;
  a.dk_out = static_cast<bf16*>(dk_out); a.lddk = lddk; a.dv_out = static_cast<bf16*>(dv_out); a.lddv = lddv; a.delta = delta_scratch; a.dq_acc = dq_acc_scratch;
  return attention_bwd(a, static_cast<cudaStream_t>(stream));
}

int b200xf_layernorm_residual_fwd(const void* x, const void* resid, const float* gamma, const float* beta, void* y, void* z_out,
                                  float* mean, float* rstd, int rows, int d, float eps, void* stream) {
  return layernorm_fwd(static_cast<const bf16*>(x), static_cast<const bf16*>(resid), gamma, beta, static_cast<bf16*>(y),
                       static_cast<bf16*>(z_out), mean, rstd, rows, d, eps, static_cast<cudaStream_t>(stream));
}
size_t b200xf_layernorm_bwd_scratch_bytes(int rows, int d) { return ln_bwd_scratch_floats(rows, d) * sizeof(float); }
int b200xf_layernorm_residual_bwd(const void* dy, const void* z, const float* mean, const float* rstd, const float* gamma, void* dz,
                                  float* dgamma, float* dbeta, float* dcol, int rows, int d, void* scratch, void* stream) {
  if (!scratch) return set_error("layernorm_bwd: scratch is null");
  return layernorm_bwd(static_cast<const bf16*>(dy), static_cast<const bf16*>(z), mean, rstd, gamma, static_cast<bf16*>(dz), dgamma,
                       dbeta, dcol, rows, d, static_cast<float*>(scratch), static_cast<cudaStream_t>(stream));
}

int b200xf_embed_pe_fwd(const int64_t* tokens, const float* table, int vocab, const float* pe, void* out, int B, int S, int d, void* stream) {
  return embed_pe_fwd(reinterpret_cast<const long long*>(tokens), table, vocab, pe, static_cast<bf16*>(out), B, S, d,
                      static_cast<cudaStream_t>(stream));
}
size_t b200xf_embed_bwd_scratch_bytes(int n_tokens) { return embed_bwd_scratch_bytes(n_tokens); }
int b200xf_embed_pe_bwd(const int64_t* tokens, const void* dx, float* dtable, int vocab, int n_tokens, int d, void* scratch,
                        size_t scratch_bytes, void* stream) {
  return embed_bwd(reinterpret_cast<const long long*>(tokens), static_cast<const bf16*>(dx), dtable, vocab, n_tokens, d, scratch,
                   scratch_bytes, static_cast<cudaStream_t>(stream));
}

namespace {
__global__ void count_nonpad_kernel(const long long* labels, int n, int* out) {
  __shared__ int wc[32];
  int c = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) c += (labels[i] != 0);
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) wc[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) { int s = 0; for (int w = 0; w < (blockDim.x >> 5); ++w) s += wc[w]; *out = s; }
}
}  // namespace

int b200xf_vocab_ce_fwd_bwd(void* logits, int ldl, const int64_t* labels, int rows, int V, int workers_count, int write_grad,
                            float* out3, void* scratch, void* stream) {
  if (!scratch || !out3 || workers_count <= 0) return set_error("vocab_ce: bad argument");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  float* row_loss = static_cast<float*>(scratch);
  int* row_correct = reinterpret_cast<int*>(row_loss + rows);
  int* count = row_correct + rows;
  const long long* lab = reinterpret_cast<const long long*>(labels);
  count_nonpad_kernel<<<1, 1024, 0, s>>>(lab, rows, count);
  count_launch();
  const float inv_w = 1.0f / float(workers_count);
  B200XF_TRY(vocab_ce_fwd_bwd(static_cast<bf16*>(logits), ldl, lab, rows, V, count, inv_w, row_loss, row_correct, write_grad, s));
  return ce_finalize(row_loss, row_correct, count, rows, inv_w, out3, s);
}

int b200xf_vocab_ce_f32(const float* logits, int ldl, const int64_t* labels, int rows, int V, int workers_count, float* out3, void* scratch,
                        void* stream) {
  if (!logits || !labels || !scratch || !out3 || workers_count <= 0) return set_error("vocab_ce_f32: bad argument");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  float* row_loss = static_cast<float*>(scratch);
  int* row_correct = reinterpret_cast<int*>(row_loss + rows);
  int* count = row_correct + rows;
  const long long* lab = reinterpret_cast<const long long*>(labels);
  count_nonpad_kernel<<<1, 1024, 0, s>>>(lab, rows, count);
  count_launch();
  B200XF_TRY(vocab_ce_f32(logits, ldl, lab, rows, V, row_loss, row_correct, s));
  return ce_finalize(row_loss, row_correct, count, rows, 1.0f / float(workers_count), out3, s);
}

int b200xf_adam_step(float* w, float* g, float* m, float* v, void* w_bf16, int64_t n, const int64_t* iterations_dev, int embedding_size,
                     int warmup_steps, float beta_1, float beta_2, float epsilon, int zero_grad, void* stream) {
  AdamHyper h;
  h.beta1 = beta_1; h.beta2 = beta_2; h.eps = epsilon;
  h.rsqrt_d_model = 1.0f / sqrtf(float(embedding_size));
  h.warmup_pow_m15 = float(pow(double(warmup_steps), -1.5));
  return adam_step(w, g, m, v, static_cast<bf16*>(w_bf16), size_t(n), reinterpret_cast<const long long*>(iterations_dev), h, zero_grad,
                   static_cast<cudaStream_t>(stream));
}

// CRC-32C (Castagnoli), slicing-by-8, for the tensor-bundle checkpoint files (host only).
uint32_t b200xf_crc32c(const void* data, size_t n, uint32_t crc) {
  static uint32_t table[8][256];
  static bool ready = false;
  if (!ready) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1) ? 0x82F63B78u : 0u);
      table[0][i] = c;
    }
    for (int k = 1; k < 8; ++k)
      for (uint32_t i = 0; i < 256; ++i) table[k][i] = (table[k - 1][i] >> 8) ^ table[0][table[k - 1][i] & 0xFF];
    ready = true;
  }
  const unsigned char* p = static_cast<const unsigned char*>(data);
  uint32_t c = crc ^ 0xFFFFFFFFu;
  while (n >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4); memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = table[7][lo & 0xFF] ^ table[6][(lo >> 8) & 0xFF] ^ table[5][(lo >> 16) & 0xFF] ^ table[4][lo >> 24] ^
        table[3][hi & 0xFF] ^ table[2][(hi >> 8) & 0xFF] ^ table[1][(hi >> 16) & 0xFF] ^ table[0][hi >> 24];
    p += 8; n -= 8;
  }
  while (n--) c = table[0][(c ^ *p++) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}

int b200xf_positional_encoding_host(float* out, int positions, int d) {
  if (!out || positions <= 0 || d <= 0) return set_error("bad argument");
  positional_encoding_host(out, positions, d);
  return 0;
}

}  // extern "C"
