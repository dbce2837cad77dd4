/* b200xf — C ABI of the B200-native data-parallel Transformer train step.
 *
 * The reference (ddrakard/tensorflow-distributed-on-gke) has no FFI: its boundary for this path is the
 * Python call surface that distributed_training_transformer/__main__.py uses.  Each entry point below
 * names the reference code it stands in for (paths relative to
 * /root/reference/distributed_training_transformer/).  A Python façade with the reference's own names
 * sits on top of this ABI; a TF custom-op wrapper would bind the same symbols.
 *
 * Conventions: every function returns 0 on success and non-zero on failure, after which
 * b200xf_last_error() (thread-local) describes the problem.  Tokens are int64, row-major [batch, seq],
 * PAD = 0 (transformer_model.py:12,22,60).  Dense kernels are [in, out] (y = xW + b).  Device work runs on
 * streams the context owns; each call first orders them after the caller's stream (the one given to
 * b200xf_ctx_create) and then orders the caller's stream after them, so the caller sees plain stream
 * semantics.  Calls are asynchronous unless stated.  A context is not thread-safe;
 * distinct contexts are independent (one per GPU / process, as one TF worker per pod in the reference).
 */
#ifndef B200XF_H_
#define B200XF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200xf_ctx b200xf_ctx;

/* Constructor arguments of transformer_model.py:319-323 `Transformer(...)` plus the optimizer/schedule
 * constants of __main__.py:58,72-73. */
typedef struct b200xf_hparams {
  int32_t layers_count;
  int32_t embedding_size;
  int32_t heads_count;
  int32_t feed_forward_hidden_size;
  int32_t input_vocabulary_size;
  int32_t target_vocabulary_size;
  int32_t maximum_input_length;
  int32_t maximum_target_length;
  float droput_rate;      /* sic — the reference's keyword (transformer_model.py:323) */
  float adam_beta_1;      /* __main__.py:73 (0.9)  */
  float adam_beta_2;      /* __main__.py:73 (0.98) */
  float adam_epsilon;     /* __main__.py:73 (1e-9) */
  int32_t warmup_steps;   /* __main__.py:58 (4000) */
  int32_t max_tokens;     /* capacity: max(batch*seq) of any call, source or target side */
} b200xf_hparams;

const char* b200xf_last_error(void);
int b200xf_version(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
long long b200xf_launch_count(void);

/* ---- model state: Transformer.__init__ + build_model (transformer_model.py:318-333,366-371) ---------- */
int b200xf_ctx_create(const b200xf_hparams* hp, int device, void* cuda_stream, b200xf_ctx** out);
int b200xf_ctx_destroy(b200xf_ctx* ctx);

/* Device memory a context with these hyper-parameters allocates (nothing is allocated by this call): state_bytes = 18 bytes
 * per parameter (fp32 weights, gradients, Adam m and v, bf16 operand copy), workspace_bytes = the activation / scratch arena
 * for max_tokens tokens per side.  The memory model behind the batch size of the long-sequence configuration. */
int b200xf_device_bytes(const b200xf_hparams* hparams, int64_t* state_bytes, int64_t* workspace_bytes);

/* model.trainable_variables / load_weights / save_weights (__main__.py:87,112-114; checkpoint.py:86):
 * the variables live in one flat fp32 device buffer; entry i is a [rows, cols] view (cols == 1-D length
 * for vectors, rows == 1) at element `offset` with row pitch `ld`.  Names are the Keras object-graph keys
 * of saved_weights/<n>/model_weights.index.  grad / adam_m / adam_v buffers share the same layout. */
int b200xf_param_count(b200xf_ctx* ctx);
int b200xf_param_info(b200xf_ctx* ctx, int index, char* name_buf, size_t name_buf_len, int64_t* rows, int64_t* cols,
                      int64_t* ld, int64_t* offset, int32_t* ndim);
int64_t b200xf_param_total(b200xf_ctx* ctx);
float* b200xf_param_ptr(b200xf_ctx* ctx);
float* b200xf_grad_ptr(b200xf_ctx* ctx);
float* b200xf_adam_m_ptr(b200xf_ctx* ctx);
float* b200xf_adam_v_ptr(b200xf_ctx* ctx);
/* Must be called after the host writes into param_ptr (refreshes the bf16 operand copy the GEMMs read). */
int b200xf_params_changed(b200xf_ctx* ctx);
/* optimizer.iterations (__main__.py:72-73,113-114) */
int b200xf_set_iterations(b200xf_ctx* ctx, int64_t iterations);
int b200xf_get_iterations(b200xf_ctx* ctx, int64_t* iterations);

/* ---- Transformer.call (transformer_model.py:335-348); training != 0 enables dropout when droput_rate > 0 ------
 * inp [B,S], tgt_in [B,T] int64 DEVICE pointers; logits_out fp32 [B,T,V] DEVICE pointer or NULL. */
int b200xf_forward(b200xf_ctx* ctx, const int64_t* inp, const int64_t* tgt_in, int B, int S, int T, int training,
                   float* logits_out);

/* The second return value of Transformer.call (transformer_model.py:312-313,348): the softmax attention weights of
 * decoder layer `decoder_layer` (0-based), block 1 = look-ahead self-attention [B,H,T,T], block 2 = cross-attention
 * [B,H,T,S], fp32 DEVICE buffer, for the batch of the most recent b200xf_forward (whose token buffers must still be
 * alive).  tester.py:51-52 is the caller in the reference. */
int b200xf_attention_weights(b200xf_ctx* ctx, int decoder_layer, int block, float* out);
/* Incremental greedy decoding for the loop of tester.py:27-42.  The reference re-runs encoder and decoder over the whole prefix
 * for every generated token; here b200xf_decode_begin runs the encoder once (inp [B,S] int64 DEVICE, must stay alive) and
 * projects the cross-attention K|V of every decoder layer, and each b200xf_decode_step runs the decoder for ONE new target
 * position t (tokens_t [B] int64 DEVICE = the target tokens at position t, t = 0 being the start token), appending that
 * position's self-attention K|V to per-layer caches of max_target_length positions.  Outputs (either may be NULL): logits_out
 * fp32 [B, V] of position t — the row Transformer.call would return at [:, t, :] — and next_tokens [B] = their argmax
 * (lowest index among equal maxima, tester.py:38).  Steps must be issued for t = 0, 1, 2, ... in order. */
int b200xf_decode_begin(b200xf_ctx* ctx, const int64_t* inp, int B, int S, int max_target_length);
int b200xf_decode_step(b200xf_ctx* ctx, const int64_t* tokens_t, int t, float* logits_out, int64_t* next_tokens);
/* tf.argmax(x, axis=-1) over fp32 DEVICE rows (tester.py:38): out[r] = lowest index of the maximum of row r. */
int b200xf_argmax_rows(const float* x, int ld, int rows, int cols, int64_t* out, void* stream);

/* On-device batch assembly: the `.to_tensor()` that ends the reference's tokenisation (english_portugese_dataset.py:36-39)
 * pads the ragged sentences of a batch with 0 (PAD) to a common length, and experimental_distribute_dataset
 * (__main__.py:48-53) hands each replica a contiguous slice of the rows.  flat_tokens: the batch's sentences back to back
 * (DEVICE int64), offsets: [n_sentences + 1] DEVICE int64 (sentence r = flat[offsets[r], offsets[r+1])); writes rows
 * first_row .. first_row + rows - 1 into out [rows, length] (DEVICE int64).  *too_long (DEVICE int, zeroed by the caller)
 * receives the longest sentence length that did not fit `length` (such rows are truncated). */
int b200xf_assemble_padded(const int64_t* flat_tokens, const int64_t* offsets, int first_row, int rows, int length, int64_t* out,
                           int* too_long, void* stream);

/* ---- train_step (__main__.py:105-117): shift, forward, local_loss_function(.., workers_count), backward,
 * gradient all-reduce when a communicator is attached (cluster/cluster.py:66), Adam + schedule in place,
 * accuracy counts.  inp [B,S], target [B,T+1] int64.  out6 (fp32) = {local loss (already / workers_count),
 * accuracy numerator, accuracy denominator, then the same three SUMmed over replicas — what
 * strategy.reduce(SUM, ...) returns at __main__.py:131-132; equal to the first three without a communicator}.
 *   _device : DEVICE token pointers, DEVICE out6, fully asynchronous.
 *   _host   : HOST token pointers (pinned for async copies), HOST out6; copies in, runs, copies the six
 *             scalars back and synchronises the stream — the reference-facing call. */
int b200xf_train_step_device(b200xf_ctx* ctx, const int64_t* inp, const int64_t* target, int B, int S, int T_plus_1,
                             int workers_count, float* out6);
int b200xf_train_step_host(b200xf_ctx* ctx, const int64_t* inp, const int64_t* target, int B, int S, int T_plus_1,
                           int workers_count, float* out6);
/* test_step (__main__.py:119-125): forward + loss(workers_count=1) + accuracy, no update.  DEVICE pointers. */
int b200xf_eval_step_device(b200xf_ctx* ctx, const int64_t* inp, const int64_t* target, int B, int S, int T_plus_1,
                            float* out6);
/* Run only forward+loss+backward (no all-reduce, no Adam): gradients are left in grad_ptr. */
int b200xf_forward_backward_device(b200xf_ctx* ctx, const int64_t* inp, const int64_t* target, int B, int S, int T_plus_1,
                                   int workers_count, float* out6);
int b200xf_synchronize(b200xf_ctx* ctx);

/* Dropout(rate) (transformer_model.py:190-191,223-225,268,299; applied at :196,202,232,238,244,276,308) is inverted
 * dropout from a counter-based generator keyed by (seed, replica rank, site, optimizer.iterations, element index): nothing is stored,
 * the backward pass recomputes the same mask.  TensorFlow's random stream cannot be matched, so the checker reads the
 * mask itself: b200xf_dropout_mask writes, for dropout site `site` at the CURRENT iterations value, a DEVICE fp32
 * [rows, cols] array holding 1/(1-rate) where the element is kept and 0 where it is dropped.  Sites: 0 / 1 encoder /
 * decoder embedding; 2+2l, 3+2l encoder layer l after attention / FFN; 2+2L+3l, +1, +2 decoder layer l after
 * self-attention / cross-attention / FFN. */
int b200xf_set_dropout_seed(b200xf_ctx* ctx, uint64_t seed);
int b200xf_dropout_mask(b200xf_ctx* ctx, int site, int rows, int cols, float* out);

/* ---- MultiWorkerMirroredStrategy's implicit SUM all-reduce (cluster/cluster.py:66; __main__.py:113-114,
 * 131-132): NCCL over NVLink, one rank per process/GPU.  unique_id is the 128-byte ncclUniqueId obtained
 * from b200xf_comm_unique_id on rank 0 and broadcast by the caller. */
int b200xf_comm_unique_id(void* unique_id_128_bytes);
int b200xf_comm_init(b200xf_ctx* ctx, const void* unique_id_128_bytes, int rank, int world);
int b200xf_comm_destroy(b200xf_ctx* ctx);
/* With more than one rank a train step reduce-scatters each gradient bucket, applies Adam to the local 1/world shard and
 * all-gathers the updated weights (same wire bytes as an all-reduce, 1/world of the optimizer traffic): every rank ends
 * with identical fp32 weights, but adam_m / adam_v are then only maintained for the local shard of each bucket.
 * B200XF_SHARDED_UPDATE=0 selects all-reduce + replicated Adam (the reference's mirrored optimizer slots, cluster/cluster.py:66).
 * b200xf_gather_optimizer_state is the collective (every rank calls it) that completes adam_m / adam_v on all ranks by
 * all-gathering the shards — call it before reading the slots for a checkpoint or before detaching the communicator;
 * b200xf_optimizer_state_complete returns 1 when the slots are current everywhere (always with one rank), else 0.
 * Bucket boundaries follow from the layout and B200XF_BUCKET_MB, so keep both unchanged for the life of a training run.
 * Each replica draws its own dropout masks: the rank is part of the mask key (see b200xf_set_dropout_seed). */
int b200xf_gather_optimizer_state(b200xf_ctx* ctx);
int b200xf_optimizer_state_complete(b200xf_ctx* ctx);
/* The same exchange as ONE kernel over NVLink / NVSwitch (nvls_update.cu): each rank reads the SUM of every replica's gradients
 * of its shard with multimem.ld_reduce (the switch adds in flight), applies Adam, and writes the new weights into every
 * replica's memory with multimem.st, inbound and outbound traffic overlapping; NCCL is then only used for the three loss /
 * accuracy scalars.  It needs the flat fp32 weights and gradients in SYMMETRIC memory with a multicast mapping, which the host
 * side allocates (torch.distributed._symmetric_memory in training.py):
 *   b200xf_use_external_state  moves weights and gradients into caller-owned DEVICE buffers (256-byte aligned, param_total
 *                              floats each; contents are carried over; b200xf_param_ptr / b200xf_grad_ptr change);
 *   b200xf_nvls_attach         after b200xf_comm_init: the multicast addresses of those two buffers and, per rank, the address
 *                              of a zeroed symmetric flag pad of >= 256 bytes as mapped into this process; world <= 8.
 *                              point_to_point != 0 selects the variant of the same kernel that reads every replica's gradients and
 *                              writes every replica's weights through the per-rank peer mappings (grads_peers / params_peers, own
 *                              buffers at [rank]) instead of multimem: (W-1)/W bucket sizes per link direction instead of ~(1 + 1/W),
 *                              the better choice below 8 replicas; the multicast addresses may then be NULL;
 *   b200xf_nvls_error          non-zero once a replica failed to arrive at a cross-replica flag within ~10 s.
 * No counterpart in the reference beyond cluster/cluster.py:66: the collective is implicit there. */
int b200xf_use_external_state(b200xf_ctx* ctx, float* params, float* grads);
int b200xf_nvls_attach(b200xf_ctx* ctx, void* grads_multicast, void* params_multicast, void* const* flag_pads, void* const* grads_peers,
                       void* const* params_peers, int point_to_point, int rank, int world);
int b200xf_nvls_error(b200xf_ctx* ctx);
/* Debug / benchmark aid: one fused update of the flat range [begin, begin + count) (count a multiple of 64 * world) on `ctas`
 * CTAs, optionally followed by the bf16 refresh / gradient clear; every replica must call it with the same arguments. */
int b200xf_debug_nvls_update(b200xf_ctx* ctx, int64_t begin, int64_t count, int ctas, int refresh);
/* All-reduce (SUM, fp32, in place) of `count` floats at DEVICE pointer on the context's comm stream. */
int b200xf_allreduce_bucket(b200xf_ctx* ctx, float* device_ptr, int64_t count);

/* ---- per-kernel entry points (unit tests, ncu).  All pointers are DEVICE pointers, `stream` a cudaStream_t. */

/* tf.keras.layers.Dense matmul and its gradients (transformer_model.py:119-122,169-175,333):
 * D[M,N] = alpha * A . B^T-like product, bf16 inputs, fp32 accumulate on tcgen05 tensor cores.
 *   a_mn=0: A stored [M][K] (ld=lda); a_mn=1: stored [K][M].   b_mn=0: B stored [N][K]; b_mn=1: stored [K][N].
 *   bias fp32 [N] or NULL; relu!=0 applies max(0,.); resid bf16 [M][ldr] with resid_op 1 (add) or 2 (zero
 *   where resid<=0); d_fp32!=0 writes fp32; accumulate!=0 atomically adds into fp32 D (split-K allowed). */
int b200xf_gemm_bf16(const void* a, int lda, int a_mn, const void* b, int ldb, int b_mn, void* d, int ldd, int d_fp32,
                     int accumulate, int M, int N, int K, const float* bias, int relu, const void* resid, int ldr,
                     int resid_op, float alpha, int force_bn, int split_k, void* stream);

/* One sub-layer tail of EncoderLayer / DecoderLayer.call (transformer_model.py:196-198,202-204,232-233,238-240,244-246):
 *   z = resid + (x W + bias)   (bf16, kept for the backward pass),   y = LayerNormalization(epsilon)(z)
 * as ONE tcgen05 GEMM whose epilogue normalises: the CTAs holding a 128-row block's column tiles form a thread-block
 * cluster and exchange per-row partial statistics through distributed shared memory.  x [rows, d_in] bf16, w [d_in, d_out]
 * bf16 (Keras kernel layout), z / y [rows, d_out] bf16, mean / rstd fp32 [rows]; d_out a multiple of 8, at most 2048. */
int b200xf_dense_layernorm_fwd(const void* x, int ldx, const void* w, int ldw, const float* bias, const void* resid, int ldr,
                               void* z, int ldz, const float* gamma, const float* beta, void* y, int ldy, float* mean,
                               float* rstd, int rows, int d_out, int d_in, float eps, void* stream);
/* Both parameter gradients of a Dense layer (tape.gradient at __main__.py:112) from one pass over dy:
 *   dw[d_in, d_out] += x^T dy (fp32, accumulated),   dbias[d_out] = sum over tokens of dy
 * the column sums are taken from the dy tiles the weight-gradient GEMM stages in shared memory anyway.  scratch: fp32
 * DEVICE buffer of at least 148 * d_out floats (one partial row per K split and row block of the weight). */
int b200xf_dense_wgrad_bias(const void* x, int ldx, const void* dy, int lddy, float* dw, int lddw, float* dbias, float* scratch,
                            size_t scratch_floats, int tokens, int d_in, int d_out, void* stream);

/* Debug aid (no counterpart in the reference): when given a DEVICE buffer of 8 * 148 uint64, every later GEMM launch
 * records %globaltimer stamps per CTA (start, set-up done, first operands landed, last MMA issued, accumulator ready,
 * epilogue done, exit).  Pass NULL to switch it off. */
int b200xf_debug_gemm_trace(void* device_u64_buffer);
/* CTA pairs (tcgen05 cta_group::2: the two SMs of a TPC share one 256-row tile and each stage only half of the B tile).
 * mode -1 = heuristic (default: wide tiles with at least three waves of work), 0 = never, 1 = whenever M > 128 and the
 * tile is at least 128 columns wide.  Process-wide; tests use it to run every shape through both kernels.
 * No counterpart in the reference (tuning / test hook). */
int b200xf_gemm_set_pair_mode(int mode);
/* Same, for a whole step: the buffer holds `slots` x 148 x 8 u64 and successive launches use successive slots (wrapping).
 * Stamp 7 of a slot's first CTA is (M << 44 | N << 24 | K << 4 | a_mn << 1 | b_mn).  Launches captured into a CUDA graph
 * keep their slot, so replaying the step refills the ring.  No counterpart in the reference (debugging aid). */
int b200xf_debug_gemm_trace_ring(void* device_u64_buffer, int slots);

/* scaled_dot_product_attention + masks + split/merge heads (transformer_model.py:56-70,73-109,124-131,160-163).
 * q [B*Sq, ldq], k/v [B*Sk, ld*], o [B*Sq, ldo] bf16, head h in columns [h*dk,(h+1)*dk); key_tokens [B,Sk] int64;
 * lse fp32 [B,H,Sq,2] = {row max, log2(row sum)} of the scaled, masked scores in the log2 domain.
 * Backward scratch: delta_scratch fp32 [B,H,Sq]; dq_acc_scratch fp32 [B*Sq, H*dk] or NULL — with it, head size 64 and
 * sequences longer than one 128-key tile run the tcgen05 multi-tile backward (dQ accumulated by TMA reduce-add), without
 * it the mma.sync kernels. */
int b200xf_attention_fwd(const void* q, int ldq, const void* k, int ldk, const void* v, int ldv, void* o, int ldo,
                         float* lse, const int64_t* key_tokens, int causal, int B, int H, int Sq, int Sk, int dk,
                         void* stream);
int b200xf_attention_bwd(const void* q, int ldq, const void* k, int ldk, const void* v, int ldv, const void* o, int ldo,
                         const float* lse, const int64_t* key_tokens, int causal, int B, int H, int Sq, int Sk, int dk,
                         const void* d_o, int lddo, void* dq, int lddq, void* dk_out, int lddk, void* dv_out, int lddv,
                         float* delta_scratch, float* dq_acc_scratch, void* stream);

/* LayerNormalization(epsilon) of (x + resid) (transformer_model.py:187-188,198,204).  bf16 rows of length d.
 * z_out (optional) receives x+resid; mean/rstd fp32 [rows]. */
int b200xf_layernorm_residual_fwd(const void* x, const void* resid, const float* gamma, const float* beta, void* y,
                                  void* z_out, float* mean, float* rstd, int rows, int d, float eps, void* stream);
/* Backward.  dcol (optional) = column sums of dz (the bias gradient of the Dense that produced x).  scratch:
 * b200xf_layernorm_bwd_scratch_bytes(rows, d) bytes of device memory; it receives per-block partial column sums that a
 * second launch adds up in a fixed order (deterministic); its previous contents do not matter. */
size_t b200xf_layernorm_bwd_scratch_bytes(int rows, int d);
int b200xf_layernorm_residual_bwd(const void* dy, const void* z, const float* mean, const float* rstd, const float* gamma,
                                  void* dz, float* dgamma, float* dbeta, float* dcol, int rows, int d, void* scratch,
                                  void* stream);

/* LayerNorm backward for callers that cannot allocate device memory (no counterpart in the reference).
 * Identical results to b200xf_layernorm_residual_bwd, but the partial sums live in a fixed 3 637 248-byte array inside the
 * library image instead of caller scratch, and the call allocates nothing.  Meant for an op hosted by a framework whose
 * memory pool already owns the GPU (TensorFlow's default), or a loop that must not hit cudaMalloc's device-wide sync.
 * Refused (error, nothing launched) when b200xf_layernorm_bwd_scratch_bytes(rows, d) exceeds that size (it holds up to
 * 9 472 rows at d = 512, 2 368 rows at d = 1024).  The first call on a device resolves the array's address: make it before any stream capture, with that
 * device current.  Calls share the array, so calls that overlap on the GPU race; nothing in it needs clearing. */
int b200xf_layernorm_residual_bwd_noalloc(const void* dy, const void* z, const float* mean, const float* rstd,
                                          const float* gamma, void* dz, float* dgamma, float* dbeta, float* dcol, int rows,
                                          int d, void* stream);

/* Embedding lookup * sqrt(d) + positional encoding (transformer_model.py:273-275,305-307) and its gradient. */
int b200xf_embed_pe_fwd(const int64_t* tokens, const float* table, int vocab, const float* pe, void* out, int B, int S,
                        int d, void* stream);
size_t b200xf_embed_bwd_scratch_bytes(int n_tokens);
int b200xf_embed_pe_bwd(const int64_t* tokens, const void* dx, float* dtable, int vocab, int n_tokens, int d,
                        void* scratch, size_t scratch_bytes, void* stream);

/* loss_object + local_loss_function + accuracy_function (transformer_model.py:7-26) over bf16 logits
 * [rows, ldl]; overwrites logits with d(loss)/d(logits) when write_grad != 0.  out3 = {loss, acc_num, acc_den}.
 * scratch: rows * 8 + 16 bytes. */
int b200xf_vocab_ce_fwd_bwd(void* logits, int ldl, const int64_t* labels, int rows, int V, int workers_count,
                            int write_grad, float* out3, void* scratch, void* stream);

/* local_loss_function / accuracy_function (transformer_model.py:11-26) on fp32 logits as Transformer.call returns them:
 * PAD-masked sparse softmax cross-entropy / (non-pad count * workers_count) and the masked argmax accuracy counts, computed
 * in fp32 from the fp32 logits (nothing is rounded to bf16), no gradient.  out3 = {loss, accuracy numerator, denominator};
 * scratch: (2 * rows + 64) * 4 bytes. */
int b200xf_vocab_ce_f32(const float* logits, int ldl, const int64_t* labels, int rows, int V, int workers_count, float* out3,
                        void* scratch, void* stream);

/* tf.keras.optimizers.Adam(TransformerLearningRateSchedule(d), beta_1, beta_2, epsilon) applied to a flat range
 * (__main__.py:56-73,113-114).  iterations_dev: DEVICE int64 holding optimizer.iterations before the step. */
int b200xf_adam_step(float* w, float* g, float* m, float* v, void* w_bf16_or_null, int64_t n,
                     const int64_t* iterations_dev, int embedding_size, int warmup_steps, float beta_1, float beta_2,
                     float epsilon, int zero_grad, void* stream);

/* CRC-32C of a host buffer (continue from `crc`, 0 to start): the per-tensor and per-block checksums of the TF
 * tensor-bundle files that model.save_weights / load_weights exchange (checkpoint.py:86, __main__.py:87). */
uint32_t b200xf_crc32c(const void* data, size_t n, uint32_t crc);

/* positional_encoding (transformer_model.py:29-53) computed on the host in fp32: out [positions, d]. */
int b200xf_positional_encoding_host(float* out, int positions, int d);

#ifdef __cplusplus
}
#endif
#endif /* B200XF_H_ */
