// Model state + one-step orchestration: the C++ counterpart of Transformer.__init__/call
// (transformer_model.py:318-348) and train_step (__main__.py:105-117) over the CUDA kernels.
//
// HBM layout.  All trainable variables live in ONE flat fp32 buffer (with identically laid out grad /
// Adam-m / Adam-v buffers and a bf16 operand copy), ordered in reverse forward order so that backward
// completes contiguous ranges from the front: final_layer, decoder layers L-1..0, decoder embedding,
// the cross-attention K|V projections of ALL decoder layers fused into one [d, L*2d] matrix, encoder
// layers L-1..0, encoder embedding.  Q|K|V of each self-attention is one fused [d, 3d] matrix.  The
// reference's 172 logical variables (saved_weights/*/model_weights.index) are exposed as strided views.
#include "context.h"

#include <dlfcn.h>
#include <nccl.h>

#include <cmath>
#include <cstdio>
#include <cstring>

namespace b200xf {

namespace {
size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr size_t kAlignElems = 64;  // every tensor starts on a 256-byte boundary (fp32)
}  // namespace

// ------------------------------------------------------------------------------------ NCCL (dlopen)
struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.handle ? &api : nullptr;
  tried = true;
  const char* env = getenv("B200XF_NCCL_LIB");
  const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    if (!n) continue;
    api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.handle) break;
  }
  if (!api.handle) return nullptr;
#define LOAD(field, sym) api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, sym))
  LOAD(GetUniqueId, "ncclGetUniqueId"); LOAD(CommInitRank, "ncclCommInitRank"); LOAD(CommDestroy, "ncclCommDestroy");
  LOAD(AllReduce, "ncclAllReduce"); LOAD(GroupStart, "ncclGroupStart"); LOAD(GroupEnd, "ncclGroupEnd");
  LOAD(GetErrorString, "ncclGetErrorString"); LOAD(ReduceScatter, "ncclReduceScatter"); LOAD(AllGather, "ncclAllGather");
#undef LOAD
  if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy) { api.handle = nullptr; return nullptr; }
  return &api;
}

#define B200XF_NCCL(expr)                                                                               \
  do {                                                                                                  \
    ncclResult_t _r = (expr);                                                                           \
    if (_r != ncclSuccess) return set_error("%s: %s", #expr, nccl_api()->GetErrorString ? nccl_api()->GetErrorString(_r) : "nccl error"); \
  } while (0)

int comm_unique_id(void* out128) {
  NcclApi* api = nccl_api();
  if (!api) return set_error("NCCL library not found (set B200XF_NCCL_LIB)");
  ncclUniqueId id;
  B200XF_NCCL(api->GetUniqueId(&id));
  memcpy(out128, &id, sizeof(id));
  return 0;
}

// ------------------------------------------------------------------------------------ layout
size_t Context::add_tensor(size_t& cursor, size_t elems) {
  const size_t off = cursor;
  cursor = align_up(cursor + elems, kAlignElems);
  return off;
}

Dense Context::add_dense(size_t& cursor, int K, int N, int ld) {
  Dense p; p.K = K; p.N = N; p.ld = ld;
  p.w = add_tensor(cursor, size_t(K) * ld);
  p.b = add_tensor(cursor, size_t(ld));
  return p;
}

void Context::add_view(const std::string& name, size_t offset, int64_t rows, int64_t cols, int64_t ld, int ndim) {
  views[name] = ParamView{name, rows, cols, ld, int64_t(offset), ndim};
}

void Context::build_layout() {
  const int L = hp.layers_count, d = hp.embedding_size, dff = hp.feed_forward_hidden_size;
  const int Vin = hp.input_vocabulary_size, Vt = hp.target_vocabulary_size;
  vt_ld = int(align_up(size_t(Vt), 8));
  size_t cur = 0;
  sections.clear();
  // every section ends on a 4096-element boundary, so any bucket of sections splits into equal 64-element-aligned
  // shards for up to 64 ranks (reduce-scatter / all-gather of the sharded update)
  auto close_section = [&](size_t begin) { cur = align_up(cur, 4096); sections.push_back({begin, cur}); };

  size_t begin = cur;
  final_layer = add_dense(cur, d, Vt, vt_ld);
  add_view("final_layer/kernel", final_layer.w, d, Vt, vt_ld, 2);
  add_view("final_layer/bias", final_layer.b, 1, Vt, vt_ld, 1);
  close_section(begin);

  dec.resize(L); enc.resize(L);
  auto ln = [&](LnP& p, const std::string& prefix) {
    p.g = add_tensor(cur, d); p.b = add_tensor(cur, d);
    add_view(prefix + "/gamma", p.g, 1, d, d, 1); add_view(prefix + "/beta", p.b, 1, d, d, 1);
  };
  auto dense_views = [&](const Dense& p, const std::string& prefix) {
    add_view(prefix + "/kernel", p.w, p.K, p.N, p.ld, 2); add_view(prefix + "/bias", p.b, 1, p.N, p.ld, 1);
  };
  auto qkv_views = [&](const Dense& p, const std::string& prefix) {
    const char* names[3] = {"query_generator_weights", "key_generator_weights", "value_generator_weights"};
    for (int i = 0; i < 3; ++i) {
      add_view(prefix + "/" + names[i] + "/kernel", p.w + size_t(i) * d, d, d, p.ld, 2);
      add_view(prefix + "/" + names[i] + "/bias", p.b + size_t(i) * d, 1, d, d, 1);
    }
  };
  for (int l = L - 1; l >= 0; --l) {
    begin = cur;
    DecLayer& y = dec[l];
    const std::string p = "decoder/decoder_layers/" + std::to_string(l);
    y.f2 = add_dense(cur, dff, d, d); dense_views(y.f2, p + "/ffn/layer_with_weights-1");
    y.f1 = add_dense(cur, d, dff, dff); dense_views(y.f1, p + "/ffn/layer_with_weights-0");
    ln(y.ln3, p + "/layernorm3");
    y.o2 = add_dense(cur, d, d, d); dense_views(y.o2, p + "/mha2/dense");
    y.q2 = add_dense(cur, d, d, d); dense_views(y.q2, p + "/mha2/query_generator_weights");
    ln(y.ln2, p + "/layernorm2");
    y.o1 = add_dense(cur, d, d, d); dense_views(y.o1, p + "/mha1/dense");
    y.qkv1 = add_dense(cur, d, 3 * d, 3 * d); qkv_views(y.qkv1, p + "/mha1");
    ln(y.ln1, p + "/layernorm1");
    close_section(begin);
  }
  begin = cur;
  dec_emb = add_tensor(cur, size_t(Vt) * d);
  add_view("decoder/embedding/embeddings", dec_emb, Vt, d, d, 2);
  kv_all = add_dense(cur, d, L * 2 * d, L * 2 * d);
  for (int l = 0; l < L; ++l) {
    const std::string p = "decoder/decoder_layers/" + std::to_string(l) + "/mha2";
    add_view(p + "/key_generator_weights/kernel", kv_all.w + size_t(l) * 2 * d, d, d, kv_all.ld, 2);
    add_view(p + "/key_generator_weights/bias", kv_all.b + size_t(l) * 2 * d, 1, d, d, 1);
    add_view(p + "/value_generator_weights/kernel", kv_all.w + size_t(l) * 2 * d + d, d, d, kv_all.ld, 2);
    add_view(p + "/value_generator_weights/bias", kv_all.b + size_t(l) * 2 * d + d, 1, d, d, 1);
  }
  close_section(begin);
  for (int l = L - 1; l >= 0; --l) {
    begin = cur;
    EncLayer& y = enc[l];
    const std::string p = "encoder/encoder_layers/" + std::to_string(l);
    y.f2 = add_dense(cur, dff, d, d); dense_views(y.f2, p + "/ffn/layer_with_weights-1");
    y.f1 = add_dense(cur, d, dff, dff); dense_views(y.f1, p + "/ffn/layer_with_weights-0");
    ln(y.ln2, p + "/layernorm2");
    y.o = add_dense(cur, d, d, d); dense_views(y.o, p + "/mha/dense");
    y.qkv = add_dense(cur, d, 3 * d, 3 * d); qkv_views(y.qkv, p + "/mha");
    ln(y.ln1, p + "/layernorm1");
    close_section(begin);
  }
  begin = cur;
  enc_emb = add_tensor(cur, size_t(Vin) * d);
  add_view("encoder/embedding/embeddings", enc_emb, Vin, d, d, 2);
  close_section(begin);
  total_params = cur;

  // logical table in the physical order of the reference's data shard
  table.clear();
  auto push = [&](const std::string& n) { table.push_back(views.at(n)); };
  push("final_layer/kernel"); push("final_layer/bias");
  push("encoder/embedding/embeddings"); push("decoder/embedding/embeddings");
  for (int l = 0; l < L; ++l)
    for (const char* s : {"layernorm1", "layernorm2"}) {
      push("encoder/encoder_layers/" + std::to_string(l) + "/" + s + "/gamma");
      push("encoder/encoder_layers/" + std::to_string(l) + "/" + s + "/beta");
    }
  for (int l = 0; l < L; ++l)
    for (const char* s : {"layernorm1", "layernorm2", "layernorm3"}) {
      push("decoder/decoder_layers/" + std::to_string(l) + "/" + s + "/gamma");
      push("decoder/decoder_layers/" + std::to_string(l) + "/" + s + "/beta");
    }
  auto mha = [&](const std::string& p) {
    for (const char* s : {"query_generator_weights", "key_generator_weights", "value_generator_weights", "dense"}) {
      push(p + "/" + s + "/kernel"); push(p + "/" + s + "/bias");
    }
  };
  auto ffn = [&](const std::string& p) {
    for (const char* s : {"layer_with_weights-0", "layer_with_weights-1"}) { push(p + "/ffn/" + s + "/kernel"); push(p + "/ffn/" + s + "/bias"); }
  };
  for (int l = 0; l < L; ++l) { const std::string p = "encoder/encoder_layers/" + std::to_string(l); mha(p + "/mha"); ffn(p); }
  for (int l = 0; l < L; ++l) { const std::string p = "decoder/decoder_layers/" + std::to_string(l); mha(p + "/mha1"); mha(p + "/mha2"); ffn(p); }
}

// ------------------------------------------------------------------------------------ create / destroy
void positional_encoding_host(float* out, int positions, int d) {
  // transformer_model.py:29-53, evaluated in fp32 like the TF ops it restates.
  for (int p = 0; p < positions; ++p)
    for (int i = 0; i < d; ++i) {
      const float expo = float(2 * (i / 2)) / float(d);
      const float rate = 1.0f / powf(10000.0f, expo);
      const float ang = float(p) * rate;
      out[size_t(p) * d + i] = (i & 1) ? cosf(ang) : sinf(ang);
    }
}

template <typename T>
int Context::arena_alloc(T** out, size_t count) {
  const size_t bytes = align_up(count * sizeof(T), 256);
  if (arena_used + bytes > arena_size) return set_error("workspace arena exhausted (%zu + %zu > %zu)", arena_used, bytes, arena_size);
  *out = reinterpret_cast<T*>(arena + arena_used);
  arena_used += bytes;
  return 0;
}

// Bytes of the activation / workspace arena for `max_tokens` tokens per side (the memory model behind the batch sizes of the
// long-sequence configuration: see b200xf_device_bytes).
size_t Context::arena_bytes_for(const b200xf_hparams& hp, int vt_ld, size_t* red_capacity_out, size_t* emb_scratch_out) {
  const int L = hp.layers_count, d = hp.embedding_size, H = hp.heads_count, dff = hp.feed_forward_hidden_size;
  const size_t N = size_t(hp.max_tokens);
  const size_t nd = N * d, nff = N * dff;
  size_t red_capacity = 0;
  // column-sum partials of one backward pass: 5L LayerNorms, per-layer QKV / FFN1 / (Q2) bias sums, K|V and vocab bias sums
  {
    const int n = int(N);
    size_t f = size_t(5 * L) * (ln_bwd_scratch_floats(n, d) + 64);
    f += size_t(2 * L) * (size_t(colsum_parts(n, 3 * d)) * 3 * d + 64) + size_t(2 * L) * (size_t(colsum_parts(n, dff)) * dff + 64);
    f += size_t(L) * (size_t(colsum_parts(n, d)) * d + 64) + size_t(colsum_parts(n, L * 2 * d)) * L * 2 * d + 64;
    f += size_t(colsum_parts(n, hp.target_vocabulary_size)) * hp.target_vocabulary_size + 64;
    red_capacity = f + 4096;
  }
  const size_t emb_scratch_bytes = embed_bwd_scratch_bytes(int(N));
  size_t bytes = 0;
  auto need = [&](size_t count, size_t elt) { bytes += align_up(count * elt, 256); };
  // encoder: x (L+1), qkv, ctx, z1, o1, h, z2, lse, 4 stats
  need(nd * (L + 1), 2);
  for (int l = 0; l < L; ++l) { need(3 * nd, 2); need(nd, 2); need(nd, 2); need(nd, 2); need(nff, 2); need(nd, 2); need(2 * N * H, 4); need(4 * N, 4); }
  // decoder: x (L+1), qkv, ctx1, z1, o1, q2, ctx2, z2, o2, h, z3, lse1, lse2, 6 stats
  need(nd * (L + 1), 2);
  for (int l = 0; l < L; ++l) { need(3 * nd, 2); for (int i = 0; i < 8; ++i) need(nd, 2); need(nff, 2); need(4 * N * H, 4); need(6 * N, 4); }
  need(N * size_t(L) * 2 * d, 2); need(N * size_t(L) * 2 * d, 2);   // kv_all, dkv_all
  need(N * size_t(vt_ld), 2);                                       // logits / dlogits
  for (int i = 0; i < 9; ++i) need(nd, 2);                          // dA dB dctx, 2 x (dz dzm dq2)
  need(2 * nff, 2); need(2 * 3 * nd, 2);                            // 2 x dh, 2 x dqkv
  need(N * H, 4); need(nd, 4);                                      // delta, fp32 dQ accumulator
  for (int i = 0; i < 4; ++i) need(N + 64, 8);                      // inp, target, tgt_in, labels
  need(N, 4); need(N, 4);                                           // row_loss, row_correct
  need(red_capacity, 4); need(emb_scratch_bytes, 1);
  need(size_t(hp.maximum_input_length) * d, 4); need(size_t(hp.maximum_target_length) * d, 4);
  need(1024, 1);
  if (red_capacity_out) *red_capacity_out = red_capacity;
  if (emb_scratch_out) *emb_scratch_out = emb_scratch_bytes;
  return bytes + (1 << 20);
}

int Context::init(const b200xf_hparams& h, int dev, cudaStream_t s) {
  hp = h; device = dev; user_stream = s;
  const int L = hp.layers_count, d = hp.embedding_size, H = hp.heads_count, dff = hp.feed_forward_hidden_size;
  if (L <= 0 || d <= 0 || H <= 0 || dff <= 0) return set_error("invalid hyper-parameters");
  if (d % H) return set_error("embedding_size %% heads_count != 0 (transformer_model.py:117)");
  const int dk = d / H;
  if (dk != 16 && dk != 32 && dk != 64) return set_error("head size %d unsupported (16, 32 or 64)", dk);
  if (d % 8 || dff % 8 || d > 2048) return set_error("embedding_size and feed_forward_hidden_size must be multiples of 8 (d <= 2048)");
  if (!(hp.droput_rate >= 0.f && hp.droput_rate < 1.f)) return set_error("dropout rate %g must be in [0, 1)", hp.droput_rate);
  if (hp.max_tokens <= 0) return set_error("max_tokens must be positive");
  B200XF_CUDA(cudaSetDevice(device));
  {
    int least = 0, greatest = 0;
    B200XF_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    B200XF_CUDA(cudaStreamCreateWithPriority(&stream, cudaStreamNonBlocking, greatest));  // the critical path outranks the side streams
    B200XF_CUDA(cudaEventCreateWithFlags(&ev_user, cudaEventDisableTiming));
    B200XF_CUDA(cudaEventCreateWithFlags(&ev_own, cudaEventDisableTiming));
  }
  build_layout();
  const size_t P = total_params;
  B200XF_CUDA(cudaMalloc(&params, P * sizeof(float)));
  B200XF_CUDA(cudaMalloc(&grads, P * sizeof(float)));
  B200XF_CUDA(cudaMalloc(&adam_m, P * sizeof(float)));
  B200XF_CUDA(cudaMalloc(&adam_v, P * sizeof(float)));
  B200XF_CUDA(cudaMalloc(&shadow, P * sizeof(bf16)));
  B200XF_CUDA(cudaMemsetAsync(params, 0, P * sizeof(float), stream));
  B200XF_CUDA(cudaMemsetAsync(grads, 0, P * sizeof(float), stream));
  B200XF_CUDA(cudaMemsetAsync(adam_m, 0, P * sizeof(float), stream));
  B200XF_CUDA(cudaMemsetAsync(adam_v, 0, P * sizeof(float), stream));
  B200XF_CUDA(cudaMemsetAsync(shadow, 0, P * sizeof(bf16), stream));
  grads_clean = true;

  // ---- workspace arena
  const size_t N = size_t(hp.max_tokens);
  const size_t nd = N * d, nff = N * dff;
  arena_size = arena_bytes_for(hp, vt_ld, &red_capacity, &emb_scratch_bytes);
  B200XF_CUDA(cudaMalloc(&arena, arena_size));
  B200XF_CUDA(cudaMemsetAsync(arena, 0, arena_size, stream));
  arena_used = 0;

  B200XF_TRY(arena_alloc(&enc_x, nd * (L + 1)));
  ea.resize(L);
  for (int l = 0; l < L; ++l) {
    EncAct& a = ea[l];
    B200XF_TRY(arena_alloc(&a.qkv, 3 * nd)); B200XF_TRY(arena_alloc(&a.ctx, nd)); B200XF_TRY(arena_alloc(&a.z1, nd));
    B200XF_TRY(arena_alloc(&a.o1, nd)); B200XF_TRY(arena_alloc(&a.h, nff)); B200XF_TRY(arena_alloc(&a.z2, nd));
    B200XF_TRY(arena_alloc(&a.lse, 2 * N * H)); B200XF_TRY(arena_alloc(&a.stats, 4 * N));
  }
  B200XF_TRY(arena_alloc(&dec_x, nd * (L + 1)));
  da.resize(L);
  for (int l = 0; l < L; ++l) {
    DecAct& a = da[l];
    B200XF_TRY(arena_alloc(&a.qkv, 3 * nd)); B200XF_TRY(arena_alloc(&a.ctx1, nd)); B200XF_TRY(arena_alloc(&a.z1, nd));
    B200XF_TRY(arena_alloc(&a.o1, nd)); B200XF_TRY(arena_alloc(&a.q2, nd)); B200XF_TRY(arena_alloc(&a.ctx2, nd));
    B200XF_TRY(arena_alloc(&a.z2, nd)); B200XF_TRY(arena_alloc(&a.o2, nd)); B200XF_TRY(arena_alloc(&a.z3, nd));
    B200XF_TRY(arena_alloc(&a.h, nff)); B200XF_TRY(arena_alloc(&a.lse1, 2 * N * H)); B200XF_TRY(arena_alloc(&a.lse2, 2 * N * H));
    B200XF_TRY(arena_alloc(&a.stats, 6 * N));
  }
  B200XF_TRY(arena_alloc(&kv, N * size_t(L) * 2 * d)); B200XF_TRY(arena_alloc(&dkv, N * size_t(L) * 2 * d));
  B200XF_TRY(arena_alloc(&logits, N * size_t(vt_ld)));
  B200XF_TRY(arena_alloc(&dA, nd)); B200XF_TRY(arena_alloc(&dB, nd)); B200XF_TRY(arena_alloc(&dctx, nd));
  for (int c = 0; c < 2; ++c) {
    B200XF_TRY(arena_alloc(&sb_dz.buf[c], nd)); B200XF_TRY(arena_alloc(&sb_dzm.buf[c], nd)); B200XF_TRY(arena_alloc(&sb_dq2.buf[c], nd));
    B200XF_TRY(arena_alloc(&sb_dh.buf[c], nff)); B200XF_TRY(arena_alloc(&sb_dqkv.buf[c], 3 * nd));
    for (SideBuf* b : {&sb_dz, &sb_dzm, &sb_dq2, &sb_dh, &sb_dqkv}) B200XF_CUDA(cudaEventCreateWithFlags(&b->done[c], cudaEventDisableTiming));
  }
  dz = sb_dz.buf[0]; dzm = sb_dzm.buf[0]; dq2 = sb_dq2.buf[0]; dh = sb_dh.buf[0]; dqkv = sb_dqkv.buf[0];
  B200XF_TRY(arena_alloc(&delta, N * H)); B200XF_TRY(arena_alloc(&dq_acc, nd));
  B200XF_TRY(arena_alloc(&tok_inp, N + 64)); B200XF_TRY(arena_alloc(&tok_target, N + 64));
  B200XF_TRY(arena_alloc(&tok_tgt_in, N + 64)); B200XF_TRY(arena_alloc(&tok_labels, N + 64));
  B200XF_TRY(arena_alloc(&row_loss, N)); B200XF_TRY(arena_alloc(&row_correct, N));
  B200XF_TRY(arena_alloc(&red_scratch, red_capacity));
  B200XF_TRY(arena_alloc(&emb_scratch, emb_scratch_bytes));
  B200XF_TRY(arena_alloc(&pe_in, size_t(hp.maximum_input_length) * d));
  B200XF_TRY(arena_alloc(&pe_tgt, size_t(hp.maximum_target_length) * d));
  char* small = nullptr;
  B200XF_TRY(arena_alloc(&small, 1024));
  iterations_dev = reinterpret_cast<long long*>(small);
  nonpad_count = reinterpret_cast<int*>(small + 64);
  metrics = reinterpret_cast<float*>(small + 256);      // 3 local + 3 reduced
  nvls_words = reinterpret_cast<unsigned*>(small + 512); // launch counter and CTA counter of the NVLS update kernels (arena is zeroed)

  {
    std::vector<float> pe(size_t(std::max(hp.maximum_input_length, hp.maximum_target_length)) * d);
    positional_encoding_host(pe.data(), hp.maximum_input_length, d);
    B200XF_CUDA(cudaMemcpyAsync(pe_in, pe.data(), size_t(hp.maximum_input_length) * d * 4, cudaMemcpyHostToDevice, stream));
    B200XF_CUDA(cudaStreamSynchronize(stream));
    positional_encoding_host(pe.data(), hp.maximum_target_length, d);
    B200XF_CUDA(cudaMemcpyAsync(pe_tgt, pe.data(), size_t(hp.maximum_target_length) * d * 4, cudaMemcpyHostToDevice, stream));
    B200XF_CUDA(cudaStreamSynchronize(stream));
  }
  if (const char* e = getenv("B200XF_OVERLAP_UPDATE")) overlap_single = atoi(e) != 0;
  adam.beta1 = hp.adam_beta_1; adam.beta2 = hp.adam_beta_2; adam.eps = hp.adam_epsilon;
  adam.rsqrt_d_model = 1.0f / sqrtf(float(d));
  adam.warmup_pow_m15 = float(pow(double(hp.warmup_steps), -1.5));
  B200XF_CUDA(cudaMallocHost(&host_out, 64));
  B200XF_CUDA(cudaStreamCreateWithFlags(&comm_stream, cudaStreamNonBlocking));
  B200XF_CUDA(cudaStreamCreateWithFlags(&upd_stream, cudaStreamNonBlocking));
  B200XF_CUDA(cudaStreamCreateWithFlags(&side_stream, cudaStreamNonBlocking));
  B200XF_CUDA(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
  B200XF_CUDA(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
  if (const char* e = getenv("B200XF_SIDE_WGRAD")) side_on = atoi(e) != 0;
  if (const char* e = getenv("B200XF_FUSE_LN")) fuse_ln = atoi(e);
  if (const char* e = getenv("B200XF_FUSE_COLSUM")) fuse_colsum = atoi(e);
  if (const char* e = getenv("B200XF_FUSE_MIN_TOKENS")) fuse_min_tokens = std::max(0, atoi(e));
  if (const char* e = getenv("B200XF_PHASE_TIMING")) phase_timing = atoi(e) != 0;
  if (const char* e = getenv("B200XF_GRAPH")) use_graph = atoi(e) != 0;
  if (const char* e = getenv("B200XF_SHARDED_UPDATE")) sharded_update = std::max(0, std::min(2, atoi(e)));
  if (const char* e = getenv("B200XF_BUCKET_JOIN")) bucket_join = atoi(e) != 0;
  if (const char* e = getenv("B200XF_PIPELINED_UPDATE")) pipelined_update = atoi(e);  // 0 off, 1 every bucket, 2 the last two bucket closes
  if (const char* e = getenv("B200XF_COMM_SMS")) comm_sms = std::max(0, atoi(e));
  if (const char* e = getenv("B200XF_BWD_GEMM_CTAS")) bwd_gemm_ctas = std::max(0, atoi(e));
  if (const char* e = getenv("B200XF_NVLS_CTAS")) nvls_ctas = std::max(1, atoi(e));
  if (const char* e = getenv("B200XF_NVLS_TAIL_CTAS")) nvls_tail_ctas = std::max(1, atoi(e));
  if (const char* e = getenv("B200XF_WEIGHT_PREFETCH")) weights_static = atoi(e);  // bit mask, see context.h
  if (const char* e = getenv("B200XF_BUCKET_MB")) { bucket_bytes = size_t(std::max(1, atoi(e))) << 20; bucket_bytes_set = true; }
  B200XF_CUDA(cudaEventCreateWithFlags(&ev_ready, cudaEventDisableTiming));
  B200XF_CUDA(cudaEventCreateWithFlags(&ev_done, cudaEventDisableTiming));
  B200XF_CUDA(cudaEventCreateWithFlags(&ev_rs, cudaEventDisableTiming));
  B200XF_CUDA(cudaEventCreateWithFlags(&ev_adam[0], cudaEventDisableTiming));
  B200XF_CUDA(cudaEventCreateWithFlags(&ev_adam[1], cudaEventDisableTiming));
  B200XF_CUDA(cudaEventCreateWithFlags(&ev_ag, cudaEventDisableTiming));
  return 0;
}

Context::~Context() {
  if (!stream && !params && !arena) return;  // never initialised (layout-only use): nothing on any device to release
  drop_graphs();
  if (phase_timing && phase_n > 0)
    fprintf(stderr, "[b200xf] phases over %ld steps (ms): forward %.3f  loss %.3f  backward %.3f  update+join %.3f\n", phase_n,
            phase_ms[0] / phase_n, phase_ms[1] / phase_n, phase_ms[2] / phase_n, phase_ms[3] / phase_n);
  if (comm) comm_destroy();
  cudaSetDevice(device);
  if (stream) { cudaStreamSynchronize(stream); cudaStreamDestroy(stream); }
  if (ev_user) cudaEventDestroy(ev_user);
  if (ev_own) cudaEventDestroy(ev_own);
  if (!external_state) { cudaFree(params); cudaFree(grads); }
  cudaFree(adam_m); cudaFree(adam_v); cudaFree(shadow); cudaFree(arena);
  if (host_out) cudaFreeHost(host_out);
  if (comm_stream) cudaStreamDestroy(comm_stream);
  if (upd_stream) cudaStreamDestroy(upd_stream);
  for (cudaEvent_t e : {ev_rs, ev_adam[0], ev_adam[1], ev_ag}) if (e) cudaEventDestroy(e);
  if (side_stream) { cudaStreamSynchronize(side_stream); cudaStreamDestroy(side_stream); }
  if (ev_fork) cudaEventDestroy(ev_fork);
  if (ev_join) cudaEventDestroy(ev_join);
  for (SideBuf* b : {&sb_dz, &sb_dzm, &sb_dq2, &sb_dh, &sb_dqkv}) for (int c = 0; c < 2; ++c) if (b->done[c]) cudaEventDestroy(b->done[c]);
  if (ev_ready) cudaEventDestroy(ev_ready);
  if (ev_done) cudaEventDestroy(ev_done);
  gemm_clear_tensor_map_cache();
}

int Context::params_changed() { return cast_f32_to_bf16(params, shadow, total_params, stream); }

int Context::set_iterations(long long it) {
  iterations_host = it;
  B200XF_CUDA(cudaMemcpyAsync(iterations_dev, &iterations_host, sizeof(long long), cudaMemcpyHostToDevice, stream));
  B200XF_CUDA(cudaStreamSynchronize(stream));
  return 0;
}
int Context::get_iterations(long long* it) {
  B200XF_CUDA(cudaMemcpyAsync(&iterations_host, iterations_dev, sizeof(long long), cudaMemcpyDeviceToHost, stream));
  B200XF_CUDA(cudaStreamSynchronize(stream));
  *it = iterations_host;
  return 0;
}

int Context::enter() {
  B200XF_CUDA(cudaSetDevice(device));
  B200XF_CUDA(cudaEventRecord(ev_user, user_stream));
  B200XF_CUDA(cudaStreamWaitEvent(stream, ev_user, 0));
  return 0;
}
int Context::leave() {
  B200XF_CUDA(cudaEventRecord(ev_own, stream));
  B200XF_CUDA(cudaStreamWaitEvent(user_stream, ev_own, 0));
  return 0;
}

// ------------------------------------------------------------------------------------ comm
int Context::comm_init(const void* id128, int rank_, int world_) {
  NcclApi* api = nccl_api();
  if (!api) return set_error("NCCL library not found (set B200XF_NCCL_LIB)");
  if (comm) return set_error("communicator already attached");
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  B200XF_CUDA(cudaSetDevice(device));
  ncclComm_t c;
  B200XF_NCCL(api->CommInitRank(&c, world_, id, rank_));
  comm = c; rank = rank_; world = world_;
  drop_graphs();  // steps captured without (or with another) communicator must not be replayed
  return 0;
}
int Context::comm_destroy() {
  if (!comm) return 0;
  NcclApi* api = nccl_api();
  cudaStreamSynchronize(comm_stream);
  if (upd_stream) cudaStreamSynchronize(upd_stream);
  drop_graphs();  // captured steps hold collectives of the communicator that goes away
  if (api) api->CommDestroy(static_cast<ncclComm_t>(comm));
  comm = nullptr; world = 1; rank = 0; nvls_on = false;
  return 0;
}
// Parameters and gradients move into buffers the caller allocated (symmetric memory with a multicast mapping, so that the
// NVLS update kernels can reduce / broadcast them across replicas); contents are carried over, the old buffers are released.
int Context::use_external_state(float* params_ext, float* grads_ext) {
  if (!params_ext || !grads_ext) return set_error("use_external_state: null buffer");
  if ((reinterpret_cast<uintptr_t>(params_ext) | reinterpret_cast<uintptr_t>(grads_ext)) & 255) return set_error("use_external_state: buffers must be 256-byte aligned");
  B200XF_CUDA(cudaSetDevice(device));
  drop_graphs();
  B200XF_CUDA(cudaMemcpyAsync(params_ext, params, total_params * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  B200XF_CUDA(cudaMemcpyAsync(grads_ext, grads, total_params * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  B200XF_CUDA(cudaStreamSynchronize(stream));
  if (!external_state) { cudaFree(params); cudaFree(grads); }
  params = params_ext; grads = grads_ext; external_state = true;
  gemm_clear_tensor_map_cache();
  return 0;
}

int Context::nvls_attach(void* grads_mc, void* params_mc, void* const* pads, void* const* grads_peers, void* const* params_peers, int p2p,
                         int rank_, int world_) {
  if (!comm || rank_ != rank || world_ != world) return set_error("nvls_attach: attach the communicator first (same rank / world)");
  if (!external_state) return set_error("nvls_attach: parameters and gradients must live in symmetric memory (use_external_state)");
  if (world_ < 2 || world_ > kNvlsMaxWorld) return set_error("nvls_attach: world of %d unsupported (2..%d)", world_, kNvlsMaxWorld);
  if (!pads) return set_error("nvls_attach: null flag pads");
  if (p2p ? (!grads_peers || !params_peers) : (!grads_mc || !params_mc)) return set_error("nvls_attach: null %s mapping", p2p ? "peer" : "multicast");
  nvls = NvlsComm();
  nvls.p2p = p2p ? 1 : 0;
  if (grads_peers && params_peers)
    for (int r = 0; r < world_; ++r) {
      if (p2p && (!grads_peers[r] || !params_peers[r])) return set_error("nvls_attach: null peer mapping");
      nvls.grads_peer[r] = static_cast<float*>(grads_peers[r]); nvls.params_peer[r] = static_cast<float*>(params_peers[r]);
    }
  if (p2p && (nvls.grads_peer[rank_] != grads || nvls.params_peer[rank_] != params)) return set_error("nvls_attach: this rank's peer mapping is not its own buffer");
  nvls.grads_mc = static_cast<float*>(grads_mc); nvls.params_mc = static_cast<float*>(params_mc);
  for (int r = 0; r < world_; ++r) { if (!pads[r]) return set_error("nvls_attach: null flag pad"); nvls.pad_peer[r] = static_cast<unsigned*>(pads[r]); }
  nvls.pad_local = nvls.pad_peer[rank_];
  nvls.rank = rank_; nvls.world = world_;
  if (!getenv("B200XF_NVLS_CTAS")) nvls_ctas = std::max(16, 128 / world_);  // a rank's shard is 1/world of a bucket: 16 CTAs keep up on 8 replicas, 4 need 32
  B200XF_CUDA(cudaMemsetAsync(nvls_words, 0, 2 * sizeof(unsigned), stream));
  B200XF_CUDA(cudaStreamSynchronize(stream));
  drop_graphs();
  nvls_on = true;
  return 0;
}

int Context::allreduce(float* ptr, size_t count, cudaStream_t s) {
  if (!comm) return set_error("no communicator attached");
  NcclApi* api = nccl_api();
  B200XF_NCCL(api->AllReduce(ptr, ptr, count, ncclFloat32, ncclSum, static_cast<ncclComm_t>(comm), s));
  return 0;
}

// ------------------------------------------------------------------------------------ dropout sites
// site ids: 0 encoder embedding, 1 decoder embedding, 2+2l / 3+2l encoder layer l after MHA / FFN,
// 2+2L+3l.. decoder layer l after MHA1 / MHA2 / FFN (transformer_model.py:196,202,232,238,244,276,308).
DropArgs Context::site_drop(int site) const {
  DropArgs d;
  if (!dropout_on) return d;
  d.step = iterations_dev;
  d.key = (dropout_seed * 0x9E3779B97F4A7C15ull) ^ (static_cast<unsigned long long>(site + 1) * 0xC2B2AE3D27D4EB4Full);
  // Each replica draws its own masks, as the workers of MultiWorkerMirroredStrategy do (every worker has its own RNG):
  // the replica's rank is part of the key.  A mask is reproduced by the same (seed, rank, site, iterations).
  d.key ^= static_cast<unsigned long long>(rank) * 0xD6E8FEB86659FD93ull;
  const double t = double(hp.droput_rate) * 4294967296.0;
  d.thresh = t >= 4294967295.0 ? 4294967295u : static_cast<unsigned>(t);
  d.scale = 1.0f / (1.0f - hp.droput_rate);
  return d;
}

// ------------------------------------------------------------------------------------ side stream (wgrad, bias sums)
int Context::sb_write(SideBuf& b, bf16** out) {
  b.cur ^= 1;
  if (b.pending[b.cur]) { B200XF_CUDA(cudaStreamWaitEvent(stream, b.done[b.cur], 0)); b.pending[b.cur] = false; }
  *out = b.buf[b.cur];
  return 0;
}
int Context::side_launch_begin(cudaStream_t* s) {
  if (!side_on) { *s = stream; return 0; }
  B200XF_CUDA(cudaEventRecord(ev_fork, stream));               // everything the compute stream has produced so far ...
  B200XF_CUDA(cudaStreamWaitEvent(side_stream, ev_fork, 0));   // ... is visible to the side stream
  side_dirty = true;
  *s = side_stream;
  return 0;
}
int Context::side_launch_end(const void* read_ptr) {
  if (!side_on) return 0;
  for (SideBuf* b : {&sb_dz, &sb_dzm, &sb_dq2, &sb_dh, &sb_dqkv})
    for (int c = 0; c < 2; ++c)
      if (read_ptr == b->buf[c]) { B200XF_CUDA(cudaEventRecord(b->done[c], side_stream)); b->pending[c] = true; }
  return 0;
}
int Context::join_side() {
  if (!side_on || !side_dirty) return 0;
  B200XF_CUDA(cudaEventRecord(ev_join, side_stream));
  B200XF_CUDA(cudaStreamWaitEvent(stream, ev_join, 0));
  side_dirty = false;
  // everything the side stream has read is now ordered before whatever the compute stream does next
  for (SideBuf* b : {&sb_dz, &sb_dzm, &sb_dq2, &sb_dh, &sb_dqkv}) b->pending[0] = b->pending[1] = false;
  return 0;
}

// ------------------------------------------------------------------------------------ GEMM helpers
// With a communicator attached, NCCL's reduce-scatter / all-gather kernels sit on a few SMs for much of the backward
// pass; a persistent GEMM grid that counts on all 148 SMs then waits for them with its last CTAs.  Leaving those SMs
// out of the backward GEMM grids costs the 128-tile problems nothing.
int Context::backward_gemm_cap() const {
  if (bwd_gemm_ctas > 0) return bwd_gemm_ctas;
  if (!comm || world <= 1 || comm_sms <= 0) return 0;
  return std::max(num_sms() / 2, num_sms() - comm_sms);
}

int Context::dense_fwd(const bf16* x, int M, const Dense& p, bf16* y, int relu, const bf16* resid, int drop_site) {
  GemmArgs g;
  g.a = x; g.lda = p.K; g.a_mn = 0;
  g.b = shadow + p.w; g.ldb = p.ld; g.b_mn = 1;
  g.d = y; g.ldd = p.ld; g.d_fp32 = 0;
  g.M = M; g.N = p.N; g.K = p.K;
  g.bias = params + p.b; g.relu = relu;
  g.resid = resid; g.ldr = p.ld; g.resid_op = resid ? 1 : 0;
  if (drop_site >= 0) g.drop = site_drop(drop_site);
  g.b_static = weights_static & 1;
  return gemm_bf16(g, stream);
}
int Context::dense_dgrad(const bf16* dy, int M, const Dense& p, bf16* dx, const bf16* resid, int resid_op) {
  GemmArgs g;
  g.a = dy; g.lda = p.ld; g.a_mn = 0;
  g.b = shadow + p.w; g.ldb = p.ld; g.b_mn = 0;  // W [K rows][N] read as [out=K][contract=N]
  g.d = dx; g.ldd = p.K; g.d_fp32 = 0;
  g.M = M; g.N = p.K; g.K = p.N;
  g.resid = resid; g.ldr = p.K; g.resid_op = resid ? resid_op : 0;
  g.b_static = weights_static & 1;
  g.max_ctas = backward_gemm_cap();
  return gemm_bf16(g, stream);
}
int Context::dense_ln_fwd(const bf16* x, int M, const Dense& p, bf16* z, const bf16* resid, int drop_site, const LnP& ln, bf16* y,
                          float* mean, float* rstd) {
  const float eps = 1e-6f;  // transformer_model.py:187-188
  if (!fused(fuse_ln, M) || !gemm_ln_fusable(M, p.N)) {
    B200XF_TRY(dense_fwd(x, M, p, z, 0, resid, drop_site));
    return layernorm_fwd(z, nullptr, params + ln.g, params + ln.b, y, nullptr, mean, rstd, M, p.N, eps, stream);
  }
  GemmArgs g;
  g.a = x; g.lda = p.K; g.a_mn = 0;
  g.b = shadow + p.w; g.ldb = p.ld; g.b_mn = 1;
  g.d = z; g.ldd = p.ld; g.d_fp32 = 0;
  g.M = M; g.N = p.N; g.K = p.K;
  g.bias = params + p.b;
  g.resid = resid; g.ldr = p.ld; g.resid_op = resid ? 1 : 0;
  if (drop_site >= 0) g.drop = site_drop(drop_site);
  g.b_static = weights_static & 1;
  g.ln_gamma = params + ln.g; g.ln_beta = params + ln.b; g.ln_y = y; g.ln_ldy = p.N; g.ln_mean = mean; g.ln_rstd = rstd; g.ln_eps = eps;
  return gemm_bf16(g, stream);
}

int Context::dense_wgrad(const bf16* x, const bf16* dy, int M, const Dense& p, bool bias_grad) {
  if (bias_grad && !fused(fuse_colsum, M)) { B200XF_TRY(colsum(dy, p.ld, M, p.N, p.b)); bias_grad = false; }
  GemmArgs g;
  g.a = x; g.lda = p.K; g.a_mn = 1;    // X [tokens][K]  == A stored [contract][M=K]
  g.b = dy; g.ldb = p.ld; g.b_mn = 1;  // dY [tokens][N] == B stored [contract][N]
  g.d = grads + p.w; g.ldd = p.ld; g.d_fp32 = 1; g.accumulate = 1;
  g.M = p.K; g.N = p.N; g.K = M;
  g.max_ctas = backward_gemm_cap();
  if (bias_grad) {
    GemmInfo info;
    B200XF_TRY(gemm_plan(g, &info));
    const int rows = info.splits * info.tiles_m;   // one partial row per (K split, row block of the weight)
    float* part = red_alloc(size_t(rows) * p.N);
    if (!part) return set_error("reduction scratch exhausted");
    g.colsum_partials = part;
    red_jobs.push_back({part, grads + p.b, rows, p.N, p.N});
  }
  cudaStream_t s;
  B200XF_TRY(side_launch_begin(&s));
  // No operand prefetch here: dy arrives over the fork event, and a launch with the programmatic attribute may begin
  // before that event has fired (the dependency is resolved at griddepcontrol.wait) -- prefetching it read stale
  // gradients; prefetching only x (a forward activation) measured slower.
  B200XF_TRY(gemm_bf16(g, s));
  return side_launch_end(dy);
}

// ------------------------------------------------------------------------------------ forward
int Context::check_shapes(int B, int S, int T) {
  if (B <= 0 || S <= 0 || T <= 0) return set_error("empty batch (B=%d S=%d T=%d)", B, S, T);
  if (size_t(B) * S > size_t(hp.max_tokens) || size_t(B) * T > size_t(hp.max_tokens))
    return set_error("batch of %d x %d / %d x %d tokens exceeds max_tokens=%d", B, S, B, T, hp.max_tokens);
  if (S > hp.maximum_input_length) return set_error("input length %d exceeds maximum_input_length %d", S, hp.maximum_input_length);
  if (T > hp.maximum_target_length) return set_error("target length %d exceeds maximum_target_length %d", T, hp.maximum_target_length);
  return 0;
}

int Context::forward_encoder(const long long* inp, int B, int S) {
  const int L = hp.layers_count, d = hp.embedding_size, H = hp.heads_count, dk = d / H;
  const int Ns = B * S;
  const size_t cap = size_t(hp.max_tokens) * d;
  // ---- encoder (transformer_model.py:270-279)
  B200XF_TRY(embed_pe_fwd(inp, params + enc_emb, hp.input_vocabulary_size, pe_in, enc_x, B, S, d, stream, site_drop(0)));
  for (int l = 0; l < L; ++l) {
    const EncLayer& w = enc[l]; EncAct& a = ea[l];
    const bf16* x = enc_x + size_t(l) * cap; bf16* out = enc_x + size_t(l + 1) * cap;
    float* st = a.stats;
    B200XF_TRY(dense_fwd(x, Ns, w.qkv, a.qkv, 0, nullptr));
    AttnArgs at;
    at.q = a.qkv; at.k = a.qkv + d; at.v = a.qkv + 2 * d; at.ldq = at.ldk = at.ldv = 3 * d;
    at.o = a.ctx; at.ldo = d; at.lse = a.lse; at.key_tokens = inp; at.causal = 0;
    at.B = B; at.H = H; at.Sq = S; at.Sk = S; at.dk = dk;
    B200XF_TRY(attention_fwd(at, stream));
    B200XF_TRY(dense_ln_fwd(a.ctx, Ns, w.o, a.z1, x, 2 + 2 * l, w.ln1, a.o1, st, st + Ns));          // z1 = x + drop(mha(x)); o1 = LN1(z1)
    B200XF_TRY(dense_fwd(a.o1, Ns, w.f1, a.h, 1, nullptr));
    B200XF_TRY(dense_ln_fwd(a.h, Ns, w.f2, a.z2, a.o1, 3 + 2 * l, w.ln2, out, st + 2 * Ns, st + 3 * Ns));  // z2 = o1 + drop(ffn(o1)); out = LN2(z2)
  }
  const bf16* enc_out = enc_x + size_t(L) * cap;
  return dense_fwd(enc_out, Ns, kv_all, kv, 0, nullptr);  // K|V of every decoder layer's cross-attention
}

int Context::forward(const long long* inp, const long long* tgt_in, int B, int S, int T, float* logits_f32, bool training) {
  B200XF_TRY(check_shapes(B, S, T));
  dropout_on = training && hp.droput_rate > 0.f;
  last_B = B; last_S = S; last_T = T; last_inp = inp; last_tgt = tgt_in;
  const int L = hp.layers_count, d = hp.embedding_size, H = hp.heads_count, dk = d / H;
  const int Nt = B * T;
  const size_t cap = size_t(hp.max_tokens) * d;
  B200XF_TRY(forward_encoder(inp, B, S));

  // ---- decoder (transformer_model.py:301-315)
  B200XF_TRY(embed_pe_fwd(tgt_in, params + dec_emb, hp.target_vocabulary_size, pe_tgt, dec_x, B, T, d, stream, site_drop(1)));
  for (int l = 0; l < L; ++l) {
    const DecLayer& w = dec[l]; DecAct& a = da[l];
    const bf16* x = dec_x + size_t(l) * cap; bf16* out = dec_x + size_t(l + 1) * cap;
    float* st = a.stats;
    B200XF_TRY(dense_fwd(x, Nt, w.qkv1, a.qkv, 0, nullptr));
    AttnArgs s1;
    s1.q = a.qkv; s1.k = a.qkv + d; s1.v = a.qkv + 2 * d; s1.ldq = s1.ldk = s1.ldv = 3 * d;
    s1.o = a.ctx1; s1.ldo = d; s1.lse = a.lse1; s1.key_tokens = tgt_in; s1.causal = 1;
    s1.B = B; s1.H = H; s1.Sq = T; s1.Sk = T; s1.dk = dk;
    B200XF_TRY(attention_fwd(s1, stream));
    B200XF_TRY(dense_ln_fwd(a.ctx1, Nt, w.o1, a.z1, x, 2 + 2 * L + 3 * l, w.ln1, a.o1, st, st + Nt));
    B200XF_TRY(dense_fwd(a.o1, Nt, w.q2, a.q2, 0, nullptr));
    AttnArgs c2;
    c2.q = a.q2; c2.ldq = d; c2.k = kv + size_t(l) * 2 * d; c2.v = kv + size_t(l) * 2 * d + d; c2.ldk = c2.ldv = kv_all.ld;
    c2.o = a.ctx2; c2.ldo = d; c2.lse = a.lse2; c2.key_tokens = inp; c2.causal = 0;
    c2.B = B; c2.H = H; c2.Sq = T; c2.Sk = S; c2.dk = dk;
    c2.early = (weights_static & 2) ? 2 : 0;  // K|V of all layers were projected right after the encoder
    B200XF_TRY(attention_fwd(c2, stream));
    B200XF_TRY(dense_ln_fwd(a.ctx2, Nt, w.o2, a.z2, a.o1, 3 + 2 * L + 3 * l, w.ln2, a.o2, st + 2 * Nt, st + 3 * Nt));
    B200XF_TRY(dense_fwd(a.o2, Nt, w.f1, a.h, 1, nullptr));
    B200XF_TRY(dense_ln_fwd(a.h, Nt, w.f2, a.z3, a.o2, 4 + 2 * L + 3 * l, w.ln3, out, st + 4 * Nt, st + 5 * Nt));
  }
  const bf16* dec_out = dec_x + size_t(L) * cap;
  // ---- final_layer (transformer_model.py:347)
  if (logits_f32) {
    GemmArgs g;
    g.a = dec_out; g.lda = d; g.a_mn = 0; g.b = shadow + final_layer.w; g.ldb = final_layer.ld; g.b_mn = 1;
    g.d = logits_f32; g.ldd = final_layer.N; g.d_fp32 = 1;
    g.M = Nt; g.N = final_layer.N; g.K = d; g.bias = params + final_layer.b;
    B200XF_TRY(gemm_bf16(g, stream));
  } else {
    B200XF_TRY(dense_fwd(dec_out, Nt, final_layer, logits, 0, nullptr));
  }
  return 0;
}

// ------------------------------------------------------------------------------------ incremental decoding
int Context::decode_begin(const long long* inp, int B, int S, int Tmax) {
  B200XF_TRY(check_shapes(B, S, Tmax));
  dropout_on = false;
  B200XF_TRY(forward_encoder(inp, B, S));
  // the decoded tokens so far, [B, Tmax], PAD (0) beyond the current position: the self-attention masks those keys out
  B200XF_CUDA(cudaMemsetAsync(tok_tgt_in, 0, size_t(B) * Tmax * sizeof(long long), stream));
  dec_B = B; dec_S = S; dec_Tmax = Tmax; dec_inp = inp;
  last_B = 0;  // the activations of the last full forward pass are gone: attention_weights needs a new one
  return 0;
}

namespace {
__global__ void put_tokens_kernel(const long long* __restrict__ tok, long long* __restrict__ all, int B, int Tmax, int t) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) all[size_t(b) * Tmax + t] = tok[b];
}
}  // namespace

// One decoder pass over position t of every sentence (transformer_model.py:301-315 restricted to the newest target
// position): tokens_t [B] are the target tokens at position t; logits_f32 [B, V] / next_tokens [B] (argmax) are optional.
int Context::decode_step(const long long* tokens_t, int t, float* logits_f32, long long* next_tokens) {
  if (dec_B <= 0) return set_error("decode_step: call decode_begin first");
  if (t < 0 || t >= dec_Tmax) return set_error("decode_step: position %d outside the cache of %d", t, dec_Tmax);
  const int L = hp.layers_count, d = hp.embedding_size, H = hp.heads_count, dk = d / H;
  const int B = dec_B, S = dec_S, Tmax = dec_Tmax;
  dropout_on = false;
  put_tokens_kernel<<<(B + 127) / 128, 128, 0, stream>>>(tokens_t, tok_tgt_in, B, Tmax, t);
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  bf16* x = dA;  // [B, d] working rows (the backward pass's scratch is free during inference)
  bf16* y = dB;
  B200XF_TRY(embed_pe_fwd(tokens_t, params + dec_emb, hp.target_vocabulary_size, pe_tgt + size_t(t) * d, x, B, 1, d, stream, DropArgs()));
  for (int l = 0; l < L; ++l) {
    const DecLayer& w = dec[l]; DecAct& a = da[l];
    float* st = a.stats;
    {  // Q|K|V of the new position into row t of each sentence's cache (rows of one sentence are Tmax * 3d apart)
      GemmArgs g;
      g.a = x; g.lda = d; g.b = shadow + w.qkv1.w; g.ldb = w.qkv1.ld; g.b_mn = 1;
      g.d = a.qkv + size_t(t) * 3 * d; g.ldd = Tmax * 3 * d;
      g.M = B; g.N = 3 * d; g.K = d; g.bias = params + w.qkv1.b; g.b_static = weights_static & 1;
      B200XF_TRY(gemm_bf16(g, stream));
    }
    AttnArgs s1;
    s1.q = a.qkv + size_t(t) * 3 * d; s1.ldq = Tmax * 3 * d; s1.k = a.qkv + d; s1.v = a.qkv + 2 * d; s1.ldk = s1.ldv = 3 * d;
    s1.o = a.ctx1; s1.ldo = d; s1.lse = a.lse1; s1.key_tokens = tok_tgt_in; s1.causal = 0;  // every cached key is <= t; later ones are PAD
    s1.B = B; s1.H = H; s1.Sq = 1; s1.Sk = Tmax; s1.dk = dk; s1.generic = 1;
    B200XF_TRY(attention_fwd(s1, stream));
    B200XF_TRY(dense_ln_fwd(a.ctx1, B, w.o1, a.z1, x, -1, w.ln1, a.o1, st, st + B));
    B200XF_TRY(dense_fwd(a.o1, B, w.q2, a.q2, 0, nullptr));
    AttnArgs c2;
    c2.q = a.q2; c2.ldq = d; c2.k = kv + size_t(l) * 2 * d; c2.v = kv + size_t(l) * 2 * d + d; c2.ldk = c2.ldv = kv_all.ld;
    c2.o = a.ctx2; c2.ldo = d; c2.lse = a.lse2; c2.key_tokens = dec_inp; c2.causal = 0;
    c2.B = B; c2.H = H; c2.Sq = 1; c2.Sk = S; c2.dk = dk; c2.generic = 1;
    B200XF_TRY(attention_fwd(c2, stream));
    B200XF_TRY(dense_ln_fwd(a.ctx2, B, w.o2, a.z2, a.o1, -1, w.ln2, a.o2, st + 2 * B, st + 3 * B));
    B200XF_TRY(dense_fwd(a.o2, B, w.f1, a.h, 1, nullptr));
    B200XF_TRY(dense_ln_fwd(a.h, B, w.f2, a.z3, a.o2, -1, w.ln3, y, st + 4 * B, st + 5 * B));
    std::swap(x, y);
  }
  float* lg = logits_f32 ? logits_f32 : reinterpret_cast<float*>(logits);  // the training logits buffer (bf16 [N, V]) holds B fp32 rows
  if (!logits_f32 && size_t(B) * final_layer.N * 4 > size_t(hp.max_tokens) * vt_ld * 2) return set_error("decode_step: batch too large for the logits buffer");
  GemmArgs g;
  g.a = x; g.lda = d; g.b = shadow + final_layer.w; g.ldb = final_layer.ld; g.b_mn = 1;
  g.d = lg; g.ldd = final_layer.N; g.d_fp32 = 1;
  g.M = B; g.N = final_layer.N; g.K = d; g.bias = params + final_layer.b;
  B200XF_TRY(gemm_bf16(g, stream));
  if (next_tokens) B200XF_TRY(argmax_rows(lg, final_layer.N, B, final_layer.N, next_tokens, stream));
  return 0;
}

// attention weights of decoder layer `layer`, block 1 (causal self-attention, [B,H,T,T]) or 2 (cross, [B,H,T,S]), for
// the batch of the LAST forward / step call (transformer_model.py:312-313)
int Context::attention_weights(int layer, int block, float* out) {
  if (last_B <= 0) return set_error("attention_weights: no forward pass has run yet");
  if (layer < 0 || layer >= hp.layers_count || (block != 1 && block != 2)) return set_error("attention_weights: bad layer/block");
  const int d = hp.embedding_size, H = hp.heads_count, dk = d / H;
  const DecAct& a = da[layer];
  AttnArgs at;
  at.B = last_B; at.H = H; at.Sq = last_T; at.dk = dk;
  if (block == 1) {
    at.q = a.qkv; at.k = a.qkv + d; at.ldq = at.ldk = 3 * d; at.lse = a.lse1; at.key_tokens = last_tgt; at.causal = 1; at.Sk = last_T;
  } else {
    at.q = a.q2; at.ldq = d; at.k = kv + size_t(layer) * 2 * d; at.ldk = kv_all.ld; at.lse = a.lse2; at.key_tokens = last_inp; at.causal = 0; at.Sk = last_S;
  }
  return attention_probs(at, out, stream);
}

// ------------------------------------------------------------------------------------ backward
float* Context::red_alloc(size_t floats) {
  floats = (floats + 63) / 64 * 64;
  if (red_used + floats > red_capacity) return nullptr;
  float* p = red_scratch + red_used;
  red_used += floats;
  return p;
}

int Context::ln_bwd(const bf16* dy, const bf16* z, const float* st_mean, const float* st_rstd, const LnP& p, size_t dcol_off,
                    int rows, int drop_site) {
  const int d = hp.embedding_size;
  const int n = ln_bwd_parts(rows, d);
  float* part = red_alloc(size_t(n) * 3 * d);
  if (!part) return set_error("reduction scratch exhausted");
  const DropArgs drop = site_drop(drop_site);
  B200XF_TRY(sb_write(sb_dz, &dz));
  if (drop.thresh) B200XF_TRY(sb_write(sb_dzm, &dzm));
  dsub = drop.thresh ? dzm : dz;  // what the Dense that fed this residual sees
  B200XF_TRY(layernorm_bwd_partial(dy, z, st_mean, st_rstd, params + p.g, dz, rows, d, part, stream, dzm, drop));
  red_jobs.push_back({part, grads + p.g, n, d, 3 * d});
  red_jobs.push_back({part + d, grads + p.b, n, d, 3 * d});
  red_jobs.push_back({part + 2 * d, grads + dcol_off, n, d, 3 * d});
  return 0;
}

int Context::colsum(const bf16* x, int ld, int rows, int cols, size_t dst_off) {
  const int n = colsum_parts(rows, cols);
  float* part = red_alloc(size_t(n) * cols);
  if (!part) return set_error("reduction scratch exhausted");
  cudaStream_t s;
  B200XF_TRY(side_launch_begin(&s));
  B200XF_TRY(colsum_partial(x, ld, rows, cols, part, s));
  red_jobs.push_back({part, grads + dst_off, n, cols, cols});
  return side_launch_end(x);
}

// One launch sums every pending set of column-sum partials (bias / gamma / beta gradients) into the flat gradient.
int Context::flush_reductions(cudaStream_t on) {
  if (red_jobs.empty()) return 0;
  B200XF_TRY(reduce_partials(red_jobs.data(), int(red_jobs.size()), on ? on : stream));
  red_jobs.clear();
  return 0;
}

int Context::backward(const long long* inp, const long long* tgt_in, int B, int S, int T) {
  const int L = hp.layers_count, d = hp.embedding_size, H = hp.heads_count, dk = d / H, dff = hp.feed_forward_hidden_size;
  const int Ns = B * S, Nt = B * T;
  const size_t cap = size_t(hp.max_tokens) * d;
  (void)dff;
  if (!grads_clean) B200XF_CUDA(cudaMemsetAsync(grads, 0, total_params * sizeof(float), stream));
  grads_clean = false;
  next_section = 0;
  red_used = 0; red_jobs.clear();

  const bf16* dec_out = dec_x + size_t(L) * cap;
  B200XF_TRY(dense_wgrad(dec_out, logits, Nt, final_layer, true));
  bf16* cur = dA; bf16* other = dB;
  B200XF_TRY(dense_dgrad(logits, Nt, final_layer, cur, nullptr, 0));
  B200XF_TRY(section_done());

  for (int l = L - 1; l >= 0; --l) {
    const DecLayer& w = dec[l]; DecAct& a = da[l];
    const bf16* x = dec_x + size_t(l) * cap;
    float* st = a.stats;
    // LN3 / FFN
    B200XF_TRY(ln_bwd(cur, a.z3, st + 4 * Nt, st + 5 * Nt, w.ln3, w.f2.b, Nt, 4 + 2 * L + 3 * l));
    B200XF_TRY(dense_wgrad(a.h, dsub, Nt, w.f2));
    B200XF_TRY(sb_write(sb_dh, &dh));
    B200XF_TRY(dense_dgrad(dsub, Nt, w.f2, dh, a.h, 2));
    B200XF_TRY(dense_wgrad(a.o2, dh, Nt, w.f1, true));
    B200XF_TRY(dense_dgrad(dh, Nt, w.f1, other, dz, 1));
    std::swap(cur, other);
    // LN2 / cross attention
    B200XF_TRY(ln_bwd(cur, a.z2, st + 2 * Nt, st + 3 * Nt, w.ln2, w.o2.b, Nt, 3 + 2 * L + 3 * l));
    B200XF_TRY(dense_wgrad(a.ctx2, dsub, Nt, w.o2));
    B200XF_TRY(dense_dgrad(dsub, Nt, w.o2, dctx, nullptr, 0));
    AttnArgs c2;
    c2.q = a.q2; c2.ldq = d; c2.k = kv + size_t(l) * 2 * d; c2.v = kv + size_t(l) * 2 * d + d; c2.ldk = c2.ldv = kv_all.ld;
    c2.o = a.ctx2; c2.ldo = d; c2.lse = a.lse2; c2.key_tokens = inp; c2.causal = 0;
    c2.B = B; c2.H = H; c2.Sq = T; c2.Sk = S; c2.dk = dk;
    B200XF_TRY(sb_write(sb_dq2, &dq2));
    c2.d_o = dctx; c2.lddo = d; c2.dq = dq2; c2.lddq = d;
    c2.dk_out = dkv + size_t(l) * 2 * d; c2.dv_out = dkv + size_t(l) * 2 * d + d; c2.lddk = c2.lddv = kv_all.ld; c2.delta = delta; c2.dq_acc = dq_acc;
    c2.early = (weights_static & 2) ? 3 : 0;
    B200XF_TRY(attention_bwd(c2, stream));
    B200XF_TRY(dense_wgrad(a.o1, dq2, Nt, w.q2, true));
    B200XF_TRY(dense_dgrad(dq2, Nt, w.q2, other, dz, 1));
    std::swap(cur, other);
    // LN1 / causal self attention
    B200XF_TRY(ln_bwd(cur, a.z1, st, st + Nt, w.ln1, w.o1.b, Nt, 2 + 2 * L + 3 * l));
    B200XF_TRY(dense_wgrad(a.ctx1, dsub, Nt, w.o1));
    B200XF_TRY(dense_dgrad(dsub, Nt, w.o1, dctx, nullptr, 0));
    AttnArgs s1;
    s1.q = a.qkv; s1.k = a.qkv + d; s1.v = a.qkv + 2 * d; s1.ldq = s1.ldk = s1.ldv = 3 * d;
    s1.o = a.ctx1; s1.ldo = d; s1.lse = a.lse1; s1.key_tokens = tgt_in; s1.causal = 1;
    s1.B = B; s1.H = H; s1.Sq = T; s1.Sk = T; s1.dk = dk;
    B200XF_TRY(sb_write(sb_dqkv, &dqkv));
    s1.d_o = dctx; s1.lddo = d; s1.dq = dqkv; s1.dk_out = dqkv + d; s1.dv_out = dqkv + 2 * d; s1.lddq = s1.lddk = s1.lddv = 3 * d;
    s1.delta = delta; s1.dq_acc = dq_acc;
    s1.early = (weights_static & 2) ? 3 : 0;
    B200XF_TRY(attention_bwd(s1, stream));
    B200XF_TRY(dense_wgrad(x, dqkv, Nt, w.qkv1, true));
    B200XF_TRY(dense_dgrad(dqkv, Nt, w.qkv1, other, dz, 1));
    std::swap(cur, other);
    B200XF_TRY(section_done());
  }
  B200XF_TRY(embed_bwd(tgt_in, cur, grads + dec_emb, hp.target_vocabulary_size, Nt, d, emb_scratch, emb_scratch_bytes, stream, site_drop(1)));
  const bf16* enc_out = enc_x + size_t(L) * cap;
  B200XF_TRY(dense_wgrad(enc_out, dkv, Ns, kv_all, true));
  B200XF_TRY(dense_dgrad(dkv, Ns, kv_all, cur, nullptr, 0));
  B200XF_TRY(section_done());

  for (int l = L - 1; l >= 0; --l) {
    const EncLayer& w = enc[l]; EncAct& a = ea[l];
    const bf16* x = enc_x + size_t(l) * cap;
    float* st = a.stats;
    B200XF_TRY(ln_bwd(cur, a.z2, st + 2 * Ns, st + 3 * Ns, w.ln2, w.f2.b, Ns, 3 + 2 * l));
    B200XF_TRY(dense_wgrad(a.h, dsub, Ns, w.f2));
    B200XF_TRY(sb_write(sb_dh, &dh));
    B200XF_TRY(dense_dgrad(dsub, Ns, w.f2, dh, a.h, 2));
    B200XF_TRY(dense_wgrad(a.o1, dh, Ns, w.f1, true));
    B200XF_TRY(dense_dgrad(dh, Ns, w.f1, other, dz, 1));
    std::swap(cur, other);
    B200XF_TRY(ln_bwd(cur, a.z1, st, st + Ns, w.ln1, w.o.b, Ns, 2 + 2 * l));
    B200XF_TRY(dense_wgrad(a.ctx, dsub, Ns, w.o));
    B200XF_TRY(dense_dgrad(dsub, Ns, w.o, dctx, nullptr, 0));
    AttnArgs at;
    at.q = a.qkv; at.k = a.qkv + d; at.v = a.qkv + 2 * d; at.ldq = at.ldk = at.ldv = 3 * d;
    at.o = a.ctx; at.ldo = d; at.lse = a.lse; at.key_tokens = inp; at.causal = 0;
    at.B = B; at.H = H; at.Sq = S; at.Sk = S; at.dk = dk;
    B200XF_TRY(sb_write(sb_dqkv, &dqkv));
    at.d_o = dctx; at.lddo = d; at.dq = dqkv; at.dk_out = dqkv + d; at.dv_out = dqkv + 2 * d; at.lddq = at.lddk = at.lddv = 3 * d;
    at.delta = delta; at.dq_acc = dq_acc;
    at.early = (weights_static & 2) ? 3 : 0;
    B200XF_TRY(attention_bwd(at, stream));
    B200XF_TRY(dense_wgrad(x, dqkv, Ns, w.qkv, true));
    B200XF_TRY(dense_dgrad(dqkv, Ns, w.qkv, other, dz, 1));
    std::swap(cur, other);
    B200XF_TRY(section_done());
  }
  B200XF_TRY(embed_bwd(inp, cur, grads + enc_emb, hp.input_vocabulary_size, Ns, d, emb_scratch, emb_scratch_bytes, stream, site_drop(0)));
  B200XF_TRY(section_done());
  B200XF_TRY(join_side());
  B200XF_TRY(flush_reductions());
  return 0;
}

// Called each time backward finishes one layout section.  In a train step, completed sections are grouped into
// buckets of >= bucket_bytes; each bucket is all-reduced (when a communicator is attached) on the side stream, after
// an event on the compute stream, and its Adam update is chained behind the reduce, so both hide under the rest of
// the backward pass (whose small kernels leave HBM bandwidth idle).
int Context::section_done() {
  const size_t idx = next_section++;
  if (!overlap_update) return 0;
  const size_t end = sections[idx].second;
  const bool last = (idx + 1 == sections.size());
  // With a communicator a bucket must amortise a collective's latency (>= 32 MB); alone, every layout section is its
  // own bucket (several small optimizer launches disturb the backward pass less than a few large ones: 4.10 vs 4.13 ms)
  // (smaller buckets towards the end of backward, to shorten what is left queued behind the collectives when backward
  // ends, measured slower on 2 GPUs: 4.63 ms against 4.46 -- more collective launches, each disturbing the GEMMs)
  const size_t want = bucket_want_bytes();
  if (!last && (end - bucket_begin) * sizeof(float) < want) return 0;
  // The bucket's gradients come from the compute stream (LayerNorm partials, embedding rows) and from the side stream
  // (weight gradients, bias partials).  Default: the compute stream waits for the side stream here and sums the partials
  // itself.  B200XF_BUCKET_JOIN=0 lets the update stream wait for both instead, so the compute stream never stops --
  // measured slower (4.19 vs 4.13 ms/step at Transformer-base): the optimizer kernel then lands on top of the next
  // layer's GEMMs, whose CTAs need whole SMs and wait for every optimizer CTA that was dispatched first.
  if (bucket_join) {  // B200XF_BUCKET_JOIN=1: the compute stream waits for the side stream and sums the partials itself
    B200XF_TRY(join_side());
    B200XF_TRY(flush_reductions());
  }
  B200XF_CUDA(cudaEventRecord(ev_ready, stream));
  B200XF_CUDA(cudaStreamWaitEvent(comm_stream, ev_ready, 0));
  if (side_on && side_dirty) {
    B200XF_CUDA(cudaEventRecord(ev_join, side_stream));
    B200XF_CUDA(cudaStreamWaitEvent(comm_stream, ev_join, 0));
  }
  B200XF_TRY(flush_reductions(comm_stream));
  const size_t n = end - bucket_begin;
  NcclApi* api = comm ? nccl_api() : nullptr;
  if (nvls_on && sharded_active() && world > 1 && n % (size_t(world) * 64) == 0) {
    // One kernel reduces this rank's shard over all replicas through the NVSwitch, applies Adam and broadcasts the new weights
    // (nvls_update.cu); a one-CTA kernel waits for the other replicas; the local bf16 refresh / gradient clear follows.
    const size_t sh = n / world, off = bucket_begin + size_t(rank) * sh;
    const bool trailing = idx + 2 >= sections.size();   // nothing (or little) of backward is left to disturb
    B200XF_TRY(nvls_reduce_adam_broadcast(nvls, params, adam_m, adam_v, off, sh, iterations_dev, adam, nvls_words, nvls_words + 1,
                                          trailing ? nvls_tail_ctas : nvls_ctas, comm_stream));
    B200XF_TRY(refresh_shadow_zero_grad(params + bucket_begin, shadow + bucket_begin, grads + bucket_begin, n, comm_stream));
  } else if (sharded_active() && api->ReduceScatter && api->AllGather && n % (size_t(world) * 64) == 0) {
    // all-reduce = reduce-scatter + all-gather: put the Adam update between the two halves, on this rank's 1/world
    // shard only (same bytes on the wire, 1/world of the optimizer's HBM traffic), then gather the updated fp32
    // weights and refresh the bf16 operand copy / clear the gradients of the whole bucket in one pass
    const size_t sh = n / world, off = bucket_begin + size_t(rank) * sh;
    ncclComm_t c = static_cast<ncclComm_t>(comm);
    if (pipelined_update == 1 || (pipelined_update == 2 && idx + 2 >= sections.size())) {
      // Collectives stay on the communicator's stream, Adam and the refresh pass go to a second one, and a bucket's
      // all-gather is issued only after the NEXT bucket's reduce-scatter:  RS(k) | RS(k+1) + Adam(k) | AG(k) + Adam(k+1) |
      // AG(k+1) + refresh(k) | refresh(k+1).  What is left when backward ends is two buckets' collectives back to back
      // instead of two full reduce-scatter -> Adam -> all-gather -> refresh chains.
      B200XF_NCCL(api->ReduceScatter(grads + bucket_begin, grads + off, sh, ncclFloat32, ncclSum, c, comm_stream));
      B200XF_CUDA(cudaEventRecord(ev_rs, comm_stream));
      B200XF_CUDA(cudaStreamWaitEvent(upd_stream, ev_rs, 0));
      B200XF_TRY(adam_step(params + off, grads + off, adam_m + off, adam_v + off, nullptr, sh, iterations_dev, adam, 0, upd_stream));
      B200XF_CUDA(cudaEventRecord(ev_adam[adam_slot], upd_stream));  // this bucket's Adam (two events, alternating per bucket)
      if (pending_gather.valid) B200XF_TRY(flush_pending_gather());  // the PREVIOUS bucket's gather, behind this bucket's reduce-scatter
      pending_gather = PendingGather{true, bucket_begin, off, sh, n, adam_slot};
      adam_slot ^= 1;
      upd_dirty = true;
    } else {
      B200XF_NCCL(api->ReduceScatter(grads + bucket_begin, grads + off, sh, ncclFloat32, ncclSum, c, comm_stream));
      B200XF_TRY(adam_step(params + off, grads + off, adam_m + off, adam_v + off, nullptr, sh, iterations_dev, adam, 0, comm_stream));
      B200XF_NCCL(api->AllGather(params + off, params + bucket_begin, sh, ncclFloat32, c, comm_stream));
      B200XF_TRY(refresh_shadow_zero_grad(params + bucket_begin, shadow + bucket_begin, grads + bucket_begin, n, comm_stream));
    }
  } else {
    if (comm) B200XF_TRY(allreduce(grads + bucket_begin, n, comm_stream));
    B200XF_TRY(adam_step(params + bucket_begin, grads + bucket_begin, adam_m + bucket_begin, adam_v + bucket_begin,
                         shadow + bucket_begin, n, iterations_dev, adam, 1, comm_stream));
  }
  if (last && pending_gather.valid) B200XF_TRY(flush_pending_gather());
  // (a bucket that did not take the pipelined path while a gather is pending cannot occur: the pipelined buckets are
  // either all of them or the trailing ones)
  bucket_begin = end;
  return 0;
}

// The optimizer buckets of a train step: consecutive layout sections grouped until a bucket holds bucket_want_bytes().
std::vector<std::pair<size_t, size_t>> Context::bucket_ranges() const {
  std::vector<std::pair<size_t, size_t>> out;
  size_t begin = 0;
  for (size_t i = 0; i < sections.size(); ++i) {
    const size_t end = sections[i].second;
    if (i + 1 == sections.size() || (end - begin) * sizeof(float) >= bucket_want_bytes()) { out.push_back({begin, end}); begin = end; }
  }
  return out;
}

int Context::gather_optimizer_state() {
  if (opt_state_complete || !sharded_active() || world == 1) { opt_state_complete = true; return 0; }
  NcclApi* api = nccl_api();
  ncclComm_t c = static_cast<ncclComm_t>(comm);
  B200XF_CUDA(cudaEventRecord(ev_ready, stream));
  B200XF_CUDA(cudaStreamWaitEvent(comm_stream, ev_ready, 0));
  for (const auto& b : bucket_ranges()) {
    const size_t n = b.second - b.first;
    if (n % (size_t(world) * 64)) continue;  // such a bucket took the replicated update: complete everywhere already
    const size_t sh = n / world, off = b.first + size_t(rank) * sh;
    B200XF_NCCL(api->AllGather(adam_m + off, adam_m + b.first, sh, ncclFloat32, c, comm_stream));
    B200XF_NCCL(api->AllGather(adam_v + off, adam_v + b.first, sh, ncclFloat32, c, comm_stream));
  }
  B200XF_CUDA(cudaEventRecord(ev_done, comm_stream));
  B200XF_CUDA(cudaStreamWaitEvent(stream, ev_done, 0));
  opt_state_complete = true;
  return 0;
}

// The all-gather of the bucket whose Adam was queued last (after that Adam), then its bf16 refresh / gradient clear.
int Context::flush_pending_gather() {
  const PendingGather g = pending_gather;
  pending_gather.valid = false;
  NcclApi* api = nccl_api();
  B200XF_CUDA(cudaStreamWaitEvent(comm_stream, ev_adam[g.slot], 0));
  B200XF_NCCL(api->AllGather(params + g.off, params + g.begin, g.sh, ncclFloat32, static_cast<ncclComm_t>(comm), comm_stream));
  B200XF_CUDA(cudaEventRecord(ev_ag, comm_stream));
  B200XF_CUDA(cudaStreamWaitEvent(upd_stream, ev_ag, 0));
  B200XF_TRY(refresh_shadow_zero_grad(params + g.begin, shadow + g.begin, grads + g.begin, g.n, upd_stream));
  upd_dirty = true;
  return 0;
}

// ------------------------------------------------------------------------------------ steps
int Context::phase_mark(int i) {
  if (!phase_timing) return 0;
  if (i == 0 && phase_n >= 0 && ev_phase[0]) {
    // fold in the previous step's intervals (its events have long completed when the next step starts)
    if (cudaEventQuery(ev_phase[4]) == cudaSuccess) {
      for (int k = 0; k < 4; ++k) { float ms = 0.f; if (cudaEventElapsedTime(&ms, ev_phase[k], ev_phase[k + 1]) == cudaSuccess) phase_ms[k] += ms; }
      ++phase_n;
    }
  }
  if (!ev_phase[i]) B200XF_CUDA(cudaEventCreate(&ev_phase[i]));
  B200XF_CUDA(cudaEventRecord(ev_phase[i], stream));
  return 0;
}

int Context::loss_and_grad(int Nt, int workers_count, int write_grad) {
  const float inv_w = 1.0f / float(workers_count);
  B200XF_TRY(vocab_ce_fwd_bwd(logits, final_layer.ld, tok_labels, Nt, final_layer.N, nonpad_count, inv_w, row_loss, row_correct,
                              write_grad, stream));
  B200XF_TRY(ce_finalize(row_loss, row_correct, nonpad_count, Nt, inv_w, metrics, stream));
  return 0;
}

void Context::drop_graphs() {
  for (auto& kv : graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  graphs.clear();
}

int Context::step_device(const long long* inp, const long long* target, int B, int S, int T1, int workers_count, int mode,
                         float* out6) {
  // mode 0: eval (fwd + loss), 1: fwd + bwd, 2: full train step
  if (T1 < 2) return set_error("target needs at least 2 positions");
  B200XF_TRY(check_shapes(B, S, T1 - 1));
  if (workers_count <= 0) return set_error("workers_count must be positive");
  B200XF_CUDA(cudaSetDevice(device));
  if (mode == 2 && use_graph && !phase_timing) {
    // the graph reads the tokens from the context's own buffers
    if (inp != tok_inp) B200XF_CUDA(cudaMemcpyAsync(tok_inp, inp, size_t(B) * S * sizeof(long long), cudaMemcpyDeviceToDevice, stream));
    if (target != tok_target) B200XF_CUDA(cudaMemcpyAsync(tok_target, target, size_t(B) * T1 * sizeof(long long), cudaMemcpyDeviceToDevice, stream));
    StepGraph& g = graphs[{B, S, T1, workers_count}];
    if (!g.exec && ++g.seen > 2) {
      if (graphs.size() > 64) { const StepGraph keep = g; drop_graphs(); graphs[{B, S, T1, workers_count}] = keep; }
      StepGraph& gg = graphs[{B, S, T1, workers_count}];
      cudaGraph_t graph = nullptr;
      const long long launches_before = launches_total();
      if (cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
        const int rc = step_body(tok_inp, tok_target, B, S, T1, workers_count, mode);
        gg.launches = int(launches_total() - launches_before);  // kernels recorded into the graph
        const cudaError_t e = cudaStreamEndCapture(stream, &graph);
        if (rc == 0 && e == cudaSuccess && graph && cudaGraphInstantiate(&gg.exec, graph, 0) == cudaSuccess) {
          cudaGraphDestroy(graph);
        } else {
          if (graph) cudaGraphDestroy(graph);
          count_launch(-gg.launches);  // nothing was recorded after all
          cudaGetLastError();       // capture is not possible here (e.g. a collective library that cannot be captured):
          gg.exec = nullptr;        // stay on stream launches from now on
          use_graph = false;
        }
      } else {
        cudaGetLastError();
        use_graph = false;
      }
    }
    StepGraph& cur = graphs[{B, S, T1, workers_count}];
    if (cur.exec) {
      // The captured step starts from cleared gradients (each step's optimizer pass leaves them so) and therefore holds
      // no memset; a forward_backward call in between leaves gradients behind, which are cleared here, outside the graph.
      if (!grads_clean) B200XF_CUDA(cudaMemsetAsync(grads, 0, total_params * sizeof(float), stream));
      B200XF_CUDA(cudaGraphLaunch(cur.exec, stream));
      if (cur.counted) count_launch(cur.launches); else cur.counted = true;  // the capture itself already counted them once
      grads_clean = true;
    } else {
      B200XF_TRY(step_body(tok_inp, tok_target, B, S, T1, workers_count, mode));
    }
  } else {
    B200XF_TRY(step_body(inp, target, B, S, T1, workers_count, mode));
  }
  if (out6) B200XF_CUDA(cudaMemcpyAsync(out6, metrics, 6 * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  if (mode == 2 && sharded_active() && world > 1) opt_state_complete = false;  // m / v now current on the local shards only
  return 0;
}

int Context::step_body(const long long* inp, const long long* target, int B, int S, int T1, int workers_count, int mode) {
  const int T = T1 - 1;
  B200XF_TRY(phase_mark(0));
  B200XF_TRY(prep_targets(target, tok_tgt_in, tok_labels, nonpad_count, B, T1, stream));
  B200XF_TRY(forward(inp, tok_tgt_in, B, S, T, nullptr, mode != 0));
  B200XF_TRY(phase_mark(1));
  B200XF_TRY(loss_and_grad(B * T, workers_count, mode != 0));
  B200XF_TRY(phase_mark(2));
  if (mode >= 1) {
    overlap_update = (mode == 2) && (comm != nullptr || overlap_single);
    bucket_begin = 0;
    pending_gather.valid = false;
    B200XF_TRY(backward(inp, tok_tgt_in, B, S, T));
  }
  B200XF_TRY(phase_mark(3));
  if (mode == 2) {
    // metrics: local values stay in metrics[0..2]; the cross-replica sums go to metrics[3..5]
    B200XF_CUDA(cudaMemcpyAsync(metrics + 3, metrics, 3 * sizeof(float), cudaMemcpyDeviceToDevice, stream));
    if (comm) {
      B200XF_CUDA(cudaEventRecord(ev_ready, stream));
      B200XF_CUDA(cudaStreamWaitEvent(comm_stream, ev_ready, 0));
      B200XF_TRY(allreduce(metrics + 3, 3, comm_stream));
    }
    if (overlap_update) {
      if (upd_dirty) {  // the second update stream rejoins too
        B200XF_CUDA(cudaEventRecord(ev_ag, upd_stream));
        B200XF_CUDA(cudaStreamWaitEvent(stream, ev_ag, 0));
        upd_dirty = false;
      }
      B200XF_CUDA(cudaEventRecord(ev_done, comm_stream));
      B200XF_CUDA(cudaStreamWaitEvent(stream, ev_done, 0));  // every bucket's update (and the metrics) done before what follows
    } else {
      B200XF_TRY(adam_step(params, grads, adam_m, adam_v, shadow, total_params, iterations_dev, adam, 1, stream));
    }
    grads_clean = true;
    B200XF_TRY(increment_counter(iterations_dev, stream));
  } else {
    B200XF_CUDA(cudaMemcpyAsync(metrics + 3, metrics, 3 * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  }
  B200XF_TRY(phase_mark(4));
  return 0;
}

int Context::step_host(const long long* inp, const long long* target, int B, int S, int T1, int workers_count, float* out6_host) {
  if (T1 < 2) return set_error("target needs at least 2 positions");
  B200XF_TRY(check_shapes(B, S, T1 - 1));
  B200XF_CUDA(cudaSetDevice(device));
  B200XF_CUDA(cudaMemcpyAsync(tok_inp, inp, size_t(B) * S * sizeof(long long), cudaMemcpyHostToDevice, stream));
  B200XF_CUDA(cudaMemcpyAsync(tok_target, target, size_t(B) * T1 * sizeof(long long), cudaMemcpyHostToDevice, stream));
  B200XF_TRY(step_device(tok_inp, tok_target, B, S, T1, workers_count, 2, nullptr));
  B200XF_CUDA(cudaMemcpyAsync(host_out, metrics, 6 * sizeof(float), cudaMemcpyDeviceToHost, stream));
  B200XF_CUDA(cudaStreamSynchronize(stream));
  memcpy(out6_host, host_out, 6 * sizeof(float));
  return 0;
}

}  // namespace b200xf
