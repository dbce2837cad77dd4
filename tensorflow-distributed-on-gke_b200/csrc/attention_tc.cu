// tcgen05 attention for sequences that fit one 128 x 128 tile (Sq, Sk <= 128, head size 64 — the Transformer-base
// shape): one CTA per (batch, head) does the WHOLE backward of scaled_dot_product_attention
// (transformer_model.py:73-109 with the masks of :56-70,:350-363) with five tensor-core products and no recomputation:
//     S  = Q K^T          dP = dO V^T                      (tcgen05.mma, accumulators in TMEM)
//     P  = softmax(S * scale + mask * -1e9)   delta_i = sum_j P_ij dP_ij   dS = P * (dP - delta)     (one thread per row)
//     dV = P^T dO         dK = scale * dS^T Q              dQ = scale * dS K
// Q, K, V, dO arrive by TMA as 128B-swizzled [row][64] tiles; P and dS are written once to shared memory in the same
// swizzled [q][key] layout and then consumed BOTH as a K-major operand (dS K) and as an MN-major operand (P^T dO, dS^T Q)
// purely through the UMMA descriptor's major bits — nothing is ever transposed.  delta is the exact fp32 row sum of
// P * dP (no rounded O involved), and softmax statistics are recomputed in-thread, so no saved LSE is needed.
#include <cuda.h>
#include <cuda_runtime.h>

#include <mutex>

#include "internal.h"
#include "ptx.cuh"

namespace b200xf {

namespace {

constexpr int kT = 128;          // tile rows (queries) and columns (keys)
constexpr int kD = 64;           // head size
constexpr int kTileBytes = kT * kD * 2;          // 16 KB: [128 rows][64 bf16], 128B swizzle
constexpr int kPBytes = kT * kT * 2;             // 32 KB: two [128 rows][64 keys] halves
constexpr int kThreads = 192;    // warp 0: TMA + MMA issue, warp 1: TMEM allocation, warps 2-5: one thread per row
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kMasked = -1.0e9f * 1.4426950408889634f;

struct TcArgs {
  const long long* key_tokens;
  bf16 *dq, *dk_out, *dv_out; int lddq, lddk, lddv;
  bf16* o; int ldo; float* lse;   // forward only
  int H, Sq, Sk, causal;
  float scale, scale_log2;
};

// smem: sQ sK sV sdO (16 KB each) | sP (32 KB) | sdS (32 KB) | barriers
constexpr int kSmemBwd = 4 * kTileBytes + 2 * kPBytes + 1024 /*alignment slack*/ + 1024 /*barriers, TMEM slot, key mask*/;

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* f) {
  uint32_t v[32];
  ptx::tmem_ld_32x32(taddr, v);
  ptx::tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
}

// Row `row` of a [128][128] bf16 operand tile kept as two 128B-swizzled [128][64] halves: write 32 columns [c0, c0+32).
__device__ __forceinline__ void store_row32(uint8_t* base, int row, int c0, const float* f) {
  const int half = c0 >> 6, chunk0 = (c0 & 63) >> 3;  // 16-byte chunk index within the 128-byte row
  uint8_t* r = base + half * kTileBytes + row * 128;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    uint4 u;
    u.x = ptx::pack_bf16x2(f[q * 8], f[q * 8 + 1]); u.y = ptx::pack_bf16x2(f[q * 8 + 2], f[q * 8 + 3]);
    u.z = ptx::pack_bf16x2(f[q * 8 + 4], f[q * 8 + 5]); u.w = ptx::pack_bf16x2(f[q * 8 + 6], f[q * 8 + 7]);
    *reinterpret_cast<uint4*>(r + (((chunk0 + q) ^ (row & 7)) << 4)) = u;
  }
}

__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                   const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_do, const TcArgs p) {
  extern __shared__ uint8_t smem_raw[];
  // round up to 1024 B with pointer arithmetic (not an integer round trip) so the compiler keeps treating every
  // access below as shared-memory (STS/LDS), not generic
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem; uint8_t* sK = sQ + kTileBytes; uint8_t* sV = sK + kTileBytes; uint8_t* sdO = sV + kTileBytes;
  uint8_t* sP = sdO + kTileBytes; uint8_t* sdS = sP + kPBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sdS + kPBytes);  // 0 load, 1 S/dP ready, 2 P/dS written, 3 outputs ready
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  int* s_mask = reinterpret_cast<int*>(bars + 6);               // [128] key mask: 0 keep, 1 PAD, 2 past Sk

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x, b = blockIdx.y;
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_q); ptx::prefetch_tmap(&tm_k); ptx::prefetch_tmap(&tm_v); ptx::prefetch_tmap(&tm_do);
    ptx::mbar_init(&bars[0], 1); ptx::mbar_init(&bars[1], 1); ptx::mbar_init(&bars[2], 128); ptx::mbar_init(&bars[3], 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) { ptx::tmem_alloc(tmem_slot, 512); ptx::tmem_relinquish(); }
  pdl_wait();   // everything above overlaps the tail of the previous kernel
  pdl_trigger();
  if (warp >= 2) {
    const int j = threadIdx.x - 64;
    s_mask[j] = (j >= p.Sk) ? 2 : (p.key_tokens[size_t(b) * p.Sk + j] == 0 ? 1 : 0);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + 128, tdQ = tmem + 256, tdK = tmem + 320, tdV = tmem + 384;

  if (warp == 0) {
    if (lane == 0) {
      ptx::mbar_expect_tx(&bars[0], 4 * kTileBytes);
      ptx::tma_load_2d(sQ, &tm_q, &bars[0], h * kD, b * p.Sq);
      ptx::tma_load_2d(sdO, &tm_do, &bars[0], h * kD, b * p.Sq);
      ptx::tma_load_2d(sK, &tm_k, &bars[0], h * kD, b * p.Sk);
      ptx::tma_load_2d(sV, &tm_v, &bars[0], h * kD, b * p.Sk);
      ptx::mbar_wait(&bars[0], 0);
      ptx::tc_fence_after();
      // S = Q K^T, dP = dO V^T : A, B K-major, 4 k-steps of 16 over the head dimension
      constexpr uint32_t idesc_kk = ptx::make_idesc_bf16(128, 128, 0, 0);
      const uint32_t aQ = ptx::smem_u32(sQ), aK = ptx::smem_u32(sK), aV = ptx::smem_u32(sV), aDO = ptx::smem_u32(sdO);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        ptx::umma_bf16(tS, ptx::make_smem_desc(aQ + k * 32, 16, 1024), ptx::make_smem_desc(aK + k * 32, 16, 1024), idesc_kk, k > 0);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        ptx::umma_bf16(tdP, ptx::make_smem_desc(aDO + k * 32, 16, 1024), ptx::make_smem_desc(aV + k * 32, 16, 1024), idesc_kk, k > 0);
      ptx::umma_commit(&bars[1]);

      ptx::mbar_wait(&bars[2], 0);  // P and dS are in shared memory
      ptx::tc_fence_after();
      const uint32_t aP = ptx::smem_u32(sP), aDS = ptx::smem_u32(sdS);
      constexpr uint32_t idesc_mm = ptx::make_idesc_bf16(128, 64, 1, 1);  // A = [k=q][m=key], B = [k=q][n=dk]
      constexpr uint32_t idesc_km = ptx::make_idesc_bf16(128, 64, 0, 1);  // A = [m=q][k=key], B = [k=key][n=dk]
#pragma unroll
      for (int k = 0; k < 8; ++k)  // dV = P^T dO : 8 k-steps of 16 query rows (2048 B apart in both operands)
        ptx::umma_bf16(tdV, ptx::make_smem_desc(aP + k * 2048, kTileBytes, 1024), ptx::make_smem_desc(aDO + k * 2048, kTileBytes, 1024), idesc_mm, k > 0);
#pragma unroll
      for (int k = 0; k < 8; ++k)  // dK = dS^T Q
        ptx::umma_bf16(tdK, ptx::make_smem_desc(aDS + k * 2048, kTileBytes, 1024), ptx::make_smem_desc(aQ + k * 2048, kTileBytes, 1024), idesc_mm, k > 0);
#pragma unroll
      for (int k = 0; k < 8; ++k)  // dQ = dS K : 8 k-steps of 16 keys; keys 64.. live in the second half of dS
        ptx::umma_bf16(tdQ, ptx::make_smem_desc(aDS + (k >> 2) * kTileBytes + (k & 3) * 32, 16, 1024),
                       ptx::make_smem_desc(aK + k * 2048, kTileBytes, 1024), idesc_km, k > 0);
      ptx::umma_commit(&bars[3]);
    }
  } else if (warp >= 2) {
    const int quarter = warp & 3;                 // TMEM lane quarter of this warp
    const int i = quarter * 32 + lane;            // query row (and, in the epilogue, key row) of this thread
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    ptx::mbar_wait(&bars[1], 0);
    ptx::tc_fence_after();
    const bool row_ok = i < p.Sq;
    // pass 1: masked, scaled scores (log2 domain) -> row max and sum
    float s[kT];
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      tmem_ld32(tS + lane_base + c * 32, s + c * 32);
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int key = c * 32 + j;
        const int mk = s_mask[key];
        float v = s[key] * p.scale_log2;
        if (mk == 1 || (p.causal && key > i)) v = kMasked;
        if (mk == 2) v = -INFINITY;
        s[key] = v;
        mx = fmaxf(mx, v);
      }
    }
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < kT; ++j) { s[j] = exp2f(s[j] - mx); sum += s[j]; }
    const float inv = row_ok ? 1.0f / sum : 0.f;  // rows past Sq contribute nothing to dK / dV
#pragma unroll
    for (int j = 0; j < kT; ++j) s[j] *= inv;
    // pass 2: delta = sum_j P dP (exact, fp32)
    float delta = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float dp[32];
      tmem_ld32(tdP + lane_base + c * 32, dp);
#pragma unroll
      for (int j = 0; j < 32; ++j) delta += s[c * 32 + j] * dp[j];
    }
    // pass 3: write P and dS = P (dP - delta) to shared memory (bf16, swizzled operand layout)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float dp[32];
      tmem_ld32(tdP + lane_base + c * 32, dp);
      store_row32(sP, i, c * 32, s + c * 32);
#pragma unroll
      for (int j = 0; j < 32; ++j) dp[j] = s[c * 32 + j] * (dp[j] - delta);
      store_row32(sdS, i, c * 32, dp);
    }
    ptx::tc_fence_before();
    ptx::fence_proxy_async();
    ptx::mbar_arrive(&bars[2]);

    ptx::mbar_wait(&bars[3], 0);
    ptx::tc_fence_after();
    // epilogue: thread i owns query row i of dQ and key row i of dK, dV (64 fp32 each in TMEM)
    // (the TMEM loads are warp-collective: every thread executes them, only rows inside the sequence store)
    auto store_row = [&](uint32_t taddr, bf16* dst, float mul, bool ok) {
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float f[32];
        tmem_ld32(taddr + lane_base + c * 32, f);
        if (ok) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 u;
            u.x = ptx::pack_bf16x2(f[q * 8] * mul, f[q * 8 + 1] * mul); u.y = ptx::pack_bf16x2(f[q * 8 + 2] * mul, f[q * 8 + 3] * mul);
            u.z = ptx::pack_bf16x2(f[q * 8 + 4] * mul, f[q * 8 + 5] * mul); u.w = ptx::pack_bf16x2(f[q * 8 + 6] * mul, f[q * 8 + 7] * mul);
            *reinterpret_cast<uint4*>(dst + c * 32 + q * 8) = u;
          }
        }
      }
    };
    const bool key_ok = i < p.Sk;
    store_row(tdQ, p.dq + (size_t(b) * p.Sq + (row_ok ? i : 0)) * p.lddq + h * kD, p.scale, row_ok);
    store_row(tdK, p.dk_out + (size_t(b) * p.Sk + (key_ok ? i : 0)) * p.lddk + h * kD, p.scale, key_ok);
    store_row(tdV, p.dv_out + (size_t(b) * p.Sk + (key_ok ? i : 0)) * p.lddv + h * kD, 1.0f, key_ok);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}

// ---------------------------------------------------------------------------------------------------------- forward
// S = Q K^T (TMEM) -> one thread per query row: mask, max, exp2, sum -> P (bf16, swizzled smem) -> O = P V (TMEM) -> O / sum.
// 80 KB of shared memory and 256 TMEM columns, so two CTAs share an SM.
constexpr int kSmemFwd = 3 * kTileBytes + kPBytes + 1024 + 1024;

__global__ void __launch_bounds__(kThreads, 2)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                   const __grid_constant__ CUtensorMap tm_v, const TcArgs p) {
  extern __shared__ uint8_t smem_raw[];
  // round up to 1024 B with pointer arithmetic (not an integer round trip) so the compiler keeps treating every
  // access below as shared-memory (STS/LDS), not generic
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem; uint8_t* sK = sQ + kTileBytes; uint8_t* sV = sK + kTileBytes; uint8_t* sP = sV + kTileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kPBytes);  // 0 load, 1 S ready, 2 P written, 3 O ready
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  int* s_mask = reinterpret_cast<int*>(bars + 6);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x, b = blockIdx.y;
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_q); ptx::prefetch_tmap(&tm_k); ptx::prefetch_tmap(&tm_v);
    ptx::mbar_init(&bars[0], 1); ptx::mbar_init(&bars[1], 1); ptx::mbar_init(&bars[2], 128); ptx::mbar_init(&bars[3], 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) { ptx::tmem_alloc(tmem_slot, 256); ptx::tmem_relinquish(); }
  pdl_wait();
  pdl_trigger();
  if (warp >= 2) {
    const int j = threadIdx.x - 64;
    s_mask[j] = (j >= p.Sk) ? 2 : (p.key_tokens[size_t(b) * p.Sk + j] == 0 ? 1 : 0);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + 128;

  if (warp == 0) {
    if (lane == 0) {
      ptx::mbar_expect_tx(&bars[0], 3 * kTileBytes);
      ptx::tma_load_2d(sQ, &tm_q, &bars[0], h * kD, b * p.Sq);
      ptx::tma_load_2d(sK, &tm_k, &bars[0], h * kD, b * p.Sk);
      ptx::tma_load_2d(sV, &tm_v, &bars[0], h * kD, b * p.Sk);
      ptx::mbar_wait(&bars[0], 0);
      ptx::tc_fence_after();
      constexpr uint32_t idesc_kk = ptx::make_idesc_bf16(128, 128, 0, 0);
      const uint32_t aQ = ptx::smem_u32(sQ), aK = ptx::smem_u32(sK), aV = ptx::smem_u32(sV), aP = ptx::smem_u32(sP);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        ptx::umma_bf16(tS, ptx::make_smem_desc(aQ + k * 32, 16, 1024), ptx::make_smem_desc(aK + k * 32, 16, 1024), idesc_kk, k > 0);
      ptx::umma_commit(&bars[1]);
      ptx::mbar_wait(&bars[2], 0);
      ptx::tc_fence_after();
      constexpr uint32_t idesc_km = ptx::make_idesc_bf16(128, 64, 0, 1);  // A = P [m=q][k=key], B = V [k=key][n=dk]
#pragma unroll
      for (int k = 0; k < 8; ++k)
        ptx::umma_bf16(tO, ptx::make_smem_desc(aP + (k >> 2) * kTileBytes + (k & 3) * 32, 16, 1024),
                       ptx::make_smem_desc(aV + k * 2048, kTileBytes, 1024), idesc_km, k > 0);
      ptx::umma_commit(&bars[3]);
    }
  } else if (warp >= 2) {
    const int quarter = warp & 3;
    const int i = quarter * 32 + lane;
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    ptx::mbar_wait(&bars[1], 0);
    ptx::tc_fence_after();
    const bool row_ok = i < p.Sq;
    float s[kT];
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      tmem_ld32(tS + lane_base + c * 32, s + c * 32);
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int key = c * 32 + j;
        const int mk = s_mask[key];
        float v = s[key] * p.scale_log2;
        if (mk == 1 || (p.causal && key > i)) v = kMasked;
        if (mk == 2) v = -INFINITY;
        s[key] = v;
        mx = fmaxf(mx, v);
      }
    }
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < kT; ++j) { s[j] = exp2f(s[j] - mx); sum += s[j]; }
#pragma unroll
    for (int c = 0; c < 4; ++c) store_row32(sP, i, c * 32, s + c * 32);
    ptx::tc_fence_before();
    ptx::fence_proxy_async();
    ptx::mbar_arrive(&bars[2]);

    ptx::mbar_wait(&bars[3], 0);
    ptx::tc_fence_after();
    const float inv = 1.0f / sum;
    bf16* dst = p.o + (size_t(b) * p.Sq + (row_ok ? i : 0)) * p.ldo + h * kD;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      float f[32];
      tmem_ld32(tO + lane_base + c * 32, f);
      if (row_ok) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 u;
          u.x = ptx::pack_bf16x2(f[q * 8] * inv, f[q * 8 + 1] * inv); u.y = ptx::pack_bf16x2(f[q * 8 + 2] * inv, f[q * 8 + 3] * inv);
          u.z = ptx::pack_bf16x2(f[q * 8 + 4] * inv, f[q * 8 + 5] * inv); u.w = ptx::pack_bf16x2(f[q * 8 + 6] * inv, f[q * 8 + 7] * inv);
          *reinterpret_cast<uint4*>(dst + c * 32 + q * 8) = u;
        }
      }
    }
    if (row_ok && p.lse) reinterpret_cast<float2*>(p.lse)[(size_t(b) * p.H + h) * p.Sq + i] = make_float2(mx, log2f(sum));
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 256); }
}

}  // namespace

bool attention_tc_supported(const AttnArgs& a) {
  static int enabled = -1;
  if (enabled < 0) { const char* e = getenv("B200XF_ATTENTION_TC"); enabled = (e && atoi(e) == 0) ? 0 : 1; }
  return enabled && a.dk == kD && a.Sq <= kT && a.Sk <= kT && a.Sq > 0 && a.Sk > 0;
}

int attention_fwd_tc(const AttnArgs& a, cudaStream_t stream) {
  CUtensorMap tq, tk, tv;
  const uint64_t cols = uint64_t(a.H) * kD;
  B200XF_TRY(tensor_map_2d(a.q, cols, uint64_t(a.B) * a.Sq, a.ldq, kT, &tq, 0));
  B200XF_TRY(tensor_map_2d(a.k, cols, uint64_t(a.B) * a.Sk, a.ldk, kT, &tk, 0));
  B200XF_TRY(tensor_map_2d(a.v, cols, uint64_t(a.B) * a.Sk, a.ldv, kT, &tv, 0));
  TcArgs p{};
  p.key_tokens = a.key_tokens; p.o = a.o; p.ldo = a.ldo; p.lse = a.lse;
  p.H = a.H; p.Sq = a.Sq; p.Sk = a.Sk; p.causal = a.causal;
  p.scale = 1.0f / sqrtf(float(a.dk)); p.scale_log2 = p.scale * kLog2e;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] { attr_err = cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemFwd); });
  if (attr_err != cudaSuccess) return set_error("cudaFuncSetAttribute(attn_fwd_tc): %s", cudaGetErrorString(attr_err));
  B200XF_CUDA(launch_k(attn_fwd_tc_kernel, dim3(a.H, a.B), kThreads, kSmemFwd, stream, tq, tk, tv, p));
  count_launch();
  return 0;
}

int attention_bwd_tc(const AttnArgs& a, cudaStream_t stream) {
  CUtensorMap tq, tk, tv, tdo;
  const uint64_t cols = uint64_t(a.H) * kD;
  B200XF_TRY(tensor_map_2d(a.q, cols, uint64_t(a.B) * a.Sq, a.ldq, kT, &tq, 0));
  B200XF_TRY(tensor_map_2d(a.k, cols, uint64_t(a.B) * a.Sk, a.ldk, kT, &tk, 0));
  B200XF_TRY(tensor_map_2d(a.v, cols, uint64_t(a.B) * a.Sk, a.ldv, kT, &tv, 0));
  B200XF_TRY(tensor_map_2d(a.d_o, cols, uint64_t(a.B) * a.Sq, a.lddo, kT, &tdo, 0));
  TcArgs p{};
  p.key_tokens = a.key_tokens;
  p.dq = a.dq; p.dk_out = a.dk_out; p.dv_out = a.dv_out; p.lddq = a.lddq; p.lddk = a.lddk; p.lddv = a.lddv;
  p.H = a.H; p.Sq = a.Sq; p.Sk = a.Sk; p.causal = a.causal;
  p.scale = 1.0f / sqrtf(float(a.dk)); p.scale_log2 = p.scale * kLog2e;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] { attr_err = cudaFuncSetAttribute(attn_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBwd); });
  if (attr_err != cudaSuccess) return set_error("cudaFuncSetAttribute(attn_bwd_tc): %s", cudaGetErrorString(attr_err));
  B200XF_CUDA(launch_k(attn_bwd_tc_kernel, dim3(a.H, a.B), kThreads, kSmemBwd, stream, tq, tk, tv, tdo, p));
  count_launch();
  return 0;
}

}  // namespace b200xf
