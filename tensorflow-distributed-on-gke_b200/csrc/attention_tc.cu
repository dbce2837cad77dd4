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

#include <algorithm>
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
  int early;   // bit 0: q, bit 1: k and v are not outputs of the kernel launched just before -> loaded ahead of the dependency wait
};

// smem: sQ sK sV sdO (16 KB each) | sP (32 KB) | sdS (32 KB) | barriers
constexpr int kSmemBwd = 4 * kTileBytes + 2 * kPBytes + 1024 /*alignment slack*/ + 4096 /*barriers, TMEM slot, key mask, row exchange*/;
constexpr int kThreads2 = 320;   // warp 0: TMA + MMA issue, warp 1: TMEM allocation, warps 2-9: two threads per row

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

__global__ void __launch_bounds__(kThreads2, 1)
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
  float* xch = reinterpret_cast<float*>(s_mask + kT);           // [3][2][128] partial max / sum / delta of the two column halves

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x, b = blockIdx.y;
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_q); ptx::prefetch_tmap(&tm_k); ptx::prefetch_tmap(&tm_v); ptx::prefetch_tmap(&tm_do);
    ptx::mbar_init(&bars[0], 1); ptx::mbar_init(&bars[1], 1); ptx::mbar_init(&bars[2], 256); ptx::mbar_init(&bars[3], 1);
    ptx::fence_barrier_init();
    ptx::mbar_expect_tx(&bars[0], 4 * kTileBytes);
    // saved forward activations can be on their way while the kernel that produces dO is still finishing
    if (p.early & 1) ptx::tma_load_2d(sQ, &tm_q, &bars[0], h * kD, b * p.Sq);
    if (p.early & 2) {
      ptx::tma_load_2d(sK, &tm_k, &bars[0], h * kD, b * p.Sk);
      ptx::tma_load_2d(sV, &tm_v, &bars[0], h * kD, b * p.Sk);
    }
  }
  if (warp == 1) { ptx::tmem_alloc(tmem_slot, 512); ptx::tmem_relinquish(); }
  pdl_wait();   // everything above overlaps the tail of the previous kernel
  pdl_trigger();
  if (warp == 0 && lane == 0) {  // the rest of the operands, before anything else can delay them
    if (!(p.early & 1)) ptx::tma_load_2d(sQ, &tm_q, &bars[0], h * kD, b * p.Sq);
    ptx::tma_load_2d(sdO, &tm_do, &bars[0], h * kD, b * p.Sq);
    if (!(p.early & 2)) {
      ptx::tma_load_2d(sK, &tm_k, &bars[0], h * kD, b * p.Sk);
      ptx::tma_load_2d(sV, &tm_v, &bars[0], h * kD, b * p.Sk);
    }
  }
  int my_mask = 0;
  if (warp >= 2 && warp < 6) {
    const int j = threadIdx.x - 64;
    my_mask = (j >= p.Sk) ? 2 : (p.key_tokens[size_t(b) * p.Sk + j] == 0 ? 1 : 0);
    s_mask[j] = my_mask;
  }
  ptx::tc_fence_before();
  const bool plain = (__syncthreads_or(my_mask != 0) == 0) && !p.causal;   // no PAD key, Sk = 128, no look-ahead mask: no per-key test needed
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + 128, tdQ = tmem + 256, tdK = tmem + 320, tdV = tmem + 384;

  if (warp == 0) {
    if (lane == 0) {
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
    // two threads per row: warps 2-5 take keys 0..63 (output columns 0..31), warps 6-9 keys 64..127 (columns 32..63);
    // the row max, sum and delta are combined through shared memory
    const int quarter = warp & 3;                 // TMEM lane quarter of this warp
    const int half = (warp - 2) >> 2;
    const int i = quarter * 32 + lane;            // query row (and, in the epilogue, key row) of this thread
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    ptx::mbar_wait(&bars[1], 0);
    ptx::tc_fence_after();
    const bool row_ok = i < p.Sq;
    float s[64], dp[64];
    float mx = -INFINITY;
    {   // this thread's 64 score and 64 dP columns: four TMEM loads behind ONE wait (a load round trip costs ~0.5 us)
      uint32_t rs[64], rp[64];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        ptx::tmem_ld_32x32(tS + lane_base + (half * 2 + c) * 32, rs + c * 32);
        ptx::tmem_ld_32x32(tdP + lane_base + (half * 2 + c) * 32, rp + c * 32);
      }
      ptx::tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 64; ++j) dp[j] = __uint_as_float(rp[j]);
      if (plain) {
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int j = 0; j < 64; ++j) { s[j] = __uint_as_float(rs[j]) * p.scale_log2; m4[j & 3] = fmaxf(m4[j & 3], s[j]); }
        mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      } else {
#pragma unroll
        for (int j = 0; j < 64; ++j) {
          const int key = half * 64 + j;
          const int mk = s_mask[key];
          float v = __uint_as_float(rs[j]) * p.scale_log2;
          if (mk == 1 || (p.causal && key > i)) v = kMasked;
          if (mk == 2) v = -INFINITY;
          s[j] = v;
          mx = fmaxf(mx, v);
        }
      }
    }
    xch[half * kT + i] = mx;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    mx = fmaxf(xch[i], xch[kT + i]);              // finite: key 0 always exists
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 64; ++j) { s[j] = ptx::fast_exp2(s[j] - mx); sum += s[j]; }
    xch[2 * kT + half * kT + i] = sum;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    sum = xch[2 * kT + i] + xch[3 * kT + i];
    const float inv = row_ok ? 1.0f / sum : 0.f;  // rows past Sq contribute nothing to dK / dV
    float delta = 0.f;                            // delta = sum_j P dP (exact, fp32)
#pragma unroll
    for (int j = 0; j < 64; ++j) { s[j] *= inv; delta += s[j] * dp[j]; }
    xch[4 * kT + half * kT + i] = delta;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    delta = xch[4 * kT + i] + xch[5 * kT + i];
    // P and dS = P (dP - delta) -> shared memory (bf16, swizzled operand layout)
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      store_row32(sP, i, half * 64 + c * 32, s + c * 32);
#pragma unroll
      for (int j = 0; j < 32; ++j) dp[c * 32 + j] = s[c * 32 + j] * (dp[c * 32 + j] - delta);
      store_row32(sdS, i, half * 64 + c * 32, dp + c * 32);
    }
    ptx::tc_fence_before();
    ptx::fence_proxy_async();
    ptx::mbar_arrive(&bars[2]);

    ptx::mbar_wait(&bars[3], 0);
    ptx::tc_fence_after();
    // epilogue: row i of dQ (query i) and of dK, dV (key i); this thread converts 32 of the 64 columns
    // (the TMEM loads are warp-collective: every thread executes them, only rows inside the sequence store)
    auto store_row = [&](uint32_t taddr, bf16* dst, float mul, bool ok) {
      float f[32];
      tmem_ld32(taddr + lane_base + half * 32, f);
      if (ok) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 u;
          u.x = ptx::pack_bf16x2(f[q * 8] * mul, f[q * 8 + 1] * mul); u.y = ptx::pack_bf16x2(f[q * 8 + 2] * mul, f[q * 8 + 3] * mul);
          u.z = ptx::pack_bf16x2(f[q * 8 + 4] * mul, f[q * 8 + 5] * mul); u.w = ptx::pack_bf16x2(f[q * 8 + 6] * mul, f[q * 8 + 7] * mul);
          *reinterpret_cast<uint4*>(dst + half * 32 + q * 8) = u;
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
    ptx::mbar_expect_tx(&bars[0], 3 * kTileBytes);
    if (p.early & 1) ptx::tma_load_2d(sQ, &tm_q, &bars[0], h * kD, b * p.Sq);
    if (p.early & 2) {  // cross-attention: K and V of every layer were projected right after the encoder
      ptx::tma_load_2d(sK, &tm_k, &bars[0], h * kD, b * p.Sk);
      ptx::tma_load_2d(sV, &tm_v, &bars[0], h * kD, b * p.Sk);
    }
  }
  if (warp == 1) { ptx::tmem_alloc(tmem_slot, 256); ptx::tmem_relinquish(); }
  pdl_wait();
  pdl_trigger();
  if (warp == 0 && lane == 0) {  // the rest of the operands, before anything else can delay them
    if (!(p.early & 1)) ptx::tma_load_2d(sQ, &tm_q, &bars[0], h * kD, b * p.Sq);
    if (!(p.early & 2)) {
      ptx::tma_load_2d(sK, &tm_k, &bars[0], h * kD, b * p.Sk);
      ptx::tma_load_2d(sV, &tm_v, &bars[0], h * kD, b * p.Sk);
    }
  }
  int my_mask = 0;
  if (warp >= 2) {
    const int j = threadIdx.x - 64;
    my_mask = (j >= p.Sk) ? 2 : (p.key_tokens[size_t(b) * p.Sk + j] == 0 ? 1 : 0);
    s_mask[j] = my_mask;
  }
  ptx::tc_fence_before();
  const bool plain = (__syncthreads_or(my_mask != 0) == 0) && !p.causal;   // no PAD key, Sk = 128, no look-ahead mask: no per-key test needed
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + 128;

  if (warp == 0) {
    if (lane == 0) {
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
    {   // the whole 128-column score row: four TMEM loads behind ONE wait
      uint32_t rs[kT];
#pragma unroll
      for (int c = 0; c < 4; ++c) ptx::tmem_ld_32x32(tS + lane_base + c * 32, rs + c * 32);
      ptx::tmem_ld_wait();
      if (plain) {
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int j = 0; j < kT; ++j) { s[j] = __uint_as_float(rs[j]) * p.scale_log2; m4[j & 3] = fmaxf(m4[j & 3], s[j]); }
        mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      } else {
#pragma unroll
        for (int key = 0; key < kT; ++key) {
          const int mk = s_mask[key];
          float v = __uint_as_float(rs[key]) * p.scale_log2;
          if (mk == 1 || (p.causal && key > i)) v = kMasked;
          if (mk == 2) v = -INFINITY;
          s[key] = v;
          mx = fmaxf(mx, v);
        }
      }
    }
    float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < kT; ++j) { s[j] = ptx::fast_exp2(s[j] - mx); s4[j & 3] += s[j]; }
    const float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
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

constexpr int kThreadsMulti = 320;  // warp 0: TMA + MMA issue, warp 1: TMEM allocation, warps 2-9: two threads per row (64 keys each)

// ---------------------------------------------------------------------------------------------------------- forward, any length
// One CTA per (128-query tile, head, batch) walks the 128-key tiles (TMA double buffered).  Per tile: S = Q K^T (TMEM) ->
// two threads per query row (64 keys each) mask / online-softmax it, exchange their partial max and sum through shared
// memory, write P (bf16) -> O_tile = P V (TMEM) -> each thread folds its 32 output columns into a register accumulator
// with the usual exp2(m_old - m_new) correction.
constexpr int kSmemFwdMulti = kTileBytes /*Q*/ + 4 * kTileBytes /*K V x2*/ + kPBytes /*P*/ + 1024 + 4096;

__global__ void __launch_bounds__(kThreadsMulti, 1)
attn_fwd_tc_multi_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                         const __grid_constant__ CUtensorMap tm_v, const TcArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem; uint8_t* sK = sQ + kTileBytes; uint8_t* sV = sK + 2 * kTileBytes; uint8_t* sP = sV + 2 * kTileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kPBytes);
  uint64_t* bar_q = bars; uint64_t* kv_full = bars + 1; uint64_t* kv_empty = bars + 3;
  uint64_t* s_ready = bars + 5; uint64_t* p_ready = bars + 6; uint64_t* o_ready = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  int* s_mask = reinterpret_cast<int*>(bars + 10);            // [2][128] per K/V stage
  float* x_max = reinterpret_cast<float*>(s_mask + 256);       // [2][128] partial row max per column half
  float* x_sum = x_max + 256;                                  // [2][128] partial row sum per column half

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int q0 = qt * kT;
  const long long* toks = p.key_tokens + size_t(b) * p.Sk;
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_q); ptx::prefetch_tmap(&tm_k); ptx::prefetch_tmap(&tm_v);
    ptx::mbar_init(bar_q, 1);
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(&kv_full[i], 1); ptx::mbar_init(&kv_empty[i], 1); }
    ptx::mbar_init(s_ready, 1); ptx::mbar_init(p_ready, 256); ptx::mbar_init(o_ready, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) { ptx::tmem_alloc(tmem_slot, 256); ptx::tmem_relinquish(); }
  pdl_wait();
  pdl_trigger();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + 128;

  int kv_end = p.Sk;
  if (p.causal && toks[0] != 0) kv_end = min(p.Sk, q0 + kT);   // later key tiles are exactly zero-weight
  const int ntiles = (kv_end + kT - 1) / kT;

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t aQ = ptx::smem_u32(sQ), aP = ptx::smem_u32(sP);
      constexpr uint32_t idesc_kk = ptx::make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_km = ptx::make_idesc_bf16(128, 64, 0, 1);
      auto load_kv = [&](int j) {
        const int st = j & 1;
        ptx::mbar_expect_tx(&kv_full[st], 2 * kTileBytes);
        ptx::tma_load_2d(sK + st * kTileBytes, &tm_k, &kv_full[st], h * kD, b * p.Sk + j * kT);
        ptx::tma_load_2d(sV + st * kTileBytes, &tm_v, &kv_full[st], h * kD, b * p.Sk + j * kT);
      };
      ptx::mbar_expect_tx(bar_q, kTileBytes);
      ptx::tma_load_2d(sQ, &tm_q, bar_q, h * kD, b * p.Sq + q0);
      load_kv(0);
      ptx::mbar_wait(bar_q, 0);
      for (int j = 0; j < ntiles; ++j) {
        const int st = j & 1;
        const uint32_t aK = ptx::smem_u32(sK + st * kTileBytes), aV = ptx::smem_u32(sV + st * kTileBytes);
        ptx::mbar_wait(&kv_full[st], (j >> 1) & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tS, ptx::make_smem_desc(aQ + k * 32, 16, 1024), ptx::make_smem_desc(aK + k * 32, 16, 1024), idesc_kk, k > 0);
        ptx::umma_commit(s_ready);
        if (j + 1 < ntiles) {
          ptx::mbar_wait(&kv_empty[(j + 1) & 1], (((j + 1) >> 1) & 1) ^ 1);
          load_kv(j + 1);
        }
        ptx::mbar_wait(p_ready, j & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k)
          ptx::umma_bf16(tO, ptx::make_smem_desc(aP + (k >> 2) * kTileBytes + (k & 3) * 32, 16, 1024),
                         ptx::make_smem_desc(aV + k * 2048, kTileBytes, 1024), idesc_km, k > 0);
        ptx::umma_commit(&kv_empty[st]);   // this K/V stage may be refilled once these MMAs retire
        ptx::umma_commit(o_ready);
      }
    }
  } else if (warp >= 2) {
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    const int i = quarter * 32 + lane;
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    const int e = threadIdx.x - 64;                              // 0..255
    const bool row_ok = q0 + i < p.Sq;
    float o[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) o[c] = 0.f;
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < ntiles; ++j) {
      const int k0 = j * kT;
      if (e < kT) s_mask[(j & 1) * kT + e] = (k0 + e >= p.Sk) ? 2 : (toks[k0 + e] == 0 ? 1 : 0);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const int* mask = s_mask + (j & 1) * kT;
      ptx::mbar_wait(s_ready, j & 1);
      ptx::tc_fence_after();
      float s[64];
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        tmem_ld32(tS + lane_base + (half * 2 + c) * 32, s + c * 32);
#pragma unroll
        for (int t = 0; t < 32; ++t) {
          const int key = (half * 2 + c) * 32 + t;
          const int mk = mask[key];
          float v = s[c * 32 + t] * p.scale_log2;
          if (mk == 1 || (p.causal && k0 + key > q0 + i)) v = kMasked;
          if (mk == 2) v = -INFINITY;
          s[c * 32 + t] = v;
          mx = fmaxf(mx, v);
        }
      }
      x_max[half * kT + i] = mx;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const float m_new = fmaxf(m, fmaxf(x_max[i], x_max[kT + i]));   // finite: the first tile always holds a real key
      const float corr = ptx::fast_exp2(m - m_new);
      float sum = 0.f;
#pragma unroll
      for (int t = 0; t < 64; ++t) { s[t] = ptx::fast_exp2(s[t] - m_new); sum += s[t]; }
      x_sum[half * kT + i] = sum;
      store_row32(sP, i, half * 64, s);
      store_row32(sP, i, half * 64 + 32, s + 32);
      ptx::tc_fence_before();
      ptx::fence_proxy_async();
      ptx::mbar_arrive(p_ready);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      l = l * corr + (x_sum[i] + x_sum[kT + i]);
      m = m_new;
      ptx::mbar_wait(o_ready, j & 1);
      ptx::tc_fence_after();
      float ot[32];
      tmem_ld32(tO + lane_base + half * 32, ot);
#pragma unroll
      for (int c = 0; c < 32; ++c) o[c] = o[c] * corr + ot[c];
      ptx::tc_fence_before();
    }
    if (row_ok) {
      const float inv = 1.0f / l;
      bf16* dst = p.o + (size_t(b) * p.Sq + q0 + i) * p.ldo + h * kD + half * 32;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = ptx::pack_bf16x2(o[q * 8] * inv, o[q * 8 + 1] * inv); u.y = ptx::pack_bf16x2(o[q * 8 + 2] * inv, o[q * 8 + 3] * inv);
        u.z = ptx::pack_bf16x2(o[q * 8 + 4] * inv, o[q * 8 + 5] * inv); u.w = ptx::pack_bf16x2(o[q * 8 + 6] * inv, o[q * 8 + 7] * inv);
        *reinterpret_cast<uint4*>(dst + q * 8) = u;
      }
      if (half == 0 && p.lse) reinterpret_cast<float2*>(p.lse)[(size_t(b) * p.H + h) * p.Sq + q0 + i] = make_float2(m, log2f(l));
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 256); }
}

// ---------------------------------------------------------------------------------------------------------- forward, any length, two query tiles
// One CTA per (256 queries, head, batch): two 128-query tiles A and B share every K/V tile that TMA brings in, and their
// softmax passes take turns with the tensor core -- while one tile's 128 threads (one per row) exponentiate, the MMA thread
// runs the other tile's P V and the next S = Q K^T.  A thread pulls its whole 128-column score row out of TMEM with four loads
// behind ONE wait (TMEM round trips, not arithmetic, were what an earlier two-pass version spent its time on), finds the row
// maximum, exponentiates, and writes P (bf16) back into TMEM over the score columns it has consumed: the P V product takes its A
// operand straight from TMEM (no shared-memory round trip, no bank conflicts, no proxy fence).  O accumulates in TMEM across the key tiles; when a row's running maximum moves, the warp
// rescales its O rows in place (tcgen05.ld / st) before the next P V is issued -- after the first few key tiles that is rare.
// TMEM: S_A | S_B (128 columns each) | O_A | O_B (64 each).
constexpr bool kPInTmem = true;   // P as the tensor core's A operand straight from TMEM (false: through shared memory)
constexpr int kThreadsPP = 320;   // warp 0: TMA + MMA issue, warp 1: TMEM allocation, warps 2-5: tile A rows, warps 6-9: tile B rows
constexpr int kSmemFwdPP = 2 * kTileBytes /*Q_A Q_B*/ + 4 * kTileBytes /*K V x2*/ + (kPInTmem ? 0 : 2 * kPBytes) /*P_A P_B*/ + 1024 + 4096;

__global__ void __launch_bounds__(kThreadsPP, 1)
attn_fwd_tc_pp_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                      const __grid_constant__ CUtensorMap tm_v, const TcArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem; uint8_t* sK = sQ + 2 * kTileBytes; uint8_t* sV = sK + 2 * kTileBytes; uint8_t* sP = sV + 2 * kTileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + (kPInTmem ? 0 : 2 * kPBytes));
  uint64_t* bar_q = bars; uint64_t* kv_full = bars + 1; uint64_t* kv_empty = bars + 3;
  uint64_t* s_ready = bars + 5; uint64_t* p_ready = bars + 7; uint64_t* o_ready = bars + 9;   // [2] each: tile A, tile B
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 11);
  int* s_mask = reinterpret_cast<int*>(bars + 12);            // [2 tiles][2 stages][128]
  int* any_flag = s_mask + 4 * kT;                             // [2 tiles][2 stages][4 warps]: some key of the tile is masked

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = (p.causal ? int(gridDim.x) - 1 - int(blockIdx.x) : int(blockIdx.x)) * 2 * kT;   // causal: the longest tiles first
  const long long* toks = p.key_tokens + size_t(b) * p.Sk;
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_q); ptx::prefetch_tmap(&tm_k); ptx::prefetch_tmap(&tm_v);
    ptx::mbar_init(bar_q, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&kv_full[i], 1); ptx::mbar_init(&kv_empty[i], 1);
      ptx::mbar_init(&s_ready[i], 1); ptx::mbar_init(&p_ready[i], 128); ptx::mbar_init(&o_ready[i], 1);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) { ptx::tmem_alloc(tmem_slot, 512); ptx::tmem_relinquish(); }
  pdl_wait();
  pdl_trigger();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  // key tiles each query tile has to visit (a causal tile stops at its own diagonal; a tile past Sq visits none)
  int nt[2];
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    const int qg = q0 + g * kT;
    int kv_end = p.Sk;
    if (p.causal && toks[0] != 0) kv_end = min(p.Sk, qg + kT);   // later key tiles are exactly zero-weight
    nt[g] = qg < p.Sq ? (kv_end + kT - 1) / kT : 0;
  }
  const int ntiles = max(nt[0], nt[1]);

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t aQ = ptx::smem_u32(sQ), aP = ptx::smem_u32(sP);
      constexpr uint32_t idesc_kk = ptx::make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_km = ptx::make_idesc_bf16(128, 64, 0, 1);
      auto load_kv = [&](int j) {
        const int st = j & 1;
        ptx::mbar_expect_tx(&kv_full[st], 2 * kTileBytes);
        ptx::tma_load_2d(sK + st * kTileBytes, &tm_k, &kv_full[st], h * kD, b * p.Sk + j * kT);
        ptx::tma_load_2d(sV + st * kTileBytes, &tm_v, &kv_full[st], h * kD, b * p.Sk + j * kT);
      };
      auto issue_s = [&](int g, int j) {   // S_g = Q_g K_j^T
        const uint32_t aK = ptx::smem_u32(sK + (j & 1) * kTileBytes);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tmem + g * 128, ptx::make_smem_desc(aQ + g * kTileBytes + k * 32, 16, 1024), ptx::make_smem_desc(aK + k * 32, 16, 1024), idesc_kk, k > 0);
        ptx::umma_commit(&s_ready[g]);
      };
      auto issue_pv = [&](int g, int j) {  // O_g (+)= P_g V_j
        const uint32_t aV = ptx::smem_u32(sV + (j & 1) * kTileBytes);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (kPInTmem)   // P_g sits in the first 64 columns of S_g's TMEM region: 16 keys = 8 columns per MMA
            ptx::umma_bf16_ts(tmem + 256 + g * 64, tmem + g * 128 + k * 8, ptx::make_smem_desc(aV + k * 2048, kTileBytes, 1024), idesc_km, (j > 0 || k > 0) ? 1u : 0u);
          else
            ptx::umma_bf16(tmem + 256 + g * 64, ptx::make_smem_desc(aP + g * kPBytes + (k >> 2) * kTileBytes + (k & 3) * 32, 16, 1024),
                           ptx::make_smem_desc(aV + k * 2048, kTileBytes, 1024), idesc_km, (j > 0 || k > 0) ? 1u : 0u);
        }
        ptx::umma_commit(&o_ready[g]);
      };
      ptx::mbar_expect_tx(bar_q, (nt[1] > 0 ? 2 : 1) * kTileBytes);
      ptx::tma_load_2d(sQ, &tm_q, bar_q, h * kD, b * p.Sq + q0);
      if (nt[1] > 0) ptx::tma_load_2d(sQ + kTileBytes, &tm_q, bar_q, h * kD, b * p.Sq + q0 + kT);
      if (ntiles > 0) load_kv(0);
      ptx::mbar_wait(bar_q, 0);
      if (ntiles > 0) {
        ptx::mbar_wait(&kv_full[0], 0);
        ptx::tc_fence_after();
        if (nt[0] > 0) issue_s(0, 0);
        if (nt[1] > 0) issue_s(1, 0);
        if (ntiles > 1) load_kv(1);
      }
      for (int j = 0; j < ntiles; ++j) {
        const bool more = j + 1 < ntiles;
        if (more) { ptx::mbar_wait(&kv_full[(j + 1) & 1], ((j + 1) >> 1) & 1); ptx::tc_fence_after(); }
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          if (j < nt[g]) {
            ptx::mbar_wait(&p_ready[g], j & 1);       // tile g has read S_g(j), rescaled O_g if needed and written P_g(j)
            ptx::tc_fence_after();
            issue_pv(g, j);
            if (j + 1 < nt[g]) issue_s(g, j + 1);      // its next S, while the other tile is in its softmax
          }
        }
        ptx::umma_commit(&kv_empty[j & 1]);            // K_j / V_j may be overwritten once the MMAs above retire
        if (j + 2 < ntiles) {
          ptx::mbar_wait(&kv_empty[j & 1], (j >> 1) & 1);
          load_kv(j + 2);
        }
      }
    }
  } else if (warp >= 2) {
    const int g = (warp - 2) >> 2;                      // 0: tile A, 1: tile B
    const int quarter = warp & 3;
    const int i = quarter * 32 + lane;                  // row within the tile
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    const int e = (threadIdx.x - 64) & 127;
    const int qg = q0 + g * kT;
    const bool row_ok = qg + i < p.Sq;
    const uint32_t tS = tmem + g * 128, tO = tmem + 256 + g * 64;
    uint8_t* sPg = sP + g * kPBytes;
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < nt[g]; ++j) {
      const int k0 = j * kT;
      int* mask = s_mask + (g * 2 + (j & 1)) * kT;
      const int mk_mine = (k0 + e >= p.Sk) ? 2 : (toks[k0 + e] == 0 ? 1 : 0);
      mask[e] = mk_mine;
      // a key tile without PAD keys, inside Sk and (when causal) wholly below this query tile's diagonal needs no per-key test
      if (lane == 0) any_flag[(g * 2 + (j & 1)) * 4 + quarter] = 0;
      __syncwarp();
      if (mk_mine != 0) any_flag[(g * 2 + (j & 1)) * 4 + quarter] = 1;
      if (g == 0) asm volatile("bar.sync 1, 128;" ::: "memory"); else asm volatile("bar.sync 2, 128;" ::: "memory");
      const int* af = any_flag + (g * 2 + (j & 1)) * 4;
      const bool plain = (af[0] | af[1] | af[2] | af[3]) == 0 && !(p.causal && k0 + kT - 1 > qg);
      ptx::mbar_wait(&s_ready[g], j & 1);
      ptx::tc_fence_after();
      uint32_t sv[128];
#pragma unroll
      for (int c = 0; c < 4; ++c) ptx::tmem_ld_32x32(tS + lane_base + c * 32, sv + c * 32);
      ptx::tmem_ld_wait();
      float mx = -INFINITY;
      if (plain) {
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // four independent chains: a single one is 64 dependent steps
#pragma unroll
        for (int key = 0; key < 128; key += 8)
#pragma unroll
          for (int u = 0; u < 4; ++u) m4[u] = fmaxf(m4[u], fmaxf(__uint_as_float(sv[key + 2 * u]), __uint_as_float(sv[key + 2 * u + 1])));
        mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * p.scale_log2;   // the scale is positive: max(s) * scale = max(s * scale)
      } else {
#pragma unroll
        for (int key = 0; key < 128; ++key) {
          const int mk = mask[key];
          float v = __uint_as_float(sv[key]) * p.scale_log2;
          if (mk == 1 || (p.causal && k0 + key > qg + i)) v = kMasked;
          if (mk == 2) v = -INFINITY;
          sv[key] = __float_as_uint(v);
          mx = fmaxf(mx, v);
        }
      }
      const float m_new = fmaxf(m, mx);                 // finite: the first tile always holds a real key
      const float corr = ptx::fast_exp2(m - m_new);
      if (j > 0) {
        ptx::mbar_wait(&o_ready[g], (j - 1) & 1);       // P V of the previous key tile is in TMEM (and has read the P buffer)
        ptx::tc_fence_after();
        if (__any_sync(0xffffffffu, corr != 1.0f)) {    // some row's maximum moved: rescale this warp's O rows in place
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t ov[32];
            ptx::tmem_ld_32x32(tO + lane_base + c * 32, ov);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int t = 0; t < 32; ++t) ov[t] = __float_as_uint(__uint_as_float(ov[t]) * corr);
            ptx::tmem_st_32x32(tO + lane_base + c * 32, ov);
          }
          ptx::tmem_st_wait();
        }
      }
      float s4[4] = {0.f, 0.f, 0.f, 0.f};               // independent partial sums (one chain would be 128 dependent adds)
      const float sc = plain ? p.scale_log2 : 1.0f;     // plain tiles still hold the raw scores
      if (kPInTmem) {
        // P (bf16 pairs) goes back into TMEM over the score columns this thread has just consumed: its own lane, no shared-memory
        // traffic (a row-per-thread store into the swizzled operand tile is a 4-way bank conflict), no proxy fence
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t pk[32];
#pragma unroll
          for (int t = 0; t < 32; ++t) {
            const float a = ptx::fast_exp2(fmaf(__uint_as_float(sv[c * 64 + 2 * t]), sc, -m_new));
            const float b2 = ptx::fast_exp2(fmaf(__uint_as_float(sv[c * 64 + 2 * t + 1]), sc, -m_new));
            s4[t & 1] += a; s4[2 + (t & 1)] += b2;
            pk[t] = ptx::pack_bf16x2(a, b2);
          }
          ptx::tmem_st_32x32(tS + lane_base + c * 32, pk);
        }
        ptx::tmem_st_wait();
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float f[32];
#pragma unroll
          for (int t = 0; t < 32; ++t) { f[t] = ptx::fast_exp2(fmaf(__uint_as_float(sv[c * 32 + t]), sc, -m_new)); s4[t & 3] += f[t]; }   // exp2(-inf) = 0 past Sk
          store_row32(sPg, i, c * 32, f);
        }
        ptx::fence_proxy_async();
      }
      const float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
      ptx::tc_fence_before();
      ptx::mbar_arrive(&p_ready[g]);
      l = l * corr + sum;
      m = m_new;
    }
    if (nt[g] > 0) {
      ptx::mbar_wait(&o_ready[g], (nt[g] - 1) & 1);
      ptx::tc_fence_after();
      const float inv = 1.0f / l;
      bf16* dst = p.o + (size_t(b) * p.Sq + qg + i) * p.ldo + h * kD;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float ot[32];
        tmem_ld32(tO + lane_base + c * 32, ot);
        if (row_ok) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 u;
            u.x = ptx::pack_bf16x2(ot[q * 8] * inv, ot[q * 8 + 1] * inv); u.y = ptx::pack_bf16x2(ot[q * 8 + 2] * inv, ot[q * 8 + 3] * inv);
            u.z = ptx::pack_bf16x2(ot[q * 8 + 4] * inv, ot[q * 8 + 5] * inv); u.w = ptx::pack_bf16x2(ot[q * 8 + 6] * inv, ot[q * 8 + 7] * inv);
            *reinterpret_cast<uint4*>(dst + c * 32 + q * 8) = u;
          }
        }
      }
      if (row_ok && p.lse) reinterpret_cast<float2*>(p.lse)[(size_t(b) * p.H + h) * p.Sq + qg + i] = make_float2(m, log2f(l));
      ptx::tc_fence_before();
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}

// ---------------------------------------------------------------------------------------------------------- backward, any length
// One CTA per (128-key tile, head, batch) walks the 128-query tiles that can see those keys.  Per (q tile, kv tile) pair:
//   S = Q K^T, dP = dO V^T -> threads: P = exp2(s - m_i - log2 l_i), dS = P (dP - delta_i) -> smem ->
//   dV += P^T dO, dK += dS^T Q (accumulated in TMEM over the q tiles), dQ_pair = dS K -> fp32 staging -> TMA reduce-add
//   into an fp32 dQ accumulator (other kv tiles add their share).  delta_i = rowsum(dO * O) and the zeroing of the
//   accumulator come from attn_bwd_prep_kernel; attn_bwd_dq_out_kernel scales and rounds dQ at the end.
constexpr int kSmemBwdMulti = 2 * kTileBytes /*K V*/ + 4 * kTileBytes /*Q dO x2*/ + 2 * kPBytes /*P dS*/ + 2 * kTileBytes /*dQ fp32 staging*/ +
                              1024 /*alignment*/ + 4096 /*barriers, stats, mask*/;

struct TcMultiArgs {
  const long long* key_tokens;
  const float* lse; const float* delta;   // [B,H,Sq,2], [B,H,Sq]
  bf16 *dk_out, *dv_out; int lddk, lddv;
  int H, Sq, Sk, causal;
  float scale, scale_log2;
};

__global__ void __launch_bounds__(kThreadsMulti, 1)
attn_bwd_tc_multi_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                         const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_do,
                         const __grid_constant__ CUtensorMap tm_dq, const TcMultiArgs p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sK = smem; uint8_t* sV = sK + kTileBytes;
  uint8_t* sQ = sV + kTileBytes;            // [2] stages
  uint8_t* sdO = sQ + 2 * kTileBytes;       // [2] stages
  uint8_t* sP = sdO + 2 * kTileBytes; uint8_t* sdS = sP + kPBytes;
  uint8_t* sdQ = sdS + kPBytes;             // two [128 rows][32 fp32] boxes
  uint64_t* bars = reinterpret_cast<uint64_t*>(sdQ + 2 * kTileBytes);
  uint64_t* bar_kv = bars; uint64_t* q_full = bars + 1; uint64_t* q_empty = bars + 3;
  uint64_t* s_ready = bars + 5; uint64_t* p_ready = bars + 6; uint64_t* dq_ready = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  int* s_mask = reinterpret_cast<int*>(bars + 10);   // [128]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int jt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int k0 = jt * kT;
  const long long* toks = p.key_tokens + size_t(b) * p.Sk;
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_q); ptx::prefetch_tmap(&tm_k); ptx::prefetch_tmap(&tm_v); ptx::prefetch_tmap(&tm_do); ptx::prefetch_tmap(&tm_dq);
    ptx::mbar_init(bar_kv, 1);
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(&q_full[i], 1); ptx::mbar_init(&q_empty[i], 1); }
    ptx::mbar_init(s_ready, 1); ptx::mbar_init(p_ready, 256); ptx::mbar_init(dq_ready, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) { ptx::tmem_alloc(tmem_slot, 512); ptx::tmem_relinquish(); }
  pdl_wait();
  pdl_trigger();
  int my_mask = 0;
  if (warp >= 2 && warp < 6) {
    const int j = threadIdx.x - 64;
    my_mask = (k0 + j >= p.Sk) ? 2 : (toks[k0 + j] == 0 ? 1 : 0);
    s_mask[j] = my_mask;
  }
  ptx::tc_fence_before();
  const bool tile_masked = __syncthreads_or(my_mask != 0) != 0;   // this CTA's key tile holds a PAD key or runs past Sk
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + 128, tdQ = tmem + 256, tdK = tmem + 320, tdV = tmem + 384;

  // query tiles that can see this key tile: all of them, except under a causal mask whose rows cannot be fully masked
  const int nq = (p.Sq + kT - 1) / kT;
  const int i_begin = (p.causal && toks[0] != 0) ? jt : 0;
  const int npairs = nq - i_begin;

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t aK = ptx::smem_u32(sK), aV = ptx::smem_u32(sV), aP = ptx::smem_u32(sP), aDS = ptx::smem_u32(sdS);
      constexpr uint32_t idesc_kk = ptx::make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_mm = ptx::make_idesc_bf16(128, 64, 1, 1);
      constexpr uint32_t idesc_km = ptx::make_idesc_bf16(128, 64, 0, 1);
      auto load_q = [&](int it) {
        const int st = it & 1, q0 = (i_begin + it) * kT;
        ptx::mbar_expect_tx(&q_full[st], 2 * kTileBytes);
        ptx::tma_load_2d(sQ + st * kTileBytes, &tm_q, &q_full[st], h * kD, b * p.Sq + q0);
        ptx::tma_load_2d(sdO + st * kTileBytes, &tm_do, &q_full[st], h * kD, b * p.Sq + q0);
      };
      ptx::mbar_expect_tx(bar_kv, 2 * kTileBytes);
      ptx::tma_load_2d(sK, &tm_k, bar_kv, h * kD, b * p.Sk + k0);
      ptx::tma_load_2d(sV, &tm_v, bar_kv, h * kD, b * p.Sk + k0);
      if (npairs > 0) load_q(0);
      ptx::mbar_wait(bar_kv, 0);
      for (int it = 0; it < npairs; ++it) {
        const int st = it & 1;
        const uint32_t aQ = ptx::smem_u32(sQ + st * kTileBytes), aDO = ptx::smem_u32(sdO + st * kTileBytes);
        ptx::mbar_wait(&q_full[st], (it >> 1) & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tS, ptx::make_smem_desc(aQ + k * 32, 16, 1024), ptx::make_smem_desc(aK + k * 32, 16, 1024), idesc_kk, k > 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tdP, ptx::make_smem_desc(aDO + k * 32, 16, 1024), ptx::make_smem_desc(aV + k * 32, 16, 1024), idesc_kk, k > 0);
        ptx::umma_commit(s_ready);
        if (it + 1 < npairs) {  // prefetch the next query tile while this pair is being processed
          ptx::mbar_wait(&q_empty[(it + 1) & 1], (((it + 1) >> 1) & 1) ^ 1);
          load_q(it + 1);
        }
        ptx::mbar_wait(p_ready, it & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k)
          ptx::umma_bf16(tdV, ptx::make_smem_desc(aP + k * 2048, kTileBytes, 1024), ptx::make_smem_desc(aDO + k * 2048, kTileBytes, 1024), idesc_mm, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          ptx::umma_bf16(tdK, ptx::make_smem_desc(aDS + k * 2048, kTileBytes, 1024), ptx::make_smem_desc(aQ + k * 2048, kTileBytes, 1024), idesc_mm, (it > 0 || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          ptx::umma_bf16(tdQ, ptx::make_smem_desc(aDS + (k >> 2) * kTileBytes + (k & 3) * 32, 16, 1024),
                         ptx::make_smem_desc(aK + k * 2048, kTileBytes, 1024), idesc_km, k > 0);
        ptx::umma_commit(dq_ready);
      }
    }
  } else if (warp >= 2) {
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;              // which 64 of the tile's 128 key columns (32 of the 64 output columns) this thread handles
    const int i = quarter * 32 + lane;             // row of this thread: a query row in the pair loop, a key row at the end
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    const size_t stat_base = (size_t(b) * p.H + h) * p.Sq;
    auto load_stats = [&](int it, float& m, float& l2, float& dl) {
      const int row = (i_begin + it) * kT + i;
      if (it < npairs && row < p.Sq) {
        const float2 ml = reinterpret_cast<const float2*>(p.lse)[stat_base + row];
        m = ml.x; l2 = ml.y; dl = p.delta[stat_base + row];
      } else { m = INFINITY; l2 = 0.f; dl = 0.f; }   // exp2(x - inf) = 0: rows past the sequence contribute nothing
    };
    float m_cur, l_cur, d_cur;
    load_stats(0, m_cur, l_cur, d_cur);
    for (int it = 0; it < npairs; ++it) {
      const int q0 = (i_begin + it) * kT;
      float m_nxt, l_nxt, d_nxt;
      load_stats(it + 1, m_nxt, l_nxt, d_nxt);   // in flight during this pair
      ptx::mbar_wait(s_ready, it & 1);
      ptx::tc_fence_after();
      // a (query tile, key tile) pair with no PAD key and wholly below the causal diagonal needs no per-key test
      const bool plain = !tile_masked && !(p.causal && k0 + kT - 1 > q0);
#pragma unroll
      for (int c = half * 2; c < half * 2 + 2; ++c) {
        float sv[32], dp[32];
        tmem_ld32(tS + lane_base + c * 32, sv);
        tmem_ld32(tdP + lane_base + c * 32, dp);
        if (plain) {
          const float ml = -m_cur - l_cur;   // (finite here: a row that sees an unmasked key has a real maximum; rows past Sq carry m = +inf -> P = 0)
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float pr = ptx::fast_exp2(fmaf(sv[j], p.scale_log2, ml));
            sv[j] = pr;
            dp[j] = pr * (dp[j] - d_cur);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int key = c * 32 + j;
            const int mk = s_mask[key];
            float v = sv[j] * p.scale_log2;
            if (mk == 1 || (p.causal && k0 + key > q0 + i)) v = kMasked;
            const float pr = (mk == 2) ? 0.f : ptx::fast_exp2((v - m_cur) - l_cur);
            sv[j] = pr;
            dp[j] = pr * (dp[j] - d_cur);
          }
        }
        store_row32(sP, i, c * 32, sv);
        store_row32(sdS, i, c * 32, dp);
      }
      ptx::tc_fence_before();
      ptx::fence_proxy_async();
      ptx::mbar_arrive(p_ready);

      ptx::mbar_wait(dq_ready, it & 1);
      ptx::tc_fence_after();
      if (threadIdx.x == 64) {
        ptx::mbar_arrive(&q_empty[it & 1]);  // every MMA that read this Q / dO stage has completed
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the previous pair's reduce-add has read the staging tile
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      {   // this pair's share of dQ row i -> fp32 staging (two 128B-swizzled [128][32] boxes, one per column half)
        const int c = half;
        float f[32];
        tmem_ld32(tdQ + lane_base + c * 32, f);
#pragma unroll
        for (int q = 0; q < 8; ++q)
          *reinterpret_cast<float4*>(sdQ + c * kTileBytes + i * 128 + ((q ^ (i & 7)) << 4)) =
              make_float4(f[q * 4] * p.scale, f[q * 4 + 1] * p.scale, f[q * 4 + 2] * p.scale, f[q * 4 + 3] * p.scale);
      }
      ptx::tc_fence_before();
      ptx::fence_proxy_async();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 64) {
        ptx::tma_reduce_add_2d(&tm_dq, sdQ, h * kD, b * p.Sq + q0);
        ptx::tma_reduce_add_2d(&tm_dq, sdQ + kTileBytes, h * kD + 32, b * p.Sq + q0);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      m_cur = m_nxt; l_cur = l_nxt; d_cur = d_nxt;
    }
    if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    // dK, dV rows of this key tile (zero when no query tile was visited)
    const bool key_ok = k0 + i < p.Sk;
    bf16* dstk = p.dk_out + (size_t(b) * p.Sk + (key_ok ? k0 + i : 0)) * p.lddk + h * kD;
    bf16* dstv = p.dv_out + (size_t(b) * p.Sk + (key_ok ? k0 + i : 0)) * p.lddv + h * kD;
    {
      const int c = half;
      float fk[32], fv[32];
      if (npairs > 0) { tmem_ld32(tdK + lane_base + c * 32, fk); tmem_ld32(tdV + lane_base + c * 32, fv); }
      else {
#pragma unroll
        for (int j = 0; j < 32; ++j) { fk[j] = 0.f; fv[j] = 0.f; }
      }
      if (key_ok) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 u, w;
          u.x = ptx::pack_bf16x2(fk[q * 8] * p.scale, fk[q * 8 + 1] * p.scale); u.y = ptx::pack_bf16x2(fk[q * 8 + 2] * p.scale, fk[q * 8 + 3] * p.scale);
          u.z = ptx::pack_bf16x2(fk[q * 8 + 4] * p.scale, fk[q * 8 + 5] * p.scale); u.w = ptx::pack_bf16x2(fk[q * 8 + 6] * p.scale, fk[q * 8 + 7] * p.scale);
          w.x = ptx::pack_bf16x2(fv[q * 8], fv[q * 8 + 1]); w.y = ptx::pack_bf16x2(fv[q * 8 + 2], fv[q * 8 + 3]);
          w.z = ptx::pack_bf16x2(fv[q * 8 + 4], fv[q * 8 + 5]); w.w = ptx::pack_bf16x2(fv[q * 8 + 6], fv[q * 8 + 7]);
          *reinterpret_cast<uint4*>(dstk + c * 32 + q * 8) = u;
          *reinterpret_cast<uint4*>(dstv + c * 32 + q * 8) = w;
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}

// delta[b,h,i] = sum_c dO[i,c] * O[i,c]; also clears the fp32 dQ accumulator.  One warp per (row, head).
__global__ void __launch_bounds__(256)
attn_bwd_prep_kernel(const bf16* __restrict__ o, int ldo, const bf16* __restrict__ d_o, int lddo, float* __restrict__ delta,
                     float* __restrict__ dq_acc, int H, int Sq, int total_rows) {
  pdl_wait();
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int idx = blockIdx.x * 8 + (threadIdx.x >> 5);   // (row, head) pair
  if (idx >= total_rows * H) return;
  const int row = idx / H, h = idx % H;
  const uint32_t uo = *reinterpret_cast<const uint32_t*>(o + size_t(row) * ldo + h * kD + lane * 2);
  const uint32_t ud = *reinterpret_cast<const uint32_t*>(d_o + size_t(row) * lddo + h * kD + lane * 2);
  float acc = ptx::bf16_lo(uo) * ptx::bf16_lo(ud) + ptx::bf16_hi(uo) * ptx::bf16_hi(ud);
#pragma unroll
  for (int s2 = 16; s2 > 0; s2 >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s2);
  const int b = row / Sq, i = row % Sq;
  if (lane == 0) delta[(size_t(b) * H + h) * Sq + i] = acc;
  *reinterpret_cast<float2*>(dq_acc + (size_t(row) * H + h) * kD + lane * 2) = make_float2(0.f, 0.f);
}

__global__ void __launch_bounds__(256)
attn_bwd_dq_out_kernel(const float* __restrict__ dq_acc, bf16* __restrict__ dq, int lddq, int cols, size_t total) {
  pdl_wait();
  pdl_trigger();
  for (size_t e = (blockIdx.x * size_t(blockDim.x) + threadIdx.x) * 4; e < total; e += size_t(gridDim.x) * blockDim.x * 4) {
    const size_t row = e / cols; const int c = int(e % cols);
    const float4 f = *reinterpret_cast<const float4*>(dq_acc + e);
    uint2 u; u.x = ptx::pack_bf16x2(f.x, f.y); u.y = ptx::pack_bf16x2(f.z, f.w);
    *reinterpret_cast<uint2*>(dq + row * lddq + c) = u;
  }
}

}  // namespace

// Sequences longer than one tile (head size 64): the two-query-tile kernel above (B200XF_ATTENTION_TC_FWD_MULTI: 1 = default,
// 2 = the earlier one-tile-per-CTA tcgen05 kernel, which only ties with the mma.sync kernel because one CTA serialises
// S -> softmax -> P V; 0 = the mma.sync kernel).
static int fwd_multi_mode() {
  static int mode = -1;  // read once
  if (mode < 0) { const char* e = getenv("B200XF_ATTENTION_TC_FWD_MULTI"); mode = e ? std::max(0, std::min(2, atoi(e))) : 1; }
  return mode;
}
bool attention_tc_fwd_multi_supported(const AttnArgs& a) {
  static int enabled = -1;
  if (enabled < 0) { const char* e = getenv("B200XF_ATTENTION_TC"); enabled = (e && atoi(e) == 0) ? 0 : 1; }
  return enabled && fwd_multi_mode() != 0 && a.dk == kD && a.Sq > 0 && a.Sk > 0;
}

int attention_fwd_tc_multi(const AttnArgs& a, cudaStream_t stream) {
  CUtensorMap tq, tk, tv;
  const uint64_t cols = uint64_t(a.H) * kD;
  B200XF_TRY(tensor_map_2d(a.q, cols, uint64_t(a.B) * a.Sq, a.ldq, kT, &tq, 0));
  B200XF_TRY(tensor_map_2d(a.k, cols, uint64_t(a.B) * a.Sk, a.ldk, kT, &tk, 0));
  B200XF_TRY(tensor_map_2d(a.v, cols, uint64_t(a.B) * a.Sk, a.ldv, kT, &tv, 0));
  TcArgs p{};
  p.key_tokens = a.key_tokens; p.o = a.o; p.ldo = a.ldo; p.lse = a.lse;
  p.H = a.H; p.Sq = a.Sq; p.Sk = a.Sk; p.causal = a.causal;
  p.scale = 1.0f / sqrtf(float(a.dk)); p.scale_log2 = p.scale * kLog2e;
  if (fwd_multi_mode() == 1) {
    B200XF_TRY(ensure_dynamic_smem(attn_fwd_tc_pp_kernel, kSmemFwdPP));
    B200XF_CUDA(launch_k(attn_fwd_tc_pp_kernel, dim3((a.Sq + 2 * kT - 1) / (2 * kT), a.H, a.B), kThreadsPP, kSmemFwdPP, stream, tq, tk, tv, p));
    count_launch();
    return 0;
  }
  B200XF_TRY(ensure_dynamic_smem(attn_fwd_tc_multi_kernel, kSmemFwdMulti));
  B200XF_CUDA(launch_k(attn_fwd_tc_multi_kernel, dim3((a.Sq + kT - 1) / kT, a.H, a.B), kThreadsMulti, kSmemFwdMulti, stream, tq, tk, tv, p));
  count_launch();
  return 0;
}

bool attention_tc_multi_supported(const AttnArgs& a) {
  static int enabled = -1;
  if (enabled < 0) { const char* e = getenv("B200XF_ATTENTION_TC"); enabled = (e && atoi(e) == 0) ? 0 : 1; }
  return enabled && a.dk == kD && a.dq_acc != nullptr && a.Sq > 0 && a.Sk > 0;
}

int attention_bwd_tc_multi(const AttnArgs& a, cudaStream_t stream) {
  CUtensorMap tq, tk, tv, tdo, tdq;
  const uint64_t cols = uint64_t(a.H) * kD;
  B200XF_TRY(tensor_map_2d(a.q, cols, uint64_t(a.B) * a.Sq, a.ldq, kT, &tq, 0));
  B200XF_TRY(tensor_map_2d(a.k, cols, uint64_t(a.B) * a.Sk, a.ldk, kT, &tk, 0));
  B200XF_TRY(tensor_map_2d(a.v, cols, uint64_t(a.B) * a.Sk, a.ldv, kT, &tv, 0));
  B200XF_TRY(tensor_map_2d(a.d_o, cols, uint64_t(a.B) * a.Sq, a.lddo, kT, &tdo, 0));
  B200XF_TRY(tensor_map_2d(a.dq_acc, cols, uint64_t(a.B) * a.Sq, cols, kT, &tdq, 1));
  const int rows = a.B * a.Sq;
  B200XF_CUDA(launch_k(attn_bwd_prep_kernel, (rows * a.H + 7) / 8, 256, 0, stream, static_cast<const bf16*>(a.o), a.ldo, a.d_o, a.lddo,
                       a.delta, a.dq_acc, a.H, a.Sq, rows));
  TcMultiArgs p{};
  p.key_tokens = a.key_tokens; p.lse = a.lse; p.delta = a.delta;
  p.dk_out = a.dk_out; p.dv_out = a.dv_out; p.lddk = a.lddk; p.lddv = a.lddv;
  p.H = a.H; p.Sq = a.Sq; p.Sk = a.Sk; p.causal = a.causal;
  p.scale = 1.0f / sqrtf(float(a.dk)); p.scale_log2 = p.scale * kLog2e;
  B200XF_TRY(ensure_dynamic_smem(attn_bwd_tc_multi_kernel, kSmemBwdMulti));
  B200XF_CUDA(launch_k(attn_bwd_tc_multi_kernel, dim3((a.Sk + kT - 1) / kT, a.H, a.B), kThreadsMulti, kSmemBwdMulti, stream, tq, tk, tv, tdo, tdq, p));
  const size_t total = size_t(rows) * cols;
  B200XF_CUDA(launch_k(attn_bwd_dq_out_kernel, int(std::min<size_t>((total / 4 + 255) / 256, 4096)), 256, 0, stream,
                       static_cast<const float*>(a.dq_acc), a.dq, a.lddq, int(cols), total));
  count_launch(3);
  return 0;
}

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
  p.early = a.early;
  B200XF_TRY(ensure_dynamic_smem(attn_fwd_tc_kernel, kSmemFwd));
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
  p.early = a.early;
  B200XF_TRY(ensure_dynamic_smem(attn_bwd_tc_kernel, kSmemBwd));
  B200XF_CUDA(launch_k(attn_bwd_tc_kernel, dim3(a.H, a.B), kThreads2, kSmemBwd, stream, tq, tk, tv, tdo, p));
  count_launch();
  return 0;
}

}  // namespace b200xf
