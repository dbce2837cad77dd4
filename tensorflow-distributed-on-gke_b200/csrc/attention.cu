// Fused multi-head attention forward / backward (flash style: no [B,H,Sq,Sk] tensor, no mask tensor,
// no head transposes).  Restates transformer_model.py:73-109 (scaled_dot_product_attention) together
// with the masks of :56-70,:350-363 and the split/merge-heads reshapes of :124-131,:160-163:
//   scores = Q K^T / sqrt(dk) + mask * (-1e9);  weights = softmax(scores);  out = weights V
//   mask(i,j) = [key_token[j] == 0]  OR  (causal AND j > i)
// Q/K/V/O are read/written as column slices [.., h*dk:(h+1)*dk] of row-major [tokens, ld] buffers, so the
// fused QKV GEMM output is consumed in place.  The additive -1e9 is emulated exactly in fp32: a masked
// score *is* -1e9 (fp32 absorbs |x| < 32), so fully-masked rows come out uniform, not NaN.
//
// Tensor-core path: mma.sync.m16n8k16 bf16 (fp32 accumulate) with ldmatrix operands; 64 query rows per
// CTA (16 per warp) against 64-key tiles.
#include <mutex>
#include <set>

#include "internal.h"
#include "ptx.cuh"

namespace b200xf {

namespace {

constexpr int kTile = 64;      // rows per Q tile / keys per KV tile
constexpr int kWarps = 4;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kMasked = -1.0e9f * 1.4426950408889634f;  // the reference's -1e9, in the log2 domain

__device__ __forceinline__ void mma_bf16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t* r, const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(ptx::smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t* r, const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(ptx::smem_u32(p)));
}

template <int DK>
struct Smem {
  static constexpr int kPitch = DK + 8;  // +16 B per row: conflict-free ldmatrix
  typedef bf16 Tile[kTile][kPitch];
};

// Cooperative load of a [kTile x DK] slice (rows row0.. of a [*, ld] buffer at column col0) into padded smem.
template <int DK>
__device__ __forceinline__ void load_tile(typename Smem<DK>::Tile& dst, const bf16* __restrict__ src, int ld, int row0,
                                          int rows_total, int col0) {
  constexpr int kChunks = DK / 8;
  for (int i = threadIdx.x; i < kTile * kChunks; i += kWarps * 32) {
    const int r = i / kChunks, c = (i % kChunks) * 8;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (row0 + r < rows_total) v = *reinterpret_cast<const uint4*>(src + size_t(row0 + r) * ld + col0 + c);
    *reinterpret_cast<uint4*>(&dst[r][c]) = v;
  }
}

// Same, with cp.async (16-byte chunks, zero fill past the end): the copy overlaps whatever the CTA does next.
template <int DK>
__device__ __forceinline__ void load_tile_async(typename Smem<DK>::Tile& dst, const bf16* __restrict__ src, int ld, int row0,
                                                int rows_total, int col0) {
  constexpr int kChunks = DK / 8;
  for (int i = threadIdx.x; i < kTile * kChunks; i += kWarps * 32) {
    const int r = i / kChunks, c = (i % kChunks) * 8;
    const bool ok = row0 + r < rows_total;
    const bf16* g = src + size_t(ok ? row0 + r : 0) * ld + col0 + c;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(ptx::smem_u32(&dst[r][c])), "l"(g), "r"(ok ? 16 : 0) : "memory");
  }
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// A fragments (16 rows x DK) of this warp's rows, one x4 per 16-wide k step.
template <int DK>
__device__ __forceinline__ void load_a_frags(uint32_t (*a)[4], typename Smem<DK>::Tile& t, int row0, int lane) {
#pragma unroll
  for (int kk = 0; kk < DK / 16; ++kk) ldsm_x4(a[kk], &t[row0 + (lane & 15)][kk * 16 + (lane >> 4) * 8]);
}

// acc[16 x 64] (+)= A[16 x DK] . T[64 x DK]^T   (T rows are the n index; contraction over DK)
template <int DK>
__device__ __forceinline__ void mma_a_tT(float (*acc)[4], const uint32_t (*a)[4], typename Smem<DK>::Tile& t, int lane) {
#pragma unroll
  for (int kk = 0; kk < DK / 16; ++kk) {
#pragma unroll
    for (int nb2 = 0; nb2 < kTile / 16; ++nb2) {
      uint32_t b[4];
      ldsm_x4(b, &t[nb2 * 16 + (lane & 7) + ((lane >> 4) << 3)][kk * 16 + ((lane >> 3) & 1) * 8]);
      mma_bf16(acc[nb2 * 2], a[kk], b[0], b[1]);
      mma_bf16(acc[nb2 * 2 + 1], a[kk], b[2], b[3]);
    }
  }
}

// out[16 x DK] += P[16 x 64] . T[64 x DK]   (P given as C-layout accumulators, contraction over the 64 tile rows)
template <int DK>
__device__ __forceinline__ void mma_p_t(float (*out)[4], const float (*p)[4], typename Smem<DK>::Tile& t, int lane) {
#pragma unroll
  for (int kk = 0; kk < kTile / 16; ++kk) {
    uint32_t a[4];
    a[0] = ptx::pack_bf16x2(p[2 * kk][0], p[2 * kk][1]);
    a[1] = ptx::pack_bf16x2(p[2 * kk][2], p[2 * kk][3]);
    a[2] = ptx::pack_bf16x2(p[2 * kk + 1][0], p[2 * kk + 1][1]);
    a[3] = ptx::pack_bf16x2(p[2 * kk + 1][2], p[2 * kk + 1][3]);
#pragma unroll
    for (int nb2 = 0; nb2 < DK / 16; ++nb2) {
      uint32_t b[4];
      ldsm_x4_t(b, &t[kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8][nb2 * 16 + (lane >> 4) * 8]);
      mma_bf16(out[nb2 * 2], a, b[0], b[1]);
      mma_bf16(out[nb2 * 2 + 1], a, b[2], b[3]);
    }
  }
}

__device__ __forceinline__ float quad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

struct KernelArgs {
  const bf16 *q, *k, *v; bf16* o; float* lse;
  const long long* key_tokens;
  int ldq, ldk, ldv, ldo;
  int H, Sq, Sk, causal;
  float scale_log2;  // log2(e) / sqrt(dk)
  float scale;       // 1 / sqrt(dk)
  const bf16* d_o; int lddo;
  bf16 *dq, *dk_out, *dv_out; int lddq, lddk, lddv;
  float* delta;
};

// ------------------------------------------------------------------------------------------------ forward
template <int DK>
struct FwdSmem { typename Smem<DK>::Tile q, k[2], v[2]; int mask[2][kTile]; };

template <int DK>
__global__ void __launch_bounds__(kWarps * 32)
attn_fwd_kernel(const KernelArgs p) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ __align__(16) uint8_t smem_raw[];
  FwdSmem<DK>& sm = *reinterpret_cast<FwdSmem<DK>*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kTile;
  const bf16* Q = p.q + size_t(b) * p.Sq * p.ldq;
  const bf16* K = p.k + size_t(b) * p.Sk * p.ldk;
  const bf16* V = p.v + size_t(b) * p.Sk * p.ldv;
  const long long* toks = p.key_tokens + size_t(b) * p.Sk;

  // Future key tiles are exactly zero-weight under the causal mask unless a row is fully masked, which
  // needs token[0] == PAD; only then must every tile be visited to reproduce the uniform row.
  int kv_end = p.Sk;
  if (p.causal && toks[0] != 0) kv_end = min(p.Sk, q0 + kTile);
  const int ntiles = (kv_end + kTile - 1) / kTile;

  auto issue_kv = [&](int it) {
    const int buf = it & 1, k0 = it * kTile;
    load_tile_async<DK>(sm.k[buf], K, p.ldk, k0, p.Sk, h * DK);
    load_tile_async<DK>(sm.v[buf], V, p.ldv, k0, p.Sk, h * DK);
    if (threadIdx.x < kTile) {
      const int j = k0 + threadIdx.x;
      sm.mask[buf][threadIdx.x] = (j >= p.Sk) ? 2 : (toks[j] == 0 ? 1 : 0);
    }
  };
  load_tile_async<DK>(sm.q, Q, p.ldq, q0, p.Sq, h * DK);
  issue_kv(0);
  cp_async_commit();

  uint32_t qa[DK / 16][4];
  float o[DK / 8][4];
#pragma unroll
  for (int i = 0; i < DK / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};
  const int r0 = q0 + warp * 16 + g;  // this thread's rows: r0 and r0 + 8

  for (int it = 0; it < ntiles; ++it) {
    const int buf = it & 1, k0 = it * kTile;
    if (it + 1 < ntiles) { issue_kv(it + 1); cp_async_commit(); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    if (it == 0) load_a_frags<DK>(qa, sm.q, warp * 16, lane);

    float s[kTile / 8][4];
#pragma unroll
    for (int i = 0; i < kTile / 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
    mma_a_tT<DK>(s, qa, sm.k[buf], lane);

    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nb = 0; nb < kTile / 8; ++nb) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int jl = nb * 8 + t4 * 2 + (e & 1);
        const int row = r0 + (e >> 1) * 8;
        const int mk = sm.mask[buf][jl];
        float v = s[nb][e] * p.scale_log2;
        if (mk == 1 || (p.causal && k0 + jl > row)) v = kMasked;
        if (mk == 2) v = -INFINITY;
        s[nb][e] = v;
        mx[e >> 1] = fmaxf(mx[e >> 1], v);
      }
    }
    float corr[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const float mn = fmaxf(m[r], quad_max(mx[r]));
      corr[r] = exp2f(m[r] - mn);
      m[r] = mn;
    }
    float rs[2] = {0.f, 0.f};
#pragma unroll
    for (int nb = 0; nb < kTile / 8; ++nb) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float pv = exp2f(s[nb][e] - m[e >> 1]);
        s[nb][e] = pv;
        rs[e >> 1] += pv;
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) l[r] = l[r] * corr[r] + quad_sum(rs[r]);
#pragma unroll
    for (int i = 0; i < DK / 8; ++i) { o[i][0] *= corr[0]; o[i][1] *= corr[0]; o[i][2] *= corr[1]; o[i][3] *= corr[1]; }
    mma_p_t<DK>(o, s, sm.v[buf], lane);
    __syncthreads();  // everyone is done with this buffer before the prefetch two tiles ahead overwrites it
  }

  const float inv[2] = {1.f / l[0], 1.f / l[1]};
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int row = r0 + r * 8;
    if (row < p.Sq) {
      bf16* dst = p.o + (size_t(b) * p.Sq + row) * p.ldo + h * DK;
#pragma unroll
      for (int nb = 0; nb < DK / 8; ++nb)
        *reinterpret_cast<uint32_t*>(dst + nb * 8 + t4 * 2) = ptx::pack_bf16x2(o[nb][r * 2] * inv[r], o[nb][r * 2 + 1] * inv[r]);
      if (t4 == 0 && p.lse) {  // kept as two terms: for a fully masked row m = -1.44e9 and would absorb log2(l)
        float2* dst2 = reinterpret_cast<float2*>(p.lse) + (size_t(b) * p.H + h) * p.Sq + row;
        *dst2 = make_float2(m[r], log2f(l[r]));
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward: dQ (+ delta)
// One CTA per 64-row Q tile; recomputes P per KV tile:  dS = P * (dO V^T - delta);  dQ = scale * dS K.
template <int DK>
struct DqSmem { typename Smem<DK>::Tile q, d_o, k[2], v[2]; int mask[2][kTile]; };

template <int DK>
__global__ void __launch_bounds__(kWarps * 32)
attn_bwd_dq_kernel(const KernelArgs p) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ __align__(16) uint8_t smem_raw[];
  DqSmem<DK>& sm = *reinterpret_cast<DqSmem<DK>*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kTile;
  const bf16* K = p.k + size_t(b) * p.Sk * p.ldk;
  const bf16* V = p.v + size_t(b) * p.Sk * p.ldv;
  const long long* toks = p.key_tokens + size_t(b) * p.Sk;
  const int r0 = q0 + warp * 16 + g;

  int kv_end = p.Sk;
  if (p.causal && toks[0] != 0) kv_end = min(p.Sk, q0 + kTile);
  const int ntiles = (kv_end + kTile - 1) / kTile;
  auto issue_kv = [&](int it) {
    const int buf = it & 1, k0 = it * kTile;
    load_tile_async<DK>(sm.k[buf], K, p.ldk, k0, p.Sk, h * DK);
    load_tile_async<DK>(sm.v[buf], V, p.ldv, k0, p.Sk, h * DK);
    if (threadIdx.x < kTile) {
      const int j = k0 + threadIdx.x;
      sm.mask[buf][threadIdx.x] = (j >= p.Sk) ? 2 : (toks[j] == 0 ? 1 : 0);
    }
  };
  load_tile_async<DK>(sm.q, p.q + size_t(b) * p.Sq * p.ldq, p.ldq, q0, p.Sq, h * DK);
  load_tile_async<DK>(sm.d_o, p.d_o + size_t(b) * p.Sq * p.lddo, p.lddo, q0, p.Sq, h * DK);
  issue_kv(0);
  cp_async_commit();

  // delta = rowsum(dO * O) for this thread's two rows (each row's DK columns are spread over the quad);
  // these global loads are in flight together with the tile copies above.
  float delta[2] = {0.f, 0.f}, lse_m[2] = {0.f, 0.f}, lse_l[2] = {0.f, 0.f};
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int row = r0 + r * 8;
    if (row < p.Sq) {
      const bf16* po = p.o + (size_t(b) * p.Sq + row) * p.ldo + h * DK;
      const bf16* pd = p.d_o + (size_t(b) * p.Sq + row) * p.lddo + h * DK;
      float acc = 0.f;
#pragma unroll
      for (int c = t4 * 4; c < DK; c += 16) {
        const uint2 uo = *reinterpret_cast<const uint2*>(po + c), ud = *reinterpret_cast<const uint2*>(pd + c);
        acc += ptx::bf16_lo(uo.x) * ptx::bf16_lo(ud.x) + ptx::bf16_hi(uo.x) * ptx::bf16_hi(ud.x) +
               ptx::bf16_lo(uo.y) * ptx::bf16_lo(ud.y) + ptx::bf16_hi(uo.y) * ptx::bf16_hi(ud.y);
      }
      delta[r] = acc;
      const float2 ml = reinterpret_cast<const float2*>(p.lse)[(size_t(b) * p.H + h) * p.Sq + row];
      lse_m[r] = ml.x; lse_l[r] = ml.y;
    }
    delta[r] = quad_sum(delta[r]);
    if (t4 == 0 && row < p.Sq) p.delta[(size_t(b) * p.H + h) * p.Sq + row] = delta[r];
  }

  uint32_t qa[DK / 16][4], doa[DK / 16][4];
  float dq[DK / 8][4];
#pragma unroll
  for (int i = 0; i < DK / 8; ++i) dq[i][0] = dq[i][1] = dq[i][2] = dq[i][3] = 0.f;

  for (int it = 0; it < ntiles; ++it) {
    const int buf = it & 1, k0 = it * kTile;
    if (it + 1 < ntiles) { issue_kv(it + 1); cp_async_commit(); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    if (it == 0) { load_a_frags<DK>(qa, sm.q, warp * 16, lane); load_a_frags<DK>(doa, sm.d_o, warp * 16, lane); }
    float s[kTile / 8][4], dp[kTile / 8][4];
#pragma unroll
    for (int i = 0; i < kTile / 8; ++i) { s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f; dp[i][0] = dp[i][1] = dp[i][2] = dp[i][3] = 0.f; }
    mma_a_tT<DK>(s, qa, sm.k[buf], lane);
    mma_a_tT<DK>(dp, doa, sm.v[buf], lane);
#pragma unroll
    for (int nb = 0; nb < kTile / 8; ++nb) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int jl = nb * 8 + t4 * 2 + (e & 1);
        const int row = r0 + (e >> 1) * 8;
        const int mk = sm.mask[buf][jl];
        float v = s[nb][e] * p.scale_log2;
        if (mk == 1 || (p.causal && k0 + jl > row)) v = kMasked;
        const float pr = (mk == 2) ? 0.f : exp2f((v - lse_m[e >> 1]) - lse_l[e >> 1]);
        s[nb][e] = pr * (dp[nb][e] - delta[e >> 1]);  // dS
      }
    }
    mma_p_t<DK>(dq, s, sm.k[buf], lane);
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int row = r0 + r * 8;
    if (row < p.Sq) {
      bf16* dst = p.dq + (size_t(b) * p.Sq + row) * p.lddq + h * DK;
#pragma unroll
      for (int nb = 0; nb < DK / 8; ++nb)
        *reinterpret_cast<uint32_t*>(dst + nb * 8 + t4 * 2) = ptx::pack_bf16x2(dq[nb][r * 2] * p.scale, dq[nb][r * 2 + 1] * p.scale);
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward: dK, dV
// One CTA per 64-key tile (16 keys per warp), looping over Q tiles, in the transposed formulation:
//   S^T = K Q^T;  P^T = exp2(S^T - lse[q]);  dV += P^T dO;  dP^T = V dO^T;  dS^T = P^T (dP^T - delta[q]);  dK += scale dS^T Q
template <int DK>
struct DkvSmem { typename Smem<DK>::Tile k, v, q[2], d_o[2]; float lse_m[2][kTile], lse_l[2][kTile], delta[2][kTile]; };

template <int DK>
__global__ void __launch_bounds__(kWarps * 32)
attn_bwd_dkv_kernel(const KernelArgs p) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ __align__(16) uint8_t smem_raw[];
  DkvSmem<DK>& sm = *reinterpret_cast<DkvSmem<DK>*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int b = blockIdx.z, h = blockIdx.y, k0 = blockIdx.x * kTile;
  const bf16* Q = p.q + size_t(b) * p.Sq * p.ldq;
  const bf16* DO = p.d_o + size_t(b) * p.Sq * p.lddo;
  const long long* toks = p.key_tokens + size_t(b) * p.Sk;
  const int key0 = k0 + warp * 16 + g;  // this thread's keys: key0, key0 + 8
  int kmask[2];
#pragma unroll
  for (int r = 0; r < 2; ++r) { const int j = key0 + r * 8; kmask[r] = (j >= p.Sk) ? 2 : (toks[j] == 0 ? 1 : 0); }

  int q_begin = 0;
  if (p.causal && toks[0] != 0) q_begin = k0;  // earlier query tiles never see these keys
  const int ntiles = (p.Sq - q_begin + kTile - 1) / kTile;
  auto issue_q = [&](int it) {
    const int buf = it & 1, q0 = q_begin + it * kTile;
    load_tile_async<DK>(sm.q[buf], Q, p.ldq, q0, p.Sq, h * DK);
    load_tile_async<DK>(sm.d_o[buf], DO, p.lddo, q0, p.Sq, h * DK);
    if (threadIdx.x < kTile) {
      const int i = q0 + threadIdx.x;
      const float2 ml = (i < p.Sq) ? reinterpret_cast<const float2*>(p.lse)[(size_t(b) * p.H + h) * p.Sq + i] : make_float2(INFINITY, 0.f);
      sm.lse_m[buf][threadIdx.x] = ml.x; sm.lse_l[buf][threadIdx.x] = ml.y;  // exp2(x - inf) = 0 for rows past Sq
      sm.delta[buf][threadIdx.x] = (i < p.Sq) ? p.delta[(size_t(b) * p.H + h) * p.Sq + i] : 0.f;
    }
  };
  load_tile_async<DK>(sm.k, p.k + size_t(b) * p.Sk * p.ldk, p.ldk, k0, p.Sk, h * DK);
  load_tile_async<DK>(sm.v, p.v + size_t(b) * p.Sk * p.ldv, p.ldv, k0, p.Sk, h * DK);
  if (ntiles > 0) issue_q(0);
  cp_async_commit();

  uint32_t ka[DK / 16][4], va[DK / 16][4];
  float dk[DK / 8][4], dv[DK / 8][4];
#pragma unroll
  for (int i = 0; i < DK / 8; ++i) { dk[i][0] = dk[i][1] = dk[i][2] = dk[i][3] = 0.f; dv[i][0] = dv[i][1] = dv[i][2] = dv[i][3] = 0.f; }

  for (int it = 0; it < ntiles; ++it) {
    const int buf = it & 1, q0 = q_begin + it * kTile;
    if (it + 1 < ntiles) { issue_q(it + 1); cp_async_commit(); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    if (it == 0) { load_a_frags<DK>(ka, sm.k, warp * 16, lane); load_a_frags<DK>(va, sm.v, warp * 16, lane); }
    float st[kTile / 8][4], dpt[kTile / 8][4];
#pragma unroll
    for (int i = 0; i < kTile / 8; ++i) { st[i][0] = st[i][1] = st[i][2] = st[i][3] = 0.f; dpt[i][0] = dpt[i][1] = dpt[i][2] = dpt[i][3] = 0.f; }
    mma_a_tT<DK>(st, ka, sm.q[buf], lane);
    mma_a_tT<DK>(dpt, va, sm.d_o[buf], lane);
    float pt[kTile / 8][4];
#pragma unroll
    for (int nb = 0; nb < kTile / 8; ++nb) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int il = nb * 8 + t4 * 2 + (e & 1);     // query within tile (column)
        const int key = key0 + (e >> 1) * 8;          // key (row)
        const int mk = kmask[e >> 1];
        float v = st[nb][e] * p.scale_log2;
        if (mk == 1 || (p.causal && key > q0 + il)) v = kMasked;
        const float pr = (mk == 2) ? 0.f : exp2f((v - sm.lse_m[buf][il]) - sm.lse_l[buf][il]);
        pt[nb][e] = pr;
        st[nb][e] = pr * (dpt[nb][e] - sm.delta[buf][il]);   // dS^T
      }
    }
    mma_p_t<DK>(dv, pt, sm.d_o[buf], lane);
    mma_p_t<DK>(dk, st, sm.q[buf], lane);
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int key = key0 + r * 8;
    if (key < p.Sk) {
      bf16* dstk = p.dk_out + (size_t(b) * p.Sk + key) * p.lddk + h * DK;
      bf16* dstv = p.dv_out + (size_t(b) * p.Sk + key) * p.lddv + h * DK;
#pragma unroll
      for (int nb = 0; nb < DK / 8; ++nb) {
        *reinterpret_cast<uint32_t*>(dstk + nb * 8 + t4 * 2) = ptx::pack_bf16x2(dk[nb][r * 2] * p.scale, dk[nb][r * 2 + 1] * p.scale);
        *reinterpret_cast<uint32_t*>(dstv + nb * 8 + t4 * 2) = ptx::pack_bf16x2(dv[nb][r * 2], dv[nb][r * 2 + 1]);
      }
    }
  }
}

int check(const AttnArgs& a, bool bwd) {
  if (a.dk != 16 && a.dk != 32 && a.dk != 64) return set_error("attention: head size %d unsupported (16, 32 or 64)", a.dk);
  if (a.B <= 0 || a.H <= 0 || a.Sq <= 0 || a.Sk <= 0) return set_error("attention: empty problem");
  if ((a.ldq | a.ldk | a.ldv | a.ldo) % 8) return set_error("attention: leading dimensions must be multiples of 8");
  if (!a.q || !a.k || !a.v || !a.o || !a.key_tokens || !a.lse) return set_error("attention: null pointer");
  if (bwd && (!a.d_o || !a.dq || !a.dk_out || !a.dv_out || !a.delta)) return set_error("attention_bwd: null pointer");
  if (bwd && ((a.lddo | a.lddq | a.lddk | a.lddv) % 8)) return set_error("attention_bwd: leading dimensions must be multiples of 8");
  return 0;
}

KernelArgs pack(const AttnArgs& a) {
  KernelArgs p;
  p.q = a.q; p.k = a.k; p.v = a.v; p.o = a.o; p.lse = a.lse; p.key_tokens = a.key_tokens;
  p.ldq = a.ldq; p.ldk = a.ldk; p.ldv = a.ldv; p.ldo = a.ldo;
  p.H = a.H; p.Sq = a.Sq; p.Sk = a.Sk; p.causal = a.causal;
  p.scale = 1.0f / sqrtf(float(a.dk));
  p.scale_log2 = p.scale * kLog2e;
  p.d_o = a.d_o; p.lddo = a.lddo; p.dq = a.dq; p.dk_out = a.dk_out; p.dv_out = a.dv_out;
  p.lddq = a.lddq; p.lddk = a.lddk; p.lddv = a.lddv; p.delta = a.delta;
  return p;
}

template <typename Kern>
int launch_attn(Kern kern, size_t smem, dim3 grid, const KernelArgs& p, cudaStream_t stream) {
  if (smem > 48 * 1024) B200XF_TRY(ensure_dynamic_smem(kern, int(smem)));
  B200XF_CUDA(launch_k(kern, grid, kWarps * 32, smem, stream, p));
  count_launch();
  return 0;
}

}  // namespace

int attention_fwd(const AttnArgs& a, cudaStream_t stream) {
  B200XF_TRY(check(a, false));
  if (!a.generic && attention_tc_supported(a)) return attention_fwd_tc(a, stream);
  if (!a.generic && attention_tc_fwd_multi_supported(a)) return attention_fwd_tc_multi(a, stream);
  const KernelArgs p = pack(a);
  const dim3 grid((a.Sq + kTile - 1) / kTile, a.H, a.B);
  switch (a.dk) {
    case 16: return launch_attn(attn_fwd_kernel<16>, sizeof(FwdSmem<16>), grid, p, stream);
    case 32: return launch_attn(attn_fwd_kernel<32>, sizeof(FwdSmem<32>), grid, p, stream);
    default: return launch_attn(attn_fwd_kernel<64>, sizeof(FwdSmem<64>), grid, p, stream);
  }
}

int attention_bwd(const AttnArgs& a, cudaStream_t stream) {
  B200XF_TRY(check(a, true));
  if (attention_tc_supported(a)) return attention_bwd_tc(a, stream);
  if (attention_tc_multi_supported(a)) return attention_bwd_tc_multi(a, stream);
  const KernelArgs p = pack(a);
  const dim3 gq((a.Sq + kTile - 1) / kTile, a.H, a.B), gk((a.Sk + kTile - 1) / kTile, a.H, a.B);
  switch (a.dk) {
    case 16:
      B200XF_TRY(launch_attn(attn_bwd_dq_kernel<16>, sizeof(DqSmem<16>), gq, p, stream));
      return launch_attn(attn_bwd_dkv_kernel<16>, sizeof(DkvSmem<16>), gk, p, stream);
    case 32:
      B200XF_TRY(launch_attn(attn_bwd_dq_kernel<32>, sizeof(DqSmem<32>), gq, p, stream));
      return launch_attn(attn_bwd_dkv_kernel<32>, sizeof(DkvSmem<32>), gk, p, stream);
    default:
      B200XF_TRY(launch_attn(attn_bwd_dq_kernel<64>, sizeof(DqSmem<64>), gq, p, stream));
      return launch_attn(attn_bwd_dkv_kernel<64>, sizeof(DkvSmem<64>), gk, p, stream);
  }
}

}  // namespace b200xf

// ------------------------------------------------------------------------------------------------ attention weight maps
// transformer_model.py:106,312-313: the softmax weights themselves, [B,H,Sq,Sk] fp32, for the inference API
// (tester.py:51-52 plots them).  Recomputed from Q, K and the row statistics the forward pass saved; not on the training path.
namespace b200xf {
namespace {
__global__ void __launch_bounds__(128)
attention_probs_kernel(const bf16* __restrict__ q, int ldq, const bf16* __restrict__ k, int ldk, const float* __restrict__ lse,
                       const long long* __restrict__ key_tokens, int causal, int H, int Sq, int Sk, int dk, float scale_log2,
                       float* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sq[128];
  const int i = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const bf16* qrow = q + (size_t(b) * Sq + i) * ldq + h * dk;
  for (int c = threadIdx.x; c < dk; c += blockDim.x) sq[c] = __bfloat162float(qrow[c]);
  __syncthreads();
  const float2 ml = reinterpret_cast<const float2*>(lse)[(size_t(b) * H + h) * Sq + i];
  float* orow = out + ((size_t(b) * H + h) * Sq + i) * Sk;
  for (int j = threadIdx.x; j < Sk; j += blockDim.x) {
    const bf16* krow = k + (size_t(b) * Sk + j) * ldk + h * dk;
    float s = 0.f;
    for (int c = 0; c < dk; c += 2) {
      const uint32_t u = *reinterpret_cast<const uint32_t*>(krow + c);
      s += sq[c] * ptx::bf16_lo(u) + sq[c + 1] * ptx::bf16_hi(u);
    }
    float v = s * scale_log2;
    if (key_tokens[size_t(b) * Sk + j] == 0 || (causal && j > i)) v = kMasked;
    orow[j] = exp2f((v - ml.x) - ml.y);
  }
}
}  // namespace

int attention_probs(const AttnArgs& a, float* out, cudaStream_t stream) {
  if (!a.q || !a.k || !a.lse || !a.key_tokens || !out) return set_error("attention_probs: null pointer");
  if (a.dk > 128 || a.dk % 2) return set_error("attention_probs: head size %d unsupported", a.dk);
  const float scale_log2 = kLog2e / sqrtf(float(a.dk));
  B200XF_CUDA(launch_k(attention_probs_kernel, dim3(a.Sq, a.H, a.B), 128, 0, stream, a.q, a.ldq, a.k, a.ldk,
                       static_cast<const float*>(a.lse), a.key_tokens, a.causal, a.H, a.Sq, a.Sk, a.dk, scale_log2, out));
  count_launch();
  return 0;
}
}  // namespace b200xf
