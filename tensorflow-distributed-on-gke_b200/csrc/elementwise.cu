// HBM-bound kernels of the train step: LayerNorm (+residual) fwd/bwd, embedding*sqrt(d)+PE fwd/bwd,
// target shift + PAD count, masked softmax cross-entropy over the vocab (fwd+bwd fused), Adam with
// the warm-up schedule over flat buckets, column sums (bias gradients), casts.
// All global accesses are 128-bit vectorised and coalesced; reductions use warp shuffles.
#include <algorithm>
#include <mutex>

#include <cub/device/device_radix_sort.cuh>

#include "internal.h"
#include "ptx.cuh"

namespace b200xf {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  f[0] = ptx::bf16_lo(u.x); f[1] = ptx::bf16_hi(u.x); f[2] = ptx::bf16_lo(u.y); f[3] = ptx::bf16_hi(u.y);
  f[4] = ptx::bf16_lo(u.z); f[5] = ptx::bf16_hi(u.z); f[6] = ptx::bf16_lo(u.w); f[7] = ptx::bf16_hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 u;
  u.x = ptx::pack_bf16x2(f[0], f[1]); u.y = ptx::pack_bf16x2(f[2], f[3]);
  u.z = ptx::pack_bf16x2(f[4], f[5]); u.w = ptx::pack_bf16x2(f[6], f[7]);
  return u;
}
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16(x)); }

// ============================================================================ LayerNorm
// One warp per row; the row lives in registers (CH chunks of 8 bf16 per lane): one HBM read, one write.
constexpr int kLnWarps = 8;

template <int CH>
__global__ void __launch_bounds__(kLnWarps * 32)
ln_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ resid, const float* __restrict__ gamma,
              const float* __restrict__ beta, bf16* __restrict__ y, bf16* __restrict__ z_out, float* __restrict__ mean_out,
              float* __restrict__ rstd_out, int rows, int d, float eps) {
  pdl_wait();
  pdl_trigger();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * kLnWarps + warp;
  if (row >= rows) return;
  const size_t base = size_t(row) * d;
  float v[CH][8];
  float sum = 0.f;
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int col = (c * 32 + lane) * 8;
    if (col < d) {
      unpack8(*reinterpret_cast<const uint4*>(x + base + col), v[c]);
      if (resid) {
        float r[8];
        unpack8(*reinterpret_cast<const uint4*>(resid + base + col), r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[c][j] = round_bf16(v[c][j] + r[j]);  // z is kept in bf16 for the backward
        if (z_out) *reinterpret_cast<uint4*>(z_out + base + col) = pack8(v[c]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += v[c][j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[c][j] = 0.f;
    }
  }
  const float mean = warp_sum(sum) / float(d);
  float sq = 0.f;
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int col = (c * 32 + lane) * 8;
    if (col < d) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float t = v[c][j] - mean; sq += t * t; }
    }
  }
  const float var = warp_sum(sq) / float(d);  // biased variance, as tf.nn.moments
  const float rstd = 1.0f / sqrtf(var + eps);
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int col = (c * 32 + lane) * 8;
    if (col < d) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + col)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + col + 4));
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + col)), b1 = __ldg(reinterpret_cast<const float4*>(beta + col + 4));
      const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[c][j] - mean) * rstd * g[j] + b[j];
      *reinterpret_cast<uint4*>(y + base + col) = pack8(o);
    }
  }
  if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
}

// Backward: dz = rstd * (g - mean(g) - xhat * mean(g*xhat)),  g = dy * gamma.
// Latency-oriented: each warp owns RPW rows whose loads are all issued up front (no serial row loop); the block's
// 8 warps then reduce their column sums (dy*xhat, dy, dz) through shared memory, all threads in parallel, and write
// ONE partial [3][d] per block (8*RPW rows).  A batched reduce_partials launch sums the partials later in a fixed
// order (deterministic, no float atomics).
template <int CH, int RPW>
__global__ void __launch_bounds__(kLnWarps * 32, (CH * RPW <= 4) ? 2 : 1)
ln_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ z, const float* __restrict__ mean_in,
              const float* __restrict__ rstd_in, const float* __restrict__ gamma, bf16* __restrict__ dz, int rows, int d,
              float* __restrict__ partials, bf16* __restrict__ dzm, const DropArgs drop) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ float red[];  // [kLnWarps][3][d]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = (blockIdx.x * kLnWarps + warp) * RPW;
  uint4 u_dy[RPW][CH], u_z[RPW][CH];
  float mean[RPW], rstd[RPW];
#pragma unroll
  for (int r = 0; r < RPW; ++r) {
    const int row = row0 + r;
    mean[r] = 0.f; rstd[r] = 0.f;
    if (row < rows) {
      const size_t base = size_t(row) * d;
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int col = (c * 32 + lane) * 8;
        if (col < d) { u_dy[r][c] = *reinterpret_cast<const uint4*>(dy + base + col); u_z[r][c] = *reinterpret_cast<const uint4*>(z + base + col); }
      }
      mean[r] = mean_in[row]; rstd[r] = rstd_in[row];
    }
  }
  float gam[CH][8];
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int col = (c * 32 + lane) * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) gam[c][j] = 0.f;
    if (col < d) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + col)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + col + 4));
      gam[c][0] = g0.x; gam[c][1] = g0.y; gam[c][2] = g0.z; gam[c][3] = g0.w; gam[c][4] = g1.x; gam[c][5] = g1.y; gam[c][6] = g1.z; gam[c][7] = g1.w;
    }
  }
  const unsigned long long dstream = drop.thresh ? drop_stream(drop) : 0ull;
  float acc_g[CH][8], acc_b[CH][8], acc_c[CH][8];
#pragma unroll
  for (int c = 0; c < CH; ++c)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_g[c][j] = acc_b[c][j] = acc_c[c][j] = 0.f;

#pragma unroll
  for (int r = 0; r < RPW; ++r) {
    const int row = row0 + r;
    if (row < rows) {  // warp-uniform
      const size_t base = size_t(row) * d;
      float xh[CH][8], g[CH][8];
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int col = (c * 32 + lane) * 8;
        if (col < d) {
          float a[8], b[8];
          unpack8(u_dy[r][c], a);
          unpack8(u_z[r][c], b);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            xh[c][j] = (b[j] - mean[r]) * rstd[r];
            g[c][j] = a[j] * gam[c][j];
            s1 += g[c][j]; s2 += g[c][j] * xh[c][j];
            acc_g[c][j] += a[j] * xh[c][j];
            acc_b[c][j] += a[j];
          }
        }
      }
      const float m1 = warp_sum(s1) / float(d), m2 = warp_sum(s2) / float(d);
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int col = (c * 32 + lane) * 8;
        if (col < d) {
          float o[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = rstd[r] * (g[c][j] - m1 - xh[c][j] * m2);
          *reinterpret_cast<uint4*>(dz + base + col) = pack8(o);
          if (drop.thresh) {  // gradient seen by the Dense whose (dropped-out) output was added to the residual
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = drop_keep(dstream, base + col + j, drop.thresh) ? o[j] * drop.scale : 0.f;
            *reinterpret_cast<uint4*>(dzm + base + col) = pack8(o);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) acc_c[c][j] += round_bf16(o[j]);
        }
      }
    }
  }
  // block reduce: every warp parks its sums in shared memory, then all threads add the 8 copies in warp order
  float* mine_s = red + size_t(warp) * 3 * d;
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int col = (c * 32 + lane) * 8;
    if (col < d) {
#pragma unroll
      for (int j = 0; j < 8; j += 4) {
        *reinterpret_cast<float4*>(mine_s + col + j) = make_float4(acc_g[c][j], acc_g[c][j + 1], acc_g[c][j + 2], acc_g[c][j + 3]);
        *reinterpret_cast<float4*>(mine_s + d + col + j) = make_float4(acc_b[c][j], acc_b[c][j + 1], acc_b[c][j + 2], acc_b[c][j + 3]);
        *reinterpret_cast<float4*>(mine_s + 2 * d + col + j) = make_float4(acc_c[c][j], acc_c[c][j + 1], acc_c[c][j + 2], acc_c[c][j + 3]);
      }
    }
  }
  __syncthreads();
  float* mine = partials + size_t(blockIdx.x) * 3 * d;
  for (int i = threadIdx.x * 4; i < 3 * d; i += blockDim.x * 4) {
    float4 s = *reinterpret_cast<const float4*>(red + i);
#pragma unroll
    for (int w = 1; w < kLnWarps; ++w) {
      const float4 t = *reinterpret_cast<const float4*>(red + size_t(w) * 3 * d + i);
      s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
    }
    *reinterpret_cast<float4*>(mine + i) = s;
  }
}

// dst[c] = sum_p src[p * stride + c] for a batch of jobs; block = 32 columns x 32 slices of the partials.
struct ReduceJobTable {
  static constexpr int kMax = 96;
  const float* src[kMax]; float* dst[kMax];
  int nparts[kMax], cols[kMax], stride[kMax], first_block[kMax + 1];
  int njobs;
};
__global__ void __launch_bounds__(1024)
reduce_partials_kernel(const __grid_constant__ ReduceJobTable t) {
  pdl_wait();
  pdl_trigger();
  __shared__ float red[32][33];
  int j = 0;
  while (j + 1 < t.njobs && int(blockIdx.x) >= t.first_block[j + 1]) ++j;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col = (blockIdx.x - t.first_block[j]) * 32 + tx;
  float acc = 0.f;
  if (col < t.cols[j]) {
    const float* s = t.src[j] + col;
    const int n = t.nparts[j], st = t.stride[j];
#pragma unroll 4
    for (int p = ty; p < n; p += 32) acc += __ldcg(s + size_t(p) * st);
  }
  red[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && col < t.cols[j]) {
    float v = red[0][tx];
#pragma unroll
    for (int i = 1; i < 32; ++i) v += red[i][tx];
    t.dst[j][col] = v;
  }
}

}  // namespace

int layernorm_fwd(const bf16* x, const bf16* resid, const float* gamma, const float* beta, bf16* y, bf16* z_out, float* mean,
                  float* rstd, int rows, int d, float eps, cudaStream_t stream) {
  if (d % 8 || d > 2048) return set_error("layernorm: d=%d must be a multiple of 8 and <= 2048", d);
  if (rows <= 0) return 0;
  const int grid = (rows + kLnWarps - 1) / kLnWarps;
  const int ch = (d + 255) / 256;
#define LN_FWD(CH) B200XF_CUDA(launch_k(ln_fwd_kernel<CH>, grid, kLnWarps * 32, 0, stream, x, resid, gamma, beta, y, z_out, mean, rstd, rows, d, eps))
  if (ch <= 1) LN_FWD(1); else if (ch <= 2) LN_FWD(2); else if (ch <= 4) LN_FWD(4); else LN_FWD(8);
#undef LN_FWD
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

namespace {
int ln_rows_per_warp(int d) { return d <= 512 ? 2 : 1; }  // keeps the rows-in-flight register budget bounded
}
int ln_bwd_parts(int rows, int d) { const int per_block = kLnWarps * ln_rows_per_warp(d); return (rows + per_block - 1) / per_block; }
size_t ln_bwd_scratch_floats(int rows, int d) { return size_t(ln_bwd_parts(rows, d)) * 3 * d; }

int layernorm_bwd_partial(const bf16* dy, const bf16* z, const float* mean, const float* rstd, const float* gamma, bf16* dz, int rows,
                          int d, float* partials, cudaStream_t stream, bf16* dzm, DropArgs drop) {
  if (drop.thresh && !dzm) return set_error("layernorm_bwd: dropout needs a dzm buffer");
  if (d % 8 || d > 2048) return set_error("layernorm: d=%d must be a multiple of 8 and <= 2048", d);
  if (rows <= 0) return set_error("layernorm_bwd: no rows");
  const int grid = ln_bwd_parts(rows, d);
  const int ch = (d + 255) / 256;
  const size_t smem = size_t(kLnWarps) * 3 * d * sizeof(float);
  auto launch = [&](auto kern) -> int {
    if (smem > 48 * 1024) B200XF_TRY(ensure_dynamic_smem(kern, int(kLnWarps * 3 * 2048 * sizeof(float))));
    B200XF_CUDA(launch_k(kern, grid, kLnWarps * 32, smem, stream, dy, z, mean, rstd, gamma, dz, rows, d, partials, dzm, drop));
    return 0;
  };
  if (ch <= 1) B200XF_TRY(launch(ln_bwd_kernel<1, 2>));
  else if (ch <= 2) B200XF_TRY(launch(ln_bwd_kernel<2, 2>));
  else if (ch <= 4) B200XF_TRY(launch(ln_bwd_kernel<4, 1>));
  else B200XF_TRY(launch(ln_bwd_kernel<8, 1>));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

int reduce_partials(const ReduceJob* jobs, int njobs, cudaStream_t stream) {
  for (int base = 0; base < njobs; base += ReduceJobTable::kMax) {
    ReduceJobTable t;
    t.njobs = std::min(ReduceJobTable::kMax, njobs - base);
    int blocks = 0;
    for (int i = 0; i < t.njobs; ++i) {
      const ReduceJob& j = jobs[base + i];
      t.src[i] = j.src; t.dst[i] = j.dst; t.nparts[i] = j.nparts; t.cols[i] = j.cols; t.stride[i] = j.stride;
      t.first_block[i] = blocks;
      blocks += (j.cols + 31) / 32;
    }
    t.first_block[t.njobs] = blocks;
    if (blocks == 0) continue;
    B200XF_CUDA(launch_k(reduce_partials_kernel, blocks, 1024, 0, stream, t));
    B200XF_CUDA(cudaGetLastError());
    count_launch();
  }
  return 0;
}

int layernorm_bwd(const bf16* dy, const bf16* z, const float* mean, const float* rstd, const float* gamma, bf16* dz,
                  float* dgamma, float* dbeta, float* dcol, int rows, int d, float* scratch, cudaStream_t stream) {
  if (rows <= 0) return 0;
  B200XF_TRY(layernorm_bwd_partial(dy, z, mean, rstd, gamma, dz, rows, d, scratch, stream));
  const int n = ln_bwd_parts(rows, d);
  ReduceJob jobs[3] = {{scratch, dgamma, n, d, 3 * d}, {scratch + d, dbeta, n, d, 3 * d}, {scratch + 2 * d, dcol, n, d, 3 * d}};
  return reduce_partials(jobs, dcol ? 3 : 2, stream);
}

// ============================================================================ embedding + PE
namespace {

// x[b,s,:] = table[token] * sqrt(d) + pe[s]   (fp32 master rows gathered bit-exactly, output bf16)
__global__ void embed_fwd_kernel(const long long* __restrict__ tokens, const float* __restrict__ table, int vocab,
                                 const float* __restrict__ pe, bf16* __restrict__ out, int S, int d, float scale, const DropArgs drop) {
  pdl_wait();
  pdl_trigger();
  const int n = blockIdx.x;
  const unsigned long long dstream = drop.thresh ? drop_stream(drop) : 0ull;
  const long long tok = tokens[n];
  const bool ok = tok >= 0 && tok < vocab;  // out-of-range ids read as a zero row (TF GPU gather semantics)
  const float* row = table + size_t(ok ? tok : 0) * d;
  const float* p = pe + size_t(n % S) * d;
  for (int c = threadIdx.x * 4; c < d; c += blockDim.x * 4) {
    float4 e = ok ? __ldg(reinterpret_cast<const float4*>(row + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 q = __ldg(reinterpret_cast<const float4*>(p + c));
    uint2 u;
    // two separately rounded fp32 ops, as `x *= sqrt(d); x += pe` in the reference (no FMA contraction)
    float v[4] = {__fadd_rn(__fmul_rn(e.x, scale), q.x), __fadd_rn(__fmul_rn(e.y, scale), q.y),
                  __fadd_rn(__fmul_rn(e.z, scale), q.z), __fadd_rn(__fmul_rn(e.w, scale), q.w)};
    if (drop.thresh) {  // transformer_model.py:276,308
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = drop_keep(dstream, size_t(n) * d + c + j, drop.thresh) ? v[j] * drop.scale : 0.f;
    }
    u.x = ptx::pack_bf16x2(v[0], v[1]);
    u.y = ptx::pack_bf16x2(v[2], v[3]);
    *reinterpret_cast<uint2*>(out + size_t(n) * d + c) = u;
  }
}

__global__ void embed_keys_kernel(const long long* __restrict__ tokens, int* __restrict__ keys, int* __restrict__ vals, int n, int vocab) {
  pdl_wait();
  pdl_trigger();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long t = tokens[i];
  keys[i] = (t >= 0 && t < vocab) ? int(t) : vocab;  // out-of-range ids sort last and are skipped
  vals[i] = i;
}

// One block per sorted position; the block at the head of a run of equal token ids sums that run's
// gradient rows in sorted (= original position) order and writes the table row once: deterministic.
__global__ void embed_bwd_kernel(const int* __restrict__ keys, const int* __restrict__ vals, const bf16* __restrict__ dx,
                                 float* __restrict__ dtable, int n, int d, int vocab, float scale, const DropArgs drop) {
  pdl_wait();
  pdl_trigger();
  const int p = blockIdx.x;
  const unsigned long long dstream = drop.thresh ? drop_stream(drop) : 0ull;
  const int key = keys[p];
  if (key >= vocab || (p > 0 && keys[p - 1] == key)) return;
  int end = p + 1;
  while (end < n && keys[end] == key) ++end;
  for (int c = threadIdx.x * 4; c < d; c += blockDim.x * 4) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    for (int q = p; q < end; ++q) {
      const size_t pos = size_t(vals[q]) * d + c;
      const uint2 u = *reinterpret_cast<const uint2*>(dx + pos);
      float g0 = ptx::bf16_lo(u.x), g1 = ptx::bf16_hi(u.x), g2 = ptx::bf16_lo(u.y), g3 = ptx::bf16_hi(u.y);
      if (drop.thresh) {
        g0 = drop_keep(dstream, pos, drop.thresh) ? g0 * drop.scale : 0.f; g1 = drop_keep(dstream, pos + 1, drop.thresh) ? g1 * drop.scale : 0.f;
        g2 = drop_keep(dstream, pos + 2, drop.thresh) ? g2 * drop.scale : 0.f; g3 = drop_keep(dstream, pos + 3, drop.thresh) ? g3 * drop.scale : 0.f;
      }
      a0 += g0; a1 += g1; a2 += g2; a3 += g3;
    }
    float4* dst = reinterpret_cast<float4*>(dtable + size_t(key) * d + c);
    float4 cur = *dst;
    cur.x += a0 * scale; cur.y += a1 * scale; cur.z += a2 * scale; cur.w += a3 * scale;
    *dst = cur;
  }
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Sort of the (token id, position) pairs of one batch side in ONE CTA, for batches of up to kSortMax (4096) tokens: the combined key
// (token << 13 | position) is unique, so a bitonic network (not stable by itself) yields the same order as a stable sort by
// token: ascending token, positions ascending within a token -- the order the gradient rows are summed in (deterministic).
// Replaces key extraction + a library radix sort (two launches) for the Transformer-base / -big batch sizes.
constexpr int kSortMax = 4096, kSortThreads = 1024;
// Element e lives in thread (e % 1024), register (e / 1024).  Compare-exchange distances below 32 run as warp shuffles, distances
// of 1024 and more inside a thread's own registers; only the distances 32..512 go through shared memory (one barrier each).
template <int R>
__device__ __forceinline__ void bitonic_sort_regs(unsigned (&x)[R], unsigned* buf) {
  const int t = threadIdx.x;
  int flip = 0;
  for (int k = 2; k <= R * kSortThreads; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= kSortThreads) {
        const int jr = j / kSortThreads;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int l = r ^ jr;
          if (l > r) {
            const bool up = ((r * kSortThreads + t) & k) == 0;
            const unsigned a = x[r], b = x[l];
            if ((a > b) == up) { x[r] = b; x[l] = a; }
          }
        }
      } else if (j >= 32) {
        unsigned* bb = buf + flip * (R * kSortThreads);
        flip ^= 1;
#pragma unroll
        for (int r = 0; r < R; ++r) bb[r * kSortThreads + t] = x[r];
        __syncthreads();
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int i = r * kSortThreads + t;
          const unsigned other = bb[i ^ j];
          const bool up = (i & k) == 0, lower = (i & j) == 0;
          const unsigned mn = min(x[r], other), mx = max(x[r], other);
          x[r] = (lower == up) ? mn : mx;
        }
      } else {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int i = r * kSortThreads + t;
          const unsigned other = __shfl_xor_sync(0xffffffffu, x[r], j);
          const bool up = (i & k) == 0, lower = (i & j) == 0;
          const unsigned mn = min(x[r], other), mx = max(x[r], other);
          x[r] = (lower == up) ? mn : mx;
        }
      }
    }
  }
}

template <int R>
__global__ void __launch_bounds__(kSortThreads)
embed_sort_kernel(const long long* __restrict__ tokens, int* __restrict__ keys_out, int* __restrict__ vals_out, int n, int vocab) {
  pdl_wait();
  pdl_trigger();
  __shared__ unsigned buf[2 * R * kSortThreads];
  unsigned x[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int i = r * kSortThreads + threadIdx.x;
    unsigned key = 0xFFFFFFFFu;  // padding sorts last
    if (i < n) {
      const long long tk = tokens[i];
      const unsigned tok = (tk >= 0 && tk < vocab) ? unsigned(tk) : unsigned(vocab);  // out-of-range ids sort after every real id and are skipped
      key = (tok << 13) | unsigned(i);
    }
    x[r] = key;
  }
  bitonic_sort_regs<R>(x, buf);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int i = r * kSortThreads + threadIdx.x;
    if (i < n) { keys_out[i] = int(x[r] >> 13); vals_out[i] = int(x[r] & 8191u); }
  }
}

}  // namespace

int embed_pe_fwd(const long long* tokens, const float* table, int vocab, const float* pe, bf16* out, int B, int S, int d,
                 cudaStream_t stream, DropArgs drop) {
  if (d % 4) return set_error("embedding: d=%d must be a multiple of 4", d);
  const int n = B * S;
  if (n <= 0) return 0;
  const int threads = min(256, max(32, (d / 4 + 31) / 32 * 32));
  B200XF_CUDA(launch_k(embed_fwd_kernel, n, threads, 0, stream, tokens, table, vocab, pe, out, S, d, sqrtf(float(d)), drop));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

bool embed_sort_is_single_cta(int n, int vocab);

size_t embed_bwd_scratch_bytes(int n) {
  size_t temp = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, temp, (const int*)nullptr, (int*)nullptr, (const int*)nullptr, (int*)nullptr, n);
  return align_up(temp, 256) + 4 * align_up(size_t(n) * sizeof(int), 256);
}

// phase 1: (token, position) pairs of the batch side sorted by token (positions ascending within a token) into the scratch
int embed_bwd_sort(const long long* tokens, int vocab, int n, void* scratch, size_t scratch_bytes, cudaStream_t stream) {
  if (n <= 0) return 0;
  if (scratch_bytes < embed_bwd_scratch_bytes(n)) return set_error("embed_bwd: scratch too small");
  const size_t seg = align_up(size_t(n) * sizeof(int), 256);
  char* base = static_cast<char*>(scratch);
  int* keys_in = reinterpret_cast<int*>(base);
  int* keys_out = reinterpret_cast<int*>(base + seg);
  int* vals_in = reinterpret_cast<int*>(base + 2 * seg);
  int* vals_out = reinterpret_cast<int*>(base + 3 * seg);
  void* temp = base + 4 * seg;
  size_t temp_bytes = scratch_bytes - 4 * seg;
  if (embed_sort_is_single_cta(n, vocab)) {
    if (n <= 1024) B200XF_CUDA(launch_k(embed_sort_kernel<1>, 1, kSortThreads, 0, stream, tokens, keys_out, vals_out, n, vocab));
    else if (n <= 2048) B200XF_CUDA(launch_k(embed_sort_kernel<2>, 1, kSortThreads, 0, stream, tokens, keys_out, vals_out, n, vocab));
    else B200XF_CUDA(launch_k(embed_sort_kernel<4>, 1, kSortThreads, 0, stream, tokens, keys_out, vals_out, n, vocab));
    count_launch(1);
  } else {  // long-sequence batches: key extraction + device-wide radix sort
    B200XF_CUDA(launch_k(embed_keys_kernel, (n + 255) / 256, 256, 0, stream, tokens, keys_in, vals_in, n, vocab));
    int bits = 1;
    while ((1ll << bits) <= vocab) ++bits;
    B200XF_CUDA(cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0, bits, stream));
    count_launch(2);
  }
  B200XF_CUDA(cudaGetLastError());
  return 0;
}

// phase 2: the head of every run of equal tokens sums the run's gradient rows in order into the table's gradient
int embed_bwd_sorted(const bf16* dx, float* dtable, int vocab, int n, int d, const void* scratch, cudaStream_t stream, DropArgs drop) {
  if (n <= 0) return 0;
  const size_t seg = align_up(size_t(n) * sizeof(int), 256);
  const char* base = static_cast<const char*>(scratch);
  const int* keys_out = reinterpret_cast<const int*>(base + seg);
  const int* vals_out = reinterpret_cast<const int*>(base + 3 * seg);
  const int threads = min(256, max(32, (d / 4 + 31) / 32 * 32));
  B200XF_CUDA(launch_k(embed_bwd_kernel, n, threads, 0, stream, keys_out, vals_out, dx, dtable, n, d, vocab, sqrtf(float(d)), drop));
  B200XF_CUDA(cudaGetLastError());
  count_launch(1);
  return 0;
}

bool embed_sort_is_single_cta(int n, int vocab) { return n <= kSortMax && vocab < (1 << 19); }

int embed_bwd(const long long* tokens, const bf16* dx, float* dtable, int vocab, int n, int d, void* scratch, size_t scratch_bytes,
              cudaStream_t stream, DropArgs drop) {
  B200XF_TRY(embed_bwd_sort(tokens, vocab, n, scratch, scratch_bytes, stream));
  return embed_bwd_sorted(dx, dtable, vocab, n, d, scratch, stream, drop);
}

namespace {
__global__ void dropout_mask_kernel(float* __restrict__ out, size_t n, const DropArgs drop) {
  const unsigned long long dstream = drop_stream(drop);
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
    out[i] = drop_keep(dstream, i, drop.thresh) ? drop.scale : 0.f;
}
}  // namespace

int dropout_mask_fill(float* out, int rows, int cols, DropArgs drop, cudaStream_t stream) {
  const size_t n = size_t(rows) * cols;
  if (n == 0) return 0;
  dropout_mask_kernel<<<int(std::min<size_t>((n + 255) / 256, 4096)), 256, 0, stream>>>(out, n, drop);
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

// ============================================================================ targets
namespace {
__global__ void prep_targets_kernel(const long long* __restrict__ target, long long* __restrict__ tgt_in,
                                    long long* __restrict__ labels, int* __restrict__ count, int B, int T1) {
  pdl_wait();
  pdl_trigger();
  __shared__ int warp_cnt[32];
  const int T = T1 - 1;
  int local = 0;
  for (int i = threadIdx.x; i < B * T; i += blockDim.x) {
    const int b = i / T, t = i % T;
    const long long a = target[size_t(b) * T1 + t], l = target[size_t(b) * T1 + t + 1];
    tgt_in[i] = a; labels[i] = l;
    local += (l != 0);
  }
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
    for (int w = 0; w < (blockDim.x >> 5); ++w) s += warp_cnt[w];
    *count = s;
  }
}
}  // namespace

int prep_targets(const long long* target, long long* tgt_in, long long* labels, int* nonpad_count, int B, int T_plus_1,
                 cudaStream_t stream) {
  if (T_plus_1 < 2) return set_error("target must have at least 2 positions");
  B200XF_CUDA(launch_k(prep_targets_kernel, 1, 1024, 0, stream, target, tgt_in, labels, nonpad_count, B, T_plus_1));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

// ============================================================================ softmax cross-entropy
namespace {
constexpr int kCeThreads = 256;

using ptx::fast_exp2;

struct MaxSum { float m, s; };
__device__ __forceinline__ MaxSum combine(MaxSum a, MaxSum b) {
  const float m = fmaxf(a.m, b.m);
  if (m == -INFINITY) return {m, 0.f};
  return {m, a.s * fast_exp2((a.m - m) * 1.4426950408889634f) + b.s * fast_exp2((b.m - m) * 1.4426950408889634f)};
}

// One block per target position: pass 1 = online max/sum-exp, first-max argmax, label logit;
// pass 2 (row now L2-resident) = dlogits = (softmax - onehot) * mask / (count * workers), in place.
constexpr float kLog2eCe = 1.4426950408889634f;

constexpr int kCeUnroll = 4;
__global__ void __launch_bounds__(kCeThreads)
vocab_ce_kernel(bf16* __restrict__ logits, int ldl, const long long* __restrict__ labels, int V,
                const int* __restrict__ nonpad_count, float inv_workers, float* __restrict__ row_loss,
                int* __restrict__ row_correct, int write_grad) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sm_m[kCeThreads / 32], sm_s[kCeThreads / 32], sm_am[kCeThreads / 32];
  __shared__ int sm_ai[kCeThreads / 32];
  __shared__ float bc[3];
  __shared__ int bci;
  const int row = blockIdx.x;
  bf16* x = logits + size_t(row) * ldl;
  const long long label = labels[row];
  const int nvec = (V + 7) / 8;
  if (label == 0) {  // PAD target: contributes nothing (transformer_model.py:12-15)
    if (write_grad)
      for (int i = threadIdx.x; i < nvec; i += kCeThreads) {
        if (i * 8 + 8 <= V) *reinterpret_cast<uint4*>(x + i * 8) = make_uint4(0, 0, 0, 0);
        else for (int j = i * 8; j < V; ++j) x[j] = __float2bfloat16(0.f);
      }
    if (threadIdx.x == 0) { row_loss[row] = 0.f; row_correct[row] = 0; }
    return;
  }
  // running (max, sum of 2^((x - max) log2 e)) per thread: the sum is rescaled only when the maximum moves, which after
  // the first few vectors is rare -- one MUFU per logit instead of 1.25
  MaxSum ms{-INFINITY, 0.f};
  float am = -INFINITY; int ai = 0x7fffffff;
  float mneg = 0.f;  // -max * log2(e), kept ready for the FMA below
  // kCeUnroll independent 16-byte loads per thread and iteration: one load in flight per thread (32 KB per SM) cannot
  // cover HBM latency at more than ~2.5 TB/s
  for (int i0 = threadIdx.x; i0 < nvec; i0 += kCeThreads * kCeUnroll) {
    uint4 u[kCeUnroll];
#pragma unroll
    for (int k = 0; k < kCeUnroll; ++k) {
      const int i = i0 + k * kCeThreads;
      if (i < nvec && i * 8 + 8 <= V) u[k] = *reinterpret_cast<const uint4*>(x + i * 8);
    }
#pragma unroll
    for (int k = 0; k < kCeUnroll; ++k) {
      const int i = i0 + k * kCeThreads;
      if (i >= nvec) break;
      float f[8];
      if (i * 8 + 8 <= V) unpack8(u[k], f);
      else for (int j = 0; j < 8; ++j) f[j] = (i * 8 + j < V) ? __bfloat162float(x[i * 8 + j]) : -INFINITY;
      float lm = f[0];
#pragma unroll
      for (int j = 1; j < 8; ++j) lm = fmaxf(lm, f[j]);
      if (lm > am) {  // strictly greater keeps the earliest index within this thread's ascending scan
#pragma unroll
        for (int j = 7; j >= 0; --j) if (f[j] == lm) ai = i * 8 + j;
        am = lm;
        ms.s *= fast_exp2((ms.m - lm) * kLog2eCe);  // 2^(-inf) = 0 the first time
        ms.m = lm; mneg = -lm * kLog2eCe;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) ms.s += fast_exp2(fmaf(f[j], kLog2eCe, mneg));
    }
  }
  // warp then block reduction
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MaxSum other{__shfl_xor_sync(0xffffffffu, ms.m, o), __shfl_xor_sync(0xffffffffu, ms.s, o)};
    ms = combine(ms, other);
    const float om = __shfl_xor_sync(0xffffffffu, am, o); const int oi = __shfl_xor_sync(0xffffffffu, ai, o);
    if (om > am || (om == am && oi < ai)) { am = om; ai = oi; }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { sm_m[warp] = ms.m; sm_s[warp] = ms.s; sm_am[warp] = am; sm_ai[warp] = ai; }
  __syncthreads();
  if (threadIdx.x == 0) {
    MaxSum t{sm_m[0], sm_s[0]}; float bm = sm_am[0]; int bi = sm_ai[0];
    for (int w = 1; w < kCeThreads / 32; ++w) {
      t = combine(t, MaxSum{sm_m[w], sm_s[w]});
      if (sm_am[w] > bm || (sm_am[w] == bm && sm_ai[w] < bi)) { bm = sm_am[w]; bi = sm_ai[w]; }
    }
    const bool label_ok = label > 0 && label < V;
    const float xl = label_ok ? __bfloat162float(x[label]) : 0.f;
    const float lse = t.m + logf(t.s);
    row_loss[row] = lse - xl;
    row_correct[row] = (bi == int(label)) ? 1 : 0;
    bc[0] = t.m; bc[1] = 1.0f / t.s;
    bc[2] = inv_workers / float(max(*nonpad_count, 1));
    bci = label_ok ? int(label) : -1;
  }
  if (!write_grad) return;
  __syncthreads();
  const float mx = bc[0], inv_s = bc[1], scale = bc[2];
  const int lab = bci;
  const float mxneg = -mx * kLog2eCe, ps = inv_s * scale;
  for (int i0 = threadIdx.x; i0 < nvec; i0 += kCeThreads * kCeUnroll) {
    uint4 u[kCeUnroll];
#pragma unroll
    for (int k = 0; k < kCeUnroll; ++k) {
      const int i = i0 + k * kCeThreads;
      if (i < nvec && i * 8 + 8 <= V) u[k] = *reinterpret_cast<const uint4*>(x + i * 8);
    }
#pragma unroll
    for (int k = 0; k < kCeUnroll; ++k) {
      const int i = i0 + k * kCeThreads;
      if (i >= nvec) break;
      float f[8];
      const bool fullv = i * 8 + 8 <= V;
      if (fullv) unpack8(u[k], f);
      else for (int j = 0; j < 8; ++j) f[j] = (i * 8 + j < V) ? __bfloat162float(x[i * 8 + j]) : -INFINITY;
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = fast_exp2(fmaf(f[j], kLog2eCe, mxneg)) * ps;   // softmax * mask / (count * workers)
      if ((lab >> 3) == i) {  // the label's vector (lab = -1 never matches): subtract the one-hot term
#pragma unroll
        for (int j = 0; j < 8; ++j) if (j == (lab & 7)) f[j] -= scale;
      }
      if (fullv) *reinterpret_cast<uint4*>(x + i * 8) = pack8(f);
      else for (int j = 0; j < 8; ++j) if (i * 8 + j < V) x[i * 8 + j] = __float2bfloat16(f[j]);
    }
  }
}

// fp32 logits, loss / accuracy only (local_loss_function / accuracy_function on the logits Transformer.call returns,
// transformer_model.py:11-26): same online max / sum and first-max argmax as above, nothing rounded to bf16, no gradient.
__global__ void __launch_bounds__(kCeThreads)
vocab_ce_f32_kernel(const float* __restrict__ logits, int ldl, const long long* __restrict__ labels, int V,
                    float* __restrict__ row_loss, int* __restrict__ row_correct) {
  __shared__ float sm_m[kCeThreads / 32], sm_s[kCeThreads / 32], sm_am[kCeThreads / 32];
  __shared__ int sm_ai[kCeThreads / 32];
  const int row = blockIdx.x;
  const float* x = logits + size_t(row) * ldl;
  const long long label = labels[row];
  if (label == 0) {
    if (threadIdx.x == 0) { row_loss[row] = 0.f; row_correct[row] = 0; }
    return;
  }
  MaxSum ms{-INFINITY, 0.f};
  float am = -INFINITY; int ai = 0x7fffffff;
  for (int i = threadIdx.x; i < V; i += kCeThreads) {
    const float f = x[i];
    if (f > am) { am = f; ai = i; ms.s *= fast_exp2((ms.m - f) * kLog2eCe); ms.m = f; }
    ms.s += fast_exp2((f - ms.m) * kLog2eCe);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MaxSum other{__shfl_xor_sync(0xffffffffu, ms.m, o), __shfl_xor_sync(0xffffffffu, ms.s, o)};
    ms = combine(ms, other);
    const float om = __shfl_xor_sync(0xffffffffu, am, o); const int oi = __shfl_xor_sync(0xffffffffu, ai, o);
    if (om > am || (om == am && oi < ai)) { am = om; ai = oi; }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { sm_m[warp] = ms.m; sm_s[warp] = ms.s; sm_am[warp] = am; sm_ai[warp] = ai; }
  __syncthreads();
  if (threadIdx.x == 0) {
    MaxSum t{sm_m[0], sm_s[0]}; float bm = sm_am[0]; int bi = sm_ai[0];
    for (int w = 1; w < kCeThreads / 32; ++w) {
      t = combine(t, MaxSum{sm_m[w], sm_s[w]});
      if (sm_am[w] > bm || (sm_am[w] == bm && sm_ai[w] < bi)) { bm = sm_am[w]; bi = sm_ai[w]; }
    }
    const float xl = (label > 0 && label < V) ? x[label] : 0.f;
    row_loss[row] = t.m + logf(t.s) - xl;
    row_correct[row] = (bi == int(label)) ? 1 : 0;
  }
}

__global__ void ce_finalize_kernel(const float* __restrict__ row_loss, const int* __restrict__ row_correct,
                                   const int* __restrict__ nonpad_count, int rows, float inv_workers, float* __restrict__ out3) {
  pdl_wait();
  pdl_trigger();
  __shared__ float wl[32]; __shared__ int wc[32];
  float l = 0.f; int c = 0;
  // fixed assignment of rows to threads and a fixed reduction tree: deterministic
  for (int i = threadIdx.x; i < rows; i += blockDim.x) { l += row_loss[i]; c += row_correct[i]; }
  for (int o = 16; o > 0; o >>= 1) { l += __shfl_xor_sync(0xffffffffu, l, o); c += __shfl_xor_sync(0xffffffffu, c, o); }
  if ((threadIdx.x & 31) == 0) { wl[threadIdx.x >> 5] = l; wc[threadIdx.x >> 5] = c; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float L = 0.f; int C = 0;
    for (int w = 0; w < (blockDim.x >> 5); ++w) { L += wl[w]; C += wc[w]; }
    const int cnt = *nonpad_count;
    out3[0] = L / float(cnt) * inv_workers;  // 0/0 -> NaN for an all-PAD batch, as the reference
    out3[1] = float(C);
    out3[2] = float(cnt);
  }
}
}  // namespace

int vocab_ce_fwd_bwd(bf16* logits, int ldl, const long long* labels, int rows, int V, const int* nonpad_count, float inv_workers,
                     float* row_loss, int* row_correct, int write_grad, cudaStream_t stream) {
  if (ldl % 8) return set_error("vocab_ce: ld=%d must be a multiple of 8", ldl);
  if (rows <= 0) return 0;
  B200XF_CUDA(launch_k(vocab_ce_kernel, rows, kCeThreads, 0, stream, logits, ldl, labels, V, nonpad_count, inv_workers, row_loss, row_correct, write_grad));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

int vocab_ce_f32(const float* logits, int ldl, const long long* labels, int rows, int V, float* row_loss, int* row_correct, cudaStream_t stream) {
  if (rows <= 0) return 0;
  vocab_ce_f32_kernel<<<rows, kCeThreads, 0, stream>>>(logits, ldl, labels, V, row_loss, row_correct);
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

int ce_finalize(const float* row_loss, const int* row_correct, const int* nonpad_count, int rows, float inv_workers, float* out3,
                cudaStream_t stream) {
  B200XF_CUDA(launch_k(ce_finalize_kernel, 1, 1024, 0, stream, row_loss, row_correct, nonpad_count, rows, inv_workers, out3));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

// ============================================================================ Adam + schedule
namespace {
// TF-form Adam (ResourceApplyAdam) with the reference's warm-up schedule evaluated on the device from
// `iterations` (the value BEFORE this step's increment, so the very first step has lr == 0).
constexpr int kAdamPerBlock = 8;
__global__ void __launch_bounds__(256)
adam_kernel(float4* __restrict__ w, float4* __restrict__ g, float4* __restrict__ m, float4* __restrict__ v,
            uint2* __restrict__ wb, size_t n4, const long long* __restrict__ iterations, AdamHyper h, int zero_grad) {
  pdl_wait();
  pdl_trigger();
  const float step = float(*iterations);
  const float lr = h.rsqrt_d_model * fminf(step * h.warmup_pow_m15, 1.0f / sqrtf(step));
  const float t = step + 1.0f;
  const float alpha = lr * sqrtf(1.0f - powf(h.beta2, t)) / (1.0f - powf(h.beta1, t));
  const float c1 = 1.0f - h.beta1, c2 = 1.0f - h.beta2;
  // short-lived CTAs (8 x 256 float4 each): the update runs on a low-priority stream under the backward pass, and the
  // block scheduler can only hand an SM back to the critical path when a CTA retires
  const size_t base = size_t(blockIdx.x) * (kAdamPerBlock * 256) + threadIdx.x;
#pragma unroll 2
  for (int kk = 0; kk < kAdamPerBlock; ++kk) {
    const size_t i = base + size_t(kk) * 256;
    if (i >= n4) break;
    float4 W = w[i], M = m[i], V = v[i];
    const float4 G = g[i];
    // each operation rounded separately, exactly as the expression is written (no FMA contraction)
#define ADAM1(c)                                                                              \
    M.c = __fadd_rn(M.c, __fmul_rn(__fsub_rn(G.c, M.c), c1));                                 \
    V.c = __fadd_rn(V.c, __fmul_rn(__fsub_rn(__fmul_rn(G.c, G.c), V.c), c2));                 \
    W.c = __fsub_rn(W.c, __fdiv_rn(__fmul_rn(M.c, alpha), __fadd_rn(__fsqrt_rn(V.c), h.eps)));
    ADAM1(x) ADAM1(y) ADAM1(z) ADAM1(w)
#undef ADAM1
    w[i] = W; m[i] = M; v[i] = V;
    if (wb) { uint2 u; u.x = ptx::pack_bf16x2(W.x, W.y); u.y = ptx::pack_bf16x2(W.z, W.w); wb[i] = u; }
    if (zero_grad) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}
__global__ void increment_kernel(long long* c) {
  pdl_wait();
  pdl_trigger(); *c += 1; }
}  // namespace

int adam_step(float* w, float* g, float* m, float* v, bf16* w_bf16, size_t n, const long long* iterations_dev, AdamHyper h,
              int zero_grad, cudaStream_t stream) {
  if (n % 4) return set_error("adam: n=%zu must be a multiple of 4", n);
  if (n == 0) return 0;
  const size_t n4 = n / 4;
  const int grid = int((n4 + size_t(kAdamPerBlock) * 256 - 1) / (size_t(kAdamPerBlock) * 256));
  B200XF_CUDA(launch_k(adam_kernel, grid, 256, 0, stream, reinterpret_cast<float4*>(w), reinterpret_cast<float4*>(g), reinterpret_cast<float4*>(m),
                       reinterpret_cast<float4*>(v), reinterpret_cast<uint2*>(w_bf16), n4, iterations_dev, h, zero_grad));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

namespace {
__global__ void __launch_bounds__(256)
refresh_kernel(const float4* __restrict__ w, uint2* __restrict__ wb, float4* __restrict__ g, size_t n4) {
  pdl_wait();
  pdl_trigger();
  const size_t base = size_t(blockIdx.x) * (kAdamPerBlock * 256) + threadIdx.x;
#pragma unroll 4
  for (int kk = 0; kk < kAdamPerBlock; ++kk) {
    const size_t i = base + size_t(kk) * 256;
    if (i >= n4) break;
    const float4 W = w[i];
    uint2 u; u.x = ptx::pack_bf16x2(W.x, W.y); u.y = ptx::pack_bf16x2(W.z, W.w);
    wb[i] = u;
    g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}
}  // namespace

int refresh_shadow_zero_grad(const float* w, bf16* shadow, float* g, size_t n, cudaStream_t stream) {
  if (n % 4) return set_error("refresh: n must be a multiple of 4");
  if (n == 0) return 0;
  const size_t n4 = n / 4;
  const int grid = int((n4 + size_t(kAdamPerBlock) * 256 - 1) / (size_t(kAdamPerBlock) * 256));
  B200XF_CUDA(launch_k(refresh_kernel, grid, 256, 0, stream, reinterpret_cast<const float4*>(w), reinterpret_cast<uint2*>(shadow),
                       reinterpret_cast<float4*>(g), n4));
  count_launch();
  return 0;
}

int increment_counter(long long* iterations_dev, cudaStream_t stream) {
  B200XF_CUDA(launch_k(increment_kernel, 1, 1, 0, stream, iterations_dev));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

// ============================================================================ helpers
namespace {
// partial[strip][c] = sum over the strip's rows of x[r][c]; block = 8 column groups (8 bf16 each) x 32 row phases.
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const bf16* __restrict__ x, int ld, int rows, int cols, int rows_per_strip, float* __restrict__ partials) {
  pdl_wait();
  pdl_trigger();
  __shared__ float red[32][65];
  const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
  const int col = blockIdx.x * 64 + tx * 8;
  const int r0 = blockIdx.y * rows_per_strip, r1 = min(rows, r0 + rows_per_strip);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (col < cols) {
    if (col + 8 <= cols) {
#pragma unroll 4
      for (int r = r0 + ty; r < r1; r += 32) {
        float f[8];
        unpack8(*reinterpret_cast<const uint4*>(x + size_t(r) * ld + col), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    } else {
      for (int r = r0 + ty; r < r1; r += 32)
        for (int j = 0; j < 8 && col + j < cols; ++j) acc[j] += __bfloat162float(x[size_t(r) * ld + col + j]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[ty][tx * 8 + j] = acc[j];
  __syncthreads();
  if (threadIdx.x < 64 && blockIdx.x * 64 + threadIdx.x < cols) {
    float v = 0.f;
#pragma unroll 8
    for (int i = 0; i < 32; ++i) v += red[i][threadIdx.x];
    partials[size_t(blockIdx.y) * cols + blockIdx.x * 64 + threadIdx.x] = v;
  }
}

__global__ void cast_kernel(const float4* __restrict__ in, uint2* __restrict__ out, size_t n4) {
  pdl_wait();
  pdl_trigger();
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n4; i += size_t(gridDim.x) * blockDim.x) {
    const float4 f = in[i];
    uint2 u; u.x = ptx::pack_bf16x2(f.x, f.y); u.y = ptx::pack_bf16x2(f.z, f.w);
    out[i] = u;
  }
}

__global__ void cast_back_kernel(const bf16* __restrict__ in, int ldi, float* __restrict__ out, int ldo, int rows, int cols) {
  const size_t total = size_t(rows) * cols;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const size_t r = i / cols, c = i % cols;
    out[r * ldo + c] = __bfloat162float(in[r * ldi + c]);
  }
}
}  // namespace

int colsum_parts(int rows, int cols) {
  const int colblocks = (cols + 63) / 64;
  // enough CTAs to fill the machine, and never more than 256 rows per CTA: these sums run on the side stream next to
  // GEMM CTAs that need a whole SM's shared memory, which the block scheduler can only hand over when every small CTA
  // on that SM has retired -- long-lived CTAs here (one strip of 4096 rows under the vocabulary projection) stalled
  // the critical path for their whole duration
  const int want = std::max((2 * num_sms() + colblocks - 1) / colblocks, (rows + 255) / 256);
  return std::max(1, std::min(want, (rows + 127) / 128));
}

int colsum_partial(const bf16* x, int ld, int rows, int cols, float* partials, cudaStream_t stream) {
  if (ld % 8 || (reinterpret_cast<uintptr_t>(x) & 15)) return set_error("colsum: x must be 16-byte aligned with ld %% 8 == 0");
  if (cols <= 0 || rows <= 0) return set_error("colsum: empty problem");
  const int strips = colsum_parts(rows, cols);
  const int rps = (rows + strips - 1) / strips;
  B200XF_CUDA(launch_k(colsum_partial_kernel, dim3((cols + 63) / 64, strips), 256, 0, stream, x, ld, rows, cols, rps, partials));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

int cast_f32_to_bf16(const float* in, bf16* out, size_t n, cudaStream_t stream) {
  if (n % 4) return set_error("cast: n must be a multiple of 4");
  if (n == 0) return 0;
  const size_t n4 = n / 4;
  const int grid = int(std::min<size_t>((n4 + 255) / 256, size_t(num_sms()) * 8));
  B200XF_CUDA(launch_k(cast_kernel, grid, 256, 0, stream, reinterpret_cast<const float4*>(in), reinterpret_cast<uint2*>(out), n4));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

namespace {
// tf.argmax over the last axis (tester.py:38, transformer_model.py:21): lowest index among equal maxima.
__global__ void __launch_bounds__(256)
argmax_rows_kernel(const float* __restrict__ x, int ld, int cols, long long* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sv[8]; __shared__ int si[8];
  const float* row = x + size_t(blockIdx.x) * ld;
  float bv = -INFINITY; int bi = 0x7fffffff;
  for (int j = threadIdx.x; j < cols; j += blockDim.x) { const float v = row[j]; if (v > bv) { bv = v; bi = j; } }
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
    out[blockIdx.x] = bi;
  }
}
}  // namespace

namespace {
// Ragged -> dense: the `.to_tensor()` at the end of the reference's tokenisation (english_portugese_dataset.py:36-39)
// pads every sentence of a batch with 0 (= PAD) to a common length.  flat holds the batch's sentences back to back,
// offsets[r] .. offsets[r+1] delimit sentence r; one warp per output row, 16-byte stores when the row pitch allows.
__global__ void __launch_bounds__(256)
assemble_padded_kernel(const long long* __restrict__ flat, const long long* __restrict__ offsets, int first_row, int rows,
                       int L, long long* __restrict__ out, int* __restrict__ too_long) {
  pdl_wait();
  pdl_trigger();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const long long b0 = offsets[first_row + warp], b1 = offsets[first_row + warp + 1];
  const long long len = b1 - b0;
  if (len > L && lane == 0) atomicMax(too_long, int(len));  // reported to the host: a sentence does not fit the row
  long long* o = out + size_t(warp) * L;
  for (int j = lane; j < L; j += 32) o[j] = (j < len) ? flat[b0 + j] : 0ll;
}
}  // namespace

int assemble_padded(const long long* flat, const long long* offsets, int first_row, int rows, int L, long long* out,
                    int* too_long_dev, cudaStream_t stream) {
  if (rows <= 0 || L <= 0) return set_error("assemble_padded: empty batch (rows=%d L=%d)", rows, L);
  if (first_row < 0) return set_error("assemble_padded: negative first row");
  B200XF_CUDA(launch_k(assemble_padded_kernel, (rows * 32 + 255) / 256, 256, 0, stream, flat, offsets, first_row, rows, L, out, too_long_dev));
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

int argmax_rows(const float* x, int ld, int rows, int cols, long long* out, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return set_error("argmax: empty problem");
  B200XF_CUDA(launch_k(argmax_rows_kernel, rows, 256, 0, stream, x, ld, cols, out));
  count_launch();
  return 0;
}

int cast_bf16_to_f32_2d(const bf16* in, int ldi, float* out, int ldo, int rows, int cols, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return 0;
  const size_t total = size_t(rows) * cols;
  const int grid = int(std::min<size_t>((total + 255) / 256, size_t(num_sms()) * 8));
  cast_back_kernel<<<grid, 256, 0, stream>>>(in, ldi, out, ldo, rows, cols);
  B200XF_CUDA(cudaGetLastError());
  count_launch();
  return 0;
}

}  // namespace b200xf

// ============================================================================ LayerNorm backward without caller scratch
// b200xf_layernorm_residual_bwd needs a scratch buffer for its per-block partial column sums, which normally means the
// caller allocates device memory.  Some callers cannot: an op hosted by a framework that has already handed the whole
// GPU to its own memory pool (TensorFlow's default allocator does this, and a C library underneath it cannot reach that
// pool), or a steady-state loop that must never run into cudaMalloc's device-wide synchronisation.  For them the
// variant below keeps the partials in a fixed array that is part of the loaded library image.
namespace b200xf {
namespace {
// Capacity: 909 312 floats (3.47 MiB) of per-block partial sums — one [3][d] partial per 16 rows (d <= 512) or per
// 8 rows (larger d), i.e. up to 9 472 rows at d = 512 or 2 368 rows at d = 1024.
// Every GPU that loads the library carries this array until the process exits.
constexpr size_t kStaticLnScratchFloats = size_t(296) * 3 * 1024;
static __device__ __align__(256) float g_ln_bwd_static_scratch[kStaticLnScratchFloats];

// In host code `g_ln_bwd_static_scratch` is only a handle for the symbol, not a device address, so the address the
// kernels need is looked up once per device (with that device current) and remembered.
int static_ln_scratch(float** out) {
  static std::mutex mu;
  static float* per_device[64] = {nullptr};
  int dev = 0;
  B200XF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return set_error("device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!per_device[dev]) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_ln_bwd_static_scratch);
    if (e != cudaSuccess || !p) return set_error("cudaGetSymbolAddress(static LayerNorm scratch): %s", cudaGetErrorString(e));
    per_device[dev] = p;
  }
  *out = per_device[dev];
  return 0;
}
}  // namespace
}  // namespace b200xf

// Same results as b200xf_layernorm_residual_bwd (same kernels, same summation order), no scratch argument and no
// allocation of any kind: no cudaMalloc / cudaFree, no allocating library call.
//  * Limit: b200xf_layernorm_bwd_scratch_bytes(rows, d) must not exceed 3 637 248 bytes (909 312 floats): up to 9 472 rows
//    at d = 512, 2 368 rows at d = 1024.  Larger problems are refused with an error, nothing is launched.
//  * The first call on each device looks the array's address up (cudaGetSymbolAddress) with that device current; make that
//    first call outside any stream capture.  `stream` must belong to the current device.
//  * All calls on a device share the one array and it keeps whatever the previous call left in it.  Nothing in it has to be
//    zero beforehand: each call first overwrites every partial slot it will later read.  For the same reason two calls that
//    overlap on the GPU (different streams, or a second thread) race on it — serialise them.
extern "C" int b200xf_layernorm_residual_bwd_noalloc(const void* dy, const void* z, const float* mean, const float* rstd,
                                                     const float* gamma, void* dz, float* dgamma, float* dbeta, float* dcol,
                                                     int rows, int d, void* stream) {
  using namespace b200xf;
  if (!dy || !z || !mean || !rstd || !gamma || !dz || !dgamma || !dbeta) return set_error("layernorm_bwd_noalloc: null pointer");
  if (rows <= 0 || d <= 0 || d % 8 || d > 2048) return set_error("layernorm_bwd_noalloc: rows=%d d=%d unsupported (d multiple of 8, <= 2048)", rows, d);
  const size_t need = ln_bwd_scratch_floats(rows, d);
  if (need > kStaticLnScratchFloats)
    return set_error("layernorm_bwd_noalloc: rows=%d d=%d needs %zu bytes of scratch, the built-in buffer holds %zu; use "
                     "b200xf_layernorm_residual_bwd with caller scratch", rows, d, need * sizeof(float), kStaticLnScratchFloats * sizeof(float));
  float* scratch = nullptr;
  B200XF_TRY(static_ln_scratch(&scratch));
  return layernorm_bwd(static_cast<const bf16*>(dy), static_cast<const bf16*>(z), mean, rstd, gamma, static_cast<bf16*>(dz), dgamma,
                       dbeta, dcol, rows, d, scratch, static_cast<cudaStream_t>(stream));
}
