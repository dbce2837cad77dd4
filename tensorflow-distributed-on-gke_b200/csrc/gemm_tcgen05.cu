// Persistent warp-specialised bf16 GEMM for sm_100a: TMA -> shared memory (128B swizzle) ->
// tcgen05.mma (accumulators in TMEM, double buffered) -> tcgen05.ld epilogue.
//
// Replaces every tf.keras.layers.Dense matmul on the reference's hot path
// (transformer_model.py:119-122,136-143,165 Q/K/V/O; :169-175 FFN; :333,347 vocab projection)
// and their tape-derived backward products (__main__.py:112): forward  Y = X W + b,
// dgrad dX = dY W^T, wgrad dW = X^T dY.  One kernel covers the three operand-layout cases through
// the UMMA "major" bits, so no activation or weight is ever transposed in HBM:
//   fwd   : A K-major  (X  [M,K]),  B MN-major (W  [K,N] as stored, [in,out])
//   dgrad : A K-major  (dY [M,N]),  B K-major  (W  [K,N] read as [out=K][contract=N])
//   wgrad : A MN-major (X  [tok,K] read as [contract=tok][out=K]), B MN-major (dY [tok,N])
//
// D[m][n] = sum_k A[m][k] * B[n][k];  fp32 accumulate; epilogue: +bias[n], ReLU, +residual or
// ReLU-mask, output bf16 or fp32 (optionally atomically accumulated for split-K).
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <map>
#include <mutex>
#include <tuple>

#include "internal.h"
#include "ptx.cuh"

namespace b200xf {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;       // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int kThreads = 320;  // warp0 TMA, warp1 MMA, warps 2-9 epilogue (two per TMEM lane quarter, half the columns each)
// Weight-gradient layout (A and B both MN-major): two more warps (10, 11) that sum the columns of the B tiles in shared memory
// while the tensor core consumes them -- the bias gradient of the same Dense, for free (GemmArgs::colsum_partials).
constexpr int kThreadsWgrad = 384;

// PAIR: two CTAs of a cluster (the two SMs of a TPC) work on one 256 x BN tile with cta_group::2 MMAs.  Each CTA stages
// its own 128 rows of A and only HALF of the B tile (BN/2 columns; the tensor core reads the other half from the peer),
// so a k-block costs each SM 16 KB + BN*64 B of L2 -> shared-memory traffic instead of 16 KB + BN*128 B: the K=512 GEMMs
// of the model, which run at the per-SM L2 bandwidth, get a third of it back.
template <int BN, bool PAIR, bool LNF = false>
struct Cfg {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = (PAIR ? BN / 2 : BN) * BK * 2;  // per CTA
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagesFull = (192 * 1024 / kStageBytes) > 8 ? 8 : (192 * 1024 / kStageBytes);  // 48 KB: 4, 32 KB: 6, 24 KB: 8
  // the fused-LayerNorm variant keeps gamma / beta slices and the per-row statistics in shared memory as well
  static constexpr int kStages = !LNF ? kStagesFull : (BN == 256 ? 3 : (kStagesFull > 5 ? 5 : kStagesFull));
  static constexpr int kLnBytes = LNF ? (2 * 256 * 4 + 8 * 2 * BM * 8) : 0;  // gamma, beta slices; {local mean, M2} per source CTA, column half and row
  static constexpr int kSmem = kStages * kStageBytes + 4 * 2 * 4096 /*store staging*/ + 1024 /*bias slice*/ + kLnBytes + 1024 /*align slack*/ + 256 /*barriers*/;
  static constexpr int kTmemCols = LNF ? BN : 2 * BN;  // double-buffered accumulator (one tile per CTA with the fused LayerNorm)
};

struct GemmParams {
  int M, N, K;
  int tiles_m, tiles_n, splits, kb_per_split;
  void* d; int ldd; int d_fp32; int accumulate; int d_vec; int d_tma;  // d_vec: rows are 16-byte aligned -> vector stores
  const float* bias; int relu;
  const __nv_bfloat16* resid; int ldr; int resid_op;  // 0 none, 1 add, 2 relu-mask
  float alpha;
  unsigned long long* trace;  // debug: 8 globaltimer stamps per CTA (null in production)
  DropArgs drop;
  int b_static;  // B is not written by the kernel launched just before this one: its first tiles may be fetched before griddepcontrol.wait
  float* colsum;  // weight-gradient layout: [splits][N] partial column sums of B (null: off)
  // fused LayerNorm (LNF instantiations)
  const float* ln_gamma; const float* ln_beta; float* ln_mean; float* ln_rstd; float ln_eps;
};

// bias add, ReLU, residual add / ReLU mask on 32 accumulator columns [col0, col0+32) of one output row.
// `rpre`: the four 16-byte residual vectors of these 32 columns, already loaded (or null: load them here).
__device__ __forceinline__ void epilogue_math(float* f, const GemmParams& p, int row, bool row_ok, int col0, const float* sbias,
                                              const uint4* rpre = nullptr) {
  if (sbias) {  // this tile's bias slice, staged in shared memory once per tile (zero past N): broadcast reads
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 b = *reinterpret_cast<const float4*>(sbias + j);
      f[j] += b.x; f[j + 1] += b.y; f[j + 2] += b.z; f[j + 3] += b.w;
    }
  }
  if (p.relu) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
  }
  if (p.drop.thresh) {
    const unsigned long long stream = drop_stream(p.drop);
    const unsigned long long base = static_cast<unsigned long long>(row) * p.N + col0;
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = drop_keep(stream, base + j, p.drop.thresh) ? f[j] * p.drop.scale : 0.f;
  }
  if (p.resid_op && row_ok) {
    const __nv_bfloat16* rp = p.resid + size_t(row) * p.ldr + col0;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float r[8];
      if (rpre != nullptr || col0 + g * 8 + 8 <= p.N) {
        const uint4 u = rpre != nullptr ? rpre[g] : *reinterpret_cast<const uint4*>(rp + g * 8);
        r[0] = ptx::bf16_lo(u.x); r[1] = ptx::bf16_hi(u.x); r[2] = ptx::bf16_lo(u.y); r[3] = ptx::bf16_hi(u.y);
        r[4] = ptx::bf16_lo(u.z); r[5] = ptx::bf16_hi(u.z); r[6] = ptx::bf16_lo(u.w); r[7] = ptx::bf16_hi(u.w);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = (col0 + g * 8 + j < p.N) ? __bfloat162float(rp[g * 8 + j]) : 0.f;
      }
      if (p.resid_op == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[g * 8 + j] += r[j];
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[g * 8 + j] = r[j] > 0.f ? f[g * 8 + j] : 0.f;
      }
    }
  }
}

__device__ __forceinline__ void stamp(const GemmParams& p, int slot) {
  if (p.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[size_t(blockIdx.x) * 8 + slot] = t;
  }
}

// Store into the shared memory of CTA `cta_rank` of the cluster, at the offset the pointer has in this CTA's.
__device__ __forceinline__ void st_cluster_f2(void* local_smem_ptr, uint32_t cta_rank, float2 v) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(ptx::smem_u32(local_smem_ptr)), "r"(cta_rank));
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(remote), "f"(v.x), "f"(v.y) : "memory");
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

template <int BN, int A_MN, int B_MN, bool PAIR, bool LNF, bool CS>
__global__ void __launch_bounds__(CS ? kThreadsWgrad : kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
            const __grid_constant__ CUtensorMap tmap_d, const __grid_constant__ CUtensorMap tmap_y, const GemmParams p) {
  using C = Cfg<BN, PAIR, LNF>;
  static_assert(!(LNF && PAIR), "the fused LayerNorm uses its cluster for the row statistics");
  static_assert(!CS || (A_MN && B_MN), "column sums of B: weight-gradient layout only");
  constexpr bool WGRAD = CS;   // the instantiations with the two column-sum warps
  constexpr int BNC = PAIR ? BN / 2 : BN;  // B columns this CTA stages
  const int rank = PAIR ? int(ptx::cluster_ctarank()) : 0;   // 0 = the pair's leader: arms the barriers' byte counts, issues the MMAs
  const int worker = PAIR ? int(blockIdx.x >> 1) : int(blockIdx.x), workers = PAIR ? int(gridDim.x >> 1) : int(gridDim.x);
  extern __shared__ uint8_t smem_raw[];
  // round up to 1024 B with pointer arithmetic (not an integer round trip) so the compiler keeps treating every
  // access below as shared-memory (STS/LDS), not generic
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* stage_base = smem + C::kStages * C::kStageBytes;  // 8 epilogue warps x 4 KB store staging
  float* sbias = reinterpret_cast<float*>(stage_base + 8 * 4096);  // [256] bias slice of the current tile
  float* sgamma = sbias + 256;                                      // LNF: [256] gamma, [256] beta, [2][BM] {mean, M2}
  float* sbeta = sgamma + 256;
  float2* sstat = reinterpret_cast<float2*>(sbeta + 256);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sbias + 256) + C::kLnBytes);
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* tmem_full = empty_bar + C::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  // CTA pairs: the loads of BOTH CTAs are credited to the leader's full barrier, so the second CTA has no local signal that a
  // stage has landed; when its column-sum warps need one, the leader's MMA thread forwards it to the peer's full2 barrier.
  uint64_t* full2_bar = tmem_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(full2_bar + C::kStages);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const bool colsum_on = WGRAD && p.colsum != nullptr;
  if (threadIdx.x == 0) {
    stamp(p, 0);
    if (p.trace && blockIdx.x == 0)  // slot 7 of CTA 0 identifies the launch: M, N, K and the operand majors
      p.trace[7] = (static_cast<unsigned long long>(p.M) << 44) | (static_cast<unsigned long long>(p.N) << 24) |
                   (static_cast<unsigned long long>(p.K) << 4) | (A_MN << 1) | B_MN;
  }

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmap_a);
    ptx::prefetch_tmap(&tmap_b);
    if (p.d_tma) ptx::prefetch_tmap(&tmap_d);
    if (LNF) ptx::prefetch_tmap(&tmap_y);
    // a stage is free again once its MMAs have retired and, when the column sums are on, both summing warps have read it
    for (int i = 0; i < C::kStages; ++i) {
      ptx::mbar_init(&full_bar[i], 1); ptx::mbar_init(&empty_bar[i], colsum_on ? 3 : 1);
      if (CS && PAIR) ptx::mbar_init(&full2_bar[i], 1);
    }
    // tmem_empty collects the epilogue warps of BOTH CTAs of a pair (the leader's MMA thread is the only waiter)
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(&tmem_full[i], 1); ptx::mbar_init(&tmem_empty[i], PAIR ? 16 : 8); }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    if (PAIR) { ptx::tmem_alloc_pair(tmem_slot, C::kTmemCols); ptx::tmem_relinquish_pair(); }
    else { ptx::tmem_alloc(tmem_slot, C::kTmemCols); ptx::tmem_relinquish(); }
  }
  ptx::tc_fence_before();
  if (PAIR) ptx::cluster_sync(); else __syncthreads();  // the peer's barriers exist before anything signals them
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int total_tiles = p.tiles_m * p.tiles_n * p.splits;
  const int total_kb = (p.K + BK - 1) / BK;
  // Tile order.  Default: consecutive CTAs walk down the rows of one column tile (they share the B tile in L2).  With the
  // fused LayerNorm the CTAs of a cluster hold the column tiles of ONE row block: tn = rank in the cluster.
  auto tile_tm = [&](int t) { return LNF ? t / p.tiles_n : t % p.tiles_m; };
  auto tile_tn = [&](int t) { return LNF ? t % p.tiles_n : (t / p.tiles_m) % p.tiles_n; };
  // This CTA's share of a k-block: its 128 rows of A, and its BNC columns of B (all of them, or its half of a pair's).
  auto load_a = [&](uint8_t* sa, uint64_t* bar, int tm, int kb) {
    const int row0 = (PAIR ? tm * 2 + rank : tm) * BM;
    if (A_MN) {
#pragma unroll
      for (int c = 0; c < BM / 64; ++c) {
        if (PAIR) ptx::tma_load_2d_pair(sa + c * 8192, &tmap_a, bar, row0 + c * 64, kb * BK);
        else ptx::tma_load_2d(sa + c * 8192, &tmap_a, bar, row0 + c * 64, kb * BK);
      }
    } else {
      if (PAIR) ptx::tma_load_2d_pair(sa, &tmap_a, bar, kb * BK, row0);
      else ptx::tma_load_2d(sa, &tmap_a, bar, kb * BK, row0);
    }
  };
  auto load_b = [&](uint8_t* sb, uint64_t* bar, int tn, int kb) {
    const int col0 = tn * BN + rank * BNC;
    if (B_MN) {
#pragma unroll
      for (int c = 0; c < BNC / 64; ++c) {
        if (PAIR) ptx::tma_load_2d_pair(sb + c * 8192, &tmap_b, bar, col0 + c * 64, kb * BK);
        else ptx::tma_load_2d(sb + c * 8192, &tmap_b, bar, col0 + c * 64, kb * BK);
      }
    } else {
      if (PAIR) ptx::tma_load_2d_pair(sb, &tmap_b, bar, kb * BK, col0);
      else ptx::tma_load_2d(sb, &tmap_b, bar, kb * BK, col0);
    }
  };
  // Weights (the B operand of the forward and data-gradient GEMMs) do not depend on the previous kernel: start
  // fetching the first tile's B blocks while that kernel is still finishing, and only the A blocks after the wait.
  int prefetched = 0;
  if (p.b_static && threadIdx.x == 0 && worker < total_tiles) {
    const int t = worker;
    const int tn = tile_tn(t), z = t / (p.tiles_m * p.tiles_n);
    const int kb0 = z * p.kb_per_split;
    const int kb1 = min(total_kb, kb0 + p.kb_per_split);
    prefetched = min(C::kStages, kb1 - kb0);
    for (int s = 0; s < prefetched; ++s) {
      uint8_t* sb = smem + s * C::kStageBytes + C::kABytes;
      if (rank == 0) ptx::mbar_expect_tx(&full_bar[s], (PAIR ? 2 : 1) * C::kStageBytes);
      load_b(sb, &full_bar[s], tn, kb0 + s);
    }
  }
  pdl_wait();     // barrier init, TMEM allocation and descriptor prefetch above overlap the previous kernel's tail
  pdl_trigger();
  if (threadIdx.x == 0) stamp(p, 1);

  // LNF: what the epilogue threads carry from their first pass (z, packed bf16) into the second (y = LayerNorm(z))
  constexpr int kChunksPerHalf = (BN / 32) / 2;
  // tiles of >= 128 columns: the two warps of a lane quarter take half of the columns each; narrower: the first takes all
  constexpr bool kLnSplit = kChunksPerHalf >= 2;
  constexpr int kLnChunks = LNF ? (kLnSplit ? kChunksPerHalf : BN / 32) : 1;
  uint32_t zp[kLnChunks][16];
  (void)zp;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (one lane)
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (int t = worker; t < total_tiles; t += workers) {
        const int tm = tile_tm(t), tn = tile_tn(t), z = t / (p.tiles_m * p.tiles_n);
        const int kb0 = z * p.kb_per_split;
        const int kb1 = min(total_kb, kb0 + p.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          const bool have_b = prefetched > 0;  // this stage was armed and its B block requested before the wait
          if (have_b) --prefetched;
          if (!have_b) ptx::mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sa = smem + s * C::kStageBytes;
          uint8_t* sb = sa + C::kABytes;
          // the leader's barrier counts the bytes of both CTAs of a pair
          if (!have_b && rank == 0) ptx::mbar_expect_tx(&full_bar[s], (PAIR ? 2 : 1) * C::kStageBytes);
          load_a(sa, &full_bar[s], tm, kb);
          if (!have_b) load_b(sb, &full_bar[s], tn, kb);
          if (++s == C::kStages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one lane)
    if (lane == 0 && rank == 0) {  // the leader issues for the pair
      constexpr uint32_t idesc = ptx::make_idesc_bf16(PAIR ? 2 * BM : BM, BN, A_MN, B_MN);
      constexpr uint32_t a_lbo = A_MN ? 8192 : 16, b_lbo = B_MN ? 8192 : 16;
      constexpr uint32_t a_kstep = A_MN ? (UMMA_K * 128) : (UMMA_K * 2);  // bytes per 16-wide k step
      constexpr uint32_t b_kstep = B_MN ? (UMMA_K * 128) : (UMMA_K * 2);
      int s = 0; uint32_t ph = 0; int it = 0;
      for (int t = worker; t < total_tiles; t += workers, ++it) {
        const int z = t / (p.tiles_m * p.tiles_n);
        const int kb0 = z * p.kb_per_split;
        const int kb1 = min(total_kb, kb0 + p.kb_per_split);
        const int as = LNF ? 0 : (it & 1); const uint32_t aph = LNF ? 0 : ((it >> 1) & 1);
        ptx::mbar_wait(&tmem_empty[as], aph ^ 1);
        ptx::tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          ptx::mbar_wait(&full_bar[s], ph);
          ptx::tc_fence_after();
          if (PAIR && WGRAD && colsum_on) ptx::mbar_arrive_peer_release(&full2_bar[s], 1);
          if (it == 0 && kb == kb0) stamp(p, 2);
          const uint32_t sa = ptx::smem_u32(smem + s * C::kStageBytes);
          const uint32_t sb = sa + C::kABytes;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t adesc = ptx::make_smem_desc(sa + k * a_kstep, a_lbo, 1024);
            const uint64_t bdesc = ptx::make_smem_desc(sb + k * b_kstep, b_lbo, 1024);
            if (PAIR) ptx::umma_bf16_pair(tmem_d, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            else ptx::umma_bf16(tmem_d, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          // frees the smem stage (in both CTAs of a pair) once these MMAs retire
          if (PAIR) ptx::umma_commit_pair(&empty_bar[s]); else ptx::umma_commit(&empty_bar[s]);
          if (++s == C::kStages) { s = 0; ph ^= 1; }
        }
        // accumulator complete -> epilogue (of both CTAs)
        if (PAIR) ptx::umma_commit_pair(&tmem_full[as]); else ptx::umma_commit(&tmem_full[as]);
        if (it == 0) stamp(p, 3);
      }
    }
  } else if (WGRAD && warp >= 10) {
    // ------------------------------------------------------------ column sums of B (the bias gradient), two warps
    // B tile of a stage: BNC / 64 chunks of [64 k rows][64 columns] bf16, rows 128 B apart, 16-byte units XOR-swizzled by
    // (k & 7).  A warp owns every second chunk; lane l owns columns 2l, 2l+1 of it: 32 consecutive words of a row per
    // access, conflict-free.
    if (colsum_on) {
      constexpr int kMyChunks = (BNC / 64 + 1) / 2;
      const int w = warp - 10;
      int s = 0; uint32_t ph = 0;
      for (int t = worker; t < total_tiles; t += workers) {
        const int tm = tile_tm(t), tn = tile_tn(t), z = t / (p.tiles_m * p.tiles_n);
        const int kb0 = z * p.kb_per_split;
        const int kb1 = min(total_kb, kb0 + p.kb_per_split);
        // The CTAs of the tiles_m row blocks of one column tile all stream the same B; each sums every tiles_m-th k-block, so
        // the extra shared-memory reads (which compete with the tensor core's operand reads) are spread evenly.
        float acc[kMyChunks][2];
#pragma unroll
        for (int c = 0; c < kMyChunks; ++c) acc[c][0] = acc[c][1] = 0.f;
        for (int kb = kb0; kb < kb1; ++kb) {
          if (PAIR && rank == 1) ptx::mbar_wait_cluster(&full2_bar[s], ph); else ptx::mbar_wait(&full_bar[s], ph);
          if ((kb - kb0) % p.tiles_m == tm) {
            const uint8_t* sb = smem + s * C::kStageBytes + C::kABytes;
#pragma unroll
            for (int c = 0; c < kMyChunks; ++c) {
              const int chunk = c * 2 + w;
              if (chunk < BNC / 64) {
                const uint8_t* base = sb + chunk * 8192 + ((lane & 3) << 2);
#pragma unroll 16
                for (int k = 0; k < 64; ++k) {
                  const uint32_t v = *reinterpret_cast<const uint32_t*>(base + k * 128 + ((((lane >> 2) ^ (k & 7))) << 4));
                  acc[c][0] += ptx::bf16_lo(v); acc[c][1] += ptx::bf16_hi(v);
                }
              }
            }
          }
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&empty_bar[s]);
          if (++s == C::kStages) { s = 0; ph ^= 1; }
        }
#pragma unroll
        for (int c = 0; c < kMyChunks; ++c) {
          const int chunk = c * 2 + w;
          const int col = tn * BN + rank * BNC + chunk * 64 + 2 * lane;
          if (chunk < BNC / 64 && col < p.N) {
            float* dst = p.colsum + (size_t(z) * p.tiles_m + tm) * p.N + col;   // one partial row per (K split, row block)
            dst[0] = acc[c][0];
            if (col + 1 < p.N) dst[1] = acc[c][1];
          }
        }
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue (8 warps, 128 TMEM lanes)
    const int quarter = warp & 3;  // TMEM lane quarter this warp may touch
    constexpr int kBiasPerThread = 1;  // 256 epilogue threads cover BN <= 256 columns
    // this warp's share of the tile's 32-column chunks: the two warps of a lane quarter split the columns
    const int halfsel = (warp - 2) >> 2;
    float bias_pre[kBiasPerThread];
#pragma unroll
    for (int k = 0; k < kBiasPerThread; ++k) {  // first tile's bias slice
      const int j = int(threadIdx.x) - 64 + k * 256;
      const int tn0 = tile_tn(worker);
      bias_pre[k] = (p.bias != nullptr && worker < total_tiles && j < BN && tn0 * BN + j < p.N) ? __ldg(p.bias + tn0 * BN + j) : 0.f;
      if (LNF) {  // gamma / beta slices of this CTA's columns
        const bool ok = j < BN && tn0 * BN + j < p.N;
        if (j < 256) { sgamma[j] = ok ? __ldg(p.ln_gamma + tn0 * BN + j) : 0.f; sbeta[j] = ok ? __ldg(p.ln_beta + tn0 * BN + j) : 0.f; }
      }
    }
    int it = 0;
    for (int t = worker; t < total_tiles; t += workers, ++it) {
      const int tn = tile_tn(t), z = t / (p.tiles_m * p.tiles_n);
      const int tm = PAIR ? tile_tm(t) * 2 + rank : tile_tm(t);  // this CTA's 128-row block
      const int as = LNF ? 0 : (it & 1); const uint32_t aph = LNF ? 0 : ((it >> 1) & 1);
      const int row = tm * BM + quarter * 32 + lane;
      const bool row_ok = row < p.M;
      // the residual of this thread's first 64 columns is requested before the accumulator is ready, so that its L2
      // latency (0.6-0.8 us) runs under the main loop instead of inside the epilogue
      uint4 rpre[8];
      bool rpre_ok = false;
      if (p.resid_op && p.d_tma && !p.d_fp32) {
        const int c_first = (kChunksPerHalf >= 2) ? halfsel * kChunksPerHalf : (halfsel ? BN / 32 : 0);
        const int colfirst = tn * BN + c_first * 32;
        if (row_ok && c_first < BN / 32 && colfirst + 64 <= p.N) {
          const uint4* rp = reinterpret_cast<const uint4*>(p.resid + size_t(row) * p.ldr + colfirst);
#pragma unroll
          for (int g = 0; g < 8; ++g) rpre[g] = rp[g];
          rpre_ok = true;
        }
      }
      ptx::mbar_wait(&tmem_full[as], aph);
      ptx::tc_fence_after();
      if (it == 0 && warp == 2 && lane == 0) stamp(p, 4);
      const uint32_t taddr = tmem_base + as * BN + (uint32_t(quarter * 32) << 16);
      const bool add_bias = p.bias != nullptr && z == 0;
      if (add_bias) {  // CTA-uniform: the four epilogue warps stage this tile's BN bias values once
        asm volatile("bar.sync 1, 256;" ::: "memory");   // everyone is done with the previous tile's slice
#pragma unroll
        for (int k = 0; k < kBiasPerThread; ++k) { const int j = int(threadIdx.x) - 64 + k * 256; if (j < BN) sbias[j] = bias_pre[k]; }
        asm volatile("bar.sync 1, 256;" ::: "memory");
      }
      if (p.bias != nullptr && t + workers < total_tiles) {  // fetch the next tile's slice now; it is needed a whole tile later
        const int tn_next = tile_tn(t + workers);
#pragma unroll
        for (int k = 0; k < kBiasPerThread; ++k) {
          const int j = int(threadIdx.x) - 64 + k * 256;
          bias_pre[k] = (j < BN && tn_next * BN + j < p.N) ? __ldg(p.bias + tn_next * BN + j) : 0.f;
        }
      }
      if constexpr (LNF) {
        // ---- fused LayerNorm, pass 1: z = epilogue(acc), rounded to bf16 (what the backward pass and the statistics see),
        // stored through the staging tile, kept packed in registers; per-thread {mean, M2} of its columns.
        uint8_t* buf = stage_base + (warp - 2) * 4096;
        const int c_begin = kLnSplit ? halfsel * kChunksPerHalf : 0;
        const bool active = kLnSplit || halfsel == 0;
        float lsum = 0.f, lsq = 0.f, pivot = 0.f; int cnt = 0;
        const bool full_tile = (tn + 1) * BN <= p.N;   // CTA-uniform: no column of this tile lies past N
        if (active) {
#pragma unroll
          for (int cc = 0; cc < kLnChunks; ++cc) {
            const int col0 = tn * BN + (c_begin + cc) * 32;
            uint32_t v[32];
            ptx::tmem_ld_32x32(taddr + (c_begin + cc) * 32, v);
            ptx::tmem_ld_wait();
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * p.alpha;
            if (col0 < p.N) epilogue_math(f, p, row, row_ok, col0, add_bias ? sbias + (c_begin + cc) * 32 : nullptr,
                                          (rpre_ok && cc < 2) ? rpre + cc * 4 : nullptr);
            if (full_tile) {
              // one pass: sums of (x - pivot) and (x - pivot)^2 with the thread's first value as pivot (no cancellation beyond
              // the spread of the row segment itself)
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const uint32_t w2 = ptx::pack_bf16x2(f[2 * j], f[2 * j + 1]);
                zp[cc][j] = w2;
                if (cc == 0 && j == 0) pivot = ptx::bf16_lo(w2);
                const float a = ptx::bf16_lo(w2) - pivot, b = ptx::bf16_hi(w2) - pivot;
                lsum += a + b;
                lsq = fmaf(a, a, lsq); lsq = fmaf(b, b, lsq);
              }
              cnt += 32;
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const bool ok0 = col0 + 2 * j < p.N, ok1 = col0 + 2 * j + 1 < p.N;
                const uint32_t w2 = ptx::pack_bf16x2(ok0 ? f[2 * j] : 0.f, ok1 ? f[2 * j + 1] : 0.f);
                zp[cc][j] = w2;
                lsum += ptx::bf16_lo(w2) + ptx::bf16_hi(w2);
                cnt += int(ok0) + int(ok1);
              }
            }
          }
        }
        float lmean, m2 = 0.f;
        if (full_tile) {
          const float dm = cnt > 0 ? lsum / float(cnt) : 0.f;   // mean of (x - pivot)
          lmean = pivot + dm;
          m2 = fmaxf(lsq - lsum * dm, 0.f);                     // sum (x - mean)^2 = sum d^2 - (sum d)^2 / n
        } else {
          lmean = cnt > 0 ? lsum / float(cnt) : 0.f;
          if (active) {
#pragma unroll
            for (int cc = 0; cc < kLnChunks; ++cc) {
              const int col0 = tn * BN + (c_begin + cc) * 32;
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const float a = ptx::bf16_lo(zp[cc][j]) - lmean, b = ptx::bf16_hi(zp[cc][j]) - lmean;
                if (col0 + 2 * j < p.N) m2 += a * a;
                if (col0 + 2 * j + 1 < p.N) m2 += b * b;
              }
            }
          }
        }
        // every CTA of the cluster receives this thread's {mean, M2} in the slot of the sending CTA: after the one cluster
        // barrier below each CTA reads only its own shared memory, and nobody touches a peer's again (no exit hazard)
        {
          const uint32_t my_rank = uint32_t(tn);
          float2* slot = &sstat[(my_rank * 2 + halfsel) * BM + quarter * 32 + lane];
          for (int r = 0; r < p.tiles_n; ++r) st_cluster_f2(slot, uint32_t(r), make_float2(lmean, m2));
        }
        cluster_arrive();
        // z out: two 32-column chunks per 128-byte staging row and TMA box
        if (active) {
#pragma unroll 1
          for (int c0 = 0; c0 < kLnChunks; c0 += 2) {
            const int colbase = tn * BN + (c_begin + c0) * 32;
            if (colbase >= p.N) break;
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
              if (c0 + cc < kLnChunks) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                  *reinterpret_cast<uint4*>(buf + lane * 128 + (((cc * 4 + q) ^ (lane & 7)) << 4)) =
                      make_uint4(zp[c0 + cc][q * 4], zp[c0 + cc][q * 4 + 1], zp[c0 + cc][q * 4 + 2], zp[c0 + cc][q * 4 + 3]);
              }
            }
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              ptx::tma_store_2d(&tmap_d, buf, colbase, tm * BM + quarter * 32);
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
          }
        }
        ptx::tc_fence_before();
      } else if (p.d_tma) {
        // TMEM -> registers -> 128B-swizzled staging tile (32 rows x 128 B, one per warp) -> TMA store /
        // TMA reduce-add: coalesced, asynchronous, clipped to the tensor bounds by the hardware.
        const int chunks = p.d_fp32 ? 1 : 2;      // fp32: 32 columns = 128 B per row; bf16 packs two 32-column chunks per box
        int c_begin = halfsel * kChunksPerHalf, c_end = c_begin + kChunksPerHalf;
        if (kChunksPerHalf < chunks) { c_begin = halfsel ? BN / 32 : 0; c_end = BN / 32; }  // too narrow to split: first warp takes all
        uint8_t* buf = stage_base + (warp - 2) * 4096;
#pragma unroll 1
        for (int c0 = c_begin; c0 < c_end; c0 += chunks) {
          const int colbase = tn * BN + c0 * 32;
          if (colbase >= p.N) break;  // warp-uniform
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the previous store has read the buffer
          __syncwarp();
          for (int cc = 0; cc < chunks; ++cc) {
            const int col0 = colbase + cc * 32;
            uint32_t v[32];
            ptx::tmem_ld_32x32(taddr + (c0 + cc) * 32, v);
            ptx::tmem_ld_wait();
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * p.alpha;
            if (col0 < p.N) epilogue_math(f, p, row, row_ok, col0, add_bias ? sbias + (c0 + cc) * 32 : nullptr,
                                          (rpre_ok && c0 == c_begin) ? rpre + cc * 4 : nullptr);
            if (p.d_fp32) {
#pragma unroll
              for (int q = 0; q < 8; ++q)  // 8 x 16-byte chunks per 128-byte row, XOR-swizzled by the row index
                *reinterpret_cast<float4*>(buf + lane * 128 + ((q ^ (lane & 7)) << 4)) = make_float4(f[q * 4], f[q * 4 + 1], f[q * 4 + 2], f[q * 4 + 3]);
            } else {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                uint4 u;
                u.x = ptx::pack_bf16x2(f[q * 8], f[q * 8 + 1]); u.y = ptx::pack_bf16x2(f[q * 8 + 2], f[q * 8 + 3]);
                u.z = ptx::pack_bf16x2(f[q * 8 + 4], f[q * 8 + 5]); u.w = ptx::pack_bf16x2(f[q * 8 + 6], f[q * 8 + 7]);
                *reinterpret_cast<uint4*>(buf + lane * 128 + (((cc * 4 + q) ^ (lane & 7)) << 4)) = u;
              }
            }
          }
          ptx::fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            const int r0 = tm * BM + quarter * 32;
            if (p.accumulate) ptx::tma_reduce_add_2d(&tmap_d, buf, colbase, r0); else ptx::tma_store_2d(&tmap_d, buf, colbase, r0);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
      } else {
#pragma unroll 1
      for (int c = (kChunksPerHalf ? halfsel * kChunksPerHalf : (halfsel ? BN / 32 : 0)); c < (kChunksPerHalf ? (halfsel + 1) * kChunksPerHalf : BN / 32); ++c) {
        const int col0 = tn * BN + c * 32;
        if (col0 >= p.N) break;  // warp-uniform
        uint32_t v[32];
        ptx::tmem_ld_32x32(taddr + c * 32, v);
        ptx::tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * p.alpha;
        epilogue_math(f, p, row, row_ok, col0, add_bias ? sbias + c * 32 : nullptr);
        if (row_ok) {
          if (p.d_fp32) {
            float* dp = reinterpret_cast<float*>(p.d) + size_t(row) * p.ldd + col0;
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              if (p.d_vec && col0 + g * 4 + 4 <= p.N) {
                if (p.accumulate) ptx::red_add_v4(dp + g * 4, f[g * 4], f[g * 4 + 1], f[g * 4 + 2], f[g * 4 + 3]);
                else *reinterpret_cast<float4*>(dp + g * 4) = make_float4(f[g * 4], f[g * 4 + 1], f[g * 4 + 2], f[g * 4 + 3]);
              } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  if (col0 + g * 4 + j < p.N) {
                    if (p.accumulate) atomicAdd(dp + g * 4 + j, f[g * 4 + j]); else dp[g * 4 + j] = f[g * 4 + j];
                  }
              }
            }
          } else {
            __nv_bfloat16* dp = reinterpret_cast<__nv_bfloat16*>(p.d) + size_t(row) * p.ldd + col0;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (p.d_vec && col0 + g * 8 + 8 <= p.N) {
                uint4 u;
                u.x = ptx::pack_bf16x2(f[g * 8], f[g * 8 + 1]); u.y = ptx::pack_bf16x2(f[g * 8 + 2], f[g * 8 + 3]);
                u.z = ptx::pack_bf16x2(f[g * 8 + 4], f[g * 8 + 5]); u.w = ptx::pack_bf16x2(f[g * 8 + 6], f[g * 8 + 7]);
                *reinterpret_cast<uint4*>(dp + g * 8) = u;
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (col0 + g * 8 + j < p.N) dp[g * 8 + j] = __float2bfloat16(f[g * 8 + j]);
              }
            }
          }
        }
      }
      }
      if constexpr (!LNF) {
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (PAIR) ptx::mbar_arrive_leader(&tmem_empty[as]); else ptx::mbar_arrive(&tmem_empty[as]); }
      }
      if (it == 0 && warp == 2 && lane == 0) stamp(p, 5);
    }
    // the staging buffers must outlive the stores' reads of them; the writes themselves complete with the grid
    if (!LNF && p.d_tma && lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }

  if constexpr (LNF) {
    // ---- fused LayerNorm, pass 2.  Every CTA of the cluster has parked its threads' {mean, M2}; combine the 2 x cluster-size
    // parts of each row (Chan's parallel variance: exact two-pass results per part, merged), normalise the register-held
    // z and store y.  The cluster barriers involve every thread of every CTA (the producer / MMA warps just pass through).
    if (warp < 2) cluster_arrive();   // the epilogue warps arrived right after publishing their statistics
    cluster_wait();
    if (threadIdx.x == 64 && blockIdx.x != 0) stamp(p, 7);  // (slot 7 of CTA 0 identifies the launch)
    if (warp >= 2 && warp < 10 && worker < total_tiles) {
      const int quarter = warp & 3, halfsel = (warp - 2) >> 2;
      const int t = worker;
      const int tm = tile_tm(t), tn = tile_tn(t);
      const int row = tm * BM + quarter * 32 + lane;
      const bool active = kLnSplit || halfsel == 0;
      constexpr int kHalfCols = kLnChunks * 32;
      constexpr int kParts = kLnSplit ? 2 : 1;
      const int csize = p.tiles_n;
      float mean = 0.f;
      float2 st[16];
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int h = 0; h < 2; ++h)
          st[r * 2 + h] = (r < csize && h < kParts) ? sstat[(r * 2 + h) * BM + quarter * 32 + lane] : make_float2(0.f, 0.f);
      auto part_cnt = [&](int r, int h) { return max(0, min(kHalfCols, p.N - (r * BN + h * kHalfCols))); };
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int h = 0; h < 2; ++h)
          if (r < csize && h < kParts) mean += st[r * 2 + h].x * float(part_cnt(r, h));
      mean /= float(p.N);
      float m2 = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int h = 0; h < 2; ++h)
          if (r < csize && h < kParts) { const float dlt = st[r * 2 + h].x - mean; m2 += st[r * 2 + h].y + float(part_cnt(r, h)) * dlt * dlt; }
      const float rstd = 1.0f / sqrtf(m2 / float(p.N) + p.ln_eps);  // biased variance, as tf.nn.moments
      if (tn == 0 && halfsel == 0 && row < p.M) { p.ln_mean[row] = mean; p.ln_rstd[row] = rstd; }
      if (active) {
        uint8_t* buf = stage_base + (warp - 2) * 4096;
        const int c_begin = kLnSplit ? halfsel * kChunksPerHalf : 0;
#pragma unroll 1
        for (int c0 = 0; c0 < kLnChunks; c0 += 2) {
          const int colbase = tn * BN + (c_begin + c0) * 32;
          if (colbase >= p.N) break;
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          __syncwarp();
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            if (c0 + cc < kLnChunks) {
              const float* g = sgamma + (c_begin + c0 + cc) * 32;
              const float* b = sbeta + (c_begin + c0 + cc) * 32;
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                uint32_t o[4];
                const float4 g0 = *reinterpret_cast<const float4*>(g + q * 8), g1 = *reinterpret_cast<const float4*>(g + q * 8 + 4);
                const float4 b0 = *reinterpret_cast<const float4*>(b + q * 8), b1 = *reinterpret_cast<const float4*>(b + q * 8 + 4);
                const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const int j = q * 4 + e;
                  const float y0 = fmaf((ptx::bf16_lo(zp[c0 + cc][j]) - mean) * rstd, gg[2 * e], bb[2 * e]);
                  const float y1 = fmaf((ptx::bf16_hi(zp[c0 + cc][j]) - mean) * rstd, gg[2 * e + 1], bb[2 * e + 1]);
                  o[e] = ptx::pack_bf16x2(y0, y1);
                }
                *reinterpret_cast<uint4*>(buf + lane * 128 + (((cc * 4 + q) ^ (lane & 7)) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
              }
            }
          }
          ptx::fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            ptx::tma_store_2d(&tmap_y, buf, colbase, tm * BM + quarter * 32);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
    }
  }

  ptx::tc_fence_before();
  if (PAIR) ptx::cluster_sync(); else __syncthreads();  // neither CTA of a pair leaves while the other may still signal it
  if (warp == 1) {
    ptx::tc_fence_after();
    if (PAIR) ptx::tmem_dealloc_pair(tmem_base, C::kTmemCols); else ptx::tmem_dealloc(tmem_base, C::kTmemCols);
  }
  if (threadIdx.x == 0) stamp(p, 6);
}

// ------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
  });
  return fn;
}

struct MapKey {
  const void* ptr; uint64_t inner, outer, ld; uint32_t box_outer; int kind;  // kind 0: bf16 operand/output (box 64 wide), 1: fp32 output (box 32 wide)
  bool operator<(const MapKey& o) const {
    return std::tie(ptr, inner, outer, ld, box_outer, kind) < std::tie(o.ptr, o.inner, o.outer, o.ld, o.box_outer, o.kind);
  }
};
unsigned long long* g_gemm_trace = nullptr;
int g_gemm_trace_slots = 1;   // ring of per-launch trace slots (148 CTAs x 8 stamps each)
int g_gemm_trace_next = 0;
std::map<MapKey, CUtensorMap> g_map_cache;
std::mutex g_map_mutex;

// 2-D bf16 tensor map: inner (contiguous) extent, outer extent, row pitch ld (elements); box = 64 x box_outer, 128B swizzle.
int get_tmap(const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_outer, CUtensorMap* out, int kind = 0) {
  MapKey key{ptr, inner, outer, ld, box_outer, kind};
  std::lock_guard<std::mutex> lock(g_map_mutex);
  auto it = g_map_cache.find(key);
  if (it != g_map_cache.end()) { *out = it->second; return 0; }
  EncodeTiledFn enc = get_encode();
  if (!enc) return set_error("cuTensorMapEncodeTiled entry point unavailable");
  const uint64_t elt = kind == 1 ? 4 : 2;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || ((ld * elt) % 16)) return set_error("gemm operand must be 16-byte aligned with a 16-byte multiple row pitch");
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * elt};
  cuuint32_t box[2] = {kind == 1 ? 32u : 64u, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, kind == 1 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error("cuTensorMapEncodeTiled failed (%d) inner=%llu outer=%llu ld=%llu", int(r),
                                          (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld);
  if (g_map_cache.size() > 8192) g_map_cache.clear();
  g_map_cache[key] = *out;
  return 0;
}

template <int BN, int A_MN, int B_MN, bool PAIR, bool LNF = false, bool CS = false>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, const CUtensorMap& ty, const GemmParams& p, int grid,
           int cluster, cudaStream_t stream) {
  auto kern = gemm_kernel<BN, A_MN, B_MN, PAIR, LNF, CS>;
  B200XF_TRY(ensure_dynamic_smem(kern, Cfg<BN, PAIR, LNF>::kSmem));
  cudaError_t e = launch_k_cluster(kern, grid, CS ? kThreadsWgrad : kThreads, Cfg<BN, PAIR, LNF>::kSmem, stream, cluster, ta, tb, td, ty, p);
  if (e != cudaSuccess) return set_error("gemm launch: %s", cudaGetErrorString(e));
  count_launch();
  return 0;
}

template <int BN, bool PAIR>
int launch_major(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, const GemmParams& p, int grid, cudaStream_t s) {
  const int cl = PAIR ? 2 : 1;
  if (!a_mn && b_mn) return launch<BN, 0, 1, PAIR>(ta, tb, td, td, p, grid, cl, s);
  if (!a_mn && !b_mn) return launch<BN, 0, 0, PAIR>(ta, tb, td, td, p, grid, cl, s);
  if (a_mn && b_mn) return p.colsum ? launch<BN, 1, 1, PAIR, false, true>(ta, tb, td, td, p, grid, cl, s) : launch<BN, 1, 1, PAIR>(ta, tb, td, td, p, grid, cl, s);
  return launch<BN, 1, 0, PAIR>(ta, tb, td, td, p, grid, cl, s);
}

// fused LayerNorm: forward (B MN-major) and data-gradient (B K-major) layouts, A K-major
template <int BN>
int launch_lnf(int b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, const CUtensorMap& ty, const GemmParams& p, int grid,
               int cluster, cudaStream_t s) {
  if (b_mn) return launch<BN, 0, 1, false, true>(ta, tb, td, ty, p, grid, cluster, s);
  return launch<BN, 0, 0, false, true>(ta, tb, td, ty, p, grid, cluster, s);
}

// tile width of the fused-LayerNorm variant: the row block's column tiles must fit one cluster (<= 8 CTAs)
int lnf_bn(int N) { return N <= 64 ? 64 : ((N + 127) / 128 <= 8 ? 128 : 256); }

std::atomic<int> g_pair_mode{-2};
int pair_mode() {  // 0 never, 1 whenever the problem has at least two 128-row blocks, -1 heuristic (B200XF_GEMM_PAIR sets the default)
  int v = g_pair_mode.load(std::memory_order_relaxed);
  if (v == -2) { const char* e = getenv("B200XF_GEMM_PAIR"); v = e ? atoi(e) : -1; g_pair_mode.store(v, std::memory_order_relaxed); }
  return v;
}

}  // namespace

bool gemm_ln_fusable(int M, int N) { return M > 0 && N > 0 && N % 8 == 0 && (N + lnf_bn(N) - 1) / lnf_bn(N) <= 8; }

static int gemm_impl(const GemmArgs& a, cudaStream_t stream, GemmInfo* info, bool plan_only);
int gemm_bf16(const GemmArgs& a, cudaStream_t stream, GemmInfo* info) { return gemm_impl(a, stream, info, false); }
int gemm_plan(const GemmArgs& a, GemmInfo* info) { return gemm_impl(a, nullptr, info, true); }

static int gemm_impl(const GemmArgs& a, cudaStream_t stream, GemmInfo* info, bool plan_only) {
  if (a.M <= 0 || a.N <= 0 || a.K <= 0) return set_error("gemm: empty problem M=%d N=%d K=%d", a.M, a.N, a.K);
  if (a.accumulate && !a.d_fp32) return set_error("gemm: accumulate requires fp32 output");
  const bool lnf = a.ln_y != nullptr;
  if (lnf) {
    if (a.d_fp32 || a.accumulate || a.a_mn || !a.ln_gamma || !a.ln_beta || !a.ln_mean || !a.ln_rstd)
      return set_error("gemm: the fused LayerNorm needs a bf16, non-accumulating, K-major-A problem with gamma / beta / mean / rstd");
    if (!gemm_ln_fusable(a.M, a.N)) return set_error("gemm: width %d cannot carry the fused LayerNorm", a.N);
  }
  if (a.colsum_partials && !(a.a_mn && a.b_mn)) return set_error("gemm: column sums of B are produced in the weight-gradient layout only");
  const int nsm = (a.max_ctas > 0 && a.max_ctas < num_sms()) ? a.max_ctas : num_sms();  // SMs this launch may fill
  const int tiles_m = (a.M + BM - 1) / BM;
  const int total_kb = (a.K + BK - 1) / BK;
  // Tile width: model each CTA's time as (k-blocks it runs) x (BN + 64 columns of fixed per-k-block cost) plus a per-tile
  // epilogue, over the waves the grid needs; split-K (fp32 atomic accumulate only) can fill idle SMs when tiles are few.
  auto plan = [&](int cand, int* splits_out) {
    const long tiles = long(tiles_m) * ((a.N + cand - 1) / cand);
    int sp = 1;
    if (a.accumulate) {
      sp = a.split_k > 0 ? a.split_k : int(std::max<long>(1, std::min<long>(nsm / tiles, total_kb / 4)));
      sp = std::max(1, std::min(sp, total_kb));
    }
    const int kb_per = (total_kb + sp - 1) / sp;
    sp = (total_kb + kb_per - 1) / kb_per;
    *splits_out = sp;
    const long ctas = tiles * sp;
    const long waves = (ctas + nsm - 1) / nsm;
    return waves * (long(kb_per) * (cand + 64) + 4L * cand);
  };
  int bn = lnf ? lnf_bn(a.N) : a.force_bn, splits = 1;
  if (bn == 0) {
    long best = -1;
    for (int cand : {256, 128, 64}) {
      if (cand > 64 && cand / 2 >= a.N) continue;  // do not waste more than half a tile
      int sp;
      const long cost = plan(cand, &sp);
      if (best < 0 || cost < best) { best = cost; bn = cand; }
    }
  }
  // CTA pairs: 256-row tiles worked by the two SMs of a TPC (needs two 128-row blocks and a B half of >= 64 columns).
  // Measured against the single-CTA kernel (scripts/gemm_bench.py --pairs): 256-wide pair tiles gain 14-18% on the
  // vocabulary and K|V projections and reach 1.43 PFLOP/s at 8192^3; back to back they also gain 3-8% on the one-to-
  // two-wave QKV / FFN1 GEMMs, but inside the step those lose it again (a pair needs both SMs of a TPC free at once, and
  // cluster scheduling plus the pair's extra synchronisation are fixed costs: 4.08 ms/step with them, 4.02 without), and
  // single-wave problems lose 10-20%; 128-wide pair tiles always lose.  Hence: at least three waves of 256-wide tiles.
  const long wide_tiles = long(tiles_m) * ((a.N + 255) / 256);
  const bool pair_auto = !lnf && a.N > 256 && wide_tiles >= 3L * nsm;
  if (!lnf && a.force_bn == 0 && a.pair >= 0 && pair_mode() < 0 && pair_auto && a.M > BM && nsm >= 2) bn = 256;
  if (bn != 64 && bn != 128 && bn != 256) return set_error("gemm: unsupported BN %d", bn);
  const int tiles_n = (a.N + bn - 1) / bn;
  plan(bn, &splits);
  const bool pair_ok = !lnf && a.pair >= 0 && bn >= 128 && a.M > BM && nsm >= 2;
  const bool pair = pair_ok && (a.pair > 0 || pair_mode() == 1 || (pair_mode() < 0 && pair_auto && bn == 256));
  int tiles_m_eff = tiles_m;
  if (pair) {
    tiles_m_eff = (a.M + 2 * BM - 1) / (2 * BM);
    const long tiles = long(tiles_m_eff) * tiles_n;
    const int workers = nsm / 2;
    int sp = 1;
    if (a.accumulate) {
      sp = a.split_k > 0 ? a.split_k : int(std::max<long>(1, std::min<long>(workers / tiles, total_kb / 4)));
      sp = std::max(1, std::min(sp, total_kb));
    }
    const int kbp = (total_kb + sp - 1) / sp;
    splits = (total_kb + kbp - 1) / kbp;
  }
  const int kb_per = (total_kb + splits - 1) / splits;

  GemmParams p;
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.tiles_m = tiles_m_eff; p.tiles_n = tiles_n; p.splits = splits; p.kb_per_split = kb_per;
  p.d = a.d; p.ldd = a.ldd; p.d_fp32 = a.d_fp32; p.accumulate = a.accumulate;
  p.d_vec = ((reinterpret_cast<uintptr_t>(a.d) & 15) == 0) && (a.ldd % (a.d_fp32 ? 4 : 8) == 0);
  if (a.bias && (reinterpret_cast<uintptr_t>(a.bias) & 15)) return set_error("gemm: bias must be 16-byte aligned");
  if (a.resid && ((reinterpret_cast<uintptr_t>(a.resid) & 15) || a.ldr % 8)) return set_error("gemm: resid must be 16-byte aligned with ldr %% 8 == 0");
  p.bias = a.bias; p.relu = a.relu;
  p.resid = reinterpret_cast<const __nv_bfloat16*>(a.resid); p.ldr = a.ldr; p.resid_op = a.resid ? a.resid_op : 0;
  p.alpha = a.alpha;
  p.trace = (g_gemm_trace && !plan_only) ? g_gemm_trace + size_t(g_gemm_trace_next++ % g_gemm_trace_slots) * (148 * 8) : nullptr;
  p.drop = a.drop;
  p.b_static = a.b_static;
  p.colsum = a.colsum_partials;
  p.ln_gamma = a.ln_gamma; p.ln_beta = a.ln_beta; p.ln_mean = a.ln_mean; p.ln_rstd = a.ln_rstd; p.ln_eps = a.ln_eps;
  if (info) { info->splits = splits; info->tiles_m = tiles_m_eff; info->tiles_n = tiles_n; info->bn = bn; info->pair = pair ? 1 : 0; info->fused_ln = lnf ? 1 : 0; }
  if (plan_only) return 0;

  CUtensorMap ta, tb;
  int rc;
  if (a.a_mn) rc = get_tmap(a.a, a.M, a.K, a.lda, 64, &ta);
  else rc = get_tmap(a.a, a.K, a.M, a.lda, BM, &ta);
  if (rc) return rc;
  if (a.b_mn) rc = get_tmap(a.b, a.N, a.K, a.ldb, 64, &tb);
  else rc = get_tmap(a.b, a.K, a.N, a.ldb, pair ? bn / 2 : bn, &tb);  // a pair's CTAs each fetch half of the B tile's rows
  if (rc) return rc;

  // output through TMA (store / reduce-add) whenever its rows are 16-byte aligned with a 16-byte multiple pitch
  CUtensorMap td = ta;
  p.d_tma = ((reinterpret_cast<uintptr_t>(a.d) & 15) == 0) && ((size_t(a.ldd) * (a.d_fp32 ? 4 : 2)) % 16 == 0) && !a.no_tma_store;
  if (p.d_tma) { rc = get_tmap(a.d, a.N, a.M, a.ldd, 32, &td, a.d_fp32 ? 1 : 0); if (rc) return rc; }

  const int total = tiles_m_eff * tiles_n * splits;
  if (lnf) {
    if (!p.d_tma) return set_error("gemm: the fused LayerNorm stores through TMA (16-byte aligned rows)");
    CUtensorMap ty;
    rc = get_tmap(a.ln_y, a.N, a.M, a.ln_ldy, 32, &ty, 0);
    if (rc) return rc;
    const int grid = tiles_m * tiles_n;  // one tile per CTA; a cluster = the tiles_n column tiles of a row block
    switch (bn) {
      case 64: return launch_lnf<64>(a.b_mn, ta, tb, td, ty, p, grid, tiles_n, stream);
      case 128: return launch_lnf<128>(a.b_mn, ta, tb, td, ty, p, grid, tiles_n, stream);
      default: return launch_lnf<256>(a.b_mn, ta, tb, td, ty, p, grid, tiles_n, stream);
    }
  }
  if (pair) {
    const int grid = 2 * min(total, nsm / 2);
    if (bn == 128) return launch_major<128, true>(a.a_mn, a.b_mn, ta, tb, td, p, grid, stream);
    return launch_major<256, true>(a.a_mn, a.b_mn, ta, tb, td, p, grid, stream);
  }
  const int grid = min(total, nsm);
  switch (bn) {
    case 64: return launch_major<64, false>(a.a_mn, a.b_mn, ta, tb, td, p, grid, stream);
    case 128: return launch_major<128, false>(a.a_mn, a.b_mn, ta, tb, td, p, grid, stream);
    default: return launch_major<256, false>(a.a_mn, a.b_mn, ta, tb, td, p, grid, stream);
  }
}

int tensor_map_2d(const void* ptr, unsigned long long inner, unsigned long long outer, unsigned long long ld, unsigned box_outer,
                  void* out_cu_tensor_map, int kind) {
  return get_tmap(ptr, inner, outer, ld, box_outer, static_cast<CUtensorMap*>(out_cu_tensor_map), kind);
}

void gemm_set_trace(unsigned long long* dev_ptr, int slots) {
  g_gemm_trace = dev_ptr; g_gemm_trace_slots = slots > 0 ? slots : 1; g_gemm_trace_next = 0;
}

void gemm_set_pair_mode(int mode) { g_pair_mode.store(mode < -1 || mode > 1 ? -1 : mode, std::memory_order_relaxed); }

void gemm_clear_tensor_map_cache() {
  std::lock_guard<std::mutex> lock(g_map_mutex);
  g_map_cache.clear();
}

}  // namespace b200xf
