// The cross-replica gradient SUM of MultiWorkerMirroredStrategy (cluster/cluster.py:66, triggered inside
// optimizer.apply_gradients at __main__.py:113-114) fused with the Adam update, as ONE kernel over NVLink / NVSwitch:
//
//   reduce-scatter : every rank owns 1/world of a bucket; it reads the SUM of all replicas' gradients of its shard with
//                    multimem.ld_reduce (the NVSwitch adds the replicas' values in flight: NVLS),
//   Adam           : applies the TF-form update (same arithmetic as adam_kernel) to its shard with its local m / v,
//   all-gather     : writes the updated fp32 weights to EVERY replica with multimem.st (the switch replicates the store).
//
// Below 8 replicas a point-to-point variant of the same kernel (plain loads from every replica's gradients, plain stores into every
// replica's weights, over the peer mappings of the same symmetric buffers) moves fewer bytes per link: (W-1)/W bucket sizes per
// direction against ~(1 + 1/W) through the switch.
// Inbound (reduce) and outbound (broadcast) traffic run concurrently on the two directions of the links, where the
// NCCL sequence reduce-scatter -> adam -> all-gather runs them one after the other.  The flat gradient and parameter buffers
// live in symmetric memory with a multicast mapping (allocated by the host side, torch.distributed._symmetric_memory).
// Replicas synchronise through flag words in symmetric memory: "ready" (my gradients of this bucket are final) before the
// reduce, "done" (my shard's new weights are in everybody's memory, and I no longer read anybody's gradients) before the
// bf16 refresh / gradient clear.  Flags carry a launch counter kept on the device, so a captured graph replays correctly.
#include <algorithm>

#include "internal.h"
#include "nvls.h"

namespace b200xf {

namespace {

__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc_addr) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float* mc_addr, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Wait until every replica's flag in this rank's own pad has reached `epoch` (wrap-safe comparison).  A bounded spin: a
// replica that never arrives (crashed process) must not hang the GPU forever -- after ~10 s the error word is set and the
// kernel goes on (its results are then garbage, the host reports the error).
__device__ __forceinline__ void wait_flags(const unsigned* pad, int world, unsigned epoch, unsigned* error_word) {
  if (int(threadIdx.x) < world) {
    const unsigned* f = pad + threadIdx.x;
    long long spins = 0;
    while (int(ld_acquire_sys(f) - epoch) < 0) {
      __nanosleep(64);
      if (++spins > (1ll << 27)) { atomicExch(error_word, 1u); break; }
    }
  }
  __syncthreads();
}

constexpr int kUnroll = 8;   // independent 16-byte multimem loads in flight per thread: a load through the switch takes microseconds
constexpr int kThreadsNvls = 512;

// grid-stride over the shard's float4s, kUnroll independent multimem loads in flight per thread
template <bool P2P>
__global__ void __launch_bounds__(kThreadsNvls)
nvls_reduce_adam_broadcast_kernel(NvlsComm c, float4* __restrict__ m, float4* __restrict__ v, const float4* __restrict__ w_local,
                                  size_t shard_off4, size_t n4, const long long* __restrict__ iterations, AdamHyper h,
                                  const unsigned* __restrict__ epoch_dev, unsigned* __restrict__ cta_counter) {
  const unsigned epoch = *epoch_dev + 1u;
  // ---- "ready": this replica's gradients of the bucket are final (they were written by earlier kernels of this stream)
  if (blockIdx.x == 0 && int(threadIdx.x) < c.world) {
    __threadfence_system();
    st_release_sys(c.pad_peer[threadIdx.x] + kNvlsReady + c.rank, epoch);
  }
  wait_flags(c.pad_local + kNvlsReady, c.world, epoch, c.pad_local + kNvlsError);

  const float step = float(*iterations);
  const float lr = h.rsqrt_d_model * fminf(step * h.warmup_pow_m15, 1.0f / sqrtf(step));
  const float t = step + 1.0f;
  const float alpha = lr * sqrtf(1.0f - powf(h.beta2, t)) / (1.0f - powf(h.beta1, t));
  const float c1 = 1.0f - h.beta1, c2 = 1.0f - h.beta2;
  const float4* g_mc = reinterpret_cast<const float4*>(c.grads_mc) + shard_off4;
  float4* w_mc = reinterpret_cast<float4*>(c.params_mc) + shard_off4;
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i0 = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i0 < n4; i0 += stride * kUnroll) {
    float4 G[kUnroll], W[kUnroll], M[kUnroll], V[kUnroll];
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      const size_t i = i0 + size_t(k) * stride;
      if (i < n4) {
        if (P2P) {   // the replicas' gradients one by one, always in rank order (a shard is summed by exactly one rank)
          float4 acc = __ldcg(reinterpret_cast<const float4*>(c.grads_peer[0]) + shard_off4 + i);
          for (int r = 1; r < c.world; ++r) {
            const float4 t4 = __ldcg(reinterpret_cast<const float4*>(c.grads_peer[r]) + shard_off4 + i);
            acc.x += t4.x; acc.y += t4.y; acc.z += t4.z; acc.w += t4.w;
          }
          G[k] = acc;
        } else {
          G[k] = multimem_ld_reduce_add(reinterpret_cast<const float*>(g_mc + i));
        }
        W[k] = w_local[shard_off4 + i]; M[k] = m[shard_off4 + i]; V[k] = v[shard_off4 + i];
      }
    }
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) {
      const size_t i = i0 + size_t(k) * stride;
      if (i >= n4) break;
      // each operation rounded separately, exactly as in adam_kernel (no FMA contraction)
#define ADAM1(c)                                                                                    \
      M[k].c = __fadd_rn(M[k].c, __fmul_rn(__fsub_rn(G[k].c, M[k].c), c1));                         \
      V[k].c = __fadd_rn(V[k].c, __fmul_rn(__fsub_rn(__fmul_rn(G[k].c, G[k].c), V[k].c), c2));      \
      W[k].c = __fsub_rn(W[k].c, __fdiv_rn(__fmul_rn(M[k].c, alpha), __fadd_rn(__fsqrt_rn(V[k].c), h.eps)));
      ADAM1(x) ADAM1(y) ADAM1(z) ADAM1(w)
#undef ADAM1
      m[shard_off4 + i] = M[k]; v[shard_off4 + i] = V[k];
      if (P2P) {
        for (int r = 0; r < c.world; ++r) __stcg(reinterpret_cast<float4*>(c.params_peer[r]) + shard_off4 + i, W[k]);
      } else {
        multimem_st(reinterpret_cast<float*>(w_mc + i), W[k]);   // every replica's copy of the weights, this one's included
      }
    }
  }
  // ---- "done": once every CTA's stores are out, tell every replica
  __threadfence_system();
  __syncthreads();
  __shared__ bool last;
  if (threadIdx.x == 0) last = (atomicAdd(cta_counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (last) {
    if (threadIdx.x == 0) *cta_counter = 0;
    if (int(threadIdx.x) < c.world) {
      __threadfence_system();
      st_release_sys(c.pad_peer[threadIdx.x] + kNvlsDone + c.rank, epoch);
    }
  }
}

// One small CTA: returns when every replica has finished the bucket (their shards of the weights have landed here, and
// nobody reads this replica's gradients any more), then moves the launch counter on.  The bf16 refresh / gradient clear
// that follows on the stream is an ordinary kernel.
__global__ void nvls_wait_done_kernel(NvlsComm c, unsigned* epoch_dev) {
  const unsigned epoch = *epoch_dev + 1u;
  wait_flags(c.pad_local + kNvlsDone, c.world, epoch, c.pad_local + kNvlsError);
  if (threadIdx.x == 0) *epoch_dev = epoch;
}

}  // namespace

int nvls_reduce_adam_broadcast(const NvlsComm& c, float* w_local, float* m, float* v, size_t shard_off, size_t n, const long long* iterations_dev,
                               AdamHyper h, unsigned* epoch_dev, unsigned* cta_counter, int max_ctas, cudaStream_t stream) {
  if (n % 4 || shard_off % 4) return set_error("nvls update: shard must be a multiple of 4 floats");
  if (n == 0) return 0;
  const size_t n4 = n / 4;
  int grid = int(std::min<size_t>((n4 + size_t(kThreadsNvls) * kUnroll - 1) / (size_t(kThreadsNvls) * kUnroll), size_t(max_ctas > 0 ? max_ctas : 64)));
  if (grid < 1) grid = 1;
  if (c.p2p)
    nvls_reduce_adam_broadcast_kernel<true><<<grid, kThreadsNvls, 0, stream>>>(c, reinterpret_cast<float4*>(m), reinterpret_cast<float4*>(v),
                                                                      reinterpret_cast<const float4*>(w_local), shard_off / 4, n4, iterations_dev, h,
                                                                      epoch_dev, cta_counter);
  else
    nvls_reduce_adam_broadcast_kernel<false><<<grid, kThreadsNvls, 0, stream>>>(c, reinterpret_cast<float4*>(m), reinterpret_cast<float4*>(v),
                                                                       reinterpret_cast<const float4*>(w_local), shard_off / 4, n4, iterations_dev, h,
                                                                       epoch_dev, cta_counter);
  B200XF_CUDA(cudaGetLastError());
  nvls_wait_done_kernel<<<1, 32, 0, stream>>>(c, epoch_dev);
  B200XF_CUDA(cudaGetLastError());
  count_launch(2);
  return 0;
}

}  // namespace b200xf
