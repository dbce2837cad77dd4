// Shared between the orchestration and the NVLS (NVLink SHARP / multimem) update kernels.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>

namespace b200xf {

constexpr int kNvlsMaxWorld = 8;
// word offsets inside a replica's flag pad (one word per sending replica in each group)
constexpr int kNvlsReady = 0, kNvlsDone = 16, kNvlsError = 32, kNvlsPadWords = 64;

struct NvlsComm {
  float* grads_mc = nullptr;    // multicast mapping of the flat gradient buffer (a load reduces over every replica's copy)
  float* params_mc = nullptr;   // multicast mapping of the flat fp32 weights (a store lands in every replica's copy)
  // the same two buffers of every replica as mapped into this process (own ones included): the point-to-point variant of the
  // update reads its shard of the gradients from each replica in turn and writes the new weights into each replica's copy
  float* grads_peer[kNvlsMaxWorld] = {nullptr};
  float* params_peer[kNvlsMaxWorld] = {nullptr};
  int p2p = 0;                                   // 1: point-to-point loads / stores instead of multimem (fewer bytes per link below 8 replicas)
  unsigned* pad_local = nullptr;                 // this replica's flag pad
  unsigned* pad_peer[kNvlsMaxWorld] = {nullptr}; // every replica's flag pad as mapped into this process (own one included)
  int rank = 0, world = 1;
};

struct AdamHyper;
// One bucket's update: reduce the shard [shard_off, shard_off + n) of the gradients over all replicas, Adam on it with the
// local m / v / weights, broadcast the new weights to every replica; then a one-CTA kernel that returns once every replica
// has done the same.  epoch_dev / cta_counter: zero-initialised device words owned by the caller.
int nvls_reduce_adam_broadcast(const NvlsComm& c, float* w_local, float* m, float* v, size_t shard_off, size_t n, const long long* iterations_dev,
                               AdamHyper h, unsigned* epoch_dev, unsigned* cta_counter, int max_ctas, cudaStream_t stream);

}  // namespace b200xf
