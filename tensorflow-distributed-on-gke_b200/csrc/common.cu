// Error reporting and small process-wide bookkeeping shared by all translation units.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "internal.h"

namespace b200xf {

namespace {
thread_local char g_error[1024] = "";
std::atomic<long long> g_launches{0};
}  // namespace

int set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
  return 1;
}
const char* last_error_string() { return g_error; }

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("B200XF_PDL"); v = (e && atoi(e) == 0) ? 0 : 1; }
  return v != 0;
}

int stream_priority(cudaStream_t stream) {
  // streams are few and long-lived; a stale entry (handle reused after destroy) only mis-states a scheduling hint
  thread_local cudaStream_t keys[8]; thread_local int vals[8]; thread_local int used = 0;
  for (int i = 0; i < used; ++i) if (keys[i] == stream) return vals[i];
  int prio = 0;
  if (cudaStreamGetPriority(stream, &prio) != cudaSuccess) { (void)cudaGetLastError(); prio = 0; }
  if (used < 8) { keys[used] = stream; vals[used] = prio; ++used; }
  return prio;
}

int ensure_dynamic_smem(const void* kernel, int bytes) {
  struct Entry { const void* kernel; unsigned long long devices; int bytes; };
  static Entry table[256];
  static int used = 0;
  static std::mutex mu;
  int dev = 0;
  B200XF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return set_error("device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  Entry* e = nullptr;
  for (int i = 0; i < used; ++i) if (table[i].kernel == kernel) { e = &table[i]; break; }
  if (!e) {
    if (used == 256) return set_error("kernel attribute table full");
    e = &table[used++]; e->kernel = kernel; e->devices = 0; e->bytes = 0;
  }
  if (((e->devices >> dev) & 1ull) && e->bytes >= bytes) return 0;
  if (bytes > e->bytes) { e->bytes = bytes; e->devices = 0; }
  const cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, e->bytes);
  if (err != cudaSuccess) return set_error("cudaFuncSetAttribute(MaxDynamicSharedMemorySize=%d): %s", e->bytes, cudaGetErrorString(err));
  e->devices |= 1ull << dev;
  return 0;
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
long long launches_total() { return g_launches.load(std::memory_order_relaxed); }

}  // namespace b200xf
