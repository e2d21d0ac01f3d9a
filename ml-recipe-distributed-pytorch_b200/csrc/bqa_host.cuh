// bqa_host.cuh — host-side helpers shared by the C-ABI translation units.
#pragma once

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include <atomic>

#include "bqa_common.cuh"

namespace bqa {

void set_error(const char* fmt, ...);
const char* last_error();

#define BQA_CUDA_CHECK(expr)                                                                   \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      bqa::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return 1;                                                                                \
    }                                                                                          \
  } while (0)

constexpr int MAX_DEVICES = 64;

// SM count of the current device; cached per device ordinal (atomics: any thread may ask)
inline int device_sm_count(int* out) {
  static std::atomic<int> cached[MAX_DEVICES];  // zero-initialised: 0 = not queried yet
  int dev = 0;
  BQA_CUDA_CHECK(cudaGetDevice(&dev));
  int n = (dev >= 0 && dev < MAX_DEVICES) ? cached[dev].load(std::memory_order_relaxed) : 0;
  if (n == 0) {
    BQA_CUDA_CHECK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    if (dev >= 0 && dev < MAX_DEVICES) cached[dev].store(n, std::memory_order_relaxed);
  }
  *out = n;
  return 0;
}

// Function attributes (cudaFuncSetAttribute) are per device: one flag per device ordinal and call site.
// needs_setup() is true until mark_done() has been called for the current device; running the set-up twice
// (two threads racing) is harmless.
struct PerDeviceOnce {
  std::atomic<bool> done[MAX_DEVICES];
  // true until mark_done(*dev) has been called for the current device; *dev = its ordinal (-1 if unknown:
  // then the set-up runs on every call)
  bool needs_setup(int* dev) {
    if (cudaGetDevice(dev) != cudaSuccess || *dev < 0 || *dev >= MAX_DEVICES) {
      *dev = -1;
      return true;
    }
    return !done[*dev].load(std::memory_order_acquire);
  }
  void mark_done(int dev) {
    if (dev >= 0) done[dev].store(true, std::memory_order_release);
  }
};

inline uint64_t splitmix64(uint64_t x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

// (seed, site) -> dropout stream key.  `site` enumerates the dropout call sites of one step
// (embeddings, 3 per layer, classifier — SURVEY.md App. A) combined with the step counter by
// the caller.  The Python oracle restates this function (oracle/bertqa_oracle.py: drop_key).
inline DropKey make_drop_key(uint64_t seed, uint32_t site, float p) {
  DropKey k;
  const uint64_t h = splitmix64(seed ^ (0xd1b54a32d192ed03ull * (static_cast<uint64_t>(site) + 1ull)));
  k.k0 = static_cast<uint32_t>(h);
  k.k1 = static_cast<uint32_t>(h >> 32);
  if (p > 0.0f) {
    k.thresh = static_cast<uint32_t>(lrintf(p * 65536.0f));
    if (k.thresh > 65535u) k.thresh = 65535u;
    k.scale = 1.0f / (1.0f - p);
  } else {
    k.thresh = 0u;
    k.scale = 1.0f;
  }
  return k;
}

}  // namespace bqa
