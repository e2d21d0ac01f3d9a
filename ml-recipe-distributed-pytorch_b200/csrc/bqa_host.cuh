// bqa_host.cuh — host-side helpers shared by the C-ABI translation units.
#pragma once

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

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

inline int device_sm_count(int* out) {
  int dev = 0;
  BQA_CUDA_CHECK(cudaGetDevice(&dev));
  static int cached_dev = -1, cached = 0;
  if (cached_dev != dev) {
    BQA_CUDA_CHECK(cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev));
    cached_dev = dev;
  }
  *out = cached;
  return 0;
}

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
