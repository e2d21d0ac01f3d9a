// optimizer.cu — global gradient-norm clip fused with the HF-AdamW update over the flat
// parameter buffer, plus the fp32 -> bf16 shadow-weight refresh.
//
// Reference: Trainer._clip_grad_norm (modules/model/trainer/trainer.py:221-225, i.e.
// torch.nn.utils.clip_grad_norm_) followed by optimizer.step() (trainer.py:286) on
// transformers.AdamW(correct_bias=False) built in modules/init.py:125-138:
//   g <- g * min(1, max_norm / (||g||_2 + 1e-6))
//   m <- b1 m + (1-b1) g ; v <- b2 v + (1-b2) g^2 ; p <- p - lr * m / (sqrt(v) + eps)
//   p <- p - lr * wd * p         (decoupled decay, applied after the Adam update, no bias correction)
// The reference runs ~10 ATen launches per tensor x 207 tensors plus 208 norm launches; here
// the step is two HBM-bound passes: sum of squares (read g), then one pass that reads
// g,p,m,v and writes p,m,v, the bf16 shadow used by the GEMMs, and zeroes g (optimizer.zero_grad).
#include "kernels.cuh"

namespace bqa {

namespace {

__global__ void __launch_bounds__(256)
sumsq_partial_kernel(const float* __restrict__ g, size_t n, float* __restrict__ partial) {
  __shared__ float red[8];
  float acc = 0.0f;
  const size_t n4 = n / 4;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float4 v = __ldg(g4 + i);
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const float v = g[n4 * 4 + threadIdx.x];
    acc += v * v;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.0f;
    for (int w = 0; w < 8; ++w) s += red[w];
    partial[blockIdx.x] = s;
  }
}

__global__ void __launch_bounds__(256)
sumsq_final_kernel(const float* __restrict__ partial, int n, double* __restrict__ accum) {
  // fixed-shape tree: the result does not depend on scheduling (deterministic given the partials)
  __shared__ double red[256];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) s += static_cast<double>(partial[i]);
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *accum += red[0];
}

__global__ void __launch_bounds__(256)
clip_adamw_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                  bf16* __restrict__ pb, const AdamWChunk* __restrict__ chunks, const double* __restrict__ sumsq,
                  float max_grad_norm, LrTable lrs, float beta1, float beta2, float eps, int zero_grad) {
  const AdamWChunk ch = chunks[blockIdx.x];
  const float lr = lrs.lr[ch.group & 7];
  float coef = 1.0f;
  if (max_grad_norm > 0.0f) {
    const float total = sqrtf(static_cast<float>(*sumsq));
    coef = fminf(max_grad_norm / (total + 1e-6f), 1.0f);
  }
  const float wd_mul = 1.0f - lr * ch.weight_decay;
  // chunk boundaries are multiples of 4 elements by construction (tensors are 128-aligned)
  for (uint64_t i = ch.begin + threadIdx.x * 4ull; i < ch.end; i += blockDim.x * 4ull) {
    float4 gg = *reinterpret_cast<const float4*>(g + i);
    float4 pp = *reinterpret_cast<const float4*>(p + i);
    float4 mm = *reinterpret_cast<const float4*>(m + i);
    float4 vv = *reinterpret_cast<const float4*>(v + i);
    float* gp = reinterpret_cast<float*>(&gg);
    float* ppp = reinterpret_cast<float*>(&pp);
    float* mp = reinterpret_cast<float*>(&mm);
    float* vp = reinterpret_cast<float*>(&vv);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gk = gp[k] * coef;
      mp[k] = beta1 * mp[k] + (1.0f - beta1) * gk;
      vp[k] = beta2 * vp[k] + (1.0f - beta2) * gk * gk;
      float x = ppp[k] - lr * (mp[k] / (sqrtf(vp[k]) + eps));
      if (ch.weight_decay > 0.0f) x = x * wd_mul;
      ppp[k] = x;
    }
    *reinterpret_cast<float4*>(p + i) = pp;
    *reinterpret_cast<float4*>(m + i) = mm;
    *reinterpret_cast<float4*>(v + i) = vv;
    if (pb != nullptr) {
      uint2 o;
      o.x = pack_bf16x2(ppp[0], ppp[1]);
      o.y = pack_bf16x2(ppp[2], ppp[3]);
      *reinterpret_cast<uint2*>(pb + i) = o;
    }
    if (zero_grad) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// AdaMod (reference optim.py:42-100) with the same chunk table, clip coefficient, zero_grad and bf16
// shadow refresh as clip_adamw_kernel.  ss = lr * sqrt(1 - beta2^t) / (1 - beta1^t) per group.
__global__ void __launch_bounds__(256)
clip_adamod_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                   float* __restrict__ s, bf16* __restrict__ pb, const AdamWChunk* __restrict__ chunks,
                   const double* __restrict__ sumsq, float max_grad_norm, LrTable lrs, float beta1, float beta2,
                   float beta3, float eps, float bias_mul, int zero_grad) {
  const AdamWChunk ch = chunks[blockIdx.x];
  const float lr = lrs.lr[ch.group & 7];
  float coef = 1.0f;
  if (max_grad_norm > 0.0f) {
    const float total = sqrtf(static_cast<float>(*sumsq));
    coef = fminf(max_grad_norm / (total + 1e-6f), 1.0f);
  }
  const float ss = lr * bias_mul;                    // step_size before the division by denom
  const float wd_mul = 1.0f - lr * ch.weight_decay;  // p.data.add_(-weight_decay * lr, p.data), before the update
  for (uint64_t i = ch.begin + threadIdx.x * 4ull; i < ch.end; i += blockDim.x * 4ull) {
    float4 gg = *reinterpret_cast<const float4*>(g + i);
    float4 pp = *reinterpret_cast<const float4*>(p + i);
    float4 mm = *reinterpret_cast<const float4*>(m + i);
    float4 vv = *reinterpret_cast<const float4*>(v + i);
    float4 sv = *reinterpret_cast<const float4*>(s + i);
    float* gp = reinterpret_cast<float*>(&gg);
    float* ppp = reinterpret_cast<float*>(&pp);
    float* mp = reinterpret_cast<float*>(&mm);
    float* vp = reinterpret_cast<float*>(&vv);
    float* sp = reinterpret_cast<float*>(&sv);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gk = gp[k] * coef;
      mp[k] = beta1 * mp[k] + (1.0f - beta1) * gk;
      vp[k] = beta2 * vp[k] + (1.0f - beta2) * gk * gk;
      float x = ppp[k];
      if (ch.weight_decay != 0.0f) x *= wd_mul;
      float eta = ss / (sqrtf(vp[k]) + eps);
      sp[k] = beta3 * sp[k] + (1.0f - beta3) * eta;
      eta = fminf(eta, sp[k]);
      ppp[k] = x - eta * mp[k];
    }
    *reinterpret_cast<float4*>(p + i) = pp;
    *reinterpret_cast<float4*>(m + i) = mm;
    *reinterpret_cast<float4*>(v + i) = vv;
    *reinterpret_cast<float4*>(s + i) = sv;
    if (pb != nullptr) {
      uint2 o;
      o.x = pack_bf16x2(ppp[0], ppp[1]);
      o.y = pack_bf16x2(ppp[2], ppp[3]);
      *reinterpret_cast<uint2*>(pb + i) = o;
    }
    if (zero_grad) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

__global__ void __launch_bounds__(256)
f32_to_bf16_kernel(const float* __restrict__ src, bf16* __restrict__ dst, size_t n) {
  const size_t n4 = n / 4;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(src) + i);
    uint2 o;
    o.x = pack_bf16x2(v.x, v.y);
    o.y = pack_bf16x2(v.z, v.w);
    reinterpret_cast<uint2*>(dst)[i] = o;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) dst[n4 * 4 + threadIdx.x] = __float2bfloat16_rn(src[n4 * 4 + threadIdx.x]);
}

}  // namespace

int sumsq_scratch_floats(int num_sms) { return num_sms * 8; }

int launch_grad_sumsq(const float* g, size_t n, double* sumsq_accum, float* partial_scratch, int num_sms,
                      cudaStream_t st) {
  if ((reinterpret_cast<uintptr_t>(g) & 15u) != 0) {
    set_error("grad_sumsq: gradient buffer must be 16-byte aligned");
    return 1;
  }
  int grid = num_sms * 8;
  const size_t need = (n / 4 + 255) / 256;
  if (static_cast<size_t>(grid) > need) grid = static_cast<int>(need ? need : 1);
  sumsq_partial_kernel<<<grid, 256, 0, st>>>(g, n, partial_scratch);
  if (check_launch("sumsq_partial")) return 1;
  sumsq_final_kernel<<<1, 256, 0, st>>>(partial_scratch, grid, sumsq_accum);
  return check_launch("sumsq_final");
}

int launch_clip_adamw(float* p, float* g, float* m, float* v, bf16* p_bf16, const AdamWChunk* chunks_dev,
                      int n_chunks, const double* sumsq, float max_grad_norm, LrTable lrs, float beta1,
                      float beta2, float eps, int zero_grad, cudaStream_t st) {
  if (n_chunks <= 0) return 0;
  clip_adamw_kernel<<<n_chunks, 256, 0, st>>>(p, g, m, v, p_bf16, chunks_dev, sumsq, max_grad_norm, lrs, beta1,
                                              beta2, eps, zero_grad);
  return check_launch("clip_adamw");
}

int launch_clip_adamod(float* p, float* g, float* m, float* v, float* s, bf16* p_bf16, const AdamWChunk* chunks_dev,
                       int n_chunks, const double* sumsq, float max_grad_norm, LrTable lrs, float beta1,
                       float beta2, float beta3, float eps, float bias_mul, int zero_grad, cudaStream_t st) {
  if (n_chunks <= 0) return 0;
  clip_adamod_kernel<<<n_chunks, 256, 0, st>>>(p, g, m, v, s, p_bf16, chunks_dev, sumsq, max_grad_norm, lrs, beta1,
                                               beta2, beta3, eps, bias_mul, zero_grad);
  return check_launch("clip_adamod");
}

int launch_f32_to_bf16(const float* src, bf16* dst, size_t n, int num_sms, cudaStream_t st) {
  int grid = num_sms * 8;
  const size_t need = (n / 4 + 255) / 256;
  if (static_cast<size_t>(grid) > need) grid = static_cast<int>(need ? need : 1);
  f32_to_bf16_kernel<<<grid, 256, 0, st>>>(src, dst, n);
  return check_launch("f32_to_bf16");
}

}  // namespace bqa

// ---------------------------------------------------------------------------------------
// Sum of squares without any allocation (extern "C", declared in include/bqa.h).
//
// bqa_grad_sumsq needs a small scratch buffer for its per-block partial sums and leaves it to the
// caller to provide one.  Some callers cannot: a serving process that hands all device memory to
// its own pool at start-up and forbids cudaMalloc afterwards, a plugin running inside a host
// application whose allocator it has no access to, or code on a latency-critical path where
// cudaMalloc's implicit device-wide synchronisation is not acceptable.  For them the same two
// kernels run with the partial sums in the fixed array below, which is part of the library image:
// every GPU that loads libbqa keeps these 8 KiB until the process exits.
// ---------------------------------------------------------------------------------------
#include <mutex>

#include "../../include/bqa.h"

namespace {

constexpr int kNoAllocPartials = 2048;                 // one float per thread block of the first pass
constexpr uint64_t kNoAllocMaxElems = 1ull << 31;      // keeps the elements behind one fp32 partial below ~2M: the grid
                                                       // is 8 blocks per SM, capped at 2048 (1184 blocks on a B200)
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) float g_noalloc_partials[kNoAllocPartials];

// In host code `g_noalloc_partials` is only a handle for the runtime, not an address a kernel could
// use; the device address is asked for once per device, with that device current.
int noalloc_scratch_address(float** out) {
  static std::mutex mu;
  static float* addr[kNoAllocMaxDevices] = {};
  int dev = 0;
  BQA_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) {
    bqa::set_error("bqa_grad_sumsq_noalloc: device ordinal %d outside [0, %d)", dev, kNoAllocMaxDevices);
    return 1;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    BQA_CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_partials));
    addr[dev] = static_cast<float*>(p);
  }
  *out = addr[dev];
  return 0;
}

}  // namespace

extern "C" int bqa_grad_sumsq_noalloc(const float* g, uint64_t n, double* sumsq, void* stream) {
  if (g == nullptr || sumsq == nullptr) {
    bqa::set_error("bqa_grad_sumsq_noalloc: null argument");
    return 1;
  }
  if (n > kNoAllocMaxElems) {
    bqa::set_error("bqa_grad_sumsq_noalloc: %llu elements exceed the built-in scratch's limit of %llu; "
                   "use bqa_grad_sumsq with a caller-provided scratch buffer",
                   (unsigned long long)n, (unsigned long long)kNoAllocMaxElems);
    return 1;
  }
  if ((reinterpret_cast<uintptr_t>(g) & 15u) != 0) {
    bqa::set_error("bqa_grad_sumsq_noalloc: gradient buffer must be 16-byte aligned");
    return 1;
  }
  float* partials = nullptr;
  if (noalloc_scratch_address(&partials)) return 1;
  int sms = 0;
  if (bqa::device_sm_count(&sms)) return 1;
  int grid = sms * 8;
  if (grid > kNoAllocPartials) grid = kNoAllocPartials;
  const uint64_t need = (n / 4 + 255) / 256;
  if (static_cast<uint64_t>(grid) > need) grid = static_cast<int>(need ? need : 1);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // Every launched block overwrites its own partial and the second pass reads exactly `grid` of them,
  // so what an earlier call left in the array is never read and nothing has to be cleared first.
  bqa::sumsq_partial_kernel<<<grid, 256, 0, st>>>(g, n, partials);
  if (bqa::check_launch("sumsq_partial (no-alloc)")) return 1;
  bqa::sumsq_final_kernel<<<1, 256, 0, st>>>(partials, grid, sumsq);
  return bqa::check_launch("sumsq_final (no-alloc)");
}
