// Issue rate of packed fp32 (FFMA2) against scalar FFMA on sm_100a, alone and mixed with MUFU.EX2.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu && ./ffma2
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(float* out, long long* clk, int iters) {
  float a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = threadIdx.x * 0.001f + i; b[i] = 1.0f + i * 1e-6f; }
  const float c = 0.999f, d = 1e-3f;
  unsigned long long p[4], q[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    asm("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(a[2 * i]), "f"(a[2 * i + 1]));
    asm("mov.b64 %0, {%1, %2};" : "=l"(q[i]) : "f"(b[2 * i]), "f"(b[2 * i + 1]));
  }
  unsigned u[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) u[i] = threadIdx.x * 2654435761u + i;
  unsigned long long cc, dd;
  asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
  asm("mov.b64 %0, {%1, %1};" : "=l"(dd) : "f"(d));
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {  // 8 independent scalar FFMA chains
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], c, d);
    } else if (MODE == 1) {  // 4 independent packed chains = the same 8 values
#pragma unroll
      for (int i = 0; i < 4; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(cc), "l"(dd));
    } else if (MODE == 2) {  // 8 scalar FFMA + 2 MUFU
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], c, d);
#pragma unroll
      for (int i = 0; i < 2; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b[i]));
    } else if (MODE == 3) {  // 4 packed + 2 MUFU
#pragma unroll
      for (int i = 0; i < 4; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(cc), "l"(dd));
#pragma unroll
      for (int i = 0; i < 2; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b[i]));
    } else if (MODE == 4) {  // 8 scalar FFMA + 8 LOP3-ish integer ops
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], c, d);
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("xor.b32 %0, %0, 0x5bd1e995;" : "+r"(u[i]));
    } else if (MODE == 5) {  // 4 packed + 8 integer ops
#pragma unroll
      for (int i = 0; i < 4; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(cc), "l"(dd));
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("xor.b32 %0, %0, 0x5bd1e995;" : "+r"(u[i]));
    }
  }
  const long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i] + b[i] + __uint_as_float(u[i] & 0x3fffffffu);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i]));
    s += lo + hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(q[i]));
    s += lo + hi;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <int MODE>
void run(const char* name, int threads) {
  float* out; long long* clk; long long h;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&clk, 8);
  const int iters = 20000;
  k<MODE><<<148, threads>>>(out, clk, iters);
  k<MODE><<<148, threads>>>(out, clk, iters);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
  printf("%-28s warps/SM %2d: %.2f clk per loop iteration (8 fp32 FMAs per thread per iteration)\n", name, threads / 32, double(h) / iters);
  cudaFree(out); cudaFree(clk);
}

int main() {
  for (int th : {128, 256, 512}) {
    run<0>("FFMA x8", th);
    run<1>("FFMA2 x4", th);
    run<2>("FFMA x8 + MUFU x2", th);
    run<3>("FFMA2 x4 + MUFU x2", th);
    run<4>("FFMA x8 + LOP x8", th);
    run<5>("FFMA2 x4 + LOP x8", th);
  }
  return 0;
}
