// microbenchmark: scalar FFMA vs packed FFMA2 throughput and issue cost on sm_100a
#include <cuda_runtime.h>
#include <cstdio>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float s) {
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
  if (MODE == 0) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], s, 0.5f);
    }
  } else {
    u64 p[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { float2 f = make_float2(a[2*i], a[2*i+1]); p[i] = *reinterpret_cast<u64*>(&f); }
    float2 sf = make_float2(s, s), hf = make_float2(0.5f, 0.5f);
    u64 s2 = *reinterpret_cast<u64*>(&sf), h2 = *reinterpret_cast<u64*>(&hf);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], s2, h2);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) { float2 f = *reinterpret_cast<float2*>(&p[i]); a[2*i] = f.x; a[2*i+1] = f.y; }
  }
  float t = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) t += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int mode = 0; mode < 2; ++mode) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) k<0><<<148 * 8, 256>>>(out, iters, 0.999f); else k<1><<<148 * 8, 256>>>(out, iters, 0.999f);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double flops = 2.0 * 16 * iters * 148 * 8 * 256;
      if (rep) printf("%s: %.3f ms  %.1f TFLOP/s fp32\n", mode ? "FFMA2 (packed)" : "FFMA (scalar)", ms, flops / ms / 1e9);
    }
  }
  return 0;
}
