// FFMA vs FFMA2 (fma.rn.f32x2) issue throughput on sm_100a: 16 independent accumulator chains per thread.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu && ./ffma2
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void k_ffma(float* out, float a, float b, int iters) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x + i;
  float x = a + threadIdx.x, w = b;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fmaf(x, w, acc[i]);
    x += 1.0f;
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void ffma2(unsigned long long& d, unsigned long long a, unsigned long long b) {
  asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}

__global__ void k_ffma2(float* out, float a, float b, int iters) {
  unsigned long long acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) { float2 v = make_float2(threadIdx.x + i, i); acc[i] = *reinterpret_cast<unsigned long long*>(&v); }
  float2 xv = make_float2(a + threadIdx.x, a), wv = make_float2(b, b);
  unsigned long long x = *reinterpret_cast<unsigned long long*>(&xv), w = *reinterpret_cast<unsigned long long*>(&wv);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) ffma2(acc[i], x, w);
    x += 1;
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) { float2 v = *reinterpret_cast<float2*>(&acc[i]); s += v.x + v.y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int mode = 0; mode < 2; ++mode)
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) k_ffma<<<148 * 8, 256>>>(out, 1.0f, 1e-7f, iters); else k_ffma2<<<148 * 8, 256>>>(out, 1.0f, 1e-7f, iters);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      const double inst = 148.0 * 8 * 256 * (double)iters * 16;
      printf("%s: %.3f ms, %.2f T thread-instr/s, %.2f TFMA/s\n", mode ? "FFMA2" : "FFMA ", ms, inst / ms / 1e9, inst * (mode ? 2 : 1) / ms / 1e9);
    }
  return 0;
}
