// HBM bandwidth for read:write mixes with plain 16-byte accesses (no staging): the ceilings the sample
// kernel (1 B read : 4 B written) can be compared with.  nvcc -arch=sm_100a -O3 -o membw membw.cu ; ./membw
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

__global__ void k_write(float4* dst, size_t n) {
  const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = v;
}
__global__ void k_write_cs(float4* dst, size_t n) {
  const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) __stcs(dst + i, v);
}
__global__ void k_read(const float4* src, size_t n, float* sink) {
  float acc = 0.f;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = __ldg(src + i);
    acc += v.x + v.y + v.z + v.w;
  }
  if (acc == 123.456f) *sink = acc;
}
__global__ void k_copy(const float4* src, float4* dst, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = __ldg(src + i);
}
// 1:4 mix: one u32 (4 pixels) in, one float4 out
__global__ void k_cast(const uint32_t* src, float4* dst, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint32_t w = __ldg(src + i);
    dst[i] = make_float4((float)(w & 255), (float)((w >> 8) & 255), (float)((w >> 16) & 255), (float)(w >> 24));
  }
}
// 1:4 mix, 16 pixels per thread: one uint4 in, four float4 out (64 contiguous bytes per thread)
__global__ void k_cast16(const uint4* src, float4* dst, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 q = __ldg(src + i);
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int j = 0; j < 4; ++j)
      dst[4 * i + j] = make_float4((float)(w[j] & 255), (float)((w[j] >> 8) & 255), (float)((w[j] >> 16) & 255), (float)(w[j] >> 24));
  }
}

template <class F>
static double best_ms(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) f();
  cudaDeviceSynchronize();
  double best = 1e9;
  for (int i = 0; i < 10; ++i) {
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  const size_t bytes = 2ull << 30;   // 2 GiB per buffer
  void *a, *b;
  cudaMalloc(&a, bytes); cudaMalloc(&b, bytes);
  cudaMemset(a, 1, bytes); cudaMemset(b, 2, bytes);
  float* sink; cudaMalloc(&sink, 4);
  const size_t n16 = bytes / 16;
  for (int ctas = 4; ctas <= 16; ctas *= 2) {
    const int grid = 148 * ctas, blk = 256;
    double w = best_ms([&] { k_write<<<grid, blk>>>((float4*)a, n16); });
    double wcs = best_ms([&] { k_write_cs<<<grid, blk>>>((float4*)a, n16); });
    double r = best_ms([&] { k_read<<<grid, blk>>>((const float4*)a, n16, sink); });
    double c = best_ms([&] { k_copy<<<grid, blk>>>((const float4*)a, (float4*)b, n16); });
    const size_t npx4 = bytes / 16;          // float4 outputs into b (2 GiB), u32 inputs from a (512 MiB)
    double m = best_ms([&] { k_cast<<<grid, blk>>>((const uint32_t*)a, (float4*)b, npx4); });
    double m16 = best_ms([&] { k_cast16<<<grid, blk>>>((const uint4*)a, (float4*)b, npx4 / 4); });
    printf("{\"ctas_per_sm\": %d, \"write_GBs\": %.1f, \"write_cs_GBs\": %.1f, \"read_GBs\": %.1f, \"copy_GBs\": %.1f, \"cast_1r4w_GBs\": %.1f, \"cast16_1r4w_GBs\": %.1f}\n",
           ctas, bytes / w / 1e6, bytes / wcs / 1e6, bytes / r / 1e6, 2.0 * bytes / c / 1e6, 1.25 * bytes / m / 1e6, 1.25 * bytes / m16 / 1e6);
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
