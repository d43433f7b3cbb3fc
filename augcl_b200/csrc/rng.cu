// Device RNG fills (Philox4x32-10).  Stand in for the reference's host draws:
//   np.random.randint            /root/reference/src/utils.py:232          -> augcl_rng_fill_indices
//   LongTensor(n).random_(0, m)  /root/reference/src/augmentations.py:404   -> augcl_rng_fill_offsets
//   kornia RandomCrop parameters /root/reference/src/augmentations.py:382   -> augcl_rng_fill_offsets
//   torch.randn(3,3,3,3)         /root/reference/src/augmentations.py:269   -> augcl_rng_fill_normal
// Element e of a fill uses 32-bit word e of the stream (block e/4 = counter (e/4, subsequence), key = seed).
#include "common.cuh"

namespace augcl {

template <typename T>
__global__ void __launch_bounds__(256) fill_randint_kernel(T* out, int64_t count, int64_t lo, uint64_t range,
                                                           uint64_t seed, uint64_t subseq) {
  const int64_t nblk = (count + 3) / 4;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nblk; b += (int64_t)gridDim.x * blockDim.x) {
    const uint4 r = philox4x32_10((uint64_t)b, subseq, seed);
    const uint32_t wds[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t e = b * 4 + j;
      if (e < count) out[e] = (T)(lo + (int64_t)(((uint64_t)wds[j] * range) >> 32));
    }
  }
}

__global__ void __launch_bounds__(256) fill_normal_kernel(float* out, int64_t count, uint64_t seed, uint64_t subseq) {
  const int64_t nblk = (count + 3) / 4;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nblk; b += (int64_t)gridDim.x * blockDim.x) {
    const uint4 r = philox4x32_10((uint64_t)b, subseq, seed);
    const uint32_t wds[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      float nc, ns;
      normal_pair(wds[2 * j], wds[2 * j + 1], nc, ns);
      const int64_t e = b * 4 + 2 * j;
      if (e < count) out[e] = nc;
      if (e + 1 < count) out[e + 1] = ns;
    }
  }
}

// byte e of a fill = byte (e % 16) of Philox block e / 16 (little-endian words)
__global__ void __launch_bounds__(256) fill_u8_kernel(uint8_t* out, int64_t count, uint64_t seed, uint64_t subseq) {
  const int64_t nblk = (count + 15) / 16;
  const bool vec = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nblk; b += (int64_t)gridDim.x * blockDim.x) {
    const uint4 r = philox4x32_10((uint64_t)b, subseq, seed);
    if (vec && b * 16 + 16 <= count) {
      reinterpret_cast<uint4*>(out)[b] = r;
    } else {
      const uint32_t wds[4] = {r.x, r.y, r.z, r.w};
      for (int j = 0; j < 16 && b * 16 + j < count; ++j) out[b * 16 + j] = (uint8_t)(wds[j >> 2] >> (8 * (j & 3)));
    }
  }
}

static inline int rng_grid(int64_t count) {
  int64_t blocks = ((count + 3) / 4 + 255) / 256;
  return (int)(blocks < 1 ? 1 : (blocks > kNumSMs * 8 ? kNumSMs * 8 : blocks));
}

}  // namespace augcl

using namespace augcl;

extern "C" {

int augcl_rng_fill_offsets(int32_t* out, int64_t count, int32_t lo, int32_t hi, uint64_t seed, uint64_t subsequence,
                           augcl_stream stream) {
  AUGCL_REQUIRE(out && count > 0, "augcl_rng_fill_offsets: out is NULL or count <= 0");
  AUGCL_REQUIRE(hi > lo, "augcl_rng_fill_offsets: empty range [%d, %d)", lo, hi);
  fill_randint_kernel<int32_t><<<rng_grid(count), 256, 0, (cudaStream_t)stream>>>(out, count, lo, (uint64_t)((int64_t)hi - lo), seed,
                                                                                  subsequence);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

int augcl_rng_fill_indices(int64_t* out, int64_t count, int64_t lo, int64_t hi, uint64_t seed, uint64_t subsequence,
                           augcl_stream stream) {
  AUGCL_REQUIRE(out && count > 0, "augcl_rng_fill_indices: out is NULL or count <= 0");
  AUGCL_REQUIRE(hi > lo && hi - lo <= (1ll << 32), "augcl_rng_fill_indices: range [%lld, %lld) empty or wider than 2^32",
                (long long)lo, (long long)hi);
  fill_randint_kernel<int64_t><<<rng_grid(count), 256, 0, (cudaStream_t)stream>>>(out, count, lo, (uint64_t)(hi - lo), seed,
                                                                                  subsequence);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

int augcl_rng_fill_bytes(uint8_t* out, int64_t count, uint64_t seed, uint64_t subsequence, augcl_stream stream) {
  AUGCL_REQUIRE(out && count > 0, "augcl_rng_fill_bytes: out is NULL or count <= 0");
  fill_u8_kernel<<<rng_grid((count + 3) / 4), 256, 0, (cudaStream_t)stream>>>(out, count, seed, subsequence);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

int augcl_rng_fill_normal(float* out, int64_t count, uint64_t seed, uint64_t subsequence, augcl_stream stream) {
  AUGCL_REQUIRE(out && count > 0, "augcl_rng_fill_normal: out is NULL or count <= 0");
  fill_normal_kernel<<<rng_grid(count), 256, 0, (cudaStream_t)stream>>>(out, count, seed, subsequence);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

}  // extern "C"
