// Shared host/device helpers for the augcl_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdio>
#include <string>

#include "augcl_b200.h"

namespace augcl {

// ---- process-wide state (state.cu) ----
extern std::atomic<int64_t> g_launches;
extern std::atomic<int> g_staging;  // 0 = tma (default), 1 = generic
void set_error(const std::string& msg);
// Opt a kernel into `bytes` of dynamic shared memory on the CURRENT device.  The attribute is per (function, device);
// the (function, device) pairs already configured are remembered, so a repeated call costs one cudaGetDevice.
// Returns AUGCL_OK or AUGCL_ERR_CUDA (message set).
int ensure_dyn_smem(const void* func, int bytes);

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

#define AUGCL_FAIL(code, ...)                            \
  do {                                                   \
    char _b[512];                                        \
    snprintf(_b, sizeof(_b), __VA_ARGS__);               \
    ::augcl::set_error(_b);                              \
    return (code);                                       \
  } while (0)

#define AUGCL_REQUIRE(cond, ...) \
  do {                           \
    if (!(cond)) AUGCL_FAIL(AUGCL_ERR_BAD_ARG, __VA_ARGS__); \
  } while (0)

#define AUGCL_CUDA(expr)                                                                   \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess)                                                                 \
      AUGCL_FAIL(AUGCL_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),   \
                 __FILE__, __LINE__);                                                      \
  } while (0)

// Checks the launch that was just enqueued and counts it.
#define AUGCL_LAUNCHED()                                   \
  do {                                                     \
    AUGCL_CUDA(cudaGetLastError());                        \
    ::augcl::g_launches.fetch_add(1, std::memory_order_relaxed); \
  } while (0)

// Launch with programmatic stream serialization (PDL): the grid's CTAs may become resident while the previous kernel of the
// stream drains.  The kernel must call pdl_wait() before it reads anything a predecessor may have written, and
// pdl_launch_dependents() to release its own successor early.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t st, bool pdl,
                              Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

// RAII: make `device` current for the scope of one API call, restore afterwards.
struct DeviceGuard {
  int prev = -1;
  bool switched = false;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != device) {
      err = cudaSetDevice(device);
      switched = (err == cudaSuccess);
    }
  }
  ~DeviceGuard() {
    if (switched) cudaSetDevice(prev);
  }
};

// ---------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------

// Byte `i` (0..3) of a 32-bit word as an exact f32, without the quarter-rate I2F pipe:
// PRMT builds the float 2^23 + byte, one FADD removes the bias.
template <int I>
__device__ __forceinline__ float byte_to_float(uint32_t w) {
  return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650 + I)) - 8388608.0f;
}

__device__ __forceinline__ float4 bytes_to_float4(uint32_t w) {
  return make_float4(byte_to_float<0>(w), byte_to_float<1>(w), byte_to_float<2>(w), byte_to_float<3>(w));
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

// Philox4x32-10 (Salmon et al., SC'11).  One call = 4 x u32 for counter block (blk, subseq).
__device__ __forceinline__ uint4 philox4x32_10(uint64_t blk, uint64_t subseq, uint64_t seed) {
  uint32_t c0 = (uint32_t)blk, c1 = (uint32_t)(blk >> 32), c2 = (uint32_t)subseq, c3 = (uint32_t)(subseq >> 32);
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

// Box-Muller on an exactly representable pair: u1 in (0,1] from w0, u2 in [0,1) from w1 (24 bits each).
// One definition for augcl_rng_fill_normal and the in-kernel conv-weight draw of the fused sampler: same values.
__device__ __forceinline__ void normal_pair(uint32_t w0, uint32_t w1, float& n_cos, float& n_sin) {
  const float u1 = (float)((w0 >> 8) + 1u) * (1.0f / 16777216.0f);
  const float u2 = (float)(w1 >> 8) * (1.0f / 16777216.0f);
  const float rad = sqrtf(-2.0f * logf(u1));
  float sn, cs;
  sincospif(2.0f * u2, &sn, &cs);
  n_cos = rad * cs;
  n_sin = rad * sn;
}

// ---- mbarrier + 1-D bulk copy (TMA engine, SASS: UBLKCP) ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {   // release at CTA scope
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// elect.sync over `mask` (every thread of `mask`, and only those, must execute it): true in exactly one of them.  Code
// under the returned predicate is single-threaded BY CONSTRUCTION for ptxas, so instructions that take uniform-register
// operands (UBLKCP) get plain R2UR moves instead of a divergence-safe loop (ELECT / R2UR / BRA.U.ANY per instruction).
__device__ __forceinline__ bool elect_in(uint32_t mask) {
  uint32_t leader;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, %1;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(leader)
      : "r"(mask));
  return leader != 0;
}
// global -> shared bulk copy; dst/src 16-B aligned, bytes a multiple of 16; completes on `bar`.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}


// 16-byte store through the global (not generic) address space: STG.E.128.
__device__ __forceinline__ void st_global_f4(float4* ptr, const float4& v) {
  asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(__cvta_generic_to_global(ptr)), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Programmatic dependent launch (PDL): a kernel launched with the programmatic-stream-serialization
// attribute may become resident while its predecessor drains; pdl_wait() blocks until the predecessor
// has completed and its writes are visible, pdl_launch_dependents() lets the successor start early.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Byte selector for replicate padding of a 4-pixel group.  The group wants columns c0..c0+3 of a row of
// width W; the word actually loaded starts at cs = clamp(c0, 0, W-4).  d = clamp(c0 - cs, -4, 4) and
// output byte j must take loaded byte clamp(j + d, 0, 3): a 16-bit PRMT selector cut out of the constant
// 0x3333'3210'0000 (d = 0 -> 0x3210 identity, d = -1 -> 0x2100, d = +1 -> 0x3321, |d| = 4 -> all edge).
__device__ __forceinline__ uint32_t edge_selector(int d) {
  return __funnelshift_rc(0x32100000u, 0x00003333u, (uint32_t)(4 * (d + 4))) & 0xFFFFu;
}

}  // namespace augcl
