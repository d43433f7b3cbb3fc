// Process-wide state of the library: last error (per thread), launch counter, runtime options.
#include <cstring>

#include "common.cuh"

namespace augcl {
std::atomic<int64_t> g_launches{0};
std::atomic<int> g_staging{0};
std::atomic<int> g_sample_stages{0};   // 0 = auto (4 when a CTA gets few units, else 2: scripts/tune_sample.py)
std::atomic<int> g_sample_ctas_per_sm{6};
std::atomic<int> g_sample_pdl{1};
std::atomic<int> g_sample_band{1};
std::atomic<int> g_sample_fused_min_batch{384};   // augcl_ring_sample*: batches >= this use the fused.cu kernel (0 = never)
extern std::atomic<int> g_fused_ctas_per_sm;   // fused.cu
extern std::atomic<int> g_fused_fast_mask;
extern std::atomic<int> g_fused_stages;
extern std::atomic<int> g_fused_batch;
extern std::atomic<int> g_fused_skip;
extern std::atomic<int> g_fused_interleave;
extern std::atomic<int> g_fused_band;
extern std::atomic<int> g_bil_tma;
extern std::atomic<int> g_conv_part;
static thread_local std::string t_error;
void set_error(const std::string& msg) { t_error = msg; }

// (function, device) pairs whose MaxDynamicSharedMemorySize has been raised.  Lock-free: a slot is claimed by CAS on its
// function pointer; the per-device bits are set only after cudaFuncSetAttribute succeeded (a racing thread at worst
// repeats the idempotent call).  More kernels than slots: the attribute is simply set on every call.
namespace {
struct SmemSlot {
  std::atomic<const void*> func{nullptr};
  std::atomic<uint64_t> bytes_ok[2];   // device bit masks (devices 0..127) configured for >= `bytes`
  std::atomic<int> bytes{0};
};
constexpr int kSmemSlots = 256;
SmemSlot g_smem_slots[kSmemSlots];
}  // namespace

int ensure_dyn_smem(const void* func, int bytes) {
  int dev = 0;
  AUGCL_CUDA(cudaGetDevice(&dev));
  SmemSlot* slot = nullptr;
  for (int i = 0; i < kSmemSlots; ++i) {
    const void* f = g_smem_slots[i].func.load(std::memory_order_acquire);
    if (f == nullptr) {
      const void* expected = nullptr;
      if (g_smem_slots[i].func.compare_exchange_strong(expected, func)) f = func; else f = expected;
    }
    if (f == func) { slot = &g_smem_slots[i]; break; }
  }
  const bool tracked = slot != nullptr && dev >= 0 && dev < 128;
  if (tracked && slot->bytes.load(std::memory_order_acquire) >= bytes &&
      ((slot->bytes_ok[dev >> 6].load(std::memory_order_acquire) >> (dev & 63)) & 1))
    return AUGCL_OK;
  AUGCL_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  if (tracked) {
    if (slot->bytes.load() < bytes) {   // a larger request invalidates what other devices were configured for
      slot->bytes_ok[0].store(0); slot->bytes_ok[1].store(0);
      slot->bytes.store(bytes, std::memory_order_release);
    }
    slot->bytes_ok[dev >> 6].fetch_or(1ull << (dev & 63), std::memory_order_release);
  }
  return AUGCL_OK;
}
}  // namespace augcl

namespace augcl {
// One thread spins on %globaltimer for `ns` nanoseconds (bounded by the host wrapper): occupies the stream so that the
// launches enqueued behind it are all in the queue when it ends (bench.py: the timed region then measures the GPU, not
// the host's launch cadence).  Waits on nothing but the clock.
__global__ void stream_delay_kernel(unsigned long long ns) {
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do {
    __nanosleep(1000);
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  } while (t - t0 < ns);
}
}  // namespace augcl

extern "C" {

int augcl_stream_delay(uint64_t nanoseconds, augcl_stream stream) {
  AUGCL_REQUIRE(nanoseconds <= 50ull * 1000 * 1000, "augcl_stream_delay: at most 50 ms");
  augcl::stream_delay_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((unsigned long long)nanoseconds);
  AUGCL_CUDA(cudaGetLastError());
  return AUGCL_OK;
}

const char* augcl_version(void) { return "augcl_b200 0.1.0 (sm_100a)"; }
const char* augcl_last_error(void) { return augcl::t_error.c_str(); }
int64_t augcl_launch_count(void) { return augcl::g_launches.load(); }

int augcl_set_option(const char* key, const char* value) {
  using namespace augcl;
  AUGCL_REQUIRE(key && value, "augcl_set_option: null key/value");
  if (!strcmp(key, "staging")) {
    if (!strcmp(value, "tma")) { g_staging = 0; return AUGCL_OK; }
    if (!strcmp(value, "generic")) { g_staging = 1; return AUGCL_OK; }
    AUGCL_FAIL(AUGCL_ERR_BAD_ARG, "augcl_set_option: staging must be tma|generic, got '%s'", value);
  }
  if (!strcmp(key, "sample_stages")) {
    int v = atoi(value);
    AUGCL_REQUIRE(v >= 0 && v <= 8, "sample_stages must be in [0,8] (0 = auto)");
    g_sample_stages = v; return AUGCL_OK;
  }
  if (!strcmp(key, "sample_ctas_per_sm")) {
    int v = atoi(value);
    AUGCL_REQUIRE(v >= 1 && v <= 16, "sample_ctas_per_sm must be in [1,16]");
    g_sample_ctas_per_sm = v; return AUGCL_OK;
  }
  if (!strcmp(key, "sample_pdl")) { g_sample_pdl = atoi(value) != 0; return AUGCL_OK; }
  if (!strcmp(key, "sample_fused_min_batch")) { g_sample_fused_min_batch = atoi(value); return AUGCL_OK; }
  if (!strcmp(key, "sample_band")) { g_sample_band = atoi(value) != 0; return AUGCL_OK; }
  if (!strcmp(key, "fused_ctas_per_sm")) {
    int v = atoi(value);
    AUGCL_REQUIRE(v >= 0 && v <= 16, "fused_ctas_per_sm must be in [0,16] (0 = auto)");
    g_fused_ctas_per_sm = v; return AUGCL_OK;
  }
  if (!strcmp(key, "fused_cost_obs") || !strcmp(key, "fused_cost_next") || !strcmp(key, "fused_dynamic") ||
      !strcmp(key, "fused_ahead") || !strcmp(key, "fused_wait_hint"))
    return AUGCL_OK;   // retired switches of earlier schedules: accepted, ignored
  if (!strcmp(key, "fused_stages")) {
    int v = atoi(value);
    AUGCL_REQUIRE(v >= 0 && v <= 4, "fused_stages must be in [0,4] (0 = auto)");
    g_fused_stages = v; return AUGCL_OK;
  }
  if (!strcmp(key, "fused_batch")) {
    const int v = atoi(value);
    AUGCL_REQUIRE(v >= 0 && v <= 4, "fused_batch must be in [0,4] (0 = auto)");
    g_fused_batch = v; return AUGCL_OK;
  }
  if (!strcmp(key, "fused_interleave")) { g_fused_interleave = atoi(value); return AUGCL_OK; }   // -1 = auto
  if (!strcmp(key, "fused_skip")) { g_fused_skip = atoi(value); return AUGCL_OK; }   // timing experiments only
  if (!strcmp(key, "fused_band")) {
    int v = atoi(value);
    AUGCL_REQUIRE(v >= 0 && v <= 14, "fused_band must be in [0,14] (0 = auto)");
    g_fused_band = v; return AUGCL_OK;
  }
  if (!strcmp(key, "conv_part")) { g_conv_part = atoi(value); return AUGCL_OK; }   // 0 off, 1 auto, >= 2 parts
  if (!strcmp(key, "bilinear_tma")) { g_bil_tma = atoi(value) != 0; return AUGCL_OK; }
  if (!strcmp(key, "fused_fast_mask")) { g_fused_fast_mask = atoi(value) != 0; return AUGCL_OK; }
  AUGCL_FAIL(AUGCL_ERR_BAD_ARG, "augcl_set_option: unknown key '%s'", key);
}

}  // extern "C"
