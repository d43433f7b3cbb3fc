// Shared between ring.cu and fused.cu: the ring handle, the draw sources (device / host-inline / Philox)
// and the scalar gather of the sample kernels.
#pragma once
#include "common.cuh"

namespace augcl {
extern std::atomic<int> g_sample_stages;
extern std::atomic<int> g_sample_ctas_per_sm;
extern std::atomic<int> g_sample_pdl;
extern std::atomic<int> g_sample_band;
extern std::atomic<int> g_sample_fused_min_batch;
}

struct augcl_ring {
  int device;
  int64_t capacity;
  int k, cpf, C, H, W, A;
  int64_t stack_bytes;  // C*H*W
  uint8_t* obs;
  uint8_t* next;
  float* act;
  float* rew;
  float* nd;
  int64_t bytes;
  // "v2" (de-duplicated) layout: obs = frame pool [pool][cpf][H][W] u8, next = NULL,
  // refs[capacity][2][k] = pool slot of every frame of (obs, next_obs)
  uint32_t* sched;   // kSchedSlots x {ticket, done, ..} (64 B each): counters of the fused sampler's unit schedule, one set per
                     // stream the ring is sampled from (zero between launches); sched_streams[i] owns slot i
  cudaStream_t sched_streams[16];
  int n_sched;
  std::atomic_flag sched_lock;
  cudaMemPool_t scratch_pool;   // stream-ordered scratch for staged HOST draws; never trimmed (the default pool gives its
                                // memory back at every synchronize: ~0.5 ms per cudaMallocAsync + ~0.8 ms per sync after)
  int dedup;
  int64_t pool;
  int32_t* refs;
};

namespace augcl {

constexpr int kSchedSlots = 16;
// Counter set of `stream` on this ring (launches of one stream are ordered, so they can share one; two streams must not).
// Returns nullptr when more than kSchedSlots distinct streams have sampled the ring.
inline uint32_t* sched_for(augcl_ring* r, cudaStream_t stream) {
  while (r->sched_lock.test_and_set(std::memory_order_acquire)) {}
  int slot = -1;
  for (int i = 0; i < r->n_sched; ++i)
    if (r->sched_streams[i] == stream) { slot = i; break; }
  if (slot < 0 && r->n_sched < kSchedSlots) { slot = r->n_sched; r->sched_streams[r->n_sched++] = stream; }
  r->sched_lock.clear(std::memory_order_release);
  return slot < 0 ? nullptr : r->sched + 16 * slot;
}

constexpr int kMaxViews = 8;

struct SampleParams {
  const uint8_t* store[2];   // v1: obs / next_obs stacks; v2: store[0] = frame pool
  const int32_t* refs;       // v2: [capacity][2][k] pool slots (NULL = v1)
  int k, cpf;
  // output "views": 0 = obs, 1 = next_obs, 2.. = further independently shifted / cropped copies of either tensor
  // (sample_curl's `pos`; the K + M augmented copies of a DrQ update)
  float* out[kMaxViews];
  const int32_t* off[kMaxViews];
  int vsrc[kMaxViews];       // ring tensor a view reads (0 obs, 1 next_obs)
  const int64_t* idx;
  const float* act;
  const float* rew;
  const float* nd;
  float* out_act;
  float* out_rew;
  float* out_nd;
  int64_t capacity;
  int B, C, H, W, Ho, Wo, A;
  int pad;        // SHIFT: pad; NONE/CROP: 0.  source coord = clamp(out coord + off - pad)
  int normalize;
  int nt;         // number of tensors written (1 or 2); slot s -> tensor tmap[s]
  int tmap[kMaxViews];
  int stage_bytes;
  int stages;
  // where the draws come from (AUGCL_SRC_*; kSrcInline = by value in the launch's parameter block, kSrcNone = identity)
  int idx_mode;
  int off_mode[kMaxViews];
  uint32_t off_range;      // PHILOX offsets are uniform in [0, off_range)
  uint64_t idx_hi;         // PHILOX indices are uniform in [0, idx_hi)
  uint64_t seed, subseq;   // PHILOX streams: indices = subseq, obs offsets = subseq + 1, next_obs offsets = subseq + 2
};

constexpr int kSrcInline = 3;
constexpr int kSrcNone = 4;
constexpr int kInlineMax = 256;   // the reference's batch sizes: 128 (src/arguments.py:30) and 256 (soda_batch_size, :60)

// Draws carried BY VALUE in the kernel's parameter block (constant bank): a host caller's indices and
// offsets reach the GPU inside the launch itself -- no separate H2D copy, no device scratch.
template <int NB>
struct InlineDraws {
  int32_t idx[NB];
  int8_t off[3][NB][2];
};
template <>
struct InlineDraws<0> {};

__device__ __forceinline__ int64_t clamp_idx(int64_t i, int64_t cap) {
  return i < 0 ? 0 : (i >= cap ? cap - 1 : i);
}

__device__ __forceinline__ uint32_t philox_word(uint64_t e, uint64_t subseq, uint64_t seed) {
  const uint4 r = philox4x32_10(e >> 2, subseq, seed);
  const uint32_t k = (uint32_t)e & 3u;
  return k == 0 ? r.x : (k == 1 ? r.y : (k == 2 ? r.z : r.w));
}

// Ring slot of sample b (always clamped into [0, capacity) for memory safety).
template <class Inl>
__device__ __forceinline__ int64_t sample_index(const SampleParams& p, const Inl& inl, int b) {
  int64_t i;
  if constexpr (sizeof(Inl) > 1) {
    if (p.idx_mode == kSrcInline) return clamp_idx((int64_t)inl.idx[b], p.capacity);
  }
  if (p.idx_mode == AUGCL_SRC_PHILOX)   // same values as augcl_rng_fill_indices(out, B, 0, idx_hi, seed, subseq)
    i = (int64_t)(((uint64_t)philox_word((uint64_t)b, p.subseq, p.seed) * p.idx_hi) >> 32);
  else
    i = p.idx[b];
  return clamp_idx(i, p.capacity);
}

// (dy, dx) of sample b for tensor t: source coordinate = output coordinate + (dy, dx), clamped.
template <class Inl>
__device__ __forceinline__ void sample_offset(const SampleParams& p, const Inl& inl, int t, int b, int& dy, int& dx) {
  const int m = p.off_mode[t];
  if (m == kSrcNone) { dy = 0; dx = 0; return; }
  int oy = 0, ox = 0;
  if (m == AUGCL_SRC_PHILOX) {          // same values as augcl_rng_fill_offsets(out, 2B, 0, off_range, seed, subseq+1+t)
    const uint4 r = philox4x32_10((uint64_t)(b >> 1), p.subseq + 1 + t, p.seed);
    oy = (int)__umulhi((b & 1) ? r.z : r.x, p.off_range);
    ox = (int)__umulhi((b & 1) ? r.w : r.y, p.off_range);
  } else {
    bool done = false;
    if constexpr (sizeof(Inl) > 1) {
      if (m == kSrcInline) { oy = inl.off[t][b][0]; ox = inl.off[t][b][1]; done = true; }
    }
    if (!done) { oy = p.off[t][2 * b]; ox = p.off[t][2 * b + 1]; }
  }
  dy = oy - p.pad; dx = ox - p.pad;
}

// actions / rewards / not_dones rows (src/utils.py:210-212): a few hundred floats, done by a grid-stride loop over
// `nthreads` threads of the grid (`first` = this thread's index among them)
template <class Inl>
__device__ __forceinline__ void gather_scalars_part(const SampleParams& p, const Inl& inl, int first, int nthreads) {
  const int per = p.A + 2;
  for (int i = first; i < p.B * per; i += nthreads) {
    const int b = i / per, j = i - b * per;
    const int64_t s = sample_index(p, inl, b);
    if (j < p.A) {
      if (p.out_act) p.out_act[(int64_t)b * p.A + j] = p.act[s * p.A + j];
    } else if (j == p.A) {
      if (p.out_rew) p.out_rew[b] = p.rew[s];
    } else {
      if (p.out_nd) p.out_nd[b] = p.nd[s];
    }
  }
}
template <class Inl>
__device__ __forceinline__ void gather_scalars(const SampleParams& p, const Inl& inl) {
  gather_scalars_part(p, inl, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}

// First byte of u8 plane `ch` of tensor t (0 obs, 1 next_obs) of ring slot `row`; P = H*W.
__device__ __forceinline__ const uint8_t* plane_ptr(const SampleParams& p, int t, int64_t row, int ch, int64_t P) {
  if (p.refs) {   // v2: one dependent 4-byte load (L2-resident table) resolves the frame
    const int f = ch / p.cpf;
    const int64_t fs = p.refs[(row * 2 + t) * p.k + f];
    return p.store[0] + (fs * p.cpf + (ch - f * p.cpf)) * P;
  }
  return p.store[t] + (row * p.C + ch) * P;
}

__device__ __forceinline__ float4 finish(uint32_t w, int normalize) {
  float4 o = bytes_to_float4(w);
  if (normalize) {
    o.x = __fdiv_rn(o.x, 255.0f); o.y = __fdiv_rn(o.y, 255.0f);
    o.z = __fdiv_rn(o.z, 255.0f); o.w = __fdiv_rn(o.w, 255.0f);
  }
  return o;
}

}  // namespace augcl
