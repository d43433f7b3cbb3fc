// Device-resident replay ring + the fused minibatch gather / u8->f32 / shift|crop kernel.
//
// Replaces the host storage and sampling of utils.ReplayBuffer
// (/root/reference/src/utils.py:139-154 storage, :216-227 add, :206-214 + :234-240 gather/H2D/cast,
//  :242-280 sample*), optionally fused with random_shift / random_crop
// (/root/reference/src/augmentations.py:378-382, 385-413).
//
// HBM layout ("v1"): obs[cap][C][H][W] u8 | next[cap][C][H][W] u8 | act[cap][A] f32 | rew[cap] | nd[cap].
//
// Sample kernel (TMA staging): work item = one u8 plane (H*W bytes, 7 056 B at 84x84) of one
// (sample, tensor).  Persistent CTAs walk items round-robin; an elected thread streams the next
// planes global->shared with cp.async.bulk (1-D TMA, completion on an mbarrier, kStages deep) while
// all threads expand the current plane: each thread builds one float4 of output from 4 source bytes
// (two aligned LDS.32 + funnel shift for the interior, clamped byte loads at the replicate border),
// converts with PRMT+FADD and stores 16 B coalesced.  Algorithmic traffic per sample:
// 2*C*H*W (u8 read) + 2*4*C*Ho*Wo (f32 write) = 635 040 B at C=9, 84x84.
#include "common.cuh"

#include "ring.cuh"

namespace augcl {

struct ItemMeta {
  int lead;  // byte offset of the plane inside its stage buffer (source aligned down to 16 B)
  int dy, dx;
  int pad_;
};

// WQ = Wo/4 as a compile-time constant (0 = runtime) so the row/column split is a multiply-shift.
template <int WQ>
__global__ void __launch_bounds__(256, 6) ring_sample_tma_kernel(const SampleParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int S = p.stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);                       // S mbarriers
  ItemMeta* meta = reinterpret_cast<ItemMeta*>(smem + 64);                   // S metas (S <= 8 -> 128 B)
  uint8_t* bufs = smem + 64 + 128;                                           // S stage buffers, 16-B aligned
  const int tid = threadIdx.x;
  const int P = p.H * p.W;
  const int total = p.B * p.nt * p.C;
  const int wq = WQ ? WQ : (p.Wo >> 2);
  const int nvec = p.Ho * wq;
  const InlineDraws<0> none{};

  if (tid == 0) {
    for (int s = 0; s < S; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncthreads();

  auto issue = [&](int item, int s) {   // tid 0 only
    const int ch = item % p.C;
    const int bt = item / p.C;
    const int slot = bt % p.nt, b = bt / p.nt;
    const int t = p.tmap[slot];
    const int64_t row = sample_index(p, none, b);
    const uint8_t* src = plane_ptr(p, p.vsrc[t], row, ch, (int64_t)P);
    const uintptr_t a = reinterpret_cast<uintptr_t>(src);
    const int lead = (int)(a & 15);
    const uint32_t bytes = (uint32_t)((lead + P + 15) & ~15);
    int dy, dx;
    sample_offset(p, none, t, b, dy, dx);
    meta[s] = ItemMeta{lead, dy, dx, 0};
    mbar_expect_tx(&bars[s], bytes);
    bulk_g2s(bufs + (size_t)s * p.stage_bytes, src - lead, bytes, &bars[s]);
  };

  if (tid == 0) {
    for (int j = 0; j < S; ++j) {
      const int item = blockIdx.x + j * gridDim.x;
      if (item < total) issue(item, j);
    }
  }
  pdl_wait();   // the read-only prologue above may overlap the previous kernel's tail (see ring_sample_band_kernel)
  pdl_launch_dependents();

  gather_scalars(p, none);

  int s = 0;
  uint32_t phase = 0;
  for (int item = blockIdx.x; item < total; item += gridDim.x) {
    mbar_wait(&bars[s], phase);
    const ItemMeta m = meta[s];
    const uint8_t* sb = bufs + (size_t)s * p.stage_bytes;
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(sb);
    const int ch = item % p.C;
    const int bt = item / p.C;
    const int slot = bt % p.nt, b = bt / p.nt;
    float4* outp = reinterpret_cast<float4*>(p.out[p.tmap[slot]] + ((int64_t)b * p.C + ch) * (int64_t)(p.Ho * p.Wo));
    for (int v = tid; v < nvec; v += blockDim.x) {
      const int r = v / wq, q4 = v - r * wq;
      const int rr = clampi(r + m.dy, 0, p.H - 1);
      const int c0 = q4 * 4 + m.dx;
      const int base = m.lead + rr * p.W;
      uint32_t w;
      if (c0 >= 0 && c0 + 3 < p.W) {
        const int a = base + c0;
        w = __funnelshift_r(sw[a >> 2], sw[(a >> 2) + 1], (a & 3) * 8);
      } else {
        const uint32_t b0 = sb[base + clampi(c0, 0, p.W - 1)], b1 = sb[base + clampi(c0 + 1, 0, p.W - 1)];
        const uint32_t b2 = sb[base + clampi(c0 + 2, 0, p.W - 1)], b3 = sb[base + clampi(c0 + 3, 0, p.W - 1)];
        w = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
      }
      outp[v] = finish(w, p.normalize);
    }
    __syncthreads();  // every thread is done reading stage s -> it may be refilled
    if (tid == 0) {
      const int nxt = item + S * gridDim.x;
      if (nxt < total) issue(nxt, s);
    }
    if (++s == S) { s = 0; phase ^= 1; }
  }
}

// ---------------------------------------------------------------------------------------------
// Band kernel: compile-time geometry (H, W source plane; HO, WO output plane).
//   * thread mapping is fixed for the whole kernel: thread = (row r0 = tid / WQ, 4-pixel group q4 = tid % WQ)
//     of a group of RPI = 256 / WQ rows; ACTIVE = RPI * WQ threads work (252 of 256 at WO = 84), and one
//     "unit" of work is one such row group of one plane (NIT = HO / RPI units per plane).
//   * the units of the whole launch are split evenly and contiguously over the CTAs (static partition:
//     every CTA gets total/grid units +-1, so all CTAs finish together even when the launch is only a
//     few planes per CTA, e.g. 2.6 at batch 128).  A CTA walks its range plane-segment by plane-segment;
//     each segment is ONE bulk copy of exactly the source rows it needs.
//   * per output float4: clamp row, 2 x LDS.32, funnel shift, PRMT (replicate-pad selector, constant per
//     segment), 4 x (PRMT + FADD), STG.128 -- no branches, no divisions.
// ---------------------------------------------------------------------------------------------
struct BandMeta {   // written by the issuing lane before the copy is started, read by everyone after it lands
  float4* out;      // output plane
  int koff;         // stage-buffer byte offset of source pixel (row 0, col 0); negative when row_lo > 0
  int dy, dx;
  int it_lo, it_hi;
  int pad_;
};
static_assert(sizeof(BandMeta) == 32, "BandMeta is 32 bytes");

template <int H, int W, int HO, int WO, int NB>
__global__ void __launch_bounds__(256, 6) ring_sample_band_kernel(const SampleParams p,
                                                               const __grid_constant__ InlineDraws<NB> inl) {
  constexpr int WQ = WO / 4, RPI = 256 / WQ, ACTIVE = RPI * WQ, NIT = HO / RPI, P = H * W;
  static_assert(WO % 4 == 0 && HO % RPI == 0, "geometry must tile into row groups");
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);               // <= 8 mbarriers
  BandMeta* meta = reinterpret_cast<BandMeta*>(smem + 64);          // <= 8 metas
  uint8_t* bufs = smem + 64 + 8 * sizeof(BandMeta);
  const int S = p.stages;
  const int tid = threadIdx.x;

  const int64_t total = (int64_t)p.B * p.nt * p.C * NIT;
  const int64_t u0 = total * blockIdx.x / gridDim.x, u1 = total * (blockIdx.x + 1) / gridDim.x;
  const int64_t pl0 = u0 / NIT;
  const int nseg = u1 > u0 ? (int)((u1 - 1) / NIT - pl0) + 1 : 0;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncthreads();

  auto issue = [&](int k, int s) {   // one lane
    const int64_t plane = pl0 + k;
    const int ch = (int)(plane % p.C);
    const int64_t bt = plane / p.C;
    const int slot = (int)(bt % p.nt), b = (int)(bt / p.nt);
    const int t = p.tmap[slot];
    const int64_t row = sample_index(p, inl, b);
    int dy, dx;
    sample_offset(p, inl, t, b, dy, dx);
    const int it_lo = k == 0 ? (int)(u0 - pl0 * NIT) : 0;
    const int64_t rem = u1 - plane * NIT;
    const int it_hi = rem < NIT ? (int)rem : NIT;
    const int row_lo = clampi(it_lo * RPI + dy, 0, H - 1), row_hi = clampi(it_hi * RPI - 1 + dy, 0, H - 1);
    const uint8_t* src = plane_ptr(p, p.vsrc[t], row, ch, (int64_t)P);
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(src + row_lo * W) & ~(uintptr_t)15;
    const uintptr_t a1 = (reinterpret_cast<uintptr_t>(src + (row_hi + 1) * W) + 15) & ~(uintptr_t)15;
    BandMeta m;
    m.out = reinterpret_cast<float4*>(p.out[t] + ((int64_t)b * p.C + ch) * (int64_t)(HO * WO));
    m.koff = (int)(reinterpret_cast<uintptr_t>(src) - a0);
    m.dy = dy; m.dx = dx; m.it_lo = it_lo; m.it_hi = it_hi; m.pad_ = 0;
    meta[s] = m;
    const uint32_t bytes = (uint32_t)(a1 - a0);
    mbar_expect_tx(&bars[s], bytes);
    bulk_g2s(bufs + (size_t)s * p.stage_bytes, reinterpret_cast<const void*>(a0), bytes, &bars[s]);
  };

  if (tid < S && tid < nseg) issue(tid, tid);   // the S prologue copies are prepared by S lanes in parallel
  // Everything above -- the draws, the slot -> address arithmetic, the first S bulk copies -- overlaps the tail of the
  // previous kernel in the stream when this one was launched programmatically (PDL): it only READS the ring, the refs
  // table and the draws, and the only kernels that let a successor start early (griddepcontrol.launch_dependents) are
  // this library's sample kernels, which write nothing but their outputs.  The first STORE waits for the predecessor.
  pdl_wait();
  pdl_launch_dependents();
  gather_scalars(p, inl);

  const int r0 = tid / WQ, q4 = tid - r0 * WQ;
  int s = 0;
  uint32_t phase = 0;
  for (int k = 0; k < nseg; ++k) {
    mbar_wait(&bars[s], phase);
    const BandMeta m = meta[s];
    if (tid < ACTIVE) {
      const uint8_t* sb = bufs + (size_t)s * p.stage_bytes;
      const int c0 = q4 * 4 + m.dx;
      const int cs = clampi(c0, 0, W - 4);
      const uint32_t sel = edge_selector(clampi(c0 - cs, -4, 4));
      const int kbase = m.koff + cs;
      float4* outp = m.out + tid;
#pragma unroll 2
      for (int it = m.it_lo; it < m.it_hi; ++it) {
        const int rr = clampi(it * RPI + r0 + m.dy, 0, H - 1);
        const int a = rr * W + kbase;
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(sb + (a & ~3));
        const uint32_t w = __byte_perm(__funnelshift_r(wp[0], wp[1], (uint32_t)a << 3), 0u, sel);
        st_global_f4(outp + it * ACTIVE, finish(w, p.normalize));
      }
    }
    __syncthreads();  // stage s fully consumed -> refill it
    if (tid == 0 && k + S < nseg) issue(k + S, s);
    if (++s == S) { s = 0; phase ^= 1; }
  }
}

// No staging, any shape: one thread per output element.  Fallback for Wo % 4 != 0 and the
// "generic" arm of the staging A/B switch.
__global__ void __launch_bounds__(256) ring_sample_generic_kernel(const SampleParams p) {
  const InlineDraws<0> none{};
  gather_scalars(p, none);
  const int64_t plane_o = (int64_t)p.Ho * p.Wo;
  const int64_t total = (int64_t)p.B * p.nt * p.C * plane_o;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int q = (int)(i % p.Wo);
    int64_t rest = i / p.Wo;
    const int r = (int)(rest % p.Ho); rest /= p.Ho;
    const int ch = (int)(rest % p.C); rest /= p.C;
    const int slot = (int)(rest % p.nt);
    const int b = (int)(rest / p.nt);
    const int t = p.tmap[slot];
    int dy, dx;
    sample_offset(p, none, t, b, dy, dx);
    const int rr = clampi(r + dy, 0, p.H - 1), cc = clampi(q + dx, 0, p.W - 1);
    const int64_t row = sample_index(p, none, b);
    float v = (float)plane_ptr(p, p.vsrc[t], row, ch, (int64_t)p.H * p.W)[rr * p.W + cc];
    if (p.normalize) v = __fdiv_rn(v, 255.0f);
    p.out[t][(((int64_t)b * p.C + ch) * p.Ho + r) * p.Wo + q] = v;
  }
}

// Synthetic content: byte j of a store = byte (j % 16) of Philox block j / 16 (little-endian words).
__global__ void __launch_bounds__(256) fill_bytes_kernel(uint4* dst, int64_t nblk, uint64_t seed, uint64_t subseq) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nblk; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = philox4x32_10((uint64_t)i, subseq, seed);
}

__global__ void __launch_bounds__(256) fill_refs_window_kernel(int32_t* refs, int64_t count, int k, int64_t pool) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count * 2 * k; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t slot = i / (2 * k);
    const int j = (int)(i - slot * 2 * k), t = j / k, f = j - t * k;
    refs[i] = (int32_t)((slot + t + f) % pool);
  }
}

__global__ void __launch_bounds__(256) fill_scalars_kernel(float* act, float* rew, float* nd, int64_t count, int A,
                                                           uint64_t seed) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    const uint4 r = philox4x32_10((uint64_t)i, 3, seed);
    rew[i] = (float)(r.x >> 8) * (1.0f / 16777216.0f);            // [0,1)
    nd[i] = (r.y & 15u) ? 1.0f : 0.0f;                            // ~6 % terminal transitions
    for (int j = 0; j < A; ++j) {
      const uint4 a = philox4x32_10((uint64_t)(i * A + j), 2, seed);
      act[i * A + j] = (float)(a.x >> 8) * (2.0f / 16777216.0f) - 1.0f;  // [-1,1)
    }
  }
}

}  // namespace augcl

using namespace augcl;

extern "C" {

static int ring_create_impl(const char* who, augcl_ring** out, int device, int64_t capacity, int k, int c_per_frame, int H,
                            int W, int action_dim, int64_t pool_frames) {
  AUGCL_REQUIRE(out, "%s: out is NULL", who);
  *out = nullptr;
  AUGCL_REQUIRE(capacity > 0 && k > 0 && c_per_frame > 0 && H > 0 && W > 0 && action_dim > 0,
                "%s: capacity, k, c_per_frame, H, W, action_dim must be positive", who);
  AUGCL_REQUIRE((int64_t)H * W < (1 << 24), "%s: frame plane too large", who);
  const bool dedup = pool_frames >= 0;
  if (dedup) AUGCL_REQUIRE(pool_frames > k && pool_frames < (1ll << 31), "%s: pool_frames must be in (k, 2^31)", who);
  int ndev = 0;
  AUGCL_CUDA(cudaGetDeviceCount(&ndev));
  AUGCL_REQUIRE(device >= 0 && device < ndev, "%s: device %d out of range (%d devices)", who, device, ndev);
  DeviceGuard g(device);
  AUGCL_CUDA(g.err);
  augcl_ring* r = new augcl_ring();
  r->device = device; r->capacity = capacity; r->k = k; r->cpf = c_per_frame;
  r->C = k * c_per_frame; r->H = H; r->W = W; r->A = action_dim;
  r->stack_bytes = (int64_t)r->C * H * W;
  r->dedup = dedup ? 1 : 0; r->pool = dedup ? pool_frames : 0;
  const int64_t frame_bytes = (int64_t)c_per_frame * H * W;
  const int64_t payload = dedup ? pool_frames * frame_bytes : capacity * r->stack_bytes;
  const int64_t store = ((payload + 255) / 256) * 256 + 256;  // slack: bulk copies round up to 16 B
  const int64_t fbytes = capacity * (int64_t)sizeof(float);
  const int64_t rbytes = dedup ? capacity * 2 * k * (int64_t)sizeof(int32_t) : 0;
  r->bytes = (dedup ? 1 : 2) * store + rbytes + fbytes * (action_dim + 2);
  cudaError_t e = cudaMalloc(&r->obs, store);
  if (e == cudaSuccess && !dedup) e = cudaMalloc(&r->next, store);
  if (e == cudaSuccess && dedup) e = cudaMalloc(&r->refs, rbytes);
  if (e == cudaSuccess && dedup) e = cudaMemset(r->refs, 0, rbytes);     // every slot points at pool frame 0 until written
  if (e == cudaSuccess) e = cudaMalloc(&r->act, fbytes * action_dim);
  if (e == cudaSuccess) e = cudaMalloc(&r->rew, fbytes);
  if (e == cudaSuccess) e = cudaMalloc(&r->nd, fbytes);
  if (e == cudaSuccess) e = cudaMalloc(&r->sched, 64 * kSchedSlots);
  if (e == cudaSuccess) e = cudaMemset(r->sched, 0, 64 * kSchedSlots);
  r->n_sched = 0;
  r->sched_lock.clear();
  if (e == cudaSuccess) {
    cudaMemPoolProps props{};
    props.allocType = cudaMemAllocationTypePinned;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    e = cudaMemPoolCreate(&r->scratch_pool, &props);
    if (e == cudaSuccess) {
      uint64_t keep = UINT64_MAX;
      e = cudaMemPoolSetAttribute(r->scratch_pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
  }
  if (e != cudaSuccess) {
    if (r->scratch_pool) cudaMemPoolDestroy(r->scratch_pool);
    cudaFree(r->sched);
    cudaFree(r->obs); cudaFree(r->next); cudaFree(r->refs); cudaFree(r->act); cudaFree(r->rew); cudaFree(r->nd);
    delete r;
    cudaGetLastError();
    AUGCL_FAIL(AUGCL_ERR_OOM, "%s: cudaMalloc of %lld bytes failed: %s", who, (long long)((dedup ? 1 : 2) * store + rbytes),
               cudaGetErrorString(e));
  }
  *out = r;
  return AUGCL_OK;
}

int augcl_ring_create(augcl_ring** out, int device, int64_t capacity, int k, int c_per_frame, int H, int W,
                      int action_dim) {
  return ring_create_impl("augcl_ring_create", out, device, capacity, k, c_per_frame, H, W, action_dim, -1);
}

int augcl_ring_create_dedup(augcl_ring** out, int device, int64_t capacity, int k, int c_per_frame, int H, int W,
                            int action_dim, int64_t pool_frames) {
  return ring_create_impl("augcl_ring_create_dedup", out, device, capacity, k, c_per_frame, H, W, action_dim, pool_frames);
}

int augcl_ring_destroy(augcl_ring* r) {
  if (!r) return AUGCL_OK;
  DeviceGuard g(r->device);
  cudaFree(r->obs); cudaFree(r->next); cudaFree(r->refs); cudaFree(r->act); cudaFree(r->rew); cudaFree(r->nd);
  cudaFree(r->sched);
  if (r->scratch_pool) cudaMemPoolDestroy(r->scratch_pool);
  delete r;
  return AUGCL_OK;
}

int64_t augcl_ring_bytes(const augcl_ring* r) { return r ? r->bytes : 0; }

static int ring_copy(augcl_ring* r, int64_t slot0, int64_t count, bool put, uint8_t* obs, uint8_t* next, float* act,
                     float* rew, float* nd, cudaStream_t st, const char* who) {
  AUGCL_REQUIRE(r, "%s: ring is NULL", who);
  AUGCL_REQUIRE(slot0 >= 0 && count >= 0 && slot0 + count <= r->capacity,
                "%s: slots [%lld, %lld) outside capacity %lld", who, (long long)slot0, (long long)(slot0 + count),
                (long long)r->capacity);
  if (count == 0) return AUGCL_OK;
  if (r->dedup && (obs || next))
    AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "%s: a de-duplicated ring holds frames, not stacks: use augcl_ring_put_frames / "
               "augcl_ring_put_refs (writes) or augcl_ring_sample (reads)", who);
  DeviceGuard g(r->device);
  AUGCL_CUDA(g.err);
  auto cp = [&](void* ring_ptr, void* user, size_t bytes) -> cudaError_t {
    if (!user) return cudaSuccess;
    return put ? cudaMemcpyAsync(ring_ptr, user, bytes, cudaMemcpyDefault, st)
               : cudaMemcpyAsync(user, ring_ptr, bytes, cudaMemcpyDefault, st);
  };
  if (!r->dedup) {
    AUGCL_CUDA(cp(r->obs + slot0 * r->stack_bytes, obs, (size_t)(count * r->stack_bytes)));
    AUGCL_CUDA(cp(r->next + slot0 * r->stack_bytes, next, (size_t)(count * r->stack_bytes)));
  }
  AUGCL_CUDA(cp(r->act + slot0 * r->A, act, (size_t)(count * r->A) * sizeof(float)));
  AUGCL_CUDA(cp(r->rew + slot0, rew, (size_t)count * sizeof(float)));
  AUGCL_CUDA(cp(r->nd + slot0, nd, (size_t)count * sizeof(float)));
  return AUGCL_OK;
}

int augcl_ring_put(augcl_ring* r, int64_t slot0, int64_t count, const uint8_t* obs, const uint8_t* next_obs,
                   const float* actions, const float* rewards, const float* not_dones, augcl_stream stream) {
  return ring_copy(r, slot0, count, true, const_cast<uint8_t*>(obs), const_cast<uint8_t*>(next_obs),
                   const_cast<float*>(actions), const_cast<float*>(rewards), const_cast<float*>(not_dones),
                   (cudaStream_t)stream, "augcl_ring_put");
}

int augcl_ring_get(augcl_ring* r, int64_t slot0, int64_t count, uint8_t* obs, uint8_t* next_obs, float* actions,
                   float* rewards, float* not_dones, augcl_stream stream) {
  return ring_copy(r, slot0, count, false, obs, next_obs, actions, rewards, not_dones, (cudaStream_t)stream,
                   "augcl_ring_get");
}

int augcl_ring_put_frames(augcl_ring* r, int64_t frame0, int64_t count, const uint8_t* frames, augcl_stream stream) {
  AUGCL_REQUIRE(r && r->dedup, "augcl_ring_put_frames: needs a ring made by augcl_ring_create_dedup");
  AUGCL_REQUIRE(frame0 >= 0 && count >= 0 && frame0 + count <= r->pool, "augcl_ring_put_frames: frames [%lld, %lld) outside the pool of %lld",
                (long long)frame0, (long long)(frame0 + count), (long long)r->pool);
  if (count == 0) return AUGCL_OK;
  AUGCL_REQUIRE(frames, "augcl_ring_put_frames: frames is NULL");
  DeviceGuard g(r->device);
  AUGCL_CUDA(g.err);
  const int64_t fb = (int64_t)r->cpf * r->H * r->W;
  AUGCL_CUDA(cudaMemcpyAsync(r->obs + frame0 * fb, frames, (size_t)(count * fb), cudaMemcpyDefault, (cudaStream_t)stream));
  return AUGCL_OK;
}

int augcl_ring_put_refs(augcl_ring* r, int64_t slot0, int64_t count, const int32_t* refs, const float* actions,
                        const float* rewards, const float* not_dones, augcl_stream stream) {
  AUGCL_REQUIRE(r && r->dedup, "augcl_ring_put_refs: needs a ring made by augcl_ring_create_dedup");
  AUGCL_REQUIRE(slot0 >= 0 && count >= 0 && slot0 + count <= r->capacity, "augcl_ring_put_refs: slots [%lld, %lld) outside capacity %lld",
                (long long)slot0, (long long)(slot0 + count), (long long)r->capacity);
  if (count == 0) return AUGCL_OK;
  DeviceGuard g(r->device);
  AUGCL_CUDA(g.err);
  cudaStream_t st = (cudaStream_t)stream;
  if (refs) {
    cudaPointerAttributes at{};
    const bool host = cudaPointerGetAttributes(&at, refs) != cudaSuccess || at.type == cudaMemoryTypeUnregistered ||
                      at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    if (host)
      for (int64_t e = 0; e < count * 2 * r->k; ++e)
        AUGCL_REQUIRE(refs[e] >= 0 && refs[e] < r->pool, "augcl_ring_put_refs: frame reference %d (element %lld) outside the pool of %lld",
                      refs[e], (long long)e, (long long)r->pool);
    AUGCL_CUDA(cudaMemcpyAsync(r->refs + slot0 * 2 * r->k, refs, (size_t)(count * 2 * r->k) * sizeof(int32_t), cudaMemcpyDefault, st));
  }
  if (actions) AUGCL_CUDA(cudaMemcpyAsync(r->act + slot0 * r->A, actions, (size_t)(count * r->A) * sizeof(float), cudaMemcpyDefault, st));
  if (rewards) AUGCL_CUDA(cudaMemcpyAsync(r->rew + slot0, rewards, (size_t)count * sizeof(float), cudaMemcpyDefault, st));
  if (not_dones) AUGCL_CUDA(cudaMemcpyAsync(r->nd + slot0, not_dones, (size_t)count * sizeof(float), cudaMemcpyDefault, st));
  return AUGCL_OK;
}

int augcl_ring_fill_synthetic(augcl_ring* r, int64_t count, uint64_t seed, augcl_stream stream) {
  AUGCL_REQUIRE(r, "augcl_ring_fill_synthetic: ring is NULL");
  AUGCL_REQUIRE(count >= 0 && count <= r->capacity, "augcl_ring_fill_synthetic: count out of range");
  if (count == 0) return AUGCL_OK;
  DeviceGuard g(r->device);
  AUGCL_CUDA(g.err);
  cudaStream_t st = (cudaStream_t)stream;
  if (r->dedup) {
    // the whole pool gets Philox bytes; slot i references frames i .. i+k-1 (obs) and i+1 .. i+k (next_obs), modulo the
    // pool: the sliding window a FrameStack environment produces
    const int64_t nblk = (r->pool * r->cpf * r->H * r->W + 15) / 16;
    fill_bytes_kernel<<<kNumSMs * 8, 256, 0, st>>>(reinterpret_cast<uint4*>(r->obs), nblk, seed, 0);
    AUGCL_LAUNCHED();
    fill_refs_window_kernel<<<kNumSMs * 2, 256, 0, st>>>(r->refs, count, r->k, r->pool);
    AUGCL_LAUNCHED();
  } else {
    const int64_t nblk = (count * r->stack_bytes + 15) / 16;  // store has >= 256 B of slack
    fill_bytes_kernel<<<kNumSMs * 8, 256, 0, st>>>(reinterpret_cast<uint4*>(r->obs), nblk, seed, 0);
    AUGCL_LAUNCHED();
    fill_bytes_kernel<<<kNumSMs * 8, 256, 0, st>>>(reinterpret_cast<uint4*>(r->next), nblk, seed, 1);
    AUGCL_LAUNCHED();
  }
  fill_scalars_kernel<<<kNumSMs * 2, 256, 0, st>>>(r->act, r->rew, r->nd, count, r->A, seed);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

// Shared implementation of augcl_ring_sample / augcl_ring_sample_ex.
static int sample_impl(const char* who, augcl_ring* r, int idx_src, const int64_t* idx, int64_t idx_hi, int B,
                       float* out_obs, float* out_next, float* out_act, float* out_rew, float* out_nd, int aug_mode,
                       int off_src, const int32_t* off_obs, const int32_t* off_next, int pad_or_size, int normalize,
                       uint64_t seed, uint64_t subsequence, cudaStream_t st, int n_extra = 0,
                       const int* extra_src = nullptr, float* const* extra_out = nullptr,
                       const int32_t* const* extra_off = nullptr) {
  AUGCL_REQUIRE(r, "%s: ring is NULL", who);
  AUGCL_REQUIRE(B > 0, "%s: B must be positive (got %d)", who, B);
  AUGCL_REQUIRE(n_extra >= 0 && n_extra <= kMaxViews - 2, "%s: at most %d extra views", who, kMaxViews - 2);
  const int32_t* off[kMaxViews] = {off_obs, off_next};
  float* outs[kMaxViews] = {out_obs, out_next};
  int vsrc[kMaxViews] = {0, 1};
  bool any_out = out_obs || out_next;
  for (int v = 0; v < n_extra; ++v) {
    AUGCL_REQUIRE(extra_src && extra_out && (extra_src[v] == 0 || extra_src[v] == 1), "%s: bad extra view %d", who, v);
    outs[2 + v] = extra_out[v]; off[2 + v] = extra_off ? extra_off[v] : nullptr; vsrc[2 + v] = extra_src[v];
    any_out = any_out || extra_out[v];
  }
  const bool has_extra = n_extra > 0;
  AUGCL_REQUIRE(any_out, "%s: every image output is NULL", who);
  AUGCL_REQUIRE(aug_mode == AUGCL_AUG_NONE || aug_mode == AUGCL_AUG_SHIFT || aug_mode == AUGCL_AUG_CROP ||
                    aug_mode == AUGCL_AUG_SHIFT_OBS, "%s: unknown aug_mode %d", who, aug_mode);
  const bool obs_only = aug_mode == AUGCL_AUG_SHIFT_OBS;
  if (obs_only) { aug_mode = AUGCL_AUG_SHIFT; off_next = nullptr; off[1] = nullptr; }
  AUGCL_REQUIRE(idx_src == AUGCL_SRC_DEVICE || idx_src == AUGCL_SRC_HOST || idx_src == AUGCL_SRC_PHILOX,
                "%s: unknown idx_src %d", who, idx_src);
  AUGCL_REQUIRE(off_src == AUGCL_SRC_DEVICE || off_src == AUGCL_SRC_HOST || off_src == AUGCL_SRC_PHILOX,
                "%s: unknown off_src %d", who, off_src);
  if (idx_src == AUGCL_SRC_PHILOX)
    AUGCL_REQUIRE(idx_hi > 0 && idx_hi <= r->capacity, "%s: idx_hi %lld outside (0, capacity]", who, (long long)idx_hi);
  else
    AUGCL_REQUIRE(idx, "%s: idx is NULL", who);
  // Larger batches go through the band-of-frames kernel of fused.cu (same outputs bit for bit, same draws): its dynamic
  // unit schedule is 5-8 % faster from batch 256 up (scripts/bench_fused.py drq_shift_both vs this kernel); the plane
  // kernel below keeps the short launches (batch 128: 15.8 vs 18.2 us).
  {
    const int min_b = g_sample_fused_min_batch.load();
    if (min_b > 0 && B >= min_b && aug_mode != AUGCL_AUG_CROP && !has_extra && g_staging.load() == 0 && r->cpf == 3 &&
        (r->W & 3) == 0 && ((r->H * r->W) & 15) == 0 && r->W <= 128 && r->W >= 8 && pad_or_size <= 63 &&
        aligned(out_obs, 16) && aligned(out_next, 16)) {
      augcl_fused_desc d{};
      d.struct_size = (int32_t)sizeof(d);
      d.B = B; d.idx_src = idx_src; d.off_src = off_src; d.idx = idx; d.idx_hi = idx_hi;
      d.pad = aug_mode == AUGCL_AUG_NONE ? 0 : pad_or_size; d.normalize = normalize; d.seed = seed; d.subsequence = subsequence;
      d.out_act = out_act; d.out_rew = out_rew; d.out_nd = out_nd;
      const bool sh_obs = aug_mode != AUGCL_AUG_NONE && (off_obs || off_src == AUGCL_SRC_PHILOX);
      const bool sh_next = aug_mode == AUGCL_AUG_SHIFT && !obs_only && (off_next || off_src == AUGCL_SRC_PHILOX);
      if (sh_obs) { d.out_obs_shift = out_obs; d.off_obs = off_obs; } else d.out_obs = out_obs;
      if (sh_next) { d.out_next_shift = out_next; d.off_next = off_next; } else d.out_next = out_next;
      return augcl_ring_sample_fused(r, &d, (augcl_stream)st);
    }
  }
  SampleParams p{};
  p.store[0] = r->obs; p.store[1] = r->next; p.refs = r->dedup ? r->refs : nullptr; p.k = r->k; p.cpf = r->cpf;
  for (int v = 0; v < kMaxViews; ++v) { p.out[v] = outs[v]; p.vsrc[v] = vsrc[v]; }
  p.idx = idx; p.act = r->act; p.rew = r->rew; p.nd = r->nd;
  p.out_act = out_act; p.out_rew = out_rew; p.out_nd = out_nd;
  p.capacity = r->capacity;
  p.B = B; p.C = r->C; p.H = r->H; p.W = r->W; p.A = r->A;
  p.Ho = r->H; p.Wo = r->W; p.pad = 0; p.normalize = normalize ? 1 : 0;
  p.idx_mode = idx_src; p.idx_hi = (uint64_t)idx_hi; p.seed = seed; p.subseq = subsequence;
  int hi_row = 0, hi_col = 0;   // valid offsets: row in [0, hi_row], col in [0, hi_col]
  if (aug_mode == AUGCL_AUG_SHIFT) {
    AUGCL_REQUIRE(pad_or_size >= 0 && pad_or_size < (1 << 15), "%s: bad pad %d", who, pad_or_size);
    p.pad = pad_or_size;
    hi_row = hi_col = 2 * pad_or_size;
    p.off_range = (uint32_t)(2 * pad_or_size + 1);         // kornia RandomCrop: [0, 2*pad]
  } else if (aug_mode == AUGCL_AUG_CROP) {
    AUGCL_REQUIRE(pad_or_size > 0 && pad_or_size <= r->H && pad_or_size <= r->W,
                  "%s: crop size %d does not fit %dx%d", who, pad_or_size, r->H, r->W);
    p.Ho = p.Wo = pad_or_size;
    // the reference indexes windows[arange, w1, h1] of view_as_windows_cuda, which has H - size (not H - size + 1)
    // window positions (src/augmentations.py:416-433): offset H - size itself raises IndexError there
    hi_row = r->H - pad_or_size - 1; hi_col = r->W - pad_or_size - 1;
    if (hi_row < 0) hi_row = 0;   // size == H: the single window at 0
    if (hi_col < 0) hi_col = 0;
    p.off_range = (uint32_t)(r->W - pad_or_size);           // random_(0, crop_max): upper bound exclusive (:404-405)
    if (off_src == AUGCL_SRC_PHILOX)
      AUGCL_REQUIRE(r->H == r->W && p.off_range > 0, "%s: PHILOX crop offsets need a square image larger than the crop", who);
  }
  for (int t = 0; t < kMaxViews; ++t) {
    if (aug_mode == AUGCL_AUG_NONE || !outs[t] || (obs_only && t == 1)) { p.off_mode[t] = kSrcNone; continue; }
    if (off_src == AUGCL_SRC_PHILOX) { p.off_mode[t] = off[t] ? AUGCL_SRC_DEVICE : AUGCL_SRC_PHILOX; p.off[t] = off[t]; continue; }
    if (!off[t]) {
      AUGCL_REQUIRE(aug_mode != AUGCL_AUG_CROP, "%s: CROP needs offsets for every tensor that is written", who);
      p.off_mode[t] = kSrcNone;     // SHIFT without offsets = identity for that tensor (SVEA: next_obs un-augmented)
      continue;
    }
    p.off_mode[t] = off_src;
    p.off[t] = off[t];
  }
  p.nt = 0;
  for (int v = 0; v < kMaxViews; ++v)
    if (outs[v]) p.tmap[p.nt++] = v;
  AUGCL_REQUIRE((int64_t)B * p.nt * p.C < (1ll << 31), "%s: batch too large", who);
  DeviceGuard g(r->device);
  AUGCL_CUDA(g.err);

  // ---- HOST draws: validate (the reference raises IndexError on a bad index), then carry them by value
  //      in the launch (small batches, band geometries) or stage them through stream-ordered scratch.
  const bool host_idx = idx_src == AUGCL_SRC_HOST;
  bool host_off = false, host_off_extra = false;
  for (int t = 0; t < kMaxViews; ++t)
    if (off_src == AUGCL_SRC_HOST && p.off_mode[t] == AUGCL_SRC_HOST) { host_off = true; host_off_extra = host_off_extra || t >= 3; }
  if (host_idx)
    for (int b = 0; b < B; ++b)
      AUGCL_REQUIRE(idx[b] >= 0 && idx[b] < r->capacity, "%s: index %lld (position %d) out of range [0, %lld)", who,
                    (long long)idx[b], b, (long long)r->capacity);
  if (host_off)
    for (int t = 0; t < kMaxViews; ++t)
      if (p.off_mode[t] == AUGCL_SRC_HOST)
        for (int e = 0; e < 2 * B; ++e)
          AUGCL_REQUIRE(off[t][e] >= 0 && off[t][e] <= ((e & 1) ? hi_col : hi_row),
                        "%s: offset %d (tensor %d, element %d) outside [0, %d]", who, off[t][e], t, e, (e & 1) ? hi_col : hi_row);

  const int kMaxSmem = 227 * 1024;
  const int P = p.H * p.W;
  p.stage_bytes = ((P + 15) & ~15) + 32;
  const int kHeader = 64 + 8 * (int)sizeof(BandMeta);   // mbarriers + per-stage metas (both kernels fit in it)
  bool outs_aligned = true;
  for (int v = 0; v < kMaxViews; ++v) outs_aligned = outs_aligned && aligned(outs[v], 16);
  const bool vec_ok = (p.Wo % 4 == 0) && outs_aligned &&
                      kHeader + p.stage_bytes <= kMaxSmem;
  const bool staged = g_staging.load() == 0 && vec_ok;
  int geom = 0;   // compile-time geometries of the reference's two image sizes (src/arguments.py:155-163)
  if (staged && g_sample_band.load()) {
    if (p.H == 84 && p.W == 84 && p.Ho == 84 && p.Wo == 84) geom = 1;
    else if (p.H == 100 && p.W == 100 && p.Ho == 100 && p.Wo == 100) geom = 2;
    else if (p.H == 100 && p.W == 100 && p.Ho == 84 && p.Wo == 84) geom = 3;
  }

  InlineDraws<kInlineMax> inl;
  bool use_inline = false;
  void* scratch = nullptr;
  if (host_idx || host_off) {
    if (geom && B <= kInlineMax && r->capacity < (1ll << 31) && hi_row <= 127 && hi_col <= 127 && !host_off_extra) {
      use_inline = true;
      if (host_idx) {
        for (int b = 0; b < B; ++b) inl.idx[b] = (int32_t)idx[b];
        p.idx_mode = kSrcInline;
      }
      for (int t = 0; t < 3; ++t)
        if (p.off_mode[t] == AUGCL_SRC_HOST) {
          for (int b = 0; b < B; ++b) { inl.off[t][b][0] = (int8_t)off[t][2 * b]; inl.off[t][b][1] = (int8_t)off[t][2 * b + 1]; }
          p.off_mode[t] = kSrcInline;
        }
    } else {
      // stream-ordered scratch: [idx int64 B][offsets of view 0 int32 2B][view 1] ...
      const size_t ib = (size_t)B * 8, ob = (size_t)B * 8;
      AUGCL_CUDA(cudaMallocFromPoolAsync(&scratch, ib + kMaxViews * ob, r->scratch_pool, st));
      uint8_t* sp = static_cast<uint8_t*>(scratch);
      cudaError_t e = cudaSuccess;
      if (host_idx) {
        e = cudaMemcpyAsync(sp, idx, ib, cudaMemcpyHostToDevice, st);
        p.idx = reinterpret_cast<const int64_t*>(sp);
        p.idx_mode = AUGCL_SRC_DEVICE;
      }
      for (int t = 0; t < kMaxViews && e == cudaSuccess; ++t)
        if (p.off_mode[t] == AUGCL_SRC_HOST) {
          e = cudaMemcpyAsync(sp + ib + t * ob, off[t], ob, cudaMemcpyHostToDevice, st);
          p.off[t] = reinterpret_cast<const int32_t*>(sp + ib + t * ob);
          p.off_mode[t] = AUGCL_SRC_DEVICE;
        }
      if (e != cudaSuccess) {
        cudaFreeAsync(scratch, st);
        AUGCL_FAIL(AUGCL_ERR_CUDA, "%s: staging the host draws failed: %s", who, cudaGetErrorString(e));
      }
    }
  }
  auto finish_call = [&](int rc) -> int {
    if (scratch) cudaFreeAsync(scratch, st);   // stream-ordered: released after the kernel that reads it
    return rc;
  };

  if (staged) {
    int stages = g_sample_stages.load();
    if (stages == 0) stages = (int64_t)B * p.nt * p.C <= 8 * kNumSMs * 6 ? 4 : 2;   // short launches: deeper prefetch hides the ramp
    int ctas_per_sm = g_sample_ctas_per_sm.load();
    while (stages > 1 && kHeader + stages * p.stage_bytes > kMaxSmem) --stages;
    const int smem = kHeader + stages * p.stage_bytes;
    p.stages = stages;
    const int fit = (228 * 1024) / (smem + 1024);
    if (ctas_per_sm > fit) ctas_per_sm = fit < 1 ? 1 : fit;
    if (ctas_per_sm > 8) ctas_per_sm = 8;   // 2048 threads / 256
    const bool pdl = g_sample_pdl.load() != 0;
    auto launch = [&](auto kern, int64_t work_items, auto... args) -> int {
      // per (kernel address, device): a static flag here would be shared by every instantiation with the same signature
      if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kern), kMaxSmem)) return rc;
      int64_t grid = (int64_t)kNumSMs * ctas_per_sm;
      if (grid > work_items) grid = work_items;
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3((unsigned)grid);
      cfg.blockDim = dim3(256);
      cfg.dynamicSmemBytes = (size_t)smem;
      cfg.stream = st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = pdl ? 1 : 0;
      AUGCL_CUDA(cudaLaunchKernelEx(&cfg, kern, args...));
      AUGCL_LAUNCHED();
      return AUGCL_OK;
    };
    const int64_t planes = (int64_t)B * p.nt * p.C;
    const InlineDraws<0> none{};
    if (use_inline) {
      if (geom == 1) return launch(ring_sample_band_kernel<84, 84, 84, 84, kInlineMax>, planes * 7, p, inl);
      if (geom == 2) return launch(ring_sample_band_kernel<100, 100, 100, 100, kInlineMax>, planes * 10, p, inl);
      return launch(ring_sample_band_kernel<100, 100, 84, 84, kInlineMax>, planes * 7, p, inl);
    }
    if (geom == 1) return finish_call(launch(ring_sample_band_kernel<84, 84, 84, 84, 0>, planes * 7, p, none));
    if (geom == 2) return finish_call(launch(ring_sample_band_kernel<100, 100, 100, 100, 0>, planes * 10, p, none));
    if (geom == 3) return finish_call(launch(ring_sample_band_kernel<100, 100, 84, 84, 0>, planes * 7, p, none));
    if (p.Wo == 84) return finish_call(launch(ring_sample_tma_kernel<21>, planes, p));
    if (p.Wo == 100) return finish_call(launch(ring_sample_tma_kernel<25>, planes, p));
    return finish_call(launch(ring_sample_tma_kernel<0>, planes, p));
  }
  const int64_t total = (int64_t)B * p.nt * p.C * p.Ho * p.Wo;
  int64_t blocks = (total + 255) / 256;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  ring_sample_generic_kernel<<<(int)blocks, 256, 0, st>>>(p);
  cudaError_t le = cudaGetLastError();
  if (le != cudaSuccess) {
    finish_call(0);
    AUGCL_FAIL(AUGCL_ERR_CUDA, "%s: launch failed: %s", who, cudaGetErrorString(le));
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return finish_call(AUGCL_OK);
}

int augcl_ring_sample(augcl_ring* r, const int64_t* idx, int B, float* out_obs, float* out_next, float* out_act,
                      float* out_rew, float* out_nd, int aug_mode, const int32_t* off_obs, const int32_t* off_next,
                      int pad_or_size, int normalize, augcl_stream stream) {
  return sample_impl("augcl_ring_sample", r, AUGCL_SRC_DEVICE, idx, 0, B, out_obs, out_next, out_act, out_rew, out_nd,
                     aug_mode, AUGCL_SRC_DEVICE, off_obs, off_next, pad_or_size, normalize, 0, 0, (cudaStream_t)stream);
}

int augcl_ring_sample_ex(augcl_ring* r, int idx_src, const int64_t* idx, int64_t idx_hi, int B, float* out_obs,
                         float* out_next, float* out_act, float* out_rew, float* out_nd, int aug_mode, int off_src,
                         const int32_t* off_obs, const int32_t* off_next, int pad_or_size, int normalize, uint64_t seed,
                         uint64_t subsequence, augcl_stream stream) {
  return sample_impl("augcl_ring_sample_ex", r, idx_src, idx, idx_hi, B, out_obs, out_next, out_act, out_rew, out_nd,
                     aug_mode, off_src, off_obs, off_next, pad_or_size, normalize, seed, subsequence, (cudaStream_t)stream);
}

int augcl_ring_sample_curl(augcl_ring* r, int idx_src, const int64_t* idx, int64_t idx_hi, int B, float* out_obs,
                           float* out_next, float* out_pos, float* out_act, float* out_rew, float* out_nd, int off_src,
                           const int32_t* off_obs, const int32_t* off_next, const int32_t* off_pos, int size, int normalize,
                           uint64_t seed, uint64_t subsequence, augcl_stream stream) {
  const int src[1] = {0};
  float* const outs[1] = {out_pos};
  const int32_t* const offs[1] = {off_pos};
  return sample_impl("augcl_ring_sample_curl", r, idx_src, idx, idx_hi, B, out_obs, out_next, out_act, out_rew, out_nd,
                     AUGCL_AUG_CROP, off_src, off_obs, off_next, size, normalize, seed, subsequence, (cudaStream_t)stream,
                     out_pos ? 1 : 0, src, outs, offs);
}

int augcl_ring_sample_views(augcl_ring* r, int idx_src, const int64_t* idx, int64_t idx_hi, int B, int n_views,
                            const int* view_src, float* const* outs, const int32_t* const* offs, float* out_act,
                            float* out_rew, float* out_nd, int aug_mode, int off_src, int pad_or_size, int normalize,
                            uint64_t seed, uint64_t subsequence, augcl_stream stream) {
  const char* who = "augcl_ring_sample_views";
  AUGCL_REQUIRE(n_views >= 1 && n_views <= kMaxViews - 2, "%s: n_views must be in [1, %d]", who, kMaxViews - 2);
  AUGCL_REQUIRE(view_src && outs, "%s: view_src/outs is NULL", who);
  AUGCL_REQUIRE(aug_mode == AUGCL_AUG_NONE || aug_mode == AUGCL_AUG_SHIFT || aug_mode == AUGCL_AUG_CROP,
                "%s: aug_mode must be NONE, SHIFT or CROP", who);
  // every requested view is an "extra" view (2 ..): its Philox offsets come from subsequence + 3 + v
  return sample_impl(who, r, idx_src, idx, idx_hi, B, nullptr, nullptr, out_act, out_rew, out_nd, aug_mode, off_src, nullptr,
                     nullptr, pad_or_size, normalize, seed, subsequence, (cudaStream_t)stream, n_views, view_src, outs, offs);
}

// ------------------------------------------------------------------------------------------------
// augcl_ring_sample_noalloc: augcl_ring_sample_ex for HOST draws without any allocation (no counterpart in the
// reference).  augcl_ring_sample_ex stages host indices / offsets of
// batches beyond the inline limit through stream-ordered scratch (cudaMallocFromPoolAsync above); here that scratch is
// a file-scope device array of fixed size, so the call is usable where allocation is forbidden.
// Layout: [idx int64 kNoAllocMaxBatch][obs offsets int32 2*kNoAllocMaxBatch][next_obs offsets int32 2*kNoAllocMaxBatch].
// ------------------------------------------------------------------------------------------------
constexpr int kNoAllocMaxBatch = AUGCL_NOALLOC_MAX_BATCH;
static __device__ __align__(256) unsigned char g_noalloc_scratch[(size_t)kNoAllocMaxBatch * 24];

// Device address of the array on the CURRENT device (the symbol's name is only a host-side stand-in): resolved once
// per device with that device current, remembered afterwards.
static int noalloc_scratch_ptr(int device, unsigned char** out) {
  static std::atomic<unsigned char*> cache[64];
  AUGCL_REQUIRE(device >= 0 && device < 64, "augcl_ring_sample_noalloc: device %d outside [0, 64)", device);
  unsigned char* ptr = cache[device].load(std::memory_order_acquire);
  if (!ptr) {
    void* sym = nullptr;
    AUGCL_CUDA(cudaGetSymbolAddress(&sym, g_noalloc_scratch));
    ptr = static_cast<unsigned char*>(sym);
    cache[device].store(ptr, std::memory_order_release);
  }
  *out = ptr;
  return AUGCL_OK;
}

int augcl_ring_sample_noalloc(augcl_ring* r, const int64_t* idx, int B, float* out_obs, float* out_next, float* out_act,
                              float* out_rew, float* out_nd, int aug_mode, const int32_t* off_obs, const int32_t* off_next,
                              int pad_or_size, int normalize, augcl_stream stream) {
  const char* who = "augcl_ring_sample_noalloc";
  AUGCL_REQUIRE(r, "%s: ring is NULL", who);
  AUGCL_REQUIRE(idx, "%s: idx is NULL", who);
  AUGCL_REQUIRE(B > 0, "%s: B must be positive (got %d)", who, B);
  if (B > kNoAllocMaxBatch)
    AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "%s: batch %d exceeds the fixed scratch (AUGCL_NOALLOC_MAX_BATCH = %d)", who, B,
               kNoAllocMaxBatch);
  AUGCL_REQUIRE(aug_mode == AUGCL_AUG_NONE || aug_mode == AUGCL_AUG_SHIFT || aug_mode == AUGCL_AUG_SHIFT_OBS,
                "%s: aug_mode must be NONE, SHIFT or SHIFT_OBS (got %d)", who, aug_mode);
  // the host draws are validated like augcl_ring_sample_ex's (the reference raises IndexError)
  for (int b = 0; b < B; ++b)
    AUGCL_REQUIRE(idx[b] >= 0 && idx[b] < r->capacity, "%s: index %lld (position %d) out of range [0, %lld)", who,
                  (long long)idx[b], b, (long long)r->capacity);
  const int32_t* off[2] = {off_obs, aug_mode == AUGCL_AUG_SHIFT ? off_next : nullptr};
  if (aug_mode != AUGCL_AUG_NONE) {
    AUGCL_REQUIRE(pad_or_size >= 0 && pad_or_size <= 63, "%s: pad %d outside [0, 63]", who, pad_or_size);
    for (int t = 0; t < 2; ++t)
      if (off[t])
        for (int e = 0; e < 2 * B; ++e)
          AUGCL_REQUIRE(off[t][e] >= 0 && off[t][e] <= 2 * pad_or_size, "%s: offset %d (tensor %d, element %d) outside [0, %d]",
                        who, off[t][e], t, e, 2 * pad_or_size);
  }
  DeviceGuard g(r->device);
  AUGCL_CUDA(g.err);
  unsigned char* sp = nullptr;
  if (int rc = noalloc_scratch_ptr(r->device, &sp)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t ib = (size_t)kNoAllocMaxBatch * 8;
  // every byte the kernel reads is rewritten here: nothing in the array has to start at zero
  AUGCL_CUDA(cudaMemcpyAsync(sp, idx, (size_t)B * 8, cudaMemcpyHostToDevice, st));
  const int32_t* d_off[2] = {nullptr, nullptr};
  if (aug_mode != AUGCL_AUG_NONE)
    for (int t = 0; t < 2; ++t)
      if (off[t]) {
        unsigned char* dst = sp + ib + (size_t)t * kNoAllocMaxBatch * 8;
        AUGCL_CUDA(cudaMemcpyAsync(dst, off[t], (size_t)B * 8, cudaMemcpyHostToDevice, st));
        d_off[t] = reinterpret_cast<const int32_t*>(dst);
      }
  // DEVICE draws from here on: the same kernels as augcl_ring_sample, no scratch, no allocation
  return sample_impl(who, r, AUGCL_SRC_DEVICE, reinterpret_cast<const int64_t*>(sp), 0, B, out_obs, out_next, out_act, out_rew,
                     out_nd, aug_mode, AUGCL_SRC_DEVICE, d_off[0], d_off[1], pad_or_size, normalize, 0, 0, st);
}

}  // extern "C"
