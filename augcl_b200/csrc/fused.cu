// Fused per-algorithm sampler: gather + u8->f32 + random_shift + {splice | CS_splice | overlay} + random_conv
// in ONE kernel, straight out of the device ring (include/augcl_b200.h: augcl_ring_sample_fused).
//
// Replaces, per update() of SVEA / AugCL (/root/reference/src/algorithms/svea.py:66-70, augcl.py:69-72), the chain
//   ReplayBuffer.sample (src/utils.py:272-280) -> random_shift (src/augmentations.py:378-382)
//   -> RAD.apply_aug (src/algorithms/rad.py:18-22) over splice / random_overlay / random_conv (:253-324)
// which in the unfused form costs four kernels and 2.2 MB of HBM traffic per sample; fused, every u8 frame is read
// once and only the final f32 tensors are written (952 560 B per sample for AugCL's splice-conv at 9x84x84).
//
// Work unit = a band of kFBand rows of one RGB frame (3 u8 planes) of one (tensor, sample).  The units of a launch
// are laid out obs-first and split over the persistent CTAs by COST (an obs unit writes several tensors and may
// convolve; a next_obs unit is a plain copy), contiguously, so a CTA walks consecutive bands of the same frame and
// re-reads its halo rows out of L2.  Per unit, one lane starts <= 6 bulk copies (cp.async.bulk, SASS UBLKCP): the
// hull of source rows of the 3 planes and, for splice / overlay, the matching rows of the 3 background planes;
// they land in a 2-deep stage ring guarded by mbarriers while the CTA works on the previous unit.
//   phase A (elementwise): thread = 4 pixels x 3 channels of one tile row.  Plain word = 1 LDS.32; shifted word =
//     2 LDS.32 + funnel shift + PRMT replicate selector (as in ring_sample_band_kernel); HSV mask / overlay in
//     registers; direct outputs are STG.128; the conv input goes to an f32 tile in shared memory.
//   phase B (conv): packed FFMA2 (fma.rn.f32x2) on ROW PAIRS (a, a+6) of the 12-row band.  The tile stores tile rows
//     y and y+6 interleaved -- element (yp, c) is the float2 (x[yp][c], x[yp+6][c]) -- so every FFMA2 operand is an
//     aligned register pair straight out of an LDS.128 (no re-packing MOVs; the weight is the scalar-broadcast
//     operand form `R.F32`), thread = 4 pixels x 2 rows x 3 output channels = 12 float2 accumulators: per input
//     channel 108 FFMA2 + 18 LDS.128 (the scalar version: 216 FFMA + 21 LDS with 4-way conflicts on its edge loads).
#include "pixel.cuh"
#include "ring.cuh"

namespace augcl {

std::atomic<int> g_fused_ctas_per_sm{0};   // 0 = auto
std::atomic<int> g_fused_cost_obs{0};    // 0 = auto
std::atomic<int> g_fused_cost_next{0};
std::atomic<int> g_fused_stages{0};      // 0 = auto
std::atomic<int> g_fused_dynamic{1};     // 1: atomic-ticket unit schedule, 0: static cost partition
std::atomic<int> g_fused_band{0};        // 0 = auto (non-conv chains: 14 or 7 rows)
std::atomic<int> g_fused_fast_mask{1};   // 0: always the float HSV path (A/B and the exhaustive equivalence test)

constexpr int kFHeader = 256 + 432 + 16;   // mbarriers + metas + conv weights + per-stage "holds a unit" flags
constexpr int kFBand = 12;        // conv chains: 12 rows + 2 halo rows = 14 tile rows = two rounds of 7 rows
constexpr int kFBandNoConv = 14;  // every other chain: 14 rows = two full rounds, 84 = 6 bands
constexpr int kFThreads = 160;   // 147 = 7 rows x 21 quads work at W = 84: a 14-row tile is exactly two rounds

struct FusedParams {
  SampleParams sp;          // stores, draw sources, scalar gather (sp.out unused)
  float* o_plain[2];        // [obs, next]
  float* o_shift[2];
  float* o_aug;
  const uint8_t* bank;
  const int64_t* bank_idx;
  uint64_t bank_subseq;     // bank_idx == nullptr: rows drawn here, same values as augcl_rng_fill_indices
  const float* weights;
  int64_t bank_len;
  SpliceThr thr;
  uint32_t thr_k, thr_c7;   // kPreSpliceFast: the byte threshold t (smallest b with b/255 > thr.val) as SWAR constants
  float oma, alpha;
  int ops;                  // AUGCL_OP_*
  int aug_shifted;          // chain input: 0 plain, 1 shifted
  int k, bands, band;       // frames per stack, bands per frame, rows per band
  int64_t units_obs, units_next;
  int cost_obs, cost_next;
  int plane_stage, bank_stage, stage_bytes, tile_rs, stages;
  uint32_t* sched;          // dynamic schedule: {next ticket, CTAs done}; NULL = static cost partition
  int normalize;
};

struct FusedMeta {
  float* o_plain;   // plane 0 of this frame in each output (nullptr = not written by this unit)
  float* o_shift;
  float* o_aug;
  int koff;         // byte offset of source pixel (row 0, col 0) inside a plane's stage buffer (may be negative)
  int gkoff;        // same for the background planes
  int dy, dx;
  int r0, rows;     // band
  int sample;
  int pad_;
};

template <bool NORM>
__device__ __forceinline__ float fin(float v) { return NORM ? div255(v) : v; }   // div255 == IEEE v / 255 (pixel.cuh)
template <bool NORM>
__device__ __forceinline__ float4 fin4(float4 v) {
  return make_float4(fin<NORM>(v.x), fin<NORM>(v.y), fin<NORM>(v.z), fin<NORM>(v.w));
}

// chain stage before the conv
enum { kPreNone = 0, kPreSplice = 1, kPreSpliceFast = 2, kPreOverlay = 3 };

// ---- conv tile: [3 ci][kFPairs row pairs][TROW floats].  Row pair yp holds tile rows yp (lane .x) and yp + kFHalf
// (lane .y) of the (kFBand + 2)-row band; tile row t = band row t - 1 (halo).  Columns c' = c + 1 (c = -1 .. W) as float2,
// split into two arrays of 16-byte chunks so that the three LDS.128 of lane q4 (c' = 4q4 .. 4q4+5) are 16 B apart
// between lanes (conflict-free):  A0[q] = {T[4q], T[4q+1]} at float 4q (q = 0 .. nq),  A1[q] = {T[4q+2], T[4q+3]} at
// float 4(nq+1) + 4q.  A thread reads A0[q4], A1[q4], A0[q4+1].
constexpr int kFHalf = kFBand / 2;           // output rows a and a + kFHalf share a thread
constexpr int kFPairs = kFHalf + 2;        // row pairs per channel
__host__ __device__ constexpr int tile_row_floats(int W) { return (2 * (W >> 2) + 1) * 4; }

// conv weights are stored pre-multiplied by -log2(e) / 255: acc = -log2(e) * conv(x / 255), so
// sigmoid(conv) * 255 = 255 / (1 + 2^acc) = rcp(fma(2^acc, 1/255, 1/255)): MUFU.EX2, FFMA, MUFU.RCP.
constexpr float kConvWScale = -1.4426950408889634f / 255.0f;
template <bool NORM>
__device__ __forceinline__ float conv_sigmoid(float acc) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(acc));
  const float d = NORM ? 1.0f + e : fmaf(e, 0.00392156862745098f, 0.00392156862745098f);
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
  return r;
}

// Default-threshold splice mask on four u8 pixels at once (one 32-bit word per channel), SWAR on the byte MSBs:
//   keep = max > min  and not (r == max and g == b)  and  max >= t
//        = (g != b  or  r < g)  and  (r >= t or g >= t or b >= t)          (g == b: r == max <=> r >= g)
// K, C7 encode the threshold t in 1 .. 255: c = 256 - t, K = (c & 0x7f) * 0x01010101, C7 = c & 0x80 ? 0x80808080 : 0;
// x >= t <=> the 8-bit add x + c carries out = maj(x7, c7, carry into bit 7).  Returns 0xff in the bytes to keep.
__device__ __forceinline__ uint32_t splice_keep_bytes(uint32_t r, uint32_t g, uint32_t b, uint32_t K, uint32_t C7) {
  const uint32_t L = 0x7f7f7f7fu, H = 0x80808080u;
  const uint32_t x = g ^ b;
  const uint32_t ne = ((x & L) + L) | x;                       // bit 7: g != b
  const uint32_t d = (r | H) - (g & L);                        // bit 7: (r & 0x7f) >= (g & 0x7f); no borrow between bytes
  const uint32_t lt = (~r & g) | (~(r ^ g) & ~d);              // bit 7: r < g
  const uint32_t sr = (r & L) + K, sg = (g & L) + K, sb = (b & L) + K;
  const uint32_t ge = ((r & sr) | (C7 & (r | sr))) | ((g & sg) | (C7 & (g | sg))) | ((b & sb) | (C7 & (b | sb)));
  uint32_t m;   // replicate each byte's MSB over the byte (prmt's sign-replicate selector bit; __byte_perm documents 3 bits only)
  asm("prmt.b32 %0, %1, 0, 0xba98;" : "=r"(m) : "r"((ne | lt) & ge));
  return m;
}

// PRE, CONV: the aug chain (compile time: every variant is straight-line code, a few hundred SASS instructions
// instead of one 5 000-instruction kernel that thrashes the instruction cache).  WC: image width (0 = p.sp.W).
template <int PRE, bool CONV, bool NORM, int WC>
__global__ void __launch_bounds__(kFThreads, 6) ring_sample_fused_kernel(const __grid_constant__ FusedParams p,
                                                                         const __grid_constant__ InlineDraws<kInlineMax> inl) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);                 // <= 4 mbarriers
  FusedMeta* meta = reinterpret_cast<FusedMeta*>(smem + 32);          // <= 4 metas (4 x 56 B = 224 B)
  float* wsm = reinterpret_cast<float*>(smem + 256);                  // [27][4] conv weights / 255
  int* valid = reinterpret_cast<int*>(smem + 256 + 432);              // [4]: stage holds a unit (dynamic schedule)
  uint8_t* stage = smem + kFHeader;                                   // NS x stage_bytes (704 = 16 * 44: 16-B aligned)
  const int NS = p.stages;
  float* tile = reinterpret_cast<float*>(stage + NS * p.stage_bytes); // [3][kFPairs][TROW] (layout above)
  const int tid = threadIdx.x;
  const int H = WC ? WC : p.sp.H, W = WC ? WC : p.sp.W, nq = W >> 2;   // WC != 0: square WC x WC frames
  // fixed thread -> (row slot, 4-pixel group) mapping: no per-task division, constant address increments
  const int RPR = kFThreads / nq;            // rows per round (7 at W = 84)
  const int row0 = tid / nq, q4 = tid - row0 * nq;
  const bool active = row0 < RPR;
  const int TROW = WC ? tile_row_floats(WC) : p.tile_rs;   // floats per tile row pair
  const int TA1 = (nq + 1) * 4;                            // float offset of chunk array A1 inside a row pair
  constexpr bool kSplice = PRE == kPreSplice || PRE == kPreSpliceFast;
  constexpr bool kBank = PRE != kPreNone;
  constexpr int hal = CONV ? 1 : 0;

  // ---- static partition of the cost line [0, T) over the CTAs; a CTA owns the units whose start falls in its range
  const int64_t T = p.units_obs * p.cost_obs + p.units_next * p.cost_next;
  const int64_t c0 = T * blockIdx.x / gridDim.x, c1 = T * (blockIdx.x + 1) / gridDim.x;
  auto first_unit_at = [&](int64_t c) -> int64_t {   // smallest unit whose start cost >= c
    const int64_t To = p.units_obs * p.cost_obs;
    if (c <= To) return (c + p.cost_obs - 1) / p.cost_obs;
    return p.units_obs + (c - To + p.cost_next - 1) / p.cost_next;
  };
  const bool dyn = p.sched != nullptr;
  const int64_t u0 = first_unit_at(c0), u1 = first_unit_at(c1);
  // dynamic schedule: units are handed out one at a time, in order (heavy obs units first, next_obs copies fill the
  // tail), by an atomic ticket counter; the loop below then runs until a stage comes back empty
  const int nun = dyn ? 0x7fffffff : (int)(u1 - u0);
  const uint32_t total_units = (uint32_t)(p.units_obs + p.units_next);

  if (tid == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  pdl_wait();

  constexpr uint32_t kFetch = 0xffffffffu;
  // one lane: decode unit u0+kk (static) / the next ticket (dynamic), write its meta, start its copies into stage s
  auto issue = [&](int kk, int s, uint32_t ticket) {
    uint32_t u = (uint32_t)(u0 + kk);          // host: units_obs + units_next < 2^31 -> 32-bit divisions
    if (dyn) {
      u = ticket != kFetch ? ticket : atomicAdd(p.sched, 1u);
      valid[s] = u < total_units;
      if (u >= total_units) return;
    }
    const int t = u >= (uint32_t)p.units_obs ? 1 : 0;
    if (t) u -= (uint32_t)p.units_obs;
    const uint32_t fr = u / (uint32_t)p.bands;  // frame row of x.reshape(B*k, 3, H, W)
    const int band = (int)(u - fr * (uint32_t)p.bands);
    const int b = (int)(fr / (uint32_t)p.k), f = (int)(fr - (uint32_t)b * (uint32_t)p.k);
    const int BAND = CONV ? kFBand : p.band;
    const int r0 = band * BAND, rows = min(BAND, H - r0);
    const int64_t slot = sample_index(p.sp, inl, b);
    int dy, dx;
    sample_offset(p.sp, inl, t, b, dy, dx);
    const bool want_aug = t == 0 && p.o_aug != nullptr;
    const bool want_shift = p.o_shift[t] != nullptr;
    const bool want_plain = p.o_plain[t] != nullptr;
    // hull of the source rows this unit touches
    int lo = H, hi = -1;
    if (want_plain || (want_aug && !p.aug_shifted)) {
      const int h2 = (want_aug && !p.aug_shifted) ? hal : 0;
      lo = min(lo, max(r0 - h2, 0)); hi = max(hi, min(r0 + rows - 1 + h2, H - 1));
    }
    if (want_shift || (want_aug && p.aug_shifted)) {
      const int h2 = (want_aug && p.aug_shifted) ? hal : 0;
      const int a = clampi(max(r0 - h2, 0) + dy, 0, H - 1), z = clampi(min(r0 + rows - 1 + h2, H - 1) + dy, 0, H - 1);
      lo = min(lo, a); hi = max(hi, z);
    }
    const int64_t P = (int64_t)H * W;
    const size_t fo = (((size_t)b * p.sp.C + (size_t)f * 3) * H) * W;
    FusedMeta m;
    m.o_plain = want_plain ? p.o_plain[t] + fo : nullptr;
    m.o_shift = want_shift ? p.o_shift[t] + fo : nullptr;
    m.o_aug = want_aug ? p.o_aug + fo : nullptr;
    m.dy = dy; m.dx = dx; m.r0 = r0; m.rows = rows; m.sample = b; m.pad_ = 0;
    const uint8_t* src = plane_ptr(p.sp, t, slot, f * 3, P);   // plane 0 of the frame (its 3 planes are contiguous)
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(src + (int64_t)lo * W) & ~(uintptr_t)15;
    const uintptr_t a1 = (reinterpret_cast<uintptr_t>(src + (int64_t)(hi + 1) * W) + 15) & ~(uintptr_t)15;
    m.koff = (int)((intptr_t)reinterpret_cast<uintptr_t>(src) - (intptr_t)a0);
    const uint32_t sbytes = (uint32_t)(a1 - a0);
    uint32_t gbytes = 0;
    uintptr_t g0 = 0;
    m.gkoff = 0;
    if (kBank && want_aug) {
      const int64_t j = kSplice ? (int64_t)(fr % (uint32_t)p.sp.B) : b;   // batch-tiled (splice) vs channel-tiled (overlay)
      int64_t row = p.bank_idx ? p.bank_idx[j]
                               : (int64_t)(((uint64_t)philox_word((uint64_t)j, p.bank_subseq, p.sp.seed) * (uint64_t)p.bank_len) >> 32);
      row = row < 0 ? 0 : (row >= p.bank_len ? p.bank_len - 1 : row);
      const uint8_t* g = p.bank + row * 3 * P;
      const int glo = max(r0 - hal, 0), ghi = min(r0 + rows - 1 + hal, H - 1);
      g0 = reinterpret_cast<uintptr_t>(g + (int64_t)glo * W) & ~(uintptr_t)15;
      const uintptr_t g1 = (reinterpret_cast<uintptr_t>(g + (int64_t)(ghi + 1) * W) + 15) & ~(uintptr_t)15;
      m.gkoff = (int)((intptr_t)reinterpret_cast<uintptr_t>(g) - (intptr_t)g0);
      gbytes = (uint32_t)(g1 - g0);
    }
    meta[s] = m;
    mbar_expect_tx(&bars[s], 3 * (sbytes + gbytes));   // P % 16 == 0: the three planes share lead and length
    uint8_t* sb = stage + (size_t)s * p.stage_bytes;
#pragma unroll
    for (int ci = 0; ci < 3; ++ci)
      bulk_g2s(sb + ci * p.plane_stage, reinterpret_cast<const void*>(a0 + (uintptr_t)(ci * P)), sbytes, &bars[s]);
    if (gbytes) {
      uint8_t* gb = sb + 3 * p.plane_stage;
#pragma unroll
      for (int ci = 0; ci < 3; ++ci)
        bulk_g2s(gb + ci * p.bank_stage, reinterpret_cast<const void*>(g0 + (uintptr_t)(ci * P)), gbytes, &bars[s]);
    }
  };

  if (tid < 32) {   // the prologue copies are prepared by NS lanes in parallel (dynamic: NS consecutive tickets, in stage order)
    uint32_t base = 0;
    if (dyn) {
      if (tid == 0) base = atomicAdd(p.sched, (uint32_t)NS);
      base = __shfl_sync(0xffffffffu, base, 0);
    }
    if (tid < NS && tid < nun) issue(tid, tid, base + tid);
  }
  if (dyn) __syncthreads();                     // valid[] of the prologue stages
  pdl_launch_dependents();
  gather_scalars(p.sp, inl);

  uint32_t phase = 0;
  int s = 0;
  auto next_stage = [&]() { if (++s == NS) { s = 0; phase ^= 1; } };
  for (int kk = 0; kk < nun; ++kk) {
    if (dyn && !valid[s]) break;   // written >= one __syncthreads ago (NS >= 2 in this mode)
    mbar_wait(&bars[s], phase);
    const FusedMeta m = meta[s];
    const uint8_t* sb = stage + (size_t)s * p.stage_bytes;
    const size_t HW = (size_t)H * W;
    const int PS = p.plane_stage;

    // one 4-pixel word of each of the 3 planes: un-shifted, and shifted (2 LDS.32 + funnel shift + replicate selector)
    auto load_plain = [&](int r, int q4, uint32_t (&w)[3]) {
      const int a = m.koff + r * W + 4 * q4;
#pragma unroll
      for (int ci = 0; ci < 3; ++ci) w[ci] = *reinterpret_cast<const uint32_t*>(sb + ci * PS + a);
    };
    auto load_shift = [&](int r, int q4, uint32_t (&w)[3]) {
      const int rr = clampi(r + m.dy, 0, H - 1);
      const int cc = 4 * q4 + m.dx;
      const int cs = clampi(cc, 0, W - 4);
      const uint32_t sel = edge_selector(clampi(cc - cs, -4, 4));
      const int a = m.koff + rr * W + cs;
#pragma unroll
      for (int ci = 0; ci < 3; ++ci) {
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(sb + ci * PS + (a & ~3));
        w[ci] = __byte_perm(__funnelshift_r(wp[0], wp[1], (uint32_t)a << 3), 0u, sel);
      }
    };
    auto store3 = [&](float* base, size_t oo, const uint32_t (&w)[3]) {
#pragma unroll
      for (int ci = 0; ci < 3; ++ci)
        st_global_f4(reinterpret_cast<float4*>(base + oo + ci * HW), fin4<NORM>(bytes_to_float4(w[ci])));
    };

    if (m.o_aug == nullptr) {
      // ---------------- copy unit (next_obs, or obs without an augmented output) ----------------
      for (int pr = active ? row0 : m.rows; pr < m.rows; pr += RPR) {
        const int r = m.r0 + pr;
        const size_t oo = (size_t)r * W + 4 * q4;
        uint32_t w[3];
        if (m.o_plain) { load_plain(r, q4, w); store3(m.o_plain, oo, w); }
        if (m.o_shift) { load_shift(r, q4, w); store3(m.o_shift, oo, w); }
      }
      __syncthreads();   // stage s consumed
      if (tid == 0 && kk + NS < nun) issue(kk + NS, s, kFetch);
      next_stage();
      continue;
    }

    // ---------------- augmented obs unit, phase A: elementwise ----------------
    const uint8_t* gb = sb + 3 * PS;
    const int trows = m.rows + 2 * hal;
    const bool aug_sh = p.aug_shifted != 0;
    if (CONV && tid < 81) {
      // weights[sample][co][ci][ky][kx] / 255 -> wsm[(ci*9 + ky*3 + kx)*4 + co]
      const int co = tid / 27, tap = tid - co * 27;
      wsm[tap * 4 + co] = p.weights[(int64_t)m.sample * 81 + tid] * kConvWScale;
    }
    for (int pr = active ? row0 : trows; pr < trows; pr += RPR) {
      const int r = clampi(m.r0 + pr - hal, 0, H - 1);
      const bool own = !CONV || (pr >= 1 && pr <= m.rows);   // halo rows only feed the conv tile
      const size_t oo = (size_t)r * W + 4 * q4;              // offset inside plane 0; planes are H*W apart
      uint32_t Aw[3], Ow[3];
      if (aug_sh) {
        load_shift(r, q4, Aw);
        if (own && m.o_shift) store3(m.o_shift, oo, Aw);
        if (own && m.o_plain) { load_plain(r, q4, Ow); store3(m.o_plain, oo, Ow); }
      } else {
        load_plain(r, q4, Aw);
        if (own && m.o_plain) store3(m.o_plain, oo, Aw);
        if (own && m.o_shift) { load_shift(r, q4, Ow); store3(m.o_shift, oo, Ow); }
      }
      float v[3][4];
      if (PRE == kPreSpliceFast) {
        // default thresholds (hue = sat = 0): the HSV test of pixel.cuh:hsv_gt only compares, and v -> v/255 is strictly
        // increasing, so mask and select run on the raw bytes, four pixels per instruction (splice_keep_bytes);
        // tests/: all 2^24 colours against the float path.
        const int ga = m.gkoff + r * W + 4 * q4;
        const uint32_t keep = splice_keep_bytes(Aw[0], Aw[1], Aw[2], p.thr_k, p.thr_c7);
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) {
          const uint32_t gw = *reinterpret_cast<const uint32_t*>(gb + ci * p.bank_stage + ga);
          const float4 t4 = bytes_to_float4((Aw[ci] & keep) | (gw & ~keep));
          // splice returns (x/255 or bg/255) * 255; as conv input the raw byte is the same value to 1 ulp
          v[ci][0] = CONV ? t4.x : div255(t4.x) * 255.0f; v[ci][1] = CONV ? t4.y : div255(t4.y) * 255.0f;
          v[ci][2] = CONV ? t4.z : div255(t4.z) * 255.0f; v[ci][3] = CONV ? t4.w : div255(t4.w) * 255.0f;
        }
      } else {
        float a[3][4];
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) {
          const float4 t4 = bytes_to_float4(Aw[ci]);
          a[ci][0] = t4.x; a[ci][1] = t4.y; a[ci][2] = t4.z; a[ci][3] = t4.w;
        }
        if (kBank) {
          float g[3][4];
          const int ga = m.gkoff + r * W + 4 * q4;
#pragma unroll
          for (int ci = 0; ci < 3; ++ci) {
            const float4 g4 = bytes_to_float4(*reinterpret_cast<const uint32_t*>(gb + ci * p.bank_stage + ga));
            g[ci][0] = g4.x; g[ci][1] = g4.y; g[ci][2] = g4.z; g[ci][3] = g4.w;
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (PRE == kPreOverlay) {
#pragma unroll
              for (int ci = 0; ci < 3; ++ci) v[ci][j] = overlay_px(a[ci][j], div255(g[ci][j]), p.oma, p.alpha);
            } else {
              const bool keep = splice_keep(div255(a[0][j]), div255(a[1][j]), div255(a[2][j]), p.thr);
#pragma unroll
              for (int ci = 0; ci < 3; ++ci) {
                const float sel = keep ? a[ci][j] : g[ci][j];
                v[ci][j] = CONV ? sel : div255(sel) * 255.0f;
              }
            }
          }
        } else {
#pragma unroll
          for (int ci = 0; ci < 3; ++ci)
#pragma unroll
            for (int j = 0; j < 4; ++j) v[ci][j] = a[ci][j];
        }
      }
      if (CONV) {
        // tile row pr -> lane .x of row pair pr (pr < kFPairs) and / or lane .y of row pair pr - kFHalf (pr >= kFHalf);
        // pixels 4q4 .. 4q4+3 are columns c' = 4q4+1 .. 4q4+4: A0[q4].hi, A1[q4].lo, A1[q4].hi, A0[q4+1].lo
        auto put = [&](float* tb) {
#pragma unroll
          for (int ci = 0; ci < 3; ++ci) {
            float* t = tb + ci * kFPairs * TROW;
            t[2] = v[ci][0]; t[TA1] = v[ci][1]; t[TA1 + 2] = v[ci][2]; t[4] = v[ci][3];
            if (q4 == 0) t[0] = v[ci][0];              // column -1: replicate
            if (q4 == nq - 1) t[6] = v[ci][3];         // column W
          }
        };
        if (pr < kFPairs) put(tile + pr * TROW + 4 * q4);
        if (pr >= kFHalf) put(tile + (pr - kFHalf) * TROW + 4 * q4 + 1);
      } else {
#pragma unroll
        for (int ci = 0; ci < 3; ++ci)
          st_global_f4(reinterpret_cast<float4*>(m.o_aug + oo + ci * HW),
                       fin4<NORM>(make_float4(v[ci][0], v[ci][1], v[ci][2], v[ci][3])));
      }
    }
    __syncthreads();   // stage s consumed (refill may start), tile + wsm complete
    if (tid == 0 && kk + NS < nun) issue(kk + NS, s, kFetch);
    // ---------------- phase B: conv on the tile ----------------
    if (CONV) {
      const int na = min(kFHalf, m.rows);
      for (int a = active ? row0 : na; a < na; a += RPR) {
        float2 acc[3][4];   // [co][pixel] = (band row a, band row a + kFHalf)
#pragma unroll
        for (int co = 0; co < 3; ++co)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[co][j] = make_float2(0.0f, 0.0f);
#pragma unroll 1
        for (int ci = 0; ci < 3; ++ci) {
#pragma unroll
          for (int ky = 0; ky < 3; ++ky) {
            const float* rb = tile + ((ci * kFPairs + a + ky) * TROW + 4 * q4);
            const float4 X0 = *reinterpret_cast<const float4*>(rb);
            const float4 X1 = *reinterpret_cast<const float4*>(rb + TA1);
            const float4 X2 = *reinterpret_cast<const float4*>(rb + 4);
            const float2 xs[6] = {make_float2(X0.x, X0.y), make_float2(X0.z, X0.w), make_float2(X1.x, X1.y),
                                  make_float2(X1.z, X1.w), make_float2(X2.x, X2.y), make_float2(X2.z, X2.w)};
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
              const float4 wv = *reinterpret_cast<const float4*>(wsm + (ci * 9 + ky * 3 + kx) * 4);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                acc[0][j] = __ffma2_rn(xs[j + kx], make_float2(wv.x, wv.x), acc[0][j]);
                acc[1][j] = __ffma2_rn(xs[j + kx], make_float2(wv.y, wv.y), acc[1][j]);
                acc[2][j] = __ffma2_rn(xs[j + kx], make_float2(wv.z, wv.z), acc[2][j]);
              }
            }
          }
        }
        float* orow = m.o_aug + (size_t)(m.r0 + a) * W + 4 * q4;
#pragma unroll
        for (int co = 0; co < 3; ++co)
          st_global_f4(reinterpret_cast<float4*>(orow + co * HW),
                       make_float4(conv_sigmoid<NORM>(acc[co][0].x), conv_sigmoid<NORM>(acc[co][1].x),
                                   conv_sigmoid<NORM>(acc[co][2].x), conv_sigmoid<NORM>(acc[co][3].x)));
        if (a + kFHalf < m.rows) {
#pragma unroll
          for (int co = 0; co < 3; ++co)
            st_global_f4(reinterpret_cast<float4*>(orow + co * HW + (size_t)kFHalf * W),
                         make_float4(conv_sigmoid<NORM>(acc[co][0].y), conv_sigmoid<NORM>(acc[co][1].y),
                                     conv_sigmoid<NORM>(acc[co][2].y), conv_sigmoid<NORM>(acc[co][3].y)));
        }
      }
      __syncthreads();   // tile and wsm are free again
    }
    next_stage();
  }
  if (dyn && tid == 0) {
    // the last CTA to run out of units re-arms the counters for the next launch on this ring (launches of one ring
    // are stream-ordered, and a PDL successor takes its first ticket only after griddepcontrol.wait)
    __threadfence();
    if (atomicAdd(p.sched + 1, 1u) == gridDim.x - 1) {
      p.sched[0] = 0;
      p.sched[1] = 0;
      __threadfence();
    }
  }
}

}  // namespace augcl

namespace augcl {
using FusedKernel = void (*)(const FusedParams, const InlineDraws<kInlineMax>);

template <int PRE, bool CONV>
static FusedKernel pick_wn(bool norm, int W) {
  if (norm) return ring_sample_fused_kernel<PRE, CONV, true, 0>;
  if (W == 84) return ring_sample_fused_kernel<PRE, CONV, false, 84>;
  return ring_sample_fused_kernel<PRE, CONV, false, 0>;
}

static FusedKernel pick_kernel(int pre, bool conv, bool norm, int W) {
  switch (pre) {
    case kPreSplice: return conv ? pick_wn<kPreSplice, true>(norm, W) : pick_wn<kPreSplice, false>(norm, W);
    case kPreSpliceFast: return conv ? pick_wn<kPreSpliceFast, true>(norm, W) : pick_wn<kPreSpliceFast, false>(norm, W);
    case kPreOverlay: return conv ? pick_wn<kPreOverlay, true>(norm, W) : pick_wn<kPreOverlay, false>(norm, W);
    default: return conv ? pick_wn<kPreNone, true>(norm, W) : pick_wn<kPreNone, false>(norm, W);
  }
}
}  // namespace augcl

using namespace augcl;

extern "C" int augcl_ring_sample_fused(augcl_ring* r, const augcl_fused_desc* d, augcl_stream stream) {
  const char* who = "augcl_ring_sample_fused";
  AUGCL_REQUIRE(d, "%s: desc is NULL", who);
  AUGCL_REQUIRE(d->struct_size == (int32_t)sizeof(augcl_fused_desc), "%s: desc.struct_size %d != %d (ABI mismatch)", who,
                d->struct_size, (int)sizeof(augcl_fused_desc));
  AUGCL_REQUIRE(r, "%s: ring is NULL", who);
  const int B = d->B;
  AUGCL_REQUIRE(B > 0, "%s: B must be positive (got %d)", who, B);
  AUGCL_REQUIRE(d->idx_src == AUGCL_SRC_DEVICE || d->idx_src == AUGCL_SRC_HOST || d->idx_src == AUGCL_SRC_PHILOX,
                "%s: unknown idx_src %d", who, d->idx_src);
  AUGCL_REQUIRE(d->off_src == AUGCL_SRC_DEVICE || d->off_src == AUGCL_SRC_HOST || d->off_src == AUGCL_SRC_PHILOX,
                "%s: unknown off_src %d", who, d->off_src);
  if (d->idx_src == AUGCL_SRC_PHILOX)
    AUGCL_REQUIRE(d->idx_hi > 0 && d->idx_hi <= r->capacity, "%s: idx_hi %lld outside (0, capacity]", who, (long long)d->idx_hi);
  else
    AUGCL_REQUIRE(d->idx, "%s: idx is NULL", who);
  AUGCL_REQUIRE(d->pad >= 0 && d->pad <= 63, "%s: pad %d outside [0, 63]", who, d->pad);
  const bool any_obs = d->out_obs || d->out_obs_shift || d->out_obs_aug;
  const bool any_next = d->out_next || d->out_next_shift;
  AUGCL_REQUIRE(any_obs || any_next, "%s: every image output is NULL", who);
  const int ops = d->out_obs_aug ? d->aug_ops : 0;
  if (d->out_obs_aug) {
    AUGCL_REQUIRE(ops != 0 && (ops & ~15) == 0, "%s: out_obs_aug needs aug_ops out of AUGCL_OP_* (got %d)", who, ops);
    const int pre = ops & (AUGCL_OP_SPLICE | AUGCL_OP_CS_SPLICE | AUGCL_OP_OVERLAY);
    AUGCL_REQUIRE((pre & (pre - 1)) == 0, "%s: at most one of SPLICE / CS_SPLICE / OVERLAY per chain", who);
    AUGCL_REQUIRE(d->aug_input == AUGCL_AUG_INPUT_PLAIN || d->aug_input == AUGCL_AUG_INPUT_SHIFTED,
                  "%s: unknown aug_input %d", who, d->aug_input);
    if (pre) {
      AUGCL_REQUIRE(d->bank && d->bank_len > 0 && d->bank_len <= (1ll << 32), "%s: splice / overlay need bank and bank_len", who);
      if (!aligned(d->bank, 16)) AUGCL_FAIL(AUGCL_ERR_BAD_ALIGN, "%s: bank must be 16-byte aligned", who);
    }
    if (ops & AUGCL_OP_CONV) AUGCL_REQUIRE(d->conv_weights, "%s: CONV needs conv_weights", who);
  }
  if (r->cpf != 3) AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "%s: needs RGB frames (c_per_frame == 3)", who);
  if ((r->W & 3) || ((r->H * r->W) & 15) || r->W > 128 || r->W < 8)
    AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "%s: needs W %% 4 == 0, 8 <= W <= 128 and (H*W) %% 16 == 0 (got %dx%d)", who, r->H, r->W);
  float* outs[5] = {d->out_obs, d->out_obs_shift, d->out_obs_aug, d->out_next, d->out_next_shift};
  for (float* o : outs)
    if (!aligned(o, 16)) AUGCL_FAIL(AUGCL_ERR_BAD_ALIGN, "%s: image outputs must be 16-byte aligned", who);

  FusedParams p{};
  SampleParams& sp = p.sp;
  sp.store[0] = r->obs; sp.store[1] = r->next; sp.refs = r->dedup ? r->refs : nullptr; sp.k = r->k; sp.cpf = r->cpf;
  sp.idx = d->idx; sp.act = r->act; sp.rew = r->rew; sp.nd = r->nd;
  sp.out_act = d->out_act; sp.out_rew = d->out_rew; sp.out_nd = d->out_nd;
  sp.capacity = r->capacity;
  sp.B = B; sp.C = r->C; sp.H = r->H; sp.W = r->W; sp.Ho = r->H; sp.Wo = r->W; sp.A = r->A;
  sp.pad = d->pad; sp.normalize = d->normalize ? 1 : 0;
  sp.idx_mode = d->idx_src; sp.idx_hi = (uint64_t)d->idx_hi; sp.seed = d->seed; sp.subseq = d->subsequence;
  sp.off_range = (uint32_t)(2 * d->pad + 1);
  // offsets: tensor 0 (obs) is shifted when out_obs_shift is written or the chain runs on the shifted tensor
  const bool sh[2] = {d->out_obs_shift != nullptr || (d->out_obs_aug && d->aug_input == AUGCL_AUG_INPUT_SHIFTED),
                      d->out_next_shift != nullptr};
  const int32_t* off[2] = {d->off_obs, d->off_next};
  for (int t = 0; t < 2; ++t) {
    if (!sh[t]) { sp.off_mode[t] = kSrcNone; continue; }
    if (d->off_src == AUGCL_SRC_PHILOX) { sp.off_mode[t] = off[t] ? AUGCL_SRC_DEVICE : AUGCL_SRC_PHILOX; sp.off[t] = off[t]; continue; }
    if (!off[t]) { sp.off_mode[t] = kSrcNone; continue; }
    sp.off_mode[t] = d->off_src; sp.off[t] = off[t];
  }
  p.o_plain[0] = d->out_obs; p.o_plain[1] = d->out_next;
  p.o_shift[0] = d->out_obs_shift; p.o_shift[1] = d->out_next_shift;
  p.o_aug = d->out_obs_aug;
  p.bank = d->bank; p.bank_idx = d->bank_idx; p.bank_subseq = d->bank_subsequence; p.bank_len = d->bank_len; p.weights = d->conv_weights;
  p.thr = SpliceThr{d->hue_thres, d->sat_thres, d->val_thres, (ops & AUGCL_OP_CS_SPLICE) ? AUGCL_CS_SPLICE : AUGCL_SPLICE};
  int pre = kPreNone;
  if (ops & AUGCL_OP_OVERLAY) pre = kPreOverlay;
  if (ops & AUGCL_OP_CS_SPLICE) pre = kPreSplice;
  if (ops & AUGCL_OP_SPLICE) {
    // default hue/sat thresholds: integer-exact mask on the raw bytes (kernel comment); anything else: float HSV
    pre = (d->hue_thres == 0.0f && d->sat_thres == 0.0f && g_fused_fast_mask.load()) ? kPreSpliceFast : kPreSplice;
    int t = 256;
    for (int b = 255; b >= 0; --b) {
      volatile float q = (float)b / 255.0f;      // IEEE division, as div255 / torch
      if (q > d->val_thres) t = b; else break;
    }
    if (t < 1 || t > 255) pre = kPreSplice;      // every / no pixel passes the value test: the float path handles it
    const uint32_t c = (uint32_t)(256 - t) & 0xffu;
    p.thr_k = (c & 0x7fu) * 0x01010101u;
    p.thr_c7 = (c & 0x80u) ? 0x80808080u : 0u;
  }
  p.alpha = d->alpha; p.oma = (float)(1.0 - (double)d->alpha);
  p.ops = ops; p.aug_shifted = d->aug_input == AUGCL_AUG_INPUT_SHIFTED ? 1 : 0;
  // non-conv chains: 14-row bands = two full rounds of 7 rows.  (7-row bands were measured for short launches, where a CTA
  // only gets ~5 units: the finer partition does not pay for the doubled per-unit cost -- 27.0 vs 24.6 us at batch 128.)
  int band = (ops & AUGCL_OP_CONV) ? kFBand : kFBandNoConv;
  if (!(ops & AUGCL_OP_CONV) && g_fused_band.load() > 0) band = g_fused_band.load();
  p.band = band;
  p.k = r->k; p.bands = (r->H + band - 1) / band;
  p.units_obs = any_obs ? (int64_t)B * r->k * p.bands : 0;
  p.units_next = any_next ? (int64_t)B * r->k * p.bands : 0;
  AUGCL_REQUIRE(p.units_obs + p.units_next < (1ll << 31), "%s: batch too large", who);
  // cost of a unit ~ the f32 planes it writes (+ the conv's arithmetic, worth about one and a half more)
  const int n_obs = (d->out_obs ? 1 : 0) + (d->out_obs_shift ? 1 : 0) + (d->out_obs_aug ? 1 : 0);
  const int n_next = (d->out_next ? 1 : 0) + (d->out_next_shift ? 1 : 0);
  p.cost_obs = g_fused_cost_obs.load() > 0 ? g_fused_cost_obs.load()
                                          : 4 * n_obs + ((ops & AUGCL_OP_CONV) ? 6 : 0) + ((ops & 7) ? 1 : 0) + 1;
  p.cost_next = g_fused_cost_next.load() > 0 ? g_fused_cost_next.load() : 4 * n_next + 1;
  if (!any_obs) p.cost_obs = 1;
  if (!any_next) p.cost_next = 1;
  p.normalize = d->normalize ? 1 : 0;
  const bool has_conv = (ops & AUGCL_OP_CONV) != 0;
  const int halo2 = has_conv ? 2 : 0;
  const int rows_max = (band + halo2 + d->pad) < r->H ? (band + halo2 + d->pad) : r->H;
  p.plane_stage = ((rows_max * r->W + 30 + 15) & ~15) + 16;
  p.bank_stage = (ops & 7) ? ((((band + halo2) * r->W + 30 + 15) & ~15) + 16) : 0;
  p.stage_bytes = 3 * (p.plane_stage + p.bank_stage);
  p.tile_rs = tile_row_floats(r->W);
  const size_t tile_bytes = has_conv ? 3 * (size_t)kFPairs * p.tile_rs * sizeof(float) : 0;
  // conv chains are issue / barrier bound: 2 stages and the smaller footprint; elementwise chains wait on HBM: deeper
  // (measured, scripts/bench_fused.py --opts: 4 stages x 6 CTAs/SM with a bank, 3 x 7 for plain copies)
  p.stages = g_fused_stages.load() > 0 ? g_fused_stages.load() : (has_conv ? 2 : ((ops & 7) ? 4 : 3));
  while (p.stages > 1 && kFHeader + (size_t)p.stages * p.stage_bytes + tile_bytes > 100 * 1024) --p.stages;
  const size_t smem = kFHeader + (size_t)p.stages * p.stage_bytes + tile_bytes;
  p.sched = (g_fused_dynamic.load() && p.stages >= 2) ? r->sched : nullptr;
  if (smem > 200 * 1024) AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "%s: geometry needs %zu B of shared memory", who, smem);

  DeviceGuard g(r->device);
  AUGCL_CUDA(g.err);
  cudaStream_t st = (cudaStream_t)stream;

  // ---- HOST draws: validate like the reference's IndexError, then stage through stream-ordered scratch
  void* scratch = nullptr;
  const bool host_idx = d->idx_src == AUGCL_SRC_HOST;
  if (host_idx)
    for (int b = 0; b < B; ++b)
      AUGCL_REQUIRE(d->idx[b] >= 0 && d->idx[b] < r->capacity, "%s: index %lld (position %d) out of range [0, %lld)", who,
                    (long long)d->idx[b], b, (long long)r->capacity);
  bool host_off = false;
  for (int t = 0; t < 2; ++t)
    if (sp.off_mode[t] == AUGCL_SRC_HOST) {
      host_off = true;
      for (int e = 0; e < 2 * B; ++e)
        AUGCL_REQUIRE(off[t][e] >= 0 && off[t][e] <= 2 * d->pad, "%s: offset %d (tensor %d, element %d) outside [0, %d]", who,
                      off[t][e], t, e, 2 * d->pad);
    }
  InlineDraws<kInlineMax> inl;
  if ((host_idx || host_off) && B <= kInlineMax && r->capacity < (1ll << 31) && 2 * d->pad <= 127) {
    // by value in the launch's parameter block: no scratch, no H2D copy (the offsets were range-checked above)
    if (host_idx) {
      for (int b = 0; b < B; ++b) inl.idx[b] = (int32_t)d->idx[b];
      sp.idx_mode = kSrcInline;
    }
    for (int t = 0; t < 2; ++t)
      if (sp.off_mode[t] == AUGCL_SRC_HOST) {
        for (int b = 0; b < B; ++b) { inl.off[t][b][0] = (int8_t)off[t][2 * b]; inl.off[t][b][1] = (int8_t)off[t][2 * b + 1]; }
        sp.off_mode[t] = kSrcInline;
      }
  } else if (host_idx || host_off) {
    const size_t ib = (size_t)B * 8, ob = (size_t)B * 8;
    AUGCL_CUDA(cudaMallocFromPoolAsync(&scratch, ib + 2 * ob, r->scratch_pool, st));
    uint8_t* spn = static_cast<uint8_t*>(scratch);
    cudaError_t e = cudaSuccess;
    if (host_idx) {
      e = cudaMemcpyAsync(spn, d->idx, ib, cudaMemcpyHostToDevice, st);
      sp.idx = reinterpret_cast<const int64_t*>(spn);
      sp.idx_mode = AUGCL_SRC_DEVICE;
    }
    for (int t = 0; t < 2 && e == cudaSuccess; ++t)
      if (sp.off_mode[t] == AUGCL_SRC_HOST) {
        e = cudaMemcpyAsync(spn + ib + t * ob, off[t], ob, cudaMemcpyHostToDevice, st);
        sp.off[t] = reinterpret_cast<const int32_t*>(spn + ib + t * ob);
        sp.off_mode[t] = AUGCL_SRC_DEVICE;
      }
    if (e != cudaSuccess) {
      cudaFreeAsync(scratch, st);
      AUGCL_FAIL(AUGCL_ERR_CUDA, "%s: staging the host draws failed: %s", who, cudaGetErrorString(e));
    }
  }

  const FusedKernel kern = pick_kernel(pre, (ops & AUGCL_OP_CONV) != 0, p.normalize != 0, (r->H == 84 && r->W == 84) ? 84 : 0);
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(kern), 200 * 1024)) {   // per (kernel, device), remembered
    if (scratch) cudaFreeAsync(scratch, st);
    return rc;
  }
  int per_sm = g_fused_ctas_per_sm.load();
  if (per_sm <= 0) per_sm = ops == 0 ? 7 : 6;
  const int fit = (int)((228 * 1024) / (smem + 1024));
  if (per_sm > fit) per_sm = fit < 1 ? 1 : fit;
  int64_t grid = (int64_t)kNumSMs * per_sm;
  const int64_t units = p.units_obs + p.units_next;
  if (grid > units) grid = units;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kFThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_sample_pdl.load() ? 1 : 0;
  cudaError_t le = cudaLaunchKernelEx(&cfg, kern, p, inl);
  if (le == cudaSuccess) le = cudaGetLastError();
  if (scratch) cudaFreeAsync(scratch, st);   // stream-ordered: released after the kernel that reads it
  if (le != cudaSuccess) AUGCL_FAIL(AUGCL_ERR_CUDA, "%s: launch failed: %s", who, cudaGetErrorString(le));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return AUGCL_OK;
}
