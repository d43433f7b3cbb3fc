// Fused per-algorithm sampler: gather + u8->f32 + random_shift + {splice | CS_splice | overlay} + random_conv
// in ONE kernel, straight out of the device ring (include/augcl_b200.h: augcl_ring_sample_fused).
//
// Replaces, per update() of SVEA / AugCL (/root/reference/src/algorithms/svea.py:66-70, augcl.py:69-72), the chain
//   ReplayBuffer.sample (src/utils.py:272-280) -> random_shift (src/augmentations.py:378-382)
//   -> RAD.apply_aug (src/algorithms/rad.py:18-22) over splice / random_overlay / random_conv (:253-324)
// which in the unfused form costs four kernels and 2.2 MB of HBM traffic per sample; fused, every u8 frame is read
// once and only the final f32 tensors are written (952 560 B per sample for AugCL's splice-conv at 9x84x84).
//
// Work unit = a band of rows (kFBand for conv chains, else p.band) of one RGB frame (3 u8 planes) of one (tensor, sample).
// Units are handed out by ticket (guided self-scheduling on a counter in the ring handle, one set per stream) to
// persistent, warp-specialised CTAs: a producer warp decodes units and starts <= 6 bulk copies per unit (cp.async.bulk,
// SASS UBLKCP: the hull of source rows of the 3 planes and, for splice / overlay, the matching rows of the 3 background
// planes) into a ring of stages; A-warps run the elementwise phase out of a stage; conv chains hand an f32 tile over to
// B-warps.  All hand-overs are mbarriers (see "CTA roles" below).
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
std::atomic<int> g_fused_stages{0};      // 0 = auto
std::atomic<int> g_fused_batch{0};       // units the producer warp decodes side by side (0 = auto)
std::atomic<int> g_fused_interleave{-1}; // obs / next_obs units alternate in ticket order (-1 = auto: conv chains only)
std::atomic<int> g_fused_skip{0};        // TIMING EXPERIMENTS ONLY (wrong results): 1 = no conv arithmetic, 2 = no tile stores, 4 = no direct obs stores
std::atomic<int> g_fused_band{0};        // 0 = auto (non-conv chains: 14 or 7 rows)
std::atomic<int> g_fused_fast_mask{1};   // 0: always the float HSV path (A/B and the exhaustive equivalence test)

constexpr int kFHeader = 1664;   // mbarriers, per-stage flags, tile metas, the unit-meta ring, conv weights per tile
constexpr int kFBand = 12;        // conv chains: 12 rows + 2 halo rows = 14 tile rows = 294 items = two rounds of 160 A-threads
constexpr int kFBandNoConv = 14;  // every other chain: 84 = 6 bands

struct FusedParams {
  SampleParams sp;          // stores, draw sources, scalar gather (sp.out unused)
  float* o_plain[2];        // [obs, next]
  float* o_shift[2];
  float* o_aug;
  const uint8_t* bank;
  const int64_t* bank_idx;
  uint64_t bank_subseq;     // bank_idx == nullptr: rows drawn here, same values as augcl_rng_fill_indices
  uint64_t conv_subseq;     // weights == nullptr: conv weights drawn here, same values as augcl_rng_fill_normal
  const float* weights;
  int64_t bank_len;
  SpliceThr thr;
  uint32_t thr_k, thr_c7;   // kPreSpliceFast: the byte threshold t (smallest b with b/255 > thr.val) as SWAR constants
  float oma, alpha;
  int ops;                  // AUGCL_OP_*
  int aug_shifted;          // chain input: 0 plain, 1 shifted
  int k, bands, band;       // frames per stack, bands per frame, rows per band
  int64_t units_obs, units_next;
  int plane_stage, bank_stage, stage_bytes, tile_rs, stages;
  uint32_t* sched;          // unit schedule: {next ticket, CTAs done}
  int normalize;
  int batch, interleave, skip;
  double rcp_bands, rcp_k, rcp_B;   // exact small-divisor division in the producer warp (udiv_small)
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
  int pad_[3];
};

template <bool NORM>
__device__ __forceinline__ float fin(float v) { return NORM ? div255(v) : v; }   // div255 == IEEE v / 255 (pixel.cuh)
template <bool NORM>
__device__ __forceinline__ float4 fin4(float4 v) {
  return make_float4(fin<NORM>(v.x), fin<NORM>(v.y), fin<NORM>(v.z), fin<NORM>(v.w));
}

static_assert(sizeof(FusedMeta) == 64, "FusedMeta travels through shared memory as four 16-byte words");
__device__ __forceinline__ void meta_store(uint4* dst, const FusedMeta& m) {
  const uint64_t a = reinterpret_cast<uint64_t>(m.o_plain), b = reinterpret_cast<uint64_t>(m.o_shift), c = reinterpret_cast<uint64_t>(m.o_aug);
  dst[0] = make_uint4((uint32_t)a, (uint32_t)(a >> 32), (uint32_t)b, (uint32_t)(b >> 32));
  dst[1] = make_uint4((uint32_t)c, (uint32_t)(c >> 32), (uint32_t)m.koff, (uint32_t)m.gkoff);
  dst[2] = make_uint4((uint32_t)m.dy, (uint32_t)m.dx, (uint32_t)m.r0, (uint32_t)m.rows);
  dst[3] = make_uint4((uint32_t)m.sample, 0u, 0u, 0u);
}
__device__ __forceinline__ FusedMeta meta_load(const uint4* src) {
  const uint4 v0 = src[0], v1 = src[1], v2 = src[2], v3 = src[3];
  FusedMeta m;
  m.o_plain = reinterpret_cast<float*>((uint64_t)v0.x | ((uint64_t)v0.y << 32));
  m.o_shift = reinterpret_cast<float*>((uint64_t)v0.z | ((uint64_t)v0.w << 32));
  m.o_aug = reinterpret_cast<float*>((uint64_t)v1.x | ((uint64_t)v1.y << 32));
  m.koff = (int)v1.z; m.gkoff = (int)v1.w;
  m.dy = (int)v2.x; m.dx = (int)v2.y; m.r0 = (int)v2.z; m.rows = (int)v2.w;
  m.sample = (int)v3.x; m.pad_[0] = 0; m.pad_[1] = 0; m.pad_[2] = 0;
  return m;
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

// ---- CTA roles (warp-specialised; every hand-over is an mbarrier, there is no CTA-wide barrier after start-up) ----
//   producer warp (the last one): takes unit tickets, decodes them, stages the conv weights of the unit's sample and starts
//     the bulk copies of stage s as soon as the A-warps have released it (empty[s]); the ticket atomic and the dependent
//     index / table loads of unit n+1 run while the other warps work on unit n.
//   A-warps: wait full[s]; elementwise phase (copy units: next_obs and un-augmented obs; aug units: phase A).  Conv chains
//     write the conv input into tile[t] (two tiles) after tempty[t], then arrive on tfull[t]; every A-warp releases the
//     stage (empty[s]) when its own reads are done.
//   B-warps (conv chains): wait tfull[t]; phase B on tile[t]; arrive on tempty[t].
// The first round-2 kernel ran both phases on all warps with two __syncthreads per unit: ncu showed 38 % of the warp
// samples parked on those barriers (24 % on the one after phase B alone, most of it waiting for warp 0, whose lane 0
// was still decoding the next unit: ~2.9 k cycles per unit), issue slots 48 % busy.
// conv chains: 5 A-warps (the 294 items of a 14-row tile = two rounds of 160), 4 B-warps (126 tasks), 3 CTAs per SM at 64
// registers; the first split was 4 + 4: ncu showed the A-warps busy 74 % (and waiting for data the rest), the B-warps 59 %.
constexpr int kFWarpsAConv = 5, kFWarpsBConv = 4, kFWarpsAPlain = 8;
constexpr int fused_threads(bool conv) { return 32 * ((conv ? kFWarpsAConv + kFWarpsBConv : kFWarpsAPlain) + 1); }
constexpr int kFMetaRing = 8;   // unit metas live in their own ring: written a pass at a time, before the stages are free
constexpr int kFMaxStages = 4;

struct TileMeta {
  float* o_aug;     // plane 0 of the frame in the augmented output
  int r0, rows;     // band
  int valid;        // 0: no more tiles (the B-warps leave)
  int sample;
};

// floor(u / d) for u < 2^31 through d's double reciprocal: (u + 0.5) / d is never within 2^-40 of an integer, so the
// rounding of the reciprocal and of the product cannot carry it across one.  Three instructions instead of the ~25 of
// a 32-bit division by a run-time divisor (the producer warp decodes every unit with three of them).
__device__ __forceinline__ uint32_t udiv_small(uint32_t u, double rcp) {
  return (uint32_t)__double2uint_rz(((double)u + 0.5) * rcp);
}

// PRE, CONV: the aug chain (compile time: every variant is straight-line code, a few hundred SASS instructions
// instead of one 5 000-instruction kernel that thrashes the instruction cache).  WC: image width (0 = p.sp.W).
template <int PRE, bool CONV, bool NORM, int WC>
__global__ void __launch_bounds__(fused_threads(CONV), CONV ? 3 : 4)
ring_sample_fused_kernel(const __grid_constant__ FusedParams p, const __grid_constant__ InlineDraws<kInlineMax> inl) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);                 // [4] stage s holds a unit (tx bytes) / end marker
  uint64_t* empty = full + kFMaxStages;                               // [4] every A-warp is done with stage s
  uint64_t* tfull = full + 2 * kFMaxStages;                           // [2] tile t written by every A-warp
  uint64_t* tempty = tfull + 2;                                       // [2] tile t read by every B-warp
  int* valid = reinterpret_cast<int*>(smem + 96);                     // [4]
  TileMeta* tmeta = reinterpret_cast<TileMeta*>(smem + 128);          // [2]
  uint4* meta = reinterpret_cast<uint4*>(smem + 192);                 // [kFMetaRing][4]: FusedMeta of unit n at slot n % 8
  float* wsm = reinterpret_cast<float*>(smem + 704);                  // [2][27][4] conv weights of tile t (scaled)
  uint8_t* stage = smem + kFHeader;                                   // NS x stage_bytes
  const int NS = p.stages;
  float* tiles = reinterpret_cast<float*>(stage + NS * p.stage_bytes); // [2][3][kFPairs][TROW] (layout above)
  constexpr int NAW = CONV ? kFWarpsAConv : kFWarpsAPlain, NBW = CONV ? kFWarpsBConv : 0;
  constexpr int NTA = NAW * 32, NTB = NBW * 32;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int H = WC ? WC : p.sp.H, W = WC ? WC : p.sp.W, nq = W >> 2;   // WC != 0: square WC x WC frames
  const int TROW = WC ? tile_row_floats(WC) : p.tile_rs;   // floats per tile row pair
  const int TA1 = (nq + 1) * 4;                            // float offset of chunk array A1 inside a row pair
  const int TILEF = 3 * kFPairs * TROW;
  constexpr bool kSplice = PRE == kPreSplice || PRE == kPreSpliceFast;
  constexpr bool kBank = PRE != kPreNone;
  constexpr int hal = CONV ? 1 : 0;
  const uint32_t total_units = (uint32_t)(p.units_obs + p.units_next);

  if (tid == 0) {
    for (int s = 0; s < kFMaxStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NAW); valid[s] = 1; }   // valid: cleared by the end marker
    for (int t = 0; t < 2; ++t) { mbar_init(&tfull[t], NAW); mbar_init(&tempty[t], NBW > 0 ? NBW : 1); }
    mbar_fence_init();
  }
  __syncthreads();
  // Programmatic dependent launch: a CTA that became resident while the previous kernel of the stream drains may READ
  // (ring, refs table, draws, background bank, conv weights: nothing a sample kernel -- the only kind that lets its
  // successor start early -- ever writes) but must not store, nor touch the ticket counter the predecessor re-arms at
  // its end, before griddepcontrol.wait.  The producer's first chunk of tickets is static, so its first copies are in
  // flight before that point; each role waits right before its first store / atomic.
  pdl_launch_dependents();

  if (warp == NAW + NBW) {
    // =============================== producer warp ===============================
    // One pass = G units (G <= NS): lane i < G owns ticket t0 + i.  A pass is prepared in two steps so that no load
    // latency sits between the release of a stage and the start of its refill:
    //   fetch  (one pass AHEAD): ticket -> (tensor, sample, frame, band); the draws that are loads -- ring slot, shift
    //          offsets, background row -- and, by all 32 lanes, three of the 81 conv weights per unit, into registers;
    //   finish (when the pass is due): address arithmetic on those registers, then for each unit in order: wait for
    //          the A-warps to release the stage, store the weights, write the meta, start the bulk copies.
    // Tickets: guided self-scheduling on one global counter.  A CTA claims a CHUNK of consecutive tickets per atomic --
    // about 1/(2 gridDim) of what is left, between one pass (G tickets: fine balance at the tail) and 8 -- and works
    // through it in passes of G; the next chunk is requested while the last pass of the current one is handed out, and
    // the first chunk of every CTA is static (no queue on the counter at launch).  The CTAs thus sweep the unit list as
    // one narrow window (neighbouring bands of a frame are fetched and written at about the same time: DRAM pages stay
    // open), with ~1/8 of the atomics.  Measured alternatives: one ticket per atomic -- a same-address atomic is served
    // every ~11 cycles, so the 43 k claims of a batch-1024 launch took as long as the kernel and every producer saw
    // ~5 k cycles of queueing per ticket; 3/4 of the passes static round-robin + a dynamic tail -- no queueing, but the
    // CTAs drift apart over the list and the elementwise chains lost 5-10 % at batch 4096 (0.88 vs 0.97 of HBM peak).
    const uint32_t G = (uint32_t)p.batch;   // 1 .. NS
    const uint32_t npair = p.interleave ? (uint32_t)min(p.units_obs, p.units_next) : 0u;
    const int64_t P = (int64_t)H * W;
    auto chunk_for = [&](uint32_t claimed) -> uint32_t {   // tickets to claim when `claimed` of them are already handed out
      const uint32_t left = claimed < total_units ? total_units - claimed : 0u;
      uint32_t c = left / (2u * gridDim.x);
      c = c > 8u ? 8u : c;
      c -= c % G;
      return c < G ? G : c;
    };
    const uint32_t c0 = chunk_for(0);
    uint32_t cur = blockIdx.x * c0, end = cur + c0;   // static first chunk; the counter hands out dyn0 + ...
    const uint32_t dyn0 = gridDim.x * c0;
    // up to two claims in flight (an atomic on the contended counter takes longer than one short pass)
    uint32_t nb0 = 0, ns0 = 0, nb1 = 0, ns1 = 0, claimed = dyn0;   // claimed: newest known value of the global counter
    int npend = 0;
    bool pdl_done = false;
    auto request = [&]() {
      const uint32_t sz = chunk_for(claimed);
      uint32_t b = 0;
      if (!pdl_done) { pdl_wait(); pdl_done = true; }   // the predecessor re-arms the counter at its very end
      if (lane == 0) b = atomicAdd(p.sched, sz);
      if (npend == 0) { nb0 = b; ns0 = sz; } else { nb1 = b; ns1 = sz; }
      ++npend;
      claimed += sz;   // a lower bound of the counter after this claim (other CTAs claim too)
    };
    auto next_pass = [&]() -> uint32_t {   // first ticket of the next pass (warp-uniform)
      if (cur == end) {
        if (npend == 0) request();   // (not reached: requests are always outstanding when the local range runs dry)
        cur = dyn0 + __shfl_sync(0xffffffffu, nb0, 0);
        end = cur + ns0;
        if (end > claimed) claimed = end;
        nb0 = nb1; ns0 = ns1;
        --npend;
      }
      const uint32_t t0 = cur;
      cur += G;
      // keep two passes' worth of tickets in hand or on the way
      while (npend < 2 && (end - cur) + (npend > 0 ? ns0 : 0u) <= G) request();
      return t0;
    };
    struct Draw {
      int live, dy, dx;
      int t, band, b, f;
      uint32_t fr;
      int64_t slot, row;
    };
    auto fetch = [&](uint32_t t0) -> Draw {
      Draw d{};
      uint32_t u = t0 + (uint32_t)lane;
      d.live = (uint32_t)lane < G && u < total_units;
      if (d.live) {
        // ticket -> unit: obs and next_obs units alternate (conv chains: the A-warps copy a next_obs band while the
        // B-warps convolve the obs band before it; neither role has a phase without work), then the longer list's rest
        if (u < 2u * npair) { d.t = (int)(u & 1u); u >>= 1; }
        else if (p.interleave) { d.t = p.units_obs > p.units_next ? 0 : 1; u -= npair; }
        else { d.t = u >= (uint32_t)p.units_obs ? 1 : 0; if (d.t) u -= (uint32_t)p.units_obs; }   // obs units first
        d.fr = udiv_small(u, p.rcp_bands);   // frame row of x.reshape(B*k, 3, H, W)
        d.band = (int)(u - d.fr * (uint32_t)p.bands);
        d.b = (int)udiv_small(d.fr, p.rcp_k);
        d.f = (int)(d.fr - (uint32_t)d.b * (uint32_t)p.k);
        const int t = d.t, b = d.b;
        // raw draws only: nothing here may USE a loaded value (that would stall the fetch on the load)
        if (p.sp.idx_mode == kSrcInline) d.slot = (int64_t)inl.idx[b];
        else if (p.sp.idx_mode == AUGCL_SRC_PHILOX)   // same values as augcl_rng_fill_indices(out, B, 0, idx_hi, seed, subseq)
          d.slot = (int64_t)(((uint64_t)philox_word((uint64_t)b, p.sp.subseq, p.sp.seed) * p.sp.idx_hi) >> 32);
        else d.slot = p.sp.idx[b];
        const int om = p.sp.off_mode[t];
        d.dy = p.sp.pad; d.dx = p.sp.pad;   // kSrcNone: no shift
        if (om == AUGCL_SRC_PHILOX) {        // same values as augcl_rng_fill_offsets(out, 2B, 0, off_range, seed, subseq+1+t)
          const uint4 r4 = philox4x32_10((uint64_t)(b >> 1), p.sp.subseq + 1 + t, p.sp.seed);
          d.dy = (int)__umulhi((b & 1) ? r4.z : r4.x, p.sp.off_range);
          d.dx = (int)__umulhi((b & 1) ? r4.w : r4.y, p.sp.off_range);
        } else if (om == kSrcInline) {
          d.dy = inl.off[t][b][0]; d.dx = inl.off[t][b][1];
        } else if (om != kSrcNone) {
          d.dy = p.sp.off[t][2 * b]; d.dx = p.sp.off[t][2 * b + 1];
        }
        if (kBank && t == 0 && p.o_aug != nullptr) {
          // batch-tiled (splice: frame row fr takes background fr mod B) vs channel-tiled (overlay: the sample's)
          const int64_t j = kSplice ? (int64_t)(d.fr - udiv_small(d.fr, p.rcp_B) * (uint32_t)p.sp.B) : (int64_t)b;
          d.row = p.bank_idx ? p.bank_idx[j]
                             : (int64_t)(((uint64_t)philox_word((uint64_t)j, p.bank_subseq, p.sp.seed) * (uint64_t)p.bank_len) >> 32);
        }
      }
      return d;
    };
    Draw nx = fetch(next_pass());
    uint32_t un = 0;   // units issued so far
    int s = 0;
    uint32_t eph = 1;   // parity of the previous phase of empty[s]: passes at once while the ring fills for the first time
    for (bool more = true; more;) {
      const Draw d = nx;
      const uint32_t lmask = __ballot_sync(0xffffffffu, d.live);
      if (pdl_done && lmask) nx = fetch(next_pass());   // the next pass's loads travel during this one
      FusedMeta m{};
      uintptr_t a0 = 0, g0 = 0;
      uint32_t sbytes = 0, gbytes = 0;
      if (d.live) {
        const int t = d.t, band = d.band, b = d.b, f = d.f;
        const int BAND = CONV ? kFBand : p.band;
        const int r0 = band * BAND, rows = min(BAND, H - r0);
        const int dy = d.dy - p.sp.pad, dx = d.dx - p.sp.pad;   // source coordinate = output coordinate + (dy, dx), clamped
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
        const size_t fo = (((size_t)b * p.sp.C + (size_t)f * 3) * H) * W;
        m.o_plain = want_plain ? p.o_plain[t] + fo : nullptr;
        m.o_shift = want_shift ? p.o_shift[t] + fo : nullptr;
        m.o_aug = want_aug ? p.o_aug + fo : nullptr;
        m.dy = dy; m.dx = dx; m.r0 = r0; m.rows = rows; m.sample = b; m.pad_[0] = 0; m.pad_[1] = 0; m.pad_[2] = 0;
        const uint8_t* src = plane_ptr(p.sp, t, clamp_idx(d.slot, p.sp.capacity), f * 3, P);   // plane 0 of the frame (its 3 planes are contiguous)
        a0 = reinterpret_cast<uintptr_t>(src + (int64_t)lo * W) & ~(uintptr_t)15;
        const uintptr_t a1 = (reinterpret_cast<uintptr_t>(src + (int64_t)(hi + 1) * W) + 15) & ~(uintptr_t)15;
        m.koff = (int)((intptr_t)reinterpret_cast<uintptr_t>(src) - (intptr_t)a0);
        sbytes = (uint32_t)(a1 - a0);
        m.gkoff = 0;
        if (kBank && want_aug) {
          const int64_t row = d.row < 0 ? 0 : (d.row >= p.bank_len ? p.bank_len - 1 : d.row);
          const uint8_t* g = p.bank + row * 3 * P;
          const int glo = max(r0 - hal, 0), ghi = min(r0 + rows - 1 + hal, H - 1);
          g0 = reinterpret_cast<uintptr_t>(g + (int64_t)glo * W) & ~(uintptr_t)15;
          const uintptr_t g1 = (reinterpret_cast<uintptr_t>(g + (int64_t)(ghi + 1) * W) + 15) & ~(uintptr_t)15;
          m.gkoff = (int)((intptr_t)reinterpret_cast<uintptr_t>(g) - (intptr_t)g0);
          gbytes = (uint32_t)(g1 - g0);
        }
      }
      // metas of the pass, all lanes at once: slot (n + i) % 8 last held unit n + i - 8, consumed long ago (G + NS <= 8)
      if (d.live) meta_store(meta + ((un + (uint32_t)lane) % kFMetaRing) * 4, m);
      un += G;
      __syncwarp();
#pragma unroll
      for (int i = 0; i < kFMaxStages; ++i) {
        if ((uint32_t)i >= G) break;
        mbar_wait(&empty[s], eph);   // every A-warp has released the stage's previous unit
        if (!((lmask >> i) & 1u)) {  // end marker
          if (lane == 0) { valid[s] = 0; mbar_arrive(&full[s]); }
          more = false;
          break;
        }
        if (lane == i && elect_in(1u << i)) {   // (single lane, and ptxas knows it: no uniformisation loop around each UBLKCP)
          mbar_expect_tx(&full[s], 3 * (sbytes + gbytes));   // P % 16 == 0: the three planes share lead and length
          uint8_t* sb = stage + (size_t)s * p.stage_bytes;
#pragma unroll
          for (int ci = 0; ci < 3; ++ci)
            bulk_g2s(sb + ci * p.plane_stage, reinterpret_cast<const void*>(a0 + (uintptr_t)(ci * P)), sbytes, &full[s]);
          if (gbytes) {
            uint8_t* gb = sb + 3 * p.plane_stage;
#pragma unroll
            for (int ci = 0; ci < 3; ++ci)
              bulk_g2s(gb + ci * p.bank_stage, reinterpret_cast<const void*>(g0 + (uintptr_t)(ci * P)), gbytes, &full[s]);
          }
        }
        __syncwarp();
        if (++s == NS) { s = 0; eph ^= 1; }
      }
      if (!pdl_done) {   // the first pass (static tickets) is on its way: now wait for the predecessor, then go on as usual
        pdl_wait();
        pdl_done = true;
        if (lmask && more) nx = fetch(next_pass());
      }
    }
    if (!pdl_done) pdl_wait();
    if (lane == 0) {
      // the last CTA to run out of units re-arms the counters for the next launch on this ring (launches of one ring
      // are stream-ordered, and a PDL successor takes its first ticket only after griddepcontrol.wait)
      __threadfence();
      if (atomicAdd(p.sched + 1, 1u) == gridDim.x - 1) {
        p.sched[0] = 0;
        p.sched[1] = 0;
        __threadfence();
      }
    }
    return;
  }

  if (CONV && warp >= NAW) {
    // =============================== B-warps: conv on tile t ===============================
    const int bt = tid - NTA;
    pdl_wait();
    gather_scalars_part(p.sp, inl, blockIdx.x * NTB + bt, gridDim.x * NTB);
    const int a0 = bt / nq, q4 = bt - a0 * nq;
    const int ARPR = NTB / nq;                      // band rows per round (6 at W = 84: one round)
    const size_t HW = (size_t)H * W;
    for (int tc = 0;; ++tc) {
      const int t = tc & 1;
      mbar_wait(&tfull[t], (uint32_t)(tc >> 1) & 1u);
      const TileMeta tm = tmeta[t];
      if (!tm.valid) break;
      const float* tile = tiles + t * TILEF;
      const float* wt = wsm + t * 108;
      if (p.weights == nullptr) {
        // the sample's 81 N(0,1) weights, element e = 81 * sample + i of the stream augcl_rng_fill_normal(.., seed,
        // conv_subseq) writes: Philox block e / 4, Box-Muller pair (e % 4) / 2, cos for even e, sin for odd
        if (bt < 81) {
          const uint64_t e = (uint64_t)tm.sample * 81u + (uint64_t)bt;
          const uint4 r4 = philox4x32_10(e >> 2, p.conv_subseq, p.sp.seed);
          float nc, ns;
          if (e & 2) normal_pair(r4.z, r4.w, nc, ns); else normal_pair(r4.x, r4.y, nc, ns);
          const int co = bt / 27, tap = bt - co * 27;
          wsm[t * 108 + tap * 4 + co] = ((e & 1) ? ns : nc) * kConvWScale;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(NTB) : "memory");   // the B-warps only
      }
      const int na = min(kFHalf, tm.rows);
      for (int a = (a0 < ARPR && !(p.skip & 1)) ? a0 : na; a < na; a += ARPR) {
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
              const float4 wv = *reinterpret_cast<const float4*>(wt + (ci * 9 + ky * 3 + kx) * 4);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                acc[0][j] = __ffma2_rn(xs[j + kx], make_float2(wv.x, wv.x), acc[0][j]);
                acc[1][j] = __ffma2_rn(xs[j + kx], make_float2(wv.y, wv.y), acc[1][j]);
                acc[2][j] = __ffma2_rn(xs[j + kx], make_float2(wv.z, wv.z), acc[2][j]);
              }
            }
          }
        }
        float* orow = tm.o_aug + (size_t)(tm.r0 + a) * W + 4 * q4;
#pragma unroll
        for (int co = 0; co < 3; ++co)
          st_global_f4(reinterpret_cast<float4*>(orow + co * HW),
                       make_float4(conv_sigmoid<NORM>(acc[co][0].x), conv_sigmoid<NORM>(acc[co][1].x),
                                   conv_sigmoid<NORM>(acc[co][2].x), conv_sigmoid<NORM>(acc[co][3].x)));
        if (a + kFHalf < tm.rows) {
#pragma unroll
          for (int co = 0; co < 3; ++co)
            st_global_f4(reinterpret_cast<float4*>(orow + co * HW + (size_t)kFHalf * W),
                         make_float4(conv_sigmoid<NORM>(acc[co][0].y), conv_sigmoid<NORM>(acc[co][1].y),
                                     conv_sigmoid<NORM>(acc[co][2].y), conv_sigmoid<NORM>(acc[co][3].y)));
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[t]);
    }
    return;
  }

  // =============================== A-warps: elementwise phase ===============================
  bool pdl_done = false;
  if (!CONV) { pdl_wait(); pdl_done = true; gather_scalars_part(p.sp, inl, blockIdx.x * NTA + tid, gridDim.x * NTA); }
  // items of a unit = (row, 4-pixel group), handed out flat over the NTA threads: no per-item division
  const int pr0 = tid / nq, q0 = tid - pr0 * nq;
  const int dpr = NTA / nq, dq = NTA - dpr * nq;
  const size_t HW = (size_t)H * W;
  const int PS = p.plane_stage;
  int s = 0, tcount = 0;
  uint32_t fph = 0, un = 0;   // un: units consumed so far
  for (;;) {
    mbar_wait(&full[s], fph);
    if (!pdl_done) { pdl_wait(); pdl_done = true; }   // first stores below
    if (!valid[s]) break;
    const FusedMeta m = meta_load(meta + (un % kFMetaRing) * 4);
    ++un;
    // conv weights[sample][co][ci][ky][kx]: one per thread, requested now and stored (scaled, transposed) after the items
    float wreg = 0.0f;
    if (CONV && p.weights != nullptr && m.o_aug != nullptr && tid < 81) wreg = __ldg(p.weights + (int64_t)m.sample * 81 + tid);
    const uint8_t* sb = stage + (size_t)s * p.stage_bytes;

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
      for (int pr = pr0, q4 = q0; pr < m.rows;) {
        const int r = m.r0 + pr;
        const size_t oo = (size_t)r * W + 4 * q4;
        uint32_t w[3];
        if (m.o_plain) { load_plain(r, q4, w); store3(m.o_plain, oo, w); }
        if (m.o_shift) { load_shift(r, q4, w); store3(m.o_shift, oo, w); }
        q4 += dq; pr += dpr;
        if (q4 >= nq) { q4 -= nq; ++pr; }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
      if (++s == NS) { s = 0; fph ^= 1; }
      continue;
    }

    // ---------------- augmented obs unit: elementwise chain (conv chains: phase A into tile t) ----------------
    const uint8_t* gb = sb + 3 * PS;
    const int trows = m.rows + 2 * hal;
    const bool aug_sh = p.aug_shifted != 0;
    const int t = tcount & 1;
    float* tile = tiles + t * TILEF;
    if (CONV) {
      if (tcount >= 2) mbar_wait(&tempty[t], (uint32_t)((tcount >> 1) - 1) & 1u);   // the B-warps have left tile t
      if (tid == 0) { TileMeta tm; tm.o_aug = m.o_aug; tm.r0 = m.r0; tm.rows = m.rows; tm.valid = 1; tm.sample = m.sample; tmeta[t] = tm; }
    }
    // one row of one 4-pixel group: the direct outputs of the row (plain / shifted) are stored, v = the chain's value
    // before the conv (conv chains) or the augmented output (other chains)
    auto row_values = [&](int pr, int q4, float (&v)[3][4], size_t& oo_out) {
        const int r = clampi(m.r0 + pr - hal, 0, H - 1);
        const bool own = (!CONV || (pr >= 1 && pr <= m.rows)) && !(p.skip & 4);   // halo rows only feed the conv tile
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
      oo_out = oo;
    };
    if (!CONV) {
      for (int pr = pr0, q4 = q0; pr < trows;) {
        float v[3][4];
        size_t oo;
        row_values(pr, q4, v, oo);
#pragma unroll
        for (int ci = 0; ci < 3; ++ci)
          st_global_f4(reinterpret_cast<float4*>(m.o_aug + oo + ci * HW),
                       fin4<NORM>(make_float4(v[ci][0], v[ci][1], v[ci][2], v[ci][3])));
        q4 += dq; pr += dpr;
        if (q4 >= nq) { q4 -= nq; ++pr; }
      }
    } else {
      // Conv chains.  An item is a 4-pixel group of TWO tile rows that share a row pair of the tile -- (g, g + kFHalf) for
      // g < kFHalf -- so the thread holds both lanes of every float2 and stores whole 16- and 8-byte pieces (A1[q4]; the
      // halves of A0[q4], A0[q4+1]): 3 stores per channel for two rows where single rows need 8 four-byte stores that
      // conflict 4-way (the tile stores were 13 % of the kernel in the timing experiments).  The last two tile rows
      // (item row kFHalf) are lane .y of row pairs kFHalf, kFHalf + 1, whose lane .x -- tile rows kFHalf, kFHalf + 1 -- is
      // stored a second time by the items that computed them as lane .y of pairs 0, 1.
      auto put = [&](float* tb, const float (&v)[3][4], int q4) {   // one lane (.x or .y) of a row pair, four-byte stores
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) {
          float* tp = tb + ci * kFPairs * TROW;
          tp[2] = v[ci][0]; tp[TA1] = v[ci][1]; tp[TA1 + 2] = v[ci][2]; tp[4] = v[ci][3];
          if (q4 == 0) tp[0] = v[ci][0];              // column -1: replicate
          if (q4 == nq - 1) tp[6] = v[ci][3];         // column W
        }
      };
      const int nitem = (kFHalf + 1) * nq;
      for (int it = tid; it < nitem; it += NTA) {
        const int g = it / nq, q4 = it - g * nq;
        const int r1 = g < kFHalf ? g : 2 * kFHalf, r2 = g < kFHalf ? g + kFHalf : 2 * kFHalf + 1;   // tile rows of the item
        float vlo[3][4], v[3][4];
        size_t oo;
        if (r1 < trows) row_values(r1, q4, vlo, oo);
        if (r2 < trows) row_values(r2, q4, v, oo);
        if (p.skip & 2) continue;
        float* tb = tile + 4 * q4;
        if (g < kFHalf) {
          if (r2 < trows) {          // both lanes of row pair g
#pragma unroll
            for (int ci = 0; ci < 3; ++ci) {
              float* tp = tb + (ci * kFPairs + g) * TROW;
              *reinterpret_cast<float2*>(tp + 2) = make_float2(vlo[ci][0], v[ci][0]);
              *reinterpret_cast<float4*>(tp + TA1) = make_float4(vlo[ci][1], v[ci][1], vlo[ci][2], v[ci][2]);
              *reinterpret_cast<float2*>(tp + 4) = make_float2(vlo[ci][3], v[ci][3]);
              if (q4 == 0) *reinterpret_cast<float2*>(tp) = make_float2(vlo[ci][0], v[ci][0]);
              if (q4 == nq - 1) *reinterpret_cast<float2*>(tp + 6) = make_float2(vlo[ci][3], v[ci][3]);
            }
            if (g < 2) put(tb + (g + kFHalf) * TROW, v, q4);      // tile row kFHalf + g is also lane .x of pair kFHalf + g
          } else if (r1 < trows) {   // a short band: lane .x only
            put(tb + g * TROW, vlo, q4);
          }
        } else {                     // tile rows 2 kFHalf, 2 kFHalf + 1: lane .y of pairs kFHalf, kFHalf + 1
          if (r1 < trows) put(tb + kFHalf * TROW + 1, vlo, q4);
          if (r2 < trows) put(tb + (kFHalf + 1) * TROW + 1, v, q4);
        }
      }
    }
    if (CONV && p.weights != nullptr && tid < 81) {   // -> [(ci*9 + ky*3 + kx)*4 + co], scaled for conv_sigmoid
      const int co = tid / 27, tap = tid - co * 27;
      wsm[t * 108 + tap * 4 + co] = wreg * kConvWScale;
    }
    __syncwarp();
    if (lane == 0) {
      mbar_arrive(&empty[s]);
      if (CONV) mbar_arrive(&tfull[t]);
    }
    ++tcount;
    if (++s == NS) { s = 0; fph ^= 1; }
  }
  if (CONV) {
    // end marker for the B-warps, in the tile slot the next unit would have used
    const int t = tcount & 1;
    if (tcount >= 2) mbar_wait(&tempty[t], (uint32_t)((tcount >> 1) - 1) & 1u);
    if (tid == 0) tmeta[t].valid = 0;
    __syncwarp();
    if (lane == 0) mbar_arrive(&tfull[t]);
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
  p.bank = d->bank; p.bank_idx = d->bank_idx; p.bank_subseq = d->bank_subsequence; p.conv_subseq = d->conv_subsequence; p.bank_len = d->bank_len; p.weights = d->conv_weights;
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
  p.normalize = d->normalize ? 1 : 0;
  const bool has_conv = (ops & AUGCL_OP_CONV) != 0;
  const int halo2 = has_conv ? 2 : 0;
  const int rows_max = (band + halo2 + d->pad) < r->H ? (band + halo2 + d->pad) : r->H;
  p.plane_stage = ((rows_max * r->W + 30 + 15) & ~15) + 16;
  p.bank_stage = (ops & 7) ? ((((band + halo2) * r->W + 30 + 15) & ~15) + 16) : 0;
  p.stage_bytes = 3 * (p.plane_stage + p.bank_stage);
  p.tile_rs = tile_row_floats(r->W);
  const size_t tile_bytes = has_conv ? 2 * 3 * (size_t)kFPairs * p.tile_rs * sizeof(float) : 0;   // two tiles
  // stage ring depth: the producer warp runs this many units ahead of the A-warps (scripts/bench_fused.py --opts)
  p.stages = g_fused_stages.load() > 0 ? g_fused_stages.load() : ((has_conv && B >= 512) ? 4 : 3);   // measured (r03h sweep)
  if (p.stages > kFMaxStages) p.stages = kFMaxStages;
  const size_t budget = (228 * 1024) / (has_conv ? 3 : 4) - 1024;   // 3 (conv) / 4 CTAs per SM
  while (p.stages > 1 && kFHeader + (size_t)p.stages * p.stage_bytes + tile_bytes > budget) --p.stages;
  const size_t smem = kFHeader + (size_t)p.stages * p.stage_bytes + tile_bytes;
  p.sched = sched_for(r, (cudaStream_t)stream);
  if (!p.sched) AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "%s: a ring can be sampled from at most %d streams", who, kSchedSlots);
  p.rcp_bands = 1.0 / (double)p.bands; p.rcp_k = 1.0 / (double)p.k; p.rcp_B = 1.0 / (double)B;
  // conv chains: pairs (an obs band for the B-warps + a next_obs band) so every pass feeds both roles; other chains: one
  p.batch = g_fused_batch.load() > 0 ? g_fused_batch.load() : (has_conv ? 2 : 1);
  if (p.batch > p.stages) p.batch = p.stages;
  p.skip = g_fused_skip.load();
  p.interleave = g_fused_interleave.load() >= 0 ? g_fused_interleave.load() : ((has_conv || B <= 512) ? 1 : 0);
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
  if (per_sm <= 0) per_sm = has_conv ? 3 : 4;
  const int fit = (int)((228 * 1024) / (smem + 1024));
  if (per_sm > fit) per_sm = fit < 1 ? 1 : fit;
  int64_t grid = (int64_t)kNumSMs * per_sm;
  const int64_t units = p.units_obs + p.units_next;
  if (grid > units) grid = units;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(fused_threads(has_conv));
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
