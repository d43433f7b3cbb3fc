/*
 * augcl_b200 — C ABI of the B200-native replay-gather + image-augmentation hot path.
 *
 * The reference (GT-RIPL/AugCL) exposes this path as Python functions, not as an FFI
 * (SURVEY.md §8b).  Each entry point below names the reference interface it replaces
 * (file:line under /root/reference); the Python modules augcl_b200/augmentations.py and
 * augcl_b200/utils.py mirror those interfaces 1:1 and bind this library through ctypes
 * (INTEGRATION.md shows the stub a reference maintainer would add).
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types cross the boundary.
 *   - every function returns 0 (AUGCL_OK) or a negative augcl_status; the message of
 *     the last failure on the calling thread is returned by augcl_last_error().
 *   - functions never throw, never allocate outputs and never synchronise the device:
 *     work is enqueued on `stream` (a cudaStream_t passed as void*; NULL = legacy default
 *     stream).  The caller owns every buffer; the library owns only a ring's storage.
 *   - image tensors are contiguous NCHW.  Pixels are f32 in 0..255 exactly like the
 *     reference's augmentations (the /255 lives in its encoder, src/algorithms/modules.py:85-90);
 *     `normalize != 0` fuses that division into the kernel (out = v / 255.0f, IEEE-rounded).
 *   - random draws are always *inputs* (indices, offsets, conv weights, background rows) so
 *     results are comparable with the reference "given identical draws"; the augcl_rng_*
 *     functions generate such draws on the device (Philox4x32-10) when the caller has none.
 *   - device pointers must belong to the calling thread's current CUDA device, except ring
 *     calls, which switch to the ring's device for the duration of the call.
 *   - threading: the stateless augmentation entry points may be called from any thread on any
 *     stream.  A ring handle is single-writer like the reference's ReplayBuffer (one Python
 *     thread, src/train.py): put / fill calls on ONE ring are issued in order on one stream.
 *     SAMPLING one ring from several streams at once is supported (up to 16 streams per ring:
 *     the fused sampler keeps one set of schedule counters per stream, AUGCL_ERR_UNSUPPORTED
 *     beyond that; tests/test_gpu_fused.py); the caller orders samples after the puts they
 *     should see.  Different rings are independent.  CUDA-graph capture of ring calls is
 *     supported (tests/test_gpu_graph.py).
 */
#ifndef AUGCL_B200_H_
#define AUGCL_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct augcl_ring augcl_ring;   /* opaque: device-resident replay ring  */
typedef void* augcl_stream;             /* cudaStream_t                          */

typedef enum augcl_status {
  AUGCL_OK = 0,
  AUGCL_ERR_BAD_ARG = -1,    /* null pointer, non-positive size, value out of range    */
  AUGCL_ERR_BAD_ALIGN = -2,  /* pointer not aligned as the entry point requires         */
  AUGCL_ERR_CUDA = -3,       /* a CUDA runtime call failed; see augcl_last_error()      */
  AUGCL_ERR_OOM = -4,        /* device allocation for a ring failed                     */
  AUGCL_ERR_UNSUPPORTED = -5 /* shape outside what the kernels handle                   */
} augcl_status;

/* aug_mode of augcl_ring_sample */
/* SHIFT_OBS: random_shift on obs only, next_obs gathered un-augmented (the SVEA / AugCL update,
 * src/algorithms/svea.py:66-70, augcl.py:69-72); off_next is ignored. */
enum { AUGCL_AUG_NONE = 0, AUGCL_AUG_SHIFT = 1, AUGCL_AUG_CROP = 2, AUGCL_AUG_SHIFT_OBS = 3 };
/* mode of augcl_random_shift */
enum { AUGCL_SHIFT_INT = 0, AUGCL_SHIFT_BILINEAR = 1 };
/* bank_dtype of overlay / splice */
enum { AUGCL_BANK_U8 = 0, AUGCL_BANK_F32 = 1 };
/* variant of augcl_splice */
enum { AUGCL_SPLICE = 0, AUGCL_CS_SPLICE = 1 };

const char* augcl_version(void);
const char* augcl_last_error(void);
/* Number of kernel launches this library has enqueued from the calling process so far
 * (bench.py's "gpu_launches" evidence). */
int64_t augcl_launch_count(void);
/* Measurement helper (no reference analogue): enqueue a one-thread kernel that keeps `stream` busy for about
 * `nanoseconds` (<= 50 ms; it waits on the GPU's clock only).  bench.py puts it in front of the timed region so that every
 * timed launch is already queued when the GPU reaches the first one.  Not counted by augcl_launch_count. */
int augcl_stream_delay(uint64_t nanoseconds, augcl_stream stream);
/* Runtime switches, for A/B measurement (csrc/state.cu holds the table; integers are passed as decimal strings, 0 = auto
 * where a count is meant).  "staging" = "tma" (default; cp.async.bulk + mbarrier pipeline through shared memory) |
 * "generic" (one thread per output element, no staging).  Sample kernels: "sample_stages", "sample_ctas_per_sm",
 * "sample_pdl" (programmatic dependent launch, default 1), "sample_band", "sample_fused_min_batch" (batches from here up
 * run on the fused sampler's copy kernel; default 384).  Fused sampler: "fused_stages" (ring depth, <= 4), "fused_batch"
 * (units the producer warp decodes side by side), "fused_ctas_per_sm", "fused_band" (rows per unit, non-conv chains),
 * "fused_interleave" (obs / next_obs units alternate; -1 auto), "fused_fast_mask" (byte-domain splice mask, default 1),
 * "fused_skip" (TIMING EXPERIMENTS ONLY -- switches parts of the conv chain off, results are wrong).  Stand-alone kernels:
 * "conv_part", "bilinear_tma". */
int augcl_set_option(const char* key, const char* value);

/* ------------------------------------------------------------------------------------------
 * Device-resident replay ring.
 * Replaces the host storage of utils.ReplayBuffer (src/utils.py:139-154): the Python list
 * `_obses` of (LazyFrames, LazyFrames) and the numpy `actions/rewards/not_dones` arrays.
 * Layout in HBM ("v1", full stacks):  obs[capacity][k*cpf][H][W] u8, next_obs[...] same,
 * actions[capacity][action_dim] f32, rewards[capacity] f32, not_dones[capacity] f32.
 * The write cursor (`idx`, `full`) stays in the caller (utils.ReplayBuffer keeps the
 * reference's attributes of those names, which src/train.py reads and writes).
 * ------------------------------------------------------------------------------------------ */
int augcl_ring_create(augcl_ring** out, int device, int64_t capacity, int k, int c_per_frame,
                      int H, int W, int action_dim);
int augcl_ring_destroy(augcl_ring* ring);
/* bytes of HBM the ring holds */
int64_t augcl_ring_bytes(const augcl_ring* ring);

/* Store `count` transitions at slots [slot0, slot0+count) (no wrap: slot0+count <= capacity).
 * Replaces ReplayBuffer.add (src/utils.py:216-227) for count==1 and the chunk re-load of
 * ReplayBuffer.load (src/utils.py:177-204) for count>1.  Source pointers may be host or device
 * memory (cudaMemcpyDefault); obs/next are count*k*cpf*H*W bytes, actions count*action_dim f32,
 * rewards / not_dones count f32.  Any source pointer may be NULL to leave that field untouched. */
int augcl_ring_put(augcl_ring* ring, int64_t slot0, int64_t count, const uint8_t* obs,
                   const uint8_t* next_obs, const float* actions, const float* rewards,
                   const float* not_dones, augcl_stream stream);
/* Inverse of put (drain for ReplayBuffer.save, src/utils.py:156-170).  NULL pointers are skipped. */
int augcl_ring_get(augcl_ring* ring, int64_t slot0, int64_t count, uint8_t* obs, uint8_t* next_obs,
                   float* actions, float* rewards, float* not_dones, augcl_stream stream);
/* De-duplicated ring ("v2" layout).  The reference stores transitions as (LazyFrames, LazyFrames) whose frame arrays are
 * SHARED between obs / next_obs and between consecutive steps (FrameStack, src/env/wrappers.py:183-203; LazyFrames,
 * src/utils.py:283-322): 2k references per transition, ~1 new frame.  This layout keeps that sharing on the device:
 *   frames[pool_frames][c_per_frame][H][W] u8   (one copy of every distinct frame)
 *   refs[capacity][2][k] int32                   (pool slot of each frame of obs / next_obs)
 * 100 000 transitions of 9x84x84 take ~2.1 GB instead of 12.7 GB and an env step uploads one 21 KB frame instead of two
 * 63 KB stacks; a sample reads k+1 distinct frames (the second read of a shared frame hits L2): compulsory traffic
 * 592 704 B per sample instead of 635 040 B.  augcl_ring_sample* / augcl_ring_sample_fused work on both layouts.
 * The caller owns the pool cursor and must not overwrite a frame a live transition still references
 * (augcl_b200.utils.ReplayBuffer(dedup=True) does that bookkeeping).  augcl_ring_put / _get of image stacks return
 * AUGCL_ERR_UNSUPPORTED on such a ring (scalars still work). */
int augcl_ring_create_dedup(augcl_ring** out, int device, int64_t capacity, int k, int c_per_frame, int H, int W,
                            int action_dim, int64_t pool_frames);
/* Store `count` frames (count*c_per_frame*H*W bytes, host or device) at pool slots [frame0, frame0+count) (no wrap). */
int augcl_ring_put_frames(augcl_ring* ring, int64_t frame0, int64_t count, const uint8_t* frames, augcl_stream stream);
/* Store the frame references (int32 [count][2][k]: obs frames then next_obs frames, oldest first) and the scalars of
 * transitions [slot0, slot0+count).  Host references are range-checked.  NULL pointers leave that field untouched. */
int augcl_ring_put_refs(augcl_ring* ring, int64_t slot0, int64_t count, const int32_t* refs, const float* actions,
                        const float* rewards, const float* not_dones, augcl_stream stream);

/* Fill slots [0, count) with Philox bytes / normals (benchmarks and tests; no reference analogue). */
int augcl_ring_fill_synthetic(augcl_ring* ring, int64_t count, uint64_t seed, augcl_stream stream);

/* Minibatch gather + u8->f32 (+ fused random_shift or random_crop).
 * Replaces ReplayBuffer._encode_obses + tensor_buffer_samples (src/utils.py:206-214,234-240),
 * i.e. the body of sample / sample_obs / sample_curl / sample_drq (src/utils.py:242-280), and,
 * when aug_mode != NONE, the random_shift / random_crop call that follows it
 * (src/augmentations.py:378-382, 385-413).
 *   idx        device int64[B], each in [0, capacity)  (clamped for memory safety)
 *   out_obs / out_next   device f32 [B, k*cpf, S, S'] with (S,S') = (H,W) for NONE/SHIFT and
 *              (size,size) for CROP; either may be NULL to skip that tensor (sample_obs)
 *   out_act [B,action_dim], out_rew [B], out_nd [B]  device f32, each may be NULL
 *   off_obs / off_next   device int32[B,2] = (row, col) offsets.  SHIFT: in [0, 2*pad], the crop
 *              origin inside the replicate-padded image (pad,pad == identity); NULL = identity for
 *              that tensor.  CROP: (w1, h1) of the reference (w1 indexes rows), in [0, H-size) x
 *              [0, W-size) -- the reference's window view has H-size positions, the last geometric one is excluded
 *              (src/augmentations.py:416-433); HOST offsets outside are AUGCL_ERR_BAD_ARG, DEVICE offsets are trusted
 *              (clamped per pixel for memory safety); must not be NULL for a tensor that is written.
 *   pad_or_size  pad (SHIFT) or size (CROP); ignored for NONE.
 * Stream ordering.  The sample kernels (this call, _ex, _curl, _views, _fused) are launched with programmatic stream
 * serialization (option "sample_pdl", default on): when the previous kernel of the stream is itself a sample kernel,
 * the read-only start of the next one -- the draws, the ring / refs / bank reads, its first bulk copies -- overlaps the
 * tail of the previous one; every store, and every access to the ring's schedule counters, waits for the previous
 * kernel to complete (griddepcontrol.wait).  After any other kernel or copy the launch is serialized as usual: a grid can
 * only start early behind a predecessor that executed griddepcontrol.launch_dependents, and the only kernels that do are
 * this library's (sample and stand-alone augmentation kernels), which write nothing but their float outputs -- never the
 * ring, the refs table, the background bank or a draw array.  Draws or ring data written by any other kernel right before
 * the call are therefore complete and visible (tests/test_gpu_replay_parity.py). */
int augcl_ring_sample(augcl_ring* ring, const int64_t* idx, int B, float* out_obs, float* out_next,
                      float* out_act, float* out_rew, float* out_nd, int aug_mode,
                      const int32_t* off_obs, const int32_t* off_next, int pad_or_size,
                      int normalize, augcl_stream stream);

/* Where augcl_ring_sample_ex takes the sampled indices / the offsets from. */
enum { AUGCL_SRC_DEVICE = 0, AUGCL_SRC_HOST = 1, AUGCL_SRC_PHILOX = 2 };

/* augcl_ring_sample with the draws taken from the HOST or generated inside the kernel.  This is the
 * entry point behind ReplayBuffer.sample*() when the caller holds numpy indices, as the reference does
 * (np.random.randint on the host, src/utils.py:229-232, then H2D of the gathered batch, :206-214).
 *   idx_src  DEVICE: idx = device int64[B] (== augcl_ring_sample)
 *            HOST:   idx = host int64[B], consumed before the call returns; every value must be in
 *                    [0, capacity) (AUGCL_ERR_BAD_ARG otherwise, like the reference's IndexError).  For
 *                    B <= 256 and the 84x84 / 100x100 geometries the indices (and HOST offsets) travel BY
 *                    VALUE in the kernel's parameter block: no separate H2D copy; otherwise they are staged
 *                    through stream-ordered scratch (cudaMallocAsync) with one cudaMemcpyAsync per array.
 *            PHILOX: idx ignored; slot of sample b = (philox_word(b; seed, subsequence) * idx_hi) >> 32,
 *                    identical to augcl_rng_fill_indices(out, B, 0, idx_hi, seed, subsequence).
 *   off_src  DEVICE / HOST: off_obs / off_next = int32[B,2] there (NULL = identity for SHIFT).
 *            PHILOX: offsets of a tensor whose pointer is NULL are drawn in the kernel, uniform in
 *                    [0, 2*pad] (SHIFT) or [0, W-size) (CROP; the reference's exclusive bound), identical to
 *                    augcl_rng_fill_offsets(out, 2*B, 0, range, seed, subsequence + 1) for obs and
 *                    (…, subsequence + 2) for next_obs; a non-NULL pointer is a DEVICE array and wins.
 * All other arguments as in augcl_ring_sample. */
int augcl_ring_sample_ex(augcl_ring* ring, int idx_src, const int64_t* idx, int64_t idx_hi, int B,
                         float* out_obs, float* out_next, float* out_act, float* out_rew, float* out_nd,
                         int aug_mode, int off_src, const int32_t* off_obs, const int32_t* off_next,
                         int pad_or_size, int normalize, uint64_t seed, uint64_t subsequence,
                         augcl_stream stream);

/* ReplayBuffer.sample_curl (src/utils.py:247-258) in ONE launch: three independent random crops of the same gather --
 * obs, next_obs and pos = random_crop(obs.clone()) (src/augmentations.py:385-413) -- each with its own (w1, h1) offsets
 * (int32[B,2]; PHILOX draws use subsequence + 1 / + 2 / + 3).  Any of the three outputs may be NULL.  Other arguments as
 * in augcl_ring_sample_ex with aug_mode = CROP. */
int augcl_ring_sample_curl(augcl_ring* ring, int idx_src, const int64_t* idx, int64_t idx_hi, int B, float* out_obs,
                           float* out_next, float* out_pos, float* out_act, float* out_rew, float* out_nd, int off_src,
                           const int32_t* off_obs, const int32_t* off_next, const int32_t* off_pos, int size,
                           int normalize, uint64_t seed, uint64_t subsequence, augcl_stream stream);

/* One gather, up to 6 independently shifted / cropped views: the K augmented copies of obs and M of next_obs that a
 * DrQ update builds with K + M calls of random_shift after sample() (src/algorithms/drq.py:67-70; K = drq_k, M = drq_m).
 *   view_src[v]  0 = obs, 1 = next_obs        outs[v]  device f32 [B, C, Ho, Wo]
 *   offs         NULL, or offs[v] = int32[B,2] offsets of view v (NULL entry: identity, or a PHILOX draw from
 *                subsequence + 3 + v when off_src == PHILOX)
 * aug_mode NONE / SHIFT / CROP; the other arguments as in augcl_ring_sample_ex.  Every u8 plane leaves HBM once; the
 * later views read it from L2. */
int augcl_ring_sample_views(augcl_ring* ring, int idx_src, const int64_t* idx, int64_t idx_hi, int B, int n_views,
                            const int* view_src, float* const* outs, const int32_t* const* offs, float* out_act,
                            float* out_rew, float* out_nd, int aug_mode, int off_src, int pad_or_size, int normalize,
                            uint64_t seed, uint64_t subsequence, augcl_stream stream);

/* ------------------------------------------------------------------------------------------
 * Fused per-algorithm sampler: ONE kernel gathers the minibatch and writes every tensor an update needs.
 * Replaces, for one update() of
 *   SVEA   (src/algorithms/svea.py:66-70, :45-52):  sample() -> obs = random_shift(obs) -> apply_aug(obs.clone())
 *   AugCL  (src/algorithms/augcl.py:69-72):         sample() -> obs_shift = random_shift(obs); obs_aug = apply_aug(obs)
 *   DrQ / RAD-style updates that also want the un-augmented observation,
 * the chain ReplayBuffer.sample (src/utils.py:272-280) + random_shift (src/augmentations.py:378-382) +
 * RAD.apply_aug (src/algorithms/rad.py:18-22) over splice / CS_splice / random_overlay / random_conv
 * (src/augmentations.py:253-324).  Each u8 frame is read from HBM once; no intermediate f32 tensor is written.
 * Algorithmic bytes per sample at 9x84x84, AugCL splice-conv: 2*63 504 + 3*21 168 + 3*254 016 = 952 560.
 *
 * aug_ops chains (applied in this order): [SPLICE | CS_SPLICE | OVERLAY] then [CONV]; 0 = no augmented output.
 * Requirements: W % 4 == 0, (H*W) % 16 == 0, W <= 128, u8 background bank; otherwise AUGCL_ERR_UNSUPPORTED
 * (the caller then chains augcl_ring_sample_ex and the stand-alone kernels).
 * ------------------------------------------------------------------------------------------ */
enum { AUGCL_OP_SPLICE = 1, AUGCL_OP_CS_SPLICE = 2, AUGCL_OP_OVERLAY = 4, AUGCL_OP_CONV = 8 };
enum { AUGCL_AUG_INPUT_PLAIN = 0, AUGCL_AUG_INPUT_SHIFTED = 1 };

typedef struct augcl_fused_desc {
  int32_t struct_size;        /* = sizeof(augcl_fused_desc); guards the ABI                                     */
  int32_t B;                  /* minibatch size                                                                  */
  int32_t idx_src;            /* AUGCL_SRC_*: where idx comes from (as in augcl_ring_sample_ex)                  */
  int32_t off_src;            /* AUGCL_SRC_*: where off_obs / off_next come from                                 */
  const int64_t* idx;         /* int64[B] (DEVICE or HOST); ignored for PHILOX                                   */
  int64_t idx_hi;             /* PHILOX: indices uniform in [0, idx_hi)                                          */
  const int32_t* off_obs;     /* int32[B,2] (row, col) in [0, 2*pad]; NULL = identity, or a draw when PHILOX     */
  const int32_t* off_next;    /* same for out_next_shift                                                         */
  int32_t pad;                /* random_shift pad                                                                */
  int32_t normalize;          /* != 0: every image output is divided by 255 (IEEE)                               */
  uint64_t seed, subsequence; /* Philox streams as in augcl_ring_sample_ex                                       */
  float* out_obs;             /* [B,C,H,W] obs, un-augmented                                 (NULL = not wanted) */
  float* out_obs_shift;       /* [B,C,H,W] random_shift(obs)                                 (NULL = not wanted) */
  float* out_obs_aug;         /* [B,C,H,W] aug_ops chain applied to the aug_input tensor     (NULL = not wanted) */
  float* out_next;            /* [B,C,H,W] next_obs, un-augmented                            (NULL = not wanted) */
  float* out_next_shift;      /* [B,C,H,W] random_shift(next_obs)                            (NULL = not wanted) */
  float* out_act;             /* [B,action_dim] / [B] / [B]; each may be NULL                                    */
  float* out_rew;
  float* out_nd;
  int32_t aug_input;          /* AUGCL_AUG_INPUT_PLAIN (AugCL) | AUGCL_AUG_INPUT_SHIFTED (SVEA)                  */
  int32_t aug_ops;            /* AUGCL_OP_* bitmask                                                              */
  const uint8_t* bank;        /* device [bank_len,3,H,W] u8 background bank (SPLICE / CS_SPLICE / OVERLAY)       */
  int64_t bank_len;
  const int64_t* bank_idx;    /* device int64[B]: overlay uses bank_idx[s] for every frame of sample s; splice   */
                              /* uses bank_idx[(s*k+f) mod B] for frame f (the reference's tiling quirks).       */
                              /* NULL = drawn in the kernel: the values augcl_rng_fill_indices(out, B, 0,        */
                              /* bank_len, seed, bank_subsequence) would write (no extra launch)                 */
  const float* conv_weights;  /* device f32 [B,3,3,3,3] (CONV).  NULL = drawn in the kernel: the values             */
                              /* augcl_rng_fill_normal(out, B*81, seed, conv_subsequence) would write               */
  float hue_thres, sat_thres, val_thres;   /* splice thresholds (src/augmentations.py:295 / :309)               */
  float alpha;                /* overlay alpha (0.5 in the reference)                                            */
  uint64_t bank_subsequence;  /* Philox subsequence of the in-kernel bank row draw (bank_idx == NULL)            */
  uint64_t conv_subsequence;  /* Philox subsequence of the in-kernel conv weight draw (conv_weights == NULL)     */
} augcl_fused_desc;

int augcl_ring_sample_fused(augcl_ring* ring, const augcl_fused_desc* desc, augcl_stream stream);

/* No counterpart in the reference: added for callers that must not allocate
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * augcl_ring_sample_ex with HOST indices / offsets stages batches beyond its inline limit through stream-ordered
 * scratch (cudaMallocFromPoolAsync); this variant stages them through a fixed file-scope device array instead
 * (24 bytes x AUGCL_NOALLOC_MAX_BATCH = 384 KiB, held by every GPU that loads the library until exit) and calls no
 * cudaMalloc / cudaFree / allocating routine.  Same outputs as augcl_ring_sample / _ex for the same draws.
 *   idx          HOST int64[B], each in [0, capacity) (AUGCL_ERR_BAD_ARG otherwise)
 *   off_obs / off_next   HOST int32[B,2] shift offsets in [0, 2*pad], NULL = identity for that tensor
 *   aug_mode     AUGCL_AUG_NONE, AUGCL_AUG_SHIFT or AUGCL_AUG_SHIFT_OBS; outputs as for augcl_ring_sample
 *   B > AUGCL_NOALLOC_MAX_BATCH is rejected with AUGCL_ERR_UNSUPPORTED.
 * All calls (on any ring of a device) share the one array, which keeps the last call's draws; every byte a call reads
 * is rewritten by that call, nothing needs clearing.  Calls that overlap on the GPU (different streams) RACE on it:
 * issue them on one stream, or order them with events. */
#define AUGCL_NOALLOC_MAX_BATCH 16384
int augcl_ring_sample_noalloc(augcl_ring* ring, const int64_t* idx, int B, float* out_obs, float* out_next,
                              float* out_act, float* out_rew, float* out_nd, int aug_mode, const int32_t* off_obs,
                              const int32_t* off_next, int pad_or_size, int normalize, augcl_stream stream);

/* ------------------------------------------------------------------------------------------
 * Stand-alone augmentations on f32 NCHW batches (in and out must not alias).
 * ------------------------------------------------------------------------------------------ */

/* random_shift (src/augmentations.py:378-382): replicate-pad by `pad`, crop (h,w) at offsets[i].
 *   mode INT:      offsets int32[n,2] = (row, col) in [0, 2*pad]; exact copy (kornia slice crop).
 *   mode BILINEAR: DrQ-v2 grid_sample formulation named by the north star: offsets = (row, col)
 *                  integer shifts; `base` = device f32[h] holding linspace(-1+1/(h+2p), 1-1/(h+2p),
 *                  h+2p)[:h] (augcl_drq_base_grid); requires h == w. */
int augcl_random_shift(const float* in, float* out, int n, int c, int h, int w, int pad,
                       const int32_t* offsets, int mode, const float* base, augcl_stream stream);
/* Host helper: writes the h base-grid values of the DrQ-v2 formulation (fp32, torch.linspace order). */
int augcl_drq_base_grid(int h, int pad, float* out_host);

/* random_crop (src/augmentations.py:385-433): out[i] = in[i, :, r:r+size, q:q+size] with
 * (r, q) = offsets[i] = (w1_i, h1_i).  in [n,c,H,W] -> out [n,c,size,size]. */
int augcl_random_crop(const float* in, float* out, int n, int c, int H, int W, int size,
                      const int32_t* offsets, augcl_stream stream);

/* random_conv (src/augmentations.py:265-274): per sample i, for each of its c/3 RGB frames,
 * out = sigmoid(conv2d(replicate_pad1(x/255), weights[i])) * 255, weights f32 [n,3(out),3(in),3,3]. */
int augcl_random_conv(const float* in, float* out, int n, int c, int h, int w, const float* weights,
                      augcl_stream stream);

/* random_overlay (src/augmentations.py:253-262): out = ((1-alpha)*(x/255) + alpha*img)*255 where
 * img = bank[bank_idx[i]] ([3,h,w]; u8 rows are divided by 255 like torchvision ToTensor, f32 rows
 * are used as they are), the same row for every frame of sample i (channel ch uses img[ch % 3]).
 * bank_idx device int64[n] in [0, bank_len). */
int augcl_random_overlay(const float* in, float* out, int n, int c, int h, int w, const void* bank,
                         int bank_dtype, int64_t bank_len, const int64_t* bank_idx, float alpha,
                         augcl_stream stream);

/* splice / CS_splice (src/augmentations.py:295-306, 309-324, mask 277-284): per RGB frame, keep the
 * pixels whose HSV (kornia.color.rgb_to_hsv of x/255) exceeds (hue,sat,val), replace the others by the
 * background pixel; output (.)*255.  Frame f of sample s (row j = s*(c/3)+f of the reshaped batch)
 * takes background bank[bank_idx[j mod n]] — the reference's batch-tiling (fact 7 of SURVEY.md). */
int augcl_splice(const float* in, float* out, int n, int c, int h, int w, const void* bank,
                 int bank_dtype, int64_t bank_len, const int64_t* bank_idx, float hue_thres,
                 float sat_thres, float val_thres, int variant, augcl_stream stream);

/* color_jitter (src/augmentations.py:225-237): torchvision ColorJitter(0, (0,2), (0,3), (-0.5,0.5)).forward on the RGB
 * frames listed in sel (device int32 [n_sel], distinct frame indices into imgs viewed as [frames][3][h][w]), IN PLACE like
 * the reference's `flat[sampled_idxs] = model(flat[sampled_idxs])`.  order = fn_idx (host int32[4], a permutation of
 * 0 brightness, 1 contrast, 2 saturation, 3 hue); enabled bit i = op i has a factor (the reference's model: 0b1110).
 * The factors are ColorJitter.get_params' draws: the caller makes them, like every other draw. */
int augcl_color_jitter(float* imgs, int frames, int h, int w, const int32_t* sel, int n_sel, const int32_t* order,
                       double brightness, double contrast, double saturation, double hue, int enabled,
                       augcl_stream stream);

/* create_hsv_mask (src/augmentations.py:277-284): rgb f32 [n,3,h,w] ALREADY in [0,1] (the reference passes x/255);
 * mask u8 [n,h,w] = 1 where kornia-HSV hue, saturation and value all exceed their thresholds (strictly). */
int augcl_hsv_mask(const float* rgb, uint8_t* mask, int n, int h, int w, float hue_thres, float sat_thres, float val_thres,
                   augcl_stream stream);

/* ------------------------------------------------------------------------------------------
 * The remaining aug_to_func entries (src/augmentations.py:436-449).  Inherited from RAD, where images are
 * in [0,1]; the reference applies them to its 0..255 tensors unchanged and so do these.
 * ------------------------------------------------------------------------------------------ */

/* random_grayscale (src/augmentations.py:111-136, grayscale :88-108): every pixel becomes
 * trunc_u8(x*255)/255 (torch's CPU float->uint8 cast: truncate, low 8 bits); samples with mask[i] != 0
 * additionally replace the three channels of each frame by trunc_u8(0.2989 r + 0.587 g + 0.114 b).
 * mask: device uint8[n] (the reference's np.random.uniform(size=n) <= p). */
int augcl_random_grayscale(const float* in, float* out, int n, int c, int h, int w, const uint8_t* mask,
                           augcl_stream stream);
/* random_flip (src/augmentations.py:139-168): horizontal flip of the samples with flip[i] != 0; device int32[n]. */
int augcl_random_flip(const float* in, float* out, int n, int c, int h, int w, const int32_t* flip,
                      augcl_stream stream);
/* random_rotation (src/augmentations.py:171-213): torch.rot90(x, quarter_turns[i], [2,3]) per sample,
 * quarter_turns device int32[n] in {0,1,2,3} (= rnd_rot * (rnd <= p)); requires h == w. */
int augcl_random_rotation(const float* in, float* out, int n, int c, int h, int w, const int32_t* quarter_turns,
                          augcl_stream stream);
/* random_cutout_color (src/augmentations.py:327-350): box[i] = (w1, h1) device int32[n,2]; rows h1..2*h1-1 and
 * cols w1..2*w1-1 (start == size: the reference's slicing) of channel ch become colors[i*c + ch] (device f32[n,c],
 * the reference's randint(0,255)/255.0). */
int augcl_random_cutout_color(const float* in, float* out, int n, int c, int h, int w, const int32_t* box,
                              const float* colors, augcl_stream stream);

/* ------------------------------------------------------------------------------------------
 * Device RNG (Philox4x32-10, counter = (block_lo, block_hi, subseq_lo, subseq_hi), key = seed).
 * Stand in for the reference's host draws: np.random.randint (src/utils.py:232),
 * LongTensor.random_ (src/augmentations.py:404-405), kornia's crop parameters (:382),
 * torch.randn(3,3,3,3) (:269).  Values: lo + ((u32 * (hi-lo)) >> 32).
 * ------------------------------------------------------------------------------------------ */
int augcl_rng_fill_offsets(int32_t* out, int64_t count, int32_t lo, int32_t hi, uint64_t seed,
                           uint64_t subsequence, augcl_stream stream);
int augcl_rng_fill_indices(int64_t* out, int64_t count, int64_t lo, int64_t hi, uint64_t seed,
                           uint64_t subsequence, augcl_stream stream);
int augcl_rng_fill_normal(float* out, int64_t count, uint64_t seed, uint64_t subsequence,
                          augcl_stream stream);
/* Uniform bytes (synthetic frames / background bank; stands in for the absent Places365 / COCO
 * images of src/augmentations.py:19-74).  Byte e = byte (e % 16) of Philox block e / 16. */
int augcl_rng_fill_bytes(uint8_t* out, int64_t count, uint64_t seed, uint64_t subsequence,
                         augcl_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* AUGCL_B200_H_ */
