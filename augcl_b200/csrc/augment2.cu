// The remaining aug_to_func entries of /root/reference/src/augmentations.py as sm_100a kernels:
//   random_grayscale :111-136 (+grayscale :88-108)   random_flip :139-168   random_rotation :171-213
//   random_cutout_color :327-350   color_jitter :225-237 (torchvision ColorJitter restated)
// Same class as the hot five: read S32, write S32 per sample, pure permutation / elementwise work, HBM-bound.
// The per-sample decisions (apply mask, quarter turns, box, colours) are inputs, like every other draw.
// These functions came from RAD, where images are in [0,1]; the reference applies them to its 0..255 tensors
// unchanged, and so do these kernels (e.g. grayscale's x*255 -> uint8 wraps modulo 256 exactly like torch's CPU cast).
#include "common.cuh"
#include "pixel.cuh"

namespace augcl {

// One CTA per (sample, RGB frame) x chunk of pixels; thread = one pixel of the three channels.
__global__ void __launch_bounds__(256) grayscale_kernel(const float* __restrict__ in, float* __restrict__ out, int k, int hw,
                                                        int chunks, const uint8_t* __restrict__ mask) {
  const int chunk = blockIdx.x % chunks;
  const int frame = blockIdx.x / chunks;          // sample * k + f
  const int px = chunk * 256 + threadIdx.x;
  if (px >= hw) return;
  const int64_t o = (int64_t)frame * 3 * hw + px;
  // images * 255.0 -> .type(torch.uint8): truncate toward zero, keep the low 8 bits
  const float r = (float)((int)(__ldg(in + o) * 255.0f) & 255);
  const float g = (float)((int)(__ldg(in + o + hw) * 255.0f) & 255);
  const float b = (float)((int)(__ldg(in + o + 2 * (int64_t)hw) * 255.0f) & 255);
  float o_r = r, o_g = g, o_b = b;
  if (mask[frame / k]) {
    // r*0.2989 + g*0.587 + b*0.114 in fp32, left to right, then .type(torch.uint8).float()   (:96-101)
    const float y = (float)((int)((r * 0.2989f + g * 0.587f) + b * 0.114f) & 255);
    o_r = o_g = o_b = y;
  }
  out[o] = __fdiv_rn(o_r, 255.0f);
  out[o + hw] = __fdiv_rn(o_g, 255.0f);
  out[o + 2 * (int64_t)hw] = __fdiv_rn(o_b, 255.0f);
}

// mode 0: flip columns where flag != 0 (random_flip); mode 1: flag = quarter turns 0..3 (random_rotation, h == w).
// One CTA per plane; reads are gathered (L1/L2 absorb the transposed access), writes are coalesced.
__global__ void __launch_bounds__(256) permute_kernel(const float* __restrict__ in, float* __restrict__ out, int planes, int c,
                                                      int h, int w, const int32_t* __restrict__ flag, int mode) {
  const int hw = h * w;
  for (int plane = blockIdx.x; plane < planes; plane += gridDim.x) {
    const int f = flag[plane / c];
    const float* src = in + (int64_t)plane * hw;
    float* dst = out + (int64_t)plane * hw;
    for (int e = threadIdx.x; e < hw; e += blockDim.x) {
      const int i = e / w, j = e - i * w;
      int si = i, sj = j;
      if (mode == 0) {
        if (f) sj = w - 1 - j;
      } else {
        // torch.rot90(x, r, [2, 3]):  r=1: out[i][j] = x[j][w-1-i];  r=2: x[h-1-i][w-1-j];  r=3: x[h-1-j][i]
        if (f == 1) { si = j; sj = w - 1 - i; }
        else if (f == 2) { si = h - 1 - i; sj = w - 1 - j; }
        else if (f == 3) { si = h - 1 - j; sj = i; }
      }
      dst[e] = __ldg(src + si * w + sj);
    }
  }
}

// box[i] = (w1, h1): rows h1..2*h1-1, cols w1..2*w1-1 of every channel ch become colors[i][ch]
__global__ void __launch_bounds__(256) cutout_kernel(const float* __restrict__ in, float* __restrict__ out, int planes, int c,
                                                     int h, int w, const int32_t* __restrict__ box,
                                                     const float* __restrict__ colors) {
  const int hw = h * w;
  for (int plane = blockIdx.x; plane < planes; plane += gridDim.x) {
    const int i = plane / c;
    const int w1 = box[2 * i], h1 = box[2 * i + 1];
    const float col = colors[plane];
    const float* src = in + (int64_t)plane * hw;
    float* dst = out + (int64_t)plane * hw;
    for (int e = threadIdx.x; e < hw; e += blockDim.x) {
      const int r = e / w, q = e - r * w;
      const bool inside = r >= h1 && r < 2 * h1 && q >= w1 && q < 2 * w1;
      dst[e] = inside ? col : __ldg(src + e);
    }
  }
}

// ---- 16-byte versions (h*w % 4 == 0, w % 4 == 0 where rows matter, 16-byte aligned tensors): same arithmetic per pixel ----
__device__ __forceinline__ float gray_q(float v) { return (float)((int)(v * 255.0f) & 255); }   // x*255 -> uint8 (wraps) -> float

// thread = 4 pixels of the three channels of one frame: 3 LDG.128 + 3 STG.128
__global__ void __launch_bounds__(256) grayscale_vec_kernel(const float4* __restrict__ in, float4* __restrict__ out, int k, int hw4,
                                                            int chunks, const uint8_t* __restrict__ mask) {
  const int chunk = blockIdx.x % chunks;
  const int frame = blockIdx.x / chunks;          // sample * k + f
  const int p4 = chunk * 256 + threadIdx.x;
  if (p4 >= hw4) return;
  const int64_t o = (int64_t)frame * 3 * hw4 + p4;
  const float4 r4 = __ldg(in + o), g4 = __ldg(in + o + hw4), b4 = __ldg(in + o + 2 * (int64_t)hw4);
  float r[4] = {gray_q(r4.x), gray_q(r4.y), gray_q(r4.z), gray_q(r4.w)};
  float g[4] = {gray_q(g4.x), gray_q(g4.y), gray_q(g4.z), gray_q(g4.w)};
  float b[4] = {gray_q(b4.x), gray_q(b4.y), gray_q(b4.z), gray_q(b4.w)};
  if (mask[frame / k]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float y = (float)((int)((r[j] * 0.2989f + g[j] * 0.587f) + b[j] * 0.114f) & 255);
      r[j] = g[j] = b[j] = y;
    }
  }
  out[o] = make_float4(div255(r[0]), div255(r[1]), div255(r[2]), div255(r[3]));
  out[o + hw4] = make_float4(div255(g[0]), div255(g[1]), div255(g[2]), div255(g[3]));
  out[o + 2 * (int64_t)hw4] = make_float4(div255(b[0]), div255(b[1]), div255(b[2]), div255(b[3]));
}

// row-preserving permutations on whole 16-byte groups: identity, column flip (random_flip) and the half turn (rot90 by 2);
// planes drawn a quarter turn (rot90 by 1 / 3, a transposition) take the scalar gather of permute_kernel, plane by plane
__global__ void __launch_bounds__(256) permute_rows_vec_kernel(const float4* __restrict__ in, float4* __restrict__ out, int planes,
                                                               int c, int h, int wq, const int32_t* __restrict__ flag, int mode,
                                                               int use_tile) {
  extern __shared__ __align__(16) float ptile[];   // quarter turns: the plane, row stride w + 1 (conflict-free column reads)
  const int n4 = h * wq;
  for (int plane = blockIdx.x; plane < planes; plane += gridDim.x) {
    const int f = flag[plane / c];
    const bool fc = mode == 0 ? f != 0 : f == 2, fr = mode == 1 && f == 2;
    const float4* src = in + (int64_t)plane * n4;
    float4* dst = out + (int64_t)plane * n4;
    if (mode == 1 && (f & 1)) {   // h == w here
      const int w = wq * 4;
      const float* s1 = reinterpret_cast<const float*>(src);
      float* d1 = reinterpret_cast<float*>(dst);
      if (use_tile) {
        // transpose through shared memory: coalesced 16-byte reads in, coalesced 16-byte writes out, the transposed access
        // (a column of the tile per output row) happens on the padded tile
        const int ts = w + 1;
        __syncthreads();                       // the previous plane's reads of the tile are done
        for (int e = threadIdx.x; e < n4; e += 256) {
          const int i = e / wq, q = e - i * wq;
          const float4 v = __ldg(src + e);
          float* t = ptile + i * ts + 4 * q;
          t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w;
        }
        __syncthreads();
        for (int e = threadIdx.x; e < n4; e += 256) {
          const int i = e / wq, j = (e - i * wq) * 4;
          // torch.rot90(x, r, [2, 3]):  r=1: out[i][j] = x[j][w-1-i];  r=3: x[h-1-j][i]
          float4 o;
          if (f == 1) {
            const float* t = ptile + j * ts + (w - 1 - i);
            o = make_float4(t[0], t[ts], t[2 * ts], t[3 * ts]);
          } else {
            const float* t = ptile + (h - 1 - j) * ts + i;
            o = make_float4(t[0], t[-ts], t[-2 * ts], t[-3 * ts]);
          }
          dst[e] = o;
        }
        continue;
      }
      for (int e = threadIdx.x; e < h * w; e += 256) {
        const int i = e / w, j = e - i * w;
        // torch.rot90(x, r, [2, 3]):  r=1: out[i][j] = x[j][w-1-i];  r=3: x[h-1-j][i]
        d1[e] = f == 1 ? __ldg(s1 + j * w + (w - 1 - i)) : __ldg(s1 + (h - 1 - j) * w + i);
      }
      continue;
    }
    for (int e0 = threadIdx.x; e0 < n4; e0 += 4 * 256) {   // four independent loads in flight per thread before the first store
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e = e0 + u * 256;
        if (e < n4) {
          const int i = e / wq, q = e - i * wq;
          v[u] = __ldg(src + (fr ? h - 1 - i : i) * wq + (fc ? wq - 1 - q : q));
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e = e0 + u * 256;
        if (e < n4) dst[e] = fc ? make_float4(v[u].w, v[u].z, v[u].y, v[u].x) : v[u];
      }
    }
  }
}

__global__ void __launch_bounds__(256) cutout_vec_kernel(const float4* __restrict__ in, float4* __restrict__ out, int planes, int c,
                                                         int h, int wq, const int32_t* __restrict__ box,
                                                         const float* __restrict__ colors) {
  const int n4 = h * wq;
  for (int plane = blockIdx.x; plane < planes; plane += gridDim.x) {
    const int i = plane / c;
    const int w1 = box[2 * i], h1 = box[2 * i + 1];
    const float col = colors[plane];
    const float4* src = in + (int64_t)plane * n4;
    float4* dst = out + (int64_t)plane * n4;
    for (int e = threadIdx.x; e < n4; e += 256) {
      const int r = e / wq, q0 = (e - r * wq) * 4;
      float4 v = __ldg(src + e);
      if (r >= h1 && r < 2 * h1) {
        if (q0 >= w1 && q0 < 2 * w1) v.x = col;
        if (q0 + 1 >= w1 && q0 + 1 < 2 * w1) v.y = col;
        if (q0 + 2 >= w1 && q0 + 2 < 2 * w1) v.z = col;
        if (q0 + 3 >= w1 && q0 + 3 < 2 * w1) v.w = col;
      }
      dst[e] = v;
    }
  }
}

static int check_args(const char* who, const float* in, float* out, int n, int c, int h, int w, const void* draw) {
  AUGCL_REQUIRE(in && out && draw, "%s: NULL pointer", who);
  AUGCL_REQUIRE(in != out, "%s: in and out must not alias", who);
  AUGCL_REQUIRE(n > 0 && c > 0 && h > 0 && w > 0, "%s: n, c, h, w must be positive", who);
  AUGCL_REQUIRE((int64_t)n * c < (1ll << 31) && (int64_t)h * w < (1ll << 28), "%s: shape too large", who);
  return AUGCL_OK;
}

// create_hsv_mask (/root/reference/src/augmentations.py:277-284): kornia HSV of an RGB image already in [0,1], all three
// channels strictly above their thresholds.  One thread per pixel; mask[n][h*w] = 0 / 1.
__global__ void __launch_bounds__(256) hsv_mask_kernel(const float* __restrict__ rgb, uint8_t* __restrict__ mask, int64_t npix,
                                                       int hw, float ht, float st, float vt) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < npix; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = e / hw;
    const int px = (int)(e - i * hw);
    const float* b = rgb + i * 3 * hw + px;
    mask[e] = hsv_gt(__ldg(b), __ldg(b + hw), __ldg(b + 2 * (int64_t)hw), ht, st, vt) ? 1 : 0;
  }
}


// color_jitter (src/augmentations.py:225-237): torchvision ColorJitter(0, (0,2), (0,3), (-0.5,0.5)) on the selected RGB
// frames, in place.  One CTA per selected frame: the frame (3*h*w floats, 85 KB at 84x84) sits in shared memory while the
// up-to-four adjustments run in the drawn order (torchvision/transforms/_functional_tensor.py: _blend, adjust_contrast,
// adjust_saturation, adjust_hue with _rgb2hsv/_hsv2rgb, rgb_to_grayscale), each restated op for op; contrast needs the
// frame's mean luminance, a block reduction (the only place the summation order differs from torch's).
struct JitterParams {
  int order[4];        // torchvision's fn_idx: 0 brightness, 1 contrast, 2 saturation, 3 hue
  float ratio[3];      // (float)factor for ops 0..2
  float omr[3];        // (float)(1.0 - factor), formed in double like Python does
  float hue;
  int enabled;         // bit i = op i has a factor
};

__device__ __forceinline__ float lum(float r, float g, float b) { return (0.2989f * r + 0.587f * g) + 0.114f * b; }
__device__ __forceinline__ float clamp01(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }

__device__ __forceinline__ void hue_px(float& r, float& g, float& b, float hf) {
  const float maxc = fmaxf(fmaxf(r, g), b), minc = fminf(fminf(r, g), b);
  const bool eqc = maxc == minc;
  const float cr = maxc - minc;
  const float s = __fdiv_rn(cr, eqc ? 1.0f : maxc);
  const float d = eqc ? 1.0f : cr;
  const float rc = __fdiv_rn(maxc - r, d), gc = __fdiv_rn(maxc - g, d), bc = __fdiv_rn(maxc - b, d);
  const float hr = (maxc == r) ? (bc - gc) : 0.0f;
  const float hg = (maxc == g && maxc != r) ? ((2.0f + rc) - bc) : 0.0f;
  const float hb = (maxc != g && maxc != r) ? ((4.0f + gc) - rc) : 0.0f;
  float h = (hr + hg) + hb;
  h = fmodf(__fdiv_rn(h, 6.0f) + 1.0f, 1.0f);
  // (h + hue_factor) % 1.0: torch.remainder = fmod, moved into [0, 1) when negative
  float m = fmodf(h + hf, 1.0f);
  if (m != 0.0f && m < 0.0f) m += 1.0f;
  const float h6 = m * 6.0f;
  const float fl = floorf(h6);
  const float f = h6 - fl;
  const int i = ((int)fl) % 6;
  const float v = maxc;
  const float pp = clamp01(v * (1.0f - s));
  const float q = clamp01(v * (1.0f - s * f));
  const float t = clamp01(v * (1.0f - s * (1.0f - f)));
  r = i == 0 ? v : i == 1 ? q : i == 2 ? pp : i == 3 ? pp : i == 4 ? t : v;
  g = i == 0 ? t : i == 1 ? v : i == 2 ? v : i == 3 ? q : i == 4 ? pp : pp;
  b = i == 0 ? pp : i == 1 ? pp : i == 2 ? t : i == 3 ? v : i == 4 ? v : q;
}

__global__ void __launch_bounds__(512, 2) color_jitter_kernel(float* __restrict__ imgs, const int32_t* __restrict__ sel, int hw,
                                                           JitterParams P) {
  extern __shared__ __align__(16) float jsm[];     // [3][hw]
  __shared__ float s_red[16];
  const int NT = blockDim.x;     // 512: two CTAs (85 KB frame each at 84x84) = 32 warps per SM behind the divisions of the hue step
  float* fr = imgs + (int64_t)sel[blockIdx.x] * 3 * hw;
  const int tid = threadIdx.x;
  for (int e = tid; e < 3 * hw; e += NT) jsm[e] = fr[e];
  __syncthreads();
  for (int step = 0; step < 4; ++step) {
    const int op = P.order[step];
    if (!((P.enabled >> op) & 1)) continue;
    if (op == 1) {
      float part = 0.0f;
      for (int px = tid; px < hw; px += NT) part += lum(jsm[px], jsm[hw + px], jsm[2 * hw + px]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
      if ((tid & 31) == 0) s_red[tid >> 5] = part;
      __syncthreads();
      float tot = 0.0f;
#pragma unroll
      for (int j = 0; j < 16; ++j) tot += (j < (NT >> 5)) ? s_red[j] : 0.0f;
      const float other = P.omr[1] * __fdiv_rn(tot, (float)hw);
      for (int e = tid; e < 3 * hw; e += NT) jsm[e] = clamp01(P.ratio[1] * jsm[e] + other);
    } else if (op == 0) {
      for (int e = tid; e < 3 * hw; e += NT) jsm[e] = clamp01(P.ratio[0] * jsm[e] + P.omr[0] * 0.0f);
    } else {
      for (int px = tid; px < hw; px += NT) {
        float r = jsm[px], g = jsm[hw + px], b = jsm[2 * hw + px];
        if (op == 2) {
          const float y = P.omr[2] * lum(r, g, b);
          r = clamp01(P.ratio[2] * r + y); g = clamp01(P.ratio[2] * g + y); b = clamp01(P.ratio[2] * b + y);
        } else {
          hue_px(r, g, b, P.hue);
        }
        jsm[px] = r; jsm[hw + px] = g; jsm[2 * hw + px] = b;
      }
    }
    __syncthreads();
  }
  for (int e = tid; e < 3 * hw; e += NT) fr[e] = jsm[e];
}
}  // namespace augcl

using namespace augcl;

extern "C" {

int augcl_hsv_mask(const float* rgb, uint8_t* mask, int n, int h, int w, float hue_thres, float sat_thres, float val_thres,
                   augcl_stream stream) {
  AUGCL_REQUIRE(rgb && mask, "augcl_hsv_mask: rgb/mask is NULL");
  AUGCL_REQUIRE(n > 0 && h > 0 && w > 0, "augcl_hsv_mask: n, h, w must be positive");
  const int64_t npix = (int64_t)n * h * w;
  int64_t blocks = (npix + 255) / 256;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  hsv_mask_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(rgb, mask, npix, h * w, hue_thres, sat_thres, val_thres);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

int augcl_random_grayscale(const float* in, float* out, int n, int c, int h, int w, const uint8_t* mask,
                           augcl_stream stream) {
  if (int rc = check_args("augcl_random_grayscale", in, out, n, c, h, w, mask)) return rc;
  AUGCL_REQUIRE(c % 3 == 0, "augcl_random_grayscale: c=%d is not a multiple of 3", c);
  const int hw = h * w, k = c / 3;
  if ((hw & 3) == 0 && aligned(in, 16) && aligned(out, 16)) {
    const int hw4 = hw / 4, chunks4 = (hw4 + 255) / 256;
    const int64_t grid4 = (int64_t)n * k * chunks4;
    AUGCL_REQUIRE(grid4 < (1ll << 31), "augcl_random_grayscale: batch too large");
    grayscale_vec_kernel<<<(unsigned)grid4, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(in), reinterpret_cast<float4*>(out),
                                                                            k, hw4, chunks4, mask);
    AUGCL_LAUNCHED();
    return AUGCL_OK;
  }
  const int chunks = (hw + 255) / 256;
  const int64_t grid = (int64_t)n * k * chunks;
  AUGCL_REQUIRE(grid < (1ll << 31), "augcl_random_grayscale: batch too large");
  grayscale_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(in, out, k, hw, chunks, mask);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

static int launch_permute(const char* who, const float* in, float* out, int n, int c, int h, int w, const int32_t* flag,
                          int mode, augcl_stream stream) {
  if (int rc = check_args(who, in, out, n, c, h, w, flag)) return rc;
  const int planes = n * c;
  const int grid = planes < kNumSMs * 8 ? planes : kNumSMs * 8;
  if ((w & 3) == 0 && aligned(in, 16) && aligned(out, 16)) {
    const size_t tile = (mode == 1 && (size_t)h * (w + 1) * 4 <= 48 * 1024) ? (size_t)h * (w + 1) * 4 : 0;   // quarter turns only
    permute_rows_vec_kernel<<<grid, 256, tile, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(in), reinterpret_cast<float4*>(out),
                                                                     planes, c, h, w / 4, flag, mode, tile ? 1 : 0);
    AUGCL_LAUNCHED();
    return AUGCL_OK;
  }
  permute_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in, out, planes, c, h, w, flag, mode);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

int augcl_random_flip(const float* in, float* out, int n, int c, int h, int w, const int32_t* flip,
                      augcl_stream stream) {
  return launch_permute("augcl_random_flip", in, out, n, c, h, w, flip, 0, stream);
}

int augcl_random_rotation(const float* in, float* out, int n, int c, int h, int w, const int32_t* quarter_turns,
                          augcl_stream stream) {
  if (h != w) AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "augcl_random_rotation: needs square images (got %dx%d)", h, w);
  return launch_permute("augcl_random_rotation", in, out, n, c, h, w, quarter_turns, 1, stream);
}

int augcl_random_cutout_color(const float* in, float* out, int n, int c, int h, int w, const int32_t* box,
                              const float* colors, augcl_stream stream) {
  if (int rc = check_args("augcl_random_cutout_color", in, out, n, c, h, w, box)) return rc;
  AUGCL_REQUIRE(colors, "augcl_random_cutout_color: colors is NULL");
  const int planes = n * c;
  const int grid = planes < kNumSMs * 8 ? planes : kNumSMs * 8;
  if ((w & 3) == 0 && aligned(in, 16) && aligned(out, 16)) {
    cutout_vec_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(in), reinterpret_cast<float4*>(out), planes, c,
                                                            h, w / 4, box, colors);
    AUGCL_LAUNCHED();
    return AUGCL_OK;
  }
  cutout_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in, out, planes, c, h, w, box, colors);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

int augcl_color_jitter(float* imgs, int frames, int h, int w, const int32_t* sel, int n_sel, const int32_t* order,
                       double brightness, double contrast, double saturation, double hue, int enabled,
                       augcl_stream stream) {
  AUGCL_REQUIRE(imgs && order, "augcl_color_jitter: imgs/order is NULL");
  AUGCL_REQUIRE(frames > 0 && h > 0 && w > 0 && n_sel >= 0 && n_sel <= frames, "augcl_color_jitter: bad sizes");
  AUGCL_REQUIRE((enabled & ~15) == 0, "augcl_color_jitter: enabled has unknown bits");
  JitterParams P;
  int seen = 0;
  for (int i = 0; i < 4; ++i) {
    AUGCL_REQUIRE(order[i] >= 0 && order[i] < 4, "augcl_color_jitter: order[%d]=%d", i, order[i]);
    seen |= 1 << order[i];
    P.order[i] = order[i];
  }
  AUGCL_REQUIRE(seen == 15, "augcl_color_jitter: order is not a permutation of 0..3");
  const double fac[3] = {brightness, contrast, saturation};
  for (int i = 0; i < 3; ++i) {
    if ((enabled >> i) & 1) AUGCL_REQUIRE(fac[i] >= 0.0, "augcl_color_jitter: factor %d is negative", i);
    P.ratio[i] = (float)fac[i];
    P.omr[i] = (float)(1.0 - fac[i]);
  }
  if (enabled & 8) AUGCL_REQUIRE(hue >= -0.5 && hue <= 0.5, "augcl_color_jitter: hue factor %f outside [-0.5, 0.5]", hue);
  P.hue = (float)hue;
  P.enabled = enabled;
  if (n_sel == 0 || enabled == 0) return AUGCL_OK;
  AUGCL_REQUIRE(sel, "augcl_color_jitter: sel is NULL");
  const size_t smem = (size_t)3 * h * w * sizeof(float);
  if (smem > 200 * 1024) AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "augcl_color_jitter: %dx%d frame does not fit shared memory", h, w);
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(color_jitter_kernel), 200 * 1024)) return rc;
  color_jitter_kernel<<<n_sel, 512, smem, (cudaStream_t)stream>>>(imgs, sel, h * w, P);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

}  // extern "C"
