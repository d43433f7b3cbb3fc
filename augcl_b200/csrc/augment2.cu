// The remaining aug_to_func entries of /root/reference/src/augmentations.py as sm_100a kernels:
//   random_grayscale :111-136 (+grayscale :88-108)   random_flip :139-168   random_rotation :171-213
//   random_cutout_color :327-350
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
  const int hw = h * w, chunks = (hw + 255) / 256, k = c / 3;
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
  cutout_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in, out, planes, c, h, w, box, colors);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

}  // extern "C"
