// Per-pixel arithmetic shared by the stand-alone augmentation kernels (augment.cu) and the fused sampler
// (fused.cu).  Every translation unit that includes this is compiled with --fmad=false: each multiply / add is
// rounded separately, like the reference's chain of torch elementwise ops.
#pragma once
#include "common.cuh"

namespace augcl {

// sigmoid(v) * 255 in five instructions: FMUL, MUFU.EX2, FADD, MUFU.RCP, FMUL.  The .ftz forms skip the
// denormal fix-up sequences (FSETP / FSEL / extra FMULs) that __expf and __fdividef carry; exp underflow flushes
// to 0 (sigmoid -> 255) and overflow gives rcp(inf) = 0, both exact limits.  Error << 1e-5 on the /255 scale (tests).
__device__ __forceinline__ float sigmoid255(float v) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(v * -1.4426950408889634f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return r * 255.0f;
}

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

// x / 255 correctly rounded (== __fdiv_rn(x, 255.0f) for every x whose quotient is a normal float) in three
// FP ops: Markstein's correction of q0 = x * fl(1/255) by the exact remainder.  tests/ checks it against IEEE
// division over all integers 0..255 and 2^24 random floats.
__device__ __forceinline__ float div255(float x) {
  const float rcp = 0.00392156862745098f;          // fl(1/255)
  const float q0 = x * rcp;
  const float rem = fmaf(-255.0f, q0, x);
  return fmaf(rem, rcp, q0);
}

__device__ __forceinline__ float overlay_px(float x, float img, float oma, float alpha) {
  // ((1 - alpha) * (x / 255.0) + alpha * imgs) * 255.0     (src/augmentations.py:262)
  return (oma * div255(x) + alpha * img) * 255.0f;
}

struct SpliceThr {
  float hue, sat, val;
  int variant;
};

// kornia.color.rgb_to_hsv (0.5.x, eps=1e-6) + the threshold test of create_hsv_mask, per pixel.
// Fast path for hue_thres == sat_thres == 0 (the reference's defaults, src/augmentations.py:295): the three
// divisions only decide signs there.  With delta = max - min >= 0 and d = delta (or 1):
//   s   = delta / (max + eps) > 0   <=>  delta > 0 and max + eps > 0
//   hue = 2pi * frac(hsel / d / 6) > 0  <=>  hsel != 0, and hsel is >= d when g or b is the first maximum,
//         so only the r-is-max branch (hsel = (max-b) - (max-g)) can be zero.
// (No quotient can underflow for inputs below 2^100.)  Other thresholds take the full formula.
__device__ __forceinline__ bool hsv_gt(float r, float g, float b, float ht, float st, float vt) {
  const float maxc = fmaxf(fmaxf(r, g), b);
  const float minc = fminf(fminf(r, g), b);
  const float delta = maxc - minc;
  const float rc = maxc - r, gc = maxc - g, bc = maxc - b;
  if (ht == 0.0f && st == 0.0f) {
    const bool hue_pos = !(r == maxc && (bc - gc) == 0.0f);
    return delta > 0.0f && (maxc + 1e-6f) > 0.0f && hue_pos && maxc > vt;
  }
  const float s = __fdiv_rn(delta, maxc + 1e-6f);
  const float d = (delta == 0.0f) ? 1.0f : delta;
  const float hsel = (r == maxc) ? (bc - gc) : ((g == maxc) ? ((2.0f * d + rc) - bc) : ((4.0f * d + gc) - rc));
  float hh = __fdiv_rn(__fdiv_rn(hsel, d), 6.0f);
  if (hh < 0.0f) hh = hh + 1.0f;                 // torch.remainder(., 1.0) for |.| < 1
  const float hue = 6.2831854820251465f * hh;    // float32(2*pi)
  return (hue > ht) && (s > st) && (maxc > vt);
}

// true = the pixel is kept (foreground), false = it is replaced by the background.  r, g, b are x/255.
__device__ __forceinline__ bool splice_keep(float r, float g, float b, const SpliceThr& t) {
  bool keep = hsv_gt(r, g, b, t.hue, t.sat, t.val);
  if (t.variant == AUGCL_CS_SPLICE) {
    // overlay[mask2] = x[mask2]; overlay[~mask] = x[~mask]   (src/augmentations.py:318-323)
    const bool mask2 = hsv_gt(r, g, b, 0.0f, t.sat, 0.6f);
    keep = mask2 || !keep;
  }
  return keep;
}

__device__ __forceinline__ void splice_px(float xr, float xg, float xb, float gr, float gg, float gb,
                                          const SpliceThr& t, float& o_r, float& o_g, float& o_b) {
  const float r = div255(xr), g = div255(xg), b = div255(xb);
  const bool keep = splice_keep(r, g, b, t);
  o_r = (keep ? r : gr) * 255.0f;
  o_g = (keep ? g : gg) * 255.0f;
  o_b = (keep ? b : gb) * 255.0f;
}

}  // namespace augcl
