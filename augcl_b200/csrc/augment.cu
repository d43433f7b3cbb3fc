// Stand-alone augmentation kernels on f32 NCHW batches (sm_100a).
//
// Reference functions replaced (all under /root/reference/src/augmentations.py):
//   random_shift :378-382   random_crop :385-433   random_conv :265-274
//   random_overlay :253-262   splice :295-306 + create_hsv_mask :277-284   CS_splice :309-324
//
// All of them are HBM-bound streaming kernels (read S32 [+ bank], write S32 per sample); none is a
// dense contraction, so no tensor cores.  This translation unit is compiled with --fmad=false: every
// multiply/add below is rounded separately, exactly like the reference's chain of torch elementwise
// ops, and FMAs appear only where written explicitly (the conv accumulation).
#include <climits>
#include "common.cuh"
#include "pixel.cuh"

namespace augcl {
extern std::atomic<int> g_sample_pdl;   // state.cu: programmatic dependent launch on / off
std::atomic<int> g_conv_part{1};  // random_conv: part-of-frame items (0 = 12-row bands, double-buffered)
std::atomic<int> g_bil_tma{1};   // bilinear shift: double-buffered cp.async.bulk tiles (0 = single tile, LDG fill)

// ---------------------------------------------------------------------------------------------
// window copy: out[i,ch,r,q] = in[i,ch,clamp(r+dy_i,0,H-1),clamp(q+dx_i,0,W-1)]
//   random_shift INT: (dy,dx) = offsets[i] - pad      random_crop: (dy,dx) = offsets[i]
// One CTA walks (sample, channel) planes; a thread emits one float4 from four source floats (the
// source window is mis-aligned by dx floats, so the loads are scalar and coalesce in L1).
// ---------------------------------------------------------------------------------------------
template <int WQ>
__global__ void __launch_bounds__(256) window_copy_vec_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                              int planes, int c, int H, int W, int Ho, int Wo,
                                                              const int32_t* __restrict__ offsets, int pad) {
  constexpr int UNR = 4;   // 16 independent loads in flight per thread before the first store
  const int wq = WQ ? WQ : (Wo >> 2);
  const int nvec = Ho * wq;
  // programmatic launch: this grid's CTAs become resident while the previous kernel of the stream drains (its output may be
  // our input: nothing is read before the wait), and the next kernel's CTAs while this one does
  pdl_wait();
  pdl_launch_dependents();
  for (int plane = blockIdx.x; plane < planes; plane += gridDim.x) {
    const int i = plane / c;
    const int dy = offsets[2 * i] - pad, dx = offsets[2 * i + 1] - pad;
    const float* src = in + (int64_t)plane * H * W;
    float4* dst = reinterpret_cast<float4*>(out + (int64_t)plane * Ho * Wo);
    for (int v0 = threadIdx.x; v0 < nvec; v0 += UNR * 256) {
      float4 o[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int v = v0 + u * 256;
        if (v < nvec) {
          const int r = v / wq, q4 = v - r * wq;
          const float* row = src + (int64_t)clampi(r + dy, 0, H - 1) * W;
          const int c0 = q4 * 4 + dx;
          o[u] = make_float4(__ldg(row + clampi(c0, 0, W - 1)), __ldg(row + clampi(c0 + 1, 0, W - 1)),
                             __ldg(row + clampi(c0 + 2, 0, W - 1)), __ldg(row + clampi(c0 + 3, 0, W - 1)));
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int v = v0 + u * 256;
        if (v < nvec) dst[v] = o[u];
      }
    }
  }
}

__global__ void __launch_bounds__(256) window_copy_generic_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                                  int64_t total, int c, int H, int W, int Ho, int Wo,
                                                                  const int32_t* __restrict__ offsets, int pad) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int q = (int)(e % Wo);
    int64_t rest = e / Wo;
    const int r = (int)(rest % Ho);
    const int64_t plane = rest / Ho;
    const int i = (int)(plane / c);
    const int rr = clampi(r + offsets[2 * i] - pad, 0, H - 1), cc = clampi(q + offsets[2 * i + 1] - pad, 0, W - 1);
    out[e] = in[(plane * H + rr) * W + cc];
  }
}

// ---------------------------------------------------------------------------------------------
// bilinear shift, DrQ-v2 formulation: F.grid_sample(replicate_pad(x), base_grid + shift,
// padding_mode='zeros', align_corners=False).  The fp32 sequence below restates ATen's CPU
// grid_sampler_2d (the oracle of record): unnormalise ((g+1)*size-1)/2 -> floor -> four corner
// weights as products of distances -> weighted sum in nw,ne,sw,se order; corners outside the padded
// image contribute zero.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) shift_bilinear_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                             int64_t total, int c, int h, int w, int pad,
                                                             const int32_t* __restrict__ offsets,
                                                             const float* __restrict__ base) {
  const int Hp = h + 2 * pad, Wp = w + 2 * pad;
  const float sc = (float)(2.0 / (double)Hp);   // python: shift *= 2.0 / (h + 2*pad)  (h == w)
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int q = (int)(e % w);
    int64_t rest = e / w;
    const int r = (int)(rest % h);
    const int64_t plane = rest / h;
    const int i = (int)(plane / c);
    const float gx = base[q] + (float)offsets[2 * i + 1] * sc;
    const float gy = base[r] + (float)offsets[2 * i] * sc;
    const float ix = ((gx + 1.0f) * (float)Wp - 1.0f) / 2.0f;
    const float iy = ((gy + 1.0f) * (float)Hp - 1.0f) / 2.0f;
    const float fx = floorf(ix), fy = floorf(iy);
    const int x0 = (int)fx, y0 = (int)fy, x1 = x0 + 1, y1 = y0 + 1;
    const float tx = ix - fx, ty = iy - fy;      // distance to the west / north corner
    const float ex = 1.0f - tx, ey = 1.0f - ty;
    const float* src = in + plane * h * w;
    auto tap = [&](int y, int x) -> float {
      if (x < 0 || y < 0 || x >= Wp || y >= Hp) return 0.0f;
      return __ldg(src + (int64_t)clampi(y - pad, 0, h - 1) * w + clampi(x - pad, 0, w - 1));
    };
    const float nw = tap(y0, x0) * (ex * ey), ne = tap(y0, x1) * (tx * ey);
    const float sw = tap(y1, x0) * (ex * ty), se = tap(y1, x1) * (tx * ty);
    out[e] = ((nw + ne) + sw) + se;
  }
}

// Same arithmetic, restructured for throughput (w % 4 == 0, 16-B aligned out): one CTA walks (sample, channel)
// planes, a thread produces 4 consecutive pixels of a row -- the row-side terms (iy, floor, weights, the two
// clamped source rows) are computed once per thread-row instead of once per pixel, all index math is 32-bit, and
// the result leaves as one STG.128.  The 16 taps of a thread are L1 hits (neighbouring threads share them).
__global__ void __launch_bounds__(256) shift_bilinear_vec_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                                 int planes, int c, int h, int w, int pad,
                                                                 const int32_t* __restrict__ offsets,
                                                                 const float* __restrict__ base) {
  const int Hp = h + 2 * pad, Wp = w + 2 * pad;
  const float sc = (float)(2.0 / (double)Hp);
  const int nq = w >> 2;
  // fixed mapping: thread = (row slot, 4-pixel column group); the column-side terms (ix, floor, weights, the two clamped
  // source columns and their validity, x 4 columns) depend only on the sample -> computed once per plane, not per pixel
  const int rpr = 256 / nq;                      // rows per round (12 at w = 84); requires nq <= 256
  const int r0 = threadIdx.x / nq, q4 = threadIdx.x - r0 * nq;
  const bool active = r0 < rpr;
  float bx[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) bx[j] = active ? __ldg(base + 4 * q4 + j) : 0.0f;
  for (int plane = blockIdx.x; plane < planes; plane += gridDim.x) {
    const int i = plane / c;
    const float sy = (float)offsets[2 * i] * sc, sx = (float)offsets[2 * i + 1] * sc;
    const float* src = in + (int64_t)plane * h * w;
    float4* dst = reinterpret_cast<float4*>(out + (int64_t)plane * h * w);
    float tx[4], ex[4];
    int c0[4], c1[4];
    bool vx0[4], vx1[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gx = bx[j] + sx;
      const float ix = ((gx + 1.0f) * (float)Wp - 1.0f) / 2.0f;
      const float fx = floorf(ix);
      const int x0 = (int)fx, x1 = x0 + 1;
      tx[j] = ix - fx; ex[j] = 1.0f - tx[j];
      vx0[j] = x0 >= 0 && x0 < Wp; vx1[j] = x1 >= 0 && x1 < Wp;
      c0[j] = clampi(x0 - pad, 0, w - 1); c1[j] = clampi(x1 - pad, 0, w - 1);
    }
    for (int r = active ? r0 : h; r < h; r += rpr) {
      const float gy = __ldg(base + r) + sy;
      const float iy = ((gy + 1.0f) * (float)Hp - 1.0f) / 2.0f;
      const float fy = floorf(iy);
      const int y0 = (int)fy, y1 = y0 + 1;
      const float ty = iy - fy, ey = 1.0f - ty;
      const bool vy0 = y0 >= 0 && y0 < Hp, vy1 = y1 >= 0 && y1 < Hp;
      const float* row0 = src + clampi(y0 - pad, 0, h - 1) * w;
      const float* row1 = src + clampi(y1 - pad, 0, h - 1) * w;
      float o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float t00 = (vy0 && vx0[j]) ? __ldg(row0 + c0[j]) : 0.0f, t01 = (vy0 && vx1[j]) ? __ldg(row0 + c1[j]) : 0.0f;
        const float t10 = (vy1 && vx0[j]) ? __ldg(row1 + c0[j]) : 0.0f, t11 = (vy1 && vx1[j]) ? __ldg(row1 + c1[j]) : 0.0f;
        const float nw = t00 * (ex[j] * ey), ne = t01 * (tx[j] * ey), sw = t10 * (ex[j] * ty), se = t11 * (tx[j] * ty);
        o[j] = ((nw + ne) + sw) + se;
      }
      dst[r * nq + q4] = make_float4(o[0], o[1], o[2], o[3]);
    }
  }
}

// Staged variant for planes that fit shared memory (h, w <= 256 and h*w*4 + h*16 <= 96 KB: 84x84 takes 29 KB, so 7 CTAs
// per SM overlap each other's fill and compute phases).  The 4-pixel-per-thread kernel above is issue bound (ncu: 69
// instructions per pixel, a third of them 64-bit address arithmetic for the 4 taps); here the plane is copied to shared
// memory with coalesced float4 loads, thread = (row chunk, column) so the 4 taps are conflict-free LDS at 32-bit
// addresses, the column terms are computed once per plane per thread and the row terms once per plane per row.
constexpr int kBilMaxH = 256;
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
constexpr int kBilSmemMax = 96 * 1024;
__global__ void __launch_bounds__(256) shift_bilinear_smem_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                                  int planes, int c, int h, int w, int pad,
                                                                  const int32_t* __restrict__ offsets,
                                                                  const float* __restrict__ base, int vec) {
  extern __shared__ __align__(16) unsigned char bsm[];
  // per output row: {byte offset of the upper source row in the tile, of the lower one, ty, flags}; flags bit 0 / 1 = the
  // upper / lower row lies outside the padded image (those taps read zero, as grid_sample's zeros padding does).  The
  // table is padded to a multiple of 4 rows (copies of the last row) so a 4-row trip never indexes past it.
  // vec == 2: two tiles, the next plane's 4*h*w bytes arrive by ONE cp.async.bulk while this plane is computed.
  const int h4 = (h + 3) & ~3;
  __shared__ uint64_t bars[2];
  int4* s_row = reinterpret_cast<int4*>(bsm);
  float* tile0 = reinterpret_cast<float*>(bsm + (size_t)h4 * 16);
  const bool tma = vec == 2;
  if (tma && threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_fence_init();
  }
  pdl_wait();                 // launched programmatically: nothing is read before the previous kernel has completed
  pdl_launch_dependents();
  const int Hp = h + 2 * pad, Wp = w + 2 * pad;
  const float sc = (float)(2.0 / (double)Hp);
  const int nchunk = 256 / w, rows_per = (((h + nchunk - 1) / nchunk) + 3) & ~3;
  const int rc = threadIdx.x / w, x = threadIdx.x - rc * w;
  const bool active = rc < nchunk && rc * rows_per < h;
  const float bx = active ? __ldg(base + x) : 0.0f;
  const int n_el = h * w;
  if (tma) {
    __syncthreads();
    if (threadIdx.x == 0 && (int)blockIdx.x < planes) {
      mbar_expect_tx(&bars[0], (uint32_t)n_el * 4);
      bulk_g2s(tile0, in + (int64_t)blockIdx.x * n_el, (uint32_t)n_el * 4, &bars[0]);
    }
  }
  int it = 0;
  for (int plane = blockIdx.x; plane < planes; plane += gridDim.x, ++it) {
    const int i = plane / c;
    const float sy = (float)offsets[2 * i] * sc, sx = (float)offsets[2 * i + 1] * sc;
    const float* src = in + (int64_t)plane * n_el;
    float* tile = tile0 + (tma ? (it & 1) * n_el : 0);
    const uint32_t tile_s = smem_u32(tile);
    __syncthreads();                               // the previous plane's taps are done
    if (tma) {
      const int nxt = plane + gridDim.x;
      if (threadIdx.x == 0 && nxt < planes) {
        mbar_expect_tx(&bars[(it + 1) & 1], (uint32_t)n_el * 4);
        bulk_g2s(tile0 + ((it + 1) & 1) * n_el, in + (int64_t)nxt * n_el, (uint32_t)n_el * 4, &bars[(it + 1) & 1]);
      }
    } else if (vec) {
      const float4* s4 = reinterpret_cast<const float4*>(src);
      float4* t4 = reinterpret_cast<float4*>(tile);
      for (int e = threadIdx.x; e < (n_el >> 2); e += 256) t4[e] = __ldg(s4 + e);
    } else {
      for (int e = threadIdx.x; e < n_el; e += 256) tile[e] = __ldg(src + e);
    }
    for (int r = threadIdx.x; r < h4; r += 256) {
      const float gy = __ldg(base + min(r, h - 1)) + sy;
      const float iy = ((gy + 1.0f) * (float)Hp - 1.0f) / 2.0f;
      const float fy = floorf(iy);
      const int y0 = (int)fy, y1 = y0 + 1;
      const int fl = ((y0 >= 0 && y0 < Hp) ? 0 : 1) | ((y1 >= 0 && y1 < Hp) ? 0 : 2);
      s_row[r] = make_int4(clampi(y0 - pad, 0, h - 1) * w * 4, clampi(y1 - pad, 0, h - 1) * w * 4,
                           __float_as_int(iy - fy), fl);
    }
    __syncthreads();
    if (!active) continue;
    if (tma) mbar_wait(&bars[it & 1], (uint32_t)(it >> 1) & 1u);
    const float gx = bx + sx;
    const float ix = ((gx + 1.0f) * (float)Wp - 1.0f) / 2.0f;
    const float fx = floorf(ix);
    const int x0 = (int)fx, x1 = x0 + 1;
    const float tx = ix - fx, ex = 1.0f - tx;
    const bool vx0 = x0 >= 0 && x0 < Wp, vx1 = x1 >= 0 && x1 < Wp;
    const int xfl = (vx0 && vx1) ? 0 : 4;
    const uint32_t a0 = tile_s + clampi(x0 - pad, 0, w - 1) * 4, a1 = tile_s + clampi(x1 - pad, 0, w - 1) * 4;
    const int r_beg = rc * rows_per, r_end = min(h, r_beg + rows_per);
    float* o = out + (int64_t)plane * n_el + (int64_t)r_beg * w + x;
    for (int rb = r_beg; rb < r_end; rb += 4, o += 4 * w) {
      float t[4][4];
      int4 q[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        q[u] = s_row[rb + u];
        t[u][0] = lds_f32(a0 + q[u].x); t[u][1] = lds_f32(a1 + q[u].x);
        t[u][2] = lds_f32(a0 + q[u].y); t[u][3] = lds_f32(a1 + q[u].y);
      }
      // a tap outside the padded image (only the first / last row or column, when a rounding puts floor() one off)
      if ((q[0].w | q[1].w | q[2].w | q[3].w | xfl) != 0) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if ((q[u].w & 1) || !vx0) t[u][0] = 0.0f;
          if ((q[u].w & 1) || !vx1) t[u][1] = 0.0f;
          if ((q[u].w & 2) || !vx0) t[u][2] = 0.0f;
          if ((q[u].w & 2) || !vx1) t[u][3] = 0.0f;
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float ty = __int_as_float(q[u].z), ey = 1.0f - ty;
        const float nw = t[u][0] * (ex * ey), ne = t[u][1] * (tx * ey), sw = t[u][2] * (ex * ty), se = t[u][3] * (tx * ty);
        if (rb + u < r_end) o[u * w] = ((nw + ne) + sw) + se;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// random_conv: CTA (128 threads) = one RGB frame x one band of kBand rows.  The band (+1-row/col replicate
// halo), already divided by 255, is staged in shared memory ([ci][row][1 + col], so the 6-wide input window
// of a 4-pixel strip is one aligned LDS.128 + one LDS.64); the 81 weights of the sample sit in shared memory
// as [27][4] so one broadcast LDS.128 feeds the three output channels of a tap.  A thread produces a 2x4
// block of pixels for all three output channels: 648 explicit FMAs from 24 input and 27 weight loads, then
// sigmoid and *255.  Per output element: 27 FMA + 2 MUFU -- the kernel sits at the FP32/HBM ridge.
// ---------------------------------------------------------------------------------------------
constexpr int kBand = 12;
constexpr int kConvThreads = 128;

__global__ void __launch_bounds__(kConvThreads, 6) random_conv_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                                      int frames, int k, int h, int w,
                                                                      const float* __restrict__ weights, int bands, int RS,
                                                                      int vec) {
  extern __shared__ __align__(16) float csm[];
  float* wsm = csm;              // [27][4]
  float* tile = csm + 27 * 4;    // [3][kBand+2][RS]
  const int tid = threadIdx.x;
  const int nq = (w + 3) >> 2;
  for (int work = blockIdx.x; work < frames * bands; work += gridDim.x) {
    const int frame = work / bands, band = work - frame * bands;
    const int sample = frame / k;
    const int r0 = band * kBand;
    const int rows = min(kBand, h - r0);
    const float* src = in + (int64_t)frame * 3 * h * w;
    __syncthreads();  // previous work item fully consumed
    if (tid < 81) {
      // weights[sample][co][ci][ky][kx] -> wsm[(ci*9 + ky*3 + kx)*4 + co]
      const int co = tid / 27, tap = tid - co * 27;
      wsm[tap * 4 + co] = weights[(int64_t)sample * 81 + tid];
    }
    const int trows = rows + 2;
    for (int e = tid; e < 3 * trows * nq; e += kConvThreads) {
      const int q = e % nq;
      const int t2 = e / nq;
      const int pr = t2 % trows, ci = t2 / trows;
      const int rr = clampi(r0 + pr - 1, 0, h - 1);
      const float* row = src + ((int64_t)ci * h + rr) * w;
      float* trow = tile + (ci * (kBand + 2) + pr) * RS;
      float v[4];
      if (vec) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(row) + q);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = __ldg(row + min(4 * q + j, w - 1));
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = div255(v[j]);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (vec || 4 * q + j < w) trow[1 + 4 * q + j] = v[j];
      if (q == 0) trow[0] = v[0];                                   // replicate pad, left
      if (q == nq - 1) trow[w + 1] = v[(w - 1) & 3];                // replicate pad, right (column w-1)
    }
    __syncthreads();
    const int ntask = ((rows + 1) >> 1) * nq;
    for (int task = tid; task < ntask; task += kConvThreads) {
      const int rp = task / nq, q4 = task - rp * nq;
      const int r = rp * 2;
      float acc[2][3][4];
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int co = 0; co < 3; ++co)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[a][co][j] = 0.0f;
#pragma unroll 1
      for (int ci = 0; ci < 3; ++ci) {
        float x[4][6];
#pragma unroll
        for (int y = 0; y < 4; ++y) {
          const float* trow = tile + (ci * (kBand + 2) + r + y) * RS + q4 * 4;
          const float4 a = *reinterpret_cast<const float4*>(trow);
          const float2 b = *reinterpret_cast<const float2*>(trow + 4);
          x[y][0] = a.x; x[y][1] = a.y; x[y][2] = a.z; x[y][3] = a.w; x[y][4] = b.x; x[y][5] = b.y;
        }
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const float4 wv = *reinterpret_cast<const float4*>(wsm + (ci * 9 + ky * 3 + kx) * 4);
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                acc[a][0][j] = fmaf(x[a + ky][j + kx], wv.x, acc[a][0][j]);
                acc[a][1][j] = fmaf(x[a + ky][j + kx], wv.y, acc[a][1][j]);
                acc[a][2][j] = fmaf(x[a + ky][j + kx], wv.z, acc[a][2][j]);
              }
          }
        }
      }
      const int col = q4 * 4;
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        if (r + a >= rows) break;
#pragma unroll
        for (int co = 0; co < 3; ++co) {
          float* dst = out + (((int64_t)frame * 3 + co) * h + (r0 + r + a)) * w + col;
          const float4 o = make_float4(sigmoid255(acc[a][co][0]), sigmoid255(acc[a][co][1]), sigmoid255(acc[a][co][2]),
                                       sigmoid255(acc[a][co][3]));
          if (vec) {
            *reinterpret_cast<float4*>(dst) = o;
          } else {
            const float ov[4] = {o.x, o.y, o.z, o.w};
            for (int j = 0; j < 4 && col + j < w; ++j) dst[j] = ov[j];
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// random_conv, TMA-staged (w % 4 == 0, 16-B aligned tensors): same 2x4 register blocking, but the band is
// brought in RAW by one cp.async.bulk per row (336 B at w = 84) into a double-buffered tile -- the next
// work item's rows are in flight while this one is computed, and no thread spends instructions on the
// fill.  The /255 moves into the 81 weights (conv(x/255, W) == conv(x, W/255) up to fp32 rounding, well
// inside the 1e-5 normalised tolerance); the replicate columns are patched by 3*(rows+2) threads after the
// copy lands.  Tile row layout: [3]=left pad, [4 .. 4+w) = pixels (16-B aligned for the bulk copy and the
// LDS.128 of the window), [4+w] = right pad.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kConvThreads, 7) random_conv_tma_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                                          int frames, int k, int h, int w,
                                                                          const float* __restrict__ weights, int bands, int RS) {
  extern __shared__ __align__(128) float ctsm[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(ctsm);      // 2 mbarriers (16 B), padded to 16 floats
  float* wsm = ctsm + 16;                                  // [27][4]
  float* tiles = ctsm + 16 + 27 * 4;                        // [2][3][kBand+2][RS]   (124 floats in: 16-B aligned)
  const int tile_floats = 3 * (kBand + 2) * RS;
  const int tid = threadIdx.x;
  const int nq = w >> 2;
  const int total = frames * bands;
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_fence_init();
  }
  __syncthreads();
  pdl_wait();

  auto issue = [&](int work, int buf) {   // threads 0 .. 3*trows-1: one row each
    const int frame = work / bands, band = work - frame * bands;
    const int r0 = band * kBand;
    const int trows = min(kBand, h - r0) + 2;
    if (tid == 0) mbar_expect_tx(&bars[buf], (uint32_t)(3 * trows * w * 4));
    if (tid < 3 * trows) {
      const int ci = tid / trows, pr = tid - ci * trows;
      const int rr = clampi(r0 + pr - 1, 0, h - 1);
      const float* src = in + (((int64_t)frame * 3 + ci) * h + rr) * w;
      float* dst = tiles + buf * tile_floats + (ci * (kBand + 2) + pr) * RS + 4;
      bulk_g2s(dst, src, (uint32_t)(w * 4), &bars[buf]);
    }
  };

  if ((int)blockIdx.x < total) issue(blockIdx.x, 0);
  pdl_launch_dependents();
  int n = 0;
  for (int work = blockIdx.x; work < total; work += gridDim.x, ++n) {
    const int buf = n & 1;
    const uint32_t phase = (uint32_t)(n >> 1) & 1u;
    if (work + (int)gridDim.x < total) issue(work + gridDim.x, buf ^ 1);   // that buffer was released by the sync below
    const int frame = work / bands, band = work - frame * bands;
    const int sample = frame / k;
    const int r0 = band * kBand;
    const int rows = min(kBand, h - r0);
    if (tid < 81) {
      // weights[sample][co][ci][ky][kx] / 255 -> wsm[(ci*9 + ky*3 + kx)*4 + co]
      const int co = tid / 27, tap = tid - co * 27;
      wsm[tap * 4 + co] = __fdiv_rn(weights[(int64_t)sample * 81 + tid], 255.0f);
    }
    mbar_wait(&bars[buf], phase);
    float* tile = tiles + buf * tile_floats;
    if (tid < 3 * (rows + 2)) {
      const int trows = rows + 2;
      const int ci = tid / trows, pr = tid - ci * trows;
      float* trow = tile + (ci * (kBand + 2) + pr) * RS;
      trow[3] = trow[4];
      trow[4 + w] = trow[3 + w];
    }
    __syncthreads();
    const int ntask = ((rows + 1) >> 1) * nq;
    for (int task = tid; task < ntask; task += kConvThreads) {
      const int rp = task / nq, q4 = task - rp * nq;
      const int r = rp * 2;
      float acc[2][3][4];
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int co = 0; co < 3; ++co)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[a][co][j] = 0.0f;
#pragma unroll 1
      for (int ci = 0; ci < 3; ++ci) {
        float x[4][6];
#pragma unroll
        for (int y = 0; y < 4; ++y) {
          const float* trow = tile + (ci * (kBand + 2) + r + y) * RS + q4 * 4;
          const float4 a = *reinterpret_cast<const float4*>(trow + 4);
          x[y][0] = trow[3]; x[y][1] = a.x; x[y][2] = a.y; x[y][3] = a.z; x[y][4] = a.w; x[y][5] = trow[8];
        }
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const float4 wv = *reinterpret_cast<const float4*>(wsm + (ci * 9 + ky * 3 + kx) * 4);
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                acc[a][0][j] = fmaf(x[a + ky][j + kx], wv.x, acc[a][0][j]);
                acc[a][1][j] = fmaf(x[a + ky][j + kx], wv.y, acc[a][1][j]);
                acc[a][2][j] = fmaf(x[a + ky][j + kx], wv.z, acc[a][2][j]);
              }
          }
        }
      }
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        if (r + a >= rows) break;
#pragma unroll
        for (int co = 0; co < 3; ++co) {
          float4* dst = reinterpret_cast<float4*>(out + (((int64_t)frame * 3 + co) * h + (r0 + r + a)) * w) + q4;
          st_global_f4(dst, make_float4(sigmoid255(acc[a][co][0]), sigmoid255(acc[a][co][1]), sigmoid255(acc[a][co][2]),
                                        sigmoid255(acc[a][co][3])));
        }
      }
    }
    __syncthreads();   // tile[buf] and wsm are free again
  }
}

// ---------------------------------------------------------------------------------------------
// random_conv, part-of-frame variant (w % 4 == 0, 16-B aligned tensors): a work item is R consecutive rows of one RGB
// frame (R = 42 at 84x84: half a frame), brought in by THREE cp.async.bulk (each plane's rows are contiguous) into one
// tile whose row stride is w itself; the replicate pad is index clamping (rows) and a select in registers (columns), so
// nothing touches the tile after the copy lands.  224 threads = 441 tasks in two rounds; one item = 7x the work of the
// 12-row band kernel between CTA barriers.  Single tile: the other CTAs of the SM compute while this one loads.
// ---------------------------------------------------------------------------------------------
constexpr int kPartThreads = 224;
__global__ void __launch_bounds__(kPartThreads, 4) random_conv_part_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                                           int frames, int k, int h, int w,
                                                                           const float* __restrict__ weights, int parts, int R) {
  extern __shared__ __align__(128) float psm[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(psm);        // 1 mbarrier; header padded to 128 floats
  float* wsm = psm + 16;                                   // [27][4]
  float* tile = psm + 128;                                 // [3][R+2][w]
  const int tid = threadIdx.x;
  const int nq = w >> 2;
  const int total = frames * parts;
  if (tid == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();
  int n = 0;
  for (int work = blockIdx.x; work < total; work += gridDim.x, ++n) {
    const int frame = work / parts, part = work - frame * parts;
    const int sample = frame / k;
    const int r0 = part * R;
    const int rows = min(R, h - r0);
    const int lo = max(r0 - 1, 0), hi = min(r0 + rows, h - 1);      // source rows [lo, hi] are staged
    const int nrows = hi - lo + 1;
    __syncthreads();                                       // the previous item no longer reads the tile / wsm
    if (tid == 0) {
      const uint32_t bytes = (uint32_t)(nrows * w * 4);
      mbar_expect_tx(bar, 3 * bytes);
#pragma unroll
      for (int ci = 0; ci < 3; ++ci)
        bulk_g2s(tile + ci * (R + 2) * w, in + (((int64_t)frame * 3 + ci) * h + lo) * w, bytes, bar);
    }
    if (tid < 81) {
      // weights[sample][co][ci][ky][kx] / 255 -> wsm[(ci*9 + ky*3 + kx)*4 + co]
      const int co = tid / 27, tap = tid - co * 27;
      wsm[tap * 4 + co] = weights[(int64_t)sample * 81 + tid] * kConvWScale;   // scaled for conv_sigmoid
    }
    __syncthreads();
    mbar_wait(bar, (uint32_t)n & 1u);
    const int ntask = ((rows + 1) >> 1) * nq;
    for (int task = tid; task < ntask; task += (int)blockDim.x) {
      const int rp = task / nq, q4 = task - rp * nq;
      const int r = rp * 2;
      const bool left = q4 == 0, right = q4 == nq - 1;
      int trow_off[4];
#pragma unroll
      for (int y = 0; y < 4; ++y) trow_off[y] = (clampi(r0 + r + y - 1, 0, h - 1) - lo) * w + q4 * 4;
      // output channels 0 and 1 share one packed FFMA2 per (pixel, tap): the weight pair (w0, w1) is an aligned register
      // pair out of the LDS.128, the pixel is FFMA2's scalar-broadcast operand -- no re-packing; channel 2 is a plain FFMA
      float2 acc01[2][4];
      float acc2[2][4];
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int j = 0; j < 4; ++j) { acc01[a][j] = make_float2(0.0f, 0.0f); acc2[a][j] = 0.0f; }
#pragma unroll 1
      for (int ci = 0; ci < 3; ++ci) {
        const float* pl = tile + ci * (R + 2) * w;
        float x[4][6];
#pragma unroll
        for (int y = 0; y < 4; ++y) {
          const float* trow = pl + trow_off[y];
          const float4 a = *reinterpret_cast<const float4*>(trow);
          const float l = trow[-1], rt = trow[4];
          x[y][0] = left ? a.x : l; x[y][1] = a.x; x[y][2] = a.y; x[y][3] = a.z; x[y][4] = a.w; x[y][5] = right ? a.w : rt;
        }
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const float4 wv = *reinterpret_cast<const float4*>(wsm + (ci * 9 + ky * 3 + kx) * 4);
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float xv = x[a + ky][j + kx];
                acc01[a][j] = __ffma2_rn(make_float2(xv, xv), make_float2(wv.x, wv.y), acc01[a][j]);
                acc2[a][j] = fmaf(xv, wv.z, acc2[a][j]);
              }
          }
        }
      }
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        if (r + a >= rows) break;
        float4* dst = reinterpret_cast<float4*>(out + (((int64_t)frame * 3) * h + (r0 + r + a)) * w) + q4;
        const int64_t pl4 = (int64_t)h * w / 4;
        st_global_f4(dst, make_float4(conv_sigmoid<false>(acc01[a][0].x), conv_sigmoid<false>(acc01[a][1].x),
                                      conv_sigmoid<false>(acc01[a][2].x), conv_sigmoid<false>(acc01[a][3].x)));
        st_global_f4(dst + pl4, make_float4(conv_sigmoid<false>(acc01[a][0].y), conv_sigmoid<false>(acc01[a][1].y),
                                            conv_sigmoid<false>(acc01[a][2].y), conv_sigmoid<false>(acc01[a][3].y)));
        st_global_f4(dst + 2 * pl4, make_float4(conv_sigmoid<false>(acc2[a][0]), conv_sigmoid<false>(acc2[a][1]),
                                                conv_sigmoid<false>(acc2[a][2]), conv_sigmoid<false>(acc2[a][3])));
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// overlay / splice: streaming elementwise kernels.  Background rows come from a device-resident bank
// ([M][3][h][w], u8 or f32 in [0,1]); a u8 pixel is divided by 255 like torchvision's ToTensor.
// ---------------------------------------------------------------------------------------------
// Thread = one 4-pixel group of one background channel (sample i, ch, p4): the background pixels are
// loaded and normalised once and reused for the k frames of the sample (channel f*3+ch of x uses img[ch]).
// blockIdx.x = (i * 3 + ch) * chunks + chunk: the decode is per CTA, not per thread.
template <bool BANK_U8, int K>
__global__ void __launch_bounds__(256) overlay_vec_kernel(const float4* __restrict__ in, float4* __restrict__ out,
                                                          int k, int hw4, int chunks, const void* __restrict__ bank,
                                                          int64_t bank_len, const int64_t* __restrict__ bank_idx,
                                                          float oma, float alpha) {
  pdl_wait();                 // launched programmatically: nothing is read before the previous kernel has completed
  pdl_launch_dependents();
  const int chunk = blockIdx.x % chunks;
  const int ic = blockIdx.x / chunks;
  const int i = ic / 3, ch = ic - 3 * i;
  const int p4 = chunk * 256 + threadIdx.x;
  if (p4 >= hw4) return;
  int64_t row = bank_idx[i];
  row = row < 0 ? 0 : (row >= bank_len ? bank_len - 1 : row);
  const int64_t boff = (row * 3 + ch) * (int64_t)hw4 + p4;
  float4 g;
  if (BANK_U8) {
    const float4 t = bytes_to_float4(__ldg(reinterpret_cast<const uint32_t*>(bank) + boff));
    g = make_float4(div255(t.x), div255(t.y), div255(t.z), div255(t.w));
  } else {
    g = __ldg(reinterpret_cast<const float4*>(bank) + boff);
  }
  const int64_t base = ((int64_t)i * 3 * k + ch) * hw4 + p4;
  if constexpr (K > 0) {   // frame_stack known at compile time: all K loads are issued before the first use
    float4 x[K > 0 ? K : 1];
#pragma unroll
    for (int f = 0; f < K; ++f) x[f] = __ldg(in + base + (int64_t)f * 3 * hw4);
#pragma unroll
    for (int f = 0; f < K; ++f)
      out[base + (int64_t)f * 3 * hw4] = make_float4(overlay_px(x[f].x, g.x, oma, alpha), overlay_px(x[f].y, g.y, oma, alpha),
                                                    overlay_px(x[f].z, g.z, oma, alpha), overlay_px(x[f].w, g.w, oma, alpha));
  } else {
    for (int f = 0; f < k; ++f) {
      const int64_t o = base + (int64_t)f * 3 * hw4;
      const float4 x = __ldg(in + o);
      out[o] = make_float4(overlay_px(x.x, g.x, oma, alpha), overlay_px(x.y, g.y, oma, alpha),
                           overlay_px(x.z, g.z, oma, alpha), overlay_px(x.w, g.w, oma, alpha));
    }
  }
}

template <bool BANK_U8>
__global__ void __launch_bounds__(256) overlay_generic_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                              int64_t total, int c, int hw, const void* __restrict__ bank,
                                                              int64_t bank_len, const int64_t* __restrict__ bank_idx,
                                                              float oma, float alpha) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t plane = e / hw;
    const int px = (int)(e - plane * hw);
    const int i = (int)(plane / c), ch = (int)(plane - (int64_t)i * c);
    int64_t row = bank_idx[i];
    row = row < 0 ? 0 : (row >= bank_len ? bank_len - 1 : row);
    const int64_t boff = (row * 3 + (ch % 3)) * (int64_t)hw + px;
    const float g = BANK_U8 ? __fdiv_rn((float)reinterpret_cast<const uint8_t*>(bank)[boff], 255.0f)
                            : reinterpret_cast<const float*>(bank)[boff];
    out[e] = overlay_px(in[e], g, oma, alpha);
  }
}

template <bool BANK_U8>
__device__ __forceinline__ float4 bank_vec(const void* bank, int64_t off4) {
  if (BANK_U8) {
    const float4 t = bytes_to_float4(__ldg(reinterpret_cast<const uint32_t*>(bank) + off4));
    return make_float4(div255(t.x), div255(t.y), div255(t.z), div255(t.w));
  }
  return __ldg(reinterpret_cast<const float4*>(bank) + off4);
}

// Thread = one 4-pixel group of one RGB frame; blockIdx.x = frame * chunks + chunk (per-CTA decode).
template <bool BANK_U8>
__global__ void __launch_bounds__(256) splice_vec_kernel(const float4* __restrict__ in, float4* __restrict__ out,
                                                         int n, int hw4, int chunks, const void* __restrict__ bank,
                                                         int64_t bank_len, const int64_t* __restrict__ bank_idx,
                                                         SpliceThr thr) {
  pdl_wait();
  pdl_launch_dependents();
  const int chunk = blockIdx.x % chunks;
  const int frame = blockIdx.x / chunks;     // row j of x.reshape(n*k, 3, h, w)
  const int p4 = chunk * 256 + threadIdx.x;
  if (p4 >= hw4) return;
  int64_t row = bank_idx[frame % n];         // overlay = batch.repeat(k,1,1,1): row j <- background j mod n
  row = row < 0 ? 0 : (row >= bank_len ? bank_len - 1 : row);
  const int64_t xo = (int64_t)frame * 3 * hw4 + p4, bo = row * 3 * hw4 + p4;
  const float4 xr = __ldg(in + xo), xg = __ldg(in + xo + hw4), xb = __ldg(in + xo + 2 * (int64_t)hw4);
  const float4 gr = bank_vec<BANK_U8>(bank, bo), gg = bank_vec<BANK_U8>(bank, bo + hw4),
               gb = bank_vec<BANK_U8>(bank, bo + 2 * (int64_t)hw4);
  float4 o_r, o_g, o_b;
  splice_px(xr.x, xg.x, xb.x, gr.x, gg.x, gb.x, thr, o_r.x, o_g.x, o_b.x);
  splice_px(xr.y, xg.y, xb.y, gr.y, gg.y, gb.y, thr, o_r.y, o_g.y, o_b.y);
  splice_px(xr.z, xg.z, xb.z, gr.z, gg.z, gb.z, thr, o_r.z, o_g.z, o_b.z);
  splice_px(xr.w, xg.w, xb.w, gr.w, gg.w, gb.w, thr, o_r.w, o_g.w, o_b.w);
  out[xo] = o_r;
  out[xo + hw4] = o_g;
  out[xo + 2 * (int64_t)hw4] = o_b;
}

template <bool BANK_U8>
__global__ void __launch_bounds__(256) splice_generic_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                             int64_t nwork, int n, int hw, const void* __restrict__ bank,
                                                             int64_t bank_len, const int64_t* __restrict__ bank_idx,
                                                             SpliceThr thr) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nwork; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t frame = v / hw;
    const int px = (int)(v - frame * hw);
    int64_t row = bank_idx[frame % n];
    row = row < 0 ? 0 : (row >= bank_len ? bank_len - 1 : row);
    const int64_t xo = frame * 3 * hw + px, bo = row * 3 * hw + px;
    float g[3];
    for (int j = 0; j < 3; ++j)
      g[j] = BANK_U8 ? __fdiv_rn((float)reinterpret_cast<const uint8_t*>(bank)[bo + (int64_t)j * hw], 255.0f)
                     : reinterpret_cast<const float*>(bank)[bo + (int64_t)j * hw];
    float o_r, o_g, o_b;
    splice_px(in[xo], in[xo + hw], in[xo + 2 * (int64_t)hw], g[0], g[1], g[2], thr, o_r, o_g, o_b);
    out[xo] = o_r;
    out[xo + hw] = o_g;
    out[xo + 2 * (int64_t)hw] = o_b;
  }
}

static inline int grid_for(int64_t work_items, int threads, int max_ctas_per_sm) {
  int64_t blocks = (work_items + threads - 1) / threads;
  const int64_t cap = (int64_t)kNumSMs * max_ctas_per_sm;
  return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

static int check_image_args(const char* who, const float* in, float* out, int n, int c, int h, int w) {
  AUGCL_REQUIRE(in && out, "%s: in/out is NULL", who);
  AUGCL_REQUIRE(in != out, "%s: in and out must not alias", who);
  AUGCL_REQUIRE(n > 0 && c > 0 && h > 0 && w > 0, "%s: n, c, h, w must be positive (got %d,%d,%d,%d)", who, n, c, h, w);
  AUGCL_REQUIRE((int64_t)n * c < (1ll << 31) && (int64_t)h * w < (1ll << 28), "%s: shape too large", who);
  if (!aligned(in, 4) || !aligned(out, 4)) AUGCL_FAIL(AUGCL_ERR_BAD_ALIGN, "%s: pointers must be 4-byte aligned", who);
  return AUGCL_OK;
}

static int launch_window_copy(const float* in, float* out, int n, int c, int H, int W, int Ho, int Wo,
                              const int32_t* offsets, int pad, cudaStream_t st) {
  const int planes = n * c;
  if (g_staging.load() == 0 && (Wo & 3) == 0 && aligned(out, 16)) {
    const int grid = planes < kNumSMs * 8 ? planes : kNumSMs * 8;
    const bool pdl = g_sample_pdl.load() != 0;
    const cudaError_t le = Wo == 84 ? launch_pdl(window_copy_vec_kernel<21>, grid, 256, 0, st, pdl, in, out, planes, c, H, W, Ho, Wo, offsets, pad)
                                    : launch_pdl(window_copy_vec_kernel<0>, grid, 256, 0, st, pdl, in, out, planes, c, H, W, Ho, Wo, offsets, pad);
    if (le != cudaSuccess) AUGCL_FAIL(AUGCL_ERR_CUDA, "window copy: launch failed: %s", cudaGetErrorString(le));
  } else {
    const int64_t total = (int64_t)planes * Ho * Wo;
    window_copy_generic_kernel<<<grid_for(total, 256, 16), 256, 0, st>>>(in, out, total, c, H, W, Ho, Wo, offsets, pad);
  }
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

}  // namespace augcl

using namespace augcl;

extern "C" {

int augcl_drq_base_grid(int h, int pad, float* out_host) {
  AUGCL_REQUIRE(out_host && h > 0 && pad >= 0, "augcl_drq_base_grid: bad arguments");
  // torch.linspace(-1+eps, 1-eps, steps, dtype=float32)[:h]: fp32 start/end/step, symmetric evaluation
  // (first half fma(step, i, start), second half fma(-step, steps-1-i, end): ATen's vectorised CPU kernel fuses).
  const int steps = h + 2 * pad;
  const double eps = 1.0 / (double)steps;
  const float start = (float)(-1.0 + eps), end = (float)(1.0 - eps);
  const float step = steps > 1 ? (end - start) / (float)(steps - 1) : 0.0f;
  const int halfway = steps / 2;
  for (int i = 0; i < h; ++i)
    out_host[i] = i < halfway ? fmaf(step, (float)i, start) : fmaf(-step, (float)(steps - i - 1), end);
  return AUGCL_OK;
}

int augcl_random_shift(const float* in, float* out, int n, int c, int h, int w, int pad, const int32_t* offsets,
                       int mode, const float* base, augcl_stream stream) {
  if (int rc = check_image_args("augcl_random_shift", in, out, n, c, h, w)) return rc;
  AUGCL_REQUIRE(offsets, "augcl_random_shift: offsets is NULL");
  AUGCL_REQUIRE(pad >= 0 && pad < (1 << 15), "augcl_random_shift: bad pad %d", pad);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == AUGCL_SHIFT_INT) return launch_window_copy(in, out, n, c, h, w, h, w, offsets, pad, st);
  AUGCL_REQUIRE(mode == AUGCL_SHIFT_BILINEAR, "augcl_random_shift: unknown mode %d", mode);
  AUGCL_REQUIRE(base, "augcl_random_shift: BILINEAR needs the base grid");
  if (h != w) AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "augcl_random_shift: BILINEAR (DrQ-v2 grid) requires h == w");
  const size_t bsmem = (size_t)((h + 3) & ~3) * 16 + (size_t)h * w * 4;
  if (g_staging.load() == 0 && w <= 256 && h <= kBilMaxH && bsmem <= (size_t)kBilSmemMax) {
    if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(shift_bilinear_smem_kernel), kBilSmemMax)) return rc;
    const int planes = n * c;
    int vec = ((h * w) & 3) == 0 && aligned(in, 16);
    size_t smem = bsmem;
    if (vec && bsmem + (size_t)h * w * 4 <= (size_t)kBilSmemMax && g_bil_tma.load()) {
      vec = 2;                                   // double-buffered tiles filled by cp.async.bulk
      smem = bsmem + (size_t)h * w * 4;
    }
    const int per_sm = (int)((228 * 1024) / (smem + 1024));
    const int cap = kNumSMs * (per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm));
    launch_pdl(shift_bilinear_smem_kernel, (unsigned)(planes < cap ? planes : cap), 256, smem, st, g_sample_pdl.load() != 0, in, out, planes, c,
               h, w, pad, offsets, base, vec);
    AUGCL_LAUNCHED();
    return AUGCL_OK;
  }
  if (g_staging.load() == 0 && (w & 3) == 0 && w <= 1024 && aligned(out, 16)) {
    const int planes = n * c;
    shift_bilinear_vec_kernel<<<planes < kNumSMs * 8 ? planes : kNumSMs * 8, 256, 0, st>>>(in, out, planes, c, h, w, pad,
                                                                                        offsets, base);
    AUGCL_LAUNCHED();
    return AUGCL_OK;
  }
  const int64_t total = (int64_t)n * c * h * w;
  shift_bilinear_kernel<<<grid_for(total, 256, 16), 256, 0, st>>>(in, out, total, c, h, w, pad, offsets, base);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

int augcl_random_crop(const float* in, float* out, int n, int c, int H, int W, int size, const int32_t* offsets,
                      augcl_stream stream) {
  if (int rc = check_image_args("augcl_random_crop", in, out, n, c, H, W)) return rc;
  AUGCL_REQUIRE(offsets, "augcl_random_crop: offsets is NULL");
  AUGCL_REQUIRE(size > 0 && size <= H && size <= W, "augcl_random_crop: size %d does not fit %dx%d", size, H, W);
  return launch_window_copy(in, out, n, c, H, W, size, size, offsets, 0, (cudaStream_t)stream);
}

int augcl_random_conv(const float* in, float* out, int n, int c, int h, int w, const float* weights,
                      augcl_stream stream) {
  if (int rc = check_image_args("augcl_random_conv", in, out, n, c, h, w)) return rc;
  AUGCL_REQUIRE(weights, "augcl_random_conv: weights is NULL");
  AUGCL_REQUIRE(c % 3 == 0, "augcl_random_conv: c=%d is not a multiple of 3", c);
  const int k = c / 3, frames = n * k;
  const int bands = (h + kBand - 1) / kBand;
  const int RS = ((w + 2 + 3) & ~3) + 4;  // padded row stride: multiple of 4 floats, room for the float2 tail
  const size_t smem = (27 * 4 + 3 * (kBand + 2) * (size_t)RS) * sizeof(float);
  if (smem > 200 * 1024) AUGCL_FAIL(AUGCL_ERR_UNSUPPORTED, "augcl_random_conv: width %d too large", w);
  if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(random_conv_kernel), 200 * 1024)) return rc;
  const int vec = ((w & 3) == 0 && aligned(in, 16) && aligned(out, 16)) ? 1 : 0;
  const int64_t work = (int64_t)frames * bands;
  AUGCL_REQUIRE(work < (1ll << 30), "augcl_random_conv: batch too large");
  if (vec && g_staging.load() == 0 && g_conv_part.load()) {
    // part-of-frame items: the fewest parts whose tile fits ~55 KB (4 CTAs per SM) -- measured best from batch 128 to
    // 2048 (scripts/tune_conv.py; conv_part >= 2 forces the count)
    int parts = 1;
    while (parts < h && (size_t)3 * ((h + parts - 1) / parts + 2) * w * 4 > 54 * 1024) ++parts;
    if (g_conv_part.load() >= 2) parts = g_conv_part.load();
    const int R = (((h + parts - 1) / parts) + 1) & ~1;          // even: row pairs never straddle two items
    parts = (h + R - 1) / R;
    // threads: the item's tasks in the fewest equal rounds of <= 224 threads
    const int ntask = (R / 2) * (w / 4);
    const int rounds = (ntask + kPartThreads - 1) / kPartThreads;
    int threads = (((ntask + rounds - 1) / rounds) + 31) & ~31;
    if (threads < 96) threads = 96;                              // 81 threads stage the weights
    const size_t smem3 = (128 + (size_t)3 * (R + 2) * w + 4) * sizeof(float);
    if (smem3 <= 100 * 1024) {
      if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(random_conv_part_kernel), 100 * 1024)) return rc;
      int per_sm = (int)((228 * 1024) / (smem3 + 1024));
      const int by_regs = 65536 / (72 * threads);
      per_sm = per_sm > by_regs ? by_regs : per_sm;
      per_sm = per_sm < 1 ? 1 : per_sm;
      const int64_t items = (int64_t)frames * parts;
      const int grid = (int)(items < kNumSMs * per_sm ? items : kNumSMs * per_sm);
      // (no programmatic launch here: measured 22.5 -> 24.3 us at batch 128 with it -- waiting successor CTAs hold 45 KB tiles)
      const cudaError_t le = launch_pdl(random_conv_part_kernel, grid, threads, smem3, (cudaStream_t)stream, false, in, out,
                                        frames, k, h, w, weights, parts, R);
      if (le != cudaSuccess) AUGCL_FAIL(AUGCL_ERR_CUDA, "augcl_random_conv: launch failed: %s", cudaGetErrorString(le));
      AUGCL_LAUNCHED();
      return AUGCL_OK;
    }
  }
  if (vec && g_staging.load() == 0) {
    const int RS2 = (w + 8) & ~3;
    const size_t smem2 = (16 + 27 * 4 + 2 * 3 * (kBand + 2) * (size_t)RS2) * sizeof(float);
    if (smem2 <= 200 * 1024) {
      if (int rc = ensure_dyn_smem(reinterpret_cast<const void*>(random_conv_tma_kernel), 200 * 1024)) return rc;
      int per_sm = (int)((228 * 1024) / (smem2 + 1024));
      per_sm = per_sm < 1 ? 1 : (per_sm > 7 ? 7 : per_sm);
      const int grid = (int)(work < kNumSMs * per_sm ? work : kNumSMs * per_sm);
      random_conv_tma_kernel<<<grid, kConvThreads, smem2, (cudaStream_t)stream>>>(in, out, frames, k, h, w, weights, bands, RS2);
      AUGCL_LAUNCHED();
      return AUGCL_OK;
    }
  }
  const int grid = (int)(work < kNumSMs * 12 ? work : kNumSMs * 12);
  random_conv_kernel<<<grid, kConvThreads, smem, (cudaStream_t)stream>>>(in, out, frames, k, h, w, weights, bands, RS, vec);
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

static int check_bank(const char* who, const void* bank, int bank_dtype, int64_t bank_len, const int64_t* bank_idx) {
  AUGCL_REQUIRE(bank && bank_idx, "%s: bank/bank_idx is NULL", who);
  AUGCL_REQUIRE(bank_len > 0, "%s: bank_len must be positive", who);
  AUGCL_REQUIRE(bank_dtype == AUGCL_BANK_U8 || bank_dtype == AUGCL_BANK_F32, "%s: unknown bank_dtype %d", who, bank_dtype);
  return AUGCL_OK;
}

int augcl_random_overlay(const float* in, float* out, int n, int c, int h, int w, const void* bank, int bank_dtype,
                         int64_t bank_len, const int64_t* bank_idx, float alpha, augcl_stream stream) {
  if (int rc = check_image_args("augcl_random_overlay", in, out, n, c, h, w)) return rc;
  if (int rc = check_bank("augcl_random_overlay", bank, bank_dtype, bank_len, bank_idx)) return rc;
  AUGCL_REQUIRE(c % 3 == 0, "augcl_random_overlay: c=%d is not a multiple of 3", c);
  cudaStream_t st = (cudaStream_t)stream;
  const float oma = (float)(1.0 - (double)alpha);
  const int hw = h * w;
  const int64_t total = (int64_t)n * c * hw;
  const bool u8 = bank_dtype == AUGCL_BANK_U8;
  const bool vec = g_staging.load() == 0 && (hw & 3) == 0 && aligned(in, 16) && aligned(out, 16) &&
                   aligned(bank, u8 ? 4 : 16);
  if (vec) {
    const int hw4 = hw / 4, chunks = (hw4 + 255) / 256;
    const int64_t grid64 = (int64_t)n * 3 * chunks;
    AUGCL_REQUIRE(grid64 < (1ll << 31), "augcl_random_overlay: batch too large");
    const unsigned grid = (unsigned)grid64;
    const int k = c / 3;
#define AUGCL_OV(U8, KK) launch_pdl(overlay_vec_kernel<U8, KK>, grid, 256, 0, st, g_sample_pdl.load() != 0, (const float4*)in, (float4*)out, k, hw4, chunks, bank, bank_len, bank_idx, oma, alpha)
    if (u8) { if (k == 3) AUGCL_OV(true, 3); else AUGCL_OV(true, 0); }
    else    { if (k == 3) AUGCL_OV(false, 3); else AUGCL_OV(false, 0); }
#undef AUGCL_OV
  } else {
    const int grid = grid_for(total, 256, 16);
    if (u8)
      overlay_generic_kernel<true><<<grid, 256, 0, st>>>(in, out, total, c, hw, bank, bank_len, bank_idx, oma, alpha);
    else
      overlay_generic_kernel<false><<<grid, 256, 0, st>>>(in, out, total, c, hw, bank, bank_len, bank_idx, oma, alpha);
  }
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

int augcl_splice(const float* in, float* out, int n, int c, int h, int w, const void* bank, int bank_dtype,
                 int64_t bank_len, const int64_t* bank_idx, float hue_thres, float sat_thres, float val_thres,
                 int variant, augcl_stream stream) {
  if (int rc = check_image_args("augcl_splice", in, out, n, c, h, w)) return rc;
  if (int rc = check_bank("augcl_splice", bank, bank_dtype, bank_len, bank_idx)) return rc;
  AUGCL_REQUIRE(c % 3 == 0, "augcl_splice: c=%d is not a multiple of 3", c);
  AUGCL_REQUIRE(variant == AUGCL_SPLICE || variant == AUGCL_CS_SPLICE, "augcl_splice: unknown variant %d", variant);
  cudaStream_t st = (cudaStream_t)stream;
  const SpliceThr thr{hue_thres, sat_thres, val_thres, variant};
  const int hw = h * w;
  const int64_t frames = (int64_t)n * (c / 3);
  const bool u8 = bank_dtype == AUGCL_BANK_U8;
  const bool vec = g_staging.load() == 0 && (hw & 3) == 0 && aligned(in, 16) && aligned(out, 16) &&
                   aligned(bank, u8 ? 4 : 16);
  if (vec) {
    const int hw4 = hw / 4, chunks = (hw4 + 255) / 256;
    const int64_t grid64 = frames * chunks;
    AUGCL_REQUIRE(grid64 < (1ll << 31), "augcl_splice: batch too large");
    const unsigned grid = (unsigned)grid64;
    if (u8)
      launch_pdl(splice_vec_kernel<true>, grid, 256, 0, st, g_sample_pdl.load() != 0, (const float4*)in, (float4*)out, n, hw4, chunks, bank, bank_len, bank_idx, thr);
    else
      launch_pdl(splice_vec_kernel<false>, grid, 256, 0, st, g_sample_pdl.load() != 0, (const float4*)in, (float4*)out, n, hw4, chunks, bank, bank_len, bank_idx, thr);
  } else {
    const int64_t nwork = frames * hw;
    const int grid = grid_for(nwork, 256, 16);
    if (u8)
      splice_generic_kernel<true><<<grid, 256, 0, st>>>(in, out, nwork, n, hw, bank, bank_len, bank_idx, thr);
    else
      splice_generic_kernel<false><<<grid, 256, 0, st>>>(in, out, nwork, n, hw, bank, bank_len, bank_idx, thr);
  }
  AUGCL_LAUNCHED();
  return AUGCL_OK;
}

}  // extern "C"
