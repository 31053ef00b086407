// Stage 1 — image pyramid with Sobel gradients, replacing the OpenCV calls of
// ConstructImgPyramide (run_dense.cpp:131-179) and the dimension padding of
// run_dense.cpp:299-312.
//
// cv::resize(.5,.5,INTER_LINEAR) is a 2x2 box mean and a cascade of l such steps on a
// u8-sourced image is the exact mean of the 2^l x 2^l block (all intermediates are dyadic
// rationals below 2^24, l <= 8), so the finest level that is actually read (sc_l) is computed
// straight from the u8 frame in one pass; levels 0..sc_l-1, which the reference builds but
// never reads (oflow.cpp:182), are skipped.  Coarser levels come from the previous one.
#include "dis_internal.h"

namespace disb200 {

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }
__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return clampi(i, 0, n - 1);
}

// One thread per (level pixel, channel).  Reads the replicate-padded u8 frame.
template <int C>
__global__ void downscale_u8_kernel(const uint8_t* __restrict__ frames, int w0, int h0, int padl, int padt,
                                    int lvl, int w, int h, int pad, int tw, size_t img_elems,
                                    float* __restrict__ img) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int pair = blockIdx.z;
  if (x >= w) return;
  const uint8_t* f = frames + (size_t)pair * w0 * h0 * C;
  const int k = 1 << lvl;
  unsigned sum[C];
#pragma unroll
  for (int c = 0; c < C; ++c) sum[c] = 0;
  for (int j = 0; j < k; ++j) {
    const int sy = clampi(y * k + j - padt, 0, h0 - 1);
    const uint8_t* row = f + (size_t)sy * w0 * C;
    for (int i = 0; i < k; ++i) {
      const int sx = clampi(x * k + i - padl, 0, w0 - 1);
#pragma unroll
      for (int c = 0; c < C; ++c) sum[c] += row[sx * C + c];
    }
  }
  const float inv = 1.0f / (float)(1u << (2 * lvl));
  float* o = img + (size_t)pair * img_elems + ((size_t)(y + pad) * tw + (x + pad)) * C;
#pragma unroll
  for (int c = 0; c < C; ++c) o[c] = (float)sum[c] * inv;
}

template <int C>
__global__ void halve_kernel(const float* __restrict__ src, int stw, int spad, size_t s_elems, float* __restrict__ dst,
                             int w, int h, int pad, int tw, size_t d_elems) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int pair = blockIdx.z;
  if (x >= w) return;
  const float* s = src + (size_t)pair * s_elems;
  float* o = dst + (size_t)pair * d_elems + ((size_t)(y + pad) * tw + (x + pad)) * C;
  const float* r0 = s + ((size_t)(2 * y + spad) * stw + (2 * x + spad)) * C;
  const float* r1 = r0 + (size_t)stw * C;
#pragma unroll
  for (int c = 0; c < C; ++c) o[c] = ((r0[c] + r0[C + c]) + (r1[c] + r1[C + c])) * 0.25f;
}

// One thread per padded pixel: replicate the image border, Sobel(k=3, 1/8, REFLECT_101) on the
// unpadded level image for the interior of dx/dy, zeros on their border.
template <int C>
__global__ void sobel_pad_kernel(float* __restrict__ img, float* __restrict__ dx, float* __restrict__ dy, int w,
                                 int h, int pad, int tw, int th, size_t elems) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int pair = blockIdx.z;
  if (x >= tw) return;
  float* I = img + (size_t)pair * elems;
  const size_t o = ((size_t)y * tw + x) * C;
  float* gx = dx + (size_t)pair * elems + o;
  float* gy = dy + (size_t)pair * elems + o;
  const int sx = x - pad, sy = y - pad;
  const bool inside = sx >= 0 && sy >= 0 && sx < w && sy < h;
  if (!inside) {
    const float* srcp = I + ((size_t)(clampi(sy, 0, h - 1) + pad) * tw + (clampi(sx, 0, w - 1) + pad)) * C;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      I[o + c] = srcp[c];
      gx[c] = 0.f;
      gy[c] = 0.f;
    }
    return;
  }
  const int xm = reflect101(sx - 1, w) + pad, xp = reflect101(sx + 1, w) + pad, xc = sx + pad;
  const int ym = reflect101(sy - 1, h) + pad, yp = reflect101(sy + 1, h) + pad, yc = sy + pad;
#define S(X, Y, c) I[((size_t)(Y)*tw + (X)) * C + (c)]
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const float r0 = S(xp, ym, c) - S(xm, ym, c);
    const float r1 = S(xp, yc, c) - S(xm, yc, c);
    const float r2 = S(xp, yp, c) - S(xm, yp, c);
    gx[c] = 0.25f * r1 + 0.125f * (r2 + r0);
    const float s0 = 0.25f * S(xc, ym, c) + 0.125f * (S(xp, ym, c) + S(xm, ym, c));
    const float s2 = 0.25f * S(xc, yp, c) + 0.125f * (S(xp, yp, c) + S(xm, yp, c));
    gy[c] = s2 - s0;
  }
#undef S
}

void launch_downscale_u8(const uint8_t* frames, int n, int w0, int h0, int C, int padl, int padt, int lvl,
                         const LevelGeom& g, float* img, cudaStream_t s, LaunchStats* st) {
  dim3 blk(128), grd((g.w + 127) / 128, g.h, n);
  if (C == 1)
    downscale_u8_kernel<1><<<grd, blk, 0, s>>>(frames, w0, h0, padl, padt, lvl, g.w, g.h, g.pad, g.tw, g.img_elems, img);
  else
    downscale_u8_kernel<3><<<grd, blk, 0, s>>>(frames, w0, h0, padl, padt, lvl, g.w, g.h, g.pad, g.tw, g.img_elems, img);
  st->launches++;
}

void launch_halve(const float* src, const LevelGeom& gs, float* dst, const LevelGeom& gd, int n, int C,
                  cudaStream_t s, LaunchStats* st) {
  dim3 blk(128), grd((gd.w + 127) / 128, gd.h, n);
  if (C == 1)
    halve_kernel<1><<<grd, blk, 0, s>>>(src, gs.tw, gs.pad, gs.img_elems, dst, gd.w, gd.h, gd.pad, gd.tw, gd.img_elems);
  else
    halve_kernel<3><<<grd, blk, 0, s>>>(src, gs.tw, gs.pad, gs.img_elems, dst, gd.w, gd.h, gd.pad, gd.tw, gd.img_elems);
  st->launches++;
}

void launch_sobel_pad(float* img, float* dx, float* dy, const LevelGeom& g, int n, int C, cudaStream_t s,
                      LaunchStats* st) {
  dim3 blk(128), grd((g.tw + 127) / 128, g.th, n);
  if (C == 1)
    sobel_pad_kernel<1><<<grd, blk, 0, s>>>(img, dx, dy, g.w, g.h, g.pad, g.tw, g.th, g.img_elems);
  else
    sobel_pad_kernel<3><<<grd, blk, 0, s>>>(img, dx, dy, g.w, g.h, g.pad, g.tw, g.th, g.img_elems);
  st->launches++;
}

}  // namespace disb200
