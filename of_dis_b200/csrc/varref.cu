// Stage 4 — variational refinement of one level (VarRefClass, refine_variational.cpp:25-342
// and the FDF1.0.1 routines it drives).  All arithmetic is fp32 in the reference's operation
// order and this file is compiled with -fmad=false, so on identical inputs the lexicographic
// mode reproduces the reference bit for bit.
//
// Kernels (all batched over frame pairs through blockIdx.z / blockIdx.y):
//   warp_kernel     image_warp + the avg/Iz part of get_derivatives   (opticalflow_aux.c:33-99)
//   deriv1_kernel   Ix Iy Ixz Iyz   (5-tap [1,-8,0,8,-1]/12, replicate border; image.cpp:467-640)
//   deriv2_kernel   Ixx Ixy Iyy
//   smooth_kernel2  diffusivities {h, v} of compute_smoothness        (opticalflow_aux.c:138-187)
//   system_kernel   compute_data / compute_data_DE + sub_laplacian + the 2x2 block inverse of
//                   sor_coupled's first sweep                          (opticalflow_aux.c:194-223,
//                   409-747; solver.c:122-128)
//   sor_*_kernel    lexicographic SOR as an anti-diagonal wavefront    (solver.c:83-471): one warp per pair
//                   (sor_warp.cuh), one warp per 32-row band (sor_band.cuh), general CTA wavefront; plus the
//                   fused per-level kernel of small batches and the red-black variant
//   finish_kernel   flow += increment (with the DE sign clamp)         (refine_variational.cpp:208-227,301-323)
// The planes read by the refinement are the unpadded interiors of the padded pyramid arrays,
// addressed in place (the reference copies them out first, refine_variational.cpp:120-149).
#ifdef SORB_STATS
#include <cstdio>
#endif
#include <algorithm>

#include "dis_internal.h"

namespace disb200 {

namespace {

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

// coefficients of convolution_new(2, {0,-8/12,1/12}, 0): {1/12, -8/12, -0, 8/12, -1/12}
__device__ __forceinline__ float tap5(float s0, float s1, float s2, float s3, float s4) {
  const float h1 = -8.0f / 12.0f, h2 = 1.0f / 12.0f;
  return h2 * s0 + ((h1 * s1 + (-0.0f) * s2) + ((-h1) * s3 + (-h2) * s4));
}
// convolution_new(1, {0,-0.5}, 0): {-0.5, -0, 0.5}
__device__ __forceinline__ float tap3(float s0, float s1, float s2) { return -0.5f * s0 + ((-0.0f) * s1 + 0.5f * s2); }

__device__ __forceinline__ float mla(float a, float b, float c) { return a + b * c; }
__device__ __forceinline__ float mls(float a, float b, float c) { return a - b * c; }

struct Dims {
  int w, h, pad, tw;
  size_t img_elems;   // padded array elements per pair
  size_t plane;       // plane capacity per pair (elements)
  int y0, y1;         // rows this launch produces (the whole level unless the pair is tiled over several GPUs)
};
__host__ __device__ inline Dims rows(Dims d, int y0, int y1) {   // the same level, rows [y0, y1) clamped to it
  d.y0 = y0 < 0 ? 0 : y0;
  d.y1 = y1 > d.h ? d.h : y1;
  if (d.y1 < d.y0) d.y1 = d.y0;
  return d;
}

// ---------------------------------------------------------------------------------------
template <int C, bool OF>
__global__ void warp_kernel(const float* __restrict__ im_a, const float* __restrict__ im_b,
                            const float* __restrict__ flow, Dims d, float* __restrict__ avg,
                            float* __restrict__ Iz, float* __restrict__ mask) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = d.y0 + blockIdx.y * blockDim.y + threadIdx.y, pair = blockIdx.z;
  if (i >= d.w || j >= d.y1) return;
  constexpr int NOP = OF ? 2 : 1;
  const size_t pix = (size_t)j * d.w + i;
  const float* fl = flow + ((size_t)pair * d.w * d.h + pix) * NOP;
  const float wx = fl[0], wy = OF ? fl[1] : 0.f;
  const float xx = (float)i + wx, yy = (float)j + wy;
  const float fx = floorf(xx), fy = floorf(yy);
  const int x = (int)fx, y = (int)fy;
  const float dx = xx - fx, dy = yy - fy;
  mask[(size_t)pair * d.plane + pix] = (xx >= 0.f && xx <= (float)(d.w - 1) && yy >= 0.f && yy <= (float)(d.h - 1)) ? 1.f : 0.f;
  const int x1 = clampi(x, 0, d.w - 1) + d.pad, x2 = clampi(x + 1, 0, d.w - 1) + d.pad;
  const int y1 = clampi(y, 0, d.h - 1) + d.pad, y2 = clampi(y + 1, 0, d.h - 1) + d.pad;
  const float* A = im_a + (size_t)pair * d.img_elems;
  const float* B = im_b + (size_t)pair * d.img_elems;
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const float s11 = __ldg(B + ((size_t)y1 * d.tw + x1) * C + c), s21 = __ldg(B + ((size_t)y1 * d.tw + x2) * C + c);
    const float s12 = __ldg(B + ((size_t)y2 * d.tw + x1) * C + c), s22 = __ldg(B + ((size_t)y2 * d.tw + x2) * C + c);
    const float w2 = s11 * (1.0f - dx) * (1.0f - dy) + s21 * dx * (1.0f - dy) + s12 * (1.0f - dx) * dy + s22 * dx * dy;
    const float i1 = __ldg(A + ((size_t)(j + d.pad) * d.tw + (i + d.pad)) * C + c);
    const size_t o = ((size_t)pair * C + c) * d.plane + pix;
    avg[o] = 0.5f * (w2 + i1);
    Iz[o] = w2 - i1;
  }
}

__device__ __forceinline__ float conv5h(const float* p, int i, int j, int w) {
  const float* r = p + (size_t)j * w;
  return tap5(r[clampi(i - 2, 0, w - 1)], r[clampi(i - 1, 0, w - 1)], r[i], r[clampi(i + 1, 0, w - 1)], r[clampi(i + 2, 0, w - 1)]);
}
__device__ __forceinline__ float conv5v(const float* p, int i, int j, int w, int h) {
  return tap5(p[(size_t)clampi(j - 2, 0, h - 1) * w + i], p[(size_t)clampi(j - 1, 0, h - 1) * w + i], p[(size_t)j * w + i],
              p[(size_t)clampi(j + 1, 0, h - 1) * w + i], p[(size_t)clampi(j + 2, 0, h - 1) * w + i]);
}

// blockIdx.z enumerates (pair, channel) planes
__global__ void deriv1_kernel(const float* __restrict__ avg, const float* __restrict__ Iz, Dims d,
                              float* __restrict__ Ix, float* __restrict__ Iy, float* __restrict__ Ixz,
                              float* __restrict__ Iyz) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = d.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= d.w || j >= d.y1) return;
  const size_t po = (size_t)blockIdx.z * d.plane, pix = (size_t)j * d.w + i;
  Ix[po + pix] = conv5h(avg + po, i, j, d.w);
  Iy[po + pix] = conv5v(avg + po, i, j, d.w, d.h);
  Ixz[po + pix] = conv5h(Iz + po, i, j, d.w);
  Iyz[po + pix] = conv5v(Iz + po, i, j, d.w, d.h);
}
__global__ void deriv2_kernel(const float* __restrict__ Ix, const float* __restrict__ Iy, Dims d,
                              float* __restrict__ Ixx, float* __restrict__ Ixy, float* __restrict__ Iyy) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = d.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= d.w || j >= d.y1) return;
  const size_t po = (size_t)blockIdx.z * d.plane, pix = (size_t)j * d.w + i;
  Ixx[po + pix] = conv5h(Ix + po, i, j, d.w);
  Ixy[po + pix] = conv5v(Ix + po, i, j, d.w, d.h);
  Iyy[po + pix] = conv5v(Iy + po, i, j, d.w, d.h);
}

// uu = wx + du (DE: clamped), read on the fly
template <bool OF>
__device__ __forceinline__ void load_uv(const float* flow, const float* du, size_t pix, int camlr, float& u, float& v) {
  if (OF) {
    const float2 f = *reinterpret_cast<const float2*>(flow + pix * 2);
    const float2 q = *reinterpret_cast<const float2*>(du + pix * 2);
    u = f.x + q.x;
    v = f.y + q.y;
  } else {
    const float t = flow[pix] + du[pix];
    u = camlr == 0 ? fminf(t, 0.f) : fmaxf(t, 0.f);
    v = 0.f;
  }
}

}  // namespace
}  // namespace disb200

namespace disb200 {
namespace {

struct DerivPtrs {
  const float *Ix, *Iy, *Iz, *Ixx, *Ixy, *Iyy, *Ixz, *Iyz;
};

#define DNORM (0.1f * 0.1f)
#define EPSC (0.001f * 0.001f)
#define EPSG (0.001f * 0.001f)

// The eight derivative values of one pixel and channel, loaded up front so that all loads of a pixel
// are in flight together.
struct Deriv8 {
  float Ix, Iy, Iz, Ixx, Ixy, Iyy, Ixz, Iyz;
};
template <int C>
__device__ __forceinline__ void load_derivs(const DerivPtrs& D, size_t o, size_t cstride, Deriv8 (&dv)[C]) {
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const size_t k = o + (size_t)c * cstride;
    dv[c].Ix = __ldg(D.Ix + k); dv[c].Iy = __ldg(D.Iy + k); dv[c].Iz = __ldg(D.Iz + k);
    dv[c].Ixx = __ldg(D.Ixx + k); dv[c].Ixy = __ldg(D.Ixy + k); dv[c].Iyy = __ldg(D.Iyy + k);
    dv[c].Ixz = __ldg(D.Ixz + k); dv[c].Iyz = __ldg(D.Iyz + k);
  }
}

// compute_data (opticalflow_aux.c:409-594), one pixel.  P(plane, c) reads channel c.
template <int C>
__device__ __forceinline__ void data_term_of(const Deriv8 (&dv)[C], float m, float u, float v,
                                             float hd3, float hg3, float& A11, float& A12, float& A22, float& B1,
                                             float& B2) {
#define P(p, c) dv[c].p
  A11 = 0.f; A12 = 0.f; A22 = 0.f; B1 = 0.f; B2 = 0.f;
  // mask == 0 (pixel warped out of the image): both robust weights are +0 and every term below is an exact
  // +0.  Skipping them is bit-identical and keeps the zero numerators off the IEEE-division slow path, which
  // a whole warp would take for one such lane.
  if (m == 0.f) return;
  float tmp, tmp2, tmp3, tmpx, tmpy, tmpxy, n1, n2, n3 = 0.f, n4 = 0.f, n5 = 0.f, n6 = 0.f;
  if (hd3 != 0.f) {
    tmpx = P(Ix, 0); tmpy = P(Iy, 0);
    tmp2 = mla(mla(P(Iz, 0), tmpx, u), tmpy, v);
    n1 = mla(mla(DNORM, tmpx, tmpx), tmpy, tmpy);
    tmp = (tmp2 * tmp2) / n1;
    if (C == 3) {
      tmpx = P(Ix, 1); tmpy = P(Iy, 1);
      tmp2 = mla(mla(P(Iz, 1), tmpx, u), tmpy, v);
      n2 = mla(mla(DNORM, tmpx, tmpx), tmpy, tmpy);
      tmp = tmp + (tmp2 * tmp2) / n2;
      tmpx = P(Ix, 2); tmpy = P(Iy, 2);
      tmp2 = mla(mla(P(Iz, 2), tmpx, u), tmpy, v);
      n3 = mla(mla(DNORM, tmpx, tmpx), tmpy, tmpy);
      tmp = tmp + (tmp2 * tmp2) / n3;
      tmp = (m * hd3) / sqrtf(EPSC + tmp);
      tmp3 = tmp / n3; tmp2 = tmp3 * tmpx; tmp3 = tmp3 * tmpy;
      A11 = mla(A11, tmp2, tmpx); A12 = mla(A12, tmp2, tmpy); A22 = mla(A22, tmp3, tmpy);
      B1 = mls(B1, tmp2, P(Iz, 2)); B2 = mls(B2, tmp3, P(Iz, 2));
      tmpx = P(Ix, 1); tmpy = P(Iy, 1);
      tmp3 = tmp / n2; tmp2 = tmp3 * tmpx; tmp3 = tmp3 * tmpy;
      A11 = mla(A11, tmp2, tmpx); A12 = mla(A12, tmp2, tmpy); A22 = mla(A22, tmp3, tmpy);
      B1 = mls(B1, tmp2, P(Iz, 1)); B2 = mls(B2, tmp3, P(Iz, 1));
      tmpx = P(Ix, 0); tmpy = P(Iy, 1);   // sic: Iy of channel 2 (opticalflow_aux.c:496)
    } else {
      tmp = (m * hd3) / sqrtf(mla(EPSC, 3.f, tmp));
    }
    tmp3 = tmp / n1; tmp2 = tmp3 * tmpx; tmp3 = tmp3 * tmpy;
    A11 = mla(A11, tmp2, tmpx); A12 = mla(A12, tmp2, tmpy); A22 = mla(A22, tmp3, tmpy);
    B1 = mls(B1, tmp2, P(Iz, 0)); B2 = mls(B2, tmp3, P(Iz, 0));
  }
  tmpx = P(Ixx, 0); tmpy = P(Iyy, 0); tmpxy = P(Ixy, 0);
  tmp2 = mla(mla(P(Ixz, 0), tmpx, u), tmpxy, v);
  tmp3 = mla(mla(P(Iyz, 0), tmpxy, u), tmpy, v);
  tmpxy = tmpxy * tmpxy;
  n1 = mla(tmpxy + DNORM, tmpx, tmpx);
  n2 = mla(tmpxy + DNORM, tmpy, tmpy);
  tmp = (tmp2 * tmp2) / n1 + (tmp3 * tmp3) / n2;
  if (C == 3) {
    tmpx = P(Ixx, 1); tmpy = P(Iyy, 1); tmpxy = P(Ixy, 1);
    tmp2 = mla(mla(P(Ixz, 1), tmpx, u), tmpxy, v);
    tmp3 = mla(mla(P(Iyz, 1), tmpxy, u), tmpy, v);
    tmpxy = tmpxy * tmpxy;
    n3 = mla(tmpxy + DNORM, tmpx, tmpx);
    n4 = mla(tmpxy + DNORM, tmpy, tmpy);
    // sic: the channel sums are assigned, not accumulated (opticalflow_aux.c:529,538)
    tmpx = P(Ixx, 2); tmpy = P(Iyy, 2); tmpxy = P(Ixy, 2);
    tmp2 = mla(mla(P(Ixz, 2), tmpx, u), tmpxy, v);
    tmp3 = mla(mla(P(Iyz, 2), tmpxy, u), tmpy, v);
    tmpxy = tmpxy * tmpxy;
    n5 = mla(tmpxy + DNORM, tmpx, tmpx);
    n6 = mla(tmpxy + DNORM, tmpy, tmpy);
    tmp = (tmp2 * tmp2) / n5 + (tmp3 * tmp3) / n6;
    tmp = (m * hg3) / sqrtf(EPSG + tmp);
    tmp2 = tmp / n6; tmp3 = tmp / n5;
    A11 = mla(mla(A11, tmp3, tmpx * tmpx), tmp2, tmpxy);
    A22 = mla(mla(A22, tmp2, tmpy * tmpy), tmp3, tmpxy);
    tmpxy = P(Ixy, 2);
    A12 = mla(A12, mla(tmp3 * tmpx, tmp2, tmpy), tmpxy);
    B1 = mls(mls(B1, tmp3 * tmpx, P(Ixz, 2)), tmp2 * tmpxy, P(Iyz, 2));
    B2 = mls(mls(B2, tmp2 * tmpy, P(Iyz, 2)), tmp3 * tmpxy, P(Ixz, 2));
    tmp2 = tmp / n4; tmp3 = tmp / n3;
    tmpx = P(Ixx, 1); tmpy = P(Iyy, 1); tmpxy = P(Ixy, 1);
    A12 = mla(A12, mla(tmp3 * tmpx, tmp2, tmpy), tmpxy);
    B1 = mls(mls(B1, tmp3 * tmpx, P(Ixz, 1)), tmp2 * tmpxy, P(Iyz, 1));
    B2 = mls(mls(B2, tmp2 * tmpy, P(Iyz, 1)), tmp3 * tmpxy, P(Ixz, 1));
    tmpxy = tmpxy * tmpxy;
    A11 = mla(mla(A11, tmp3, tmpx * tmpx), tmp2, tmpxy);
    A22 = mla(mla(A22, tmp2, tmpy * tmpy), tmp3, tmpxy);
    tmpx = P(Ixx, 0); tmpy = P(Iyy, 0);
  } else {
    tmp = (m * hg3) / sqrtf(mla(EPSG, 3.f, tmp));
  }
  tmp2 = tmp / n2; tmp3 = tmp / n1;
  tmpxy = P(Ixy, 0);
  A12 = mla(A12, mla(tmp3 * tmpx, tmp2, tmpy), tmpxy);
  B1 = mls(mls(B1, tmp3 * tmpx, P(Ixz, 0)), tmp2 * tmpxy, P(Iyz, 0));
  B2 = mls(mls(B2, tmp2 * tmpy, P(Iyz, 0)), tmp3 * tmpxy, P(Ixz, 0));
  tmpxy = tmpxy * tmpxy;
  A11 = mla(mla(A11, tmp3, tmpx * tmpx), tmp2, tmpxy);
  A22 = mla(mla(A22, tmp2, tmpy * tmpy), tmp3, tmpxy);
  if (C == 1) { A11 *= 3.f; A12 *= 3.f; A22 *= 3.f; B1 *= 3.f; B2 *= 3.f; }
  (void)n4;
#undef P
}

// compute_data_DE (opticalflow_aux.c:602-747), one pixel
template <int C>
__device__ __forceinline__ void data_term_de(const Deriv8 (&dv)[C], float m, float u,
                                             float hd3, float hg3, float& A11, float& B1) {
#define P(p, c) dv[c].p
  A11 = 0.f; B1 = 0.f;
  if (m == 0.f) return;   // see data_term_of
  float tmp, tmp2, tmp3, tmpx, tmpy, tmpxy, n1, n2, n3 = 0.f, n4 = 0.f, n5 = 0.f, n6 = 0.f;
  if (hd3 != 0.f) {
    tmpx = P(Ix, 0); tmpy = P(Iy, 0);
    tmp2 = mla(P(Iz, 0), tmpx, u);
    n1 = mla(mla(DNORM, tmpy, tmpy), tmpx, tmpx);
    tmp = (tmp2 * tmp2) / n1;
    if (C == 3) {
      tmpx = P(Ix, 1); tmpy = P(Iy, 1);
      tmp2 = mla(P(Iz, 1), tmpx, u);
      n2 = mla(mla(DNORM, tmpy, tmpy), tmpx, tmpx);
      tmp = tmp + (tmp2 * tmp2) / n2;
      tmpx = P(Ix, 2); tmpy = P(Iy, 2);
      tmp2 = mla(P(Iz, 2), tmpx, u);
      n3 = mla(mla(DNORM, tmpy, tmpy), tmpx, tmpx);
      tmp = tmp + (tmp2 * tmp2) / n3;
      tmp = (m * hd3) / sqrtf(EPSC + tmp);
      tmp2 = (tmp / n3) * tmpx;
      A11 = mla(A11, tmp2, tmpx); B1 = mls(B1, tmp2, P(Iz, 2));
      tmpx = P(Ix, 1);
      tmp2 = (tmp / n2) * tmpx;
      A11 = mla(A11, tmp2, tmpx); B1 = mls(B1, tmp2, P(Iz, 1));
      tmpx = P(Ix, 0);
    } else {
      tmp = (m * hd3) / sqrtf(mla(EPSC, 3.f, tmp));
    }
    tmp2 = (tmp / n1) * tmpx;
    A11 = mla(A11, tmp2, tmpx); B1 = mls(B1, tmp2, P(Iz, 0));
  }
  tmpx = P(Ixx, 0); tmpy = P(Iyy, 0); tmpxy = P(Ixy, 0);
  tmp2 = mla(P(Iyz, 0), tmpxy, u);
  tmpxy = mla(DNORM, tmpxy, tmpxy);
  n1 = mla(tmpxy, tmpx, tmpx);
  n2 = mla(tmpxy, tmpy, tmpy);
  tmp = (tmp2 * tmp2) / n2;
  tmp2 = mla(P(Ixz, 0), tmpx, u);
  tmp = tmp + (tmp2 * tmp2) / n1;
  if (C == 3) {
    tmpx = P(Ixx, 1); tmpy = P(Iyy, 1); tmpxy = P(Ixy, 1);
    tmp2 = mla(P(Iyz, 1), tmpxy, u);
    tmpxy = mla(DNORM, tmpxy, tmpxy);
    n3 = mla(tmpxy, tmpx, tmpx);
    n4 = mla(tmpxy, tmpy, tmpy);
    tmp = tmp + (tmp2 * tmp2) / n4;
    tmp2 = mla(P(Ixz, 1), tmpx, u);
    tmp = tmp + (tmp2 * tmp2) / n3;
    tmpx = P(Ixx, 2); tmpy = P(Iyy, 2); tmpxy = P(Ixy, 2);
    tmp2 = mla(P(Iyz, 2), tmpxy, u);
    tmpxy = mla(DNORM, tmpxy, tmpxy);
    n5 = mla(tmpxy, tmpx, tmpx);
    n6 = mla(tmpxy, tmpy, tmpy);
    tmp = tmp + (tmp2 * tmp2) / n6;
    tmp2 = mla(P(Ixz, 2), tmpx, u);
    tmp = tmp + (tmp2 * tmp2) / n5;
    tmp = (m * hg3) / sqrtf(EPSG + tmp);
    tmpxy = P(Ixy, 2);
    tmp2 = (tmp / n6) * tmpxy; tmp3 = (tmp / n5) * tmpx;
    A11 = mla(mla(A11, tmp3, tmpx), tmp2, tmpxy);
    B1 = mls(mls(B1, tmp3, P(Ixz, 2)), tmp2, P(Iyz, 2));
    tmpx = P(Ixx, 1); tmpy = P(Iyy, 1); tmpxy = P(Ixy, 1);
    tmp2 = (tmp / n4) * tmpxy; tmp3 = (tmp / n3) * tmpx;
    A11 = mla(mla(A11, tmp3, tmpx), tmp2, tmpxy);
    B1 = mls(mls(B1, tmp3, P(Ixz, 1)), tmp2, P(Iyz, 1));
    tmpx = P(Ixx, 0); tmpy = P(Iyy, 0);
  } else {
    tmp = (m * hg3) / sqrtf(mla(EPSG, 3.f, tmp));
  }
  tmpxy = P(Ixy, 0);
  tmp2 = (tmp / n2) * tmpxy; tmp3 = (tmp / n1) * tmpx;
  A11 = mla(mla(A11, tmp3, tmpx), tmp2, tmpxy);
  B1 = mls(mls(B1, tmp3, P(Ixz, 0)), tmp2, P(Iyz, 0));
  if (C == 1) { A11 *= 3.f; B1 *= 3.f; }
  (void)tmpy;
#undef P
}

// Data term + laplacian of the current flow into the right-hand side + (OF) inverse of the
// 2x2 diagonal block, for one pixel.  `sp` (diffusivity plane) and `q` (flow increment) may
// point to shared or global memory.  Output pack, two float4 per pixel:
//   OF: cA = {h, v, vt, i12}   cB = {i11, i22, b1, b2}
//   DE: cA = {h, v, vt, a11 + sum(psi)}   cB = {b1, 0, 0, 0}
// (hl of pixel i is h of pixel i-1, which the solver already holds.)
// RAW (OF only; the one-pixel-wide levels that sor_coupled hands to sor_coupled_slow_but_readable, solver.c:86-89): the
// data term as compute_data leaves it, cA = {h, v, vt, a12}, cB = {a11, a22, b1, b2}.
template <int C, bool OF, bool RAW = false>
__device__ __forceinline__ void system_pixel(const DerivPtrs& D, const float* __restrict__ mask, const float* sp,
                                             const float* __restrict__ fl, const float* q, int i, int j, int w,
                                             int h, size_t cstride, size_t o, float hd3, float hg3, float4& cA,
                                             float4& cB) {
  constexpr int NOP = OF ? 2 : 1;
  const int pix = j * w + i;
  // Every load of the pixel is issued before the first use (one exposed memory latency instead of one per
  // branch).  Neighbours outside the image are clamped to the pixel itself and get weight 0, which adds the
  // same +0 the reference's skipped branches do not add: x -+ 0 * (w - w) == x.
  const bool hasl = i > 0, hasr = i < w - 1, hast = j > 0, hasb = j < h - 1;
  const int pl = hasl ? pix - 1 : pix, pr = hasr ? pix + 1 : pix, pt = hast ? pix - w : pix, pb = hasb ? pix + w : pix;
  Deriv8 dv[C];
  load_derivs<C>(D, o + pix, cstride, dv);
  const float m = mask[pix];
  const float sc = sp[pix], sl = sp[pl], sr = sp[pr], st = sp[pt], sb = sp[pb];
  float wc[NOP], wl[NOP], wr[NOP], wt[NOP], wb[NOP], qc[NOP];
  if (OF) {
    const float2* f2 = reinterpret_cast<const float2*>(fl);
    const float2 c2 = f2[pix], l2 = f2[pl], r2 = f2[pr], t2 = f2[pt], b2 = f2[pb];
    const float2 q2 = reinterpret_cast<const float2*>(q)[pix];
    wc[0] = c2.x; wc[NOP - 1] = c2.y; wl[0] = l2.x; wl[NOP - 1] = l2.y; wr[0] = r2.x; wr[NOP - 1] = r2.y;
    wt[0] = t2.x; wt[NOP - 1] = t2.y; wb[0] = b2.x; wb[NOP - 1] = b2.y; qc[0] = q2.x; qc[NOP - 1] = q2.y;
  } else {
    wc[0] = fl[pix]; wl[0] = fl[pl]; wr[0] = fl[pr]; wt[0] = fl[pt]; wb[0] = fl[pb]; qc[0] = q[pix];
  }
  // smooth_horiz / smooth_vert (opticalflow_aux.c:161-184): h(i) = s(i)+s(i+1), 0 in the last column
  const float hh = hasr ? sc + sr : 0.f;
  const float hl = hasl ? sl + sc : 0.f;
  const float vv = hasb ? sc + sb : 0.f;
  const float vt = hast ? st + sc : 0.f;
  float b[NOP];
  float A11, A12 = 0.f, A22 = 0.f;
  if (OF) {
    data_term_of<C>(dv, m, qc[0], qc[NOP - 1], hd3, hg3, A11, A12, A22, b[0], b[NOP - 1]);
  } else {
    data_term_de<C>(dv, m, qc[0], hd3, hg3, A11, b[0]);
  }
  // sub_laplacian (opticalflow_aux.c:194-223): b = (((b - t(i-1)) + t(i)) - tv(j-1)) + tv(j)
#pragma unroll
  for (int k = 0; k < NOP; ++k) {
    float bb = b[k];
    bb -= hl * (wc[k] - wl[k]);
    bb += hh * (wr[k] - wc[k]);
    bb -= vt * (wc[k] - wt[k]);
    bb += vv * (wb[k] - wc[k]);
    b[k] = bb;
  }
  if (OF && RAW) {
    cA = make_float4(hh, vv, vt, A12);
    cB = make_float4(A11, A22, b[0], b[NOP - 1]);
  } else if (OF) {
    // solver.c:122-128,184,244
    const float dfirst = hl + (hh + vv), dmid = (hl + hh) + (vt + vv), dlast = hl + (hh + vt);
    const float dpsis = !hast ? dfirst : (hasb ? dmid : dlast);
    const float I11 = A22 + dpsis, I22 = A11 + dpsis;
    const float det = I11 * I22 - A12 * A12;
    cA = make_float4(hh, vv, vt, A12 / (0.f - det));
    cB = make_float4(I11 / det, I22 / det, b[0], b[NOP - 1]);
  } else {
    // solver.c:447-465: sum_dpsis accumulates top, left, bottom, right (absent ones are +0)
    const float sum = (((0.f + vt) + hl) + vv) + hh;
    cA = make_float4(hh, vv, vt, A11 + sum);
    cB = make_float4(b[0], 0.f, 0.f, 0.f);
  }
}

template <int C, bool OF>
__global__ void system_kernel(DerivPtrs D, const float* __restrict__ mask, const float* __restrict__ s,
                              const float* __restrict__ flow, const float* __restrict__ du, Dims d, float hd3,
                              float hg3, float4* __restrict__ cA, float4* __restrict__ cB) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = d.y0 + blockIdx.y * blockDim.y + threadIdx.y, pair = blockIdx.z;
  if (i >= d.w || j >= d.y1) return;
  constexpr int NOP = OF ? 2 : 1;
  const size_t po = (size_t)pair * d.plane;
  float4 a, b;
  system_pixel<C, OF>(D, mask + po, s + po, flow + (size_t)pair * d.w * d.h * NOP, du + po * NOP, i, j, d.w, d.h, d.plane,
                      (size_t)pair * C * d.plane, hd3, hg3, a, b);
  cA[po + (size_t)j * d.w + i] = a;
  cB[po + (size_t)j * d.w + i] = b;
}

// compute_smoothness, first half, for one pixel (uu = wx + du; DE: clamped; first iteration: uu = wx)
template <bool OF>
__device__ __forceinline__ float smooth_pixel(const float* __restrict__ fl, const float* q, int i, int j, int w, int h,
                                              int camlr, float quarter_alpha, bool first_iter) {
  constexpr int NOP = OF ? 2 : 1;
  const int xm = max(i - 1, 0), xp = min(i + 1, w - 1), ym = max(j - 1, 0), yp = min(j + 1, h - 1);
  float ul, vl, uc, vc, ur, vr, ut, vt, ub, vb;
  auto ld = [&](int x, int y, float& u, float& v) {
    const int pix = y * w + x;
    if (OF) {
      u = fl[pix * NOP];
      v = fl[pix * NOP + NOP - 1];
      if (!first_iter) { u = u + q[pix * NOP]; v = v + q[pix * NOP + NOP - 1]; }
    } else {
      u = fl[pix];
      if (!first_iter) { const float t = u + q[pix]; u = camlr == 0 ? fminf(t, 0.f) : fmaxf(t, 0.f); }
      v = 0.f;
    }
  };
  ld(xm, j, ul, vl); ld(i, j, uc, vc); ld(xp, j, ur, vr); ld(i, ym, ut, vt); ld(i, yp, ub, vb);
  const float ux = tap3(ul, uc, ur), vx = tap3(vl, vc, vr), uy = tap3(ut, uc, ub), vy = tap3(vt, vc, vb);
  const float eps = 0.001f * 0.001f;
  const float t = eps + ((ux * ux + uy * uy) + (vx * vx + vy * vy));
  return quarter_alpha / sqrtf(t);
}

// One lexicographic SOR update (solver.c:129-141,191-203,251-262 for OF; solver.c:446-467 for
// DE) from values already in registers: cur = current value of the pixel, r / t / bt = right,
// top, bottom neighbours (zero where there is none), left = value just written at (i-1, j),
// hl = h(i-1, j).  Returns the new value.
template <bool OF>
__device__ __forceinline__ float2 sor_update(const float4 cA, const float4 cB, float hl, bool has_left, int rowtype,
                                             float omega, float2 left, float2 cur, float2 r, float2 t, float2 bt) {
  float2 nw = make_float2(0.f, 0.f);
  if (OF) {
    const float hh = cA.x, vv = cA.y, vt = cA.z, i12 = cA.w, i11 = cB.x, i22 = cB.y, b1 = cB.z, b2 = cB.w;
    float s1, s2;
    if (rowtype == 0) {
      s1 = mla(mla(b1, hh, r.x), vv, bt.x);
      s2 = mla(mla(b2, hh, r.y), vv, bt.y);
    } else if (rowtype == 1) {
      s1 = mla(hh * r.x, vt, t.x) + mla(b1, vv, bt.x);
      s2 = mla(hh * r.y, vt, t.y) + mla(b2, vv, bt.y);
    } else {
      s1 = mla(mla(b1, hh, r.x), vt, t.x);
      s2 = mla(mla(b2, hh, r.y), vt, t.y);
    }
    float B1 = s1, B2 = s2;
    if (has_left) { B1 = hl * left.x + s1; B2 = hl * left.y + s2; }
    nw.x = cur.x + omega * (i11 * B1 + i12 * B2 - cur.x);
    nw.y = cur.y + omega * (i12 * B1 + i22 * B2 - cur.y);
  } else {
    float sigma = 0.f;
    if (rowtype != 0) sigma -= cA.z * t.x;
    if (has_left) sigma -= hl * left.x;
    if (rowtype != 2) sigma -= cA.y * bt.x;
    sigma -= cA.x * r.x;   // h is 0 in the last column, r is 0 there too: adds -0 like the skipped branch adds nothing
    const float B1 = cB.x - sigma;
    nw.x = (1.0f - omega) * cur.x + omega * (B1 / cA.w);
  }
  return nw;
}

template <bool OF>
__device__ __forceinline__ float2 ld_du(const float* du, int pix) {
  if (OF) return reinterpret_cast<const float2*>(du)[pix];
  return make_float2(du[pix], 0.f);
}
template <bool OF>
__device__ __forceinline__ void st_du(float* du, int pix, float2 v) {
  if (OF) reinterpret_cast<float2*>(du)[pix] = v;
  else du[pix] = v.x;
}

// Skewed wavefront over rows [j0, j0 + rows) for `nsweeps` pipelined sweeps (nsweeps > 1 only
// when the band is the whole image).  Pixel (i, j) of sweep s needs (i-1, j) and (i, j-1) of
// sweep s and (i+1, j), (i, j+1) of sweep s-1.  Thread (s, r) owns row r of sweep s and, at
// macro-step m, updates the K columns of chunk  m - r - 2 s  left to right, so every dependency
// (including column i+1 of the next chunk in sweep s-1) was produced at least one barrier
// earlier, and no value is overwritten before its last reader of the previous sweep.  K columns
// per barrier amortise the barrier and let the neighbour loads of a chunk issue together.
// The per-pixel coefficients do not depend on neighbours: they are loaded two chunks ahead
// into a register ring whose slot is m % 2 (static after unrolling), which takes the L2 round
// trip off the critical path.  Plain (coherent) loads, because in the fused kernel the
// coefficients were written by this CTA.  All threads of the CTA must call this.
template <bool OF>
__device__ __forceinline__ void sor_wavefront(float* du, const float4* cA, const float4* cB, int w, int h, int j0,
                                              int rows, int nsweeps, float omega) {
  constexpr int K = 4;
  const int tid = threadIdx.x;
  const int s = tid / rows, r = tid - s * rows, j = j0 + r;
  const bool act = s < nsweeps;
  const int off = r + 2 * s;
  const int nchunks = (w + K - 1) / K;
  const int nsteps = nchunks + rows - 1 + 2 * (nsweeps - 1);
  const int rowtype = j == 0 ? 0 : (j < h - 1 ? 1 : 2);
  const int rowbase = (act ? j : 0) * w;
  const float4* pa = cA + rowbase;
  const float4* pb = cB + rowbase;
  float4 ra[2][K], rb[2][K];
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int c = ((u - off % 2) + 2) % 2;   // the one chunk in {0, 1} that is consumed from slot u
#pragma unroll
    for (int k = 0; k < K; ++k) {
      ra[u][k] = make_float4(0.f, 0.f, 0.f, 0.f);
      rb[u][k] = ra[u][k];
      const int i = c * K + k;
      if (act && i < w) { ra[u][k] = pa[i]; rb[u][k] = pb[i]; }
    }
  }
  const float2 zero2 = make_float2(0.f, 0.f);
  float2 left = zero2;
  float hprev = 0.f;
  for (int m0 = 0; m0 < nsteps; m0 += 2) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int c = m0 + u - off;
      if (act && c >= 0 && c < nchunks) {
        const int i0 = c * K;
        float2 own[K + 1], top[K], bot[K];
#pragma unroll
        for (int k = 0; k <= K; ++k) own[k] = (i0 + k < w) ? ld_du<OF>(du, rowbase + i0 + k) : zero2;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const bool in = i0 + k < w;
          top[k] = (in && rowtype != 0) ? ld_du<OF>(du, rowbase + i0 + k - w) : zero2;
          bot[k] = (in && rowtype != 2) ? ld_du<OF>(du, rowbase + i0 + k + w) : zero2;
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
          if (i0 + k < w) {
            const float2 nw = sor_update<OF>(ra[u][k], rb[u][k], hprev, i0 + k > 0, rowtype, omega, left, own[k],
                                             own[k + 1], top[k], bot[k]);
            st_du<OF>(du, rowbase + i0 + k, nw);
            left = nw;
            hprev = ra[u][k].x;
          }
        }
        // refill the slot only after its last use, so the loads target the ring registers directly
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const int i = i0 + 2 * K + k;
          if (i < w) { ra[u][k] = pa[i]; rb[u][k] = pb[i]; }
        }
      }
      __syncthreads();
    }
  }
}

#include "sor_warp.cuh"
#include "sor_fused.cuh"
#include "sor_band.cuh"

// General path: one CTA per frame pair, increment du in global memory.  When rows x sweeps
// exceed the CTA, sweeps run one after another over row bands (same lexicographic order).
template <bool OF>
__global__ void __launch_bounds__(512)
sor_wavefront_kernel(float* du_all, const float4* __restrict__ cA_all, const float4* __restrict__ cB_all, Dims d,
                     int iters, float omega, int pipelined) {
  constexpr int NOP = OF ? 2 : 1;
  const int pair = blockIdx.x, w = d.w, h = d.h;
  float* du = du_all + (size_t)pair * d.plane * NOP;
  const float4* cA = cA_all + (size_t)pair * d.plane;
  const float4* cB = cB_all + (size_t)pair * d.plane;
  if (pipelined) {
    sor_wavefront<OF>(du, cA, cB, w, h, 0, h, iters, omega);
  } else {
    const int R = blockDim.x;
    for (int it = 0; it < iters; ++it)
      for (int j0 = 0; j0 < h; j0 += R) sor_wavefront<OF>(du, cA, cB, w, h, j0, min(R, h - j0), 1, omega);
  }
}

// Red-black SOR (DISB200_SOR_REDBLACK): one thread per pixel of one colour ((i + j) & 1 == colour), same
// per-pixel formulas as the lexicographic sweep but every neighbour is read at its current value.  Not the
// reference's iteration order: results differ from solver.c (oracle/dis_oracle.cpp restates this order for
// the tests) and are reported with their own EPE, never as the parity result.
template <bool OF>
__global__ void sor_redblack_kernel(float* __restrict__ du_all, const float4* __restrict__ cA_all,
                                    const float4* __restrict__ cB_all, Dims d, float omega, int colour) {
  constexpr int NOP = OF ? 2 : 1;
  const int j = blockIdx.y * blockDim.y + threadIdx.y, pair = blockIdx.z;
  const int i = 2 * (blockIdx.x * blockDim.x + threadIdx.x) + ((j + colour) & 1);
  if (i >= d.w || j >= d.h) return;
  const int w = d.w, h = d.h, pix = j * w + i;
  float* du = du_all + (size_t)pair * d.plane * NOP;
  const float4 cA = cA_all[(size_t)pair * d.plane + pix], cB = cB_all[(size_t)pair * d.plane + pix];
  const float2 z = make_float2(0.f, 0.f);
  const float2 cur = ld_du<OF>(du, pix);
  const float2 left = i > 0 ? ld_du<OF>(du, pix - 1) : z, r = i < w - 1 ? ld_du<OF>(du, pix + 1) : z;
  const float2 t = j > 0 ? ld_du<OF>(du, pix - w) : z, bt = j < h - 1 ? ld_du<OF>(du, pix + w) : z;
  const float hl = i > 0 ? cA_all[(size_t)pair * d.plane + pix - 1].x : 0.f;
  const int rowtype = j == 0 ? 0 : (j < h - 1 ? 1 : 2);
  st_du<OF>(du, pix, sor_update<OF>(cA, cB, hl, i > 0, rowtype, omega, left, cur, r, t, bt));
}

// The same diffusivity for four consecutive pixels of a row per thread (rows whose width is a multiple of 4):
// the centre row is read once for the six horizontal neighbours, vector loads replace ten scalar ones per
// pixel, and the per-pixel arithmetic is literally smooth_pixel's.
template <bool OF>
__global__ void smooth_kernel4(const float* __restrict__ flow, const float* __restrict__ du, Dims d, int camlr,
                               float quarter_alpha, int first_iter, float* __restrict__ s) {
  constexpr int NOP = OF ? 2 : 1;
  const int i0 = 4 * (blockIdx.x * blockDim.x + threadIdx.x), j = d.y0 + blockIdx.y * blockDim.y + threadIdx.y, pair = blockIdx.z;
  if (i0 >= d.w || j >= d.y1) return;
  const float* fl = flow + (size_t)pair * d.w * d.h * NOP;
  const float* q = du + (size_t)pair * d.plane * NOP;
  const bool fi = first_iter != 0;
  // uu / vv of the four pixels starting at (x, y); x is a multiple of 4
  auto ld4 = [&](int x, int y, float (&u)[4], float (&v)[4]) {
    const size_t pix = (size_t)y * d.w + x;
    if (OF) {
      const float4 a = *reinterpret_cast<const float4*>(fl + pix * 2), b = *reinterpret_cast<const float4*>(fl + pix * 2 + 4);
      u[0] = a.x; v[0] = a.y; u[1] = a.z; v[1] = a.w; u[2] = b.x; v[2] = b.y; u[3] = b.z; v[3] = b.w;
      if (!fi) {
        const float4 c = *reinterpret_cast<const float4*>(q + pix * 2), e = *reinterpret_cast<const float4*>(q + pix * 2 + 4);
        u[0] = u[0] + c.x; v[0] = v[0] + c.y; u[1] = u[1] + c.z; v[1] = v[1] + c.w;
        u[2] = u[2] + e.x; v[2] = v[2] + e.y; u[3] = u[3] + e.z; v[3] = v[3] + e.w;
      }
    } else {
      const float4 a = *reinterpret_cast<const float4*>(fl + pix);
      u[0] = a.x; u[1] = a.y; u[2] = a.z; u[3] = a.w;
      if (!fi) {
        const float4 c = *reinterpret_cast<const float4*>(q + pix);
        const float t0 = u[0] + c.x, t1 = u[1] + c.y, t2 = u[2] + c.z, t3 = u[3] + c.w;
        u[0] = camlr == 0 ? fminf(t0, 0.f) : fmaxf(t0, 0.f);
        u[1] = camlr == 0 ? fminf(t1, 0.f) : fmaxf(t1, 0.f);
        u[2] = camlr == 0 ? fminf(t2, 0.f) : fmaxf(t2, 0.f);
        u[3] = camlr == 0 ? fminf(t3, 0.f) : fmaxf(t3, 0.f);
      }
      v[0] = v[1] = v[2] = v[3] = 0.f;
    }
  };
  auto ld1 = [&](int x, int y, float& u, float& v) {
    const int pix = y * d.w + x;
    if (OF) {
      const float2 a = *reinterpret_cast<const float2*>(fl + (size_t)pix * 2);
      u = a.x; v = a.y;
      if (!fi) { const float2 c = *reinterpret_cast<const float2*>(q + (size_t)pix * 2); u = u + c.x; v = v + c.y; }
    } else {
      u = fl[pix];
      if (!fi) { const float t = u + q[pix]; u = camlr == 0 ? fminf(t, 0.f) : fmaxf(t, 0.f); }
      v = 0.f;
    }
  };
  float uc[6], vc[6], ut[4], vt[4], ub[4], vb[4];
  {
    float u4[4], v4[4];
    ld4(i0, j, u4, v4);
#pragma unroll
    for (int k = 0; k < 4; ++k) { uc[k + 1] = u4[k]; vc[k + 1] = v4[k]; }
  }
  if (i0 > 0) ld1(i0 - 1, j, uc[0], vc[0]);
  else { uc[0] = uc[1]; vc[0] = vc[1]; }                    // xm = max(i - 1, 0)
  if (i0 + 4 < d.w) ld1(i0 + 4, j, uc[5], vc[5]);
  else { uc[5] = uc[4]; vc[5] = vc[4]; }                    // xp = min(i + 1, w - 1)
  ld4(i0, max(j - 1, 0), ut, vt);
  ld4(i0, min(j + 1, d.h - 1), ub, vb);
  const float eps = 0.001f * 0.001f;
  float4 out;
  float* o = &out.x;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float ux = tap3(uc[k], uc[k + 1], uc[k + 2]), vx = tap3(vc[k], vc[k + 1], vc[k + 2]);
    const float uy = tap3(ut[k], uc[k + 1], ub[k]), vy = tap3(vt[k], vc[k + 1], vb[k]);
    const float t = eps + ((ux * ux + uy * uy) + (vx * vx + vy * vy));
    o[k] = quarter_alpha / sqrtf(t);
  }
  *reinterpret_cast<float4*>(s + (size_t)pair * d.plane + (size_t)j * d.w + i0) = out;
}

template <bool OF>
__global__ void smooth_kernel2(const float* __restrict__ flow, const float* __restrict__ du, Dims d, int camlr,
                               float quarter_alpha, int first_iter, float* __restrict__ s) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = d.y0 + blockIdx.y * blockDim.y + threadIdx.y, pair = blockIdx.z;
  if (i >= d.w || j >= d.y1) return;
  constexpr int NOP = OF ? 2 : 1;
  s[(size_t)pair * d.plane + (size_t)j * d.w + i] =
      smooth_pixel<OF>(flow + (size_t)pair * d.w * d.h * NOP, du + (size_t)pair * d.plane * NOP, i, j, d.w, d.h, camlr,
                       quarter_alpha, first_iter != 0);
}

template <bool OF>
__global__ void finish_kernel(float* __restrict__ flow, const float* __restrict__ du, Dims d, int camlr) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = d.y0 + blockIdx.y * blockDim.y + threadIdx.y, pair = blockIdx.z;
  if (i >= d.w || j >= d.y1) return;
  constexpr int NOP = OF ? 2 : 1;
  const size_t pix = (size_t)j * d.w + i;
  float u, v;
  load_uv<OF>(flow + (size_t)pair * d.w * d.h * NOP, du + (size_t)pair * d.plane * NOP, pix, camlr, u, v);
  float* o = flow + ((size_t)pair * d.w * d.h + pix) * NOP;
  o[0] = u;
  if (OF) o[1] = v;
}

// Small levels: one CTA per frame pair runs ALL fixed-point iterations of the level —
// diffusivity, linear system, wavefront SOR, flow update — with the flow increment and the
// diffusivity plane resident in shared memory; only the coefficient pack goes through global
// memory (L2).  Replaces 3 launches per fixed-point iteration by one launch per level.
template <int C, bool OF>
__global__ void __launch_bounds__(512)
tv_fused_kernel(DerivPtrs D, const float* __restrict__ mask_all, float* __restrict__ flow_all, Dims d, int camlr,
                float quarter_alpha, float hd3, float hg3, int n_inner, int iters, float omega,
                float4* __restrict__ cA_all, float4* __restrict__ cB_all) {
  constexpr int NOP = OF ? 2 : 1;
  extern __shared__ float4 smem_raw[];
  const int pair = blockIdx.x, w = d.w, h = d.h, npix = w * h;
  float* du = reinterpret_cast<float*>(smem_raw);          // [npix][NOP]
  float* sp = du + (size_t)npix * NOP;                      // [npix]
  float* fl = flow_all + (size_t)pair * npix * NOP;
  const float* mask = mask_all + (size_t)pair * d.plane;
  float4* cA = cA_all + (size_t)pair * d.plane;
  float4* cB = cB_all + (size_t)pair * d.plane;
  const size_t o = (size_t)pair * C * d.plane;
  const int tid = threadIdx.x, T = blockDim.x;
  for (int k = tid; k < npix * NOP; k += T) du[k] = 0.f;
  __syncthreads();
  for (int it = 0; it < n_inner; ++it) {
    for (int pix = tid; pix < npix; pix += T) {
      const int j = pix / w, i = pix - j * w;
      sp[pix] = smooth_pixel<OF>(fl, du, i, j, w, h, camlr, quarter_alpha, it == 0);
    }
    __syncthreads();
    for (int pix = tid; pix < npix; pix += T) {
      const int j = pix / w, i = pix - j * w;
      float4 a, b;
      system_pixel<C, OF>(D, mask, sp, fl, du, i, j, w, h, d.plane, o, hd3, hg3, a, b);
      cA[pix] = a;
      cB[pix] = b;
    }
    __syncthreads();
    sor_wavefront<OF>(du, cA, cB, w, h, 0, h, iters, omega);
  }
  // wx = uu (refine_variational.cpp:226-227, 323)
  for (int pix = tid; pix < npix; pix += T) {
    float u, v;
    load_uv<OF>(fl, du, pix, camlr, u, v);
    fl[pix * NOP] = u;
    if (OF) fl[pix * NOP + 1] = v;
  }
}

// Levels one pixel wide (w < 2, any h >= 2; h < 2 crashes the reference itself and is refused at create): sor_coupled
// falls back to sor_coupled_slow_but_readable (solver.c:34-78,86-89), a point-wise update - du from the OLD dv, then dv
// from the NEW du - instead of the 2x2 block solve.  A handful of pixels per pair: one thread per pair runs the whole
// level sequentially (diffusivity, data term + laplacian, the sweeps), in the reference's order.
template <int C, bool OF>
__global__ void tv_narrow_kernel(DerivPtrs D, const float* __restrict__ mask_all, float* __restrict__ flow_all,
                                 float* __restrict__ du_all, float* __restrict__ s_all, Dims d, int camlr,
                                 float quarter_alpha, float hd3, float hg3, int n_inner, int iters, float omega,
                                 float4* __restrict__ cA_all, float4* __restrict__ cB_all, int n) {
  constexpr int NOP = OF ? 2 : 1;
  const int pair = blockIdx.x * blockDim.x + threadIdx.x;
  if (pair >= n) return;
  const int w = d.w, h = d.h, npix = w * h;
  float* fl = flow_all + (size_t)pair * npix * NOP;
  float* du = du_all + (size_t)pair * d.plane * NOP;
  float* sp = s_all + (size_t)pair * d.plane;
  const float* mask = mask_all + (size_t)pair * d.plane;
  float4* cA = cA_all + (size_t)pair * d.plane;
  float4* cB = cB_all + (size_t)pair * d.plane;
  const size_t o = (size_t)pair * C * d.plane;
  for (int k = 0; k < npix * NOP; ++k) du[k] = 0.f;
  for (int it = 0; it < n_inner; ++it) {
    for (int pix = 0; pix < npix; ++pix) sp[pix] = smooth_pixel<OF>(fl, du, pix % w, pix / w, w, h, camlr, quarter_alpha, it == 0);
    for (int pix = 0; pix < npix; ++pix) {
      float4 a, b;
      system_pixel<C, OF, true>(D, mask, sp, fl, du, pix % w, pix / w, w, h, d.plane, o, hd3, hg3, a, b);
      cA[pix] = a;
      cB[pix] = b;
    }
    for (int sw = 0; sw < iters; ++sw)
      for (int j = 0; j < h; ++j)
        for (int i = 0; i < w; ++i) {
          const int pix = j * w + i;
          const float4 a = cA[pix], b = cB[pix];
          const float hl = i > 0 ? cA[pix - 1].x : 0.f;
          if (OF) {
            float2* q = reinterpret_cast<float2*>(du);
            float su = 0.f, sv = 0.f, sum = 0.f;                     // solver.c:42-63
            if (j > 0) { su -= a.z * q[pix - w].x; sv -= a.z * q[pix - w].y; sum += a.z; }
            if (i > 0) { su -= hl * q[pix - 1].x; sv -= hl * q[pix - 1].y; sum += hl; }
            if (j < h - 1) { su -= a.y * q[pix + w].x; sv -= a.y * q[pix + w].y; sum += a.y; }
            if (i < w - 1) { su -= a.x * q[pix + 1].x; sv -= a.x * q[pix + 1].y; sum += a.x; }
            const float A11 = b.x + sum, A12 = a.w, A22 = b.y + sum;
            const float B1 = b.z - su, B2 = b.w - sv;
            float2 c = q[pix];
            c.x = (1.0f - omega) * c.x + omega / A11 * (B1 - A12 * c.y);   // solver.c:73-74
            c.y = (1.0f - omega) * c.y + omega / A22 * (B2 - A12 * c.x);
            q[pix] = c;
          } else {
            const float2 z = make_float2(0.f, 0.f);
            const float2 left = i > 0 ? make_float2(du[pix - 1], 0.f) : z, r = i < w - 1 ? make_float2(du[pix + 1], 0.f) : z;
            const float2 t = j > 0 ? make_float2(du[pix - w], 0.f) : z, bt = j < h - 1 ? make_float2(du[pix + w], 0.f) : z;
            const int rowtype = j == 0 ? 0 : (j < h - 1 ? 1 : 2);
            du[pix] = sor_update<false>(a, b, hl, i > 0, rowtype, omega, left, make_float2(du[pix], 0.f), r, t, bt).x;
          }
        }
  }
  for (int pix = 0; pix < npix; ++pix) {
    float u, v;
    load_uv<OF>(fl, du, pix, camlr, u, v);
    fl[pix * NOP] = u;
    if (OF) fl[pix * NOP + 1] = v;
  }
}

// One level of refinement for one or two "lanes" (sub-batches of a batch, each with its own stream and planes).
// Two lanes run in ANTI-PHASE: the fixed-point iteration alternates a throughput-bound part (diffusivity + system
// assembly: full-grid streaming kernels) and a latency-bound part (the one-warp-per-pair SOR: six warps per SM waiting on
// their own dependency chains), so lane 1 assembles while lane 0 solves and vice versa, kept in step by two events per
// iteration.  The assembly kernels ask for the SOR's shared-memory carve-out so that both can be resident on one SM.
template <int C, bool OF>
void varref_level_lanes(TvLane* lanes, int nl, const LevelGeom& g, const DisParams& p, int camlr) {
  constexpr int NOP = OF ? 2 : 1;
  const int n_inner = p.tv_innerit * (g.level + 1);                       // refine_variational.cpp:36
  if (n_inner <= 0) return;
  const float qa = 0.25f * p.tv_alpha, hg3 = p.tv_gamma * 0.5f / 3.0f, hd3 = p.tv_delta * 0.5f / 3.0f;
  // 32 x 4 pixel tiles: every warp of a block is a full row segment at all level widths
  const dim3 blk(32, 4);
  const bool pipelined = g.h * p.tv_solverit <= 512;
  const size_t fused_smem = (size_t)g.w * g.h * (NOP + 1) * sizeof(float);
  const int nbands = (g.h + 31) / 32;
  const int sorb_rows = p.tv_solverit + (p.tv_solverit > 1 ? p.tv_solverit - 1 : 1);   // exchange rows per band
  const size_t sorw_smem = sorw_smem_bytes(g.h, p.tv_solverit);
  int threads = pipelined ? g.h * p.tv_solverit : (g.h < 512 ? g.h : 512);
  threads = ((threads + 31) / 32) * 32;

#define DISB200_TEAM_SELECT (sched == 0 || sched == 5)   // DisTuning::tv_schedule = 2 keeps one warp per pair (A/B)
  struct Plan {
    Dims d;
    dim3 grd, grdc, grd4;
    DerivPtrs D;
    float4 *cA, *cB;
    bool done = false, smooth4 = false;
    void (*sorw)(float*, const float4*, const float4*, Dims, float) = nullptr;
    void (*sort)(float*, const float4*, const float4*, Dims, float, unsigned) = nullptr;   // team schedule: one warp per band, one CTA per pair
    int sort_threads = 0;
    void (*sorb)(float*, const float4*, const float4*, float4*, float4*, float4*, float4*, unsigned*, unsigned, unsigned, Dims, int, int, int, float) = nullptr;
    size_t sorb_smem = 0;
    // system assembly inside the SOR kernel (sor_fused.cuh): replaces system_kernel + sorw
    void (*solvef)(DerivPtrs, const float*, const float*, const float*, float*, Dims, float, float, float) = nullptr;
    int solvef_threads = 0;
  } pl[2];

  // ---- per level: warp, derivatives; levels that one kernel handles whole (narrow, zero sweeps, fused) end here ----
  for (int l = 0; l < nl; ++l) {
    TvLane& L = lanes[l];
    Plan& P = pl[l];
    const VarRefBuffers& vb = L.vb;
    const int n = L.n;
    cudaStream_t st = L.st;
    LaunchStats* ls = L.ls;
    P.d = Dims{g.w, g.h, g.pad, g.tw, g.img_elems, vb.plane_cap, 0, g.h};
    P.grd = dim3((g.w + 31) / 32, (g.h + 3) / 4, n);
    P.grdc = dim3((g.w + 31) / 32, (g.h + 3) / 4, n * C);
    P.grd4 = dim3((g.w / 4 + 31) / 32, (g.h + 3) / 4, n);
    warp_kernel<C, OF><<<P.grd, blk, 0, st>>>(L.im_a, L.im_b, L.flow, P.d, vb.avg, vb.Iz, vb.mask);
    deriv1_kernel<<<P.grdc, blk, 0, st>>>(vb.avg, vb.Iz, P.d, vb.Ix, vb.Iy, vb.Ixz, vb.Iyz);
    deriv2_kernel<<<P.grdc, blk, 0, st>>>(vb.Ix, vb.Iy, P.d, vb.Ixx, vb.Ixy, vb.Iyy);
    ls->launches += 3;
    P.D = DerivPtrs{vb.Ix, vb.Iy, vb.Iz, vb.Ixx, vb.Ixy, vb.Iyy, vb.Ixz, vb.Iyz};
    P.cA = reinterpret_cast<float4*>(vb.coef);
    P.cB = reinterpret_cast<float4*>(vb.inv);
    if (g.w < 2) {   // one-pixel-wide level: the reference's point-wise fallback (solver.c:86-89)
      tv_narrow_kernel<C, OF><<<(n + 31) / 32, 32, 0, st>>>(P.D, vb.mask, L.flow, vb.du, vb.smooth, P.d, camlr, qa, hd3, hg3, n_inner,
                                                            p.tv_solverit, p.tv_sor, P.cA, P.cB, n);
      ls->launches++;
      P.done = true;
      continue;
    }
    if (p.tv_solverit < 1) {
      // sor_coupled / sor_coupled_slow_but_readable_DE with zero sweeps leave the increment at 0: the level's result is
      // uu = wx + 0, which for depth still applies the sign clamp (refine_variational.cpp:301-323)
      cudaMemsetAsync(vb.du, 0, sizeof(float) * vb.plane_cap * NOP * n, st);
      finish_kernel<OF><<<P.grd, blk, 0, st>>>(L.flow, vb.du, P.d, camlr);
      ls->launches++;
      P.done = true;
      continue;
    }
    const int sched = vb.tv_schedule;   // 0 auto, 1 fused, 2 split, 3 split + general wavefront SOR (DisTuning)
    // "fused": one CTA per pair runs the whole level; "split": diffusivity / system / SOR kernels per fixed-point
    // iteration, whose one-warp-per-band SOR keeps thousands of pairs resident.  Measured with the team SOR
    // (profiles/r02u_fused_vs_team.txt, batches 1 ... 296): optical flow is faster split at EVERY batch size (one C1 pair
    // 0.83 -> 0.71 ms, 296 pairs 168 k -> 184 k pairs/s), depth - a division per update in the chain - stays faster fused
    // while every pair of the batch is resident (148 C3 pairs: 130 k fused, 100 k split).  DisTuning::tv_schedule overrides.
    const bool sorw_level = g.h <= 96 && p.tv_solverit <= 5;   // the one-warp / team SOR takes this level
    const bool auto_split = (OF && sorw_level) || n > 2 * 148;
    const bool want_split = p.sor_mode == DISB200_SOR_REDBLACK || (sched ? sched != 1 : auto_split);
    if (!want_split && pipelined && fused_smem <= 100 * 1024) {
      int fthreads = ((g.h * p.tv_solverit + 31) / 32) * 32;
      if (fthreads < 256) fthreads = 256;
      cudaFuncSetAttribute(tv_fused_kernel<C, OF>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
      if (ls->mark) ls->mark(ls->ctx, 5, 1);
      tv_fused_kernel<C, OF><<<n, fthreads, fused_smem, st>>>(P.D, vb.mask, L.flow, P.d, camlr, qa, hd3, hg3, n_inner,
                                                              p.tv_solverit, p.tv_sor, P.cA, P.cB);
      if (ls->mark) ls->mark(ls->ctx, 5, 0);
      ls->launches++;
      P.done = true;
      continue;
    }
    cudaMemsetAsync(vb.du, 0, sizeof(float) * vb.plane_cap * NOP * n, st);
    // SOR launcher: one warp per pair (levels of up to 96 rows, up to 5 sweeps), else the general wavefront
    if (g.h <= 96 && sched != 3) {
      const int nb = (g.h + 31) / 32;
      switch (p.tv_solverit) {
#define DISB200_SORW_CASE(NS_) case NS_: P.sorw = nb == 3 ? sor_warp_kernel<OF, NS_, 3> : (nb == 2 ? sor_warp_kernel<OF, NS_, 2> : sor_warp_kernel<OF, NS_, 1>); break;
        DISB200_SORW_CASE(1) DISB200_SORW_CASE(2) DISB200_SORW_CASE(3) DISB200_SORW_CASE(4) DISB200_SORW_CASE(5)
#undef DISB200_SORW_CASE
        default: break;
      }
    }
    // tall levels: one warp per band of 32 rows, bands chained through tagged boundary rows (sor_band.cuh)
    // (latency mode: measured 1.35-1.5x over the general wavefront while all bands of the batch are resident, 2 per
    //  SM — one 4K pair: 27.6 -> 19.8 ms of SOR; with more bands than that the general wavefront, one CTA per
    //  pair, has the better throughput: 64 4K pairs 33 vs 45 ms)
    // vb.sorb_max = co-resident CTAs of the band kernel on this device (occupancy API, engine.cu) or less: bands spin on
    // their neighbours, so every band of the launch must be resident - never more CTAs than that
    if (!P.sorw && vb.sorb_ctl && g.h > 96 && g.w >= 8 && nbands <= 64 && p.tv_solverit <= 3 && n * nbands <= vb.sorb_max &&
        (size_t)nbands * g.w * sorb_rows * 4 <= vb.plane_cap * NOP && sched != 3) {
      switch (p.tv_solverit) {
        case 1: P.sorb = sor_band_kernel<OF, 1>; P.sorb_smem = SorbCfg<1>::BYTES; break;
        case 2: P.sorb = sor_band_kernel<OF, 2>; P.sorb_smem = SorbCfg<2>::BYTES; break;
        default: P.sorb = sor_band_kernel<OF, 3>; P.sorb_smem = SorbCfg<3>::BYTES; break;
      }
      cudaFuncSetAttribute(P.sorb, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P.sorb_smem);
    }
    if (P.sorw) cudaFuncSetAttribute(P.sorw, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sorw_smem);
    // levels of two or three bands: the same schedule with one WARP per band (sor_team_kernel) - same bits, NS updates per
    // step and warp instead of NB x NS, NB times the warps per SM for the same shared memory
    if (P.sorw && g.h > 32 && p.tv_solverit <= 3 && DISB200_TEAM_SELECT) {
      const int nb = (g.h + 31) / 32;
      switch (p.tv_solverit) {
#define DISB200_SORT_CASE(NS_) case NS_: P.sort = nb == 3 ? sor_team_kernel<OF, NS_, 3> : sor_team_kernel<OF, NS_, 2>; break;
        DISB200_SORT_CASE(1) DISB200_SORT_CASE(2) DISB200_SORT_CASE(3)
#undef DISB200_SORT_CASE
        default: break;
      }
      P.sort_threads = 32 * nb;
      if (P.sort) cudaFuncSetAttribute(P.sort, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sorw_smem + sorw_team_bytes(p.tv_solverit)));
    }
    // DisTuning::tv_schedule = 4: the one-warp SOR with the system assembled by producer warps of its CTA (sor_fused.cuh;
    // three sweeps).  Bit-identical, removes 64 B per pixel and iteration of HBM traffic - and is 2.9x SLOWER than
    // system_kernel + the stand-alone SOR (C1, 1776 pairs: 8.57 ms against 1.41 + 1.59; profiles/r02f_solvef.txt): one
    // producer warp per band cannot hide the latency of 20 dependent loads and 13 divisions per pixel inside the five
    // steps of slack a 16-pixel ring gives it, and neither registers (128 per thread, five CTAs per SM) nor shared
    // memory (35 KB of rings per pair) leave room for the dozens of warps the stand-alone kernel hides it with.
    // Kept as a tested schedule for that evidence; never chosen automatically.
    if (P.sorw && p.tv_solverit == 3 && sched == 4) {
      const int nb = (g.h + 31) / 32;
      P.solvef = nb == 3 ? tv_solve_fused_kernel<C, OF, 3, 3> : (nb == 2 ? tv_solve_fused_kernel<C, OF, 3, 2> : tv_solve_fused_kernel<C, OF, 3, 1>);
      P.solvef_threads = 32 * (1 + nb);
      cudaFuncSetAttribute(P.solvef, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sorf_smem_bytes(g.h, 3));
    }
    // four pixels per thread need 16-byte aligned rows in every plane: w % 4 == 0 (plane_cap is a multiple of 4)
    P.smooth4 = (g.w & 3) == 0 && (vb.plane_cap & 3) == 0;
  }
  const bool antiphase = nl == 2 && !pl[0].done && !pl[1].done && pl[0].sorw && pl[1].sorw;
  if (antiphase) {   // assembly next to the SOR warps on one SM: same carve-out for both
    cudaFuncSetAttribute(system_kernel<C, OF>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(smooth_kernel4<OF>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(smooth_kernel2<OF>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  }
  auto assemble = [&](int l, int it) {
    TvLane& L = lanes[l];
    Plan& P = pl[l];
    LaunchStats* ls = L.ls;
    // (a fused diffusivity + system kernel over shared-memory tiles was measured slower: 2.00 vs 1.88 ms per
    //  1480-pair step; the halo re-evaluation and the barrier cost more than the 8 B / pixel they save)
    if (ls->mark) ls->mark(ls->ctx, 7, 1);
    if (P.smooth4) smooth_kernel4<OF><<<P.grd4, blk, 0, L.st>>>(L.flow, L.vb.du, P.d, camlr, qa, it == 0, L.vb.smooth);
    else smooth_kernel2<OF><<<P.grd, blk, 0, L.st>>>(L.flow, L.vb.du, P.d, camlr, qa, it == 0, L.vb.smooth);
    if (ls->mark) ls->mark(ls->ctx, 7, 0);
    ls->launches++;
    if (P.solvef) return;   // the system is assembled inside the solve kernel
    if (ls->mark) ls->mark(ls->ctx, 6, 1);
    system_kernel<C, OF><<<P.grd, blk, 0, L.st>>>(P.D, L.vb.mask, L.vb.smooth, L.flow, L.vb.du, P.d, hd3, hg3, P.cA, P.cB);
    if (ls->mark) ls->mark(ls->ctx, 6, 0);
    ls->launches++;
  };
  auto solve = [&](int l) {
    TvLane& L = lanes[l];
    Plan& P = pl[l];
    LaunchStats* ls = L.ls;
    const VarRefBuffers& vb = L.vb;
    const int n = L.n;
    if (ls->mark) ls->mark(ls->ctx, 5, 1);
    if (p.sor_mode == DISB200_SOR_REDBLACK) {
      dim3 rblk(32, 4), rgrd((g.w / 2 + 1 + 31) / 32, (g.h + 3) / 4, n);
      for (int sw = 0; sw < 2 * p.tv_solverit; ++sw)
        sor_redblack_kernel<OF><<<rgrd, rblk, 0, L.st>>>(vb.du, P.cA, P.cB, P.d, p.tv_sor, sw & 1);
      ls->launches += 2 * p.tv_solverit - 1;
    } else if (P.solvef) {
      P.solvef<<<n, P.solvef_threads, sorf_smem_bytes(g.h, 3), L.st>>>(P.D, vb.mask, vb.smooth, L.flow, vb.du, P.d, hd3, hg3, p.tv_sor);
    } else if (P.sort) P.sort<<<n, P.sort_threads, sorw_smem + sorw_team_bytes(p.tv_solverit), L.st>>>(vb.du, P.cA, P.cB, P.d, p.tv_sor, (unsigned)sorw_smem);
    else if (P.sorw) P.sorw<<<n, 32, sorw_smem, L.st>>>(vb.du, P.cA, P.cB, P.d, p.tv_sor);
    else if (P.sorb) {
      float4* sorb_xt = reinterpret_cast<float4*>(vb.rhs);
      float4* sorb_xb = sorb_xt + (size_t)n * nbands * p.tv_solverit * g.w;
      const unsigned epoch = ++vb.sorb_host[0], base = vb.sorb_host[1];
      vb.sorb_host[1] += (unsigned)(n * nbands);
      P.sorb<<<n * nbands, 32, P.sorb_smem, L.st>>>(vb.du, P.cA, P.cB, sorb_xt, sorb_xb, sorb_xt, sorb_xb, vb.sorb_ctl, base, epoch, P.d,
                                                    nbands, 0, nbands, p.tv_sor);
    }
    else sor_wavefront_kernel<OF><<<n, threads, 0, L.st>>>(vb.du, P.cA, P.cB, P.d, p.tv_solverit, p.tv_sor, pipelined ? 1 : 0);
    if (ls->mark) ls->mark(ls->ctx, 5, 0);
    ls->launches++;
  };
  for (int it = 0; it < n_inner; ++it) {
    for (int l = 0; l < nl; ++l) {
      if (pl[l].done) continue;
      if (antiphase) {
        // lane 0 assembles iteration `it` once lane 1 has assembled it-1 (and therefore solves it-1 meanwhile);
        // lane 1 assembles `it` once lane 0 has (and solves `it` meanwhile)
        if (l == 0 && it > 0) cudaStreamWaitEvent(lanes[0].st, lanes[1].ev, 0);
        if (l == 1) cudaStreamWaitEvent(lanes[1].st, lanes[0].ev, 0);
      }
      assemble(l, it);
      if (antiphase) cudaEventRecord(lanes[l].ev, lanes[l].st);
      solve(l);
    }
  }
  for (int l = 0; l < nl; ++l) {
    if (pl[l].done) continue;
    finish_kernel<OF><<<pl[l].grd, blk, 0, lanes[l].st>>>(lanes[l].flow, lanes[l].vb.du, pl[l].d, camlr);
    lanes[l].ls->launches++;
  }
}

// One level, ONE pair, rows split over ndev devices (see TileDev in dis_internal.h).  The same kernels as the one-device
// path, each on its row range; per fixed-point iteration the only traffic between neighbours is two rows of the
// increment (after the SOR) and, inside the SOR, the tagged boundary rows of the bands on either side of a cut.
template <int C, bool OF>
int varref_level_tiled_impl(TileDev* dv, int ndev, const DisParams& p, int camlr) {
  constexpr int NOP = OF ? 2 : 1;
  const LevelGeom& g0 = dv[0].g;
  const int n_inner = p.tv_innerit * (g0.level + 1);                      // refine_variational.cpp:36
  if (n_inner <= 0) return 0;
  const float qa = 0.25f * p.tv_alpha, hg3 = p.tv_gamma * 0.5f / 3.0f, hd3 = p.tv_delta * 0.5f / 3.0f;
  const int w = g0.w, h = g0.h, nbands = (h + 31) / 32, NS = p.tv_solverit;
  const dim3 blk(32, 4);
  auto grid = [&](const Dims& d, int cols, int planes) { return dim3((cols + 31) / 32, (d.y1 - d.y0 + 3) / 4, planes); };
#define TILE_CK(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return (int)e__; } while (0)
  void (*sorb)(float*, const float4*, const float4*, float4*, float4*, float4*, float4*, unsigned*, unsigned, unsigned, Dims, int, int, int, float) =
      NS == 1 ? sor_band_kernel<OF, 1> : (NS == 2 ? sor_band_kernel<OF, 2> : sor_band_kernel<OF, 3>);
  const size_t sorb_smem = NS == 1 ? SorbCfg<1>::BYTES : (NS == 2 ? SorbCfg<2>::BYTES : SorbCfg<3>::BYTES);
  // ---- per level: warp, derivatives (halo rows recomputed locally), increment = 0 ----
  for (int k = 0; k < ndev; ++k) {
    TileDev& t = dv[k];
    TILE_CK(cudaSetDevice(t.device));
    const Dims d{w, h, t.g.pad, t.g.tw, t.g.img_elems, t.vb.plane_cap, 0, h};
    const Dims dw = rows(d, t.g.row0 - 4, t.g.row1 + 4), d1 = rows(d, t.g.row0 - 2, t.g.row1 + 2), d2 = rows(d, t.g.row0, t.g.row1);
    warp_kernel<C, OF><<<grid(dw, w, 1), blk, 0, t.stream>>>(t.im_a, t.im_b, t.flow, dw, t.vb.avg, t.vb.Iz, t.vb.mask);
    deriv1_kernel<<<grid(d1, w, C), blk, 0, t.stream>>>(t.vb.avg, t.vb.Iz, d1, t.vb.Ix, t.vb.Iy, t.vb.Ixz, t.vb.Iyz);
    deriv2_kernel<<<grid(d2, w, C), blk, 0, t.stream>>>(t.vb.Ix, t.vb.Iy, d2, t.vb.Ixx, t.vb.Ixy, t.vb.Iyy);
    TILE_CK(cudaMemsetAsync(t.vb.du, 0, sizeof(float) * (size_t)w * h * NOP, t.stream));
    TILE_CK(cudaFuncSetAttribute(sorb, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sorb_smem));
    t.st->launches += 3;
  }
  const bool smooth4 = (w & 3) == 0 && (dv[0].vb.plane_cap & 3) == 0;
  const size_t rowb = (size_t)w * NOP * sizeof(float);
  for (int it = 0; it < n_inner; ++it) {
    // ---- halo rows of the increment from the neighbours' last SOR, then diffusivity and linear system ----
    for (int k = 0; k < ndev; ++k) {
      TileDev& t = dv[k];
      TILE_CK(cudaSetDevice(t.device));
      if (it > 0) {
        if (k + 1 < ndev) {   // rows row1, row1+1 belong to the next device
          TILE_CK(cudaStreamWaitEvent(t.stream, dv[k + 1].ev_sor, 0));
          const int r0 = t.g.row1, r1 = std::min(h, t.g.row1 + 2);
          if (r1 > r0) TILE_CK(cudaMemcpyAsync(t.vb.du + (size_t)r0 * w * NOP, dv[k + 1].vb.du + (size_t)r0 * w * NOP, rowb * (r1 - r0), cudaMemcpyDefault, t.stream));
        }
        if (k > 0) {          // rows row0-2, row0-1 belong to the previous device
          TILE_CK(cudaStreamWaitEvent(t.stream, dv[k - 1].ev_sor, 0));
          const int r0 = std::max(0, t.g.row0 - 2), r1 = t.g.row0;
          if (r1 > r0) TILE_CK(cudaMemcpyAsync(t.vb.du + (size_t)r0 * w * NOP, dv[k - 1].vb.du + (size_t)r0 * w * NOP, rowb * (r1 - r0), cudaMemcpyDefault, t.stream));
        }
        TILE_CK(cudaEventRecord(t.ev_halo, t.stream));
      }
      const Dims d{w, h, t.g.pad, t.g.tw, t.g.img_elems, t.vb.plane_cap, 0, h};
      const Dims ds = rows(d, t.g.row0 - 1, t.g.row1 + 1), dy = rows(d, t.g.row0, t.g.row1);
      if (smooth4) smooth_kernel4<OF><<<dim3((w / 4 + 31) / 32, (ds.y1 - ds.y0 + 3) / 4, 1), blk, 0, t.stream>>>(t.flow, t.vb.du, ds, camlr, qa, it == 0, t.vb.smooth);
      else smooth_kernel2<OF><<<grid(ds, w, 1), blk, 0, t.stream>>>(t.flow, t.vb.du, ds, camlr, qa, it == 0, t.vb.smooth);
      DerivPtrs D{t.vb.Ix, t.vb.Iy, t.vb.Iz, t.vb.Ixx, t.vb.Ixy, t.vb.Iyy, t.vb.Ixz, t.vb.Iyz};
      system_kernel<C, OF><<<grid(dy, w, 1), blk, 0, t.stream>>>(D, t.vb.mask, t.vb.smooth, t.flow, t.vb.du, dy, hd3, hg3,
                                                                 reinterpret_cast<float4*>(t.vb.coef), reinterpret_cast<float4*>(t.vb.inv));
      t.st->launches += 2;
    }
    // ---- the SOR: every device runs its bands; a cut between two devices is a band boundary like any other, only that
    //      the tagged rows crossing it are written into the neighbour's memory ----
    // one launch epoch for all devices, ahead of every epoch any of their exchange arrays has seen
    unsigned epoch = 0;
    for (int k = 0; k < ndev; ++k) epoch = std::max(epoch, dv[k].vb.sorb_host[0]);
    ++epoch;
    for (int k = 0; k < ndev; ++k) dv[k].vb.sorb_host[0] = epoch;
    for (int k = 0; k < ndev; ++k) {
      TileDev& t = dv[k];
      TILE_CK(cudaSetDevice(t.device));
      if (it > 0) {   // the neighbours have copied their halo rows out of this device's increment before it is overwritten
        if (k + 1 < ndev) TILE_CK(cudaStreamWaitEvent(t.stream, dv[k + 1].ev_halo, 0));
        if (k > 0) TILE_CK(cudaStreamWaitEvent(t.stream, dv[k - 1].ev_halo, 0));
      }
      const Dims d{w, h, t.g.pad, t.g.tw, t.g.img_elems, t.vb.plane_cap, 0, h};
      const int band0 = t.g.row0 / 32, nbl = (t.g.row1 - t.g.row0 + 31) / 32;
      float4* xt = reinterpret_cast<float4*>(t.vb.rhs);
      float4* xb = xt + (size_t)nbands * NS * w;
      float4* xt_next = k + 1 < ndev ? reinterpret_cast<float4*>(dv[k + 1].vb.rhs) : xt;
      float4* xb_prev = k > 0 ? reinterpret_cast<float4*>(dv[k - 1].vb.rhs) + (size_t)nbands * NS * w : xb;
      const unsigned base = t.vb.sorb_host[1];
      t.vb.sorb_host[1] += (unsigned)nbl;
      if (nbl > 0)
        sorb<<<nbl, 32, sorb_smem, t.stream>>>(t.vb.du, reinterpret_cast<const float4*>(t.vb.coef), reinterpret_cast<const float4*>(t.vb.inv), xt, xb,
                                               xt_next, xb_prev, t.vb.sorb_ctl, base, epoch, d, nbands, band0, nbl, p.tv_sor);
      TILE_CK(cudaEventRecord(t.ev_sor, t.stream));
      t.st->launches++;
    }
  }
  for (int k = 0; k < ndev; ++k) {
    TileDev& t = dv[k];
    TILE_CK(cudaSetDevice(t.device));
    const Dims d{w, h, t.g.pad, t.g.tw, t.g.img_elems, t.vb.plane_cap, 0, h};
    const Dims dy = rows(d, t.g.row0, t.g.row1);
    finish_kernel<OF><<<grid(dy, w, 1), blk, 0, t.stream>>>(t.flow, t.vb.du, dy, camlr);
    t.st->launches++;
    TILE_CK(cudaGetLastError());
  }
#undef TILE_CK
  return 0;
}

}  // namespace

bool varref_level_tileable(const LevelGeom& g, const DisParams& p, int ndev, size_t plane_cap, int nop) {
  const int nbands = (g.h + 31) / 32;
  const int sorb_rows = p.tv_solverit + (p.tv_solverit > 1 ? p.tv_solverit - 1 : 1);
  return ndev > 1 && p.sor_mode == DISB200_SOR_LEXICOGRAPHIC && g.h > 96 && g.w >= 8 && nbands <= 64 && nbands >= ndev &&
         p.tv_solverit >= 1 && p.tv_solverit <= 3 && (size_t)nbands * g.w * sorb_rows * 4 <= plane_cap * nop;
}

int launch_varref_level_tiled(TileDev* dv, int ndev, const DisParams& p, int camlr) {
  const bool of = p.mode == DISB200_MODE_OF;
  if (p.channels == 1) return of ? varref_level_tiled_impl<1, true>(dv, ndev, p, camlr) : varref_level_tiled_impl<1, false>(dv, ndev, p, camlr);
  return of ? varref_level_tiled_impl<3, true>(dv, ndev, p, camlr) : varref_level_tiled_impl<3, false>(dv, ndev, p, camlr);
}

int sor_band_resident_ctas(bool of, int solverit) {
  int dev = 0, sms = 0, per_sm = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  const void* fn = nullptr;
  size_t smem = 0;
  switch (solverit) {
    case 1: fn = of ? (const void*)sor_band_kernel<true, 1> : (const void*)sor_band_kernel<false, 1>; smem = SorbCfg<1>::BYTES; break;
    case 2: fn = of ? (const void*)sor_band_kernel<true, 2> : (const void*)sor_band_kernel<false, 2>; smem = SorbCfg<2>::BYTES; break;
    case 3: fn = of ? (const void*)sor_band_kernel<true, 3> : (const void*)sor_band_kernel<false, 3>; smem = SorbCfg<3>::BYTES; break;
    default: return 0;
  }
  if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { cudaGetLastError(); return 0; }
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 32, smem) != cudaSuccess) { cudaGetLastError(); return 0; }
  return per_sm * sms;
}

size_t varref_plane_count(int C, int nop) {
  // avg, Iz, Ix, Iy, Ixx, Ixy, Iyy, Ixz, Iyz (C each) + mask + s + du (nop) + coefficient pack (2 x float4) + spare (nop)
  return (size_t)9 * C + 2 + nop + 4 + 4 + nop;
}

void launch_varref_level_lanes(TvLane* lanes, int nl, const LevelGeom& g, const DisParams& p, int camlr) {
  const bool of = p.mode == DISB200_MODE_OF;
  if (p.channels == 1) {
    if (of) varref_level_lanes<1, true>(lanes, nl, g, p, camlr);
    else varref_level_lanes<1, false>(lanes, nl, g, p, camlr);
  } else {
    if (of) varref_level_lanes<3, true>(lanes, nl, g, p, camlr);
    else varref_level_lanes<3, false>(lanes, nl, g, p, camlr);
  }
}

void launch_varref_level(const float* im_a, const float* im_b, float* flow, const LevelGeom& g, const DisParams& p,
                         int camlr, const VarRefBuffers& vb, int n, cudaStream_t s, LaunchStats* st) {
  TvLane lane{im_a, im_b, flow, vb, n, s, st, nullptr};
  launch_varref_level_lanes(&lane, 1, g, p, camlr);
}

}  // namespace disb200
