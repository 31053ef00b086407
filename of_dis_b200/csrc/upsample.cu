// Stage 5 — final upsample and crop (run_dense.cpp:407-415):
//   flowout *= 2^sc_l ; cv::resize(INTER_LINEAR, x 2^sc_l) ; crop the divisibility padding.
// cv::resize semantics (pinned against python-cv2 4.13, tests/golden): half-pixel centres,
// source index clamped with the fraction forced to 0 at the borders, horizontal pass then
// vertical pass.  This is the HBM-heavy kernel of the path (80 % of the compulsory bytes):
// each thread produces two horizontally adjacent output pixels and issues 128-bit (OF) or
// 64-bit (DE) stores; the low-resolution source stays in L1/L2.
#include "dis_internal.h"

namespace disb200 {

namespace {

__device__ __forceinline__ void src_coord(int d, int pad, double inv, int n, int& s0, int& s1, float& f) {
  f = (float)(((double)(d + pad) + 0.5) * inv - 0.5);
  const float fl = floorf(f);
  s0 = (int)fl;
  f -= fl;
  if (s0 < 0) { s0 = 0; f = 0.f; }
  if (s0 >= n - 1) { s0 = n - 1; f = 0.f; }
  s1 = min(s0 + 1, n - 1);
}

template <int NOP>
__global__ void upsample_crop_kernel(const float* __restrict__ flow, int wl, int hl, int sc_l, int padl, int padt,
                                     int w0, int h0, float* __restrict__ out) {
  const int xp = (blockIdx.x * blockDim.x + threadIdx.x) * 2, y = blockIdx.y, pair = blockIdx.z;
  if (xp >= w0) return;
  const float* src = flow + (size_t)pair * wl * hl * NOP;
  float* dst = out + ((size_t)pair * w0 * h0 + (size_t)y * w0 + xp) * NOP;
  float res[2][NOP];
  if (sc_l == 0) {
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
      for (int c = 0; c < NOP; ++c)
        res[e][c] = (xp + e < w0) ? __ldg(src + ((size_t)(y + padt) * wl + (xp + e + padl)) * NOP + c) : 0.f;
  } else {
    const double inv = 1.0 / (double)(1 << sc_l);
    const float sc = (float)(1 << sc_l);
    int sy0, sy1;
    float fy;
    src_coord(y, padt, inv, hl, sy0, sy1, fy);
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      int sx0, sx1;
      float fx;
      src_coord(min(xp + e, w0 - 1), padl, inv, wl, sx0, sx1, fx);
#pragma unroll
      for (int c = 0; c < NOP; ++c) {
        const float a = __ldg(src + ((size_t)sy0 * wl + sx0) * NOP + c) * sc, b = __ldg(src + ((size_t)sy0 * wl + sx1) * NOP + c) * sc;
        const float cc = __ldg(src + ((size_t)sy1 * wl + sx0) * NOP + c) * sc, dd = __ldg(src + ((size_t)sy1 * wl + sx1) * NOP + c) * sc;
        const float r0 = a * (1.f - fx) + b * fx;
        const float r1 = cc * (1.f - fx) + dd * fx;
        res[e][c] = r0 * (1.f - fy) + r1 * fy;
      }
    }
  }
  const bool two = xp + 1 < w0;
  const bool aligned = (((size_t)dst) & (sizeof(float) * 2 * NOP - 1)) == 0;
  if (two && aligned) {
    if (NOP == 2) *reinterpret_cast<float4*>(dst) = make_float4(res[0][0], res[0][NOP - 1], res[1][0], res[1][NOP - 1]);
    else *reinterpret_cast<float2*>(dst) = make_float2(res[0][0], res[1][0]);
  } else {
#pragma unroll
    for (int c = 0; c < NOP; ++c) dst[c] = res[0][c];
    if (two)
#pragma unroll
      for (int c = 0; c < NOP; ++c) dst[NOP + c] = res[1][c];
  }
}

}  // namespace

void launch_upsample_crop(const float* flow, const LevelGeom& g, int nop, int sc_l, int padl, int padt, int w0,
                          int h0, float* out, int n, cudaStream_t s, LaunchStats* st) {
  dim3 blk(128), grd(((w0 + 1) / 2 + 127) / 128, h0, n);
  if (nop == 2) upsample_crop_kernel<2><<<grd, blk, 0, s>>>(flow, g.w, g.h, sc_l, padl, padt, w0, h0, out);
  else upsample_crop_kernel<1><<<grd, blk, 0, s>>>(flow, g.w, g.h, sc_l, padl, padt, w0, h0, out);
  st->launches++;
}

}  // namespace disb200
