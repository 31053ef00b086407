// Stage 5 — final upsample and crop (run_dense.cpp:407-415):
//   flowout *= 2^sc_l ; cv::resize(INTER_LINEAR, x 2^sc_l) ; crop the divisibility padding.
// cv::resize semantics (pinned against python-cv2 4.13, tests/golden): half-pixel centres,
// source index clamped with the fraction forced to 0 at the borders, horizontal pass then
// vertical pass.  This is the HBM-heavy kernel of the path (80 % of the compulsory bytes).
// Each thread owns 4 adjacent output columns and walks down ROWS output rows: the horizontal
// pass on the two source rows is redone only when the source row changes (once per 2^sc_l
// output rows), the vertical pass is one lerp per output value, and stores are 128-bit.
// The source coordinates are computed in fp32: (d + pad + 0.5) * 2^-sc_l - 0.5 is exact in
// fp32 for a power-of-two scale, so it equals OpenCV's double computation rounded to float.
#include <algorithm>

#include "dis_internal.h"

namespace disb200 {

namespace {

constexpr int UPS_ROWS = 8;
constexpr int UPS_COLS = 4;

__device__ __forceinline__ void src_coord(int d, int pad, float inv, int n, int& s0, int& s1, float& f) {
  f = ((float)(d + pad) + 0.5f) * inv - 0.5f;
  const float fl = floorf(f);
  s0 = (int)fl;
  f -= fl;
  if (s0 < 0) { s0 = 0; f = 0.f; }
  if (s0 >= n - 1) { s0 = n - 1; f = 0.f; }
  s1 = min(s0 + 1, n - 1);
}

template <int NOP>
__global__ void __launch_bounds__(128)
upsample_crop_kernel(const float* __restrict__ flow, int wl, int hl, int sc_l, int padl, int padt, int w0, int h0,
                     float* __restrict__ out, int oy0, int oy1) {
  const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * UPS_COLS, y0 = oy0 + blockIdx.y * UPS_ROWS, pair = blockIdx.z;
  if (x0 >= w0) return;
  const float* src = flow + (size_t)pair * wl * hl * NOP;
  const float inv = 1.0f / (float)(1 << sc_l), sc = (float)(1 << sc_l);
  int sx0[UPS_COLS], sx1[UPS_COLS];
  float fx[UPS_COLS];
#pragma unroll
  for (int e = 0; e < UPS_COLS; ++e) src_coord(min(x0 + e, w0 - 1), padl, inv, wl, sx0[e], sx1[e], fx[e]);
  float r0[UPS_COLS][NOP], r1[UPS_COLS][NOP];
  int cur_sy = -1;
  const int ncol = min(UPS_COLS, w0 - x0);
#pragma unroll 1
  for (int r = 0; r < UPS_ROWS; ++r) {
    const int y = y0 + r;
    if (y >= oy1) break;
    int sy0, sy1;
    float fy;
    src_coord(y, padt, inv, hl, sy0, sy1, fy);
    if (sy0 != cur_sy) {   // horizontal pass on the two source rows
      cur_sy = sy0;
      const float* p0 = src + (size_t)sy0 * wl * NOP;
      const float* p1 = src + (size_t)sy1 * wl * NOP;
#pragma unroll
      for (int e = 0; e < UPS_COLS; ++e)
#pragma unroll
        for (int c = 0; c < NOP; ++c) {
          const float a = __ldg(p0 + sx0[e] * NOP + c) * sc, b = __ldg(p0 + sx1[e] * NOP + c) * sc;
          const float cc = __ldg(p1 + sx0[e] * NOP + c) * sc, dd = __ldg(p1 + sx1[e] * NOP + c) * sc;
          r0[e][c] = a * (1.f - fx[e]) + b * fx[e];
          r1[e][c] = cc * (1.f - fx[e]) + dd * fx[e];
        }
    }
    float v[UPS_COLS * NOP];
#pragma unroll
    for (int e = 0; e < UPS_COLS; ++e)
#pragma unroll
      for (int c = 0; c < NOP; ++c) v[e * NOP + c] = r0[e][c] * (1.f - fy) + r1[e][c] * fy;
    float* dst = out + ((size_t)pair * w0 * h0 + (size_t)y * w0 + x0) * NOP;
    if (ncol == UPS_COLS && (((size_t)dst) & 15) == 0) {
#pragma unroll
      for (int q = 0; q < UPS_COLS * NOP / 4; ++q)
        __stcs(reinterpret_cast<float4*>(dst) + q, make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]));
    } else {
      for (int e = 0; e < ncol; ++e)
#pragma unroll
        for (int c = 0; c < NOP; ++c) dst[e * NOP + c] = v[e * NOP + c];
    }
  }
}

// sc_l == 0: the flow is already at full resolution, only the crop remains (run_dense.cpp:408,415)
template <int NOP>
__global__ void crop_kernel(const float* __restrict__ flow, int wl, int hl, int padl, int padt, int w0, int h0,
                            float* __restrict__ out, int oy0) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = oy0 + blockIdx.y, pair = blockIdx.z;
  if (x >= w0) return;
  const float* s = flow + ((size_t)pair * wl * hl + (size_t)(y + padt) * wl + (x + padl)) * NOP;
  float* d = out + ((size_t)pair * w0 * h0 + (size_t)y * w0 + x) * NOP;
#pragma unroll
  for (int c = 0; c < NOP; ++c) d[c] = __ldg(s + c);
}

// Temporal warm start: the level-sc_l flow of the previous pair -> the `initflow` array OFClass takes
// (oflow.h:106, oflow.cpp:213-217: (h_c / 2) x (w_c / 2) x nop at half the coarsest resolution, in units of that grid),
// by the rule of the reference's disabled caller (run_dense.cpp:370-373: flow *= scale; cv::resize(INTER_AREA) = box mean).
template <int NOP>
__global__ void warm_init_kernel(const float* __restrict__ flow, int wl, int k, float scale, int wi, int hi,
                                 float* __restrict__ init) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= wi || y >= hi) return;
  float acc[NOP];
#pragma unroll
  for (int c = 0; c < NOP; ++c) acc[c] = 0.f;
  for (int r = 0; r < k; ++r)
    for (int q = 0; q < k; ++q)
#pragma unroll
      for (int c = 0; c < NOP; ++c) acc[c] += flow[((size_t)(y * k + r) * wl + (x * k + q)) * NOP + c];
  const float m = scale / (float)(k * k);
#pragma unroll
  for (int c = 0; c < NOP; ++c) init[((size_t)y * wi + x) * NOP + c] = acc[c] * m;
}

}  // namespace

void launch_warm_init(const float* flow, int wl, int nop, int k, float scale, int wi, int hi, float* init, cudaStream_t s,
                      LaunchStats* st) {
  if (wi < 1 || hi < 1) return;
  dim3 blk(64), grd((wi + 63) / 64, hi);
  if (nop == 2) warm_init_kernel<2><<<grd, blk, 0, s>>>(flow, wl, k, scale, wi, hi, init);
  else warm_init_kernel<1><<<grd, blk, 0, s>>>(flow, wl, k, scale, wi, hi, init);
  st->launches++;
}

void launch_upsample_crop(const float* flow, const LevelGeom& g, int nop, int sc_l, int padl, int padt, int w0,
                          int h0, float* out, int n, cudaStream_t s, LaunchStats* st, int out_y0, int out_ny) {
  const int oy0 = out_y0, oy1 = out_ny < 0 ? h0 : std::min(h0, out_y0 + out_ny), rows = oy1 - oy0;
  if (rows <= 0) return;
  if (sc_l == 0) {
    dim3 blk(128), grd((w0 + 127) / 128, rows, n);
    if (nop == 2) crop_kernel<2><<<grd, blk, 0, s>>>(flow, g.w, g.h, padl, padt, w0, h0, out, oy0);
    else crop_kernel<1><<<grd, blk, 0, s>>>(flow, g.w, g.h, padl, padt, w0, h0, out, oy0);
  } else {
    const int per_block = 128 * UPS_COLS;
    dim3 blk(128), grd((w0 + per_block - 1) / per_block, (rows + UPS_ROWS - 1) / UPS_ROWS, n);
    if (nop == 2) upsample_crop_kernel<2><<<grd, blk, 0, s>>>(flow, g.w, g.h, sc_l, padl, padt, w0, h0, out, oy0, oy1);
    else upsample_crop_kernel<1><<<grd, blk, 0, s>>>(flow, g.w, g.h, sc_l, padl, padt, w0, h0, out, oy0, oy1);
  }
  st->launches++;
}

}  // namespace disb200
