// Stage 1 — image pyramid with Sobel gradients, replacing the OpenCV calls of
// ConstructImgPyramide (run_dense.cpp:131-179) and the dimension padding of
// run_dense.cpp:299-312.
//
// cv::resize(.5,.5,INTER_LINEAR) is a 2x2 box mean and a cascade of l such steps on a
// u8-sourced image is the exact mean of the 2^l x 2^l block (all intermediates are dyadic
// rationals below 2^24, l <= 8), so the finest level that is actually read (sc_l) is computed
// straight from the u8 frame in one pass; levels 0..sc_l-1, which the reference builds but
// never reads (oflow.cpp:182), are skipped.  Coarser levels come from the previous one.
// Both frames of every pair go through the same launch (blockIdx.z enumerates the images of role A, then B;
// a video sequence is role A only: each frame's pyramid is built once and serves two pairs) and
// one launch computes the Sobel gradients and border fills of all levels.
#include "dis_internal.h"

namespace disb200 {

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }
__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return clampi(i, 0, n - 1);
}

// One thread per level pixel.  Reads the replicate-padded u8 frame; interior grey blocks whose
// rows are 4-byte aligned are summed with 32-bit loads + dp4a, everything else byte by byte.
template <int C>
__global__ void __launch_bounds__(128)
downscale_u8_kernel(const uint8_t* __restrict__ framesA, const uint8_t* __restrict__ framesB, int w0, int h0, int padl,
                    int padt, int lvl, int w, int h, int pad, int tw, size_t img_elems, float* __restrict__ imgA,
                    float* __restrict__ imgB, int nA) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  // blockIdx.z enumerates images: the nA images of role A (first frames / all frames of a sequence), then role B
  const int frame = (int)blockIdx.z >= nA, pair = frame ? blockIdx.z - nA : blockIdx.z;
  if (x >= w) return;
  const uint8_t* f = (frame ? framesB : framesA) + (size_t)pair * w0 * h0 * C;
  float* img = frame ? imgB : imgA;
  const int k = 1 << lvl;
  unsigned sum[C];
#pragma unroll
  for (int c = 0; c < C; ++c) sum[c] = 0;
  const int xs = x * k - padl;   // first source column of the block
  const bool fast = C == 1 && k >= 4 && xs >= 0 && xs + k <= w0 && ((xs & 3) == 0) && ((w0 & 3) == 0) &&
                    ((((size_t)f) & 3) == 0);
  const bool fast8 = fast && k >= 8 && ((xs & 7) == 0) && ((w0 & 7) == 0) && ((((size_t)f) & 7) == 0);
  if (fast8 && k == 8) {           // the 8x8 block of level 3: all eight 64-bit loads in flight at once
    uint2 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      v[j] = __ldg(reinterpret_cast<const uint2*>(f + (size_t)clampi(y * 8 + j - padt, 0, h0 - 1) * w0 + xs));
    unsigned s1 = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sum[0] = __dp4a(v[j].x, 0x01010101u, sum[0]);
      s1 = __dp4a(v[j].y, 0x01010101u, s1);
    }
    sum[0] += s1;
  } else if (fast8) {              // 64-bit loads: one per 8 source pixels
    unsigned s1 = 0;
#pragma unroll 4
    for (int j = 0; j < k; ++j) {
      const int sy = clampi(y * k + j - padt, 0, h0 - 1);
      const uint2* row = reinterpret_cast<const uint2*>(f + (size_t)sy * w0 + xs);
      for (int i = 0; i < k / 8; ++i) {
        const uint2 v = __ldg(row + i);
        sum[0] = __dp4a(v.x, 0x01010101u, sum[0]);
        s1 = __dp4a(v.y, 0x01010101u, s1);
      }
    }
    sum[0] += s1;
  } else if (fast) {
    for (int j = 0; j < k; ++j) {
      const int sy = clampi(y * k + j - padt, 0, h0 - 1);
      const unsigned* row = reinterpret_cast<const unsigned*>(f + (size_t)sy * w0 + xs);
      for (int i = 0; i < k / 4; ++i) sum[0] = __dp4a(__ldg(row + i), 0x01010101u, sum[0]);
    }
  } else if (C == 3 && k >= 4 && xs >= 0 && xs + k <= w0 && ((xs & 3) == 0) && ((w0 & 3) == 0) && ((((size_t)f) & 3) == 0)) {
    // BGR: a block row is 3 k consecutive bytes starting on a pixel boundary, read as 32-bit words; byte i of word
    // wi belongs to channel (4 wi + i) % 3 = (wi + i) % 3, so one dp4a per word and channel with the selector
    // {b0 + b3, b1, b2} rotated by wi % 3 sums it.
    const unsigned sel[3] = {0x01000001u, 0x00000100u, 0x00010000u};
    if (k == 8) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {   // 4 rows x 6 words in flight
        unsigned v[4][6];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const unsigned* row = reinterpret_cast<const unsigned*>(f + ((size_t)clampi(y * 8 + half * 4 + j - padt, 0, h0 - 1) * w0 + xs) * 3);
#pragma unroll
          for (int wi = 0; wi < 6; ++wi) v[j][wi] = __ldg(row + wi);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int wi = 0; wi < 6; ++wi)
#pragma unroll
            for (int c = 0; c < 3; ++c) sum[c] = __dp4a(v[j][wi], sel[(c + 3 - wi % 3) % 3], sum[c]);
      }
    } else {
      for (int j = 0; j < k; ++j) {
        const unsigned* row = reinterpret_cast<const unsigned*>(f + ((size_t)clampi(y * k + j - padt, 0, h0 - 1) * w0 + xs) * 3);
        for (int w3 = 0; w3 < 3 * k / 4; w3 += 3) {   // three words = four pixels
#pragma unroll
          for (int wi = 0; wi < 3; ++wi) {
            const unsigned v = __ldg(row + w3 + wi);
#pragma unroll
            for (int c = 0; c < 3; ++c) sum[c] = __dp4a(v, sel[(c + 3 - wi) % 3], sum[c]);
          }
        }
      }
    }
  } else {
    for (int j = 0; j < k; ++j) {
      const int sy = clampi(y * k + j - padt, 0, h0 - 1);
      const uint8_t* row = f + (size_t)sy * w0 * C;
      for (int i = 0; i < k; ++i) {
        const int sx = clampi(xs + i, 0, w0 - 1);
#pragma unroll
        for (int c = 0; c < C; ++c) sum[c] += row[sx * C + c];
      }
    }
  }
  const float inv = 1.0f / (float)(1u << (2 * lvl));
  float* o = img + (size_t)pair * img_elems + ((size_t)(y + pad) * tw + (x + pad)) * C;
#pragma unroll
  for (int c = 0; c < C; ++c) o[c] = (float)sum[c] * inv;
}

template <int C>
__global__ void halve_kernel(const float* __restrict__ srcA, const float* __restrict__ srcB, int stw, int spad,
                             size_t s_elems, float* __restrict__ dstA, float* __restrict__ dstB, int w, int h, int pad,
                             int tw, size_t d_elems, int nA) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int frame = (int)blockIdx.z >= nA, pair = frame ? blockIdx.z - nA : blockIdx.z;
  if (x >= w) return;
  const float* s = (frame ? srcB : srcA) + (size_t)pair * s_elems;
  float* o = (frame ? dstB : dstA) + (size_t)pair * d_elems + ((size_t)(y + pad) * tw + (x + pad)) * C;
  const float* r0 = s + ((size_t)(2 * y + spad) * stw + (2 * x + spad)) * C;
  const float* r1 = r0 + (size_t)stw * C;
  // OpenCV's own 2x2 mean takes four output pixels per 128-bit step as (a+b)+(d+e) and sums the w % 4 tail
  // columns of a row one by one; the order only shows on non-dyadic data (gradient-magnitude pyramid).
  const bool tail = C == 1 && x >= (w & ~3);
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const float ab = r0[c] + r0[C + c];
    o[c] = (tail ? (ab + r1[c]) + r1[C + c] : ab + (r1[c] + r1[C + c])) * 0.25f;
  }
}

// SELECTCHANNEL == 2 (run_dense.cpp:139-148): level 0 is sqrt(dx^2 + dy^2) of the Sobel/8 gradients (REFLECT_101
// at the borders of the dimension-padded frame) and every coarser level a cv::resize(.5) of the previous one.
// One thread per pixel of level `lvl` walks its 2^lvl x 2^lvl block of level 0 in Morton order and folds the
// cascade on the fly (three pending values per cascade level), so no finer level is ever written.  The gradients
// are integers / 8: 8 gx = 2 r1 + r0 + r2, (8 gx)^2 + (8 gy)^2 < 2^24 is exact in fp32 and sqrt(s / 64) =
// sqrt(s) / 8 exactly, which leaves cv::sqrt's correct rounding as the only rounding of level 0.
__global__ void __launch_bounds__(128)
downscale_gradmag_kernel(const uint8_t* __restrict__ framesA, const uint8_t* __restrict__ framesB, int w0, int h0, int W,
                         int H, int padl, int padt, int lvl, int w, int h, int pad, int tw, size_t img_elems,
                         float* __restrict__ imgA, float* __restrict__ imgB, int nA) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int frame = (int)blockIdx.z >= nA, pair = frame ? blockIdx.z - nA : blockIdx.z;
  if (x >= w) return;
  const uint8_t* f = (frame ? framesB : framesA) + (size_t)pair * w0 * h0;
  float* img = frame ? imgB : imgA;
  float acc[8][3];
  float v = 0.f;
  const int total = 1 << (2 * lvl);
  for (int m = 0; m < total; ++m) {
    int ix = 0, iy = 0;
    for (int b = 0; b < lvl; ++b) { ix |= ((m >> (2 * b)) & 1) << b; iy |= ((m >> (2 * b + 1)) & 1) << b; }
    const int X = (x << lvl) + ix, Y = (y << lvl) + iy;
    int col[3], row[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      col[d] = clampi(reflect101(X + d - 1, W) - padl, 0, w0 - 1);
      row[d] = clampi(reflect101(Y + d - 1, H) - padt, 0, h0 - 1) * w0;
    }
    int s[3][3];
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
      for (int i = 0; i < 3; ++i) s[j][i] = __ldg(f + row[j] + col[i]);
    const int gx = 2 * (s[1][2] - s[1][0]) + (s[0][2] - s[0][0]) + (s[2][2] - s[2][0]);
    const int gy = 2 * (s[2][1] - s[0][1]) + (s[2][0] - s[0][0]) + (s[2][2] - s[0][2]);
    v = sqrtf((float)(gx * gx + gy * gy)) * 0.125f;
    int mm = m;
    for (int j = 0; j < lvl; ++j, mm >>= 2) {
      const int q = mm & 3;
      if (q < 3) { acc[j][q] = v; break; }
      const int wj = W >> (j + 1);
      const bool tail = (X >> (j + 1)) >= (wj & ~3);
      const float ab = acc[j][0] + acc[j][1];
      v = (tail ? (ab + acc[j][2]) + v : ab + (acc[j][2] + v)) * 0.25f;
    }
  }
  img[(size_t)pair * img_elems + (size_t)(y + pad) * tw + (x + pad)] = v;
}

struct SobelLevels {
  int nlev, rows;
  int w[9], h[9], tw[9], th[9], row0[9];   // row0: first grid row of the level
  size_t elems[9];
  float* img[9][2];
  float* dx[9][2];
  float* dy[9][2];
};

// One thread per FOUR adjacent padded pixels of a row, all levels in one launch: replicate the image border,
// Sobel(k=3, 1/8, REFLECT_101) on the unpadded level image for the interior of dx/dy, zeros on their border.
// The 3 x 6 neighbourhood of the four pixels is loaded once (18 loads for 8 gradient values).
constexpr int SOBEL_PX = 4;
template <int C>
__global__ void __launch_bounds__(128) sobel_pad_kernel(SobelLevels L, int pad, int nA) {
  const int yg = blockIdx.y * blockDim.y + threadIdx.y;   // row over all levels
  if (yg >= L.rows) return;
  int lv = 0;
#pragma unroll 1
  while (lv + 1 < L.nlev && yg >= L.row0[lv + 1]) ++lv;
  const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * SOBEL_PX;
  const int y = yg - L.row0[lv];
  const int frame = (int)blockIdx.z >= nA, pair = frame ? blockIdx.z - nA : blockIdx.z;
  const int w = L.w[lv], h = L.h[lv], tw = L.tw[lv];
  if (x0 >= tw) return;
  const size_t elems = L.elems[lv];
  float* I = L.img[lv][frame] + (size_t)pair * elems;
  float* gx = L.dx[lv][frame] + (size_t)pair * elems;
  float* gy = L.dy[lv][frame] + (size_t)pair * elems;
  const int sy = y - pad, sx0 = x0 - pad;
  const int npx = min(SOBEL_PX, tw - x0);
  const size_t rowo = (size_t)y * tw;
  if (sy < 0 || sy >= h) {                       // border row: replicate the image, zero gradients
    const float* srow = I + (size_t)(clampi(sy, 0, h - 1) + pad) * tw * C;
    for (int k = 0; k < npx; ++k) {
      const int cx = clampi(sx0 + k, 0, w - 1) + pad;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const size_t o = (rowo + x0 + k) * C + c;
        I[o] = srow[cx * C + c];
        gx[o] = 0.f;
        gy[o] = 0.f;
      }
    }
    return;
  }
  int col[SOBEL_PX + 2];
#pragma unroll
  for (int e = 0; e < SOBEL_PX + 2; ++e) col[e] = (reflect101(sx0 + e - 1, w) + pad) * C;
  const float* rm = I + (size_t)(reflect101(sy - 1, h) + pad) * tw * C;
  const float* rc = I + (size_t)(sy + pad) * tw * C;
  const float* rp = I + (size_t)(reflect101(sy + 1, h) + pad) * tw * C;
#pragma unroll
  for (int c = 0; c < C; ++c) {
    float vm[SOBEL_PX + 2], vc[SOBEL_PX + 2], vp[SOBEL_PX + 2];
#pragma unroll
    for (int e = 0; e < SOBEL_PX + 2; ++e) { vm[e] = rm[col[e] + c]; vc[e] = rc[col[e] + c]; vp[e] = rp[col[e] + c]; }
    float ox[SOBEL_PX], oy[SOBEL_PX];
    bool in[SOBEL_PX];
#pragma unroll
    for (int k = 0; k < SOBEL_PX; ++k) {
      const int sx = sx0 + k;
      in[k] = sx >= 0 && sx < w;
      const float r0 = vm[k + 2] - vm[k], r1 = vc[k + 2] - vc[k], r2 = vp[k + 2] - vp[k];
      const float s0 = 0.25f * vm[k + 1] + 0.125f * (vm[k + 2] + vm[k]);
      const float s2 = 0.25f * vp[k + 1] + 0.125f * (vp[k + 2] + vp[k]);
      ox[k] = in[k] ? 0.25f * r1 + 0.125f * (r2 + r0) : 0.f;
      oy[k] = in[k] ? s2 - s0 : 0.f;
    }
    const size_t o0 = (rowo + x0) * C + c;
    if (C == 1 && npx == SOBEL_PX && (((size_t)(gx + o0) | (size_t)(gy + o0)) & 15) == 0) {
      *reinterpret_cast<float4*>(gx + o0) = make_float4(ox[0], ox[1], ox[2], ox[3]);
      *reinterpret_cast<float4*>(gy + o0) = make_float4(oy[0], oy[1], oy[2], oy[3]);
    } else {
#pragma unroll
      for (int k = 0; k < SOBEL_PX; ++k)
        if (k < npx) { gx[o0 + (size_t)k * C] = ox[k]; gy[o0 + (size_t)k * C] = oy[k]; }
    }
    // left / right border of an interior row: replicate the nearest interior pixel (vc of the clamped column)
#pragma unroll
    for (int k = 0; k < SOBEL_PX; ++k)
      if (k < npx && !in[k]) I[o0 + (size_t)k * C] = rc[(clampi(sx0 + k, 0, w - 1) + pad) * C + c];
  }
}

void launch_downscale_u8(const uint8_t* framesA, const uint8_t* framesB, int nA, int nB, int w0, int h0, int C, int padl,
                         int padt, int lvl, const LevelGeom& g, float* imgA, float* imgB, cudaStream_t s,
                         LaunchStats* st) {
  dim3 blk(128), grd((g.w + 127) / 128, g.h, nA + nB);
  if (C == 1)
    downscale_u8_kernel<1><<<grd, blk, 0, s>>>(framesA, framesB, w0, h0, padl, padt, lvl, g.w, g.h, g.pad, g.tw, g.img_elems, imgA, imgB, nA);
  else
    downscale_u8_kernel<3><<<grd, blk, 0, s>>>(framesA, framesB, w0, h0, padl, padt, lvl, g.w, g.h, g.pad, g.tw, g.img_elems, imgA, imgB, nA);
  st->launches++;
}

void launch_downscale_gradmag(const uint8_t* framesA, const uint8_t* framesB, int nA, int nB, int w0, int h0, int W, int H,
                              int padl, int padt, int lvl, const LevelGeom& g, float* imgA, float* imgB, cudaStream_t s,
                              LaunchStats* st) {
  dim3 blk(128), grd((g.w + 127) / 128, g.h, nA + nB);
  downscale_gradmag_kernel<<<grd, blk, 0, s>>>(framesA, framesB, w0, h0, W, H, padl, padt, lvl, g.w, g.h, g.pad, g.tw,
                                               g.img_elems, imgA, imgB, nA);
  st->launches++;
}

void launch_halve(const float* srcA, const float* srcB, const LevelGeom& gs, float* dstA, float* dstB,
                  const LevelGeom& gd, int nA, int nB, int C, cudaStream_t s, LaunchStats* st) {
  dim3 blk(128), grd((gd.w + 127) / 128, gd.h, nA + nB);
  if (C == 1)
    halve_kernel<1><<<grd, blk, 0, s>>>(srcA, srcB, gs.tw, gs.pad, gs.img_elems, dstA, dstB, gd.w, gd.h, gd.pad, gd.tw, gd.img_elems, nA);
  else
    halve_kernel<3><<<grd, blk, 0, s>>>(srcA, srcB, gs.tw, gs.pad, gs.img_elems, dstA, dstB, gd.w, gd.h, gd.pad, gd.tw, gd.img_elems, nA);
  st->launches++;
}

void launch_sobel_pad_levels(const PyrLevelPtrs* lv, const LevelGeom* g, int nlev, int nA, int nB, int C, cudaStream_t s,
                             LaunchStats* st) {
  SobelLevels L;
  L.nlev = nlev;
  int rows = 0, maxtw = 0;
  for (int i = 0; i < nlev; ++i) {
    L.w[i] = g[i].w; L.h[i] = g[i].h; L.tw[i] = g[i].tw; L.th[i] = g[i].th; L.elems[i] = g[i].img_elems;
    L.row0[i] = rows;
    rows += g[i].th;
    maxtw = g[i].tw > maxtw ? g[i].tw : maxtw;
    for (int f = 0; f < 2; ++f) { L.img[i][f] = lv[i].img[f]; L.dx[i][f] = lv[i].dx[f]; L.dy[i][f] = lv[i].dy[f]; }
  }
  L.rows = rows;
  const int per_block = 16 * SOBEL_PX;   // 16 x 8 threads: 64 pixels x 8 rows per block
  dim3 blk(16, 8), grd((maxtw + per_block - 1) / per_block, (rows + 7) / 8, nA + nB);
  if (C == 1) sobel_pad_kernel<1><<<grd, blk, 0, s>>>(L, g[0].pad, nA);
  else sobel_pad_kernel<3><<<grd, blk, 0, s>>>(L, g[0].pad, nA);
  st->launches++;
}

}  // namespace disb200
