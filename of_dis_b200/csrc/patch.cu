// Stage 2 — per-patch inverse-compositional Gauss-Newton search.
//
// One kernel does, per patch, what the reference spreads over
//   PatClass::InitializePatch / getPatchStaticNNGrad / ComputeHessian   (patch.cpp:55-86,298-343)
//   PatGridClass::InitializeFromCoarserOF                               (patchgrid.cpp:195-211)
//   PatClass::OptimizeStart / OptimizeIter / OptimizeComputeErrImg      (patch.cpp:118-210,275-295)
//   getPatchStaticBil / LossComputeErrorImage                           (patch.cpp:221-273,346-413)
// A sub-warp of G lanes owns one patch; value n of the patch (row-major, channels
// interleaved) lives in lane n % G, slot n / G, so that a load instruction of the group
// touches consecutive floats of the target image.  8x8 grey patches (op1 / op2) use the
// column-pair mapping instead: 4 lanes per patch, a lane owns two adjacent columns, three
// taps per row serve both of them and the lower taps of one row are the upper taps of the
// next, so an iteration loads 3 (PSZ + 1) values for 16 results (a plain mapping loads 64)
// and the scalar part of the iteration (solve, bounds, bookkeeping) is shared by 8 patches
// per warp instead of 4: 31 % fewer warp instructions per patch than the 8-lane row mapping.  Template and gradients stay in registers; the residual is consumed on the fly
// (its projections on the steepest-descent images and its weight sum are all the next
// iteration needs) and the per-pixel weights are written once, by the evaluation that ends
// the search.  Reductions are xor-butterflies, which leave every lane with the bit-identical
// sum, so all scalar state is replicated and the loop is uniform inside a group.  Groups of
// one warp may need different trip counts (min_iter < max_iter): the loop runs until every
// group of the warp has converged and finished groups stop committing state.
#include <climits>

#include "dis_internal.h"

namespace disb200 {

#ifndef DISB200_STAGE12
#define DISB200_STAGE12 true   // 12 x 12 grey search with staged tap windows (false: loads from global memory, A/B)
#endif
#ifndef DISB200_STAGE_BGR
#define DISB200_STAGE_BGR true   // 8 x 8 BGR search with staged tap windows (false: loads from global memory, A/B)
#endif
#ifndef DISB200_STAGE_BGR12
#define DISB200_STAGE_BGR12 false   // 12 x 12 BGR search staged: 244 registers (two CTAs per SM); see DESIGN.md 3.0
#endif
#ifndef DISB200_STAGE_MAX_STEP
#define DISB200_STAGE_MAX_STEP 64   // largest grid step the 8 x 8 grey search is staged for (A/B: 4 was the rule before the window skip)
#endif

// sum / NV; for a power-of-two NV the product with the exact reciprocal is the same bits as the division
template <int NV, bool APPROX = false>
__device__ __forceinline__ float mean_of(float sum) {
  if (APPROX || (NV & (NV - 1)) == 0) return sum * (1.0f / (float)NV);
  return sum / (float)NV;
}

template <int G>
__device__ __forceinline__ float gsum(float v) {
#pragma unroll
  for (int m = G / 2; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  return v;
}

// Partial sums of one lane and their reduction over the group.  CP (column-pair mapping: a lane owns two
// adjacent columns of the patch, slot k = 2 * row + parity) keeps one accumulator per column and combines
// them in the order of the 8-lane butterfly — (c, c^4), then (c, c^2), then even + odd — so that a
// 4-lane group produces the same bits as the 8-lane mapping it replaces (and as the reference's packet sums).
template <int G, bool CP>
struct PSum {
  float a0 = 0.f, a1 = 0.f;
  __device__ __forceinline__ void add(int k, float v) {
    if (CP && (k & 1)) a1 += v;
    else a0 += v;
  }
  // a0 += x * y; FM: as one fused multiply-add (only where the result is compared with a tolerance)
  template <bool FM>
  __device__ __forceinline__ void mad(int k, float x, float y) {
    if (FM && !CP) a0 = fmaf(x, y, a0);
    else add(k, x * y);
  }
  __device__ __forceinline__ float total() const {
    if (!CP) return gsum<G>(a0);
    float r0 = a0, r1 = a1;
#pragma unroll
    for (int m = G / 2; m > 0; m >>= 1) {
      r0 += __shfl_xor_sync(0xffffffffu, r0, m);
      r1 += __shfl_xor_sync(0xffffffffu, r1, m);
    }
    return r0 + r1;
  }
};

// Packed fp32 pairs (FMUL2 / FADD2 / FFMA2, sm_100): one issue slot for two values.  ptxas 12.9 contracts a packed multiply
// feeding a packed add into FFMA2 whatever the rounding suffix and -fmad say, so these are only used where the result is
// compared with a tolerance (the FM mappings).
__device__ __forceinline__ float2 f2fma(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 f2mul(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 f2add(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 f2sub(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }
__device__ __forceinline__ float2 f2dup(float a) { return make_float2(a, a); }

// LossComputeErrorImage for one value (patch.cpp:221-273): d = query - template
template <int COST>
__device__ __forceinline__ void loss(float d, float bsq, float bsq2, float& pd, float& pw) {
  if (COST == 0) {          // L2, patch.cpp:228-241
    pd = d;
    pw = fabsf(d);
  } else if (COST == 1) {   // L1, :242-257
    pw = sqrtf(fabsf(d));
    pd = copysignf(pw, d);
  } else {                  // pseudo-Huber, :258-272
    pw = sqrtf((sqrtf(1.0f + (d * d) / bsq) - 1.0f) * bsq2);
    pd = copysignf(pw, d);
  }
}

// LPR > 0 selects the block mapping (12 x 12 patches, op3 / op4): the G lanes of a patch form LPR columns x G / LPR
// row groups and a lane owns a block of RPL rows x CW consecutive floats of a row (channels interleaved).  The
// bilinear taps of a block row are CW + C consecutive floats and the lower taps of one row are the upper taps of the
// next, so an iteration loads (RPL + 1)(CW + C) values for RPL CW results instead of four per result:
// 28 for 18 (12 x 12 grey, 8 lanes) and 48 for 27 (12 x 12 BGR, 16 lanes).
//
// STAGE (8 x 8 grey column-pair mapping and packed 12 x 12 grey mapping; DisTuning::patch_map = 1 selects the loads from
// global memory): the tap window of each of the warp's patches - 9 x 9 floats for eight 8 x 8 patches, 13 x 13 for four
// 12 x 12 patches - is copied to shared memory by the WHOLE warp (4-byte cp.async, each warp instruction covering 32
// consecutive window elements = 3-4 image rows), and only when the window's integer origin moved since the evaluation
// before; every lane then reads its taps from there.  Same values, same arithmetic: bit-identical to the direct loads.
// Measurements: DESIGN.md 3.0 / 3.0c.
template <int PSZ, int C, int G, bool OF, int LPR = 0, bool FM = false, bool STAGE = false>
// (0 = no occupancy request: an explicit 1 made ptxas spend up to 228 registers on the other mappings, -15 %.  Four CTAs per SM
//  = 128 registers for the staged 8 x 8 and the packed 12 x 12 grey mappings, which fit without spilling inside the loop;
//  capping the column-pair variant at 96 registers for 5 CTAs per SM spills there: 1.60 -> 2.00 ms)
__global__ void __launch_bounds__(128, (C == 1 && (STAGE || (PSZ == 12 && G == 8 && FM))) ? 4 : ((C == 3 && STAGE && PSZ == 8) ? 3 : 0))
patch_search_kernel(PatchArgs a, LevelGeom g, OptConst o) {
  constexpr int NV = PSZ * PSZ * C;
  constexpr int VPL = (NV + G - 1) / G;
  constexpr int ROWLEN = PSZ * C;
  constexpr int LB = -PSZ / 2;
  constexpr int NOP = OF ? 2 : 1;
  constexpr bool FULL = (NV % G) == 0;       // every slot of every lane holds a value
  constexpr bool ROWMAP = ROWLEN == G;       // slot k == patch row k
  constexpr bool CP = C == 1 && ROWLEN == 2 * G && FULL;   // slot k == (row k / 2, column 2 lane + k % 2)
  constexpr bool BLK = LPR > 0;
#ifdef DISB200_FM_EXACT_DIV
  constexpr bool RCP = false;
#else
  constexpr bool RCP = FM;                   // reciprocals instead of divisions in the 2 x 2 solve (FM mappings are tolerance-compared)
#endif
  constexpr int CW = BLK ? ROWLEN / LPR : 1;            // floats of a row per lane
  constexpr int RPL = BLK ? PSZ / (G / LPR) : 1;        // rows per lane
  static_assert(!BLK || (FULL && !CP && !ROWMAP && CW * LPR == ROWLEN && RPL * (G / LPR) == PSZ && CW * RPL == VPL),
                "block mapping must tile the patch exactly");
  // PK (12 x 12 grey, 6 rows x 3 floats per lane): slots k and k + VPL / 2 - rows r and r + 3 of the lane's block - travel
  // as one packed pair through the interpolation, the residual and the three projections
#ifdef DISB200_NO_PACKED
  constexpr bool PK = false;
#else
  constexpr bool PK = BLK && FM && C == 1 && RPL == 6 && CW == 3;
#endif
  constexpr int HV = VPL / 2;
  typedef PSum<G, CP> Sum;
  static_assert(!STAGE || (CP && PSZ == 8 && G == 4) || (PSZ == 12 && C == 1 && G == 8 && FM && LPR == 4) || (C == 3 && BLK),
                "tap-window staging is written for the 8 x 8 grey column-pair, the packed 12 x 12 grey and the BGR block mappings");
  // per warp, 8 x 8: 8 windows of 81 floats, two per 200-float slab at offsets 0 and 97 (bank-conflict-free tap reads)
  constexpr int PKW = PSZ + 1, PKSLOT = PKW * PKW + 24;   // staged 12 x 12 grey: 13 x 13 window, one slot per patch
  // staged BGR block mappings: window of (PSZ + 1) rows x (ROWLEN + C) floats, rows padded by one float, slots = 1 mod 32
  constexpr int BWW = ROWLEN + C, BWWP = BWW + 1, BWH = PSZ + 1, BWN = BWW * BWH;
  constexpr int BSLOT = BWH * BWWP + (33 - (BWH * BWWP) % 32) % 32;
  constexpr int WARP_WIN = (C == 3) ? (32 / G) * BSLOT : ((PSZ == 12) ? 4 * PKSLOT : 800);
  __shared__ float win_all[STAGE ? 4 * WARP_WIN : 1];
  float* const win = win_all + (STAGE ? (threadIdx.x >> 5) * WARP_WIN : 0);
  const int win_base = ((threadIdx.x >> 3) & 3) * 200 + ((threadIdx.x >> 2) & 1) * 97;
  const int lane32 = threadIdx.x & 31;
#ifdef DISB200_NO_STAGE_SKIP
  constexpr bool STAGE_SKIP = false;
#else
  constexpr bool STAGE_SKIP = true;
#endif
  int win_off = INT_MIN;                                  // origin of the window this group's slot holds (STAGE)

  const int lane = threadIdx.x % G;
  const int blk_col = BLK ? (lane % LPR) * CW : 0, blk_row = BLK ? (lane / LPR) * RPL : 0;   // first float / row of the lane's block
  // this device's patches: grid rows gy0 .. gy0+ngy-1 of every grid column (the whole grid unless the pair is tiled)
  const int ip_raw = blockIdx.x * (blockDim.x / G) + threadIdx.x / G;
  const int nloc = g.nopw * g.ngy;
  const bool active = ip_raw < nloc;
  const int ipl = active ? ip_raw : nloc - 1;
  const int gx = ipl / g.ngy, gy = g.gy0 + (ipl - gx * g.ngy);
  const int ip = gx * g.noph + gy;   // patch index is column-major: ip = x * noph + y (patchgrid.cpp:64)
  const int pair = blockIdx.y;
  const int rowstride = g.tw * C;

  const float* timg = a.tmpl_img + (size_t)pair * g.img_elems;
  const float* tdxp = a.tmpl_dx + (size_t)pair * g.img_elems;
  const float* tdyp = a.tmpl_dy + (size_t)pair * g.img_elems;
  const float* qimg = a.tgt_img + (size_t)pair * g.img_elems;

  const float ptrx = (float)(gx * g.steps + g.offw);
  const float ptry = (float)(gy * g.steps + g.offh);

  int off[VPL];
  bool val[VPL];
#pragma unroll
  for (int k = 0; k < VPL; ++k) {
    const int n = k * G + lane;
    val[k] = FULL || n < NV;
    const int nn = val[k] ? n : 0;
    const int row = nn / ROWLEN;
    off[k] = ROWMAP ? k * rowstride + lane : CP ? (k >> 1) * rowstride + 2 * lane + (k & 1)
           : BLK ? (blk_row + k / CW) * rowstride + blk_col + k % CW : row * rowstride + (nn - row * ROWLEN);
  }
  // image offset of window element e = lane32 + 32 k (row e / 9, column e % 9 of the 9 x 9 window)
  int stage_off[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int e = lane32 + 32 * k;
    stage_off[k] = (e / 9) * rowstride + e % 9;
  }
  // index of slot k in the row-major patch (pweight layout)
  auto slot_index = [&](int k) {
    return CP ? (k >> 1) * ROWLEN + 2 * lane + (k & 1) : BLK ? (blk_row + k / CW) * ROWLEN + blk_col + k % CW : k * G + lane;
  };

  // ---- template, gradients, Hessian (patch.cpp:298-343, 69-86) ----
  float t[VPL], dx[VPL], dy[VPL];
  {
    const int px = (int)ptrx + g.pad + LB, py = (int)ptry + g.pad + LB;
    const int base = (py * g.tw + px) * C;
    Sum s;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      t[k] = val[k] ? __ldg(timg + base + off[k]) : 0.f;
      dx[k] = val[k] ? __ldg(tdxp + base + off[k]) : 0.f;
      dy[k] = (OF && val[k]) ? __ldg(tdyp + base + off[k]) : 0.f;
      s.add(k, t[k]);
    }
    if (o.patnorm > 0 || o.costfct == 10) {
      const float m = mean_of<NV, RCP>(s.total());
#pragma unroll
      for (int k = 0; k < VPL; ++k) t[k] = val[k] ? t[k] - m : 0.f;
    }
  }
  // NCC (costfct 10; defined by this build, the reference has no branch for it: DESIGN.md "NCC"): |t^|
  float tnorm = 0.f;
  if (o.costfct == 10) {
    Sum s2;
#pragma unroll
    for (int k = 0; k < VPL; ++k) s2.add(k, t[k] * t[k]);
    tnorm = sqrtf(s2.total());
  }
  float H00, H01 = 0.f, H11 = 0.f;
  {
    Sum h00, h01, h11;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      h00.add(k, dx[k] * dx[k]);
      if (OF) { h01.add(k, dx[k] * dy[k]); h11.add(k, dy[k] * dy[k]); }
    }
    H00 = h00.total();
    if (OF) { H01 = h01.total(); H11 = h11.total(); }
  }
  float l00, l10 = 0.f, l11 = 1.f;
  if (OF) {
    if (H00 * H11 - H01 * H01 == 0.f) { H00 = (float)((double)H00 + 1e-10); H11 = (float)((double)H11 + 1e-10); }
    // Eigen LLT, unblocked: stop at the first non-positive pivot, rest left unfactored
    l00 = H00; l10 = H01; l11 = H11;
    if (l00 > 0.f) {
      l00 = sqrtf(l00);
      l10 = l10 / l00;
      const float x = l11 - l10 * l10;
      if (x > 0.f) l11 = sqrtf(x);
    }
  } else {
    if (H00 == 0.f) H00 = (float)((double)H00 + 1e-10);
    l00 = H00;
    if (l00 > 0.f) l00 = sqrtf(l00);
  }

  const float il00 = RCP ? 1.0f / l00 : 0.f, il11 = (RCP && OF) ? 1.0f / l11 : 0.f;

  // ---- start (patchgrid.cpp:195-211, patch.cpp:97-154) ----
  float p_in0 = 0.f, p_in1 = 0.f;
  if (a.flow_prev) {
    const int x = (int)floorf(ptrx / 2), y = (int)floorf(ptry / 2);
    const float* fp = a.flow_prev + (size_t)pair * a.flow_prev_stride + (size_t)(y * (g.w / 2) + x) * NOP;
    p_in0 = fp[0] * 2;
    if (OF) p_in1 = fp[1] * 2;
  }
  float p0 = p_in0, p1 = p_in1;
  float ptx = ptrx + p0, pty = OF ? ptry + p1 : ptry;
  const float stx = ptx, sty = pty;
  float dp0 = 0.f, dp1 = 0.f;          // delta_p
  float dpsq_init = 1e-10f, mares = 1e5f;
  int cnt = 0;
  const bool start_ok = !(ptx < g.lb || pty < g.lb || ptx > g.ubw || pty > g.ubh);
  bool conv = !start_ok;
  float s0 = 0.f, s1 = 0.f;            // projections of the current residual on the sd images

  const size_t pi = (size_t)pair * g.npatch + ip;
  float* wout = a.pweight + pi * NV;
  if (active && !start_ok) {
    // pweight of a patch whose start is invalid is never written by the reference (patch.cpp:133-139);
    // defined as 0
#pragma unroll
    for (int k = 0; k < VPL; ++k)
      if (val[k]) wout[slot_index(k)] = 0.f;
  }

  const float bsq = o.normoutlier * o.normoutlier, bsq2 = bsq * 2.0f;

  bool first = true;
  while (true) {
    const bool upd = first ? start_ok : !conv;
    float np0 = p0, np1 = p1, nptx = ptx, npty = pty, ndp0 = dp0, ndp1 = dp1;
    bool nconv = conv;
    int ncnt = cnt;
    if (!first) {
      ncnt = cnt + 1;
      // delta_p = Hes.llt().solve(delta_p) (patch.cpp:175-182)
      if (OF && RCP) {       // tolerance-compared mappings: the four divisions of the solve (a serial chain) as products
        const float y0 = s0 * il00;
        const float y1 = (s1 - l10 * y0) * il11;
        const float x1 = y1 * il11;
        const float x0 = (y0 - l10 * x1) * il00;
        ndp0 = x0; ndp1 = x1;
        np0 = p0 - ndp0; np1 = p1 - ndp1;
      } else if (OF) {
        const float y0 = s0 / l00;
        const float y1 = (s1 - l10 * y0) / l11;
        const float x1 = y1 / l11;
        const float x0 = (y0 - l10 * x1) / l00;
        ndp0 = x0; ndp1 = x1;
        np0 = p0 - ndp0; np1 = p1 - ndp1;
      } else {
        ndp0 = RCP ? (s0 * il00) * il00 : (s0 / l00) / l00;
        np0 = p0 - ndp0;
        np0 = (a.camlr == 0) ? fminf(np0, 0.f) : fmaxf(np0, 0.f);      // patch.cpp:186-191
      }
      nptx = ptrx + np0;
      if (OF) npty = ptry + np1;
      const float ddx = stx - nptx, ddy = sty - npty;
      if ((RCP ? ddx * ddx + ddy * ddy > o.outlierthresh * o.outlierthresh : sqrtf(ddx * ddx + ddy * ddy) > o.outlierthresh) || nptx < g.lb || npty < g.lb || nptx > g.ubw || npty > g.ubh) {
        np0 = p_in0; np1 = p_in1;                                    // patch.cpp:202-205
        nptx = ptrx + np0;
        if (OF) npty = ptry + np1;
        nconv = true;
      }
    }
    // sample position: finished / invalid groups sample their reference position (always in range)
    const float ex = upd ? nptx : ptrx, ey = upd ? npty : ptry;
    // getPatchStaticBil (patch.cpp:346-413)
    const int ix1 = (int)ceilf(ex + .00001f), iy1 = (int)ceilf(ey + .00001f);
    const float rx = ex - floorf(ex), ry = ey - floorf(ey);
    const float w0 = rx * ry, w1 = (1.f - rx) * ry, w2 = rx * (1.f - ry), w3 = (1.f - rx) * (1.f - ry);
    const float* qb = qimg + ((iy1 + g.pad + LB) * g.tw + (ix1 + g.pad + LB)) * C;
    float q[VPL];
    Sum qs;
    if (CP && STAGE) {
      const int my_off = (int)(qb - qimg) - rowstride - 1;       // the window's first element (row -1, column -1)
      // a window is copied again only when its integer origin moved: late iterations step by fractions of a pixel, and
      // finished patches sample a fixed position
      const unsigned moved = STAGE_SKIP ? __ballot_sync(0xffffffffu, my_off != win_off) : 0xffffffffu;
      win_off = my_off;
      if (moved) {
#pragma unroll
        for (int gp = 0; gp < 8; ++gp) {
          if ((moved >> (4 * gp)) & 1u) {
            const float* src = qimg + __shfl_sync(0xffffffffu, my_off, gp * 4);
            float* dst = win + (gp >> 1) * 200 + (gp & 1) * 97 + lane32;
#pragma unroll
            for (int k = 0; k < 3; ++k)
              if (k < 2 || lane32 + 64 < 81) {
                const unsigned sa = (unsigned)__cvta_generic_to_shared(dst + 32 * k);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(src + stage_off[k]) : "memory");
              }
          }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      __syncwarp();
      const float* wp = win + win_base + 2 * lane;
      float u0 = wp[0], u1 = wp[1], u2 = wp[2];
#pragma unroll
      for (int r = 0; r < VPL / 2; ++r) {
        const float l0 = wp[9 * (r + 1)], l1 = wp[9 * (r + 1) + 1], l2 = wp[9 * (r + 1) + 2];
        q[2 * r] = w0 * l1 + w1 * l0 + w2 * u1 + w3 * u0;
        q[2 * r + 1] = w0 * l2 + w1 * l1 + w2 * u2 + w3 * u1;
        qs.add(2 * r, q[2 * r]);
        qs.add(2 * r + 1, q[2 * r + 1]);
        u0 = l0; u1 = l1; u2 = l2;
      }
      __syncwarp();                                              // all taps read before the next evaluation overwrites them
    } else if (CP) {
      // three taps per row serve both columns of the lane; the lower row of one step is the upper row of the next
      const float* p = qb + 2 * lane;
      float u0 = __ldg(p - rowstride - 1), u1 = __ldg(p - rowstride), u2 = __ldg(p - rowstride + 1);
#pragma unroll
      for (int r = 0; r < VPL / 2; ++r) {
        const float l0 = __ldg(p + r * rowstride - 1), l1 = __ldg(p + r * rowstride), l2 = __ldg(p + r * rowstride + 1);
        q[2 * r] = w0 * l1 + w1 * l0 + w2 * u1 + w3 * u0;
        q[2 * r + 1] = w0 * l2 + w1 * l1 + w2 * u2 + w3 * u1;
        qs.add(2 * r, q[2 * r]);
        qs.add(2 * r + 1, q[2 * r + 1]);
        u0 = l0; u1 = l1; u2 = l2;
      }
    } else if (PK && STAGE) {
      // the 13 x 13 tap window of each of the warp's four patches in shared memory, copied only when it moved (op4 runs 128
      // forced iterations, most of them by fractions of a pixel); taps at fixed offsets from one per-lane pointer
      const int my_off = (int)(qb - qimg) - rowstride - 1;
      const unsigned moved = __ballot_sync(0xffffffffu, my_off != win_off);
      win_off = my_off;
      if (moved) {
#pragma unroll
        for (int gp = 0; gp < 4; ++gp) {
          if ((moved >> (8 * gp)) & 1u) {
            const float* src = qimg + __shfl_sync(0xffffffffu, my_off, gp * 8);
            float* dst = win + gp * PKSLOT;
#pragma unroll
            for (int k = 0; k < (PKW * PKW + 31) / 32; ++k) {
              const int e = lane32 + 32 * k;
              if (e < PKW * PKW) {
                const int er = e / PKW;
                const unsigned sa = (unsigned)__cvta_generic_to_shared(dst + e);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(src + er * rowstride + (e - er * PKW)) : "memory");
              }
            }
          }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      __syncwarp();
      const float* wp = win + ((threadIdx.x >> 3) & 3) * PKSLOT + blk_row * PKW + blk_col;   // window row 0 = patch row -1
      const float2 W0 = f2dup(w0), W1 = f2dup(w1), W2 = f2dup(w2), W3 = f2dup(w3);
      float2 u[CW + C], qs2 = make_float2(0.f, 0.f);
#pragma unroll
      for (int i = 0; i < CW + C; ++i) u[i] = make_float2(wp[i], wp[(RPL / 2) * PKW + i]);
#pragma unroll
      for (int r = 0; r < RPL / 2; ++r) {
        float2 l[CW + C];
#pragma unroll
        for (int i = 0; i < CW + C; ++i) l[i] = make_float2(wp[(r + 1) * PKW + i], wp[(r + 1 + RPL / 2) * PKW + i]);
#pragma unroll
        for (int c = 0; c < CW; ++c) {
          const float2 v = f2fma(W0, l[c + C], f2fma(W1, l[c], f2fma(W2, u[c + C], f2mul(W3, u[c]))));
          q[r * CW + c] = v.x;
          q[r * CW + c + HV] = v.y;
          qs2 = f2add(qs2, v);
        }
#pragma unroll
        for (int i = 0; i < CW + C; ++i) u[i] = l[i];
      }
      qs.a0 = qs2.x + qs2.y;
      __syncwarp();                                              // all taps read before the next evaluation overwrites them
    } else if (PK) {
      const float* p = qb + blk_row * rowstride + blk_col - C;   // left tap of the block's first value
      const float2 W0 = f2dup(w0), W1 = f2dup(w1), W2 = f2dup(w2), W3 = f2dup(w3);
      float2 u[CW + C], qs2 = make_float2(0.f, 0.f);
#pragma unroll
      for (int i = 0; i < CW + C; ++i) u[i] = make_float2(__ldg(p - rowstride + i), __ldg(p + (RPL / 2 - 1) * rowstride + i));
#pragma unroll
      for (int r = 0; r < RPL / 2; ++r) {
        float2 l[CW + C];
#pragma unroll
        for (int i = 0; i < CW + C; ++i) l[i] = make_float2(__ldg(p + r * rowstride + i), __ldg(p + (r + RPL / 2) * rowstride + i));
#pragma unroll
        for (int c = 0; c < CW; ++c) {
          const float2 v = f2fma(W0, l[c + C], f2fma(W1, l[c], f2fma(W2, u[c + C], f2mul(W3, u[c]))));
          q[r * CW + c] = v.x;
          q[r * CW + c + HV] = v.y;
          qs2 = f2add(qs2, v);
        }
#pragma unroll
        for (int i = 0; i < CW + C; ++i) u[i] = l[i];
      }
      qs.a0 = qs2.x + qs2.y;
    } else if (BLK && STAGE) {
      // BGR block mappings: the (PSZ + 1) x (ROWLEN + C) tap window of each of the warp's 32 / G patches in shared memory,
      // copied only when it moved; the taps of a lane at fixed offsets from one pointer
      const int my_off = (int)(qb - qimg) - rowstride - C;
      const unsigned moved = __ballot_sync(0xffffffffu, my_off != win_off);
      win_off = my_off;
      if (moved) {
#pragma unroll
        for (int gp = 0; gp < 32 / G; ++gp) {
          if ((moved >> (G * gp)) & 1u) {
            const float* src = qimg + __shfl_sync(0xffffffffu, my_off, gp * G);
            float* dst = win + gp * BSLOT;
#pragma unroll
            for (int k = 0; k < (BWN + 31) / 32; ++k) {
              const int e = lane32 + 32 * k;
              if (e < BWN) {
                const int er = e / BWW, ec = e - er * BWW;
                const unsigned sa = (unsigned)__cvta_generic_to_shared(dst + er * BWWP + ec);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(src + er * rowstride + ec) : "memory");
              }
            }
          }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      __syncwarp();
      const float* wp = win + (lane32 / G) * BSLOT + blk_row * BWWP + blk_col;   // window row 0 = patch row -1, float 0 = column -1
      float u[CW + C];
#pragma unroll
      for (int i = 0; i < CW + C; ++i) u[i] = wp[i];
#pragma unroll
      for (int r = 0; r < RPL; ++r) {
        float l[CW + C];
#pragma unroll
        for (int i = 0; i < CW + C; ++i) l[i] = wp[(r + 1) * BWWP + i];
#pragma unroll
        for (int c = 0; c < CW; ++c) {
          q[r * CW + c] = FM ? fmaf(w0, l[c + C], fmaf(w1, l[c], fmaf(w2, u[c + C], w3 * u[c])))
                             : w0 * l[c + C] + w1 * l[c] + w2 * u[c + C] + w3 * u[c];
          qs.add(r * CW + c, q[r * CW + c]);
        }
#pragma unroll
        for (int i = 0; i < CW + C; ++i) u[i] = l[i];
      }
      __syncwarp();                                              // all taps read before the next evaluation overwrites them
    } else if (BLK) {
      const float* p = qb + blk_row * rowstride + blk_col - C;   // left tap of the block's first value
      float u[CW + C];
#pragma unroll
      for (int i = 0; i < CW + C; ++i) u[i] = __ldg(p - rowstride + i);
#pragma unroll
      for (int r = 0; r < RPL; ++r) {
        float l[CW + C];
#pragma unroll
        for (int i = 0; i < CW + C; ++i) l[i] = __ldg(p + r * rowstride + i);
#pragma unroll
        for (int c = 0; c < CW; ++c) {
          q[r * CW + c] = FM ? fmaf(w0, l[c + C], fmaf(w1, l[c], fmaf(w2, u[c + C], w3 * u[c])))
                             : w0 * l[c + C] + w1 * l[c] + w2 * u[c + C] + w3 * u[c];
          qs.add(r * CW + c, q[r * CW + c]);
        }
#pragma unroll
        for (int i = 0; i < CW + C; ++i) u[i] = l[i];
      }
    } else if (ROWMAP) {
      const float* p = qb + lane;
      float vc = __ldg(p - rowstride), vd = __ldg(p - rowstride - C);
#pragma unroll
      for (int k = 0; k < VPL; ++k) {
        const float va = __ldg(p + k * rowstride), vb = __ldg(p + k * rowstride - C);
        q[k] = w0 * va + w1 * vb + w2 * vc + w3 * vd;
        qs.add(k, q[k]);
        vc = va;
        vd = vb;
      }
    } else {
#pragma unroll
      for (int k = 0; k < VPL; ++k) {
        const float* p = qb + off[k];
        const float va = __ldg(p), vb = __ldg(p - C), vc = __ldg(p - rowstride), vd = __ldg(p - rowstride - C);
        q[k] = val[k] ? w0 * va + w1 * vb + w2 * vc + w3 * vd : 0.f;
        qs.add(k, q[k]);
      }
    }
    float m = 0.f;
    if (o.patnorm > 0 || o.costfct == 10) m = mean_of<NV, RCP>(qs.total());
    Sum as0, as1, aws;
    float npw[VPL];
#define DISB200_RESIDUAL(COST, QV)                                 \
  _Pragma("unroll") for (int k = 0; k < VPL; ++k) {              \
    float vpd, vpw;                                              \
    loss<COST>(QV - t[k], bsq, bsq2, vpd, vpw);                  \
    if (!FULL && !val[k]) { vpd = 0.f; vpw = 0.f; }              \
    npw[k] = vpw;                                                \
    as0.template mad<FM>(k, dx[k], vpd);                         \
    if (OF) as1.template mad<FM>(k, dy[k], vpd);                 \
    aws.add(k, fabsf(vpw));                                      \
  }
    float qscale = 1.0f;
    if (o.costfct == 10) {           // r = q^ * (|t^| / |q^|) - t^
      Sum s2;
#pragma unroll
      for (int k = 0; k < VPL; ++k) s2.add(k, val[k] ? (q[k] - m) * (q[k] - m) : 0.f);
      const float qn = sqrtf(s2.total());
      qscale = qn > 0.f ? tnorm / qn : 0.f;
    }
    if (PK && o.costfct == 0) {
      const float2 M = f2dup(m);
      float2 a0 = make_float2(0.f, 0.f), a1 = a0, aw = a0;
#pragma unroll
      for (int k = 0; k < HV; ++k) {
        const float2 d = f2sub(f2sub(make_float2(q[k], q[k + HV]), M), make_float2(t[k], t[k + HV]));
        const float2 ad = make_float2(fabsf(d.x), fabsf(d.y));
        npw[k] = ad.x;
        npw[k + HV] = ad.y;
        a0 = f2fma(make_float2(dx[k], dx[k + HV]), d, a0);
        if (OF) a1 = f2fma(make_float2(dy[k], dy[k + HV]), d, a1);
        aw = f2add(aw, ad);
      }
      as0.a0 = a0.x + a0.y;
      as1.a0 = a1.x + a1.y;
      aws.a0 = aw.x + aw.y;
    } else if (o.costfct == 0) { DISB200_RESIDUAL(0, (q[k] - m)) }
    else if (o.costfct == 10) { DISB200_RESIDUAL(0, (q[k] - m) * qscale) }
    else if (o.costfct == 1) { DISB200_RESIDUAL(1, (q[k] - m)) }
    else { DISB200_RESIDUAL(2, (q[k] - m)) }
#undef DISB200_RESIDUAL
    const float ns0 = as0.total();
    const float ns1 = OF ? as1.total() : 0.f;
    const float wsum = aws.total();
    if (upd) {
      // OptimizeComputeErrImg bookkeeping (patch.cpp:283-293)
      const float dpsq = OF ? ndp0 * ndp0 + ndp1 * ndp1 : ndp0 * ndp0;
      if (ncnt == 1) dpsq_init = dpsq;
      const float mares_old = mares;
      const float nmares = mean_of<NV, RCP>(wsum);
      // patch.cpp:290-293; the two rate tests (one division each) only matter once min_iter is reached
      bool go_on = (ncnt < o.max_iter) & (nmares > o.res_thresh);
      if (go_on && ncnt >= o.min_iter)
        go_on = (dpsq / dpsq_init >= o.dp_thresh_sq) & (nmares / mares_old <= o.dr_thresh);
      if (!go_on)
        nconv = true;
      mares = nmares;
      p0 = np0; p1 = np1; ptx = nptx; pty = npty; dp0 = ndp0; dp1 = ndp1; cnt = ncnt; conv = nconv;
      s0 = ns0; s1 = ns1;
      if (nconv && active) {   // this evaluation is the patch's last one: its weights are the result
#pragma unroll
        for (int k = 0; k < VPL; ++k)
          if (val[k]) wout[slot_index(k)] = npw[k];
      }
    }
    first = false;
    if (!__any_sync(0xffffffffu, !conv)) break;
  }

  if (active && lane == 0) {
    a.p_iter[pi * NOP] = p0;
    if (OF) a.p_iter[pi * NOP + 1] = p1;
    if (a.pt_iter) { a.pt_iter[pi * 2] = ptx; a.pt_iter[pi * 2 + 1] = pty; }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Any patch size (run-time psz, 2..16) and channel count: one warp per patch, value n of the patch vector in lane
// n % 32, slot n / 32.  This is the reference's arithmetic INCLUDING what it does for odd patch sizes
// (patch.cpp:316-317,377-378: lb = -psz/2, ub = psz/2 - 1 sample PS = 2 (psz/2) rows and columns, while every vector
// has novals = psz^2 C entries, oflow.cpp:91):
//   * entries n >= NE = PS^2 C are never sampled: template and gradients stay 0 there (zero-initialised vectors,
//     SURVEY.md 8c) and the query vector keeps what the PREVIOUS evaluation left in them;
//   * the mean subtraction (patch.cpp:342,412) shifts all novals entries and divides by novals;
//   * LossComputeErrorImage (patch.cpp:226-272) touches the first 4 (novals / 4) entries only: the last novals % 4
//     entries of pdiff stay "query minus mean" and their weights stay 0.
// Used for every (psz, channels) without a specialised instantiation; compared with a tolerance like the other
// kernels whose reduction order differs from Eigen's packets.
template <bool OF, int VPL>
__global__ void __launch_bounds__(128)
patch_search_generic_kernel(PatchArgs a, LevelGeom g, OptConst o) {
  constexpr int NOP = OF ? 2 : 1, G = 32;
  const int C = o.noc, psz = o.psz, NV = o.novals;
  const int lbi = -(psz / 2), PS = 2 * (psz / 2), rowlen = PS * C, NE = PS * rowlen, NB4 = (NV / 4) * 4;
  const int lane = threadIdx.x % G;
  // this device's patches: grid rows gy0 .. gy0+ngy-1 of every grid column (the whole grid unless the pair is tiled)
  const int ip_raw = blockIdx.x * (blockDim.x / G) + threadIdx.x / G;
  const int nloc = g.nopw * g.ngy;
  const bool active = ip_raw < nloc;
  const int ipl = active ? ip_raw : nloc - 1;
  const int gx = ipl / g.ngy, gy = g.gy0 + (ipl - gx * g.ngy);
  const int ip = gx * g.noph + gy;   // patch index is column-major: ip = x * noph + y (patchgrid.cpp:64)
  const int pair = blockIdx.y;
  const int rowstride = g.tw * C;
  const float fnv = (float)NV;

  const float* timg = a.tmpl_img + (size_t)pair * g.img_elems;
  const float* tdxp = a.tmpl_dx + (size_t)pair * g.img_elems;
  const float* tdyp = a.tmpl_dy + (size_t)pair * g.img_elems;
  const float* qimg = a.tgt_img + (size_t)pair * g.img_elems;
  const float ptrx = (float)(gx * g.steps + g.offw), ptry = (float)(gy * g.steps + g.offh);

  int off[VPL];
  bool smp[VPL], val[VPL];   // sampled (n < NE), exists (n < NV)
#pragma unroll
  for (int k = 0; k < VPL; ++k) {
    const int n = k * G + lane;
    val[k] = n < NV;
    smp[k] = n < NE;
    const int nn = smp[k] ? n : 0, row = nn / rowlen;
    off[k] = row * rowstride + (nn - row * rowlen);
  }
  float t[VPL], dx[VPL], dy[VPL], pd[VPL];
  {
    const int base = (((int)ptry + g.pad + lbi) * g.tw + ((int)ptrx + g.pad + lbi)) * C;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      t[k] = smp[k] ? __ldg(timg + base + off[k]) : 0.f;
      dx[k] = smp[k] ? __ldg(tdxp + base + off[k]) : 0.f;
      dy[k] = (OF && smp[k]) ? __ldg(tdyp + base + off[k]) : 0.f;
      pd[k] = 0.f;
      s += t[k];
    }
    if (o.patnorm > 0 || o.costfct == 10) {
      const float m = gsum<G>(s) / fnv;
#pragma unroll
      for (int k = 0; k < VPL; ++k) t[k] = val[k] ? t[k] - m : 0.f;
    }
  }
  float tnorm = 0.f;
  if (o.costfct == 10) {
    float s2 = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) s2 += t[k] * t[k];
    tnorm = sqrtf(gsum<G>(s2));
  }
  float H00 = 0.f, H01 = 0.f, H11 = 0.f;
#pragma unroll
  for (int k = 0; k < VPL; ++k) {
    H00 += dx[k] * dx[k];
    if (OF) { H01 += dx[k] * dy[k]; H11 += dy[k] * dy[k]; }
  }
  H00 = gsum<G>(H00);
  if (OF) { H01 = gsum<G>(H01); H11 = gsum<G>(H11); }
  float l00, l10 = 0.f, l11 = 1.f;
  if (OF) {
    if (H00 * H11 - H01 * H01 == 0.f) { H00 = (float)((double)H00 + 1e-10); H11 = (float)((double)H11 + 1e-10); }
    l00 = H00; l10 = H01; l11 = H11;
    if (l00 > 0.f) {
      l00 = sqrtf(l00);
      l10 = l10 / l00;
      const float x = l11 - l10 * l10;
      if (x > 0.f) l11 = sqrtf(x);
    }
  } else {
    if (H00 == 0.f) H00 = (float)((double)H00 + 1e-10);
    l00 = H00;
    if (l00 > 0.f) l00 = sqrtf(l00);
  }

  float p_in0 = 0.f, p_in1 = 0.f;
  if (a.flow_prev) {
    const int x = (int)floorf(ptrx / 2), y = (int)floorf(ptry / 2);
    const float* fp = a.flow_prev + (size_t)pair * a.flow_prev_stride + (size_t)(y * (g.w / 2) + x) * NOP;
    p_in0 = fp[0] * 2;
    if (OF) p_in1 = fp[1] * 2;
  }
  float p0 = p_in0, p1 = p_in1;
  float ptx = ptrx + p0, pty = OF ? ptry + p1 : ptry;
  const float stx = ptx, sty = pty;
  float dp0 = 0.f, dp1 = 0.f, dpsq_init = 1e-10f, mares = 1e5f;
  int cnt = 0;
  const bool start_ok = !(ptx < g.lb || pty < g.lb || ptx > g.ubw || pty > g.ubh);
  bool conv = !start_ok;
  float s0 = 0.f, s1 = 0.f;
  const size_t pi = (size_t)pair * g.npatch + ip;
  float* wout = a.pweight + pi * NV;
  if (active && !start_ok) {
#pragma unroll
    for (int k = 0; k < VPL; ++k)
      if (val[k]) wout[k * G + lane] = 0.f;
  }
  const float bsq = o.normoutlier * o.normoutlier, bsq2 = bsq * 2.0f;

  bool first = true;
  while (true) {
    const bool upd = first ? start_ok : !conv;
    float np0 = p0, np1 = p1, nptx = ptx, npty = pty, ndp0 = dp0, ndp1 = dp1;
    bool nconv = conv;
    int ncnt = cnt;
    if (!first) {
      ncnt = cnt + 1;
      if (OF) {
        const float y0 = s0 / l00;
        const float y1 = (s1 - l10 * y0) / l11;
        const float x1 = y1 / l11;
        const float x0 = (y0 - l10 * x1) / l00;
        ndp0 = x0; ndp1 = x1;
        np0 = p0 - ndp0; np1 = p1 - ndp1;
      } else {
        ndp0 = (s0 / l00) / l00;
        np0 = p0 - ndp0;
        np0 = (a.camlr == 0) ? fminf(np0, 0.f) : fmaxf(np0, 0.f);
      }
      nptx = ptrx + np0;
      if (OF) npty = ptry + np1;
      const float ddx = stx - nptx, ddy = sty - npty;
      if (sqrtf(ddx * ddx + ddy * ddy) > o.outlierthresh || nptx < g.lb || npty < g.lb || nptx > g.ubw || npty > g.ubh) {
        np0 = p_in0; np1 = p_in1;
        nptx = ptrx + np0;
        if (OF) npty = ptry + np1;
        nconv = true;
      }
    }
    const float ex = upd ? nptx : ptrx, ey = upd ? npty : ptry;
    const int ix1 = (int)ceilf(ex + .00001f), iy1 = (int)ceilf(ey + .00001f);
    const float rx = ex - floorf(ex), ry = ey - floorf(ey);
    const float w0 = rx * ry, w1 = (1.f - rx) * ry, w2 = rx * (1.f - ry), w3 = (1.f - rx) * (1.f - ry);
    const float* qb = qimg + ((iy1 + g.pad + lbi) * g.tw + (ix1 + g.pad + lbi)) * C;
    float q[VPL];
    float qs = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      const float* p = qb + off[k];
      if (smp[k]) {
        const float va = __ldg(p), vb = __ldg(p - C), vc = __ldg(p - rowstride), vd = __ldg(p - rowstride - C);
        q[k] = w0 * va + w1 * vb + w2 * vc + w3 * vd;
      } else {
        q[k] = pd[k];          // never sampled: what the previous evaluation left there (0 before the first, and 0 beyond novals)
      }
      qs += q[k];
    }
    if (o.patnorm > 0 || o.costfct == 10) {
      const float m = gsum<G>(qs) / fnv;
#pragma unroll
      for (int k = 0; k < VPL; ++k) q[k] = val[k] ? q[k] - m : 0.f;
    }
    float qscale = 1.0f;
    if (o.costfct == 10) {
      float s2 = 0.f;
#pragma unroll
      for (int k = 0; k < VPL; ++k) s2 += q[k] * q[k];
      const float qn = sqrtf(gsum<G>(s2));
      qscale = qn > 0.f ? tnorm / qn : 0.f;
    }
    float npd[VPL], npw[VPL];
    float as0 = 0.f, as1 = 0.f, aws = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      const int n = k * G + lane;
      float vpd = q[k], vpw = 0.f;            // tail beyond 4 (novals / 4): pdiff keeps the shifted query, the weight is never written
      if (n < NB4) {
        const float d = (o.costfct == 10 ? q[k] * qscale : q[k]) - t[k];
        if (o.costfct == 0 || o.costfct == 10) loss<0>(d, bsq, bsq2, vpd, vpw);
        else if (o.costfct == 1) loss<1>(d, bsq, bsq2, vpd, vpw);
        else loss<2>(d, bsq, bsq2, vpd, vpw);
      }
      if (!val[k]) { vpd = 0.f; vpw = 0.f; }
      npd[k] = vpd; npw[k] = vpw;
      as0 += dx[k] * vpd;
      if (OF) as1 += dy[k] * vpd;
      aws += fabsf(vpw);
    }
    const float ns0 = gsum<G>(as0);
    const float ns1 = OF ? gsum<G>(as1) : 0.f;
    const float wsum = gsum<G>(aws);
    if (upd) {
      const float dpsq = OF ? ndp0 * ndp0 + ndp1 * ndp1 : ndp0 * ndp0;
      if (ncnt == 1) dpsq_init = dpsq;
      const float mares_old = mares;
      const float nmares = wsum / fnv;
      bool go_on = (ncnt < o.max_iter) & (nmares > o.res_thresh);
      if (go_on && ncnt >= o.min_iter)
        go_on = (dpsq / dpsq_init >= o.dp_thresh_sq) & (nmares / mares_old <= o.dr_thresh);
      if (!go_on) nconv = true;
      mares = nmares;
      p0 = np0; p1 = np1; ptx = nptx; pty = npty; dp0 = ndp0; dp1 = ndp1; cnt = ncnt; conv = nconv;
      s0 = ns0; s1 = ns1;
#pragma unroll
      for (int k = 0; k < VPL; ++k) pd[k] = npd[k];
      if (nconv && active) {
#pragma unroll
        for (int k = 0; k < VPL; ++k)
          if (val[k]) wout[k * G + lane] = npw[k];
      }
    }
    first = false;
    if (!__any_sync(0xffffffffu, !conv)) break;
  }
  if (active && lane == 0) {
    a.p_iter[pi * NOP] = p0;
    if (OF) a.p_iter[pi * NOP + 1] = p1;
    if (a.pt_iter) { a.pt_iter[pi * 2] = ptx; a.pt_iter[pi * 2 + 1] = pty; }
  }
}

template <int VPL>
static void launch_generic(const PatchArgs& a, const LevelGeom& g, const OptConst& o, int n, cudaStream_t s) {
  const int threads = 128, per_block = threads / 32;
  dim3 grd((g.nopw * g.ngy + per_block - 1) / per_block, n);
  if (o.mode == DISB200_MODE_OF) patch_search_generic_kernel<true, VPL><<<grd, threads, 0, s>>>(a, g, o);
  else patch_search_generic_kernel<false, VPL><<<grd, threads, 0, s>>>(a, g, o);
}

template <int PSZ, int C, int G, int LPR = 0, bool FM = false, bool STAGE = false>
static void launch_one(const PatchArgs& a, const LevelGeom& g, const OptConst& o, int n, cudaStream_t s) {
  const int threads = 128, per_block = threads / G;
  dim3 grd((g.nopw * g.ngy + per_block - 1) / per_block, n);
  if (o.mode == DISB200_MODE_OF)
    patch_search_kernel<PSZ, C, G, true, LPR, FM, STAGE><<<grd, threads, 0, s>>>(a, g, o);
  else
    patch_search_kernel<PSZ, C, G, false, LPR, FM, STAGE><<<grd, threads, 0, s>>>(a, g, o);
}

bool launch_patch_search(const PatchArgs& a, const LevelGeom& g, const OptConst& o, int n, cudaStream_t s,
                         LaunchStats* st) {
  if (g.ngy <= 0) return true;   // tiled pair: no patch row of this level belongs to this device
  const int key = o.psz * 10 + o.noc;
  // A/B switch (DisTuning::patch_map = 1): the value-interleaved mapping (value n in lane n % G) wherever the block mapping
  // is the default, and taps read straight from global memory where the staged windows are (8 x 8 grey).  The 12 x 12 kernels (op3 / op4, compared with a tolerance anyway: different reduction order) also
  // contract the interpolation and the two projections into FMAs: op4 588 -> 728 pairs/s.  8 x 8 BGR (op2 of run_*_RGB)
  // keeps separate multiplies and adds: FMA gains 3 % of that kernel there.
  const bool blockmap = o.patch_map != 1;
  switch (key) {
    case 41: launch_one<4, 1, 8>(a, g, o, n, s); break;
    case 81:
      // tap windows staged in shared memory; DISB200_STAGE_MAX_STEP bounds the grid step it is used for
      if (blockmap && g.steps <= DISB200_STAGE_MAX_STEP) launch_one<8, 1, 4, 0, false, true>(a, g, o, n, s);
      else launch_one<8, 1, 4>(a, g, o, n, s);
      break;
    case 121:   // 4 columns x 2 row groups: 6 rows x 3 floats per lane (4 x 1 with 12 rows per lane: 233 registers, 6 % slower)
      if (blockmap) launch_one<12, 1, 8, 4, true, DISB200_STAGE12>(a, g, o, n, s);
      else launch_one<12, 1, 16>(a, g, o, n, s);
      break;
    case 161: launch_one<16, 1, 32>(a, g, o, n, s); break;
    case 43: launch_one<4, 3, 8>(a, g, o, n, s); break;
    case 83:    // 4 columns x 2 row groups: 4 rows x 6 floats per lane
      if (blockmap) launch_one<8, 3, 8, 4, false, DISB200_STAGE_BGR>(a, g, o, n, s);
      else launch_one<8, 3, 32>(a, g, o, n, s);
      break;
    case 123:   // 4 columns x 4 row groups: 3 rows x 9 floats per lane
      if (blockmap) launch_one<12, 3, 16, 4, true, DISB200_STAGE_BGR12>(a, g, o, n, s);
      else launch_one<12, 3, 32>(a, g, o, n, s);
      break;
    case 163: launch_one<16, 3, 32>(a, g, o, n, s); break;
    default: {
      // every other patch size the reference accepts (oflow.cpp:90-91 take any p_samp_s): the run-time-size kernel
      if (o.psz < 2 || o.psz > 16 || (o.noc != 1 && o.noc != 3)) return false;
      const int slots = (o.novals + 31) / 32;
      if (slots <= 4) launch_generic<4>(a, g, o, n, s);
      else if (slots <= 8) launch_generic<8>(a, g, o, n, s);
      else if (slots <= 16) launch_generic<16>(a, g, o, n, s);
      else launch_generic<24>(a, g, o, n, s);
    }
  }
  st->launches++;
  return true;
}

}  // namespace disb200
