// Lean lexicographic SOR for small levels (included by varref.cu, inside its namespaces).
//
// Same update order and fp32 operation order as sor_coupled (solver.c:83-433); the differences
// to the general sor_wavefront are purely mechanical:
//   * the flow increment lives in shared memory with one zero row above and below, so top /
//     bottom neighbours are always readable and the first / last rows need no address
//     predicates (their vertical weights vt / v are 0 by construction);
//   * w % 4 == 0, so a chunk of 4 columns is two aligned 128-bit shared loads per row;
//   * the three row formulas of solver.c (first / middle / last line differ in association
//     order) are folded into one expression with per-thread-constant selects;
//   * pixel 0 has no special case: its left weight and left value are 0 (adds +0).
// One CTA per frame pair, thread (s, r) = sweep s of row r, one barrier per 4 columns.
#ifndef DISB200_SOR_LEAN_CUH_
#define DISB200_SOR_LEAN_CUH_

#ifdef SOR_PROFILE
__device__ long long g_sor_prof[8];
#define SOR_PROF(k) if (blockIdx.x == 0 && threadIdx.x == 0) g_sor_prof[k] = clock64();
#else
#define SOR_PROF(k)
#endif

__device__ __forceinline__ float2 sor_update_lean(const float4 cA, const float4 cB, float hl, bool mid, bool first,
                                                  float omega, float2 left, float2 cur, float2 r, float2 t,
                                                  float2 bt) {
  const float hh = cA.x, vv = cA.y, vt = cA.z, i12 = cA.w, i11 = cB.x, i22 = cB.y, b1 = cB.z, b2 = cB.w;
  // x = h*r, y = vt*t, z = v*bt;  middle: (x + y) + (b + z)   first: (b + x) + z   last: (b + x) + y
  const float x1 = hh * r.x, x2 = hh * r.y, y1 = vt * t.x, y2 = vt * t.y, z1 = vv * bt.x, z2 = vv * bt.y;
  const float p1 = mid ? x1 : b1, p2 = mid ? x2 : b2;
  const float q1 = mid ? y1 : x1, q2 = mid ? y2 : x2;
  float s1, s2;
  if (mid) {
    s1 = (p1 + q1) + (b1 + z1);
    s2 = (p2 + q2) + (b2 + z2);
  } else {
    s1 = (p1 + q1) + (first ? z1 : y1);
    s2 = (p2 + q2) + (first ? z2 : y2);
  }
  const float B1 = hl * left.x + s1, B2 = hl * left.y + s2;
  float2 nw;
  nw.x = cur.x + omega * (i11 * B1 + i12 * B2 - cur.x);
  nw.y = cur.y + omega * (i12 * B1 + i22 * B2 - cur.y);
  return nw;
}

// du_all: [pair][plane][2]; cA/cB: [pair][plane] float4 (see system_pixel).  blockDim.x >= h * iters.
// Dynamic shared memory: ((h + 2) * w + 4) * sizeof(float2).
__global__ void __launch_bounds__(512)
sor_lean_kernel(float* __restrict__ du_all, const float4* __restrict__ cA_all, const float4* __restrict__ cB_all,
                Dims d, int iters, float omega) {
  extern __shared__ float4 lean_smem[];
  float2* sdu = reinterpret_cast<float2*>(lean_smem);
  const int pair = blockIdx.x, w = d.w, h = d.h, tid = threadIdx.x, T = blockDim.x;
  float2* gdu = reinterpret_cast<float2*>(du_all) + (size_t)pair * d.plane;
  const float4* cA = cA_all + (size_t)pair * d.plane;
  const float4* cB = cB_all + (size_t)pair * d.plane;
  const int npix = w * h;
  for (int k = tid; k < w + 4; k += T) {            // zero rows above and below (+ 4 slack elements)
    if (k < w) sdu[k] = make_float2(0.f, 0.f);
    sdu[(h + 1) * w + k] = make_float2(0.f, 0.f);
  }
  for (int k = tid; k < npix; k += T) sdu[w + k] = gdu[k];

  const int s = tid / h, r = tid - s * h;
  const bool act = s < iters;
  const int off = r + 2 * s;
  const int nchunks = w >> 2;
  const int nsteps = nchunks + h - 1 + 2 * (iters - 1);
  const bool mid = r > 0 && r < h - 1, first = r == 0;
  const int rowoff = (act ? r : 0) * w;
  const float4* pa = cA + rowoff;
  const float4* pb = cB + rowoff;
  float2* rowp = sdu + w + rowoff;
  float4 ra[2][4], rb[2][4];
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int c = ((u - off % 2) + 2) % 2;   // the one chunk in {0, 1} that is consumed from slot u
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      ra[u][k] = make_float4(0.f, 0.f, 0.f, 0.f);
      rb[u][k] = ra[u][k];
      if (act && c < nchunks) { ra[u][k] = __ldg(pa + c * 4 + k); rb[u][k] = __ldg(pb + c * 4 + k); }
    }
  }
  float2 left = make_float2(0.f, 0.f);
  float hprev = 0.f;
  SOR_PROF(0)
  __syncthreads();
  SOR_PROF(1)
  for (int m0 = 0; m0 < nsteps; m0 += 2) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int c = m0 + u - off;
      if (act && c >= 0 && c < nchunks) {
        float2* p = rowp + c * 4;
        const float4 o01 = *reinterpret_cast<const float4*>(p), o23 = *reinterpret_cast<const float4*>(p + 2);
        const float2 o4 = p[4];
        const float4 t01 = *reinterpret_cast<const float4*>(p - w), t23 = *reinterpret_cast<const float4*>(p - w + 2);
        const float4 b01 = *reinterpret_cast<const float4*>(p + w), b23 = *reinterpret_cast<const float4*>(p + w + 2);
        const float2 own[5] = {make_float2(o01.x, o01.y), make_float2(o01.z, o01.w), make_float2(o23.x, o23.y),
                               make_float2(o23.z, o23.w), o4};
        const float2 top[4] = {make_float2(t01.x, t01.y), make_float2(t01.z, t01.w), make_float2(t23.x, t23.y),
                               make_float2(t23.z, t23.w)};
        const float2 bot[4] = {make_float2(b01.x, b01.y), make_float2(b01.z, b01.w), make_float2(b23.x, b23.y),
                               make_float2(b23.z, b23.w)};
        float2 nw[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          nw[k] = sor_update_lean(ra[u][k], rb[u][k], hprev, mid, first, omega, left, own[k], own[k + 1], top[k], bot[k]);
          left = nw[k];
          hprev = ra[u][k].x;
        }
        *reinterpret_cast<float4*>(p) = make_float4(nw[0].x, nw[0].y, nw[1].x, nw[1].y);
        *reinterpret_cast<float4*>(p + 2) = make_float4(nw[2].x, nw[2].y, nw[3].x, nw[3].y);
        if (c + 2 < nchunks) {   // refill after the last use: the loads target the ring registers directly
#pragma unroll
          for (int k = 0; k < 4; ++k) { ra[u][k] = __ldg(pa + (c + 2) * 4 + k); rb[u][k] = __ldg(pb + (c + 2) * 4 + k); }
        }
      }
      __syncthreads();
    }
  }
  SOR_PROF(2)
  for (int k = tid; k < npix; k += T) gdu[k] = sdu[w + k];
  SOR_PROF(3)
}

#endif
