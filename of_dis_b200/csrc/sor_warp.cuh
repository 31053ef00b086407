// Single-warp lexicographic SOR for levels of up to 32 * NB rows, NB <= 3 (included by varref.cu, inside its
// namespaces).
//
// Same update order and fp32 operation order as sor_coupled / sor_coupled_slow_but_readable_DE
// (solver.c:83-471).  Pixel (i, j) of sweep s needs (i-1, j), (i, j-1) of sweep s and (i+1, j),
// (i, j+1) of sweep s-1, so at step t sweep s may process pixel  i = t - j - 2 s  of row j: every
// dependency was produced at step t-1 or t-2.
//
// Mapping: ONE warp per frame pair, lane = row (NB bands: rows `lane`, `lane + 32`, ...), and the
// NS sweeps are interleaved inside each thread, so one step is NB x NS independent pixel updates
// per lane and nothing ever crosses a warp: the values a pixel needs from its neighbours at
// step t are exactly the results the neighbouring lanes (or this lane's previous sweep)
// produced at step t-1 / t-2, so they travel through warp shuffles and registers — no barrier,
// no shared-memory round trip inside a step:
//     left   (i-1, j  ) sweep s    own result of sweep s        at step t-1
//     top    (i,   j-1) sweep s    lane-1's result of sweep s   at step t-1        (shfl up)
//     right  (i+1, j  ) sweep s-1  own result of sweep s-1      at step t-1
//     centre (i,   j  ) sweep s-1  own result of sweep s-1      at step t-2
//     bottom (i,   j+1) sweep s-1  lane+1's result of sweep s-1 at step t-1        (shfl down)
// Everything a row reads from memory — the two coefficient float4 of system_pixel and, for
// sweep 0, the increment left by the previous fixed-point iteration — is a sequential stream
// along the row that does not depend on the wavefront: each lane streams its row from global
// memory into a private shared ring with cp.async, SORW_D pixels ahead, and moves the values of
// the next step into registers one step early.  The last sweep stores the result in place,
// always behind every location still to be read.  A warp needs 10-40 KB of shared memory and
// no CTA-wide resource, so the SOR of thousands of frame pairs is resident at once and the
// schedulers interleave them; throughput, not the dependency chain of one pair, sets the time.
#ifndef DISB200_SOR_WARP_CUH_
#define DISB200_SOR_WARP_CUH_

// Row streams are cut into chunks of CH pixels.  Chunk q of a row is requested when the row's sweep-0 pixel reaches
// CH q - LEADPX; at every step the rows that reach such a point are j = (t + LEADPX) mod CH + CH g, 32 / CH per band, so ONE
// cp.async per stream serves them all: lane = CH g + e copies pixel CH q + e of row j — fully used 128-byte segments
// per instruction instead of 32 scattered sectors.  The owner lane of a row reads its pixels from the row's ring of RPX
// pixels one step before use.  The ring must hold the pixels between the last sweep (2 (NS - 1) behind sweep 0) and the
// chunk being filled (LEADPX + CH ahead).  Shared memory per pair bounds the resident warps (6 per SM at 56 rows with
// 16-pixel rings).  A 12-pixel ring of three 4-pixel chunks (-DDISB200_SORW_RING12: 27 KB, 8 pairs per SM) was measured:
// the SOR stage gains 5 % at batches that fill 8 per SM (2368 pairs: 255.5 k against 251.9 k pairs/s) and LOSES 9 % at the
// 1776-pair batch (twice as many 64-byte row segments per step); not adopted (profiles/r02i_ring.txt).
template <int NS>
struct SorwCfg {
  static constexpr int H = 2 * (NS - 1);             // pixels the last sweep trails sweep 0
#ifdef DISB200_SORW_RING12                           // A/B build: 12-pixel ring, three 4-pixel chunks (27 KB at 56 rows, 8 pairs / SM)
  static constexpr int CH = NS <= 3 ? 4 : 8;         // pixels per chunk
  static constexpr int CHS = NS <= 3 ? 2 : 3;        // log2(CH)
  static constexpr int NCH = NS <= 3 ? 3 : 4;        // chunk slots per row
#else
  static constexpr int CH = 8, CHS = 3;              // 16-pixel ring, two 8-pixel chunks (35 KB at 56 rows, 6 pairs / SM)
  static constexpr int NCH = NS <= 3 ? 2 : 4;
#endif
  static constexpr int RPX = NCH * CH;               // ring pixels per row (12: not a power of two, see sorw_ring)
  static constexpr int LEADPX = NS <= 3 ? 4 : 8;     // request distance in pixels
  static constexpr int WAITN = LEADPX - 2;           // cp.async groups that may stay in flight
  static constexpr int ROW_BYTES = RPX * (16 + 16 + 8);   // one row's rings: cA, cB (float4) and the increment
  static_assert(H + LEADPX + CH <= RPX, "ring too short for the sweeps in flight");
};
// ring slot of pixel i (i >= 0)
template <int RPX>
__device__ __forceinline__ unsigned sorw_ring(unsigned i) {
  if ((RPX & (RPX - 1)) == 0) return i & (RPX - 1);
  return i - RPX * __umulhi(i, (unsigned)(0x100000000ull / RPX) + 1u);   // i mod RPX, exact for i < 2^31 / RPX
}
// Rings exist for the h rows of the level only (not for whole 32-row bands): three regions, [h][RPX] each.
__host__ __device__ inline size_t sorw_smem_bytes(int h, int ns) {
  return (size_t)h * (ns <= 3 ? SorwCfg<3>::ROW_BYTES : SorwCfg<5>::ROW_BYTES);
}

__device__ __forceinline__ void sorw_cp16(unsigned smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void sorw_cp8(unsigned smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void sorw_cp4(unsigned smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ float4 sorw_lds128(unsigned addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
template <bool OF>
__device__ __forceinline__ float2 sorw_lds_du(unsigned addr) {
  float2 v = make_float2(0.f, 0.f);
  if (OF) asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  else asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v.x) : "r"(addr));
  return v;
}

// One OF update, branch-free (selects on the per-lane row type); same values as the expression trees of
// solver.c:122-141 (first row), :184-203 (middle rows), :244-262 (last row).
__device__ __forceinline__ float2 sorw_update_of(const float4 cA, const float4 cB, float hl, bool mid, float omega, float2 left, float2 cur, float2 r, float2 t, float2 bt) {
  const float hh = cA.x, vv = cA.y, vt = cA.z, i12 = cA.w, i11 = cB.x, i22 = cB.y, b1 = cB.z, b2 = cB.w;
  // x = h*r, y = vt*t, z = v*bt;  middle: (x + y) + (b + z)   first: (b + x) + z   last: (b + x) + y
  const float x1 = hh * r.x, x2 = hh * r.y, y1 = vt * t.x, y2 = vt * t.y, z1 = vv * bt.x, z2 = vv * bt.y;
  // (b + x) == (x + b) bit for bit, and in the first / last row the absent term (y / z) is +-0 because its
  // weight is 0, so  first: (x + b) + (y + z) == (b + x) + z  and  last: (x + b) + (y + z) == (b + x) + y
  const float q1 = mid ? y1 : b1, q2 = mid ? y2 : b2;
  const float e1 = (mid ? b1 : y1) + z1, e2 = (mid ? b2 : y2) + z2;
  const float s1 = (x1 + q1) + e1, s2 = (x2 + q2) + e2;
  const float B1 = hl * left.x + s1, B2 = hl * left.y + s2;
  float2 nw;
  nw.x = cur.x + omega * (i11 * B1 + i12 * B2 - cur.x);
  nw.y = cur.y + omega * (i12 * B1 + i22 * B2 - cur.y);
  return nw;
}
// One DE update (solver.c:446-467): sigma accumulates top, left, bottom, right; absent neighbours
// have weight 0 and a finite value here, and x - 0*y == x bit for bit (sigma is never -0).
__device__ __forceinline__ float sorw_update_de(const float4 cA, float b1, float hl, float omega, float left, float cur,
                                                float r, float t, float bt) {
  float sigma = 0.f;
  sigma -= cA.z * t;
  sigma -= hl * left;
  sigma -= cA.y * bt;
  sigma -= cA.x * r;
  const float B1 = b1 - sigma;
  return (1.0f - omega) * cur + omega * (B1 / cA.w);
}

template <bool OF>
__device__ __forceinline__ float2 sorw_shfl(float2 v, int src) {
  float2 r;
  r.x = __shfl_sync(0xffffffffu, v.x, src);
  r.y = OF ? __shfl_sync(0xffffffffu, v.y, src) : 0.f;
  return r;
}

// du: global increment plane of this pair, [h][w] float2 (OF) / float (DE), updated in place.
// smem: sorw_smem_bytes(h, NS) bytes owned by this warp.  cA / cB: global coefficient planes of
// this pair (row stride w).  Called by exactly one full warp; h <= 32 * NB.
//
// FUSED (sor_fused.cuh): the rings are not streamed from global memory by this warp but FILLED by producer warps of the
// same CTA, which evaluate the linear system of each 8-pixel chunk (system_pixel) straight into its ring slots - the
// coefficient planes never exist in memory.  sync[0] = number of fetch phases this warp has finished, sync[1 + b] = last
// production step band b's producer has finished; production step tn holds exactly what request(tn) would have copied.
//
// TEAM (sor_team_kernel): the NBT = NB bands of the pair are run by NBT WARPS of one CTA, warp b = band b, in lock-step - same
// step counter, same registers per band, same order of operations as the one-warp schedule, so the result is the same
// bits.  What the one-warp schedule moves between bands with a shuffle inside the warp (row 32 b - 1 to the first lane of
// band b, row 32 (b + 1) to the last lane of band b) goes through a double-buffered mailbox in shared memory, one CTA
// barrier per step.  A step is NS updates instead of NB x NS, and an SM holds NB times as many warps for the same
// shared memory: twice to three times the warps to hide the latency of the dependent chain with.
template <bool OF, int NS, int NBT, bool FUSED = false, bool TEAM = false>
__device__ __forceinline__ void sor_warp(float* __restrict__ du, unsigned smem, const float4* __restrict__ cA,
                                         const float4* __restrict__ cB, int w, int h, float omega,
                                         volatile int* sync = nullptr, float2* mbox = nullptr) {
  typedef SorwCfg<NS> Cfg;
  constexpr int NOP = OF ? 2 : 1, RPX = Cfg::RPX, LEADPX = Cfg::LEADPX, CH = Cfg::CH, CHS = Cfg::CHS;
  constexpr int NB = TEAM ? 1 : NBT;                       // bands of THIS warp
  const int b0 = TEAM ? (int)(threadIdx.x >> 5) : 0;      // first band of this warp
  // mailbox [parity][band][top | bottom][sweep]
  auto mb = [&](int par, int band, int which, int sw) -> float2* { return mbox + ((par * NBT + band) * 2 + which) * NS + sw; };
  const int lane = threadIdx.x & 31;
  const unsigned uw = (unsigned)w;
  const unsigned offB = (unsigned)h * (RPX * 16u);       // cB region; the increment region starts at 2 offB
  const int T = w + h - 1 + 2 * (NS - 1);
  const int up = (lane + 31) & 31, dn = (lane + 1) & 31;
  const int cg = lane >> CHS, ce = lane & (CH - 1);   // cooperative copies: row group, pixel in chunk

  int i0[NB];                  // pixel of sweep 0 at the current step (row j = b * 32 + lane)
  bool mid[NB];
  unsigned rA[NB], rD[NB];     // shared addresses of the rings of this lane's own row: cA (cB at + 32*RPX*16), du
  float* gD[NB];               // &du[j][i0]
  float2 R[NB][NS], R2[NB][NS];
  float hp[NB][NS];
  float4 a[NB][NS], c[NB][NS];
  float2 cur0[NB], nr0[NB];
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    const int j = (b0 + b) * 32 + lane;
    const bool row = j < h;
    i0[b] = row ? -j : -(1 << 28);
    mid[b] = j > 0 && j < h - 1;
    rA[b] = smem + (unsigned)j * (RPX * 16u);            // rows >= h own no ring and never touch one
    rD[b] = smem + 2u * offB + (unsigned)j * (RPX * 8u);
    const int jr = row ? j : 0;
    gD[b] = du + ((size_t)jr * w - jr) * NOP;
    cur0[b] = nr0[b] = make_float2(0.f, 0.f);
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      R[b][s] = R2[b][s] = make_float2(0.f, 0.f);
      hp[b][s] = 0.f;
      a[b][s] = c[b][s] = make_float4(0.f, 0.f, 0.f, 1.f);
    }
  }
  // cooperative copy of chunk q of row `r` (of band b) — this lane moves pixel 8 q + ce
  auto copy_chunk = [&](int b, int r, int q) {
    const int j = (b0 + b) * 32 + r, px = CH * q + ce;
    if (q >= 0 && j < h && px < w) {
      const size_t g = (size_t)j * w + px;
      const unsigned slot = (unsigned)(j * RPX) + sorw_ring<RPX>((unsigned)px);
      sorw_cp16(smem + slot * 16u, cA + g);
      sorw_cp16(smem + offB + slot * 16u, cB + g);
      if (OF) sorw_cp8(smem + 2u * offB + slot * 8u, du + g * 2);
      else sorw_cp4(smem + 2u * offB + slot * 8u, du + g);
    }
  };
  // rows whose sweep-0 pixel is at a request point when the step counter is tn
  auto request = [&](int tn, int qmin) {
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const int r = ((tn + LEADPX) & (CH - 1)) + CH * cg;
      const int q = (tn - ((b0 + b) * 32 + r) + LEADPX) >> CHS;
      if (q >= qmin) copy_chunk(b, r, q);
    }
  };
  auto fetch = [&](int b) {                // registers of the step whose sweep-0 pixel is i0[b]
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const int i = i0[b] - 2 * s;
      if ((unsigned)i < uw) {
        const unsigned sc = rA[b] + sorw_ring<RPX>((unsigned)i) * 16u;
        a[b][s] = sorw_lds128(sc);
        c[b][s] = sorw_lds128(sc + offB);
      }
    }
    // sweep 0: right neighbour (j, i0 + 1); pixel 0 also seeds the centre carry
    if ((unsigned)(i0[b] + 1) < uw) nr0[b] = sorw_lds_du<OF>(rD[b] + sorw_ring<RPX>((unsigned)(i0[b] + 1)) * 8u);
    if (i0[b] == 0) cur0[b] = sorw_lds_du<OF>(rD[b]);
  };
  // the fetch that prepares step t1 reads what production steps <= t1 - 3 delivered (first pixel of a chunk: the right
  // neighbour of sweep 0, LEADPX - 1 steps after its request)
  auto wait_produced = [&](int t1) {
    if (FUSED) {
#pragma unroll
      for (int b = 0; b < NB; ++b)
        while (sync[1 + b] < t1 - 3) __nanosleep(64);
      __threadfence_block();
    }
  };
  if (FUSED) {
    wait_produced(0);
  } else {
    // prologue: chunk 0 of every row, plus whatever is due at step 0
#pragma unroll
    for (int rr = 0; rr < CH; ++rr) {
#pragma unroll
      for (int b = 0; b < NB; ++b) copy_chunk(b, rr + CH * cg, 0);
    }
    request(0, 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
  }
#pragma unroll
  for (int b = 0; b < NB; ++b) fetch(b);
  if (FUSED) { __syncwarp(); if (lane == 0) { __threadfence_block(); sync[0] = 0; } }
  // TEAM: offer the rows the neighbouring bands read at the next step (the values the one-warp schedule shuffles)
  auto post = [&](int par) {
    if (TEAM) {
      if (lane == 31 && b0 + 1 < NBT) {
#pragma unroll
        for (int s = 0; s < NS; ++s) *mb(par, b0 + 1, 0, s) = R[0][s];
      }
      if (lane == 0 && b0 > 0) {
#pragma unroll
        for (int s = 0; s < NS; ++s) *mb(par, b0 - 1, 1, s) = s > 0 ? R[0][s > 0 ? s - 1 : 0] : nr0[0];
      }
      __syncthreads();
    }
  };
  post(0);

#pragma unroll 2
  for (int t = 0; t < T; ++t) {
    // ---- neighbour exchange: results of step t-1 (sweep 0: the streamed increment) ----
    float2 top[NB][NS], bot[NB][NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) {
#pragma unroll
      for (int b = 0; b < NB; ++b) {
        // top of row b*32+lane: lane-1 of the same band; lane 0 of band b takes the last row of band b-1
        // (lane 31 of band b-1 is read by nobody else in this shuffle, so it offers that value instead)
        float2 st = R[b][s];
        if (b > 0 && lane == 31) st = R[b > 0 ? b - 1 : 0][s];
        top[b][s] = sorw_shfl<OF>(st, up);
        // bottom: lane+1's previous sweep (sweep 0: its right neighbour, which is (i0, j+1))
        float2 sb = s > 0 ? R[b][s - 1] : nr0[b];
        if (b + 1 < NB && lane == 0) sb = s > 0 ? R[b + 1 < NB ? b + 1 : b][s - 1] : nr0[b + 1 < NB ? b + 1 : b];
        bot[b][s] = sorw_shfl<OF>(sb, dn);
        if (TEAM) {
          if (lane == 0 && b0 > 0) top[b][s] = *mb(t & 1, b0, 0, s);
          if (lane == 31 && b0 + 1 < NBT) bot[b][s] = *mb(t & 1, b0, 1, s);
        }
      }
    }
    // ---- updates ----
    float2 N[NB][NS];
#pragma unroll
    for (int b = 0; b < NB; ++b) {
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        const int i = i0[b] - 2 * s;
        const bool act = (unsigned)i < uw;
        const float2 cur = s == 0 ? cur0[b] : R2[b][s - 1];
        const float2 right = s == 0 ? nr0[b] : R[b][s - 1];
        float2 nw;
        if (OF) {
          nw = sorw_update_of(a[b][s], c[b][s], hp[b][s], mid[b], omega, R[b][s], cur, right, top[b][s],
                              bot[b][s]);
        } else {
          nw.x = sorw_update_de(a[b][s], c[b][s].x, hp[b][s], omega, R[b][s].x, cur.x, right.x, top[b][s].x,
                                bot[b][s].x);
          nw.y = 0.f;
        }
        N[b][s] = act ? nw : R[b][s];
        hp[b][s] = a[b][s].x;   // stale (zero before the row starts) when the pixel is out of range: multiplies a zero / unused left
        if (s == NS - 1 && act) {
          if (OF) *reinterpret_cast<float2*>(gD[b] - 4 * (NS - 1)) = nw;
          else gD[b][-2 * (NS - 1)] = nw.x;
        }
      }
    }
    // ---- rotate, advance, stream pixel i0 + D, fetch the registers of step t+1 ----
#pragma unroll
    for (int b = 0; b < NB; ++b) {
#pragma unroll
      for (int s = 0; s < NS; ++s) { R2[b][s] = R[b][s]; R[b][s] = N[b][s]; }
      if ((unsigned)(i0[b] + 1) < uw) cur0[b] = nr0[b];   // centre of the next pixel = this step's right neighbour
      ++i0[b]; gD[b] += NOP;
    }
    if (FUSED) {
      wait_produced(t + 1);
    } else {
      request(t + 1, 1);
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group %0;" ::"n"(Cfg::WAITN) : "memory");
      __syncwarp();
    }
#pragma unroll
    for (int b = 0; b < NB; ++b) fetch(b);
    if (FUSED) { __syncwarp(); if (lane == 0) { __threadfence_block(); sync[0] = t + 1; } }   // ring slots read by this fetch are free from now on
    post((t + 1) & 1);
  }
}

// mailbox bytes of the team schedule (behind the rings)
__host__ __device__ inline size_t sorw_team_bytes(int ns) { return (size_t)2 * 3 * 2 * ns * sizeof(float2); }

// One CTA of NB warps per frame pair, warp b = band b (TEAM above).
template <bool OF, int NS, int NB>
__global__ void __launch_bounds__(32 * NB)
sor_team_kernel(float* __restrict__ du_all, const float4* __restrict__ cA_all, const float4* __restrict__ cB_all, Dims d,
                float omega, unsigned mbox_off) {
  extern __shared__ float4 sorw_smem[];
  constexpr int NOP = OF ? 2 : 1;
  const int pair = blockIdx.x;
  sor_warp<OF, NS, NB, false, true>(du_all + (size_t)pair * d.plane * NOP, (unsigned)__cvta_generic_to_shared(sorw_smem),
                                    cA_all + (size_t)pair * d.plane, cB_all + (size_t)pair * d.plane, d.w, d.h, omega, nullptr,
                                    reinterpret_cast<float2*>(reinterpret_cast<char*>(sorw_smem) + mbox_off));
}

// One warp (= one CTA) per frame pair.
template <bool OF, int NS, int NB>
__global__ void __launch_bounds__(32)
sor_warp_kernel(float* __restrict__ du_all, const float4* __restrict__ cA_all, const float4* __restrict__ cB_all, Dims d,
                float omega) {
  extern __shared__ float4 sorw_smem[];
  constexpr int NOP = OF ? 2 : 1;
  const int pair = blockIdx.x;
  sor_warp<OF, NS, NB>(du_all + (size_t)pair * d.plane * NOP, (unsigned)__cvta_generic_to_shared(sorw_smem),
                       cA_all + (size_t)pair * d.plane, cB_all + (size_t)pair * d.plane, d.w, d.h, omega);
}

#endif
