// Banded lexicographic SOR for tall levels (more than 96 rows): one warp per band of 32 rows, bands chained
// through tagged boundary rows in global memory (included by varref.cu after sor_warp.cuh).
//
// Same arithmetic and update order as sor_warp.cuh (solver.c:83-471): inside a band, lane = row, the NS sweeps
// are interleaved in the thread and neighbours travel through shuffles and registers.  What changes is the
// schedule.  sor_warp runs sweep s of row r at pixel  i = t - r - 2 s : every dependency is one or two steps
// old, which would force two bands living in different warps into lock-step.  Here consecutive sweeps of a row
// are SIG = 26 steps apart,  i = t - r - SIG s  (t = the band's own step counter), so
//     top    (i, r-1) sweep s    lane-1's result of step t-1                    (shfl up; lane 0: band k-1)
//     left   (i-1, r) sweep s    own result of step t-1
//     right  (i+1, r) sweep s-1  own result of step t-SIG+1                     (delay ring in shared memory)
//     centre (i,   r) sweep s-1  own result of step t-SIG  = last step's right
//     bottom (i, r+1) sweep s-1  lane+1's right of this step                    (shfl down; lane 31: band k+1)
// and the two values that cross a band boundary have slack: band k+1 needs row 31 of band k at k's step
// t+31, band k needs row 0 of band k+1 (previous sweep) at k+1's step t-31-SIG.  Band k+1 may therefore trail
// band k by anything between 32 and 30+SIG steps (it settles 32+LAG behind, leaving SIG-LAG-2 steps the other
// way; both margins exceed the prefetch distance plus an L2 round trip); the bands synchronise by data flow alone:
//   * the last lane of a band publishes its result of every sweep, the first lane its result of every sweep but
//     the last, as 16-byte entries {x, tag, y, tag} (tag = launch epoch; both 8-byte halves carry it, so a torn
//     read cannot pass) — plain stores, no fence, no flag;
//   * the consumer prefetches the entry it needs PF steps ahead with cp.async (same commit groups as the
//     coefficient streams) and checks the tags when it moves the value into registers; only if the producer was
//     late does it poll global memory (volatile 16-byte loads).  A bounded poll count turns a scheduling
//     deadlock into an error flag instead of a hang.
// CTAs (one warp each) take their (pair, band) from an atomic ticket, so a running band implies every lower
// band of the pair is running or done; with at most 64 bands per pair and >= 296 resident CTAs the band a
// consumer waits for always gets a slot.
// Coefficients stay LEADPX ahead of sweep 0 in a ring of RPX pixels per row that still holds the pixel the
// last sweep works on (SIG (NS-1) behind); the sweep-0 bottom of lane 31 is the next band's first row of the
// incoming increment, streamed like an exchange entry; the last sweep stores in place, behind every reader.
#ifndef DISB200_SOR_BAND_CUH_
#define DISB200_SOR_BAND_CUH_

template <int NS>
struct SorbCfg {
  static constexpr int SIG = 26;                          // steps between consecutive sweeps of a row
  static constexpr int RPX = NS == 1 ? 16 : 64;           // coefficient ring, pixels per row (>= SIG (NS-1) + LEADPX + 8)
  static constexpr int DPX = 16;                          // increment ring, pixels per row
  static constexpr int LEADPX = 4, WAITN = 2, PF = 3;     // PF = LEADPX - 1: requested with chunk group tn, usable at tn + PF
  static constexpr int XS = 8;                            // staged exchange entries per stream
  static constexpr int DL = 32;                           // delay ring slots (>= SIG)
  static constexpr int LAG = 10;                          // extra steps a band that caught up falls behind its producer
  static constexpr unsigned OFF_B = 32u * RPX * 16u;
  static constexpr unsigned OFF_DU = 32u * RPX * 32u;
  static constexpr unsigned OFF_DELAY = OFF_DU + 32u * DPX * 8u;
  static constexpr unsigned OFF_XT = OFF_DELAY + (unsigned)(NS - 1) * DL * 32u * 8u;
  static constexpr unsigned OFF_XB = OFF_XT + (unsigned)NS * XS * 16u;   // stream 0: next band's first row of du
  static constexpr unsigned BYTES = OFF_XB + (unsigned)NS * XS * 16u;
  static_assert(NS >= 1 && NS <= 3, "ring sizes cover up to three sweeps");
};
constexpr int SORB_POLL_LIMIT = 1 << 22;

__device__ __forceinline__ float4 sorb_ld_volatile(const float4* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void sorb_sts64(unsigned addr, float2 v) {
  asm volatile("st.shared.v2.f32 [%0], {%1,%2};" ::"r"(addr), "f"(v.x), "f"(v.y) : "memory");
}
__device__ __forceinline__ float2 sorb_lds64(unsigned addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ bool sorb_tag_ok(const float4 v, unsigned epoch) {
  return __float_as_uint(v.y) == epoch && __float_as_uint(v.w) == epoch;
}

// xt: [pair][band][NS][w] entries written by the last row of a band; xb: [pair][band][NS-1 (min 1)][w] entries
// written by its first row.  xt_next_dev / xb_prev_dev: the same arrays on the device that runs band0+nb_local / band0-1
// (spatial tiling of one pair, tiled.cu); equal to xt_all / xb_all when every band runs here.  ctl[0]: ticket counter (monotonic, `ticket_base` = its value before this launch),
// ctl[1]: error flag.  nb = bands per pair = ceil(h / 32).
template <bool OF, int NS>
__global__ void __launch_bounds__(32)
sor_band_kernel(float* __restrict__ du_all, const float4* __restrict__ cA_all, const float4* __restrict__ cB_all,
                float4* xt_all, float4* xb_all, float4* xt_next_dev, float4* xb_prev_dev, unsigned* ctl, unsigned ticket_base,
                unsigned epoch, Dims d, int nb, int band0, int nb_local, float omega) {
  typedef SorbCfg<NS> Cfg;
  extern __shared__ float4 sorb_smem[];
  constexpr int NOP = OF ? 2 : 1, RPX = Cfg::RPX, DPX = Cfg::DPX, LEADPX = Cfg::LEADPX, SIG = Cfg::SIG, PF = Cfg::PF;
  constexpr int NXB = NS > 1 ? NS - 1 : 1;
  const unsigned smem = (unsigned)__cvta_generic_to_shared(sorb_smem);
  const int lane = threadIdx.x;
  unsigned tk = 0;
  if (lane == 0) tk = atomicAdd(ctl, 1u) - ticket_base;
  tk = __shfl_sync(0xffffffffu, tk, 0);
  // this launch runs bands band0 .. band0+nb_local-1 of every pair (all nb of them unless the pair is tiled over GPUs)
  const int pair = (int)tk / nb_local, k = band0 + ((int)tk - pair * nb_local);
  const int w = d.w, h = d.h;
  const unsigned uw = (unsigned)w;
  float* du = du_all + (size_t)pair * d.plane * NOP;
  const float4* cA = cA_all + (size_t)pair * d.plane;
  const float4* cB = cB_all + (size_t)pair * d.plane;
  const int j0 = 32 * k, j = j0 + lane;
  const bool has_prev = k > 0, has_next = j0 + 32 < h;
  // exchange rows: read the previous band's last row / the next band's first row, write this band's own
  const float4* xt_in = xt_all + ((size_t)pair * nb + (has_prev ? k - 1 : 0)) * NS * w;
  const float4* xb_in = xb_all + ((size_t)pair * nb + (has_next ? k + 1 : k)) * NXB * w;
  // A band whose neighbour lives on another GPU publishes straight into THAT device's exchange array (same layout, peer
  // memory over NVLink): plain 16-byte stores again, the tags make them self-validating, the consumer polls its own L2.
  float4* xt_out = (k == band0 + nb_local - 1 ? xt_next_dev : xt_all) + ((size_t)pair * nb + k) * NS * w;
  float4* xb_out = (k == band0 ? xb_prev_dev : xb_all) + ((size_t)pair * nb + k) * NXB * w;
  const float* du_next = du + (size_t)(has_next ? j0 + 32 : 0) * w * NOP;
  const float tagf = __uint_as_float(epoch);

  const int T = w + 31 + SIG * (NS - 1);
  const int up = (lane + 31) & 31, dn = (lane + 1) & 31;
  const int cg = lane >> 3, ce = lane & 7;
  const bool row = j < h;
  const bool mid = j > 0 && j < h - 1;
  int i0 = row ? -lane : -(1 << 28);
  const unsigned rA = smem + lane * (RPX * 16u);
  const unsigned rD = smem + Cfg::OFF_DU + lane * (DPX * 8u);
  const unsigned rL = smem + Cfg::OFF_DELAY + lane * 8u;          // delay ring: [sweep][slot][lane] float2
  float* gD = du + ((size_t)(row ? j : 0) * w - (row ? lane : 0)) * NOP;

  // zero the delay ring and the exchange staging (values read before their first write must be finite)
  for (unsigned o = Cfg::OFF_DELAY + lane * 16u; o < Cfg::BYTES; o += 32u * 16u)
    asm volatile("st.shared.v4.f32 [%0], {%1,%1,%1,%1};" ::"r"(smem + o), "f"(0.f) : "memory");
  __syncwarp();

  float2 R[NS], rgt[NS], ctr[NS], xtop[NS], xbot[NS];
  float hp[NS];
  float4 a[NS], c[NS];
  float2 cur0 = make_float2(0.f, 0.f), nr0 = make_float2(0.f, 0.f);
#pragma unroll
  for (int s = 0; s < NS; ++s) {
    R[s] = rgt[s] = ctr[s] = xtop[s] = xbot[s] = make_float2(0.f, 0.f);
    hp[s] = 0.f;
    a[s] = c[s] = make_float4(0.f, 0.f, 0.f, 1.f);
  }
  bool dead = false;   // a poll ran out: stop waiting, the error flag is set
#ifdef SORB_STATS
  unsigned st_miss = 0, st_polls = 0;
  const long long st_t0 = clock64();
#endif

  auto copy_chunk = [&](int r, int q) {
    const int jj = j0 + r, px = 8 * q + ce;
    if (q >= 0 && jj < h && px < w) {
      const size_t g = (size_t)jj * w + px;
      const unsigned sa = (unsigned)(r * RPX + (px & (RPX - 1))) * 16u;
      sorw_cp16(smem + sa, cA + g);
      sorw_cp16(smem + Cfg::OFF_B + sa, cB + g);
      const unsigned sd = smem + Cfg::OFF_DU + (unsigned)(r * DPX + (px & (DPX - 1))) * 8u;
      if (OF) sorw_cp8(sd, du + g * 2);
      else sorw_cp4(sd, du + g);
    }
  };
  // everything step tn + PF needs from outside the band, requested together with the chunks of step tn
  auto request = [&](int tn, int qmin) {
    const int r = ((tn + LEADPX) & 7) + 8 * cg;
    const int q = (tn - r + LEADPX) >> 3;
    if (q >= qmin) copy_chunk(r, q);
    const unsigned slot = (unsigned)((tn + PF) & (Cfg::XS - 1)) * 16u;
    if (lane == 0 && has_prev) {
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        const int i = tn + PF - SIG * s;
        if ((unsigned)i < uw) sorw_cp16(smem + Cfg::OFF_XT + (unsigned)s * Cfg::XS * 16u + slot, xt_in + (size_t)s * w + i);
      }
    }
    if (lane == 31 && has_next) {
      const int i = tn + PF - 31;
      if ((unsigned)i < uw) {
        if (OF) sorw_cp8(smem + Cfg::OFF_XB + slot, du_next + (size_t)i * 2);
        else sorw_cp4(smem + Cfg::OFF_XB + slot, du_next + i);
      }
#pragma unroll
      for (int s = 1; s < NS; ++s) {
        const int is = i - SIG * s;
        if ((unsigned)is < uw) sorw_cp16(smem + Cfg::OFF_XB + (unsigned)s * Cfg::XS * 16u + slot, xb_in + (size_t)(s - 1) * w + is);
      }
    }
  };
  // staged entry -> value; polls global memory when the producer had not written it at prefetch time
  // `ahead` > 0 (top stream): the consumer caught up with its producer; it first waits for the entry `ahead`
  // pixels further on, which puts it far enough behind for the following prefetches to find their data
  auto take_entry = [&](unsigned saddr, const float4* gsrc, int ahead) {
    float4 v = sorw_lds128(saddr);
    if (!sorb_tag_ok(v, epoch) && !dead) {
      int polls = 0;
#ifdef SORB_STATS
      ++st_miss;
#endif
      if (ahead > 0) {
        while (!sorb_tag_ok(sorb_ld_volatile(gsrc + ahead), epoch) && ++polls < SORB_POLL_LIMIT) {}
      }
      do {
        v = sorb_ld_volatile(gsrc);
      } while (!sorb_tag_ok(v, epoch) && ++polls < SORB_POLL_LIMIT);
      if (polls >= SORB_POLL_LIMIT) { dead = true; atomicExch(ctl + 1, 1u); }
#ifdef SORB_STATS
      st_polls += polls;
#endif
    }
    return make_float2(v.x, v.z);
  };
  // registers of step tn (i0 = its sweep-0 pixel)
  auto fetch = [&](int tn) {
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const int i = i0 - SIG * s;
      if ((unsigned)i < uw) {
        const unsigned sc = rA + ((unsigned)i & (RPX - 1)) * 16u;
        a[s] = sorw_lds128(sc);
        c[s] = sorw_lds128(sc + Cfg::OFF_B);
      }
      if (s > 0) {
        ctr[s] = rgt[s];
        rgt[s] = sorb_lds64(rL + (unsigned)((s - 1) * Cfg::DL + ((tn - SIG + 1) & (Cfg::DL - 1))) * 256u);
      }
    }
    if ((unsigned)(i0 + 1) < uw) nr0 = sorw_lds_du<OF>(rD + ((unsigned)(i0 + 1) & (DPX - 1)) * 8u);
    if (i0 == 0) cur0 = sorw_lds_du<OF>(rD);
    const unsigned slot = (unsigned)(tn & (Cfg::XS - 1)) * 16u;
    if (lane == 0 && has_prev) {
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        const int i = tn - SIG * s;
        if ((unsigned)i < uw)
          xtop[s] = take_entry(smem + Cfg::OFF_XT + (unsigned)s * Cfg::XS * 16u + slot, xt_in + (size_t)s * w + i,
                               min(Cfg::LAG, w - 1 - i));
      }
    }
    if (lane == 31 && has_next) {
      const int i = tn - 31;
      if ((unsigned)i < uw) xbot[0] = sorw_lds_du<OF>(smem + Cfg::OFF_XB + slot);
#pragma unroll
      for (int s = 1; s < NS; ++s) {
        const int is = i - SIG * s;
        if ((unsigned)is < uw)
          xbot[s] = take_entry(smem + Cfg::OFF_XB + (unsigned)s * Cfg::XS * 16u + slot, xb_in + (size_t)(s - 1) * w + is, 0);
      }
    }
  };

  // prologue: chunk 0 of every row and whatever steps 0 .. PF need from the neighbouring bands
#pragma unroll
  for (int rr = 0; rr < 8; ++rr) copy_chunk(rr + 8 * cg, 0);
  for (int tn = -PF; tn <= 0; ++tn) request(tn, 1);
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncwarp();
  fetch(0);

  for (int t = 0; t < T; ++t) {
    float2 top[NS], bot[NS], N[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      top[s] = sorw_shfl<OF>(R[s], up);
      bot[s] = sorw_shfl<OF>(s == 0 ? nr0 : rgt[s], dn);
      if (lane == 0 && has_prev) top[s] = xtop[s];
      if (lane == 31 && has_next) bot[s] = xbot[s];
    }
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const int i = i0 - SIG * s;
      const bool act = (unsigned)i < uw;
      const float2 cur = s == 0 ? cur0 : ctr[s];
      const float2 right = s == 0 ? nr0 : rgt[s];
      float2 nw;
      if (OF) {
        nw = sorw_update_of(a[s], c[s], hp[s], mid, omega, R[s], cur, right, top[s], bot[s]);
      } else {
        nw.x = sorw_update_de(a[s], c[s].x, hp[s], omega, R[s].x, cur.x, right.x, top[s].x, bot[s].x);
        nw.y = 0.f;
      }
      N[s] = act ? nw : R[s];
      hp[s] = a[s].x;
      if (act) {
        if (s == NS - 1) {
          if (OF) *reinterpret_cast<float2*>(gD - 2 * SIG * (NS - 1)) = nw;
          else gD[-SIG * (NS - 1)] = nw.x;
        }
        const float4 e = make_float4(nw.x, tagf, nw.y, tagf);
        if (lane == 31 && has_next) xt_out[(size_t)s * w + i] = e;
        if (s < NS - 1 && lane == 0 && has_prev) xb_out[(size_t)s * w + i] = e;
      }
      if (s < NS - 1) sorb_sts64(rL + (unsigned)(s * Cfg::DL + (t & (Cfg::DL - 1))) * 256u, N[s]);
    }
#pragma unroll
    for (int s = 0; s < NS; ++s) R[s] = N[s];
    if ((unsigned)(i0 + 1) < uw) cur0 = nr0;
    ++i0; gD += NOP;
    request(t + 1, 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group %0;" ::"n"(Cfg::WAITN) : "memory");
    __syncwarp();
    fetch(t + 1);
  }
#ifdef SORB_STATS
  if ((lane == 0 || lane == 31) && pair == 0)
    printf("band %d lane %d: T %d cycles %lld (%.0f / step) misses %u polls %u\n", k, lane, T, clock64() - st_t0,
           (double)(clock64() - st_t0) / T, st_miss, st_polls);
#endif
}

#endif
