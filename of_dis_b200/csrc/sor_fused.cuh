// The linear system assembled INSIDE the SOR kernel (included by varref.cu after sor_warp.cuh).
//
// One CTA per frame pair: warp 0 is sor_warp's consumer (the lexicographic wavefront, unchanged), warps 1 .. NB are
// producers, one per band of 32 rows.  Production step tn of band b covers the four rows r = (tn + LEADPX) mod 8 + 8 g and
// chunk q = (tn - row + LEADPX) / 8 of each - the 32 pixels the stand-alone kernel would have requested from the
// coefficient planes at step tn (sor_warp.cuh) - and evaluates system_pixel for them (data term, laplacian of the
// current flow, inverse of the 2x2 block: compute_data + sub_laplacian + sor_coupled's first-sweep inversion,
// opticalflow_aux.c:194-223,409-747, solver.c:122-128) straight into the rows' shared-memory rings, together with the
// pixel's current increment.  The two float4 coefficient packs per pixel are never written to or read from memory:
// per pixel and fixed-point iteration the refinement moves 64 bytes less through HBM (32 written by system_kernel, 32
// read back by the SOR) and one launch fewer.  Same values, same order: bit-identical to the split path.
//
// Hand-shake (shared memory, CTA scope): the consumer publishes the number of fetch phases it has finished (ring slots a
// fetch has read are free afterwards), a producer publishes the last production step it has finished.  A producer may
// run ahead by the ring depth (it waits before it reuses a slot); the consumer waits only if a producer is late.
#ifndef DISB200_SOR_FUSED_CUH_
#define DISB200_SOR_FUSED_CUH_

// Registers: the consumer loop alone needs 126 (sor_warp_kernel); 128 keeps five CTAs of three warps per SM resident
// at level 3 of the C1 workload (two bands), without spilling inside the wavefront loop.
#ifndef DISB200_SORF_REGS
#define DISB200_SORF_REGS 128
#endif
template <int C, bool OF, int NS, int NB>
__global__ void __maxnreg__(DISB200_SORF_REGS)
tv_solve_fused_kernel(DerivPtrs D, const float* __restrict__ mask_all, const float* __restrict__ s_all,
                      const float* __restrict__ flow_all, float* __restrict__ du_all, Dims d, float hd3, float hg3, float omega) {
  typedef SorwCfg<NS> Cfg;
  constexpr int NOP = OF ? 2 : 1, RPX = Cfg::RPX, LEADPX = Cfg::LEADPX, NCH = Cfg::NCH, CH = Cfg::CH, CHS = Cfg::CHS;
  extern __shared__ float4 sorf_smem[];
  const int pair = blockIdx.x, w = d.w, h = d.h, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned smem = (unsigned)__cvta_generic_to_shared(sorf_smem);
  const size_t ring_bytes = sorw_smem_bytes(h, NS);
  volatile int* sync = reinterpret_cast<volatile int*>(reinterpret_cast<char*>(sorf_smem) + ring_bytes);
  if (threadIdx.x <= NB) sync[threadIdx.x] = -(1 << 20);
  __syncthreads();
  float* du = du_all + (size_t)pair * d.plane * NOP;
  if (warp == 0) {
    sor_warp<OF, NS, NB, true>(du, smem, nullptr, nullptr, w, h, omega, sync);
    return;
  }
  // ---- producer of band `band` ----
  const int band = warp - 1;
  const int cg = lane >> CHS, ce = lane & (CH - 1);
  const float* mask = mask_all + (size_t)pair * d.plane;
  const float* sp = s_all + (size_t)pair * d.plane;
  const float* fl = flow_all + (size_t)pair * w * h * NOP;
  const size_t o = (size_t)pair * C * d.plane;
  const unsigned offB = (unsigned)h * (RPX * 16u);
  const int T = w + h - 1 + 2 * (NS - 1);
  for (int tn = -LEADPX; tn <= T - 3; ++tn) {
    const int r = ((tn + LEADPX) & (CH - 1)) + CH * cg;
    const int j = band * 32 + r, q = (tn - j + LEADPX) >> CHS, px = CH * q + ce;
    const bool valid = q >= 0 && j < h && px < w;
    if (__any_sync(0xffffffffu, valid)) {
      // a chunk beyond the ring depth overwrites the slot of chunk q - NCH: wait until the consumer's last sweep has read it
      if (__any_sync(0xffffffffu, valid && q >= NCH)) {
        while (sync[0] < tn - 1) __nanosleep(64);
        __threadfence_block();
      }
      if (valid) {
        float4 a, b;
        system_pixel<C, OF>(D, mask, sp, fl, du, px, j, w, h, d.plane, o, hd3, hg3, a, b);
        const unsigned slot = (unsigned)(j * RPX) + sorw_ring<RPX>((unsigned)px);
        asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(smem + slot * 16u), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w) : "memory");
        asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(smem + offB + slot * 16u), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
        const size_t g = (size_t)j * w + px;
        if (OF) {
          const float2 q2 = reinterpret_cast<const float2*>(du)[g];
          asm volatile("st.shared.v2.f32 [%0], {%1,%2};" ::"r"(smem + 2u * offB + slot * 8u), "f"(q2.x), "f"(q2.y) : "memory");
        } else {
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(smem + 2u * offB + slot * 8u), "f"(du[g]) : "memory");
        }
      }
    }
    __syncwarp();
    if (lane == 0) {
      __threadfence_block();
      sync[1 + band] = tn;
    }
  }
}

// dynamic shared memory of the fused kernel: the rings of sor_warp plus the hand-shake words
__host__ __device__ inline size_t sorf_smem_bytes(int h, int ns) { return sorw_smem_bytes(h, ns) + 64; }

#endif
