// Stage 3 — densification, PatGridClass::AggregateFlowDense (patchgrid.cpp:213-397), recast
// from the reference's sequential scatter into a deterministic per-pixel gather.
//
// Forward part: the patches covering pixel (x, y) sit on a regular grid, so the pixel
// enumerates them directly, in ascending patch index ip = gx * noph + gy — the order in which
// the reference's scatter loop adds them — and divides by the weight sum in the same thread
// (the normalisation loop of patchgrid.cpp:377-394 is fused).
//
// Forward-backward merging (cg != null): patches of the complementary grid are splatted
// around their END positions, which are irregular.  They are first binned by the cell
// (cell edge = patch size) holding their top-left tap — count, exclusive scan per pair, fill:
// a CSR layout with exactly npatch entries per pair, so any distribution of end positions fits
// (a contracting or occluded region may put hundreds of patches into one cell) — the per-cell
// lists are sorted by patch index, and each pixel then visits the (at most) four cells that
// can reach it, in a fixed cell order.  The result is deterministic; its summation order
// differs from the reference's global patch order only by fp32 rounding.
#include <mutex>

#include "engine.h"

namespace disb200 {

namespace {

// RGB pointer desync of the reference (patchgrid.cpp:243,255-258 and 320,333-334): the weight
// pointer advances by one per patch pixel plus two more only for pixels that pass the
// in-image test.  Returns the weight offset of patch pixel (r, c) given the rectangle
// [c_lo, c_hi] x [r_lo, r_hi] of patch pixels that pass the test (empty if lo > hi).
__device__ __forceinline__ int rgb_weight_offset(int r, int c, int ps, int c_lo, int c_hi, int r_lo, int r_hi) {
  const int ncols = max(0, c_hi - c_lo + 1);
  const int rows_before = max(0, min(r - 1, r_hi) - r_lo + 1);   // fully processed in-range rows above r
  int inside_before = rows_before * ncols;
  if (r >= r_lo && r <= r_hi) inside_before += max(0, min(c - 1, c_hi) - c_lo + 1);
  return (r * ps + c) + 2 * inside_before;
}

template <int C, int NOP>
__global__ void densify_kernel(DensifyArgs a, LevelGeom g, OptConst o) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = a.y0 + blockIdx.y, pair = blockIdx.z;
  if (x >= g.w) return;
  // the patch window is lb..ub on both axes: ps = 2 (psz / 2) pixels, = psz for even sizes; the weight pointer runs
  // over exactly those ps x ps pixels (patchgrid.cpp:239-243), whatever novals is
  const int psz = o.psz, lb = -(psz / 2), ub = psz / 2 - 1, ps = ub - lb + 1, nv = o.novals;
  const float* P = a.p_iter + (size_t)pair * g.npatch * NOP;
  const float* Wt = a.pweight + (size_t)pair * g.npatch * nv;
  float we = 0.f, f0 = 0.f, f1 = 0.f;

  // grid columns gx with  gx*steps + offw + lb <= x <= gx*steps + offw + ub
  const int ax = x - g.offw, ay = y - g.offh;
  int gx_lo = (ax - ub + g.steps - 1 >= 0) ? (ax - ub + g.steps - 1) / g.steps : 0;
  int gx_hi = (ax - lb >= 0) ? (ax - lb) / g.steps : -1;
  int gy_lo = (ay - ub + g.steps - 1 >= 0) ? (ay - ub + g.steps - 1) / g.steps : 0;
  int gy_hi = (ay - lb >= 0) ? (ay - lb) / g.steps : -1;
  gx_hi = min(gx_hi, g.nopw - 1);
  gy_hi = min(gy_hi, g.noph - 1);
  for (int gx = gx_lo; gx <= gx_hi; ++gx) {
    const int px = gx * g.steps + g.offw;
    const int c = x - px - lb;
    for (int gy = gy_lo; gy <= gy_hi; ++gy) {
      const int py = gy * g.steps + g.offh;
      const int r = y - py - lb;
      const int ip = gx * g.noph + gy;
      const float* pw = Wt + (size_t)ip * nv;
      float absw;
      if (C == 1) {
        absw = 1.0f / fmaxf(o.minerrval, __ldg(pw + r * ps + c));
      } else {
        const int c_lo = max(0, -(px + lb)), c_hi = min(ps - 1, g.w - 1 - (px + lb));
        const int r_lo = max(0, -(py + lb)), r_hi = min(ps - 1, g.h - 1 - (py + lb));
        const int off = rgb_weight_offset(r, c, ps, c_lo, c_hi, r_lo, r_hi);
        absw = fmaxf(o.minerrval, __ldg(pw + off));
        absw += fmaxf(o.minerrval, __ldg(pw + off + 1));
        absw += fmaxf(o.minerrval, __ldg(pw + off + 2));
        absw = 1.0f / absw;
      }
      we += absw;
      f0 += __ldg(P + (size_t)ip * NOP) * absw;
      if (NOP == 2) f1 += __ldg(P + (size_t)ip * NOP + 1) * absw;
    }
  }

  if (a.cg_p_iter) {
    // backward merge: gather from the binned complementary patches (patchgrid.cpp:278-375)
    const float* CP = a.cg_p_iter + (size_t)pair * g.npatch * NOP;
    const float* CT = a.cg_pt_iter + (size_t)pair * g.npatch * 2;
    const float* CW = a.cg_pweight + (size_t)pair * g.npatch * nv;
    const int ncx = (g.w + 2 * psz) / psz + 2, ncy = (g.h + 2 * psz) / psz + 2;
    const int* cbeg = a.cell_start + (size_t)pair * ncx * ncy;
    const int* cend = a.cell_end + (size_t)pair * ncx * ncy;
    const int* lst = a.cell_list + (size_t)pair * g.npatch;
    // anchor (top-left tap) ax0 = pos0 + lb - 1 reaches pixels ax0 .. ax0 + psz
    const int cx1 = (x + 2 * psz) / psz, cx0 = (x - psz + 2 * psz) / psz;
    const int cy1 = (y + 2 * psz) / psz, cy0 = (y - psz + 2 * psz) / psz;
    for (int cy = cy0; cy <= cy1; ++cy)
      for (int cx = cx0; cx <= cx1; ++cx) {
        if (cx < 0 || cy < 0 || cx >= ncx || cy >= ncy) continue;
        const int cell = cy * ncx + cx;
        const int e1 = cend[cell];
        for (int e = cbeg[cell]; e < e1; ++e) {
          const int ip = lst[e];
          const float rx = __ldg(CT + 2 * ip), ry = __ldg(CT + 2 * ip + 1);
          const int pos0 = (int)ceil((double)rx + .00001), pos1 = (int)ceil((double)ry + .00001);
          const float fx = floorf(rx), fy = floorf(ry);
          const float r0 = rx - fx, r1 = ry - fy;
          const float wb[4] = {r0 * r1, (1.f - r0) * r1, r0 * (1.f - r1), (1.f - r0) * (1.f - r1)};
          // the four taps of patch pixel (xt, yt): cc (xt,yt), fc (xt-1,yt), cf (xt,yt-1), ff (xt-1,yt-1)
          const int c_lo = max(0, 1 - (pos0 + lb)), c_hi = min(ps - 1, g.w - 2 - (pos0 + lb));
          const int r_lo = max(0, 1 - (pos1 + lb)), r_hi = min(ps - 1, g.h - 2 - (pos1 + lb));
          const float* pw = CW + (size_t)ip * nv;
          const float p0 = __ldg(CP + (size_t)ip * NOP), p1 = NOP == 2 ? __ldg(CP + (size_t)ip * NOP + 1) : 0.f;
#pragma unroll
          for (int tpi = 0; tpi < 4; ++tpi) {
            const int xt = x + (tpi & 1), yt = y + (tpi >> 1);      // tap 0 cc, 1 fc, 2 cf, 3 ff
            const int c = xt - pos0 - lb, r = yt - pos1 - lb;
            if (c < c_lo || c > c_hi || r < r_lo || r > r_hi) continue;
            float absw;
            if (C == 1) {
              absw = 1.0f / fmaxf(o.minerrval, __ldg(pw + r * ps + c));
            } else {
              const int off = rgb_weight_offset(r, c, ps, c_lo, c_hi, r_lo, r_hi);
              absw = fmaxf(o.minerrval, __ldg(pw + off));
              absw += fmaxf(o.minerrval, __ldg(pw + off + 1));
              absw += fmaxf(o.minerrval, __ldg(pw + off + 2));
              absw = 1.0f / absw;
            }
            we += wb[tpi] * absw;
            f0 -= wb[tpi] * (p0 * absw);
            if (NOP == 2) f1 -= wb[tpi] * (p1 * absw);
          }
        }
      }
  }

  if (we > 0.f) { f0 /= we; f1 /= we; }
  float* out = a.flow + ((size_t)pair * g.w * g.h + (size_t)y * g.w + x) * NOP;
  out[0] = f0;
  if (NOP == 2) out[1] = f1;
}

// ---- binning of the complementary grid's end positions (CSR: count, scan, fill, sort) ----
__device__ __forceinline__ int bin_cell_of(const float* __restrict__ pt_iter, size_t pi, int psz, int w, int h) {
  const float rx = pt_iter[pi * 2], ry = pt_iter[pi * 2 + 1];
  const int ncx = (w + 2 * psz) / psz + 2, ncy = (h + 2 * psz) / psz + 2;
  const int pos0 = (int)ceil((double)rx + .00001), pos1 = (int)ceil((double)ry + .00001);
  const int ax = pos0 - psz / 2 - 1 + 2 * psz, ay = pos1 - psz / 2 - 1 + 2 * psz;
  if (ax < 0 || ay < 0) return -1;   // cannot reach the image
  const int cx = ax / psz, cy = ay / psz;
  if (cx >= ncx || cy >= ncy) return -1;
  return cy * ncx + cx;
}

__global__ void bin_count_kernel(const float* __restrict__ pt_iter, int npatch, int psz, int w, int h, int ncell,
                                 int* __restrict__ cell_count) {
  const int ip = blockIdx.x * blockDim.x + threadIdx.x, pair = blockIdx.y;
  if (ip >= npatch) return;
  const int cell = bin_cell_of(pt_iter, (size_t)pair * npatch + ip, psz, w, h);
  if (cell >= 0) atomicAdd(cell_count + (size_t)pair * ncell + cell, 1);
}

// one CTA per pair: exclusive scan of the cell populations -> cell_start; cell_end starts as a copy (the fill cursor)
__global__ void __launch_bounds__(1024)
bin_scan_kernel(const int* __restrict__ cell_count, int ncell, int* __restrict__ cell_start, int* __restrict__ cell_end) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int* cnt = cell_count + (size_t)pair * ncell;
  if (tid == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < ncell; base += 1024) {
    const int i = base + tid;
    const int v = i < ncell ? cnt[i] : 0;
    int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
      int t = warp_tot[lane], inc2 = t;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, inc2, d);
        if (lane >= d) inc2 += u;
      }
      warp_tot[lane] = inc2 - t;   // exclusive prefix of the warp totals
    }
    __syncthreads();
    const int excl = carry + warp_tot[wid] + incl - v;
    if (i < ncell) {
      cell_start[(size_t)pair * ncell + i] = excl;
      cell_end[(size_t)pair * ncell + i] = excl;
    }
    __syncthreads();
    if (tid == 1023) carry = excl + v;
    __syncthreads();
  }
}

__global__ void bin_fill_kernel(const float* __restrict__ pt_iter, int npatch, int psz, int w, int h, int ncell,
                                int* __restrict__ cell_end, int* __restrict__ cell_list) {
  const int ip = blockIdx.x * blockDim.x + threadIdx.x, pair = blockIdx.y;
  if (ip >= npatch) return;
  const int cell = bin_cell_of(pt_iter, (size_t)pair * npatch + ip, psz, w, h);
  if (cell < 0) return;
  const int slot = atomicAdd(cell_end + (size_t)pair * ncell + cell, 1);
  cell_list[(size_t)pair * npatch + slot] = ip;
}

__global__ void bin_sort_kernel(const int* __restrict__ cell_start, const int* __restrict__ cell_end, int* cell_list,
                                int ncell, int npatch) {
  const int cell = blockIdx.x * blockDim.x + threadIdx.x, pair = blockIdx.y;
  if (cell >= ncell) return;
  const int b = cell_start[(size_t)pair * ncell + cell], m = cell_end[(size_t)pair * ncell + cell] - b;
  int* l = cell_list + (size_t)pair * npatch + b;
  for (int i = 1; i < m; ++i) {   // insertion sort: lists are a handful of entries unless the flow collapses a region
    const int v = l[i];
    int k = i - 1;
    while (k >= 0 && l[k] > v) { l[k + 1] = l[k]; --k; }
    l[k + 1] = v;
  }
}

}  // namespace

void launch_densify(const DensifyArgs& a, const LevelGeom& g, const OptConst& o, int n, cudaStream_t s,
                    LaunchStats* st) {
  if (a.cg_p_iter) {
    const int ncx = (g.w + 2 * o.psz) / o.psz + 2, ncy = (g.h + 2 * o.psz) / o.psz + 2, ncell = ncx * ncy;
    cudaMemsetAsync(a.cell_count, 0, sizeof(int) * (size_t)ncell * n, s);
    const dim3 pg((g.npatch + 127) / 128, n);
    bin_count_kernel<<<pg, 128, 0, s>>>(a.cg_pt_iter, g.npatch, o.psz, g.w, g.h, ncell, a.cell_count);
    bin_scan_kernel<<<n, 1024, 0, s>>>(a.cell_count, ncell, a.cell_start, a.cell_end);
    bin_fill_kernel<<<pg, 128, 0, s>>>(a.cg_pt_iter, g.npatch, o.psz, g.w, g.h, ncell, a.cell_end, a.cell_list);
    bin_sort_kernel<<<dim3((ncell + 127) / 128, n), 128, 0, s>>>(a.cell_start, a.cell_end, a.cell_list, ncell, g.npatch);
    st->launches += 4;
  }
  const int ny = a.ny < 0 ? g.h : a.ny;
  if (ny <= 0) return;
  dim3 blk(128), grd((g.w + 127) / 128, ny, n);
  if (o.noc == 1) {
    if (o.nop == 2) densify_kernel<1, 2><<<grd, blk, 0, s>>>(a, g, o);
    else densify_kernel<1, 1><<<grd, blk, 0, s>>>(a, g, o);
  } else {
    if (o.nop == 2) densify_kernel<3, 2><<<grd, blk, 0, s>>>(a, g, o);
    else densify_kernel<3, 1><<<grd, blk, 0, s>>>(a, g, o);
  }
  st->launches++;
}

}  // namespace disb200

// ---------------------------------------------------------------------------------------------------------
// Allocation-free variant of the AggregateFlowDense cut (disb200_aggregate_dense_static, include/disb200.h).  No
// counterpart in the reference: added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  Inputs and outputs are the CALLER's device
// arrays; the CSR binning scratch of the forward-backward merge (cell counts, starts, fill cursors, the cell lists)
// lives in the file-scope device array below - DISB200_STATIC_SCRATCH_BYTES on every GPU that loads the library, until
// the process exits.  Every call on a device shares that array and finds the previous call's contents in it: the counts
// are cleared at the start of every call (launch_densify), starts / cursors / lists are fully rewritten before they are
// read.  Calls that overlap on one GPU (different streams) race on it.
namespace {
constexpr size_t kStaticScratchInts = DISB200_STATIC_SCRATCH_BYTES / sizeof(int);
static __device__ __align__(256) int g_static_scratch[kStaticScratchInts];
constexpr int kMaxDevices = 64;
int* g_static_scratch_dev[kMaxDevices] = {};   // device address of g_static_scratch per device (the symbol itself is a host-side stand-in)
std::mutex g_static_scratch_mu;
}  // namespace

extern "C" int disb200_aggregate_dense_static(const DisParams* params, int w_l, int h_l, const float* d_p_iter,
                                              const float* d_pweight, const float* d_cg_p_iter, const float* d_cg_pt_iter,
                                              const float* d_cg_pweight, float* d_flow, void* stream) {
  using namespace disb200;
  if (!params || !d_p_iter || !d_pweight || !d_flow) return fail(DISB200_EINVAL, "null argument");
  int rc = validate(*params);
  if (rc) return rc;
  if (w_l < 1 || h_l < 1) return fail(DISB200_EINVAL, "bad level size");
  const bool fb = d_cg_p_iter || d_cg_pt_iter || d_cg_pweight;
  if (fb && !(d_cg_p_iter && d_cg_pt_iter && d_cg_pweight)) return fail(DISB200_EINVAL, "complementary grid: all three arrays or none");
  DisParams q = *params;
  q.sc_f = q.sc_l = 0;
  q.usetvref = 0;
  q.usefbcon = fb ? 1 : 0;
  const OptConst o = make_optconst(q);
  const LevelGeom g = make_levelgeom(q, w_l, h_l, 0);
  // the layout launch_densify expects: cell_count | cell_start | cell_end (ncell each) | cell_list (npatch)
  const size_t ncell = (size_t)((g.w + 2 * o.psz) / o.psz + 2) * (size_t)((g.h + 2 * o.psz) / o.psz + 2);
  const size_t need = fb ? 3 * ncell + (size_t)g.npatch : 0;
  if (need > kStaticScratchInts)
    return fail(DISB200_EINVAL, "disb200_aggregate_dense_static: level needs " + std::to_string(need * sizeof(int)) +
                                    " bytes of scratch (3 * cells + patches ints), the static array holds " +
                                    std::to_string((size_t)DISB200_STATIC_SCRATCH_BYTES) + " (DISB200_STATIC_SCRATCH_BYTES); use disb200_aggregate_dense");
  int dev = -1;
  CU(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(DISB200_EINVAL, "device ordinal out of range");
  int* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_static_scratch_mu);
    if (!g_static_scratch_dev[dev]) {   // once per device, with that device current
      void* ptr = nullptr;
      CU(cudaGetSymbolAddress(&ptr, g_static_scratch));
      g_static_scratch_dev[dev] = static_cast<int*>(ptr);
    }
    scratch = g_static_scratch_dev[dev];
  }
  DensifyArgs a;
  a.p_iter = d_p_iter; a.pweight = d_pweight;
  a.cg_p_iter = d_cg_p_iter; a.cg_pt_iter = d_cg_pt_iter; a.cg_pweight = d_cg_pweight;
  a.flow = d_flow;
  a.cell_count = scratch; a.cell_start = scratch + ncell; a.cell_end = scratch + 2 * ncell; a.cell_list = scratch + 3 * ncell;
  LaunchStats st;
  launch_densify(a, g, o, 1, static_cast<cudaStream_t>(stream), &st);
  CU(cudaGetLastError());
  return DISB200_OK;
}
