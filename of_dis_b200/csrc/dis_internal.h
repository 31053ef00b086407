// Internal declarations shared by the kernels and the engine.
#ifndef DISB200_INTERNAL_H_
#define DISB200_INTERNAL_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/disb200.h"

namespace disb200 {

// Derived constants of one pyramid level: camparam (oflow.h:30-43, oflow.cpp:139-157) and the
// patch grid of PatGridClass::PatGridClass (patchgrid.cpp:31-76).
struct LevelGeom {
  int level;
  int w, h;          // level size (of the padded-to-divisible frame)
  int pad;           // imgpadding == patch size
  int tw, th;        // w + 2 pad, h + 2 pad
  int nopw, noph, npatch;
  int offw, offh;    // grid offsets
  int steps;
  float lb, ubw, ubh;
  size_t img_elems;  // tw*th*C
  // Spatial tiling of one pair over several GPUs (tiled.cu): the part of the level this device works on.  Arrays are
  // full-size and globally indexed on every device; a device computes patch rows gy0 .. gy0+ngy-1 (patches are owned by
  // the row of their reference point) and pixel rows row0 .. row1-1 (plus the halo rows each stage needs, recomputed
  // locally).  Untiled: the whole level.
  int gy0, ngy;
  int row0, row1;
};

// optparam (oflow.h:45-91, oflow.cpp:75-107)
struct OptConst {
  int psz, noc, nop, novals, steps;
  int max_iter, min_iter, patnorm, costfct, mode;
  float dp_thresh_sq, dr_thresh, res_thresh, outlierthresh;
  float minerrval, normoutlier;
  int patch_map = 0;   // DisTuning::patch_map
};

OptConst make_optconst(const DisParams& p);
LevelGeom make_levelgeom(const DisParams& p, int W, int H, int level);

// Launch bookkeeping (gpu_launches in bench.py is read from here).
struct LaunchStats {
  int launches;
  // optional stage marks for the CUDA-event stopwatch (begin = 1 / end = 0)
  void (*mark)(void* ctx, int stage, int begin);
  void* ctx;
};

// ---- stage 1: pyramid -------------------------------------------------------------------
// u8 frames [n][h0][w0][C] -> padded level image of level `lvl` (2^lvl x 2^lvl box mean of
// the replicate-padded frame), then Sobel/8 + border fill.  Exact for any lvl <= 8.
struct PyrLevelPtrs {
  float* img[2];   // [frame]: padded level arrays of frame A / B, [n][th][tw][C]
  float* dx[2];
  float* dy[2];
};
// All images per launch: blockIdx.z enumerates the nA images of role A, then the nB of role B.
void launch_downscale_u8(const uint8_t* framesA, const uint8_t* framesB, int nA, int nB, int w0, int h0, int C, int padl,
                         int padt, int lvl, const LevelGeom& g, float* imgA, float* imgB, cudaStream_t s,
                         LaunchStats* st);
// SELECTCHANNEL == 2: level `lvl` of the gradient-magnitude pyramid straight from the u8 frames
void launch_downscale_gradmag(const uint8_t* framesA, const uint8_t* framesB, int nA, int nB, int w0, int h0, int W, int H,
                              int padl, int padt, int lvl, const LevelGeom& g, float* imgA, float* imgB, cudaStream_t s,
                              LaunchStats* st);
// level image (padded array, interior valid) -> next coarser level image interior (2x2 mean)
void launch_halve(const float* srcA, const float* srcB, const LevelGeom& gs, float* dstA, float* dstB,
                  const LevelGeom& gd, int nA, int nB, int C, cudaStream_t s, LaunchStats* st);
// interiors valid -> replicate borders of img, dx / dy (zero border), all levels in one launch
void launch_sobel_pad_levels(const PyrLevelPtrs* lv, const LevelGeom* g, int nlev, int nA, int nB, int C, cudaStream_t s,
                             LaunchStats* st);

// ---- stage 2: inverse search ------------------------------------------------------------
struct PatchArgs {
  const float* tmpl_img;   // template frame: padded level image / dx / dy   [n][th][tw][C]
  const float* tmpl_dx;
  const float* tmpl_dy;
  const float* tgt_img;    // target frame padded level image
  const float* flow_prev;  // coarser-level dense flow [n][h/2][w/2][nop] or null
  size_t flow_prev_stride; // elements per pair
  float* p_iter;           // [n][npatch][nop]
  float* pt_iter;          // [n][npatch][2]   (may be null)
  float* pweight;          // [n][npatch][novals]
  int camlr;
};
// returns false if (psz, channels) has no kernel instantiation
bool launch_patch_search(const PatchArgs& a, const LevelGeom& g, const OptConst& o, int n, cudaStream_t s,
                         LaunchStats* st);

// ---- stage 3: densification ---------------------------------------------------------------
struct DensifyArgs {
  const float* p_iter;     // own grid
  const float* pweight;
  const float* cg_p_iter;  // complementary grid (forward-backward merging) or null
  const float* cg_pt_iter;
  const float* cg_pweight;
  float* flow;             // [n][h][w][nop]
  // scratch for the backward merge (CSR cell lists), may be null when cg is null
  int* cell_count;         // [n][ncell]
  int* cell_start;         // [n][ncell]  exclusive scan of cell_count
  int* cell_end;           // [n][ncell]  fill cursor, = start + count once filled
  int* cell_list;          // [n][npatch] patch indices grouped by cell, ascending inside a cell
  int y0 = 0, ny = -1;     // pixel rows to produce (ny < 0: the whole level)
};
void launch_densify(const DensifyArgs& a, const LevelGeom& g, const OptConst& o, int n, cudaStream_t s,
                    LaunchStats* st);

// ---- stage 4: variational refinement ------------------------------------------------------
struct VarRefBuffers {
  // all planes are [n][h][w] (stride == w), sized for the largest level
  float* avg;    // [C] planes
  float* Iz;     // [C]
  float* mask;
  float* Ix; float* Iy; float* Ixx; float* Ixy; float* Iyy; float* Ixz; float* Iyz;   // [C] each
  float* du;     // [nop] interleaved
  float* smooth; // float2 {h, v}
  float* coef;   // float4 {hl, h, vt, v}
  float* inv;    // float4 {i11, i12, i22, 0}   (DE: {a11+sum,0,0,0})
  float* rhs;    // spare [nop] planes: boundary-row exchange of the banded SOR (sor_band.cuh), zeroed at create
  unsigned* sorb_ctl = nullptr;   // device: {ticket counter, error flag} of the banded SOR; null = banded SOR off
  unsigned* sorb_host = nullptr;  // host:   {launch epoch, tickets issued so far}
  size_t plane_cap;  // elements per pair per plane (>= w*h of the largest level)
  int tv_schedule = 0;   // DisTuning::tv_schedule
  int sorb_max = 0;      // banded SOR: largest n * bands (<= co-resident CTAs of sor_band_kernel on this device; 0 = off)
};
// co-resident CTAs of the banded SOR kernel on the current device (occupancy API x SM count), for `solverit` sweeps
int sor_band_resident_ctas(bool of, int solverit);
void launch_varref_level(const float* im_a, const float* im_b, float* flow, const LevelGeom& g, const DisParams& p,
                         int camlr, const VarRefBuffers& vb, int n, cudaStream_t s, LaunchStats* st);
size_t varref_plane_count(int C, int nop);  // number of float planes VarRefBuffers needs
// The same for one or two sub-batches ("lanes") at once; two lanes run their fixed-point iterations in anti-phase (one
// assembles its linear system while the other solves, varref.cu).  ev: an event of the lane's device (two-lane calls).
struct TvLane {
  const float* im_a; const float* im_b; float* flow;
  VarRefBuffers vb;
  int n;
  cudaStream_t st;
  LaunchStats* ls;
  cudaEvent_t ev;
};
void launch_varref_level_lanes(TvLane* lanes, int nl, const LevelGeom& g, const DisParams& p, int camlr);

// ---- stage 4, one pair whose rows are split over several devices (tiled.cu) ----------------------
// Every device holds the full-size, globally indexed arrays of the level and works on the band-aligned strip
// g.row0 .. g.row1-1; halo rows of the per-level planes are recomputed locally, the increment's halo rows travel between
// neighbours after every SOR call, and the lexicographic SOR itself is the banded kernel (sor_band.cuh) whose tagged
// boundary rows are stored straight into the neighbouring device's memory (NVLink P2P).
struct TileDev {
  int device;
  cudaStream_t stream;
  const float* im_a;        // padded level images of this direction's template / target frame, on this device
  const float* im_b;
  float* flow;              // [h][w][nop], full size; rows row0-4 .. row1+3 valid on entry, rows row0 .. row1-1 refined on exit
  VarRefBuffers vb;
  LevelGeom g;              // row0 / row1: this device's strip (multiples of 32 rows)
  LaunchStats* st;
  cudaEvent_t ev_sor, ev_halo;
};
bool varref_level_tileable(const LevelGeom& g, const DisParams& p, int ndev, size_t plane_cap, int nop);
// returns a cudaError_t value (0 = ok)
int launch_varref_level_tiled(TileDev* dv, int ndev, const DisParams& p, int camlr);

// ---- stage 5: upsample + crop ---------------------------------------------------------------
// out_y0 / out_ny: output rows to produce (out_ny < 0: all h0 of them); `out` always points at row 0 of pair 0
void launch_upsample_crop(const float* flow, const LevelGeom& g, int nop, int sc_l, int padl, int padt, int w0,
                          int h0, float* out, int n, cudaStream_t s, LaunchStats* st, int out_y0 = 0, int out_ny = -1);

// level-sc_l flow of one pair -> initflow of the next (box mean over k x k, times scale)
void launch_warm_init(const float* flow, int wl, int nop, int k, float scale, int wi, int hi, float* init, cudaStream_t s,
                      LaunchStats* st);

}  // namespace disb200

#endif
