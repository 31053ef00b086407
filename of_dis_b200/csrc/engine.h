// The per-device engine: device workspace + scale loop of OFC::OFClass (oflow.cpp:31-338) for up to max_batch pairs of
// one geometry.  Shared by engine.cu (handles, C-ABI) and tiled.cu (one pair split over several devices).
#ifndef DISB200_ENGINE_H_
#define DISB200_ENGINE_H_

#include <algorithm>
#include <cmath>
#include <string>
#include <vector>

#include "dis_internal.h"

namespace disb200 {

int fail(int code, const std::string& msg);        // records the thread's last error text, returns code
int validate(const DisParams& p);
extern DisTuning g_tuning;                         // disb200_set_tuning; copied into every handle / engine at create
// internal: the banded SOR lost co-residency; the caller reruns the call once (bands are off from then on)
static const int DISB200_RETRY_NO_BANDS = -77;

#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e__ = (call);                                                             \
    if (e__ != cudaSuccess)                                                               \
      return fail(e__ == cudaErrorMemoryAllocation ? DISB200_ENOMEM : DISB200_ECUDA,      \
                  std::string(#call) + ": " + cudaGetErrorString(e__));                   \
  } while (0)

// Device workspace + scale loop for n <= max_batch pairs of one geometry.
struct Engine {
  DisParams p;
  OptConst o;
  int W = 0, H = 0, C = 1, nop = 2, max_batch = 0, device = 0;
  std::vector<LevelGeom> geom;                       // indexed by level
  std::vector<float*> pyr[2][3];                     // [frame][img|dx|dy][level]
  std::vector<float*> flow[2];                       // [dir][level]  [n][h][w][nop]
  float* p_iter[2] = {nullptr, nullptr};
  float* pt_iter[2] = {nullptr, nullptr};
  float* pweight[2] = {nullptr, nullptr};
  float* initflow = nullptr;                         // optional warm start (oflow.cpp:215-217)
  bool has_initflow = false;
  float* vpool = nullptr;
  VarRefBuffers vb;
  int* cell_count = nullptr;      // fwd/bwd merge: CSR binning of the complementary grid's end positions
  int* cell_start = nullptr;
  int* cell_end = nullptr;
  int* cell_list = nullptr;
  unsigned* sorb_ctl = nullptr;     // banded SOR: device {ticket counter, error flag}
  unsigned sorb_host[2] = {0, 0};   //             host {launch epoch, tickets issued}
  size_t ncell_max = 0;
  std::vector<void*> allocs;
  cudaStream_t stream = nullptr;
  LaunchStats stats{0, nullptr, nullptr};
  // stage timing
  bool stage_timing = false;
  float stage_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  std::vector<cudaEvent_t> ev_pool;
  struct Mark { int stage; cudaEvent_t a, b; int level; };
  float scale_ms[17][3] = {};     // [level][inverse search | densification | refinement] of the last timed run (oflow.cpp:297)
  std::vector<Mark> marks;
  size_t ev_used = 0;

  template <class T>
  int dalloc(T** ptr, size_t count) {
    void* q = nullptr;
    CU(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
    allocs.push_back(q);
    *ptr = static_cast<T*>(q);
    return DISB200_OK;
  }

  int init(const DisParams& prm, int W_, int H_, int max_batch_, int device_) {
    p = prm;
    o = make_optconst(p);
    o.patch_map = g_tuning.patch_map;
    W = W_; H = H_; C = p.channels; nop = o.nop; max_batch = max_batch_; device = device_;
    if (W % (1 << p.sc_f) || H % (1 << p.sc_f)) return fail(DISB200_EINVAL, "W and H must be divisible by 2^sc_f (oflow.h:102)");
    CU(cudaSetDevice(device));
    CU(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    geom.resize(p.sc_f + 1);
    for (int f = 0; f < 2; ++f)
      for (int a = 0; a < 3; ++a) pyr[f][a].assign(p.sc_f + 1, nullptr);
    flow[0].assign(p.sc_f + 1, nullptr);
    flow[1].assign(p.sc_f + 1, nullptr);
    size_t max_patch = 0, max_pix = 0;
    for (int l = p.sc_l; l <= p.sc_f; ++l) {
      geom[l] = make_levelgeom(p, W, H, l);
      const LevelGeom& g = geom[l];
      if (g.w < 1 || g.h < 1) return fail(DISB200_EINVAL, "coarsest level is empty: W and H must be at least 2^sc_f");
      // (one-pixel-wide levels run: sor_coupled's fallback, solver.c:86-89.  A level ONE ROW high with the refinement on
      //  crashes the reference itself - segmentation fault in its 5-tap column filter - so there is nothing to match)
      if (p.usetvref && g.h < 2) return fail(DISB200_EUNSUPPORTED, "a pyramid level one row high cannot be refined (the reference crashes on it)");
      for (int f = 0; f < 2; ++f)
        for (int a = 0; a < 3; ++a) {
          int rc = dalloc(&pyr[f][a][l], g.img_elems * (max_batch + (f == 0 ? 1 : 0)));
          if (rc) return rc;
        }
      for (int d = 0; d < (p.usefbcon ? 2 : 1); ++d) {
        int rc = dalloc(&flow[d][l], (size_t)g.w * g.h * nop * max_batch);
        if (rc) return rc;
      }
      max_patch = std::max(max_patch, (size_t)g.npatch);
      max_pix = std::max(max_pix, (size_t)g.w * g.h);
      ncell_max = std::max(ncell_max, (size_t)((g.w + 2 * o.psz) / o.psz + 2) * ((g.h + 2 * o.psz) / o.psz + 2));
    }
    for (int d = 0; d < (p.usefbcon ? 2 : 1); ++d) {
      int rc = dalloc(&p_iter[d], max_patch * nop * max_batch);
      if (!rc) rc = dalloc(&pt_iter[d], max_patch * 2 * max_batch);
      if (!rc) rc = dalloc(&pweight[d], max_patch * o.novals * max_batch);
      if (rc) return rc;
    }
    if (p.usefbcon) {
      int rc = dalloc(&cell_count, ncell_max * max_batch);
      if (!rc) rc = dalloc(&cell_start, ncell_max * max_batch);
      if (!rc) rc = dalloc(&cell_end, ncell_max * max_batch);
      if (!rc) rc = dalloc(&cell_list, max_patch * max_batch);
      if (rc) return rc;
    }
    {
      const LevelGeom& gc = geom[p.sc_f];
      int rc = dalloc(&initflow, (size_t)(gc.w / 2) * (gc.h / 2) * nop * max_batch);
      if (rc) return rc;
    }
    if (p.usetvref) {
      const size_t plane = (max_pix + 3) & ~(size_t)3;   // keeps float4 planes 16-byte aligned
      const size_t nplanes = varref_plane_count(C, nop);
      int rc = dalloc(&vpool, plane * nplanes * max_batch);
      if (rc) return rc;
      float* q = vpool;
      auto take = [&](size_t k) { float* r = q; q += plane * k * max_batch; return r; };
      vb.plane_cap = plane;
      vb.coef = take(4); vb.inv = take(4);          // float4 planes first (alignment)
      vb.du = take(nop); vb.rhs = take(nop);
      vb.avg = take(C); vb.Iz = take(C); vb.Ix = take(C); vb.Iy = take(C);
      vb.Ixx = take(C); vb.Ixy = take(C); vb.Iyy = take(C); vb.Ixz = take(C); vb.Iyz = take(C);
      vb.mask = take(1); vb.smooth = take(1);
      // banded SOR of tall levels: tagged exchange rows live in the spare planes and must never hold a live epoch
      CU(cudaMemset(vpool, 0, sizeof(float) * plane * nplanes * max_batch));
      rc = dalloc(&sorb_ctl, 2);
      if (rc) return rc;
      CU(cudaMemset(sorb_ctl, 0, 2 * sizeof(unsigned)));
      sorb_host[0] = sorb_host[1] = 0;
      vb.sorb_ctl = sorb_ctl; vb.sorb_host = sorb_host;
      vb.tv_schedule = g_tuning.tv_schedule;
      // Bands wait for their neighbours: all bands of a launch must be co-resident.  The bound is what THIS device holds
      // of THIS kernel (occupancy API), not a constant; a handle that shares the GPU with other work can still fall short
      // of it, which the bounded poll turns into a flag and the entry points into a rerun without bands (band_off).
      const int resident = sor_band_resident_ctas(p.mode == DISB200_MODE_OF, std::min(std::max(p.tv_solverit, 1), 3));
      vb.sorb_max = g_tuning.sorb_max > 0 ? std::min(g_tuning.sorb_max, resident) : resident;
    }
    return DISB200_OK;
  }

  void destroy() {
    if (stream) cudaStreamSynchronize(stream);
    for (void* q : allocs) cudaFree(q);
    allocs.clear();
    for (cudaEvent_t e : ev_pool) cudaEventDestroy(e);
    ev_pool.clear();
    if (stream) cudaStreamDestroy(stream);
    stream = nullptr;
  }

  // banded SOR (sor_band.cuh): a band that waited too long for its neighbour raised the flag
  int check_sorb() {
    if (!sorb_ctl || sorb_host[0] == 0) return DISB200_OK;
    unsigned v = 0;
    CU(cudaMemcpy(&v, sorb_ctl + 1, sizeof(unsigned), cudaMemcpyDeviceToHost));
    if (v) {
      cudaMemset(sorb_ctl + 1, 0, sizeof(unsigned));
      vb.sorb_max = 0;   // this engine stops using bands: its GPU does not keep them co-resident (shared device)
      return fail(DISB200_RETRY_NO_BANDS, "banded SOR: a band timed out waiting for its neighbour (result invalid)");
    }
    return DISB200_OK;
  }

  cudaEvent_t next_event() {
    if (ev_used == ev_pool.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      ev_pool.push_back(e);
    }
    return ev_pool[ev_used++];
  }
  struct StageScope {
    Engine* e; int stage, level; cudaEvent_t a;
    StageScope(Engine* e_, int s, int lvl = -1) : e(e_), stage(s), level(lvl), a(nullptr) {
      if (e->stage_timing) { a = e->next_event(); cudaEventRecord(a, e->stream); }
    }
    ~StageScope() {
      if (e->stage_timing) { cudaEvent_t b = e->next_event(); cudaEventRecord(b, e->stream); e->marks.push_back({stage, a, b, level}); }
    }
  };
  cudaEvent_t open_mark[8] = {nullptr};
  static void mark_cb(void* ctx, int stage, int begin) {
    Engine* e = static_cast<Engine*>(ctx);
    cudaEvent_t ev = e->next_event();
    cudaEventRecord(ev, e->stream);
    if (begin) e->open_mark[stage] = ev;
    else e->marks.push_back({stage, e->open_mark[stage], ev, -1});
  }
  void begin_run() {
    stats.launches = 0;
    stats.mark = stage_timing ? &Engine::mark_cb : nullptr;
    stats.ctx = this;
    ev_used = 0;
    marks.clear();
  }
  void collect_stage_times() {
    if (!stage_timing) return;
    cudaStreamSynchronize(stream);
    for (float& v : stage_ms) v = 0.f;
    for (auto& row : scale_ms) row[0] = row[1] = row[2] = 0.f;
    for (const Mark& m : marks) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, m.a, m.b);
      stage_ms[m.stage] += ms;
      if (m.level >= 0 && m.level < 17 && m.stage >= 1 && m.stage <= 3) scale_ms[m.level][m.stage - 1] += ms;
    }
  }

  // `initflow` is indexed  floor(y / 2) * (w_c / 2) + floor(x / 2)  over the coarsest level's patch positions
  // (patchgrid.cpp:199-201), which stays inside a (h_c / 2) x (w_c / 2) array only when w_c and h_c are even: the
  // reference's (disabled) caller pads the frames to a multiple of 2^(sc_f + 1) for that reason (run_dense.cpp:301).
  bool initflow_ok() const { return geom[p.sc_f].w % 2 == 0 && geom[p.sc_f].h % 2 == 0; }
  static const char* initflow_why() { return "initflow / temporal warm start needs W and H divisible by 2^(sc_f + 1) (run_dense.cpp:301)"; }

  // Sequence mode (video): pair i = frames (i, i+1); the n + 1 frame pyramids live in the role-A arrays and the
  // second frame of pair i is simply the first frame of pair i + 1.
  bool seq = false;
  int pair0 = 0;   // sequence mode: first pair of the current run_levels call (warm-started sequences run pair by pair)
  float* P(int f, int a, int l) const { return seq ? pyr[0][a][l] + (size_t)(f + pair0) * geom[l].img_elems : pyr[f][a][l]; }

  // a1 + a2: u8 frames (device) -> padded pyramids of both frames (dB == nullptr: dA holds n + 1 consecutive frames)
  void build_pyramids(const uint8_t* dA, const uint8_t* dB, int n, int w0, int h0, int padl, int padt) {
    StageScope sc(this, 0);
    seq = dB == nullptr;
    const int nA = seq ? n + 1 : n, nB = seq ? 0 : n;
    if (p.gradmag)   // SELECTCHANNEL == 2: level 0 is the gradient magnitude, the cascade runs on arbitrary floats
      launch_downscale_gradmag(dA, dB, nA, nB, w0, h0, W, H, padl, padt, p.sc_l, geom[p.sc_l], pyr[0][0][p.sc_l], pyr[1][0][p.sc_l], stream, &stats);
    else
      launch_downscale_u8(dA, dB, nA, nB, w0, h0, C, padl, padt, p.sc_l, geom[p.sc_l], pyr[0][0][p.sc_l], pyr[1][0][p.sc_l], stream, &stats);
    for (int l = p.sc_l + 1; l <= p.sc_f; ++l)
      launch_halve(pyr[0][0][l - 1], pyr[1][0][l - 1], geom[l - 1], pyr[0][0][l], pyr[1][0][l], geom[l], nA, nB, C, stream, &stats);
    PyrLevelPtrs lv[9];
    LevelGeom gl[9];
    int nlev = 0;
    for (int l = p.sc_l; l <= p.sc_f; ++l, ++nlev) {
      gl[nlev] = geom[l];
      for (int f = 0; f < 2; ++f) { lv[nlev].img[f] = pyr[f][0][l]; lv[nlev].dx[f] = pyr[f][1][l]; lv[nlev].dy[f] = pyr[f][2][l]; }
    }
    launch_sobel_pad_levels(lv, gl, nlev, nA, nB, C, stream, &stats);
  }

  // One scale of the loop of oflow.cpp:182-330, in three pieces so that tiled.cu can run each on part of the level.
  // `g`: the level's geometry, possibly restricted to this device's patch rows (gy0 / ngy).
  int search_level(int sl, int n, const LevelGeom& g) {
    const int ndir = p.usefbcon ? 2 : 1;
    StageScope sc(this, 1, sl);
    for (int d = 0; d < ndir; ++d) {                                   // oflow.cpp:191-232
      PatchArgs a;
      a.tmpl_img = P(d, 0, sl); a.tmpl_dx = P(d, 1, sl); a.tmpl_dy = P(d, 2, sl);
      a.tgt_img = P(1 - d, 0, sl);
      a.flow_prev = nullptr; a.flow_prev_stride = 0;
      if (sl < p.sc_f) {
        a.flow_prev = flow[d][sl + 1];
        a.flow_prev_stride = (size_t)geom[sl + 1].w * geom[sl + 1].h * nop;
      } else if (d == 0 && has_initflow) {
        a.flow_prev = initflow;
        a.flow_prev_stride = (size_t)(g.w / 2) * (g.h / 2) * nop;
      }
      a.p_iter = p_iter[d]; a.pt_iter = pt_iter[d]; a.pweight = pweight[d];
      a.camlr = d;
      if (!launch_patch_search(a, g, o, n, stream, &stats))
        return fail(DISB200_EUNSUPPORTED, "no inverse-search kernel for this patch size / channel count");
    }
    return DISB200_OK;
  }
  // rows y0 .. y0+ny-1 of the dense flow of every direction this level needs (ny < 0: all rows)
  void densify_level(int sl, int n, int y0 = 0, int ny = -1) {
    const LevelGeom& g = geom[sl];
    const int ndir = p.usefbcon ? 2 : 1;
    StageScope sc(this, 2, sl);
    for (int d = 0; d < ndir; ++d) {                                   // oflow.cpp:259-267
      if (d == 1 && sl == p.sc_l) break;
      DensifyArgs a;
      a.p_iter = p_iter[d]; a.pweight = pweight[d];
      a.cg_p_iter = p.usefbcon ? p_iter[1 - d] : nullptr;
      a.cg_pt_iter = p.usefbcon ? pt_iter[1 - d] : nullptr;
      a.cg_pweight = p.usefbcon ? pweight[1 - d] : nullptr;
      a.flow = flow[d][sl];
      a.cell_count = cell_count; a.cell_start = cell_start; a.cell_end = cell_end; a.cell_list = cell_list;
      a.y0 = y0; a.ny = ny;
      launch_densify(a, g, o, n, stream, &stats);
    }
  }
  void varref_level(int sl, int n) {                                   // oflow.cpp:281-290
    if (!p.usetvref) return;
    const LevelGeom& g = geom[sl];
    StageScope sc(this, 3, sl);
    launch_varref_level(P(0, 0, sl), P(1, 0, sl), flow[0][sl], g, p, 0, vb, n, stream, &stats);
    if (p.usefbcon && sl > p.sc_l)
      launch_varref_level(P(1, 0, sl), P(0, 0, sl), flow[1][sl], g, p, 1, vb, n, stream, &stats);
  }

  // The scale loop, oflow.cpp:182-330.  Result: flow[0][sc_l].
  int run_levels(int n) {
    for (int sl = p.sc_f; sl >= p.sc_l; --sl) {
      int rc = search_level(sl, n, geom[sl]);
      if (rc) return rc;
      densify_level(sl, n);
      varref_level(sl, n);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(DISB200_ECUDA, std::string("kernel launch: ") + cudaGetErrorString(e));
    return DISB200_OK;
  }
};

}  // namespace disb200

#endif
