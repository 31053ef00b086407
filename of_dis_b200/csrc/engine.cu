// Host orchestration: the scale loop of OFC::OFClass (oflow.cpp:31-338) and the driver steps of
// run_dense.cpp:297-415 around it, for a batch of independent frame pairs, behind the C-ABI of
// include/disb200.h.  Every stage is a CUDA kernel (pyramid.cu, patch.cu, densify.cu,
// varref.cu, upsample.cu); there is no CPU implementation to fall back to.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "engine.h"

namespace disb200 {

static thread_local std::string g_err;
DisTuning g_tuning = {0, 0, 0, 0, 0, 0};
int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
OptConst make_optconst(const DisParams& p) {
  OptConst o;
  o.mode = p.mode;
  o.nop = p.mode == DISB200_MODE_OF ? 2 : 1;                              // oflow.cpp:75-79
  o.psz = p.patchsz;
  o.outlierthresh = (float)p.patchsz / 2;                                 // oflow.cpp:81
  o.max_iter = p.max_iter;
  o.min_iter = p.min_iter;
  o.dp_thresh_sq = p.dp_thresh * p.dp_thresh;                             // oflow.cpp:87
  o.dr_thresh = p.dr_thresh;
  o.res_thresh = p.res_thresh;
  o.steps = std::max(1, (int)std::floor(p.patchsz * (1 - p.patove)));     // oflow.cpp:90
  o.noc = p.channels;
  o.novals = p.channels * p.patchsz * p.patchsz;                          // oflow.cpp:91
  o.patnorm = p.patnorm;
  o.costfct = p.costfct;
  o.minerrval = 2.0f;                                                     // oflow.h:76
  o.normoutlier = 5.0f;                                                   // oflow.h:77
  return o;
}

LevelGeom make_levelgeom(const DisParams& p, int W, int H, int level) {
  LevelGeom g;
  const OptConst o = make_optconst(p);
  const float sc = std::pow(2.0f, -level);                                // oflow.cpp:142-151
  g.level = level;
  g.h = (int)(H * sc);
  g.w = (int)(W * sc);
  g.pad = p.patchsz;
  g.tw = g.w + 2 * g.pad;
  g.th = g.h + 2 * g.pad;
  g.lb = -(float)p.patchsz / 2;
  g.ubw = (float)(g.w + p.patchsz / 2 - 2);
  g.ubh = (float)(g.h + p.patchsz / 2 - 2);
  g.steps = o.steps;
  g.nopw = (int)std::ceil((float)g.w / (float)g.steps);                   // patchgrid.cpp:43-46
  g.noph = (int)std::ceil((float)g.h / (float)g.steps);
  g.offw = (g.w - (g.nopw - 1) * g.steps) / 2;
  g.offh = (g.h - (g.noph - 1) * g.steps) / 2;
  g.npatch = g.nopw * g.noph;
  g.img_elems = (size_t)g.tw * g.th * p.channels;
  g.gy0 = 0; g.ngy = g.noph;
  g.row0 = 0; g.row1 = g.h;
  return g;
}

int validate(const DisParams& p) {
  if (p.mode != DISB200_MODE_OF && p.mode != DISB200_MODE_DE) return fail(DISB200_EINVAL, "mode must be 1 (OF) or 2 (DE)");
  if (p.channels != 1 && p.channels != 3) return fail(DISB200_EINVAL, "channels must be 1 or 3");
  if (p.gradmag != 0 && p.gradmag != 1) return fail(DISB200_EINVAL, "gradmag must be 0 or 1");
  if (p.gradmag && p.channels != 1) return fail(DISB200_EINVAL, "gradmag (SELECTCHANNEL 2) is a single-band mode: channels must be 1");
  if (p.sc_l < 0 || p.sc_f < p.sc_l || p.sc_f > 8) return fail(DISB200_EINVAL, "need 0 <= sc_l <= sc_f <= 8");
  if (p.patove < 0.f || p.patove >= 1.f) return fail(DISB200_EINVAL, "patove must be in [0,1)");
  if (p.patchsz < 2 || p.patchsz > 16) return fail(DISB200_EUNSUPPORTED, "patch size must be in 2..16");
  if (p.costfct != 0 && p.costfct != 1 && p.costfct != 2 && p.costfct != 10)
    return fail(DISB200_EUNSUPPORTED, "costfct must be 0 (L2), 1 (L1), 2 (pseudo-Huber) or 10 (NCC)");
  if (p.max_iter < 0 || p.min_iter < 0) return fail(DISB200_EINVAL, "iteration counts must be >= 0");
  if (p.sor_mode != DISB200_SOR_LEXICOGRAPHIC && p.sor_mode != DISB200_SOR_REDBLACK) return fail(DISB200_EINVAL, "sor_mode must be 0 (lexicographic) or 1 (red-black)");
  if (p.usetvref && (p.tv_innerit < 0 || p.tv_solverit < 0)) return fail(DISB200_EINVAL, "tv_innerit >= 0 and tv_solverit >= 0 required");
  return DISB200_OK;
}

}  // namespace disb200

// ===========================================================================================
// C-ABI
// ===========================================================================================
using namespace disb200;

struct DisHandle {
  Engine e;                       // sub-batch engine 0; owns the handle's main stream
  std::vector<Engine*> extra;     // sub-batch engines 1.. (own stream each), created on demand
  int nsub = 1;                   // sub-batches run concurrently on separate streams
  int host_chunk = 0;             // DisTuning::host_chunk
  bool duo = false;               // two sub-batches enqueued side by side, refinements in anti-phase (DisTuning::streams == -2)
  cudaEvent_t ev_lane[2] = {nullptr, nullptr};
  int cap = 0;                    // pairs per sub-batch engine
  int max_batch = 0;              // pairs per call
  int last_per = 0;               // sub-batch size of the most recent run (for disb200_fetch)
  int last_n = 0;                 // pairs of the most recent enqueue
  cudaEvent_t ev_fork = nullptr;
  std::vector<cudaEvent_t> ev_join;
  int launches = 0;
  int w0 = 0, h0 = 0, padl = 0, padt = 0;
  Engine& sub(int i) { return i == 0 ? e : *extra[i - 1]; }
  // double-buffered device staging for the host-pointer entry points
  uint8_t* dA[2] = {nullptr, nullptr};
  uint8_t* dB[2] = {nullptr, nullptr};
  float* dOut[2] = {nullptr, nullptr};
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  cudaEvent_t ev_h2d[2], ev_comp[2], ev_d2h[2];
  cudaEvent_t t0 = nullptr, t1 = nullptr;
  bool staging = false;
  size_t in_bytes = 0, out_bytes = 0;   // per pair
  size_t low_bytes = 0;                 // per pair, flow at level sc_l (what OFClass writes to outflow, oflow.h:105)
  bool lowres = false;                  // the current call returns the sc_l flow instead of the upsampled one

  int ensure_staging() {
    if (staging) return DISB200_OK;
    for (int k = 0; k < 2; ++k) {
      int rc = e.dalloc(&dA[k], in_bytes * (max_batch + 1));   // + 1: a sequence chunk of m pairs is m + 1 frames
      if (!rc) rc = e.dalloc(&dB[k], in_bytes * max_batch);
      if (!rc) rc = e.dalloc(&dOut[k], std::max(out_bytes, low_bytes) / sizeof(float) * max_batch);   // sc_l = 0: the padded level exceeds the crop
      if (rc) return rc;
      CU(cudaEventCreateWithFlags(&ev_h2d[k], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&ev_comp[k], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&ev_d2h[k], cudaEventDisableTiming));
    }
    CU(cudaStreamCreateWithFlags(&s_h2d, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&s_d2h, cudaStreamNonBlocking));
    staging = true;
    return DISB200_OK;
  }

  int ensure_subs() {
    while ((int)extra.size() < nsub - 1) {
      Engine* x = new Engine();
      int rc = x->init(e.p, e.W, e.H, cap, e.device);
      if (rc) { x->destroy(); delete x; return rc; }
      extra.push_back(x);
      cudaEvent_t ev;
      CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
      ev_join.push_back(ev);
    }
    if (!ev_fork) CU(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
    for (int k = 0; k < 2; ++k)
      if (!ev_lane[k]) CU(cudaEventCreateWithFlags(&ev_lane[k], cudaEventDisableTiming));
    return DISB200_OK;
  }

  int enqueue_one(Engine& E, const uint8_t* a, const uint8_t* b, int n, float* out) {
    E.build_pyramids(a, b, n, w0, h0, padl, padt);
    int rc = E.run_levels(n);
    if (rc) return rc;
    if (lowres) {   // OFClass cut: hand out the level-sc_l flow as it is (no x 2^sc_l, no resize, no crop)
      CU(cudaMemcpyAsync(out, E.flow[0][E.p.sc_l], low_bytes * n, cudaMemcpyDeviceToDevice, E.stream));
    } else {
      Engine::StageScope sc(&E, 4);
      launch_upsample_crop(E.flow[0][E.p.sc_l], E.geom[E.p.sc_l], E.nop, E.p.sc_l, padl, padt, w0, h0, out, n, E.stream, &E.stats);
    }
    return DISB200_OK;
  }

  // Two sub-batches through the scale loop together (same stages as enqueue_one, level by level for both): the host
  // order lets the two refinements of a level interleave their iterations (launch_varref_level_lanes, anti-phase).
  int enqueue_two(Engine& A, const uint8_t* aA, const uint8_t* bA, int nA, float* outA, Engine& B, const uint8_t* aB,
                  const uint8_t* bB, int nB, float* outB) {
    Engine* E[2] = {&A, &B};
    const uint8_t* fa[2] = {aA, aB};
    const uint8_t* fb[2] = {bA, bB};
    const int nn[2] = {nA, nB};
    float* out[2] = {outA, outB};
    for (int k = 0; k < 2; ++k) E[k]->build_pyramids(fa[k], fb[k], nn[k], w0, h0, padl, padt);
    const DisParams& p = A.p;
    for (int sl = p.sc_f; sl >= p.sc_l; --sl) {
      for (int k = 0; k < 2; ++k) {
        int rc = E[k]->search_level(sl, nn[k], E[k]->geom[sl]);
        if (rc) return rc;
        E[k]->densify_level(sl, nn[k]);
      }
      if (p.usetvref) {
        const int ndir = (p.usefbcon && sl > p.sc_l) ? 2 : 1;
        for (int d = 0; d < ndir; ++d) {
          TvLane lanes[2];
          for (int k = 0; k < 2; ++k)
            lanes[k] = TvLane{E[k]->P(d, 0, sl), E[k]->P(1 - d, 0, sl), E[k]->flow[d][sl], E[k]->vb, nn[k], E[k]->stream, &E[k]->stats, ev_lane[k]};
          launch_varref_level_lanes(lanes, 2, A.geom[sl], p, d);
        }
      }
    }
    for (int k = 0; k < 2; ++k) {
      if (lowres) {
        CU(cudaMemcpyAsync(out[k], E[k]->flow[0][p.sc_l], low_bytes * nn[k], cudaMemcpyDeviceToDevice, E[k]->stream));
      } else {
        launch_upsample_crop(E[k]->flow[0][p.sc_l], E[k]->geom[p.sc_l], E[k]->nop, p.sc_l, padl, padt, w0, h0, out[k], nn[k], E[k]->stream, &E[k]->stats);
      }
    }
    return DISB200_OK;
  }

  // Enqueue the whole per-batch pipeline.  The batch is cut into nsub sub-batches that run on
  // their own streams (forked from / joined into e.stream), so that the latency-bound kernels
  // of one sub-batch (coarse-level SOR wavefronts) overlap the throughput-bound kernels of
  // another (inverse search, upsample).
  int enqueue(const uint8_t* a, const uint8_t* b, int n, float* out) {
    const int S = std::min(nsub, n);
    const bool serial = e.stage_timing;              // the per-stage stopwatch needs a single stream
    const int per = (n + S - 1) / S;
    if (per > cap) return fail(DISB200_EINVAL, "internal: sub-batch exceeds engine capacity");
    last_per = per;
    last_n = n;
    launches = 0;
    if (S > 1 && !serial) {
      int rc = ensure_subs();
      if (rc) return rc;
      CU(cudaEventRecord(ev_fork, e.stream));
    }
    const size_t out_elems = (lowres ? low_bytes : out_bytes) / sizeof(float);
    if (S == 2 && !serial && duo && n - per > 0) {
      // two sub-batches, enqueued level by level side by side so that their refinements can run in anti-phase
      Engine& A = sub(0);
      Engine& B = sub(1);
      B.begin_run();
      CU(cudaStreamWaitEvent(B.stream, ev_fork, 0));
      A.stats.launches = B.stats.launches = 0;
      int rc = enqueue_two(A, a, b, per, out, B, a + (size_t)per * in_bytes, b ? b + (size_t)per * in_bytes : nullptr, n - per,
                           out + (size_t)per * out_elems);
      if (rc) return rc;
      launches = A.stats.launches + B.stats.launches;
      CU(cudaEventRecord(ev_join[0], B.stream));
      CU(cudaStreamWaitEvent(e.stream, ev_join[0], 0));
      cudaError_t ce2 = cudaGetLastError();
      if (ce2 != cudaSuccess) return fail(DISB200_ECUDA, std::string("kernel launch: ") + cudaGetErrorString(ce2));
      return DISB200_OK;
    }
    for (int i = 0; i < S; ++i) {
      const int lo = i * per, m = std::min(per, n - lo);
      if (m <= 0) break;
      Engine& E = serial ? e : sub(i);
      if (i > 0 && !serial) {
        E.begin_run();
        CU(cudaStreamWaitEvent(E.stream, ev_fork, 0));
      }
      E.stats.launches = 0;
      int rc = enqueue_one(E, a + (size_t)lo * in_bytes, b ? b + (size_t)lo * in_bytes : nullptr, m, out + (size_t)lo * out_elems);
      if (rc) return rc;
      launches += E.stats.launches;     // kernels of this enqueue() call (all sub-batches)
      if (i > 0 && !serial) {
        CU(cudaEventRecord(ev_join[i - 1], E.stream));
        CU(cudaStreamWaitEvent(e.stream, ev_join[i - 1], 0));
      }
    }
    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) return fail(DISB200_ECUDA, std::string("kernel launch: ") + cudaGetErrorString(ce));
    return DISB200_OK;
  }

  // Sequence with temporal warm start (oflow.cpp:213-217; the caller the reference ships disabled, run_dense.cpp:354-379):
  // every frame's pyramid is built once, then the pairs run ONE AFTER ANOTHER on the handle's stream - pair i starts its
  // coarsest level from pair i-1's result (its level-sc_l flow, box-averaged down to half the coarsest resolution and
  // rescaled to that grid's units), pair 0 from zero like a cold start.
  int enqueue_warm(const uint8_t* frames, int n, float* out) {
    Engine& E = e;
    if (!E.initflow_ok()) return fail(DISB200_EINVAL, E.initflow_why());
    if (n > cap) return fail(DISB200_EINVAL, "warm-started sequence chunk exceeds the engine capacity");
    E.stats.launches = 0;
    E.build_pyramids(frames, nullptr, n, w0, h0, padl, padt);
    const LevelGeom& gl = E.geom[E.p.sc_l];
    const LevelGeom& gc = E.geom[E.p.sc_f];
    const int k = 1 << (E.p.sc_f + 1 - E.p.sc_l);
    const size_t out_elems = (lowres ? low_bytes : out_bytes) / sizeof(float);
    int rc = DISB200_OK;
    for (int i = 0; i < n && !rc; ++i) {
      E.pair0 = i;
      E.has_initflow = warm_valid;
      rc = E.run_levels(1);
      if (rc) break;
      launch_warm_init(E.flow[0][E.p.sc_l], gl.w, E.nop, k, 1.0f / (float)k, gc.w / 2, gc.h / 2, E.initflow, E.stream, &E.stats);
      warm_valid = true;
      if (lowres) {
        CU(cudaMemcpyAsync(out + i * out_elems, E.flow[0][E.p.sc_l], low_bytes, cudaMemcpyDeviceToDevice, E.stream));
      } else {
        Engine::StageScope sc(&E, 4);
        launch_upsample_crop(E.flow[0][E.p.sc_l], gl, E.nop, E.p.sc_l, padl, padt, w0, h0, out + i * out_elems, 1, E.stream, &E.stats);
      }
    }
    E.pair0 = 0;
    E.has_initflow = false;
    launches = E.stats.launches;
    return rc;
  }
  // the most recent device-pointer call, kept so that disb200_sync can rerun it without bands (DISB200_RETRY_NO_BANDS)
  struct { int kind = 0; const uint8_t* a = nullptr; const uint8_t* b = nullptr; int n = 0; float* out = nullptr; } last_dev;
  bool warm_valid = false;   // E.initflow holds the previous pair's result (within one warm-started sequence)

  // device-side error flags, read after the work has been synchronised
  int check_device_flags() {
    for (int i = 0; i <= (int)extra.size(); ++i) {
      Engine& E = sub(i);
      int rc = E.check_sorb();
      if (rc) return rc;
    }
    return DISB200_OK;
  }
};

// clears DisHandle::lowres when the entry point returns, whatever the path
struct LowresScope {
  DisHandle* h;
  LowresScope(DisHandle* h_, bool on) : h(h_) { if (h) h->lowres = on; }
  ~LowresScope() { if (h) h->lowres = false; }
};

// The host-pointer pipeline shared by every host entry point.  Chunks are pipelined over three streams (upload, compute,
// download) with two staging buffers, so the PCIe transfers of neighbouring chunks hide the compute and each other.
// About eight chunks per call keep the un-overlapped head (first upload + first compute) small; chunks below 16 pairs
// would starve the one-CTA-per-pair kernels.  `upload(i0, m, b, stream)` enqueues the host->device copies of pairs
// i0 .. i0+m-1 into staging buffer b, `download` the device->host copies of their results; seq: dA[b] holds m + 1 frames.
template <class Up, class Down>
static int run_pipelined(DisHandle* h, int n, bool lowres, bool seq, Up upload, Down download) {
  LowresScope ls(h, lowres);
  CU(cudaSetDevice(h->e.device));
  int rc = h->ensure_staging();
  if (rc) return rc;
  h->e.begin_run();
  const int want = h->host_chunk > 0 ? h->host_chunk : std::max(16, (n + 7) / 8);
  const int mb = std::max(1, std::min(h->max_batch, want));
  int k = 0, total_launches = 0;
  for (int i0 = 0; i0 < n; i0 += mb, ++k) {
    const int m = std::min(mb, n - i0), b = k & 1;
    // input buffer b is free once the compute of chunk k-2 is done
    if (k >= 2) CU(cudaStreamWaitEvent(h->s_h2d, h->ev_comp[b], 0));
    rc = upload(i0, m, b, h->s_h2d);
    if (rc) return rc;
    CU(cudaEventRecord(h->ev_h2d[b], h->s_h2d));
    CU(cudaStreamWaitEvent(h->e.stream, h->ev_h2d[b], 0));
    // output buffer b is free once the download of chunk k-2 is done
    if (k >= 2) CU(cudaStreamWaitEvent(h->e.stream, h->ev_d2h[b], 0));
    rc = h->enqueue(h->dA[b], seq ? nullptr : h->dB[b], m, h->dOut[b]);
    if (rc) return rc;
    total_launches += h->launches;
    CU(cudaEventRecord(h->ev_comp[b], h->e.stream));
    CU(cudaStreamWaitEvent(h->s_d2h, h->ev_comp[b], 0));
    rc = download(i0, m, b, h->s_d2h);
    if (rc) return rc;
    CU(cudaEventRecord(h->ev_d2h[b], h->s_d2h));
  }
  CU(cudaStreamSynchronize(h->s_d2h));
  CU(cudaStreamSynchronize(h->e.stream));
  h->launches = total_launches;
  h->e.collect_stage_times();
  return h->check_device_flags();
}

// a host entry point whose banded SOR lost co-residency runs once more (the engine has switched bands off by then)
template <class F>
static int with_band_retry(F call) {
  int rc = call();
  if (rc == DISB200_RETRY_NO_BANDS) rc = call();
  return rc == DISB200_RETRY_NO_BANDS ? fail(DISB200_ECUDA, "banded SOR timed out twice") : rc;
}

extern "C" {

const char* disb200_last_error(void) { return g_err.c_str(); }
const char* disb200_version(void) { return "disb200 0.1 (sm_100a)"; }

int disb200_set_tuning(const DisTuning* t) {
  if (!t) { g_tuning = DisTuning{0, 0, 0, 0, 0, 0}; return DISB200_OK; }
  if (t->streams < 0 || t->host_chunk < 0 || t->tv_schedule < 0 || t->tv_schedule > 5 || t->sorb_max < 0 || t->patch_map < 0 || t->patch_map > 1 || t->antiphase < 0 || t->antiphase > 1)
    return fail(DISB200_EINVAL, "bad tuning value");
  g_tuning = *t;
  return DISB200_OK;
}

int disb200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int disb200_operating_point(int op, int width, int mode, int channels, DisParams* p) {
  if (!p) return fail(DISB200_EINVAL, "null params");
  std::memset(p, 0, sizeof(*p));
  p->dp_thresh = 0.05f; p->dr_thresh = 0.95f; p->res_thresh = 0.0f;      // run_dense.cpp:228-232
  p->usefbcon = 0; p->patnorm = 1; p->costfct = 0;
  p->tv_alpha = 10.0f; p->tv_gamma = 10.0f; p->tv_delta = 5.0f;
  p->tv_innerit = 1; p->tv_solverit = 3; p->tv_sor = 1.6f;
  p->verbosity = 2; p->mode = mode; p->channels = channels; p->sor_mode = DISB200_SOR_LEXICOGRAPHIC;
  const int fratio = 5;
  auto first_scale = [&](int psz) {                                       // run_dense.cpp:181-184
    return std::max(0, (int)std::floor(log2((2.0f * (float)width) / ((float)fratio * (float)psz))));
  };
  switch (op) {                                                           // run_dense.cpp:238-268
    case 1: p->patchsz = 8; p->patove = 0.3f; p->sc_f = first_scale(8); p->sc_l = std::max(p->sc_f - 2, 0); p->max_iter = p->min_iter = 16; p->usetvref = 0; break;
    case 3: p->patchsz = 12; p->patove = 0.75f; p->sc_f = first_scale(12); p->sc_l = std::max(p->sc_f - 4, 0); p->max_iter = p->min_iter = 16; p->usetvref = 1; break;
    case 4: p->patchsz = 12; p->patove = 0.75f; p->sc_f = first_scale(12); p->sc_l = std::max(p->sc_f - 5, 0); p->max_iter = p->min_iter = 128; p->usetvref = 1; break;
    default: p->patchsz = 8; p->patove = 0.4f; p->sc_f = first_scale(8); p->sc_l = std::max(p->sc_f - 2, 0); p->max_iter = p->min_iter = 12; p->usetvref = 1; break;
  }
  return DISB200_OK;
}

int disb200_padded_dims(int w0, int h0, int sc_f, int* W, int* H, int* padl, int* padt) {
  if (w0 <= 0 || h0 <= 0 || sc_f < 0 || sc_f > 16) return fail(DISB200_EINVAL, "bad dimensions");
  const int scfct = 1 << sc_f;                                            // run_dense.cpp:299-305
  int padw = 0, padh = 0;
  if (w0 % scfct) padw = scfct - w0 % scfct;
  if (h0 % scfct) padh = scfct - h0 % scfct;
  if (W) *W = w0 + padw;
  if (H) *H = h0 + padh;
  if (padl) *padl = padw / 2;
  if (padt) *padt = padh / 2;
  return DISB200_OK;
}

int disb200_create(const DisParams* params, int w0, int h0, int max_batch, int device, DisHandle** out) {
  if (!params || !out) return fail(DISB200_EINVAL, "null argument");
  if (max_batch < 1) return fail(DISB200_EINVAL, "max_batch must be >= 1");
  int rc = validate(*params);
  if (rc) return rc;
  if (disb200_device_count() <= device) return fail(DISB200_ECUDA, "no CUDA device " + std::to_string(device) + " (this library has no CPU path)");
  DisHandle* h = new DisHandle();
  int W, H;
  rc = disb200_padded_dims(w0, h0, params->sc_f, &W, &H, &h->padl, &h->padt);
  h->max_batch = max_batch;
  {
    // two sub-batches on two streams overlap the tail of each kernel with the head of the next one of the other
    // half: +3 % at 1480 pairs (234.7 k -> 242.5 k pairs/s); more streams or smaller batches do not pay, and
    // staggering the sub-batches so that they run DIFFERENT kernels side by side loses 3-14 % (profiles/r02a_stagger_sweep.txt)
    const int want = g_tuning.streams > 0 ? g_tuning.streams : (max_batch >= 1184 ? 2 : 1);
    h->nsub = std::max(1, std::min(std::min(want, 16), max_batch));
    h->host_chunk = g_tuning.host_chunk;
    h->duo = g_tuning.antiphase == 1 && h->nsub == 2;
  }
  h->cap = (max_batch + h->nsub - 1) / h->nsub;
  if (!rc) rc = h->e.init(*params, W, H, h->cap, device);
  if (!rc) {
    h->w0 = w0; h->h0 = h0;
    h->in_bytes = (size_t)w0 * h0 * params->channels;
    h->out_bytes = (size_t)w0 * h0 * h->e.nop * sizeof(float);
    h->low_bytes = (size_t)(W >> params->sc_l) * (H >> params->sc_l) * h->e.nop * sizeof(float);
    if (cudaEventCreate(&h->t0) != cudaSuccess || cudaEventCreate(&h->t1) != cudaSuccess) rc = fail(DISB200_ECUDA, "cudaEventCreate failed");
  }
  if (rc) { h->e.destroy(); delete h; return rc; }
  *out = h;
  return DISB200_OK;
}

int disb200_destroy(DisHandle* h) {
  if (!h) return DISB200_OK;
  cudaSetDevice(h->e.device);
  if (h->s_h2d) { cudaStreamSynchronize(h->s_h2d); cudaStreamDestroy(h->s_h2d); }
  if (h->s_d2h) { cudaStreamSynchronize(h->s_d2h); cudaStreamDestroy(h->s_d2h); }
  if (h->staging)
    for (int k = 0; k < 2; ++k) { cudaEventDestroy(h->ev_h2d[k]); cudaEventDestroy(h->ev_comp[k]); cudaEventDestroy(h->ev_d2h[k]); }
  if (h->t0) cudaEventDestroy(h->t0);
  if (h->t1) cudaEventDestroy(h->t1);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  for (int k = 0; k < 2; ++k)
    if (h->ev_lane[k]) cudaEventDestroy(h->ev_lane[k]);
  for (cudaEvent_t ev : h->ev_join) cudaEventDestroy(ev);
  for (Engine* x : h->extra) { x->destroy(); delete x; }
  h->e.destroy();
  delete h;
  return DISB200_OK;
}

int disb200_run_batch_device(DisHandle* h, const uint8_t* dA, const uint8_t* dB, int n, float* dOut) {
  if (!h || !dA || !dB || !dOut) return fail(DISB200_EINVAL, "null argument");
  if (n < 1 || n > h->max_batch) return fail(DISB200_EINVAL, "n must be in 1..max_batch");
  CU(cudaSetDevice(h->e.device));
  h->e.begin_run();
  h->last_dev.kind = 1; h->last_dev.a = dA; h->last_dev.b = dB; h->last_dev.n = n; h->last_dev.out = dOut;
  return h->enqueue(dA, dB, n, dOut);
}

int disb200_sync(DisHandle* h) {
  if (!h) return fail(DISB200_EINVAL, "null handle");
  CU(cudaSetDevice(h->e.device));
  CU(cudaStreamSynchronize(h->e.stream));
  h->e.collect_stage_times();
  int rc = h->check_device_flags();
  if (rc == DISB200_RETRY_NO_BANDS && h->last_dev.kind) {   // bands are off now: the same call again, general wavefront SOR
    h->e.begin_run();
    h->warm_valid = false;
    rc = h->last_dev.kind == 1 ? h->enqueue(h->last_dev.a, h->last_dev.b, h->last_dev.n, h->last_dev.out)
                               : h->enqueue_warm(h->last_dev.a, h->last_dev.n, h->last_dev.out);
    if (rc) return rc;
    CU(cudaStreamSynchronize(h->e.stream));
    h->e.collect_stage_times();
    rc = h->check_device_flags();
  }
  return rc == DISB200_RETRY_NO_BANDS ? fail(DISB200_ECUDA, "banded SOR timed out twice") : rc;
}

static int run_batch_host_impl(DisHandle* h, const uint8_t* A, const uint8_t* B, int n, float* out, bool lowres) {
  if (!h || !A || !B || !out) return fail(DISB200_EINVAL, "null argument");
  if (n < 1) return fail(DISB200_EINVAL, "n must be >= 1");
  const size_t obytes = lowres ? h->low_bytes : h->out_bytes;
  return run_pipelined(h, n, lowres, false,
    [&](int i0, int m, int b, cudaStream_t st) -> int {
      CU(cudaMemcpyAsync(h->dA[b], A + (size_t)i0 * h->in_bytes, h->in_bytes * m, cudaMemcpyHostToDevice, st));
      CU(cudaMemcpyAsync(h->dB[b], B + (size_t)i0 * h->in_bytes, h->in_bytes * m, cudaMemcpyHostToDevice, st));
      return DISB200_OK;
    },
    [&](int i0, int m, int b, cudaStream_t st) -> int {
      CU(cudaMemcpyAsync(out + (size_t)i0 * (obytes / sizeof(float)), h->dOut[b], obytes * m, cudaMemcpyDeviceToHost, st));
      return DISB200_OK;
    });
}

int disb200_run_batch_host(DisHandle* h, const uint8_t* A, const uint8_t* B, int n, float* out) {
  return with_band_retry([&] { return run_batch_host_impl(h, A, B, n, out, false); });
}
int disb200_run_batch_host_lowres(DisHandle* h, const uint8_t* A, const uint8_t* B, int n, float* out_low) {
  return with_band_retry([&] { return run_batch_host_impl(h, A, B, n, out_low, true); });
}

int disb200_lowres_dims(const DisHandle* h, int* w_l, int* h_l, int* nop) {
  if (!h) return fail(DISB200_EINVAL, "null handle");
  const LevelGeom& g = h->e.geom[h->e.p.sc_l];
  if (w_l) *w_l = g.w;
  if (h_l) *h_l = g.h;
  if (nop) *nop = h->e.nop;
  return DISB200_OK;
}

int disb200_run_sequence_device(DisHandle* h, const uint8_t* dFrames, int nframes, float* dOut) {
  if (!h || !dFrames || !dOut) return fail(DISB200_EINVAL, "null argument");
  if (nframes < 2 || nframes - 1 > h->max_batch) return fail(DISB200_EINVAL, "nframes - 1 must be in 1..max_batch");
  CU(cudaSetDevice(h->e.device));
  h->e.begin_run();
  h->last_dev.kind = 1; h->last_dev.a = dFrames; h->last_dev.b = nullptr; h->last_dev.n = nframes - 1; h->last_dev.out = dOut;
  return h->enqueue(dFrames, nullptr, nframes - 1, dOut);
}

static int run_sequence_host_impl(DisHandle* h, const uint8_t* frames, int nframes, float* out, bool lowres) {
  if (!h || !frames || !out) return fail(DISB200_EINVAL, "null argument");
  if (nframes < 2) return fail(DISB200_EINVAL, "a sequence needs at least two frames");
  const size_t obytes = lowres ? h->low_bytes : h->out_bytes;
  return run_pipelined(h, nframes - 1, lowres, true,
    [&](int i0, int m, int b, cudaStream_t st) -> int {
      // pairs i0 .. i0+m-1 need frames i0 .. i0+m (the last one is uploaded again with the next chunk)
      CU(cudaMemcpyAsync(h->dA[b], frames + (size_t)i0 * h->in_bytes, h->in_bytes * (m + 1), cudaMemcpyHostToDevice, st));
      return DISB200_OK;
    },
    [&](int i0, int m, int b, cudaStream_t st) -> int {
      CU(cudaMemcpyAsync(out + (size_t)i0 * (obytes / sizeof(float)), h->dOut[b], obytes * m, cudaMemcpyDeviceToHost, st));
      return DISB200_OK;
    });
}

int disb200_run_sequence_host(DisHandle* h, const uint8_t* frames, int nframes, float* out) {
  return with_band_retry([&] { return run_sequence_host_impl(h, frames, nframes, out, false); });
}
int disb200_run_sequence_host_lowres(DisHandle* h, const uint8_t* frames, int nframes, float* out_low) {
  return with_band_retry([&] { return run_sequence_host_impl(h, frames, nframes, out_low, true); });
}

int disb200_run_sequence_warm_device(DisHandle* h, const uint8_t* dFrames, int nframes, float* dOut) {
  if (!h || !dFrames || !dOut) return fail(DISB200_EINVAL, "null argument");
  if (nframes < 2 || nframes - 1 > h->cap) return fail(DISB200_EINVAL, "nframes - 1 must be in 1..max_batch (one sub-batch)");
  CU(cudaSetDevice(h->e.device));
  h->e.begin_run();
  h->warm_valid = false;
  h->last_dev.kind = 2; h->last_dev.a = dFrames; h->last_dev.b = nullptr; h->last_dev.n = nframes - 1; h->last_dev.out = dOut;
  return h->enqueue_warm(dFrames, nframes - 1, dOut);
}

static int run_sequence_warm_host_impl(DisHandle* h, const uint8_t* frames, int nframes, float* out) {
  if (!h || !frames || !out) return fail(DISB200_EINVAL, "null argument");
  if (nframes < 2) return fail(DISB200_EINVAL, "a sequence needs at least two frames");
  CU(cudaSetDevice(h->e.device));
  int rc = h->ensure_staging();
  if (rc) return rc;
  h->e.begin_run();
  h->warm_valid = false;
  const int n = nframes - 1, mb = h->cap;
  int total_launches = 0;
  // the pairs are serial by construction (pair i needs pair i-1's result): chunks only bound the device staging
  for (int i0 = 0; i0 < n; i0 += mb) {
    const int m = std::min(mb, n - i0);
    CU(cudaMemcpyAsync(h->dA[0], frames + (size_t)i0 * h->in_bytes, h->in_bytes * (m + 1), cudaMemcpyHostToDevice, h->e.stream));
    rc = h->enqueue_warm(h->dA[0], m, h->dOut[0]);
    if (rc) return rc;
    total_launches += h->launches;
    CU(cudaMemcpyAsync(out + (size_t)i0 * (h->out_bytes / sizeof(float)), h->dOut[0], h->out_bytes * m, cudaMemcpyDeviceToHost, h->e.stream));
    CU(cudaStreamSynchronize(h->e.stream));
  }
  h->launches = total_launches;
  h->e.collect_stage_times();
  return h->check_device_flags();
}

int disb200_run_sequence_warm_host(DisHandle* h, const uint8_t* frames, int nframes, float* out) {
  return with_band_retry([&] { return run_sequence_warm_host_impl(h, frames, nframes, out); });
}

// Pointer lists (the entry the four run_* CLIs and INTEGRATION.md use): same three-stream pipeline, one copy per image.
int disb200_run_batch(DisHandle* h, const uint8_t* const* imgA, const uint8_t* const* imgB, int n, float* const* out) {
  if (!h || !imgA || !imgB || !out) return fail(DISB200_EINVAL, "null argument");
  if (n < 1) return fail(DISB200_EINVAL, "n must be >= 1");
  for (int i = 0; i < n; ++i)
    if (!imgA[i] || !imgB[i] || !out[i]) return fail(DISB200_EINVAL, "null image / output pointer in the list");
  return with_band_retry([&] {
    return run_pipelined(h, n, false, false,
      [&](int i0, int m, int b, cudaStream_t st) -> int {
        for (int i = 0; i < m; ++i) {
          CU(cudaMemcpyAsync(h->dA[b] + (size_t)i * h->in_bytes, imgA[i0 + i], h->in_bytes, cudaMemcpyHostToDevice, st));
          CU(cudaMemcpyAsync(h->dB[b] + (size_t)i * h->in_bytes, imgB[i0 + i], h->in_bytes, cudaMemcpyHostToDevice, st));
        }
        return DISB200_OK;
      },
      [&](int i0, int m, int b, cudaStream_t st) -> int {
        for (int i = 0; i < m; ++i)
          CU(cudaMemcpyAsync(out[i0 + i], h->dOut[b] + (size_t)i * (h->out_bytes / sizeof(float)), h->out_bytes, cudaMemcpyDeviceToHost, st));
        return DISB200_OK;
      });
  });
}

int disb200_timer_begin(DisHandle* h) {
  if (!h) return fail(DISB200_EINVAL, "null handle");
  CU(cudaSetDevice(h->e.device));
  CU(cudaEventRecord(h->t0, h->e.stream));
  return DISB200_OK;
}
int disb200_timer_end(DisHandle* h, float* ms) {
  if (!h || !ms) return fail(DISB200_EINVAL, "null argument");
  CU(cudaSetDevice(h->e.device));
  CU(cudaEventRecord(h->t1, h->e.stream));
  CU(cudaEventSynchronize(h->t1));
  CU(cudaEventElapsedTime(ms, h->t0, h->t1));
  return DISB200_OK;
}
int disb200_last_launch_count(const DisHandle* h) { return h ? h->launches : 0; }
int disb200_set_stage_timing(DisHandle* h, int on) {
  if (!h) return fail(DISB200_EINVAL, "null handle");
  h->e.stage_timing = on != 0;
  return DISB200_OK;
}
int disb200_stage_ms(const DisHandle* h, int stage, float* ms) {
  if (!h || !ms || stage < 0 || stage > 7) return fail(DISB200_EINVAL, "bad argument");
  *ms = h->e.stage_ms[stage];
  return DISB200_OK;
}

int disb200_scale_ms(const DisHandle* h, int level, float* search_ms, float* densify_ms, float* refine_ms, int* npatches) {
  if (!h) return fail(DISB200_EINVAL, "null handle");
  if (level < h->e.p.sc_l || level > h->e.p.sc_f) return fail(DISB200_EINVAL, "level outside sc_l..sc_f");
  if (search_ms) *search_ms = h->e.scale_ms[level][0];
  if (densify_ms) *densify_ms = h->e.scale_ms[level][1];
  if (refine_ms) *refine_ms = h->e.scale_ms[level][2];
  if (npatches) *npatches = h->e.geom[level].npatch;
  return DISB200_OK;
}

int disb200_fetch(DisHandle* h, int kind, int which, int level, int pair, void* dst, size_t nbytes) {
  if (!h || !dst) return fail(DISB200_EINVAL, "null argument");
  if (pair < 0 || pair >= h->max_batch || which < 0 || which > 1 || h->last_per < 1) return fail(DISB200_EINVAL, "bad pair / which");
  int si = pair / h->last_per;
  if (si > (int)h->extra.size()) return fail(DISB200_EINVAL, "pair was not part of the last run");
  pair -= si * h->last_per;
  if (h->e.stage_timing && h->nsub > 1) {
    // the per-stage stopwatch runs the sub-batches one after another on engine 0: only the last one's intermediates exist
    const int last_si = (h->last_n - 1) / h->last_per;
    if (si != last_si) return fail(DISB200_EINVAL, "with stage timing on, only pairs of the last sub-batch can be fetched");
    si = 0;
  }
  CU(cudaSetDevice(h->e.device));
  CU(cudaStreamSynchronize(h->e.stream));
  Engine& e = h->sub(si);
  const void* src = nullptr;
  size_t bytes = 0;
  if (kind >= 0 && kind <= 3) {
    if (level < e.p.sc_l || level > e.p.sc_f) return fail(DISB200_EINVAL, "level outside sc_l..sc_f");
    const LevelGeom& g = e.geom[level];
    if (kind <= 2) {
      bytes = g.img_elems * sizeof(float);
      src = e.P(which, kind, level) + (size_t)pair * g.img_elems;
    } else {
      if (!e.flow[which][level]) return fail(DISB200_EINVAL, "no backward flow (usefbcon = 0)");
      bytes = (size_t)g.w * g.h * e.nop * sizeof(float);
      src = e.flow[which][level] + (size_t)pair * g.w * g.h * e.nop;
    }
  } else if (kind >= 4 && kind <= 6) {
    const LevelGeom& g = e.geom[e.p.sc_l];
    if (!e.p_iter[which]) return fail(DISB200_EINVAL, "no backward grid (usefbcon = 0)");
    if (kind == 4) { bytes = (size_t)g.npatch * e.nop * sizeof(float); src = e.p_iter[which] + (size_t)pair * g.npatch * e.nop; }
    if (kind == 5) { bytes = (size_t)g.npatch * 2 * sizeof(float); src = e.pt_iter[which] + (size_t)pair * g.npatch * 2; }
    if (kind == 6) { bytes = (size_t)g.npatch * e.o.novals * sizeof(float); src = e.pweight[which] + (size_t)pair * g.npatch * e.o.novals; }
  } else {
    return fail(DISB200_EINVAL, "unknown kind");
  }
  if (bytes != nbytes) return fail(DISB200_EINVAL, "nbytes is " + std::to_string(nbytes) + ", object has " + std::to_string(bytes));
  CU(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
  return DISB200_OK;
}

// OFClass on an existing engine (pair slot 0): upload the padded pyramids, run the scale loop, download outflow
static int ofclass_on_engine(Engine& e, const float* const* im_ao, const float* const* im_ao_dx, const float* const* im_ao_dy,
                             const float* const* im_bo, const float* const* im_bo_dx, const float* const* im_bo_dy,
                             float* outflow, const float* initflow) {
  const DisParams& p = e.p;
  const float* const* src[2][3] = {{im_ao, im_ao_dx, im_ao_dy}, {im_bo, im_bo_dx, im_bo_dy}};
  for (int l = p.sc_l; l <= p.sc_f; ++l)
    for (int f = 0; f < 2; ++f)
      for (int a = 0; a < 3; ++a) {
        if (!src[f][a][l]) return fail(DISB200_EINVAL, "pyramid level " + std::to_string(l) + " is null");
        CU(cudaMemcpyAsync(e.pyr[f][a][l], src[f][a][l], e.geom[l].img_elems * sizeof(float), cudaMemcpyHostToDevice, e.stream));
      }
  if (initflow) {
    if (!e.initflow_ok()) return fail(DISB200_EINVAL, Engine::initflow_why());
    const LevelGeom& gc = e.geom[p.sc_f];
    CU(cudaMemcpyAsync(e.initflow, initflow, (size_t)(gc.w / 2) * (gc.h / 2) * e.nop * sizeof(float), cudaMemcpyHostToDevice, e.stream));
  }
  int rc = DISB200_OK;
  for (int attempt = 0; attempt < 2; ++attempt) {   // second attempt only after a banded-SOR timeout (bands off by then)
    e.seq = false;
    e.pair0 = 0;
    e.has_initflow = initflow != nullptr;
    e.begin_run();
    rc = e.run_levels(1);
    e.has_initflow = false;
    if (rc) return rc;
    const LevelGeom& g = e.geom[p.sc_l];
    CU(cudaMemcpyAsync(outflow, e.flow[0][p.sc_l], (size_t)g.w * g.h * e.nop * sizeof(float), cudaMemcpyDeviceToHost, e.stream));
    CU(cudaStreamSynchronize(e.stream));
    rc = e.check_sorb();
    if (rc != DISB200_RETRY_NO_BANDS) break;
  }
  return rc == DISB200_RETRY_NO_BANDS ? fail(DISB200_ECUDA, "banded SOR timed out twice") : rc;
}

int disb200_ofclass(const DisParams* params, const float* const* im_ao, const float* const* im_ao_dx,
                    const float* const* im_ao_dy, const float* const* im_bo, const float* const* im_bo_dx,
                    const float* const* im_bo_dy, int W, int H, float* outflow, const float* initflow) {
  if (!params || !im_ao || !im_ao_dx || !im_ao_dy || !im_bo || !im_bo_dx || !im_bo_dy || !outflow)
    return fail(DISB200_EINVAL, "null argument");
  int rc = validate(*params);
  if (rc) return rc;
  if (disb200_device_count() < 1) return fail(DISB200_ECUDA, "no CUDA device (this library has no CPU path)");
  Engine e;
  rc = e.init(*params, W, H, 1, 0);
  if (!rc) rc = ofclass_on_engine(e, im_ao, im_ao_dx, im_ao_dy, im_bo, im_bo_dx, im_bo_dy, outflow, initflow);
  e.destroy();
  return rc;
}

int disb200_ofclass_handle(DisHandle* h, const float* const* im_ao, const float* const* im_ao_dx, const float* const* im_ao_dy,
                           const float* const* im_bo, const float* const* im_bo_dx, const float* const* im_bo_dy,
                           float* outflow, const float* initflow) {
  if (!h || !im_ao || !im_ao_dx || !im_ao_dy || !im_bo || !im_bo_dx || !im_bo_dy || !outflow)
    return fail(DISB200_EINVAL, "null argument");
  CU(cudaSetDevice(h->e.device));
  CU(cudaStreamSynchronize(h->e.stream));
  h->last_dev.kind = 0;
  return ofclass_on_engine(h->e, im_ao, im_ao_dx, im_ao_dy, im_bo, im_bo_dx, im_bo_dy, outflow, initflow);
}

int disb200_aggregate_dense(const DisParams* params, int w_l, int h_l, const float* p_iter_h, const float* pweight_h,
                            const float* cg_p_iter, const float* cg_pt_iter, const float* cg_pweight, float* flow,
                            int* npatch_out) {
  if (!params) return fail(DISB200_EINVAL, "null argument");
  int rc = validate(*params);
  if (rc) return rc;
  if (w_l < 1 || h_l < 1) return fail(DISB200_EINVAL, "bad level size");
  DisParams q = *params;
  q.sc_f = q.sc_l = 0;
  q.usetvref = 0;
  const bool fb = cg_p_iter || cg_pt_iter || cg_pweight;
  q.usefbcon = fb ? 1 : 0;
  const LevelGeom g0 = make_levelgeom(q, w_l, h_l, 0);
  if (npatch_out) *npatch_out = g0.npatch;
  if (!p_iter_h && !pweight_h && !flow && !fb) return DISB200_OK;   // size query
  if (!p_iter_h || !pweight_h || !flow || (fb && !(cg_p_iter && cg_pt_iter && cg_pweight)))
    return fail(DISB200_EINVAL, "null argument");
  if (disb200_device_count() < 1) return fail(DISB200_ECUDA, "no CUDA device (this library has no CPU path)");
  Engine e;
  rc = e.init(q, w_l, h_l, 1, 0);
  if (!rc) {
    const LevelGeom& g = e.geom[0];
    const size_t np = g.npatch, nv = e.o.novals;
    cudaError_t ce = cudaMemcpyAsync(e.p_iter[0], p_iter_h, np * e.nop * sizeof(float), cudaMemcpyHostToDevice, e.stream);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(e.pweight[0], pweight_h, np * nv * sizeof(float), cudaMemcpyHostToDevice, e.stream);
    if (fb && ce == cudaSuccess) ce = cudaMemcpyAsync(e.p_iter[1], cg_p_iter, np * e.nop * sizeof(float), cudaMemcpyHostToDevice, e.stream);
    if (fb && ce == cudaSuccess) ce = cudaMemcpyAsync(e.pt_iter[1], cg_pt_iter, np * 2 * sizeof(float), cudaMemcpyHostToDevice, e.stream);
    if (fb && ce == cudaSuccess) ce = cudaMemcpyAsync(e.pweight[1], cg_pweight, np * nv * sizeof(float), cudaMemcpyHostToDevice, e.stream);
    if (ce == cudaSuccess) {
      e.begin_run();
      DensifyArgs a;
      a.p_iter = e.p_iter[0]; a.pweight = e.pweight[0];
      a.cg_p_iter = fb ? e.p_iter[1] : nullptr; a.cg_pt_iter = fb ? e.pt_iter[1] : nullptr; a.cg_pweight = fb ? e.pweight[1] : nullptr;
      a.flow = e.flow[0][0];
      a.cell_count = e.cell_count; a.cell_start = e.cell_start; a.cell_end = e.cell_end; a.cell_list = e.cell_list;
      launch_densify(a, g, e.o, 1, e.stream, &e.stats);
      ce = cudaGetLastError();
    }
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(flow, e.flow[0][0], (size_t)w_l * h_l * e.nop * sizeof(float), cudaMemcpyDeviceToHost, e.stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(e.stream);
    if (ce != cudaSuccess) rc = fail(DISB200_ECUDA, std::string("aggregate_dense: ") + cudaGetErrorString(ce));
  }
  e.destroy();
  return rc;
}

int disb200_varref_level(const DisParams* params, const float* im_a, const float* im_b, int w_l, int h_l, int level,
                         int camlr, float* flow) {
  if (!params || !im_a || !im_b || !flow) return fail(DISB200_EINVAL, "null argument");
  int rc = validate(*params);
  if (rc) return rc;
  if (disb200_device_count() < 1) return fail(DISB200_ECUDA, "no CUDA device (this library has no CPU path)");
  if (w_l < 1 || h_l < 2) return fail(DISB200_EUNSUPPORTED, "level must be at least 1 x 2 pixels (the reference crashes on one-row levels)");
  // an engine whose only level is `level`-like: reuse the allocator with sc_f = sc_l = 0 on a w_l x h_l frame
  DisParams q = *params;
  q.sc_f = q.sc_l = 0;
  q.usefbcon = 0;
  q.usetvref = 1;
  Engine e;
  rc = e.init(q, w_l, h_l, 1, 0);
  if (!rc) {
    LevelGeom g = e.geom[0];
    g.level = level;
    const size_t fb = (size_t)w_l * h_l * e.nop * sizeof(float);
    cudaError_t ce = cudaMemcpyAsync(e.pyr[0][0][0], im_a, g.img_elems * sizeof(float), cudaMemcpyHostToDevice, e.stream);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(e.pyr[1][0][0], im_b, g.img_elems * sizeof(float), cudaMemcpyHostToDevice, e.stream);
    float* d_in = nullptr;   // the flow is refined in place: keep the input for a rerun without bands
    if (ce == cudaSuccess) ce = cudaMalloc(&d_in, fb);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(d_in, flow, fb, cudaMemcpyHostToDevice, e.stream);
    for (int attempt = 0; attempt < 2 && ce == cudaSuccess; ++attempt) {
      ce = cudaMemcpyAsync(e.flow[0][0], d_in, fb, cudaMemcpyDeviceToDevice, e.stream);
      if (ce != cudaSuccess) break;
      e.begin_run();
      launch_varref_level(e.pyr[0][0][0], e.pyr[1][0][0], e.flow[0][0], g, *params, camlr, e.vb, 1, e.stream, &e.stats);
      ce = cudaGetLastError();
      if (ce == cudaSuccess) ce = cudaStreamSynchronize(e.stream);
      if (ce != cudaSuccess) break;
      rc = e.check_sorb();
      if (rc != DISB200_RETRY_NO_BANDS) break;
    }
    if (ce == cudaSuccess && !rc) ce = cudaMemcpy(flow, e.flow[0][0], fb, cudaMemcpyDeviceToHost);
    if (d_in) cudaFree(d_in);
    if (ce != cudaSuccess) rc = fail(DISB200_ECUDA, std::string("varref_level: ") + cudaGetErrorString(ce));
    if (rc == DISB200_RETRY_NO_BANDS) rc = fail(DISB200_ECUDA, "banded SOR timed out twice");
  }
  e.destroy();
  return rc;
}

}  // extern "C"
