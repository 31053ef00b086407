// Spatial tiling of ONE frame pair over several GPUs of one node (north_star: "4K/8K frames are
// additionally tiled spatially with NVLink P2P halo exchange").  One host thread drives ndev devices, one Engine (engine.h)
// per device, peer access enabled between all of them.
//
// Layout: every device holds the FULL-size, globally indexed arrays of every level (pyramids, patch results, flows,
// refinement planes) - a 4K level is a few hundred MB of the 180 GB - and computes only its part; what another device
// needs is copied into the same place of that device's arrays (cudaMemcpyAsync over NVLink), so no kernel ever
// translates an index.  Per level:
//   * levels of up to 96 rows (one-warp SOR) or with fewer bands than devices are REPLICATED: every device computes the
//     whole level from identical inputs, nothing is exchanged;
//   * taller levels are TILED into strips of whole 32-row bands:
//       inverse search      patches owned by the row of their reference point (patch rows gy0 .. gy0+ngy-1), then an
//                           all-gather of p_iter / pt_iter / pweight (the forward-backward merge splats complementary
//                           patches around their END positions, which may lie in any strip)
//       densification       the device's rows plus the 4 halo rows the refinement's stencils reach (recomputed, not sent)
//       refinement          launch_varref_level_tiled (varref.cu): halo rows of the planes recomputed locally, two rows of
//                           the increment exchanged per fixed-point iteration, the lexicographic SOR as ONE wavefront
//                           over all devices - the banded kernel, whose tagged boundary rows cross a device cut as plain
//                           stores into the neighbour's memory
//       all-gather of the refined flow rows (the next finer level starts its patches from them)
//   * upsample + crop by output rows, each device sends its rows of the result to the host.
// Every value is computed by the same kernel, from the same inputs, in the same order as on one device: the result is
// bit-identical to the one-device handle (tests/test_parity_gpu.py, tests/test_tiled_multi_gpu.py).
// `devices` may name one physical device several times ("virtual" strips on one GPU): same code path, used by the
// single-GPU test suite.
#include <cstring>

#include "engine.h"

using namespace disb200;

struct DisTiled {
  DisParams p;
  int w0 = 0, h0 = 0, W = 0, H = 0, padl = 0, padt = 0, nop = 2, ndev = 0;
  std::vector<int> devices;
  std::vector<Engine*> eng;
  std::vector<uint8_t*> dA, dB;
  std::vector<float*> dOut;
  std::vector<cudaEvent_t> ev_bar, ev_sor, ev_halo;
  cudaEvent_t t0 = nullptr, t1 = nullptr;
  struct Plan {
    bool tiled = false;
    std::vector<int> row;   // ndev + 1 strip boundaries (pixel rows, multiples of 32 except the last)
    std::vector<int> gy;    // ndev + 1 patch-row boundaries
  };
  std::vector<Plan> plan;   // indexed by level
  int launches = 0;
  bool frames_resident = false;

  Engine& E(int k) { return *eng[k]; }

  // every device waits until all devices have reached this point of their streams
  int barrier() {
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      CU(cudaEventRecord(ev_bar[k], E(k).stream));
    }
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      for (int j = 0; j < ndev; ++j)
        if (j != k) CU(cudaStreamWaitEvent(E(k).stream, ev_bar[j], 0));
    }
    return DISB200_OK;
  }

  void make_plan() {
    plan.assign(p.sc_f + 1, Plan());
    for (int l = p.sc_l; l <= p.sc_f; ++l) {
      const LevelGeom& g = E(0).geom[l];
      Plan& pl = plan[l];
      pl.row.assign(ndev + 1, 0);
      pl.gy.assign(ndev + 1, 0);
      const bool tv_ok = !p.usetvref || varref_level_tileable(g, p, ndev, E(0).vb.plane_cap, nop);
      const int nbands = (g.h + 31) / 32;
      pl.tiled = ndev > 1 && g.h > 96 && nbands >= ndev && tv_ok;
      for (int k = 0; k <= ndev; ++k) {
        pl.row[k] = pl.tiled ? std::min(g.h, (int)((long long)k * nbands / ndev) * 32) : (k == 0 ? 0 : g.h);
        if (k == ndev) pl.row[k] = g.h;
        // patches are owned by the strip that holds their reference row gy * steps + offh
        int b = (pl.row[k] - g.offh + g.steps - 1) / g.steps;
        if (pl.row[k] - g.offh <= 0) b = 0;
        pl.gy[k] = k == 0 ? 0 : (k == ndev ? g.noph : std::max(0, std::min(g.noph, b)));
      }
    }
  }

  // pull from every other device what it computed: patch rows gy[j] .. gy[j+1]-1 of the column-major patch arrays
  int gather_patches(int sl) {
    const Plan& pl = plan[sl];
    const LevelGeom& g = E(0).geom[sl];
    const int ndir = p.usefbcon ? 2 : 1;
    const size_t nv = E(0).o.novals;
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      for (int j = 0; j < ndev; ++j) {
        if (j == k) continue;
        const int gy0 = pl.gy[j], ngy = pl.gy[j + 1] - pl.gy[j];
        if (ngy <= 0) continue;
        for (int d = 0; d < ndir; ++d) {
          struct { float* dst; const float* src; size_t e; } arr[3] = {
              {E(k).p_iter[d], E(j).p_iter[d], (size_t)nop}, {E(k).pt_iter[d], E(j).pt_iter[d], 2}, {E(k).pweight[d], E(j).pweight[d], nv}};
          for (auto& a : arr) {
            const size_t pitch = (size_t)g.noph * a.e * sizeof(float), off = (size_t)gy0 * a.e;
            CU(cudaMemcpy2DAsync(a.dst + off, pitch, a.src + off, pitch, (size_t)ngy * a.e * sizeof(float), g.nopw, cudaMemcpyDefault, E(k).stream));
          }
        }
      }
    }
    return DISB200_OK;
  }

  int gather_flow(int sl) {
    const Plan& pl = plan[sl];
    const LevelGeom& g = E(0).geom[sl];
    const int ndir = (p.usefbcon && sl > p.sc_l) ? 2 : 1;
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      for (int j = 0; j < ndev; ++j) {
        if (j == k) continue;
        const int r0 = pl.row[j], r1 = pl.row[j + 1];
        if (r1 <= r0) continue;
        for (int d = 0; d < ndir; ++d) {
          const size_t off = (size_t)r0 * g.w * nop;
          CU(cudaMemcpyAsync(E(k).flow[d][sl] + off, E(j).flow[d][sl] + off, (size_t)(r1 - r0) * g.w * nop * sizeof(float), cudaMemcpyDefault, E(k).stream));
        }
      }
    }
    return DISB200_OK;
  }

  int run_level_tiled(int sl) {
    const Plan& pl = plan[sl];
    int rc;
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      LevelGeom g = E(k).geom[sl];
      g.gy0 = pl.gy[k]; g.ngy = pl.gy[k + 1] - pl.gy[k];
      if ((rc = E(k).search_level(sl, 1, g))) return rc;
    }
    if ((rc = barrier())) return rc;
    if ((rc = gather_patches(sl))) return rc;
    const int halo = p.usetvref ? 4 : 0;
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      const int y0 = std::max(0, pl.row[k] - halo), y1 = std::min(E(k).geom[sl].h, pl.row[k + 1] + halo);
      E(k).densify_level(sl, 1, y0, y1 - y0);
    }
    if (p.usetvref) {
      const int ndir = (p.usefbcon && sl > p.sc_l) ? 2 : 1;
      for (int d = 0; d < ndir; ++d) {
        std::vector<TileDev> dv(ndev);
        for (int k = 0; k < ndev; ++k) {
          Engine& e = E(k);
          dv[k].device = devices[k];
          dv[k].stream = e.stream;
          dv[k].im_a = e.P(d, 0, sl);
          dv[k].im_b = e.P(1 - d, 0, sl);
          dv[k].flow = e.flow[d][sl];
          dv[k].vb = e.vb;
          dv[k].g = e.geom[sl];
          dv[k].g.row0 = pl.row[k];
          dv[k].g.row1 = pl.row[k + 1];
          dv[k].st = &e.stats;
          dv[k].ev_sor = ev_sor[k];
          dv[k].ev_halo = ev_halo[k];
        }
        // no device may start this call's SOR (a new launch epoch in the shared exchange rows) while a neighbour's bands
        // of the previous call (the other direction) are still running
        if ((rc = barrier())) return rc;
        const int ce = launch_varref_level_tiled(dv.data(), ndev, p, d);
        if (ce) return fail(DISB200_ECUDA, std::string("tiled refinement: ") + cudaGetErrorString((cudaError_t)ce));
      }
    }
    if ((rc = barrier())) return rc;
    if ((rc = gather_flow(sl))) return rc;
    return barrier();   // nobody overwrites patch results / flows another device is still pulling
  }

  // the whole pipeline on frames already in dA / dB of every device; result rows in dOut of every device
  int enqueue() {
    int rc;
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      E(k).begin_run();
      E(k).build_pyramids(dA[k], dB[k], 1, w0, h0, padl, padt);
    }
    for (int sl = p.sc_f; sl >= p.sc_l; --sl) {
      if (plan[sl].tiled) {
        if ((rc = run_level_tiled(sl))) return rc;
      } else {
        for (int k = 0; k < ndev; ++k) {
          CU(cudaSetDevice(devices[k]));
          Engine& e = E(k);
          if ((rc = e.search_level(sl, 1, e.geom[sl]))) return rc;
          e.densify_level(sl, 1);
          e.varref_level(sl, 1);
        }
      }
    }
    launches = 0;
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      Engine& e = E(k);
      const int o0 = out_row(k), o1 = out_row(k + 1);
      Engine::StageScope sc(&e, 4);
      launch_upsample_crop(e.flow[0][p.sc_l], e.geom[p.sc_l], nop, p.sc_l, padl, padt, w0, h0, dOut[k], 1, e.stream, &e.stats, o0, o1 - o0);
      CU(cudaGetLastError());
      launches += e.stats.launches;
    }
    return DISB200_OK;
  }
  int out_row(int k) const { return k >= ndev ? h0 : (int)((long long)k * ((h0 + 7) / 8) / ndev) * 8; }

  int sync_and_check() {
    int rc = DISB200_OK;
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      CU(cudaStreamSynchronize(E(k).stream));
    }
    for (int k = 0; k < ndev; ++k) {
      CU(cudaSetDevice(devices[k]));
      const int r = E(k).check_sorb();
      if (r == DISB200_RETRY_NO_BANDS) rc = fail(DISB200_ECUDA, "tiled SOR: a band timed out waiting for its neighbour (result invalid)");
      else if (r && !rc) rc = r;
    }
    return rc;
  }
};

extern "C" {

int disb200_tiled_create(const DisParams* params, int w0, int h0, int ndev, const int* devices, DisTiled** out) {
  if (!params || !devices || !out) return fail(DISB200_EINVAL, "null argument");
  if (ndev < 1 || ndev > 16) return fail(DISB200_EINVAL, "ndev must be in 1..16");
  int rc = validate(*params);
  if (rc) return rc;
  const int have = disb200_device_count();
  for (int k = 0; k < ndev; ++k)
    if (devices[k] < 0 || devices[k] >= have) return fail(DISB200_ECUDA, "no CUDA device " + std::to_string(devices[k]) + " (this library has no CPU path)");
  // NVLink P2P between every two distinct devices of the set
  for (int i = 0; i < ndev; ++i)
    for (int j = 0; j < ndev; ++j) {
      if (devices[i] == devices[j]) continue;
      int can = 0;
      CU(cudaDeviceCanAccessPeer(&can, devices[i], devices[j]));
      if (!can) return fail(DISB200_EUNSUPPORTED, "devices " + std::to_string(devices[i]) + " and " + std::to_string(devices[j]) + " have no peer access");
      CU(cudaSetDevice(devices[i]));
      const cudaError_t pe = cudaDeviceEnablePeerAccess(devices[j], 0);
      cudaGetLastError();
      if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled)
        return fail(DISB200_ECUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(pe));
    }
  DisTiled* t = new DisTiled();
  t->p = *params;
  t->w0 = w0; t->h0 = h0; t->ndev = ndev;
  t->nop = params->mode == DISB200_MODE_OF ? 2 : 1;
  t->devices.assign(devices, devices + ndev);
  rc = disb200_padded_dims(w0, h0, params->sc_f, &t->W, &t->H, &t->padl, &t->padt);
  const size_t in_bytes = (size_t)w0 * h0 * params->channels, out_elems = (size_t)w0 * h0 * t->nop;
  for (int k = 0; k < ndev && !rc; ++k) {
    Engine* e = new Engine();
    t->eng.push_back(e);
    rc = e->init(*params, t->W, t->H, 1, devices[k]);
    uint8_t *a = nullptr, *b = nullptr;
    float* o = nullptr;
    if (!rc) rc = e->dalloc(&a, in_bytes);
    if (!rc) rc = e->dalloc(&b, in_bytes);
    if (!rc) rc = e->dalloc(&o, out_elems);
    t->dA.push_back(a); t->dB.push_back(b); t->dOut.push_back(o);
    cudaEvent_t e1 = nullptr, e2 = nullptr, e3 = nullptr;
    if (!rc && (cudaEventCreateWithFlags(&e1, cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&e2, cudaEventDisableTiming) != cudaSuccess ||
                cudaEventCreateWithFlags(&e3, cudaEventDisableTiming) != cudaSuccess))
      rc = fail(DISB200_ECUDA, "cudaEventCreate failed");
    t->ev_bar.push_back(e1); t->ev_sor.push_back(e2); t->ev_halo.push_back(e3);
  }
  if (!rc) {
    cudaSetDevice(devices[0]);
    if (cudaEventCreate(&t->t0) != cudaSuccess || cudaEventCreate(&t->t1) != cudaSuccess) rc = fail(DISB200_ECUDA, "cudaEventCreate failed");
  }
  if (!rc) t->make_plan();
  if (rc) { disb200_tiled_destroy(t); return rc; }
  *out = t;
  return DISB200_OK;
}

int disb200_tiled_destroy(DisTiled* t) {
  if (!t) return DISB200_OK;
  for (size_t k = 0; k < t->eng.size(); ++k) {
    cudaSetDevice(t->devices[k]);
    if (t->eng[k]->stream) cudaStreamSynchronize(t->eng[k]->stream);
  }
  for (size_t k = 0; k < t->eng.size(); ++k) {
    cudaSetDevice(t->devices[k]);
    if (k < t->ev_bar.size() && t->ev_bar[k]) cudaEventDestroy(t->ev_bar[k]);
    if (k < t->ev_sor.size() && t->ev_sor[k]) cudaEventDestroy(t->ev_sor[k]);
    if (k < t->ev_halo.size() && t->ev_halo[k]) cudaEventDestroy(t->ev_halo[k]);
    t->eng[k]->destroy();
    delete t->eng[k];
  }
  if (t->t0) cudaEventDestroy(t->t0);
  if (t->t1) cudaEventDestroy(t->t1);
  delete t;
  return DISB200_OK;
}

int disb200_tiled_run_host(DisTiled* t, const uint8_t* A, const uint8_t* B, float* out) {
  if (!t || !A || !B || !out) return fail(DISB200_EINVAL, "null argument");
  const size_t in_bytes = (size_t)t->w0 * t->h0 * t->p.channels;
  for (int k = 0; k < t->ndev; ++k) {
    CU(cudaSetDevice(t->devices[k]));
    CU(cudaMemcpyAsync(t->dA[k], A, in_bytes, cudaMemcpyHostToDevice, t->E(k).stream));
    CU(cudaMemcpyAsync(t->dB[k], B, in_bytes, cudaMemcpyHostToDevice, t->E(k).stream));
  }
  t->frames_resident = true;
  int rc = t->enqueue();
  if (rc) return rc;
  const size_t rowf = (size_t)t->w0 * t->nop;
  for (int k = 0; k < t->ndev; ++k) {
    CU(cudaSetDevice(t->devices[k]));
    const int o0 = t->out_row(k), o1 = t->out_row(k + 1);
    if (o1 > o0)
      CU(cudaMemcpyAsync(out + (size_t)o0 * rowf, t->dOut[k] + (size_t)o0 * rowf, (size_t)(o1 - o0) * rowf * sizeof(float), cudaMemcpyDeviceToHost, t->E(k).stream));
  }
  return t->sync_and_check();
}

int disb200_tiled_time(DisTiled* t, int reps, float* ms_per_pair) {
  if (!t || !ms_per_pair || reps < 1) return fail(DISB200_EINVAL, "bad argument");
  if (!t->frames_resident) return fail(DISB200_EINVAL, "call disb200_tiled_run_host once first: the timed runs reuse its frames on the devices");
  int rc = t->sync_and_check();
  if (rc) return rc;
  // device time on the first device's stream, between two all-device barriers
  if ((rc = t->barrier())) return rc;
  CU(cudaSetDevice(t->devices[0]));
  CU(cudaEventRecord(t->t0, t->E(0).stream));
  for (int r = 0; r < reps; ++r)
    if ((rc = t->enqueue())) return rc;
  if ((rc = t->barrier())) return rc;
  CU(cudaSetDevice(t->devices[0]));
  CU(cudaEventRecord(t->t1, t->E(0).stream));
  if ((rc = t->sync_and_check())) return rc;
  float ms = 0.f;
  CU(cudaEventElapsedTime(&ms, t->t0, t->t1));
  *ms_per_pair = ms / reps;
  return DISB200_OK;
}

int disb200_tiled_plan(const DisTiled* t, int level, int* tiled, int* rows, int* patch_rows) {
  if (!t) return fail(DISB200_EINVAL, "null handle");
  if (level < t->p.sc_l || level > t->p.sc_f) return fail(DISB200_EINVAL, "level outside sc_l..sc_f");
  const DisTiled::Plan& pl = t->plan[level];
  if (tiled) *tiled = pl.tiled ? 1 : 0;
  for (int k = 0; k <= t->ndev; ++k) {
    if (rows) rows[k] = pl.row[k];
    if (patch_rows) patch_rows[k] = pl.gy[k];
  }
  return DISB200_OK;
}

int disb200_tiled_last_launch_count(const DisTiled* t) { return t ? t->launches : 0; }

}  // extern "C"
