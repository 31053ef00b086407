// TEST INFRASTRUCTURE ONLY — not part of the product.
//
// extern "C" wrapper around the UNMODIFIED reference sources, which
// oracle/Makefile compiles by path from /root/reference (never copied) against
// oracle/eigen_shim.  Built four times (SELECTMODE x SELECTCHANNEL) into
// oracle/_ref/libref_{OF,DE}_{INT,RGB}.so.  Entry points:
//   ref_ofclass         -> OFC::OFClass ctor                 (oflow.h:99-126)  the whole path a3..a22
//   ref_patchgrid_level -> PatGridClass on one level, with per-patch dumps
//                          (patchgrid.h:19-44, patch.h:55-87)                  rows a4..a12
//   ref_varref_level    -> VarRefClass on one level          (refine_variational.h:35-37)  rows a13..a22
// The per-level entry points rebuild camparam/optparam exactly as the ctor does
// (oflow.cpp:75-107,139-157).
#include <xmmintrin.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <memory>
#include <string>
#include <thread>
#include <valarray>
#include <vector>

#include <Eigen/Core>

// Test-only access to PatGridClass::pat / PatClass::pc for the per-patch dumps.
#define private public
#include "oflow.h"
#include "patch.h"
#include "patchgrid.h"
#include "refine_variational.h"
#undef private

namespace {

void fill_op(OFC::optparam& op, int sc_f, int sc_l, int max_iter, int min_iter, float dp, float dr,
             float res, int psz, float patove, int usefbcon, int costfct, int noc, int patnorm,
             int usetvref, float tv_alpha, float tv_gamma, float tv_delta, int tv_innerit,
             int tv_solverit, float tv_sor) {
#if (SELECTMODE == 1)
  op.nop = 2;
#else
  op.nop = 1;
#endif
  op.p_samp_s = psz;
  op.outlierthresh = (float)op.p_samp_s / 2;
  op.patove = patove;
  op.sc_f = sc_f;
  op.sc_l = sc_l;
  op.max_iter = max_iter;
  op.min_iter = min_iter;
  op.dp_thresh = dp * dp;
  op.dr_thresh = dr;
  op.res_thresh = res;
  op.steps = std::max(1, (int)floor(op.p_samp_s * (1 - op.patove)));
  op.novals = noc * psz * psz;
  op.usefbcon = usefbcon;
  op.costfct = costfct;
  op.noc = noc;
  op.patnorm = patnorm;
  op.verbosity = 0;
  op.noscales = op.sc_f - op.sc_l + 1;
  op.usetvref = usetvref;
  op.tv_alpha = tv_alpha;
  op.tv_gamma = tv_gamma;
  op.tv_delta = tv_delta;
  op.tv_innerit = tv_innerit;
  op.tv_solverit = tv_solverit;
  op.tv_sor = tv_sor;
  op.normoutlier_tmpbsq = _mm_set1_ps(op.normoutlier * op.normoutlier);
  op.normoutlier_tmp2bsq = _mm_mul_ps(op.normoutlier_tmpbsq, op.twos);
  op.normoutlier_tmp4bsq = _mm_mul_ps(op.normoutlier_tmpbsq, op.fours);
}

void fill_cp(OFC::camparam& cp, int w_l, int h_l, int pad, int psz, int level, int camlr) {
  cp.sc_fct = std::pow(2.0f, -level);
  cp.height = h_l;
  cp.width = w_l;
  cp.imgpadding = pad;
  cp.tmp_lb = -(float)psz / 2;
  cp.tmp_ubw = (float)(w_l + psz / 2 - 2);
  cp.tmp_ubh = (float)(h_l + psz / 2 - 2);
  cp.tmp_w = w_l + 2 * pad;
  cp.tmp_h = h_l + 2 * pad;
  cp.curr_lv = level;
  cp.camlr = camlr;
}

void dump_grid(const OFC::PatGridClass& g, const OFC::optparam& op, float* p_iter, float* pt_iter,
               float* pweight, int* cnt) {
  for (int i = 0; i < g.nopatches; ++i) {
    const OFC::patchstate* pc = g.pat[i]->pc;
    if (p_iter)
      for (int k = 0; k < op.nop; ++k) p_iter[i * op.nop + k] = pc->p_iter[k];
    if (pt_iter) {
      pt_iter[2 * i] = pc->pt_iter[0];
      pt_iter[2 * i + 1] = pc->pt_iter[1];
    }
    if (pweight) std::memcpy(pweight + (size_t)i * op.novals, pc->pweight.data(), sizeof(float) * op.novals);
    if (cnt) cnt[i] = pc->cnt;
  }
}

}  // namespace

extern "C" {

int ref_mode() { return SELECTMODE; }
int ref_channels() { return SELECTCHANNEL; }

void ref_ofclass(const float** a, const float** adx, const float** ady, const float** b,
                 const float** bdx, const float** bdy, int imgpadding, float* outflow,
                 const float* initflow, int width, int height, int sc_f, int sc_l, int max_iter,
                 int min_iter, float dp, float dr, float res, int psz, float patove, int usefbcon,
                 int costfct, int noc, int patnorm, int usetvref, float tv_alpha, float tv_gamma,
                 float tv_delta, int tv_innerit, int tv_solverit, float tv_sor, int verbosity) {
  OFC::OFClass ofc(a, adx, ady, b, bdx, bdy, imgpadding, outflow, initflow, width, height, sc_f,
                   sc_l, max_iter, min_iter, dp, dr, res, psz, patove, usefbcon != 0, costfct, noc,
                   patnorm, usetvref != 0, tv_alpha, tv_gamma, tv_delta, tv_innerit, tv_solverit,
                   tv_sor, verbosity);
}

// One level of steps 1-4 of Algorithm 1 (grid init, init from coarser flow,
// inverse search, densification), forward grid and optionally the backward
// grid with forward-backward merging, mirroring oflow.cpp:191-267.
// Images are the padded level arrays.  flow_prev_fw/bw may be null (zero init).
// Returns the number of patches.
int ref_patchgrid_level(const float* a, const float* adx, const float* ady, const float* b,
                        const float* bdx, const float* bdy, int w_l, int h_l, int level,
                        int max_iter, int min_iter, float dp, float dr, float res, int psz,
                        float patove, int usefbcon, int costfct, int noc, int patnorm,
                        const float* flow_prev_fw, const float* flow_prev_bw, float* p_iter_fw,
                        float* pt_iter_fw, float* pweight_fw, int* cnt_fw, float* dense_fw,
                        float* p_iter_bw, float* pt_iter_bw, float* pweight_bw, float* dense_bw) {
  OFC::optparam op;
  fill_op(op, level, level, max_iter, min_iter, dp, dr, res, psz, patove, usefbcon, costfct, noc,
          patnorm, 0, 10, 10, 5, 1, 3, 1.6f);
  OFC::camparam cpl, cpr;
  fill_cp(cpl, w_l, h_l, psz, psz, level, 0);
  cpr = cpl;
  cpr.camlr = 1;
  OFC::PatGridClass gfw(&cpl, &cpr, &op);
  std::unique_ptr<OFC::PatGridClass> gbw;
  if (usefbcon) {
    gbw.reset(new OFC::PatGridClass(&cpr, &cpl, &op));
    gfw.SetComplGrid(gbw.get());
    gbw->SetComplGrid(&gfw);
  }
  gfw.InitializeGrid(a, adx, ady);
  gfw.SetTargetImage(b, bdx, bdy);
  if (usefbcon) {
    gbw->InitializeGrid(b, bdx, bdy);
    gbw->SetTargetImage(a, adx, ady);
  }
  if (flow_prev_fw) gfw.InitializeFromCoarserOF(flow_prev_fw);
  if (usefbcon && flow_prev_bw) gbw->InitializeFromCoarserOF(flow_prev_bw);
  gfw.Optimize();
  if (usefbcon) gbw->Optimize();
  dump_grid(gfw, op, p_iter_fw, pt_iter_fw, pweight_fw, cnt_fw);
  if (usefbcon) dump_grid(*gbw, op, p_iter_bw, pt_iter_bw, pweight_bw, nullptr);
  if (dense_fw) gfw.AggregateFlowDense(dense_fw);
  if (usefbcon && dense_bw) gbw->AggregateFlowDense(dense_bw);
  return gfw.GetNoPatches();
}

// Step 5 of Algorithm 1 on one level; flow is refined in place
// (refine_variational.cpp:25-116).  a / b are the padded level images (the
// gradient arguments are unused by the reference and passed as null).
void ref_varref_level(const float* a, const float* b, int w_l, int h_l, int pad, int level,
                      int camlr, int noc, float tv_alpha, float tv_gamma, float tv_delta,
                      int tv_innerit, int tv_solverit, float tv_sor, float* flow) {
  OFC::optparam op;
  fill_op(op, level, level, 1, 1, .05f, .95f, 0.f, pad, 0.f, 0, 0, noc, 1, 1, tv_alpha, tv_gamma,
          tv_delta, tv_innerit, tv_solverit, tv_sor);
  OFC::camparam cpt, cpo;
  fill_cp(cpt, w_l, h_l, pad, pad, level, camlr);
  cpo = cpt;
  cpo.camlr = 1 - camlr;
  OFC::VarRefClass v(a, nullptr, nullptr, b, nullptr, nullptr, &cpt, &cpo, &op, flow);
}

}  // extern "C"
