// TEST INFRASTRUCTURE ONLY — the parity oracle, never the product path.
//
// Plain C++ CPU restatement of the lordnn/OF_DIS per-frame-pair hot path
// (SURVEY.md 8a rows a1..a23).  Every function cites the reference file:line
// it restates.  Nothing here is linked into, imported by or called from the
// product (of_dis_b200/); only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load it.
//
// Pinning: validated in this container (a) bit-for-bit against the reference's
// own sources compiled unmodified into oracle/_ref (tests/test_oracle_cpu.py)
// and (b) for the OpenCV steps the reference delegates (resize, Sobel,
// copyMakeBorder, final resize; run_dense.cpp:131-179,299-312,407-415) against
// python-cv2 4.13 golden vectors in tests/golden/ (made by
// tests/golden/make_golden.py; the gradient-magnitude input of SELECTCHANNEL == 2,
// whose pyramid runs on arbitrary floats, by tests/golden/make_gradmag_golden.py).
// The reference itself ships no test vectors.
//
// Build: make -C oracle port   (g++ -O2 -ffp-contract=off, no FMA).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

extern "C" {
struct OrcParams {
  int sc_f, sc_l, max_iter, min_iter;
  float dp_thresh, dr_thresh, res_thresh;
  int psz;
  float patove;
  int usefbcon, patnorm, costfct, usetvref;
  float tv_alpha, tv_gamma, tv_delta;
  int tv_innerit, tv_solverit;
  float tv_sor;
  int verbosity;
  int mode;  // 1 optical flow (SELECTMODE 1), 2 depth from stereo (SELECTMODE 2)
  int noc;   // 1 grey (SELECTCHANNEL 1 or 2), 3 BGR (SELECTCHANNEL 3)
  int gradmag;  // 1: the grey frame is replaced by its gradient magnitude (SELECTCHANNEL 2, run_dense.cpp:139-148)
};
}

namespace {

typedef std::vector<float> fvec;

// Same accumulation order as oracle/eigen_shim (Eigen's SSE linear redux).
template <class F>
inline float packet_redux(size_t n, F at) {
  if (n == 0) return 0.f;
  const size_t ps = 4, asz2 = (n / 8) * 8, asz = (n / 4) * 4;
  float res;
  if (asz) {
    float p0[4], p1[4];
    for (int k = 0; k < 4; ++k) p0[k] = at(k);
    if (asz > ps) {
      for (int k = 0; k < 4; ++k) p1[k] = at(ps + k);
      for (size_t i = 2 * ps; i < asz2; i += 2 * ps) {
        for (int k = 0; k < 4; ++k) p0[k] += at(i + k);
        for (int k = 0; k < 4; ++k) p1[k] += at(i + ps + k);
      }
      for (int k = 0; k < 4; ++k) p0[k] += p1[k];
      if (asz > asz2)
        for (int k = 0; k < 4; ++k) p0[k] += at(asz2 + k);
    }
    res = (p0[0] + p0[2]) + (p0[1] + p0[3]);
    for (size_t i = asz; i < n; ++i) res += at(i);
  } else {
    res = at(0);
    for (size_t i = 1; i < n; ++i) res += at(i);
  }
  return res;
}

inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
inline int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
  return i;
}

// ---------------------------------------------------------------------------
// a1  dimension padding, run_dense.cpp:299-312
// ---------------------------------------------------------------------------
void pad_dims(int w0, int h0, int sc_f, int* W, int* H, int* padl, int* padt) {
  int scfct = 1 << sc_f;
  int padw = 0, padh = 0;
  if (w0 % scfct) padw = scfct - w0 % scfct;
  if (h0 % scfct) padh = scfct - h0 % scfct;
  *W = w0 + padw;
  *H = h0 + padh;
  *padl = padw / 2;  // floor(padw/2) left, ceil right (run_dense.cpp:309)
  *padt = padh / 2;
}

// u8 (h0,w0,C) -> f32 (H,W,C), BORDER_REPLICATE; convertTo CV_32F (run_dense.cpp:309,327)
void pad_to_float(const uint8_t* img, int w0, int h0, int C, int W, int H, int padl, int padt, fvec& out) {
  out.resize((size_t)W * H * C);
  for (int y = 0; y < H; ++y) {
    int sy = clampi(y - padt, 0, h0 - 1);
    for (int x = 0; x < W; ++x) {
      int sx = clampi(x - padl, 0, w0 - 1);
      for (int c = 0; c < C; ++c) out[((size_t)y * W + x) * C + c] = (float)img[((size_t)sy * w0 + sx) * C + c];
    }
  }
}

// a2  cv::resize(.5,.5,INTER_LINEAR) == 2x2 box mean (run_dense.cpp:151)
void halve(const fvec& src, int w, int h, int C, fvec& dst) {
  int w2 = w / 2, h2 = h / 2;
  dst.resize((size_t)w2 * h2 * C);
  for (int y = 0; y < h2; ++y)
    for (int x = 0; x < w2; ++x)
      for (int c = 0; c < C; ++c) {
        const float a = src[((size_t)(2 * y) * w + 2 * x) * C + c], b = src[((size_t)(2 * y) * w + 2 * x + 1) * C + c];
        const float d = src[((size_t)(2 * y + 1) * w + 2 * x) * C + c], e = src[((size_t)(2 * y + 1) * w + 2 * x + 1) * C + c];
        // OpenCV's own code (ResizeAreaFast, which INTER_LINEAR at exactly .5 dispatches to): four output pixels
        // per 128-bit step as (a+b)+(d+e), the w2 % 4 tail columns of a row summed one by one.  Only visible
        // on data that is not dyadic (the gradient-magnitude pyramid); pinned by tests/golden/gradmag.npz.
        const bool tail = C == 1 && x >= (w2 & ~3);
        dst[((size_t)y * w2 + x) * C + c] = (tail ? ((a + b) + d) + e : (a + b) + (d + e)) * 0.25f;
      }
}

// a2  cv::Sobel(CV_32F, k=3, scale 1/8, BORDER_DEFAULT=REFLECT_101) (run_dense.cpp:157-158)
void sobel(const fvec& src, int w, int h, int C, fvec& dx, fvec& dy) {
  dx.resize(src.size());
  dy.resize(src.size());
  auto S = [&](int x, int y, int c) { return src[((size_t)reflect101(y, h) * w + reflect101(x, w)) * C + c]; };
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x)
      for (int c = 0; c < C; ++c) {
        // dx: row [-1 0 1], column [1 2 1]/8
        float r0 = S(x + 1, y - 1, c) - S(x - 1, y - 1, c);
        float r1 = S(x + 1, y, c) - S(x - 1, y, c);
        float r2 = S(x + 1, y + 1, c) - S(x - 1, y + 1, c);
        dx[((size_t)y * w + x) * C + c] = 0.25f * r1 + 0.125f * (r2 + r0);
        // dy: row [1 2 1]/8, column [-1 0 1]
        float s0 = 0.25f * S(x, y - 1, c) + 0.125f * (S(x + 1, y - 1, c) + S(x - 1, y - 1, c));
        float s2 = 0.25f * S(x, y + 1, c) + 0.125f * (S(x + 1, y + 1, c) + S(x - 1, y + 1, c));
        dy[((size_t)y * w + x) * C + c] = s2 - s0;
      }
}

// a2, SELECTCHANNEL == 2 (run_dense.cpp:139-148): level 0 = sqrt(dx*dx + dy*dy) of the Sobel/8
// gradients of the padded float frame.  dx, dy are multiples of 1/8 below 2^7, so the squares and
// their sum are exact and the only rounding is cv::sqrt's (correctly rounded, like sqrtf).
void gradient_magnitude(const fvec& src, int w, int h, fvec& dst) {
  fvec gx, gy;
  sobel(src, w, h, 1, gx, gy);
  dst.resize(src.size());
  for (size_t i = 0; i < src.size(); ++i) dst[i] = std::sqrt(gx[i] * gx[i] + gy[i] * gy[i]);
}

// a2  copyMakeBorder by `pad` on all sides (run_dense.cpp:165-178)
void pad_level(const fvec& src, int w, int h, int C, int pad, bool replicate, float* out) {
  int tw = w + 2 * pad, th = h + 2 * pad;
  for (int y = 0; y < th; ++y)
    for (int x = 0; x < tw; ++x) {
      int sx = x - pad, sy = y - pad;
      bool inside = sx >= 0 && sy >= 0 && sx < w && sy < h;
      for (int c = 0; c < C; ++c) {
        float v = 0.f;
        if (inside || replicate) v = src[((size_t)clampi(sy, 0, h - 1) * w + clampi(sx, 0, w - 1)) * C + c];
        out[((size_t)y * tw + x) * C + c] = v;
      }
    }
}

// a23  flow *= 2^sc_l ; cv::resize(INTER_LINEAR, x 2^sc_l) ; crop (run_dense.cpp:407-415)
void upsample_crop(const float* flow, int w_l, int h_l, int nop, int sc_l, int padl, int padt, int w0, int h0, float* out) {
  const int k = 1 << sc_l;
  if (sc_l == 0) {
    for (int y = 0; y < h0; ++y)
      for (int x = 0; x < w0; ++x)
        for (int c = 0; c < nop; ++c) out[((size_t)y * w0 + x) * nop + c] = flow[((size_t)(y + padt) * w_l + x + padl) * nop + c];
    return;
  }
  const double inv = 1.0 / k;
  const float sc = (float)k;
  // per-column source coordinates do not depend on the row: computed once (same expressions, same bits)
  std::vector<int> sx0(w0), sx1v(w0);
  std::vector<float> fxv(w0);
  for (int x = 0; x < w0; ++x) {
    float fx = (float)((x + padl + 0.5) * inv - 0.5);
    int sx = (int)std::floor(fx);
    fx -= sx;
    if (sx < 0) { sx = 0; fx = 0.f; }
    if (sx >= w_l - 1) { sx = w_l - 1; fx = 0.f; }
    sx0[x] = sx; sx1v[x] = std::min(sx + 1, w_l - 1); fxv[x] = fx;
  }
  for (int y = 0; y < h0; ++y) {
    float fy = (float)((y + padt + 0.5) * inv - 0.5);
    int sy = (int)std::floor(fy);
    fy -= sy;
    if (sy < 0) { sy = 0; fy = 0.f; }
    if (sy >= h_l - 1) { sy = h_l - 1; fy = 0.f; }
    int sy1 = std::min(sy + 1, h_l - 1);
    const float* row0 = flow + (size_t)sy * w_l * nop;
    const float* row1 = flow + (size_t)sy1 * w_l * nop;
    float* o = out + (size_t)y * w0 * nop;
    for (int x = 0; x < w0; ++x) {
      const float fx = fxv[x];
      const int a = sx0[x] * nop, b = sx1v[x] * nop;
      for (int c = 0; c < nop; ++c) {
        // horizontal pass on rows sy, sy1, then vertical pass
        float r0 = (row0[a + c] * sc) * (1.f - fx) + (row0[b + c] * sc) * fx;
        float r1 = (row1[a + c] * sc) * (1.f - fx) + (row1[b + c] * sc) * fx;
        o[x * nop + c] = r0 * (1.f - fy) + r1 * fy;
      }
    }
  }
}

}  // namespace

// ===========================================================================
// Patch grid, inverse search, densification (rows a3..a12)
// ===========================================================================
namespace {

// a3  camparam / optparam derived constants, oflow.cpp:75-107,139-157
struct Level {
  int w, h, pad, tmp_w, level, camlr;
  float lb, ubw, ubh;
};
struct Opt {
  int psz, noc, nop, novals, steps, max_iter, min_iter, patnorm, costfct, mode;
  float dp_thresh_sq, dr_thresh, res_thresh, outlierthresh;
  float minerrval, normoutlier;
};

Opt make_opt(const OrcParams& p) {
  Opt o;
  o.mode = p.mode;
  o.nop = p.mode == 1 ? 2 : 1;                                      // oflow.cpp:75-79
  o.psz = p.psz;
  o.outlierthresh = (float)p.psz / 2;                               // oflow.cpp:81
  o.max_iter = p.max_iter;
  o.min_iter = p.min_iter;
  o.dp_thresh_sq = p.dp_thresh * p.dp_thresh;                       // oflow.cpp:87
  o.dr_thresh = p.dr_thresh;
  o.res_thresh = p.res_thresh;
  o.steps = std::max(1, (int)std::floor(p.psz * (1 - p.patove)));   // oflow.cpp:90 (float arithmetic)
  o.noc = p.noc;
  o.novals = p.noc * p.psz * p.psz;                                 // oflow.cpp:91
  o.patnorm = p.patnorm;
  o.costfct = p.costfct;
  o.minerrval = 2.0f;                                               // oflow.h:76
  o.normoutlier = 5.0f;                                             // oflow.h:77
  return o;
}

Level make_level(int W, int H, int level, int pad, int psz, int camlr) {
  Level L;
  float sc = std::pow(2.0f, -level);                                // oflow.cpp:142
  L.h = (int)(H * sc);
  L.w = (int)(W * sc);
  L.pad = pad;
  L.lb = -(float)psz / 2;                                           // oflow.cpp:147
  L.ubw = (float)(L.w + psz / 2 - 2);                               // oflow.cpp:148
  L.ubh = (float)(L.h + psz / 2 - 2);
  L.tmp_w = L.w + 2 * pad;
  L.level = level;
  L.camlr = camlr;
  return L;
}

// a4  PatGridClass::PatGridClass, patchgrid.cpp:31-76
struct Grid {
  int nopw, noph, n;
  fvec ptx, pty;        // pt_ref
  fvec p_init, p_iter;  // n*nop
  fvec pt_iter;         // n*2
  fvec pweight;         // n*novals  (zero-initialised: see eigen_shim note)
  std::vector<int> cnt;
  void build(const Level& L, const Opt& o) {
    nopw = (int)std::ceil((float)L.w / (float)o.steps);
    noph = (int)std::ceil((float)L.h / (float)o.steps);
    const int offw = (L.w - (nopw - 1) * o.steps) / 2;
    const int offh = (L.h - (noph - 1) * o.steps) / 2;
    n = nopw * noph;
    ptx.assign(n, 0.f); pty.assign(n, 0.f);
    p_init.assign((size_t)n * o.nop, 0.f);
    p_iter.assign((size_t)n * o.nop, 0.f);
    pt_iter.assign((size_t)n * 2, 0.f);
    pweight.assign((size_t)n * o.novals, 0.f);
    cnt.assign(n, 0);
    for (int x = 0; x < nopw; ++x)
      for (int y = 0; y < noph; ++y) {
        int i = x * noph + y;                                       // column-major patch index
        ptx[i] = (float)(x * o.steps + offw);
        pty[i] = (float)(y * o.steps + offh);
      }
  }
};

// a6  PatGridClass::InitializeFromCoarserOF, patchgrid.cpp:195-211
void init_from_coarser(Grid& g, const Level& L, const Opt& o, const float* flow_prev) {
  for (int ip = 0; ip < g.n; ++ip) {
    int x = (int)std::floor(g.ptx[ip] / 2);
    int y = (int)std::floor(g.pty[ip] / 2);
    int i = y * (L.w / 2) + x;
    if (o.nop == 2) {
      g.p_init[2 * ip] = flow_prev[2 * i] * 2;
      g.p_init[2 * ip + 1] = flow_prev[2 * i + 1] * 2;
    } else {
      g.p_init[ip] = flow_prev[i] * 2;
    }
  }
}

// a9  getPatchStaticBil, patch.cpp:346-413
void patch_bilinear(const float* img, const Level& L, const Opt& o, float mx, float my, float* out) {
  int pos0 = (int)std::ceil(mx + .00001f);   // float add on purpose (patch.cpp:356)
  int pos1 = (int)std::ceil(my + .00001f);
  int pos2 = (int)std::floor(mx);
  int pos3 = (int)std::floor(my);
  float r0 = mx - (float)pos2, r1 = my - (float)pos3;
  float we0 = r0 * r1, we1 = (1 - r0) * r1, we2 = r0 * (1 - r1), we3 = (1 - r0) * (1 - r1);
  pos0 += L.pad;
  pos1 += L.pad;
  const int C = o.noc, lb = -o.psz / 2, ub = o.psz / 2 - 1;
  float* t = out;
  for (int y = pos1 + lb; y <= pos1 + ub; ++y) {
    const float* a = img + ((size_t)y * L.tmp_w + (pos0 + lb)) * C;       // (x1, y1)
    const float* c = img + ((size_t)(y - 1) * L.tmp_w + (pos0 + lb)) * C; // (x1, y0)
    const float* b = a - C;                                                // (x0, y1)
    const float* d = c - C;                                                // (x0, y0)
    // columns lb..ub like the rows (patch.cpp:388): an odd patch size samples (psz - 1)^2 positions and the last
    // novals - (psz - 1)^2 C entries of the vector keep what they held (patch.cpp:411 still shifts all novals of them)
    for (int k = 0; k < (ub - lb + 1) * C; ++k) *t++ = we0 * a[k] + we1 * b[k] + we2 * c[k] + we3 * d[k];
  }
  if (o.patnorm > 0 || o.costfct == 10) {                                  // patch.cpp:411-412 (NCC: always)
    float m = packet_redux(o.novals, [out](size_t i) { return out[i]; }) / o.novals;
    for (int i = 0; i < o.novals; ++i) out[i] -= m;
  }
}

// a10  LossComputeErrorImage, patch.cpp:221-273 (novals/4 SIMD blocks; tail untouched)
void loss_error_image(const Opt& o, float* pdiff, float* pweight, const float* tmpl) {
  const int nblk = (o.novals / 4) * 4;
  const float bsq = o.normoutlier * o.normoutlier, bsq2 = bsq * 2.0f;
  if (o.costfct == 10) {
    // NCC.  NO REFERENCE SEMANTICS: README.md:79 advertises costfct 10 but LossComputeErrorImage has no branch
    // for it (patch.cpp:228,242,258) and leaves pweight uninitialised.  This is the definition the product
    // implements (DESIGN.md "NCC"): zero-mean patches, query rescaled to the template's norm,
    //   r = q^ * (|t^| / |q^|) - t^ ,  pdiff = r,  pweight = |r|   (r = -t^ when the query is flat).
    const float nt = std::sqrt(packet_redux(o.novals, [tmpl](size_t i) { return tmpl[i] * tmpl[i]; }));
    const float nq = std::sqrt(packet_redux(o.novals, [pdiff](size_t i) { return pdiff[i] * pdiff[i]; }));
    const float scale = nq > 0.f ? nt / nq : 0.f;
    for (int i = 0; i < nblk; ++i) {
      const float d = pdiff[i] * scale - tmpl[i];
      pdiff[i] = d;
      pweight[i] = std::fabs(d);
    }
    return;
  }
  for (int i = 0; i < nblk; ++i) {
    float d = pdiff[i] - tmpl[i];
    if (o.costfct == 0) {
      pdiff[i] = d;
      pweight[i] = std::fabs(d);
    } else if (o.costfct == 1) {
      float w = std::sqrt(std::fabs(d));
      pweight[i] = w;
      pdiff[i] = std::signbit(d) ? -w : w;   // or(and(-0,d), w): sign bit of d, magnitude w
    } else if (o.costfct == 2) {
      float w = std::sqrt((std::sqrt(1.0f + (d * d) / bsq) - 1.0f) * bsq2);
      pweight[i] = w;
      pdiff[i] = std::signbit(d) ? -w : w;
    }
    // any other costfct: the reference has no branch; nothing is written.
  }
}

struct PatchWork {
  fvec tmpl, dx, dy, pdiff;
};

// a5,a7,a8,a11  InitializePatch + OptimizeIter(untilconv=true), patch.cpp:55-210,275-295
void optimize_patch(Grid& g, int ip, const Level& L, const Opt& o, const float* a, const float* adx,
                    const float* ady, const float* b, PatchWork& wk) {
  const int C = o.noc, nv = o.novals, lb = -o.psz / 2, ub = o.psz / 2 - 1;
  // every PatClass owns its vectors (patch.cpp:31-41); entries no loop writes (odd patch sizes) start as zero, as in
  // oracle/_ref where the Eigen shim zero-initialises on resize (heap garbage upstream)
  wk.tmpl.assign(nv, 0.f); wk.dx.assign(nv, 0.f); wk.dy.assign(nv, 0.f); wk.pdiff.assign(nv, 0.f);
  float *tmpl = wk.tmpl.data(), *tdx = wk.dx.data(), *tdy = wk.dy.data(), *pdiff = wk.pdiff.data();
  float* pweight = &g.pweight[(size_t)ip * nv];
  const float ptrx = g.ptx[ip], ptry = g.pty[ip];
  // getPatchStaticNNGrad, patch.cpp:298-343
  {
    int px = (int)std::round(ptrx) + L.pad, py = (int)std::round(ptry) + L.pad, n = 0;
    for (int j = lb; j <= ub; ++j)
      for (int i = lb; i <= ub; ++i) {
        size_t idx = ((size_t)(py + j) * L.tmp_w + (px + i)) * C;
        for (int c = 0; c < C; ++c, ++n) { tmpl[n] = a[idx + c]; tdx[n] = adx[idx + c]; tdy[n] = ady[idx + c]; }
      }
    if (o.patnorm > 0 || o.costfct == 10) {
      float m = packet_redux(nv, [tmpl](size_t i) { return tmpl[i]; }) / nv;
      for (int i = 0; i < nv; ++i) tmpl[i] -= m;
    }
  }
  // ComputeHessian, patch.cpp:69-86
  float H00 = packet_redux(nv, [tdx](size_t i) { return tdx[i] * tdx[i]; });
  float H01 = 0.f, H11 = 0.f;
  if (o.mode == 1) {
    H01 = packet_redux(nv, [tdx, tdy](size_t i) { return tdx[i] * tdy[i]; });
    H11 = packet_redux(nv, [tdy](size_t i) { return tdy[i] * tdy[i]; });
    if (H00 * H11 - H01 * H01 == 0) { H00 = (float)(H00 + 1e-10); H11 = (float)(H11 + 1e-10); }
  } else {
    if (H00 == 0) H00 = (float)(H00 + 1e-10);
  }
  // Hes.llt(): Eigen unblocked LLT, stop at the first non-positive pivot
  float l00 = H00, l10 = H01, l11 = H11;
  if (o.mode == 1) {
    if (l00 > 0.f) {
      l00 = std::sqrt(l00);
      l10 = l10 / l00;
      float x = l11 - l10 * l10;
      if (x > 0.f) l11 = std::sqrt(x);
    }
  } else {
    if (l00 > 0.f) l00 = std::sqrt(l00);
  }
  // ResetPatch + OptimizeStart, patch.cpp:97-154
  float p_in[2] = {0.f, 0.f}, p_iter[2], delta_p[2] = {0.f, 0.f};
  for (int k = 0; k < o.nop; ++k) p_in[k] = g.p_init[(size_t)ip * o.nop + k];
  p_iter[0] = p_in[0]; p_iter[1] = p_in[1];
  float ptx = ptrx + p_iter[0];
  float pty = (o.mode == 1) ? ptry + p_iter[1] : ptry;              // paramtopt, patch.cpp:212-219
  const float pt_stx = ptx, pt_sty = pty;
  bool conv = false;
  int cnt = 0;
  float dpsq = 1e-10f, dpsq_init = 1e-10f, mares = 1e20f, mares_old = 1e20f;

  auto compute_err = [&]() {                                        // OptimizeComputeErrImg, patch.cpp:275-295
    patch_bilinear(b, L, o, ptx, pty, pdiff);
    loss_error_image(o, pdiff, pweight, tmpl);
    dpsq = (o.mode == 1) ? delta_p[0] * delta_p[0] + delta_p[1] * delta_p[1] : delta_p[0] * delta_p[0];
    if (cnt == 1) dpsq_init = dpsq;
    mares_old = mares;
    mares = packet_redux(nv, [pweight](size_t i) { return std::fabs(pweight[i]); }) / nv;
    if (!((cnt < o.max_iter) & (mares > o.res_thresh) &
          ((cnt < o.min_iter) | (dpsq / dpsq_init >= o.dp_thresh_sq)) &
          ((cnt < o.min_iter) | (mares / mares_old <= o.dr_thresh))))
      conv = true;
  };

  if (ptx < L.lb || pty < L.lb || ptx > L.ubw || pty > L.ubh) {
    conv = true;  // pdiff = tmp; pweight untouched (zero)
  } else {
    mares = 1e5f;
    compute_err();
  }
  while (!conv) {                                                   // OptimizeIter, patch.cpp:170-209
    cnt++;
    delta_p[0] = packet_redux(nv, [tdx, pdiff](size_t i) { return tdx[i] * pdiff[i]; });
    if (o.mode == 1) {
      delta_p[1] = packet_redux(nv, [tdy, pdiff](size_t i) { return tdy[i] * pdiff[i]; });
      // L y = b ; L^T x = y
      float y0 = delta_p[0] / l00;
      float y1 = (delta_p[1] - l10 * y0) / l11;
      float x1 = y1 / l11;
      float x0 = (y0 - l10 * x1) / l00;
      delta_p[0] = x0; delta_p[1] = x1;
      p_iter[0] -= delta_p[0];
      p_iter[1] -= delta_p[1];
    } else {
      delta_p[0] = (delta_p[0] / l00) / l00;
      p_iter[0] -= delta_p[0];
      if (L.camlr == 0) p_iter[0] = std::min(p_iter[0], 0.0f);      // patch.cpp:186-191
      else p_iter[0] = std::max(p_iter[0], 0.0f);
    }
    ptx = ptrx + p_iter[0];
    if (o.mode == 1) pty = ptry + p_iter[1];
    float ddx = pt_stx - ptx, ddy = pt_sty - pty;
    if (std::sqrt(ddx * ddx + ddy * ddy) > o.outlierthresh || ptx < L.lb || pty < L.lb || ptx > L.ubw || pty > L.ubh) {
      p_iter[0] = p_in[0]; p_iter[1] = p_in[1];                     // reset, patch.cpp:202-205
      ptx = ptrx + p_iter[0];
      if (o.mode == 1) pty = ptry + p_iter[1];
      conv = true;
    }
    compute_err();
  }
  for (int k = 0; k < o.nop; ++k) g.p_iter[(size_t)ip * o.nop + k] = p_iter[k];
  g.pt_iter[2 * ip] = ptx;
  g.pt_iter[2 * ip + 1] = pty;
  g.cnt[ip] = cnt;
}

void optimize_grid(Grid& g, const Level& L, const Opt& o, const float* a, const float* adx, const float* ady, const float* b) {
  PatchWork wk;
  for (int ip = 0; ip < g.n; ++ip) optimize_patch(g, ip, L, o, a, adx, ady, b, wk);
}

// a12  PatGridClass::AggregateFlowDense, patchgrid.cpp:213-397 (cg = complementary grid or null)
void aggregate_dense(const Grid& g, const Grid* cg, const Level& L, const Opt& o, float* flowout) {
  const int W = L.w, H = L.h, nop = o.nop, lb = -o.psz / 2, ub = o.psz / 2 - 1;
  fvec we((size_t)W * H, 0.f);
  std::memset(flowout, 0, sizeof(float) * (size_t)nop * W * H);
  for (int ip = 0; ip < g.n; ++ip) {
    const float* fl = &g.p_iter[(size_t)ip * nop];
    const float* pw = &g.pweight[(size_t)ip * o.novals];
    for (int y = lb; y <= ub; ++y)
      for (int x = lb; x <= ub; ++x, ++pw) {
        int yt = (int)(y + g.pty[ip]), xt = (int)(x + g.ptx[ip]);
        if (xt >= 0 && yt >= 0 && xt < W && yt < H) {
          int i = yt * W + xt;
          float absw;
          if (o.noc == 1) {
            absw = 1.0f / std::max(o.minerrval, *pw);
          } else {  // RGB: the two extra ++pweight are inside the in-image test (patchgrid.cpp:255-258)
            absw = std::max(o.minerrval, *pw); ++pw;
            absw += std::max(o.minerrval, *pw); ++pw;
            absw += std::max(o.minerrval, *pw);
            absw = 1.0f / absw;
          }
          we[i] += absw;
          for (int k = 0; k < nop; ++k) flowout[(size_t)nop * i + k] += fl[k] * absw;
        }
      }
  }
  if (cg) {                                                         // patchgrid.cpp:278-375
    for (int ip = 0; ip < cg->n; ++ip) {
      const float* fl = &cg->p_iter[(size_t)ip * nop];
      const float* pw = &cg->pweight[(size_t)ip * o.novals];
      const float rx = cg->pt_iter[2 * ip], ry = cg->pt_iter[2 * ip + 1];
      int pos0 = (int)std::ceil(rx + .00001);   // double add here (patchgrid.cpp:304)
      int pos1 = (int)std::ceil(ry + .00001);
      int pos2 = (int)std::floor(rx), pos3 = (int)std::floor(ry);
      float r0 = rx - pos2, r1 = ry - pos3;
      float wb0 = r0 * r1, wb1 = (1 - r0) * r1, wb2 = r0 * (1 - r1), wb3 = (1 - r0) * (1 - r1);
      for (int y = lb; y <= ub; ++y)
        for (int x = lb; x <= ub; ++x, ++pw) {
          int yt = y + pos1, xt = x + pos0;
          if (xt >= 1 && yt >= 1 && xt < W - 1 && yt < H - 1) {
            float absw;
            if (o.noc == 1) {
              absw = 1.0f / std::max(o.minerrval, *pw);
            } else {
              absw = std::max(o.minerrval, *pw); ++pw;
              absw += std::max(o.minerrval, *pw); ++pw;
              absw += std::max(o.minerrval, *pw);
              absw = 1.0f / absw;
            }
            int icc = xt + yt * W, ifc = (xt - 1) + yt * W, icf = xt + (yt - 1) * W, iff = (xt - 1) + (yt - 1) * W;
            we[icc] += wb0 * absw; we[ifc] += wb1 * absw; we[icf] += wb2 * absw; we[iff] += wb3 * absw;
            for (int k = 0; k < nop; ++k) {
              float fn = fl[k] * absw;
              flowout[(size_t)nop * icc + k] -= wb0 * fn;
              flowout[(size_t)nop * ifc + k] -= wb1 * fn;
              flowout[(size_t)nop * icf + k] -= wb2 * fn;
              flowout[(size_t)nop * iff + k] -= wb3 * fn;
            }
          }
        }
    }
  }
  for (int i = 0; i < W * H; ++i)                                    // patchgrid.cpp:377-394
    if (we[i] > 0)
      for (int k = 0; k < nop; ++k) flowout[(size_t)nop * i + k] /= we[i];
}

}  // namespace

// ===========================================================================
// Variational refinement (rows a13..a22).  Planes are stored with stride ==
// width: the reference pads rows to a multiple of 4 floats (image.cpp:31) but
// the padding lanes never feed a valid pixel.
// ===========================================================================
namespace {

struct Plane {
  int w, h;
  fvec d;
  Plane() : w(0), h(0) {}
  Plane(int w_, int h_) : w(w_), h(h_), d((size_t)w_ * h_, 0.f) {}
  float& at(int x, int y) { return d[(size_t)y * w + x]; }
  float at(int x, int y) const { return d[(size_t)y * w + x]; }
};

// a16  convolve_horiz_fast_5 / convolve_vert_fast_5 with deriv = [1,-8,0,8,-1]/12
// (image.cpp:467-526,594-640; coefficients from convolution_new, refine_variational.cpp:45-46:
//  coeffs = {1/12, -8/12, -0, 8/12, -1/12}); sum order c0*s0 + ((c1*s1 + c2*s2) + (c3*s3 + c4*s4)).
struct Deriv5 {
  float c[5];
  Deriv5() {
    const float h1 = -8.0f / 12.0f, h2 = 1.0f / 12.0f;
    c[0] = h2; c[1] = h1; c[2] = -0.0f; c[3] = -h1; c[4] = -h2;
  }
};
inline float tap5(const Deriv5& k, float s0, float s1, float s2, float s3, float s4) {
  return k.c[0] * s0 + ((k.c[1] * s1 + k.c[2] * s2) + (k.c[3] * s3 + k.c[4] * s4));
}
void conv5_h(const Plane& s, Plane& d) {
  static const Deriv5 k;
  for (int y = 0; y < s.h; ++y)
    for (int x = 0; x < s.w; ++x) {
      auto S = [&](int i) { return s.at(clampi(i, 0, s.w - 1), y); };
      d.at(x, y) = tap5(k, S(x - 2), S(x - 1), S(x), S(x + 1), S(x + 2));
    }
}
void conv5_v(const Plane& s, Plane& d) {
  static const Deriv5 k;
  for (int y = 0; y < s.h; ++y)
    for (int x = 0; x < s.w; ++x) {
      auto S = [&](int j) { return s.at(x, clampi(j, 0, s.h - 1)); };
      d.at(x, y) = tap5(k, S(y - 2), S(y - 1), S(y), S(y + 1), S(y + 2));
    }
}
// convolve_*_fast_3 with deriv_flow = {-0.5, -0, 0.5} (image.cpp:419-465,556-592; refine_variational.cpp:47-48)
inline float tap3(float s0, float s1, float s2) { return -0.5f * s0 + (-0.0f * s1 + 0.5f * s2); }

// a13  copyimage: strip padding, interleaved -> planar (refine_variational.cpp:120-149)
void strip_to_planes(const float* img, int w, int h, int pad, int C, std::vector<Plane>& out) {
  out.assign(C, Plane(w, h));
  const int tw = w + 2 * pad;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x)
      for (int c = 0; c < C; ++c) out[c].at(x, y) = img[((size_t)(y + pad) * tw + (x + pad)) * C + c];
}

// a14  image_warp, opticalflow_aux.c:33-75
void image_warp(const std::vector<Plane>& src, const Plane& wx, const Plane& wy, std::vector<Plane>& dst, Plane& mask) {
  const int w = wx.w, h = wx.h;
  for (int j = 0; j < h; ++j)
    for (int i = 0; i < w; ++i) {
      float xx = i + wx.at(i, j), yy = j + wy.at(i, j);
      int x = (int)std::floor(xx), y = (int)std::floor(yy);
      float dx = xx - x, dy = yy - y;
      mask.at(i, j) = (xx >= 0 && xx <= w - 1 && yy >= 0 && yy <= h - 1) ? 1.f : 0.f;
      int x1 = clampi(x, 0, w - 1), x2 = clampi(x + 1, 0, w - 1), y1 = clampi(y, 0, h - 1), y2 = clampi(y + 1, 0, h - 1);
      for (size_t c = 0; c < src.size(); ++c)
        dst[c].at(i, j) = src[c].at(x1, y1) * (1.0f - dx) * (1.0f - dy) + src[c].at(x2, y1) * dx * (1.0f - dy) +
                          src[c].at(x1, y2) * (1.0f - dx) * dy + src[c].at(x2, y2) * dx * dy;
    }
}

struct Derivs {
  std::vector<Plane> Ix, Iy, Iz, Ixx, Ixy, Iyy, Ixz, Iyz;
};
// a15  get_derivatives, opticalflow_aux.c:79-132
void get_derivatives(const std::vector<Plane>& im1, const std::vector<Plane>& im2w, Derivs& D) {
  const size_t C = im1.size();
  const int w = im1[0].w, h = im1[0].h;
  for (auto* v : {&D.Ix, &D.Iy, &D.Iz, &D.Ixx, &D.Ixy, &D.Iyy, &D.Ixz, &D.Iyz}) v->assign(C, Plane(w, h));
  for (size_t c = 0; c < C; ++c) {
    Plane avg(w, h);
    for (size_t i = 0; i < avg.d.size(); ++i) {
      avg.d[i] = 0.5f * (im2w[c].d[i] + im1[c].d[i]);
      D.Iz[c].d[i] = im2w[c].d[i] - im1[c].d[i];
    }
    conv5_h(avg, D.Ix[c]);
    conv5_v(avg, D.Iy[c]);
    conv5_h(D.Ix[c], D.Ixx[c]);
    conv5_v(D.Ix[c], D.Ixy[c]);
    conv5_v(D.Iy[c], D.Iyy[c]);
    conv5_h(D.Iz[c], D.Ixz[c]);
    conv5_v(D.Iz[c], D.Iyz[c]);
  }
}

// a17  compute_smoothness, opticalflow_aux.c:138-187
void compute_smoothness(const Plane& uu, const Plane& vv, float quarter_alpha, Plane& sh, Plane& sv) {
  const int w = uu.w, h = uu.h;
  const float eps = 0.001f * 0.001f;
  Plane s(w, h);
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      int xm = clampi(x - 1, 0, w - 1), xp = clampi(x + 1, 0, w - 1), ym = clampi(y - 1, 0, h - 1), yp = clampi(y + 1, 0, h - 1);
      float ux = tap3(uu.at(xm, y), uu.at(x, y), uu.at(xp, y));
      float vx = tap3(vv.at(xm, y), vv.at(x, y), vv.at(xp, y));
      float uy = tap3(uu.at(x, ym), uu.at(x, y), uu.at(x, yp));
      float vy = tap3(vv.at(x, ym), vv.at(x, y), vv.at(x, yp));
      float t = eps + ((ux * ux + uy * uy) + (vx * vx + vy * vy));
      s.at(x, y) = quarter_alpha / std::sqrt(t);
    }
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      sh.at(x, y) = (x < w - 1) ? s.at(x, y) + s.at(x + 1, y) : 0.f;
      sv.at(x, y) = (y < h - 1) ? s.at(x, y) + s.at(x, y + 1) : 0.f;
    }
}

// a19  sub_laplacian, opticalflow_aux.c:194-223 (horizontal pass over the whole image, then vertical)
void sub_laplacian(Plane& dst, const Plane& src, const Plane& wh, const Plane& wv) {
  const int w = src.w, h = src.h;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w - 1; ++x) {
      float t = wh.at(x, y) * (src.at(x + 1, y) - src.at(x, y));
      dst.at(x, y) += t;
      dst.at(x + 1, y) -= t;
    }
  for (int y = 0; y < h - 1; ++y)
    for (int x = 0; x < w; ++x) {
      float t = wv.at(x, y) * (src.at(x, y + 1) - src.at(x, y));
      dst.at(x, y) += t;
      dst.at(x, y + 1) -= t;
    }
}

const float DNORM = 0.1f * 0.1f, EPSC = 0.001f * 0.001f, EPSG = 0.001f * 0.001f;  // opticalflow_aux.c:25-29
inline float mla(float a, float b, float c) { return a + b * c; }  // _vmla_ps
inline float mls(float a, float b, float c) { return a - b * c; }  // _vmls_ps

// a18  compute_data, opticalflow_aux.c:409-594 (grey and RGB, including the RGB channel mix-ups)
void compute_data(const Derivs& D, const Plane& mask, const Plane& du, const Plane& dv, float hd3, float hg3,
                  Plane& a11, Plane& a12, Plane& a22, Plane& b1, Plane& b2) {
  const size_t n = mask.d.size();
  const bool rgb = D.Ix.size() == 3;
  for (size_t i = 0; i < n; ++i) {
    float A11 = 0, A12 = 0, A22 = 0, B1 = 0, B2 = 0;
    const float u = du.d[i], v = dv.d[i], m = mask.d[i];
    float tmp, tmp2, tmp3, tmpx, tmpy, tmpxy, n1, n2, n3 = 0, n4 = 0, n5 = 0, n6 = 0;
    auto P = [i](const std::vector<Plane>& p, int c) { return p[c].d[i]; };
    if (hd3) {
      tmpx = P(D.Ix, 0); tmpy = P(D.Iy, 0);
      tmp2 = mla(mla(P(D.Iz, 0), tmpx, u), tmpy, v);
      n1 = mla(mla(DNORM, tmpx, tmpx), tmpy, tmpy);
      tmp = (tmp2 * tmp2) / n1;
      if (rgb) {
        tmpx = P(D.Ix, 1); tmpy = P(D.Iy, 1);
        tmp2 = mla(mla(P(D.Iz, 1), tmpx, u), tmpy, v);
        n2 = mla(mla(DNORM, tmpx, tmpx), tmpy, tmpy);
        tmp = tmp + (tmp2 * tmp2) / n2;
        tmpx = P(D.Ix, 2); tmpy = P(D.Iy, 2);
        tmp2 = mla(mla(P(D.Iz, 2), tmpx, u), tmpy, v);
        n3 = mla(mla(DNORM, tmpx, tmpx), tmpy, tmpy);
        tmp = tmp + (tmp2 * tmp2) / n3;
        tmp = (m * hd3) / std::sqrt(EPSC + tmp);
        tmp3 = tmp / n3; tmp2 = tmp3 * tmpx; tmp3 = tmp3 * tmpy;
        A11 = mla(A11, tmp2, tmpx); A12 = mla(A12, tmp2, tmpy); A22 = mla(A22, tmp3, tmpy);
        B1 = mls(B1, tmp2, P(D.Iz, 2)); B2 = mls(B2, tmp3, P(D.Iz, 2));
        tmpx = P(D.Ix, 1); tmpy = P(D.Iy, 1);
        tmp3 = tmp / n2; tmp2 = tmp3 * tmpx; tmp3 = tmp3 * tmpy;
        A11 = mla(A11, tmp2, tmpx); A12 = mla(A12, tmp2, tmpy); A22 = mla(A22, tmp3, tmpy);
        B1 = mls(B1, tmp2, P(D.Iz, 1)); B2 = mls(B2, tmp3, P(D.Iz, 1));
        tmpx = P(D.Ix, 0); tmpy = P(D.Iy, 1);   // sic: Iy of channel 2 (opticalflow_aux.c:496)
      } else {
        tmp = (m * hd3) / std::sqrt(mla(EPSC, 3.f, tmp));
      }
      tmp3 = tmp / n1; tmp2 = tmp3 * tmpx; tmp3 = tmp3 * tmpy;
      A11 = mla(A11, tmp2, tmpx); A12 = mla(A12, tmp2, tmpy); A22 = mla(A22, tmp3, tmpy);
      B1 = mls(B1, tmp2, P(D.Iz, 0)); B2 = mls(B2, tmp3, P(D.Iz, 0));
    }
    // gradient constancy
    tmpx = P(D.Ixx, 0); tmpy = P(D.Iyy, 0); tmpxy = P(D.Ixy, 0);
    tmp2 = mla(mla(P(D.Ixz, 0), tmpx, u), tmpxy, v);
    tmp3 = mla(mla(P(D.Iyz, 0), tmpxy, u), tmpy, v);
    tmpxy = tmpxy * tmpxy;
    n1 = mla(tmpxy + DNORM, tmpx, tmpx);
    n2 = mla(tmpxy + DNORM, tmpy, tmpy);
    tmp = (tmp2 * tmp2) / n1 + (tmp3 * tmp3) / n2;
    if (rgb) {
      tmpx = P(D.Ixx, 1); tmpy = P(D.Iyy, 1); tmpxy = P(D.Ixy, 1);
      tmp2 = mla(mla(P(D.Ixz, 1), tmpx, u), tmpxy, v);
      tmp3 = mla(mla(P(D.Iyz, 1), tmpxy, u), tmpy, v);
      tmpxy = tmpxy * tmpxy;
      n3 = mla(tmpxy + DNORM, tmpx, tmpx);
      n4 = mla(tmpxy + DNORM, tmpy, tmpy);
      tmp = (tmp2 * tmp2) / n3 + (tmp3 * tmp3) / n4;   // sic: assigned, not accumulated (:529)
      tmpx = P(D.Ixx, 2); tmpy = P(D.Iyy, 2); tmpxy = P(D.Ixy, 2);
      tmp2 = mla(mla(P(D.Ixz, 2), tmpx, u), tmpxy, v);
      tmp3 = mla(mla(P(D.Iyz, 2), tmpxy, u), tmpy, v);
      tmpxy = tmpxy * tmpxy;
      n5 = mla(tmpxy + DNORM, tmpx, tmpx);
      n6 = mla(tmpxy + DNORM, tmpy, tmpy);
      tmp = (tmp2 * tmp2) / n5 + (tmp3 * tmp3) / n6;   // sic (:538)
      tmp = (m * hg3) / std::sqrt(EPSG + tmp);
      tmp2 = tmp / n6; tmp3 = tmp / n5;
      A11 = mla(mla(A11, tmp3, tmpx * tmpx), tmp2, tmpxy);
      A22 = mla(mla(A22, tmp2, tmpy * tmpy), tmp3, tmpxy);
      tmpxy = P(D.Ixy, 2);
      A12 = mla(A12, mla(tmp3 * tmpx, tmp2, tmpy), tmpxy);
      B1 = mls(mls(B1, tmp3 * tmpx, P(D.Ixz, 2)), tmp2 * tmpxy, P(D.Iyz, 2));
      B2 = mls(mls(B2, tmp2 * tmpy, P(D.Iyz, 2)), tmp3 * tmpxy, P(D.Ixz, 2));
      tmp2 = tmp / n4; tmp3 = tmp / n3;
      tmpx = P(D.Ixx, 1); tmpy = P(D.Iyy, 1); tmpxy = P(D.Ixy, 1);
      A12 = mla(A12, mla(tmp3 * tmpx, tmp2, tmpy), tmpxy);
      B1 = mls(mls(B1, tmp3 * tmpx, P(D.Ixz, 1)), tmp2 * tmpxy, P(D.Iyz, 1));
      B2 = mls(mls(B2, tmp2 * tmpy, P(D.Iyz, 1)), tmp3 * tmpxy, P(D.Ixz, 1));
      tmpxy = tmpxy * tmpxy;
      A11 = mla(mla(A11, tmp3, tmpx * tmpx), tmp2, tmpxy);
      A22 = mla(mla(A22, tmp2, tmpy * tmpy), tmp3, tmpxy);
      tmpx = P(D.Ixx, 0); tmpy = P(D.Iyy, 0);
    } else {
      tmp = (m * hg3) / std::sqrt(mla(EPSG, 3.f, tmp));
    }
    tmp2 = tmp / n2; tmp3 = tmp / n1;
    tmpxy = P(D.Ixy, 0);
    A12 = mla(A12, mla(tmp3 * tmpx, tmp2, tmpy), tmpxy);
    B1 = mls(mls(B1, tmp3 * tmpx, P(D.Ixz, 0)), tmp2 * tmpxy, P(D.Iyz, 0));
    B2 = mls(mls(B2, tmp2 * tmpy, P(D.Iyz, 0)), tmp3 * tmpxy, P(D.Ixz, 0));
    tmpxy = tmpxy * tmpxy;
    A11 = mla(mla(A11, tmp3, tmpx * tmpx), tmp2, tmpxy);
    A22 = mla(mla(A22, tmp2, tmpy * tmpy), tmp3, tmpxy);
    if (!rgb) { A11 *= 3.f; A12 *= 3.f; A22 *= 3.f; B1 *= 3.f; B2 *= 3.f; }
    a11.d[i] = A11; a12.d[i] = A12; a22.d[i] = A22; b1.d[i] = B1; b2.d[i] = B2;
  }
}

// a18  compute_data_DE, opticalflow_aux.c:602-747
void compute_data_DE(const Derivs& D, const Plane& mask, const Plane& du, float hd3, float hg3, Plane& a11, Plane& b1) {
  const size_t n = mask.d.size();
  const bool rgb = D.Ix.size() == 3;
  for (size_t i = 0; i < n; ++i) {
    float A11 = 0, B1 = 0;
    const float u = du.d[i], m = mask.d[i];
    float tmp, tmp2, tmp3, tmpx, tmpy, tmpxy, n1, n2, n3 = 0, n4 = 0, n5 = 0, n6 = 0;
    auto P = [i](const std::vector<Plane>& p, int c) { return p[c].d[i]; };
    if (hd3) {
      tmpx = P(D.Ix, 0); tmpy = P(D.Iy, 0);
      tmp2 = mla(P(D.Iz, 0), tmpx, u);
      n1 = mla(mla(DNORM, tmpy, tmpy), tmpx, tmpx);
      tmp = (tmp2 * tmp2) / n1;
      if (rgb) {
        tmpx = P(D.Ix, 1); tmpy = P(D.Iy, 1);
        tmp2 = mla(P(D.Iz, 1), tmpx, u);
        n2 = mla(mla(DNORM, tmpy, tmpy), tmpx, tmpx);
        tmp = tmp + (tmp2 * tmp2) / n2;
        tmpx = P(D.Ix, 2); tmpy = P(D.Iy, 2);
        tmp2 = mla(P(D.Iz, 2), tmpx, u);
        n3 = mla(mla(DNORM, tmpy, tmpy), tmpx, tmpx);
        tmp = tmp + (tmp2 * tmp2) / n3;
        tmp = (m * hd3) / std::sqrt(EPSC + tmp);
        tmp2 = (tmp / n3) * tmpx;
        A11 = mla(A11, tmp2, tmpx); B1 = mls(B1, tmp2, P(D.Iz, 2));
        tmpx = P(D.Ix, 1);
        tmp2 = (tmp / n2) * tmpx;
        A11 = mla(A11, tmp2, tmpx); B1 = mls(B1, tmp2, P(D.Iz, 1));
        tmpx = P(D.Ix, 0);
      } else {
        tmp = (m * hd3) / std::sqrt(mla(EPSC, 3.f, tmp));
      }
      tmp2 = (tmp / n1) * tmpx;
      A11 = mla(A11, tmp2, tmpx); B1 = mls(B1, tmp2, P(D.Iz, 0));
    }
    tmpx = P(D.Ixx, 0); tmpy = P(D.Iyy, 0); tmpxy = P(D.Ixy, 0);
    tmp2 = mla(P(D.Iyz, 0), tmpxy, u);
    tmpxy = mla(DNORM, tmpxy, tmpxy);
    n1 = mla(tmpxy, tmpx, tmpx);
    n2 = mla(tmpxy, tmpy, tmpy);
    tmp = (tmp2 * tmp2) / n2;
    tmp2 = mla(P(D.Ixz, 0), tmpx, u);
    tmp = tmp + (tmp2 * tmp2) / n1;
    if (rgb) {
      tmpx = P(D.Ixx, 1); tmpy = P(D.Iyy, 1); tmpxy = P(D.Ixy, 1);
      tmp2 = mla(P(D.Iyz, 1), tmpxy, u);
      tmpxy = mla(DNORM, tmpxy, tmpxy);
      n3 = mla(tmpxy, tmpx, tmpx);
      n4 = mla(tmpxy, tmpy, tmpy);
      tmp = tmp + (tmp2 * tmp2) / n4;
      tmp2 = mla(P(D.Ixz, 1), tmpx, u);
      tmp = tmp + (tmp2 * tmp2) / n3;
      tmpx = P(D.Ixx, 2); tmpy = P(D.Iyy, 2); tmpxy = P(D.Ixy, 2);
      tmp2 = mla(P(D.Iyz, 2), tmpxy, u);
      tmpxy = mla(DNORM, tmpxy, tmpxy);
      n5 = mla(tmpxy, tmpx, tmpx);
      n6 = mla(tmpxy, tmpy, tmpy);
      tmp = tmp + (tmp2 * tmp2) / n6;
      tmp2 = mla(P(D.Ixz, 2), tmpx, u);
      tmp = tmp + (tmp2 * tmp2) / n5;
      tmp = (m * hg3) / std::sqrt(EPSG + tmp);
      tmpxy = P(D.Ixy, 2);
      tmp2 = (tmp / n6) * tmpxy; tmp3 = (tmp / n5) * tmpx;
      A11 = mla(mla(A11, tmp3, tmpx), tmp2, tmpxy);
      B1 = mls(mls(B1, tmp3, P(D.Ixz, 2)), tmp2, P(D.Iyz, 2));
      tmpx = P(D.Ixx, 1); tmpy = P(D.Iyy, 1); tmpxy = P(D.Ixy, 1);
      tmp2 = (tmp / n4) * tmpxy; tmp3 = (tmp / n3) * tmpx;
      A11 = mla(mla(A11, tmp3, tmpx), tmp2, tmpxy);
      B1 = mls(mls(B1, tmp3, P(D.Ixz, 1)), tmp2, P(D.Iyz, 1));
      tmpx = P(D.Ixx, 0); tmpy = P(D.Iyy, 0);
    } else {
      tmp = (m * hg3) / std::sqrt(mla(EPSG, 3.f, tmp));
    }
    tmpxy = P(D.Ixy, 0);
    tmp2 = (tmp / n2) * tmpxy; tmp3 = (tmp / n1) * tmpx;
    A11 = mla(mla(A11, tmp3, tmpx), tmp2, tmpxy);
    B1 = mls(mls(B1, tmp3, P(D.Ixz, 0)), tmp2, P(D.Iyz, 0));
    if (!rgb) { A11 *= 3.f; B1 *= 3.f; }
    (void)tmpy;
    a11.d[i] = A11; b1.d[i] = B1;
  }
}

// Pixel order of one SOR sweep.  0: raster order (the reference, solver.c).  1: red-black — NOT in the
// reference; the restatement of the product's DISB200_SOR_REDBLACK mode (all pixels with even i + j, then all
// with odd i + j, same update formulas), so that mode too has a CPU oracle.
int g_sor_order = 0;
bool g_pyr_gradmag = false;
template <class F>
inline void sweep_pixels(int w, int h, F f) {
  if (g_sor_order == 0) {
    for (int j = 0; j < h; ++j)
      for (int i = 0; i < w; ++i) f(i, j);
  } else {
    for (int colour = 0; colour < 2; ++colour)
      for (int j = 0; j < h; ++j)
        for (int i = (j + colour) & 1; i < w; i += 2) f(i, j);
  }
}

// sor_coupled_slow_but_readable, solver.c:34-78 (used by sor_coupled for degenerate sizes, solver.c:86-89)
void sor_readable(Plane& du, Plane& dv, const Plane& a11, const Plane& a12, const Plane& a22, const Plane& b1,
                  const Plane& b2, const Plane& sh, const Plane& sv, int iters, float omega) {
  const int w = du.w, h = du.h;
  for (int it = 0; it < iters; ++it)
    for (int j = 0; j < h; ++j)
      for (int i = 0; i < w; ++i) {
        float su = 0.f, svv = 0.f, sum = 0.f;
        if (j > 0) { su -= sv.at(i, j - 1) * du.at(i, j - 1); svv -= sv.at(i, j - 1) * dv.at(i, j - 1); sum += sv.at(i, j - 1); }
        if (i > 0) { su -= sh.at(i - 1, j) * du.at(i - 1, j); svv -= sh.at(i - 1, j) * dv.at(i - 1, j); sum += sh.at(i - 1, j); }
        if (j < h - 1) { su -= sv.at(i, j) * du.at(i, j + 1); svv -= sv.at(i, j) * dv.at(i, j + 1); sum += sv.at(i, j); }
        if (i < w - 1) { su -= sh.at(i, j) * du.at(i + 1, j); svv -= sh.at(i, j) * dv.at(i + 1, j); sum += sh.at(i, j); }
        float A11 = a11.at(i, j) + sum, A12 = a12.at(i, j), A22 = a22.at(i, j) + sum;
        float B1 = b1.at(i, j) - su, B2 = b2.at(i, j) - svv;
        du.at(i, j) = (1.0f - omega) * du.at(i, j) + omega / A11 * (B1 - A12 * dv.at(i, j));
        dv.at(i, j) = (1.0f - omega) * dv.at(i, j) + omega / A22 * (B2 - A12 * du.at(i, j));
      }
}

// a20  sor_coupled, solver.c:83-433.  a11/a12/a22 are overwritten by the inverse 2x2 blocks in sweep 1.
void sor_coupled(Plane& du, Plane& dv, Plane& a11, Plane& a12, Plane& a22, const Plane& b1, const Plane& b2,
                 const Plane& sh, const Plane& sv, int iters, float omega) {
  const int w = du.w, h = du.h;
  if (w < 2 || h < 2 || iters < 1) { sor_readable(du, dv, a11, a12, a22, b1, b2, sh, sv, iters, omega); return; }
  for (int it = 0; it < iters; ++it)
    sweep_pixels(w, h, [&](int i, int j) {
        const float hl = i > 0 ? sh.at(i - 1, j) : 0.f;     // f1[0] = 0 (solver.c:102)
        const float hh = sh.at(i, j), vv = sv.at(i, j);      // 0 in last column / row by construction
        const float dur = i < w - 1 ? du.at(i + 1, j) : 0.f, dvr = i < w - 1 ? dv.at(i + 1, j) : 0.f;
        if (it == 0) {
          float dpsis;
          if (j == 0) dpsis = hl + (hh + vv);                               // solver.c:122
          else if (j < h - 1) dpsis = (hl + hh) + (sv.at(i, j - 1) + vv);   // solver.c:184
          else dpsis = hl + (hh + sv.at(i, j - 1));                         // solver.c:244
          const float A11 = a22.at(i, j) + dpsis, A22 = a11.at(i, j) + dpsis, A12 = a12.at(i, j);
          const float det = A11 * A22 - A12 * A12;
          a11.at(i, j) = A11 / det;
          a22.at(i, j) = A22 / det;
          a12.at(i, j) = A12 / (0.f - det);
        }
        float s1, s2;
        if (j == 0) {                                                       // solver.c:131
          s1 = mla(mla(b1.at(i, j), hh, dur), vv, du.at(i, j + 1));
          s2 = mla(mla(b2.at(i, j), hh, dvr), vv, dv.at(i, j + 1));
        } else if (j < h - 1) {                                             // solver.c:193
          const float vt = sv.at(i, j - 1);
          s1 = mla(hh * dur, vt, du.at(i, j - 1)) + mla(b1.at(i, j), vv, du.at(i, j + 1));
          s2 = mla(hh * dvr, vt, dv.at(i, j - 1)) + mla(b2.at(i, j), vv, dv.at(i, j + 1));
        } else {                                                            // solver.c:253
          const float vt = sv.at(i, j - 1);
          s1 = mla(mla(b1.at(i, j), hh, dur), vt, du.at(i, j - 1));
          s2 = mla(mla(b2.at(i, j), hh, dvr), vt, dv.at(i, j - 1));
        }
        float B1 = s1, B2 = s2;
        if (i > 0) { B1 = hl * du.at(i - 1, j) + s1; B2 = hl * dv.at(i - 1, j) + s2; }
        const float nu = du.at(i, j) + omega * (a11.at(i, j) * B1 + a12.at(i, j) * B2 - du.at(i, j));
        const float nv = dv.at(i, j) + omega * (a12.at(i, j) * B1 + a22.at(i, j) * B2 - dv.at(i, j));
        du.at(i, j) = nu;
        dv.at(i, j) = nv;
      });
}

// a21  sor_coupled_slow_but_readable_DE, solver.c:439-471
void sor_DE(Plane& du, const Plane& a11, const Plane& b1, const Plane& sh, const Plane& sv, int iters, float omega) {
  const int w = du.w, h = du.h;
  for (int it = 0; it < iters; ++it)
    sweep_pixels(w, h, [&](int i, int j) {
        float su = 0.f, sum = 0.f;
        if (j > 0) { su -= sv.at(i, j - 1) * du.at(i, j - 1); sum += sv.at(i, j - 1); }
        if (i > 0) { su -= sh.at(i - 1, j) * du.at(i - 1, j); sum += sh.at(i - 1, j); }
        if (j < h - 1) { su -= sv.at(i, j) * du.at(i, j + 1); sum += sv.at(i, j); }
        if (i < w - 1) { su -= sh.at(i, j) * du.at(i + 1, j); sum += sh.at(i, j); }
        float A11 = a11.at(i, j) + sum, B1 = b1.at(i, j) - su;
        du.at(i, j) = (1.0f - omega) * du.at(i, j) + omega * (B1 / A11);
      });
}

// a13 + a22  VarRefClass ctor + RefLevelOF / RefLevelDE, refine_variational.cpp:25-116,153-342
void varref_level(const float* a, const float* b, const Level& L, const OrcParams& p, float* flow) {
  const int w = L.w, h = L.h, nop = p.mode == 1 ? 2 : 1;
  const int n_inner = p.tv_innerit * (L.level + 1);                 // refine_variational.cpp:36
  const float qa = 0.25f * p.tv_alpha, hg3 = p.tv_gamma * 0.5f / 3.0f, hd3 = p.tv_delta * 0.5f / 3.0f;
  Plane wx(w, h), wy(w, h);
  for (int i = 0; i < w * h; ++i) {
    wx.d[i] = flow[(size_t)i * nop];
    if (nop == 2) wy.d[i] = flow[(size_t)i * nop + 1];
  }
  std::vector<Plane> im1, im2, w2(p.noc, Plane(w, h));
  strip_to_planes(a, w, h, L.pad, p.noc, im1);
  strip_to_planes(b, w, h, L.pad, p.noc, im2);
  Plane mask(w, h), du(w, h), dv(w, h), uu = wx, vv = wy, sh(w, h), sv(w, h);
  Plane a11(w, h), a12(w, h), a22(w, h), b1(w, h), b2(w, h);
  Derivs D;
  image_warp(im2, wx, wy, w2, mask);
  get_derivatives(im1, w2, D);
  for (int it = 0; it < n_inner; ++it) {
    compute_smoothness(uu, vv, qa, sh, sv);
    if (p.mode == 1) {
      compute_data(D, mask, du, dv, hd3, hg3, a11, a12, a22, b1, b2);
      sub_laplacian(b1, wx, sh, sv);
      sub_laplacian(b2, wy, sh, sv);
      sor_coupled(du, dv, a11, a12, a22, b1, b2, sh, sv, p.tv_solverit, p.tv_sor);
      for (int i = 0; i < w * h; ++i) { uu.d[i] = wx.d[i] + du.d[i]; vv.d[i] = wy.d[i] + dv.d[i]; }
    } else {
      compute_data_DE(D, mask, du, hd3, hg3, a11, b1);
      sub_laplacian(b1, wx, sh, sv);
      sor_DE(du, a11, b1, sh, sv, p.tv_solverit, p.tv_sor);
      for (int i = 0; i < w * h; ++i) {                               // refine_variational.cpp:305-320
        float t = wx.d[i] + du.d[i];
        uu.d[i] = L.camlr == 0 ? std::min(t, 0.f) : std::max(t, 0.f);
      }
    }
  }
  for (int i = 0; i < w * h; ++i) {
    flow[(size_t)i * nop] = uu.d[i];
    if (nop == 2) flow[(size_t)i * nop + 1] = vv.d[i];
  }
}

}  // namespace

// ===========================================================================
// Scale loop (OFClass ctor, oflow.cpp:31-338) and the run_dense driver
// ===========================================================================
namespace {

void ofclass(const OrcParams& p, const float* const* a, const float* const* adx, const float* const* ady,
             const float* const* b, const float* const* bdx, const float* const* bdy, int W, int H,
             float* outflow, const float* initflow) {
  const Opt o = make_opt(p);
  const int ns = p.sc_f - p.sc_l + 1;
  std::vector<fvec> flow_fw(ns), flow_bw(ns);
  for (int sl = p.sc_f; sl >= p.sc_l; --sl) {                       // oflow.cpp:182
    const int ii = sl - p.sc_l;
    const Level Ll = make_level(W, H, sl, p.psz, p.psz, 0), Lr = make_level(W, H, sl, p.psz, p.psz, 1);
    Grid gf, gb;
    gf.build(Ll, o);
    if (p.usefbcon) gb.build(Lr, o);
    if (sl < p.sc_f) {                                              // oflow.cpp:207-214
      init_from_coarser(gf, Ll, o, flow_fw[ii + 1].data());
      if (p.usefbcon) init_from_coarser(gb, Lr, o, flow_bw[ii + 1].data());
    } else if (initflow) {                                          // oflow.cpp:215-217
      init_from_coarser(gf, Ll, o, initflow);
    }
    optimize_grid(gf, Ll, o, a[sl], adx[sl], ady[sl], b[sl]);        // oflow.cpp:229-232
    if (p.usefbcon) optimize_grid(gb, Lr, o, b[sl], bdx[sl], bdy[sl], a[sl]);
    flow_fw[ii].resize((size_t)o.nop * Ll.w * Ll.h);
    float* dst = (sl == p.sc_l) ? outflow : flow_fw[ii].data();     // oflow.cpp:259-262
    aggregate_dense(gf, p.usefbcon ? &gb : nullptr, Ll, o, dst);
    if (p.usefbcon && sl > p.sc_l) {                                // oflow.cpp:266-267
      flow_bw[ii].resize((size_t)o.nop * Lr.w * Lr.h);
      aggregate_dense(gb, &gf, Lr, o, flow_bw[ii].data());
    }
    if (p.usetvref) {                                               // oflow.cpp:281-290
      varref_level(a[sl], b[sl], Ll, p, dst);
      if (p.usefbcon && sl > p.sc_l) varref_level(b[sl], a[sl], Lr, p, flow_bw[ii].data());
    }
  }
}

struct Pyramid {
  int W, H, padl, padt;
  std::vector<fvec> img, dx, dy;   // padded level arrays, index = level
  std::vector<const float*> pi, px, py;
};

// a1 + a2  (run_dense.cpp:299-312,131-179).  Only levels sc_l..sc_f are kept; finer
// levels are built by the reference but never read (oflow.cpp:182).
void build_pyramid(const uint8_t* im, int w0, int h0, int C, int sc_f, int sc_l, int pad, Pyramid& P, bool gradmag = false) {
  pad_dims(w0, h0, sc_f, &P.W, &P.H, &P.padl, &P.padt);
  P.img.assign(sc_f + 1, fvec()); P.dx.assign(sc_f + 1, fvec()); P.dy.assign(sc_f + 1, fvec());
  fvec cur, nxt, gx, gy;
  pad_to_float(im, w0, h0, C, P.W, P.H, P.padl, P.padt, cur);
  int w = P.W, h = P.H;
  if (gradmag && C == 1) { gradient_magnitude(cur, w, h, nxt); cur.swap(nxt); }
  for (int l = 0; l <= sc_f; ++l) {
    if (l > 0) { halve(cur, w, h, C, nxt); cur.swap(nxt); w /= 2; h /= 2; }
    if (l >= sc_l) {
      sobel(cur, w, h, C, gx, gy);
      size_t n = (size_t)(w + 2 * pad) * (h + 2 * pad) * C;
      P.img[l].resize(n); P.dx[l].resize(n); P.dy[l].resize(n);
      pad_level(cur, w, h, C, pad, true, P.img[l].data());
      pad_level(gx, w, h, C, pad, false, P.dx[l].data());
      pad_level(gy, w, h, C, pad, false, P.dy[l].data());
    }
  }
  P.pi.assign(sc_f + 1, nullptr); P.px.assign(sc_f + 1, nullptr); P.py.assign(sc_f + 1, nullptr);
  for (int l = sc_l; l <= sc_f; ++l) { P.pi[l] = P.img[l].data(); P.px[l] = P.dx[l].data(); P.py[l] = P.dy[l].data(); }
}

}  // namespace

extern "C" {

// run_dense.cpp:181-184,226-268: operating points 1..4 (anything else = 2)
void orc_operating_point(int op, int width, int mode, int noc, OrcParams* p) {
  p->dp_thresh = 0.05f; p->dr_thresh = 0.95f; p->res_thresh = 0.0f;
  p->usefbcon = 0; p->patnorm = 1; p->costfct = 0;
  p->tv_alpha = 10.0f; p->tv_gamma = 10.0f; p->tv_delta = 5.0f;
  p->tv_innerit = 1; p->tv_solverit = 3; p->tv_sor = 1.6f;
  p->verbosity = 2; p->mode = mode; p->noc = noc; p->gradmag = 0;
  const int fratio = 5;
  auto first = [&](int psz) { return std::max(0, (int)std::floor(log2((2.0f * (float)width) / ((float)fratio * (float)psz)))); };
  switch (op) {
    case 1: p->psz = 8; p->patove = 0.3f; p->sc_f = first(8); p->sc_l = std::max(p->sc_f - 2, 0); p->max_iter = p->min_iter = 16; p->usetvref = 0; break;
    case 3: p->psz = 12; p->patove = 0.75f; p->sc_f = first(12); p->sc_l = std::max(p->sc_f - 4, 0); p->max_iter = p->min_iter = 16; p->usetvref = 1; break;
    case 4: p->psz = 12; p->patove = 0.75f; p->sc_f = first(12); p->sc_l = std::max(p->sc_f - 5, 0); p->max_iter = p->min_iter = 128; p->usetvref = 1; break;
    default: p->psz = 8; p->patove = 0.4f; p->sc_f = first(8); p->sc_l = std::max(p->sc_f - 2, 0); p->max_iter = p->min_iter = 12; p->usetvref = 1; break;
  }
}

// 0 raster order (reference), 1 red-black (product's throughput mode); process-global, tests reset it to 0
void orc_set_sor_order(int order) { g_sor_order = order ? 1 : 0; }
// SELECTCHANNEL == 2 for the stage-level entry points orc_pyramid_level (run_dense entry points read OrcParams::gradmag).
void orc_set_pyramid_gradmag(int on) { g_pyr_gradmag = on != 0; }
void orc_gradient_magnitude(const float* src, int w, int h, float* dst) {
  fvec s(src, src + (size_t)w * h), d;
  gradient_magnitude(s, w, h, d);
  std::memcpy(dst, d.data(), d.size() * sizeof(float));
}
void orc_pad_dims(int w0, int h0, int sc_f, int* W, int* H, int* padl, int* padt) { pad_dims(w0, h0, sc_f, W, H, padl, padt); }

// One padded pyramid level (img, dx, dy) from a u8 image; out arrays hold (h_l+2pad)*(w_l+2pad)*C floats.
void orc_pyramid_level(const uint8_t* im, int w0, int h0, int C, int sc_f, int level, int pad, float* img, float* dx, float* dy) {
  Pyramid P;
  build_pyramid(im, w0, h0, C, std::max(sc_f, level), level, pad, P, g_pyr_gradmag);
  // note: padding is decided by sc_f, levels by `level`
  std::memcpy(img, P.img[level].data(), P.img[level].size() * sizeof(float));
  std::memcpy(dx, P.dx[level].data(), P.dx[level].size() * sizeof(float));
  std::memcpy(dy, P.dy[level].data(), P.dy[level].size() * sizeof(float));
}

// All padded levels sc_l..sc_f of one frame from ONE cascade (img[l], dx[l], dy[l] caller-allocated for those levels).
void orc_pyramid_all(const uint8_t* im, int w0, int h0, int C, int sc_f, int sc_l, int pad, float* const* img, float* const* dx,
                     float* const* dy) {
  Pyramid P;
  build_pyramid(im, w0, h0, C, sc_f, sc_l, pad, P, g_pyr_gradmag);
  for (int l = sc_l; l <= sc_f; ++l) {
    std::memcpy(img[l], P.img[l].data(), P.img[l].size() * sizeof(float));
    std::memcpy(dx[l], P.dx[l].data(), P.dx[l].size() * sizeof(float));
    std::memcpy(dy[l], P.dy[l].data(), P.dy[l].size() * sizeof(float));
  }
}

// Float-source variants of the OpenCV steps (for the cv2 golden vectors).
void orc_halve(const float* src, int w, int h, int C, float* dst) {
  fvec s(src, src + (size_t)w * h * C), d;
  halve(s, w, h, C, d);
  std::memcpy(dst, d.data(), d.size() * sizeof(float));
}
void orc_sobel(const float* src, int w, int h, int C, float* dx, float* dy) {
  fvec s(src, src + (size_t)w * h * C), gx, gy;
  sobel(s, w, h, C, gx, gy);
  std::memcpy(dx, gx.data(), gx.size() * sizeof(float));
  std::memcpy(dy, gy.data(), gy.size() * sizeof(float));
}
void orc_upsample_crop(const float* flow, int w_l, int h_l, int nop, int sc_l, int padl, int padt, int w0, int h0, float* out) {
  upsample_crop(flow, w_l, h_l, nop, sc_l, padl, padt, w0, h0, out);
}

void orc_ofclass(const OrcParams* p, const float* const* a, const float* const* adx, const float* const* ady,
                 const float* const* b, const float* const* bdx, const float* const* bdy, int W, int H,
                 float* outflow, const float* initflow) {
  ofclass(*p, a, adx, ady, b, bdx, bdy, W, H, outflow, initflow);
}

// Same contract as ref_patchgrid_level in oracle/ref_wrap.cpp.
int orc_patchgrid_level(const OrcParams* p, const float* a, const float* adx, const float* ady, const float* b,
                        const float* bdx, const float* bdy, int w_l, int h_l, int level,
                        const float* flow_prev_fw, const float* flow_prev_bw, float* p_iter_fw, float* pt_iter_fw,
                        float* pweight_fw, int* cnt_fw, float* dense_fw, float* p_iter_bw, float* pt_iter_bw,
                        float* pweight_bw, float* dense_bw) {
  const Opt o = make_opt(*p);
  Level Ll = make_level(w_l, h_l, 0, p->psz, p->psz, 0);
  Ll.level = level;
  Level Lr = Ll;
  Lr.camlr = 1;
  Grid gf, gb;
  gf.build(Ll, o);
  if (p->usefbcon) gb.build(Lr, o);
  if (flow_prev_fw) init_from_coarser(gf, Ll, o, flow_prev_fw);
  if (p->usefbcon && flow_prev_bw) init_from_coarser(gb, Lr, o, flow_prev_bw);
  optimize_grid(gf, Ll, o, a, adx, ady, b);
  if (p->usefbcon) optimize_grid(gb, Lr, o, b, bdx, bdy, a);
  auto dump = [&](const Grid& g, float* pi, float* pt, float* pw, int* cnt) {
    if (pi) std::memcpy(pi, g.p_iter.data(), g.p_iter.size() * sizeof(float));
    if (pt) std::memcpy(pt, g.pt_iter.data(), g.pt_iter.size() * sizeof(float));
    if (pw) std::memcpy(pw, g.pweight.data(), g.pweight.size() * sizeof(float));
    if (cnt) std::memcpy(cnt, g.cnt.data(), g.cnt.size() * sizeof(int));
  };
  dump(gf, p_iter_fw, pt_iter_fw, pweight_fw, cnt_fw);
  if (p->usefbcon) dump(gb, p_iter_bw, pt_iter_bw, pweight_bw, nullptr);
  if (dense_fw) aggregate_dense(gf, p->usefbcon ? &gb : nullptr, Ll, o, dense_fw);
  if (p->usefbcon && dense_bw) aggregate_dense(gb, &gf, Lr, o, dense_bw);
  return gf.n;
}

// The AggregateFlowDense cut (patchgrid.cpp:213-397): patch results of one level in, dense flow out.  cg_* = results of
// the complementary grid (forward-backward merging) or null.  Arrays are [npatch][nop], [npatch][2], [npatch][novals].
int orc_aggregate_dense(const OrcParams* p, int w_l, int h_l, const float* p_iter, const float* pweight,
                        const float* cg_p_iter, const float* cg_pt_iter, const float* cg_pweight, float* flowout) {
  const Opt o = make_opt(*p);
  const Level L = make_level(w_l, h_l, 0, p->psz, p->psz, 0);
  Grid g, cg;
  g.build(L, o);
  g.p_iter.assign(p_iter, p_iter + (size_t)g.n * o.nop);
  g.pweight.assign(pweight, pweight + (size_t)g.n * o.novals);
  if (cg_p_iter) {
    cg.build(L, o);
    cg.p_iter.assign(cg_p_iter, cg_p_iter + (size_t)cg.n * o.nop);
    cg.pt_iter.assign(cg_pt_iter, cg_pt_iter + (size_t)cg.n * 2);
    cg.pweight.assign(cg_pweight, cg_pweight + (size_t)cg.n * o.novals);
  }
  aggregate_dense(g, cg_p_iter ? &cg : nullptr, L, o, flowout);
  return g.n;
}

void orc_varref_level(const OrcParams* p, const float* a, const float* b, int w_l, int h_l, int level, int camlr, float* flow) {
  Level L = make_level(w_l, h_l, 0, p->psz, p->psz, camlr);
  L.level = level;
  varref_level(a, b, L, *p, flow);
}

// The whole driver: u8 frames in, full-resolution cropped flow out (run_dense.cpp:186-415 minus file I/O).
// out holds h0*w0*nop floats.  If lowres != null it also receives the sc_l flow (H/2^sc_l x W/2^sc_l x nop).
void orc_run_dense(const OrcParams* p, const uint8_t* A, const uint8_t* B, int w0, int h0, float* out, float* lowres) {
  Pyramid PA, PB;
  build_pyramid(A, w0, h0, p->noc, p->sc_f, p->sc_l, p->psz, PA, p->gradmag != 0);
  build_pyramid(B, w0, h0, p->noc, p->sc_f, p->sc_l, p->psz, PB, p->gradmag != 0);
  const int nop = p->mode == 1 ? 2 : 1, k = 1 << p->sc_l;
  const int wl = PA.W / k, hl = PA.H / k;
  fvec flow((size_t)wl * hl * nop);
  ofclass(*p, PA.pi.data(), PA.px.data(), PA.py.data(), PB.pi.data(), PB.px.data(), PB.py.data(), PA.W, PA.H, flow.data(), nullptr);
  if (lowres) std::memcpy(lowres, flow.data(), flow.size() * sizeof(float));
  upsample_crop(flow.data(), wl, hl, nop, p->sc_l, PA.padl, PA.padt, w0, h0, out);
}

// Timing split for the CPU baseline: pyramid+upsample ("pre/post") vs OFClass ("core"), seconds.
void orc_run_dense_timed(const OrcParams* p, const uint8_t* A, const uint8_t* B, int w0, int h0, float* out, double* t_core, double* t_prepost);

}  // extern "C"

#include <chrono>
extern "C" void orc_run_dense_timed(const OrcParams* p, const uint8_t* A, const uint8_t* B, int w0, int h0, float* out, double* t_core, double* t_prepost) {
  typedef std::chrono::steady_clock clk;
  auto t0 = clk::now();
  Pyramid PA, PB;
  build_pyramid(A, w0, h0, p->noc, p->sc_f, p->sc_l, p->psz, PA, p->gradmag != 0);
  build_pyramid(B, w0, h0, p->noc, p->sc_f, p->sc_l, p->psz, PB, p->gradmag != 0);
  const int nop = p->mode == 1 ? 2 : 1, k = 1 << p->sc_l;
  const int wl = PA.W / k, hl = PA.H / k;
  fvec flow((size_t)wl * hl * nop);
  auto t1 = clk::now();
  ofclass(*p, PA.pi.data(), PA.px.data(), PA.py.data(), PB.pi.data(), PB.px.data(), PB.py.data(), PA.W, PA.H, flow.data(), nullptr);
  auto t2 = clk::now();
  upsample_crop(flow.data(), wl, hl, nop, p->sc_l, PA.padl, PA.padt, w0, h0, out);
  auto t3 = clk::now();
  *t_core = std::chrono::duration<double>(t2 - t1).count();
  *t_prepost = std::chrono::duration<double>(t1 - t0).count() + std::chrono::duration<double>(t3 - t2).count();
}
