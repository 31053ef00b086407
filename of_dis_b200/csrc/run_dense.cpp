// run_OF_INT / run_OF_RGB / run_DE_INT / run_DE_RGB — the four command lines of the reference
// (CMakeLists.txt:35-61), as thin mains over libdisb200.so.  SELECTMODE / SELECTCHANNEL are set
// by the build (of_dis_b200/build.py) exactly like the reference's four targets.
//
// Grammar (run_dense.cpp:186-295, README.md:46-86):
//   run_*_* img1 img2 out                  operating point 2
//   run_*_* img1 img2 out X                operating point X in 1..4
//   run_*_* img1 img2 out p1 ... p20       sc_f sc_l max_iter min_iter dp_thresh dr_thresh res_thresh
//                                          patchsz overlap usefbcon patnorm costfct usetvref tv_alpha
//                                          tv_gamma tv_delta tv_innerit tv_solverit tv_sor verbosity
// Inputs are binary PGM / PPM or PNG (disb200_read_image; there is no OpenCV here: SURVEY.md 2 #12);
// output is .flo for optical flow and .pfm for depth-from-stereo (run_dense.cpp:418-422).
// Unlike the reference, a short or malformed argument list is an error (the reference reads past
// argv for argc in 6..23, run_dense.cpp:272-292) and every failure returns a non-zero status.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/disb200.h"

#ifndef SELECTMODE
#define SELECTMODE 1
#endif
#ifndef SELECTCHANNEL
#define SELECTCHANNEL 1
#endif

static double ms_since(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

static int die(const char* what, int rc) {
  std::fprintf(stderr, "%s: %s (code %d)\n", what, disb200_last_error(), rc);
  return rc ? rc : 1;
}

int main(int argc, char** argv) {
  const auto t_all = std::chrono::steady_clock::now();
  if (argc != 4 && argc != 5 && argc != 24) {
    std::fprintf(stderr,
                 "usage: %s img1.{pgm,ppm,png} img2.{pgm,ppm,png} out.%s [oppoint 1..4 | sc_f sc_l max_iter min_iter dp_thresh "
                 "dr_thresh res_thresh patchsz overlap usefbcon patnorm costfct usetvref tv_alpha tv_gamma tv_delta "
                 "tv_innerit tv_solverit tv_sor verbosity]\n",
                 argv[0], SELECTMODE == 1 ? "flo" : "pfm");
    return 1;
  }
  const int channels = SELECTCHANNEL == 3 ? 3 : 1;
  auto t0 = std::chrono::steady_clock::now();
  uint8_t *A = nullptr, *B = nullptr;
  int w0 = 0, h0 = 0, w1 = 0, h1 = 0;
  int rc = disb200_read_image(argv[1], channels, &A, &w0, &h0);
  if (rc) { std::fprintf(stderr, "cannot read %s (binary PGM / PPM or non-interlaced PNG expected)\n", argv[1]); return rc; }
  rc = disb200_read_image(argv[2], channels, &B, &w1, &h1);
  if (rc) { std::fprintf(stderr, "cannot read %s (binary PGM / PPM or non-interlaced PNG expected)\n", argv[2]); return rc; }
  if (w0 != w1 || h0 != h1) { std::fprintf(stderr, "image sizes differ: %dx%d vs %dx%d\n", w0, h0, w1, h1); return 1; }
  const double ms_load = ms_since(t0);

  DisParams p;
  if (argc <= 5) {
    rc = disb200_operating_point(argc == 5 ? std::atoi(argv[4]) : 2, w0, SELECTMODE, channels, &p);
    if (rc) return die("operating point", rc);
  } else {
    disb200_operating_point(2, w0, SELECTMODE, channels, &p);   // fills mode / channels / sor_mode
    int a = 4;
    p.sc_f = std::atoi(argv[a++]); p.sc_l = std::atoi(argv[a++]);
    p.max_iter = std::atoi(argv[a++]); p.min_iter = std::atoi(argv[a++]);
    p.dp_thresh = (float)std::atof(argv[a++]); p.dr_thresh = (float)std::atof(argv[a++]);
    p.res_thresh = (float)std::atof(argv[a++]);
    p.patchsz = std::atoi(argv[a++]); p.patove = (float)std::atof(argv[a++]);
    p.usefbcon = std::atoi(argv[a++]); p.patnorm = std::atoi(argv[a++]); p.costfct = std::atoi(argv[a++]);
    p.usetvref = std::atoi(argv[a++]);
    p.tv_alpha = (float)std::atof(argv[a++]); p.tv_gamma = (float)std::atof(argv[a++]);
    p.tv_delta = (float)std::atof(argv[a++]);
    p.tv_innerit = std::atoi(argv[a++]); p.tv_solverit = std::atoi(argv[a++]); p.tv_sor = (float)std::atof(argv[a++]);
    p.verbosity = std::atoi(argv[a++]);
  }
  p.gradmag = SELECTCHANNEL == 2;   // gradient-magnitude input (run_dense.cpp:139-148)
  if (const char* e = std::getenv("DISB200_SOR")) p.sor_mode = (e[0] == 'r') ? DISB200_SOR_REDBLACK : DISB200_SOR_LEXICOGRAPHIC;

  if (p.verbosity > 1) {
    int W, H, pl, pt;
    disb200_padded_dims(w0, h0, p.sc_f, &W, &H, &pl, &pt);
    std::printf("TIME (Image loading     ) (ms): %3g\n", ms_load);
    std::printf("Frame %dx%dx%d padded to %dx%d, scales %d..%d, patch %d, overlap %g\n", w0, h0, channels, W, H, p.sc_f,
                p.sc_l, p.patchsz, p.patove);
  }
  t0 = std::chrono::steady_clock::now();
  DisHandle* h = nullptr;
  rc = disb200_create(&p, w0, h0, 1, 0, &h);
  if (rc) return die("disb200_create", rc);
  const double ms_create = ms_since(t0);

  const int nop = SELECTMODE == 1 ? 2 : 1;
  std::vector<float> out((size_t)w0 * h0 * nop);
  const uint8_t* pa[1] = {A};
  const uint8_t* pb[1] = {B};
  float* po[1] = {out.data()};
  if (p.verbosity > 1) disb200_set_stage_timing(h, 1);
  t0 = std::chrono::steady_clock::now();
  rc = disb200_run_batch(h, pa, pb, 1, po);
  if (rc) return die("disb200_run_batch", rc);
  const double ms_run = ms_since(t0);
  if (p.verbosity > 1) {
    static const char* names[5] = {"Pyramide+Gradients", "Inverse search    ", "Densification     ",
                                   "Variational refin.", "Upsample+crop     "};
    std::printf("TIME (Device setup       ) (ms): %3g\n", ms_create);
    // the reference's per-scale line (oflow.cpp:297); template extraction, start values and the search are ONE kernel
    // here, so its time is reported as poptim and pconst / pinit are 0
    for (int sl = p.sc_f; sl >= p.sc_l; --sl) {
      float se = 0.f, de = 0.f, tv = 0.f;
      int np = 0;
      disb200_scale_ms(h, sl, &se, &de, &tv, &np);
      std::printf("TIME (Sc: %i, #p:%6i, pconst, pinit, poptim, cflow, tvopt, total): %8.2f %8.2f %8.2f %8.2f %8.2f -> %8.2f ms.\n", sl, np,
                  0.0, 0.0, se, de, tv, se + de + tv);
    }
    for (int s = 0; s < 5; ++s) {
      float ms = 0.f;
      disb200_stage_ms(h, s, &ms);
      std::printf("TIME (%s ) (ms): %3g\n", names[s], ms);
    }
  }
  if (p.verbosity > 0) std::printf("TIME (O.Flow Run-Time   ) (ms): %3g   [%d kernel launches]\n", ms_run, disb200_last_launch_count(h));

  t0 = std::chrono::steady_clock::now();
  rc = SELECTMODE == 1 ? disb200_save_flo(argv[3], out.data(), w0, h0, 2) : disb200_save_pfm(argv[3], out.data(), w0, h0);
  if (rc) { std::fprintf(stderr, "cannot write %s\n", argv[3]); return rc; }
  if (p.verbosity > 1) std::printf("TIME (Saving flow file  ) (ms): %3g\n", ms_since(t0));
  disb200_destroy(h);
  disb200_free(A);
  disb200_free(B);
  if (p.verbosity > 0) std::printf("TIME (Total              ) (ms): %3g\n", ms_since(t_all));
  return 0;
}
