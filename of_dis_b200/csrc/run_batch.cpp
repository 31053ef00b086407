// disb200_batch — many frame pairs through one handle: the throughput-side caller of the hot path
// (decode, batched flow, asynchronous writers).  The reference has no such
// tool; its main() (run_dense.cpp:186-426) handles one pair per process.  Same parameter grammar:
//   disb200_batch [--de] [--rgb] [--batch B] [--threads T] [--device D] [--video] list.txt [X | p1 ... p20]
// list.txt holds one job per line: "img1 img2 out" — or, with --video, "frame out": consecutive lines are
// consecutive frames, flow i goes from frame i to frame i+1 and is written to the `out` of line i (the last
// line's `out` is unused), each frame's pyramid being built once (disb200_run_sequence_host).
// While the GPU works on one batch, T host threads decode the next one and the writer pool stores the last.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "../../include/disb200.h"

struct Job { std::string a, b, out; };

struct Chunk {
  std::vector<uint8_t> A, B;       // n (video: n + 1 in A) frames, contiguous
  int n = 0, rc = 0;
  std::string err;
};

static void decode_chunk(const std::vector<Job>& jobs, size_t lo, int n, bool video, int channels, int w0, int h0,
                         int threads, Chunk* c) {
  const size_t fb = (size_t)w0 * h0 * channels;
  const int nframes = video ? n + 1 : 2 * n;
  c->n = n;
  c->A.resize(fb * (video ? n + 1 : n));
  if (!video) c->B.resize(fb * n);
  std::atomic<int> next(0), bad(0);
  auto work = [&] {
    for (int k = next++; k < nframes; k = next++) {
      const bool second = !video && k >= n;
      const std::string& path = video ? jobs[lo + k].a : (second ? jobs[lo + k - n].b : jobs[lo + k].a);
      uint8_t* px = nullptr;
      int w = 0, h = 0;
      int rc = disb200_read_image(path.c_str(), channels, &px, &w, &h);
      if (!rc && (w != w0 || h != h0)) rc = DISB200_EINVAL;
      if (rc) {
        if (!bad.exchange(1)) { c->rc = rc; c->err = path; }
      } else {
        std::memcpy((second ? c->B.data() + fb * (k - n) : c->A.data() + fb * k), px, fb);
      }
      disb200_free(px);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work);
  work();
  for (std::thread& t : pool) t.join();
}

int main(int argc, char** argv) {
  int mode = DISB200_MODE_OF, channels = 1, batch = 64, threads = 8, device = 0, gradmag = 0;
  bool video = false;
  int ai = 1;
  for (; ai < argc && argv[ai][0] == '-' && argv[ai][1] == '-'; ++ai) {
    const std::string f = argv[ai];
    if (f == "--de") mode = DISB200_MODE_DE;
    else if (f == "--rgb") channels = 3;
    else if (f == "--gradmag") gradmag = 1;   // SELECTCHANNEL == 2 (run_dense.cpp:139-148)
    else if (f == "--video") video = true;
    else if (f == "--batch" && ai + 1 < argc) batch = std::atoi(argv[++ai]);
    else if (f == "--threads" && ai + 1 < argc) threads = std::atoi(argv[++ai]);
    else if (f == "--device" && ai + 1 < argc) device = std::atoi(argv[++ai]);
    else { std::fprintf(stderr, "unknown option %s\n", f.c_str()); return 1; }
  }
  const int rest = argc - ai;
  if ((rest != 1 && rest != 2 && rest != 21) || batch < 1 || threads < 1) {
    std::fprintf(stderr, "usage: %s [--de] [--rgb] [--gradmag] [--video] [--batch B] [--threads T] [--device D] list.txt "
                         "[oppoint 1..4 | 20 parameters in the order of run_*_*]\n", argv[0]);
    return 1;
  }
  std::vector<Job> jobs;
  {
    std::ifstream in(argv[ai]);
    if (!in) { std::fprintf(stderr, "cannot open %s\n", argv[ai]); return 1; }
    std::string line;
    while (std::getline(in, line)) {
      std::istringstream ss(line);
      Job j;
      if (!(ss >> j.a)) continue;
      if (j.a[0] == '#') continue;
      if (video) { ss >> j.out; }
      else if (!(ss >> j.b >> j.out)) { std::fprintf(stderr, "malformed line: %s\n", line.c_str()); return 1; }
      jobs.push_back(j);
    }
  }
  const size_t npairs = video ? (jobs.empty() ? 0 : jobs.size() - 1) : jobs.size();
  if (npairs == 0) { std::fprintf(stderr, "nothing to do\n"); return 1; }
  // geometry from the first image
  int w0 = 0, h0 = 0;
  {
    uint8_t* px = nullptr;
    int rc = disb200_read_image(jobs[0].a.c_str(), channels, &px, &w0, &h0);
    disb200_free(px);
    if (rc) { std::fprintf(stderr, "cannot read %s\n", jobs[0].a.c_str()); return rc; }
  }
  DisParams p{};
  int rc;
  if (rest == 21) {
    char** a = argv + ai + 1;
    p.sc_f = std::atoi(a[0]); p.sc_l = std::atoi(a[1]); p.max_iter = std::atoi(a[2]); p.min_iter = std::atoi(a[3]);
    p.dp_thresh = (float)std::atof(a[4]); p.dr_thresh = (float)std::atof(a[5]); p.res_thresh = (float)std::atof(a[6]);
    p.patchsz = std::atoi(a[7]); p.patove = (float)std::atof(a[8]); p.usefbcon = std::atoi(a[9]);
    p.patnorm = std::atoi(a[10]); p.costfct = std::atoi(a[11]); p.usetvref = std::atoi(a[12]);
    p.tv_alpha = (float)std::atof(a[13]); p.tv_gamma = (float)std::atof(a[14]); p.tv_delta = (float)std::atof(a[15]);
    p.tv_innerit = std::atoi(a[16]); p.tv_solverit = std::atoi(a[17]); p.tv_sor = (float)std::atof(a[18]);
    p.verbosity = std::atoi(a[19]);
    p.mode = mode; p.channels = channels; p.sor_mode = DISB200_SOR_LEXICOGRAPHIC;
  } else {
    rc = disb200_operating_point(rest == 2 ? std::atoi(argv[ai + 1]) : 2, w0, mode, channels, &p);
    if (rc) { std::fprintf(stderr, "operating point: %s\n", disb200_last_error()); return rc; }
  }
  p.gradmag = gradmag;
  if ((size_t)batch > npairs) batch = (int)npairs;
  DisHandle* h = nullptr;
  rc = disb200_create(&p, w0, h0, batch, device, &h);
  if (rc) { std::fprintf(stderr, "disb200_create: %s\n", disb200_last_error()); return rc; }
  DisWriter* wr = nullptr;
  rc = disb200_writer_create(threads, (size_t)4 << 30, &wr);
  if (rc) { disb200_destroy(h); return rc; }
  const int nop = mode == DISB200_MODE_OF ? 2 : 1;
  const size_t ob = (size_t)w0 * h0 * nop;
  std::vector<float> out(ob * batch);
  const auto t0 = std::chrono::steady_clock::now();

  Chunk cur, nxt;
  decode_chunk(jobs, 0, (int)std::min<size_t>(batch, npairs), video, channels, w0, h0, threads, &cur);
  size_t done = 0;
  while (done < npairs && !rc) {
    if (cur.rc) { std::fprintf(stderr, "cannot read %s (or its size differs from %dx%d)\n", cur.err.c_str(), w0, h0); rc = cur.rc; break; }
    const size_t lo_next = done + cur.n;
    const int n_next = (int)std::min<size_t>(batch, npairs - std::min(npairs, lo_next));
    std::thread prefetch;
    if (n_next > 0) prefetch = std::thread(decode_chunk, std::cref(jobs), lo_next, n_next, video, channels, w0, h0, std::max(1, threads - 1), &nxt);
    rc = video ? disb200_run_sequence_host(h, cur.A.data(), cur.n + 1, out.data())
               : disb200_run_batch_host(h, cur.A.data(), cur.B.data(), cur.n, out.data());
    if (rc) std::fprintf(stderr, "run: %s\n", disb200_last_error());
    for (int i = 0; i < cur.n && !rc; ++i) {
      const std::string& path = jobs[done + i].out;
      if (path.empty()) continue;
      rc = mode == DISB200_MODE_OF ? disb200_writer_submit_flo(wr, path.c_str(), out.data() + ob * i, w0, h0, 2)
                                   : disb200_writer_submit_pfm(wr, path.c_str(), out.data() + ob * i, w0, h0);
    }
    if (prefetch.joinable()) prefetch.join();
    done += cur.n;
    std::swap(cur, nxt);
  }
  const int wrc = disb200_writer_wait(wr);
  if (!rc && wrc) { std::fprintf(stderr, "writing an output file failed\n"); rc = wrc; }
  const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (!rc) std::fprintf(stderr, "%zu pairs of %dx%d in %.3f s (%.1f pairs/s, decode + flow + write)\n", done, w0, h0, s, done / s);
  disb200_writer_destroy(wr);
  disb200_destroy(h);
  return rc;
}
