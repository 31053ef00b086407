// Host-side file I/O of the drop-in: the reference's writers (run_dense.cpp:17-82) and a
// PGM/PPM reader standing in for cv::imread (run_dense.cpp:209-210; OpenCV is not available,
// SURVEY.md 2 #12).  Plain C++, no device code.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/disb200.h"

extern "C" {

// SaveFlowFile, run_dense.cpp:17-58: "PIEH", int32 width, int32 height, row-major interleaved f32.
int disb200_save_flo(const char* path, const float* flow, int w, int h, int nc) {
  if (!path || !flow || w <= 0 || h <= 0 || (nc != 1 && nc != 2 && nc != 4)) return DISB200_EINVAL;
  FILE* f = std::fopen(path, "wb");
  if (!f) return DISB200_EIO;
  bool ok = std::fwrite("PIEH", 1, 4, f) == 4;
  ok = ok && std::fwrite(&w, sizeof(int), 1, f) == 1 && std::fwrite(&h, sizeof(int), 1, f) == 1;
  const size_t n = (size_t)w * h * nc;
  ok = ok && std::fwrite(flow, sizeof(float), n, f) == n;
  ok = (std::fclose(f) == 0) && ok;
  return ok ? DISB200_OK : DISB200_EIO;
}

// SavePFMFile, run_dense.cpp:61-82: "Pf\n<w> <h>\n-1.000000\n", rows bottom-up, value = -disparity.
int disb200_save_pfm(const char* path, const float* disp, int w, int h) {
  if (!path || !disp || w <= 0 || h <= 0) return DISB200_EINVAL;
  FILE* f = std::fopen(path, "wb");
  if (!f) return DISB200_EIO;
  bool ok = std::fprintf(f, "Pf\n%d %d\n%f\n", w, h, (float)-1.0f) > 0;
  std::vector<float> row(w);
  for (int y = h - 1; y >= 0 && ok; --y) {
    for (int x = 0; x < w; ++x) row[x] = -disp[(size_t)y * w + x];
    ok = std::fwrite(row.data(), sizeof(float), w, f) == (size_t)w;
  }
  ok = (std::fclose(f) == 0) && ok;
  return ok ? DISB200_OK : DISB200_EIO;
}

static bool pnm_token(FILE* f, int* v) {
  int c = std::fgetc(f);
  while (c != EOF) {
    if (c == '#') { while (c != EOF && c != '\n') c = std::fgetc(f); }
    else if (c == ' ' || c == '\t' || c == '\n' || c == '\r') c = std::fgetc(f);
    else break;
  }
  if (c == EOF || c < '0' || c > '9') return false;
  int x = 0;
  while (c >= '0' && c <= '9') { x = x * 10 + (c - '0'); c = std::fgetc(f); }
  *v = x;
  return true;  // the single whitespace after the token has been consumed
}

// Binary PGM (P5) / PPM (P6), maxval 255.  want_channels 1: grey (colour input is converted
// with OpenCV's 8-bit BGR2GRAY fixed-point weights, as imread(IMREAD_GRAYSCALE) does);
// want_channels 3: BGR interleaved (grey input is replicated), matching cv::imread's order.
int disb200_read_pnm(const char* path, int want_channels, uint8_t** data, int* w, int* h) {
  if (!path || !data || !w || !h || (want_channels != 1 && want_channels != 3)) return DISB200_EINVAL;
  FILE* f = std::fopen(path, "rb");
  if (!f) return DISB200_EIO;
  char magic[3] = {0, 0, 0};
  int W = 0, H = 0, maxv = 0;
  bool ok = std::fread(magic, 1, 2, f) == 2 && magic[0] == 'P' && (magic[1] == '5' || magic[1] == '6');
  ok = ok && pnm_token(f, &W) && pnm_token(f, &H) && pnm_token(f, &maxv) && W > 0 && H > 0 && maxv == 255;
  if (!ok) { std::fclose(f); return DISB200_EIO; }
  const int sc = magic[1] == '6' ? 3 : 1;
  std::vector<uint8_t> raw((size_t)W * H * sc);
  ok = std::fread(raw.data(), 1, raw.size(), f) == raw.size();
  std::fclose(f);
  if (!ok) return DISB200_EIO;
  uint8_t* out = (uint8_t*)std::malloc((size_t)W * H * want_channels);
  if (!out) return DISB200_ENOMEM;
  for (size_t i = 0; i < (size_t)W * H; ++i) {
    if (sc == 1 && want_channels == 1) out[i] = raw[i];
    else if (sc == 1) out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = raw[i];
    else if (want_channels == 3) { out[3 * i] = raw[3 * i + 2]; out[3 * i + 1] = raw[3 * i + 1]; out[3 * i + 2] = raw[3 * i]; }
    else out[i] = (uint8_t)((raw[3 * i] * 4899 + raw[3 * i + 1] * 9617 + raw[3 * i + 2] * 1868 + 8192) >> 14);
  }
  *data = out; *w = W; *h = H;
  return DISB200_OK;
}

void disb200_free(void* p) { std::free(p); }

}  // extern "C"
