// Host-side file I/O of the drop-in: the reference's writers (run_dense.cpp:17-82), a PGM/PPM/PNG
// reader standing in for cv::imread (run_dense.cpp:209-210; OpenCV is not available, SURVEY.md 2 #12;
// PNG through zlib's inflate only) and an asynchronous writer pool (8f rank 2:
// at >10 k pairs/s the 3.5 MB per pair of output must not be written by the thread that drives the GPU).
// Plain C++, no device code.
#include <zlib.h>

#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <new>
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

// ReadFlowFile, run_dense.cpp:84-128 (the reader of the disabled warm-start block, run_dense.cpp:356-379): 4-byte tag,
// int32 width, int32 height, then width * height * nc f32.  The file does not store nc: like the reference
// (img.channels()) the caller says what it expects.  Where the reference only prints ("problem reading", "file is too
// short", "file is too long") this returns DISB200_EIO; the tag is checked against "PIEH" as well.
int disb200_read_flo(const char* path, int nc, float** data, int* w, int* h) {
  if (!path || !data || !w || !h || (nc != 1 && nc != 2 && nc != 4)) return DISB200_EINVAL;
  *data = nullptr;
  FILE* f = std::fopen(path, "rb");
  if (!f) return DISB200_EIO;
  char tag[4];
  int ww = 0, hh = 0;
  bool ok = std::fread(tag, 1, 4, f) == 4 && std::memcmp(tag, "PIEH", 4) == 0;
  ok = ok && std::fread(&ww, sizeof(int), 1, f) == 1 && std::fread(&hh, sizeof(int), 1, f) == 1;
  ok = ok && ww > 0 && hh > 0 && (size_t)ww * hh <= ((size_t)1 << 31);
  float* buf = nullptr;
  if (ok) {
    const size_t n = (size_t)ww * hh * nc;
    buf = static_cast<float*>(std::malloc(n * sizeof(float)));
    if (!buf) { std::fclose(f); return DISB200_ENOMEM; }
    ok = std::fread(buf, sizeof(float), n, f) == n;        // "file is too short"
    ok = ok && std::fgetc(f) == EOF;                        // "file is too long"
  }
  std::fclose(f);
  if (!ok) { std::free(buf); return DISB200_EIO; }
  *data = buf; *w = ww; *h = hh;
  return DISB200_OK;
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

// Largest image the readers accept (w * h), like disb200_read_flo: bounds every allocation derived from a file header.
static const long long DISB200_MAX_PIXELS = 1ll << 31;

static bool pnm_token(FILE* f, int* v) {
  int c = std::fgetc(f);
  while (c != EOF) {
    if (c == '#') { while (c != EOF && c != '\n') c = std::fgetc(f); }
    else if (c == ' ' || c == '\t' || c == '\n' || c == '\r') c = std::fgetc(f);
    else break;
  }
  if (c == EOF || c < '0' || c > '9') return false;
  long long x = 0;
  while (c >= '0' && c <= '9') {
    x = x * 10 + (c - '0');
    if (x > (1ll << 30)) return false;   // no header field is that large: refuse instead of overflowing
    c = std::fgetc(f);
  }
  *v = (int)x;
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
  if (ok && (long long)W * H > DISB200_MAX_PIXELS) { std::fclose(f); return DISB200_EUNSUPPORTED; }
  if (!ok) { std::fclose(f); return DISB200_EIO; }
  const int sc = magic[1] == '6' ? 3 : 1;
  std::vector<uint8_t> raw;
  try { raw.resize((size_t)W * H * sc); } catch (const std::bad_alloc&) { std::fclose(f); return DISB200_ENOMEM; }
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

// ------------------------------------------------------------------------------------ PNG
namespace {

inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

inline int paeth(int a, int b, int c) {
  const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// Decodes a non-interlaced PNG of any colour type / bit depth to 8-bit samples: `ch` = 1 (grey) or 3 (RGB).
// 16-bit samples keep their high byte and alpha is dropped, as OpenCV's decoder does for 8-bit reads
// (png_set_strip_16 / png_set_strip_alpha); tRNS is ignored.  libpng converts colour to grey BEFORE it drops
// the low byte, so 16-bit colour read as grey (`want_grey`) is reduced here, at 16 bits and with libpng's
// rounding for that depth: ((9797 R + 19234 G + 3737 B + 16384) >> 15) >> 8.
int decode_png(const std::vector<uint8_t>& file, bool want_grey, std::vector<uint8_t>* pix, int* W, int* H, int* ch) {
  static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  if (file.size() < 8 + 25 || std::memcmp(file.data(), sig, 8) != 0) return DISB200_EIO;
  size_t pos = 8;
  uint32_t w = 0, h = 0;
  int depth = 0, ctype = -1, interlace = 0;
  std::vector<uint8_t> idat, plte;
  bool end = false;
  while (!end && pos + 12 <= file.size()) {
    const uint32_t len = be32(&file[pos]);
    const uint8_t* type = &file[pos + 4];
    if (pos + 12 + (size_t)len > file.size()) return DISB200_EIO;
    const uint8_t* body = &file[pos + 8];
    if (!std::memcmp(type, "IHDR", 4)) {
      if (len != 13) return DISB200_EIO;
      w = be32(body); h = be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12];
      if (body[10] != 0 || body[11] != 0) return DISB200_EIO;
    } else if (!std::memcmp(type, "PLTE", 4)) {
      plte.assign(body, body + len);
    } else if (!std::memcmp(type, "IDAT", 4)) {
      idat.insert(idat.end(), body, body + len);
    } else if (!std::memcmp(type, "IEND", 4)) {
      end = true;
    }
    pos += 12 + (size_t)len;
  }
  if (ctype < 0 || w == 0 || h == 0 || w > (1u << 20) || h > (1u << 20) || idat.empty()) return DISB200_EIO;
  if ((long long)w * h > DISB200_MAX_PIXELS) return DISB200_EUNSUPPORTED;
  if (interlace != 0) return DISB200_EUNSUPPORTED;     // Adam7
  int spp;                                              // samples per pixel in the file
  switch (ctype) {
    case 0: spp = 1; break;
    case 2: spp = 3; break;
    case 3: spp = 1; break;
    case 4: spp = 2; break;
    case 6: spp = 4; break;
    default: return DISB200_EIO;
  }
  const bool depth_ok = ctype == 0 ? (depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)
                      : ctype == 3 ? (depth == 1 || depth == 2 || depth == 4 || depth == 8)
                                   : (depth == 8 || depth == 16);
  if (!depth_ok || (ctype == 3 && plte.size() < 3)) return DISB200_EIO;
  const size_t rowbytes = ((size_t)w * spp * depth + 7) / 8;
  const int bpp = std::max(1, spp * depth / 8);         // filter distance in bytes
  std::vector<uint8_t> raw((rowbytes + 1) * h);
  {
    z_stream z;
    std::memset(&z, 0, sizeof(z));
    if (inflateInit(&z) != Z_OK) return DISB200_ENOMEM;
    // zlib counts in 32-bit uInt: feed and drain in chunks so that buffers beyond 4 GiB are not truncated
    size_t in_pos = 0, out_pos = 0;
    int zr = Z_OK;
    const size_t CH = (size_t)1 << 30;
    while (zr == Z_OK || zr == Z_BUF_ERROR) {
      if (z.avail_in == 0 && in_pos < idat.size()) {
        const size_t n = std::min(CH, idat.size() - in_pos);
        z.next_in = idat.data() + in_pos; z.avail_in = (uInt)n; in_pos += n;
      }
      if (z.avail_out == 0 && out_pos < raw.size()) {
        const size_t n = std::min(CH, raw.size() - out_pos);
        z.next_out = raw.data() + out_pos; z.avail_out = (uInt)n; out_pos += n;
      }
      const uInt in_before = z.avail_in, out_before = z.avail_out;
      zr = inflate(&z, Z_NO_FLUSH);
      if (zr == Z_BUF_ERROR && in_before == z.avail_in && out_before == z.avail_out &&
          (in_pos >= idat.size() || z.avail_in > 0) && (out_pos >= raw.size() || z.avail_out > 0))
        break;   // no progress possible: truncated stream or more data than the image holds
    }
    const bool full = (size_t)(out_pos - z.avail_out) == raw.size();
    inflateEnd(&z);
    if (zr != Z_STREAM_END || !full) return DISB200_EIO;
  }
  // undo the row filters in place (PNG spec 9.2)
  std::vector<uint8_t> zero(rowbytes, 0);
  for (uint32_t y = 0; y < h; ++y) {
    uint8_t* cur = &raw[(rowbytes + 1) * y + 1];
    const uint8_t* up = y ? &raw[(rowbytes + 1) * (y - 1) + 1] : zero.data();
    const int ft = cur[-1];
    for (size_t i = 0; i < rowbytes; ++i) {
      const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, b = up[i], c = i >= (size_t)bpp ? up[i - bpp] : 0;
      int v = cur[i];
      switch (ft) {
        case 0: break;
        case 1: v += a; break;
        case 2: v += b; break;
        case 3: v += (a + b) >> 1; break;
        case 4: v += paeth(a, b, c); break;
        default: return DISB200_EIO;
      }
      cur[i] = (uint8_t)v;
    }
  }
  const bool colour = ctype == 2 || ctype == 6 || ctype == 3;
  const bool grey16 = want_grey && colour && depth == 16;
  *ch = colour && !grey16 ? 3 : 1;
  *W = (int)w; *H = (int)h;
  pix->resize((size_t)w * h * *ch);
  const int step = depth == 16 ? 2 : 1;
  for (uint32_t y = 0; y < h; ++y) {
    const uint8_t* r = &raw[(rowbytes + 1) * y + 1];
    uint8_t* o = &(*pix)[(size_t)y * w * *ch];
    for (uint32_t x = 0; x < w; ++x) {
      if (depth >= 8) {
        const uint8_t* s = r + (size_t)x * spp * step;
        if (ctype == 3) { const size_t k = (size_t)s[0] * 3; for (int c = 0; c < 3; ++c) o[3 * x + c] = k + 2 < plte.size() ? plte[k + c] : 0; }
        else if (grey16) {
          const uint32_t r = (s[0] << 8) | s[1], g = (s[2] << 8) | s[3], b = (s[4] << 8) | s[5];
          o[x] = (uint8_t)(((r * 9797u + g * 19234u + b * 3737u + 16384u) >> 15) >> 8);
        }
        else if (colour) { o[3 * x] = s[0]; o[3 * x + 1] = s[step]; o[3 * x + 2] = s[2 * step]; }
        else o[x] = s[0];
      } else {
        const int per = 8 / depth, sh = (per - 1 - (int)(x % per)) * depth;
        const int v = (r[x / per] >> sh) & ((1 << depth) - 1);
        if (ctype == 3) { const size_t k = (size_t)v * 3; for (int c = 0; c < 3; ++c) o[3 * x + c] = k + 2 < plte.size() ? plte[k + c] : 0; }
        else o[x] = (uint8_t)(v * 255 / ((1 << depth) - 1));
      }
    }
  }
  return DISB200_OK;
}

}  // namespace

extern "C" {

// cv::imread stand-in for PGM / PPM / PNG (run_dense.cpp:209-210), same output conventions as
// disb200_read_pnm.  Colour PNG -> grey is libpng's rgb_to_gray as OpenCV's PNG decoder sets it up for
// IMREAD_GRAYSCALE, png_set_rgb_to_gray(1, 0.299, 0.587): (9797 R + 19234 G + 3737 B) >> 15, truncated —
// not cvtColor's rounding (pinned bit for bit against cv2 4.13: tests/golden/make_png_golden.py).
int disb200_read_image(const char* path, int want_channels, uint8_t** data, int* w, int* h) {
  if (!path || !data || !w || !h || (want_channels != 1 && want_channels != 3)) return DISB200_EINVAL;
  FILE* f = std::fopen(path, "rb");
  if (!f) return DISB200_EIO;
  uint8_t magic[2] = {0, 0};
  const bool got = std::fread(magic, 1, 2, f) == 2;
  if (!got) { std::fclose(f); return DISB200_EIO; }
  if (magic[0] == 'P') { std::fclose(f); return disb200_read_pnm(path, want_channels, data, w, h); }
  std::fseek(f, 0, SEEK_END);
  const long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  if (sz <= 0) { std::fclose(f); return DISB200_EIO; }
  std::vector<uint8_t> file, pix;
  int W = 0, H = 0, ch = 0, rc = DISB200_OK;
  try {   // no C++ exception may cross the C ABI: an image too large for this host is an error code
    file.resize((size_t)sz);
    const bool ok = std::fread(file.data(), 1, file.size(), f) == file.size();
    std::fclose(f);
    f = nullptr;
    if (!ok) return DISB200_EIO;
    rc = decode_png(file, want_channels == 1, &pix, &W, &H, &ch);
  } catch (const std::bad_alloc&) {
    if (f) std::fclose(f);
    return DISB200_ENOMEM;
  }
  if (rc) return rc;
  uint8_t* out = (uint8_t*)std::malloc((size_t)W * H * want_channels);
  if (!out) return DISB200_ENOMEM;
  for (size_t i = 0; i < (size_t)W * H; ++i) {
    if (ch == 1 && want_channels == 1) out[i] = pix[i];
    else if (ch == 1) out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = pix[i];
    else if (want_channels == 3) { out[3 * i] = pix[3 * i + 2]; out[3 * i + 1] = pix[3 * i + 1]; out[3 * i + 2] = pix[3 * i]; }
    else out[i] = (uint8_t)((pix[3 * i] * 9797u + pix[3 * i + 1] * 19234u + pix[3 * i + 2] * 3737u) >> 15);
  }
  *data = out; *w = W; *h = H;
  return DISB200_OK;
}

// ------------------------------------------------------------------------------------ writer pool
struct DisWriter {
  struct Job { std::string path; std::vector<float> data; int w, h, nc, kind; };
  std::vector<std::thread> threads;
  std::deque<Job> queue;
  std::mutex mu;
  std::condition_variable cv_work, cv_space, cv_idle;
  size_t queued_bytes = 0, max_bytes = 0;
  int running = 0, first_error = DISB200_OK, failed = 0;
  bool stop = false;

  void worker() {
    for (;;) {
      Job job;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv_work.wait(lk, [&] { return stop || !queue.empty(); });
        if (queue.empty()) return;
        job = std::move(queue.front());
        queue.pop_front();
        ++running;
      }
      const int rc = job.kind == 0 ? disb200_save_flo(job.path.c_str(), job.data.data(), job.w, job.h, job.nc)
                                   : disb200_save_pfm(job.path.c_str(), job.data.data(), job.w, job.h);
      {
        std::lock_guard<std::mutex> lk(mu);
        queued_bytes -= job.data.size() * sizeof(float);
        --running;
        if (rc) { ++failed; if (!first_error) first_error = rc; }
      }
      cv_space.notify_all();
      cv_idle.notify_all();
    }
  }
};

// Asynchronous .flo / .pfm writers (SaveFlowFile / SavePFMFile, run_dense.cpp:17-82, off the caller's thread).
// submit copies the array and returns; it blocks only while more than max_queued_bytes are waiting.
int disb200_writer_create(int threads, size_t max_queued_bytes, DisWriter** out) {
  if (!out || threads < 1 || threads > 256) return DISB200_EINVAL;
  DisWriter* w = new DisWriter();
  w->max_bytes = max_queued_bytes ? max_queued_bytes : ((size_t)1 << 30);
  for (int i = 0; i < threads; ++i) w->threads.emplace_back([w] { w->worker(); });
  *out = w;
  return DISB200_OK;
}

static int writer_submit(DisWriter* w, int kind, const char* path, const float* data, int wd, int h, int nc) {
  if (!w || !path || !data || wd <= 0 || h <= 0 || (nc != 1 && nc != 2 && nc != 4)) return DISB200_EINVAL;
  DisWriter::Job job;
  job.path = path; job.w = wd; job.h = h; job.nc = nc; job.kind = kind;
  job.data.assign(data, data + (size_t)wd * h * nc);
  const size_t bytes = job.data.size() * sizeof(float);
  {
    std::unique_lock<std::mutex> lk(w->mu);
    w->cv_space.wait(lk, [&] { return w->queued_bytes == 0 || w->queued_bytes + bytes <= w->max_bytes; });
    w->queued_bytes += bytes;
    w->queue.push_back(std::move(job));
  }
  w->cv_work.notify_one();
  return DISB200_OK;
}
int disb200_writer_submit_flo(DisWriter* w, const char* path, const float* flow, int wd, int h, int nc) {
  return writer_submit(w, 0, path, flow, wd, h, nc);
}
int disb200_writer_submit_pfm(DisWriter* w, const char* path, const float* disp, int wd, int h) {
  return writer_submit(w, 1, path, disp, wd, h, 1);
}
// Waits until every submitted file is on disk; returns the first error any of them hit (and clears it).
int disb200_writer_wait(DisWriter* w) {
  if (!w) return DISB200_EINVAL;
  std::unique_lock<std::mutex> lk(w->mu);
  w->cv_idle.wait(lk, [&] { return w->queue.empty() && w->running == 0; });
  const int rc = w->first_error;
  w->first_error = DISB200_OK;
  w->failed = 0;
  return rc;
}
void disb200_writer_destroy(DisWriter* w) {
  if (!w) return;
  {
    std::lock_guard<std::mutex> lk(w->mu);
    w->stop = true;
  }
  w->cv_work.notify_all();
  for (std::thread& t : w->threads) t.join();
  delete w;
}

}  // extern "C"
