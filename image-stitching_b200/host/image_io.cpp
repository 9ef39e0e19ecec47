// image_io.cpp -- see image_io.hpp.
#include "image_io.hpp"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cctype>
#include <cstdlib>
#include <cstring>

#ifdef BS_WITH_NVJPEG
#include <cuda_runtime.h>
#include <nvjpeg.h>
#endif

namespace bsio {

namespace {
void *(*g_alloc)(size_t) = nullptr;
void (*g_release)(void *) = nullptr;
}  // namespace

void set_pixel_allocator(void *(*alloc)(size_t), void (*release)(void *)) {
  g_alloc = alloc;
  g_release = release;
}
void *pixel_alloc(size_t bytes) { return g_alloc ? g_alloc(bytes) : malloc(bytes); }
void pixel_free(void *p) {
  if (g_release) g_release(p);
  else free(p);
}

namespace {

void set_err(std::string *err, const std::string &m) {
  if (err) *err = m;
}

std::string lower_ext(const std::string &p) {
  auto d = p.rfind('.');
  std::string e = d == std::string::npos ? "" : p.substr(d + 1);
  std::transform(e.begin(), e.end(), e.begin(), [](unsigned char c) { return (char)tolower(c); });
  return e;
}

// reads one whitespace/comment-delimited header integer of a PNM file
bool pnm_int(FILE *f, int &v) {
  int c = fgetc(f);
  for (;;) {
    while (c != EOF && isspace(c)) c = fgetc(f);
    if (c == '#') {
      while (c != EOF && c != '\n') c = fgetc(f);
      continue;
    }
    break;
  }
  if (c == EOF || !isdigit(c)) return false;
  long n = 0;
  while (c != EOF && isdigit(c)) {
    n = n * 10 + (c - '0');
    if (n > 1000000) return false;
    c = fgetc(f);
  }
  v = (int)n;   // the single whitespace after the last header field has been consumed
  return true;
}

bool read_ppm_frame(FILE *f, Image &out, std::string *err) {
  int c0 = fgetc(f);
  while (c0 != EOF && isspace(c0)) c0 = fgetc(f);
  if (c0 == EOF) {
    set_err(err, "end of stream");
    return false;
  }
  int c1 = fgetc(f);
  if (c0 != 'P' || c1 != '6') {
    set_err(err, "not a binary PPM (P6)");
    return false;
  }
  int w, h, maxv;
  if (!pnm_int(f, w) || !pnm_int(f, h) || !pnm_int(f, maxv) || maxv != 255 || w < 1 || h < 1) {
    set_err(err, "bad PPM header");
    return false;
  }
  out.w = w;
  out.h = h;
  out.bgr.resize((size_t)w * h * 3);
  if (fread(out.bgr.data(), 1, out.bgr.size(), f) != out.bgr.size()) {
    set_err(err, "truncated PPM");
    return false;
  }
  for (size_t i = 0; i + 2 < out.bgr.size(); i += 3) std::swap(out.bgr[i], out.bgr[i + 2]);   // RGB -> BGR
  return true;
}

#ifdef BS_WITH_NVJPEG
bool read_file(const std::string &path, std::vector<unsigned char> &data) {
  FILE *f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  data.resize(n > 0 ? (size_t)n : 0);
  bool ok = n > 0 && fread(data.data(), 1, data.size(), f) == data.size();
  fclose(f);
  return ok;
}

bool read_jpeg(const std::string &path, Image &out, std::string *err) {
  std::vector<unsigned char> data;
  if (!read_file(path, data)) {
    set_err(err, "cannot read " + path);
    return false;
  }
  nvjpegHandle_t h = nullptr;
  nvjpegJpegState_t st = nullptr;
  bool ok = false;
  unsigned char *dbuf = nullptr;
  do {
    if (nvjpegCreateSimple(&h) != NVJPEG_STATUS_SUCCESS) { set_err(err, "nvjpegCreateSimple failed (no GPU?)"); break; }
    if (nvjpegJpegStateCreate(h, &st) != NVJPEG_STATUS_SUCCESS) { set_err(err, "nvjpegJpegStateCreate failed"); break; }
    int nc = 0, ws[NVJPEG_MAX_COMPONENT], hs[NVJPEG_MAX_COMPONENT];
    nvjpegChromaSubsampling_t ss;
    if (nvjpegGetImageInfo(h, data.data(), data.size(), &nc, &ss, ws, hs) != NVJPEG_STATUS_SUCCESS) { set_err(err, "not a JPEG: " + path); break; }
    out.w = ws[0];
    out.h = hs[0];
    size_t bytes = (size_t)out.w * out.h * 3;
    if (cudaMalloc(&dbuf, bytes) != cudaSuccess) { set_err(err, "cudaMalloc failed"); break; }
    nvjpegImage_t img;
    memset(&img, 0, sizeof img);
    img.channel[0] = dbuf;
    img.pitch[0] = (size_t)out.w * 3;
    if (nvjpegDecode(h, st, data.data(), data.size(), NVJPEG_OUTPUT_BGRI, &img, 0) != NVJPEG_STATUS_SUCCESS) { set_err(err, "nvjpegDecode failed: " + path); break; }
    out.bgr.resize(bytes);
    if (cudaMemcpy(out.bgr.data(), dbuf, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) { set_err(err, "cudaMemcpy failed"); break; }
    ok = true;
  } while (false);
  if (dbuf) cudaFree(dbuf);
  if (st) nvjpegJpegStateDestroy(st);
  if (h) nvjpegDestroy(h);
  return ok;
}

bool write_jpeg(const std::string &path, const Image &im, std::string *err) {
  nvjpegHandle_t h = nullptr;
  nvjpegEncoderState_t es = nullptr;
  nvjpegEncoderParams_t ep = nullptr;
  unsigned char *dbuf = nullptr;
  bool ok = false;
  do {
    if (nvjpegCreateSimple(&h) != NVJPEG_STATUS_SUCCESS) { set_err(err, "nvjpegCreateSimple failed (no GPU?)"); break; }
    if (nvjpegEncoderStateCreate(h, &es, 0) != NVJPEG_STATUS_SUCCESS) { set_err(err, "nvjpegEncoderStateCreate failed"); break; }
    if (nvjpegEncoderParamsCreate(h, &ep, 0) != NVJPEG_STATUS_SUCCESS) { set_err(err, "nvjpegEncoderParamsCreate failed"); break; }
    nvjpegEncoderParamsSetQuality(ep, 95, 0);                       // cv::imwrite's default JPEG quality
    nvjpegEncoderParamsSetSamplingFactors(ep, NVJPEG_CSS_420, 0);   // and its default 4:2:0 subsampling
    size_t bytes = im.bgr.size();
    if (cudaMalloc(&dbuf, bytes) != cudaSuccess) { set_err(err, "cudaMalloc failed"); break; }
    if (cudaMemcpy(dbuf, im.bgr.data(), bytes, cudaMemcpyHostToDevice) != cudaSuccess) { set_err(err, "cudaMemcpy failed"); break; }
    nvjpegImage_t img;
    memset(&img, 0, sizeof img);
    img.channel[0] = dbuf;
    img.pitch[0] = im.pitch();
    if (nvjpegEncodeImage(h, es, ep, &img, NVJPEG_INPUT_BGRI, im.w, im.h, 0) != NVJPEG_STATUS_SUCCESS) { set_err(err, "nvjpegEncodeImage failed"); break; }
    size_t len = 0;
    if (nvjpegEncodeRetrieveBitstream(h, es, nullptr, &len, 0) != NVJPEG_STATUS_SUCCESS) { set_err(err, "nvjpeg bitstream size failed"); break; }
    cudaStreamSynchronize(0);
    std::vector<unsigned char> bits(len);
    if (nvjpegEncodeRetrieveBitstream(h, es, bits.data(), &len, 0) != NVJPEG_STATUS_SUCCESS) { set_err(err, "nvjpeg bitstream failed"); break; }
    cudaStreamSynchronize(0);
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) { set_err(err, "cannot write " + path); break; }
    ok = fwrite(bits.data(), 1, len, f) == len;
    fclose(f);
  } while (false);
  if (dbuf) cudaFree(dbuf);
  if (ep) nvjpegEncoderParamsDestroy(ep);
  if (es) nvjpegEncoderStateDestroy(es);
  if (h) nvjpegDestroy(h);
  return ok;
}
#endif

}  // namespace

bool is_jpeg_path(const std::string &path) {
  const std::string e = lower_ext(path);
  return e == "jpg" || e == "jpeg";
}

bool read_bytes(const std::string &path, std::vector<uint8_t> &data) {
  FILE *f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  const long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  data.resize(n > 0 ? (size_t)n : 0);
  const bool ok = n > 0 && fread(data.data(), 1, data.size(), f) == data.size();
  fclose(f);
  return ok;
}

bool write_bytes(const std::string &path, const uint8_t *data, size_t n) {
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) return false;
  const bool ok = fwrite(data, 1, n, f) == n;
  fclose(f);
  return ok;
}

// Walk the marker segments up to the first frame header (baseline, extended or progressive DCT).
bool jpeg_size(const uint8_t *d, size_t n, int *w, int *h) {
  if (n < 4 || d[0] != 0xFF || d[1] != 0xD8) return false;
  size_t i = 2;
  while (i + 4 <= n) {
    if (d[i] != 0xFF) return false;
    const uint8_t m = d[i + 1];
    if (m == 0xFF) { ++i; continue; }                            // fill byte
    if (m == 0x01 || (m >= 0xD0 && m <= 0xD8)) { i += 2; continue; }   // markers without a length
    const size_t len = ((size_t)d[i + 2] << 8) | d[i + 3];
    if (len < 2 || i + 2 + len > n) return false;
    if (m >= 0xC0 && m <= 0xC2) {
      if (len < 7) return false;
      *h = (d[i + 5] << 8) | d[i + 6];
      *w = (d[i + 7] << 8) | d[i + 8];
      return *w > 0 && *h > 0;
    }
    if (m == 0xDA) return false;                                 // scan data before any frame header
    i += 2 + len;
  }
  return false;
}

bool is_mjpeg_path(const std::string &path) {
  const std::string e = lower_ext(path);
  return e == "mjpeg" || e == "mjpg";
}

JpegStream::~JpegStream() {
  if (base_) munmap(const_cast<uint8_t *>(base_), size_);
}

bool JpegStream::open(const std::string &path) {
  const int fd = ::open(path.c_str(), O_RDONLY);
  if (fd < 0) return false;
  struct stat st;
  if (fstat(fd, &st) != 0 || st.st_size <= 0) {
    ::close(fd);
    return false;
  }
  void *p = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
  ::close(fd);
  if (p == MAP_FAILED) return false;
  base_ = static_cast<const uint8_t *>(p);
  size_ = (size_t)st.st_size;
  pos_ = 0;
  return true;
}

// One frame = SOI ... EOI.  Marker segments are skipped by their length; inside entropy-coded data a 0xFF is
// followed by 0x00 (stuffing) or RSTn, anything else is the next marker.
bool JpegStream::next(const uint8_t **frame, size_t *bytes) {
  const uint8_t *d = base_;
  size_t i = pos_;
  while (i + 1 < size_ && !(d[i] == 0xFF && d[i + 1] == 0xD8)) ++i;      // padding between frames
  if (i + 4 > size_) return false;
  const size_t start = i;
  i += 2;
  while (i + 1 < size_) {
    if (d[i] != 0xFF) return false;
    const uint8_t m = d[i + 1];
    if (m == 0xFF) { ++i; continue; }
    if (m == 0xD9) {
      *frame = d + start;
      *bytes = i + 2 - start;
      pos_ = i + 2;
      return true;
    }
    if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { i += 2; continue; }
    if (i + 3 >= size_) return false;
    const size_t len = ((size_t)d[i + 2] << 8) | d[i + 3];
    if (len < 2 || i + 2 + len > size_) return false;
    i += 2 + len;
    if (m == 0xDA) {                                            // entropy-coded data follows the scan header
      while (i + 1 < size_ && !(d[i] == 0xFF && d[i + 1] != 0x00 && !(d[i + 1] >= 0xD0 && d[i + 1] <= 0xD7))) ++i;
    }
  }
  return false;
}

bool have_jpeg() {
#ifdef BS_WITH_NVJPEG
  return true;
#else
  return false;
#endif
}

bool read_image(const std::string &path, Image &out, std::string *err) {
  const std::string e = lower_ext(path);
  if (e == "jpg" || e == "jpeg") {
#ifdef BS_WITH_NVJPEG
    return read_jpeg(path, out, err);
#else
    set_err(err, "JPEG support not built in");
    return false;
#endif
  }
  FILE *f = fopen(path.c_str(), "rb");
  if (!f) {
    set_err(err, "cannot open " + path);
    return false;
  }
  bool ok = read_ppm_frame(f, out, err);
  fclose(f);
  return ok;
}

bool write_image(const std::string &path, const Image &img, std::string *err) {
  const std::string e = lower_ext(path);
  if (e == "jpg" || e == "jpeg") {
#ifdef BS_WITH_NVJPEG
    return write_jpeg(path, img, err);
#else
    set_err(err, "JPEG support not built in");
    return false;
#endif
  }
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) {
    set_err(err, "cannot write " + path);
    return false;
  }
  fprintf(f, "P6\n%d %d\n255\n", img.w, img.h);
  std::vector<uint8_t> row((size_t)img.w * 3);
  bool ok = true;
  for (int y = 0; y < img.h && ok; ++y) {
    const uint8_t *s = img.bgr.data() + (size_t)y * img.pitch();
    for (int x = 0; x < img.w; ++x) {
      row[3 * x + 0] = s[3 * x + 2];
      row[3 * x + 1] = s[3 * x + 1];
      row[3 * x + 2] = s[3 * x + 0];
    }
    ok = fwrite(row.data(), 1, row.size(), f) == row.size();
  }
  fclose(f);
  return ok;
}

FrameStream::~FrameStream() {
  if (f_) fclose(f_);
}

bool FrameStream::open(const std::string &path) {
  if (f_) fclose(f_);
  f_ = fopen(path.c_str(), "rb");
  return f_ != nullptr;
}

bool FrameStream::read(Image &out) {
  if (!f_) return false;
  return read_ppm_frame(f_, out, nullptr);
}

}  // namespace bsio
