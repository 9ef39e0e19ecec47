// image_io.hpp -- frame input/output of the stitch host.
// The reference reads with cv::imread / cv::VideoCapture and writes with cv::imwrite
// (/root/reference/tools/stitch/main.cpp:107,114,125,127-147); OpenCV's C++ SDK is not available
// to this build, so the host reads/writes binary PPM (P6) itself and uses nvJPEG for JPEG.
// All images are packed BGR, 8 bits per channel, like a continuous CV_8UC3 cv::Mat.
#ifndef B200_STITCH_HOST_IMAGE_IO_HPP
#define B200_STITCH_HOST_IMAGE_IO_HPP

#include <cstdint>
#include <cstdio>
#include <new>
#include <string>
#include <vector>

namespace bsio {

// Pixel storage goes through a pair of hooks so that the stitch tool can keep its frames and panoramas in
// pinned host memory (bs_host_alloc: the engine then copies by DMA straight from / to them); the default is malloc.
void set_pixel_allocator(void *(*alloc)(size_t), void (*release)(void *));
void *pixel_alloc(size_t bytes);
void pixel_free(void *p);

template <class T>
struct PixelAllocator {
  using value_type = T;
  PixelAllocator() = default;
  template <class U> PixelAllocator(const PixelAllocator<U> &) {}
  T *allocate(size_t n) {
    void *p = pixel_alloc(n * sizeof(T));
    if (!p) throw std::bad_alloc();
    return static_cast<T *>(p);
  }
  void deallocate(T *p, size_t) { pixel_free(p); }
  template <class U> bool operator==(const PixelAllocator<U> &) const { return true; }
  template <class U> bool operator!=(const PixelAllocator<U> &) const { return false; }
};
using Pixels = std::vector<uint8_t, PixelAllocator<uint8_t>>;

struct Image {
  int w = 0, h = 0;
  Pixels bgr;                 // h rows of w*3 bytes
  bool empty() const { return w == 0 || h == 0; }
  size_t pitch() const { return (size_t)w * 3; }
};

// PPM (.ppm/.pnm) and, when built with nvJPEG, JPEG (.jpg/.jpeg).  Returns false on failure.
bool read_image(const std::string &path, Image &out, std::string *err = nullptr);
bool write_image(const std::string &path, const Image &img, std::string *err = nullptr);

// A "video": a stream of concatenated binary PPM frames (what
// `ffmpeg -i in.mp4 -f image2pipe -vcodec ppm out.ppms` produces).  Stands in for cv::VideoCapture.
class FrameStream {
 public:
  ~FrameStream();
  bool open(const std::string &path);
  bool isOpened() const { return f_ != nullptr; }
  bool read(Image &out);   // false at end of stream
 private:
  FILE *f_ = nullptr;
};

bool have_jpeg();

// For the compressed path of the stitch tool (bs_submit_jpeg takes the bitstreams as they are):
bool is_jpeg_path(const std::string &path);                               // .jpg / .jpeg, any case
bool read_bytes(const std::string &path, std::vector<uint8_t> &data);     // the whole file
bool jpeg_size(const uint8_t *data, size_t n, int *w, int *h);            // from the frame header (SOF0..SOF2), no decoder involved
bool write_bytes(const std::string &path, const uint8_t *data, size_t n);

// A compressed "video": concatenated JPEG frames (Motion-JPEG elementary stream, `ffmpeg -i in.mp4 -c:v mjpeg -f mjpeg
// out.mjpeg`), mapped read-only; next() hands out the frames' bitstreams in place, for bs_submit_jpeg.
bool is_mjpeg_path(const std::string &path);                              // .mjpeg / .mjpg, any case
class JpegStream {
 public:
  ~JpegStream();
  bool open(const std::string &path);
  bool isOpened() const { return base_ != nullptr; }
  bool next(const uint8_t **frame, size_t *bytes);   // false at end of stream (or at bytes that are not a JPEG frame)
 private:
  const uint8_t *base_ = nullptr;
  size_t size_ = 0, pos_ = 0;
};

}  // namespace bsio
#endif
