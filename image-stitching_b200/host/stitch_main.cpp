// stitch_main.cpp -- drop-in for the reference's `stitch` tool
// (/root/reference/tools/stitch/main.cpp): same command line, same stitch.conf.xml, same exit
// codes and messages; the per-frame OpenCV calls (:116-117, :148-151) are replaced by the CUDA
// engine behind include/b200_stitch.h.
//
//   stitch [--s-conf=stitch.conf.xml] [-i] [-v=N]        (reference flags)
//          [--out=PATH] [--frames=N] [--device=N] [--devices=0-7] [--legacy-border]   (extensions)
//
// Differences a user can observe, all forced by what this build can link against:
//   * images are read as JPEG through nvJPEG or as binary PPM (no libjpeg / OpenCV here);
//   * a "video" input is a stream of concatenated PPM frames; at end of input the tool exits 0
//     (the reference loops forever and dies on an OpenCV assertion at EOF, main.cpp:143-155);
//   * uncovered canvas pixels are 0 (the reference leaves them uninitialised, main.cpp:108,144);
//   * -i opens no window (no display on a GPU node): displayResult is a headless no-op;
//   * image mode uploads every photo once and downloads the canvas once (one fused launch for all of
//     them); the reference's one line of milliseconds per image is kept, each line being that image's
//     share of the fused time;
//   * --devices=... runs video mode frame-parallel over several GPUs (frame f on GPU f mod G, panoramas
//     in frame order); frames and panoramas live in pinned host memory.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "../../include/b200_stitch.h"
#include "file_storage.hpp"
#include "image_io.hpp"
#include "opts.hpp"

extern command_line_opts opts;

namespace {

// lib/Calibrate/utils.cpp:208-218 shows a window when opts.interactive; there is no display here.
void displayResult(const std::string &, const bsio::Image &, bool = false) {}

bool read_matrix_seq(const bsfs::Node &n, std::vector<bsfs::Matrix> &out) {
  for (size_t i = 0; i < n.size() && i < out.size(); ++i)
    if (!n[i].asMatrix(out[i])) out[i] = bsfs::Matrix();
  return true;
}

int fail_bs(const char *what, bs_handle *h) {
  std::cout << what << ": " << bs_last_error(h) << std::endl;
  return 6;
}

}  // namespace

int main(int argc, char *argv[]) {
  if (!parse_command_line_opts(argc, argv)) {
    std::cout << "Usage: " << argv[0] << " -c=/path/to/conf.xml";
    return 1;
  }

  bsfs::FileStorage fs;
  fs.open(opts.stitch_config);
  if (!fs.isOpened()) {
    std::cout << "Failed to open configuration file " << opts.stitch_config << std::endl;
    return 2;
  }

  opts.video = fs["video"].asInt(opts.video ? 1 : 0) != 0;
  const bsfs::Node &fNames = fs["file_paths"];
  if (!fNames.isSeq()) {
    std::cout << "file_paths is not a sequence!" << std::endl;
    return 4;
  }
  opts.file_paths.resize(fNames.size());
  for (size_t i = 0; i < fNames.size(); ++i) opts.file_paths[i] = fNames[i].asString();
  if (opts.file_paths.empty()) {
    std::cout << "Usage: " << argv[0] << " /path/to/img1.jpg /path/to/img2.jpg";
    return 3;
  }

  int result_w = 0, result_h = 0;
  const bsfs::Node &fSize = fs["result_size"];
  if (fSize.isSeq() && fSize.size() >= 2) {
    result_w = fSize[0].asInt();
    result_h = fSize[1].asInt();
  }

  const size_t n = opts.file_paths.size();
  std::vector<bsfs::Matrix> H(n), cameraMatrix(n), distCoeffs(n);
  const bsfs::Node &fH = fs["H"];
  if (!fH.isSeq()) {
    std::cout << "H is not a sequence!" << std::endl;
    return 4;
  }
  read_matrix_seq(fH, H);
  const bsfs::Node &fK = fs["cameraMatrix"];
  if (!fK.isSeq()) {
    std::cout << "cameraMatrix is not a sequence!" << std::endl;
    return 4;
  }
  read_matrix_seq(fK, cameraMatrix);
  const bsfs::Node &fD = fs["distCoeffs"];
  if (!fD.isSeq()) {
    std::cout << "distCoeffs is not a sequence!" << std::endl;
    return 4;
  }
  read_matrix_seq(fD, distCoeffs);

  for (size_t i = 0; i < n; ++i)
    if (H[i].rows != 3 || H[i].cols != 3) {
      std::cout << "H[" << i << "] is not a 3x3 matrix" << std::endl;
      return 4;
    }
  if (result_w < 1 || result_h < 1) {
    std::cout << "result_size is missing or empty" << std::endl;
    return 4;
  }
  const unsigned flags = opts.legacy_border ? BS_FLAG_LEGACY_BORDER : 0u;

  // frames and panoramas in pinned memory: the engine's copies then go by DMA straight from / to them
  if (bs_device_count() > 0) bsio::set_pixel_allocator(bs_host_alloc, bs_host_free);

  bool all_jpeg = bsio::have_jpeg() && !opts.video;
  for (size_t i = 0; i < n && all_jpeg; ++i) all_jpeg = bsio::is_jpeg_path(opts.file_paths[i]);
  if (all_jpeg) {
    // JPEG photos (imread at main.cpp:107,114, imwrite at :125): the bitstreams go to the GPU as they are and are
    // decoded there into the handle's frames; when the result is a JPEG too, it is encoded from the device canvas
    // and only its bitstream comes back (bs_submit_jpeg).
    std::vector<std::vector<uint8_t>> bits(n);
    std::vector<int> sw(n), sh(n);
    std::vector<double> Hall(9 * n);
    for (size_t i = 0; i < n; ++i) {
      if (!bsio::read_bytes(opts.file_paths[i], bits[i]) || !bsio::jpeg_size(bits[i].data(), bits[i].size(), &sw[i], &sh[i])) {
        std::cout << "Failed to read " << opts.file_paths[i] << ": not a JPEG file" << std::endl;
        return 5;
      }
      memcpy(&Hall[9 * i], H[i].data.data(), 9 * sizeof(double));
    }
    const std::string out = opts.output.empty() ? "final.jpg" : opts.output;
    const bool jpeg_out = bsio::is_jpeg_path(out);
    bs_handle *h = nullptr;
    if (bs_create(opts.device, (int)n, sw.data(), sh.data(), result_w, result_h, Hall.data(), flags, &h) != BS_OK)
      return fail_bs("bs_create failed", nullptr);
    bsio::Image result;
    std::vector<uint8_t> jbuf;
    bs_jpeg_out jo = {nullptr, 0, 0, 0};       // quality 0 = 95, cv::imwrite's default
    if (jpeg_out) {
      jbuf.resize((size_t)result_w * result_h * 3 / 2 + (1u << 16));
      jo.data = jbuf.data();
      jo.capacity = jbuf.size();
    } else {
      result.w = result_w;
      result.h = result_h;
      result.bgr.assign((size_t)result_w * result_h * 3, 0);
    }
    std::vector<const uint8_t *> ptrs(n);
    std::vector<size_t> sizes(n);
    for (size_t i = 0; i < n; ++i) {
      ptrs[i] = bits[i].data();
      sizes[i] = bits[i].size();
    }
    auto start = std::chrono::steady_clock::now();
    int64_t ticket = -1;
    int rc = bs_submit_jpeg(h, ptrs.data(), sizes.data(), jpeg_out ? nullptr : result.bgr.data(), jpeg_out ? 0 : result.pitch(),
                            jpeg_out ? &jo : nullptr, &ticket);
    if (rc == BS_OK) rc = bs_wait(h, ticket);
    auto end = std::chrono::steady_clock::now();
    if (rc != BS_OK) {
      rc = fail_bs("warpPerspective failed", h);
      bs_destroy(h);
      return rc;
    }
    bs_destroy(h);
    for (size_t i = 0; i < n; ++i) std::cout << std::chrono::duration<double, std::milli>(end - start).count() / (double)n << std::endl;
    std::string err;
    const bool ok = jpeg_out ? bsio::write_bytes(out, jo.data, jo.size) : bsio::write_image(out, result, &err);
    if (!ok) {
      std::cout << "Failed to write " << out << ": " << err << std::endl;
      return 6;
    }
    return 0;
  }

  if (!opts.video) {
    // One handle for all images: every photo is uploaded once, one launch warps and composites them in
    // the reference's order (later images overwrite earlier ones, main.cpp:113-117), the canvas comes
    // back once.
    std::vector<bsio::Image> images(n);
    std::vector<int> sw(n), sh(n);
    std::vector<double> Hall(9 * n);
    for (size_t i = 0; i < n; ++i) {
      std::string err;
      if (!bsio::read_image(opts.file_paths[i], images[i], &err)) {
        std::cout << "Failed to read " << opts.file_paths[i] << ": " << err << std::endl;
        return 5;
      }
      sw[i] = images[i].w;
      sh[i] = images[i].h;
      memcpy(&Hall[9 * i], H[i].data.data(), 9 * sizeof(double));
    }
    bsio::Image result;
    result.w = result_w;
    result.h = result_h;
    result.bgr.assign((size_t)result_w * result_h * 3, 0);
    bs_handle *h = nullptr;
    if (bs_create(opts.device, (int)n, sw.data(), sh.data(), result_w, result_h, Hall.data(), flags, &h) != BS_OK)
      return fail_bs("bs_create failed", nullptr);
    std::vector<const uint8_t *> ptrs(n);
    std::vector<size_t> pitches(n);
    for (size_t i = 0; i < n; ++i) {
      ptrs[i] = images[i].bgr.data();
      pitches[i] = images[i].pitch();
    }
    auto start = std::chrono::steady_clock::now();
    int rc = bs_stitch(h, ptrs.data(), pitches.data(), result.bgr.data(), result.pitch());
    auto end = std::chrono::steady_clock::now();
    if (rc != BS_OK) {
      rc = fail_bs("warpPerspective failed", h);
      bs_destroy(h);
      return rc;
    }
    bs_destroy(h);
    for (size_t i = 0; i < n; ++i) {
      std::cout << std::chrono::duration<double, std::milli>(end - start).count() / (double)n << std::endl;
      displayResult("temp", result, true);
    }
    displayResult("Final", result, true);
    std::string out = opts.output.empty() ? (bsio::have_jpeg() ? "final.jpg" : "final.ppm") : opts.output;
    std::string err;
    if (!bsio::write_image(out, result, &err)) {
      std::cout << "Failed to write " << out << ": " << err << std::endl;
      return 6;
    }
    return 0;
  }

  // ---- video mode, compressed streams -------------------------------------------------------------
  bool all_mjpeg = bsio::have_jpeg();
  for (size_t i = 0; i < n && all_mjpeg; ++i) all_mjpeg = bsio::is_mjpeg_path(opts.file_paths[i]);
  if (all_mjpeg) {
    // Motion-JPEG streams (what cv::VideoCapture >> frame decodes on the host at main.cpp:147): every frame's
    // bitstream goes to the GPU as it is (bs_submit_jpeg), up to bs_pipeline_depth frames in flight; with a .jpg
    // --out pattern the panoramas are encoded on the GPU as well.  One GPU (--device); the streams are mapped.
    std::vector<bsio::JpegStream> streams(n);
    std::vector<int> sw(n), sh(n);
    std::vector<double> Hall(9 * n);
    for (size_t i = 0; i < n; ++i) {
      const uint8_t *f0 = nullptr;
      size_t b0 = 0;
      // the reference reads and drops the first frame of every stream (main.cpp:136-141)
      if (!streams[i].open(opts.file_paths[i]) || !streams[i].next(&f0, &b0) || !bsio::jpeg_size(f0, b0, &sw[i], &sh[i])) {
        std::cout << "Failed to open file " << opts.file_paths[i] << "!" << std::endl;
        return 5;
      }
      memcpy(&Hall[9 * i], H[i].data.data(), 9 * sizeof(double));
    }
    bs_handle *h = nullptr;
    if (bs_create(opts.device, (int)n, sw.data(), sh.data(), result_w, result_h, Hall.data(), flags, &h) != BS_OK)
      return fail_bs("bs_create failed", nullptr);
    for (size_t i = 0; i < n; ++i) {
      const bsfs::Matrix &K = cameraMatrix[i], &d = distCoeffs[i];
      if (K.empty()) continue;
      if (K.rows != 3 || K.cols != 3 || bs_set_undistort(h, (int)i, K.data.data(), d.empty() ? nullptr : d.data.data(), (int)d.data.size()) != BS_OK) {
        std::cout << "bs_set_undistort failed: " << bs_last_error(h) << std::endl;
        bs_destroy(h);
        return 6;
      }
    }
    const int depth = bs_pipeline_depth(h);
    const bool want_out = !opts.output.empty(), jpeg_out = want_out && bsio::is_jpeg_path(opts.output);
    std::vector<bsio::Image> results(depth);
    std::vector<std::vector<uint8_t>> jbufs(depth);
    std::vector<bs_jpeg_out> jouts(depth);
    std::vector<int64_t> tickets(depth, -1);
    for (int k = 0; k < depth; ++k) {
      if (jpeg_out) {
        jbufs[k].resize((size_t)result_w * result_h * 3 / 2 + (1u << 16));
        jouts[k] = bs_jpeg_out{jbufs[k].data(), jbufs[k].size(), 0, 0};
      } else {
        results[k].w = result_w;
        results[k].h = result_h;
        results[k].bgr.assign((size_t)result_w * result_h * 3, 0);
      }
    }
    auto emit = [&](int slot, long index) -> bool {
      if (bs_wait(h, tickets[slot]) != BS_OK) return false;
      tickets[slot] = -1;
      if (!jpeg_out) displayResult("Final", results[slot]);
      if (want_out) {
        char name[4096];
        snprintf(name, sizeof name, opts.output.c_str(), index);   // e.g. --out=pano_%05ld.jpg
        std::string err;
        const bool ok = jpeg_out ? bsio::write_bytes(name, jouts[slot].data, jouts[slot].size) : bsio::write_image(name, results[slot], &err);
        if (!ok) std::cout << "Failed to write " << name << ": " << err << std::endl;
      }
      return true;
    };
    long produced = 0, emitted = 0;
    auto t0 = std::chrono::steady_clock::now();
    bool ok = true;
    while (ok && (opts.max_frames < 0 || produced < opts.max_frames)) {
      const int slot = (int)(produced % depth);
      if (tickets[slot] >= 0) {
        if (!emit(slot, emitted)) { ok = false; break; }
        ++emitted;
      }
      std::vector<const uint8_t *> ptrs(n);
      std::vector<size_t> sizes(n);
      bool eof = false;
      for (size_t i = 0; i < n && !eof; ++i) eof = !streams[i].next(&ptrs[i], &sizes[i]);
      if (eof) break;
      if (bs_submit_jpeg(h, ptrs.data(), sizes.data(), jpeg_out ? nullptr : results[slot].bgr.data(), jpeg_out ? 0 : results[slot].pitch(),
                         jpeg_out ? &jouts[slot] : nullptr, &tickets[slot]) != BS_OK) {
        ok = false;
        break;
      }
      ++produced;
    }
    while (ok && emitted < produced) {
      const int slot = (int)(emitted % depth);
      if (!emit(slot, emitted)) { ok = false; break; }
      ++emitted;
    }
    auto t1 = std::chrono::steady_clock::now();
    int rc = 0;
    if (!ok) {
      std::cout << "stitching failed: " << bs_last_error(h) << std::endl;
      rc = 6;
    } else if (opts.verbosity > 0) {
      std::cout << produced << " frames on 1 GPU(s), " << std::chrono::duration<double, std::milli>(t1 - t0).count() / (produced ? produced : 1)
                << " ms/frame" << std::endl;
    }
    bs_destroy(h);
    return rc;
  }

  // ---- video mode ------------------------------------------------------------------------------
  std::vector<bsio::FrameStream> videos(n);
  for (size_t i = 0; i < n; ++i) {
    videos[i].open(opts.file_paths[i]);
    if (!videos[i].isOpened()) {
      std::cout << "Failed to open file " << opts.file_paths[i] << "!" << std::endl;
      return 5;
    }
  }
  // the reference reads and drops the first frame of every stream (main.cpp:136-141)
  std::vector<bsio::Image> first(n);
  for (size_t i = 0; i < n; ++i)
    if (!videos[i].read(first[i])) {
      std::cout << "Failed to open file " << opts.file_paths[i] << "!" << std::endl;
      return 5;
    }

  std::vector<int> sw(n), sh(n);
  std::vector<double> Hall(9 * n);
  for (size_t i = 0; i < n; ++i) {
    sw[i] = first[i].w;
    sh[i] = first[i].h;
    memcpy(&Hall[9 * i], H[i].data.data(), 9 * sizeof(double));
  }
  // one replica of the calibration per GPU; frame f runs on replica f mod G (a single GPU is a pool of one)
  std::vector<int> devices = opts.devices.empty() ? std::vector<int>{opts.device} : opts.devices;
  bs_pool *pool = nullptr;
  if (bs_pool_create((int)devices.size(), devices.data(), (int)n, sw.data(), sh.data(), result_w, result_h, Hall.data(), flags,
                     &pool) != BS_OK)
    return fail_bs("bs_create failed", nullptr);
  for (size_t i = 0; i < n; ++i) {
    // cv::undistort(frame, undistorted, cameraMatrix[i], distCoeffs[i]) of main.cpp:148-149 runs on the
    // GPU in front of the warp.  Configs written in image mode carry 0x0 matrices: nothing to undo.
    const bsfs::Matrix &K = cameraMatrix[i], &d = distCoeffs[i];
    if (K.empty()) continue;
    if (K.rows != 3 || K.cols != 3 ||
        bs_pool_set_undistort(pool, (int)i, K.data.data(), d.empty() ? nullptr : d.data.data(), (int)d.data.size()) != BS_OK) {
      std::cout << "bs_set_undistort failed: " << bs_pool_last_error(pool) << std::endl;
      bs_pool_destroy(pool);
      return 6;
    }
  }

  const int depth = bs_pool_depth(pool);
  std::vector<std::vector<bsio::Image>> frames(depth, std::vector<bsio::Image>(n));
  std::vector<bsio::Image> results(depth);
  std::vector<int64_t> tickets(depth, -1);
  for (auto &r : results) {
    r.w = result_w;
    r.h = result_h;
    r.bgr.assign((size_t)result_w * result_h * 3, 0);
  }
  auto emit = [&](int slot, long index) -> bool {
    if (bs_pool_wait(pool, tickets[slot]) != BS_OK) return false;
    tickets[slot] = -1;
    displayResult("Final", results[slot]);
    if (!opts.output.empty()) {
      char name[4096];
      snprintf(name, sizeof name, opts.output.c_str(), index);   // e.g. --out=pano_%05ld.ppm
      std::string err;
      if (!bsio::write_image(name, results[slot], &err)) std::cout << "Failed to write " << name << ": " << err << std::endl;
    }
    return true;
  };

  long produced = 0, emitted = 0;
  auto t0 = std::chrono::steady_clock::now();
  bool ok = true;
  while (ok && (opts.max_frames < 0 || produced < opts.max_frames)) {
    const int slot = (int)(produced % depth);
    if (tickets[slot] >= 0) {
      if (!emit(slot, emitted)) { ok = false; break; }
      ++emitted;
    }
    bool eof = false;
    for (size_t i = 0; i < n; ++i) {
      if (!videos[i].read(frames[slot][i]) || frames[slot][i].w != sw[i] || frames[slot][i].h != sh[i]) {
        eof = true;
        break;
      }
    }
    if (eof) break;
    std::vector<const uint8_t *> ptrs(n);
    std::vector<size_t> pitches(n);
    for (size_t i = 0; i < n; ++i) {
      ptrs[i] = frames[slot][i].bgr.data();
      pitches[i] = frames[slot][i].pitch();
    }
    if (bs_pool_submit(pool, ptrs.data(), pitches.data(), results[slot].bgr.data(), results[slot].pitch(), &tickets[slot]) !=
        BS_OK) {
      ok = false;
      break;
    }
    ++produced;
  }
  while (ok && emitted < produced) {
    const int slot = (int)(emitted % depth);
    if (!emit(slot, emitted)) { ok = false; break; }
    ++emitted;
  }
  auto t1 = std::chrono::steady_clock::now();
  int rc = 0;
  if (!ok) {
    std::cout << "stitching failed: " << bs_pool_last_error(pool) << std::endl;
    rc = 6;
  } else if (opts.verbosity > 0) {
    std::cout << produced << " frames on " << devices.size() << " GPU(s), "
              << std::chrono::duration<double, std::milli>(t1 - t0).count() / (produced ? produced : 1) << " ms/frame" << std::endl;
  }
  bs_pool_destroy(pool);
  return rc;
}
