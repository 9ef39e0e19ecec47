// host_selftest.cpp -- exposes the host-side parsers to tests/test_host.py (no GPU needed).
//   host_selftest opts <args...>   -> one line of JSON with the parsed command_line_opts
//   host_selftest conf <file.xml>  -> one line of JSON describing the parsed stitch config
//   host_selftest ppm <in> <out>   -> read an image and write it back
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>

#include "file_storage.hpp"
#include "image_io.hpp"
#include "opts.hpp"

extern command_line_opts opts;

static std::string js(const std::string &s) {
  std::string o = "\"";
  for (char c : s) {
    if (c == '"' || c == '\\') o += '\\';
    o += c;
  }
  return o + "\"";
}

int main(int argc, char **argv) {
  if (argc < 2) return 64;
  const std::string cmd = argv[1];
  if (cmd == "opts") {
    bool ok = parse_command_line_opts(argc - 1, argv + 1);
    std::cout << "{\"valid\": " << (ok ? "true" : "false") << ", \"verbosity\": " << opts.verbosity
              << ", \"interactive\": " << (opts.interactive ? "true" : "false") << ", \"video\": " << (opts.video ? "true" : "false")
              << ", \"delay\": " << opts.delay << ", \"number_of_frames\": " << opts.number_of_frames
              << ", \"board_width\": " << opts.board_width << ", \"board_height\": " << opts.board_height
              << ", \"angle\": " << opts.angle << ", \"mode\": " << (opts.mode == StitchingMode::ChainOfTargets ? "\"chain\"" : "\"extend\"")
              << ", \"calibrate_config\": " << js(opts.calibrate_config) << ", \"stitch_config\": " << js(opts.stitch_config)
              << ", \"output\": " << js(opts.output) << ", \"device\": " << opts.device << ", \"devices\": [";
    for (size_t i = 0; i < opts.devices.size(); ++i) std::cout << (i ? ", " : "") << opts.devices[i];
    std::cout << "], \"file_paths\": [";
    for (size_t i = 0; i < opts.file_paths.size(); ++i) std::cout << (i ? ", " : "") << js(opts.file_paths[i]);
    std::cout << "]}" << std::endl;
    return 0;
  }
  if (cmd == "conf" && argc >= 3) {
    bsfs::FileStorage fs;
    if (!fs.open(argv[2])) {
      std::cout << "{\"opened\": false, \"error\": " << js(fs.error()) << "}" << std::endl;
      return 0;
    }
    std::cout << "{\"opened\": true";
    for (const char *key : {"video", "file_paths", "result_size", "H", "cameraMatrix", "distCoeffs", "missing"}) {
      const bsfs::Node &n = fs[key];
      std::cout << ", " << js(key) << ": {\"kind\": " << (int)n.kind << ", \"size\": " << n.size();
      if (n.kind == bsfs::Node::SEQ) {
        std::cout << ", \"items\": [";
        for (size_t i = 0; i < n.size(); ++i) {
          std::cout << (i ? ", " : "");
          bsfs::Matrix m;
          if (n[i].asMatrix(m)) {
            std::cout << "{\"rows\": " << m.rows << ", \"cols\": " << m.cols << ", \"data\": [";
            for (size_t k = 0; k < m.data.size(); ++k) {
              char buf[64];
              snprintf(buf, sizeof buf, "%.17g", m.data[k]);
              std::cout << (k ? ", " : "") << buf;
            }
            std::cout << "]}";
          } else {
            std::cout << js(n[i].asString());
          }
        }
        std::cout << "]";
      } else if (n.kind != bsfs::Node::NONE) {
        std::cout << ", \"text\": " << js(n.asString());
      }
      std::cout << "}";
    }
    std::cout << "}" << std::endl;
    return 0;
  }
  if (cmd == "ppm" && argc >= 4) {
    bsio::Image im;
    std::string err;
    if (!bsio::read_image(argv[2], im, &err)) {
      std::cout << "read failed: " << err << std::endl;
      return 2;
    }
    if (!bsio::write_image(argv[3], im, &err)) {
      std::cout << "write failed: " << err << std::endl;
      return 3;
    }
    std::cout << im.w << " " << im.h << std::endl;
    return 0;
  }
  if (cmd == "mjpeg" && argc >= 3) {
    // the frames of a Motion-JPEG stream as the stitch tool sees them: "offset bytes width height" per frame
    bsio::JpegStream st;
    if (!st.open(argv[2])) {
      std::cout << "open failed" << std::endl;
      return 2;
    }
    const uint8_t *f = nullptr, *first = nullptr;
    size_t n = 0;
    while (st.next(&f, &n)) {
      if (!first) first = f;
      int w = 0, h = 0;
      const bool ok = bsio::jpeg_size(f, n, &w, &h);
      std::cout << (f - first) << " " << n << " " << (ok ? w : -1) << " " << (ok ? h : -1) << std::endl;
    }
    return 0;
  }
  return 64;
}
