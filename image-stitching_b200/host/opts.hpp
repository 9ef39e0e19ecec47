// opts.hpp -- command-line options of the stitch / calibrate tools.
// Mirrors the reference's interface (same struct name, field names, defaults and entry point:
// /root/reference/include/opts.hpp:8-33) so code written against it compiles unchanged.
#ifndef B200_STITCH_HOST_OPTS_HPP
#define B200_STITCH_HOST_OPTS_HPP

#include <string>
#include <vector>

enum class StitchingMode { OneCommonTarget, ChainOfTargets };

struct command_line_opts {
  unsigned verbosity = 0;
  bool interactive = false;
  std::vector<std::string> file_paths;
  bool video = false;
  int delay = 300;
  int number_of_frames = 30;
  int board_width = 5;
  int board_height = 3;
  int angle = 1;
  StitchingMode mode = StitchingMode::ChainOfTargets;
  std::string calibrate_config;
  std::string stitch_config = "stitch.conf.xml";

  // --- extensions of this build (not in the reference) ---
  std::string output;      // --out=PATH : where the panorama(s) go ("" = final.jpg / final.ppm like the reference)
  int device = 0;          // --device=N
  std::vector<int> devices;    // --devices=0-3 or --devices=0,2,5 : video mode runs frame-parallel on these GPUs (frame f -> f mod G)
  int max_frames = -1;     // --frames=N : stop video mode after N panoramas (-1 = until end of input)
  bool legacy_border = false;  // --legacy-border : OpenCV 2.4/3.x rule on the last source row/column
};

// Fills the global `opts`; returns false on a malformed flag (the tools then exit with 1).
bool parse_command_line_opts(int argc, char *argv[]);

#endif
