// opts.cpp -- flag parser with the reference's observable behaviour
// (/root/reference/lib/CommandLine/opts.cpp:6-106):
//   * flags are recognised by PREFIX ("-v", "-b", "--c-conf", ...), so "-verbose=2" sets verbosity;
//   * a valued flag without '=' makes the whole parse fail;
//   * anything unrecognised is taken as an input path.
// Written table-driven rather than as the reference's if-chain.
#include "opts.hpp"

#include <algorithm>
#include <cstdlib>
#include <functional>

command_line_opts opts;

namespace {

struct ValuedFlag {
  std::vector<const char *> prefixes;
  std::function<bool(const std::string &)> apply;   // receives the text after '='
};

bool starts_with(const std::string &s, const char *p) { return s.rfind(p, 0) == 0; }

}  // namespace

bool parse_command_line_opts(int argc, char *argv[]) {
  const std::vector<ValuedFlag> valued = {
      {{"--verbosity", "-v"}, [](const std::string &v) { opts.verbosity = (unsigned)atoi(v.c_str()); return true; }},
      {{"--board", "-b"},
       [](const std::string &v) {
         const auto x = v.find('x');
         if (x == std::string::npos) return false;
         opts.board_width = atoi(v.substr(0, x).c_str());
         opts.board_height = atoi(v.substr(x + 1).c_str());
         return true;
       }},
      {{"--c-conf"}, [](const std::string &v) { opts.calibrate_config = v; return true; }},
      {{"--s-conf"}, [](const std::string &v) { opts.stitch_config = v; return true; }},
      {{"--delay"}, [](const std::string &v) { opts.delay = atoi(v.c_str()); return true; }},
      {{"--num"}, [](const std::string &v) { opts.number_of_frames = atoi(v.c_str()); return true; }},
      {{"--angle"}, [](const std::string &v) { opts.angle = atoi(v.c_str()); return true; }},
      {{"--mode"},
       [](const std::string &v) {
         if (v == "extend") opts.mode = StitchingMode::OneCommonTarget;
         else if (v == "chain") opts.mode = StitchingMode::ChainOfTargets;
         else return false;
         return true;
       }},
      // extensions
      {{"--out"}, [](const std::string &v) { opts.output = v; return true; }},
      {{"--devices"},      // before "--device": flags match by prefix
       [](const std::string &v) {
         opts.devices.clear();
         size_t pos = 0;
         while (pos <= v.size()) {
           const size_t comma = std::min(v.find(',', pos), v.size());
           const std::string tok = v.substr(pos, comma - pos);
           const size_t dash = tok.find('-');
           if (tok.empty()) return false;
           const int a = atoi(tok.substr(0, dash).c_str()), b = dash == std::string::npos ? a : atoi(tok.substr(dash + 1).c_str());
           if (a < 0 || b < a || b > 1023) return false;
           for (int d = a; d <= b; ++d) opts.devices.push_back(d);
           pos = comma + 1;
         }
         return !opts.devices.empty();
       }},
      {{"--device"}, [](const std::string &v) { opts.device = atoi(v.c_str()); return true; }},
      {{"--frames"}, [](const std::string &v) { opts.max_frames = atoi(v.c_str()); return true; }},
  };

  for (int i = 1; i < argc; ++i) {
    const std::string arg = argv[i];
    if (arg == "--interactive" || arg == "-i") {
      opts.interactive = true;
      continue;
    }
    bool handled = false;
    for (const ValuedFlag &f : valued) {
      bool match = false;
      for (const char *p : f.prefixes) match = match || starts_with(arg, p);
      if (!match) continue;
      const auto eq = arg.find('=');
      if (eq == std::string::npos) return false;
      if (!f.apply(arg.substr(eq + 1))) return false;
      handled = true;
      break;
    }
    if (handled) continue;
    if (arg == "--video") {
      opts.video = true;
    } else if (arg == "--legacy-border") {
      opts.legacy_border = true;
    } else {
      opts.file_paths.push_back(arg);   // assumed to be an input path
    }
  }
  return true;
}
