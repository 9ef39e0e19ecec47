// file_storage.hpp -- reader/writer for the subset of OpenCV's FileStorage XML that
// stitch.conf.xml uses (written at /root/reference/tools/calibrate/main.cpp:497-522, read at
// /root/reference/tools/stitch/main.cpp:26-104; exact text in SURVEY.md Appendix C).
// OpenCV's C++ SDK is not available to this build, so the format is parsed here.
#ifndef B200_STITCH_HOST_FILE_STORAGE_HPP
#define B200_STITCH_HOST_FILE_STORAGE_HPP

#include <map>
#include <string>
#include <vector>

namespace bsfs {

struct Matrix {
  int rows = 0, cols = 0;
  std::string dt = "d";
  std::vector<double> data;
  bool empty() const { return rows == 0 || cols == 0; }
};

// Node kinds follow cv::FileNode::type() as the reference tests it (`type() != FileNode::SEQ`):
// a text node with ONE token is a scalar, with two or more it is a sequence, with none it is NONE;
// a node whose children are <_> elements is a sequence whatever their number.
struct Node {
  enum Kind { NONE = 0, INT = 1, REAL = 2, STRING = 3, SEQ = 4, MAP = 5 };
  Kind kind = NONE;
  std::string text;                  // scalar value
  std::vector<Node> items;           // SEQ
  std::map<std::string, Node> fields;  // MAP
  std::string type_id;               // e.g. "opencv-matrix"

  bool isSeq() const { return kind == SEQ; }
  bool isNone() const { return kind == NONE; }
  size_t size() const { return kind == SEQ ? items.size() : (kind == NONE ? 0 : 1); }
  const Node &operator[](size_t i) const { return items.at(i); }
  int asInt(int dflt = 0) const;
  double asReal(double dflt = 0) const;
  std::string asString() const { return text; }
  bool asMatrix(Matrix &m) const;    // false if this is not an opencv-matrix map
};

class FileStorage {
 public:
  bool open(const std::string &path);          // READ
  bool parse(const std::string &xml_text);
  bool isOpened() const { return opened_; }
  const Node &operator[](const std::string &key) const;
  const std::string &error() const { return error_; }

 private:
  bool opened_ = false;
  std::string error_;
  std::map<std::string, Node> top_;
  Node none_;
};

// Writer for the same subset (used by tests and by tools that produce configs).
struct Config {
  bool video = false;
  std::vector<std::string> file_paths;
  int result_w = 0, result_h = 0;
  std::vector<Matrix> H, cameraMatrix, distCoeffs;
};
std::string to_xml(const Config &c);

}  // namespace bsfs
#endif
