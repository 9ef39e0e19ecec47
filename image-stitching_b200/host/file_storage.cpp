// file_storage.cpp -- see file_storage.hpp.
#include "file_storage.hpp"

#include <cctype>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <fstream>
#include <sstream>

namespace bsfs {

namespace {

struct Parser {
  const std::string &s;
  size_t p = 0;
  std::string err;
  explicit Parser(const std::string &text) : s(text) {}

  bool fail(const std::string &m) {
    if (err.empty()) err = m + " at byte " + std::to_string(p);
    return false;
  }
  void skip_ws() {
    while (p < s.size() && isspace((unsigned char)s[p])) ++p;
  }
  bool starts(const char *lit) const { return s.compare(p, strlen(lit), lit) == 0; }

  // skips <?...?>, <!-- ... -->
  void skip_misc() {
    for (;;) {
      skip_ws();
      if (starts("<?")) {
        size_t e = s.find("?>", p);
        p = e == std::string::npos ? s.size() : e + 2;
      } else if (starts("<!--")) {
        size_t e = s.find("-->", p);
        p = e == std::string::npos ? s.size() : e + 3;
      } else {
        return;
      }
    }
  }

  static std::string unescape(const std::string &t) {
    std::string o;
    for (size_t i = 0; i < t.size(); ++i) {
      if (t[i] == '&') {
        if (t.compare(i, 4, "&lt;") == 0) { o += '<'; i += 3; continue; }
        if (t.compare(i, 4, "&gt;") == 0) { o += '>'; i += 3; continue; }
        if (t.compare(i, 5, "&amp;") == 0) { o += '&'; i += 4; continue; }
        if (t.compare(i, 6, "&quot;") == 0) { o += '"'; i += 5; continue; }
        if (t.compare(i, 6, "&apos;") == 0) { o += '\''; i += 5; continue; }
      }
      o += t[i];
    }
    return o;
  }

  // whitespace-separated tokens; double-quoted tokens may contain spaces
  static std::vector<std::pair<std::string, bool>> tokens(const std::string &t) {
    std::vector<std::pair<std::string, bool>> out;
    size_t i = 0;
    while (i < t.size()) {
      while (i < t.size() && isspace((unsigned char)t[i])) ++i;
      if (i >= t.size()) break;
      if (t[i] == '"') {
        std::string v;
        ++i;
        while (i < t.size() && t[i] != '"') {
          if (t[i] == '\\' && i + 1 < t.size()) ++i;
          v += t[i++];
        }
        ++i;
        out.emplace_back(unescape(v), true);
      } else {
        size_t b = i;
        while (i < t.size() && !isspace((unsigned char)t[i])) ++i;
        out.emplace_back(unescape(t.substr(b, i - b)), false);
      }
    }
    return out;
  }

  static Node scalar(const std::string &tok, bool quoted) {
    Node n;
    n.text = tok;
    if (quoted) {
      n.kind = Node::STRING;
      return n;
    }
    char *end = nullptr;
    const char *c = tok.c_str();
    (void)strtol(c, &end, 10);
    if (end != c && *end == 0) {
      n.kind = Node::INT;
      return n;
    }
    (void)strtod(c, &end);
    if (end != c && *end == 0) {
      n.kind = Node::REAL;
      return n;
    }
    // OpenCV prints "1." and ".5"-style reals; strtod accepts both.  Anything else is a string.
    n.kind = Node::STRING;
    return n;
  }

  // parses one element starting at '<'; returns its tag name and node
  bool element(std::string &name, Node &node) {
    if (p >= s.size() || s[p] != '<') return fail("expected '<'");
    ++p;
    size_t b = p;
    while (p < s.size() && !isspace((unsigned char)s[p]) && s[p] != '>' && s[p] != '/') ++p;
    name = s.substr(b, p - b);
    if (name.empty()) return fail("empty tag name");
    // attributes
    for (;;) {
      skip_ws();
      if (p >= s.size()) return fail("unterminated tag");
      if (s[p] == '>') { ++p; break; }
      if (s[p] == '/' && p + 1 < s.size() && s[p + 1] == '>') { p += 2; node.kind = Node::NONE; return true; }
      size_t ab = p;
      while (p < s.size() && s[p] != '=' && !isspace((unsigned char)s[p]) && s[p] != '>') ++p;
      std::string an = s.substr(ab, p - ab);
      skip_ws();
      if (p < s.size() && s[p] == '=') {
        ++p;
        skip_ws();
        if (p >= s.size() || (s[p] != '"' && s[p] != '\'')) return fail("attribute value not quoted");
        char q = s[p++];
        size_t vb = p;
        while (p < s.size() && s[p] != q) ++p;
        std::string av = s.substr(vb, p - vb);
        ++p;
        if (an == "type_id") node.type_id = av;
      }
    }
    // content: text and/or child elements until </name>
    std::string text;
    std::vector<std::pair<std::string, Node>> children;
    for (;;) {
      if (p >= s.size()) return fail("missing </" + name + ">");
      if (s[p] == '<') {
        if (starts("<!--")) {
          size_t e = s.find("-->", p);
          p = e == std::string::npos ? s.size() : e + 3;
          continue;
        }
        if (starts("</")) {
          size_t e = s.find('>', p);
          if (e == std::string::npos) return fail("unterminated closing tag");
          std::string cn = s.substr(p + 2, e - p - 2);
          while (!cn.empty() && isspace((unsigned char)cn.back())) cn.pop_back();
          if (cn != name) return fail("mismatched </" + cn + "> for <" + name + ">");
          p = e + 1;
          break;
        }
        std::string cname;
        Node child;
        if (!element(cname, child)) return false;
        children.emplace_back(cname, child);
      } else {
        text += s[p++];
      }
    }
    if (!children.empty()) {
      bool all_anon = true;
      for (auto &c : children) all_anon = all_anon && c.first == "_";
      if (all_anon) {
        node.kind = Node::SEQ;
        for (auto &c : children) node.items.push_back(c.second);
      } else {
        node.kind = Node::MAP;
        for (auto &c : children) node.fields[c.first] = c.second;
      }
      return true;
    }
    auto toks = tokens(text);
    if (toks.empty()) {
      node.kind = Node::NONE;
    } else if (toks.size() == 1) {
      std::string tid = node.type_id;
      node = scalar(toks[0].first, toks[0].second);
      node.type_id = tid;
    } else {
      node.kind = Node::SEQ;
      for (auto &t : toks) node.items.push_back(scalar(t.first, t.second));
    }
    return true;
  }
};

}  // namespace

int Node::asInt(int dflt) const {
  if (kind == INT || kind == REAL) return (int)strtod(text.c_str(), nullptr);
  return dflt;
}

double Node::asReal(double dflt) const {
  if (kind == INT || kind == REAL) return strtod(text.c_str(), nullptr);
  return dflt;
}

bool Node::asMatrix(Matrix &m) const {
  if (kind != MAP || type_id != "opencv-matrix") return false;
  auto r = fields.find("rows"), c = fields.find("cols"), dt = fields.find("dt"), d = fields.find("data");
  if (r == fields.end() || c == fields.end() || d == fields.end()) return false;
  m.rows = r->second.asInt();
  m.cols = c->second.asInt();
  m.dt = dt == fields.end() ? "d" : dt->second.text;
  m.data.clear();
  const Node &dn = d->second;
  if (dn.kind == SEQ)
    for (const Node &v : dn.items) m.data.push_back(v.asReal());
  else if (dn.kind == INT || dn.kind == REAL)
    m.data.push_back(dn.asReal());
  return (size_t)m.rows * (size_t)m.cols == m.data.size();
}

bool FileStorage::parse(const std::string &xml) {
  opened_ = false;
  top_.clear();
  Parser ps(xml);
  ps.skip_misc();
  std::string name;
  Node root;
  if (!ps.element(name, root)) {
    error_ = ps.err;
    return false;
  }
  if (name != "opencv_storage") {
    error_ = "root element is <" + name + ">, expected <opencv_storage>";
    return false;
  }
  if (root.kind == Node::MAP) top_ = root.fields;
  opened_ = true;
  return true;
}

bool FileStorage::open(const std::string &path) {
  opened_ = false;
  std::ifstream f(path, std::ios::binary);
  if (!f) {
    error_ = "cannot open " + path;
    return false;
  }
  std::stringstream ss;
  ss << f.rdbuf();
  return parse(ss.str());
}

const Node &FileStorage::operator[](const std::string &key) const {
  auto it = top_.find(key);
  return it == top_.end() ? none_ : it->second;
}

// ---- writer ----------------------------------------------------------------------------------
namespace {

std::string fmt_double(double v) {
  // OpenCV prints 17 significant digits and "1." for integral values
  char buf[64];
  if (v == (double)(long long)v && v > -1e15 && v < 1e15) {
    snprintf(buf, sizeof buf, "%lld.", (long long)v);
    return buf;
  }
  snprintf(buf, sizeof buf, "%.17g", v);
  return buf;
}

void write_matrix_seq(std::ostringstream &o, const char *name, const std::vector<Matrix> &ms) {
  o << "<" << name << ">\n";
  for (size_t k = 0; k < ms.size(); ++k) {
    const Matrix &m = ms[k];
    o << "  <_ type_id=\"opencv-matrix\">\n    <rows>" << m.rows << "</rows>\n    <cols>" << m.cols
      << "</cols>\n    <dt>" << m.dt << "</dt>\n    <data>";
    for (size_t i = 0; i < m.data.size(); ++i) {
      o << (i % 3 == 0 ? "\n      " : " ") << fmt_double(m.data[i]);
    }
    o << "</data></_>" << (k + 1 == ms.size() ? "" : "\n");
  }
  o << "</" << name << ">\n";
}

}  // namespace

std::string to_xml(const Config &c) {
  std::ostringstream o;
  o << "<?xml version=\"1.0\"?>\n<opencv_storage>\n";
  o << "<video>" << (c.video ? 1 : 0) << "</video>\n";
  o << "<file_paths>\n ";
  for (const std::string &p : c.file_paths) {
    bool q = p.find(' ') != std::string::npos;
    o << " " << (q ? "\"" : "") << p << (q ? "\"" : "");
  }
  o << "</file_paths>\n";
  o << "<result_size>\n  " << c.result_w << " " << c.result_h << "</result_size>\n";
  write_matrix_seq(o, "H", c.H);
  write_matrix_seq(o, "cameraMatrix", c.cameraMatrix);
  write_matrix_seq(o, "distCoeffs", c.distCoeffs);
  o << "</opencv_storage>\n";
  return o.str();
}

}  // namespace bsfs
