// yaml_lite.h -- the subset of YAML that hypre_app.yaml-style input files use (/root/reference/etc/hypre_app.yaml and the
// keys read in /root/reference/src/HypreSystem.cpp:16-47,119-326,390-404,1021-1081,1613-1665; schema in SURVEY.md App. B):
// block mappings nested by indentation, plain / quoted scalars, '#' comments, flow sequences [a, b, c] and block sequences
// of scalars ("- item").  The interface mirrors the handful of yaml-cpp calls the reference makes
// (node["key"], if (node["key"]), node.as<T>(), YAML::LoadFile) so the host code reads like the reference's;
// yaml-cpp itself is not available in this image.
#pragma once
#include <cstdlib>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace YAML {

class Node {
 public:
  enum Kind { Null, Scalar, Map, Seq };
  Node() : kind_(Null) {}
  explicit Node(const std::string &s) : kind_(Scalar), scalar_(s) {}

  explicit operator bool() const { return kind_ != Null; }
  bool IsDefined() const { return kind_ != Null; }
  bool IsMap() const { return kind_ == Map; }
  bool IsSequence() const { return kind_ == Seq; }
  bool IsScalar() const { return kind_ == Scalar; }
  size_t size() const { return kind_ == Seq ? seq_.size() : (kind_ == Map ? keys_.size() : 0); }

  // missing keys give a Null node (tests false), like yaml-cpp's const access
  Node operator[](const std::string &key) const {
    if (kind_ != Map) return Node();
    auto it = map_.find(key);
    return it == map_.end() ? Node() : *it->second;
  }
  Node operator[](const char *key) const { return (*this)[std::string(key)]; }
  Node operator[](size_t i) const { return (kind_ == Seq && i < seq_.size()) ? *seq_[i] : Node(); }
  Node operator[](int i) const { return (*this)[(size_t)i]; }
  const std::vector<std::string> &keys() const { return keys_; }

  template <class T>
  T as() const;

  // ---- construction (parser) -----------------------------------------------------------------------------------------
  void make_map() { kind_ = Map; }
  void make_seq() { kind_ = Seq; }
  void set(const std::string &key, const Node &v) {
    if (!map_.count(key)) keys_.push_back(key);
    map_[key] = std::make_shared<Node>(v);
  }
  void push(const Node &v) { seq_.push_back(std::make_shared<Node>(v)); }
  const std::string &scalar() const { return scalar_; }

 private:
  Kind kind_;
  std::string scalar_;
  std::map<std::string, std::shared_ptr<Node>> map_;
  std::vector<std::string> keys_;
  std::vector<std::shared_ptr<Node>> seq_;
};

namespace detail {
inline std::string trim(const std::string &s) {
  size_t b = s.find_first_not_of(" \t\r\n"), e = s.find_last_not_of(" \t\r\n");
  return b == std::string::npos ? std::string() : s.substr(b, e - b + 1);
}
inline std::string strip_comment(const std::string &s) {
  bool sq = false, dq = false;
  for (size_t i = 0; i < s.size(); ++i) {
    char c = s[i];
    if (c == '\'' && !dq) sq = !sq;
    else if (c == '"' && !sq) dq = !dq;
    else if (c == '#' && !sq && !dq && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) return s.substr(0, i);
  }
  return s;
}
inline std::string unquote(const std::string &s) {
  if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\''))) return s.substr(1, s.size() - 2);
  return s;
}
inline Node parse_value(const std::string &raw) {
  std::string v = trim(raw);
  if (v.size() >= 2 && v.front() == '[' && v.back() == ']') {
    Node n;
    n.make_seq();
    std::string inner = v.substr(1, v.size() - 2), item;
    std::stringstream ss(inner);
    while (std::getline(ss, item, ',')) {
      item = trim(item);
      if (!item.empty()) n.push(Node(unquote(item)));
    }
    return n;
  }
  if (v == "~" || v == "null") return Node();
  return Node(unquote(v));
}
struct Line { int indent; std::string text; int lineno; };

inline Node parse_block(const std::vector<Line> &L, size_t &pos, int indent) {
  Node node;
  if (pos >= L.size()) return node;
  if (L[pos].text.compare(0, 2, "- ") == 0 || L[pos].text == "-") {
    node.make_seq();
    while (pos < L.size() && L[pos].indent == indent && (L[pos].text.compare(0, 2, "- ") == 0 || L[pos].text == "-")) {
      std::string item = trim(L[pos].text.substr(1));
      ++pos;
      if (item.empty() && pos < L.size() && L[pos].indent > indent) node.push(parse_block(L, pos, L[pos].indent));
      else node.push(parse_value(item));
    }
    return node;
  }
  node.make_map();
  while (pos < L.size() && L[pos].indent == indent) {
    const std::string &t = L[pos].text;
    size_t colon = std::string::npos;
    bool sq = false, dq = false;
    for (size_t i = 0; i < t.size(); ++i) {
      char c = t[i];
      if (c == '\'' && !dq) sq = !sq;
      else if (c == '"' && !sq) dq = !dq;
      else if (c == ':' && !sq && !dq && (i + 1 == t.size() || t[i + 1] == ' ' || t[i + 1] == '\t')) { colon = i; break; }
    }
    if (colon == std::string::npos)
      throw std::runtime_error("yaml: line " + std::to_string(L[pos].lineno) + ": expected 'key: value'");
    std::string key = unquote(trim(t.substr(0, colon))), rest = trim(t.substr(colon + 1));
    ++pos;
    if (rest.empty()) {
      if (pos < L.size() && L[pos].indent > indent) node.set(key, parse_block(L, pos, L[pos].indent));
      else if (pos < L.size() && L[pos].indent == indent && L[pos].text.compare(0, 2, "- ") == 0) node.set(key, parse_block(L, pos, indent));
      else node.set(key, Node());
    } else {
      node.set(key, parse_value(rest));
    }
  }
  if (pos < L.size() && L[pos].indent > indent)
    throw std::runtime_error("yaml: line " + std::to_string(L[pos].lineno) + ": unexpected indentation");
  return node;
}
}  // namespace detail

inline Node Load(const std::string &text) {
  std::vector<detail::Line> lines;
  std::stringstream ss(text);
  std::string raw;
  int lineno = 0;
  while (std::getline(ss, raw)) {
    ++lineno;
    std::string s = detail::strip_comment(raw);
    if (detail::trim(s).empty()) continue;
    if (detail::trim(s) == "---" || detail::trim(s) == "...") continue;
    int indent = 0;
    while (indent < (int)s.size() && s[indent] == ' ') ++indent;
    if (indent < (int)s.size() && s[indent] == '\t') throw std::runtime_error("yaml: line " + std::to_string(lineno) + ": tab indentation");
    lines.push_back({indent, detail::trim(s), lineno});
  }
  size_t pos = 0;
  if (lines.empty()) return Node();
  Node root = detail::parse_block(lines, pos, lines[0].indent);
  if (pos != lines.size()) throw std::runtime_error("yaml: line " + std::to_string(lines[pos].lineno) + ": inconsistent indentation");
  return root;
}
inline Node LoadFile(const std::string &path) {
  std::ifstream f(path);
  if (!f) throw std::runtime_error("yaml: cannot open " + path);
  std::stringstream ss;
  ss << f.rdbuf();
  return Load(ss.str());
}

template <>
inline std::string Node::as<std::string>() const {
  if (kind_ != Scalar) throw std::runtime_error("yaml: scalar expected");
  return scalar_;
}
template <>
inline double Node::as<double>() const {
  if (kind_ != Scalar) throw std::runtime_error("yaml: scalar expected");
  char *end = nullptr;
  double v = std::strtod(scalar_.c_str(), &end);
  if (end == scalar_.c_str() || *end) throw std::runtime_error("yaml: '" + scalar_ + "' is not a number");
  return v;
}
template <>
inline bool Node::as<bool>() const {
  if (kind_ != Scalar) throw std::runtime_error("yaml: scalar expected");
  const std::string &s = scalar_;
  if (s == "true" || s == "True" || s == "TRUE" || s == "yes" || s == "on" || s == "1") return true;
  if (s == "false" || s == "False" || s == "FALSE" || s == "no" || s == "off" || s == "0") return false;
  throw std::runtime_error("yaml: '" + s + "' is not a boolean");
}
template <>
inline long long Node::as<long long>() const {
  if (kind_ != Scalar) throw std::runtime_error("yaml: scalar expected");
  if (scalar_ == "true") return 1;
  if (scalar_ == "false") return 0;
  char *end = nullptr;
  long long v = std::strtoll(scalar_.c_str(), &end, 10);
  if (end == scalar_.c_str() || *end) {
    double d = as<double>();  // "1e3"-style integers
    return (long long)d;
  }
  return v;
}
template <>
inline int Node::as<int>() const { return (int)as<long long>(); }
template <>
inline std::vector<int> Node::as<std::vector<int>>() const {
  std::vector<int> out;
  for (size_t i = 0; i < size(); ++i) out.push_back((*this)[i].as<int>());
  return out;
}
template <>
inline std::vector<double> Node::as<std::vector<double>>() const {
  std::vector<double> out;
  for (size_t i = 0; i < size(); ++i) out.push_back((*this)[i].as<double>());
  return out;
}

}  // namespace YAML
