// mm_text.h -- text-file primitives of the MatrixMarket / HYPRE-IJ loaders (app/HypreSystem.cpp): a read-only memory map
// with a line cursor, bounded number parsing, the MatrixMarket banner and size line.  Written fresh; accepts what the
// reference's reader accepts (/root/reference/src/mmio.c:96-217 mm_read_banner / mm_read_mtx_crd_size, used from
// /root/reference/src/HypreSystem.cpp:1670-1969) -- tests/test_mmio_crosscheck.py compares the two on odd banners.
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <sstream>
#include <stdexcept>
#include <string>

namespace nalu {
namespace mmtext {

// read-only memory map of a text file with a line cursor
class TextFile {
 public:
  explicit TextFile(const std::string &path, const char *what) : path_(path) {
    fd_ = open(path.c_str(), O_RDONLY);
    if (fd_ < 0) throw std::runtime_error(std::string("Cannot open ") + what + " file: " + path);
    struct stat st;
    if (fstat(fd_, &st) != 0) { close(fd_); throw std::runtime_error("Cannot stat " + path); }
    size_ = (size_t)st.st_size;
    if (size_) {
      void *p = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
      if (p == MAP_FAILED) { close(fd_); throw std::runtime_error("Cannot map " + path); }
      data_ = (const char *)p;
    }
  }
  ~TextFile() {
    if (data_) munmap((void *)data_, size_);
    if (fd_ >= 0) close(fd_);
  }
  // next line as [b, e) without the terminator; false at end of file
  bool next(const char *&b, const char *&e) {
    if (pos_ >= size_) return false;
    b = data_ + pos_;
    const char *nl = (const char *)memchr(b, '\n', size_ - pos_);
    e = nl ? nl : data_ + size_;
    pos_ = (size_t)(e - data_) + (nl ? 1 : 0);
    if (e > b && e[-1] == '\r') --e;
    return true;
  }
  const std::string &path() const { return path_; }

 private:
  std::string path_;
  int fd_ = -1;
  size_t size_ = 0, pos_ = 0;
  const char *data_ = nullptr;
};

inline bool blank(const char *b, const char *e) {
  for (; b < e; ++b) if (*b != ' ' && *b != '\t') return false;
  return true;
}
// numbers are parsed from a bounded copy of the line (the mapping is not NUL-terminated)
struct LineParser {
  char buf[512];
  char *cur = nullptr;
  void set(const char *b, const char *e) {
    size_t n = std::min<size_t>((size_t)(e - b), sizeof(buf) - 1);
    memcpy(buf, b, n);
    buf[n] = 0;
    cur = buf;
  }
  bool integer(long long &v) {
    char *end;
    v = strtoll(cur, &end, 10);
    if (end == cur) return false;
    cur = end;
    return true;
  }
  bool real(double &v) {
    char *end;
    v = strtod(cur, &end);
    if (end == cur) return false;
    cur = end;
    return true;
  }
};

struct MMBanner {
  bool coordinate = false, array = false, complex_field = false, pattern = false, integer_field = false;
  char symmetry = 'G';   // G general, S symmetric, K skew-symmetric, H hermitian
};
// "%%MatrixMarket matrix <coordinate|array> <real|integer|complex|pattern> <general|symmetric|skew-symmetric|hermitian>".
// Acceptance follows the reference's reader (mmio.c:96-183 mm_read_banner + :86-94 mm_is_valid): the banner token is matched
// with its case ("%%MatrixMarket" as a prefix), the four fields are case-insensitive, all five tokens must be present.
inline MMBanner read_mm_banner(TextFile &f, const char *what) {
  const char *b, *e;
  if (!f.next(b, e)) throw std::runtime_error(std::string("Cannot read ") + what + " banner");
  std::string line(b, e), tok[5];
  std::stringstream ss(line);
  for (auto &t : tok) ss >> t;
  if (tok[4].empty()) throw std::runtime_error(std::string("Cannot read ") + what + " banner");
  for (int k = 1; k < 5; ++k) std::transform(tok[k].begin(), tok[k].end(), tok[k].begin(), ::tolower);
  if (tok[0].compare(0, 14, "%%MatrixMarket") != 0) throw std::runtime_error(std::string("Cannot read ") + what + " banner");
  MMBanner m;
  m.coordinate = tok[2] == "coordinate";
  m.array = tok[2] == "array";
  m.complex_field = tok[3] == "complex";
  m.pattern = tok[3] == "pattern";
  m.integer_field = tok[3] == "integer";
  const bool real_field = tok[3] == "real";
  const bool field_ok = real_field || m.integer_field || m.complex_field || m.pattern;
  bool sym_ok = true;
  if (tok[4] == "general") m.symmetry = 'G';
  else if (tok[4] == "symmetric") m.symmetry = 'S';
  else if (tok[4] == "skew-symmetric") m.symmetry = 'K';
  else if (tok[4] == "hermitian") m.symmetry = 'H';
  else sym_ok = false;
  if (tok[1] != "matrix" || !(m.coordinate || m.array) || !field_ok || !sym_ok)
    throw std::runtime_error(std::string("Cannot read ") + what + " banner");
  // mm_is_valid
  if ((m.array && m.pattern) || (real_field && m.symmetry == 'H') || (m.pattern && (m.symmetry == 'H' || m.symmetry == 'K')))
    throw std::runtime_error("Invalid matrix market file encountered");
  return m;
}
// first non-comment, non-blank line after the banner
inline void read_size_line(TextFile &f, LineParser &lp) {
  const char *b, *e;
  while (f.next(b, e)) {
    if (b < e && *b == '%') continue;
    if (blank(b, e)) continue;
    lp.set(b, e);
    return;
  }
  throw std::runtime_error("Cannot read sizes in file: " + f.path());
}


}  // namespace mmtext
}  // namespace nalu
