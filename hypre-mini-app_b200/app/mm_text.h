// mm_text.h -- text-file primitives of the MatrixMarket / HYPRE-IJ loaders (app/HypreSystem.cpp): a read-only memory map
// with a line cursor, bounded number parsing, a multi-threaded reader of body lines, the MatrixMarket banner and size line.  Written fresh; accepts what the
// reference's reader accepts (/root/reference/src/mmio.c:96-217 mm_read_banner / mm_read_mtx_crd_size, used from
// /root/reference/src/HypreSystem.cpp:1670-1969) -- tests/test_mmio_crosscheck.py compares the two on odd banners.
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <sched.h>
#include <algorithm>
#include <charconv>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

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
  // the unread rest of the file as one byte range; the cursor moves to the end (read_lines_parallel)
  void take_rest(const char *&b, const char *&e) {
    b = data_ + std::min(pos_, size_);
    e = data_ + size_;
    pos_ = size_;
  }

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

// Numbers read straight from the mapped line [p, e) with std::from_chars: no copy, no locale, no errno.  A token that from_chars does not
// take exactly like strtoll / strtod would (out-of-range values, hexadecimal floats, anything not followed by white space or the end of the
// line) makes the caller fall back to LineParser for that line, so every line is accepted, rejected and valued as before.
struct LineCursor {
  const char *p, *e;
  bool clean = true;   // false: re-parse this line with LineParser
  void skip() {
    while (p < e && (*p == ' ' || *p == '\t')) ++p;
    if (p + 1 < e && *p == '+' && ((p[1] >= '0' && p[1] <= '9') || p[1] == '.')) ++p;
  }
  bool delimited() const { return p == e || *p == ' ' || *p == '\t'; }
  bool integer(long long &v) {
    skip();
    auto r = std::from_chars(p, e, v);
    if (r.ec != std::errc() || r.ptr == p) { clean = false; return false; }
    p = r.ptr;
    if (!delimited()) clean = false;
    return clean;
  }
  bool real(double &v) {
    skip();
    auto r = std::from_chars(p, e, v);
    if (r.ec != std::errc() || r.ptr == p) { clean = false; return false; }
    p = r.ptr;
    if (!delimited()) clean = false;
    return clean;
  }
};

// One body line "i j [re [im]]" of a coordinate file (MatrixMarket entries, HYPRE-IJ matrix rows).  Returns false for a blank line and,
// when percent_comments is set, for a '%' comment; throws on a malformed line.  Missing value fields keep what the caller put there.
inline bool parse_entry_line(const char *b, const char *e, bool percent_comments, bool pattern, bool cplx, const std::string &path,
                             long long &i, long long &j, double &re, double &im) {
  if ((percent_comments && b < e && *b == '%') || blank(b, e)) return false;
  LineCursor c{b, e};
  if (c.integer(i) && c.integer(j) && (pattern || c.real(re)) && cplx) c.real(im);
  if (!c.clean) {   // anything from_chars does not take like strtoll / strtod: the bounded-copy parser decides
    LineParser slow;
    slow.set(b, e);
    if (!slow.integer(i) || !slow.integer(j)) throw std::runtime_error("Malformed entry line in " + path);
    if (!pattern && !slow.real(re)) throw std::runtime_error("Malformed entry line in " + path);
    if (cplx && !slow.real(im)) throw std::runtime_error("complex_numbers: missing imaginary part in " + path);
  }
  return true;
}

// host threads a loader may use: the cores this process may run on, divided by the ranks that read at the same time, at most 32
inline int loader_threads(int ranks_on_host) {
  int avail = (int)std::thread::hardware_concurrency();
  cpu_set_t set;
  if (sched_getaffinity(0, sizeof(set), &set) == 0) avail = CPU_COUNT(&set);
  return std::max(1, std::min(avail / std::max(1, ranks_on_host), 32));
}

// Body lines of a text file from the cursor to the end, parsed by `threads` host threads: the byte range is cut at line boundaries,
// piece k is handed line by line to per_line(k, b, e) (line without its terminator; CR of a CRLF pair removed), and the caller appends
// what the pieces produced in the order k = 0, 1, ... -- exactly what one reader walking the file would have produced.  The text
// loaders of the mini-app are single-threaded fscanf loops (/root/reference/src/HypreSystem.cpp:1181-1247,1717-1850); a 50 M-row system
// is tens of GB of text.  An exception thrown by per_line in any piece is rethrown (the one of the earliest piece).
template <class PerLine>
void read_lines_parallel(TextFile &f, int threads, PerLine per_line) {
  const char *b, *e;
  f.take_rest(b, e);
  const size_t total = (size_t)(e - b);
  threads = (int)std::max<size_t>(1, std::min<size_t>((size_t)threads, total / (1 << 16) + 1));   // small files: one piece
  std::vector<const char *> cut((size_t)threads + 1, e);
  cut[0] = b;
  for (int k = 1; k < threads; ++k) {
    const char *q = b + total * (size_t)k / (size_t)threads;
    if (q < cut[k - 1]) q = cut[k - 1];
    const char *nl = (const char *)memchr(q, '\n', (size_t)(e - q));
    cut[k] = nl ? nl + 1 : e;
  }
  std::vector<std::exception_ptr> err((size_t)threads);
  auto work = [&](int k) {
    try {
      const char *p = cut[k], *end = cut[k + 1];
      while (p < end) {
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        const char *le = nl ? nl : end;
        const char *next = nl ? nl + 1 : end;
        if (le > p && le[-1] == '\r') --le;
        per_line(k, p, le);
        p = next;
      }
    } catch (...) {
      err[k] = std::current_exception();
    }
  };
  std::vector<std::thread> th;
  for (int k = 1; k < threads; ++k) th.emplace_back(work, k);
  work(0);
  for (auto &t : th) t.join();
  for (auto &x : err)
    if (x) std::rethrow_exception(x);
}

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
