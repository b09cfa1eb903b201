// prints the entries that the product's multi-threaded body reader (app/mm_text.h: read_lines_parallel + parse_entry_line, the code
// path of build_mm_matrix / build_ij_matrix in app/HypreSystem.cpp) makes of a file of "i j [re [im]]" lines, in file order
//   mm_entries_probe <file> <threads> <pattern 0|1> <complex 0|1> <percent-comments 0|1> [quiet]
#include "mm_text.h"
#include <cstdio>
int main(int argc, char **argv) {
  using namespace nalu::mmtext;
  if (argc < 6) return 2;
  const int threads = atoi(argv[2]);
  const bool pattern = atoi(argv[3]) != 0, cplx = atoi(argv[4]) != 0, pct = atoi(argv[5]) != 0;
  struct E { long long i, j; double re, im; };
  try {
    TextFile f(argv[1], "matrix");
    std::vector<std::vector<E>> piece((size_t)threads);
    const std::string path = argv[1];
    read_lines_parallel(f, threads, [&](int k, const char *b, const char *e) {
      long long i = 0, j = 0;
      double re = 1.0, im = 0.0;
      if (!parse_entry_line(b, e, pct, pattern, cplx, path, i, j, re, im)) return;
      piece[(size_t)k].push_back({i, j, re, im});
    });
    if (argc > 6) {   // any 6th argument: only the count and a checksum (timing runs)
      size_t n = 0;
      double sum = 0.0;
      for (auto &p : piece) { n += p.size(); for (auto &x : p) sum += x.re + (double)(x.i % 7); }
      printf("entries %zu checksum %.17g\n", n, sum);
      return 0;
    }
    for (auto &p : piece)
      for (auto &x : p) printf("%lld %lld %.17g %.17g\n", x.i, x.j, x.re, x.im);
  } catch (const std::exception &e) {
    printf("error %s\n", e.what());
  }
  return 0;
}
