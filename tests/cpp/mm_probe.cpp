// prints what the product's MatrixMarket text primitives (app/mm_text.h) make of a file: banner flags and the size line
#include "mm_text.h"
#include <cstdio>
int main(int argc, char **argv) {
  using namespace nalu::mmtext;
  try {
    TextFile f(argv[1], "matrix");
    MMBanner m = read_mm_banner(f, "matrix");
    LineParser lp;
    read_size_line(f, lp);
    long long a = 0, b = 0, c = 0;
    bool ok = lp.integer(a) && lp.integer(b) && (m.array || lp.integer(c));
    if (!ok) { printf("error sizes\n"); return 0; }
    printf("ok %c%c%c %lld %lld %lld\n", m.coordinate ? 'C' : 'A',
           m.complex_field ? 'C' : (m.pattern ? 'P' : (m.integer_field ? 'I' : 'R')), m.symmetry, a, b, c);
  } catch (const std::exception &e) {
    printf("error %s\n", e.what());
  }
  return 0;
}
