// host harness: prints the rows of the two synthetic workloads as the product's row functions (csrc/workload_rows.cuh) build
// them, for comparison with the numpy builders of tests/workloads.py.  usage: workload_rows <poisson|momentum> <n> <seed>
#include "workload_rows.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
int main(int argc, char **argv) {
  const bool poisson = strcmp(argv[1], "poisson") == 0;
  const int n = atoi(argv[2]);
  const unsigned long long seed = strtoull(argv[3], nullptr, 10);
  const long long rows = poisson ? (long long)n * n * n : 3LL * n * n * n;
  long long cols[32];
  double vals[32], rhs;
  for (long long r = 0; r < rows; ++r) {
    const int c = poisson ? b200::wl::poisson27_row(n, r, seed, 0.2, cols, vals, &rhs) : b200::wl::momentum3_row(n, r, seed, cols, vals, &rhs);
    for (int k = 0; k < c; ++k) printf("%lld %lld %.17g\n", r, cols[k], vals[k]);
    printf("b %lld %.17g\n", r, rhs);
  }
  return 0;
}
