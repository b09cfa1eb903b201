// Host check of rows::qsort2abs(v, w, n, keep) (hypre-mini-app_b200/csrc/amg_rows.cuh): the first `keep` entries must be
// exactly those of the full recursive hypre_qsort2abs, ties included.  Built and run by tests/test_host_rows.py.
#include "amg_rows.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace b200;

static void full_sort(int *v, double *w, int left, int right) {
  if (left >= right) return;
  rows::swap2(v, w, left, (left + right) / 2);
  int last = left;
  for (int i = left + 1; i <= right; ++i)
    if (fabs(w[i]) > fabs(w[left])) rows::swap2(v, w, ++last, i);
  rows::swap2(v, w, left, last);
  full_sort(v, w, left, last - 1);
  full_sort(v, w, last + 1, right);
}

int main() {
  srand(20261104);
  long mismatches = 0, cases = 0;
  for (int it = 0; it < 100000; ++it) {
    const int n = 1 + rand() % 160, keep = 1 + rand() % 9;
    const int distinct = 1 + rand() % 7;   // few magnitudes: many ties, as in structured problems
    std::vector<int> v(n), v2;
    std::vector<double> w(n), w2;
    for (int k = 0; k < n; ++k) {
      v[k] = k;
      w[k] = ((rand() % distinct) + 1) * ((rand() & 1) ? -0.125 : 0.125);
    }
    v2 = v;
    w2 = w;
    full_sort(v.data(), w.data(), 0, n - 1);
    rows::qsort2abs(v2.data(), w2.data(), n, keep);
    for (int k = 0; k < keep && k < n; ++k)
      if (v[k] != v2[k] || w[k] != w2[k]) { ++mismatches; break; }
    // keep >= n: the whole array is sorted
    if (keep >= n)
      for (int k = 0; k < n; ++k)
        if (v[k] != v2[k]) { ++mismatches; break; }
    ++cases;
  }
  printf("cases %ld mismatches %ld\n", cases, mismatches);
  return mismatches != 0;
}
