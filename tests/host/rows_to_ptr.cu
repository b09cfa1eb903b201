// Host check of host_rows_to_ptr (hypre-mini-app_b200/csrc/ij.cu): the pass that turns the row ids of a row-sorted triplet list into a
// local row pointer on the host threads while cols / vals are in flight ("End-to-end path").  The function lives in
// the anonymous namespace of ij.cu, so this harness includes that translation unit and links the rest from libb200hypre.so; it makes
// no CUDA call.  Reads cases from stdin: "n ilower iupper" then n row ids; prints "bad" or the nloc + 1 pointer entries.
#include "../../hypre-mini-app_b200/csrc/ij.cu"
#include <iostream>

int main() {
  long long n, lo, hi;
  while (std::cin >> n >> lo >> hi) {
    std::vector<b200::i64> rows((size_t)n);
    for (auto &r : rows) std::cin >> r;
    const long long nloc = hi - lo + 1;
    std::vector<int> ptr((size_t)nloc + 1, -7);
    const bool ok = b200::host_rows_to_ptr(rows.data(), n, lo, hi, ptr.data());
    if (!ok) { std::cout << "bad\n"; continue; }
    for (size_t k = 0; k < ptr.size(); ++k) std::cout << ptr[k] << (k + 1 < ptr.size() ? ' ' : '\n');
  }
  return 0;
}
