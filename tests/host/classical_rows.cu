// Host run of rows::classical_row and rows::interp_row (hypre-mini-app_b200/csrc/amg_rows.cuh, __host__ __device__): reads one level
// (A, S, CF marker, trunc_factor, P_max_elmts) as raw little-endian arrays, builds every interpolation row with the product's row
// function and writes P the same way.  tests/test_host_rows.py compares the result bit for bit with the oracle's P.
//   usage: classical_rows <in.bin> <out.bin> <interp_type>
// in : int32 n, nnzA, nnzS, pmax; float64 tf; int32 Aptr[n+1], Acol[nnzA]; float64 Aval[nnzA]; int32 Sptr[n+1], Scol[nnzS], cf[n]
// out: int32 nnzP; int32 Pptr[n+1], Pcol[nnzP]; float64 Pval[nnzP]
#include "amg_rows.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace b200;

template <class T>
static void rd(FILE *f, T *p, size_t n) {
  if (n && fread(p, sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
}

int main(int argc, char **argv) {
  if (argc < 4) return 2;
  FILE *f = fopen(argv[1], "rb");
  if (!f) return 2;
  int hdr[4];
  double tf;
  rd(f, hdr, 4);
  rd(f, &tf, 1);
  const int n = hdr[0], nnzA = hdr[1], nnzS = hdr[2], pmax = hdr[3], itype = atoi(argv[3]);
  std::vector<int> Aptr(n + 1), Acol(nnzA), Sptr(n + 1), Scol(nnzS), cf(n), f2c(n, -1);
  std::vector<double> Aval(nnzA);
  rd(f, Aptr.data(), n + 1); rd(f, Acol.data(), nnzA); rd(f, Aval.data(), nnzA);
  rd(f, Sptr.data(), n + 1); rd(f, Scol.data(), nnzS); rd(f, cf.data(), n);
  fclose(f);
  int nc = 0;
  for (int i = 0; i < n; ++i)
    if (cf[i] >= 0) f2c[i] = nc++;
  rows::Csr A{n, n, Aptr.data(), Acol.data(), Aval.data()};
  std::vector<int> Pptr(n + 1, 0), Pcol;
  std::vector<double> Pval;
  const int cap = 4096;
  std::vector<int> lc(cap), oc(cap);
  std::vector<double> lv(cap), ov(cap);
  for (int i = 0; i < n; ++i) {
    int L;
    if (itype == 0) {
      if (cf[i] >= 0) { L = 1; oc[0] = f2c[i]; ov[0] = 1.0; }
      else if (cf[i] == -3) L = 0;
      else L = rows::classical_row(A, Sptr.data(), Scol.data(), cf.data(), f2c.data(), n, i, tf, pmax, oc.data(), ov.data(), lc.data(), lv.data());
    } else {
      L = rows::interp_row(A, Sptr.data(), Scol.data(), cf.data(), f2c.data(), n, i, tf, pmax, false, oc.data(), ov.data(), lc.data(), lv.data(), cap);
      if (L < 0) { fprintf(stderr, "scratch overflow\n"); return 3; }
    }
    for (int k = 0; k < L; ++k) { Pcol.push_back(oc[k]); Pval.push_back(ov[k]); }
    Pptr[i + 1] = Pptr[i] + L;
  }
  f = fopen(argv[2], "wb");
  const int nnzP = (int)Pcol.size();
  fwrite(&nnzP, 4, 1, f);
  fwrite(Pptr.data(), 4, n + 1, f);
  fwrite(Pcol.data(), 4, nnzP, f);
  fwrite(Pval.data(), 8, nnzP, f);
  fclose(f);
  return 0;
}
