// amg_rows.cuh -- per-row device functions of the BoomerAMG setup (strength, PMIS steps, extended+i interpolation
// with truncation, l1 norms).  One GPU thread owns one row and walks it in storage order, so every
// floating-point value that later feeds a discrete decision (strength test, truncation ties) is produced by a
// fixed sequential operation order; the translation unit is compiled with -fmad=false.
//
// The functions are __host__ __device__ only so that tests/emul (a host harness that loops over rows) can
// exercise the same logic in this GPU-less container; the product never runs them on the host.
//
// Algorithms restated from hypre's published definitions (not in /root/reference):
//   strength_row   hypre_BoomerAMGCreateS            SURVEY.md A.3
//   pmis_*         hypre_BoomerAMGCoarsenPMIS        SURVEY.md A.4   (seed: /root/reference/src/main.cpp:169)
//   interp_row     hypre_BoomerAMGBuildExtPIInterp + hypre_BoomerAMGInterpTruncation   SURVEY.md A.5
// reached from HYPRE_BoomerAMGSetup (/root/reference/src/HypreSystem.cpp:323 via :692).
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define B200_HD __host__ __device__ __forceinline__
#else
#define B200_HD inline
#endif

namespace b200 {
namespace rows {

struct Csr {  // same layout as b200::CsrView
  int nr, nc;
  const int *ptr;
  const int *col;
  const double *val;
};

// ---- A.3 strength of connection.  Returns the number of strong entries of row i; writes their columns to out
// (A's storage order) when out != nullptr.  Diagonal is the first entry of the row.
B200_HD int strength_row(const Csr &A, int i, double theta, double max_row_sum, int *out) {
  int b = A.ptr[i], e = A.ptr[i + 1];
  if (e == b) return 0;
  double diag = A.val[b];
  double row_scale = 0.0, row_sum = diag;
  for (int k = b + 1; k < e; ++k) {
    double a = A.val[k];
    if (diag < 0) row_scale = (a > row_scale) ? a : row_scale;
    else row_scale = (a < row_scale) ? a : row_scale;
    row_sum = row_sum + a;
  }
  if ((fabs(row_sum) > fabs(diag) * max_row_sum) && (max_row_sum < 1.0)) return 0;
  double thr = theta * row_scale;
  int cnt = 0;
  for (int k = b + 1; k < e; ++k) {
    double a = A.val[k];
    bool strong = (diag < 0) ? (a > thr) : (a < thr);
    if (strong) {
      if (out) out[cnt] = A.col[k];
      ++cnt;
    }
  }
  return cnt;
}

// ---- A.2 hypre_Rand stream value number (idx+1) of seed: seed * 16807^(idx+1) mod (2^31 - 1), as a double in (0,1)
B200_HD double hypre_rand_at(int seed, long long idx) {
  const unsigned long long m = 2147483647ULL;
  unsigned long long base = 16807ULL, r = (unsigned long long)seed % m;
  unsigned long long e = (unsigned long long)(idx + 1);
  while (e) {
    if (e & 1ULL) r = (r * base) % m;
    base = (base * base) % m;
    e >>= 1;
  }
  return (double)(int)r / 2147483647.0;
}

// ---- A.4 PMIS steps (cf: 0 undecided, +1 C, -1 F, -3 isolated) ---------------------------------------------
B200_HD void pmis_knock_row(const int *Sptr, const int *Scol, const double *measure, const int *cf, signed char *is, int i) {
  if (cf[i] != 0) return;
  double mi = measure[i];
  if (!(mi > 1.0)) return;
  for (int k = Sptr[i]; k < Sptr[i + 1]; ++k) {
    int j = Scol[k];
    double mj = measure[j];
    if (mj > 1.0) {
      if (mi > mj) is[j] = 0;
      else if (mj > mi) is[i] = 0;
    }
  }
}
B200_HD void pmis_setf_row(const int *Sptr, const int *Scol, int *cf, int i) {
  if (cf[i] != 0) return;
  for (int k = Sptr[i]; k < Sptr[i + 1]; ++k)
    if (cf[Scol[k]] > 0) { cf[i] = -1; return; }
}

// ---- hypre_qsort2abs without recursion (same swap sequence inside every sub-range; disjoint sub-ranges commute)
B200_HD void swap2(int *v, double *w, int i, int j) {
  int t = v[i]; v[i] = v[j]; v[j] = t;
  double d = w[i]; w[i] = w[j]; w[j] = d;
}
// Only positions [0, keep) of the result are used by the caller (P_max_elmts truncation): a sub-range that starts at or
// beyond `keep` is left unsorted.  Sub-ranges are partitioned independently of each other, so the first `keep` entries are
// exactly those of the full sort.
B200_HD void qsort2abs(int *v, double *w, int n, int keep) {
  int stk[64];
  int sp = 0;
  stk[sp++] = 0;
  stk[sp++] = n - 1;
  while (sp > 0) {
    int right = stk[--sp];
    int left = stk[--sp];
    while (left < right && left < keep) {
      swap2(v, w, left, (left + right) / 2);
      int last = left;
      for (int i = left + 1; i <= right; ++i)
        if (fabs(w[i]) > fabs(w[left])) swap2(v, w, ++last, i);
      swap2(v, w, left, last);
      const int l1 = left, r1 = last - 1, l2 = last + 1, r2 = right;
      const bool do1 = l1 < r1, do2 = l2 < r2 && l2 < keep;
      if (do1 && do2) {  // continue with the shorter part, stack the longer one (bounds the stack by log2 n entries)
        if (r1 - l1 < r2 - l2) { stk[sp++] = l2; stk[sp++] = r2; left = l1; right = r1; }
        else { stk[sp++] = l1; stk[sp++] = r1; left = l2; right = r2; }
      } else if (do1) { left = l1; right = r1; }
      else if (do2) { left = l2; right = r2; }
      else break;
    }
  }
}

B200_HD int list_find(const int *lc, int L, int key) {
  for (int k = 0; k < L; ++k)
    if (lc[k] == key) return k;
  return -1;
}

// ---- hypre_BoomerAMGInterpTruncation on one row whose columns are still fine ids (lc) with weights lv: relative threshold, then
// the p_max largest by qsort2abs' order, each with its rescaling; own-range columns (< n_own) stay in front.  Writes the coarse
// columns f2c[lc[k]] and the weights to oc / ov (may alias lc / lv) and returns the final length.
B200_HD int trunc_emit(int *lc, double *lv, int L, bool any_off, int n_own, const int *f2c, double trunc_factor, int p_max, int *oc,
                       double *ov) {
  // keep fine ids for the own/off test during re-splitting; translate at the end
  int len = L;
  if (trunc_factor > 0.0 && len > 0) {
    double max_coef = 0.0, row_sum = 0.0, scale = 0.0;
    for (int k = 0; k < len; ++k) { double a = fabs(lv[k]); max_coef = (a > max_coef) ? a : max_coef; }
    max_coef = max_coef * trunc_factor;
    int w = 0;
    for (int k = 0; k < len; ++k) {
      row_sum = row_sum + lv[k];
      if (fabs(lv[k]) >= max_coef) { scale = scale + lv[k]; lc[w] = lc[k]; lv[w] = lv[k]; ++w; }
    }
    len = w;
    if (scale != 0.0 && scale != row_sum) {
      double f = row_sum / scale;
      for (int k = 0; k < len; ++k) lv[k] = lv[k] * f;
    }
  }
  if (p_max > 0 && len > p_max) {
    double row_sum = 0.0, scale = 0.0;
    for (int k = 0; k < len; ++k) row_sum = row_sum + lv[k];
    qsort2abs(lc, lv, len, p_max);
    for (int k = 0; k < p_max; ++k) scale = scale + lv[k];
    len = p_max;
    if (scale != 0.0 && scale != row_sum) {
      double f = row_sum / scale;
      for (int k = 0; k < len; ++k) lv[k] = lv[k] * f;
    }
    if (any_off) {  // re-split, each part keeping the sorted order
      for (int k = 1; k < len; ++k) {
        int v = lc[k];
        double x = lv[k];
        if (v < n_own) {
          int p = k;
          while (p > 0 && lc[p - 1] >= n_own) { lc[p] = lc[p - 1]; lv[p] = lv[p - 1]; --p; }
          lc[p] = v;
          lv[p] = x;
        }
      }
    }
  }
  for (int k = 0; k < len; ++k) { oc[k] = f2c[lc[k]]; ov[k] = lv[k]; }
  return len;
}

// ---- A.5 extended+i interpolation row + truncation.
// A, S in local (extended) numbering; rows/cols < n_own are owned by this rank.  f2c maps a fine point to its coarse
// column index in the same (extended) coarse numbering; a coarse column is "own" iff its fine point is < n_own.
// lc/lv: scratch of capacity cap.  count_only: stop after discovery (valid when trunc_factor == 0).
// Returns the final row length, or -1 when the scratch overflowed (the caller re-runs the row with a larger one).
B200_HD int interp_row(const Csr &A, const int *Sptr, const int *Scol, const int *cf, const int *f2c, int n_own, int i,
                       double trunc_factor, int p_max, bool count_only, int *oc, double *ov, int *lc, double *lv, int cap) {
  if (cf[i] >= 0) {
    if (!count_only) { oc[0] = f2c[i]; ov[0] = 1.0; }
    return 1;
  }
  if (cf[i] == -3) return 0;
  int L = 0;
  bool any_off = false;
  // discovery of C-hat_i in hypre's order
  for (int kk = Sptr[i]; kk < Sptr[i + 1]; ++kk) {
    int i1 = Scol[kk];
    if (cf[i1] >= 0) {
      if (list_find(lc, L, i1) < 0) {
        if (L == cap) return -1;
        lc[L++] = i1;
        any_off |= (i1 >= n_own);
      }
    } else if (cf[i1] != -3) {
      for (int k2 = Sptr[i1]; k2 < Sptr[i1 + 1]; ++k2) {
        int k1 = Scol[k2];
        if (cf[k1] >= 0 && list_find(lc, L, k1) < 0) {
          if (L == cap) return -1;
          lc[L++] = k1;
          any_off |= (k1 >= n_own);
        }
      }
    }
  }
  if (any_off) {  // stable partition: own-range C points first (P_diag before P_offd)
    for (int k = 1; k < L; ++k) {
      int v = lc[k];
      if (v < n_own) {
        int p = k;
        while (p > 0 && lc[p - 1] >= n_own) { lc[p] = lc[p - 1]; --p; }
        lc[p] = v;
      }
    }
  }
  if (count_only) return (p_max > 0 && L > p_max) ? p_max : L;

  for (int k = 0; k < L; ++k) lv[k] = 0.0;
  double diagonal = A.val[A.ptr[i]];
  int sp = Sptr[i], se = Sptr[i + 1];
  for (int jj = A.ptr[i] + 1; jj < A.ptr[i + 1]; ++jj) {
    int i1 = A.col[jj];
    double aij = A.val[jj];
    bool strong = (sp < se) && (Scol[sp] == i1);
    if (strong) ++sp;
    int pos = list_find(lc, L, i1);
    if (pos >= 0) {
      lv[pos] = lv[pos] + aij;
    } else if (strong && cf[i1] < 0 && cf[i1] != -3) {
      double sum = 0.0;
      int b1 = A.ptr[i1], e1 = A.ptr[i1 + 1];
      double sgn = (A.val[b1] < 0) ? -1.0 : 1.0;
      for (int j1 = b1 + 1; j1 < e1; ++j1) {
        int i2 = A.col[j1];
        double a = A.val[j1];
        if ((sgn * a) < 0 && (i2 == i || list_find(lc, L, i2) >= 0)) sum = sum + a;
      }
      if (sum != 0.0) {
        double distribute = aij / sum;
        for (int j1 = b1 + 1; j1 < e1; ++j1) {
          int i2 = A.col[j1];
          double a = A.val[j1];
          if ((sgn * a) < 0) {
            int p2 = list_find(lc, L, i2);
            if (p2 >= 0) { double t = distribute * a; lv[p2] = lv[p2] + t; }
            if (i2 == i) { double t = distribute * a; diagonal = diagonal + t; }
          }
        }
      } else {
        diagonal = diagonal + aij;
      }
    } else if (cf[i1] != -3) {
      diagonal = diagonal + aij;
    }
  }
  if (diagonal != 0.0) {
    double md = -diagonal;
    for (int k = 0; k < L; ++k) lv[k] = lv[k] / md;
  }
  return trunc_emit(lc, lv, L, any_off, n_own, f2c, trunc_factor, p_max, oc, ov);
}

// ---- classical "modified" interpolation row (interp_type 0, hypre_BoomerAMGBuildInterp) + truncation: the reference's sample input
// selects it (/root/reference/etc/hypre_app.yaml:41 through src/HypreSystem.cpp:195-198).  Same conventions as interp_row.  The
// interpolatory set is C_i = the strong C neighbours of i in the order of the strength row; a strong F neighbour k is distributed over
// C_i in proportion to the a_km (m in C_i) whose sign is opposite to a_kk, or goes to the diagonal when there is none; weak neighbours
// go to the diagonal.  lc / lv: scratch of at least (length of the strength row of i) entries; oc / ov may alias them.
// Rows of C points (one entry, 1.0) and of special points (empty) are the caller's: call this for cf[i] == -1 only.
B200_HD int classical_row(const Csr &A, const int *Sptr, const int *Scol, const int *cf, const int *f2c, int n_own, int i,
                          double trunc_factor, int p_max, int *oc, double *ov, int *lc, double *lv) {
  int L = 0;
  bool any_off = false;
  for (int kk = Sptr[i]; kk < Sptr[i + 1]; ++kk) {
    const int j = Scol[kk];
    if (cf[j] >= 0) { lc[L++] = j; any_off |= (j >= n_own); }
  }
  if (any_off) {  // own-range C points first (P_diag before P_offd); a no-op for rows stored diag | offd
    for (int k = 1; k < L; ++k) {
      const int v = lc[k];
      if (v < n_own) {
        int p = k;
        while (p > 0 && lc[p - 1] >= n_own) { lc[p] = lc[p - 1]; --p; }
        lc[p] = v;
      }
    }
  }
  for (int k = 0; k < L; ++k) lv[k] = 0.0;
  // membership filter for "is column m in C_i": one bit per (column mod 64); a row of a neighbour has few members of C_i among its
  // entries, and the filter rejects most of the others without the list search
  unsigned long long filter = 0ull;
  for (int k = 0; k < L; ++k) filter |= 1ull << (lc[k] & 63);
  double diagonal = A.val[A.ptr[i]];
  int sp = Sptr[i];
  const int se = Sptr[i + 1];
  for (int jj = A.ptr[i] + 1; jj < A.ptr[i + 1]; ++jj) {
    const int j = A.col[jj];
    const double aij = A.val[jj];
    const bool strong = (sp < se) && (Scol[sp] == j);
    if (strong) ++sp;
    if (strong && cf[j] >= 0) {
      const int pos = list_find(lc, L, j);
      lv[pos] = lv[pos] + aij;
    } else if (strong && cf[j] != -3) {
      const int b1 = A.ptr[j], e1 = A.ptr[j + 1];
      const double sgn = (A.val[b1] < 0) ? -1.0 : 1.0;
      // one walk over row j: the sum in storage order, and the (slot, value) pairs that will receive a share, in the same order
      constexpr int HB = 16;
      int hpos[HB];
      double hval[HB];
      int nh = 0;
      double sum = 0.0;
      for (int j1 = b1 + 1; j1 < e1; ++j1) {
        const double a = A.val[j1];
        if ((sgn * a) < 0) {
          const int m = A.col[j1];
          if ((filter >> (m & 63)) & 1ull) {
            const int p2 = list_find(lc, L, m);
            if (p2 >= 0) {
              sum = sum + a;
              if (nh < HB) { hpos[nh] = p2; hval[nh] = a; }
              ++nh;
            }
          }
        }
      }
      if (sum != 0.0) {
        const double distribute = aij / sum;
        if (nh <= HB) {
          for (int h = 0; h < nh; ++h) { const double t = distribute * hval[h]; lv[hpos[h]] = lv[hpos[h]] + t; }
        } else {  // more members than the buffer holds: walk the row again
          for (int j1 = b1 + 1; j1 < e1; ++j1) {
            const double a = A.val[j1];
            if ((sgn * a) < 0) {
              const int p2 = list_find(lc, L, A.col[j1]);
              if (p2 >= 0) { const double t = distribute * a; lv[p2] = lv[p2] + t; }
            }
          }
        }
      } else {
        diagonal = diagonal + aij;
      }
    } else if (cf[j] != -3) {
      diagonal = diagonal + aij;
    }
  }
  if (diagonal == 0.0) {
    for (int k = 0; k < L; ++k) lv[k] = 0.0;
  } else {
    const double md = -diagonal;
    for (int k = 0; k < L; ++k) lv[k] = lv[k] / md;
  }
  return trunc_emit(lc, lv, L, any_off, n_own, f2c, trunc_factor, p_max, oc, ov);
}

// ---- smoother diagonal: relax 7 -> a_ii ; relax 18 -> l1 norm (sign of a_ii), SURVEY.md A.7
B200_HD double smoother_diag_row(const Csr &A, int i, int relax_type) {
  int b = A.ptr[i], e = A.ptr[i + 1];
  if (b == e) return 1.0;
  double d = A.val[b];
  if (relax_type == 7) return d;
  double s = 0.0;
  for (int k = b; k < e; ++k) s = s + fabs(A.val[k]);
  return (d < 0) ? -s : s;
}

}  // namespace rows
}  // namespace b200
