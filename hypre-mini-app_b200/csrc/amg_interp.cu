// amg_interp.cu -- extended+i interpolation (interp_type 6) with truncation, one thread GROUP per row.
// Replaces hypre_BoomerAMGBuildExtPIInterp + hypre_BoomerAMGInterpTruncation reached from HYPRE_BoomerAMGSetup
// (/root/reference/src/HypreSystem.cpp:197,212,232 set the options, :323 via :692 runs it); formula and entry order:
// SURVEY.md A.5.  hypre itself is not in /root/reference.
//
// The row function rows::interp_row (amg_rows.cuh) is the sequential definition: C-hat_i is discovered in the order
// "strong neighbours of i in storage order, and for a strong F neighbour its strong C neighbours in storage order";
// every weight is a sum whose terms arrive in (entry of row i, entry of the neighbour row) order.  The group kernel
// below produces bit-identical rows with G lanes working on one row:
//   * discovery: the outer loop over the strong neighbours stays sequential; the lanes run across the strength row of
//     one F neighbour, membership is a shared-memory hash probe, and new C points take their list positions from a
//     ballot prefix, i.e. in storage order;
//   * weights: the loop over the entries of row i stays sequential; inside one entry the lanes run across the
//     neighbour row, where every term goes to a different C-hat slot (columns of one row are distinct), so each slot
//     still receives its terms in the sequential order and no floating-point atomics are needed.  The per-neighbour
//     denominator is summed in storage order by walking the ballot mask of the contributing lanes;
//   * truncation (trunc_factor, P_max_elmts with hypre's qsort2abs tie order) is done by the first lane on the
//     shared-memory list.
// Rows are tiered by |C-hat|: (G, list capacity) = (8, 32) when strength rows are short, then (32, 128); the few rows
// beyond that run rows::interp_row itself with a global-memory scratch list.  Compiled with -fmad=false (the
// operation order of every value that feeds a discrete decision is part of the parity contract).
#include "amg.h"
#include "amg_rows.cuh"
#include <cub/cub.cuh>
#include <cstdlib>

namespace b200 {

namespace {

#include "amg_interp_flat.cuh"

template <int G>
__device__ __forceinline__ unsigned group_mask() {
  if (G >= 32) return 0xffffffffu;
  return ((1u << (G & 31)) - 1u) << (((threadIdx.x & 31) / G) * G);
}

template <int H>
__device__ __forceinline__ unsigned slot_of(int key) {
  return (((unsigned)key * 2654435761u) >> 7) & (unsigned)(H - 1);
}
template <int H>
__device__ __forceinline__ int h_probe(const int *keys, const short *spos, int key) {
  unsigned h = slot_of<H>(key);
  while (true) {
    int k = keys[h];
    if (k == key) return spos[h];
    if (k == -1) return -1;
    h = (h + 1) & (unsigned)(H - 1);
  }
}
template <int H>
__device__ __forceinline__ void h_insert(int *keys, short *spos, int key, int pos) {  // key is known to be absent
  unsigned h = slot_of<H>(key);
  while (true) {
    int old = atomicCAS(keys + h, -1, key);
    if (old == -1) { spos[h] = (short)pos; return; }
    h = (h + 1) & (unsigned)(H - 1);
  }
}

// group-relative ballot: bit t = predicate of lane t of this group
template <int G>
__device__ __forceinline__ unsigned gballot(unsigned gm, int first_lane, bool pred) {
  unsigned b = __ballot_sync(gm, pred) & gm;
  return b >> first_lane;
}

// lraw: |C-hat_i| before truncation; -1 = not computed yet (the row did not fit the tiers tried so far).
// MODE 0 (count):  rows with lraw == -1; discovery only (valid when trunc_factor == 0); writes lraw / len.
// MODE 1 (fill):   rows with lo < lraw <= CAP; writes the row at Pptr[i].
// MODE 2 (single): rows with lraw == -1; whole row in one pass, written at i * pmax (needs pmax > 0: the truncated row has
//                  at most pmax entries), lraw / len recorded; a compaction kernel builds the CSR afterwards.
// Scf[k] = cf[Scol[k]] (signed char), precomputed so that a strength row and the markers of its columns arrive together.
//
// Latency structure: the strong neighbours of i (discovery) and the entries of row i (weights) are loaded G at a time, one
// per lane, together with everything that only depends on them (marker, row extents, hash position); the loops that fix
// the operation order then run over shuffled registers, and the first G entries of the NEXT neighbour row are requested
// before the current one is processed.
template <int G, int H, int CAP, int MODE>
__global__ void __launch_bounds__(256, 5) k_interp_grp(rows::Csr A, const int *__restrict__ Sptr, const int *__restrict__ Scol,
                                                    const signed char *__restrict__ Scf, const int *__restrict__ cf,
                                                    const int *__restrict__ f2c, int n_own, int n_rows, double tf, int pmax, int lo,
                                                    int *lraw, int *len, const int *__restrict__ Pptr, int *Pcol, double *Pval,
                                                    const int *__restrict__ row_list) {
  constexpr int GROUPS = 256 / G;
  constexpr bool COUNT = (MODE == 0);
  __shared__ int s_keys[GROUPS][H];
  __shared__ short s_pos[GROUPS][H];
  __shared__ int s_lc[GROUPS][CAP];
  __shared__ double s_lv[GROUPS][CAP];
  const int g = threadIdx.x / G, lane = threadIdx.x % G;
  const int slot = blockIdx.x * GROUPS + g;
  if (slot >= n_rows) return;
  const int i = row_list ? row_list[slot] : slot;   // n_rows = length of the list when one is given
  const unsigned gm = group_mask<G>();
  const unsigned lane_lt = gm & ((1u << (threadIdx.x & 31)) - 1u);
  const int first_lane = (threadIdx.x & 31) - lane;
  if (MODE != 1) {
    if (lraw[i] != -1) return;
  } else {
    int l = lraw[i];
    if (!(l > lo && l <= CAP)) return;
  }
  const int cfi = cf[i];
  if (cfi >= 0) {
    if (lane == 0) {
      if (MODE != 1) { lraw[i] = 0; len[i] = 1; }
      if (MODE != 0) { size_t b = (MODE == 1) ? (size_t)Pptr[i] : (size_t)i * pmax; Pcol[b] = f2c[i]; Pval[b] = 1.0; }
    }
    return;
  }
  if (cfi == -3) {
    if (MODE != 1 && lane == 0) { lraw[i] = 0; len[i] = 0; }
    return;
  }
  if (MODE == 1 && Pptr[i + 1] == Pptr[i]) return;
  int *keys = s_keys[g];
  short *spos = s_pos[g];
  int *lc = s_lc[g];
  double *lv = s_lv[g];
  for (int s = lane; s < H; s += G) keys[s] = -1;
  __syncwarp(gm);

  // ---- discovery of C-hat_i ---------------------------------------------------------------------------------------
  int L = 0;
  bool any_off = false, ovf = false;
  const int sb = Sptr[i], se = Sptr[i + 1];
  for (int cb = sb; cb < se && !ovf; cb += G) {
    const int kk = cb + lane;
    int i1_l = -1, c1_l = -3, b2_l = 0, e2_l = 0;
    if (kk < se) {
      i1_l = Scol[kk];
      c1_l = Scf[kk];
      if (c1_l < 0 && c1_l != -3) { b2_l = Sptr[i1_l]; e2_l = Sptr[i1_l + 1]; }
    }
    const int nstep = min(G, se - cb);
    const unsigned fm = gballot<G>(gm, first_lane, c1_l < 0 && c1_l != -3);   // strong F neighbours of this chunk
    int tnext = fm ? (__ffs(fm) - 1) : -1;
    int b2n = 0, e2n = 0, k1n = -1, cfkn = 0;
    if (tnext >= 0) {
      b2n = __shfl_sync(gm, b2_l, tnext, G);
      e2n = __shfl_sync(gm, e2_l, tnext, G);
      if (b2n + lane < e2n) { k1n = Scol[b2n + lane]; cfkn = Scf[b2n + lane]; }
    }
    for (int t = 0; t < nstep && !ovf; ++t) {
      const int c1 = __shfl_sync(gm, c1_l, t, G);
      if (c1 >= 0) {
        const int i1 = __shfl_sync(gm, i1_l, t, G);
        if (h_probe<H>(keys, spos, i1) < 0) {
          if (L == CAP) { ovf = true; break; }
          __syncwarp(gm);  // every lane has finished probing before the table changes
          if (lane == 0) { h_insert<H>(keys, spos, i1, L); lc[L] = i1; }
          any_off |= (i1 >= n_own);
          ++L;
          __syncwarp(gm);
        }
      } else if (c1 != -3) {
        const int b2 = b2n, e2 = e2n;
        int k1 = k1n, cfk = cfkn;
        const unsigned rest = (t >= 31) ? 0u : (fm & ~((2u << t) - 1u));
        tnext = rest ? (__ffs(rest) - 1) : -1;
        k1n = -1;
        if (tnext >= 0) {   // request the next F neighbour's strength row before working on this one
          b2n = __shfl_sync(gm, b2_l, tnext, G);
          e2n = __shfl_sync(gm, e2_l, tnext, G);
          if (b2n + lane < e2n) { k1n = Scol[b2n + lane]; cfkn = Scf[b2n + lane]; }
        }
        for (int base = b2; base < e2; base += G) {
          if (base != b2) {
            k1 = -1;
            if (base + lane < e2) { k1 = Scol[base + lane]; cfk = Scf[base + lane]; }
          }
          bool isnew = false;
          if (k1 >= 0 && cfk >= 0) isnew = (h_probe<H>(keys, spos, k1) < 0);
          const unsigned nm = __ballot_sync(gm, isnew) & gm;
          const int cnt = __popc(nm);
          if (cnt) {
            if (L + cnt > CAP) { ovf = true; break; }
            if (isnew) {
              int pos = L + __popc(nm & lane_lt);
              h_insert<H>(keys, spos, k1, pos);
              lc[pos] = k1;
            }
            any_off |= ((__ballot_sync(gm, isnew && k1 >= n_own) & gm) != 0u);
            L += cnt;
            __syncwarp(gm);
          }
        }
      }
    }
  }
  if (ovf) return;  // lraw stays -1: the next tier takes the row
  const bool full = !COUNT || tf != 0.0;
  if (!full) {
    if (lane == 0) { lraw[i] = L; len[i] = (pmax > 0 && L > pmax) ? pmax : L; }
    return;
  }
  if (any_off) {  // stable partition: own-range C points first (P_diag before P_offd); positions change -> rebuild the table
    int *tmp = reinterpret_cast<int *>(lv);
    int nown = 0;
    for (int base = 0; base < L; base += G) {
      int k = base + lane;
      bool own = (k < L) && (lc[k] < n_own);
      nown += __popc(__ballot_sync(gm, own) & gm);
    }
    int co = 0, cx = 0;
    for (int base = 0; base < L; base += G) {
      int k = base + lane;
      bool valid = k < L;
      int v = valid ? lc[k] : 0;
      bool own = valid && v < n_own, off = valid && !own;
      unsigned mo = __ballot_sync(gm, own) & gm, mx = __ballot_sync(gm, off) & gm;
      if (own) tmp[co + __popc(mo & lane_lt)] = v;
      if (off) tmp[nown + cx + __popc(mx & lane_lt)] = v;
      co += __popc(mo);
      cx += __popc(mx);
    }
    __syncwarp(gm);
    for (int s = lane; s < H; s += G) keys[s] = -1;
    __syncwarp(gm);
    for (int k = lane; k < L; k += G) { int v = tmp[k]; lc[k] = v; }
    __syncwarp(gm);
    for (int k = lane; k < L; k += G) h_insert<H>(keys, spos, lc[k], k);
    __syncwarp(gm);
  }

  // ---- weights --------------------------------------------------------------------------------------------------------
  for (int k = lane; k < L; k += G) lv[k] = 0.0;
  __syncwarp(gm);
  const int ab = A.ptr[i], ae = A.ptr[i + 1];
  double diagonal = A.val[ab];
  for (int cb = ab + 1; cb < ae; cb += G) {
    const int jj = cb + lane;
    const bool valid = jj < ae;
    int i1_l = -1, pos_l = -1, kind_l = 3, b1_l = 0, e1_l = 0;
    double aij_l = 0.0;
    if (valid) { i1_l = A.col[jj]; aij_l = A.val[jj]; }
    // strong_l: is the column of this entry in the strength row of i (S rows are subsequences of A rows)
    bool strong_l = false;
    for (int scb = sb; scb < se; scb += G) {
      const int s_l = (scb + lane < se) ? Scol[scb + lane] : -1;
      const int ns = min(G, se - scb);
      for (int u = 0; u < ns; ++u) strong_l |= (__shfl_sync(gm, s_l, u, G) == i1_l);
    }
    if (valid) {
      pos_l = h_probe<H>(keys, spos, i1_l);
      if (pos_l >= 0) kind_l = 0;                       // direct: a_ij goes to its own C point
      else {
        const int c1 = cf[i1_l];
        if (strong_l && c1 < 0 && c1 != -3) { kind_l = 1; b1_l = A.ptr[i1_l]; e1_l = A.ptr[i1_l + 1]; }   // distribute over C-hat
        else if (c1 != -3) kind_l = 2;                  // lumped into the diagonal
      }
    }
    const int nstep = min(G, ae - cb);
    const unsigned fm = gballot<G>(gm, first_lane, kind_l == 1);
    int tnext = fm ? (__ffs(fm) - 1) : -1;
    int b1n = 0, e1n = 0, i2n = -1;
    double an = 0.0;
    if (tnext >= 0) {
      b1n = __shfl_sync(gm, b1_l, tnext, G);
      e1n = __shfl_sync(gm, e1_l, tnext, G);
      if (b1n + lane < e1n) { i2n = A.col[b1n + lane]; an = A.val[b1n + lane]; }
    }
    for (int t = 0; t < nstep; ++t) {
      const int kind = __shfl_sync(gm, kind_l, t, G);
      const double aij = __shfl_sync(gm, aij_l, t, G);
      if (kind == 0) {
        const int pos = __shfl_sync(gm, pos_l, t, G);
        if (lane == 0) lv[pos] = lv[pos] + aij;
        __syncwarp(gm);
      } else if (kind == 1) {
        const int b1 = b1n, e1 = e1n;
        int i2 = i2n;
        double a = an;
        const unsigned rest = (t >= 31) ? 0u : (fm & ~((2u << t) - 1u));
        tnext = rest ? (__ffs(rest) - 1) : -1;
        i2n = -1;
        if (tnext >= 0) {   // request the next neighbour row before working on this one
          b1n = __shfl_sync(gm, b1_l, tnext, G);
          e1n = __shfl_sync(gm, e1_l, tnext, G);
          if (b1n + lane < e1n) { i2n = A.col[b1n + lane]; an = A.val[b1n + lane]; }
        }
        // lane 0 of the first chunk holds the neighbour's diagonal entry
        const double dsgn = __shfl_sync(gm, a, 0, G);
        const double sgn = (dsgn < 0) ? -1.0 : 1.0;
        const bool one_chunk = (e1 - b1) <= G;
        double sum = 0.0;
        int p2 = -1;
        bool is_i = false, m = false;
        for (int base = b1; base < e1; base += G) {
          const int j1 = base + lane;
          if (base != b1) { i2 = -1; a = 0.0; if (j1 < e1) { i2 = A.col[j1]; a = A.val[j1]; } }
          m = false; p2 = -1; is_i = false;
          if (j1 < e1 && j1 > b1 && (sgn * a) < 0) {
            is_i = (i2 == i);
            p2 = h_probe<H>(keys, spos, i2);
            m = is_i || (p2 >= 0);
          }
          unsigned mm = __ballot_sync(gm, m) & gm;
          while (mm) {  // storage-order sum of the contributing entries
            int src = __ffs(mm) - 1;
            mm &= mm - 1;
            sum = sum + __shfl_sync(gm, a, src);
          }
        }
        if (sum != 0.0) {
          const double distribute = aij / sum;
          for (int base = b1; base < e1; base += G) {
            const int j1 = base + lane;
            if (!one_chunk) {
              i2 = -1; a = 0.0; m = false; p2 = -1; is_i = false;
              if (j1 < e1) { i2 = A.col[j1]; a = A.val[j1]; }
              if (j1 < e1 && j1 > b1 && (sgn * a) < 0) {
                is_i = (i2 == i);
                p2 = h_probe<H>(keys, spos, i2);
                m = is_i || (p2 >= 0);
              }
            }
            double tt = 0.0;
            if (m) {
              tt = distribute * a;
              if (p2 >= 0) lv[p2] = lv[p2] + tt;
            }
            unsigned mi = __ballot_sync(gm, m && is_i) & gm;
            if (mi) diagonal = diagonal + __shfl_sync(gm, tt, __ffs(mi) - 1);
          }
        } else {
          diagonal = diagonal + aij;
        }
        __syncwarp(gm);
      } else if (kind == 2) {
        diagonal = diagonal + aij;
      }
    }
  }
  __syncwarp(gm);
  if (diagonal != 0.0) {
    const double md = -diagonal;
    for (int k = lane; k < L; k += G) lv[k] = lv[k] / md;
  }
  __syncwarp(gm);

  // ---- truncation (sequential on the first lane: the sums and the sort are order dependent) -----------------------------
  int n_out = L;
  if (lane == 0) {
    if (tf > 0.0 && n_out > 0) {
      double max_coef = 0.0, row_sum = 0.0, scale = 0.0;
      for (int k = 0; k < n_out; ++k) { double a = fabs(lv[k]); max_coef = (a > max_coef) ? a : max_coef; }
      max_coef = max_coef * tf;
      int w = 0;
      for (int k = 0; k < n_out; ++k) {
        row_sum = row_sum + lv[k];
        if (fabs(lv[k]) >= max_coef) { scale = scale + lv[k]; lc[w] = lc[k]; lv[w] = lv[k]; ++w; }
      }
      n_out = w;
      if (scale != 0.0 && scale != row_sum) {
        double f = row_sum / scale;
        for (int k = 0; k < n_out; ++k) lv[k] = lv[k] * f;
      }
    }
    if (pmax > 0 && n_out > pmax) {
      double row_sum = 0.0, scale = 0.0;
      for (int k = 0; k < n_out; ++k) row_sum = row_sum + lv[k];
      rows::qsort2abs(lc, lv, n_out, pmax);
      for (int k = 0; k < pmax; ++k) scale = scale + lv[k];
      n_out = pmax;
      if (scale != 0.0 && scale != row_sum) {
        double f = row_sum / scale;
        for (int k = 0; k < n_out; ++k) lv[k] = lv[k] * f;
      }
      if (any_off) {  // re-split, each part keeping the sorted order
        for (int k = 1; k < n_out; ++k) {
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
  }
  __syncwarp(gm);
  n_out = __shfl_sync(gm, n_out, first_lane);
  if (MODE != 1) {
    if (lane == 0) { lraw[i] = L; len[i] = n_out; }
    if (MODE == 0) return;
  }
  const size_t b = (MODE == 1) ? (size_t)Pptr[i] : (size_t)i * pmax;
  for (int k = lane; k < n_out; k += G) {
    Pcol[b + k] = f2c[lc[k]];
    Pval[b + k] = lv[k];
  }
}

// rows that did not fit any tier: the sequential row function with a global-memory scratch list of `cap` entries per row
__global__ void k_interp_ovf(rows::Csr A, const int *Sptr, const int *Scol, const int *cf, const int *f2c, int n_own, double tf,
                             int pmax, const int *rows_list, int nrows, int *sc, double *sv, int cap, int *len, const int *Pptr,
                             int *Pcol, double *Pval, int *err) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nrows) return;
  int i = rows_list[t];
  int *lc = sc + (size_t)t * cap;
  double *lv = sv + (size_t)t * cap;
  if (Pptr == nullptr) {
    int L = rows::interp_row(A, Sptr, Scol, cf, f2c, n_own, i, tf, pmax, tf == 0.0, lc, lv, lc, lv, cap);
    if (L < 0) { *err = 1; L = 0; }
    len[i] = L;
  } else {
    int b = Pptr[i];
    if (Pptr[i + 1] > b) rows::interp_row(A, Sptr, Scol, cf, f2c, n_own, i, tf, pmax, false, Pcol + b, Pval + b, lc, lv, cap);
  }
}
__global__ void k_collect_todo(int n, const int *lraw, int *count, int *list, int cap) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && lraw[i] == -1) {
    int p = atomicAdd(count, 1);
    if (p < cap) list[p] = i;
  }
}

template <class T>
T fetch1(const T *dev) {
  T h;
  B200_CUDA(cudaMemcpyAsync(&h, dev, sizeof(T), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  return h;
}

__global__ void k_scf(i64 nnz, const int *__restrict__ Scol, const int *__restrict__ cf, signed char *__restrict__ Scf) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nnz) Scf[k] = (signed char)cf[Scol[k]];
}
// single-pass rows live at i * pmax in the scratch arrays: copy them to their CSR position
__global__ void k_compact_p(int n_rows, int pmax, const int *__restrict__ lraw, const int *__restrict__ Pptr, const int *__restrict__ tcol,
                            const double *__restrict__ tval, int *__restrict__ Pcol, double *__restrict__ Pval) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows || lraw[i] == -1) return;
  const int b = Pptr[i], n = Pptr[i + 1] - b;
  const size_t src = (size_t)i * pmax;
  for (int k = 0; k < n; ++k) { Pcol[b + k] = tcol[src + k]; Pval[b + k] = tval[src + k]; }
}

bool use_v1() {
  static int v = -1;
  if (v < 0) { const char *e = getenv("B200_INTERP_V1"); v = (e && atoi(e) != 0) ? 1 : 0; }
  return v == 1;
}

// tiers: (G, H, CAP, STG) = (8, 64, 32, 64) when strength rows are short, then (32, 256, 128, 192)
template <int G, int H, int CAP, int STG, int MODE>
void launch_flat(int grid_rows, const rows::Csr &A, const int *Sptr, const int *Scol, const signed char *Scf, const int *cf, const int *f2c,
                 int n_own, int n_rows, double tf, int pmax, int lo, int *lraw, int *len, const int *Pptr, int *Pcol, double *Pval,
                 const int *list) {
  auto kern = flat::k_interp_flat<G, H, CAP, STG, MODE>;
  constexpr size_t smem = sizeof(flat::Grp<G, H, CAP, STG>) * (256 / G);
  static bool attr_set = false;   // per instantiation
  if (!attr_set) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  static int timing = -1;
  if (timing < 0) { const char *e = getenv("B200_INTERP_TIMING"); timing = (e && atoi(e) != 0) ? 1 : 0; }
  DVec<unsigned long long> ts;
  if (timing) { ts.alloc(8); ts.zero(); }
  // MODE 2: rows that need truncation are handed to k_interp_trunc through a scratch area of CAP entries per row; the rows are
  // processed in chunks so that the scratch stays at about 1.5 GB whatever the level size
  const int chunk = (MODE == 2) ? std::max(256, (int)((size_t)(128u << 20) / CAP) / 256 * 256) : std::max(grid_rows, 1);
  DVec<int> sc;
  DVec<double> sv;
  if (MODE == 2) {
    const size_t rows_in_chunk = (size_t)std::min(chunk, std::max(grid_rows, 1));
    sc.alloc(rows_in_chunk * CAP);
    sv.alloc(rows_in_chunk * CAP);
  }
  for (int slot0 = 0; slot0 < grid_rows; slot0 += chunk) {
    const int cnt = std::min(chunk, grid_rows - slot0);
    kern<<<div_up(cnt, 256 / G), 256, smem, stream()>>>(A, Sptr, Scol, Scf, cf, f2c, n_own, std::min(n_rows, slot0 + cnt), tf, pmax, lo, lraw,
                                                       len, Pptr, Pcol, Pval, list, slot0, sc.p, sv.p, ts.p);
    count_launch();
    if (MODE == 2) {
      flat::k_interp_trunc<<<div_up(cnt, 256), 256, 0, stream()>>>(slot0, std::min(n_rows, slot0 + cnt), list, CAP, sc.p, sv.p, lraw, len, f2c,
                                                                  n_own, tf, pmax, Pcol, Pval);
      count_launch();
    }
  }
  if (timing) {   // debugging aid: share of the warp-cycles of this launch per phase
    std::vector<unsigned long long> h = ts.to_host();
    double tot = 0;
    for (int k = 0; k < 6; ++k) tot += (double)h[k];
    fprintf(stderr, "[interp timing] G=%d MODE=%d rows=%d  discovery %.1f%%  weights %.1f%%  wait1 %.1f%%  trunc %.1f%%  wait2 %.1f%%  write %.1f%%  (%.3g warp-cycles)\n",
            G, MODE, grid_rows, 100 * h[0] / tot, 100 * h[1] / tot, 100 * h[2] / tot, 100 * h[3] / tot, 100 * h[4] / tot, 100 * h[5] / tot, tot);
  }
}
template <int MODE, bool SMALL>
void launch_one(int grid_rows, const rows::Csr &A, const int *Sptr, const int *Scol, const signed char *Scf, const int *cf, const int *f2c,
                int n_own, int n_rows, double tf, int pmax, int lo, int *lraw, int *len, const int *Pptr, int *Pcol, double *Pval,
                const int *list) {
  if (use_v1()) {
    if (SMALL)
      k_interp_grp<8, 64, 32, MODE><<<div_up(grid_rows, 32), 256, 0, stream()>>>(A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lo, lraw,
                                                                               len, Pptr, Pcol, Pval, list);
    else
      k_interp_grp<32, 256, 128, MODE><<<div_up(grid_rows, 8), 256, 0, stream()>>>(A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lo, lraw,
                                                                                 len, Pptr, Pcol, Pval, list);
    count_launch();
    return;
  }
  if (SMALL) launch_flat<8, 64, 32, 64, MODE>(grid_rows, A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lo, lraw, len, Pptr, Pcol, Pval, list);
  else launch_flat<32, 256, 128, 192, MODE>(grid_rows, A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lo, lraw, len, Pptr, Pcol, Pval, list);
}

template <int MODE>
void launch_tiers(bool small_tier, const rows::Csr &A, const int *Sptr, const int *Scol, const signed char *Scf, const int *cf,
                  const int *f2c, int n_own, int n_rows, double tf, int pmax, int *lraw, int *len, const int *Pptr, int *Pcol,
                  double *Pval) {
  int lo = -1;
  if (small_tier) {
    launch_one<MODE, true>(n_rows, A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lo, lraw, len, Pptr, Pcol, Pval, nullptr);
    lo = 32;
    if (MODE != 1) {
      // the rows the small tier could not hold are few (often none): run the large tier on their list only
      constexpr int CAP = 1 << 20;
      DVec<int> cnt(1), list((size_t)CAP);
      cnt.zero();
      k_collect_todo<<<div_up(n_rows, 256), 256, 0, stream()>>>(n_rows, lraw, cnt.p, list.p, CAP);
      count_launch();
      const int left = fetch1(cnt.p);
      if (left == 0) return;
      if (left <= CAP) {
        launch_one<MODE, false>(left, A, Sptr, Scol, Scf, cf, f2c, n_own, left, tf, pmax, lo, lraw, len, Pptr, Pcol, Pval, list.p);
        B200_KERNEL_CHECK();
        return;
      }
    }
  }
  launch_one<MODE, false>(n_rows, A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lo, lraw, len, Pptr, Pcol, Pval, nullptr);
  B200_KERNEL_CHECK();
}

}  // namespace

// Interpolation rows [0, n_rows) of the (extended) level: A, S, cf, f2c are indexed by extended fine ids, ids below
// n_own are this rank's own points; f2c gives the output column of a C point.  P.nc must be set by the caller.
// With P_max_elmts > 0 every truncated row has at most pmax entries, so the tiers run ONCE writing fixed-stride rows that
// are compacted afterwards; otherwise a counting pass (discovery only) precedes the fill pass.
void build_interp(const DCsr &Aext, const int *Sptr, const int *Scol, i64 S_nnz, const int *cf, const int *f2c, int n_own, int n_rows,
                  double tf, int pmax, DCsr &P) {
  rows::Csr A{Aext.nr, Aext.nc, Aext.ptr.p, Aext.col.p, Aext.val.p};
  P.nr = n_rows;
  P.ptr.alloc((size_t)n_rows + 1);
  if (n_rows == 0) { P.nnz = 0; P.ptr.zero(); return; }
  bool small_tier = (double)S_nnz <= 8.0 * (double)Aext.nr;
  if (const char *e = getenv("B200_INTERP_SMALL_TIER")) small_tier = small_tier && atoi(e) != 0;   // benchmarking override
  const bool single = pmax > 0;
  DVec<signed char> Scf((size_t)std::max<i64>(S_nnz, 1));
  if (S_nnz > 0) {
    k_scf<<<div_up(S_nnz, 256), 256, 0, stream()>>>(S_nnz, Scol, cf, Scf.p);
    count_launch();
  }
  DVec<int> lraw((size_t)n_rows), len((size_t)n_rows + 1);
  B200_CUDA(cudaMemsetAsync(lraw.p, 0xFF, sizeof(int) * (size_t)n_rows, stream()));
  B200_CUDA(cudaMemsetAsync(len.p + n_rows, 0, sizeof(int), stream()));
  DVec<int> tcol;
  DVec<double> tval;
  // algorithmic bytes of the row kernels: A and S read once, markers, and the (at most pmax wide) rows of P written once
  {
  ProfScope prof("interp_rows", 12.0 * (double)Aext.nnz + 4.0 * (double)S_nnz + 5.0 * (double)S_nnz / 4.0 + 16.0 * (double)Aext.nr +
                                    12.0 * (double)n_rows * (pmax > 0 ? pmax : 4) * 0.75);
  if (single) {
    tcol.alloc((size_t)n_rows * pmax);
    tval.alloc((size_t)n_rows * pmax);
    launch_tiers<2>(small_tier, A, Sptr, Scol, Scf.p, cf, f2c, n_own, n_rows, tf, pmax, lraw.p, len.p, nullptr, tcol.p, tval.p);
  } else {
    launch_tiers<0>(small_tier, A, Sptr, Scol, Scf.p, cf, f2c, n_own, n_rows, tf, pmax, lraw.p, len.p, nullptr, nullptr, nullptr);
  }
  }
  // rows beyond the last tier
  DVec<int> todo_count(1), todo((size_t)65536), err(1);
  todo_count.zero();
  err.zero();
  k_collect_todo<<<div_up(n_rows, 256), 256, 0, stream()>>>(n_rows, lraw.p, todo_count.p, todo.p, 65536);
  count_launch();
  int novf = fetch1(todo_count.p);
  B200_REQUIRE(novf <= 65536, "interpolation: too many rows exceed the shared-memory C-hat capacity");
  const int ocap = 4096;
  DVec<int> sc;
  DVec<double> sv;
  if (novf > 0) {
    sc.alloc((size_t)novf * ocap);
    sv.alloc((size_t)novf * ocap);
    k_interp_ovf<<<div_up(novf, 64), 64, 0, stream()>>>(A, Sptr, Scol, cf, f2c, n_own, tf, pmax, todo.p, novf, sc.p, sv.p, ocap, len.p,
                                                        nullptr, nullptr, nullptr, err.p);
    count_launch();
    B200_REQUIRE(fetch1(err.p) == 0, "interpolation: C-hat larger than 4096 entries");
  }
  {
    size_t tb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, len.p, P.ptr.p, n_rows + 1, stream());
    DVec<char> tmp(tb);
    cub::DeviceScan::ExclusiveSum(tmp.p, tb, len.p, P.ptr.p, n_rows + 1, stream());
    count_launch(2);
  }
  P.nnz = checked_total(len.p, n_rows, "build_interp");
  P.col.alloc((size_t)std::max<i64>(P.nnz, 1));
  P.val.alloc((size_t)std::max<i64>(P.nnz, 1));
  if (single) {
    k_compact_p<<<div_up(n_rows, 256), 256, 0, stream()>>>(n_rows, pmax, lraw.p, P.ptr.p, tcol.p, tval.p, P.col.p, P.val.p);
    count_launch();
  } else {
    launch_tiers<1>(small_tier, A, Sptr, Scol, Scf.p, cf, f2c, n_own, n_rows, tf, pmax, lraw.p, len.p, P.ptr.p, P.col.p, P.val.p);
  }
  if (novf > 0) {
    k_interp_ovf<<<div_up(novf, 64), 64, 0, stream()>>>(A, Sptr, Scol, cf, f2c, n_own, tf, pmax, todo.p, novf, sc.p, sv.p, ocap, len.p,
                                                        P.ptr.p, P.col.p, P.val.p, err.p);
    count_launch();
  }
  B200_KERNEL_CHECK();
}

}  // namespace b200
