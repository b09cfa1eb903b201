// amg_interp.cu -- extended+i interpolation (interp_type 6) with truncation: tier driver; classical interpolation (interp_type 0).
// Replaces hypre_BoomerAMGBuildExtPIInterp + hypre_BoomerAMGInterpTruncation reached from HYPRE_BoomerAMGSetup
// (/root/reference/src/HypreSystem.cpp:197,212,232 set the options, :323 via :692 runs it); formula and entry order:
// SURVEY.md A.5.  hypre itself is not in /root/reference.
//
// rows::interp_row (amg_rows.cuh) is the sequential definition: C-hat_i is discovered in the order "strong neighbours of i in
// storage order, and for a strong F neighbour its strong C neighbours in storage order"; every weight is a sum whose terms
// arrive in (entry of row i, entry of the neighbour row) order.  The group kernel k_interp_flat (amg_interp_flat.cuh, design
// notes there and in DESIGN.md section 4.0) produces bit-identical rows with G lanes per row.  Rows are tiered by |C-hat| and
// hash load: (G, capacity) = (8, 32) when strength rows are short, (16, 64) when matrix rows average at most 40 entries, then
// (32, 128); a row that overflows its tier is taken by the next, and the few rows beyond the last run rows::interp_row itself with
// a global-memory scratch list.  The second half of the file is classical modified interpolation (interp_type 0).  Compiled with -fmad=false (the operation order of every value
// that feeds a discrete decision is part of the parity contract).
#include "amg.h"
#include "amg_rows.cuh"
#include <cub/cub.cuh>
#include <cstdlib>

namespace b200 {

namespace {

#include "amg_interp_flat.cuh"

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

// C-compressed copies (flat::CComp): for every non-isolated F row the off-diagonal entries of A, and the entries of S, whose
// column is a C point.  FILL = false counts (cnt arrays of n + 1 entries, the last one zero), FILL = true writes.
template <bool FILL>
__global__ void k_ccomp(rows::Csr A, const int *__restrict__ Sptr, const int *__restrict__ Scol, const signed char *__restrict__ Scf,
                        const int *__restrict__ cf, int *ac_cnt, int *sc_cnt, const int *__restrict__ acptr, int *accol, int *acpos,
                        double *acval, const int *__restrict__ scptr, int *sccol) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > A.nr) return;
  if (i == A.nr) { if (!FILL) { ac_cnt[i] = 0; sc_cnt[i] = 0; } return; }
  const int c = cf[i];
  int na = 0, ns = 0;
  if (c < 0 && c != -3) {   // only rows of F points are ever walked as a neighbour row
    int wa = FILL ? acptr[i] : 0, ws = FILL ? scptr[i] : 0;
    for (int k = A.ptr[i] + 1; k < A.ptr[i + 1]; ++k) {
      const int col = A.col[k];
      if (cf[col] >= 0) {
        if (FILL) { accol[wa] = col; acpos[wa] = k; acval[wa] = A.val[k]; ++wa; }
        ++na;
      }
    }
    for (int k = Sptr[i]; k < Sptr[i + 1]; ++k)
      if (Scf[k] >= 0) {
        if (FILL) sccol[ws++] = Scol[k];
        ++ns;
      }
  }
  if (!FILL) { ac_cnt[i] = na; sc_cnt[i] = ns; }
}
template <class T>
void exscan_int(const T *in, T *out, i64 n) {
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, (int)n, stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, in, out, (int)n, stream());
  count_launch(2);
}
flat::CComp g_cc{};   // compressed copies of the level being interpolated (set by build_interp around the tier launches)

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

// tiers: (G, H, CAP, STG) = (8, 64, 32, 48) when strength rows are short, then (32, 256, 128, 192)
template <int G, int H, int CAP, int STG, int MODE>
void launch_flat(int grid_rows, const rows::Csr &A, const int *Sptr, const int *Scol, const signed char *Scf, const int *cf, const int *f2c,
                 int n_own, int n_rows, double tf, int pmax, int lo, int *lraw, int *len, const int *Pptr, int *Pcol, double *Pval,
                 const int *list) {
  auto kern = flat::k_interp_flat<G, H, CAP, STG, MODE>;
  constexpr size_t smem = sizeof(flat::Grp<G, H, CAP, STG>) * (flat::BT / G);
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
    kern<<<div_up(cnt, flat::BT / G), flat::BT, smem, stream()>>>(A, Sptr, Scol, Scf, g_cc, cf, f2c, n_own, std::min(n_rows, slot0 + cnt), tf, pmax, lo, lraw,
                                                       len, Pptr, Pcol, Pval, list, slot0, sc.p, sv.p, ts.p);
    count_launch();
    if (MODE == 2) {
      flat::k_interp_trunc<<<div_up(cnt, flat::TRUNC_THREADS), flat::TRUNC_THREADS, 0, stream()>>>(slot0, std::min(n_rows, slot0 + cnt), list, CAP, sc.p, sv.p, lraw, len, f2c,
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
// tiers (G lanes per row, hash slots, C-hat capacity, staging): rows that exceed a tier (C-hat or hash load) go to the next one
//   8  : <8, 64, 32, 48>      4 rows per warp  -- short strength rows (7-point level 0)
//   16 : <16, 128, 64, 96>    2 rows per warp  -- rows of up to ~40 entries (the coarser levels of the 7-point hierarchy: 24 - 30
//                                                 entries and ~50 compressed items per row keep at most 16 lanes busy)
//   32 : <32, 256, 128, 192>  1 row per warp   -- everything else
template <int MODE>
void launch_one(int tier, int grid_rows, const rows::Csr &A, const int *Sptr, const int *Scol, const signed char *Scf, const int *cf,
                const int *f2c, int n_own, int n_rows, double tf, int pmax, int *lraw, int *len, const int *Pptr, int *Pcol, double *Pval,
                const int *list) {
  const int lo = -1;   // (unused by the kernels since the fill pass marks filled rows)
  if (tier == 8) launch_flat<8, 64, 32, 48, MODE>(grid_rows, A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lo, lraw, len, Pptr, Pcol, Pval, list);
  else if (tier == 16) launch_flat<16, 128, 64, 96, MODE>(grid_rows, A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lo, lraw, len, Pptr, Pcol, Pval, list);
  else launch_flat<32, 256, 128, 192, MODE>(grid_rows, A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lo, lraw, len, Pptr, Pcol, Pval, list);
}

template <int MODE>
void launch_tiers(int first_tier, const rows::Csr &A, const int *Sptr, const int *Scol, const signed char *Scf, const int *cf,
                  const int *f2c, int n_own, int n_rows, double tf, int pmax, int *lraw, int *len, const int *Pptr, int *Pcol,
                  double *Pval) {
  if (first_tier != 32) {
    launch_one<MODE>(first_tier, n_rows, A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lraw, len, Pptr, Pcol, Pval, nullptr);
    if (MODE != 1) {
      // the rows the first tier could not hold are few (often none): run the large tier on their list only
      constexpr int CAP = 1 << 20;
      DVec<int> cnt(1), list((size_t)CAP);
      cnt.zero();
      k_collect_todo<<<div_up(n_rows, 256), 256, 0, stream()>>>(n_rows, lraw, cnt.p, list.p, CAP);
      count_launch();
      const int left = fetch1(cnt.p);
      if (left == 0) return;
      if (left <= CAP) {
        launch_one<MODE>(32, left, A, Sptr, Scol, Scf, cf, f2c, n_own, left, tf, pmax, lraw, len, Pptr, Pcol, Pval, list.p);
        B200_KERNEL_CHECK();
        return;
      }
    }
  }
  launch_one<MODE>(32, n_rows, A, Sptr, Scol, Scf, cf, f2c, n_own, n_rows, tf, pmax, lraw, len, Pptr, Pcol, Pval, nullptr);
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
  // first tier from the mean lengths of the strength rows and of the matrix rows (see launch_one)
  int first_tier = 32;
  if ((double)S_nnz <= 8.0 * (double)Aext.nr) first_tier = 8;
  else if ((double)Aext.nnz <= 40.0 * (double)Aext.nr) first_tier = 16;
  if (const char *e = getenv("B200_INTERP_MID_TIER")) {   // benchmarking override: 0 = no 16-lane tier
    if (atoi(e) == 0 && first_tier == 16) first_tier = 32;
  }
  const bool single = pmax > 0;
  DVec<signed char> Scf((size_t)std::max<i64>(S_nnz, 1));
  if (S_nnz > 0) {
    k_scf<<<div_up(S_nnz, 256), 256, 0, stream()>>>(S_nnz, Scol, cf, Scf.p);
    count_launch();
  }
  // C-compressed copies of A and S for the row kernels (amg_interp_flat.cuh)
  DVec<int> acptr((size_t)Aext.nr + 1), scptr((size_t)Aext.nr + 1), accol, acpos, sccol;
  DVec<double> acval;
  {
    DVec<int> ac_cnt((size_t)Aext.nr + 1), sc_cnt((size_t)Aext.nr + 1);
    const int g1 = div_up((i64)Aext.nr + 1, 256);
    k_ccomp<false><<<g1, 256, 0, stream()>>>(A, Sptr, Scol, Scf.p, cf, ac_cnt.p, sc_cnt.p, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
    count_launch();
    exscan_int(ac_cnt.p, acptr.p, (i64)Aext.nr + 1);
    exscan_int(sc_cnt.p, scptr.p, (i64)Aext.nr + 1);
    int tot[2];
    B200_CUDA(cudaMemcpyAsync(&tot[0], acptr.p + Aext.nr, sizeof(int), cudaMemcpyDeviceToHost, stream()));
    B200_CUDA(cudaMemcpyAsync(&tot[1], scptr.p + Aext.nr, sizeof(int), cudaMemcpyDeviceToHost, stream()));
    B200_CUDA(cudaStreamSynchronize(stream()));
    accol.alloc((size_t)std::max(tot[0], 1));
    acpos.alloc((size_t)std::max(tot[0], 1));
    acval.alloc((size_t)std::max(tot[0], 1));
    sccol.alloc((size_t)std::max(tot[1], 1));
    k_ccomp<true><<<g1, 256, 0, stream()>>>(A, Sptr, Scol, Scf.p, cf, nullptr, nullptr, acptr.p, accol.p, acpos.p, acval.p, scptr.p, sccol.p);
    count_launch();
    g_cc = flat::CComp{acptr.p, accol.p, acpos.p, acval.p, scptr.p, sccol.p};
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
    launch_tiers<2>(first_tier, A, Sptr, Scol, Scf.p, cf, f2c, n_own, n_rows, tf, pmax, lraw.p, len.p, nullptr, tcol.p, tval.p);
  } else {
    launch_tiers<0>(first_tier, A, Sptr, Scol, Scf.p, cf, f2c, n_own, n_rows, tf, pmax, lraw.p, len.p, nullptr, nullptr, nullptr);
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
    launch_tiers<1>(first_tier, A, Sptr, Scol, Scf.p, cf, f2c, n_own, n_rows, tf, pmax, lraw.p, len.p, P.ptr.p, P.col.p, P.val.p);
  }
  if (novf > 0) {
    k_interp_ovf<<<div_up(novf, 64), 64, 0, stream()>>>(A, Sptr, Scol, cf, f2c, n_own, tf, pmax, todo.p, novf, sc.p, sv.p, ocap, len.p,
                                                        P.ptr.p, P.col.p, P.val.p, err.p);
    count_launch();
  }
  B200_KERNEL_CHECK();
}

// ---- classical "modified" interpolation (interp_type 0) ------------------------------------------------------------------------------
// Replaces hypre_BoomerAMGBuildInterp + hypre_BoomerAMGInterpTruncation, which the reference reaches when its input sets
// interp_type: 0 (/root/reference/src/HypreSystem.cpp:195-198; its sample input etc/hypre_app.yaml:41 does).  rows::classical_row is
// the definition.  A row only touches its own entries and the rows of its strong F neighbours, and its interpolatory set is a subset of
// its strength row, so one thread per row works in place in a scratch laid out like the strength pattern (entry k of row i at
// Sptr[i] + k): no capacity tiers, no overflow path.  The second kernel packs the rows.
namespace {

// CAP: strength rows of up to CAP entries build their row in thread-local arrays (registers / L1-resident local memory) and write
// it to the scratch once; with the list in global memory every `w_j += ...` of the walk is a dependent round trip to L2 (measured
// at 256^3: the phase took as long as the whole ext+i interpolation).  Longer rows work in place in the scratch.
template <int CAP>
__global__ void __launch_bounds__(128) k_classical_rows(rows::Csr A, const int *__restrict__ Sptr, const int *__restrict__ Scol,
                                                        const int *__restrict__ cf, const int *__restrict__ f2c, int n_own, int n_rows,
                                                        double tf, int pmax, int *sc, double *sv, int *len) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  int L;
  if (cf[i] >= 0) L = 1;
  else if (cf[i] == -3) L = 0;
  else {
    const int sb = Sptr[i];
    if (Sptr[i + 1] - sb <= CAP) {
      int lc[CAP];
      double lv[CAP];
      L = rows::classical_row(A, Sptr, Scol, cf, f2c, n_own, i, tf, pmax, sc + sb, sv + sb, lc, lv);
    } else {
      L = rows::classical_row(A, Sptr, Scol, cf, f2c, n_own, i, tf, pmax, sc + sb, sv + sb, sc + sb, sv + sb);
    }
  }
  len[i] = L;
}

__global__ void k_classical_pack(int n_rows, const int *__restrict__ Sptr, const int *__restrict__ cf, const int *__restrict__ f2c,
                                 const int *__restrict__ sc, const double *__restrict__ sv, const int *__restrict__ Pptr, int *Pcol,
                                 double *Pval) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  const int b = Pptr[i], L = Pptr[i + 1] - b;
  if (cf[i] >= 0) {
    Pcol[b] = f2c[i];
    Pval[b] = 1.0;
    return;
  }
  const int s = Sptr[i];
  for (int k = 0; k < L; ++k) {
    Pcol[b + k] = sc[s + k];
    Pval[b + k] = sv[s + k];
  }
}

}  // namespace

void build_interp_classical(const DCsr &Aext, const int *Sptr, const int *Scol, i64 S_nnz, const int *cf, const int *f2c, int n_own,
                            int n_rows, double tf, int pmax, DCsr &P) {
  rows::Csr A{Aext.nr, Aext.nc, Aext.ptr.p, Aext.col.p, Aext.val.p};
  P.nr = n_rows;
  P.ptr.alloc((size_t)n_rows + 1);
  if (n_rows == 0) { P.nnz = 0; P.ptr.zero(); return; }
  DVec<int> sc((size_t)std::max<i64>(S_nnz, 1)), len((size_t)n_rows + 1);
  DVec<double> sv((size_t)std::max<i64>(S_nnz, 1));
  B200_CUDA(cudaMemsetAsync(len.p + n_rows, 0, sizeof(int), stream()));
  {
    // algorithmic bytes: A and S read once, markers, the rows of P written once
    ProfScope prof("interp_rows", 12.0 * (double)Aext.nnz + 4.0 * (double)S_nnz + 16.0 * (double)Aext.nr + 12.0 * (double)S_nnz * 0.5);
    k_classical_rows<32><<<div_up(n_rows, 128), 128, 0, stream()>>>(A, Sptr, Scol, cf, f2c, n_own, n_rows, tf, pmax, sc.p, sv.p, len.p);
    count_launch();
  }
  {
    size_t tb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, len.p, P.ptr.p, n_rows + 1, stream());
    DVec<char> tmp(tb);
    cub::DeviceScan::ExclusiveSum(tmp.p, tb, len.p, P.ptr.p, n_rows + 1, stream());
    count_launch(2);
  }
  P.nnz = checked_total(len.p, n_rows, "build_interp_classical");
  P.col.alloc((size_t)std::max<i64>(P.nnz, 1));
  P.val.alloc((size_t)std::max<i64>(P.nnz, 1));
  k_classical_pack<<<div_up(n_rows, 128), 128, 0, stream()>>>(n_rows, Sptr, cf, f2c, sc.p, sv.p, P.ptr.p, P.col.p, P.val.p);
  count_launch();
  B200_KERNEL_CHECK();
}

}  // namespace b200
