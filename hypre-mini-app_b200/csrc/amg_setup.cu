// amg_setup.cu -- BoomerAMG setup entirely on the device (no CPU fallback): strength of connection, PMIS
// coarsening with hypre's seeded random measure, coarse numbering, extended+i or classical interpolation with truncation,
// transposition and the Galerkin product RAP as two two-pass hash SpGEMMs, smoother data, coarsest dense LU.
// Replaces HYPRE_BoomerAMGSetup (/root/reference/src/HypreSystem.cpp:323, reached through
// HYPRE_ParCSRGMRESSetup at :692); algorithm definitions: SURVEY.md Appendix A.3-A.9 (hypre is not in /root/reference).
//
// Compiled with -fmad=false: the operation order of every value that feeds a discrete decision is part of the
// parity contract with the CPU oracle.
//
// Interpolation lives in amg_interp.cu, transposition and the Galerkin SpGEMM in spgemm.cu.
#include "amg.h"
#include "amg_rows.cuh"
#include <cub/cub.cuh>
#include <algorithm>
#include <cmath>

namespace b200 {

bool coarsen_level_dist(Amg &amg, AmgLevel &L, AmgLevel &Lc);   // amg_dist.cu
void smoother_diag_dist(ParCsr &A, int relax_type, double *d);

namespace {

inline rows::Csr rv(const DCsr &A) { return rows::Csr{A.nr, A.nc, A.ptr.p, A.col.p, A.val.p}; }

struct PhaseTimer {  // CUDA-event timer for one setup phase
  Amg &amg;
  const char *name;
  cudaEvent_t e0, e1;
  PhaseTimer(Amg &a, const char *n) : amg(a), name(n) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, stream());
  }
  ~PhaseTimer() {
    cudaEventRecord(e1, stream());
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    amg.setup_phase_ms[name] += ms;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
  }
};

// ---- small CUB wrappers ---------------------------------------------------------------------------------------
template <class T>
void exclusive_scan(const T *in, T *out, i64 n) {
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, (int)n, stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, in, out, (int)n, stream());
  count_launch(2);
}
template <class T>
T fetch(const T *dev) {
  T h;
  B200_CUDA(cudaMemcpyAsync(&h, dev, sizeof(T), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  return h;
}

// ---- strength ---------------------------------------------------------------------------------------------------
__global__ void k_strength_count(rows::Csr A, double theta, double mrs, int *cnt) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < A.nr) cnt[i] = rows::strength_row(A, i, theta, mrs, nullptr);
  if (i == A.nr) cnt[i] = 0;
}
__global__ void k_strength_fill(rows::Csr A, double theta, double mrs, const int *Sptr, int *Scol) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < A.nr && Sptr[i + 1] > Sptr[i]) rows::strength_row(A, i, theta, mrs, Scol + Sptr[i]);
}

// ---- PMIS ---------------------------------------------------------------------------------------------------------
__global__ void k_col_count(const int *Scol, i64 nnz, int *cnt) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nnz) atomicAdd(cnt + Scol[k], 1);
}
__global__ void k_pmis_init(int n, const int *Sptr, const int *cnt, int seed, i64 first_index, double *measure, int *cf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (Sptr[i + 1] == Sptr[i]) { cf[i] = -3; measure[i] = 0.0; return; }
  double m = (double)cnt[i] + rows::hypre_rand_at(seed, first_index + i);
  if (m < 1.0) { cf[i] = -1; measure[i] = 0.0; }
  else { cf[i] = 0; measure[i] = m; }
}
__global__ void k_is_knock(int n, const int *Sptr, const int *Scol, const double *measure, const int *cf, signed char *is) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rows::pmis_knock_row(Sptr, Scol, measure, cf, is, i);
}
__global__ void k_set_c(int n, const signed char *is, int *cf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && is[i]) cf[i] = 1;
}
// end of a PMIS sweep in one pass over the rows: an undecided point that strongly depends on a C point becomes F
// (pmis_setf_row), decided points leave the graph (measure := 0), the points still undecided are counted, and the candidate flag
// of the next sweep's independent-set selection is written (is = undecided and measure > 1).  Row i only reads the markers of
// its neighbours, which this kernel changes from 0 to -1 at most -- invisible to the test `> 0` -- so one pass is race free.
__global__ void k_pmis_finish_sweep(int n, const int *Sptr, const int *Scol, int *cf, double *measure, signed char *is, int *counter) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int und = 0;
  if (i < n) {
    int c = cf[i];
    if (c == 0) {
      for (int k = Sptr[i]; k < Sptr[i + 1]; ++k)
        if (cf[Scol[k]] > 0) { c = -1; break; }
      if (c != 0) cf[i] = c;
    }
    if (c != 0) { measure[i] = 0.0; is[i] = 0; }
    else { und = 1; is[i] = (measure[i] > 1.0) ? 1 : 0; }
  }
  unsigned b = __ballot_sync(0xffffffffu, und);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(counter, __popc(b));
}
__global__ void k_c_flags(int n, const int *cf, int *flag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = cf[i] > 0 ? 1 : 0;
  if (i == n) flag[i] = 0;
}
__global__ void k_fold_sf(int n, const int *cf_raw, int *cf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) cf[i] = (cf_raw[i] == -3) ? -1 : cf_raw[i];
}

// ---- two-stage Gauss-Seidel (relax 11 / 12): strictly lower (upper = 0) or upper triangle of the own diag block ------------
__global__ void k_tri_count(CsrView A, int upper, int *cnt) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > A.nr) return;
  int c = 0;
  if (i < A.nr)
    for (int k = A.ptr[i]; k < A.ptr[i + 1]; ++k) c += (upper ? A.col[k] > i : A.col[k] < i) ? 1 : 0;
  cnt[i] = c;
}
__global__ void k_tri_fill(CsrView A, int upper, const int *tptr, int *tcol, double *tval) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.nr) return;
  int w = tptr[i];
  for (int k = A.ptr[i]; k < A.ptr[i + 1]; ++k)
    if (upper ? A.col[k] > i : A.col[k] < i) { tcol[w] = A.col[k]; tval[w] = A.val[k]; ++w; }
}

// ---- smoother diagonal / dense coarse matrix ------------------------------------------------------------------------
__global__ void k_smoother_diag(rows::Csr A, int relax_type, double *d) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < A.nr) d[i] = rows::smoother_diag_row(A, i, relax_type);
}
__global__ void k_inv_sqrt_diag(rows::Csr A, double *d) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < A.nr) d[i] = 1.0 / sqrt(A.val[A.ptr[i]]);
}
// rows of this rank into the dense global matrix (gn x gn): diag block columns are col_start + c, offd through the column map
__global__ void k_dense_fill_par(int n, CsrView Dg, CsrView Of, i64 row_start, i64 col_start, const i64 *colmap, int gn, double *D) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  size_t base = (size_t)(row_start + i) * gn;
  for (int k = Dg.ptr[i]; k < Dg.ptr[i + 1]; ++k) D[base + (size_t)(col_start + Dg.col[k])] = Dg.val[k];
  if (Of.ptr)
    for (int k = Of.ptr[i]; k < Of.ptr[i + 1]; ++k) D[base + (size_t)colmap[Of.col[k]]] = Of.val[k];
}
// unpivoted Gaussian elimination run once: multipliers are stored below the diagonal so that the
// solve applies exactly the operations of the per-solve elimination to the right-hand side.
__global__ void k_ge_factor(double *A, int n) {
  for (int k = 0; k < n - 1; ++k) {
    double akk = A[(size_t)k * n + k];
    if (akk != 0.0) {
      for (int j = k + 1 + threadIdx.x; j < n; j += blockDim.x) {
        double ajk = A[(size_t)j * n + k];
        double f = 0.0;
        if (ajk != 0.0) {
          f = ajk / akk;
          for (int m = k + 1; m < n; ++m) A[(size_t)j * n + m] -= f * A[(size_t)k * n + m];
        }
        A[(size_t)j * n + k] = f;
      }
    } else {
      for (int j = k + 1 + threadIdx.x; j < n; j += blockDim.x) A[(size_t)j * n + k] = 0.0;
    }
    __syncthreads();
  }
}

}  // namespace

// =====================================================================================================================
// setup driver
// =====================================================================================================================
namespace {

void tridiag_extreme_eigs(const std::vector<double> &d, const std::vector<double> &e, double &lmin, double &lmax) {
  // Sturm-sequence bisection on the symmetric tridiagonal (diag d[0..m), off-diag e[1..m)); host scalars only
  int m = (int)d.size();
  double lo = d[0], hi = d[0];
  for (int i = 0; i < m; ++i) {
    double r = (i > 0 ? std::fabs(e[i]) : 0.0) + (i + 1 < m ? std::fabs(e[i + 1]) : 0.0);
    lo = std::min(lo, d[i] - r);
    hi = std::max(hi, d[i] + r);
  }
  auto count_below = [&](double x) {
    int c = 0;
    double q = d[0] - x;
    if (q < 0) ++c;
    for (int i = 1; i < m; ++i) {
      double qq = (q == 0.0) ? 1e-300 : q;
      q = d[i] - x - e[i] * e[i] / qq;
      if (q < 0) ++c;
    }
    return c;
  };
  auto kth = [&](int k) {  // k-th smallest eigenvalue (0-based)
    double a = lo, b = hi;
    for (int it = 0; it < 200; ++it) {
      double mid = 0.5 * (a + b);
      if (mid == a || mid == b) break;
      if (count_below(mid) > k) b = mid;
      else a = mid;
    }
    return 0.5 * (a + b);
  };
  lmin = kth(0);
  lmax = kth(m - 1);
}

__global__ void k_cheby_rand(int n, int seed, i64 first, double *r) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) r[i] = 2.0 * rows::hypre_rand_at(seed, first + i) - 1.0;
}

void cheby_setup(Amg &amg, AmgLevel &L) {
  ParCsr &A = *L.A;
  int n = A.n;
  const AmgOptions &o = amg.opt;
  L.dsm.alloc((size_t)std::max(n, 1));
  // a rank may own no rows of a coarse level: no launch then (an empty grid is an invalid configuration), but it still takes part in
  // the inner products and exchanges below
  if (n) { k_inv_sqrt_diag<<<div_up(n, 256), 256, 0, stream()>>>(rv(A.diag), L.dsm.p); count_launch(); }
  double lmax = 0, lmin = 0;
  i64 gn = A.g_rows;
  int m = (int)std::min<i64>(o.cheby_eig_est, gn);
  if (m > 0) {
    i64 ld = ((i64)n + 1) & ~1LL;
    DVec<double> r(ld), p(ld), s(ld);
    if (n) { k_cheby_rand<<<div_up(n, 256), 256, 0, stream()>>>(n, ctx().rank + 1, 0, r.p); count_launch(); }
    std::vector<double> td, te;
    double gamma = dot_host(amg.dw, r.p, r.p, n), gamma_old, alpha, alpha_old = 0, beta = 0;
    vec_copy(p.p, r.p, n);
    int it = 0;
    while (it < m) {
      SpmvArgs a;  // s = D^{-1/2} A D^{-1/2} p  (CHEB epilogue with a0 = 0: y = d_i * d_i * a1 * t -> need single d: use xs + RES-free form)
      a.x = p.p; a.xs = L.dsm.p; a.y = s.p;
      par_spmv(A, SPMV_AX, a);
      vec_pointwise_mul(s.p, s.p, L.dsm.p, 1.0, n);
      double sdotp = dot_host(amg.dw, s.p, p.p, n);
      if (sdotp == 0.0) break;
      alpha = gamma / sdotp;
      vec_axpy(-alpha, s.p, r.p, n);
      gamma_old = gamma;
      gamma = dot_host(amg.dw, r.p, r.p, n);
      double beta_new = gamma / gamma_old;
      if (it == 0) { td.push_back(1.0 / alpha); te.push_back(0.0); }
      else { td.push_back(1.0 / alpha + beta / alpha_old); te.push_back(std::sqrt(beta) / alpha_old); }
      alpha_old = alpha;
      beta = beta_new;
      ++it;
      if (gamma == 0.0) break;
      vec_axpby(1.0, r.p, beta, p.p, n);
    }
    if (!td.empty()) tridiag_extreme_eigs(td, te, lmin, lmax);
  }
  L.cheby_max = lmax;
  L.cheby_min = lmin;
  double upper = 1.1 * lmax;
  double lower = (upper - lmin) * o.cheby_fraction + lmin;
  double theta = (upper + lower) / 2, delta = (upper - lower) / 2;
  int ord = std::max(1, std::min(4, o.cheby_order));
  if (ord == 1) L.cheby_coef = {1.0 / theta};
  else if (ord == 2) {
    double den = delta * delta - 2 * theta * theta;
    L.cheby_coef = {-4 * theta / den, 2 / den};
  } else if (ord == 3) {
    double den = 3 * (delta * delta) * theta - 4 * (theta * theta * theta);
    L.cheby_coef = {(3 * delta * delta - 12 * theta * theta) / den, 12 * theta / den, -4 / den};
  } else {
    double den = -(delta * delta * delta * delta) + 8 * delta * delta * theta * theta - 8 * theta * theta * theta * theta;
    L.cheby_coef = {(32 * theta * theta * theta - 16 * delta * delta * theta) / den, (8 * delta * delta - 48 * theta * theta) / den,
                    32 * theta / den, -8 / den};
  }
}

// one coarsening step on level L (single-rank domain: merged matrix == diag block).  Returns false when the level
// cannot be coarsened further.
bool coarsen_level(Amg &amg, AmgLevel &L, AmgLevel &Lc) {
  const AmgOptions &o = amg.opt;
  ParCsr &Ap = *L.A;
  const DCsr &A = Ap.diag;
  int n = A.nr;
  rows::Csr a = rv(A);
  int g = div_up(n, 256), g1 = div_up((i64)n + 1, 256);
  // strength
  {
    PhaseTimer t(amg, "strength");
    DVec<int> cnt((size_t)n + 1);
    k_strength_count<<<g1, 256, 0, stream()>>>(a, o.strong_threshold, o.max_row_sum, cnt.p);
    L.Sptr.alloc((size_t)n + 1);
    exclusive_scan(cnt.p, L.Sptr.p, (i64)n + 1);
    int nnzS = fetch(L.Sptr.p + n);
    L.S_nnz = nnzS;
    L.Scol.alloc((size_t)std::max(nnzS, 1));
    k_strength_fill<<<g, 256, 0, stream()>>>(a, o.strong_threshold, o.max_row_sum, L.Sptr.p, L.Scol.p);
    count_launch(2);
    B200_KERNEL_CHECK();
  }
  // PMIS
  int nc = 0;
  DVec<int> f2c((size_t)n + 1);
  {
    PhaseTimer t(amg, "pmis");
    int nnzS = fetch(L.Sptr.p + n);
    DVec<int> cnt((size_t)n);
    cnt.zero();
    if (nnzS) k_col_count<<<div_up(nnzS, 256), 256, 0, stream()>>>(L.Scol.p, nnzS, cnt.p);
    DVec<double> measure((size_t)n);
    L.cf_raw.alloc((size_t)n);
    k_pmis_init<<<g, 256, 0, stream()>>>(n, L.Sptr.p, cnt.p, 2747 + ctx().rank, 0, measure.p, L.cf_raw.p);
    DVec<signed char> is((size_t)n);
    DVec<int> counter(1);
    count_launch(2);
    // sweep = knock-out (independent set among the candidates) -> selected points become C -> k_pmis_finish_sweep.
    // The count of undecided points and the candidate flags of the first sweep come from the same kernel run on the
    // initial markers (no C point yet, so it decides nothing).
    counter.zero();
    k_pmis_finish_sweep<<<g, 256, 0, stream()>>>(n, L.Sptr.p, L.Scol.p, L.cf_raw.p, measure.p, is.p, counter.p);
    count_launch();
    for (int sweep = 0; sweep < 10000; ++sweep) {
      i64 und = fetch(counter.p);
      allreduce_sum_host_i64(&und, 1);
      if (und == 0) break;
      k_is_knock<<<g, 256, 0, stream()>>>(n, L.Sptr.p, L.Scol.p, measure.p, L.cf_raw.p, is.p);
      k_set_c<<<g, 256, 0, stream()>>>(n, is.p, L.cf_raw.p);
      counter.zero();
      k_pmis_finish_sweep<<<g, 256, 0, stream()>>>(n, L.Sptr.p, L.Scol.p, L.cf_raw.p, measure.p, is.p, counter.p);
      count_launch(3);
    }
    B200_KERNEL_CHECK();
    // coarse numbering: exclusive scan of the C flags
    DVec<int> flag((size_t)n + 1);
    k_c_flags<<<g1, 256, 0, stream()>>>(n, L.cf_raw.p, flag.p);
    exclusive_scan(flag.p, f2c.p, (i64)n + 1);
    nc = fetch(f2c.p + n);
    count_launch();
  }
  i64 gnc = nc, gn = n;
  allreduce_sum_host_i64(&gnc, 1);
  allreduce_sum_host_i64(&gn, 1);
  if (gnc == 0 || gnc == gn || gnc < o.min_coarse_size) return false;
  // interpolation (extended+i) + truncation
  {
    PhaseTimer t(amg, "interp");
    DCsr &P = L.P.diag;
    P.nc = nc;
    if (o.interp_type == 0) build_interp_classical(A, L.Sptr.p, L.Scol.p, L.S_nnz, L.cf_raw.p, f2c.p, n, n, o.trunc_factor, o.p_max_elmts, P);
    else build_interp(A, L.Sptr.p, L.Scol.p, L.S_nnz, L.cf_raw.p, f2c.p, n, n, o.trunc_factor, o.p_max_elmts, P);
    L.cf.alloc((size_t)n);
    k_fold_sf<<<g, 256, 0, stream()>>>(n, L.cf_raw.p, L.cf.p);
    count_launch();
    L.P.n = n;
    L.P.n_cols_local = nc;
    L.P.g_rows = gn;
    L.P.g_cols = gnc;
  }
  // R = P^T, Q = A P, A_c = R Q
  {
    PhaseTimer t(amg, "rap");
    {
      PhaseTimer t1(amg, "rap.transpose");
      csr_transpose(L.P.diag, L.R.diag);
    }
    L.R.n = nc;
    L.R.n_cols_local = n;
    L.R.g_rows = gnc;
    L.R.g_cols = gn;
    DCsr Q;
    {
      PhaseTimer t2(amg, "rap.AP");
      csr_spgemm(A, L.P.diag, Q, false, false);
    }
    Lc.A_own.reset(new ParCsr());
    Lc.A = Lc.A_own.get();
    PhaseTimer t3(amg, "rap.RQ");
    csr_spgemm(L.R.diag, Q, Lc.A->diag, true);
    Lc.A->n = nc;
    Lc.A->n_cols_local = nc;
    Lc.A->g_rows = Lc.A->g_cols = gnc;
  }
  return true;
}

}  // namespace

__global__ void k_rand_stream(int n, int seed, double *out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = rows::hypre_rand_at(seed, i);
}
void test_rand(int seed, int n, double *out_dev) {  // value i of hypre_Rand's stream for `seed`, evaluated independently per i
  if (n <= 0) return;
  k_rand_stream<<<div_up(n, 256), 256, 0, stream()>>>(n, seed, out_dev);
  count_launch();
}

Amg::~Amg() {
  if (graph_exec) cudaGraphExecDestroy(graph_exec);
}

void amg_setup(Amg &amg, ParCsr &A) {
  ensure_init();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0, stream());
  const AmgOptions &o = amg.opt;
  B200_REQUIRE(o.coarsen_type == 8 || o.coarsen_type == 10 || o.coarsen_type == 6 || o.max_levels == 1,
               "BoomerAMG: coarsen_type must be 8 (PMIS); 10 (HMIS) and 6 (Falgout) are accepted and run as PMIS with a warning; got " +
                   std::to_string(o.coarsen_type));
  B200_REQUIRE(interp_type_built(o.interp_type) || o.max_levels == 1,
               "BoomerAMG: interp_type must be 6 (extended+i) or 0 (classical modified); got " + std::to_string(o.interp_type));
  // coarsest level: Gaussian elimination (relax_type[3] = 9, what HYPRE_BoomerAMGSetRelaxType leaves there); any other
  // coarse_relax_type (HYPRE_BoomerAMGSetCycleRelaxType(.., 3), HypreSystem.cpp:135-136) is refused rather than ignored
  B200_REQUIRE(o.relax_type[3] == 9 || o.max_levels == 1,
               "BoomerAMG: coarse_relax_type must be 9 (Gaussian elimination on the coarsest level); got " + std::to_string(o.relax_type[3]));
  for (int k = 0; k < 3; ++k)
    B200_REQUIRE(o.relax_type[k] == 18 || o.relax_type[k] == 7 || o.relax_type[k] == 16 || is_gs_relax(o.relax_type[k]) ||
                     is_two_stage_gs(o.relax_type[k]),
                 "BoomerAMG: relax_type must be 7 (Jacobi), 18 (l1-Jacobi), 16 (Chebyshev), hybrid Gauss-Seidel (3, 4, 6, 8, 13, 14) or "
                 "two-stage Gauss-Seidel (11, 12)");
  B200_REQUIRE(o.relax_type[1] == o.relax_type[2] || (is_two_stage_gs(o.relax_type[1]) && is_two_stage_gs(o.relax_type[2])) ||
                   (is_gs_relax(o.relax_type[1]) && is_gs_relax(o.relax_type[2]) &&
                    is_l1_gs_relax(o.relax_type[1]) == is_l1_gs_relax(o.relax_type[2])),
               "BoomerAMG: down and up smoothers must share their diagonal scaling (same relax_type family)");
  if ((o.coarsen_type == 10 || o.coarsen_type == 6) && o.max_levels > 1 && ctx().rank == 0) {
    static bool warned[2] = {false, false};
    const int w = o.coarsen_type == 10 ? 0 : 1;
    if (!warned[w])
      fprintf(stderr, w == 0 ? "[b200hypre] warning: coarsen_type 10 (HMIS) is run as coarsen_type 8 (PMIS): the Ruge-Stueben first pass of HMIS "
                               "is not built; the hierarchy is the PMIS one\n"
                             : "[b200hypre] warning: coarsen_type 6 (Falgout) is run as coarsen_type 8 (PMIS): the Ruge-Stueben passes and CLJP "
                               "are not built; the hierarchy is the PMIS one\n");
    warned[w] = true;
  }
  if (amg.graph_exec) { cudaGraphExecDestroy(amg.graph_exec); amg.graph_exec = nullptr; }
  amg.setup_phase_ms.clear();
  amg.lev.clear();
  amg.lev.emplace_back(new AmgLevel());
  amg.lev[0]->A = &A;
  int level = 0;
  while (true) {
    AmgLevel &L = *amg.lev[level];
    i64 fine = L.A->g_rows;
    bool last = (level >= o.max_levels - 1) || (fine <= o.max_coarse_size);
    if (!last) {
      std::unique_ptr<AmgLevel> Lc(new AmgLevel());
      const std::string lname = "level." + std::to_string(level);   // whole coarsening step of this level (all phases)
      PhaseTimer tl(amg, lname.c_str());
      const bool ok = (ctx().nranks > 1) ? coarsen_level_dist(amg, L, *Lc) : coarsen_level(amg, L, *Lc);
      if (ok) amg.lev.push_back(std::move(Lc));
      else last = true;
    }
    if (last) break;
    ++level;
  }
  int nl = (int)amg.lev.size();
  {
    PhaseTimer t(amg, "smoother_data");
    for (int l = 0; l < nl; ++l) {
      AmgLevel &L = *amg.lev[l];
      ParCsr &Al = *L.A;
      int n = Al.n;
      Al.finalize_spmv();
      if (l < nl - 1) { L.P.finalize_spmv(); L.R.finalize_spmv(); }
      int rt = o.relax_type[1];
      if (rt == 16) cheby_setup(amg, L);
      else if (is_gs_relax(rt)) {
        L.dsm.alloc((size_t)std::max(n, 1));
        gs_smoother_diag(Al, rt, L.dsm.p);
      } else {
        L.dsm.alloc((size_t)std::max(n, 1));
        const int drt = is_two_stage_gs(rt) ? 7 : rt;   // two-stage Gauss-Seidel scales with a_ii
        if (ctx().nranks > 1) smoother_diag_dist(Al, drt, L.dsm.p);
        else if (n) { k_smoother_diag<<<div_up(n, 256), 256, 0, stream()>>>(rv(Al.diag), drt, L.dsm.p); count_launch(); }
        for (int side = 0; side < 2; ++side) {
          const bool need = (o.relax_type[1] == 11 + side) || (o.relax_type[2] == 11 + side);
          if (!need || n == 0) continue;
          DCsr &T = L.tri[side];
          T.nr = n; T.nc = Al.diag.nc;
          DVec<int> cnt((size_t)n + 1);
          T.ptr.alloc((size_t)n + 1);
          k_tri_count<<<div_up((i64)n + 1, 256), 256, 0, stream()>>>(Al.diag.view(), side, cnt.p);
          exclusive_scan(cnt.p, T.ptr.p, (i64)n + 1);
          T.nnz = fetch(T.ptr.p + n);
          T.col.alloc((size_t)std::max<i64>(T.nnz, 1));
          T.val.alloc((size_t)std::max<i64>(T.nnz, 1));
          k_tri_fill<<<div_up(n, 256), 256, 0, stream()>>>(Al.diag.view(), side, T.ptr.p, T.col.p, T.val.p);
          count_launch(2);
          L.tri_plan[side] = make_plan(T, nullptr);
        }
      }
      i64 ld = ((i64)n + 1) & ~1LL;
      L.u.alloc(ld); L.f.alloc(ld); L.r.alloc(ld); L.t0.alloc(ld); L.t1.alloc(ld);
      L.u.zero(); L.f.zero(); L.r.zero(); L.t0.zero(); L.t1.zero();
      if (L.cf.n == 0 && n) { L.cf.alloc(n); std::vector<int> ones(n, 1); L.cf.from_host(ones); }
      if (is_gs_relax(rt) || is_gs_relax(o.relax_type[2])) {
        // wavefront schedules of the Gauss-Seidel passes this level will run (gs.cu); the coarsest level is only smoothed
        // when it is not solved directly
        const bool smoothed = (l < nl - 1) || (nl == 1 && o.max_levels == 1) || Al.g_rows > 2048 || Al.g_rows == 0;
        if (smoothed && n) {
          DCsr pat, T;
          pat.nr = Al.diag.nr; pat.nc = Al.diag.nc; pat.nnz = Al.diag.nnz;
          pat.ptr.alloc(Al.diag.ptr.n); pat.col.alloc(Al.diag.col.n);
          B200_CUDA(cudaMemcpyAsync(pat.ptr.p, Al.diag.ptr.p, sizeof(int) * Al.diag.ptr.n, cudaMemcpyDeviceToDevice, stream()));
          B200_CUDA(cudaMemcpyAsync(pat.col.p, Al.diag.col.p, sizeof(int) * Al.diag.col.n, cudaMemcpyDeviceToDevice, stream()));
          csr_transpose(pat, T);
          bool need_dir[2] = {false, false};
          for (int k = 1; k <= 2; ++k) {
            int t = o.relax_type[k];
            if (!is_gs_relax(t)) continue;
            if (t != 4 && t != 14) need_dir[0] = true;
            if (t == 4 || t == 14 || t == 6 || t == 8) need_dir[1] = true;
          }
          for (int ps = 0; ps < 3; ++ps) {
            if ((o.relax_order == 1) != (ps != 0)) continue;
            for (int dir = 0; dir < 2; ++dir)
              if (need_dir[dir]) {
                build_gs_plan(Al, T, L.cf.p, L.dsm.p, ps == 0 ? 0 : (ps == 1 ? 1 : -1), dir, L.gs[ps][dir]);
                L.gs[ps][dir].built = true;
              }
          }
        }
      }
    }
    // coarsest: dense LU
    AmgLevel &C = *amg.lev[nl - 1];
    int n = C.A->n;
    i64 gn = C.A->g_rows;
    amg.ge_n = 0;
    amg.ge_row_start = 0;
    if (!(nl == 1 && o.max_levels == 1) && gn > 0 && gn <= 2048) {
      // every rank holds the whole (tiny) coarsest operator densely: rows are disjoint, so the sum over ranks is exact
      amg.ge_n = (int)gn;
      amg.ge_row_start = (int)C.A->row_start;
      amg.ge_lu.alloc((size_t)gn * gn);
      amg.ge_lu.zero();
      amg.ge_f.alloc((size_t)gn);
      DVec<i64> dmap;
      dmap.from_host(C.A->col_map_offd.data(), C.A->col_map_offd.size());
      if (n) {
        k_dense_fill_par<<<div_up(n, 128), 128, 0, stream()>>>(n, C.A->diag.view(),
                                                             C.A->offd.nnz > 0 ? C.A->offd.view() : CsrView{0, 0, nullptr, nullptr, nullptr},
                                                             C.A->row_start, C.A->col_start, dmap.p, (int)gn, amg.ge_lu.p);
        count_launch();
      }
      for (i64 off = 0; off < gn * gn; off += 1 << 20) allreduce_sum_dev(amg.ge_lu.p + off, (int)std::min<i64>(1 << 20, gn * gn - off));
      k_ge_factor<<<1, 256, 0, stream()>>>(amg.ge_lu.p, (int)gn);
      count_launch();
    }
    B200_KERNEL_CHECK();
  }
  amg.a_array.clear();
  for (auto &l : amg.lev) amg.a_array.push_back(l->A);
  amg.setup_done = true;
  cudaEventRecord(e1, stream());
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  amg.setup_seconds = ms * 1e-3;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
}

}  // namespace b200
