// ilu.cu -- hypre's ILU solver / preconditioner, type 0 with level of fill 0: block-Jacobi ILU(0).
// Replaces HYPRE_ILUSetup / HYPRE_ILUSolve as reached from `preconditioner: ilu` (setup_ilu_precond,
// /root/reference/src/HypreSystem.cpp:328-370) and `method: ilu` (setup_ilu, :457-497) with the mini-app's defaults
// (ilu_type 0, fill 0, local_reordering 0, trisolve 1); semantics restated in oracle/krylov.inc (ilu0_setup, ilu_apply,
// ilu_solver).  Every rank factors its own diagonal block; couplings to other ranks are ignored by the factorisation
// and enter only through the residual b - A x.
//
// Factorisation: row i depends on the rows k < i of its pattern -- the same DAG as a forward Gauss-Seidel sweep -- so
// the wavefront schedule of gs.cu (Kahn peeling in one persistent kernel) orders the rows, and one kernel per wavefront
// eliminates them: a thread owns a row, walks its strictly lower entries in ascending column order
// (l_ik = a_ik * (1 / u_kk), then a_ij -= l_ik u_kj for the (i, j) inside the pattern, found by binary search in the
// sorted row), exactly the sequential IKJ order, with unfused multiply / subtract.  Pivots below 1e-14 become 1e-6.
// Solves: L (unit diagonal) and U are stored as two CSR matrices with the diagonal first, so that the forward and the
// backward substitution are the wavefront sweeps of gs.cu (uout_i = (f_i - sum_done a_ij uout_j) / a_ii).
#include "amg.h"
#include <cub/cub.cuh>
#include <cmath>

namespace b200 {

namespace {

__global__ void k_ilu0_level(int cnt, const int *__restrict__ rows, const int *__restrict__ ptr, const int *__restrict__ col,
                             double *val, double *dinv) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= cnt) return;
  const int i = rows[t];
  const int b = ptr[i], e = ptr[i + 1];   // slot b: diagonal; b + 1 .. e: ascending columns
  for (int q = b + 1; q < e; ++q) {
    const int k = col[q];
    if (k > i) break;
    const double dpiv = __dmul_rn(val[q], dinv[k]);
    val[q] = dpiv;
    for (int p2 = ptr[k] + 1; p2 < ptr[k + 1]; ++p2) {
      const int j = col[p2];
      if (j <= k) continue;               // U part of row k
      int pos = -1;
      if (j == i) pos = b;
      else {
        int lo = b + 1, hi = e;
        while (lo < hi) {
          int mid = lo + ((hi - lo) >> 1);
          if (col[mid] < j) lo = mid + 1;
          else hi = mid;
        }
        if (lo < e && col[lo] == j) pos = lo;
      }
      if (pos >= 0) val[pos] = __dsub_rn(val[pos], __dmul_rn(dpiv, val[p2]));
    }
  }
  double d = val[b];
  if (fabs(d) < 1e-14) { d = 1e-6; val[b] = d; }
  dinv[i] = 1.0 / d;
}

__global__ void k_check_diag_first(int n, const int *ptr, const int *col, int *bad) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && (ptr[i + 1] == ptr[i] || col[ptr[i]] != i)) *bad = 1;
}
__global__ void k_lu_counts(int n, const int *ptr, const int *col, int *lcnt, int *ucnt) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == n) { lcnt[n] = 0; ucnt[n] = 0; }
  if (i >= n) return;
  int l = 1, u = 1;
  for (int q = ptr[i] + 1; q < ptr[i + 1]; ++q) (col[q] < i ? l : u) += 1;
  lcnt[i] = l;
  ucnt[i] = u;
}
__global__ void k_lu_split(int n, const int *ptr, const int *col, const double *val, const int *lptr, int *lcol, double *lval,
                           const int *uptr, int *ucol, double *uval) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int pl = lptr[i], pu = uptr[i];
  lcol[pl] = i; lval[pl] = 1.0; ++pl;
  ucol[pu] = i; uval[pu] = val[ptr[i]]; ++pu;
  for (int q = ptr[i] + 1; q < ptr[i + 1]; ++q) {
    if (col[q] < i) { lcol[pl] = col[q]; lval[pl] = val[q]; ++pl; }
    else { ucol[pu] = col[q]; uval[pu] = val[q]; ++pu; }
  }
}

void exscan_int(const int *in, int *out, i64 n) {
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, (int)n, stream());
  DVec<char> tmp(tb);
  cub::DeviceScan::ExclusiveSum(tmp.p, tb, in, out, (int)n, stream());
  count_launch();
}

void set_local_shape(ParCsr &M, const ParCsr &A) {
  M.n = A.n; M.n_cols_local = A.n; M.n_halo = 0;
  M.g_rows = A.g_rows; M.g_cols = A.g_rows; M.row_start = A.row_start; M.col_start = A.row_start;
  M.diag.nr = A.n; M.diag.nc = A.n;
  M.offd.nr = A.n; M.offd.nc = 0; M.offd.nnz = 0;
}

}  // namespace

void ilu_setup(Ilu &ilu, ParCsr &A) {
  ensure_init();
  B200_REQUIRE(ilu.type == 0 && ilu.fill == 0, "ILU: only ilu_type 0 with fill 0 (block-Jacobi ILU(0)) is built");
  B200_REQUIRE(ilu.reordering == 0, "ILU: only local_reordering 0 is built");
  B200_REQUIRE(ilu.tri_solve == 1 && ilu.iter_setup_type == 0, "ILU: only direct triangular solves and the non-iterative setup are built");
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0, stream());
  const int n = A.n;
  ilu.n = n;
  const DCsr &D = A.diag;
  {
    DVec<int> bad(1);
    bad.zero();
    if (n) k_check_diag_first<<<div_up(n, 256), 256, 0, stream()>>>(n, D.ptr.p, D.col.p, bad.p);
    count_launch();
    B200_REQUIRE(bad.to_host()[0] == 0, "ILU: a row of the diagonal block has no diagonal entry");
  }
  // factor values in the pattern of A.diag
  DVec<double> val((size_t)std::max<i64>(D.nnz, 1)), dinv((size_t)std::max(n, 1)), ones((size_t)std::max(n, 1));
  if (D.nnz) B200_CUDA(cudaMemcpyAsync(val.p, D.val.p, sizeof(double) * D.nnz, cudaMemcpyDeviceToDevice, stream()));
  vec_fill(ones.p, 1.0, n);
  // wavefront schedules of the block: forward (factorisation, L solve) and backward (U solve)
  DCsr T;
  csr_transpose(D, T);
  build_gs_plan(A, T, nullptr, ones.p, 0, 0, ilu.plan_l);
  build_gs_plan(A, T, nullptr, ones.p, 0, 1, ilu.plan_u);
  ilu.plan_l.built = ilu.plan_u.built = true;
  if (n) {
    std::vector<int> off((size_t)ilu.plan_l.depth + 1);
    B200_CUDA(cudaMemcpyAsync(off.data(), ilu.plan_l.lvl_off.p, sizeof(int) * off.size(), cudaMemcpyDeviceToHost, stream()));
    B200_CUDA(cudaStreamSynchronize(stream()));
    for (int l = 0; l < ilu.plan_l.depth; ++l) {
      int cnt = off[l + 1] - off[l];
      if (cnt <= 0) continue;
      k_ilu0_level<<<div_up(cnt, 128), 128, 0, stream()>>>(cnt, ilu.plan_l.order.p + off[l], D.ptr.p, D.col.p, val.p, dinv.p);
    }
    count_launch(ilu.plan_l.depth);
    B200_KERNEL_CHECK();
  }
  // L (unit diagonal first) and U (pivot first) as separate CSR matrices for the substitution sweeps
  DVec<int> lcnt((size_t)n + 1), ucnt((size_t)n + 1);
  k_lu_counts<<<div_up((i64)n + 1, 256), 256, 0, stream()>>>(n, D.ptr.p, D.col.p, lcnt.p, ucnt.p);
  set_local_shape(ilu.Lm, A);
  set_local_shape(ilu.Um, A);
  ilu.Lm.diag.ptr.alloc((size_t)n + 1);
  ilu.Um.diag.ptr.alloc((size_t)n + 1);
  exscan_int(lcnt.p, ilu.Lm.diag.ptr.p, (i64)n + 1);
  exscan_int(ucnt.p, ilu.Um.diag.ptr.p, (i64)n + 1);
  int tot[2] = {0, 0};
  B200_CUDA(cudaMemcpyAsync(&tot[0], ilu.Lm.diag.ptr.p + n, sizeof(int), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaMemcpyAsync(&tot[1], ilu.Um.diag.ptr.p + n, sizeof(int), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  ilu.Lm.diag.nnz = tot[0];
  ilu.Um.diag.nnz = tot[1];
  ilu.Lm.diag.col.alloc((size_t)std::max(tot[0], 1)); ilu.Lm.diag.val.alloc((size_t)std::max(tot[0], 1));
  ilu.Um.diag.col.alloc((size_t)std::max(tot[1], 1)); ilu.Um.diag.val.alloc((size_t)std::max(tot[1], 1));
  if (n)
    k_lu_split<<<div_up(n, 256), 256, 0, stream()>>>(n, D.ptr.p, D.col.p, val.p, ilu.Lm.diag.ptr.p, ilu.Lm.diag.col.p, ilu.Lm.diag.val.p,
                                                     ilu.Um.diag.ptr.p, ilu.Um.diag.col.p, ilu.Um.diag.val.p);
  count_launch(4);
  B200_KERNEL_CHECK();
  ilu.fval = std::move(val);
  ilu.dinv = std::move(dinv);
  i64 ld = ((i64)n + 1) & ~1LL;
  ilu.r.alloc((size_t)std::max<i64>(ld, 2));
  ilu.y.alloc((size_t)std::max<i64>(ld, 2));
  ilu.z.alloc((size_t)std::max<i64>(ld, 2));
  ilu.setup_done = true;
  cudaEventRecord(e1, stream());
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  ilu.setup_seconds = ms * 1e-3;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
}

// z = U^{-1} L^{-1} r on this rank's block
void ilu_apply(Ilu &ilu, const double *r, double *z) {
  B200_REQUIRE(ilu.setup_done, "ILU: Solve called before Setup");
  gs_sweep(ilu.Lm, ilu.plan_l, false, nullptr, r, r, ilu.y.p);
  gs_sweep(ilu.Um, ilu.plan_u, false, nullptr, ilu.y.p, ilu.y.p, z);
}

// hypre_ILUSolve: x += (LU)^{-1} (b - A x) while ||b - A x|| / ||b|| >= tol (at least once) and iterations < max_iter
void ilu_solve(Ilu &ilu, ParCsr &A, const double *b, double *x, bool zero_guess) {
  B200_REQUIRE(ilu.setup_done && ilu.n == A.n, "ILU: Solve called before Setup");
  const int n = A.n;
  if (zero_guess && ilu.max_iter == 1 && ilu.tol == 0.0) {  // preconditioner use: x = (LU)^{-1} b
    ilu_apply(ilu, b, x);
    ilu.num_iterations = 1;
    return;
  }
  if (zero_guess) vec_fill(x, 0.0, n);   // preconditioner use with max_iter > 1 or tol > 0: the output counts as zero
  auto residual = [&]() { SpmvArgs a; a.x = x; a.y = ilu.r.p; a.b = b; par_spmv(A, SPMV_RES, a); };
  double rhs_norm = 0.0, rel = 1.0;
  residual();
  if (ilu.tol > 0.0) {
    rhs_norm = std::sqrt(dot_host(ilu.dw, b, b, n));
    double resid = std::sqrt(dot_host(ilu.dw, ilu.r.p, ilu.r.p, n));
    rel = rhs_norm > 2.2e-16 ? resid / rhs_norm : resid;
    if (ilu.print_level > 1 && ctx().rank == 0) printf("    ILU iter  0   %e    %e\n", resid, rel);
  }
  int iter = 0;
  while ((rel >= ilu.tol || iter < 1) && iter < ilu.max_iter) {
    ilu_apply(ilu, ilu.r.p, ilu.z.p);
    vec_axpy(1.0, ilu.z.p, x, n);
    ++iter;
    if (ilu.tol > 0.0 || iter < ilu.max_iter) residual();
    if (ilu.tol > 0.0) {
      double resid = std::sqrt(dot_host(ilu.dw, ilu.r.p, ilu.r.p, n));
      rel = rhs_norm > 2.2e-16 ? resid / rhs_norm : resid;
      if (ilu.print_level > 1 && ctx().rank == 0) printf("    ILU iter %2d   %e    %e\n", iter, resid, rel);
    }
  }
  ilu.num_iterations = iter;
  ilu.rel_resid = rel;
}

}  // namespace b200
