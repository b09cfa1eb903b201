// krylov.cu -- the short-recurrence Krylov solvers the mini-app can select besides GMRES: preconditioned CG
// (`method: cg`, /root/reference/src/HypreSystem.cpp:440-455 -> HYPRE_ParCSRPCGSetup/Solve) and BiCGSTAB
// (`method: bicg`, :423-438 -> HYPRE_ParCSRBiCGSTABSetup/Solve).  Control flow, stopping tests and breakdown guards
// follow hypre_PCGSolve / hypre_BiCGSTABSolve (restated in oracle/krylov.inc); the vector work per iteration is
//   PCG      : SpMV, <s,p>, fused {x += a p ; r -= a s}, preconditioner, <r,s>, p = s + b p
//   BiCGSTAB : 2 preconditioner calls, 2 SpMV, <r0,q>, fused {x += a v ; r -= a q}, one pass for {<r,s>, <s,s>},
//              fused {x += g v ; r -= g s ; <r,r> ; <r0,r>}, fused p update
// i.e. every vector is streamed once per use with 128-bit loads and the reductions ride on the update kernels
// (vecops.cu).  Scalars come back to the host once per reduction, like hypre's InnerProd.
#include "amg.h"
#include <cmath>

namespace b200 {

void krylov_setup(Krylov &k, ParCsr &A, ParVec &b, ParVec &x) {
  ensure_init();
  k.n = A.n;
  k.ld = ((i64)A.n + 1) & ~1LL;
  if (k.ld == 0) k.ld = 2;
  int nvec = (k.method == Krylov::PCG) ? 3 : 6;
  k.W.alloc((size_t)k.ld * nvec);
  k.W.zero();
  k.hdev.alloc(16);
  if (k.precond_setup) {
    int rc = k.precond_setup(k.precond_data, &A, &b, &x);
    B200_REQUIRE(rc == 0, "preconditioner setup failed (see the error reported above)");
  }
}

namespace {

struct Ops {
  Krylov &k;
  ParCsr &A;
  ParVec pin, pout;
  i64 n;
  double hbuf[4];
  Ops(Krylov &k_, ParCsr &A_, ParVec &b) : k(k_), A(A_), n(k_.n) {
    pin.g_n = pout.g_n = b.g_n; pin.start = pout.start = b.start; pin.n = pout.n = (int)n;
  }
  void precond(double *in, double *out) {  // out = M^{-1} in (hypre clears out first; our AMG is told to treat it as zero)
    if (k.precond) {
      pin.p = in; pout.p = out;
      apply_precond(k.precond, k.precond_data, A, pin, pout, "Krylov solver");
    } else vec_copy(out, in, n);
  }
  void matvec(const double *x, double *y) { SpmvArgs a; a.x = x; a.y = y; par_spmv(A, SPMV_AX, a); }
  void residual(const double *x, const double *b, double *r) { SpmvArgs a; a.x = x; a.y = r; a.b = b; par_spmv(A, SPMV_RES, a); }
  double dot(const double *a, const double *b) { return dot_host(k.dw, a, b, n); }
  // results of a fused reduction left at k.hdev[0..cnt): all-reduce, copy to the host
  const double *fetch(int cnt) {
    allreduce_sum_dev(k.hdev.p, cnt);
    B200_CUDA(cudaMemcpyAsync(hbuf, k.hdev.p, sizeof(double) * cnt, cudaMemcpyDeviceToHost, stream()));
    B200_CUDA(cudaStreamSynchronize(stream()));
    return hbuf;
  }
};

int pcg_solve(Krylov &k, ParCsr &A, ParVec &bvec, ParVec &xvec) {
  Ops op(k, A, bvec);
  const i64 n = k.n, ld = k.ld;
  const double *b = bvec.p;
  double *x = xvec.p, *p = k.W.p, *s = k.W.p + ld, *r = k.W.p + 2 * ld;
  const bool root = ctx().rank == 0;
  double bi_prod;
  if (k.two_norm) bi_prod = op.dot(b, b);
  else { op.precond(const_cast<double *>(b), p); bi_prod = op.dot(p, b); }
  B200_REQUIRE(bi_prod == bi_prod && !std::isinf(bi_prod), "PCG: rhs contains NaN/Inf");
  double eps = k.tol * k.tol;
  k.history.clear();
  k.converged = 0;
  k.num_iterations = 0;
  if (bi_prod > 0.0) eps = std::max(k.tol * k.tol, k.atol * k.atol / bi_prod);
  else {  // zero right-hand side: the solution is x = b = 0
    vec_copy(x, b, n);
    k.converged = 1;
    k.rel_residual_norm = 0.0;
    return 0;
  }
  op.residual(x, b, r);
  op.precond(r, p);
  double gamma = op.dot(r, p);
  B200_REQUIRE(gamma == gamma && !std::isinf(gamma), "PCG: initial residual contains NaN/Inf");
  double i_prod = k.two_norm ? op.dot(r, r) : gamma;
  k.history.push_back(std::sqrt(std::fabs(i_prod)));
  if (k.print_level > 1 && root) printf("\n\nIters       ||r||_%c     conv.rate  ||r||_%c/||b||_%c\n-----    ------------    ---------  ------------ \n", k.two_norm ? '2' : 'C', k.two_norm ? '2' : 'C', k.two_norm ? '2' : 'C');
  int i = 0, rc = 0;
  while (i + 1 <= k.max_iter) {
    ++i;
    op.matvec(p, s);
    double sdotp = op.dot(s, p);
    if (sdotp == 0.0) { rc = 256; break; }
    double alpha = gamma / sdotp, gamma_old = gamma;
    if (k.two_norm) {
      fused_update2(k.dw, alpha, p, s, x, r, nullptr, n, k.hdev.p);
      i_prod = op.fetch(1)[0];
    } else fused_update2(k.dw, alpha, p, s, x, r, nullptr, n, nullptr);
    op.precond(r, s);
    gamma = op.dot(r, s);
    if (!k.two_norm) i_prod = gamma;
    double prev = k.history.back(), cur = std::sqrt(std::fabs(i_prod));
    k.history.push_back(cur);
    if (k.print_level > 1 && root) printf("% 5d    %e    %f     %e\n", i, cur, prev > 0 ? cur / prev : 0.0, std::sqrt(std::fabs(i_prod) / bi_prod));
    if (i_prod / bi_prod < eps) { k.converged = 1; break; }
    if (gamma < 1.0e-292 && -gamma < 1.0e-292) { rc = 256; break; }
    double beta = gamma / gamma_old;
    vec_axpby(1.0, s, beta, p, n);  // p = s + beta p
  }
  k.num_iterations = i;
  k.rel_residual_norm = std::sqrt(std::fabs(i_prod) / bi_prod);
  if (!k.converged && i >= k.max_iter && eps > 0) rc = 256;
  return rc;
}

int bicgstab_solve(Krylov &k, ParCsr &A, ParVec &bvec, ParVec &xvec) {
  Ops op(k, A, bvec);
  const i64 n = k.n, ld = k.ld;
  const double epsmac = 1.e-128;
  const double *b = bvec.p;
  double *x = xvec.p;
  // r and s adjacent: {<r,s>, <s,s>} is one multi_dot over the pair
  double *r = k.W.p, *s = k.W.p + ld, *r0 = k.W.p + 2 * ld, *p = k.W.p + 3 * ld, *v = k.W.p + 4 * ld, *q = k.W.p + 5 * ld;
  const bool root = ctx().rank == 0;
  k.history.clear();
  k.converged = 0;
  op.residual(x, b, r0);
  vec_copy(r, r0, n);
  vec_copy(p, r0, n);
  double b_norm = std::sqrt(op.dot(b, b));
  B200_REQUIRE(b_norm == b_norm && !std::isinf(b_norm), "BiCGSTAB: rhs contains NaN/Inf");
  double res = op.dot(r0, r0);
  double r_norm = std::sqrt(res);
  B200_REQUIRE(r_norm == r_norm && !std::isinf(r_norm), "BiCGSTAB: initial residual contains NaN/Inf");
  k.history.push_back(r_norm);
  double den_norm = (b_norm > 0.0) ? b_norm : r_norm;
  double epsilon = std::max(k.atol, k.tol * den_norm);
  int iter = 0, rc = 0;
  if (k.print_level > 1 && root) printf("=============================================\n\nIters     resid.norm     conv.rate  rel.res.norm\n-----    ------------    ---------- ------------\n");
  if (r_norm == 0.0 || (r_norm <= epsilon && iter >= k.min_iter)) {
    k.converged = 1;
    k.num_iterations = 0;
    k.rel_residual_norm = (b_norm > 0.0) ? r_norm / b_norm : r_norm;
    return 0;
  }
  while (iter < k.max_iter) {
    ++iter;
    op.precond(p, v);
    op.matvec(v, q);
    double temp = op.dot(r0, q);
    if (std::fabs(temp) < epsmac) { rc = 256; break; }
    double alpha = res / temp;
    fused_update2(k.dw, alpha, v, q, x, r, nullptr, n, nullptr);
    op.precond(r, v);
    op.matvec(v, s);
    multi_dot(k.dw, r, ld, 2, s, n, k.hdev.p);  // <r,s>, <s,s>
    const double *h = op.fetch(2);
    double gnum = h[0], gden = h[1];
    double gamma = (gnum == 0.0 && gden == 0.0) ? 0.0 : gnum / gden;
    fused_update2(k.dw, gamma, v, s, x, r, r0, n, k.hdev.p);  // + <r,r>, <r0,r>
    h = op.fetch(2);
    double prev = r_norm, res_new = h[1];
    r_norm = std::sqrt(h[0]);
    k.history.push_back(r_norm);
    if (k.print_level > 1 && root) printf("% 5d    %e    %f   %e\n", iter, r_norm, prev > 0 ? r_norm / prev : 0.0, b_norm > 0 ? r_norm / b_norm : r_norm);
    if (r_norm <= epsilon && iter >= k.min_iter) {
      op.residual(x, b, r);
      r_norm = std::sqrt(op.dot(r, r));
      if (r_norm <= epsilon) { k.converged = 1; break; }
      res_new = op.dot(r0, r);  // continue from the true residual
    }
    if (std::fabs(res) < epsmac) { rc = 256; break; }
    double beta = 1.0 / res;
    res = res_new;
    beta *= res;
    if (std::fabs(gamma) < epsmac) { vec_axpy(-gamma, q, p, n); rc = 256; break; }
    // p = (beta alpha / gamma) (p - gamma q) + r
    double c = beta * alpha / gamma;
    vec_axpby(-gamma * c, q, c, p, n);
    vec_axpy(1.0, r, p, n);
  }
  k.num_iterations = iter;
  k.rel_residual_norm = (b_norm > 0.0) ? r_norm / b_norm : r_norm;
  if (!k.converged && iter >= k.max_iter && epsilon > 0) rc = 256;
  return rc;
}

}  // namespace

int krylov_solve(Krylov &k, ParCsr &A, ParVec &b, ParVec &x) {
  ensure_init();
  int nvec = (k.method == Krylov::PCG) ? 3 : 6;
  B200_REQUIRE(k.W.n >= (size_t)k.ld * nvec && k.n == A.n, "Krylov solver: Solve called before Setup");
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0, stream());
  int rc = (k.method == Krylov::PCG) ? pcg_solve(k, A, b, x) : bicgstab_solve(k, A, b, x);
  cudaEventRecord(e1, stream());
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  k.solve_seconds = ms * 1e-3;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (k.print_level > 1 && ctx().rank == 0) printf("\n\nFinal L2 norm of residual: %e\n\n", k.rel_residual_norm);
  return rc;
}

}  // namespace b200
