// abi.cu -- hypre-named C ABI of libb200hypre (include/HYPRE.h): global state, IJ matrix / vector, ParCSR GMRES.
// Every function cites the reference call site it serves in include/HYPRE.h; semantics follow SURVEY.md §8b.
#include "abi_util.h"
#include <cstring>
#include <mutex>

namespace b200 {

namespace {
int g_error_flag = 0;
std::string g_last_error;
}  // namespace

int set_error(int code, const char *what) {
  g_error_flag |= code;
  g_last_error = what ? what : "";
  fprintf(stderr, "[b200hypre] error %d: %s\n", code, g_last_error.c_str());
  return g_error_flag;
}
GmresHandle *as_gmres(HYPRE_Solver s) {
  SolverBase *b = as_solver(s);
  B200_REQUIRE(b && b->kind == SolverBase::GMRES, "handle is not a ParCSR GMRES solver");
  return static_cast<GmresHandle *>(b);
}
KrylovHandle *as_krylov(HYPRE_Solver s, int method) {
  SolverBase *b = as_solver(s);
  B200_REQUIRE(b && b->kind == SolverBase::KRYLOV && static_cast<KrylovHandle *>(b)->k.method == method,
               method == Krylov::PCG ? "handle is not a ParCSR PCG solver" : "handle is not a ParCSR BiCGSTAB solver");
  return static_cast<KrylovHandle *>(b);
}
IluHandle *as_ilu(HYPRE_Solver s) {
  SolverBase *b = as_solver(s);
  B200_REQUIRE(b && b->kind == SolverBase::ILU, "handle is not an ILU solver");
  return static_cast<IluHandle *>(b);
}
int accumulate_error(int code) {
  g_error_flag |= code;
  return g_error_flag;
}
AmgHandle *as_amg(HYPRE_Solver s) {
  SolverBase *b = as_solver(s);
  B200_REQUIRE(b && b->kind == SolverBase::AMG, "handle is not a BoomerAMG solver");
  return static_cast<AmgHandle *>(b);
}

namespace {


// hypre's IJ text grammar: one file per rank "<name>.%05d"; header then "row col value" lines
void print_parcsr(const ParCsr &A, i64 base_i, i64 base_j, const char *filename, bool ij_header_with_cols) {
  char fn[4096];
  snprintf(fn, sizeof(fn), "%s.%05d", filename, ctx().rank);
  FILE *f = fopen(fn, "w");
  B200_REQUIRE(f, std::string("cannot open ") + fn);
  std::vector<int> dp = A.diag.ptr.to_host(), dc = A.diag.col.to_host();
  std::vector<double> dv = A.diag.val.to_host();
  std::vector<int> op, oc;
  std::vector<double> ov;
  if (A.offd.nnz > 0) { op = A.offd.ptr.to_host(); oc = A.offd.col.to_host(); ov = A.offd.val.to_host(); }
  (void)ij_header_with_cols;
  fprintf(f, "%lld %lld %lld %lld\n", A.row_start + base_i, A.row_start + A.n - 1 + base_i, A.col_start + base_j,
          A.col_start + A.n_cols_local - 1 + base_j);
  for (int i = 0; i < A.n; ++i) {
    i64 gi = A.row_start + i + base_i;
    for (int k = dp[i]; k < dp[i + 1]; ++k) fprintf(f, "%lld %lld %.14e\n", gi, A.col_start + dc[k] + base_j, dv[k]);
    if (!op.empty())
      for (int k = op[i]; k < op[i + 1]; ++k) fprintf(f, "%lld %lld %.14e\n", gi, A.col_map_offd[oc[k]] + base_j, ov[k]);
  }
  fclose(f);
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

HYPRE_Int HYPRE_GetError(void) { return g_error_flag; }
HYPRE_Int HYPRE_ClearAllErrors(void) { g_error_flag = 0; g_last_error.clear(); return 0; }
const char *b200hypre_last_error(void) { return g_last_error.c_str(); }

HYPRE_Int HYPRE_Init(void) {
  API_BEGIN
  ensure_init();
  API_END
}
HYPRE_Int HYPRE_Finalize(void) {
  API_BEGIN
  if (ctx().initialised) B200_CUDA(cudaStreamSynchronize(stream()));
  API_END
}
HYPRE_Int HYPRE_SetMemoryLocation(HYPRE_MemoryLocation loc) {
  API_ARG(loc == HYPRE_MEMORY_DEVICE, "HYPRE_SetMemoryLocation: libb200hypre keeps all solver data in device memory (no CPU path)");
  return 0;
}
HYPRE_Int HYPRE_SetExecutionPolicy(HYPRE_ExecutionPolicy pol) {
  API_ARG(pol == HYPRE_EXEC_DEVICE, "HYPRE_SetExecutionPolicy: libb200hypre executes on the device only (no CPU path)");
  return 0;
}
HYPRE_Int HYPRE_SetSpGemmUseVendor(HYPRE_Int) { return 0; }
HYPRE_Int HYPRE_SetSpMVUseVendor(HYPRE_Int) { return 0; }
HYPRE_Int HYPRE_SetSpTransUseVendor(HYPRE_Int) { return 0; }
HYPRE_Int HYPRE_SetGPUMemoryPoolSize(HYPRE_Int, HYPRE_Int, HYPRE_Int, size_t) { return 0; }
HYPRE_Int hypre_ResetDeviceRandGenerator(unsigned long long seed, unsigned long long offset) {
  // recorded for completeness; the PMIS measure uses hypre's host LCG stream (seed 2747 + rank) generated on the device,
  // which is what the parity contract names, so results do not depend on this call
  ctx().rng_seed = seed;
  ctx().rng_offset = offset;
  return 0;
}
void *b200hypre_malloc(size_t bytes, HYPRE_MemoryLocation loc) {
  if (loc == HYPRE_MEMORY_HOST) return malloc(bytes ? bytes : 1);
  void *p = nullptr;
  try {
    ensure_init();
    if (cudaMalloc(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); set_error(HYPRE_ERROR_MEMORY, "device malloc failed"); return nullptr; }
  } catch (const std::exception &e) { set_error(HYPRE_ERROR_GENERIC, e.what()); return nullptr; }
  return p;
}
void *b200hypre_calloc(size_t bytes, HYPRE_MemoryLocation loc) {
  void *p = b200hypre_malloc(bytes, loc);
  if (!p) return p;
  if (loc == HYPRE_MEMORY_HOST) memset(p, 0, bytes);
  else cudaMemset(p, 0, bytes);
  return p;
}
void b200hypre_free(void *p, HYPRE_MemoryLocation loc) {
  if (!p) return;
  if (loc == HYPRE_MEMORY_HOST) { free(p); return; }
  // device arrays handed out by the library come from the stream-ordered pool or from cudaMalloc; cudaFree handles both
  if (ctx().initialised) cudaStreamSynchronize(stream());
  cudaFree(p);
}
void b200hypre_memcpy(void *dst, const void *src, size_t bytes, HYPRE_MemoryLocation, HYPRE_MemoryLocation) {
  if (bytes) cudaMemcpy(dst, src, bytes, cudaMemcpyDefault);
}

// ---- IJ matrix -----------------------------------------------------------------------------------------------------------
HYPRE_Int HYPRE_IJMatrixCreate(MPI_Comm, HYPRE_BigInt ilower, HYPRE_BigInt iupper, HYPRE_BigInt jlower, HYPRE_BigInt jupper, HYPRE_IJMatrix *matrix) {
  API_ARG(matrix, "HYPRE_IJMatrixCreate: null output");
  API_ARG(iupper >= ilower - 1 && jupper >= jlower - 1, "HYPRE_IJMatrixCreate: bad range");
  API_ARG(iupper - ilower + 1 < (1LL << 31) - 2, "HYPRE_IJMatrixCreate: more than 2^31 local rows");
  API_BEGIN
  IJMatrix *m = new IJMatrix();
  m->ilower = ilower; m->iupper = iupper; m->jlower = jlower; m->jupper = jupper;
  *matrix = reinterpret_cast<HYPRE_IJMatrix>(m);
  API_END
}
HYPRE_Int HYPRE_IJMatrixSetObjectType(HYPRE_IJMatrix, HYPRE_Int type) {
  API_ARG(type == HYPRE_PARCSR, "HYPRE_IJMatrixSetObjectType: only HYPRE_PARCSR is supported");
  return 0;
}
HYPRE_Int HYPRE_IJMatrixInitialize(HYPRE_IJMatrix matrix) {
  API_ARG(matrix, "HYPRE_IJMatrixInitialize: null matrix");
  API_BEGIN
  ensure_init();
  IJMatrix *m = as_ijm(matrix);
  m->staged.clear();
  m->assembled = false;
  API_END
}
HYPRE_Int HYPRE_IJMatrixGetObject(HYPRE_IJMatrix matrix, void **object) {
  API_ARG(matrix && object, "HYPRE_IJMatrixGetObject: null argument");
  *object = &as_ijm(matrix)->par;   // borrowed: owned by the IJ matrix (HypreSystem.cpp:555,606)
  return 0;
}
HYPRE_Int HYPRE_IJMatrixSetConstantValues(HYPRE_IJMatrix matrix, HYPRE_Complex value) {
  API_ARG(matrix, "HYPRE_IJMatrixSetConstantValues: null matrix");
  API_BEGIN
  IJMatrix *m = as_ijm(matrix);
  if (m->assembled) {
    vec_fill(m->par.diag.val.p, value, m->par.diag.nnz);
    vec_fill(m->par.offd.val.p, value, m->par.offd.nnz);
    m->par.sell = DSell();
    m->par.finalize_spmv();
  } else {
    B200_REQUIRE(m->staged.empty() || value == 0.0, "HYPRE_IJMatrixSetConstantValues before Assemble is only supported on an empty matrix");
  }
  API_END
}
static HYPRE_Int ij_set(HYPRE_IJMatrix matrix, HYPRE_Int nrows, HYPRE_Int *ncols, const HYPRE_BigInt *rows, const HYPRE_Int *row_indexes,
                        const HYPRE_BigInt *cols, const HYPRE_Complex *values, int op) {
  API_ARG(matrix && (nrows == 0 || (rows && cols && values)), "HYPRE_IJMatrixSetValues: null argument");
  API_ARG(nrows >= 0, "HYPRE_IJMatrixSetValues: negative nrows");
  API_BEGIN
  ij_matrix_set(*as_ijm(matrix), nrows, ncols, rows, row_indexes, cols, values, op);
  API_END
}
HYPRE_Int HYPRE_IJMatrixSetValues(HYPRE_IJMatrix matrix, HYPRE_Int nrows, HYPRE_Int *ncols, const HYPRE_BigInt *rows, const HYPRE_BigInt *cols, const HYPRE_Complex *values) {
  API_ARG(ncols, "HYPRE_IJMatrixSetValues: ncols is required (use SetValues2 with ncols == NULL for one entry per row)");
  return ij_set(matrix, nrows, ncols, rows, nullptr, cols, values, 0);
}
HYPRE_Int HYPRE_IJMatrixSetValues2(HYPRE_IJMatrix matrix, HYPRE_Int nrows, HYPRE_Int *ncols, const HYPRE_BigInt *rows, const HYPRE_Int *row_indexes, const HYPRE_BigInt *cols, const HYPRE_Complex *values) {
  return ij_set(matrix, nrows, ncols, rows, row_indexes, cols, values, 0);
}
HYPRE_Int HYPRE_IJMatrixAddToValues2(HYPRE_IJMatrix matrix, HYPRE_Int nrows, HYPRE_Int *ncols, const HYPRE_BigInt *rows, const HYPRE_Int *row_indexes, const HYPRE_BigInt *cols, const HYPRE_Complex *values) {
  return ij_set(matrix, nrows, ncols, rows, row_indexes, cols, values, 1);
}
HYPRE_Int HYPRE_IJMatrixAssemble(HYPRE_IJMatrix matrix) {
  API_ARG(matrix, "HYPRE_IJMatrixAssemble: null matrix");
  API_BEGIN
  ij_matrix_assemble(*as_ijm(matrix));
  API_END
}
HYPRE_Int HYPRE_IJMatrixGetLocalRange(HYPRE_IJMatrix matrix, HYPRE_BigInt *ilower, HYPRE_BigInt *iupper, HYPRE_BigInt *jlower, HYPRE_BigInt *jupper) {
  API_ARG(matrix, "HYPRE_IJMatrixGetLocalRange: null matrix");
  IJMatrix *m = as_ijm(matrix);
  if (ilower) *ilower = m->ilower;
  if (iupper) *iupper = m->iupper;
  if (jlower) *jlower = m->jlower;
  if (jupper) *jupper = m->jupper;
  return 0;
}
HYPRE_Int HYPRE_IJMatrixPrint(HYPRE_IJMatrix matrix, const char *filename) {
  API_ARG(matrix && filename, "HYPRE_IJMatrixPrint: null argument");
  API_BEGIN
  IJMatrix *m = as_ijm(matrix);
  B200_REQUIRE(m->assembled, "HYPRE_IJMatrixPrint: matrix is not assembled");
  print_parcsr(m->par, 0, 0, filename, true);
  API_END
}
HYPRE_Int HYPRE_IJMatrixDestroy(HYPRE_IJMatrix matrix) {
  API_BEGIN
  if (matrix) {
    if (ctx().initialised) B200_CUDA(cudaStreamSynchronize(stream()));
    delete as_ijm(matrix);
  }
  API_END
}
HYPRE_Int hypre_ParCSRMatrixPrintIJ(const hypre_ParCSRMatrix *matrix, HYPRE_Int base_i, HYPRE_Int base_j, const char *filename) {
  API_ARG(matrix && filename, "hypre_ParCSRMatrixPrintIJ: null argument");
  API_BEGIN
  print_parcsr(*reinterpret_cast<const ParCsr *>(matrix), base_i, base_j, filename, true);
  API_END
}

// ---- IJ vector -----------------------------------------------------------------------------------------------------------
HYPRE_Int HYPRE_IJVectorCreate(MPI_Comm, HYPRE_BigInt jlower, HYPRE_BigInt jupper, HYPRE_IJVector *vector) {
  API_ARG(vector, "HYPRE_IJVectorCreate: null output");
  API_ARG(jupper >= jlower - 1 && jupper - jlower + 1 < (1LL << 31) - 2, "HYPRE_IJVectorCreate: bad range");
  API_BEGIN
  IJVector *v = new IJVector();
  v->jlower = jlower;
  v->jupper = jupper;
  *vector = reinterpret_cast<HYPRE_IJVector>(v);
  API_END
}
HYPRE_Int HYPRE_IJVectorSetObjectType(HYPRE_IJVector, HYPRE_Int type) {
  API_ARG(type == HYPRE_PARCSR, "HYPRE_IJVectorSetObjectType: only HYPRE_PARCSR is supported");
  return 0;
}
HYPRE_Int HYPRE_IJVectorSetNumComponents(HYPRE_IJVector vector, HYPRE_Int num_components) {
  API_ARG(vector, "HYPRE_IJVectorSetNumComponents: null vector");
  API_ARG(num_components == 1, "HYPRE_IJVectorSetNumComponents: multi-component vectors are not built yet (segregated solves use one vector per component)");
  as_ijv(vector)->num_components = num_components;
  return 0;
}
HYPRE_Int HYPRE_IJVectorSetComponent(HYPRE_IJVector vector, HYPRE_Int component) {
  API_ARG(vector, "HYPRE_IJVectorSetComponent: null vector");
  API_ARG(component == 0, "HYPRE_IJVectorSetComponent: only component 0 exists");
  as_ijv(vector)->component = component;
  return 0;
}
HYPRE_Int HYPRE_IJVectorInitialize(HYPRE_IJVector vector) {
  API_ARG(vector, "HYPRE_IJVectorInitialize: null vector");
  API_BEGIN
  ij_vector_init(*as_ijv(vector));
  API_END
}
HYPRE_Int HYPRE_IJVectorGetObject(HYPRE_IJVector vector, void **object) {
  API_ARG(vector && object, "HYPRE_IJVectorGetObject: null argument");
  *object = &as_ijv(vector)->par;
  return 0;
}
HYPRE_Int HYPRE_IJVectorSetValues(HYPRE_IJVector vector, HYPRE_Int nvalues, const HYPRE_BigInt *indices, const HYPRE_Complex *values) {
  API_ARG(vector && (nvalues == 0 || values), "HYPRE_IJVectorSetValues: null argument");
  API_BEGIN
  ij_vector_set(*as_ijv(vector), nvalues, indices, values, 0);
  API_END
}
HYPRE_Int HYPRE_IJVectorAddToValues(HYPRE_IJVector vector, HYPRE_Int nvalues, const HYPRE_BigInt *indices, const HYPRE_Complex *values) {
  API_ARG(vector && (nvalues == 0 || values), "HYPRE_IJVectorAddToValues: null argument");
  API_BEGIN
  ij_vector_set(*as_ijv(vector), nvalues, indices, values, 1);
  API_END
}
HYPRE_Int HYPRE_IJVectorGetValues(HYPRE_IJVector vector, HYPRE_Int nvalues, const HYPRE_BigInt *indices, HYPRE_Complex *values) {
  API_ARG(vector && (nvalues == 0 || values), "HYPRE_IJVectorGetValues: null argument");
  API_BEGIN
  ij_vector_get(*as_ijv(vector), nvalues, indices, values);
  API_END
}
HYPRE_Int HYPRE_IJVectorAssemble(HYPRE_IJVector vector) {
  API_ARG(vector, "HYPRE_IJVectorAssemble: null vector");
  return 0;
}
HYPRE_Int HYPRE_IJVectorPrint(HYPRE_IJVector vector, const char *filename) {
  API_ARG(vector && filename, "HYPRE_IJVectorPrint: null argument");
  API_BEGIN
  IJVector *v = as_ijv(vector);
  char fn[4096];
  snprintf(fn, sizeof(fn), "%s.%05d", filename, ctx().rank);
  FILE *f = fopen(fn, "w");
  B200_REQUIRE(f, std::string("cannot open ") + fn);
  std::vector<double> h(v->par.n);
  if (v->par.n) B200_CUDA(cudaMemcpy(h.data(), v->par.p, sizeof(double) * v->par.n, cudaMemcpyDeviceToHost));
  fprintf(f, "%lld %lld\n", v->jlower, v->jupper);
  for (int i = 0; i < v->par.n; ++i) fprintf(f, "%lld %.14e\n", v->jlower + i, h[i]);
  fclose(f);
  API_END
}
HYPRE_Int HYPRE_IJVectorDestroy(HYPRE_IJVector vector) {
  API_BEGIN
  if (vector) {
    if (ctx().initialised) B200_CUDA(cudaStreamSynchronize(stream()));
    delete as_ijv(vector);
  }
  API_END
}
HYPRE_Int HYPRE_ParVectorSetConstantValues(HYPRE_ParVector vector, HYPRE_Complex value) {
  API_ARG(vector, "HYPRE_ParVectorSetConstantValues: null vector");
  API_BEGIN
  ParVec *v = as_vec(vector);
  vec_fill(v->p, value, v->n);
  API_END
}
HYPRE_Int HYPRE_ParVectorInnerProd(HYPRE_ParVector x, HYPRE_ParVector y, HYPRE_Real *prod) {
  API_ARG(x && y && prod, "HYPRE_ParVectorInnerProd: null argument");
  API_BEGIN
  static DotWork w;
  B200_REQUIRE(as_vec(x)->n == as_vec(y)->n, "HYPRE_ParVectorInnerProd: size mismatch");
  *prod = dot_host(w, as_vec(x)->p, as_vec(y)->p, as_vec(x)->n);
  API_END
}
HYPRE_Int HYPRE_ParCSRMatrixMatvec(HYPRE_Complex alpha, HYPRE_ParCSRMatrix A, HYPRE_ParVector x, HYPRE_Complex beta, HYPRE_ParVector y) {
  API_ARG(A && x && y, "HYPRE_ParCSRMatrixMatvec: null argument");
  API_BEGIN
  ParCsr *M = as_par(A);
  ParVec *vx = as_vec(x), *vy = as_vec(y);
  B200_REQUIRE(vx->n == M->n_cols_local && vy->n == M->n, "HYPRE_ParCSRMatrixMatvec: size mismatch");
  if (alpha == 1.0 && beta == 0.0) {
    SpmvArgs a; a.x = vx->p; a.y = vy->p;
    par_spmv(*M, SPMV_AX, a);
  } else {
    DVec<double> t(((size_t)M->n + 1) & ~(size_t)1);
    SpmvArgs a; a.x = vx->p; a.y = t.p;
    par_spmv(*M, SPMV_AX, a);
    vec_axpby(alpha, t.p, beta, vy->p, M->n);
  }
  API_END
}

// ---- ParCSR GMRES ----------------------------------------------------------------------------------------------------------
HYPRE_Int HYPRE_ParCSRGMRESCreate(MPI_Comm, HYPRE_Solver *solver) {
  API_ARG(solver, "HYPRE_ParCSRGMRESCreate: null output");
  API_BEGIN
  *solver = reinterpret_cast<HYPRE_Solver>(static_cast<SolverBase *>(new GmresHandle()));
  API_END
}
HYPRE_Int HYPRE_ParCSRGMRESDestroy(HYPRE_Solver solver) {
  API_BEGIN
  if (solver) {
    if (ctx().initialised) B200_CUDA(cudaStreamSynchronize(stream()));
    delete as_gmres(solver);
  }
  API_END
}
HYPRE_Int HYPRE_ParCSRGMRESSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {
  API_ARG(solver && A && b && x, "HYPRE_ParCSRGMRESSetup: null argument");
  API_BEGIN
  gmres_setup(as_gmres(solver)->g, *as_par(A), *as_vec(b), *as_vec(x));
  API_END
}
HYPRE_Int HYPRE_ParCSRGMRESSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {
  API_ARG(solver && A && b && x, "HYPRE_ParCSRGMRESSolve: null argument");
  int rc = 0;
  try {
    rc = gmres_solve(as_gmres(solver)->g, *as_par(A), *as_vec(b), *as_vec(x));
  } catch (const std::exception &e) { return set_error(HYPRE_ERROR_GENERIC, e.what()); }
  if (rc) { g_error_flag |= rc; return g_error_flag; }
  return 0;
}
HYPRE_Int HYPRE_ParCSRGMRESSetKDim(HYPRE_Solver solver, HYPRE_Int k_dim) {
  API_ARG(solver && k_dim >= 1 && k_dim <= 120, "HYPRE_ParCSRGMRESSetKDim: k_dim must be in [1, 120]");
  API_BEGIN as_gmres(solver)->g.k_dim = k_dim; API_END
}
HYPRE_Int HYPRE_ParCSRGMRESSetTol(HYPRE_Solver solver, HYPRE_Real tol) {
  API_ARG(solver && tol >= 0, "HYPRE_ParCSRGMRESSetTol: bad argument");
  API_BEGIN as_gmres(solver)->g.tol = tol; API_END
}
HYPRE_Int HYPRE_ParCSRGMRESSetAbsoluteTol(HYPRE_Solver solver, HYPRE_Real a_tol) {
  API_ARG(solver && a_tol >= 0, "HYPRE_ParCSRGMRESSetAbsoluteTol: bad argument");
  API_BEGIN as_gmres(solver)->g.atol = a_tol; API_END
}
HYPRE_Int HYPRE_ParCSRGMRESSetMinIter(HYPRE_Solver solver, HYPRE_Int min_iter) {
  API_ARG(solver && min_iter >= 0, "HYPRE_ParCSRGMRESSetMinIter: bad argument");
  API_BEGIN as_gmres(solver)->g.min_iter = min_iter; API_END
}
HYPRE_Int HYPRE_ParCSRGMRESSetMaxIter(HYPRE_Solver solver, HYPRE_Int max_iter) {
  API_ARG(solver && max_iter >= 0, "HYPRE_ParCSRGMRESSetMaxIter: bad argument");
  API_BEGIN as_gmres(solver)->g.max_iter = max_iter; API_END
}
HYPRE_Int HYPRE_ParCSRGMRESSetPrecond(HYPRE_Solver solver, HYPRE_PtrToParSolverFcn precond, HYPRE_PtrToParSolverFcn precond_setup, HYPRE_Solver precond_solver) {
  API_ARG(solver, "HYPRE_ParCSRGMRESSetPrecond: null solver");
  API_BEGIN
  Gmres &g = as_gmres(solver)->g;
  g.precond = reinterpret_cast<PrecondFn>(precond);
  g.precond_setup = reinterpret_cast<PrecondFn>(precond_setup);
  g.precond_data = precond_solver;
  API_END
}
HYPRE_Int HYPRE_ParCSRGMRESSetLogging(HYPRE_Solver solver, HYPRE_Int logging) {
  API_ARG(solver, "HYPRE_ParCSRGMRESSetLogging: null solver");
  API_BEGIN as_gmres(solver)->g.logging = logging; API_END
}
HYPRE_Int HYPRE_ParCSRGMRESSetPrintLevel(HYPRE_Solver solver, HYPRE_Int print_level) {
  API_ARG(solver, "HYPRE_ParCSRGMRESSetPrintLevel: null solver");
  API_BEGIN as_gmres(solver)->g.print_level = print_level; API_END
}
HYPRE_Int HYPRE_ParCSRGMRESGetNumIterations(HYPRE_Solver solver, HYPRE_Int *num_iterations) {
  API_ARG(solver && num_iterations, "HYPRE_ParCSRGMRESGetNumIterations: null argument");
  API_BEGIN *num_iterations = as_gmres(solver)->g.num_iterations; API_END
}
HYPRE_Int HYPRE_ParCSRGMRESGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *norm) {
  API_ARG(solver && norm, "HYPRE_ParCSRGMRESGetFinalRelativeResidualNorm: null argument");
  API_BEGIN *norm = as_gmres(solver)->g.rel_residual_norm; API_END
}
HYPRE_Int HYPRE_ParCSRGMRESGetConverged(HYPRE_Solver solver, HYPRE_Int *converged) {
  API_ARG(solver && converged, "HYPRE_ParCSRGMRESGetConverged: null argument");
  API_BEGIN *converged = as_gmres(solver)->g.converged; API_END
}

}  // extern "C"
