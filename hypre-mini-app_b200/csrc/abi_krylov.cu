// abi_krylov.cu -- hypre-named C ABI of the sibling Krylov solvers (include/HYPRE.h): ParCSR FlexGMRES
// (/root/reference/src/HypreSystem.cpp:406-421), COGMRES (:372-388), PCG (:440-455), BiCGSTAB (:423-438).
// FlexGMRES and COGMRES are GMRES handles with `variant` set (gmres.cu); PCG and BiCGSTAB are Krylov handles
// (krylov.cu).  Defaults at Create follow hypre's *Create functions.
#include "abi_util.h"

using namespace b200;

namespace {

int run_gmres(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {
  int rc = 0;
  try {
    rc = gmres_solve(as_gmres(solver)->g, *as_par(A), *as_vec(b), *as_vec(x));
  } catch (const std::exception &e) { return set_error(HYPRE_ERROR_GENERIC, e.what()); }
  return rc ? accumulate_error(rc) : 0;
}
int run_krylov(HYPRE_Solver solver, int method, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {
  int rc = 0;
  try {
    rc = krylov_solve(as_krylov(solver, method)->k, *as_par(A), *as_vec(b), *as_vec(x));
  } catch (const std::exception &e) { return set_error(HYPRE_ERROR_GENERIC, e.what()); }
  return rc ? accumulate_error(rc) : 0;
}
void destroy_solver(HYPRE_Solver solver) {
  if (!solver) return;
  if (ctx().initialised) B200_CUDA(cudaStreamSynchronize(stream()));
  delete as_solver(solver);
}

}  // namespace

extern "C" {

// ---- the GMRES family: FlexGMRES (variant 1), COGMRES (variant 2) ---------------------------------------------------
#define GMRES_FAMILY(NAME, VARIANT, KDIM, MAXIT)                                                                                \
  HYPRE_Int HYPRE_ParCSR##NAME##Create(MPI_Comm, HYPRE_Solver *solver) {                                                        \
    API_ARG(solver, "HYPRE_ParCSR" #NAME "Create: null output");                                                                \
    API_BEGIN                                                                                                                   \
    GmresHandle *h = new GmresHandle();                                                                                         \
    h->g.variant = VARIANT; h->g.k_dim = KDIM; h->g.max_iter = MAXIT;                                                           \
    if (VARIANT == 1) h->g.orth = 0; /* hypre's FlexGMRES orthogonalises with modified Gram-Schmidt */                          \
    *solver = reinterpret_cast<HYPRE_Solver>(static_cast<SolverBase *>(h));                                                     \
    API_END                                                                                                                     \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##Destroy(HYPRE_Solver solver) { API_BEGIN if (solver) { as_gmres(solver); destroy_solver(solver); } API_END } \
  HYPRE_Int HYPRE_ParCSR##NAME##Setup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {        \
    API_ARG(solver && A && b && x, "HYPRE_ParCSR" #NAME "Setup: null argument");                                                \
    API_BEGIN gmres_setup(as_gmres(solver)->g, *as_par(A), *as_vec(b), *as_vec(x)); API_END                                     \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##Solve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {        \
    API_ARG(solver && A && b && x, "HYPRE_ParCSR" #NAME "Solve: null argument");                                                \
    return run_gmres(solver, A, b, x);                                                                                          \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetKDim(HYPRE_Solver solver, HYPRE_Int v) {                                                     \
    API_ARG(solver && v >= 1 && v <= 120, "HYPRE_ParCSR" #NAME "SetKDim: k_dim must be in [1, 120]");                           \
    API_BEGIN as_gmres(solver)->g.k_dim = v; API_END                                                                            \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetTol(HYPRE_Solver solver, HYPRE_Real v) {                                                     \
    API_ARG(solver && v >= 0, "HYPRE_ParCSR" #NAME "SetTol: bad argument");                                                     \
    API_BEGIN as_gmres(solver)->g.tol = v; API_END                                                                              \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetAbsoluteTol(HYPRE_Solver solver, HYPRE_Real v) {                                             \
    API_ARG(solver && v >= 0, "HYPRE_ParCSR" #NAME "SetAbsoluteTol: bad argument");                                             \
    API_BEGIN as_gmres(solver)->g.atol = v; API_END                                                                             \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetMinIter(HYPRE_Solver solver, HYPRE_Int v) {                                                  \
    API_ARG(solver && v >= 0, "HYPRE_ParCSR" #NAME "SetMinIter: bad argument");                                                 \
    API_BEGIN as_gmres(solver)->g.min_iter = v; API_END                                                                         \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetMaxIter(HYPRE_Solver solver, HYPRE_Int v) {                                                  \
    API_ARG(solver && v >= 0, "HYPRE_ParCSR" #NAME "SetMaxIter: bad argument");                                                 \
    API_BEGIN as_gmres(solver)->g.max_iter = v; API_END                                                                         \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetPrecond(HYPRE_Solver solver, HYPRE_PtrToParSolverFcn precond,                                \
                                           HYPRE_PtrToParSolverFcn precond_setup, HYPRE_Solver precond_solver) {                \
    API_ARG(solver, "HYPRE_ParCSR" #NAME "SetPrecond: null solver");                                                            \
    API_BEGIN                                                                                                                   \
    Gmres &g = as_gmres(solver)->g;                                                                                             \
    g.precond = reinterpret_cast<PrecondFn>(precond);                                                                           \
    g.precond_setup = reinterpret_cast<PrecondFn>(precond_setup);                                                               \
    g.precond_data = precond_solver;                                                                                            \
    API_END                                                                                                                     \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetLogging(HYPRE_Solver solver, HYPRE_Int v) {                                                  \
    API_ARG(solver, "HYPRE_ParCSR" #NAME "SetLogging: null solver");                                                            \
    API_BEGIN as_gmres(solver)->g.logging = v; API_END                                                                          \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetPrintLevel(HYPRE_Solver solver, HYPRE_Int v) {                                               \
    API_ARG(solver, "HYPRE_ParCSR" #NAME "SetPrintLevel: null solver");                                                         \
    API_BEGIN as_gmres(solver)->g.print_level = v; API_END                                                                      \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##GetNumIterations(HYPRE_Solver solver, HYPRE_Int *out) {                                         \
    API_ARG(solver && out, "HYPRE_ParCSR" #NAME "GetNumIterations: null argument");                                             \
    API_BEGIN *out = as_gmres(solver)->g.num_iterations; API_END                                                                \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##GetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *out) {                            \
    API_ARG(solver && out, "HYPRE_ParCSR" #NAME "GetFinalRelativeResidualNorm: null argument");                                 \
    API_BEGIN *out = as_gmres(solver)->g.rel_residual_norm; API_END                                                             \
  }

GMRES_FAMILY(FlexGMRES, 1, 20, 20)
GMRES_FAMILY(COGMRES, 2, 5, 1000)
#undef GMRES_FAMILY

HYPRE_Int HYPRE_ParCSRCOGMRESSetCGS(HYPRE_Solver solver, HYPRE_Int cgs) {
  API_ARG(solver && cgs >= 0, "HYPRE_ParCSRCOGMRESSetCGS: bad argument");
  API_BEGIN as_gmres(solver)->g.cgs = cgs; API_END
}
HYPRE_Int HYPRE_ParCSRCOGMRESSetUnroll(HYPRE_Solver solver, HYPRE_Int unroll) {
  API_ARG(solver, "HYPRE_ParCSRCOGMRESSetUnroll: null solver");
  API_BEGIN as_gmres(solver)->g.unroll = unroll; API_END  // a CPU loop-unrolling hint in hypre; no effect here
}

// ---- PCG and BiCGSTAB ---------------------------------------------------------------------------------------------------
#define KRYLOV_COMMON(NAME, METHOD)                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##Create(MPI_Comm, HYPRE_Solver *solver) {                                                        \
    API_ARG(solver, "HYPRE_ParCSR" #NAME "Create: null output");                                                                \
    API_BEGIN                                                                                                                   \
    KrylovHandle *h = new KrylovHandle();                                                                                       \
    h->k.method = METHOD;                                                                                                       \
    *solver = reinterpret_cast<HYPRE_Solver>(static_cast<SolverBase *>(h));                                                     \
    API_END                                                                                                                     \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##Destroy(HYPRE_Solver solver) { API_BEGIN if (solver) { as_krylov(solver, METHOD); destroy_solver(solver); } API_END } \
  HYPRE_Int HYPRE_ParCSR##NAME##Setup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {        \
    API_ARG(solver && A && b && x, "HYPRE_ParCSR" #NAME "Setup: null argument");                                                \
    API_BEGIN krylov_setup(as_krylov(solver, METHOD)->k, *as_par(A), *as_vec(b), *as_vec(x)); API_END                           \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##Solve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {        \
    API_ARG(solver && A && b && x, "HYPRE_ParCSR" #NAME "Solve: null argument");                                                \
    return run_krylov(solver, METHOD, A, b, x);                                                                                 \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetTol(HYPRE_Solver solver, HYPRE_Real v) {                                                     \
    API_ARG(solver && v >= 0, "HYPRE_ParCSR" #NAME "SetTol: bad argument");                                                     \
    API_BEGIN as_krylov(solver, METHOD)->k.tol = v; API_END                                                                     \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetAbsoluteTol(HYPRE_Solver solver, HYPRE_Real v) {                                             \
    API_ARG(solver && v >= 0, "HYPRE_ParCSR" #NAME "SetAbsoluteTol: bad argument");                                             \
    API_BEGIN as_krylov(solver, METHOD)->k.atol = v; API_END                                                                    \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetMaxIter(HYPRE_Solver solver, HYPRE_Int v) {                                                  \
    API_ARG(solver && v >= 0, "HYPRE_ParCSR" #NAME "SetMaxIter: bad argument");                                                 \
    API_BEGIN as_krylov(solver, METHOD)->k.max_iter = v; API_END                                                                \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetPrecond(HYPRE_Solver solver, HYPRE_PtrToParSolverFcn precond,                                \
                                           HYPRE_PtrToParSolverFcn precond_setup, HYPRE_Solver precond_solver) {                \
    API_ARG(solver, "HYPRE_ParCSR" #NAME "SetPrecond: null solver");                                                            \
    API_BEGIN                                                                                                                   \
    Krylov &k = as_krylov(solver, METHOD)->k;                                                                                   \
    k.precond = reinterpret_cast<PrecondFn>(precond);                                                                           \
    k.precond_setup = reinterpret_cast<PrecondFn>(precond_setup);                                                               \
    k.precond_data = precond_solver;                                                                                            \
    API_END                                                                                                                     \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetLogging(HYPRE_Solver solver, HYPRE_Int v) {                                                  \
    API_ARG(solver, "HYPRE_ParCSR" #NAME "SetLogging: null solver");                                                            \
    API_BEGIN as_krylov(solver, METHOD)->k.logging = v; API_END                                                                 \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##SetPrintLevel(HYPRE_Solver solver, HYPRE_Int v) {                                               \
    API_ARG(solver, "HYPRE_ParCSR" #NAME "SetPrintLevel: null solver");                                                         \
    API_BEGIN as_krylov(solver, METHOD)->k.print_level = v; API_END                                                             \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##GetNumIterations(HYPRE_Solver solver, HYPRE_Int *out) {                                         \
    API_ARG(solver && out, "HYPRE_ParCSR" #NAME "GetNumIterations: null argument");                                             \
    API_BEGIN *out = as_krylov(solver, METHOD)->k.num_iterations; API_END                                                       \
  }                                                                                                                             \
  HYPRE_Int HYPRE_ParCSR##NAME##GetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *out) {                            \
    API_ARG(solver && out, "HYPRE_ParCSR" #NAME "GetFinalRelativeResidualNorm: null argument");                                 \
    API_BEGIN *out = as_krylov(solver, METHOD)->k.rel_residual_norm; API_END                                                    \
  }

KRYLOV_COMMON(PCG, Krylov::PCG)
KRYLOV_COMMON(BiCGSTAB, Krylov::BICGSTAB)
#undef KRYLOV_COMMON

HYPRE_Int HYPRE_ParCSRPCGSetTwoNorm(HYPRE_Solver solver, HYPRE_Int two_norm) {
  API_ARG(solver, "HYPRE_ParCSRPCGSetTwoNorm: null solver");
  API_BEGIN as_krylov(solver, Krylov::PCG)->k.two_norm = two_norm ? 1 : 0; API_END
}
HYPRE_Int HYPRE_ParCSRPCGSetRelChange(HYPRE_Solver solver, HYPRE_Int rel_change) {
  API_ARG(solver && rel_change == 0, "HYPRE_ParCSRPCGSetRelChange: only rel_change = 0 (hypre's default) is built");
  API_BEGIN as_krylov(solver, Krylov::PCG); API_END
}
HYPRE_Int HYPRE_ParCSRBiCGSTABSetMinIter(HYPRE_Solver solver, HYPRE_Int min_iter) {
  API_ARG(solver && min_iter >= 0, "HYPRE_ParCSRBiCGSTABSetMinIter: bad argument");
  API_BEGIN as_krylov(solver, Krylov::BICGSTAB)->k.min_iter = min_iter; API_END
}
HYPRE_Int HYPRE_ParCSRBiCGSTABSetStopCrit(HYPRE_Solver solver, HYPRE_Int stop_crit) {
  API_ARG(solver && stop_crit == 0, "HYPRE_ParCSRBiCGSTABSetStopCrit: only stop_crit = 0 (relative to ||b||, hypre's default) is built");
  API_BEGIN as_krylov(solver, Krylov::BICGSTAB); API_END
}


// ---- ILU (HypreSystem.cpp:328-370 as preconditioner, :457-497 as solver): block-Jacobi ILU(0), ilu.cu --------------------
HYPRE_Int HYPRE_ILUCreate(HYPRE_Solver *solver) {
  API_ARG(solver, "HYPRE_ILUCreate: null output");
  API_BEGIN *solver = reinterpret_cast<HYPRE_Solver>(static_cast<SolverBase *>(new IluHandle())); API_END
}
HYPRE_Int HYPRE_ILUDestroy(HYPRE_Solver solver) { API_BEGIN if (solver) { as_ilu(solver); destroy_solver(solver); } API_END }
HYPRE_Int HYPRE_ILUSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector, HYPRE_ParVector) {
  API_ARG(solver && A, "HYPRE_ILUSetup: null argument");
  API_BEGIN ilu_setup(as_ilu(solver)->ilu, *as_par(A)); API_END
}
HYPRE_Int HYPRE_ILUSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {
  API_ARG(solver && A && b && x, "HYPRE_ILUSolve: null argument");
  API_BEGIN ilu_solve(as_ilu(solver)->ilu, *as_par(A), as_vec(b)->p, as_vec(x)->p, precond_zero_guess_hint()); API_END
}
#define ILU_SETTER(NAME, TYPE, COND, STMT)                                  \
  HYPRE_Int NAME(HYPRE_Solver solver, TYPE v) {                             \
    API_ARG(solver && (COND), #NAME ": bad argument");                      \
    API_BEGIN Ilu &u = as_ilu(solver)->ilu; (void)u; STMT; API_END          \
  }
ILU_SETTER(HYPRE_ILUSetType, HYPRE_Int, v >= 0, u.type = v)                              /* only 0 is built: checked at Setup */
ILU_SETTER(HYPRE_ILUSetMaxIter, HYPRE_Int, v >= 0, u.max_iter = v)
ILU_SETTER(HYPRE_ILUSetTol, HYPRE_Real, v >= 0, u.tol = v)
ILU_SETTER(HYPRE_ILUSetLocalReordering, HYPRE_Int, v >= 0, u.reordering = v)           /* only 0 is built */
ILU_SETTER(HYPRE_ILUSetPrintLevel, HYPRE_Int, true, u.print_level = v)
ILU_SETTER(HYPRE_ILUSetLogging, HYPRE_Int, true, u.logging = v)
ILU_SETTER(HYPRE_ILUSetLevelOfFill, HYPRE_Int, v >= 0, u.fill = v)                      /* only 0 is built */
ILU_SETTER(HYPRE_ILUSetMaxNnzPerRow, HYPRE_Int, true, (void)v)                          /* ILUT parameter: no effect on ILU(0) */
ILU_SETTER(HYPRE_ILUSetDropThreshold, HYPRE_Real, true, (void)v)                        /* ILUT parameter: no effect on ILU(0) */
ILU_SETTER(HYPRE_ILUSetIterativeSetupType, HYPRE_Int, v >= 0, u.iter_setup_type = v)    /* only 0 (non-iterative) is built */
ILU_SETTER(HYPRE_ILUSetIterativeSetupOption, HYPRE_Int, true, (void)v)
ILU_SETTER(HYPRE_ILUSetIterativeSetupMaxIter, HYPRE_Int, true, (void)v)
ILU_SETTER(HYPRE_ILUSetIterativeSetupTolerance, HYPRE_Real, true, (void)v)
ILU_SETTER(HYPRE_ILUSetLowerJacobiIters, HYPRE_Int, true, (void)v)                      /* used by trisolve 0 only */
ILU_SETTER(HYPRE_ILUSetUpperJacobiIters, HYPRE_Int, true, (void)v)
#undef ILU_SETTER
HYPRE_Int HYPRE_ILUGetNumIterations(HYPRE_Solver solver, HYPRE_Int *out) {
  API_ARG(solver && out, "HYPRE_ILUGetNumIterations: null argument");
  API_BEGIN *out = as_ilu(solver)->ilu.num_iterations; API_END
}
HYPRE_Int HYPRE_ILUGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *out) {
  API_ARG(solver && out, "HYPRE_ILUGetFinalRelativeResidualNorm: null argument");
  API_BEGIN *out = as_ilu(solver)->ilu.rel_resid; API_END
}
// inspection (tests): factor values in the pattern / order of the diagonal block, and the inverted pivots
HYPRE_Int b200hypre_ilu_get_factors(HYPRE_Solver solver, HYPRE_Real *val, HYPRE_Real *dinv) {
  API_ARG(solver && val && dinv, "b200hypre_ilu_get_factors: null argument");
  API_BEGIN
  Ilu &u = as_ilu(solver)->ilu;
  B200_REQUIRE(u.setup_done, "ILU: factors requested before Setup");
  B200_CUDA(cudaMemcpyAsync(val, u.fval.p, sizeof(double) * (size_t)(u.Lm.diag.nnz + u.Um.diag.nnz - u.n), cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaMemcpyAsync(dinv, u.dinv.p, sizeof(double) * (size_t)u.n, cudaMemcpyDeviceToHost, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  API_END
}
HYPRE_Real b200hypre_ilu_setup_seconds(HYPRE_Solver solver) {
  try { return as_ilu(solver)->ilu.setup_seconds; } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); return -1.0; }
}

}  // extern "C"
