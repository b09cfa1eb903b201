// abi_amg.cu -- hypre-named C ABI of the BoomerAMG preconditioner (include/HYPRE.h; option surface of
// /root/reference/src/HypreSystem.cpp:119-326) and the level-dump internals used at HypreSystem.cpp:700-714.
#include "abi_util.h"

using namespace b200;

#define AMG_SETTER(NAME, TYPE, CHECK, STMT)                       \
  HYPRE_Int NAME(HYPRE_Solver solver, TYPE v) {                   \
    API_ARG(solver, #NAME ": null solver");                       \
    API_ARG(CHECK, #NAME ": value out of range");                 \
    API_BEGIN                                                     \
    AmgHandle *h = as_amg(solver);                                \
    AmgOptions &o = h->amg.opt;                                   \
    (void)o;                                                      \
    STMT;                                                         \
    API_END                                                       \
  }

// the Krylov solvers skip clearing the preconditioner's output only for entry points that honour the zero-guess hint
namespace b200 {
bool precond_honours_zero_hint(PrecondFn f) {
  return f == reinterpret_cast<PrecondFn>(&HYPRE_BoomerAMGSolve) || f == reinterpret_cast<PrecondFn>(&HYPRE_ILUSolve);
}
}  // namespace b200

extern "C" {

HYPRE_Int HYPRE_BoomerAMGCreate(HYPRE_Solver *solver) {
  API_ARG(solver, "HYPRE_BoomerAMGCreate: null output");
  API_BEGIN
  AmgHandle *h = new AmgHandle();
  // hypre's own defaults; the mini-app overrides most of them (HypreSystem.cpp:123-158)
  AmgOptions &o = h->amg.opt;
  o.coarsen_type = 10;
  o.interp_type = 6;
  for (int k = 0; k < 3; ++k) o.relax_type[k] = 13;
  o.relax_type[3] = 9;
  o.relax_order = 0;
  o.max_levels = 25;
  o.max_iter = 20;
  o.tol = 1e-7;
  o.strong_threshold = 0.25;
  *solver = reinterpret_cast<HYPRE_Solver>(static_cast<SolverBase *>(h));
  API_END
}
HYPRE_Int HYPRE_BoomerAMGDestroy(HYPRE_Solver solver) {
  API_BEGIN
  if (solver) {
    if (ctx().initialised) B200_CUDA(cudaStreamSynchronize(stream()));
    delete as_amg(solver);
  }
  API_END
}
HYPRE_Int HYPRE_BoomerAMGSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector, HYPRE_ParVector) {
  API_ARG(solver && A, "HYPRE_BoomerAMGSetup: null argument");
  API_BEGIN
  AmgHandle *h = as_amg(solver);
  B200_REQUIRE(h->smooth_type == 0 || h->smooth_num_levels == 0, "BoomerAMG: complex smoothers (smooth_type, incl. ILU) are not built yet");
  B200_REQUIRE(h->agg_num_levels == 0, "BoomerAMG: aggressive coarsening (agg_num_levels > 0) is not built yet");
  B200_REQUIRE(h->non_galerkin_tol == 0.0, "BoomerAMG: non-Galerkin coarse operators are not built yet");
  amg_setup(h->amg, *as_par(A));
  API_END
}
HYPRE_Int HYPRE_BoomerAMGSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x) {
  API_ARG(solver && A && b && x, "HYPRE_BoomerAMGSolve: null argument");
  API_BEGIN
  AmgHandle *h = as_amg(solver);
  ParVec *vb = as_vec(b), *vx = as_vec(x);
  const AmgOptions &o = h->amg.opt;
  if (o.max_iter == 1 && o.tol == 0.0) {  // preconditioner use (HypreSystem.cpp:154-155): exactly one cycle, no norms
    if (precond_zero_guess_hint()) {
      amg_vcycle(h->amg, vb->p, vx->p);
    } else {  // x is a genuine initial guess: x += M^-1 (b - A x)
      int n = as_par(A)->n;
      DVec<double> r(((size_t)n + 1) & ~(size_t)1), z(((size_t)n + 1) & ~(size_t)1);
      SpmvArgs a; a.x = vx->p; a.y = r.p; a.b = vb->p;
      par_spmv(*as_par(A), SPMV_RES, a);
      amg_vcycle(h->amg, r.p, z.p);
      vec_axpy(1.0, z.p, vx->p, n);
    }
    h->num_iterations = 1;
  } else {
    // used as a preconditioner with max_iter > 1 or tol > 0: the Krylov solver's output vector counts as zero
    if (precond_zero_guess_hint()) vec_fill(vx->p, 0.0, as_par(A)->n);
    amg_solve(h->amg, *as_par(A), vb->p, vx->p, &h->num_iterations, &h->rel_resid);
  }
  API_END
}

AMG_SETTER(HYPRE_BoomerAMGSetPrintLevel, HYPRE_Int, true, o.print_level = v)
AMG_SETTER(HYPRE_BoomerAMGSetDebugFlag, HYPRE_Int, true, o.debug_flag = v)
AMG_SETTER(HYPRE_BoomerAMGSetCoarsenType, HYPRE_Int, v >= 0, o.coarsen_type = v)
AMG_SETTER(HYPRE_BoomerAMGSetCycleType, HYPRE_Int, v == 1, o.cycle_type = v)  // V-cycle only
AMG_SETTER(HYPRE_BoomerAMGSetRelaxType, HYPRE_Int, v >= 0, { o.relax_type[0] = o.relax_type[1] = o.relax_type[2] = v; o.relax_type[3] = 9; })
AMG_SETTER(HYPRE_BoomerAMGSetNumSweeps, HYPRE_Int, v >= 0, { for (int k = 0; k < 4; ++k) o.num_sweeps[k] = v; o.num_sweeps[3] = 1; })
AMG_SETTER(HYPRE_BoomerAMGSetSmoothNumSweeps, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetTol, HYPRE_Real, v >= 0, o.tol = v)
AMG_SETTER(HYPRE_BoomerAMGSetMaxIter, HYPRE_Int, v >= 0, o.max_iter = v)
AMG_SETTER(HYPRE_BoomerAMGSetRelaxOrder, HYPRE_Int, v == 0 || v == 1, o.relax_order = v)
AMG_SETTER(HYPRE_BoomerAMGSetMaxLevels, HYPRE_Int, v >= 1, o.max_levels = v)
AMG_SETTER(HYPRE_BoomerAMGSetStrongThreshold, HYPRE_Real, v >= 0 && v <= 1, o.strong_threshold = v)
AMG_SETTER(HYPRE_BoomerAMGSetMaxRowSum, HYPRE_Real, v > 0 && v <= 1, o.max_row_sum = v)
AMG_SETTER(HYPRE_BoomerAMGSetNonGalerkinTol, HYPRE_Real, v >= 0, h->non_galerkin_tol = v)
AMG_SETTER(HYPRE_BoomerAMGSetVariant, HYPRE_Int, true, h->variant = v)
AMG_SETTER(HYPRE_BoomerAMGSetRAP2, HYPRE_Int, true, o.rap2 = v)  // the product is always formed as R (A P) in Gustavson order
AMG_SETTER(HYPRE_BoomerAMGSetKeepTranspose, HYPRE_Int, true, o.keep_transpose = v)  // R = P^T is always stored
AMG_SETTER(HYPRE_BoomerAMGSetInterpType, HYPRE_Int, v >= 0, o.interp_type = v)
AMG_SETTER(HYPRE_BoomerAMGSetMinCoarseSize, HYPRE_Int, v >= 0, o.min_coarse_size = v)
AMG_SETTER(HYPRE_BoomerAMGSetMaxCoarseSize, HYPRE_Int, v >= 1, o.max_coarse_size = v)
AMG_SETTER(HYPRE_BoomerAMGSetPMaxElmts, HYPRE_Int, v >= 0, o.p_max_elmts = v)
AMG_SETTER(HYPRE_BoomerAMGSetAggPMaxElmts, HYPRE_Int, v >= 0, (void)v)  // only used on aggressive levels (none are built)
AMG_SETTER(HYPRE_BoomerAMGSetAggNumLevels, HYPRE_Int, v >= 0, h->agg_num_levels = v)
AMG_SETTER(HYPRE_BoomerAMGSetAggInterpType, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetTruncFactor, HYPRE_Real, v >= 0 && v < 1, o.trunc_factor = v)
AMG_SETTER(HYPRE_BoomerAMGSetSmoothType, HYPRE_Int, true, h->smooth_type = v)
AMG_SETTER(HYPRE_BoomerAMGSetSmoothNumLevels, HYPRE_Int, v >= 0, h->smooth_num_levels = v)
AMG_SETTER(HYPRE_BoomerAMGSetRelaxWt, HYPRE_Real, true, o.relax_weight = v)
AMG_SETTER(HYPRE_BoomerAMGSetChebyOrder, HYPRE_Int, v >= 1 && v <= 4, o.cheby_order = v)
AMG_SETTER(HYPRE_BoomerAMGSetChebyFraction, HYPRE_Real, v > 0 && v <= 1, o.cheby_fraction = v)
AMG_SETTER(HYPRE_BoomerAMGSetChebyEigEst, HYPRE_Int, v >= 0, o.cheby_eig_est = v)
// ILU-smoother knobs: recorded nowhere, rejected at Setup through smooth_type (see include/HYPRE.h)
AMG_SETTER(HYPRE_BoomerAMGSetILUType, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILULevel, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILULocalReordering, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILUMaxRowNnz, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILUMaxIter, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILUDroptol, HYPRE_Real, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILUTriSolve, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILULowerJacobiIters, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILUUpperJacobiIters, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILUIterSetupType, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILUIterSetupOption, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILUIterSetupMaxIter, HYPRE_Int, true, (void)v)
AMG_SETTER(HYPRE_BoomerAMGSetILUIterSetupTolerance, HYPRE_Real, true, (void)v)

HYPRE_Int HYPRE_ILUSetTriSolve(HYPRE_Solver solver, HYPRE_Int v) {
  // the mini-app also calls this on the BoomerAMG handle (HypreSystem.cpp:306): harmless no-op there
  API_ARG(solver, "HYPRE_ILUSetTriSolve: null solver");
  API_BEGIN
  if (as_solver(solver)->kind == SolverBase::ILU) as_ilu(solver)->ilu.tri_solve = v;
  API_END
}
HYPRE_Int HYPRE_BoomerAMGSetCycleRelaxType(HYPRE_Solver solver, HYPRE_Int relax_type, HYPRE_Int k) {
  API_ARG(solver && k >= 1 && k <= 3 && relax_type >= 0, "HYPRE_BoomerAMGSetCycleRelaxType: bad argument");
  API_BEGIN as_amg(solver)->amg.opt.relax_type[k] = relax_type; API_END
}
HYPRE_Int HYPRE_BoomerAMGSetCycleNumSweeps(HYPRE_Solver solver, HYPRE_Int num_sweeps, HYPRE_Int k) {
  API_ARG(solver && k >= 1 && k <= 3 && num_sweeps >= 0, "HYPRE_BoomerAMGSetCycleNumSweeps: bad argument");
  API_BEGIN as_amg(solver)->amg.opt.num_sweeps[k] = num_sweeps; API_END
}
HYPRE_Int HYPRE_BoomerAMGSetLevelNonGalerkinTol(HYPRE_Solver solver, HYPRE_Real tol, HYPRE_Int) {
  API_ARG(solver && tol >= 0, "HYPRE_BoomerAMGSetLevelNonGalerkinTol: bad argument");
  API_BEGIN if (tol > 0) as_amg(solver)->non_galerkin_tol = tol; API_END
}
HYPRE_Int HYPRE_BoomerAMGGetNumIterations(HYPRE_Solver solver, HYPRE_Int *num_iterations) {
  API_ARG(solver && num_iterations, "HYPRE_BoomerAMGGetNumIterations: null argument");
  API_BEGIN *num_iterations = as_amg(solver)->num_iterations; API_END
}
HYPRE_Int HYPRE_BoomerAMGGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *rel_resid_norm) {
  API_ARG(solver && rel_resid_norm, "HYPRE_BoomerAMGGetFinalRelativeResidualNorm: null argument");
  API_BEGIN *rel_resid_norm = as_amg(solver)->rel_resid; API_END
}

hypre_ParCSRMatrix **hypre_ParAMGDataAArray(hypre_ParAMGData *amg_data) {
  try {
    return reinterpret_cast<hypre_ParCSRMatrix **>(as_amg(amg_data)->amg.a_array.data());
  } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); return nullptr; }
}
HYPRE_Int hypre_ParAMGDataNumLevels(hypre_ParAMGData *amg_data) {
  try {
    return (HYPRE_Int)as_amg(amg_data)->amg.lev.size();
  } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); return 0; }
}

}  // extern "C"
