// HypreSystem.cpp -- see HypreSystem.h.  Behavioural twin of /root/reference/src/HypreSystem.cpp over libb200hypre's C ABI.
//
// Documented differences from the reference (SURVEY.md §0.4 "quirks"):
//   * check_solution reduces the per-rank verdict with the send/receive buffers the right way round (the reference's
//     MPI_Reduce at HypreSystem.cpp:827 has them swapped, so its verdict is whatever rank 0 computed);
//   * `method: boomeramg` applies its options to the solver handle (the reference applies them to the still-null
//     preconditioner handle, HypreSystem.cpp:99-110);
//   * build_27pt_stencil / build_7pt_stencil call the device generator of libb200hypre with true-neighbour halo columns
//     (the reference glue at HypreSystem.cpp:1555 maps halo indices to fictitious global columns; the input key
//     linear_system.halo_map: reference reproduces that numbering);
//   * ILU: only the default combination (ilu_type 0, fill 0, local_reordering 0, trisolve 1 = block-Jacobi ILU(0)) is built;
//     other values are passed on and rejected by HYPRE_ILUSetup.
#include "HypreSystem.h"
#include "mm_text.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <limits>
#include <stdexcept>

namespace nalu {

namespace {
using clk = std::chrono::system_clock;
double since(clk::time_point t0) { return std::chrono::duration<double>(clk::now() - t0).count(); }

struct IntSetter { const char *key; HYPRE_Int (*fn)(HYPRE_Solver, HYPRE_Int); };
struct RealSetter { const char *key; HYPRE_Int (*fn)(HYPRE_Solver, HYPRE_Real); };
}  // namespace

HypreSystem::HypreSystem(MPI_Comm comm, YAML::Node &inpfile) : comm_(comm), inpfile_(inpfile) {
  MPI_Comm_rank(comm, &iproc_);
  MPI_Comm_size(comm, &nproc_);
}

// ---- HypreSystem.cpp:16-47 ------------------------------------------------------------------------------------------------
void HypreSystem::load() {
  YAML::Node linsys = inpfile_["linear_system"];
  writeAmgMatrices_ = get_optional<bool>(linsys, "write_amg_matrices", false);
  std::string mat_format = get_optional<std::string>(linsys, "type", "matrix_market");
  if (iproc_ == 0) printf("%s : Using %s mat_format\n", __FUNCTION__, mat_format.c_str());
  if (mat_format == "matrix_market") load_matrix_market();
  else if (mat_format == "hypre_ij") load_hypre_format();
  else if (mat_format == "build_27pt_stencil") build_stencil(27);
  else if (mat_format == "build_7pt_stencil") build_stencil(7);
  else throw std::runtime_error("Invalid linear system format option: " + mat_format);
  outputSystem_ = get_optional<bool>(linsys, "write_outputs", false);
  outputSolution_ = get_optional<bool>(linsys, "write_solution", false);
}

// ---- HypreSystem.cpp:49-89 ------------------------------------------------------------------------------------------------
void HypreSystem::setup_precon_and_solver() {
  YAML::Node solver = inpfile_["solver_settings"];
  if (!solver["method"] || !solver["preconditioner"])
    throw std::runtime_error("solver_settings: 'method' and 'preconditioner' are required");
  std::string method = solver["method"].as<std::string>();
  std::string preconditioner = solver["preconditioner"].as<std::string>();
  if (iproc_ == 0) printf("%s : Using %s solver with %s preconditioner\n", __FUNCTION__, method.c_str(), preconditioner.c_str());

  if (preconditioner == "boomeramg") setup_boomeramg_precond();
  else if (preconditioner == "none") usePrecond_ = false;
  else if (preconditioner == "ilu") setup_ilu_precond();
  else throw std::runtime_error("Invalid option for preconditioner provided" + preconditioner);

  if (method == "gmres") setup_gmres();
  else if (method == "boomeramg") setup_boomeramg_solver();
  else if (method == "cg") setup_cg();
  else if (method == "bicg") setup_bicg();
  else if (method == "fgmres") setup_fgmres();
  else if (method == "cogmres") setup_cogmres();
  else if (method == "ilu") setup_ilu();
  else throw std::runtime_error("Invalid option for solver method provided: " + method);
  MPI_Barrier(comm_);
  fflush(stdout);
}

// ---- HypreSystem.cpp:91-117 -----------------------------------------------------------------------------------------------
void HypreSystem::setup_boomeramg_solver() {
  YAML::Node node = inpfile_["solver_settings"];
  HYPRE_BoomerAMGCreate(&solver_);
  HYPRE_BoomerAMGSetTol(solver_, get_optional(node, "tolerance", 1.0e-5));
  HYPRE_BoomerAMGSetMaxIter(solver_, get_optional(node, "max_iterations", 1000));
  HYPRE_BoomerAMGSetPrintLevel(solver_, get_optional(node, "print_level", 4));
  HYPRE_BoomerAMGSetCoarsenType(solver_, get_optional(node, "coarsen_type", 8));
  HYPRE_BoomerAMGSetCycleType(solver_, get_optional(node, "cycle_type", 1));
  HYPRE_BoomerAMGSetRelaxType(solver_, get_optional(node, "relax_type", 6));
  HYPRE_BoomerAMGSetNumSweeps(solver_, get_optional(node, "num_sweeps", 1));
  HYPRE_BoomerAMGSetSmoothNumSweeps(solver_, get_optional(node, "smooth_num_sweeps", 1));
  HYPRE_BoomerAMGSetRelaxOrder(solver_, get_optional(node, "relax_order", 1));
  HYPRE_BoomerAMGSetMaxLevels(solver_, get_optional(node, "max_levels", 20));
  HYPRE_BoomerAMGSetStrongThreshold(solver_, get_optional(node, "strong_threshold", 0.57));
  if (node["interp_type"]) HYPRE_BoomerAMGSetInterpType(solver_, node["interp_type"].as<int>());
  solverDestroyPtr_ = &HYPRE_BoomerAMGDestroy;
  solverSetupPtr_ = &HYPRE_BoomerAMGSetup;
  solverPrecondPtr_ = nullptr;
  solverSolvePtr_ = &HYPRE_BoomerAMGSolve;
  solverItersPtr_ = &HYPRE_BoomerAMGGetNumIterations;
  solverRelresPtr_ = &HYPRE_BoomerAMGGetFinalRelativeResidualNorm;
  usePrecond_ = false;
}

// ---- HypreSystem.cpp:119-326: YAML key -> setter, defaults as coded in the reference ---------------------
void HypreSystem::setup_boomeramg_precond() {
  YAML::Node node = inpfile_["boomeramg_settings"];
  HYPRE_BoomerAMGCreate(&precond_);
  HYPRE_BoomerAMGSetPrintLevel(precond_, get_optional(node, "print_level", 1));
  HYPRE_BoomerAMGSetDebugFlag(precond_, get_optional(node, "debug_flag", 1));
  HYPRE_BoomerAMGSetCoarsenType(precond_, get_optional(node, "coarsen_type", 8));
  HYPRE_BoomerAMGSetCycleType(precond_, get_optional(node, "cycle_type", 1));

  // per-cycle-position relaxation types / sweep counts are all-or-nothing (:129-151)
  const char *rkeys[3] = {"down_relax_type", "up_relax_type", "coarse_relax_type"};
  if (node[rkeys[0]] && node[rkeys[1]] && node[rkeys[2]]) {
    for (int k = 0; k < 3; ++k) HYPRE_BoomerAMGSetCycleRelaxType(precond_, node[rkeys[k]].as<int>(), k + 1);
  } else {
    HYPRE_BoomerAMGSetRelaxType(precond_, get_optional(node, "relax_type", 8));
  }
  const char *skeys[3] = {"num_down_sweeps", "num_up_sweeps", "num_coarse_sweeps"};
  if (node[skeys[0]] && node[skeys[1]] && node[skeys[2]]) {
    for (int k = 0; k < 3; ++k) HYPRE_BoomerAMGSetCycleNumSweeps(precond_, node[skeys[k]].as<int>(), k + 1);
  } else {
    HYPRE_BoomerAMGSetNumSweeps(precond_, get_optional(node, "num_sweeps", 1));
  }
  HYPRE_BoomerAMGSetSmoothNumSweeps(precond_, get_optional(node, "smooth_num_sweeps", 1));
  HYPRE_BoomerAMGSetTol(precond_, get_optional(node, "tolerance", 0.0));
  HYPRE_BoomerAMGSetMaxIter(precond_, get_optional(node, "max_iterations", 1));
  HYPRE_BoomerAMGSetRelaxOrder(precond_, get_optional(node, "relax_order", 1));
  HYPRE_BoomerAMGSetMaxLevels(precond_, get_optional(node, "max_levels", 20));
  HYPRE_BoomerAMGSetStrongThreshold(precond_, get_optional(node, "strong_threshold", 0.57));

  if (node["non_galerkin_tol"]) {   // :161-178
    HYPRE_BoomerAMGSetNonGalerkinTol(precond_, node["non_galerkin_tol"].as<double>());
    if (node["non_galerkin_level_tols"]) {
      YAML::Node ng = node["non_galerkin_level_tols"];
      std::vector<int> levels = ng["levels"].as<std::vector<int>>();
      std::vector<double> tol = ng["tolerances"].as<std::vector<double>>();
      if (levels.size() != tol.size()) throw std::runtime_error("Hypre Config:: Invalid non_galerkin_level_tols");
      for (size_t i = 0; i < levels.size(); i++) HYPRE_BoomerAMGSetLevelNonGalerkinTol(precond_, tol[i], levels[i]);
    }
  }
  // setters applied only when the key is present (:180-320), in the reference's order
  static const IntSetter int_opts[] = {
      {"variant", HYPRE_BoomerAMGSetVariant},
      {"rap2", HYPRE_BoomerAMGSetRAP2},
      {"keep_transpose", HYPRE_BoomerAMGSetKeepTranspose},
      {"interp_type", HYPRE_BoomerAMGSetInterpType},
      {"min_coarse_size", HYPRE_BoomerAMGSetMinCoarseSize},
      {"max_coarse_size", HYPRE_BoomerAMGSetMaxCoarseSize},
      {"pmax_elmts", HYPRE_BoomerAMGSetAggPMaxElmts},   // sic: the reference routes pmax_elmts to the Agg setter (:212)
      {"agg_num_levels", HYPRE_BoomerAMGSetAggNumLevels},
      {"agg_interp_type", HYPRE_BoomerAMGSetAggInterpType},
      {"agg_pmax_elmts", HYPRE_BoomerAMGSetAggPMaxElmts},
  };
  for (const IntSetter &o : int_opts)
    if (node[o.key]) o.fn(precond_, node[o.key].as<int>());
  if (node["trunc_factor"]) HYPRE_BoomerAMGSetTruncFactor(precond_, node["trunc_factor"].as<double>());
  static const IntSetter smooth_opts[] = {
      {"smooth_type", HYPRE_BoomerAMGSetSmoothType},
      {"smooth_num_sweeps", HYPRE_BoomerAMGSetSmoothNumSweeps},
      {"smooth_num_levels", HYPRE_BoomerAMGSetSmoothNumLevels},
      {"ilu_type", HYPRE_BoomerAMGSetILUType},
      {"ilu_level", HYPRE_BoomerAMGSetILULevel},
      {"ilu_reordering_type", HYPRE_BoomerAMGSetILULocalReordering},
      {"ilu_max_row_nnz", HYPRE_BoomerAMGSetILUMaxRowNnz},
      {"ilu_max_iter", HYPRE_BoomerAMGSetILUMaxIter},
      {"iterative_ilu_algorithm_type", HYPRE_BoomerAMGSetILUIterSetupType},
      {"iterative_ilu_setup_option", HYPRE_BoomerAMGSetILUIterSetupOption},
      {"iterative_ilu_max_iterations", HYPRE_BoomerAMGSetILUIterSetupMaxIter},
      {"ilu_tri_solve", HYPRE_BoomerAMGSetILUTriSolve},
      {"ilu_lower_jacobi_iters", HYPRE_BoomerAMGSetILULowerJacobiIters},
      {"ilu_upper_jacobi_iters", HYPRE_BoomerAMGSetILUUpperJacobiIters},
  };
  for (const IntSetter &o : smooth_opts)
    if (node[o.key]) o.fn(precond_, node[o.key].as<int>());
  static const RealSetter real_opts[] = {
      {"ilu_drop_tol", HYPRE_BoomerAMGSetILUDroptol},
      {"iterative_ilu_tolerance", HYPRE_BoomerAMGSetILUIterSetupTolerance},
  };
  for (const RealSetter &o : real_opts)
    if (node[o.key]) o.fn(precond_, node[o.key].as<double>());
  // extensions of this implementation (hypre names): Chebyshev smoother knobs, interpolation row cap, max_row_sum
  if (node["p_max_elmts"]) HYPRE_BoomerAMGSetPMaxElmts(precond_, node["p_max_elmts"].as<int>());
  if (node["cheby_order"]) HYPRE_BoomerAMGSetChebyOrder(precond_, node["cheby_order"].as<int>());
  if (node["cheby_fraction"]) HYPRE_BoomerAMGSetChebyFraction(precond_, node["cheby_fraction"].as<double>());
  if (node["max_row_sum"]) HYPRE_BoomerAMGSetMaxRowSum(precond_, node["max_row_sum"].as<double>());

  precondSetupPtr_ = &HYPRE_BoomerAMGSetup;
  precondSolvePtr_ = &HYPRE_BoomerAMGSolve;
  precondDestroyPtr_ = &HYPRE_BoomerAMGDestroy;
}

// ---- HypreSystem.cpp:390-404 ----------------------------------------------------------------------------------------------
void HypreSystem::setup_gmres() {
  YAML::Node node = inpfile_["solver_settings"];
  HYPRE_ParCSRGMRESCreate(comm_, &solver_);
  HYPRE_ParCSRGMRESSetTol(solver_, get_optional(node, "tolerance", 1.0e-5));
  HYPRE_ParCSRGMRESSetMaxIter(solver_, get_optional(node, "max_iterations", 1000));
  HYPRE_ParCSRGMRESSetKDim(solver_, get_optional(node, "kspace", 10));
  HYPRE_ParCSRGMRESSetPrintLevel(solver_, get_optional(node, "print_level", 4));
  if (node["orthogonalization"]) {   // extension: "mgs" = hypre's loop, "cgs2" = fused two-pass Gram-Schmidt (default)
    std::string o = node["orthogonalization"].as<std::string>();
    if (o != "mgs" && o != "cgs2") throw std::runtime_error("solver_settings.orthogonalization must be mgs or cgs2");
    b200hypre_gmres_set_orthogonalization(solver_, o == "mgs" ? 0 : 1);
  }
  solverDestroyPtr_ = &HYPRE_ParCSRGMRESDestroy;
  solverSetupPtr_ = &HYPRE_ParCSRGMRESSetup;
  solverPrecondPtr_ = &HYPRE_ParCSRGMRESSetPrecond;
  solverSolvePtr_ = &HYPRE_ParCSRGMRESSolve;
  solverItersPtr_ = &HYPRE_ParCSRGMRESGetNumIterations;
  solverRelresPtr_ = &HYPRE_ParCSRGMRESGetFinalRelativeResidualNorm;
}

// ---- HypreSystem.cpp:328-370 (preconditioner) and :457-497 (solver): one option surface, two YAML nodes and defaults ---------
static void ilu_options(HYPRE_Solver h, YAML::Node node, int max_iter_default, int print_level_default) {
  HYPRE_ILUSetType(h, get_optional(node, "ilu_type", 0));
  HYPRE_ILUSetMaxIter(h, get_optional(node, "max_iterations", max_iter_default));
  HYPRE_ILUSetTol(h, get_optional(node, "tolerance", 0.0));
  HYPRE_ILUSetLocalReordering(h, get_optional(node, "local_reordering", 0));
  HYPRE_ILUSetPrintLevel(h, get_optional(node, "print_level", print_level_default));
  HYPRE_ILUSetLevelOfFill(h, get_optional(node, "fill", 0));
  HYPRE_ILUSetMaxNnzPerRow(h, get_optional(node, "max_nnz_per_row", 1000));
  HYPRE_ILUSetDropThreshold(h, get_optional(node, "drop_threshold", 1.0e-2));
  HYPRE_ILUSetIterativeSetupType(h, get_optional(node, "iterative_algorithm_type", 0));
  HYPRE_ILUSetIterativeSetupOption(h, get_optional(node, "iterative_setup_option", 2));
  HYPRE_ILUSetIterativeSetupMaxIter(h, get_optional(node, "iterative_ilu_max_iterations", 1));
  HYPRE_ILUSetIterativeSetupTolerance(h, get_optional(node, "iterative_ilu_tolerance", 1e-5));
  HYPRE_ILUSetTriSolve(h, get_optional(node, "trisolve", 1));
  HYPRE_ILUSetLowerJacobiIters(h, get_optional(node, "lower_jacobi_iters", 5));
  HYPRE_ILUSetUpperJacobiIters(h, get_optional(node, "upper_jacobi_iters", 5));
}
void HypreSystem::setup_ilu_precond() {
  HYPRE_ILUCreate(&precond_);
  ilu_options(precond_, inpfile_["ilu_preconditioner_settings"], 1, 1);
  precondSetupPtr_ = &HYPRE_ILUSetup;
  precondSolvePtr_ = &HYPRE_ILUSolve;
  precondDestroyPtr_ = &HYPRE_ILUDestroy;
}
void HypreSystem::setup_ilu() {
  HYPRE_ILUCreate(&solver_);
  ilu_options(solver_, inpfile_["solver_settings"], 20, 4);
  solverDestroyPtr_ = &HYPRE_ILUDestroy;
  solverSetupPtr_ = &HYPRE_ILUSetup;
  solverPrecondPtr_ = nullptr;
  solverSolvePtr_ = &HYPRE_ILUSolve;
  solverItersPtr_ = &HYPRE_ILUGetNumIterations;
  solverRelresPtr_ = &HYPRE_ILUGetFinalRelativeResidualNorm;
  usePrecond_ = false;
}

// ---- HypreSystem.cpp:372-388, 406-421, 423-438, 440-455: the sibling Krylov methods -----------------------------------------
#define B200_KRYLOV_SETUP(FN, NAME, KDIM_LINE, EXTRA)                                         \
  void HypreSystem::FN() {                                                                    \
    YAML::Node node = inpfile_["solver_settings"];                                            \
    HYPRE_ParCSR##NAME##Create(comm_, &solver_);                                              \
    HYPRE_ParCSR##NAME##SetTol(solver_, get_optional(node, "tolerance", 1.0e-5));             \
    HYPRE_ParCSR##NAME##SetMaxIter(solver_, get_optional(node, "max_iterations", 1000));      \
    KDIM_LINE;                                                                                \
    HYPRE_ParCSR##NAME##SetPrintLevel(solver_, get_optional(node, "print_level", 4));         \
    EXTRA;                                                                                    \
    solverDestroyPtr_ = &HYPRE_ParCSR##NAME##Destroy;                                         \
    solverSetupPtr_ = &HYPRE_ParCSR##NAME##Setup;                                             \
    solverPrecondPtr_ = &HYPRE_ParCSR##NAME##SetPrecond;                                      \
    solverSolvePtr_ = &HYPRE_ParCSR##NAME##Solve;                                             \
    solverItersPtr_ = &HYPRE_ParCSR##NAME##GetNumIterations;                                  \
    solverRelresPtr_ = &HYPRE_ParCSR##NAME##GetFinalRelativeResidualNorm;                     \
  }
B200_KRYLOV_SETUP(setup_cogmres, COGMRES, HYPRE_ParCSRCOGMRESSetKDim(solver_, get_optional(node, "kspace", 10)),
                  HYPRE_ParCSRCOGMRESSetCGS(solver_, get_optional(node, "cgs", 0)))
B200_KRYLOV_SETUP(setup_fgmres, FlexGMRES, HYPRE_ParCSRFlexGMRESSetKDim(solver_, get_optional(node, "kspace", 10)), (void)0)
B200_KRYLOV_SETUP(setup_bicg, BiCGSTAB, (void)0 /* kspace is commented out at :431 */, (void)0)
B200_KRYLOV_SETUP(setup_cg, PCG, (void)0 /* kspace is commented out at :448 */, (void)0)
#undef B200_KRYLOV_SETUP

// ---- HypreSystem.cpp:499-523: the caller owns and destroys IJ objects, solver and preconditioner separately -------------------
void HypreSystem::destroy_system() {
  if (mat_) HYPRE_IJMatrixDestroy(mat_);
  for (auto &v : rhs_) if (v) HYPRE_IJVectorDestroy(v);
  for (auto &v : sln_) if (v) HYPRE_IJVectorDestroy(v);
  for (auto &v : slnRef_) if (v) HYPRE_IJVectorDestroy(v);
  if (solver_) solverDestroyPtr_(solver_);
  if (precond_) precondDestroyPtr_(precond_);
  mat_ = nullptr;
  solver_ = precond_ = nullptr;
  rhs_.clear(); sln_.clear(); slnRef_.clear();
  if (d_vector_indices_) hypre_TFree(d_vector_indices_, HYPRE_MEMORY_DEVICE);
  if (d_vector_vals_) hypre_TFree(d_vector_vals_, HYPRE_MEMORY_DEVICE);
  if (d_rows_) hypre_TFree(d_rows_, HYPRE_MEMORY_DEVICE);
  if (d_cols_) hypre_TFree(d_cols_, HYPRE_MEMORY_DEVICE);
  if (d_vals_) hypre_TFree(d_vals_, HYPRE_MEMORY_DEVICE);
}

// ---- HypreSystem.cpp:525-544: block rows, the first (N mod P) ranks get one extra row ------------------------------------------
void HypreSystem::init_row_decomposition() {
  if (iproc_ == 0) printf("\tComputing row decomposition\n");
  HYPRE_BigInt starts[2];
  HYPRE_BigInt q = totalRows_ / nproc_, r = totalRows_ % nproc_;
  starts[0] = q * iproc_ + std::min<HYPRE_BigInt>(iproc_, r);
  starts[1] = q * (iproc_ + 1) + std::min<HYPRE_BigInt>(iproc_ + 1, r);
  iLower_ = starts[0];
  iUpper_ = starts[1] - 1;
  jLower_ = iLower_;
  jUpper_ = iUpper_;
  numRows_ = (HYPRE_Int)(iUpper_ - iLower_ + 1);
  MPI_Barrier(comm_);
  std::cout << "\tRank: " << std::setw(4) << iproc_ << " :: iLower = " << std::setw(9) << iLower_ << "; iUpper = " << std::setw(9)
            << iUpper_ << "; numRows = " << numRows_ << std::endl;
  MPI_Barrier(comm_);
  fflush(stdout);
}

// ---- HypreSystem.cpp:546-598 ------------------------------------------------------------------------------------------------
void HypreSystem::init_system() {
  MPI_Barrier(comm_);
  auto start = clk::now();
  if (iproc_ == 0) printf("\tInitializing HYPRE data structures\n");
  HYPRE_IJMatrixCreate(comm_, iLower_, iUpper_, jLower_, jUpper_, &mat_);
  HYPRE_IJMatrixSetObjectType(mat_, HYPRE_PARCSR);
  HYPRE_IJMatrixInitialize(mat_);
  HYPRE_IJMatrixGetObject(mat_, (void **)&parMat_);
  HYPRE_IJMatrixSetConstantValues(mat_, 0.0);
  auto make = [&](HYPRE_IJVector &v, HYPRE_ParVector &pv) {
    HYPRE_IJVectorCreate(comm_, iLower_, iUpper_, &v);
    HYPRE_IJVectorSetObjectType(v, HYPRE_PARCSR);
    HYPRE_IJVectorSetNumComponents(v, numVectors_);
    HYPRE_IJVectorInitialize(v);
    HYPRE_IJVectorGetObject(v, (void **)&pv);
    HYPRE_ParVectorSetConstantValues(pv, 0.0);
  };
  rhs_.assign(numSolves_, nullptr); parRhs_.assign(numSolves_, nullptr);
  sln_.assign(numSolves_, nullptr); parSln_.assign(numSolves_, nullptr);
  if (checkSolution_) { slnRef_.assign(numSolves_, nullptr); parSlnRef_.assign(numSolves_, nullptr); }
  for (int i = 0; i < numSolves_; ++i) {
    make(rhs_[i], parRhs_[i]);
    make(sln_[i], parSln_[i]);
    if (checkSolution_) make(slnRef_[i], parSlnRef_[i]);
  }
  MPI_Barrier(comm_);
  timers_.emplace_back("Initialize system", since(start));
  MPI_Barrier(comm_);
  fflush(stdout);
}

// ---- HypreSystem.cpp:600-636 ------------------------------------------------------------------------------------------------
void HypreSystem::assemble_system() {
  auto start = clk::now();
  if (iproc_ == 0) printf("Assembling HYPRE data structures\n");
  HYPRE_IJMatrixAssemble(mat_);
  HYPRE_IJMatrixGetObject(mat_, (void **)&parMat_);
  for (int i = 0; i < numSolves_; ++i) {
    HYPRE_IJVectorAssemble(rhs_[i]);
    HYPRE_IJVectorAssemble(sln_[i]);
    HYPRE_IJVectorGetObject(rhs_[i], (void **)&(parRhs_[i]));
    HYPRE_IJVectorGetObject(sln_[i], (void **)&(parSln_[i]));
    if (checkSolution_) {
      HYPRE_IJVectorAssemble(slnRef_[i]);
      HYPRE_IJVectorGetObject(slnRef_[i], (void **)&(parSlnRef_[i]));
    }
  }
  MPI_Barrier(comm_);
  timers_.emplace_back("Assemble system", since(start));
  MPI_Barrier(comm_);
  // staging buffers are caller-owned and released after Assemble (:626-633)
  if (d_vector_indices_) hypre_TFree(d_vector_indices_, HYPRE_MEMORY_DEVICE);
  if (d_vector_vals_) hypre_TFree(d_vector_vals_, HYPRE_MEMORY_DEVICE);
  if (d_rows_) hypre_TFree(d_rows_, HYPRE_MEMORY_DEVICE);
  if (d_cols_) hypre_TFree(d_cols_, HYPRE_MEMORY_DEVICE);
  if (d_vals_) hypre_TFree(d_vals_, HYPRE_MEMORY_DEVICE);
  checkMemory();
}

// ---- HypreSystem.cpp:638-671 ------------------------------------------------------------------------------------------------
void HypreSystem::checkMemory() {
  double free_gb = 0, total_gb = 0;
  char name[256] = "";
  int device = 0, count = 0, major = 0, minor = 0;
  b200hypre_device_info(name, (HYPRE_Int)sizeof(name), &device, &count, &major, &minor, &free_gb, &total_gb);
  if (iproc_ == 0)
    printf("\trank=%d : %s %s %d : %s (cc=%d.%d): device=%d of %d : free memory=%1.8g GB, total memory=%1.8g GB\n", iproc_,
           __FUNCTION__, __FILE__, __LINE__, name, major, minor, device, count, free_gb, total_gb);
}

// ---- HypreSystem.cpp:673-737: "Preconditioner setup" = SetPrecond + solver Setup (contains the whole AMG setup),
// "Solve" = solver Solve; both summed over the segregated components ------------------------------------------------------------
void HypreSystem::solve() {
  assemble_system();
  double setup = 0, write_operators = 0, solve = 0;
  for (int i = 0; i < numSolves_; ++i) {
    if (iproc_ == 0) printf("Setting up preconditioner\n");
    auto start = clk::now();
    if (usePrecond_) solverPrecondPtr_(solver_, precondSolvePtr_, precondSetupPtr_, precond_);
    checkMemory();
    if (iproc_ == 0) printf("Setting up solver\n");
    solverSetupPtr_(solver_, parMat_, parRhs_[i], parSln_[i]);
    checkMemory();
    MPI_Barrier(comm_);
    auto stop1 = clk::now();
    setup += std::chrono::duration<double>(stop1 - start).count();
    MPI_Barrier(comm_);
    fflush(stdout);

    if (writeAmgMatrices_ && precond_) {   // :700-714: one IJ file set per level, "<matrix stem>_level_<l>.IJ.<rank>"
      YAML::Node linsys = inpfile_["linear_system"];
      std::string matfile = get_optional<std::string>(linsys, "matrix_file", "amg.mm");
      std::string stem = matfile.substr(0, matfile.rfind("."));
      hypre_ParAMGData *amg_data = (hypre_ParAMGData *)precond_;
      hypre_ParCSRMatrix **A_array = hypre_ParAMGDataAArray(amg_data);
      int num_levels = hypre_ParAMGDataNumLevels(amg_data);
      for (int l = 0; l < num_levels; ++l) {
        std::string fname = stem + "_level_" + std::to_string(l) + ".IJ";
        hypre_ParCSRMatrixPrintIJ(A_array[l], 0, 0, fname.c_str());
      }
      MPI_Barrier(comm_);
    }
    auto stop2 = clk::now();
    write_operators += std::chrono::duration<double>(stop2 - stop1).count();
    MPI_Barrier(comm_);
    fflush(stdout);

    if (iproc_ == 0) printf("Solving the system\n");
    solverSolvePtr_(solver_, parMat_, parRhs_[i], parSln_[i]);
    MPI_Barrier(comm_);
    solve += std::chrono::duration<double>(clk::now() - stop2).count();
    MPI_Barrier(comm_);
    fflush(stdout);
    solverItersPtr_(solver_, &iterations_);
    solverRelresPtr_(solver_, &relres_);
    if (iproc_ == 0) printf("Solve %d: iterations = %d, final relative residual norm = %e\n", i, iterations_, relres_);
  }
  timers_.emplace_back("Preconditioner setup", setup);
  if (writeAmgMatrices_) timers_.emplace_back("Write AMG Matrices", write_operators);
  timers_.emplace_back("Solve", solve);
  solveComplete_ = true;
}

// ---- HypreSystem.cpp:739-769 ------------------------------------------------------------------------------------------------
void HypreSystem::output_linear_system() {
  if (!outputSystem_ && !outputSolution_) return;
  auto start = clk::now();
  if (outputSystem_) {
    HYPRE_IJMatrixPrint(mat_, "IJM.mat");
    for (int i = 0; i < numSolves_; ++i) {
      HYPRE_IJVectorPrint(rhs_[i], ("IJV" + std::to_string(i) + ".rhs").c_str());
      HYPRE_IJVectorPrint(sln_[i], ("IJV" + std::to_string(i) + ".sln").c_str());
    }
  }
  if (outputSolution_) {
    for (int i = 0; i < numSolves_; ++i)
      for (int j = 0; j < numVectors_; ++j) {
        HYPRE_IJVectorSetComponent(sln_[i], j);
        HYPRE_IJVectorPrint(sln_[i], ("IJV" + std::to_string(std::max(i, j)) + ".sln").c_str());
      }
  }
  MPI_Barrier(comm_);
  timers_.emplace_back("Output system", since(start));
}

// ---- HypreSystem.cpp:771-845: element-wise all-close, |x - ref| < max(rtol * max(|x|, |ref|), atol) -------------------------------
void HypreSystem::check_solution() {
  if (!checkSolution_) {
    if (iproc_ == 0) std::cout << "Reference solution not provided; skipping error check." << std::endl;
    return;
  }
  if (!solveComplete_) throw std::runtime_error("Solve was not called before check_solution");
  auto start = clk::now();
  HYPRE_Int n = (HYPRE_Int)(iUpper_ - iLower_ + 1);
  HYPRE_Complex *dsln = hypre_TAlloc(HYPRE_Complex, n, HYPRE_MEMORY_DEVICE);
  HYPRE_Complex *dslnRef = hypre_TAlloc(HYPRE_Complex, n, HYPRE_MEMORY_DEVICE);
  std::vector<HYPRE_Complex> hsln(n), hslnRef(n);
  allClose_ = true;
  for (int j = 0; j < numSolves_; ++j) {
    for (HYPRE_Int i = 0; i < numVectors_; i++) {
      HYPRE_IJVectorSetComponent(sln_[j], i);
      HYPRE_IJVectorSetComponent(slnRef_[j], i);
      HYPRE_IJVectorGetValues(sln_[j], n, NULL, dsln);
      HYPRE_IJVectorGetValues(slnRef_[j], n, NULL, dslnRef);
      hypre_TMemcpy(hsln.data(), dsln, HYPRE_Complex, n, HYPRE_MEMORY_HOST, HYPRE_MEMORY_DEVICE);
      hypre_TMemcpy(hslnRef.data(), dslnRef, HYPRE_Complex, n, HYPRE_MEMORY_HOST, HYPRE_MEMORY_DEVICE);
      HYPRE_Int printCount = 0;
      bool isClose = true;
      for (HYPRE_Int k = 0; k < n; k++) {
        double diff = std::fabs(hsln[k] - hslnRef[k]);
        double bound = std::max(rtol_ * std::max(std::fabs(hsln[k]), std::fabs(hslnRef[k])), atol_);
        if (diff >= bound) {
          isClose = false;
          if (printCount < 20 && iproc_ == (nproc_ > 1 ? 1 : 0)) {
            std::cout << k + iLower_ << ":" << hsln[k] << " " << hslnRef[k] << " " << diff << " " << bound << std::endl;
            printCount++;
          }
        }
      }
      HYPRE_Int test = isClose ? 1 : 0, testAll = 1;
      MPI_Allreduce(&test, &testAll, 1, MPI_INT, MPI_MIN, comm_);
      allClose_ = allClose_ && testAll;
      if (iproc_ == 0)
        std::cout << "Solve " << j << " comp " << i << " atol=" << atol_ << " rtol=" << rtol_ << " allClose=" << testAll << std::endl;
    }
  }
  hypre_TFree(dsln, HYPRE_MEMORY_DEVICE);
  hypre_TFree(dslnRef, HYPRE_MEMORY_DEVICE);
  MPI_Barrier(comm_);
  timers_.emplace_back("Check solution", since(start));
}

// ---- HypreSystem.cpp:847-889 ------------------------------------------------------------------------------------------------
void HypreSystem::retrieve_timers(std::vector<std::string> &names, std::vector<std::vector<double>> &data) {
  if (iproc_ != 0) return;
  if (names.empty()) {
    for (auto &t : timers_) { names.push_back(t.first); data.push_back(std::vector<double>(1, t.second)); }
    return;
  }
  for (auto &t : timers_) {
    auto it = std::find(names.begin(), names.end(), t.first);
    if (it != names.end()) data[it - names.begin()].push_back(t.second);
  }
}
void HypreSystem::summarize_timers() {
  if (iproc_ != 0) return;
  std::cout << "\nTimer summary: " << std::endl;
  for (auto &t : timers_) std::cout << "    " << std::setw(25) << std::left << t.first << t.second << " seconds" << std::endl;
}

// ---- HypreSystem.cpp:897-955: the staged COO triplets go to the device and are handed to SetValues2 with ncols = NULL
// ("one column per row entry") ------------------------------------------------------------------------------------------------
void HypreSystem::hypre_matrix_set_values() {
  if (iproc_ == 0) printf("%s : loading matrix into HYPRE_IJMatrix\n", __FUNCTION__);
  auto start = clk::now();
  size_t nnz_this_rank = rows_.size();
  if (nnz_this_rank >= (size_t)std::numeric_limits<HYPRE_Int>::max())
    throw std::runtime_error("more than 2^31 - 1 matrix entries on one rank: use more ranks");
  d_rows_ = hypre_TAlloc(HYPRE_BigInt, nnz_this_rank, HYPRE_MEMORY_DEVICE);
  d_cols_ = hypre_TAlloc(HYPRE_BigInt, nnz_this_rank, HYPRE_MEMORY_DEVICE);
  d_vals_ = hypre_TAlloc(HYPRE_Complex, nnz_this_rank, HYPRE_MEMORY_DEVICE);
  hypre_TMemcpy(d_rows_, rows_.data(), HYPRE_BigInt, nnz_this_rank, HYPRE_MEMORY_DEVICE, HYPRE_MEMORY_HOST);
  hypre_TMemcpy(d_cols_, cols_.data(), HYPRE_BigInt, nnz_this_rank, HYPRE_MEMORY_DEVICE, HYPRE_MEMORY_HOST);
  hypre_TMemcpy(d_vals_, vals_.data(), HYPRE_Complex, nnz_this_rank, HYPRE_MEMORY_DEVICE, HYPRE_MEMORY_HOST);
  printf("rank=%d : %s %s %d : nnz_this_rank=%zu\n", iproc_, __FILE__, __FUNCTION__, __LINE__, nnz_this_rank);
  HYPRE_IJMatrixSetValues2(mat_, (HYPRE_Int)nnz_this_rank, NULL, d_rows_, NULL, d_cols_, d_vals_);
  MPI_Barrier(comm_);
  timers_.emplace_back("Build HYPRE matrix", since(start));
  MPI_Barrier(comm_);
  fflush(stdout);
  std::vector<HYPRE_BigInt>().swap(rows_);
  std::vector<HYPRE_BigInt>().swap(cols_);
  std::vector<HYPRE_Complex>().swap(vals_);
}

// ---- HypreSystem.cpp:957-1015 ------------------------------------------------------------------------------------------------
void HypreSystem::hypre_vector_set_values(HYPRE_IJVector &v, int component) {
  if (iproc_ == 0) printf("%s : loading vector into HYPRE_IJVector\n", __FUNCTION__);
  auto start = clk::now();
  HYPRE_IJVectorSetComponent(v, component);
  size_t N = vector_vals_.size();
  if (d_vector_indices_) hypre_TFree(d_vector_indices_, HYPRE_MEMORY_DEVICE);
  if (d_vector_vals_) hypre_TFree(d_vector_vals_, HYPRE_MEMORY_DEVICE);
  d_vector_indices_ = hypre_TAlloc(HYPRE_BigInt, N, HYPRE_MEMORY_DEVICE);
  d_vector_vals_ = hypre_TAlloc(HYPRE_Complex, N, HYPRE_MEMORY_DEVICE);
  hypre_TMemcpy(d_vector_indices_, vector_indices_.data(), HYPRE_BigInt, N, HYPRE_MEMORY_DEVICE, HYPRE_MEMORY_HOST);
  hypre_TMemcpy(d_vector_vals_, vector_vals_.data(), HYPRE_Complex, N, HYPRE_MEMORY_DEVICE, HYPRE_MEMORY_HOST);
  HYPRE_IJVectorSetValues(v, (HYPRE_Int)N, d_vector_indices_, d_vector_vals_);
  MPI_Barrier(comm_);
  timers_.emplace_back("Build HYPRE vector", since(start));
  MPI_Barrier(comm_);
  fflush(stdout);
}

// =====================================================================================================================
// Loaders.  File grammars: SURVEY.md App. C (MatrixMarket coordinate / array, HYPRE-IJ text).  Every rank scans the whole
// file and keeps the entries of its own block rows, as the reference does (HypreSystem.cpp:1804, :1219-1221).
// =====================================================================================================================
using namespace mmtext;

void HypreSystem::read_common_linsys_keys(std::vector<std::string> &rhsfile, std::vector<std::string> &slnfile) {
  YAML::Node linsys = inpfile_["linear_system"];
  numComps_ = get_optional(linsys, "num_components", 1);
  segregatedSolve_ = get_optional(linsys, "segregated_solve", 1) != 0;
  numSolves_ = segregatedSolve_ ? numComps_ : 1;
  numVectors_ = segregatedSolve_ ? 1 : numComps_;
  if (numVectors_ > 1)
    throw std::runtime_error("segregated_solve: 0 (multi-vector solves) is outside the built path; use segregated_solve: 1");
  rtol_ = get_optional(linsys, "rtol", 1.0e-6);
  atol_ = get_optional(linsys, "atol", 1.0e-8);
  rhsfile.assign(numComps_, "");
  slnfile.assign(numComps_, "");
  if (numComps_ == 1 && linsys["rhs_file"]) {
    rhsfile[0] = linsys["rhs_file"].as<std::string>();
    if (linsys["sln_file"]) { slnfile[0] = linsys["sln_file"].as<std::string>(); checkSolution_ = true; }
  } else {
    int count = 0;
    for (int i = 0; i < numComps_; ++i) {
      YAML::Node r = linsys["rhs_file" + std::to_string(i)];
      if (!r) throw std::runtime_error("linear_system: rhs_file" + std::to_string(i) + " is missing");
      rhsfile[i] = r.as<std::string>();
      YAML::Node s = linsys["sln_file" + std::to_string(i)];
      if (s) { slnfile[i] = s.as<std::string>(); count++; }
    }
    checkSolution_ = (count == numComps_);
  }
}

// ---- HypreSystem.cpp:1613-1665 ----------------------------------------------------------------------------------------------
void HypreSystem::load_matrix_market() {
  YAML::Node linsys = inpfile_["linear_system"];
  std::vector<std::string> rhsfile, slnfile;
  read_common_linsys_keys(rhsfile, slnfile);
  if (!linsys["matrix_file"]) throw std::runtime_error("linear_system: matrix_file is missing");
  std::string matfile = linsys["matrix_file"].as<std::string>();
  complexNumbers_ = get_optional<bool>(linsys, "complex_numbers", false);
  determine_mm_system_sizes(matfile);
  init_row_decomposition();
  init_system();
  build_mm_matrix(matfile);
  build_mm_vector(rhsfile, rhs_);
  if (checkSolution_) build_mm_vector(slnfile, slnRef_);
}

// ---- HypreSystem.cpp:1670-1712 ----------------------------------------------------------------------------------------------
void HypreSystem::determine_mm_system_sizes(const std::string &matfile) {
  MPI_Barrier(comm_);
  auto start = clk::now();
  TextFile f(matfile, "matrix");
  MMBanner m = read_mm_banner(f, "matrix");
  if (!m.coordinate) throw std::runtime_error("Invalid matrix market file encountered");
  LineParser lp;
  read_size_line(f, lp);
  long long msize, nsize, nnz;
  if (!lp.integer(msize) || !lp.integer(nsize) || !lp.integer(nnz)) throw std::runtime_error("Cannot read matrix sizes in file: " + matfile);
  const int mult = complexNumbers_ ? 2 : 1;   // complex entries become real 2x2 blocks (:1698-1702)
  totalRows_ = M_ = mult * msize;
  N_ = mult * nsize;
  nnz_ = (long long)mult * mult * nnz;
  MPI_Barrier(comm_);
  timers_.emplace_back("Matrix market : determine system size", since(start));
  MPI_Barrier(comm_);
  fflush(stdout);
}

// ---- HypreSystem.cpp:1717-1850: "i j v" (1-based) or "i j re im" -> (2i,2j,re) (2i,2j+1,-im) (2i+1,2j,im) (2i+1,2j+1,re) ------
void HypreSystem::build_mm_matrix(const std::string &matfile) {
  MPI_Barrier(comm_);
  auto start = clk::now();
  if (iproc_ == 0) printf("%s : Reading from %s into HYPRE_IJMatrix\n", __FUNCTION__, matfile.c_str());
  TextFile f(matfile, "matrix");
  MMBanner m = read_mm_banner(f, "matrix");
  LineParser lp;
  read_size_line(f, lp);
  rows_.clear(); cols_.clear(); vals_.clear();
  // the body is parsed by all the host threads this rank may use (mm_text.h: read_lines_parallel); the pieces are appended in file
  // order, so rows_ / cols_ / vals_ are what a single reader would have produced (duplicates keep their order: Set semantics)
  struct Piece { std::vector<HYPRE_BigInt> rows, cols; std::vector<HYPRE_Complex> vals; long long seen = 0; };
  const int threads = mmtext::loader_threads(nproc_);
  std::vector<Piece> piece((size_t)threads);
  const bool pattern = m.pattern, cplx = complexNumbers_;
  const long long lo = iLower_, hi = iUpper_;
  mmtext::read_lines_parallel(f, threads, [&](int k, const char *b, const char *e) {
    long long i = 0, j = 0;
    double re = 1.0, im = 0.0;
    if (!mmtext::parse_entry_line(b, e, true, pattern, cplx, matfile, i, j, re, im)) return;
    --i; --j;
    Piece &P = piece[(size_t)k];
    ++P.seen;
    auto keep = [&](long long r, long long cc, double v) {
      if (r >= lo && r <= hi) { P.rows.push_back(r); P.cols.push_back(cc); P.vals.push_back(v); }
    };
    if (!cplx) keep(i, j, re);
    else {
      keep(2 * i, 2 * j, re);
      keep(2 * i, 2 * j + 1, -im);
      keep(2 * i + 1, 2 * j, im);
      keep(2 * i + 1, 2 * j + 1, re);
    }
  });
  long long seen = 0;
  size_t kept = 0;
  for (auto &P : piece) { seen += P.seen; kept += P.rows.size(); }
  rows_.reserve(kept); cols_.reserve(kept); vals_.reserve(kept);
  for (auto &P : piece) {
    rows_.insert(rows_.end(), P.rows.begin(), P.rows.end());
    cols_.insert(cols_.end(), P.cols.begin(), P.cols.end());
    vals_.insert(vals_.end(), P.vals.begin(), P.vals.end());
    Piece().rows.swap(P.rows); Piece().cols.swap(P.cols); Piece().vals.swap(P.vals);   // release the piece
  }
  if (seen * (complexNumbers_ ? 4 : 1) != nnz_ && iproc_ == 0)
    std::cout << "WARNING: " << matfile << " declares " << nnz_ << " entries, found " << seen * (complexNumbers_ ? 4 : 1) << std::endl;
  MPI_Barrier(comm_);
  timers_.emplace_back("Matrix market : read matrix", since(start));
  hypre_matrix_set_values();
  MPI_Barrier(comm_);
  fflush(stdout);
}

// ---- HypreSystem.cpp:1855-1969: array format, one value per line (complex: "re im" -> rows i, i+1) -------------------------------
void HypreSystem::build_mm_vector(std::vector<std::string> &mmfiles, std::vector<HYPRE_IJVector> &vec) {
  MPI_Barrier(comm_);
  auto start = clk::now();
  for (int j = 0; j < numComps_; ++j) {
    const std::string &mmfile = mmfiles[j];
    if (iproc_ == 0) printf("%s : Reading from %s into HYPRE_IJVector\n", __FUNCTION__, mmfile.c_str());
    TextFile f(mmfile, "vector");
    MMBanner m = read_mm_banner(f, "array");
    if (!m.array) throw std::runtime_error("Invalid matrix market file encountered");
    LineParser lp;
    read_size_line(f, lp);
    long long msize, nsize;
    if (!lp.integer(msize) || !lp.integer(nsize)) throw std::runtime_error("Cannot read array sizes in file: " + mmfile);
    if (!complexNumbers_ && msize != M_) throw std::runtime_error("Inconsistent sizes for Matrix and Vector");
    if (complexNumbers_ && 2 * msize != M_) throw std::runtime_error("Inconsistent sizes for Complex Matrix and Vector");
    vector_indices_.clear();
    vector_vals_.clear();
    const char *b, *e;
    long long i = 0;
    while (f.next(b, e)) {
      if ((b < e && *b == '%') || blank(b, e)) continue;
      bool first_in = i >= iLower_ && i <= iUpper_, second_in = complexNumbers_ && (i + 1) >= iLower_ && (i + 1) <= iUpper_;
      if (first_in || second_in) {
        lp.set(b, e);
        double re = 0, im = 0;
        if (!lp.real(re)) throw std::runtime_error("Malformed value line in " + mmfile);
        if (complexNumbers_ && !lp.real(im)) throw std::runtime_error("complex_numbers: missing imaginary part in " + mmfile);
        if (first_in) { vector_indices_.push_back(i); vector_vals_.push_back(re); }
        if (second_in) { vector_indices_.push_back(i + 1); vector_vals_.push_back(im); }
      }
      i += complexNumbers_ ? 2 : 1;
    }
    if (numSolves_ == 1) hypre_vector_set_values(vec[0], j);
    else hypre_vector_set_values(vec[j], 0);
  }
  MPI_Barrier(comm_);
  timers_.emplace_back("Matrix market : read and build vector", since(start));
  MPI_Barrier(comm_);
  fflush(stdout);
}

// ---- HypreSystem.cpp:1021-1081 ----------------------------------------------------------------------------------------------
void HypreSystem::load_hypre_format() {
  YAML::Node linsys = inpfile_["linear_system"];
  int nfiles = get_optional(linsys, "num_partitions", nproc_);
  std::vector<std::string> rhsfile, slnfile;
  read_common_linsys_keys(rhsfile, slnfile);
  if (!linsys["matrix_file"]) throw std::runtime_error("linear_system: matrix_file is missing");
  std::string matfile = linsys["matrix_file"].as<std::string>();
  determine_ij_system_sizes(matfile, nfiles);
  init_row_decomposition();
  init_system();
  build_ij_matrix(matfile, nfiles);
  build_ij_vector(rhsfile, nfiles, rhs_);
  if (checkSolution_) build_ij_vector(slnfile, nfiles, slnRef_);
}

namespace {
std::string ij_part_name(const std::string &stem, int part) {
  char suffix[16];
  snprintf(suffix, sizeof(suffix), ".%05d", part);
  return stem + suffix;
}
}  // namespace

// ---- HypreSystem.cpp:1138-1176: headers "ilower iupper jlower jupper" of every partition file ------------------------------------
void HypreSystem::determine_ij_system_sizes(const std::string &matfile, int nfiles) {
  auto start = clk::now();
  long long lo = 0, hi = -1;
  for (int p = 0; p < nfiles; ++p) {
    TextFile f(ij_part_name(matfile, p), "matrix");
    LineParser lp;
    read_size_line(f, lp);
    long long il, iu, jl, ju;
    if (!lp.integer(il) || !lp.integer(iu) || !lp.integer(jl) || !lp.integer(ju)) throw std::runtime_error("Bad IJ matrix header in " + f.path());
    lo = std::min(lo, il);
    hi = std::max(hi, iu);
  }
  totalRows_ = M_ = N_ = hi - lo + 1;
  MPI_Barrier(comm_);
  timers_.emplace_back("IJ : determine system size", since(start));
}

// ---- HypreSystem.cpp:1181-1247: body lines "row col value"; files whose row range misses this rank are skipped ----------------------
void HypreSystem::build_ij_matrix(const std::string &matfile, int nfiles) {
  auto start = clk::now();
  rows_.clear(); cols_.clear(); vals_.clear();
  for (int p = 0; p < nfiles; ++p) {
    TextFile f(ij_part_name(matfile, p), "matrix");
    if (iproc_ == 0) printf("%s : Reading from %s into HYPRE_IJMatrix\n", __FUNCTION__, f.path().c_str());
    LineParser lp;
    read_size_line(f, lp);
    long long il, iu, jl, ju;
    if (!lp.integer(il) || !lp.integer(iu) || !lp.integer(jl) || !lp.integer(ju)) throw std::runtime_error("Bad IJ matrix header in " + f.path());
    if (iu < iLower_ || il > iUpper_) continue;
    struct Piece { std::vector<HYPRE_BigInt> rows, cols; std::vector<HYPRE_Complex> vals; };
    const int threads = mmtext::loader_threads(nproc_);
    std::vector<Piece> piece((size_t)threads);
    const long long lo = iLower_, hi = iUpper_;
    const std::string path = f.path();
    mmtext::read_lines_parallel(f, threads, [&](int k, const char *b, const char *e) {
      long long r = 0, c = 0;
      double v = 0.0, unused = 0.0;
      if (!mmtext::parse_entry_line(b, e, false, false, false, path, r, c, v, unused)) return;
      if (r >= lo && r <= hi) { piece[(size_t)k].rows.push_back(r); piece[(size_t)k].cols.push_back(c); piece[(size_t)k].vals.push_back(v); }
    });
    for (auto &P : piece) {
      rows_.insert(rows_.end(), P.rows.begin(), P.rows.end());
      cols_.insert(cols_.end(), P.cols.begin(), P.cols.end());
      vals_.insert(vals_.end(), P.vals.begin(), P.vals.end());
    }
  }
  MPI_Barrier(comm_);
  timers_.emplace_back("IJ : read matrix", since(start));
  hypre_matrix_set_values();
}

// ---- HypreSystem.cpp:1252-1318: header "ilower iupper", body "row value" ---------------------------------------------------------
void HypreSystem::build_ij_vector(std::vector<std::string> &files, int nfiles, std::vector<HYPRE_IJVector> &vec) {
  auto start = clk::now();
  for (int j = 0; j < numComps_; ++j) {
    vector_indices_.clear();
    vector_vals_.clear();
    for (int p = 0; p < nfiles; ++p) {
      TextFile f(ij_part_name(files[j], p), "vector");
      if (iproc_ == 0) printf("%s : Reading from %s into HYPRE_IJVector\n", __FUNCTION__, f.path().c_str());
      LineParser lp;
      read_size_line(f, lp);
      long long il, iu;
      if (!lp.integer(il) || !lp.integer(iu)) throw std::runtime_error("Bad IJ vector header in " + f.path());
      if (iu < iLower_ || il > iUpper_) continue;
      const char *b, *e;
      while (f.next(b, e)) {
        if (blank(b, e)) continue;
        lp.set(b, e);
        long long r;
        double v;
        if (!lp.integer(r) || !lp.real(v)) throw std::runtime_error("Malformed value line in " + f.path());
        if (r >= iLower_ && r <= iUpper_) { vector_indices_.push_back(r); vector_vals_.push_back(v); }
      }
    }
    if (numSolves_ == 1) hypre_vector_set_values(vec[0], j);
    else hypre_vector_set_values(vec[j], 0);
  }
  MPI_Barrier(comm_);
  timers_.emplace_back("IJ : read and build vector", since(start));
}

// ---- HypreSystem.cpp:1476-1607: generated system, one nx x ny x nz box per rank on the generator's process grid -------------------
void HypreSystem::build_stencil(int stencil) {
  YAML::Node linsys = inpfile_["linear_system"];
  int nx = get_optional(linsys, "nx", 128), ny = get_optional(linsys, "ny", 128), nz = get_optional(linsys, "nz", 128);
  numComps_ = numSolves_ = numVectors_ = 1;
  auto start = clk::now();
  HYPRE_Int px, py, pz;
  b200hypre_process_grid(nproc_, &px, &py, &pz);
  if (iproc_ == 0) printf("%s : %d-pt stencil, %d x %d x %d rows per rank on a %d x %d x %d process grid\n", __FUNCTION__, stencil, nx, ny, nz, px, py, pz);
  HYPRE_BigInt *rows = nullptr, *cols = nullptr, nnz = 0, nloc = 0, rstart = 0, grows = 0;
  HYPRE_Complex *vals = nullptr, *rhs = nullptr;
  if (b200hypre_generate_laplace(nx, ny, nz, px, py, pz, iproc_, stencil, &rows, &cols, &vals, &rhs, &nnz, &nloc, &rstart, &grows))
    throw std::runtime_error(std::string("generator failed: ") + b200hypre_last_error());
  const std::string halo_map = get_optional(linsys, "halo_map", std::string("true"));
  if (halo_map == "reference") {   // the surrogate off-rank numbering of HypreSystem.cpp:1553-1556 instead of the true neighbours
    if (b200hypre_apply_reference_halo_map(cols, nnz, rstart, rstart + nloc - 1, iproc_, NULL))
      throw std::runtime_error(std::string("halo_map failed: ") + b200hypre_last_error());
  } else if (halo_map != "true") {
    throw std::runtime_error("linear_system.halo_map must be 'true' or 'reference'");
  }
  totalRows_ = M_ = N_ = grows;
  nnz_ = nnz;
  iLower_ = jLower_ = rstart;
  iUpper_ = jUpper_ = rstart + nloc - 1;
  numRows_ = (HYPRE_Int)nloc;
  init_system();
  // the triplets are device resident already (HypreSystem.cpp:1567,1596 pass device pointers too)
  HYPRE_IJMatrixSetValues2(mat_, (HYPRE_Int)nnz, NULL, rows, NULL, cols, vals);
  HYPRE_IJVectorSetValues(rhs_[0], (HYPRE_Int)nloc, NULL, rhs);
  d_rows_ = rows; d_cols_ = cols; d_vals_ = vals; d_vector_vals_ = rhs;   // released after Assemble
  MPI_Barrier(comm_);
  timers_.emplace_back(stencil == 27 ? "Build 27Pt Stencil HYPRE matrix" : "Build 7Pt Stencil HYPRE matrix", since(start));
}

}  // namespace nalu
