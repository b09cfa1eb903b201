// HypreSystem.h -- host side of the mini-app over the C ABI of libb200hypre (include/HYPRE.h).
//
// Mirrors nalu::HypreSystem of the reference (/root/reference/src/HypreSystem.h:66-298): same public life cycle
// (setup_precon_and_solver -> load -> solve -> check_solution -> output_linear_system -> summarize_timers ->
// destroy_system, driven by /root/reference/src/main.cpp:172-192), same YAML keys and defaults, same
// timer names, the seven solver / preconditioner function pointers (HypreSystem.h:265-277) and the COO staging
// vectors rows_/cols_/vals_ (HypreSystem.h:174-227).  Written from scratch against the behaviour documented in
// SURVEY.md; methods gmres | cogmres | fgmres | bicg | cg | boomeramg with BoomerAMG / no preconditioner are wired, and
// the ILU preconditioner / solver of HypreSystem.cpp:61,82 in its default form (block-Jacobi ILU(0)).
#pragma once
#include <string>
#include <utility>
#include <vector>

#include "HYPRE.h"
#include "mpi.h"
#include "yaml_lite.h"

namespace nalu {

template <class T>
T get_optional(const YAML::Node &node, const std::string &key, T def) {
  YAML::Node v = node[key];
  return v ? v.as<T>() : def;
}



class HypreSystem {
 public:
  HypreSystem(MPI_Comm comm, YAML::Node &inpfile);

  void load();                      // HypreSystem.cpp:16-47
  void setup_precon_and_solver();   // :49-89
  void solve();                     // :673-737
  void check_solution();            // :771-845
  void output_linear_system();      // :739-769
  void summarize_timers();          // :879-889
  void retrieve_timers(std::vector<std::string> &names, std::vector<std::vector<double>> &data);  // :847-877
  void destroy_system();            // :499-523
  void checkMemory();               // :638-671

  // results of the last solve (harness access; the reference only prints them through hypre's print levels)
  int iterations() const { return iterations_; }
  double final_relative_residual() const { return relres_; }
  bool all_close() const { return allClose_; }

 private:
  // loaders
  void load_matrix_market();                                          // :1613-1665
  void load_hypre_format();                                           // :1021-1081
  void build_stencil(int stencil);                                    // :1476-1607 (device generator of libb200hypre)
  void determine_mm_system_sizes(const std::string &matfile);         // :1670-1712
  void build_mm_matrix(const std::string &matfile);                   // :1717-1850
  void build_mm_vector(std::vector<std::string> &files, std::vector<HYPRE_IJVector> &vec);   // :1855-1969
  void determine_ij_system_sizes(const std::string &matfile, int nfiles);                    // :1138-1176
  void build_ij_matrix(const std::string &matfile, int nfiles);                              // :1181-1247
  void build_ij_vector(std::vector<std::string> &files, int nfiles, std::vector<HYPRE_IJVector> &vec);   // :1252-1318
  void read_common_linsys_keys(std::vector<std::string> &rhsfile, std::vector<std::string> &slnfile);

  void init_row_decomposition();   // :525-544
  void init_system();              // :546-598
  void assemble_system();          // :600-636
  void hypre_matrix_set_values();  // :897-955
  void hypre_vector_set_values(HYPRE_IJVector &vec, int component);   // :957-1015

  // solver / preconditioner option surfaces
  void setup_boomeramg_precond();  // :119-326
  void setup_boomeramg_solver();   // :91-117
  void setup_gmres();              // :390-404
  void setup_cogmres();            // :372-388
  void setup_fgmres();             // :406-421
  void setup_bicg();               // :423-438
  void setup_cg();                 // :440-455
  void setup_ilu_precond();        // :328-370
  void setup_ilu();                // :457-497

  MPI_Comm comm_;
  YAML::Node &inpfile_;
  int iproc_ = 0, nproc_ = 1;

  // COO staging (HypreSystem.h:174-227): global row / column ids and values in file order; vector staging
  std::vector<HYPRE_BigInt> rows_, cols_;
  std::vector<HYPRE_Complex> vals_;
  std::vector<HYPRE_BigInt> vector_indices_;
  std::vector<HYPRE_Complex> vector_vals_;
  HYPRE_BigInt *d_rows_ = nullptr, *d_cols_ = nullptr, *d_vector_indices_ = nullptr;
  HYPRE_Complex *d_vals_ = nullptr, *d_vector_vals_ = nullptr;

  std::vector<std::pair<std::string, double>> timers_;

  HYPRE_IJMatrix mat_ = nullptr;
  HYPRE_ParCSRMatrix parMat_ = nullptr;
  std::vector<HYPRE_IJVector> rhs_, sln_, slnRef_;
  std::vector<HYPRE_ParVector> parRhs_, parSln_, parSlnRef_;
  HYPRE_Solver solver_ = nullptr, precond_ = nullptr;

  // the seven function pointers of HypreSystem.h:265-277
  HYPRE_Int (*solverDestroyPtr_)(HYPRE_Solver) = nullptr;
  HYPRE_Int (*solverSetupPtr_)(HYPRE_Solver, HYPRE_ParCSRMatrix, HYPRE_ParVector, HYPRE_ParVector) = nullptr;
  HYPRE_Int (*solverPrecondPtr_)(HYPRE_Solver, HYPRE_PtrToParSolverFcn, HYPRE_PtrToParSolverFcn, HYPRE_Solver) = nullptr;
  HYPRE_Int (*solverSolvePtr_)(HYPRE_Solver, HYPRE_ParCSRMatrix, HYPRE_ParVector, HYPRE_ParVector) = nullptr;
  HYPRE_Int (*precondSetupPtr_)(HYPRE_Solver, HYPRE_ParCSRMatrix, HYPRE_ParVector, HYPRE_ParVector) = nullptr;
  HYPRE_Int (*precondSolvePtr_)(HYPRE_Solver, HYPRE_ParCSRMatrix, HYPRE_ParVector, HYPRE_ParVector) = nullptr;
  HYPRE_Int (*precondDestroyPtr_)(HYPRE_Solver) = nullptr;

  HYPRE_BigInt iLower_ = 0, iUpper_ = -1, jLower_ = 0, jUpper_ = -1, totalRows_ = 0, M_ = 0, N_ = 0, nnz_ = 0;
  HYPRE_Int numRows_ = 0;
  int numComps_ = 1, numSolves_ = 1, numVectors_ = 1;
  bool segregatedSolve_ = true;
  double rtol_ = 1.0e-6, atol_ = 1.0e-8;
  bool solveComplete_ = false, checkSolution_ = false, outputSystem_ = false, outputSolution_ = false, usePrecond_ = true;
  bool writeAmgMatrices_ = false, complexNumbers_ = false;
  // result accessors of the selected solver (reported after every solve; not in the reference, which prints nothing)
  HYPRE_Int (*solverItersPtr_)(HYPRE_Solver, HYPRE_Int *) = nullptr;
  HYPRE_Int (*solverRelresPtr_)(HYPRE_Solver, HYPRE_Real *) = nullptr;
  int iterations_ = 0;
  double relres_ = 0.0;
  bool allClose_ = true;
};

}  // namespace nalu
