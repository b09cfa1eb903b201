// abi_util.h -- helpers shared by the C-ABI translation units: error-flag plumbing (no exception crosses the ABI) and
// the handle layout.  HYPRE_IJMatrix = IJMatrix*, HYPRE_IJVector = IJVector*, HYPRE_ParCSRMatrix = ParCsr*,
// HYPRE_ParVector = ParVec*, HYPRE_Solver = SolverBase*.
#pragma once
#include "../../include/b200hypre.h"
#include "amg.h"

namespace b200 {

int set_error(int code, const char *what);

struct SolverBase {
  enum Kind { GMRES = 1, AMG = 2, KRYLOV = 3, ILU = 4 } kind;
  explicit SolverBase(Kind k) : kind(k) {}
  virtual ~SolverBase() {}
};
struct GmresHandle : SolverBase {
  Gmres g;
  GmresHandle() : SolverBase(GMRES) {}
};
struct KrylovHandle : SolverBase {   // PCG / BiCGSTAB
  Krylov k;
  KrylovHandle() : SolverBase(KRYLOV) {}
};
struct IluHandle : SolverBase {
  Ilu ilu;
  IluHandle() : SolverBase(ILU) {}
};
struct AmgHandle : SolverBase {
  Amg amg;
  int num_iterations = 0;
  double rel_resid = 0.0;
  // accepted-but-unused knobs are recorded so that Setup can reject the ones that would change results
  int smooth_type = 0, smooth_num_levels = 0, agg_num_levels = 0, variant = 0;
  double non_galerkin_tol = 0.0;
  AmgHandle() : SolverBase(AMG) {}
};

inline ParCsr *as_par(HYPRE_ParCSRMatrix A) { return reinterpret_cast<ParCsr *>(A); }
inline ParVec *as_vec(HYPRE_ParVector v) { return reinterpret_cast<ParVec *>(v); }
inline IJMatrix *as_ijm(HYPRE_IJMatrix m) { return reinterpret_cast<IJMatrix *>(m); }
inline IJVector *as_ijv(HYPRE_IJVector v) { return reinterpret_cast<IJVector *>(v); }
inline SolverBase *as_solver(HYPRE_Solver s) { return reinterpret_cast<SolverBase *>(s); }
GmresHandle *as_gmres(HYPRE_Solver s);
AmgHandle *as_amg(HYPRE_Solver s);
KrylovHandle *as_krylov(HYPRE_Solver s, int method);
IluHandle *as_ilu(HYPRE_Solver s);
int accumulate_error(int code);  // OR a solver return code (e.g. HYPRE_ERROR_CONV) into the global flag, return the flag

bool precond_zero_guess_hint();
void comm_get_unique_id(void *out128);
void comm_init(int rank, int nranks, const void *id128);
void comm_init_shm(int rank, int nranks, const char *name, size_t slot_bytes);
void comm_init_from_env();
void comm_finalize();
void process_grid(int nprocs, int *px, int *py, int *pz);
void block_partition(i64 n, int np, i64 *starts);
void plan_recv(const i64 *col_map, int n_halo, const i64 *col_starts, int nranks, std::vector<int> &recv_rank, std::vector<int> &recv_off);
void relax_level(Amg &amg, int level, double *u, const double *f);
void test_rand(int seed, int n, double *out_dev);

}  // namespace b200

#define API_BEGIN try {
#define API_END                                                                 \
  }                                                                             \
  catch (const std::bad_alloc &e) { return ::b200::set_error(HYPRE_ERROR_MEMORY, e.what()); } \
  catch (const std::exception &e) { return ::b200::set_error(HYPRE_ERROR_GENERIC, e.what()); } \
  catch (...) { return ::b200::set_error(HYPRE_ERROR_GENERIC, "unknown exception"); }         \
  return 0;
#define API_ARG(cond, msg) \
  if (!(cond)) return ::b200::set_error(HYPRE_ERROR_ARG, msg)
