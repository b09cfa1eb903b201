// amg.h -- internal data structures of the BoomerAMG hierarchy, GMRES solver and ParCSR objects.
#pragma once
#include "common.h"
#include <map>
#include <memory>
#include <string>

namespace b200 {

void maybe_build_sell(const DCsr &A, DSell &S);

// ---- ParCSR matrix: block row of the global matrix owned by this rank --------------------------
// diag: columns inside [col_start, col_start + n_cols_local) in local numbering, diagonal entry first in each row.
// offd: remaining columns compressed through the sorted col_map_offd (global ids); halo x values are gathered into
// x_halo (length n_halo) before the offd product.  One rank: offd is empty.
struct HaloPlan {                 // receive lists grouped by owner (contiguous), send lists in requester order
  std::vector<int> recv_rank, recv_off;  // recv_off has size recv_rank.size()+1 (offsets into the halo buffer)
  std::vector<int> send_rank, send_off;
  DVec<int> send_idx;                     // local indices to pack, concatenated per destination
  DVec<double> send_buf;
  int n_send = 0;
  // peer-memory exchange (comm_peer.cu): region of this rank's arena = [flags | halo buffer 0 | halo buffer 1]
  bool p2p = false;
  i64 region = -1;                        // arena offset, -1: none
  DVec<unsigned long long> state;         // device: [0] exchanges completed on this plan (exchange s uses buffer s & 1),
                                          //         [1] address of the halo buffer of the exchange in flight (read by consumers)
  unsigned long long *flags = nullptr;    // one per sending neighbour (recv_rank order), written by the peers
  double *buf[2] = {nullptr, nullptr};
  DVec<int> send_off_dev;                 // send_off on the device
  DVec<unsigned long long> push_dst;      // per destination: address of my segment in its buffer 0, buffer 1, and of my flag
  DVec<unsigned int> ticket;
  HaloPlan() {}
  HaloPlan(const HaloPlan &) = delete;
  HaloPlan &operator=(const HaloPlan &) = delete;
  HaloPlan(HaloPlan &&o) noexcept { *this = std::move(o); }
  HaloPlan &operator=(HaloPlan &&o) noexcept;
  ~HaloPlan() { release(); }
  void release();
};

struct ParCsr {
  i64 g_rows = 0, g_cols = 0, row_start = 0, col_start = 0;
  int n = 0, n_cols_local = 0, n_halo = 0;
  DCsr diag, offd;
  DVec<int> offd_rows;            // rows that own at least one offd entry (the offd product runs over these only)
  std::vector<i64> col_map_offd;  // sorted global ids of the halo columns
  std::vector<i64> row_starts, col_starts;  // partition over ranks (nranks + 1)
  HaloPlan halo;
  DVec<double> x_halo;            // NCCL / shm transport: receive buffer
  // device location holding the address of the halo values of the exchange in flight: x_halo.p, or the peer-memory buffer
  // the last wait selected.  Kept on the device so that exchanges can be replayed inside a CUDA graph.
  const double *const *xhp() const { return reinterpret_cast<const double *const *>(halo.state.p + 1); }
  DSell sell;
  SpmvPlan plan;
  void finalize_spmv();           // pick kernels / build SELL after diag is final
};

struct ParVec {   // local part of a distributed vector; p may point into own or into foreign storage (a view)
  i64 g_n = 0, start = 0;
  int n = 0;
  double *p = nullptr;
  DVec<double> own;
  void alloc(int n_) { n = n_; own.alloc(((size_t)n_ + 1) & ~(size_t)1); own.zero(); p = own.p; }
};

// y-type operation on a ParCsr: local diag product with the fused epilogue, plus halo exchange + offd correction
// when nranks > 1.
void par_spmv(ParCsr &A, SpmvMode mode, const SpmvArgs &args);

// ---- BoomerAMG ----------------------------------------------------------------------------------------------------
struct AmgOptions {
  int print_level = 0, debug_flag = 0;
  int coarsen_type = 8;        // HYPRE default 10 (HMIS); mini-app default 8 (HypreSystem.cpp:125-126)
  int interp_type = 6, relax_type[4] = {18, 18, 18, 9}, num_sweeps[4] = {1, 1, 1, 1};
  int relax_order = 0, max_levels = 25, cycle_type = 1, max_iter = 20, p_max_elmts = 4;
  int max_coarse_size = 9, min_coarse_size = 0, cheby_order = 2, cheby_eig_est = 10, keep_transpose = 1, rap2 = 1;
  int agg_num_levels = 0, smooth_type = 0;
  double tol = 1e-7, strong_threshold = 0.25, max_row_sum = 0.9, trunc_factor = 0.0, cheby_fraction = 0.3, relax_weight = 1.0;
  double non_galerkin_tol = 0.0;
};

// wavefront schedule of one Gauss-Seidel pass (gs.cu): rows of the set sorted by dependency level
struct GsPlan {
  int n = 0, dir = 0, depth = 0, width = 0, rows = 0, cap = 1, last_blocks = 0;
  DVec<signed char> mask;         // 1 = row is updated by this pass
  DVec<int> order, lvl_off;       // rows by wavefront; wavefront l = order[lvl_off[l] .. lvl_off[l+1])
  DVec<unsigned long long> bar;   // grid-barrier counter
  bool built = false;
};

struct AmgLevel {
  ParCsr *A = nullptr;            // level 0: borrowed from the caller; others: owned (A_own)
  std::unique_ptr<ParCsr> A_own;
  ParCsr P, R;
  DVec<int> Sptr, Scol;           // strength pattern (kept for inspection)
  i64 S_nnz = 0;
  DVec<int> cf, cf_raw;           // +1 C / -1 F ; cf_raw keeps -3 for isolated points
  DVec<double> dsm;               // smoother diagonal: l1 norm, a_ii, or Chebyshev D^{-1/2}
  double cheby_max = 0, cheby_min = 0;
  std::vector<double> cheby_coef;
  DVec<double> u, f, r, t0, t1;   // work vectors of the cycle
  GsPlan gs[3][2];                // [all | C | F points][forward | backward], built for the Gauss-Seidel smoothers only
  DCsr tri[2];                    // two-stage Gauss-Seidel (relax 11 / 12): strictly lower / upper triangle of the own diag block
  SpmvPlan tri_plan[2];
};

struct Amg {
  AmgOptions opt;
  std::vector<std::unique_ptr<AmgLevel>> lev;
  // coarsest level: dense LU (unpivoted Gaussian elimination multipliers + U), SURVEY.md A.9
  DVec<double> ge_lu, ge_f;       // ge_f: gathered right-hand side of the coarsest level (multi-rank)
  int ge_n = 0, ge_row_start = 0;
  bool setup_done = false;
  DotWork dw;
  double setup_seconds = 0;
  std::map<std::string, double> setup_phase_ms;   // CUDA-event time per setup phase (strength, pmis, interp, rap ...)
  // opaque array handed to hypre_ParAMGDataAArray()
  std::vector<void *> a_array;
  // levels >= 1 of the V-cycle captured as a CUDA graph (all their buffers are internal, hence pointer-stable)
  cudaGraphExec_t graph_exec = nullptr;
  int graph_launches = 0;
  bool graph_failed = false;
  ~Amg();
};

inline bool is_two_stage_gs(int relax_type) { return relax_type == 11 || relax_type == 12; }
bool is_gs_relax(int relax_type);
bool is_l1_gs_relax(int relax_type);
void gs_smoother_diag(ParCsr &A, int relax_type, double *d);
void build_gs_plan(ParCsr &A, const DCsr &T, const int *cf, const double *dsm, int points, int dir, GsPlan &plan);
void gs_sweep(ParCsr &A, GsPlan &plan, bool l1, const double *dsm, const double *f, const double *uin, double *uout);
void gs_halo(ParCsr &A, const double *u);
void amg_setup(Amg &amg, ParCsr &A);
void amg_vcycle(Amg &amg, const double *rhs, double *z);   // z = M^{-1} rhs, zero initial guess
void amg_solve(Amg &amg, ParCsr &A, const double *b, double *x, int *iters, double *relres);  // BoomerAMG as a solver

// ---- GMRES ----------------------------------------------------------------------------------------------------------
typedef int (*PrecondFn)(void *solver, void *A, void *b, void *x);  // HYPRE_PtrToParSolverFcn shape
struct Gmres {
  double tol = 1e-6, atol = 0.0;
  int max_iter = 1000, min_iter = 0, k_dim = 5, print_level = 0, logging = 0;
  int orth = 1;                   // 0 = modified Gram-Schmidt (hypre's), 1 = fused CGS2 (default), 2 = fused one-pass CGS
  int variant = 0;                // 0 = GMRES, 1 = FlexGMRES (keeps z_j = M^{-1} v_j), 2 = COGMRES (orth from `cgs`)
  int cgs = 1, unroll = 0;        // COGMRES knobs (HYPRE_ParCSRCOGMRESSetCGS / SetUnroll)
  PrecondFn precond = nullptr, precond_setup = nullptr;
  void *precond_data = nullptr;
  // workspace
  DVec<double> V;                 // (k_dim + 1) basis vectors, leading dimension ld
  DVec<double> Z;                 // FlexGMRES only: k_dim preconditioned vectors
  i64 ld = 0;
  DVec<double> r, w, hdev;
  DotWork dw;
  int n = 0;
  // results
  int num_iterations = 0, converged = 0;
  double rel_residual_norm = 0.0;
  std::vector<double> history;
  double solve_seconds = 0;
};
void gmres_setup(Gmres &g, ParCsr &A, ParVec &b, ParVec &x);
int gmres_solve(Gmres &g, ParCsr &A, ParVec &b, ParVec &x);

// ---- PCG / BiCGSTAB (krylov.cu) ---------------------------------------------------------------------------------------
struct Krylov {
  enum Method { PCG = 0, BICGSTAB = 1 } method = PCG;
  double tol = 1e-6, atol = 0.0;
  int max_iter = 1000, min_iter = 0, print_level = 0, logging = 0, two_norm = 0;
  PrecondFn precond = nullptr, precond_setup = nullptr;
  void *precond_data = nullptr;
  DVec<double> W;                 // work vectors, leading dimension ld
  i64 ld = 0;
  DVec<double> hdev;
  DotWork dw;
  int n = 0;
  int num_iterations = 0, converged = 0;
  double rel_residual_norm = 0.0;
  std::vector<double> history;
  double solve_seconds = 0;
};
bool precond_honours_zero_hint(PrecondFn f);  // f is one of the library's own Solve entry points (abi_amg.cu)
void set_precond_zero_guess_hint(bool v);
void apply_precond(PrecondFn fn, void *data, ParCsr &A, ParVec &in, ParVec &out, const char *who);  // gmres.cu  // the next preconditioner call may treat its output vector as zero (gmres.cu)
void krylov_setup(Krylov &k, ParCsr &A, ParVec &b, ParVec &x);
int krylov_solve(Krylov &k, ParCsr &A, ParVec &b, ParVec &x);

// ---- ILU (ilu.cu): hypre ILU type 0 / fill 0 = block-Jacobi ILU(0) ------------------------------------------------------------
struct Ilu {
  int type = 0, fill = 0, reordering = 0, tri_solve = 1, iter_setup_type = 0;   // the only combination built (checked at Setup)
  int max_iter = 20, print_level = 0, logging = 0;
  double tol = 1e-7;
  int n = 0;
  ParCsr Lm, Um;                  // L with unit diagonal first / U with the pivot first, local block only
  GsPlan plan_l, plan_u;          // forward / backward wavefront schedules (gs.cu)
  DVec<double> fval, dinv;        // factor values in the pattern of A.diag (inspection) and inverted pivots
  DVec<double> r, y, z;
  DotWork dw;
  bool setup_done = false;
  int num_iterations = 0;
  double rel_resid = 0.0, setup_seconds = 0.0;
};
void ilu_setup(Ilu &ilu, ParCsr &A);
void ilu_apply(Ilu &ilu, const double *r, double *z);
void ilu_solve(Ilu &ilu, ParCsr &A, const double *b, double *x, bool zero_guess);

// ---- IJ interface (ij.cu) --------------------------------------------------------------------------------------------
struct IJMatrix {
  i64 ilower = 0, iupper = -1, jlower = 0, jupper = -1;
  bool assembled = false;
  // staged COO chunks in insertion order; op: 0 = set, 1 = add
  struct Chunk { DVec<i64> rows, cols; DVec<double> vals; int op; i64 n; };
  std::vector<Chunk> staged;
  ParCsr par;
};
struct IJVector {
  i64 jlower = 0, jupper = -1;
  int component = 0, num_components = 1;
  ParVec par;
};
void ij_matrix_set(IJMatrix &M, i64 nrows, const int *ncols, const i64 *rows, const int *row_indexes, const i64 *cols,
                   const double *vals, int op);
void ij_matrix_assemble(IJMatrix &M);
unsigned long long ij_h2d_bytes();   // bytes the IJ interface has copied from caller host buffers to the device so far
void ij_vector_init(IJVector &v);
void ij_vector_set(IJVector &v, i64 n, const i64 *idx, const double *vals, int op);
void ij_vector_get(IJVector &v, i64 n, const i64 *idx, double *out);

// ---- setup building blocks (amg_setup.cu), exposed for tests through include/b200hypre.h ----------------------------
i64 checked_total(const int *counts, i64 n, const char *what);   // 64-bit sum of per-row counts, fails beyond int32 (spgemm.cu)
void csr_transpose(const DCsr &A, DCsr &T);
// sorted = false: rows of C are written in no particular order (enough for the intermediate Q = A P of the Galerkin product)
void csr_spgemm(const DCsr &A, const DCsr &B, DCsr &C, bool diag_first, bool sorted = true);
void build_interp(const DCsr &Aext, const int *Sptr, const int *Scol, i64 S_nnz, const int *cf, const int *f2c, int n_own, int n_rows,
                  double tf, int pmax, DCsr &P);
// interp_type 0 (classical modified interpolation), same contract
void build_interp_classical(const DCsr &Aext, const int *Sptr, const int *Scol, i64 S_nnz, const int *cf, const int *f2c, int n_own, int n_rows,
                  double tf, int pmax, DCsr &P);
inline bool interp_type_built(int t) { return t == 6 || t == 0; }

// ---- generator (generator.cu): laplace_3d_weak_scaling semantics --------------------------------------------------------
// fills device COO (global ids) + rhs for this rank's box; returns nnz.  Arrays are allocated by the callee.
struct GenOut { DVec<i64> rows, cols; DVec<double> vals, rhs; i64 nnz = 0, n_local = 0, row_start = 0, g_rows = 0; };
void generate_laplace(int nx, int ny, int nz, int px, int py, int pz, int rank, int stencil, GenOut &out);
i64 apply_reference_halo_map(i64 *cols_dev, i64 nnz, i64 i_lower, i64 i_upper, int rank);   // HypreSystem.cpp:1553-1556; returns n_halo
// rows [row_lo, row_hi] of poisson27_hex(n) (kind 0) or momentum3(n) (kind 1): workload_rows.cuh
void generate_workload(int kind, int n, i64 row_lo, i64 row_hi, unsigned long long seed, GenOut &out);

}  // namespace b200
