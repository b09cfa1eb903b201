// abi_ext.cu -- C ABI extensions declared in include/b200hypre.h (rank wiring, generator, hierarchy inspection,
// instrumentation, building-block test hooks) and the MPI shim of include/mpi.h.
#include "abi_util.h"
#include "../../include/mpi.h"
#include <cstring>

using namespace b200;
namespace b200 {
bool peer_active();
bool shm_active();
}

namespace {

// merged local rows of a ParCsr with global column ids: diag entries then offd entries (storage order of each block)
void parcsr_to_host(const ParCsr &A, int *ptr, i64 *col, double *val) {
  std::vector<int> dp = A.diag.ptr.to_host(), dc = A.diag.col.to_host();
  std::vector<double> dv = A.diag.val.to_host();
  std::vector<int> op, oc;
  std::vector<double> ov;
  if (A.offd.nnz > 0) { op = A.offd.ptr.to_host(); oc = A.offd.col.to_host(); ov = A.offd.val.to_host(); }
  i64 p = 0;
  for (int i = 0; i < A.n; ++i) {
    if (ptr) ptr[i] = (int)p;
    for (int k = dp[i]; k < dp[i + 1]; ++k, ++p) {
      if (col) col[p] = A.col_start + dc[k];
      if (val) val[p] = dv[k];
    }
    if (!op.empty())
      for (int k = op[i]; k < op[i + 1]; ++k, ++p) {
        if (col) col[p] = A.col_map_offd[oc[k]];
        if (val) val[p] = ov[k];
      }
  }
  if (ptr) ptr[A.n] = (int)p;
}
void parcsr_sizes(const ParCsr &A, i64 *sizes) {
  sizes[0] = A.n;
  sizes[1] = A.g_cols;
  sizes[2] = A.diag.nnz + A.offd.nnz;
  sizes[3] = A.g_rows;
}
ParCsr *pick(AmgHandle *h, int level, int which) {
  Amg &amg = h->amg;
  B200_REQUIRE(level >= 0 && level < (int)amg.lev.size(), "level out of range");
  AmgLevel &L = *amg.lev[level];
  if (which == 'A') return L.A;
  B200_REQUIRE(level + 1 < (int)amg.lev.size(), "the coarsest level has no P / R / S");
  if (which == 'P') return &L.P;
  if (which == 'R') return &L.R;
  B200_REQUIRE(false, "which must be 'A', 'P', 'R' or 'S'");
  return nullptr;
}
template <class T>
T *host_dup(const std::vector<T> &v) {
  T *p = (T *)malloc(sizeof(T) * (v.size() ? v.size() : 1));
  if (!v.empty()) memcpy(p, v.data(), sizeof(T) * v.size());
  return p;
}
void upload(DCsr &D, int nr, int nc, const int *ptr, const int *col, const double *val) {
  D.nr = nr;
  D.nc = nc;
  D.nnz = ptr[nr];
  D.ptr.from_host(ptr, (size_t)nr + 1);
  D.col.from_host(col, (size_t)D.nnz);
  D.val.from_host(val, (size_t)D.nnz);
}

}  // namespace

extern "C" {

// ---- ranks ---------------------------------------------------------------------------------------------------------------
HYPRE_Int b200hypre_get_unique_id(void *out128) {
  API_ARG(out128, "b200hypre_get_unique_id: null output");
  API_BEGIN comm_get_unique_id(out128); API_END
}
HYPRE_Int b200hypre_comm_init(HYPRE_Int rank, HYPRE_Int nranks, const void *id128) {
  API_ARG(nranks == 1 || id128, "b200hypre_comm_init: the NCCL id is required when nranks > 1");
  API_BEGIN comm_init(rank, nranks, id128); API_END
}
HYPRE_Int b200hypre_comm_init_shm(HYPRE_Int rank, HYPRE_Int nranks, const char *name, size_t slot_bytes) {
  API_ARG(nranks == 1 || name, "b200hypre_comm_init_shm: segment name required when nranks > 1");
  API_BEGIN comm_init_shm(rank, nranks, name, slot_bytes); API_END
}
HYPRE_Int b200hypre_device_info(char *name, HYPRE_Int name_cap, int *device, int *count, int *cc_major, int *cc_minor, double *free_gb,
                                double *total_gb) {
  API_BEGIN
  ensure_init();
  cudaDeviceProp prop;
  int dev = 0, cnt = 0;
  size_t fr = 0, tot = 0;
  B200_CUDA(cudaGetDevice(&dev));
  B200_CUDA(cudaGetDeviceCount(&cnt));
  B200_CUDA(cudaGetDeviceProperties(&prop, dev));
  B200_CUDA(cudaMemGetInfo(&fr, &tot));
  if (name && name_cap > 0) { strncpy(name, prop.name, name_cap - 1); name[name_cap - 1] = 0; }
  if (device) *device = dev;
  if (count) *count = cnt;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  if (free_gb) *free_gb = fr / 1.e9;
  if (total_gb) *total_gb = tot / 1.e9;
  API_END
}
HYPRE_Int b200hypre_comm_init_from_env(void) { API_BEGIN comm_init_from_env(); API_END }
HYPRE_Int b200hypre_comm_finalize(void) { API_BEGIN comm_finalize(); API_END }
HYPRE_Int b200hypre_comm_rank(void) { return ctx().rank; }
HYPRE_Int b200hypre_comm_size(void) { return ctx().nranks; }
HYPRE_Int b200hypre_barrier(void) { API_BEGIN ensure_init(); barrier(); API_END }
HYPRE_Int b200hypre_allreduce_sum(double *values, HYPRE_Int count) {
  API_BEGIN ensure_init(); B200_CUDA(cudaStreamSynchronize(stream())); allreduce_sum_host(values, count); API_END
}
HYPRE_Int b200hypre_allreduce_max(double *values, HYPRE_Int count) {
  API_BEGIN
  ensure_init();
  // max(x) over ranks through the integer max reduction of the order-preserving bit pattern
  std::vector<i64> bits(count);
  for (int k = 0; k < count; ++k) {
    i64 b;
    memcpy(&b, &values[k], 8);
    bits[k] = (b < 0) ? (b ^ 0x7fffffffffffffffLL) : b;
  }
  allreduce_max_host_i64(bits.data(), count);
  for (int k = 0; k < count; ++k) {
    i64 b = bits[k];
    b = (b < 0) ? (b ^ 0x7fffffffffffffffLL) : b;
    memcpy(&values[k], &b, 8);
  }
  API_END
}

// ---- generator / partition ----------------------------------------------------------------------------------------------
void b200hypre_process_grid(HYPRE_Int nprocs, HYPRE_Int *px, HYPRE_Int *py, HYPRE_Int *pz) {
  try { process_grid(nprocs, px, py, pz); } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); }
}
void b200hypre_block_partition(HYPRE_BigInt n, HYPRE_Int nranks, HYPRE_BigInt *starts) { block_partition(n, nranks, starts); }
HYPRE_Int b200hypre_plan_recv(const HYPRE_BigInt *col_map, HYPRE_Int n_halo, const HYPRE_BigInt *col_starts, HYPRE_Int nranks,
                              HYPRE_Int *recv_rank, HYPRE_Int *recv_off) {
  std::vector<int> rr, ro;
  plan_recv(col_map, n_halo, col_starts, nranks, rr, ro);
  for (size_t k = 0; k < rr.size(); ++k) recv_rank[k] = rr[k];
  for (size_t k = 0; k < ro.size(); ++k) recv_off[k] = ro[k];
  return (HYPRE_Int)rr.size();
}
HYPRE_Int b200hypre_generate_laplace(HYPRE_Int nx, HYPRE_Int ny, HYPRE_Int nz, HYPRE_Int px, HYPRE_Int py, HYPRE_Int pz,
                                     HYPRE_Int rank, HYPRE_Int stencil, HYPRE_BigInt **rows, HYPRE_BigInt **cols,
                                     HYPRE_Complex **vals, HYPRE_Complex **rhs, HYPRE_BigInt *nnz, HYPRE_BigInt *n_local,
                                     HYPRE_BigInt *row_start, HYPRE_BigInt *global_rows) {
  API_ARG(rows && cols && vals && rhs && nnz, "b200hypre_generate_laplace: null output");
  API_BEGIN
  GenOut g;
  generate_laplace(nx, ny, nz, px, py, pz, rank, stencil, g);
  *rows = g.rows.p; g.rows.p = nullptr;   // ownership moves to the caller (b200hypre_free(..., HYPRE_MEMORY_DEVICE))
  *cols = g.cols.p; g.cols.p = nullptr;
  *vals = g.vals.p; g.vals.p = nullptr;
  *rhs = g.rhs.p; g.rhs.p = nullptr;
  *nnz = g.nnz;
  if (n_local) *n_local = g.n_local;
  if (row_start) *row_start = g.row_start;
  if (global_rows) *global_rows = g.g_rows;
  API_END
}

HYPRE_Int b200hypre_apply_reference_halo_map(HYPRE_BigInt *cols, HYPRE_BigInt nnz, HYPRE_BigInt ilower, HYPRE_BigInt iupper, HYPRE_Int rank,
                                             HYPRE_BigInt *n_halo) {
  API_ARG(cols || nnz == 0, "b200hypre_apply_reference_halo_map: null columns");
  API_BEGIN
  const i64 nh = apply_reference_halo_map(cols, nnz, ilower, iupper, rank);
  if (n_halo) *n_halo = nh;
  API_END
}

HYPRE_Int b200hypre_generate_workload(HYPRE_Int kind, HYPRE_Int n, HYPRE_BigInt ilower, HYPRE_BigInt iupper, unsigned long long seed,
                                      HYPRE_BigInt **rows, HYPRE_BigInt **cols, HYPRE_Complex **vals, HYPRE_Complex **rhs,
                                      HYPRE_BigInt *nnz, HYPRE_BigInt *global_rows) {
  API_ARG(rows && cols && vals && rhs && nnz, "b200hypre_generate_workload: null output");
  API_BEGIN
  GenOut g;
  generate_workload(kind, n, ilower, iupper, seed, g);
  *rows = g.rows.p; g.rows.p = nullptr;   // ownership moves to the caller (b200hypre_free(..., HYPRE_MEMORY_DEVICE))
  *cols = g.cols.p; g.cols.p = nullptr;
  *vals = g.vals.p; g.vals.p = nullptr;
  *rhs = g.rhs.p; g.rhs.p = nullptr;
  *nnz = g.nnz;
  if (global_rows) *global_rows = g.g_rows;
  API_END
}

// ---- inspection ---------------------------------------------------------------------------------------------------------
HYPRE_Int b200hypre_parcsr_sizes(HYPRE_ParCSRMatrix A, HYPRE_BigInt *sizes) {
  API_ARG(A && sizes, "b200hypre_parcsr_sizes: null argument");
  API_BEGIN parcsr_sizes(*as_par(A), sizes); API_END
}
HYPRE_Int b200hypre_parcsr_get(HYPRE_ParCSRMatrix A, HYPRE_Int *ptr, HYPRE_BigInt *col, HYPRE_Complex *val) {
  API_ARG(A, "b200hypre_parcsr_get: null argument");
  API_BEGIN parcsr_to_host(*as_par(A), ptr, col, val); API_END
}
HYPRE_Int b200hypre_amg_num_levels(HYPRE_Solver amg) {
  try { return (HYPRE_Int)as_amg(amg)->amg.lev.size(); } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); return 0; }
}
HYPRE_Int b200hypre_amg_level_sizes(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Int which, HYPRE_BigInt *sizes) {
  API_ARG(amg && sizes, "b200hypre_amg_level_sizes: null argument");
  API_BEGIN
  AmgHandle *h = as_amg(amg);
  if (which == 'S') {
    B200_REQUIRE(level >= 0 && level + 1 < (int)h->amg.lev.size(), "level out of range for S");
    AmgLevel &L = *h->amg.lev[level];
    sizes[0] = L.A->n; sizes[1] = L.A->g_cols; sizes[2] = (i64)L.S_nnz; sizes[3] = L.A->g_rows;
  } else parcsr_sizes(*pick(h, level, which), sizes);
  API_END
}
HYPRE_Int b200hypre_amg_level_get(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Int which, HYPRE_Int *ptr, HYPRE_BigInt *col, HYPRE_Complex *val) {
  API_ARG(amg, "b200hypre_amg_level_get: null argument");
  API_BEGIN
  AmgHandle *h = as_amg(amg);
  if (which == 'S') {
    B200_REQUIRE(level >= 0 && level + 1 < (int)h->amg.lev.size(), "level out of range for S");
    AmgLevel &L = *h->amg.lev[level];
    std::vector<int> sp = L.Sptr.to_host();
    if (ptr) memcpy(ptr, sp.data(), sizeof(int) * sp.size());
    if (col) {
      // merged numbering: own columns first, then the halo columns of A (multi-rank)
      std::vector<int> sc = L.Scol.to_host();
      const int nown = L.A->n_cols_local;
      for (i64 k = 0; k < L.S_nnz; ++k) col[k] = sc[k] < nown ? L.A->col_start + sc[k] : L.A->col_map_offd[sc[k] - nown];
    }
  } else parcsr_to_host(*pick(h, level, which), ptr, col, val);
  API_END
}
HYPRE_Int b200hypre_amg_level_cf(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Int *cf) {
  API_ARG(amg && cf, "b200hypre_amg_level_cf: null argument");
  API_BEGIN
  AmgHandle *h = as_amg(amg);
  B200_REQUIRE(level >= 0 && level < (int)h->amg.lev.size(), "level out of range");
  AmgLevel &L = *h->amg.lev[level];
  const DVec<int> &src = L.cf_raw.n ? L.cf_raw : L.cf;
  std::vector<int> v = src.to_host();
  memcpy(cf, v.data(), sizeof(int) * (size_t)L.A->n);
  API_END
}
HYPRE_Int b200hypre_amg_level_smoother(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Complex *d) {
  API_ARG(amg && d, "b200hypre_amg_level_smoother: null argument");
  API_BEGIN
  AmgHandle *h = as_amg(amg);
  B200_REQUIRE(level >= 0 && level < (int)h->amg.lev.size(), "level out of range");
  std::vector<double> v = h->amg.lev[level]->dsm.to_host();
  memcpy(d, v.data(), sizeof(double) * std::min<size_t>(v.size(), (size_t)h->amg.lev[level]->A->n));
  API_END
}
HYPRE_Int b200hypre_amg_level_cheby(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Real *out6) {
  API_ARG(amg && out6, "b200hypre_amg_level_cheby: null argument");
  API_BEGIN
  AmgHandle *h = as_amg(amg);
  B200_REQUIRE(level >= 0 && level < (int)h->amg.lev.size(), "level out of range");
  AmgLevel &L = *h->amg.lev[level];
  out6[0] = L.cheby_max; out6[1] = L.cheby_min;
  for (size_t k = 0; k < 4; ++k) out6[2 + k] = k < L.cheby_coef.size() ? L.cheby_coef[k] : 0.0;
  API_END
}
// wavefront schedule of a Gauss-Seidel pass: which = 0 all points, 1 C points, 2 F points; dir = 0 forward, 1 backward.
// out5 = { built, depth (wavefronts), rows in the set, widest wavefront, blocks of the last sweep launch }
HYPRE_Int b200hypre_amg_gs_plan(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Int which, HYPRE_Int dir, HYPRE_Int *out5) {
  API_ARG(amg && out5 && which >= 0 && which < 3 && dir >= 0 && dir < 2, "b200hypre_amg_gs_plan: bad argument");
  API_BEGIN
  AmgHandle *h = as_amg(amg);
  B200_REQUIRE(level >= 0 && level < (int)h->amg.lev.size(), "level out of range");
  const GsPlan &p = h->amg.lev[level]->gs[which][dir];
  out5[0] = p.built ? 1 : 0; out5[1] = p.depth; out5[2] = p.rows; out5[3] = p.width; out5[4] = p.last_blocks;
  API_END
}
HYPRE_Real b200hypre_amg_setup_seconds(HYPRE_Solver amg) {
  try { return as_amg(amg)->amg.setup_seconds; } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); return -1.0; }
}
HYPRE_Int b200hypre_amg_setup_phases(HYPRE_Solver amg, char *names, HYPRE_Int names_cap, HYPRE_Real *ms, HYPRE_Int cap) {
  int n = 0;
  try {
    std::string joined;
    for (auto &kv : as_amg(amg)->amg.setup_phase_ms) {
      if (n < cap && ms) ms[n] = kv.second;
      joined += (n ? ";" : "") + kv.first;
      ++n;
    }
    if (names && names_cap > 0) { strncpy(names, joined.c_str(), names_cap - 1); names[names_cap - 1] = 0; }
  } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); return 0; }
  return n;
}
HYPRE_Int b200hypre_amg_relax(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Complex *u, const HYPRE_Complex *f) {
  API_ARG(amg && u && f, "b200hypre_amg_relax: null argument");
  API_BEGIN relax_level(as_amg(amg)->amg, level, u, f); API_END
}
HYPRE_Int b200hypre_amg_vcycle(HYPRE_Solver amg, const HYPRE_Complex *r, HYPRE_Complex *z) {
  API_ARG(amg && r && z, "b200hypre_amg_vcycle: null argument");
  API_BEGIN amg_vcycle(as_amg(amg)->amg, r, z); API_END
}

// which transport the halo exchanges of this communicator use
const char *b200hypre_halo_transport(void) {
  if (ctx().nranks <= 1) return "single-rank";
  if (shm_active()) return "host-staged shm (ranks share one GPU)";
  return peer_active() ? "NVLink peer memory (IPC arena, device flags) + NCCL all-reduce" : "NCCL send/recv + NCCL all-reduce";
}

// out = M^-1 in through a HYPRE_PtrToParSolverFcn, applied exactly as the Krylov solvers apply their preconditioner
HYPRE_Int b200hypre_precond_apply(HYPRE_PtrToParSolverFcn fn, HYPRE_Solver precond, HYPRE_ParCSRMatrix A, const HYPRE_Complex *in,
                                  HYPRE_Complex *out) {
  API_ARG(fn && A && in && out, "b200hypre_precond_apply: null argument");
  API_BEGIN
  ParCsr *M = as_par(A);
  ParVec pin, pout;
  pin.g_n = pout.g_n = M->g_rows; pin.start = pout.start = M->row_start; pin.n = pout.n = M->n;
  pin.p = const_cast<double *>(in); pout.p = out;
  apply_precond(reinterpret_cast<PrecondFn>(fn), precond, *M, pin, pout, "b200hypre_precond_apply");
  API_END
}

// ---- GMRES extras / instrumentation -------------------------------------------------------------------------------------
HYPRE_Int b200hypre_gmres_set_orthogonalization(HYPRE_Solver gmres, HYPRE_Int kind) {
  API_ARG(gmres && kind >= 0 && kind <= 2, "b200hypre_gmres_set_orthogonalization: kind must be 0 (MGS), 1 (CGS2) or 2 (CGS)");
  API_BEGIN as_gmres(gmres)->g.orth = kind; API_END
}
HYPRE_Int b200hypre_gmres_get_history(HYPRE_Solver solver, HYPRE_Real *out, HYPRE_Int cap) {  // any Krylov handle
  try {
    SolverBase *b = as_solver(solver);
    B200_REQUIRE(b && (b->kind == SolverBase::GMRES || b->kind == SolverBase::KRYLOV), "handle is not a Krylov solver");
    const std::vector<double> &h = b->kind == SolverBase::GMRES ? static_cast<GmresHandle *>(b)->g.history : static_cast<KrylovHandle *>(b)->k.history;
    int n = (int)h.size();
    for (int i = 0; i < n && i < cap; ++i) out[i] = h[i];
    return n;
  } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); return -1; }
}
HYPRE_Real b200hypre_gmres_solve_seconds(HYPRE_Solver solver) {  // any Krylov handle
  try {
    SolverBase *b = as_solver(solver);
    B200_REQUIRE(b && (b->kind == SolverBase::GMRES || b->kind == SolverBase::KRYLOV), "handle is not a Krylov solver");
    return b->kind == SolverBase::GMRES ? static_cast<GmresHandle *>(b)->g.solve_seconds : static_cast<KrylovHandle *>(b)->k.solve_seconds;
  } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); return -1.0; }
}
HYPRE_Int b200hypre_solver_converged(HYPRE_Solver solver) {  // any Krylov handle: 1 if the last Solve met its stopping test
  try {
    SolverBase *b = as_solver(solver);
    B200_REQUIRE(b && (b->kind == SolverBase::GMRES || b->kind == SolverBase::KRYLOV), "handle is not a Krylov solver");
    return b->kind == SolverBase::GMRES ? static_cast<GmresHandle *>(b)->g.converged : static_cast<KrylovHandle *>(b)->k.converged;
  } catch (const std::exception &e) { set_error(HYPRE_ERROR_ARG, e.what()); return -1; }
}
unsigned long long b200hypre_h2d_bytes(void) { return ij_h2d_bytes(); }
unsigned long long b200hypre_launch_count(void) { return ctx().launches; }
// out[0..3] = bytes the library's stream-ordered pool holds from the driver (now / high-water) and bytes in use (now / high)
HYPRE_Int b200hypre_pool_stats(unsigned long long *out) {
  API_BEGIN
  ensure_init();
  cudaMemPool_t pool;
  B200_CUDA(cudaDeviceGetDefaultMemPool(&pool, ctx().device));
  unsigned long long v[4] = {0, 0, 0, 0};
  B200_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &v[0]));
  B200_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemHigh, &v[1]));
  B200_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &v[2]));
  B200_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemHigh, &v[3]));
  for (int k = 0; k < 4; ++k) out[k] = v[k];
  API_END
}
// grows the pool to at least `bytes` in one piece (the memory stays with the library: release threshold = max)
HYPRE_Int b200hypre_pool_reserve(unsigned long long bytes) {
  API_BEGIN
  ensure_init();
  void *p = nullptr;
  B200_CUDA(cudaMallocAsync(&p, (size_t)bytes, stream()));
  B200_CUDA(cudaFreeAsync(p, stream()));
  B200_CUDA(cudaStreamSynchronize(stream()));
  API_END
}
HYPRE_Int b200hypre_profile_enable(HYPRE_Int on, double min_bytes) {
  API_BEGIN ensure_init(); prof_enable(on != 0, min_bytes); API_END
}
HYPRE_Int b200hypre_profile_summary(char *names, HYPRE_Int names_cap, double *ms, double *calls, double *bytes, HYPRE_Int cap) {
  int n = 0;
  try {
    std::vector<std::string> nm;
    std::vector<double> t, c, b;
    n = prof_summary(nm, t, c, b);
    std::string joined;
    for (int k = 0; k < n; ++k) {
      if (k < cap) { if (ms) ms[k] = t[k]; if (calls) calls[k] = c[k]; if (bytes) bytes[k] = b[k]; }
      joined += (k ? ";" : "") + nm[k];
    }
    if (names && names_cap > 0) { strncpy(names, joined.c_str(), names_cap - 1); names[names_cap - 1] = 0; }
  } catch (const std::exception &e) { set_error(HYPRE_ERROR_GENERIC, e.what()); return 0; }
  return n;
}
void *b200hypre_stream(void) {
  try { ensure_init(); } catch (const std::exception &e) { set_error(HYPRE_ERROR_GENERIC, e.what()); return nullptr; }
  return (void *)stream();
}
HYPRE_Int b200hypre_device_synchronize(void) {
  API_BEGIN ensure_init(); B200_CUDA(cudaStreamSynchronize(stream())); B200_CUDA(cudaStreamSynchronize(ctx().comm_stream)); API_END
}
void *b200hypre_vector_data(HYPRE_ParVector v) { return v ? as_vec(v)->p : nullptr; }
HYPRE_Int b200hypre_vector_local_size(HYPRE_ParVector v) { return v ? as_vec(v)->n : 0; }
const char *b200hypre_spmv_kernel_name(HYPRE_ParCSRMatrix A) {
  static char buf[32];
  if (!A) return "";
  ParCsr *M = as_par(A);
  if (M->plan.use_sell) snprintf(buf, sizeof(buf), "sell32");
  else snprintf(buf, sizeof(buf), "csr_vec<%d>", M->plan.lanes);
  return buf;
}

// ---- test hooks -----------------------------------------------------------------------------------------------------------
HYPRE_Int b200hypre_test_spgemm(HYPRE_Int a_nr, HYPRE_Int a_nc, const HYPRE_Int *a_ptr, const HYPRE_Int *a_col, const HYPRE_Complex *a_val,
                                HYPRE_Int b_nc, const HYPRE_Int *b_ptr, const HYPRE_Int *b_col, const HYPRE_Complex *b_val,
                                HYPRE_Int diag_first, HYPRE_Int **c_ptr, HYPRE_Int **c_col, HYPRE_Complex **c_val) {
  API_BEGIN
  ensure_init();
  DCsr A, B, C;
  upload(A, a_nr, a_nc, a_ptr, a_col, a_val);
  upload(B, a_nc, b_nc, b_ptr, b_col, b_val);
  csr_spgemm(A, B, C, diag_first != 0);
  *c_ptr = host_dup(C.ptr.to_host());
  *c_col = host_dup(C.col.to_host());
  *c_val = host_dup(C.val.to_host());
  API_END
}
HYPRE_Int b200hypre_test_transpose(HYPRE_Int nr, HYPRE_Int nc, const HYPRE_Int *ptr, const HYPRE_Int *col, const HYPRE_Complex *val,
                                   HYPRE_Int **t_ptr, HYPRE_Int **t_col, HYPRE_Complex **t_val) {
  API_BEGIN
  ensure_init();
  DCsr A, T;
  upload(A, nr, nc, ptr, col, val);
  csr_transpose(A, T);
  *t_ptr = host_dup(T.ptr.to_host());
  *t_col = host_dup(T.col.to_host());
  *t_val = host_dup(T.val.to_host());
  API_END
}
HYPRE_Int b200hypre_test_rand(HYPRE_Int seed, HYPRE_Int n, HYPRE_Real *out) {
  API_BEGIN
  ensure_init();
  DVec<double> d((size_t)(n > 0 ? n : 1));
  test_rand(seed, n, d.p);
  std::vector<double> h = d.to_host();
  memcpy(out, h.data(), sizeof(double) * (size_t)n);
  API_END
}
void b200hypre_host_free(void *p) { free(p); }

// ---- MPI shim (include/mpi.h) ---------------------------------------------------------------------------------------------
int MPI_Init(int *, char ***) { return b200hypre_comm_init_from_env(); }
int MPI_Finalize(void) { return b200hypre_comm_finalize(); }
int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = ctx().rank; return 0; }
int MPI_Comm_size(MPI_Comm, int *size) { *size = ctx().nranks; return 0; }
int MPI_Barrier(MPI_Comm) { return ctx().nranks > 1 ? b200hypre_barrier() : 0; }
int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm) {
  API_BEGIN
  std::vector<i64> w(count);
  for (int k = 0; k < count; ++k) {
    if (type == MPI_INT) w[k] = ((const int *)sendbuf)[k];
    else if (type == MPI_LONG_LONG) w[k] = ((const long long *)sendbuf)[k];
    else B200_REQUIRE(false, "MPI shim: integer reductions only (use b200hypre_allreduce_sum/max for doubles)");
    if (op == MPI_MIN) w[k] = -w[k];
  }
  if (ctx().nranks > 1) {
    ensure_init();
    if (op == MPI_SUM) allreduce_sum_host_i64(w.data(), count);
    else allreduce_max_host_i64(w.data(), count);
  }
  for (int k = 0; k < count; ++k) {
    i64 v = (op == MPI_MIN) ? -w[k] : w[k];
    if (type == MPI_INT) ((int *)recvbuf)[k] = (int)v;
    else ((long long *)recvbuf)[k] = v;
  }
  API_END
}
// getDevice() of the reference (/root/reference/src/main.cpp:13-18): all ranks live on one node
int MPI_Comm_split_type(MPI_Comm comm, int, int, MPI_Info, MPI_Comm *newcomm) {
  if (newcomm) *newcomm = comm;
  return 0;
}
int MPI_Comm_free(MPI_Comm *comm) {
  if (comm) *comm = MPI_COMM_NULL;
  return 0;
}
int MPI_Reduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, int, MPI_Comm comm) {
  return MPI_Allreduce(sendbuf, recvbuf, count, type, op, comm);
}

}  // extern "C"
