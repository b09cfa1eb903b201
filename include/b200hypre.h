/* b200hypre.h -- extensions of libb200hypre beyond hypre's public API: rank wiring without MPI, the device generator,
 * hierarchy inspection (the parity artefacts of SURVEY.md §8d), instrumentation used by bench.py, and a few test
 * hooks for the building blocks.  Plain C ABI: pointers and sizes only. */
#ifndef B200HYPRE_EXT_H
#define B200HYPRE_EXT_H
#include "HYPRE.h"
#ifdef __cplusplus
extern "C" {
#endif

/* ---- ranks: one process per GPU (replaces MPI_Init / MPI_Comm_rank / MPI_Comm_size, main.cpp:33-45) ------------------ */
HYPRE_Int b200hypre_get_unique_id(void *out128);                     /* rank 0: 128-byte NCCL id to ship to the others */
HYPRE_Int b200hypre_comm_init(HYPRE_Int rank, HYPRE_Int nranks, const void *id128);   /* id may be NULL when nranks == 1 */
/* N ranks sharing ONE GPU (NCCL refuses that): host-staged shared-memory transport, for parity tests of the N-rank path on a
 * single-GPU box; `name` is a POSIX shm name ("/b200hypre-<id>") common to all ranks, slot_bytes the per-rank mailbox (0 = 64 MiB) */
HYPRE_Int b200hypre_comm_init_shm(HYPRE_Int rank, HYPRE_Int nranks, const char *name, size_t slot_bytes);
HYPRE_Int b200hypre_comm_init_from_env(void);                        /* RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT */
HYPRE_Int b200hypre_comm_finalize(void);
HYPRE_Int b200hypre_comm_rank(void);
HYPRE_Int b200hypre_comm_size(void);
HYPRE_Int b200hypre_barrier(void);
HYPRE_Int b200hypre_allreduce_max(double *values, HYPRE_Int count);  /* host values, in place */
HYPRE_Int b200hypre_allreduce_sum(double *values, HYPRE_Int count);

/* ---- generator: laplace_3d_weak_scaling semantics (laplace_3d_weak_scaling.hpp:604-1090, HypreSystem.cpp:1476-1607) --- */
void b200hypre_process_grid(HYPRE_Int nprocs, HYPRE_Int *px, HYPRE_Int *py, HYPRE_Int *pz);   /* hpp:98-169; host only */
/* Fills DEVICE arrays owned by the caller afterwards (release with b200hypre_free(p, HYPRE_MEMORY_DEVICE)):
 * COO triplets with global ids for this rank's box and the rhs.  stencil: 27 (reference) or 7. */
HYPRE_Int b200hypre_generate_laplace(HYPRE_Int nx, HYPRE_Int ny, HYPRE_Int nz, HYPRE_Int px, HYPRE_Int py, HYPRE_Int pz,
                                     HYPRE_Int rank, HYPRE_Int stencil, HYPRE_BigInt **rows, HYPRE_BigInt **cols,
                                     HYPRE_Complex **vals, HYPRE_Complex **rhs, HYPRE_BigInt *nnz, HYPRE_BigInt *n_local,
                                     HYPRE_BigInt *row_start, HYPRE_BigInt *global_rows);
/* Synthetic nalu-wind-like systems of BASELINE.json configs[3] / [4] generated on the device: rows [ilower, iupper] of
 * kind 0 = perturbed-hex 27-point Poisson on n^3 nodes, kind 1 = 3-DOF upwind momentum on n^3 nodes (3 n^3 rows), as COO triplets
 * (row order, ascending columns) + rhs = A 1; same ownership rule as b200hypre_generate_laplace.  The systems then go through the
 * IJ interface like the reference's file-loaded systems (HypreSystem.cpp:897-1015). */
HYPRE_Int b200hypre_generate_workload(HYPRE_Int kind, HYPRE_Int n, HYPRE_BigInt ilower, HYPRE_BigInt iupper, unsigned long long seed,
                                      HYPRE_BigInt **rows, HYPRE_BigInt **cols, HYPRE_Complex **vals, HYPRE_Complex **rhs,
                                      HYPRE_BigInt *nnz, HYPRE_BigInt *global_rows);
/* `halo_map: reference` -- rewrites, in place, the off-rank columns (outside [ilower, iupper]) of DEVICE COO columns into the numbering
 * the reference glue hands to hypre (/root/reference/src/HypreSystem.cpp:1553-1556: generator-local halo index + (rank == 0 ? iUpper + 1
 * : 0)); the halo index is the position of the true column among the rank's sorted distinct off-rank columns
 * (src/laplace_3d_weak_scaling.hpp:495-531,966-1056).  n_halo (may be NULL) receives the number of halo points. */
HYPRE_Int b200hypre_apply_reference_halo_map(HYPRE_BigInt *cols, HYPRE_BigInt nnz, HYPRE_BigInt ilower, HYPRE_BigInt iupper, HYPRE_Int rank,
                                             HYPRE_BigInt *n_halo);
void b200hypre_block_partition(HYPRE_BigInt n, HYPRE_Int nranks, HYPRE_BigInt *starts /* nranks + 1 */);  /* HypreSystem.cpp:525-544; host only */
/* receive side of the halo plan for sorted halo column ids; returns the number of neighbours; host only */
HYPRE_Int b200hypre_plan_recv(const HYPRE_BigInt *col_map, HYPRE_Int n_halo, const HYPRE_BigInt *col_starts, HYPRE_Int nranks,
                              HYPRE_Int *recv_rank /* cap nranks */, HYPRE_Int *recv_off /* cap nranks + 1 */);

/* ---- matrix / hierarchy inspection ------------------------------------------------------------------------------------------
 * which: 'A' level operator, 'P' interpolation, 'R' restriction (stored transpose), 'S' strength pattern (no values).
 * sizes[0] = local rows, sizes[1] = global columns, sizes[2] = local nnz (diag + offd), sizes[3] = global rows.
 * get: rows in local order; every row lists its diag-block entries then its offd-block entries, columns as GLOBAL ids. */
HYPRE_Int b200hypre_parcsr_sizes(HYPRE_ParCSRMatrix A, HYPRE_BigInt *sizes);
HYPRE_Int b200hypre_parcsr_get(HYPRE_ParCSRMatrix A, HYPRE_Int *ptr, HYPRE_BigInt *col, HYPRE_Complex *val);
HYPRE_Int b200hypre_amg_num_levels(HYPRE_Solver amg);
HYPRE_Int b200hypre_amg_level_sizes(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Int which, HYPRE_BigInt *sizes);
HYPRE_Int b200hypre_amg_level_get(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Int which, HYPRE_Int *ptr, HYPRE_BigInt *col, HYPRE_Complex *val);
HYPRE_Int b200hypre_amg_level_cf(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Int *cf);              /* +1 C, -1 F, -3 isolated */
HYPRE_Int b200hypre_amg_level_smoother(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Complex *d);     /* l1 norms / a_ii / D^-1/2 */
HYPRE_Int b200hypre_amg_level_cheby(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Real *out6);        /* max, min, c0..c3 */
/* Gauss-Seidel wavefront schedule of a level: which = 0 all / 1 C / 2 F points, dir = 0 forward / 1 backward;
   out5 = { built, wavefronts, rows, widest wavefront, blocks of the last sweep launch } */
HYPRE_Int b200hypre_amg_gs_plan(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Int which, HYPRE_Int dir, HYPRE_Int *out5);
HYPRE_Real b200hypre_amg_setup_seconds(HYPRE_Solver amg);
HYPRE_Int b200hypre_amg_setup_phases(HYPRE_Solver amg, char *names, HYPRE_Int names_cap, HYPRE_Real *ms, HYPRE_Int cap); /* ';'-joined names */
/* one relaxation sweep of the level's smoother: u <- relax(u, f) (device pointers of the level's local length) */
HYPRE_Int b200hypre_amg_relax(HYPRE_Solver amg, HYPRE_Int level, HYPRE_Complex *u, const HYPRE_Complex *f);
HYPRE_Int b200hypre_amg_vcycle(HYPRE_Solver amg, const HYPRE_Complex *r, HYPRE_Complex *z);      /* z = M^-1 r, device pointers */
const char *b200hypre_halo_transport(void);   /* what the halo exchanges of the current communicator run over */
/* out = M^-1 in through a preconditioner callback, applied as HYPRE_ParCSRGMRESSolve applies it (output counts as zero) */
HYPRE_Int b200hypre_precond_apply(HYPRE_PtrToParSolverFcn fn, HYPRE_Solver precond, HYPRE_ParCSRMatrix A, const HYPRE_Complex *in,
                                  HYPRE_Complex *out);

/* ---- GMRES extras ---------------------------------------------------------------------------------------------------------- */
HYPRE_Int b200hypre_gmres_set_orthogonalization(HYPRE_Solver gmres, HYPRE_Int kind);   /* 0 = MGS (hypre), 1 = fused CGS2 (default) */
HYPRE_Int b200hypre_gmres_get_history(HYPRE_Solver gmres, HYPRE_Real *out, HYPRE_Int cap);   /* returns the history length */
HYPRE_Real b200hypre_gmres_solve_seconds(HYPRE_Solver gmres);                                 /* history / seconds accept any Krylov handle */
HYPRE_Int b200hypre_solver_converged(HYPRE_Solver krylov);
HYPRE_Int b200hypre_ilu_get_factors(HYPRE_Solver ilu, HYPRE_Real *val, HYPRE_Real *dinv);   /* host arrays: nnz(A_diag), n */
HYPRE_Real b200hypre_ilu_setup_seconds(HYPRE_Solver ilu);

/* device name / compute capability / memory of the device this rank runs on (the checkMemory() print, HypreSystem.cpp:638-671) */
HYPRE_Int b200hypre_device_info(char *name, HYPRE_Int name_cap, int *device, int *count, int *cc_major, int *cc_minor, double *free_gb,
                                double *total_gb);

/* ---- instrumentation ------------------------------------------------------------------------------------------------------- */
unsigned long long b200hypre_launch_count(void);   /* kernels launched by the library so far */
/* bytes the IJ interface (SetValues2 / AddToValues2 / IJVectorSetValues) has copied from caller HOST buffers to the device so far.
 * Row ids of a triplet list sorted by row cross the link as a row pointer (nloc + 1 ints instead of nnz 64-bit words). */
unsigned long long b200hypre_h2d_bytes(void);
/* device memory pool of the library (stream-ordered, never shrinks): out[0..3] = reserved now / high, used now / high */
HYPRE_Int b200hypre_pool_stats(unsigned long long *out);
/* pre-grow the pool to `bytes` in one piece (what HYPRE_SetGPUMemoryPoolSize is for, /root/reference/src/main.cpp:104) */
HYPRE_Int b200hypre_pool_reserve(unsigned long long bytes);
/* kernel-class timing with CUDA events on the library stream: scopes whose algorithmic byte count is >= min_bytes are
 * recorded while enabled; summary returns ';'-joined class names with total ms, call count and algorithmic bytes */
HYPRE_Int b200hypre_profile_enable(HYPRE_Int on, double min_bytes);
HYPRE_Int b200hypre_profile_summary(char *names, HYPRE_Int names_cap, double *ms, double *calls, double *bytes, HYPRE_Int cap);
void *b200hypre_stream(void);                      /* cudaStream_t every kernel of the library is launched on */
HYPRE_Int b200hypre_device_synchronize(void);
void *b200hypre_vector_data(HYPRE_ParVector v);    /* device pointer of the local part */
HYPRE_Int b200hypre_vector_local_size(HYPRE_ParVector v);
const char *b200hypre_spmv_kernel_name(HYPRE_ParCSRMatrix A);   /* "sell32" or "csr_vec<G>" chosen for the diag block */

/* ---- allocation-free local inner products ---------------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  out[j] = <V_j, x> over n local entries for j < k,
 * V_j = V + j * ld; V, x and out are DEVICE pointers of the caller (V and x 16-byte aligned, ld even and >= n when k > 1).  It is the
 * multi-dot kernel of the GMRES path (what HYPRE_ParVectorInnerProd runs for k = 1) with the same launch grid, but its per-block
 * partial sums and its arrival counter live in a file-scope device array of 128 KiB + 256 B inside the library (held by every
 * GPU that loads the library until the process exits) instead of an allocated work buffer; the call makes no cudaMalloc /
 * cudaFree / pool allocation and creates no stream or event.  Limits, rejected with an error flag and a message
 * (b200hypre_last_error): k <= 64 vectors per call, a launch grid of at most 2048 blocks (the grid is min(ceil(n / 2048), 8 x SM
 * count): 1184 on B200, any n).  HYPRE_Init must have run (it creates the library stream); the first call on a device resolves the
 * array's address with cudaGetSymbolAddress, so make that call outside a capture.  The work is ordered on b200hypre_stream() and
 * NOT synchronised; local to the rank (no all-reduce).  All calls share the one array and it keeps the last call's contents (the
 * counter is cleared at the start of every call, the partials are written before they are read): calls that overlap on the GPU
 * -- issued from several host threads or replayed from graphs on different streams -- race; issue them in stream order. */
HYPRE_Int b200hypre_multi_dot_noalloc(const HYPRE_Complex *V, HYPRE_BigInt ld, HYPRE_Int k, const HYPRE_Complex *x, HYPRE_BigInt n,
                                      HYPRE_Complex *out);

/* ---- test hooks: building blocks on host CSR arrays (int32 indices), results returned in malloc'ed host arrays that the
 * caller releases with b200hypre_host_free ---------------------------------------------------------------------------------- */
HYPRE_Int b200hypre_test_spgemm(HYPRE_Int a_nr, HYPRE_Int a_nc, const HYPRE_Int *a_ptr, const HYPRE_Int *a_col, const HYPRE_Complex *a_val,
                                HYPRE_Int b_nc, const HYPRE_Int *b_ptr, const HYPRE_Int *b_col, const HYPRE_Complex *b_val,
                                HYPRE_Int diag_first, HYPRE_Int **c_ptr, HYPRE_Int **c_col, HYPRE_Complex **c_val);
HYPRE_Int b200hypre_test_transpose(HYPRE_Int nr, HYPRE_Int nc, const HYPRE_Int *ptr, const HYPRE_Int *col, const HYPRE_Complex *val,
                                   HYPRE_Int **t_ptr, HYPRE_Int **t_col, HYPRE_Complex **t_val);
HYPRE_Int b200hypre_test_rand(HYPRE_Int seed, HYPRE_Int n, HYPRE_Real *out);   /* device evaluation of the jumpable hypre_Rand stream */
void b200hypre_host_free(void *p);

#ifdef __cplusplus
}
#endif
#endif
