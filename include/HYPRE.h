/* HYPRE.h -- C ABI of libb200hypre: the subset of hypre's public C API that Exawind/hypre-mini-app drives for its
 * GMRES + BoomerAMG hot path, implemented from scratch for NVIDIA B200 (sm_100a).
 *
 * Every entry point keeps hypre's name, argument order and return convention (HYPRE_Int error flag, 0 = success) so that
 * /root/reference/src/HypreSystem.{h,cpp} and /root/reference/src/main.cpp can be compiled against this header and linked
 * with -lb200hypre instead of -lHYPRE.  Each declaration cites the reference call site it serves.
 * The companion headers HYPRE_parcsr_ls.h, HYPRE_parcsr_mv.h, HYPRE_IJ_mv.h, HYPRE_krylov.h, krylov.h,
 * _hypre_parcsr_ls.h, _hypre_parcsr_mv.h, _hypre_utilities.h and mpi.h in this directory all forward to this file
 * (HypreSystem.h:18-24 includes them by those names).
 *
 * Build configuration mirrored: HYPRE_MIXEDINT (HYPRE_Int = 32-bit local indices, HYPRE_BigInt = 64-bit global ids;
 * HypreSystem.h:174-219), HYPRE_Complex = HYPRE_Real = double (HypreSystem.h:223), HYPRE_USING_CUDA / HYPRE_USING_GPU.
 * Pointers handed to SetValues-type calls may be host or device pointers (HypreSystem.cpp:907-945 passes device copies).
 * C++ exceptions never cross this boundary; a failing call prints the reason to stderr, sets the global error flag and
 * returns it (HYPRE_GetError).  There is no CPU fallback: without a CUDA device every compute call fails loudly. */
#ifndef B200_HYPRE_H
#define B200_HYPRE_H

#include <stddef.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HYPRE_MIXEDINT 1
#define HYPRE_USING_CUDA 1
#define HYPRE_USING_GPU 1
#define HYPRE_RELEASE_NAME "b200hypre"
#define HYPRE_RELEASE_VERSION "0.1.0"

typedef int HYPRE_Int;
typedef long long HYPRE_BigInt;
typedef double HYPRE_Real;
typedef double HYPRE_Complex;
typedef unsigned int hypre_uint;

/* ---- MPI shim (the box has no MPI; ranks are one process per GPU, wired with NCCL) --------------------------------- */
#ifndef B200_HAVE_REAL_MPI
typedef int MPI_Comm;
#define MPI_COMM_WORLD 0
#define MPI_COMM_NULL (-1)
#endif

/* ---- error flags (hypre's values) ------------------------------------------------------------------------------------ */
#define HYPRE_ERROR_GENERIC 1
#define HYPRE_ERROR_MEMORY 2
#define HYPRE_ERROR_ARG 4
#define HYPRE_ERROR_CONV 256
HYPRE_Int HYPRE_GetError(void);
HYPRE_Int HYPRE_ClearAllErrors(void);
const char *b200hypre_last_error(void);

/* ---- global state: /root/reference/src/main.cpp:82-156,169,218 -------------------------------------------------------- */
typedef enum { HYPRE_MEMORY_UNDEFINED = -1, HYPRE_MEMORY_HOST = 0, HYPRE_MEMORY_DEVICE = 1 } HYPRE_MemoryLocation;
typedef enum { HYPRE_EXEC_UNDEFINED = -1, HYPRE_EXEC_HOST = 0, HYPRE_EXEC_DEVICE = 1 } HYPRE_ExecutionPolicy;
HYPRE_Int HYPRE_Init(void);                                             /* main.cpp:82  */
HYPRE_Int HYPRE_Finalize(void);                                         /* main.cpp:218 */
HYPRE_Int HYPRE_SetMemoryLocation(HYPRE_MemoryLocation loc);            /* main.cpp:122 (DEVICE only) */
HYPRE_Int HYPRE_SetExecutionPolicy(HYPRE_ExecutionPolicy pol);          /* main.cpp:125 (DEVICE only) */
HYPRE_Int HYPRE_SetSpGemmUseVendor(HYPRE_Int use);                      /* main.cpp:130 (accepted; own kernels) */
HYPRE_Int HYPRE_SetSpMVUseVendor(HYPRE_Int use);                        /* main.cpp:140 */
HYPRE_Int HYPRE_SetSpTransUseVendor(HYPRE_Int use);                     /* main.cpp:150 */
HYPRE_Int HYPRE_SetGPUMemoryPoolSize(HYPRE_Int bin_growth, HYPRE_Int min_bin, HYPRE_Int max_bin, size_t max_cached_bytes); /* main.cpp:104 */
HYPRE_Int hypre_ResetDeviceRandGenerator(unsigned long long seed, unsigned long long offset);  /* main.cpp:169 */
/* hypre_TAlloc / hypre_TFree / hypre_TMemcpy (HypreSystem.cpp:793-810,907-926,516-522) */
void *b200hypre_malloc(size_t bytes, HYPRE_MemoryLocation loc);
void b200hypre_free(void *p, HYPRE_MemoryLocation loc);
void b200hypre_memcpy(void *dst, const void *src, size_t bytes, HYPRE_MemoryLocation dst_loc, HYPRE_MemoryLocation src_loc);
#define hypre_TAlloc(type, count, loc) ((type *)b200hypre_malloc(sizeof(type) * (size_t)(count), loc))
#define hypre_CTAlloc(type, count, loc) ((type *)b200hypre_calloc(sizeof(type) * (size_t)(count), loc))
#define hypre_TFree(ptr, loc) (b200hypre_free((void *)(ptr), loc), (ptr) = NULL)
#define hypre_TMemcpy(dst, src, type, count, dloc, sloc) b200hypre_memcpy((void *)(dst), (const void *)(src), sizeof(type) * (size_t)(count), dloc, sloc)
void *b200hypre_calloc(size_t bytes, HYPRE_MemoryLocation loc);

/* ---- opaque handles ------------------------------------------------------------------------------------------------------ */
struct hypre_IJMatrix_struct;
struct hypre_IJVector_struct;
struct hypre_ParCSRMatrix_struct;
struct hypre_ParVector_struct;
struct hypre_Solver_struct;
typedef struct hypre_IJMatrix_struct *HYPRE_IJMatrix;
typedef struct hypre_IJVector_struct *HYPRE_IJVector;
typedef struct hypre_ParCSRMatrix_struct *HYPRE_ParCSRMatrix;
typedef struct hypre_ParVector_struct *HYPRE_ParVector;
typedef struct hypre_Solver_struct *HYPRE_Solver;
typedef struct hypre_ParCSRMatrix_struct hypre_ParCSRMatrix;
typedef HYPRE_Int (*HYPRE_PtrToParSolverFcn)(HYPRE_Solver, HYPRE_ParCSRMatrix, HYPRE_ParVector, HYPRE_ParVector); /* HypreSystem.h:265-277 */
#define HYPRE_PARCSR 5555

/* ---- IJ matrix: HypreSystem.cpp:552-556,605-606,942-945,1567-1572,746,501 ------------------------------------------------ */
HYPRE_Int HYPRE_IJMatrixCreate(MPI_Comm comm, HYPRE_BigInt ilower, HYPRE_BigInt iupper, HYPRE_BigInt jlower, HYPRE_BigInt jupper, HYPRE_IJMatrix *matrix);
HYPRE_Int HYPRE_IJMatrixSetObjectType(HYPRE_IJMatrix matrix, HYPRE_Int type);
HYPRE_Int HYPRE_IJMatrixInitialize(HYPRE_IJMatrix matrix);
HYPRE_Int HYPRE_IJMatrixGetObject(HYPRE_IJMatrix matrix, void **object);
HYPRE_Int HYPRE_IJMatrixSetConstantValues(HYPRE_IJMatrix matrix, HYPRE_Complex value);
HYPRE_Int HYPRE_IJMatrixSetValues(HYPRE_IJMatrix matrix, HYPRE_Int nrows, HYPRE_Int *ncols, const HYPRE_BigInt *rows, const HYPRE_BigInt *cols, const HYPRE_Complex *values);
HYPRE_Int HYPRE_IJMatrixSetValues2(HYPRE_IJMatrix matrix, HYPRE_Int nrows, HYPRE_Int *ncols, const HYPRE_BigInt *rows, const HYPRE_Int *row_indexes, const HYPRE_BigInt *cols, const HYPRE_Complex *values);
HYPRE_Int HYPRE_IJMatrixAddToValues2(HYPRE_IJMatrix matrix, HYPRE_Int nrows, HYPRE_Int *ncols, const HYPRE_BigInt *rows, const HYPRE_Int *row_indexes, const HYPRE_BigInt *cols, const HYPRE_Complex *values);
HYPRE_Int HYPRE_IJMatrixAssemble(HYPRE_IJMatrix matrix);
HYPRE_Int HYPRE_IJMatrixGetLocalRange(HYPRE_IJMatrix matrix, HYPRE_BigInt *ilower, HYPRE_BigInt *iupper, HYPRE_BigInt *jlower, HYPRE_BigInt *jupper);
HYPRE_Int HYPRE_IJMatrixPrint(HYPRE_IJMatrix matrix, const char *filename);
HYPRE_Int HYPRE_IJMatrixDestroy(HYPRE_IJMatrix matrix);

/* ---- IJ vector: HypreSystem.cpp:567-580,609,749,758,804,967,1002-1005,1596,504 --------------------------------------------- */
HYPRE_Int HYPRE_IJVectorCreate(MPI_Comm comm, HYPRE_BigInt jlower, HYPRE_BigInt jupper, HYPRE_IJVector *vector);
HYPRE_Int HYPRE_IJVectorSetObjectType(HYPRE_IJVector vector, HYPRE_Int type);
HYPRE_Int HYPRE_IJVectorSetNumComponents(HYPRE_IJVector vector, HYPRE_Int num_components);
HYPRE_Int HYPRE_IJVectorSetComponent(HYPRE_IJVector vector, HYPRE_Int component);
HYPRE_Int HYPRE_IJVectorInitialize(HYPRE_IJVector vector);
HYPRE_Int HYPRE_IJVectorGetObject(HYPRE_IJVector vector, void **object);
HYPRE_Int HYPRE_IJVectorSetValues(HYPRE_IJVector vector, HYPRE_Int nvalues, const HYPRE_BigInt *indices, const HYPRE_Complex *values);
HYPRE_Int HYPRE_IJVectorAddToValues(HYPRE_IJVector vector, HYPRE_Int nvalues, const HYPRE_BigInt *indices, const HYPRE_Complex *values);
HYPRE_Int HYPRE_IJVectorGetValues(HYPRE_IJVector vector, HYPRE_Int nvalues, const HYPRE_BigInt *indices, HYPRE_Complex *values);
HYPRE_Int HYPRE_IJVectorAssemble(HYPRE_IJVector vector);
HYPRE_Int HYPRE_IJVectorPrint(HYPRE_IJVector vector, const char *filename);
HYPRE_Int HYPRE_IJVectorDestroy(HYPRE_IJVector vector);
HYPRE_Int HYPRE_ParVectorSetConstantValues(HYPRE_ParVector vector, HYPRE_Complex value);   /* HypreSystem.cpp:579 */
HYPRE_Int HYPRE_ParVectorInnerProd(HYPRE_ParVector x, HYPRE_ParVector y, HYPRE_Real *prod);
HYPRE_Int HYPRE_ParCSRMatrixMatvec(HYPRE_Complex alpha, HYPRE_ParCSRMatrix A, HYPRE_ParVector x, HYPRE_Complex beta, HYPRE_ParVector y);

/* ---- ParCSR GMRES: HypreSystem.cpp:392-403 (called at :692,:723) --------------------------------------------------------- */
HYPRE_Int HYPRE_ParCSRGMRESCreate(MPI_Comm comm, HYPRE_Solver *solver);
HYPRE_Int HYPRE_ParCSRGMRESDestroy(HYPRE_Solver solver);
HYPRE_Int HYPRE_ParCSRGMRESSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRGMRESSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRGMRESSetKDim(HYPRE_Solver solver, HYPRE_Int k_dim);
HYPRE_Int HYPRE_ParCSRGMRESSetTol(HYPRE_Solver solver, HYPRE_Real tol);
HYPRE_Int HYPRE_ParCSRGMRESSetAbsoluteTol(HYPRE_Solver solver, HYPRE_Real a_tol);
HYPRE_Int HYPRE_ParCSRGMRESSetMinIter(HYPRE_Solver solver, HYPRE_Int min_iter);
HYPRE_Int HYPRE_ParCSRGMRESSetMaxIter(HYPRE_Solver solver, HYPRE_Int max_iter);
HYPRE_Int HYPRE_ParCSRGMRESSetPrecond(HYPRE_Solver solver, HYPRE_PtrToParSolverFcn precond, HYPRE_PtrToParSolverFcn precond_setup, HYPRE_Solver precond_solver);
HYPRE_Int HYPRE_ParCSRGMRESSetLogging(HYPRE_Solver solver, HYPRE_Int logging);
HYPRE_Int HYPRE_ParCSRGMRESSetPrintLevel(HYPRE_Solver solver, HYPRE_Int print_level);
HYPRE_Int HYPRE_ParCSRGMRESGetNumIterations(HYPRE_Solver solver, HYPRE_Int *num_iterations);
HYPRE_Int HYPRE_ParCSRGMRESGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *norm);
HYPRE_Int HYPRE_ParCSRGMRESGetConverged(HYPRE_Solver solver, HYPRE_Int *converged);

/* ---- ParCSR FlexGMRES: HypreSystem.cpp:406-421 (setup_fgmres; Setup/Solve through the function pointers at :692,:723) ---- */
HYPRE_Int HYPRE_ParCSRFlexGMRESCreate(MPI_Comm comm, HYPRE_Solver *solver);
HYPRE_Int HYPRE_ParCSRFlexGMRESDestroy(HYPRE_Solver solver);
HYPRE_Int HYPRE_ParCSRFlexGMRESSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRFlexGMRESSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRFlexGMRESSetTol(HYPRE_Solver solver, HYPRE_Real tol);
HYPRE_Int HYPRE_ParCSRFlexGMRESSetAbsoluteTol(HYPRE_Solver solver, HYPRE_Real a_tol);
HYPRE_Int HYPRE_ParCSRFlexGMRESSetMaxIter(HYPRE_Solver solver, HYPRE_Int max_iter);
HYPRE_Int HYPRE_ParCSRFlexGMRESSetPrecond(HYPRE_Solver solver, HYPRE_PtrToParSolverFcn precond, HYPRE_PtrToParSolverFcn precond_setup, HYPRE_Solver precond_solver);
HYPRE_Int HYPRE_ParCSRFlexGMRESSetLogging(HYPRE_Solver solver, HYPRE_Int logging);
HYPRE_Int HYPRE_ParCSRFlexGMRESSetPrintLevel(HYPRE_Solver solver, HYPRE_Int print_level);
HYPRE_Int HYPRE_ParCSRFlexGMRESGetNumIterations(HYPRE_Solver solver, HYPRE_Int *num_iterations);
HYPRE_Int HYPRE_ParCSRFlexGMRESGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *norm);
HYPRE_Int HYPRE_ParCSRFlexGMRESSetKDim(HYPRE_Solver solver, HYPRE_Int k_dim);
HYPRE_Int HYPRE_ParCSRFlexGMRESSetMinIter(HYPRE_Solver solver, HYPRE_Int min_iter);

/* ---- ParCSR COGMRES: HypreSystem.cpp:372-388 (setup_cogmres) ----------------------------------------------------------- */
HYPRE_Int HYPRE_ParCSRCOGMRESCreate(MPI_Comm comm, HYPRE_Solver *solver);
HYPRE_Int HYPRE_ParCSRCOGMRESDestroy(HYPRE_Solver solver);
HYPRE_Int HYPRE_ParCSRCOGMRESSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRCOGMRESSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRCOGMRESSetTol(HYPRE_Solver solver, HYPRE_Real tol);
HYPRE_Int HYPRE_ParCSRCOGMRESSetAbsoluteTol(HYPRE_Solver solver, HYPRE_Real a_tol);
HYPRE_Int HYPRE_ParCSRCOGMRESSetMaxIter(HYPRE_Solver solver, HYPRE_Int max_iter);
HYPRE_Int HYPRE_ParCSRCOGMRESSetPrecond(HYPRE_Solver solver, HYPRE_PtrToParSolverFcn precond, HYPRE_PtrToParSolverFcn precond_setup, HYPRE_Solver precond_solver);
HYPRE_Int HYPRE_ParCSRCOGMRESSetLogging(HYPRE_Solver solver, HYPRE_Int logging);
HYPRE_Int HYPRE_ParCSRCOGMRESSetPrintLevel(HYPRE_Solver solver, HYPRE_Int print_level);
HYPRE_Int HYPRE_ParCSRCOGMRESGetNumIterations(HYPRE_Solver solver, HYPRE_Int *num_iterations);
HYPRE_Int HYPRE_ParCSRCOGMRESGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *norm);
HYPRE_Int HYPRE_ParCSRCOGMRESSetKDim(HYPRE_Solver solver, HYPRE_Int k_dim);
HYPRE_Int HYPRE_ParCSRCOGMRESSetMinIter(HYPRE_Solver solver, HYPRE_Int min_iter);
HYPRE_Int HYPRE_ParCSRCOGMRESSetCGS(HYPRE_Solver solver, HYPRE_Int cgs);        /* :384; 0/1 = one classical Gram-Schmidt pass, 2 = two */
HYPRE_Int HYPRE_ParCSRCOGMRESSetUnroll(HYPRE_Solver solver, HYPRE_Int unroll);

/* ---- ParCSR PCG: HypreSystem.cpp:440-455 (setup_cg) --------------------------------------------------------------------- */
HYPRE_Int HYPRE_ParCSRPCGCreate(MPI_Comm comm, HYPRE_Solver *solver);
HYPRE_Int HYPRE_ParCSRPCGDestroy(HYPRE_Solver solver);
HYPRE_Int HYPRE_ParCSRPCGSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRPCGSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRPCGSetTol(HYPRE_Solver solver, HYPRE_Real tol);
HYPRE_Int HYPRE_ParCSRPCGSetAbsoluteTol(HYPRE_Solver solver, HYPRE_Real a_tol);
HYPRE_Int HYPRE_ParCSRPCGSetMaxIter(HYPRE_Solver solver, HYPRE_Int max_iter);
HYPRE_Int HYPRE_ParCSRPCGSetPrecond(HYPRE_Solver solver, HYPRE_PtrToParSolverFcn precond, HYPRE_PtrToParSolverFcn precond_setup, HYPRE_Solver precond_solver);
HYPRE_Int HYPRE_ParCSRPCGSetLogging(HYPRE_Solver solver, HYPRE_Int logging);
HYPRE_Int HYPRE_ParCSRPCGSetPrintLevel(HYPRE_Solver solver, HYPRE_Int print_level);
HYPRE_Int HYPRE_ParCSRPCGGetNumIterations(HYPRE_Solver solver, HYPRE_Int *num_iterations);
HYPRE_Int HYPRE_ParCSRPCGGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *norm);
HYPRE_Int HYPRE_ParCSRPCGSetTwoNorm(HYPRE_Solver solver, HYPRE_Int two_norm);
HYPRE_Int HYPRE_ParCSRPCGSetRelChange(HYPRE_Solver solver, HYPRE_Int rel_change);

/* ---- ParCSR BiCGSTAB: HypreSystem.cpp:423-438 (setup_bicg) ------------------------------------------------------------- */
HYPRE_Int HYPRE_ParCSRBiCGSTABCreate(MPI_Comm comm, HYPRE_Solver *solver);
HYPRE_Int HYPRE_ParCSRBiCGSTABDestroy(HYPRE_Solver solver);
HYPRE_Int HYPRE_ParCSRBiCGSTABSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRBiCGSTABSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);
HYPRE_Int HYPRE_ParCSRBiCGSTABSetTol(HYPRE_Solver solver, HYPRE_Real tol);
HYPRE_Int HYPRE_ParCSRBiCGSTABSetAbsoluteTol(HYPRE_Solver solver, HYPRE_Real a_tol);
HYPRE_Int HYPRE_ParCSRBiCGSTABSetMaxIter(HYPRE_Solver solver, HYPRE_Int max_iter);
HYPRE_Int HYPRE_ParCSRBiCGSTABSetPrecond(HYPRE_Solver solver, HYPRE_PtrToParSolverFcn precond, HYPRE_PtrToParSolverFcn precond_setup, HYPRE_Solver precond_solver);
HYPRE_Int HYPRE_ParCSRBiCGSTABSetLogging(HYPRE_Solver solver, HYPRE_Int logging);
HYPRE_Int HYPRE_ParCSRBiCGSTABSetPrintLevel(HYPRE_Solver solver, HYPRE_Int print_level);
HYPRE_Int HYPRE_ParCSRBiCGSTABGetNumIterations(HYPRE_Solver solver, HYPRE_Int *num_iterations);
HYPRE_Int HYPRE_ParCSRBiCGSTABGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *norm);
HYPRE_Int HYPRE_ParCSRBiCGSTABSetMinIter(HYPRE_Solver solver, HYPRE_Int min_iter);
HYPRE_Int HYPRE_ParCSRBiCGSTABSetStopCrit(HYPRE_Solver solver, HYPRE_Int stop_crit);

/* ---- ILU solver / preconditioner: HypreSystem.cpp:328-370 (setup_ilu_precond), :457-497 (setup_ilu).  Built: ilu_type 0 with
 * level of fill 0 (block-Jacobi ILU(0)), local_reordering 0, direct triangular solves; other values are accepted by the
 * setters and rejected by HYPRE_ILUSetup with HYPRE_ERROR_GENERIC. ---------------------------------------------------------- */
HYPRE_Int HYPRE_ILUCreate(HYPRE_Solver *solver);                                       /* :331,:459 */
HYPRE_Int HYPRE_ILUDestroy(HYPRE_Solver solver);                                       /* :369,:493 */
HYPRE_Int HYPRE_ILUSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);   /* :367,:494 */
HYPRE_Int HYPRE_ILUSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);   /* :368,:496 */
HYPRE_Int HYPRE_ILUSetType(HYPRE_Solver solver, HYPRE_Int ilu_type);                    /* :332 */
HYPRE_Int HYPRE_ILUSetMaxIter(HYPRE_Solver solver, HYPRE_Int max_iter);                 /* :333 */
HYPRE_Int HYPRE_ILUSetTol(HYPRE_Solver solver, HYPRE_Real tol);                         /* :334 */
HYPRE_Int HYPRE_ILUSetLocalReordering(HYPRE_Solver solver, HYPRE_Int reordering_type);  /* :335 */
HYPRE_Int HYPRE_ILUSetPrintLevel(HYPRE_Solver solver, HYPRE_Int print_level);           /* :336 */
HYPRE_Int HYPRE_ILUSetLogging(HYPRE_Solver solver, HYPRE_Int logging);
HYPRE_Int HYPRE_ILUSetLevelOfFill(HYPRE_Solver solver, HYPRE_Int lfil);                 /* :339 */
HYPRE_Int HYPRE_ILUSetMaxNnzPerRow(HYPRE_Solver solver, HYPRE_Int nzmax);               /* :342 */
HYPRE_Int HYPRE_ILUSetDropThreshold(HYPRE_Solver solver, HYPRE_Real threshold);         /* :343 */
HYPRE_Int HYPRE_ILUSetIterativeSetupType(HYPRE_Solver solver, HYPRE_Int v);             /* :351 */
HYPRE_Int HYPRE_ILUSetIterativeSetupOption(HYPRE_Solver solver, HYPRE_Int v);           /* :353 */
HYPRE_Int HYPRE_ILUSetIterativeSetupMaxIter(HYPRE_Solver solver, HYPRE_Int v);          /* :354 */
HYPRE_Int HYPRE_ILUSetIterativeSetupTolerance(HYPRE_Solver solver, HYPRE_Real v);       /* :356 */
HYPRE_Int HYPRE_ILUSetLowerJacobiIters(HYPRE_Solver solver, HYPRE_Int v);               /* :364 */
HYPRE_Int HYPRE_ILUSetUpperJacobiIters(HYPRE_Solver solver, HYPRE_Int v);               /* :365 */
HYPRE_Int HYPRE_ILUGetNumIterations(HYPRE_Solver solver, HYPRE_Int *num_iterations);
HYPRE_Int HYPRE_ILUGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *res_norm);

/* ---- BoomerAMG: HypreSystem.cpp:119-326 ------------------------------------------------------------------------------------ */
HYPRE_Int HYPRE_BoomerAMGCreate(HYPRE_Solver *solver);                                                  /* :122 */
HYPRE_Int HYPRE_BoomerAMGDestroy(HYPRE_Solver solver);                                                  /* :325 */
HYPRE_Int HYPRE_BoomerAMGSetup(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);  /* :323 */
HYPRE_Int HYPRE_BoomerAMGSolve(HYPRE_Solver solver, HYPRE_ParCSRMatrix A, HYPRE_ParVector b, HYPRE_ParVector x);  /* :324 */
HYPRE_Int HYPRE_BoomerAMGSetPrintLevel(HYPRE_Solver solver, HYPRE_Int print_level);                     /* :123 */
HYPRE_Int HYPRE_BoomerAMGSetDebugFlag(HYPRE_Solver solver, HYPRE_Int debug_flag);                       /* :124 */
HYPRE_Int HYPRE_BoomerAMGSetCoarsenType(HYPRE_Solver solver, HYPRE_Int coarsen_type);                   /* :125 */
HYPRE_Int HYPRE_BoomerAMGSetCycleType(HYPRE_Solver solver, HYPRE_Int cycle_type);                       /* :127 */
HYPRE_Int HYPRE_BoomerAMGSetCycleRelaxType(HYPRE_Solver solver, HYPRE_Int relax_type, HYPRE_Int k);     /* :131-136 */
HYPRE_Int HYPRE_BoomerAMGSetRelaxType(HYPRE_Solver solver, HYPRE_Int relax_type);                       /* :138 */
HYPRE_Int HYPRE_BoomerAMGSetCycleNumSweeps(HYPRE_Solver solver, HYPRE_Int num_sweeps, HYPRE_Int k);     /* :143-148 */
HYPRE_Int HYPRE_BoomerAMGSetNumSweeps(HYPRE_Solver solver, HYPRE_Int num_sweeps);                       /* :150 */
HYPRE_Int HYPRE_BoomerAMGSetSmoothNumSweeps(HYPRE_Solver solver, HYPRE_Int smooth_num_sweeps);          /* :152 */
HYPRE_Int HYPRE_BoomerAMGSetTol(HYPRE_Solver solver, HYPRE_Real tol);                                   /* :154 */
HYPRE_Int HYPRE_BoomerAMGSetMaxIter(HYPRE_Solver solver, HYPRE_Int max_iter);                           /* :155 */
HYPRE_Int HYPRE_BoomerAMGSetRelaxOrder(HYPRE_Solver solver, HYPRE_Int relax_order);                     /* :156 */
HYPRE_Int HYPRE_BoomerAMGSetMaxLevels(HYPRE_Solver solver, HYPRE_Int max_levels);                       /* :157 */
HYPRE_Int HYPRE_BoomerAMGSetStrongThreshold(HYPRE_Solver solver, HYPRE_Real strong_threshold);          /* :158 */
HYPRE_Int HYPRE_BoomerAMGSetMaxRowSum(HYPRE_Solver solver, HYPRE_Real max_row_sum);
HYPRE_Int HYPRE_BoomerAMGSetNonGalerkinTol(HYPRE_Solver solver, HYPRE_Real tol);                        /* :163 */
HYPRE_Int HYPRE_BoomerAMGSetLevelNonGalerkinTol(HYPRE_Solver solver, HYPRE_Real tol, HYPRE_Int level);  /* :175 */
HYPRE_Int HYPRE_BoomerAMGSetVariant(HYPRE_Solver solver, HYPRE_Int variant);                            /* :182 */
HYPRE_Int HYPRE_BoomerAMGSetRAP2(HYPRE_Solver solver, HYPRE_Int rap2);                                  /* :187 */
HYPRE_Int HYPRE_BoomerAMGSetKeepTranspose(HYPRE_Solver solver, HYPRE_Int keep);                         /* :192 */
HYPRE_Int HYPRE_BoomerAMGSetInterpType(HYPRE_Solver solver, HYPRE_Int interp_type);                     /* :197 */
HYPRE_Int HYPRE_BoomerAMGSetMinCoarseSize(HYPRE_Solver solver, HYPRE_Int min_coarse_size);              /* :202 */
HYPRE_Int HYPRE_BoomerAMGSetMaxCoarseSize(HYPRE_Solver solver, HYPRE_Int max_coarse_size);              /* :207 */
HYPRE_Int HYPRE_BoomerAMGSetPMaxElmts(HYPRE_Solver solver, HYPRE_Int p_max_elmts);
HYPRE_Int HYPRE_BoomerAMGSetAggPMaxElmts(HYPRE_Solver solver, HYPRE_Int agg_p_max_elmts);               /* :212,:227 */
HYPRE_Int HYPRE_BoomerAMGSetAggNumLevels(HYPRE_Solver solver, HYPRE_Int agg_num_levels);                /* :217 */
HYPRE_Int HYPRE_BoomerAMGSetAggInterpType(HYPRE_Solver solver, HYPRE_Int agg_interp_type);              /* :222 */
HYPRE_Int HYPRE_BoomerAMGSetTruncFactor(HYPRE_Solver solver, HYPRE_Real trunc_factor);                  /* :232 */
HYPRE_Int HYPRE_BoomerAMGSetSmoothType(HYPRE_Solver solver, HYPRE_Int smooth_type);                     /* :239 */
HYPRE_Int HYPRE_BoomerAMGSetSmoothNumLevels(HYPRE_Solver solver, HYPRE_Int smooth_num_levels);          /* :247 */
HYPRE_Int HYPRE_BoomerAMGSetRelaxWt(HYPRE_Solver solver, HYPRE_Real relax_weight);
HYPRE_Int HYPRE_BoomerAMGSetChebyOrder(HYPRE_Solver solver, HYPRE_Int order);
HYPRE_Int HYPRE_BoomerAMGSetChebyFraction(HYPRE_Solver solver, HYPRE_Real ratio);
HYPRE_Int HYPRE_BoomerAMGSetChebyEigEst(HYPRE_Solver solver, HYPRE_Int eig_est);
HYPRE_Int HYPRE_BoomerAMGGetNumIterations(HYPRE_Solver solver, HYPRE_Int *num_iterations);
HYPRE_Int HYPRE_BoomerAMGGetFinalRelativeResidualNorm(HYPRE_Solver solver, HYPRE_Real *rel_resid_norm);
/* ILU-smoother knobs of BoomerAMG (HypreSystem.cpp:253-320): accepted and recorded; using an ILU smoother
 * (smooth_type 5) makes Setup fail with HYPRE_ERROR_ARG (next) */
HYPRE_Int HYPRE_BoomerAMGSetILUType(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILULevel(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILULocalReordering(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILUMaxRowNnz(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILUMaxIter(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILUDroptol(HYPRE_Solver solver, HYPRE_Real v);
HYPRE_Int HYPRE_BoomerAMGSetILUTriSolve(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILULowerJacobiIters(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILUUpperJacobiIters(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILUIterSetupType(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILUIterSetupOption(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILUIterSetupMaxIter(HYPRE_Solver solver, HYPRE_Int v);
HYPRE_Int HYPRE_BoomerAMGSetILUIterSetupTolerance(HYPRE_Solver solver, HYPRE_Real v);
HYPRE_Int HYPRE_ILUSetTriSolve(HYPRE_Solver solver, HYPRE_Int v);   /* HypreSystem.cpp:361 on an ILU handle; :306 on the AMG handle (no-op) */

/* ---- internals the mini-app touches for the AMG level dump: HypreSystem.cpp:700-714 ------------------------------------------ */
typedef struct hypre_Solver_struct hypre_ParAMGData;
hypre_ParCSRMatrix **hypre_ParAMGDataAArray(hypre_ParAMGData *amg_data);
HYPRE_Int hypre_ParAMGDataNumLevels(hypre_ParAMGData *amg_data);
HYPRE_Int hypre_ParCSRMatrixPrintIJ(const hypre_ParCSRMatrix *matrix, HYPRE_Int base_i, HYPRE_Int base_j, const char *filename);

#ifdef __cplusplus
}
#endif
#endif /* B200_HYPRE_H */
