/* mpi.h -- minimal MPI shim for host code written against /root/reference/src (HypreSystem.h:24, main.cpp:33-35).
 * The GPU box has no MPI; ranks are one process per GPU wired with NCCL by libb200hypre (b200hypre.h).  Only the
 * calls the mini-app makes are provided. */
#ifndef B200_MPI_SHIM_H
#define B200_MPI_SHIM_H
#include "b200hypre.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_INT 1
#define MPI_LONG_LONG 2
#define MPI_DOUBLE 3
#define HYPRE_MPI_INT MPI_INT
#define HYPRE_MPI_BIG_INT MPI_LONG_LONG
#define MPI_MIN 1
#define MPI_MAX 2
#define MPI_SUM 3
#define MPI_SUCCESS 0
int MPI_Init(int *argc, char ***argv);          /* b200hypre_comm_init_from_env */
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Barrier(MPI_Comm comm);
int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Reduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm);
/* rank -> GPU binding of the reference, getDevice() at /root/reference/src/main.cpp:13-18.  One node: every rank is in the
 * shared-memory communicator, so the split returns a communicator with the ranks and order of `comm`. */
typedef int MPI_Info;
#define MPI_INFO_NULL 0
#define MPI_COMM_TYPE_SHARED 1
int MPI_Comm_split_type(MPI_Comm comm, int split_type, int key, MPI_Info info, MPI_Comm *newcomm);
int MPI_Comm_free(MPI_Comm *comm);
#ifdef __cplusplus
}
/* /root/reference/src/HypreSystem.cpp:1219-1220,1291-1292 call std::min / std::max with one HYPRE_BigInt and one HYPRE_Int
 * argument, which only compiles when both are the same type (a non-mixed-int hypre build).  This library mirrors the
 * mixed-int build (HypreSystem.h:174-219), so the two overloads below let those four lines compile unchanged; define
 * B200_NO_MIXED_MINMAX to leave namespace std alone and patch the four lines instead. */
#ifndef B200_NO_MIXED_MINMAX
#include <algorithm>
namespace std {
inline HYPRE_BigInt min(HYPRE_BigInt a, HYPRE_Int b) { return a < (HYPRE_BigInt)b ? a : (HYPRE_BigInt)b; }
inline HYPRE_BigInt max(HYPRE_BigInt a, HYPRE_Int b) { return a > (HYPRE_BigInt)b ? a : (HYPRE_BigInt)b; }
}
#endif
#endif
#endif
