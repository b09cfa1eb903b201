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
#ifdef __cplusplus
}
#endif
#endif
