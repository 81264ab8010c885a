/* Minimal single-host MPI stand-in so the reference's UNMODIFIED cnnmpi.c can be compiled and timed
 * where no MPI exists.  It provides exactly the five calls the reference makes
 * (/root/reference/cnnmpi.c:419-422, :490, :558) over POSIX shared memory; ranks come from
 * RANK / WORLD_SIZE (what torchrun and our launcher export) or MPI_SHIM_RANK / MPI_SHIM_SIZE.
 * This is baseline tooling, not part of the product. */
#ifndef B200CNN_MPI_SHIM_H
#define B200CNN_MPI_SHIM_H
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_COMM_WORLD 0
#define MPI_DOUBLE 1
#define MPI_SUM 1
#define MPI_IN_PLACE ((void*)1)
#define MPI_SUCCESS 0
#ifdef __cplusplus
extern "C" {
#endif
int MPI_Init(int* argc, char*** argv);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Finalize(void);
#ifdef __cplusplus
}
#endif
#endif
