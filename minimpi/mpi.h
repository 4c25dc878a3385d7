/* mpi.h — the subset of MPI that the reference's mpi/reduce.c (and our tree_reduce.c) use, implemented by libminimpi
 * over POSIX shared memory for ranks on ONE host. There is no MPI on the target image; this shim lets
 * the reference file compile and run UNMODIFIED (`b2r_mpirun -np 4 bin/reduce_ref`), which is BASELINE.json config 1
 * ("mpi/reduce.c int and double sum, custom tree vs MPI_Reduce, mpirun -np 2/4 on CPU — plumbing, no GPU").
 * MPI_Reduce runs the library's own host algorithms (src/coll/host_algos.h): MINIMPI_ALGO=oneshot|twoshot|tree. */
#ifndef B2R_MINIMPI_H_
#define B2R_MINIMPI_H_

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_ERR_OTHER 15
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_ANY_TAG (-1)

#define MPI_INT 1
#define MPI_FLOAT 2
#define MPI_DOUBLE 3
#define MPI_BYTE 4

#define MPI_MAX 1
#define MPI_MIN 2
#define MPI_SUM 3

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Barrier(MPI_Comm comm);
int MPI_Abort(MPI_Comm comm, int code);
double MPI_Wtime(void);
int MPI_Reduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Status* status);

#ifdef __cplusplus
}
#endif
#endif
