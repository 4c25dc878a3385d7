// minimpi.cpp — single-host MPI subset over POSIX shared memory (see mpi.h). Ranks are processes started by b2r_mpirun,
// which creates the segment and passes its name, the rank and the size through the environment. All the work is done by
// the library's multi-process host transport (src/coll/host_shm.cpp): collectives copy a chunk of the send buffer into
// the rank's window and run the same algorithm code as the device collectives restated for the host
// (src/coll/host_algos.h) behind a process-shared barrier; point-to-point uses one full/empty slot per ordered pair
// (blocking, in order), enough for a hand-written tree.
#include <time.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../src/coll/host_shm.h"
#include "mpi.h"

namespace {

b2r::HostShmComm* g_comm = nullptr;
b2r_algo_t g_algo = B2R_ALGO_TWOSHOT;

bool to_b2r(MPI_Datatype t, MPI_Op o, b2r_dtype_t* dt, b2r_op_t* op) {
  switch (t) {
    case MPI_INT: *dt = B2R_INT32; break;
    case MPI_FLOAT: *dt = B2R_FLOAT32; break;
    case MPI_DOUBLE: *dt = B2R_FLOAT64; break;
    default: return false;
  }
  switch (o) {
    case MPI_SUM: *op = B2R_SUM; break;
    case MPI_MIN: *op = B2R_MIN; break;
    case MPI_MAX: *op = B2R_MAX; break;
    default: return false;
  }
  return true;
}

size_t type_size(MPI_Datatype t) { return t == MPI_DOUBLE ? 8 : (t == MPI_BYTE ? 1 : 4); }

int reduce_impl(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op mop, int root) {
  b2r_dtype_t dt;
  b2r_op_t op;
  if (!g_comm || count < 0 || !to_b2r(type, mop, &dt, &op)) return MPI_ERR_OTHER;
  return g_comm->reduce(sendbuf, recvbuf, (uint64_t)count, dt, op, root, g_algo) == B2R_OK ? MPI_SUCCESS : MPI_ERR_OTHER;
}

}  // namespace

extern "C" {

int MPI_Init(int*, char***) {
  const char* name = getenv("MINIMPI_SHM");
  const char* rk = getenv("MINIMPI_RANK");
  const char* sz = getenv("MINIMPI_SIZE");
  if (!name || !rk || !sz) {
    fprintf(stderr, "minimpi: not started by b2r_mpirun (MINIMPI_SHM/RANK/SIZE missing); running as a single rank is not supported\n");
    return MPI_ERR_OTHER;
  }
  g_comm = new b2r::HostShmComm();
  if (g_comm->open(name, atoi(rk), atoi(sz), 0, /*create=*/false, /*unlink_on_close=*/false) != B2R_OK) {
    fprintf(stderr, "minimpi: %s\n", b2r_last_error_string());
    return MPI_ERR_OTHER;
  }
  const char* a = getenv("MINIMPI_ALGO");
  if (a && !strcmp(a, "oneshot")) g_algo = B2R_ALGO_ONESHOT;
  else if (a && !strcmp(a, "tree")) g_algo = B2R_ALGO_TREE;
  else if (a && !strcmp(a, "ring")) g_algo = B2R_ALGO_RING;
  else g_algo = B2R_ALGO_TWOSHOT;
  g_comm->barrier();
  return MPI_SUCCESS;
}

int MPI_Finalize(void) {
  if (g_comm) {
    g_comm->barrier();
    delete g_comm;
    g_comm = nullptr;
  }
  return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm, int* rank) { *rank = g_comm ? g_comm->rank() : 0; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm, int* size) { *size = g_comm ? g_comm->size() : 1; return MPI_SUCCESS; }
int MPI_Barrier(MPI_Comm) { if (g_comm) g_comm->barrier(); return MPI_SUCCESS; }
int MPI_Abort(MPI_Comm, int code) {
  if (g_comm) g_comm->abort();
  _exit(code ? code : 1);
}
double MPI_Wtime(void) {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int MPI_Reduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm) {
  if (!g_comm || root < 0 || root >= g_comm->size()) return MPI_ERR_OTHER;
  return reduce_impl(sendbuf, recvbuf, count, type, op, root);
}
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm) {
  return reduce_impl(sendbuf, recvbuf, count, type, op, -1);
}

int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm) {
  if (!g_comm) return MPI_ERR_OTHER;
  return g_comm->send(buf, (size_t)count * type_size(type), dest, tag) == B2R_OK ? MPI_SUCCESS : MPI_ERR_OTHER;
}

int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm, MPI_Status* status) {
  if (!g_comm) return MPI_ERR_OTHER;
  int seen = tag;
  if (g_comm->recv(buf, (size_t)count * type_size(type), source, tag, &seen) != B2R_OK) {
    fprintf(stderr, "minimpi: %s\n", b2r_last_error_string());
    MPI_Abort(0, 4);
  }
  if (status) {
    status->MPI_SOURCE = source;
    status->MPI_TAG = seen;
    status->MPI_ERROR = MPI_SUCCESS;
  }
  return MPI_SUCCESS;
}

}  // extern "C"
