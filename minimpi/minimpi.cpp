// minimpi.cpp — single-host MPI subset over POSIX shared memory (see mpi.h). Ranks are processes started by b2r_mpirun,
// which creates the segment and passes its name, the rank and the size through the environment.
//
// Segment layout: Control | P windows (collective staging, WINDOW bytes each) | P result windows | P*P point-to-point slots.
// Collectives: every rank copies a chunk of its send buffer into its window, then all ranks run the SAME algorithm code
// as the library's device collectives restated for the host (host_collective_rank in src/coll/host_algos.h) with a
// process-shared barrier; the root copies the result window out. Point-to-point: one slot per ordered pair with a
// full/empty flag (blocking, in order), enough for a hand-written tree.
#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../src/coll/host_algos.h"
#include "mpi.h"
#include "minimpi_shm.h"

namespace {

using namespace minimpi;

Control* g_ctl = nullptr;
char* g_base = nullptr;
size_t g_map_bytes = 0;
int g_rank = -1, g_size = 0;
b2r_algo_t g_algo = B2R_ALGO_TWOSHOT;

char* window(int r) { return g_base + g_ctl->off_windows + (size_t)r * g_ctl->window_bytes; }
char* result(int r) { return g_base + g_ctl->off_results + (size_t)r * g_ctl->window_bytes; }
char* work(int r) { return g_base + g_ctl->off_work + (size_t)r * g_ctl->window_bytes; }
P2PSlot* slot(int src, int dst) { return (P2PSlot*)(g_base + g_ctl->off_p2p + ((size_t)src * g_size + dst) * g_ctl->p2p_stride); }

void barrier() {
  // sense-reversing counter barrier in shared memory (pthread barriers are awkward to tear down across processes)
  const unsigned gen = g_ctl->bar_gen.load(std::memory_order_acquire);
  if (g_ctl->bar_count.fetch_add(1, std::memory_order_acq_rel) + 1 == (unsigned)g_size) {
    g_ctl->bar_count.store(0, std::memory_order_relaxed);
    g_ctl->bar_gen.fetch_add(1, std::memory_order_release);
  } else {
    unsigned spins = 0;
    while (g_ctl->bar_gen.load(std::memory_order_acquire) == gen) {
      if (g_ctl->abort_flag.load(std::memory_order_relaxed)) _exit(3);
      if (++spins > 2000) sched_yield();
    }
  }
}

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
  if (!g_ctl || count < 0 || !to_b2r(type, mop, &dt, &op) || root < -1 || root >= g_size) return MPI_ERR_OTHER;
  const size_t esz = b2r_dtype_size(dt);
  const size_t chunk_elems = g_ctl->window_bytes / esz;
  std::vector<const void*> send(g_size);
  std::vector<void*> recv(g_size), wk(g_size);
  for (int r = 0; r < g_size; ++r) {
    send[r] = window(r);
    recv[r] = result(r);
    wk[r] = work(r);
  }
  for (size_t first = 0; first < (size_t)count; first += chunk_elems) {
    const size_t n = std::min(chunk_elems, (size_t)count - first);
    memcpy(window(g_rank), (const char*)sendbuf + first * esz, n * esz);
    b2r::HostRankView v{g_rank, g_size, send.data(), recv.data(), wk.data()};
    b2r::host_collective_rank(v, n, dt, op, root, g_algo, [] { barrier(); });
    if (root < 0 || root == g_rank) memcpy((char*)recvbuf + first * esz, result(g_rank), n * esz);
    barrier();  // windows may be overwritten by the next chunk / call
  }
  return MPI_SUCCESS;
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
  g_rank = atoi(rk);
  g_size = atoi(sz);
  int fd = shm_open(name, O_RDWR, 0600);
  if (fd < 0) { perror("minimpi: shm_open"); return MPI_ERR_OTHER; }
  struct stat st;
  fstat(fd, &st);
  g_map_bytes = (size_t)st.st_size;
  g_base = (char*)mmap(nullptr, g_map_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (g_base == MAP_FAILED) { perror("minimpi: mmap"); return MPI_ERR_OTHER; }
  g_ctl = (Control*)g_base;
  const char* a = getenv("MINIMPI_ALGO");
  if (a && !strcmp(a, "oneshot")) g_algo = B2R_ALGO_ONESHOT;
  else if (a && !strcmp(a, "tree")) g_algo = B2R_ALGO_TREE;
  else g_algo = B2R_ALGO_TWOSHOT;
  barrier();
  return MPI_SUCCESS;
}

int MPI_Finalize(void) {
  if (g_ctl) barrier();
  if (g_base) munmap(g_base, g_map_bytes);
  g_base = nullptr;
  g_ctl = nullptr;
  return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm, int* rank) { *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm, int* size) { *size = g_size; return MPI_SUCCESS; }
int MPI_Barrier(MPI_Comm) { barrier(); return MPI_SUCCESS; }
int MPI_Abort(MPI_Comm, int code) {
  if (g_ctl) g_ctl->abort_flag.store(1);
  _exit(code ? code : 1);
}
double MPI_Wtime(void) {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int MPI_Reduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm) {
  return reduce_impl(sendbuf, recvbuf, count, type, op, root);
}
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm) {
  return reduce_impl(sendbuf, recvbuf, count, type, op, -1);
}

int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm) {
  if (!g_ctl || dest < 0 || dest >= g_size || dest == g_rank) return MPI_ERR_OTHER;
  P2PSlot* s = slot(g_rank, dest);
  const size_t total = (size_t)count * type_size(type), cap = g_ctl->p2p_stride - sizeof(P2PSlot);
  size_t done = 0;
  do {
    const size_t n = std::min(cap, total - done);
    unsigned spins = 0;
    while (s->full.load(std::memory_order_acquire)) {
      if (g_ctl->abort_flag.load(std::memory_order_relaxed)) _exit(3);
      if (++spins > 2000) sched_yield();
    }
    memcpy(s->data(), (const char*)buf + done, n);
    s->bytes = n;
    s->tag = tag;
    s->full.store(1, std::memory_order_release);
    done += n;
  } while (done < total);
  return MPI_SUCCESS;
}

int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm, MPI_Status* status) {
  if (!g_ctl || source < 0 || source >= g_size || source == g_rank) return MPI_ERR_OTHER;
  P2PSlot* s = slot(source, g_rank);
  const size_t total = (size_t)count * type_size(type);
  size_t done = 0;
  int seen_tag = tag;
  do {
    unsigned spins = 0;
    while (!s->full.load(std::memory_order_acquire)) {
      if (g_ctl->abort_flag.load(std::memory_order_relaxed)) _exit(3);
      if (++spins > 2000) sched_yield();
    }
    const size_t n = s->bytes;
    if (done + n > total || (tag != MPI_ANY_TAG && s->tag != tag)) {
      fprintf(stderr, "minimpi: rank %d received an unexpected message from %d (tag %d, %zu bytes)\n", g_rank, source, s->tag, n);
      MPI_Abort(0, 4);
    }
    seen_tag = s->tag;
    memcpy((char*)buf + done, s->data(), n);
    s->full.store(0, std::memory_order_release);
    done += n;
  } while (done < total);
  if (status) {
    status->MPI_SOURCE = source;
    status->MPI_TAG = seen_tag;
    status->MPI_ERROR = MPI_SUCCESS;
  }
  return MPI_SUCCESS;
}

}  // extern "C"
