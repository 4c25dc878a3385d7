// host_shm.cpp — see host_shm.h.
#include "host_shm.h"

#include <chrono>
#include <errno.h>
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "../util/check.h"
#include "host_algos.h"

namespace b2r {

namespace {
size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
}  // namespace

void shm_layout(ShmControl* c, int nranks, size_t window_bytes, size_t p2p_payload) {
  c->nranks = nranks;
  c->window_bytes = window_bytes;
  c->off_windows = round_up(sizeof(ShmControl), 4096);
  c->off_results = c->off_windows + (size_t)nranks * window_bytes;
  c->off_work = c->off_results + (size_t)nranks * window_bytes;
  c->off_p2p = c->off_work + (size_t)nranks * window_bytes;
  c->p2p_stride = round_up(sizeof(ShmP2PSlot) + p2p_payload, 4096);
  c->total_bytes = c->off_p2p + (size_t)nranks * nranks * c->p2p_stride;
}

HostShmComm::~HostShmComm() { close(); }

b2r_status_t HostShmComm::open(const std::string& name, int rank, int nranks, size_t window_bytes, bool create, bool unlink_on_close) {
  B2R_REQUIRE(nranks >= 1 && nranks <= 256 && rank >= 0 && rank < nranks, "host comm: bad rank %d / %d", rank, nranks);
  name_ = name;
  rank_ = rank;
  size_ = nranks;
  unlink_ = unlink_on_close;
  int fd = -1;
  if (create) {
    ShmControl tmp;
    shm_layout(&tmp, nranks, window_bytes, 4 << 20);
    fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) return fail(B2R_ERR_SYSTEM, "shm_open(%s, create): %s", name.c_str(), strerror(errno));
    if (ftruncate(fd, (off_t)tmp.total_bytes) != 0) {
      ::close(fd);
      shm_unlink(name.c_str());
      return fail(B2R_ERR_SYSTEM, "ftruncate(%zu): %s", tmp.total_bytes, strerror(errno));
    }
    map_bytes_ = tmp.total_bytes;
  } else {
    // the creator may not have sized the segment yet: wait until it has a Control block in it
    for (int tries = 0;; ++tries) {
      fd = shm_open(name.c_str(), O_RDWR, 0600);
      struct stat st;
      if (fd >= 0 && fstat(fd, &st) == 0 && (size_t)st.st_size >= sizeof(ShmControl)) {
        map_bytes_ = (size_t)st.st_size;
        break;
      }
      if (fd >= 0) ::close(fd);
      if (tries > 4000) return fail(B2R_ERR_TIMEOUT, "host comm: segment %s never appeared", name.c_str());
      usleep(5000);
    }
  }
  void* p = mmap(nullptr, map_bytes_, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  ::close(fd);
  if (p == MAP_FAILED) return fail(B2R_ERR_SYSTEM, "mmap(%zu): %s", map_bytes_, strerror(errno));
  base_ = (char*)p;
  ctl_ = (ShmControl*)base_;
  if (create) {
    // A fresh POSIX shm segment is zero-filled, so the barrier words and the abort flag already hold their initial value:
    // they are NOT stored to here (an opener that raced ahead may already have arrived at the first barrier, and a late
    // store of 0 would wipe its arrival). Only the layout is written, then `ready` is published with release semantics.
    shm_layout(ctl_, nranks, window_bytes, 4 << 20);
    ctl_->ready.store(1, std::memory_order_release);
  } else {
    for (int tries = 0; ctl_->ready.load(std::memory_order_acquire) == 0; ++tries) {  // creator still initialising
      if (tries > 4000) return fail(B2R_ERR_TIMEOUT, "host comm: segment %s was never initialised", name.c_str());
      usleep(5000);
    }
    if (ctl_->nranks != nranks) return fail(B2R_ERR_INVALID, "host comm: segment %s was created for %d ranks, not %d", name.c_str(), ctl_->nranks, nranks);
  }
  return B2R_OK;
}

void HostShmComm::close() {
  if (base_) munmap(base_, map_bytes_);
  if (unlink_ && !name_.empty()) shm_unlink(name_.c_str());
  base_ = nullptr;
  ctl_ = nullptr;
  unlink_ = false;
}

bool HostShmComm::aborted() const { return ctl_ && ctl_->abort_flag.load(std::memory_order_relaxed) != 0; }
void HostShmComm::abort() {
  if (ctl_) ctl_->abort_flag.store(1);
}

void HostShmComm::barrier() {
  // generation-counting barrier in shared memory (process-shared pthread barriers are awkward to tear down)
  const unsigned gen = ctl_->bar_gen.load(std::memory_order_acquire);
  if (ctl_->bar_count.fetch_add(1, std::memory_order_acq_rel) + 1 == (unsigned)size_) {
    ctl_->bar_count.store(0, std::memory_order_relaxed);
    ctl_->bar_gen.fetch_add(1, std::memory_order_release);
  } else {
    // deadline: a peer that died (or never came) must not hang the survivors forever (B2R_HOST_BARRIER_TIMEOUT_S, default 600)
    static const double limit_s = [] { const char* v = getenv("B2R_HOST_BARRIER_TIMEOUT_S"); return v && *v ? atof(v) : 600.0; }();
    const auto t0 = std::chrono::steady_clock::now();
    unsigned spins = 0;
    while (ctl_->bar_gen.load(std::memory_order_acquire) == gen) {
      if (aborted()) _exit(3);
      if (++spins > 2000) {
        sched_yield();
        if ((spins & 0xfff) == 0 && limit_s > 0 && std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > limit_s) {
          fprintf(stderr, "b2r host comm: rank %d waited %.0f s in a barrier; a peer is gone, aborting the job\n", rank_, limit_s);
          abort();
          _exit(3);
        }
      }
    }
  }
}

b2r_status_t HostShmComm::reduce(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root, b2r_algo_t algo) {
  B2R_REQUIRE(ctl_ != nullptr, "host comm: not open");
  B2R_REQUIRE(root >= -1 && root < size_, "host comm: bad root %d", root);
  B2R_REQUIRE(dtype >= B2R_INT32 && dtype <= B2R_BFLOAT16 && op >= B2R_SUM && op <= B2R_MAX, "host comm: bad dtype/op");
  B2R_REQUIRE(sendbuf != nullptr || count == 0, "host comm: sendbuf is NULL");
  if (algo == B2R_ALGO_AUTO || algo == B2R_ALGO_NVLS) algo = B2R_ALGO_TWOSHOT;
  const size_t esz = b2r_dtype_size(dtype);
  const size_t chunk_elems = ctl_->window_bytes / esz;
  std::vector<const void*> send(size_);
  std::vector<void*> recv(size_), wk(size_);
  for (int r = 0; r < size_; ++r) {
    send[r] = window(r);
    recv[r] = result(r);
    wk[r] = work(r);
  }
  for (uint64_t first = 0; first < count; first += chunk_elems) {
    const uint64_t n = std::min<uint64_t>(chunk_elems, count - first);
    memcpy(window(rank_), (const char*)sendbuf + first * esz, n * esz);
    HostRankView v{rank_, size_, send.data(), recv.data(), wk.data()};
    host_collective_rank(v, n, dtype, op, root, algo, [this] { barrier(); });
    if ((root < 0 || root == rank_) && recvbuf) memcpy((char*)recvbuf + first * esz, result(rank_), n * esz);
    barrier();  // windows may be overwritten by the next chunk / call
  }
  return B2R_OK;
}

b2r_status_t HostShmComm::send(const void* buf, size_t total, int dest, int tag) {
  B2R_REQUIRE(ctl_ && dest >= 0 && dest < size_ && dest != rank_, "host comm: bad destination %d", dest);
  ShmP2PSlot* s = slot(rank_, dest);
  const size_t cap = ctl_->p2p_stride - sizeof(ShmP2PSlot);
  size_t done = 0;
  do {
    const size_t n = std::min(cap, total - done);
    unsigned spins = 0;
    while (s->full.load(std::memory_order_acquire)) {
      if (aborted()) _exit(3);
      if (++spins > 2000) sched_yield();
    }
    memcpy(s->data(), (const char*)buf + done, n);
    s->bytes = n;
    s->tag = tag;
    s->full.store(1, std::memory_order_release);
    done += n;
  } while (done < total);
  return B2R_OK;
}

b2r_status_t HostShmComm::recv(void* buf, size_t total, int source, int tag, int* tag_out) {
  B2R_REQUIRE(ctl_ && source >= 0 && source < size_ && source != rank_, "host comm: bad source %d", source);
  ShmP2PSlot* s = slot(source, rank_);
  size_t done = 0;
  int seen = tag;
  do {
    unsigned spins = 0;
    while (!s->full.load(std::memory_order_acquire)) {
      if (aborted()) _exit(3);
      if (++spins > 2000) sched_yield();
    }
    const size_t n = s->bytes;
    if (done + n > total || (tag >= 0 && s->tag != tag))
      return fail(B2R_ERR_INVALID, "host comm: rank %d received an unexpected message from %d (tag %d, %zu bytes)", rank_, source, s->tag, n);
    seen = s->tag;
    memcpy((char*)buf + done, s->data(), n);
    s->full.store(0, std::memory_order_release);
    done += n;
  } while (done < total);
  if (tag_out) *tag_out = seen;
  return B2R_OK;
}

}  // namespace b2r

// ------------------------------------------------------------------------------------------------ C API
using namespace b2r;

struct b2r_host_comm {
  HostShmComm impl;
};

extern "C" {

b2r_status_t b2r_host_comm_init(b2r_host_comm_t* out, const char* shm_name, int rank, int nranks, uint64_t window_bytes) {
  B2R_REQUIRE(out && shm_name, "b2r_host_comm_init: NULL argument");
  b2r_host_comm* c = new b2r_host_comm();
  // rank 0 creates (and unlinks on close); the others open
  b2r_status_t st = c->impl.open(shm_name, rank, nranks, window_bytes ? (size_t)window_bytes : ((size_t)16 << 20), rank == 0, rank == 0);
  if (st != B2R_OK) {
    delete c;
    return st;
  }
  c->impl.barrier();
  *out = c;
  return B2R_OK;
}

b2r_status_t b2r_host_comm_destroy(b2r_host_comm_t c) {
  if (!c) return B2R_OK;
  c->impl.barrier();  // nobody may still be reading a window of a rank that unmaps
  delete c;
  return B2R_OK;
}

b2r_status_t b2r_host_comm_reduce(b2r_host_comm_t c, const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root,
                                  b2r_algo_t algo) {
  B2R_REQUIRE(c, "b2r_host_comm_reduce: NULL communicator");
  return c->impl.reduce(sendbuf, recvbuf, count, dtype, op, root, algo);
}

b2r_status_t b2r_host_comm_barrier(b2r_host_comm_t c) {
  B2R_REQUIRE(c, "b2r_host_comm_barrier: NULL communicator");
  c->impl.barrier();
  return B2R_OK;
}

}  // extern "C"
