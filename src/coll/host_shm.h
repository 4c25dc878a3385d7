// host_shm.h — CPU ranks that are separate PROCESSES sharing one POSIX shared-memory segment: the host loopback
// transport of SURVEY §4 in its multi-process form. One implementation serves the bundled mini-MPI (minimpi/) and the
// library's host communicator (b2r_host_comm_*, Python HostCommunicator under torch.distributed/gloo).
// Layout: Control | P windows (send staging) | P result windows | P work windows | P*P point-to-point slots.
#pragma once
#include <atomic>
#include <cstddef>
#include <string>

#include "b2r.h"

namespace b2r {

struct ShmControl {
  std::atomic<unsigned> bar_count, bar_gen, abort_flag;
  std::atomic<unsigned> ready;  // 0 in a fresh (zero-filled) segment; the creator release-stores 1 as the LAST step of its set-up
  int nranks;
  size_t window_bytes;                                 // per-rank staging for collectives
  size_t off_windows, off_results, off_work, off_p2p;  // byte offsets from the segment base
  size_t p2p_stride;                                   // bytes per (src, dst) slot incl. header
  size_t total_bytes;
};

struct ShmP2PSlot {
  std::atomic<unsigned> full;
  int tag;
  size_t bytes;
  char pad[48];
  char* data() { return (char*)(this + 1); }
};

// Fills the layout fields of `c` (does not touch the atomics).
void shm_layout(ShmControl* c, int nranks, size_t window_bytes, size_t p2p_payload);

class HostShmComm {
 public:
  HostShmComm() = default;
  ~HostShmComm();
  HostShmComm(const HostShmComm&) = delete;
  HostShmComm& operator=(const HostShmComm&) = delete;

  // create = true: make and size the segment (rank 0, or a launcher); false: open an existing one (retries for a while)
  b2r_status_t open(const std::string& name, int rank, int nranks, size_t window_bytes, bool create, bool unlink_on_close);
  void close();
  int rank() const { return rank_; }
  int size() const { return size_; }
  bool aborted() const;
  void abort();

  void barrier();
  // root >= 0: reduce to root (recvbuf may be null elsewhere); root == -1: allreduce
  b2r_status_t reduce(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root, b2r_algo_t algo);
  b2r_status_t send(const void* buf, size_t bytes, int dest, int tag);
  b2r_status_t recv(void* buf, size_t bytes, int source, int tag, int* tag_out);

 private:
  char* window(int r) const { return base_ + ctl_->off_windows + (size_t)r * ctl_->window_bytes; }
  char* result(int r) const { return base_ + ctl_->off_results + (size_t)r * ctl_->window_bytes; }
  char* work(int r) const { return base_ + ctl_->off_work + (size_t)r * ctl_->window_bytes; }
  ShmP2PSlot* slot(int src, int dst) const { return (ShmP2PSlot*)(base_ + ctl_->off_p2p + ((size_t)src * size_ + dst) * ctl_->p2p_stride); }

  ShmControl* ctl_ = nullptr;
  char* base_ = nullptr;
  size_t map_bytes_ = 0;
  int rank_ = -1, size_ = 0;
  std::string name_;
  bool unlink_ = false;
};

}  // namespace b2r
