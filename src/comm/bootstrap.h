// bootstrap.h — out-of-band rendezvous of the ranks of one node over an abstract Unix-domain socket.
// The reference relies on mpirun + MPI_Init for this (mpi/reduce.c:32-34); there is no MPI on the target image, and the
// things to exchange are file descriptors (VMM shareable handles), which only SCM_RIGHTS ancillary messages can carry.
// Star topology: rank 0 listens on "\0b2r-<id>", everybody else connects; every collective below goes through rank 0.
#pragma once
#include <cstddef>
#include <string>
#include <vector>

#include "b2r.h"

namespace b2r {

class Bootstrap {
 public:
  Bootstrap() = default;
  ~Bootstrap();
  Bootstrap(const Bootstrap&) = delete;
  Bootstrap& operator=(const Bootstrap&) = delete;

  b2r_status_t init(const std::string& id, int rank, int nranks, int timeout_ms);
  // all[r*bytes .. (r+1)*bytes) = rank r's `mine`
  b2r_status_t allgather(const void* mine, void* all, size_t bytes);
  // fds_out[r] = a descriptor in THIS process referring to rank r's `my_fd` (fds_out[rank] == my_fd)
  b2r_status_t allgather_fd(int my_fd, std::vector<int>* fds_out);
  // rank 0's `fd` duplicated into every process (returned through *fd on the others)
  b2r_status_t bcast_fd(int* fd);
  b2r_status_t barrier();
  int rank() const { return rank_; }
  int nranks() const { return nranks_; }

 private:
  int rank_ = 0, nranks_ = 1;
  int listen_fd_ = -1;
  std::vector<int> peers_;  // rank 0: socket to rank r (peers_[0] unused); others: peers_[0] = socket to rank 0
};

}  // namespace b2r
