#include "bootstrap.h"

#include <errno.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <thread>

#include "../util/check.h"

namespace b2r {

namespace {

b2r_status_t send_all(int fd, const void* buf, size_t n) {
  const char* p = (const char*)buf;
  while (n) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k < 0) {
      if (errno == EINTR) continue;
      return fail(B2R_ERR_SYSTEM, "bootstrap send: %s", strerror(errno));
    }
    p += k;
    n -= (size_t)k;
  }
  return B2R_OK;
}

b2r_status_t recv_all(int fd, void* buf, size_t n) {
  char* p = (char*)buf;
  while (n) {
    ssize_t k = ::recv(fd, p, n, 0);
    if (k < 0) {
      if (errno == EINTR) continue;
      if (errno == EAGAIN || errno == EWOULDBLOCK) return fail(B2R_ERR_TIMEOUT, "bootstrap recv: the peer sent nothing before the deadline");
      return fail(B2R_ERR_SYSTEM, "bootstrap recv: %s", strerror(errno));
    }
    if (k == 0) return fail(B2R_ERR_SYSTEM, "bootstrap recv: peer closed the connection");
    p += k;
    n -= (size_t)k;
  }
  return B2R_OK;
}

// one descriptor + a 4-byte tag per message
b2r_status_t send_fd(int sock, int fd, int tag) {
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  iovec io;
  io.iov_base = &tag;
  io.iov_len = sizeof(tag);
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  for (;;) {
    ssize_t k = ::sendmsg(sock, &msg, MSG_NOSIGNAL);
    if (k < 0 && errno == EINTR) continue;
    if (k < 0) return fail(B2R_ERR_SYSTEM, "bootstrap sendmsg(SCM_RIGHTS): %s", strerror(errno));
    return B2R_OK;
  }
}

b2r_status_t recv_fd(int sock, int* fd, int* tag) {
  msghdr msg;
  memset(&msg, 0, sizeof(msg));
  iovec io;
  io.iov_base = tag;
  io.iov_len = sizeof(*tag);
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  for (;;) {
    ssize_t k = ::recvmsg(sock, &msg, 0);
    if (k < 0 && errno == EINTR) continue;
    if (k < 0) return fail(B2R_ERR_SYSTEM, "bootstrap recvmsg: %s", strerror(errno));
    if (k == 0) return fail(B2R_ERR_SYSTEM, "bootstrap recvmsg: peer closed the connection");
    break;
  }
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS) return fail(B2R_ERR_SYSTEM, "bootstrap: message carried no descriptor");
  memcpy(fd, CMSG_DATA(c), sizeof(int));
  return B2R_OK;
}

// receive timeout of a socket; 0 = block for ever
void set_recv_timeout(int fd, int ms) {
  timeval tv;
  tv.tv_sec = ms / 1000;
  tv.tv_usec = (ms % 1000) * 1000;
  ::setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
}

// The rendezvous socket lives in the abstract namespace, which any local process may connect to; the descriptors that
// travel over it grant access to GPU memory. Only peers running as the same user are let in.
bool same_user(int fd) {
  ucred cred;
  socklen_t len = sizeof(cred);
  if (::getsockopt(fd, SOL_SOCKET, SO_PEERCRED, &cred, &len) != 0) return false;
  return cred.uid == ::geteuid();
}

socklen_t make_addr(const std::string& id, sockaddr_un* a) {
  memset(a, 0, sizeof(*a));
  a->sun_family = AF_UNIX;
  std::string name = "b2r-" + id;
  if (name.size() > sizeof(a->sun_path) - 2) name.resize(sizeof(a->sun_path) - 2);
  a->sun_path[0] = '\0';  // abstract namespace: no file to clean up, vanishes with the process
  memcpy(a->sun_path + 1, name.data(), name.size());
  return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + name.size());
}

}  // namespace

Bootstrap::~Bootstrap() {
  for (int fd : peers_)
    if (fd >= 0) ::close(fd);
  if (listen_fd_ >= 0) ::close(listen_fd_);
}

b2r_status_t Bootstrap::init(const std::string& id, int rank, int nranks, int timeout_ms) {
  rank_ = rank;
  nranks_ = nranks;
  if (nranks == 1) return B2R_OK;
  if (timeout_ms <= 0) timeout_ms = 60000;
  sockaddr_un addr;
  socklen_t alen = make_addr(id, &addr);
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  if (rank == 0) {
    peers_.assign(nranks, -1);
    listen_fd_ = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (listen_fd_ < 0) return fail(B2R_ERR_SYSTEM, "bootstrap socket: %s", strerror(errno));
    if (::bind(listen_fd_, (sockaddr*)&addr, alen) < 0) return fail(B2R_ERR_SYSTEM, "bootstrap bind(%s): %s", id.c_str(), strerror(errno));
    if (::listen(listen_fd_, nranks) < 0) return fail(B2R_ERR_SYSTEM, "bootstrap listen: %s", strerror(errno));
    for (int got = 0; got < nranks - 1;) {
      pollfd p = {listen_fd_, POLLIN, 0};
      int left = (int)std::chrono::duration_cast<std::chrono::milliseconds>(deadline - std::chrono::steady_clock::now()).count();
      if (left <= 0 || ::poll(&p, 1, left) <= 0) return fail(B2R_ERR_TIMEOUT, "bootstrap: only %d of %d ranks connected", got + 1, nranks);
      int s = ::accept4(listen_fd_, nullptr, nullptr, SOCK_CLOEXEC);
      if (s < 0) continue;
      if (!same_user(s)) {  // not one of ours: drop it and keep waiting
        ::close(s);
        continue;
      }
      int r = -1;
      // a rank sends its number right after connecting; a client that connects and then says nothing is dropped after
      // two seconds instead of hanging rank 0, and the wait for the real ranks goes on
      set_recv_timeout(s, std::max(1, std::min(left, 2000)));
      b2r_status_t st = recv_all(s, &r, sizeof(r));
      set_recv_timeout(s, 0);
      if (st != B2R_OK) {
        ::close(s);
        continue;
      }
      if (r <= 0 || r >= nranks || peers_[r] != -1) {
        ::close(s);
        return fail(B2R_ERR_SYSTEM, "bootstrap: bad or duplicate rank %d", r);
      }
      peers_[r] = s;
      ++got;
    }
  } else {
    peers_.assign(1, -1);
    for (;;) {
      int s = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
      if (s < 0) return fail(B2R_ERR_SYSTEM, "bootstrap socket: %s", strerror(errno));
      if (::connect(s, (sockaddr*)&addr, alen) == 0) {
        peers_[0] = s;
        break;
      }
      ::close(s);
      if (std::chrono::steady_clock::now() > deadline) return fail(B2R_ERR_TIMEOUT, "bootstrap: rank 0 never listened on %s", id.c_str());
      std::this_thread::sleep_for(std::chrono::milliseconds(5));
    }
    B2R_CHECK(send_all(peers_[0], &rank, sizeof(rank)));
  }
  return barrier();
}

b2r_status_t Bootstrap::allgather(const void* mine, void* all, size_t bytes) {
  if (nranks_ == 1) {
    memcpy(all, mine, bytes);
    return B2R_OK;
  }
  if (rank_ == 0) {
    memcpy(all, mine, bytes);
    for (int r = 1; r < nranks_; ++r) B2R_CHECK(recv_all(peers_[r], (char*)all + (size_t)r * bytes, bytes));
    for (int r = 1; r < nranks_; ++r) B2R_CHECK(send_all(peers_[r], all, bytes * nranks_));
  } else {
    B2R_CHECK(send_all(peers_[0], mine, bytes));
    B2R_CHECK(recv_all(peers_[0], all, bytes * nranks_));
  }
  return B2R_OK;
}

b2r_status_t Bootstrap::barrier() {
  char c = 0;
  std::vector<char> all((size_t)nranks_);
  return allgather(&c, all.data(), 1);
}

b2r_status_t Bootstrap::allgather_fd(int my_fd, std::vector<int>* out) {
  out->assign(nranks_, -1);
  (*out)[rank_] = my_fd;
  if (nranks_ == 1) return B2R_OK;
  if (rank_ == 0) {
    for (int r = 1; r < nranks_; ++r) {
      int tag = -1;
      B2R_CHECK(recv_fd(peers_[r], &(*out)[r], &tag));
      if (tag != r) return fail(B2R_ERR_SYSTEM, "bootstrap: descriptor tag %d from rank %d", tag, r);
    }
    for (int r = 1; r < nranks_; ++r)
      for (int q = 0; q < nranks_; ++q)
        if (q != r) B2R_CHECK(send_fd(peers_[r], (*out)[q], q));
  } else {
    B2R_CHECK(send_fd(peers_[0], my_fd, rank_));
    for (int i = 0; i < nranks_ - 1; ++i) {
      int fd = -1, tag = -1;
      B2R_CHECK(recv_fd(peers_[0], &fd, &tag));
      if (tag < 0 || tag >= nranks_ || tag == rank_) return fail(B2R_ERR_SYSTEM, "bootstrap: bad descriptor tag %d", tag);
      (*out)[tag] = fd;
    }
  }
  return B2R_OK;
}

b2r_status_t Bootstrap::bcast_fd(int* fd) {
  if (nranks_ == 1) return B2R_OK;
  if (rank_ == 0) {
    for (int r = 1; r < nranks_; ++r) B2R_CHECK(send_fd(peers_[r], *fd, 0));
  } else {
    int tag = -1;
    B2R_CHECK(recv_fd(peers_[0], fd, &tag));
  }
  return B2R_OK;
}

}  // namespace b2r

// Self-test hook (used by tests/test_bootstrap.py on CPU-only machines): every rank contributes one value and one pipe
// descriptor; afterwards every rank writes its rank into every other rank's pipe through the RECEIVED descriptors and
// reads nranks-1 bytes from its own pipe. Returns B2R_OK when both the byte all-gather and the SCM_RIGHTS path work.
extern "C" b2r_status_t b2r_bootstrap_selftest(const char* id, int rank, int nranks) {
  using namespace b2r;
  Bootstrap b;
  B2R_CHECK(b.init(id, rank, nranks, 20000));
  std::vector<int> vals(nranks, -1);
  int mine = 1000 + rank;
  B2R_CHECK(b.allgather(&mine, vals.data(), sizeof(int)));
  for (int r = 0; r < nranks; ++r)
    if (vals[r] != 1000 + r) return fail(B2R_ERR_SYSTEM, "selftest: allgather slot %d holds %d", r, vals[r]);
  int pfd[2];
  if (::pipe(pfd) != 0) return fail(B2R_ERR_SYSTEM, "selftest: pipe: %s", strerror(errno));
  std::vector<int> wfds;
  B2R_CHECK(b.allgather_fd(pfd[1], &wfds));
  for (int r = 0; r < nranks; ++r) {
    if (r == rank) continue;
    unsigned char byte = (unsigned char)rank;
    if (::write(wfds[r], &byte, 1) != 1) return fail(B2R_ERR_SYSTEM, "selftest: write to rank %d's pipe: %s", r, strerror(errno));
  }
  B2R_CHECK(b.barrier());
  int seen = 0;
  for (int i = 0; i < nranks - 1; ++i) {
    unsigned char byte = 0xff;
    if (::read(pfd[0], &byte, 1) != 1) return fail(B2R_ERR_SYSTEM, "selftest: short read");
    if (byte >= nranks || byte == rank) return fail(B2R_ERR_SYSTEM, "selftest: unexpected byte %d", (int)byte);
    seen |= 1 << byte;
  }
  if (seen != (((1 << nranks) - 1) & ~(1 << rank))) return fail(B2R_ERR_SYSTEM, "selftest: missing writers (mask %x)", seen);
  int bfd = rank == 0 ? pfd[1] : -1;
  B2R_CHECK(b.bcast_fd(&bfd));
  if (rank != 0) {
    unsigned char byte = (unsigned char)(100 + rank);
    if (::write(bfd, &byte, 1) != 1) return fail(B2R_ERR_SYSTEM, "selftest: write through broadcast descriptor failed");
  }
  B2R_CHECK(b.barrier());
  if (rank == 0) {
    for (int i = 0; i < nranks - 1; ++i) {
      unsigned char byte = 0;
      if (::read(pfd[0], &byte, 1) != 1 || byte <= 100) return fail(B2R_ERR_SYSTEM, "selftest: broadcast descriptor path broken");
    }
  }
  for (int r = 0; r < nranks; ++r)
    if (r != rank && wfds[r] >= 0) ::close(wfds[r]);
  if (rank != 0 && bfd >= 0) ::close(bfd);
  ::close(pfd[0]);
  ::close(pfd[1]);
  return B2R_OK;
}
