// b2r_mpirun — launcher of the bundled mini-MPI: `b2r_mpirun -np N [--window-mb M] prog [args...]`.
// Stands in for the reference's `mpirun -mode VN|CO` under SLURM (mpi/ccni_vn.sh:7-9, mpi/submit_all.sh:3-5) on one host:
// creates the shared segment, forks N ranks with MINIMPI_RANK/SIZE/SHM set, waits, propagates the worst exit code.
#include <fcntl.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../src/coll/host_shm.h"

int main(int argc, char** argv) {
  int np = 2;
  size_t window_mb = 32;
  int i = 1;
  for (; i < argc; ++i) {
    if ((!strcmp(argv[i], "-np") || !strcmp(argv[i], "-n")) && i + 1 < argc) np = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--window-mb") && i + 1 < argc) window_mb = (size_t)atol(argv[++i]);
    else break;
  }
  if (i >= argc || np < 1 || np > 256) {
    fprintf(stderr, "usage: b2r_mpirun -np N [--window-mb M] prog [args...]\n");
    return 2;
  }
  b2r::ShmControl tmp;
  b2r::shm_layout(&tmp, np, window_mb << 20, 4 << 20);
  std::string name = "/b2r-minimpi-" + std::to_string(getpid());
  int fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
  if (fd < 0) { perror("b2r_mpirun: shm_open"); return 2; }
  if (ftruncate(fd, (off_t)tmp.total_bytes) != 0) { perror("b2r_mpirun: ftruncate"); shm_unlink(name.c_str()); return 2; }
  void* base = mmap(nullptr, tmp.total_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (base == MAP_FAILED) { perror("b2r_mpirun: mmap"); shm_unlink(name.c_str()); return 2; }
  // the segment is fresh and zero-filled (barrier words, abort flag): write the layout, then publish `ready` (host_shm.cpp)
  b2r::ShmControl* ctl = (b2r::ShmControl*)base;
  b2r::shm_layout(ctl, np, window_mb << 20, 4 << 20);
  ctl->ready.store(1, std::memory_order_release);

  std::vector<pid_t> pids;
  for (int r = 0; r < np; ++r) {
    pid_t p = fork();
    if (p == 0) {
      setenv("MINIMPI_SHM", name.c_str(), 1);
      setenv("MINIMPI_RANK", std::to_string(r).c_str(), 1);
      setenv("MINIMPI_SIZE", std::to_string(np).c_str(), 1);
      execvp(argv[i], argv + i);
      perror("b2r_mpirun: exec");
      _exit(127);
    }
    pids.push_back(p);
  }
  int worst = 0;
  for (size_t k = 0; k < pids.size(); ++k) {
    int st = 0;
    waitpid(pids[k], &st, 0);
    int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
    if (code != 0) {
      ctl->abort_flag.store(1);  // unblock ranks spinning on a barrier with a dead peer
      if (code > worst) worst = code;
    }
  }
  munmap(base, tmp.total_bytes);
  shm_unlink(name.c_str());
  return worst;
}
