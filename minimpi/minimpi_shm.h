// minimpi_shm.h — layout of the shared segment (shared by libminimpi and b2r_mpirun).
#pragma once
#include <atomic>
#include <cstddef>

namespace minimpi {

struct Control {
  std::atomic<unsigned> bar_count, bar_gen, abort_flag;
  int nranks;
  size_t window_bytes;                                  // per-rank staging for collectives
  size_t off_windows, off_results, off_work, off_p2p;   // byte offsets from the segment base
  size_t p2p_stride;                                    // bytes per (src, dst) slot incl. header
  size_t total_bytes;
};

struct P2PSlot {
  std::atomic<unsigned> full;
  int tag;
  size_t bytes;
  char pad[48];
  char* data() { return (char*)(this + 1); }
};

inline size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

inline void layout(Control* c, int nranks, size_t window_bytes, size_t p2p_payload) {
  c->nranks = nranks;
  c->window_bytes = window_bytes;
  c->off_windows = round_up(sizeof(Control), 4096);
  c->off_results = c->off_windows + (size_t)nranks * window_bytes;
  c->off_work = c->off_results + (size_t)nranks * window_bytes;
  c->off_p2p = c->off_work + (size_t)nranks * window_bytes;
  c->p2p_stride = round_up(sizeof(P2PSlot) + p2p_payload, 4096);
  c->total_bytes = c->off_p2p + (size_t)nranks * nranks * c->p2p_stride;
}

}  // namespace minimpi
