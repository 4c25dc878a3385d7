// workspace.h — per-stream scratch of the single-GPU reduction (internal).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace b2r {
constexpr int kMaxCtas = 4096;    // upper bound on partials (CTAs or clusters) of one launch
constexpr int kStageBufs = 3;     // staging ring of b2r_reduce_from_host
constexpr int kMaxChunks = 4096;  // chunks per b2r_reduce_from_host call
constexpr int kTraceRing = 8;     // traced launches kept (a ring: the stamps of the last kTraceRing launches can be read back)
}  // namespace b2r

struct b2r_workspace {
  int device = 0;
  int sm_count = 0;
  void* d_partials = nullptr;             // kMaxCtas x 8 bytes
  unsigned* d_ticket = nullptr;           // arrival counter, zero between launches
  unsigned* d_tile_ctr = nullptr;         // dynamic tail: next unclaimed tile, zero between launches (nullptr: no dynamic tail)
  unsigned long long* d_acc64 = nullptr;  // int32 SUM: packed {arrivals, running sum}, zero between launches (nullptr: not offered)
  int max_partials = 0;                   // capacity of d_partials in 8-byte entries
  unsigned long long* d_trace = nullptr;  // kTraceRing x kMaxCtas x 4 stamps
  unsigned trace_seq = 0;                 // traced launches so far (slot = trace_seq % kTraceRing)
  int trace_ctas[b2r::kTraceRing] = {0};  // CTAs of the launch in each slot
  bool trace_enabled = false;
  int last_ctas = 0;
  int last_geom[4] = {0, 0, 0, 0};
  int last_dyn_tiles = 0;
  const void* last_out = nullptr;         // output pointer of the previous launch (PDL overlap guard)
  // b2r_reduce_from_host
  cudaStream_t copy_stream = nullptr;
  void* d_stage[b2r::kStageBufs] = {nullptr, nullptr, nullptr};
  uint64_t stage_bytes = 0;
  cudaEvent_t ev_copied[b2r::kStageBufs] = {nullptr, nullptr, nullptr};
  cudaEvent_t ev_consumed[b2r::kStageBufs] = {nullptr, nullptr, nullptr};
  void* d_chunk_out = nullptr;
  void* h_chunk_out = nullptr;
};
