// comm.h — communicator internals: symmetric heap layout, device-visible context, single-process groups.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <memory>
#include <mutex>
#include <vector>

#include "../core/workspace.h"
#include "b2r.h"
#include "bootstrap.h"
#include "coll_ctx.h"

namespace b2r {

enum CommMode { kModeAll = 0, kModeRank = 1, kModeLoopback = 2 };

struct PendingCall;  // coll.cu

// State shared by the communicators of one process (init_all / loopback).
struct CommGroup {
  int nranks = 0;
  CommMode mode = kModeAll;
  std::vector<b2r_comm*> members;
  // VMM resources (kModeAll / kModeRank: one group object per process)
  std::vector<CUmemGenericAllocationHandle> handles;  // imported or created, index = rank
  CUdeviceptr va_base = 0;
  size_t va_size = 0;
  CUmemGenericAllocationHandle mc_handle = 0;
  CUdeviceptr mc_va = 0;
  std::vector<int> devices;
  std::vector<void*> loop_heaps;  // kModeLoopback: cudaMalloc'ed heaps
  // loopback: collective calls recorded until every virtual rank has called
  std::mutex mu;
  std::vector<std::shared_ptr<PendingCall>> pending;
  int npending = 0;
  int alive = 0;
  ~CommGroup();
};

}  // namespace b2r

struct b2r_comm {
  b2r::CommMode mode = b2r::kModeAll;
  int rank = 0, nranks = 1, device = 0, sm_count = 0;
  b2r::CollCtx ctx;
  size_t heap_bytes = 0, stage_bytes = 0, work_bytes = 0, user_off = 0, bump = 0;
  size_t user_end = 0;                       // first byte past the user allocations (== heap_bytes without an LL-big region)
  size_t ll_big_off = 0, ll_big_bytes = 0;   // optional large LL region behind the user heap (B2R_LL_BIG_MB at creation)
  uint32_t epoch = 0;   // barrier epoch counter (same sequence on every rank)
  uint32_t xg_seq = 0;  // scalar-sharded call counter
  uint32_t push_seq = 0;  // push-algorithm call counter (flag values are seq << 16 | round)
  uint32_t ll_seq = 0;    // LL path call counter (consecutive: slot parity and the n-2 acknowledgement rule depend on it)
  uint32_t ll_big_seq = 0;  // the same for the large LL region (separate slots and acknowledgements, separate numbering)
  uint32_t posted_seq = 0;  // posted-reduce call counter (slot parity = seq & 1)
  int posted_root[2] = {-1, -1};  // root of the last posted call of each parity: whose acknowledgement frees the slot
  b2r_workspace* ws = nullptr;
  int max_ctas = 0, timeout_ms = 0;
  bool has_mc = false;
  bool symmetric_only = false;  // caller promised symmetric buffers only: no staging, no chunking
  b2r::Bootstrap* boot = nullptr;
  std::shared_ptr<b2r::CommGroup> group;

  char* my_heap() const { return ctx.heap[rank]; }
  char* stage_send() const { return my_heap() + b2r::kCtrlBytes; }
  char* stage_recv() const { return my_heap() + b2r::kCtrlBytes + stage_bytes; }
  char* stage_work() const { return my_heap() + b2r::kCtrlBytes + 2 * stage_bytes; }
  // offset of p inside my heap if [p, p+bytes) lies in the user/staging part and is 16-byte aligned, else -1
  int64_t heap_offset(const void* p, size_t bytes) const {
    const char* c = (const char*)p;
    if (c < my_heap() + b2r::kCtrlBytes || c + bytes > my_heap() + heap_bytes) return -1;
    const int64_t off = c - my_heap();
    return (off % 16) ? -1 : off;
  }
};
