// coll_ctx.h — what device code needs to know about a communicator: heap layout, context, mailbox, signal helpers.
// Synchronisation uses monotonically increasing epoch values written with st.release.sys and polled with
// ld.acquire.sys (no resets -> no ABA), and every wait has a %globaltimer deadline that raises an error
// word instead of hanging the GPU.
#pragma once
#include <cstdint>

#include "b2r.h"

namespace b2r {

// ---- layout of every rank's symmetric heap (identical offsets on all ranks)
constexpr int kCollMaxCtas = 256;                    // CTAs per collective kernel (per rank)
constexpr size_t kOffFlags = 0;                      // uint32 flags[kCollMaxCtas][B2R_MAX_RANKS]: per-CTA epoch flags
constexpr size_t kOffMailbox = 32 * 1024;            // XgMailbox (scalar-sharded reduce)
constexpr size_t kOffError = 40 * 1024;              // uint32 error word (device-side timeouts)
constexpr size_t kOffCtx = 44 * 1024;                // device copy of CollCtx (read by the scalar-sharded kernels)
constexpr size_t kOffMeta = 64 * 1024;               // PeerMeta meta[kCollMaxCtas][B2R_MAX_RANKS]: per-collective buffer offsets
constexpr size_t kOffPushData = 128 * 1024;          // uint32 [kCollMaxCtas][B2R_MAX_RANKS]: push algorithm, "block of round i from rank s has landed"
constexpr size_t kOffPostedReady = 144 * 1024;       // uint32 [kPostedCtas][B2R_MAX_RANKS]: posted reduce, at the root: "CTA b of rank s has staged call n"
constexpr size_t kOffPostedAck = 152 * 1024;         // uint32 [B2R_MAX_RANKS][kPostedCtas]: posted reduce, "CTA b of root r has consumed call n"
constexpr size_t kOffPushAck = 160 * 1024;           // uint32 [kCollMaxCtas][B2R_MAX_RANKS]: push algorithm, "worker w has consumed round i"
constexpr size_t kOffLLAck = 176 * 1024;             // uint32 [B2R_MAX_RANKS]: LL path, "rank r has finished LL call n" (flow control)
constexpr size_t kOffLL = 192 * 1024;                // LL path: packet slots [call parity][source rank][ll_slot_bytes(P)], up to kOffPushFlags
constexpr size_t kOffPushFlags = 1 << 20;            // push algorithm: uint32 [B2R_MAX_RANKS][kPushChunkCap], "chunk k of source s landed in call n"
constexpr size_t kPushChunkCap = 16384;              //   chunks per slice the flag table can name (x 16 ranks x 4 bytes = 1 MiB)
constexpr size_t kCtrlBytes = 2 << 20;               // control region size; staging starts here

// LL ("low latency") small-message path: every 16-byte vector travels as four 8-byte packets {32 data bits, 32-bit call
// number}, so a slot holds half its size in payload. One CTA of up to kLLThreads threads, kLLVecPerThread vectors each.
constexpr int kLLThreads = 512;
constexpr int kLLVecPerThread = 4;
constexpr size_t ll_slot_bytes(int nranks) { return ((kOffPushFlags - kOffLL) / (2 * (size_t)nranks)) & ~(size_t)255; }
constexpr size_t ll_max_msg_bytes(int nranks) {
  return ll_slot_bytes(nranks) / 2 < (size_t)kLLThreads * kLLVecPerThread * 16 ? ll_slot_bytes(nranks) / 2 : (size_t)kLLThreads * kLLVecPerThread * 16;
}
// The optional large instance of the same layout (communicators created with B2R_LL_BIG_MB=n get n MiB behind the user heap):
// kLLBigCtas CTAs own kLLBigVecPerCta vectors each, the acknowledgement words sit at the end of the region.
// Round 2, measured on 8 GPUs: with 16 CTAs x 16 KiB a 16-64 KiB message kept ONE to FOUR CTAs busy and needed 12.3-12.6 us
// (NCCL: 10.8-11.5); the partition is now 64 CTAs x 4 KiB (one vector per thread of a 256-thread CTA), so a 16 KiB message
// already spreads over four SMs. Messages of at most kLLSmallMaxBytes stay on the one-CTA instance in the control region.
constexpr int kLLBigCtas = 64;
constexpr int kLLBigVecPerCta = 256;   // 4 KiB of payload per CTA
constexpr int kLLBigThreads = 256;
constexpr size_t kLLSmallMaxBytes = 8192;
constexpr size_t kLLBigAckBytes = (size_t)kLLBigCtas * B2R_MAX_RANKS * sizeof(unsigned int);
constexpr size_t ll_big_slot_bytes(size_t region_bytes, int nranks) {
  return region_bytes <= kLLBigAckBytes ? 0 : ((region_bytes - kLLBigAckBytes) / (2 * (size_t)nranks)) & ~(size_t)255;
}
constexpr size_t ll_big_max_msg_bytes(size_t region_bytes, int nranks) {
  return ll_big_slot_bytes(region_bytes, nranks) / 2 < (size_t)kLLBigCtas * kLLBigVecPerCta * 16 ? ll_big_slot_bytes(region_bytes, nranks) / 2
                                                                                                   : (size_t)kLLBigCtas * kLLBigVecPerCta * 16;
}

// Posted reduce (coll_posted.cuh): a FIXED grid of kPostedCtas CTAs per rank; chunk k of kPostedChunkVecs 16-byte vectors belongs to
// CTA k % grid on every rank in every call, so one ready word per (CTA, source) and one acknowledgement word per (root, CTA) is
// all the flow control it needs. The two staging slots (call parity) live at the start of the work region.
// Geometry, second version. The first (64 CTAs x 256 threads, 16 KiB chunks, four vectors per thread) kept 1 MiB of multimem
// requests in flight from the root and measured 245 GB/s for the 16-byte forms on 4 GPUs (110 for the 4- and 8-byte forms):
// latency-bound. Now 128 CTAs x 512 threads; a chunk is one vector per thread (8 KiB), so a 1 MiB message already occupies
// every CTA, and a thread keeps four chunks in flight.
constexpr int kPostedCtas = 128;
constexpr int kPostedThreads = 512;
constexpr unsigned int kPostedChunkVecs = 512;  // 8 KiB
constexpr size_t kPostedMaxSlotBytes = (size_t)64 << 20;
static_assert(kOffPostedReady + (size_t)kPostedCtas * B2R_MAX_RANKS * 4 <= kOffPostedAck, "posted: ready words overlap the ack words");
static_assert(kOffPostedAck + (size_t)B2R_MAX_RANKS * kPostedCtas * 4 <= kOffPushAck, "posted: ack words overlap the ring's");

// Where rank r keeps its send/recv buffer for the collective in flight, published to every peer inside the start
// barrier, so buffers do NOT have to sit at the same heap offset on every rank (only the NVLS path needs that).
struct PeerMeta {
  unsigned long long send_off, recv_off;
};

// What a kernel needs to know about the communicator. Plain data: passed by value in kernel parameters.
struct CollCtx {
  char* heap[B2R_MAX_RANKS];  // heap[p]: rank p's symmetric heap as mapped in THIS rank's address space
  char* mc;                   // multicast (NVLS) mapping of the same heap, or nullptr
  int rank, nranks;
  unsigned long long timeout_ns;  // 0 = wait forever
  unsigned int fault_delay_us;    // fault injection: sleep before the first signal of a collective
  int fault_skip_signal;          // fault injection: never signal (peers must time out)
};

// Scalar-sharded reduce: partials travel as "packets" that carry their own arrival flag (the LL idea applied to one scalar):
// a 16-byte slot = two 8-byte words {low 32 bits of the value | call number << 32} and {high 32 bits | call number << 32},
// written with ONE 16-byte store. Each 8-byte word is a single-copy-atomic access, so a reader that sees the call number it is
// waiting for in both words has the whole value: no separate flag, no release fence on the sender's exit path (the first
// version's st.release.sys cost one NVLink round trip per call). Slots rotate over kXgSlots calls, so a rank may run up to
// kXgSlots calls ahead of the root (the two-slot version let it run one call ahead, which put the root's acknowledgement
// round trip on every step's critical path: 0.81 / 0.74 / 0.76 weak-scaling efficiency at 2 / 4 / 8 GPUs in round 1).
constexpr int kXgSlots = 16;
struct alignas(16) XgPacket {
  unsigned long long lo, hi;
};
struct XgMailbox {
  XgPacket val[kXgSlots][B2R_MAX_RANKS];  // [call % kXgSlots][source rank]: partial scalar in accumulator precision
  XgPacket result[kXgSlots];              // [call % kXgSlots]: final scalar pushed by the root (scalar allreduce)
  unsigned int ack;                       // last call number the root has finished reading (flow control, monotonic)
  unsigned int calls;                     // scalar-sharded calls this rank has made (written by this rank's own kernels only):
                                          // the call number lives on the device, so a captured CUDA graph of calls can be replayed
};
static_assert(sizeof(XgMailbox) <= kOffError - kOffMailbox, "XgMailbox outgrew its slice of the control region");

#if defined(__CUDACC__)
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int ld_volatile_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long coll_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Spins until *flag >= want (wrap-safe). Returns false if the deadline passed or another waiter already gave up; the
// error word lives in this rank's own control region and is read by b2r_comm_check(). The first ~10 us are a pure spin
// (the expected wait is one NVLink hop, 1-2 us; the first version slept 64 ns after every 128 polls, which added up to
// half a microsecond of reaction time to every barrier); only a wait that is clearly long backs off with __nanosleep.
__device__ __forceinline__ bool wait_flag_ge(const unsigned int* flag, unsigned int want, const CollCtx& c) {
  unsigned int* err = (unsigned int*)(c.heap[c.rank] + kOffError);
  unsigned long long t0 = 0;
  unsigned int spins = 0;
  bool slow = false;
  while ((int)(ld_acquire_sys(flag) - want) < 0) {
    if (++spins >= 32 || slow) {
      spins = 0;
      if (ld_volatile_u32(err)) return false;
      const unsigned long long now = coll_globaltimer();
      if (t0 == 0) t0 = now;
      else if (c.timeout_ns && now - t0 > c.timeout_ns) {
        atomicExch(err, 1u);
        return false;
      }
      if (now - t0 > 10000ull) {
        slow = true;
        __nanosleep(100);
      }
    }
  }
  return true;
}

// The same wait with relaxed polls (no acquire fence): for counters that only gate the REUSE of a buffer the waiter is about to
// overwrite — nothing is read on the strength of the value, so there is nothing to acquire.
__device__ __forceinline__ unsigned int ld_relaxed_sys_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool wait_flag_ge_relaxed(const unsigned int* flag, unsigned int want, const CollCtx& c) {
  unsigned int* err = (unsigned int*)(c.heap[c.rank] + kOffError);
  unsigned long long t0 = 0;
  unsigned int spins = 0;
  bool slow = false;
  while ((int)(ld_relaxed_sys_u32(flag) - want) < 0) {
    if (++spins >= 32 || slow) {
      spins = 0;
      if (ld_volatile_u32(err)) return false;
      const unsigned long long now = coll_globaltimer();
      if (t0 == 0) t0 = now;
      else if (c.timeout_ns && now - t0 > c.timeout_ns) {
        atomicExch(err, 1u);
        return false;
      }
      if (now - t0 > 10000ull) {
        slow = true;
        __nanosleep(100);
      }
    }
  }
  return true;
}

// Scalar-sharded reduce packets (see XgMailbox).
__device__ __forceinline__ void xg_post(XgPacket* slot, unsigned long long bits, unsigned int seq) {
  const unsigned long long tag = (unsigned long long)seq << 32;
  const unsigned long long lo = tag | (bits & 0xffffffffull), hi = tag | (bits >> 32);
  asm volatile("st.relaxed.sys.global.v2.b64 [%0], {%1,%2};" ::"l"(slot), "l"(lo), "l"(hi) : "memory");
}
// Polls until the packet of call `seq` has landed; false on time-out (error word set) or when another waiter gave up.
__device__ __forceinline__ bool xg_wait(const XgPacket* slot, unsigned int seq, const CollCtx& c, unsigned long long* bits) {
  unsigned int* err = (unsigned int*)(c.heap[c.rank] + kOffError);
  unsigned long long t0 = 0;
  unsigned int spins = 0;
  bool slow = false;
  for (;;) {
    unsigned long long lo, hi;
    asm volatile("ld.relaxed.sys.global.v2.b64 {%0,%1}, [%2];" : "=l"(lo), "=l"(hi) : "l"(slot) : "memory");
    if ((unsigned int)(lo >> 32) == seq && (unsigned int)(hi >> 32) == seq) {
      *bits = (lo & 0xffffffffull) | (hi << 32);
      return true;
    }
    if (++spins >= 32 || slow) {
      spins = 0;
      if (ld_volatile_u32(err)) return false;
      const unsigned long long now = coll_globaltimer();
      if (t0 == 0) t0 = now;
      else if (c.timeout_ns && now - t0 > c.timeout_ns) {
        atomicExch(err, 1u);
        return false;
      }
      if (now - t0 > 10000ull) {
        slow = true;
        __nanosleep(100);
      }
    }
  }
}

__device__ __forceinline__ void fault_delay(const CollCtx& c) {
  if (c.fault_delay_us == 0) return;
  const unsigned long long t0 = coll_globaltimer();
  while (coll_globaltimer() - t0 < (unsigned long long)c.fault_delay_us * 1000ull) __nanosleep(1000);
}
#endif

}  // namespace b2r
