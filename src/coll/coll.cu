// coll.cu — host side of the collectives: algorithm selection, staging / chunking of non-symmetric buffers, launch
// (one kernel per rank; ONE cooperative kernel for all virtual ranks in loopback mode), and the scalar-sharded reduce.
// API counterpart: MPI_Reduce(sendbuf, recvbuf, count, type, op, root, comm) as used by mpi/reduce.c:76,90.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <vector>

#include "../comm/comm.h"
#include "../core/reduce_internal.h"
#include "../util/check.h"
#include "../util/env.h"
#include "../util/nvtx.h"
#include "coll_kernels.cuh"

namespace b2r {

enum CallKind { kCallReduce = 0, kCallAllreduce = 1, kCallScalar = 2, kCallBarrier = 3 };

struct PendingCall {
  int kind;
  const void* send;
  void* recv;
  uint64_t count;
  b2r_dtype_t dtype;
  b2r_op_t op;
  int root, all;
  b2r_algo_t algo;
  cudaStream_t stream;
  b2r_launch_cfg_t cfg;  // scalar-sharded only
  bool has_cfg;
};

namespace {

using CollKernel = void (*)(const CollLaunch);

template <int D, int O, int UNR>
CollKernel pick_kind(int kind) {
  switch (kind) {
    case kCollOneshot: return coll_kernel<D, O, kCollOneshot, UNR>;
    case kCollTwoshot: return coll_kernel<D, O, kCollTwoshot, UNR>;
    case kCollTree: return coll_kernel<D, O, kCollTree, UNR>;
    case kCollNvls: return Nvls<D, O>::kSupported ? coll_kernel<D, O, kCollNvls, UNR> : nullptr;
    case kCollBarrier: return coll_kernel<D, O, kCollBarrier, 4>;
    case kCollPush: return coll_kernel<D, O, kCollPush, 4>;
    case kCollRing: return coll_kernel<D, O, kCollRing, 4>;
    case kCollLL: return coll_kernel<D, O, kCollLL, 4>;
    case kCollPosted: return coll_kernel<D, O, kCollPosted, 4>;
  }
  return nullptr;
}
template <int D>
CollKernel pick_op(int op, int kind, int unroll) {
  switch (op) {
    case B2R_SUM: return unroll == 8 ? pick_kind<D, B2R_SUM, 8>(kind) : pick_kind<D, B2R_SUM, 4>(kind);
    case B2R_MIN: return unroll == 8 ? pick_kind<D, B2R_MIN, 8>(kind) : pick_kind<D, B2R_MIN, 4>(kind);
    case B2R_MAX: return unroll == 8 ? pick_kind<D, B2R_MAX, 8>(kind) : pick_kind<D, B2R_MAX, 4>(kind);
  }
  return nullptr;
}
CollKernel pick_kernel(b2r_dtype_t dtype, int op, int kind, int unroll) {
  switch (dtype) {
    case B2R_INT32: return pick_op<B2R_INT32>(op, kind, unroll);
    case B2R_FLOAT32: return pick_op<B2R_FLOAT32>(op, kind, unroll);
    case B2R_FLOAT64: return pick_op<B2R_FLOAT64>(op, kind, unroll);
    case B2R_BFLOAT16: return pick_op<B2R_BFLOAT16>(op, kind, unroll);
  }
  return nullptr;
}

// Launch shape per algorithm (identical on every rank: a pure function of the algorithm, the message size, the
// communicator config and the environment). Measured on 8x B200 (profiles/collectives.md):
//   NVLS   wants FEW threads: 32 CTAs x 256 threads x 8 vectors in flight reach 843 GB/s bus bandwidth at 1 GiB, while
//          128 x 512 threads drop to 530-650 (the switch's reduction engine is the resource, not SM-side parallelism);
//   pulls  (one-shot / two-shot / tree) want many: 128 CTAs x 512 threads to cover the NVLink load latency;
//   push   is bounded by its staging ring (see run_collective).
struct Shape {
  int ctas, threads, unroll;
};
Shape coll_shape(const b2r_comm* c, int kind, uint64_t bytes, int nvirtual, int unit_bytes = 16) {
  Shape s;
  int cap;
  uint64_t per_cta;
  if (kind == kCollNvls) {
    s.threads = 256;
    s.unroll = bytes >= (8u << 20) ? 8 : 4;
    cap = 32;
    per_cta = 32768;
    // The switch wants a certain number of BYTES in flight, and a multimem instruction moves 16 (fp32 x4, bf16 x8), 8 (fp64) or 4
    // (32-bit integers) of them: the narrow forms get proportionally more CTAs and always eight requests per thread. int32 at 4
    // GPUs, 1 GiB: 32 x 256 x 4 -> 198 GB/s, 32 x 256 x 8 -> 342, 128 x 256 x 8 -> 558 (= the fp32 form); at 8 GPUs each rank
    // owns half as much of the message and 32 x 256 x 8 was already enough up to 256 MiB, not beyond (profiles/raw/r2_4gpu, r2_8gpu_b).
    if (unit_bytes < 16) {
      cap = std::min(128, 32 * 16 / std::max(4, unit_bytes));
      if (bytes >= (1u << 20)) s.unroll = 8;
    }
  } else if (kind == kCollPush || kind == kCollRing) {
    s.threads = 512;
    s.unroll = 4;
    cap = 128;
    per_cta = 65536;
  } else {
    s.threads = 512;
    // four vectors in flight per thread: with eight the source-double-buffered loops spill (one-shot int32 at 2 GPUs: 417 GB/s
    // with 8, 619 with 4)
    (void)bytes;
    s.unroll = 4;  // two-shot reduce at 8 GPUs: 62.7 / 208 / 685 us with four against 71.6 / 301 / 760 with eight (16 / 64 / 256 MiB)
    cap = 128;
    per_cta = 65536;
  }
  const int et = B2R_ENV_INT("B2R_COLL_THREADS", 0), eu = B2R_ENV_INT("B2R_COLL_UNROLL", 0), ec = B2R_ENV_INT("B2R_COLL_CTAS", 0);
  if (et >= 64 && et <= 512) s.threads = et - et % 32;
  if (eu == 4 || eu == 8) s.unroll = eu;
  if (ec > 0) cap = ec;
  if (c->max_ctas > 0) cap = std::min(cap, c->max_ctas);
  cap = std::max(1, std::min(cap, kCollMaxCtas));
  if (nvirtual > 1) cap = std::max(1, std::min(cap, c->sm_count / nvirtual));  // cooperative launch: all ranks co-resident
  s.ctas = (int)std::max<uint64_t>(1, std::min<uint64_t>((bytes + per_cta - 1) / per_cta, (uint64_t)cap));
  return s;
}

bool kernel_exists(b2r_dtype_t dtype, int op, int kind) { return pick_kernel(dtype, op, kind, 4) != nullptr; }

bool nvls_supported(b2r_dtype_t dtype, b2r_op_t op) {
  if (dtype == B2R_FLOAT32 || dtype == B2R_FLOAT64) return op == B2R_SUM;
  return true;  // bf16x2 add/min/max, 32-bit integer add/min/max
}
// fp32 / fp64 min and max go through the switch as integer keys: the send buffer is first staged (encoded) into the work region
bool nvls_keyed(b2r_dtype_t dtype, b2r_op_t op) { return (dtype == B2R_FLOAT32 || dtype == B2R_FLOAT64) && op != B2R_SUM; }

// Largest message the LL path can carry on this communicator: the small instance in the control region, or the large one
// behind the user heap when the communicator was created with B2R_LL_BIG_MB.
uint64_t ll_limit(const b2r_comm* c) {
  const uint64_t small = ll_max_msg_bytes(c->nranks);
  const uint64_t big = c->ll_big_bytes ? ll_big_max_msg_bytes(c->ll_big_bytes, c->nranks) : 0;
  return std::max(small, big);
}

// Posted reduce: one staging slot per call parity at the start of the work region; a message must fit one slot.
// The two slots take at most the first half of the work region; what lies behind them is the staging area of the keyed NVLS
// form. (Everything else that uses the work region — tree, push, ring — starts with a cross-GPU barrier, so it cannot overwrite a
// slot that a lagging root is still reading; the keyed form's encoding pass has no barrier in front of it, hence its own area.)
uint64_t posted_slot_bytes(const b2r_comm* c) { return std::min<uint64_t>(c->work_bytes / 4, kPostedMaxSlotBytes) & ~(uint64_t)255; }
uint64_t keyed_off(const b2r_comm* c) { return 2 * posted_slot_bytes(c); }
uint64_t keyed_bytes(const b2r_comm* c) { return (c->work_bytes - keyed_off(c)) & ~(uint64_t)255; }
// The grid of the posted path is a property of the communicator (chunk ownership depends on it): kPostedCtas, fewer if the
// communicator caps CTAs; a loopback group runs all its virtual ranks in one cooperative launch, one CTA per SM in total.
int posted_ctas(const b2r_comm* c) {
  int g = c->max_ctas > 0 ? std::min(c->max_ctas, kPostedCtas) : kPostedCtas;
  if (c->mode == kModeLoopback && c->nranks > 1) g = std::min(g, c->sm_count / c->nranks);
  return std::max(1, g);
}

// Which LL instance carries a message: the one-CTA instance in the control region up to kLLSmallMaxBytes (or whenever there is
// no large region), the 64-CTA instance behind the user heap beyond.
bool ll_use_small(const b2r_comm* c, uint64_t bytes) {
  if (!c->ll_big_bytes || bytes > ll_big_max_msg_bytes(c->ll_big_bytes, c->nranks)) return true;
  return bytes <= kLLSmallMaxBytes && bytes <= ll_max_msg_bytes(c->nranks);
}

// What AUTO resolves to, on cached environment values (the public b2r_select_algo below refreshes them first). The crossovers
// were read off one 8-GPU measurement of every candidate against NCCL (profiles/collectives_8gpu.md, raw rows in
// profiles/raw/r2_8gpu/) and a 2-GPU one (profiles/multi_gpu.md):
//   * <= 8 KiB: the LL path (flag-in-data packets, no barrier at either end) — reduce 8.3 us where the barrier-framed kernels
//     need 14 (NVLS) to 21 (one-shot) and NCCL 8.2; allreduce 10.3 us against NCCL's 25. Beyond 8 KiB its single CTA is the limit.
//   * from 4 GPUs up, every (dtype, op) with a multimem form goes through the switch (NVLS) at every larger size — including
//     int32, whose 4-byte multimem ops were assumed too slow in round 1 and measure the same as the 16-byte fp32 form
//     (allreduce 256 MiB: 576 us either way, two-shot 891, NCCL 727).
//   * what has no multimem form (fp32 / fp64 min and max) stays on peer loads: one-shot while the whole message is small,
//     two-shot beyond; and on stores (push) from the crossovers below up.
//   * 2 GPUs: reduce = the root pulls the peer's message (one-shot, 630 GB/s at 64 MiB vs NCCL 580); allreduce = two-shot.
// measured crossovers (MiB) of the push algorithm, see above; 0 = AUTO never picks it
constexpr int kAutoPushReduceMB = 0, kAutoPushAllreduceMB = 0, kAutoPushReduceNvlsMB = 0;
// LL: 4 GPUs, 64-CTA instance: reduce 8.2-10.5 us up to 256 KiB (NCCL 7.8-11.0, NVLS 12.5-15.7); allreduce ahead of NVLS up to
// 128 KiB at 4 GPUs — every consumer receives (P-1) x 2 x the message, so the crossover is set lower for all sizes of P
constexpr int kAutoLLAllreduceBytes = 65536, kAutoLLReduceBytes = 262144;
// posted reduce (stage locally, return; the root alone reduces the staged copies through the switch): from the LL crossover up
// to these sizes. Measured on 8 GPUs (profiles/raw/r2_posted/): one GPU's multimem request stream sustains ~290 GB/s whatever the
// form (once the 4- and 8-byte requests are coalesced across the warp), and a call costs 12.9 us up to 1 MiB. Against the
// barrier-framed NVLS kernel (native forms) it is ahead up to 8 MiB (1 MiB: 13 vs 18 us, 4 MiB: 22 vs 25-26, 8 MiB: 34-36 vs
// 35-36); against two-shot by peer loads (fp32 / fp64 min and max, which travel as 32- / 64-bit integer keys) up to 32 MiB
// (fp32: 16 MiB 60 vs 63 us) and 8 MiB (fp64: 38 vs 46; 16 MiB 68 vs 64).
constexpr int kAutoNvlsKeys = 0;
constexpr int kAutoPosted2GpuMaxBytes = 2 << 20;
constexpr int kAutoPostedMaxBytes = 8 << 20, kAutoPostedKey4MaxBytes = 32 << 20, kAutoPostedKey8MaxBytes = 8 << 20;

b2r_algo_t select_algo_cached(int is_allreduce, uint64_t bytes, b2r_dtype_t dtype, b2r_op_t op, int nranks, int has_multicast, uint64_t ll_max,
                              uint64_t posted_max) {
  if (nranks <= 1) return B2R_ALGO_ONESHOT;
  const int forced = B2R_ENV_INT("B2R_ALGO", 0);
  // a forced LL only applies to messages that fit its slots; everything else falls through to the normal choice
  if (forced == B2R_ALGO_LL && nranks <= B2R_MAX_RANKS && bytes <= ll_max) return B2R_ALGO_LL;
  if (forced >= B2R_ALGO_ONESHOT && forced <= B2R_ALGO_PUSH) return (b2r_algo_t)forced;
  if (forced == B2R_ALGO_RING && !is_allreduce) return B2R_ALGO_RING;
  if (forced == B2R_ALGO_POSTED && !is_allreduce && bytes <= posted_max) return B2R_ALGO_POSTED;
  // LL crossovers: allreduce B2R_AUTO_LL_MAX_BYTES, reduce-to-root B2R_AUTO_LL_REDUCE_MAX_BYTES (only the root consumes, so the
  // 64-CTA instance stays ahead of the barrier-framed kernels for longer); both capped by what the communicator's slots hold
  const uint64_t auto_ll = (uint64_t)std::max(0, is_allreduce ? B2R_ENV_INT("B2R_AUTO_LL_MAX_BYTES", kAutoLLAllreduceBytes)
                                                             : B2R_ENV_INT("B2R_AUTO_LL_REDUCE_MAX_BYTES", kAutoLLReduceBytes));
  // two GPUs: a consumer receives from ONE peer, so LL stays ahead of the barrier-framed one-shot (16.6 us at 256 KiB) for as long as its slots reach
  const uint64_t ll_cut = (is_allreduce && nranks == 2 && auto_ll) ? std::max<uint64_t>(auto_ll, 262144) : auto_ll;
  if (ll_cut && nranks <= B2R_MAX_RANKS && bytes <= std::min<uint64_t>(ll_cut, ll_max)) return B2R_ALGO_LL;
  const uint64_t auto_ring = (uint64_t)std::max(0, B2R_ENV_INT("B2R_AUTO_RING_MIN_MB", 0));
  if (auto_ring && !is_allreduce && nranks > 2 && bytes >= (auto_ring << 20)) return B2R_ALGO_RING;
  const bool switch_ok = has_multicast && nranks >= 4 && !B2R_ENV_INT("B2R_DISABLE_NVLS", 0);
  const bool nvls_ok = switch_ok && nvls_supported(dtype, op);
  // fp32 / fp64 min and max through the switch as integer keys (one extra local pass to encode): B2R_AUTO_NVLS_KEYS=0 keeps them on peer loads
  const bool keyed_ok = switch_ok && nvls_keyed(dtype, op) && B2R_ENV_INT("B2R_AUTO_NVLS_KEYS", kAutoNvlsKeys);
  // push (stores only, two hops): from how many MiB on it replaces the pulls / the switch; 0 = never. Three crossovers:
  //   B2R_AUTO_PUSH_REDUCE_MB        reduce-to-root of a (dtype, op) without a multimem form (fp32 / fp64 min and max)
  //   B2R_AUTO_PUSH_ALLREDUCE_MB     allreduce of the same
  //   B2R_AUTO_PUSH_REDUCE_NVLS_MB   reduce-to-root of what NVLS serves (the switch's ~520 GB/s ceiling loses to NCCL's ring)
  if (nranks > 2) {
    const uint64_t mb = (uint64_t)std::max(0, nvls_ok ? (is_allreduce ? 0 : B2R_ENV_INT("B2R_AUTO_PUSH_REDUCE_NVLS_MB", kAutoPushReduceNvlsMB))
                                                      : (is_allreduce ? B2R_ENV_INT("B2R_AUTO_PUSH_ALLREDUCE_MB", kAutoPushAllreduceMB)
                                                                      : B2R_ENV_INT("B2R_AUTO_PUSH_REDUCE_MB", kAutoPushReduceMB)));
    if (mb && bytes >= (mb << 20)) return B2R_ALGO_PUSH;
  }
  // posted: mid-size reduce-to-root wherever the switch can do the reading (natively, or on integer keys for fp min / max)
  if (switch_ok && !is_allreduce) {
    // B2R_AUTO_POSTED_MAX_BYTES: pairs with a multimem form; B2R_AUTO_POSTED_KEY_MAX_BYTES_4 / _8: fp32 / fp64 min and max; 0 = never
    const int cut_i = nvls_ok ? B2R_ENV_INT("B2R_AUTO_POSTED_MAX_BYTES", kAutoPostedMaxBytes)
                              : (dtype == B2R_FLOAT64 ? B2R_ENV_INT("B2R_AUTO_POSTED_KEY_MAX_BYTES_8", kAutoPostedKey8MaxBytes)
                                                      : B2R_ENV_INT("B2R_AUTO_POSTED_KEY_MAX_BYTES_4", kAutoPostedKey4MaxBytes));
    const uint64_t cut = (uint64_t)std::max(0, cut_i);
    if (bytes <= std::min(cut, posted_max)) return B2R_ALGO_POSTED;
  }
  // 32-bit integers through the switch collapse for reduce-to-root at >= 1 GiB on 8 GPUs (2.1-4.8 ms at 1 GiB where fp32 needs 2.0 and
  // 256 MiB is on par: 4-byte multimem loads AND 4-byte stores to the root); the store-forwarding ring is steady there (2.04 ms)
  if (nvls_ok && !is_allreduce && dtype == B2R_INT32 && bytes >= (512ull << 20)) return B2R_ALGO_RING;
  if (nvls_ok || keyed_ok) return B2R_ALGO_NVLS;
  if (!is_allreduce && nranks == 2) {
    // Two GPUs: the root pulls (one-shot: 630 GB/s at 64 MiB, parity with NCCL from 4 MiB), but between the LL range and ~2 MiB its two
    // barriers make it 18.5-19.3 us at 1 MiB against NCCL's 10.8-12.3 (profiles/collectives_2_4gpu.md). The posted reduce has no
    // barrier: 11.4-12.9 us per call at 1 MiB on 4 and 8 GPUs whatever the type. Verified on 2 GPUs (tests/dist_worker.py), its timing
    // there is inferred from those floors, not measured: B2R_AUTO_POSTED_2GPU_MAX_BYTES=0 restores the one-shot.
    const uint64_t cut2 = (uint64_t)std::max(0, B2R_ENV_INT("B2R_AUTO_POSTED_2GPU_MAX_BYTES", kAutoPosted2GpuMaxBytes));
    if (has_multicast && !B2R_ENV_INT("B2R_DISABLE_NVLS", 0) && bytes <= std::min(cut2, posted_max)) return B2R_ALGO_POSTED;
    return B2R_ALGO_ONESHOT;
  }
  // One-shot reads the WHOLE message from every peer: fine for two GPUs or tiny messages, hopeless beyond.
  const uint64_t oneshot_max = (uint64_t)std::max(0, B2R_ENV_INT("B2R_ONESHOT_MAX_KB", nranks <= 2 ? 512 : 16)) << 10;
  if (bytes <= oneshot_max) return B2R_ALGO_ONESHOT;
  return B2R_ALGO_TWOSHOT;
}

// B2R_WEAK_DATA=1: switch the payload accesses of the collective kernels to weak loads/stores (see coll_kernels.cuh). The
// constant lives per device; it is written once per device before the first collective launched there.
b2r_status_t apply_data_flavour(int device) {
  static std::mutex mu;
  static bool done[64] = {false};
  const int weak = B2R_ENV_INT("B2R_WEAK_DATA", 0) ? 1 : 0;
  if (!weak || device < 0 || device >= 64) return B2R_OK;
  std::lock_guard<std::mutex> lk(mu);
  if (done[device]) return B2R_OK;
  B2R_CUDA_CHECK(cudaMemcpyToSymbol(c_weak_data, &weak, sizeof(int)));
  done[device] = true;
  return B2R_OK;
}

struct RankCall {
  b2r_comm* c;
  const void* send;
  void* recv;
  cudaStream_t stream;
};

// Runs one element-wise collective for the `n` calls in rc (n == 1: this process drives one rank; n == P: loopback).
b2r_status_t run_collective(RankCall* rc, int n, int kind_call, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root, b2r_algo_t algo) {
  b2r_comm* c0 = rc[0].c;
  const int P = c0->nranks;
  const size_t esz = b2r_dtype_size(dtype);
  const bool is_all = kind_call == kCallAllreduce;
  NvtxRange nvtx(is_all ? "b2r_allreduce" : "b2r_reduce");
  B2R_REQUIRE(dtype >= B2R_INT32 && dtype <= B2R_BFLOAT16, "collective: bad dtype %d", (int)dtype);
  B2R_REQUIRE(op >= B2R_SUM && op <= B2R_MAX, "collective: bad op %d", (int)op);
  B2R_REQUIRE(is_all || (root >= 0 && root < P), "reduce: root %d out of range [0,%d)", root, P);
  for (int i = 0; i < n; ++i) {
    b2r_comm* c = rc[i].c;
    const bool gets_result = is_all || c->rank == root;
    B2R_REQUIRE(rc[i].send != nullptr || count == 0, "collective: sendbuf is NULL on rank %d", c->rank);
    B2R_REQUIRE(!gets_result || rc[i].recv != nullptr || count == 0, "collective: recvbuf is NULL on rank %d", c->rank);
    B2R_REQUIRE(((uintptr_t)rc[i].send % esz) == 0 && ((uintptr_t)rc[i].recv % esz) == 0, "collective: buffer not aligned to the element size");
  }
  if (count == 0) return B2R_OK;

  int prev_dev = -1;
  B2R_CUDA_CHECK(cudaGetDevice(&prev_dev));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev_dev};
  B2R_CUDA_CHECK(cudaSetDevice(c0->device));

  if (P == 1) {
    if (rc[0].recv != rc[0].send) B2R_CUDA_CHECK(cudaMemcpyAsync(rc[0].recv, rc[0].send, count * esz, cudaMemcpyDeviceToDevice, rc[0].stream));
    return B2R_OK;
  }
  B2R_CHECK(apply_data_flavour(c0->device));

  const uint64_t total_bytes = count * esz;
  if (algo == B2R_ALGO_AUTO) algo = select_algo_cached(is_all ? 1 : 0, total_bytes, dtype, op, P, c0->has_mc ? 1 : 0, ll_limit(c0), posted_slot_bytes(c0));
  if (algo == B2R_ALGO_NVLS) {
    if (!c0->has_mc) return fail(B2R_ERR_UNSUPPORTED, "NVLS requested but this communicator has no multicast mapping");
    if (!nvls_supported(dtype, op) && !nvls_keyed(dtype, op))
      return fail(B2R_ERR_UNSUPPORTED, "NVLS has no %s %s instruction", b2r_dtype_name(dtype), b2r_op_name(op));
  }
  int kind = algo == B2R_ALGO_ONESHOT ? kCollOneshot : algo == B2R_ALGO_TWOSHOT ? kCollTwoshot : algo == B2R_ALGO_TREE ? kCollTree
             : algo == B2R_ALGO_PUSH ? kCollPush : algo == B2R_ALGO_RING ? kCollRing : algo == B2R_ALGO_LL ? kCollLL
             : algo == B2R_ALGO_POSTED ? kCollPosted : kCollNvls;
  if (kind == kCollRing && is_all) return fail(B2R_ERR_UNSUPPORTED, "the ring algorithm implements reduce-to-root only");
  if (kind == kCollPosted) {
    if (is_all) return fail(B2R_ERR_UNSUPPORTED, "the posted algorithm implements reduce-to-root only");
    if (((total_bytes + 15) & ~15ull) > posted_slot_bytes(c0))
      return fail(B2R_ERR_INVALID, "the posted path carries at most %llu bytes per rank on this communicator (got %llu)",
                  (unsigned long long)posted_slot_bytes(c0), (unsigned long long)total_bytes);
  }
  if (kind == kCollLL) {
    // LL never touches a peer's send/recv buffer: any device pointer works, nothing is chunked, and only a buffer that is not
    // 16-byte aligned is bounced through this rank's own staging (a purely local decision: ranks need not agree on it)
    if (total_bytes > ll_limit(c0))
      return fail(B2R_ERR_INVALID, "the LL path carries at most %llu bytes per rank on this communicator of %d ranks (got %llu)",
                  (unsigned long long)ll_limit(c0), P, (unsigned long long)total_bytes);
  }
  B2R_REQUIRE(kernel_exists(dtype, (int)op, kind), "collective: no kernel for dtype %d op %d algo %d", (int)dtype, (int)op, (int)algo);

  // Chunking rule (identical on every rank): messages larger than one staging region are cut into stage-sized chunks,
  // unless the communicator was created with symmetric_only (then nothing is ever staged and one launch does it all).
  const uint64_t stage_elems = (c0->stage_bytes / 16) * 16 / esz;
  const uint64_t chunk_elems = c0->symmetric_only ? count : std::min<uint64_t>(count, stage_elems);
  // the tree keeps its running partial in the work region, and push stages P whole slices there: both are bounded per launch
  const uint64_t work_elems = (c0->work_bytes / 16) * 16 / esz;
  uint64_t chunk = chunk_elems;
  if (kind == kCollTree) chunk = std::min<uint64_t>(chunk_elems, work_elems);
  const bool keyed = kind == kCollNvls && nvls_keyed(dtype, op);  // the encoded copy of a piece lives in the work region
  if (keyed) chunk = std::min<uint64_t>(chunk_elems, keyed_bytes(c0) / esz);
  if (kind == kCollPosted) chunk = count;  // like LL: peers' buffers are never touched, nothing is staged per chunk
  if (kind == kCollPush) {
    const uint64_t W = is_all ? P : P - 1;
    const uint64_t slice_bytes = std::min<uint64_t>(c0->work_bytes / P - 64, (uint64_t)kPushChunkCap * 512 * 4 * 16 - 64);
    chunk = std::min<uint64_t>(chunk_elems, (slice_bytes / 16) * 16 * W / esz);
  }

  for (uint64_t first = 0; first < count; first += chunk) {
    const uint64_t cnt = std::min<uint64_t>(chunk, count - first);
    const uint64_t bytes = cnt * esz;
    CollLaunch L;
    memset(&L, 0, sizeof(L));
    std::vector<bool> recv_staged(n, false);
    for (int i = 0; i < n; ++i) {
      b2r_comm* c = rc[i].c;
      const bool gets_result = is_all || c->rank == root;
      const char* send = (const char*)rc[i].send + first * esz;
      char* recv = rc[i].recv ? (char*)rc[i].recv + first * esz : nullptr;
      if (kind == kCollLL || kind == kCollPosted) {
        CollArgs& A = L.a[n == 1 ? 0 : c->rank];
        A.ctx = c->ctx;
        if (((uintptr_t)send % 16) || (gets_result && ((uintptr_t)recv % 16)))
          B2R_REQUIRE(bytes <= c->stage_bytes, "rank %d: a buffer that is not 16-byte aligned is bounced through staging (%zu bytes), got %llu",
                      c->rank, c->stage_bytes, (unsigned long long)bytes);
        if ((uintptr_t)send % 16) {
          B2R_CUDA_CHECK(cudaMemcpyAsync(c->stage_send(), send, bytes, cudaMemcpyDeviceToDevice, rc[i].stream));
          send = c->stage_send();
        }
        if (gets_result && ((uintptr_t)recv % 16)) {
          recv = c->stage_recv();
          recv_staged[i] = true;
        }
        A.send_ptr = send;
        A.recv_ptr = gets_result ? recv : nullptr;
        A.count = cnt;
        A.epoch = c->epoch;
        A.root = is_all ? -1 : root;
        if (kind == kCollPosted) {
          A.seq = ++c->posted_seq;
          A.work_off = (unsigned long long)(c->stage_work() - c->my_heap());
          A.recv_off = posted_slot_bytes(c);                 // slot stride (the recv buffer travels as recv_ptr)
          A.ring_depth = c->posted_root[A.seq & 1u];         // whose acknowledgement frees this parity's slot
          c->posted_root[A.seq & 1u] = root;
          continue;
        }
        // small instance (control region, one CTA) when the message fits it, else the large one; each has its own slots,
        // acknowledgement words and consecutive call numbering (the slot parity and the n-2 rule depend on it)
        LLLayout lay;
        if (ll_use_small(c, bytes)) {
          lay.region_off = kOffLL;
          lay.ack_off = kOffLLAck;
          lay.slot_bytes = (unsigned int)ll_slot_bytes(P);
          lay.vec_per_cta = (unsigned int)(kLLThreads * kLLVecPerThread);
          A.seq = ++c->ll_seq;
        } else {
          lay.slot_bytes = (unsigned int)ll_big_slot_bytes(c->ll_big_bytes, P);
          lay.region_off = c->ll_big_off;
          lay.ack_off = c->ll_big_off + 2ull * P * lay.slot_bytes;
          lay.vec_per_cta = (unsigned int)kLLBigVecPerCta;
          A.seq = ++c->ll_big_seq;
        }
        ll_set_layout(A, lay);
        continue;
      }
      char* const key_stage = c->stage_work() + keyed_off(c);
      int64_t soff = keyed ? (int64_t)(key_stage - c->my_heap()) : c->heap_offset(send, bytes);
      if (keyed) {  // any readable device pointer will do: the switch reads the encoded copy, never the send buffer itself
        B2R_REQUIRE(n == 1, "keyed NVLS needs a multicast mapping, which loopback groups do not have");
        const uint64_t nunits = bytes / esz;
        const int ethreads = 512;
        const int ectas = (int)std::max<uint64_t>(1, std::min<uint64_t>((nunits * esz / 16 + ethreads * 4 - 1) / (ethreads * 4), (uint64_t)c->sm_count * 4));
        if (esz == 4) encode_keys_kernel<4><<<ectas, ethreads, 0, rc[i].stream>>>(send, key_stage, nunits);
        else encode_keys_kernel<8><<<ectas, ethreads, 0, rc[i].stream>>>(send, key_stage, nunits);
        B2R_CUDA_CHECK(cudaGetLastError());
      }
      int64_t roff = (gets_result && recv) ? c->heap_offset(recv, bytes) : -1;
      // in place is only safe zero-copy for two-shot / NVLS (a slice is read and written by its single owner)
      const bool inplace_hazard = (send == recv) && (kind == kCollOneshot || kind == kCollTree);
      if (soff < 0 || inplace_hazard) {
        if (c->symmetric_only) return fail(B2R_ERR_INVALID, "rank %d: sendbuf is not a 16-byte aligned symmetric allocation (symmetric_only)", c->rank);
        B2R_CUDA_CHECK(cudaMemcpyAsync(c->stage_send(), send, bytes, cudaMemcpyDeviceToDevice, rc[i].stream));
        soff = c->stage_send() - c->my_heap();
      }
      if (roff < 0) {
        if (gets_result && c->symmetric_only) return fail(B2R_ERR_INVALID, "rank %d: recvbuf is not a 16-byte aligned symmetric allocation (symmetric_only)", c->rank);
        roff = c->stage_recv() - c->my_heap();
        recv_staged[i] = gets_result;
      }
      CollArgs& A = L.a[n == 1 ? 0 : c->rank];
      A.ctx = c->ctx;
      A.send_off = (unsigned long long)soff;
      A.recv_off = (unsigned long long)roff;
      A.work_off = (unsigned long long)(c->stage_work() - c->my_heap());
      A.count = cnt;
      A.epoch = c->epoch;
      A.root = is_all ? -1 : root;
      c->epoch += 8;  // start + end + up to 4 tree rounds, rounded up; same increment on every rank
      if (kind == kCollPush || kind == kCollRing) A.seq = ++c->push_seq;
    }
    const Shape shape = coll_shape(c0, kind, bytes, n, (dtype == B2R_INT32 || (keyed && dtype == B2R_FLOAT32)) ? 4 : (dtype == B2R_FLOAT64 ? 8 : 16));
    int ctas = shape.ctas;
    int threads = shape.threads;
    if (kind == kCollLL) {
      if (ll_use_small(c0, bytes)) {  // one CTA; just enough threads for one vector each, up to kLLThreads x kLLVecPerThread vectors
        const uint64_t nvec = (bytes + 15) / 16;
        ctas = 1;
        threads = nvec <= 128 ? 128 : nvec <= 256 ? 256 : kLLThreads;
      } else {  // the large instance: always the same grid, whatever the length (fixed partition, see ll_body)
        ctas = kLLBigCtas;
        // loopback: all virtual ranks' CTAs must be co-resident in ONE cooperative launch (8 x 64 CTAs): half-size CTAs, two
        // vectors per thread, same partition
        threads = n > 1 ? kLLBigThreads / 2 : kLLBigThreads;
      }
    }
    if (kind == kCollPosted) {  // the same grid in every call (fixed partition, see coll_posted.cuh); loopback: half-size CTAs
      ctas = posted_ctas(c0);
      threads = n > 1 ? kPostedThreads / 2 : kPostedThreads;
    }
    if (kind == kCollRing) {
      // staging ring in the work region: depth x ctas steps of sub_blocks x (threads x 8) vectors; at least 2 deep.
      // Large messages use big steps (default 8 sub-blocks = 512 KiB with 512 threads) to amortise the per-step fence.
      const uint64_t sub_bytes = (uint64_t)threads * 8 * 16;
      int subs = std::max(1, B2R_ENV_INT("B2R_RING_SUBBLOCKS", 8));
      int cap = std::max(1, B2R_ENV_INT("B2R_RING_CTAS", 32));
      if (n > 1) cap = std::max(1, std::min(cap, c0->sm_count / n));
      while (subs > 1 && bytes < (uint64_t)cap * sub_bytes * subs * 2) subs /= 2;  // small message: keep all CTAs busy
      const uint64_t step_bytes = sub_bytes * subs;
      ctas = (int)std::max<uint64_t>(1, std::min<uint64_t>((bytes + step_bytes - 1) / step_bytes, (uint64_t)cap));
      ctas = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)ctas, c0->work_bytes / (2ull * step_bytes)));
      const int depth = (int)std::min<uint64_t>(8, c0->work_bytes / ((uint64_t)ctas * step_bytes));
      for (int i = 0; i < n; ++i) {
        L.a[n == 1 ? 0 : rc[i].c->rank].ring_depth = depth;
        L.a[n == 1 ? 0 : rc[i].c->rank].sub_blocks = subs;
        L.a[n == 1 ? 0 : rc[i].c->rank].stagger_ns = (unsigned int)std::max(0, B2R_ENV_INT("B2R_RING_STAGGER_US", 0)) * 1000u;
      }
    }
    if (kind == kCollNvls && !is_all) {
      // reduce: share of the message (in 64ths) that the root pulls through the switch itself; 0 = equal slices for all ranks
      const int share = std::max(0, std::min(64, B2R_ENV_INT("B2R_NVLS_ROOT_SHARE", 0)));
      for (int i = 0; i < n; ++i) L.a[n == 1 ? 0 : rc[i].c->rank].sub_blocks = share;
    }
    if (kind == kCollPush) {
      // whole-slice staging in the work region: P sources x slice_cap vectors (coll_push.cuh). The chunk = threads x 4 vectors:
      // small slices take small chunks so that enough CTAs get work; the flag table names at most kPushChunkCap chunks per slice.
      const int W = is_all ? P : P - 1;
      const uint64_t nvec = bytes / 16, slice_cap = (nvec + W - 1) / W + 1;
      threads = 512;
      while (threads > 128 && slice_cap / ((uint64_t)threads * 4) < 96) threads /= 2;
      const int et = B2R_ENV_INT("B2R_COLL_THREADS", 0);
      if (et >= 64 && et <= 512) threads = et - et % 32;
      while (threads < 512 && (slice_cap + (uint64_t)threads * 4 - 1) / ((uint64_t)threads * 4) > kPushChunkCap) threads *= 2;
      const uint64_t nchunks = (slice_cap + (uint64_t)threads * 4 - 1) / ((uint64_t)threads * 4);
      B2R_REQUIRE(nchunks <= kPushChunkCap && (uint64_t)P * slice_cap * 16 <= c0->work_bytes,
                  "push: a piece of %llu bytes does not fit the work region (%zu bytes) / flag table", (unsigned long long)bytes, c0->work_bytes);
      int cap = std::max(1, B2R_ENV_INT("B2R_PUSH_CTAS", 128));
      if (c0->max_ctas > 0) cap = std::min(cap, c0->max_ctas);
      cap = std::min(cap, kCollMaxCtas);
      if (n > 1) cap = std::max(1, std::min(cap, c0->sm_count / n));
      ctas = (int)std::max<uint64_t>(1, std::min<uint64_t>(nchunks, (uint64_t)cap));
    }
    CollKernel kernel = pick_kernel(dtype, (int)op, kind, shape.unroll);
    if (n == 1) {
      kernel<<<dim3((unsigned)ctas, 1), threads, 0, rc[0].stream>>>(L);
      B2R_CUDA_CHECK(cudaGetLastError());
    } else {
      void* args[] = {(void*)&L};
      B2R_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)kernel, dim3((unsigned)ctas, (unsigned)n), dim3((unsigned)threads), args, 0, rc[0].stream));
    }
    for (int i = 0; i < n; ++i) {
      if (!recv_staged[i]) continue;
      b2r_comm* c = rc[i].c;
      B2R_CUDA_CHECK(cudaMemcpyAsync((char*)rc[i].recv + first * esz, c->stage_recv(), bytes, cudaMemcpyDeviceToDevice, n == 1 ? rc[i].stream : rc[0].stream));
    }
  }
  return B2R_OK;
}

b2r_status_t run_barrier(RankCall* rc, int n) {
  b2r_comm* c0 = rc[0].c;
  if (c0->nranks == 1) return B2R_OK;
  int prev_dev = -1;
  B2R_CUDA_CHECK(cudaGetDevice(&prev_dev));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev_dev};
  B2R_CUDA_CHECK(cudaSetDevice(c0->device));
  CollLaunch L;
  memset(&L, 0, sizeof(L));
  for (int i = 0; i < n; ++i) {
    b2r_comm* c = rc[i].c;
    CollArgs& A = L.a[n == 1 ? 0 : c->rank];
    A.ctx = c->ctx;
    A.epoch = c->epoch;
    A.root = -1;
    c->epoch += 8;
  }
  CollKernel kernel = coll_kernel<B2R_INT32, B2R_SUM, kCollBarrier, 4>;
  if (n == 1) {
    kernel<<<dim3(1, 1), 32, 0, rc[0].stream>>>(L);
    B2R_CUDA_CHECK(cudaGetLastError());
  } else {
    void* args[] = {(void*)&L};
    B2R_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)kernel, dim3(1, (unsigned)n), dim3(32), args, 0, rc[0].stream));
  }
  return B2R_OK;
}

// Scalar-sharded reduce: the local streaming reduction with the cross-GPU mailbox combine fused into its tail.
b2r_status_t run_scalar(RankCall* rc, int n, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root, int all, const b2r_launch_cfg_t* ucfg) {
  b2r_comm* c0 = rc[0].c;
  const int P = c0->nranks;
  B2R_REQUIRE(root >= 0 && root < P, "reduce_scalar_sharded: root %d out of range [0,%d)", root, P);
  for (int i = 0; i < n; ++i) {
    const bool gets = all || rc[i].c->rank == root;
    B2R_REQUIRE(!gets || rc[i].recv != nullptr, "reduce_scalar_sharded: out is NULL on rank %d", rc[i].c->rank);
  }
  if (n == 1) {
    b2r_comm* c = rc[0].c;
    XgCall xg;
    xg.ctx = (const CollCtx*)(c->my_heap() + kOffCtx);
    xg.root = root;
    xg.all = all;
    xg.seq = ++c->xg_seq;
    void* out = rc[0].recv ? rc[0].recv : (void*)c->ws->d_partials;  // non-root without `all`: never written
    if (P == 1) return reduce_local_impl(rc[0].send, out, count, dtype, op, c->ws, rc[0].stream, ucfg, 0, nullptr);
    return reduce_local_impl(rc[0].send, out, count, dtype, op, c->ws, rc[0].stream, ucfg, 0, &xg);
  }
  // loopback: every virtual rank in one cooperative launch (LDG128 variant, same grid for all ranks)
  int prev_dev = -1;
  B2R_CUDA_CHECK(cudaGetDevice(&prev_dev));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev_dev};
  B2R_CUDA_CHECK(cudaSetDevice(c0->device));
  MultiLocalParams mp;
  memset(&mp, 0, sizeof(mp));
  const int threads = 256;
  int ctas = 1;
  {
    const uint64_t bytes = count * b2r_dtype_size(dtype);
    const uint64_t want = (bytes + 32767) / 32768;
    ctas = (int)std::max<uint64_t>(1, std::min<uint64_t>(want, (uint64_t)std::max(1, c0->sm_count * 2 / n)));
  }
  for (int i = 0; i < n; ++i) {
    b2r_comm* c = rc[i].c;
    b2r_launch_cfg_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.variant = B2R_VARIANT_LDG128;
    cfg.threads = threads;
    cfg.ctas = ctas;
    cfg.cluster = 1;
    cfg.unroll = 4;
    LaunchGeom g;
    void* out = rc[i].recv ? rc[i].recv : (void*)c->ws->d_partials;
    B2R_CHECK(prepare_local(rc[i].send, out, count, dtype, op, c->ws, &cfg, 0, &g, &mp.p[c->rank]));
    LocalParams& p = mp.p[c->rank];
    p.xg_ctx = (const CollCtx*)(c->my_heap() + kOffCtx);
    p.xg_root = root;
    p.xg_all = all;
    p.xg_seq = ++c->xg_seq;
    p.pdl = 0;
  }
  return launch_multi_any(dtype, (int)op, n, ctas, threads, mp, rc[0].stream);
}

// Loopback: record the call; the last rank of the group to arrive launches for everybody.
b2r_status_t loopback_submit(b2r_comm* c, const PendingCall& call) {
  CommGroup* g = c->group.get();
  std::lock_guard<std::mutex> lk(g->mu);
  if (g->pending[c->rank]) return fail(B2R_ERR_INVALID, "loopback rank %d issued a second collective before the group completed the first", c->rank);
  g->pending[c->rank] = std::make_shared<PendingCall>(call);
  if (++g->npending < g->nranks) return B2R_ERR_PENDING;
  std::vector<RankCall> rc(g->nranks);
  const PendingCall& ref = *g->pending[0];
  b2r_status_t st = B2R_OK;
  for (int r = 0; r < g->nranks; ++r) {
    const PendingCall& pc = *g->pending[r];
    if (pc.kind != ref.kind || pc.count != ref.count || pc.dtype != ref.dtype || pc.op != ref.op || pc.root != ref.root || pc.all != ref.all ||
        pc.algo != ref.algo)
      st = fail(B2R_ERR_INVALID, "loopback: rank %d called with different arguments than rank 0", r);
    rc[r].c = g->members[r];
    rc[r].send = pc.send;
    rc[r].recv = pc.recv;
    rc[r].stream = call.stream;  // everything runs on the last caller's stream
  }
  if (st == B2R_OK) {
    if (ref.kind == kCallScalar) st = run_scalar(rc.data(), g->nranks, ref.count, ref.dtype, ref.op, ref.root, ref.all, nullptr);
    else if (ref.kind == kCallBarrier) st = run_barrier(rc.data(), g->nranks);
    else st = run_collective(rc.data(), g->nranks, ref.kind, ref.count, ref.dtype, ref.op, ref.root, ref.algo);
  }
  for (auto& p : g->pending) p.reset();
  g->npending = 0;
  return st;
}

b2r_status_t submit(b2r_comm* c, const PendingCall& call) {
  B2R_REQUIRE(c != nullptr, "collective: communicator is NULL");
  if (c->mode == kModeLoopback && c->nranks > 1) return loopback_submit(c, call);
  RankCall rc{c, call.send, call.recv, call.stream};
  if (call.kind == kCallScalar) return run_scalar(&rc, 1, call.count, call.dtype, call.op, call.root, call.all, call.has_cfg ? &call.cfg : nullptr);
  if (call.kind == kCallBarrier) return run_barrier(&rc, 1);
  return run_collective(&rc, 1, call.kind, call.count, call.dtype, call.op, call.root, call.algo);
}

}  // namespace
}  // namespace b2r

using namespace b2r;

extern "C" {

b2r_status_t b2r_comm_ll_max_bytes(b2r_comm_t comm, uint64_t* bytes) {
  B2R_REQUIRE(comm && bytes, "b2r_comm_ll_max_bytes: NULL argument");
  *bytes = comm->nranks >= 2 ? ll_limit(comm) : 0;
  return B2R_OK;
}

b2r_status_t b2r_comm_posted_max_bytes(b2r_comm_t comm, uint64_t* bytes) {
  B2R_REQUIRE(comm && bytes, "b2r_comm_posted_max_bytes: NULL argument");
  *bytes = comm->nranks >= 2 ? posted_slot_bytes(comm) : 0;
  return B2R_OK;
}

uint64_t b2r_ll_max_bytes(int nranks) { return (nranks >= 2 && nranks <= B2R_MAX_RANKS) ? (uint64_t)ll_max_msg_bytes(nranks) : 0; }

// Size thresholds measured on 8x B200 (profiles/collectives_8gpu.md). One-shot costs one NVLink round trip but every
// rank pulls the WHOLE message from every peer; two-shot / NVLS move 1/P of it per rank.
// Public, for tests and harnesses: looks at the environment as it is now (the launch path uses cached values, util/env.h).
b2r_algo_t b2r_select_algo(int is_allreduce, uint64_t bytes, b2r_dtype_t dtype, b2r_op_t op, int nranks, int has_multicast) {
  env_refresh();
  // the LL limit of a communicator created with the default large LL region (B2R_LL_BIG_MB unset = 16)
  const uint64_t ll_max = nranks <= B2R_MAX_RANKS
                              ? std::max<uint64_t>(ll_max_msg_bytes(std::max(nranks, 1)), ll_big_max_msg_bytes((size_t)16 << 20, std::max(nranks, 1)))
                              : 0;
  // and with the default work region (B2R_WORK_MB unset = 320): posted slots of 64 MiB
  return select_algo_cached(is_allreduce, bytes, dtype, op, nranks, has_multicast, ll_max, kPostedMaxSlotBytes);
}

b2r_status_t b2r_reduce_ex(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root, b2r_comm_t comm,
                           void* stream, b2r_algo_t algo) {
  PendingCall pc{kCallReduce, sendbuf, recvbuf, count, dtype, op, root, 0, algo, (cudaStream_t)stream};
  return submit(comm, pc);
}
b2r_status_t b2r_allreduce_ex(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, b2r_comm_t comm, void* stream,
                              b2r_algo_t algo) {
  PendingCall pc{kCallAllreduce, sendbuf, recvbuf, count, dtype, op, -1, 0, algo, (cudaStream_t)stream};
  return submit(comm, pc);
}
b2r_status_t b2r_reduce(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root, b2r_comm_t comm,
                        void* stream) {
  return b2r_reduce_ex(sendbuf, recvbuf, count, dtype, op, root, comm, stream, B2R_ALGO_AUTO);
}
b2r_status_t b2r_allreduce(const void* sendbuf, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, b2r_comm_t comm, void* stream) {
  return b2r_allreduce_ex(sendbuf, recvbuf, count, dtype, op, comm, stream, B2R_ALGO_AUTO);
}
b2r_status_t b2r_reduce_scalar_sharded_ex(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root, int all,
                                          b2r_comm_t comm, void* stream, const b2r_launch_cfg_t* cfg) {
  PendingCall pc{kCallScalar, in, out, count, dtype, op, root, all ? 1 : 0, B2R_ALGO_AUTO, (cudaStream_t)stream};
  pc.has_cfg = cfg != nullptr;
  if (cfg) pc.cfg = *cfg;
  return submit(comm, pc);
}
b2r_status_t b2r_reduce_scalar_sharded(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root, int all,
                                       b2r_comm_t comm, void* stream) {
  return b2r_reduce_scalar_sharded_ex(in, out, count, dtype, op, root, all, comm, stream, nullptr);
}
// End-to-end flavour of the sharded scalar reduce (what the vendored simpleMultiGPU does with cudaMemcpyAsync + a host loop,
// simpleMultiGPU.cpp:227-293): this rank's shard lives in (pinned) HOST memory. It is streamed to the GPU in chunks by the copy
// engine while earlier chunks are being reduced; the per-chunk scalars are then folded, and combined across the GPUs over
// NVLink, by ONE more launch of the fused kernel; the ranks that own the result read it back. Synchronous on `stream`.
b2r_status_t b2r_reduce_scalar_sharded_from_host(const void* host_in, void* out_host, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root,
                                                 int all, b2r_comm_t comm, void* stream_) {
  NvtxRange nvtx("b2r_reduce_scalar_sharded_from_host");
  B2R_REQUIRE(comm != nullptr, "b2r_reduce_scalar_sharded_from_host: communicator is NULL");
  B2R_REQUIRE(!(comm->mode == kModeLoopback && comm->nranks > 1), "b2r_reduce_scalar_sharded_from_host: not available on loopback groups");
  B2R_REQUIRE(root >= 0 && root < comm->nranks, "b2r_reduce_scalar_sharded_from_host: root %d out of range [0,%d)", root, comm->nranks);
  const bool gets = all || comm->rank == root;
  B2R_REQUIRE(!gets || out_host != nullptr, "b2r_reduce_scalar_sharded_from_host: out_host is NULL on a rank that receives the result");
  cudaStream_t stream = (cudaStream_t)stream_;
  b2r_workspace* ws = comm->ws;
  int prev = 0;
  B2R_CUDA_CHECK(cudaGetDevice(&prev));
  B2R_CUDA_CHECK(cudaSetDevice(comm->device));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  uint64_t nchunks = 0;
  B2R_CHECK(stream_chunks_from_host(host_in, count, dtype, op, ws, stream, &nchunks));
  // fold the chunk scalars (accumulator precision, chunk order) and combine across ranks in the same launch
  const b2r_dtype_t acc_dtype = dtype == B2R_BFLOAT16 ? B2R_FLOAT32 : dtype;
  const size_t acc_sz = dtype == B2R_FLOAT64 ? 8 : 4;
  char* d_final = (char*)ws->d_chunk_out + (size_t)kMaxChunks * 8;
  PendingCall pc{kCallScalar, ws->d_chunk_out, d_final, nchunks, acc_dtype, op, root, all ? 1 : 0, B2R_ALGO_AUTO, stream};
  pc.has_cfg = false;
  B2R_CHECK(submit(comm, pc));
  unsigned char acc[8] = {0};
  if (gets) B2R_CUDA_CHECK(cudaMemcpyAsync(ws->h_chunk_out, d_final, acc_sz, cudaMemcpyDeviceToHost, stream));
  B2R_CUDA_CHECK(cudaStreamSynchronize(stream));
  if (gets) {
    memcpy(acc, ws->h_chunk_out, acc_sz);
    acc_to_dtype(acc, dtype, out_host);
  }
  return B2R_OK;
}

b2r_status_t b2r_barrier(b2r_comm_t comm, void* stream) {
  PendingCall pc{kCallBarrier, nullptr, nullptr, 0, B2R_INT32, B2R_SUM, -1, 0, B2R_ALGO_AUTO, (cudaStream_t)stream};
  return submit(comm, pc);
}

}  // extern "C"
