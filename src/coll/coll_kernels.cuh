// coll_kernels.cuh — element-wise reduce / allreduce across GPUs (MPI_Reduce semantics, mpi/reduce.c:76,90) as ONE
// kernel per rank that moves data itself over NVLink: peer loads/stores on the mapped symmetric heaps, or
// multimem.ld_reduce / multimem.st through the NVSwitch. No NCCL, no separate local-reduce kernel, no host in the loop.
//
// Common skeleton (gridDim.x CTAs per rank; CTA b of every rank talks to CTA b of its peers):
//   start barrier  : "my kernel has started" == everything ordered before it in my stream is done (send buffers ready)
//   data phase     : algorithm specific, 16-byte vectors, fixed rank order for floating-point sums
//   end barrier    : every peer is done reading my send buffer / writing my recv buffer
// In loopback mode (tests on one GPU) the same kernel is launched ONCE, cooperatively, with gridDim.y = number of
// virtual ranks and blockIdx.y selecting the rank's arguments.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../comm/coll_ctx.h"
#include "../core/primitives.cuh"

namespace b2r {

struct CollArgs {
  CollCtx ctx;
  unsigned long long send_off, recv_off, work_off;  // byte offsets inside every rank's heap (16-byte aligned)
  unsigned long long count;                         // elements
  unsigned int epoch;                               // barrier values epoch+1, epoch+2, ...
  int root;                                         // >= 0: reduce to root; -1: allreduce
  unsigned int seq;                                 // push / ring: call number (flag values are seq<<16 | round)
  int ring_depth;                                   // push / ring: staging slots per (source, CTA)
  int sub_blocks;                                   // ring: sub-blocks of (threads x 8) vectors moved per flag/fence
  unsigned int stagger_ns;                          // ring: delay of the odd channels at the head of the chain (0 = none)
  const char* send_ptr;                             // LL: this rank's buffers as plain device pointers (peers never touch them,
  char* recv_ptr;                                   //     so they need not live in the heap); 16-byte aligned
};
constexpr int kLoopMaxRanks = 8;
struct CollLaunch {
  CollArgs a[kLoopMaxRanks];  // a[blockIdx.y]; real multi-GPU launches use a[0] only
};

// ------------------------------------------------------------------------------------------------ memory ops on peer / multicast addresses
// Bulk data accesses come in two flavours, chosen per process by B2R_WEAK_DATA (0 = default):
//   strong: ld/st.relaxed.sys  (SASS LDG/STG ... STRONG.SYS) — what every number recorded in profiles/ was measured with;
//   weak  : plain ld/st.global (SASS LDG.E.128 / STG.E.128)  — what NCCL-style kernels use for payload; ordering is still
//           provided by the st.release.sys / ld.acquire.sys flags around the data phase (the acquire also invalidates L1).
// The weak flavour is a candidate for the pull/push paths that plateau below link rate (profiles/multi_gpu.md §3); it is
// correct under the PTX memory model but not yet measured on NVLink, hence opt-in.
__constant__ int c_weak_data = 0;

__device__ __forceinline__ V16 ld_peer(const char* p) {  // may be local or remote; never through the read-only path
  V16 v;
  if (c_weak_data) {
    asm volatile("ld.global.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3])
                 : "l"(p)
                 : "memory");
  } else {
    asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3])
                 : "l"(p)
                 : "memory");
  }
  return v;
}
__device__ __forceinline__ void st_peer(char* p, const V16& v) {
  if (c_weak_data) {
    asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]), "r"(v.w[3]) : "memory");
  } else {
    asm volatile("st.relaxed.sys.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]), "r"(v.w[3]) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------ vector combine per (dtype, op)
template <int D, int O>
struct VecOps;

template <int O>
struct VecOps<B2R_INT32, O> {
  static constexpr int kElem = 4;
  struct Acc { int32_t v[4]; };
  __device__ __forceinline__ static Acc unpack(const V16& x) { Acc a; for (int j = 0; j < 4; ++j) a.v[j] = (int32_t)x.w[j]; return a; }
  __device__ __forceinline__ static void combine(Acc& a, const V16& x) { for (int j = 0; j < 4; ++j) a.v[j] = OpI32<O>::apply(a.v[j], (int32_t)x.w[j]); }
  __device__ __forceinline__ static V16 pack(const Acc& a) { V16 x; for (int j = 0; j < 4; ++j) x.w[j] = (uint32_t)a.v[j]; return x; }
};
template <int O>
struct VecOps<B2R_FLOAT32, O> {
  static constexpr int kElem = 4;
  struct Acc { float v[4]; };
  __device__ __forceinline__ static Acc unpack(const V16& x) { Acc a; for (int j = 0; j < 4; ++j) a.v[j] = __uint_as_float(x.w[j]); return a; }
  __device__ __forceinline__ static void combine(Acc& a, const V16& x) { for (int j = 0; j < 4; ++j) a.v[j] = OpF32<O>::apply(a.v[j], __uint_as_float(x.w[j])); }
  __device__ __forceinline__ static V16 pack(const Acc& a) { V16 x; for (int j = 0; j < 4; ++j) x.w[j] = __float_as_uint(a.v[j]); return x; }
};
template <int O>
struct VecOps<B2R_FLOAT64, O> {
  static constexpr int kElem = 8;
  struct Acc { double v[2]; };
  __device__ __forceinline__ static Acc unpack(const V16& x) {
    Acc a;
    a.v[0] = __hiloint2double((int)x.w[1], (int)x.w[0]);
    a.v[1] = __hiloint2double((int)x.w[3], (int)x.w[2]);
    return a;
  }
  __device__ __forceinline__ static void combine(Acc& a, const V16& x) {
    a.v[0] = OpF64<O>::apply(a.v[0], __hiloint2double((int)x.w[1], (int)x.w[0]));
    a.v[1] = OpF64<O>::apply(a.v[1], __hiloint2double((int)x.w[3], (int)x.w[2]));
  }
  __device__ __forceinline__ static V16 pack(const Acc& a) {
    V16 x;
    x.w[0] = (uint32_t)__double2loint(a.v[0]);
    x.w[1] = (uint32_t)__double2hiint(a.v[0]);
    x.w[2] = (uint32_t)__double2loint(a.v[1]);
    x.w[3] = (uint32_t)__double2hiint(a.v[1]);
    return x;
  }
};
// bf16 travels as bf16 but all P contributions are combined in fp32 and rounded ONCE (better than P-1 bf16 roundings).
template <int O>
struct VecOps<B2R_BFLOAT16, O> {
  static constexpr int kElem = 2;
  struct Acc { float v[8]; };
  __device__ __forceinline__ static Acc unpack(const V16& x) {
    Acc a;
    for (int j = 0; j < 4; ++j) {
      a.v[2 * j] = __uint_as_float(x.w[j] << 16);
      a.v[2 * j + 1] = __uint_as_float(x.w[j] & 0xffff0000u);
    }
    return a;
  }
  __device__ __forceinline__ static void combine(Acc& a, const V16& x) {
    for (int j = 0; j < 4; ++j) {
      a.v[2 * j] = OpF32<O>::apply(a.v[2 * j], __uint_as_float(x.w[j] << 16));
      a.v[2 * j + 1] = OpF32<O>::apply(a.v[2 * j + 1], __uint_as_float(x.w[j] & 0xffff0000u));
    }
  }
  __device__ __forceinline__ static V16 pack(const Acc& a) {
    V16 x;
    for (int j = 0; j < 4; ++j) {
      __nv_bfloat162 h = __floats2bfloat162_rn(a.v[2 * j], a.v[2 * j + 1]);
      x.w[j] = *reinterpret_cast<uint32_t*>(&h);
    }
    return x;
  }
};

// scalar helpers for the (< 16 byte) tail: same accumulator precision and single final rounding as the vector path
template <int D>
struct TailAcc { using type = float; };
template <>
struct TailAcc<B2R_INT32> { using type = int32_t; };
template <>
struct TailAcc<B2R_FLOAT64> { using type = double; };

template <int D>
__device__ __forceinline__ typename TailAcc<D>::type tail_load(const char* p) {
  if constexpr (D == B2R_INT32) return *(const volatile int32_t*)p;
  else if constexpr (D == B2R_FLOAT32) return *(const volatile float*)p;
  else if constexpr (D == B2R_FLOAT64) return *(const volatile double*)p;
  else return __uint_as_float(((uint32_t) * (const volatile uint16_t*)p) << 16);
}
template <int D, int O>
__device__ __forceinline__ typename TailAcc<D>::type tail_apply(typename TailAcc<D>::type a, typename TailAcc<D>::type b) {
  if constexpr (D == B2R_INT32) return OpI32<O>::apply(a, b);
  else if constexpr (D == B2R_FLOAT64) return OpF64<O>::apply(a, b);
  else return OpF32<O>::apply(a, b);
}
template <int D>
__device__ __forceinline__ void tail_store(char* p, typename TailAcc<D>::type v) {
  if constexpr (D == B2R_INT32) *(volatile int32_t*)p = v;
  else if constexpr (D == B2R_FLOAT32) *(volatile float*)p = v;
  else if constexpr (D == B2R_FLOAT64) *(volatile double*)p = v;
  else {
    __nv_bfloat16 h = __float2bfloat16_rn(v);
    *(volatile uint16_t*)p = *reinterpret_cast<uint16_t*>(&h);
  }
}

// ------------------------------------------------------------------------------------------------ CTA-to-CTA barrier across ranks
struct PeerOffsets {  // shared-memory copy of what every peer published for this collective
  unsigned long long send[B2R_MAX_RANKS], recv[B2R_MAX_RANKS];
};

__device__ __forceinline__ void cta_barrier(const CollArgs& A, unsigned int value) {
  const CollCtx& c = A.ctx;
  // bar.sync orders every thread's data stores before the signalling threads' st.release.sys (cumulativity)
  __syncthreads();
  if (threadIdx.x < (unsigned)c.nranks) {
    const int peer = (int)threadIdx.x;
    if (!c.fault_skip_signal) {
      unsigned int* remote = (unsigned int*)(c.heap[peer] + kOffFlags) + (size_t)blockIdx.x * B2R_MAX_RANKS + c.rank;
      st_release_sys(remote, value);
    }
    const unsigned int* local = (const unsigned int*)(c.heap[c.rank] + kOffFlags) + (size_t)blockIdx.x * B2R_MAX_RANKS + peer;
    wait_flag_ge(local, value, c);
  }
  __syncthreads();
}

// Start barrier: additionally publishes my buffer offsets to every peer (plain stores ordered by the release store of
// the flag) and collects theirs. "My kernel has started" implies my send buffer is ready (stream order).
__device__ __forceinline__ void start_barrier(const CollArgs& A, PeerOffsets* po) {
  const CollCtx& c = A.ctx;
  __syncthreads();
  if (threadIdx.x < (unsigned)c.nranks) {
    const int peer = (int)threadIdx.x;
    fault_delay(c);
    if (!c.fault_skip_signal) {
      PeerMeta* rm = (PeerMeta*)(c.heap[peer] + kOffMeta) + (size_t)blockIdx.x * B2R_MAX_RANKS + c.rank;
      st_relaxed_sys_u64(&rm->send_off, A.send_off);
      st_relaxed_sys_u64(&rm->recv_off, A.recv_off);
      unsigned int* remote = (unsigned int*)(c.heap[peer] + kOffFlags) + (size_t)blockIdx.x * B2R_MAX_RANKS + c.rank;
      st_release_sys(remote, A.epoch + 1);
    }
    const unsigned int* local = (const unsigned int*)(c.heap[c.rank] + kOffFlags) + (size_t)blockIdx.x * B2R_MAX_RANKS + peer;
    if (wait_flag_ge(local, A.epoch + 1, c)) {
      const PeerMeta* lm = (const PeerMeta*)(c.heap[c.rank] + kOffMeta) + (size_t)blockIdx.x * B2R_MAX_RANKS + peer;
      po->send[peer] = ld_relaxed_sys_u64(&lm->send_off);
      po->recv[peer] = ld_relaxed_sys_u64(&lm->recv_off);
    } else {  // timed out: keep the kernel memory-safe, the error word tells the host the result is garbage
      po->send[peer] = A.send_off;
      po->recv[peer] = A.recv_off;
    }
  }
  __syncthreads();
}

// Streams vectors [lo, hi) of `nsrc` sources (combined in the order given), writes the result to `ndst` destinations.
// UNR independent vectors per thread are in flight, and the loads of source s+1 are issued before source s is combined.
template <int D, int O, int UNR>
__device__ __forceinline__ void combine_range(unsigned long long lo, unsigned long long hi, const char* const* src, int nsrc, char* const* dst,
                                              int ndst) {
  using V = VecOps<D, O>;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long base = lo + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; base < hi; base += stride * UNR) {
    typename V::Acc acc[UNR];
    V16 cur[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const unsigned long long i = base + u * stride;
      if (i < hi) cur[u] = ld_peer(src[0] + i * 16);
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) acc[u] = V::unpack(cur[u]);
    if (nsrc > 1) {
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const unsigned long long i = base + u * stride;
        if (i < hi) cur[u] = ld_peer(src[1] + i * 16);
      }
      for (int s = 1; s < nsrc; ++s) {
        V16 nxt[UNR];
        if (s + 1 < nsrc) {
#pragma unroll
          for (int u = 0; u < UNR; ++u) {
            const unsigned long long i = base + u * stride;
            if (i < hi) nxt[u] = ld_peer(src[s + 1] + i * 16);
          }
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) V::combine(acc[u], cur[u]);
#pragma unroll
        for (int u = 0; u < UNR; ++u) cur[u] = nxt[u];
      }
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const unsigned long long i = base + u * stride;
      if (i < hi) {
        const V16 r = V::pack(acc[u]);
        for (int d = 0; d < ndst; ++d) st_peer(dst[d] + i * 16, r);
      }
    }
  }
}

// the tail elements after the last full 16-byte vector, done by one thread per element in CTA 0
template <int D, int O>
__device__ __forceinline__ void combine_tail(unsigned long long first_elem, unsigned long long count, const char* const* src, int nsrc,
                                             char* const* dst, int ndst) {
  constexpr int E = VecOps<D, O>::kElem;
  if (blockIdx.x != 0) return;
  for (unsigned long long e = first_elem + threadIdx.x; e < count; e += blockDim.x) {
    typename TailAcc<D>::type acc = tail_load<D>(src[0] + e * E);
    for (int s = 1; s < nsrc; ++s) acc = tail_apply<D, O>(acc, tail_load<D>(src[s] + e * E));
    for (int d = 0; d < ndst; ++d) tail_store<D>(dst[d] + e * E, acc);
  }
}


// ------------------------------------------------------------------------------------------------ one-shot
// Every rank (allreduce) or only the root (reduce) reads all P send buffers and combines them in rank order.
template <int D, int O, int UNR>
__device__ void oneshot_body(const CollArgs& A) {
  const CollCtx& c = A.ctx;
  constexpr int E = VecOps<D, O>::kElem;
  const unsigned long long nvec = A.count * E / 16;
  __shared__ PeerOffsets po;
  start_barrier(A, &po);
  if (A.root < 0 || A.root == c.rank) {
    __shared__ const char* src[B2R_MAX_RANKS];  // pointer tables live in shared memory: indexed at run time in the hot loop
    __shared__ char* dst[1];
    if (threadIdx.x == 0) {
      for (int p = 0; p < c.nranks; ++p) src[p] = c.heap[p] + po.send[p];
      dst[0] = c.heap[c.rank] + A.recv_off;
    }
    __syncthreads();
    combine_range<D, O, UNR>(0, nvec, src, c.nranks, dst, 1);
    combine_tail<D, O>(nvec * (16 / E), A.count, src, c.nranks, dst, 1);
  }
  cta_barrier(A, A.epoch + 2);
}

// ------------------------------------------------------------------------------------------------ two-shot
// allreduce: rank r pulls slice r from every peer (reduce-scatter by loads), combines, and pushes the result into every
//            peer's recv buffer (all-gather by stores) — per GPU 2(P-1)/P of the message in each direction.
// reduce   : the P-1 non-root ranks each own a slice and push only to the root, so the root's ingress is exactly one
//            message (not 2(P-1)/P of it) and every link carries the same load.
template <int D, int O, int UNR>
__device__ void twoshot_body(const CollArgs& A) {
  const CollCtx& c = A.ctx;
  constexpr int E = VecOps<D, O>::kElem;
  const int P = c.nranks;
  const unsigned long long nvec = A.count * E / 16;
  __shared__ PeerOffsets po;
  start_barrier(A, &po);
  int nworkers, me;
  if (A.root < 0) {
    nworkers = P;
    me = c.rank;
  } else {
    nworkers = P - 1;
    me = c.rank == A.root ? -1 : (c.rank < A.root ? c.rank : c.rank - 1);
  }
  if (me >= 0) {
    const unsigned long long lo = nvec * (unsigned long long)me / nworkers, hi = nvec * (unsigned long long)(me + 1) / nworkers;
    __shared__ const char* src[B2R_MAX_RANKS];
    __shared__ char* dst[B2R_MAX_RANKS];
    const int ndst = A.root < 0 ? P : 1;
    if (threadIdx.x == 0) {
      for (int p = 0; p < P; ++p) src[p] = c.heap[p] + po.send[p];
      if (A.root < 0) {
        for (int p = 0; p < P; ++p) dst[p] = c.heap[p] + po.recv[p];
      } else {
        dst[0] = c.heap[A.root] + po.recv[A.root];
      }
    }
    __syncthreads();
    combine_range<D, O, UNR>(lo, hi, src, P, dst, ndst);
    if (me == nworkers - 1) combine_tail<D, O>(nvec * (16 / E), A.count, src, P, dst, ndst);
  }
  cta_barrier(A, A.epoch + 2);
}

// ------------------------------------------------------------------------------------------------ hand-written binomial tree
// ceil(log2 P) rounds; in round k the rank at virtual position v (v % 2^(k+1) == 0) pulls the running partial of
// position v + 2^k and combines it into its own work buffer; the last round writes the final destination. This is the
// "custom tree vs library reduce" arm of the assignment; on a uniform NVSwitch it moves (P-1) whole messages through
// log2 P dependent rounds, so it is expected to lose to two-shot for large messages.
template <int D, int O, int UNR>
__device__ void tree_body(const CollArgs& A) {
  const CollCtx& c = A.ctx;
  constexpr int E = VecOps<D, O>::kElem;
  const int P = c.nranks;
  const int root = A.root < 0 ? 0 : A.root;
  const int v = (c.rank - root + P) % P;
  const unsigned long long nvec = A.count * E / 16;
  int rounds = 0;
  while ((1 << rounds) < P) ++rounds;
  __shared__ PeerOffsets po;
  start_barrier(A, &po);
  for (int k = 0; k < rounds; ++k) {
    const int bit = 1 << k;
    if (v % (2 * bit) == 0 && v + bit < P) {
      const int partner = (v + bit + root) % P;
      const bool last = (k == rounds - 1);
      // a position has a running partial in `work` once it has received at least once (positions are even for k >= 1)
      const bool mine_in_work = k >= 1 && v + 1 < P;
      const bool peer_in_work = k >= 1 && (v + bit) + 1 < P;
      __shared__ const char* src[2];
      __shared__ char* dst[B2R_MAX_RANKS];
      const int ndst = (last && A.root < 0) ? P : 1;
      if (threadIdx.x == 0) {
        src[0] = c.heap[c.rank] + (mine_in_work ? A.work_off : A.send_off);
        src[1] = c.heap[partner] + (peer_in_work ? A.work_off : po.send[partner]);
        if (!last) dst[0] = c.heap[c.rank] + A.work_off;
        else if (A.root >= 0) dst[0] = c.heap[c.rank] + A.recv_off;
        else
          for (int p = 0; p < P; ++p) dst[p] = c.heap[p] + po.recv[p];  // allreduce: the root broadcasts by stores
      }
      __syncthreads();
      combine_range<D, O, UNR>(0, nvec, src, 2, dst, ndst);
      combine_tail<D, O>(nvec * (16 / E), A.count, src, 2, dst, ndst);
    }
    cta_barrier(A, A.epoch + 2 + k);
  }
}

}  // namespace b2r
#include "coll_push.cuh"  // push: scatter by stores into whole-slice staging, combine locally, store the result
namespace b2r {

// ------------------------------------------------------------------------------------------------ ring (store-forwarding chain to the root)
// What NCCL's reduce does, restated for this runtime: the ranks form a chain that ends at the root (position p = distance
// to the root); the message is cut into blocks; CTA b of every rank owns blocks b, b+G, ... ("channels"). For each block
// the rank at position p waits for the running partial from position p+1 in its own staging slot (local memory), combines
// it with its own block and STORES the result into the slot of position p-1 (the root stores into recv). Every GPU
// receives one message and sends one message, all of it posted stores to exactly ONE neighbour: no remote loads, no
// incast. Slots are a ring of `depth` entries per CTA with data flags (downstream) and acks (upstream), values
// (call << 16 | step): monotonic, never reset. Combination order = chain order (farthest rank first), the host restatement
// (host_algos.h, B2R_ALGO_RING) folds in the same order, so results are bit-identical to it. Reduce only.
template <int D, int O, int UNR>
__device__ void ring_body(const CollArgs& A) {
  using V = VecOps<D, O>;
  const CollCtx& c = A.ctx;
  constexpr int E = V::kElem;
  constexpr int VPT = 8;  // vectors per thread per sub-block
  const int P = c.nranks, r = c.rank, root = A.root, G = (int)gridDim.x, b = (int)blockIdx.x;
  const int pos = (r - root + P) % P;
  const int pred = pos + 1 < P ? (root + pos + 1) % P : -1;  // sends to me
  const int succ = pos > 0 ? (root + pos - 1) % P : -1;      // I send to it
  const unsigned long long nvec = A.count * E / 16;
  // One step = sub_blocks sub-blocks of (threads x 8) vectors, then ONE release-store of the data flag: the fence in
  // front of that store waits for the step's outstanding remote stores, so its cost is amortised over the whole step
  // (64 KiB steps measured 407 GB/s at 2 GPUs where the link sustains 630+).
  const unsigned int Sv = blockDim.x * VPT;
  const unsigned int Cv = Sv * (unsigned int)max(1, A.sub_blocks);
  const unsigned long long nblocks = (nvec + Cv - 1) / Cv;
  const unsigned long long steps = (unsigned long long)b < nblocks ? (nblocks - b + G - 1) / G : 0;
  const int depth = A.ring_depth;
  const unsigned int base_val = A.seq << 16;

  __shared__ PeerOffsets po;
  start_barrier(A, &po);
  const char* send = c.heap[r] + A.send_off;
  auto slot = [&](int rank, unsigned long long step) {
    return c.heap[rank] + A.work_off + (((step % depth) * G + b) * (unsigned long long)Cv) * 16ull;
  };
  // All channels of a rank run the same step at the same time — stream, then sit in the release fence while the link drains,
  // then wait one hop for the next flag — so the link idles in the latency part of every step (measured: 30 us per 512 KiB step,
  // 21 of them link time). Delaying the ODD channels by about half a step at the head of the chain (the delay then travels down
  // the chain with the flags) makes one half stream while the other half waits.
  if (A.stagger_ns && (b & 1) && pred < 0 && threadIdx.x == 0) {
    const unsigned long long t0 = coll_globaltimer();
    while (coll_globaltimer() - t0 < (unsigned long long)A.stagger_ns) __nanosleep(500);
  }
  if (A.stagger_ns && (b & 1) && pred < 0) __syncthreads();
  for (unsigned long long i = 0; i < steps; ++i) {
    const unsigned long long lo = (b + i * G) * (unsigned long long)Cv;
    const unsigned int len = (unsigned int)min((unsigned long long)Cv, nvec - lo);
    if (threadIdx.x == 0 && pred >= 0) {
      const unsigned int* flag = (const unsigned int*)(c.heap[r] + kOffPushData) + (size_t)b * B2R_MAX_RANKS + pred;
      wait_flag_ge(flag, base_val + (unsigned int)(i + 1), c);
    }
    if (threadIdx.x == 32 && succ >= 0 && i >= (unsigned long long)depth) {
      const unsigned int* ack = (const unsigned int*)(c.heap[r] + kOffPushAck) + (size_t)b * B2R_MAX_RANKS + succ;
      wait_flag_ge(ack, base_val + (unsigned int)(i - depth + 1), c);
    }
    __syncthreads();
    char* dst = succ >= 0 ? slot(succ, i) : c.heap[r] + A.recv_off + lo * 16;
    const char* in = slot(r, i);
    for (unsigned int sb = 0; sb < len; sb += Sv) {
      V16 own[VPT], got[VPT];
#pragma unroll
      for (int u = 0; u < VPT; ++u) {
        const unsigned int j = sb + threadIdx.x + u * blockDim.x;
        if (j < len) {
          own[u] = ld_peer(send + (lo + j) * 16);
          if (pred >= 0) got[u] = ld_peer(in + (unsigned long long)j * 16);
        }
      }
#pragma unroll
      for (int u = 0; u < VPT; ++u) {
        const unsigned int j = sb + threadIdx.x + u * blockDim.x;
        if (j < len) {
          if (pred >= 0) {
            typename V::Acc acc = V::unpack(got[u]);
            V::combine(acc, own[u]);
            st_peer(dst + (unsigned long long)j * 16, V::pack(acc));
          } else {
            st_peer(dst + (unsigned long long)j * 16, own[u]);
          }
        }
      }
    }
    __syncthreads();
    if (threadIdx.x == 0 && succ >= 0 && !c.fault_skip_signal) {
      unsigned int* flag = (unsigned int*)(c.heap[succ] + kOffPushData) + (size_t)b * B2R_MAX_RANKS + r;
      st_release_sys(flag, base_val + (unsigned int)(i + 1));
    }
    if (threadIdx.x == 32 && pred >= 0 && !c.fault_skip_signal) {
      unsigned int* ack = (unsigned int*)(c.heap[pred] + kOffPushAck) + (size_t)b * B2R_MAX_RANKS + r;
      st_release_sys(ack, base_val + (unsigned int)(i + 1));
    }
  }
  // the (< 16 byte) tail: the root pulls it, folding in chain order and — like the vector path, whose partials travel as
  // bf16 — rounding bf16 after every hop
  if (pos == 0 && nvec * (16 / E) < A.count && blockIdx.x == 0) {
    for (unsigned long long e = nvec * (16 / E) + threadIdx.x; e < A.count; e += blockDim.x) {
      typename TailAcc<D>::type acc = tail_load<D>(c.heap[(root + P - 1) % P] + po.send[(root + P - 1) % P] + e * E);
      for (int q = 1; q < P; ++q) {
        const int rank_q = (root + (P - 1 - q)) % P;
        acc = tail_apply<D, O>(acc, tail_load<D>(c.heap[rank_q] + po.send[rank_q] + e * E));
        if constexpr (D == B2R_BFLOAT16) acc = __bfloat162float(__float2bfloat16_rn(acc));
      }
      tail_store<D>(c.heap[r] + A.recv_off + e * E, acc);
    }
  }
  cta_barrier(A, A.epoch + 2);
}

// ------------------------------------------------------------------------------------------------ NVLS (multimem)
// The switch performs the combine: multimem.ld_reduce on the multicast address returns op over all ranks' copies;
// multimem.st writes all ranks' copies. Supported by the hardware for: fp32 add (v4), bf16x2 add/min/max (v4, fp32
// accumulation for add), fp64 add (scalar), 32-bit integer add/min/max (scalar). fp32/fp64 min/max do not exist.
template <int D, int O>
struct Nvls {
  static constexpr bool kSupported = false;
  static constexpr int kBytes = 16;
  __device__ static void ld_reduce(const char*, uint32_t*) {}
  __device__ static void st(char*, const uint32_t*) {}
};
// fp32 / fp64 min and max have no multimem form, but signed-integer min / max do, and IEEE values order like the integers
// k = b ^ ((b >> 31) & 0x7fff...) (the map is its own inverse; -0 orders below +0, a NaN beyond the infinity of its sign). The host
// stages the send buffer as keys (encode_keys_kernel, a local copy into the work region), the switch reduces the keys, and the
// value is decoded before it is stored — so these specialisations read KEYS and write VALUES.
template <int D, int O>
struct NvlsKeyed {
  static constexpr bool value = (D == B2R_FLOAT32 || D == B2R_FLOAT64) && O != B2R_SUM;
};
__device__ __forceinline__ uint32_t key32(uint32_t b) { return b ^ ((uint32_t)((int32_t)b >> 31) & 0x7fffffffu); }
__device__ __forceinline__ void key64(uint32_t& lo, uint32_t& hi) {
  const uint32_t m = (uint32_t)((int32_t)hi >> 31);
  lo ^= m;
  hi ^= m & 0x7fffffffu;
}
#define B2R_NVLS_KEY32(OP, PTXOP)                                                                                           \
  template <>                                                                                                               \
  struct Nvls<B2R_FLOAT32, OP> {                                                                                            \
    static constexpr bool kSupported = true;                                                                                \
    static constexpr int kBytes = 4;                                                                                        \
    __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) {                                          \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." PTXOP " %0, [%1];" : "=r"(w[0]) : "l"(p) : "memory");           \
      w[0] = key32(w[0]);                                                                                                   \
    }                                                                                                                       \
    __device__ __forceinline__ static void st(char* p, const uint32_t* w) {                                                 \
      asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(w[0]) : "memory");                          \
    }                                                                                                                       \
  };
B2R_NVLS_KEY32(B2R_MIN, "min.s32")
B2R_NVLS_KEY32(B2R_MAX, "max.s32")
#undef B2R_NVLS_KEY32
#define B2R_NVLS_KEY64(OP, PTXOP)                                                                                           \
  template <>                                                                                                               \
  struct Nvls<B2R_FLOAT64, OP> {                                                                                            \
    static constexpr bool kSupported = true;                                                                                \
    static constexpr int kBytes = 8;                                                                                        \
    __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) {                                          \
      unsigned long long d;                                                                                                 \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." PTXOP " %0, [%1];" : "=l"(d) : "l"(p) : "memory");              \
      w[0] = (uint32_t)(d & 0xffffffffull);                                                                                 \
      w[1] = (uint32_t)(d >> 32);                                                                                           \
      key64(w[0], w[1]);                                                                                                    \
    }                                                                                                                       \
    __device__ __forceinline__ static void st(char* p, const uint32_t* w) {                                                 \
      const unsigned long long d = (unsigned long long)w[0] | ((unsigned long long)w[1] << 32);                             \
      asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(d) : "memory");                             \
    }                                                                                                                       \
  };
B2R_NVLS_KEY64(B2R_MIN, "min.s64")
B2R_NVLS_KEY64(B2R_MAX, "max.s64")
#undef B2R_NVLS_KEY64

// The staging pass of the keyed forms: dst[i] = key(src[i]) over `nunits` units of W bytes (4 or 8); 16-byte vectors when both
// pointers allow it. Purely local.
template <int W>
__global__ void __launch_bounds__(512) encode_keys_kernel(const char* __restrict__ src, char* __restrict__ dst, unsigned long long nunits) {
  const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (unsigned long long)gridDim.x * blockDim.x;
  unsigned long long done = 0;
  if ((((unsigned long long)src | (unsigned long long)dst) & 15ull) == 0) {
    const unsigned long long nvec = nunits * W / 16;
    for (unsigned long long v = tid; v < nvec; v += nthreads) {
      uint4 x = *reinterpret_cast<const uint4*>(src + v * 16);
      if (W == 4) {
        x.x = key32(x.x); x.y = key32(x.y); x.z = key32(x.z); x.w = key32(x.w);
      } else {
        key64(x.x, x.y);
        key64(x.z, x.w);
      }
      *reinterpret_cast<uint4*>(dst + v * 16) = x;
    }
    done = nvec * (16 / W);
  }
  for (unsigned long long i = done + tid; i < nunits; i += nthreads) {
    if (W == 4) {
      *reinterpret_cast<uint32_t*>(dst + i * 4) = key32(*reinterpret_cast<const uint32_t*>(src + i * 4));
    } else {
      uint2 x = *reinterpret_cast<const uint2*>(src + i * 8);
      key64(x.x, x.y);
      *reinterpret_cast<uint2*>(dst + i * 8) = x;
    }
  }
}
template <>
struct Nvls<B2R_FLOAT32, B2R_SUM> {
  static constexpr bool kSupported = true;
  static constexpr int kBytes = 16;
  __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p) : "memory");
  }
  __device__ __forceinline__ static void st(char* p, const uint32_t* w) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
  }
};
#define B2R_NVLS_BF16(OP, PTXOP)                                                                                                  \
  template <>                                                                                                                     \
  struct Nvls<B2R_BFLOAT16, OP> {                                                                                                 \
    static constexpr bool kSupported = true;                                                                                      \
    static constexpr int kBytes = 16;                                                                                             \
    __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) {                                                \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." PTXOP ".v4.bf16x2 {%0,%1,%2,%3}, [%4];"                               \
                   : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3])                                                               \
                   : "l"(p)                                                                                                       \
                   : "memory");                                                                                                   \
    }                                                                                                                             \
    __device__ __forceinline__ static void st(char* p, const uint32_t* w) {                                                       \
      asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) \
                   : "memory");                                                                                                   \
    }                                                                                                                             \
  };
B2R_NVLS_BF16(B2R_SUM, "add.acc::f32")
B2R_NVLS_BF16(B2R_MIN, "min")
B2R_NVLS_BF16(B2R_MAX, "max")
#undef B2R_NVLS_BF16
template <>
struct Nvls<B2R_FLOAT64, B2R_SUM> {
  static constexpr bool kSupported = true;
  static constexpr int kBytes = 8;
  __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) {
    double d;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(d) : "l"(p) : "memory");
    w[0] = (uint32_t)__double2loint(d);
    w[1] = (uint32_t)__double2hiint(d);
  }
  __device__ __forceinline__ static void st(char* p, const uint32_t* w) {
    double d = __hiloint2double((int)w[1], (int)w[0]);
    asm volatile("multimem.st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(d) : "memory");
  }
};
#define B2R_NVLS_I32(OP, PTXOP)                                                                                             \
  template <>                                                                                                               \
  struct Nvls<B2R_INT32, OP> {                                                                                              \
    static constexpr bool kSupported = true;                                                                                \
    static constexpr int kBytes = 4;                                                                                        \
    __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) {                                          \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." PTXOP " %0, [%1];" : "=r"(w[0]) : "l"(p) : "memory");           \
    }                                                                                                                       \
    __device__ __forceinline__ static void st(char* p, const uint32_t* w) {                                                 \
      asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(w[0]) : "memory");                          \
    }                                                                                                                       \
  };
B2R_NVLS_I32(B2R_SUM, "add.u32")
B2R_NVLS_I32(B2R_MIN, "min.s32")
B2R_NVLS_I32(B2R_MAX, "max.s32")
#undef B2R_NVLS_I32

template <int D, int O, int UNR>
__device__ void nvls_body(const CollArgs& A) {
  using N = Nvls<D, O>;
  const CollCtx& c = A.ctx;
  constexpr int E = VecOps<D, O>::kElem;
  constexpr int B = N::kBytes;  // bytes per multimem instruction
  const int P = c.nranks;
  __shared__ PeerOffsets po;
  start_barrier(A, &po);
  // one multicast address reaches the same offset of every heap, so the buffers MUST be symmetric here
  bool symmetric = true;
  for (int p = 0; p < P; ++p) symmetric = symmetric && po.send[p] == A.send_off && (A.root >= 0 || po.recv[p] == A.recv_off);
  if (!symmetric && threadIdx.x == 0) atomicExch((unsigned int*)(c.heap[c.rank] + kOffError), 2u);
  if (N::kSupported && symmetric) {
    // units of B bytes; count*E is a multiple of B for every supported (dtype, op) except bf16 tails (< 16 bytes),
    // which rank P-1 / the root finishes with peer loads.
    const unsigned long long nunit = A.count * E / B;
    // Every rank owns a slice and lets the switch reduce it (ingress 1/P of the message per GPU). allreduce: the result
    // goes back out through the multicast address (multimem.st reaches all P copies). reduce: the result is stored
    // straight into the ROOT's recv buffer over NVLink — issuing the multimem loads from all P GPUs instead of from the
    // root alone keeps every GPU's request stream short (a root-only variant tops out near 385 GB/s).
    unsigned long long lo = nunit * (unsigned long long)c.rank / P, hi = nunit * (unsigned long long)(c.rank + 1) / P;
    if (A.root >= 0 && A.sub_blocks > 0) {
      // reduce with an uneven split: the root pulls the first sub_blocks/64 of the message itself (its result needs no second
      // hop: egress of the other GPUs drops from 1 + 1/P towards 1 message), the other ranks share the rest equally
      const unsigned long long cut = nunit / 64 * (unsigned)A.sub_blocks;
      if (c.rank == A.root) {
        lo = 0;
        hi = cut;
      } else {
        const unsigned long long w = (unsigned long long)(c.rank < A.root ? c.rank : c.rank - 1);
        lo = cut + (nunit - cut) * w / (P - 1);
        hi = cut + (nunit - cut) * (w + 1) / (P - 1);
      }
    }
    {
      const char* msrc = c.mc + A.send_off;
      char* mdst = c.mc + A.recv_off;
      char* rdst = A.root < 0 ? nullptr : c.heap[A.root] + po.recv[A.root];
      char* ldst = c.heap[c.rank] + A.recv_off;
      const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
      constexpr int U = UNR;
      for (unsigned long long base = lo + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; base < hi; base += stride * U) {
        uint32_t w[U][4];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long i = base + u * stride;
          if (i < hi) N::ld_reduce(msrc + i * B, w[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long i = base + u * stride;
          if (i < hi) {
            if (A.root < 0) N::st(mdst + i * B, w[u]);
            else if (B == 16) st_peer(rdst + i * 16, *reinterpret_cast<const V16*>(w[u]));
            else if (B == 8) asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1,%2};" ::"l"(rdst + i * 8), "r"(w[u][0]), "r"(w[u][1]) : "memory");
            else asm volatile("st.relaxed.sys.global.b32 [%0], %1;" ::"l"(rdst + i * 4), "r"(w[u][0]) : "memory");
          }
        }
      }
      const bool tail_owner = A.root < 0 ? (c.rank == P - 1) : (c.rank == A.root);
      if (tail_owner && nunit * B < A.count * E) {
        __shared__ const char* src[B2R_MAX_RANKS];
        __shared__ char* dst[B2R_MAX_RANKS];
        const int ndst = A.root < 0 ? P : 1;
        if (threadIdx.x == 0) {
          for (int p = 0; p < P; ++p) src[p] = c.heap[p] + po.send[p];
          if (A.root < 0) for (int p = 0; p < P; ++p) dst[p] = c.heap[p] + po.recv[p];
          else dst[0] = ldst;
        }
        __syncthreads();
        combine_tail<D, O>(nunit * B / E, A.count, src, P, dst, ndst);
      }
    }
  }
  cta_barrier(A, A.epoch + 2);
}

}  // namespace b2r
#include "coll_experimental.cuh"  // LL small-message path
#include "coll_posted.cuh"        // posted reduce-to-root (stage locally, the root reduces through the switch)
namespace b2r {

// barrier-only kernel body
__device__ inline void barrier_body(const CollArgs& A) {
  __shared__ PeerOffsets po;
  start_barrier(A, &po);
}

enum CollKind { kCollOneshot = 1, kCollTwoshot = 2, kCollTree = 3, kCollNvls = 4, kCollBarrier = 5, kCollPush = 6, kCollRing = 7, kCollLL = 8, kCollPosted = 9 };

template <int D, int O, int KIND, int UNR>
__global__ void __launch_bounds__(512) coll_kernel(const __grid_constant__ CollLaunch L) {
  const CollArgs& A = L.a[blockIdx.y];
  if (KIND == kCollOneshot) oneshot_body<D, O, UNR>(A);
  else if (KIND == kCollTwoshot) twoshot_body<D, O, UNR>(A);
  else if (KIND == kCollTree) tree_body<D, O, UNR>(A);
  else if (KIND == kCollNvls) nvls_body<D, O, UNR>(A);
  else if (KIND == kCollPush) push_body<D, O, UNR>(A);
  else if (KIND == kCollRing) ring_body<D, O, UNR>(A);
  else if (KIND == kCollLL) ll_body<D, O>(A);
  else if (KIND == kCollPosted) posted_body<D, O>(A);
  else barrier_body(A);
}

}  // namespace b2r
