// reduce_kernels.cuh — the single-GPU kernel family (one launch, arbitrary 64-bit n, deterministic).
//
// Pipeline of one launch (replaces sumreduce6/minreduce6/maxreduce6 + the second in-place pass,
// reduction_kernel.cu:74-253 and reduction.cpp:344-357, and generalises the vendored single-pass
// threadFenceReduction, threadFenceReduction_kernel.cu:116-171):
//   stream   : every CTA owns one contiguous, 128-byte-aligned slice of the body and streams it either with
//              unrolled 128/256-bit LDGs (reduce_ldg_kernel) or with a cp.async.bulk ring fed by one producer
//              lane (reduce_tma_kernel); head/tail elements that do not fill a vector are folded in by CTA 0.
//   warp     : redux.sync (int32, fp32 min/max) or shfl.sync butterflies.
//   block    : one shared-memory hop, warp 0 finishes.
//   cluster  : (optional) non-leader CTAs store their partial into the leader's shared memory (DSMEM),
//              barrier.cluster, leader combines in rank order.
//   grid     : leader stores the partial, takes a ticket with atom.acq_rel.gpu; the last arriver re-reduces all
//              partials in index order, writes the scalar and re-arms the ticket.
#pragma once
#include <cooperative_groups.h>

#include "../comm/coll_ctx.h"
#include "primitives.cuh"

namespace b2r {

namespace cg = cooperative_groups;

// TMA variant: every consumer thread owns this many 16-byte vectors of a tile (tile = consumers x 4 x 16 bytes).
constexpr int kTmaVecsPerThread = 4;

struct LocalParams {
  const void* in;             // element-aligned base pointer
  void* out;                  // one element of the input dtype
  uint64_t n;                 // elements
  uint64_t head;              // scalar elements before the first aligned vector
  uint64_t nvec;              // vectors (16 or 32 bytes) in the body
  uint64_t tail;              // scalar elements after the body
  void* partials;             // acc_t[gridDim.x / cluster]
  unsigned* ticket;           // zero between launches
  unsigned long long* trace;  // nullable: 4 stamps per CTA
  const CollCtx* xg_ctx;      // nullable: scalar-sharded reduce across GPUs (device copy of the communicator context)
  int xg_root;                //   root rank
  int xg_all;                 //   1 = every rank receives the scalar
  unsigned xg_seq;            //   host-side call count (diagnostics only: the kernels number their calls on the device, see xgpu_finish)
  int stages;                 // TMA ring depth
  int tile_vecs;              // TMA tile size in 16-byte vectors
  int out_as_acc;             // store the scalar as acc_t (fp32 for bf16) instead of the input dtype
  int pdl;                    // 0 plain; 1 programmatic launch, wait before the first load; 2 wait only before the workspace
                              // (caller declared the input complete; see b2r_launch_cfg_t.pdl)
  unsigned* tile_ctr;         // dynamic tail: next unclaimed tile (zero between launches, re-armed with the ticket)
  uint64_t dyn_first;         // dynamic tail: first vector of the dynamically scheduled region (== nvec: none)
  unsigned dyn_tile_vecs;     // dynamic tail: vectors per tile (multiple of blockDim.x)
  unsigned dyn_ntiles;        // dynamic tail: number of tiles
  unsigned long long* acc64;  // int32 SUM: packed {arrivals << 48 | running sum} accumulator, zero between launches (nullptr: use partials)
  unsigned prefetch_lines;    // pdl == 1: 128-byte lines per thread to pull into L2 while waiting for the previous kernel (0 = none)
};

// CTA c of G gets vectors [split(c), split(c+1)); boundaries are multiples of `align` vectors (128 bytes).
__device__ __forceinline__ uint64_t split_point(uint64_t nvec, unsigned G, unsigned c, unsigned align) {
  if (c >= G) return nvec;
  // nvec < 2^60 / 16 and c < 2^16: the product fits comfortably in 64 bits for any addressable buffer
  uint64_t s = (nvec / G) * c + ((nvec % G) * c) / G;
  return s - (s % align);
}

// Programmatic dependent launch: let the next kernel of the stream be scheduled while we are still running ...
__device__ __forceinline__ void pdl_launch_dependents(const LocalParams& p) {
  if (p.pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
// ... and, as the dependent, wait for the previous kernel (and the visibility of its writes). Default flavour (pdl == 1):
// before the FIRST LOAD, so an input written by the preceding kernel is read correctly whatever that kernel is — only launch
// latency and the prologue overlap. Flavour 2 (the caller declared the input complete before the previous kernel started):
// the input is streamed first and the wait comes just before the shared workspace (partials, ticket) and `out` are touched.
__device__ __forceinline__ void pdl_wait_before_input(const LocalParams& p) {
  if (p.pdl == 1) asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait_prior(const LocalParams& p) {
  if (p.pdl == 2) asm volatile("griddepcontrol.wait;" ::: "memory");
}

template <class R>
__device__ __forceinline__ void store_result(const LocalParams& p, typename R::acc_t v) {
  if (p.out_as_acc) *(typename R::acc_t*)p.out = v;
  else R::store(p.out, v);
}

// Scalar-sharded reduce across GPUs, fused into the tail of the local reduction (one kernel per rank; the vendored
// simpleMultiGPU copies per-GPU partials to the host and sums there, simpleMultiGPU.cpp:264-293). Called by ALL threads
// of the CTA that holds this rank's scalar (`a` is valid in thread 0):
//   post   : thread 0 stores the partial into the root's mailbox over NVLink as ONE self-flagged 16-byte packet (XgPacket)
//            and is done — no fence, no wait on the exit path of a non-root rank;
//   root   : thread r polls rank r's packet of this call (P waits in parallel, all on local memory), thread 0 combines the
//            partials in rank order (deterministic) and writes `out`, then thread r pushes the result packet (allreduce
//            flavour) and the acknowledgement to rank r — P stores in parallel cost one NVLink trip, not P;
//   others : with `all`, thread 0 waits for the result packet.
// Slots rotate over kXgSlots calls; a rank may run up to kXgSlots calls ahead of the root's acknowledgement, so in a stream of
// back-to-back calls the root finds the packets already there and nobody waits for anybody.
template <class R>
__device__ __forceinline__ void xgpu_finish(const LocalParams& p, typename R::acc_t a) {
  using acc_t = typename R::acc_t;
  using F = typename R::F;
  __shared__ unsigned long long s_vals[B2R_MAX_RANKS];
  __shared__ unsigned long long s_res;
  __shared__ int s_fail;
  __shared__ unsigned s_seq;
  const CollCtx& c = *p.xg_ctx;
  const int P = c.nranks, me = c.rank, root = p.xg_root;
  XgMailbox* my_mb = (XgMailbox*)(c.heap[me] + kOffMailbox);
  // The call number is kept in this rank's own mailbox and advanced here, by the one thread of the one CTA that finishes the
  // rank's launch; launches of one communicator are stream-ordered, and every rank makes the same sequence of calls, so the
  // numbers agree across ranks without any host state — a captured graph of calls replays correctly.
  unsigned seq = 0;
  if (threadIdx.x == 0) {
    seq = my_mb->calls + 1u;
    my_mb->calls = seq;
    s_seq = seq;
  }
  unsigned slot = seq % (unsigned)kXgSlots;
  auto to_bits = [](acc_t v) -> unsigned long long {
    if (sizeof(acc_t) == 8) return (unsigned long long)__double_as_longlong((double)v);
    if (R::kIsInt) return (unsigned long long)(unsigned int)(int)v;
    return (unsigned long long)__float_as_uint((float)v);
  };
  auto from_bits = [](unsigned long long b) -> acc_t {
    if (sizeof(acc_t) == 8) return (acc_t)__longlong_as_double((long long)b);
    if (R::kIsInt) return (acc_t)(int)(unsigned int)b;
    return (acc_t)__uint_as_float((unsigned int)b);
  };
  if (threadIdx.x == 0) {
    s_fail = 0;
    fault_delay(c);
    // the slot was last used by call seq - kXgSlots: the root must have read it
    // (relaxed poll: the value only says the slot may be overwritten; an acquire fence at system scope costs microseconds)
    if (seq > (unsigned)kXgSlots && !wait_flag_ge_relaxed(&my_mb->ack, seq - (unsigned)kXgSlots, c)) s_fail = 1;
    else if (!c.fault_skip_signal) {
      XgMailbox* root_mb = (XgMailbox*)(c.heap[root] + kOffMailbox);
      xg_post(&root_mb->val[slot][me], to_bits(a), seq);
    }
  }
  if (me != root && !p.xg_all) return;
  __syncthreads();
  if (s_fail) return;
  seq = s_seq;
  slot = seq % (unsigned)kXgSlots;
  if (me == root) {
    if (threadIdx.x < (unsigned)P) {
      unsigned long long bits = 0;
      if (xg_wait(&my_mb->val[slot][threadIdx.x], seq, c, &bits)) s_vals[threadIdx.x] = bits;
      else s_fail = 1;
    }
    __syncthreads();
    if (s_fail) return;
    if (threadIdx.x == 0) {
      acc_t acc = from_bits(s_vals[0]);
      for (int r = 1; r < P; ++r) acc = F::apply(acc, from_bits(s_vals[r]));
      store_result<R>(p, acc);
      s_res = to_bits(acc);
    }
    __syncthreads();
    if (threadIdx.x < (unsigned)P) {
      XgMailbox* mb = (XgMailbox*)(c.heap[threadIdx.x] + kOffMailbox);
      if (p.xg_all) xg_post(&mb->result[slot], s_res, seq);
      // The acknowledgement only allows rank r to OVERWRITE its slot kXgSlots calls from now. The reads of the slots were
      // performed before the two bar.syncs above (a polling loop cannot leave before its load has returned), so a plain
      // system-scope store is enough; a release here is a MEMBAR.SYS on the exit path of every root launch (measured: the
      // root's last CTA spent 4-9 us between its combine and its exit with it, see profiles/multi_gpu.md).
      st_relaxed_sys_u32(&mb->ack, seq);
    }
  } else if (threadIdx.x == 0) {
    unsigned long long bits = 0;
    if (xg_wait(&my_mb->result[slot], seq, c, &bits)) store_result<R>(p, from_bits(bits));
  }
}

// `v` is valid in thread 0; every thread of the CTA calls this.
template <class R>
__device__ __forceinline__ void finish_scalar(const LocalParams& p, typename R::acc_t v) {
  if (p.xg_ctx) xgpu_finish<R>(p, v);
  else if (threadIdx.x == 0) store_result<R>(p, v);
}

// int32 SUM only: the grid stage in ONE atomic. The accumulator word carries the arrival count in its top 16 bits and the
// running sum of the partials (each taken as an unsigned 32-bit value; at most 2^16 of them, so the sum stays below 2^48 and
// never carries into the count) below. The returned old value tells the last arriver both that it is last and what the
// others added up to; the wrapped 32-bit total is the low word. Integer addition is exact in any order, so this is still
// bit-exact. It replaces store-partial / ticket / re-read all partials / second block reduce: one L2 round trip instead of
// three on the exit path of every launch. The values travel inside the atomic itself, so it can be relaxed.
__device__ __forceinline__ unsigned long long atom_add_relaxed_gpu_u64(unsigned long long* p, unsigned long long v) {
  unsigned long long old;
  asm volatile("atom.relaxed.gpu.global.add.u64 %0, [%1], %2;" : "=l"(old) : "l"(p), "l"(v) : "memory");
  return old;
}
template <class R>
__device__ __forceinline__ void finish_grid_packed(const LocalParams& p, unsigned narrivals, int v) {
  __shared__ int s_last, s_total;
  if (threadIdx.x == 0) {
    const unsigned long long mine = (1ull << 48) | (unsigned long long)(unsigned int)v;
    const unsigned long long old = atom_add_relaxed_gpu_u64(p.acc64, mine);
    s_last = (unsigned)(old >> 48) == narrivals - 1;
    if (s_last) {
      s_total = (int)(unsigned int)(old + mine);
      *p.acc64 = 0;  // re-arm for the next (stream-ordered) launch
      if (p.tile_ctr) *p.tile_ctr = 0;
    }
  }
  __syncthreads();
  if (s_last) finish_scalar<R>(p, (typename R::acc_t)s_total);
  if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 3] = globaltimer_ns();
}

// Safe programmatic launch (pdl == 1): the CTA is resident before the previous kernel of the stream has finished and has
// nothing to do until griddepcontrol.wait returns. Pulling the head of its slice into L2 meanwhile is harmless whatever the
// previous kernel is doing to the buffer — L2 is the coherence point of device memory, a line prefetched early simply
// receives the producer's later writes — and it puts the HBM channels to work during the previous kernel's exit path.
__device__ __forceinline__ void prefetch_l2_range(const char* base, uint64_t bytes, unsigned lines_per_thread) {
  for (unsigned j = 0; j < lines_per_thread; ++j) {
    const uint64_t off = ((uint64_t)threadIdx.x + (uint64_t)j * blockDim.x) * 128ull;
    if (off < bytes) asm volatile("prefetch.global.L2 [%0];" ::"l"(base + off));
  }
}

// Grid stage: every CTA (or cluster leader) that calls this has already stored its partial(s); it takes a ticket with
// atom.acq_rel.gpu (release ours, acquire everybody else's if we turn out to be last), and the last arriver re-reduces
// partials[0, nparts) in index order, re-arms the ticket (and the dynamic-tile counter) and finishes the scalar.
template <class R>
__device__ __forceinline__ void finish_grid(const LocalParams& p, unsigned nparts, unsigned narrivals, typename R::acc_t* s_scratch) {
  using acc_t = typename R::acc_t;
  using F = typename R::F;
  __shared__ int s_is_last;
  acc_t* partials = (acc_t*)p.partials;
  if (threadIdx.x == 0) {
    unsigned t = atom_add_acq_rel_gpu(p.ticket, 1u);
    s_is_last = (t == narrivals - 1);
  }
  __syncthreads();
  if (!s_is_last) {
    if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 3] = globaltimer_ns();
    return;
  }
  acc_t a = F::identity();
  for (unsigned i = threadIdx.x; i < nparts; i += blockDim.x) a = F::apply(a, ld_cg(partials + i));
  a = block_reduce<R>(a, s_scratch);
  if (threadIdx.x == 0) {  // re-arm for the next (stream-ordered) launch
    *p.ticket = 0;
    if (p.tile_ctr) *p.tile_ctr = 0;
  }
  finish_scalar<R>(p, a);
  if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 3] = globaltimer_ns();
}

// Everything after the streaming loop. `tv` is the per-thread value.
template <class R, bool CLUSTER>
__device__ __forceinline__ void finish_block(typename R::acc_t tv, const LocalParams& p) {
  using acc_t = typename R::acc_t;
  using F = typename R::F;
  __shared__ acc_t s_scratch[32];
  __shared__ acc_t s_cluster[8];

  acc_t v = block_reduce<R>(tv, s_scratch);
  pdl_wait_prior(p);
  if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 2] = globaltimer_ns();

  unsigned nparts = gridDim.x, part = blockIdx.x;
  if (CLUSTER) {
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned cr = cluster.block_rank(), cs = cluster.num_blocks();
    if (threadIdx.x == 0) {
      acc_t* leader = cluster.map_shared_rank(s_cluster, 0);  // DSMEM address in the leader CTA
      leader[cr] = v;
    }
    cluster.sync();  // barrier.cluster.arrive.release + wait.acquire
    if (cr != 0) return;
    if (threadIdx.x == 0) {
      v = s_cluster[0];
      for (unsigned i = 1; i < cs; ++i) v = F::apply(v, s_cluster[i]);
    }
    nparts = gridDim.x / cs;
    part = blockIdx.x / cs;
  }

  if (nparts == 1) {
    finish_scalar<R>(p, v);
    if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 3] = globaltimer_ns();
    return;
  }
  if constexpr (R::kPackedSum) {
    if (p.acc64) {
      finish_grid_packed<R>(p, nparts, (int)v);
      return;
    }
  }
  if (threadIdx.x == 0) ((acc_t*)p.partials)[part] = v;
  finish_grid<R>(p, nparts, nparts, s_scratch);
}

// Head/tail scalars (fewer than one vector each) are folded in by CTA 0.
template <class R>
__device__ __forceinline__ void add_edges(typename R::Thread& acc, const LocalParams& p, uint64_t body_elems) {
  if (blockIdx.x != 0) return;
  const uint64_t edge = p.head + p.tail;
  for (uint64_t i = threadIdx.x; i < edge; i += blockDim.x) {
    const uint64_t idx = i < p.head ? i : p.head + body_elems + (i - p.head);
    acc.add_elem(p.in, idx);
  }
}

// ------------------------------------------------------------------------------------------------ LDG variant
// One pass over vectors [lo, hi) of the body: UNROLL independent vector loads are issued back to back, then consumed
// (UNROLL * sizeof(Vec) bytes in flight per thread).
template <class R, class Vec, int VECW, int UNROLL>
__device__ __forceinline__ void stream_range(typename R::Thread& acc, const Vec* __restrict__ vp, uint64_t lo, uint64_t hi) {
  const unsigned nt = blockDim.x;
  uint64_t i = lo + threadIdx.x;
  for (; i + (uint64_t)(UNROLL - 1) * nt < hi; i += (uint64_t)UNROLL * nt) {
    Vec v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) v[u] = ldg_stream(vp + i + (uint64_t)u * nt);
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      acc.add16(v[u].w);
      if (VECW == 8) acc.add16(v[u].w + 4);
    }
  }
  for (; i < hi; i += nt) {
    Vec v = ldg_stream(vp + i);
    acc.add16(v.w);
    if (VECW == 8) acc.add16(v.w + 4);
  }
}

// DYN = false: every CTA owns one contiguous slice (balanced static split) and that is all.
// DYN = true : the static split covers only vectors [0, dyn_first); the rest is cut into tiles that CTAs claim with an atomic
//   counter as they run out of work, so CTAs that were scheduled late or sit behind a slow HBM channel do not set the kernel's
//   duration (trace of the static version at 64 MiB: streaming ended between 9.2 and 12.3 us across CTAs). Most of the input
//   still streams without any block-wide barrier; only the tail pays one bar.sync per tile, and the next tile index is always
//   requested one tile ahead so the atomic's latency is hidden.
//   Determinism: integer results do not depend on who adds what, so an integer CTA keeps accumulating in its registers. A
//   floating-point CTA writes one partial for its static slice and one partial PER TILE (partials[G + tile]); the last
//   arriver folds all of them in index order, so the bits do not depend on which CTA happened to claim which tile.
template <int D, int O, int VECW, int UNROLL, bool CLUSTER, bool DYN>
__device__ __forceinline__ void reduce_ldg_body(const LocalParams& p) {
  using R = Reducer<D, O>;
  using acc_t = typename R::acc_t;
  using Vec = typename std::conditional<VECW == 8, V32, V16>::type;
  constexpr unsigned kAlign = 128 / sizeof(Vec);

  pdl_launch_dependents(p);
  const Vec* __restrict__ vp = (const Vec*)((const char*)p.in + p.head * R::kElemBytes);
  const uint64_t nstatic = DYN ? p.dyn_first : p.nvec;
  const uint64_t lo = split_point(nstatic, gridDim.x, blockIdx.x, kAlign);
  const uint64_t hi = split_point(nstatic, gridDim.x, blockIdx.x + 1, kAlign);
  if (p.pdl == 1 && p.prefetch_lines) prefetch_l2_range((const char*)(vp + lo), (hi - lo) * sizeof(Vec), p.prefetch_lines);
  pdl_wait_before_input(p);
  if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 0] = globaltimer_ns();

  typename R::Thread acc;
  acc.init();

  if (!DYN) {
    stream_range<R, Vec, VECW, UNROLL>(acc, vp, lo, hi);
    add_edges<R>(acc, p, p.nvec * (sizeof(Vec) / R::kElemBytes));
    if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 1] = globaltimer_ns();
    finish_block<R, CLUSTER>(acc.collapse(), p);
    return;
  }

  __shared__ acc_t s_scratch[32];
  __shared__ unsigned s_next[2];
  const unsigned G = gridDim.x, ntiles = p.dyn_ntiles;
  // claim the first tile now: the answer is needed only after the static slice
  if (threadIdx.x == 0) s_next[0] = atomicAdd(p.tile_ctr, 1u);
  stream_range<R, Vec, VECW, UNROLL>(acc, vp, lo, hi);
  add_edges<R>(acc, p, p.nvec * (sizeof(Vec) / R::kElemBytes));
  acc_t* partials = (acc_t*)p.partials;
  if (!R::kIsInt) {
    const acc_t v = block_reduce<R>(acc.collapse(), s_scratch);
    if (threadIdx.x == 0) partials[blockIdx.x] = v;
    acc.init();
  }
  __syncthreads();  // publishes s_next[0]
  unsigned tile = s_next[0];
  int buf = 1;
  while (tile < ntiles) {
    if (threadIdx.x == 0) s_next[buf] = atomicAdd(p.tile_ctr, 1u);
    const uint64_t tlo = p.dyn_first + (uint64_t)tile * p.dyn_tile_vecs;
    const uint64_t thi = tile + 1 == ntiles ? p.nvec : tlo + (uint64_t)p.dyn_tile_vecs;  // the last tile takes the alignment slack
    stream_range<R, Vec, VECW, UNROLL>(acc, vp, tlo, thi);
    if (!R::kIsInt) {
      const acc_t v = block_reduce<R>(acc.collapse(), s_scratch);
      if (threadIdx.x == 0) partials[G + tile] = v;
      acc.init();
    }
    __syncthreads();
    tile = s_next[buf];
    buf ^= 1;
  }
  if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 1] = globaltimer_ns();
  unsigned nparts = G + ntiles;
  if (R::kIsInt) {
    const acc_t v = block_reduce<R>(acc.collapse(), s_scratch);
    if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 2] = globaltimer_ns();
    if constexpr (R::kPackedSum) {
      if (p.acc64) {
        finish_grid_packed<R>(p, G, (int)v);
        return;
      }
    }
    if (threadIdx.x == 0) partials[blockIdx.x] = v;
    nparts = G;
  } else if (p.trace && threadIdx.x == 0) {
    p.trace[4 * blockIdx.x + 2] = globaltimer_ns();
  }
  finish_grid<R>(p, nparts, G, s_scratch);
}

template <int D, int O, int VECW, int UNROLL, bool CLUSTER>
__global__ void __launch_bounds__(512) reduce_ldg_kernel(const __grid_constant__ LocalParams p) {
  reduce_ldg_body<D, O, VECW, UNROLL, CLUSTER, false>(p);
}

// The same streaming loop with the dynamically scheduled tail (plain stream order, no cluster).
template <int D, int O, int VECW, int UNROLL>
__global__ void __launch_bounds__(512) reduce_ldg_dyn_kernel(const __grid_constant__ LocalParams p) {
  reduce_ldg_body<D, O, VECW, UNROLL, false, true>(p);
}

// Loopback flavour: gridDim.y virtual ranks in ONE (cooperative) launch, blockIdx.y picks the rank's parameters, so the
// cross-rank mailbox protocol of xgpu_finish can be exercised on a single GPU without co-scheduling separate kernels.
constexpr int kMultiMaxRanks = 8;
struct MultiLocalParams {
  LocalParams p[kMultiMaxRanks];
};
template <int D, int O>
__global__ void __launch_bounds__(512) reduce_ldg_multi_kernel(const __grid_constant__ MultiLocalParams mp) {
  reduce_ldg_body<D, O, 4, 4, false, false>(mp.p[blockIdx.y]);
}

// ------------------------------------------------------------------------------------------------ TMA (bulk-copy) variant
// Warp 0 is the producer: one elected lane keeps `stages` cp.async.bulk transfers in flight into a shared-memory
// ring (full/empty mbarriers). Warps 1.. are consumers: conflict-free LDS.128 + register accumulate. No register
// is tied up by data in flight, so few resident threads suffice to cover the HBM latency-bandwidth product.
template <int D, int O, bool CLUSTER>
__global__ void __launch_bounds__(512) reduce_tma_kernel(const __grid_constant__ LocalParams p) {
  using R = Reducer<D, O>;
  extern __shared__ __align__(128) unsigned char smem[];

  pdl_launch_dependents(p);
  pdl_wait_before_input(p);
  if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 0] = globaltimer_ns();

  const int stages = p.stages;
  const uint32_t tile_vecs = (uint32_t)p.tile_vecs;
  const uint32_t tile_bytes = tile_vecs * 16u;
  const uint32_t ring = smem_u32(smem);
  const uint32_t bars = ring + (uint32_t)stages * tile_bytes;  // full[stages], empty[stages]
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned ncw = (blockDim.x >> 5) - 1;  // consumer warps

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(bars + 8u * s, 1u);               // producer's arrive.expect_tx
      mbar_init(bars + 8u * (stages + s), ncw);   // one arrive per consumer warp
    }
    mbar_fence_init();
  }
  __syncthreads();

  typename R::Thread acc;
  acc.init();

  const V16* __restrict__ vp = (const V16*)((const char*)p.in + p.head * R::kElemBytes);
  const uint64_t lo = split_point(p.nvec, gridDim.x, blockIdx.x, 8);
  const uint64_t hi = split_point(p.nvec, gridDim.x, blockIdx.x + 1, 8);
  const uint64_t nv = hi - lo;
  const uint64_t ntiles = (nv + tile_vecs - 1) / tile_vecs;

  if (warp == 0) {
    if (lane == 0) {
      int s = 0;
      uint32_t k = 0;  // how many times the ring has wrapped
      for (uint64_t t = 0; t < ntiles; ++t) {
        mbar_wait(bars + 8u * (stages + s), (k & 1u) ^ 1u);  // slot free (passes immediately on first use)
        const uint64_t off = t * tile_vecs;
        const uint32_t vecs = (uint32_t)min((uint64_t)tile_vecs, nv - off);
        mbar_arrive_expect_tx(bars + 8u * s, vecs * 16u);
        bulk_g2s(ring + (uint32_t)s * tile_bytes, vp + lo + off, vecs * 16u, bars + 8u * s);
        if (++s == stages) { s = 0; ++k; }
      }
    }
    __syncwarp();
  } else {
    const unsigned ct = threadIdx.x - 32, nct = blockDim.x - 32;
    int s = 0;
    uint32_t k = 0;
    for (uint64_t t = 0; t < ntiles; ++t) {
      mbar_wait(bars + 8u * s, k & 1u);  // bytes have landed
      const uint64_t off = t * tile_vecs;
      const uint32_t vecs = (uint32_t)min((uint64_t)tile_vecs, nv - off);
      const uint32_t base = ring + (uint32_t)s * tile_bytes;
      if (vecs == tile_vecs) {
        // full tile: kTmaVecsPerThread conflict-free LDS.128 issued back to back, then consumed
        V16 v[kTmaVecsPerThread];
#pragma unroll
        for (int u = 0; u < kTmaVecsPerThread; ++u) v[u] = lds128(base + (ct + (uint32_t)u * nct) * 16u);
#pragma unroll
        for (int u = 0; u < kTmaVecsPerThread; ++u) acc.add16(v[u].w);
      } else {
        for (uint32_t j = ct; j < vecs; j += nct) {
          V16 v = lds128(base + j * 16u);
          acc.add16(v.w);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bars + 8u * (stages + s));  // slot may be refilled
      if (++s == stages) { s = 0; ++k; }
    }
  }
  add_edges<R>(acc, p, p.nvec * (16 / R::kElemBytes));
  if (p.trace && threadIdx.x == 0) p.trace[4 * blockIdx.x + 1] = globaltimer_ns();

  finish_block<R, CLUSTER>(acc.collapse(), p);
}

// ------------------------------------------------------------------------------------------------ host-side launcher (per dtype TU)
struct LaunchGeom {
  int variant;   // b2r_variant_t (resolved, never AUTO)
  int threads;
  int ctas;
  int cluster;   // >= 1
  int unroll;    // 4 or 8
  int smem_bytes;
  int pdl;       // 0 plain, 1 programmatic launch (wait before input), 2 programmatic launch with early streaming
  int dyn_pct;   // share of the body (percent) scheduled dynamically at the end; 0 = static split only
};

template <int D>
b2r_status_t launch_local(int op, const LaunchGeom& g, const LocalParams& p, cudaStream_t stream);
template <int D>
b2r_status_t launch_local_multi(int op, int nranks, int ctas, int threads, const MultiLocalParams& mp, cudaStream_t stream);

}  // namespace b2r
