// coll_push.cuh — "push": scatter by stores, combine locally, store the result. Two hops, nothing ever loaded from a peer.
//
// Why it exists. On 8 B200s the NVSwitch's in-switch reduction (multimem.ld_reduce) tops out near 520 GB/s of reduced output
// whatever the launch shape, a store-forwarding ring pays its per-hop latency seven times (pipeline fill), and pulls (two-shot)
// plateau near 400 GB/s; posted STORES run closest to link rate, and an all-to-all of stores is what an NVSwitch is built
// for. So the message is cut into one slice per worker (reduce: the P-1 non-root ranks; allreduce: all P ranks) and
//   1. every rank stores slice w of ITS send buffer into worker w's staging area [source rank][offset in slice];
//   2. worker w combines its own slice with the P-1 staged ones in rank order (all local reads — the bits equal the
//      one-shot / two-shot kernels' and the host restatement's) and stores the result into the root's (allreduce:
//      everybody's) recv buffer.
// Per GPU and direction: one message for reduce, 2(P-1)/P of one for allreduce — the same as the best pull schedule, all of it
// stores, and only two hops deep.
//
// How it keeps the link busy. Work is cut into chunks of (threads x 4) vectors; CTA b owns chunks b, b+G, ... of EVERY slice. Per
// chunk the CTA pushes its piece to every worker, release-stores one flag per worker (value = call number, location =
// [source][chunk] in the worker's control region), and then — if this rank is a worker — combines the chunk it pushed one
// iteration earlier, whose seven peers' flags have had a whole iteration to arrive. The staging area holds whole slices, so
// there is no slot reuse inside a call, no acknowledgements and no flow control: CTAs never wait for one another across ranks
// except for data that is already on its way, and they drift apart naturally, so while one CTA sits in its release fence the
// others keep issuing. (The first version of this algorithm recycled a small staging ring with acks per 32 KiB and ran in
// lockstep rounds: 243 GB/s at 64 MiB on 8 GPUs.)
// Re-use across calls is safe because the end barrier is all-to-all per CTA index: when a rank's kernel has finished, every
// peer CTA has passed its last combine. In place (send == recv) is safe: a piece of recv is written only by the worker that
// has already received — or, for its own slice, already read — the matching piece of send.
#pragma once

namespace b2r {

// flag [source rank][chunk] of the call number in which that chunk of that source's contribution landed here
__device__ __forceinline__ unsigned int* push_flag(char* heap, int src, unsigned long long chunk) {
  return (unsigned int*)(heap + kOffPushFlags) + (size_t)src * kPushChunkCap + chunk;
}

template <int D, int O, int UNR>
__device__ void push_body(const CollArgs& A) {
  using V = VecOps<D, O>;
  const CollCtx& c = A.ctx;
  constexpr int E = V::kElem;
  constexpr int VPT = 4;  // vectors per thread per chunk
  const int P = c.nranks, r = c.rank, G = (int)gridDim.x, b = (int)blockIdx.x;
  const int W = A.root < 0 ? P : P - 1;
  const int my_w = A.root < 0 ? r : (r == A.root ? -1 : (r < A.root ? r : r - 1));
  const unsigned long long nvec = A.count * E / 16;
  const unsigned int Cv = blockDim.x * VPT;                                  // vectors per chunk
  const unsigned long long slice_cap = (nvec + W - 1) / W + 1;               // vectors reserved per (worker, source) in staging
  const unsigned long long nchunks = (slice_cap + Cv - 1) / Cv;              // chunks per slice (same on every rank; host: <= kPushChunkCap)
  const unsigned int seq = A.seq;

  __shared__ PeerOffsets po;
  __shared__ char* s_dst[B2R_MAX_RANKS];
  start_barrier(A, &po);
  if (threadIdx.x < (unsigned)P) s_dst[threadIdx.x] = c.heap[threadIdx.x] + po.recv[threadIdx.x];
  __syncthreads();

  auto worker_rank = [&](int w) { return A.root < 0 ? w : (w < A.root ? w : w + 1); };
  auto slice_lo = [&](int w) { return nvec * (unsigned long long)w / W; };
  const char* send = c.heap[r] + A.send_off;
  // staging of worker-rank `wr` for source `src`: whole slices, laid out [src][vector]
  auto stage = [&](int wr, int src) { return c.heap[wr] + A.work_off + (unsigned long long)src * slice_cap * 16ull; };

  // ---- combine chunk k of my slice (workers only): wait for the P-1 peers' flags, fold in rank order, store the result
  auto combine_chunk = [&](unsigned long long k) {
    if (threadIdx.x < (unsigned)P && (int)threadIdx.x != r)
      wait_flag_ge(push_flag(c.heap[r], (int)threadIdx.x, k), seq, c);
    __syncthreads();
    const unsigned long long s0 = slice_lo(my_w), s1 = slice_lo(my_w + 1);
    const unsigned long long off = k * Cv;
    if (s0 + off < s1) {
      const unsigned int len = (unsigned int)min((unsigned long long)Cv, s1 - s0 - off);
      auto src_ptr = [&](int sidx) { return sidx == r ? send + (s0 + off) * 16 : stage(r, sidx) + off * 16; };
      typename V::Acc acc[VPT];
      V16 cur[VPT], nxt[VPT];
      {
        const char* sp = src_ptr(0);
#pragma unroll
        for (int u = 0; u < VPT; ++u) {
          const unsigned int j = threadIdx.x + u * blockDim.x;
          if (j < len) cur[u] = ld_peer(sp + (unsigned long long)j * 16);
        }
#pragma unroll
        for (int u = 0; u < VPT; ++u) acc[u] = V::unpack(cur[u]);
        sp = src_ptr(1 < P ? 1 : 0);
#pragma unroll
        for (int u = 0; u < VPT; ++u) {
          const unsigned int j = threadIdx.x + u * blockDim.x;
          if (j < len) cur[u] = ld_peer(sp + (unsigned long long)j * 16);
        }
      }
      for (int sidx = 1; sidx < P; ++sidx) {
        if (sidx + 1 < P) {
          const char* sp = src_ptr(sidx + 1);
#pragma unroll
          for (int u = 0; u < VPT; ++u) {
            const unsigned int j = threadIdx.x + u * blockDim.x;
            if (j < len) nxt[u] = ld_peer(sp + (unsigned long long)j * 16);
          }
        }
#pragma unroll
        for (int u = 0; u < VPT; ++u) V::combine(acc[u], cur[u]);
#pragma unroll
        for (int u = 0; u < VPT; ++u) cur[u] = nxt[u];
      }
#pragma unroll
      for (int u = 0; u < VPT; ++u) {
        const unsigned int j = threadIdx.x + u * blockDim.x;
        if (j < len) {
          const V16 res = V::pack(acc[u]);
          const unsigned long long at = (s0 + off + j) * 16;
          if (A.root < 0) {
            for (int p = 0; p < P; ++p) st_peer(s_dst[p] + at, res);
          } else {
            st_peer(s_dst[A.root] + at, res);
          }
        }
      }
    }
  };

  bool have_prev = false;
  unsigned long long prev = 0;
  for (unsigned long long k = (unsigned long long)b; k < nchunks; k += (unsigned long long)G) {
    // ---- 1. push my piece of chunk k to every worker; the loads for worker w+1 are in flight while w is being stored
    const unsigned long long off = k * Cv;
    V16 v[VPT], vn[VPT];
    auto piece_len = [&](int w) -> unsigned int {
      const unsigned long long s0 = slice_lo(w), s1 = slice_lo(w + 1);
      return (worker_rank(w) == r || s0 + off >= s1) ? 0u : (unsigned int)min((unsigned long long)Cv, s1 - s0 - off);
    };
    auto load_piece = [&](int w, unsigned int len, V16* dstv) {
      const char* sp = send + (slice_lo(w) + off) * 16;
#pragma unroll
      for (int u = 0; u < VPT; ++u) {
        const unsigned int j = threadIdx.x + u * blockDim.x;
        if (j < len) dstv[u] = ld_peer(sp + (unsigned long long)j * 16);
      }
    };
    unsigned int len = piece_len(0);
    load_piece(0, len, v);
    for (int w = 0; w < W; ++w) {
      const unsigned int len_n = w + 1 < W ? piece_len(w + 1) : 0u;
      if (w + 1 < W) load_piece(w + 1, len_n, vn);
      if (len) {
        char* dp = stage(worker_rank(w), r) + off * 16;
#pragma unroll
        for (int u = 0; u < VPT; ++u) {
          const unsigned int j = threadIdx.x + u * blockDim.x;
          if (j < len) st_peer(dp + (unsigned long long)j * 16, v[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < VPT; ++u) v[u] = vn[u];
      len = len_n;
    }
    __syncthreads();
    // ---- 2. one flag per worker: "chunk k of rank r has landed" (the release also publishes the results stored by the
    //         previous iteration's combine)
    if (threadIdx.x < (unsigned)W && !c.fault_skip_signal) {
      const int wr = worker_rank((int)threadIdx.x);
      if (wr != r) st_release_sys(push_flag(c.heap[wr], r, k), seq);
    }
    // ---- 3. worker: combine the chunk pushed one iteration ago (its peers' pieces have had an iteration to arrive)
    if (my_w >= 0 && have_prev) combine_chunk(prev);
    prev = k;
    have_prev = true;
  }
  if (my_w >= 0 && have_prev) combine_chunk(prev);

  // the (< 16 byte) tail: pulled by the last worker
  if (my_w == W - 1 && nvec * (16 / E) < A.count) {
    __shared__ const char* src[B2R_MAX_RANKS];
    __shared__ char* dst[B2R_MAX_RANKS];
    const int ndst = A.root < 0 ? P : 1;
    if (threadIdx.x == 0) {
      for (int p = 0; p < P; ++p) src[p] = c.heap[p] + po.send[p];
      if (A.root < 0) for (int p = 0; p < P; ++p) dst[p] = s_dst[p];
      else dst[0] = s_dst[A.root];
    }
    __syncthreads();
    combine_tail<D, O>(nvec * (16 / E), A.count, src, P, dst, ndst);
  }
  cta_barrier(A, A.epoch + 2);
}

}  // namespace b2r
