// coll_posted.cuh — "posted" reduce-to-root: every rank stages its contribution in its OWN heap and returns; the root alone
// reduces the staged copies through the switch.
//
// Why it exists. Between 512 KiB and a few MiB the barrier-framed switch reduction (nvls_body) is all overhead: two NVLink
// round trips frame 2-20 us of data movement, and every call of a back-to-back series pays them again because no rank may
// leave before the root has read its send buffer (8 GPUs, 1 MiB: 18 us where NCCL's pipelined chain needs 12.6). Here the
// send buffer is copied (16-byte vectors, local HBM/L2 traffic only) into one of two staging slots of the work region, a
// per-CTA ready word is released to the root, and the rank is done: its next call — posted or not — can start while the root is
// still working on this one. The root waits for the ready words (in a series they arrived long ago), issues
// multimem.ld_reduce over the slot (one instruction fetches and combines all P copies), writes its recv buffer locally and
// acknowledges. No barrier at either end; the only cross-GPU wait on a non-root rank is for the acknowledgement of the call
// two back (same slot), which is off the critical path.
//   partition : FIXED. gridDim.x is the same in every call on every rank; chunk k (kPostedChunkVecs vectors) belongs to CTA
//               k % gridDim.x. So slot bytes are only ever touched by the same CTA index, and one word per (CTA, source) /
//               (root, CTA) carries the whole protocol, with call numbers that only grow.
//   roots     : may change from call to call; the acknowledgements a rank waits for are those of the ranks that read its slot
//               in the call two back (the host passes them in as a bit mask), and words are indexed by the acknowledging rank.
//   fp min/max: have no multimem form, but signed-integer min/max do, and IEEE floats order like sign-magnitude integers.
//               Staging maps every value to the two's-complement integer of the same rank (k = b ^ ((b >> 31) & 0x7fff...)), the
//               switch reduces integers, the root maps back (the map is its own inverse). -0 orders below +0; a NaN orders
//               beyond the infinity of its sign (the host restatement and the pull kernels use a < b ? a : b).
//   no switch : (loopback groups, or a (dtype, op) pair without any multimem form) the root folds the staged copies in rank
//               order with peer loads — same protocol, same bits as the one-shot kernel; this is what the loopback tests
//               exercise. AUTO only ever picks the algorithm where the switch does the reading.
#pragma once

namespace b2r {

__device__ __forceinline__ unsigned int* posted_ready(char* heap, int cta, int src) {
  return (unsigned int*)(heap + kOffPostedReady) + (size_t)cta * B2R_MAX_RANKS + src;
}
__device__ __forceinline__ unsigned int* posted_ack(char* heap, int acker, int cta) {
  return (unsigned int*)(heap + kOffPostedAck) + (size_t)acker * kPostedCtas + cta;
}

// 64-bit signed min/max through the switch (the keys of fp64 min/max)
template <int O>
struct NvlsS64 {
  __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) {
    long long d;
    if (O == B2R_MIN) asm volatile("multimem.ld_reduce.relaxed.sys.global.min.s64 %0, [%1];" : "=l"(d) : "l"(p) : "memory");
    else asm volatile("multimem.ld_reduce.relaxed.sys.global.max.s64 %0, [%1];" : "=l"(d) : "l"(p) : "memory");
    w[0] = (uint32_t)((unsigned long long)d & 0xffffffffull);
    w[1] = (uint32_t)((unsigned long long)d >> 32);
  }
};

// How a (dtype, op) pair travels: which multimem form reads the slots, and what staging does to the bits.
template <int D, int O>
struct PostedCodec {
  static constexpr bool kSwitch = Nvls<D, O>::kSupported;
  static constexpr int kBytes = Nvls<D, O>::kBytes;
  __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) { Nvls<D, O>::ld_reduce(p, w); }
  __device__ __forceinline__ static V16 code(const V16& v) { return v; }
};
template <int O>
struct PostedKey32 {
  static constexpr bool kSwitch = true;
  static constexpr int kBytes = 4;
  __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) { Nvls<B2R_INT32, O>::ld_reduce(p, w); }
  __device__ __forceinline__ static V16 code(const V16& v) {
    V16 k;
#pragma unroll
    for (int j = 0; j < 4; ++j) k.w[j] = v.w[j] ^ ((uint32_t)((int32_t)v.w[j] >> 31) & 0x7fffffffu);
    return k;
  }
};
template <int O>
struct PostedKey64 {
  static constexpr bool kSwitch = true;
  static constexpr int kBytes = 8;
  __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) { NvlsS64<O>::ld_reduce(p, w); }
  __device__ __forceinline__ static V16 code(const V16& v) {  // little endian: w[1], w[3] hold the sign and the high bits
    V16 k;
#pragma unroll
    for (int j = 0; j < 4; j += 2) {
      const uint32_t m = (uint32_t)((int32_t)v.w[j + 1] >> 31);
      k.w[j] = v.w[j] ^ m;
      k.w[j + 1] = v.w[j + 1] ^ (m & 0x7fffffffu);
    }
    return k;
  }
};
template <> struct PostedCodec<B2R_FLOAT32, B2R_MIN> : PostedKey32<B2R_MIN> {};
template <> struct PostedCodec<B2R_FLOAT32, B2R_MAX> : PostedKey32<B2R_MAX> {};
template <> struct PostedCodec<B2R_FLOAT64, B2R_MIN> : PostedKey64<B2R_MIN> {};
template <> struct PostedCodec<B2R_FLOAT64, B2R_MAX> : PostedKey64<B2R_MAX> {};

template <int D, int O>
__device__ void posted_body(const CollArgs& A) {
  using V = VecOps<D, O>;
  using C = PostedCodec<D, O>;
  const CollCtx& c = A.ctx;
  constexpr int E = V::kElem;
  constexpr int U = 4;  // vectors in flight per thread
  const int P = c.nranks, me = c.rank, root = A.root, b = (int)blockIdx.x, G = (int)gridDim.x;
  const unsigned int seq = A.seq;
  const unsigned int ack_mask = (unsigned int)A.sub_blocks;  // ranks that read this parity's slots in call seq-2 (bit p)
  const unsigned long long bytes = A.count * E;
  const unsigned long long nfull = bytes / 16;
  const unsigned int tail_bytes = (unsigned int)(bytes % 16);
  const unsigned long long nvec = nfull + (tail_bytes ? 1ull : 0ull);
  const unsigned long long slot_off = A.work_off + (unsigned long long)(seq & 1u) * A.recv_off;  // recv_off carries the slot stride
  const bool via_switch = C::kSwitch && c.mc != nullptr;
  const bool has_work = (unsigned long long)b * kPostedChunkVecs < nvec;
  char* my_slot = c.heap[me] + slot_off;

  auto load_send = [&](unsigned long long v) -> V16 {
    return v < nfull ? ld_peer(A.send_ptr + v * 16) : ll_load_tail<E>(A.send_ptr + nfull * 16, tail_bytes);
  };

  // Work of CTA b: chunks b, b+G, b+2G, ...; a thread handles vector j of a chunk (j = threadIdx.x when the CTA has
  // kPostedChunkVecs threads) and keeps the same vector of U consecutive owned chunks in flight.
  const unsigned long long nchunks = (nvec + kPostedChunkVecs - 1) / kPostedChunkVecs;
  auto vec_of = [&](unsigned long long k0, int u, unsigned int j) -> unsigned long long {  // ~0 = nothing there
    const unsigned long long k = k0 + (unsigned long long)u * G;
    const unsigned long long v = k * kPostedChunkVecs + j;
    return (k < nchunks && v < nvec) ? v : ~0ull;
  };

  if (threadIdx.x == 0) fault_delay(c);
  // ---- stage my contribution (a root that pulls reads its own send buffer directly and stages nothing)
  if (has_work && (me != root || via_switch)) {
    if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me && ((ack_mask >> threadIdx.x) & 1u))
      wait_flag_ge_relaxed(posted_ack(c.heap[me], (int)threadIdx.x, b), seq - 2, c);
    __syncthreads();
    for (unsigned long long k0 = (unsigned long long)b; k0 < nchunks; k0 += (unsigned long long)G * U) {
      for (unsigned int j = threadIdx.x; j < kPostedChunkVecs; j += blockDim.x) {
        V16 x[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long v = vec_of(k0, u, j);
          if (v != ~0ull) x[u] = load_send(v);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long v = vec_of(k0, u, j);
          if (v != ~0ull) st_peer(my_slot + v * 16, via_switch ? C::code(x[u]) : x[u]);
        }
      }
    }
    __syncthreads();
    if (me != root) {
      if (threadIdx.x == 0 && !c.fault_skip_signal) st_release_sys(posted_ready(c.heap[root], b, me), seq);
    } else {
      // my own staged copy is read back through the switch: make it visible at system scope first
      if (threadIdx.x == 0) __threadfence_system();
      __syncthreads();
    }
  }
  if (me != root) return;

  // ---- the root: wait for every peer's CTA b, reduce the chunks of CTA b, acknowledge
  if (has_work) {
    if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me) wait_flag_ge(posted_ready(c.heap[me], b, (int)threadIdx.x), seq, c);
    __syncthreads();
    const char* mslot = via_switch ? c.mc + slot_off : nullptr;
    for (unsigned long long k0 = (unsigned long long)b; k0 < nchunks; k0 += (unsigned long long)G * U) {
      for (unsigned int j = threadIdx.x; j < kPostedChunkVecs; j += blockDim.x) {
        V16 r[U];
        if (via_switch) {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long v = vec_of(k0, u, j);
            if (v != ~0ull) {
#pragma unroll
              for (int q = 0; q < 16 / C::kBytes; ++q) C::ld_reduce(mslot + v * 16 + q * C::kBytes, &r[u].w[q * C::kBytes / 4]);
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) r[u] = C::code(r[u]);
        } else {
          typename V::Acc acc[U];
          for (int p = 0; p < P; ++p) {
            V16 x[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const unsigned long long v = vec_of(k0, u, j);
              if (v != ~0ull) x[u] = p == me ? load_send(v) : ld_peer(c.heap[p] + slot_off + v * 16);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
              if (p == 0) acc[u] = V::unpack(x[u]);
              else V::combine(acc[u], x[u]);
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) r[u] = V::pack(acc[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long v = vec_of(k0, u, j);
          if (v != ~0ull) {
            if (v < nfull) st_peer(A.recv_ptr + v * 16, r[u]);
            else ll_store_tail<E>(A.recv_ptr + nfull * 16, r[u], tail_bytes);
          }
        }
      }
    }
  }
  // Every CTA acknowledges every call, with or without work, so that "ack[root][b] >= n" always means "the root's CTA b is past
  // call n". Relaxed: the loads above have returned (their values were stored), and a release here would make the root wait
  // for a system-wide fence on its exit path (measured at 4-9 us for the scalar-sharded kernel's acknowledgement).
  __syncthreads();
  if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me && !c.fault_skip_signal)
    st_relaxed_sys_u32(posted_ack(c.heap[threadIdx.x], me, b), seq);
}

// ------------------------------------------------------------------------------------------------ assisted reduce-to-root
// posted_body leaves all the multimem requests to the root, and one GPU's request stream is the limit: ~400 GB/s for the
// 16-byte forms, ~215 for the 8-byte, ~120 for the 4-byte ones (int32, fp32 min/max keys) — fine up to a MiB or a few, hopeless
// beyond. Here the P-1 non-root ranks ("helpers") do the reducing, and the root still never appears on their critical path:
//   * the root's staging slot permanently holds the IDENTITY of the operation (zeros; the largest / smallest key), so a
//     multimem.ld_reduce over all P slots returns the combination of the helpers' contributions only — the root's own data is
//     not needed until the very end;
//   * every helper stages its contribution, tells the other helpers (one release per CTA and peer), reduces its share of every
//     chunk through the switch and stores the result into a RESULT slot in the root's work region (not into recv: the root may
//     not have started this call yet), then releases a done word to the root and leaves;
//   * the root waits for the done words (in a series they are there already), combines the result slot with its own send
//     buffer — a purely local pass at HBM speed — writes recv and acknowledges.
// Helpers run in step with one another but up to two calls ahead of the root; per call the root pays a launch and a local pass,
// a helper one NVLink hop and 1/(P-1) of the switch traffic.
// The identity has to be (re)written when the root, the (dtype, op) class, or the length changes, or after anything else used
// the work region: the host keeps track and flags such a call ("clean"); then the root fills its slot first and the helpers
// also wait for the root's ready word — one hop, like a barrier. A series with a fixed root pays that once.
// Order of combination: helpers in rank order, the root's own contribution last (host restatement: host_algos.h). Without a
// switch (loopback) a helper folds the other helpers' slots with peer loads — same protocol, bit-identical to the restatement.
__device__ __forceinline__ unsigned int* assist_done(char* heap, int cta, int src) {
  return (unsigned int*)(heap + kOffAssistDone) + (size_t)cta * B2R_MAX_RANKS + src;
}

template <int D, int O>
__device__ void assist_body(const CollArgs& A) {
  using V = VecOps<D, O>;
  using C = PostedCodec<D, O>;
  const CollCtx& c = A.ctx;
  constexpr int E = V::kElem;
  constexpr int U = 4;
  constexpr unsigned int CH = kPostedChunkVecs;
  const int P = c.nranks, me = c.rank, root = A.root, b = (int)blockIdx.x, G = (int)gridDim.x;
  const int W = P - 1, my_w = me == root ? -1 : (me < root ? me : me - 1);
  const unsigned int seq = A.seq;
  const unsigned int ack_mask = (unsigned int)A.sub_blocks;  // ranks that read this parity's slots in call seq-2 (bit p)
  const bool clean = A.stagger_ns != 0;                      // the root rewrites the identity into its slot in this call
  const unsigned long long ident = A.send_off;               // 8 bytes of the identity pattern (send travels as send_ptr)
  const unsigned long long bytes = A.count * E;
  const unsigned long long nfull = bytes / 16;
  const unsigned int tail_bytes = (unsigned int)(bytes % 16);
  const unsigned long long nvec = nfull + (tail_bytes ? 1ull : 0ull);
  const unsigned long long slot_off = A.work_off + (unsigned long long)(seq & 1u) * A.recv_off;       // staged contributions
  const unsigned long long res_off = A.work_off + (unsigned long long)(2u + (seq & 1u)) * A.recv_off;  // results, at the root
  const bool via_switch = C::kSwitch && c.mc != nullptr;
  const unsigned long long nchunks = (nvec + CH - 1) / CH;
  const bool has_work = (unsigned long long)b < nchunks;
  char* my_slot = c.heap[me] + slot_off;

  auto load_send = [&](unsigned long long v) -> V16 {
    return v < nfull ? ld_peer(A.send_ptr + v * 16) : ll_load_tail<E>(A.send_ptr + nfull * 16, tail_bytes);
  };
  auto vec_of = [&](unsigned long long k0, int u, unsigned int j) -> unsigned long long {  // ~0 = nothing there
    const unsigned long long k = k0 + (unsigned long long)u * G;
    const unsigned long long v = k * CH + j;
    return (k < nchunks && v < nvec) ? v : ~0ull;
  };

  if (threadIdx.x == 0) fault_delay(c);
  // ---- whoever writes a slot of this parity (helpers: their contribution and the root's result slot; a cleaning root: its
  //      identity) first makes sure that the readers of call seq-2 are done with it
  if (has_work && (me != root || clean)) {
    if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me && ((ack_mask >> threadIdx.x) & 1u))
      wait_flag_ge_relaxed(posted_ack(c.heap[me], (int)threadIdx.x, b), seq - 2, c);
    __syncthreads();
  }

  if (me != root) {
    if (has_work) {
      // ---- 1. stage my contribution, tell the other helpers
      for (unsigned long long k0 = (unsigned long long)b; k0 < nchunks; k0 += (unsigned long long)G * U) {
        for (unsigned int j = threadIdx.x; j < CH; j += blockDim.x) {
          V16 x[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long v = vec_of(k0, u, j);
            if (v != ~0ull) x[u] = load_send(v);
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long v = vec_of(k0, u, j);
            if (v != ~0ull) st_peer(my_slot + v * 16, via_switch ? C::code(x[u]) : x[u]);
          }
        }
      }
      __syncthreads();
      if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me && (int)threadIdx.x != root && !c.fault_skip_signal)
        st_release_sys(posted_ready(c.heap[threadIdx.x], b, me), seq);
      // ---- 2. wait for the other helpers (and for a root that had to clean its slot)
      if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me && ((int)threadIdx.x != root || clean))
        wait_flag_ge(posted_ready(c.heap[me], b, (int)threadIdx.x), seq, c);
      __syncthreads();
      // ---- 3. my share of every chunk of CTA b: vectors [j_lo, j_hi) of the chunk
      const unsigned int sub = (CH + W - 1) / W;
      const unsigned int j_lo = min(CH, (unsigned int)my_w * sub), j_hi = min(CH, j_lo + sub), len = j_hi - j_lo;
      const unsigned int rounds = (unsigned int)((nchunks - b + G - 1) / G);
      const unsigned int items = rounds * len;
      const char* mslot = via_switch ? c.mc + slot_off : nullptr;
      char* res = c.heap[root] + res_off;
      auto item_vec = [&](unsigned int q) -> unsigned long long {
        if (q >= items) return ~0ull;
        const unsigned int i = q / len, j = j_lo + q % len;
        const unsigned long long v = ((unsigned long long)b + (unsigned long long)i * G) * CH + j;
        return v < nvec ? v : ~0ull;
      };
      for (unsigned int q0 = threadIdx.x; q0 < items; q0 += blockDim.x * U) {
        V16 r[U];
        if (via_switch) {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long v = item_vec(q0 + u * blockDim.x);
            if (v != ~0ull) {
#pragma unroll
              for (int q = 0; q < 16 / C::kBytes; ++q) C::ld_reduce(mslot + v * 16 + q * C::kBytes, &r[u].w[q * C::kBytes / 4]);
            }
          }
        } else {
          typename V::Acc acc[U];
          bool first = true;
          for (int p = 0; p < P; ++p) {
            if (p == root) continue;
            V16 x[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const unsigned long long v = item_vec(q0 + u * blockDim.x);
              if (v != ~0ull) x[u] = ld_peer(c.heap[p] + slot_off + v * 16);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
              if (first) acc[u] = V::unpack(x[u]);
              else V::combine(acc[u], x[u]);
            }
            first = false;
          }
#pragma unroll
          for (int u = 0; u < U; ++u) r[u] = V::pack(acc[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long v = item_vec(q0 + u * blockDim.x);
          if (v != ~0ull) st_peer(res + v * 16, r[u]);  // keys stay keys: the root decodes
        }
      }
      // ---- 4. delivered
      __syncthreads();
      if (threadIdx.x == 0 && !c.fault_skip_signal) st_release_sys(assist_done(c.heap[root], b, me), seq);
    }
  } else {
    if (has_work && clean) {
      V16 idv;
      idv.w[0] = idv.w[2] = (uint32_t)(ident & 0xffffffffull);
      idv.w[1] = idv.w[3] = (uint32_t)(ident >> 32);
      for (unsigned long long k = (unsigned long long)b; k < nchunks; k += (unsigned long long)G)
        for (unsigned int j = threadIdx.x; j < CH; j += blockDim.x) {
          const unsigned long long v = k * CH + j;
          if (v < nvec) st_peer(my_slot + v * 16, idv);
        }
      __syncthreads();
      if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me && !c.fault_skip_signal) st_release_sys(posted_ready(c.heap[threadIdx.x], b, me), seq);
    }
    if (has_work) {
      if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me) wait_flag_ge(assist_done(c.heap[me], b, (int)threadIdx.x), seq, c);
      __syncthreads();
      const char* res = c.heap[me] + res_off;
      for (unsigned long long k0 = (unsigned long long)b; k0 < nchunks; k0 += (unsigned long long)G * U) {
        for (unsigned int j = threadIdx.x; j < CH; j += blockDim.x) {
          V16 x[U], own[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long v = vec_of(k0, u, j);
            if (v != ~0ull) {
              x[u] = ld_peer(res + v * 16);
              own[u] = load_send(v);
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long v = vec_of(k0, u, j);
            if (v != ~0ull) {
              typename V::Acc acc = V::unpack(via_switch ? C::code(x[u]) : x[u]);
              V::combine(acc, own[u]);
              const V16 r = V::pack(acc);
              if (v < nfull) st_peer(A.recv_ptr + v * 16, r);
              else ll_store_tail<E>(A.recv_ptr + nfull * 16, r, tail_bytes);
            }
          }
        }
      }
    }
  }
  // every CTA of every rank acknowledges every call to everybody (see posted_body for why relaxed is enough)
  __syncthreads();
  if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me && !c.fault_skip_signal)
    st_relaxed_sys_u32(posted_ack(c.heap[threadIdx.x], me, b), seq);
}

}  // namespace b2r
