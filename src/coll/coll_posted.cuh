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
//   roots     : may change from call to call; the acknowledgement a rank waits for is that of the root of the call two back
//               (the host passes it in), and words are indexed by the acknowledging rank.
//   fp min/max: have no multimem form, but signed-integer min/max do, and IEEE floats order like sign-magnitude integers.
//               Staging maps every value to the two's-complement integer of the same rank (k = b ^ ((b >> 31) & 0x7fff...)), the
//               switch reduces integers, the root maps back (the map is its own inverse). -0 orders below +0; a NaN orders
//               beyond the infinity of its sign (the host restatement and the pull kernels use a < b ? a : b).
//   history   : a variant in which the P-1 non-root ranks did the switch reductions around an identity kept in the root's slot
//               and delivered into a result slot at the root ("assist") was built and measured on 4 GPUs: correct, and no
//               faster than this one at any size (profiles/raw/r2_posted/table_4gpu_assist.txt). Removed.
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
    for (int j = 0; j < 4; ++j) k.w[j] = key32(v.w[j]);
    return k;
  }
};
template <int O>
struct PostedKey64 {
  static constexpr bool kSwitch = true;
  static constexpr int kBytes = 8;
  __device__ __forceinline__ static void ld_reduce(const char* p, uint32_t* w) { NvlsS64<O>::ld_reduce(p, w); }
  __device__ __forceinline__ static V16 code(const V16& v) {  // little endian: w[1], w[3] hold the sign and the high bits
    V16 k = v;
    key64(k.w[0], k.w[1]);
    key64(k.w[2], k.w[3]);
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
  const int prev_root = A.ring_depth;  // root of posted call seq-2, the last reader of this parity's slots (-1: none yet)
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
    if (prev_root >= 0 && prev_root != me) {
      if (threadIdx.x == 0) wait_flag_ge_relaxed(posted_ack(c.heap[me], prev_root, b), seq - 2, c);
      __syncthreads();
    }
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
        if (via_switch && C::kBytes < 16) {
          // 4- and 8-byte multimem forms: a thread does NOT take the 16 / B units of "its" vector (a warp's instruction would
          // then touch 32 units 16 bytes apart: four times the transactions, and the first version measured 120 GB/s for int32
          // against 390 for fp32). Unit m of the thread sits at j + m x (units of a quarter chunk), so every instruction of a
          // warp covers 32 consecutive units. Messages are whole units (count x element size), so no vector tail here.
          constexpr int B = C::kBytes, M = 16 / B;
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long k = k0 + (unsigned long long)u * G;
#pragma unroll
            for (int m = 0; m < M; ++m) {
              const unsigned long long off = k * kPostedChunkVecs * 16ull + ((unsigned long long)j + (unsigned long long)m * kPostedChunkVecs) * B;
              if (k < nchunks && off < bytes) C::ld_reduce(mslot + off, &r[u].w[m * B / 4]);
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) r[u] = C::code(r[u]);  // element-wise (32- or 64-bit), so any unit may sit in any lane
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long k = k0 + (unsigned long long)u * G;
#pragma unroll
            for (int m = 0; m < M; ++m) {
              const unsigned long long off = k * kPostedChunkVecs * 16ull + ((unsigned long long)j + (unsigned long long)m * kPostedChunkVecs) * B;
              if (k < nchunks && off < bytes) {
                if (B == 8) asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1,%2};" ::"l"(A.recv_ptr + off), "r"(r[u].w[2 * m]), "r"(r[u].w[(2 * m + 1) & 3]) : "memory");
                else asm volatile("st.relaxed.sys.global.b32 [%0], %1;" ::"l"(A.recv_ptr + off), "r"(r[u].w[m]) : "memory");
              }
            }
          }
          continue;
        }
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

}  // namespace b2r
