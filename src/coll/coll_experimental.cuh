// coll_experimental.cuh — the LL ("low latency") small-message path. Included by coll_kernels.cuh (needs its helpers: VecOps,
// ld_peer / st_peer, barriers); tests/test_experimental_gpu.py (loopback), tests/dist_worker.py (real NVLink) and
// tests/test_ll_protocol_model.py (CPU model of the slot / acknowledgement protocol) cover it.
// History: this file also held a pipelined ring (waiter / data / signal warps) and a ring allreduce (P chains + all-gather by
// stores). Both were measured on 8 B200s in round 2 and removed: the pipelined ring was slower than the plain ring at every
// size (884 vs 669 us at 256 MiB) and failed its bf16 check on hardware; the ring allreduce lost to NVLS and to two-shot at
// every size (802 vs 576 / 891 us at 256 MiB int32). profiles/raw/r2_8gpu/ has the rows.
#pragma once

namespace b2r {

// ------------------------------------------------------------------------------------------------ LL (small messages, no barriers)  
// The barrier-framed kernels above pay two NVLink round trips (start + end barrier) whatever the size: 12.5-14.5 us for a
// 1 KiB-256 KiB reduce at 8 GPUs where NCCL's LL protocol needs 8-12. Here the synchronisation rides with the data, as in
// that protocol: every 16-byte vector is pushed into the consumer's slot as four 8-byte packets {32 data bits, call number};
// a packet is valid as soon as its upper half equals the call number the consumer is waiting for, so arrival needs no
// separate flag and no fence, and a sender is done once its stores are issued. One NVLink one-way trip end to end.
//   consumers : reduce -> the root; allreduce -> every rank (each combines all P contributions itself, in rank order, so the
//               bits equal the one-shot kernel's and the host restatement's).
//   slots     : [call parity][source rank] in the consumer's control region — two calls may be in flight; before writing
//               call n a sender checks that the consumer has acknowledged call n-2 (every rank acknowledges every call to
//               every peer when its own work for that call is finished; the check is off the critical path).
//   geometry  : a FIXED partition — CTA b of every rank owns vectors [b x vpc, (b+1) x vpc) of the message whatever its
//               length, and every call launches the same number of CTAs (those past the end of a short message only take
//               part in the acknowledgements). So CTA b's slots are only ever touched by CTA b of the peers, and one
//               acknowledgement word per (CTA, peer) is all the flow control needs. Two instances of this layout exist:
//               the small one in the control region (1 CTA, <= 26 KiB at 8 ranks) and, when the communicator was created
//               with B2R_LL_BIG_MB set, a large one behind the user heap (16 CTAs x 16 KiB = 256 KiB). The host passes the
//               layout in fields of CollArgs that LL does not otherwise need (see ll_layout below).
// Packets are written and read with 8-byte-element vector accesses (each element is one scalar access in the PTX memory
// model, hence never torn). Call numbers are 32 bits and never reset: 2^32 LL calls per communicator.
__device__ __forceinline__ void ll_store_vec(char* p, const V16& v, unsigned int seq) {
  const unsigned long long hi = (unsigned long long)seq << 32;
  const unsigned long long a = hi | v.w[0], b = hi | v.w[1], c = hi | v.w[2], d = hi | v.w[3];
  asm volatile("st.relaxed.sys.global.v2.b64 [%0], {%1,%2};" ::"l"(p), "l"(a), "l"(b) : "memory");
  asm volatile("st.relaxed.sys.global.v2.b64 [%0], {%1,%2};" ::"l"(p + 16), "l"(c), "l"(d) : "memory");
}
__device__ __forceinline__ void ll_load_vec(const char* p, unsigned long long* pk) {
  asm volatile("ld.relaxed.sys.global.v2.b64 {%0,%1}, [%2];" : "=l"(pk[0]), "=l"(pk[1]) : "l"(p) : "memory");
  asm volatile("ld.relaxed.sys.global.v2.b64 {%0,%1}, [%2];" : "=l"(pk[2]), "=l"(pk[3]) : "l"(p + 16) : "memory");
}
__device__ __forceinline__ bool ll_valid(const unsigned long long* pk, unsigned int seq) {
  return (unsigned int)(pk[0] >> 32) == seq && (unsigned int)(pk[1] >> 32) == seq && (unsigned int)(pk[2] >> 32) == seq &&
         (unsigned int)(pk[3] >> 32) == seq;
}
// the last (< 16 byte) piece of a message as a zero-padded vector, and back; E = element size (2, 4 or 8)
template <int E>
__device__ __forceinline__ V16 ll_load_tail(const char* p, unsigned int tail_bytes) {
  V16 v;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t w = 0;
    if (4u * j + 4u <= tail_bytes) w = *(const volatile uint32_t*)(p + 4 * j);
    else if (E == 2 && 4u * j + 2u <= tail_bytes) w = *(const volatile uint16_t*)(p + 4 * j);
    v.w[j] = w;
  }
  return v;
}
template <int E>
__device__ __forceinline__ void ll_store_tail(char* p, const V16& v, unsigned int tail_bytes) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (4u * j + 4u <= tail_bytes) *(volatile uint32_t*)(p + 4 * j) = v.w[j];
    else if (E == 2 && 4u * j + 2u <= tail_bytes) *(volatile uint16_t*)(p + 4 * j) = (uint16_t)(v.w[j] & 0xffffu);
  }
}

// Where an LL call keeps its slots and acknowledgement words. Carried in CollArgs fields the LL path has no other use for, so
// that sizeof(CollArgs) — and with it the code of every other kernel — stays what the GPU test-suite has passed on.
struct LLLayout {
  unsigned long long region_off;  // A.work_off : slots [call parity][source rank][slot_bytes]
  unsigned long long ack_off;     // A.send_off : uint32 ack[CTA][B2R_MAX_RANKS]
  unsigned int slot_bytes;        // A.sub_blocks
  unsigned int vec_per_cta;       // A.ring_depth
};
__host__ __device__ inline LLLayout ll_layout(const CollArgs& A) {
  LLLayout l;
  l.region_off = A.work_off;
  l.ack_off = A.send_off;
  l.slot_bytes = (unsigned int)A.sub_blocks;
  l.vec_per_cta = (unsigned int)A.ring_depth;
  return l;
}
inline void ll_set_layout(CollArgs& A, const LLLayout& l) {
  A.work_off = l.region_off;
  A.send_off = l.ack_off;
  A.sub_blocks = (int)l.slot_bytes;
  A.ring_depth = (int)l.vec_per_cta;
}

template <int D, int O>
__device__ void ll_body(const CollArgs& A) {
  using V = VecOps<D, O>;
  const CollCtx& c = A.ctx;
  constexpr int E = V::kElem;
  const int P = c.nranks, me = c.rank, root = A.root;
  const unsigned int seq = A.seq;
  const bool consumer = root < 0 || root == me;
  const unsigned long long bytes = A.count * E;
  const unsigned int nfull = (unsigned int)(bytes / 16), tail_bytes = (unsigned int)(bytes % 16);
  const unsigned int nvec = nfull + (tail_bytes ? 1u : 0u);
  const LLLayout lay = ll_layout(A);
  const size_t slot = lay.slot_bytes;
  const size_t parity_off = lay.region_off + (size_t)(seq & 1u) * P * slot;
  // this CTA's share of the message: [v0, vend), at most blockDim.x * kLLVecPerThread vectors (the host sizes vec_per_cta so)
  const unsigned int v0 = blockIdx.x * lay.vec_per_cta;
  const unsigned int vend = min(nvec, v0 + lay.vec_per_cta);
  const size_t my_acks = lay.ack_off + (size_t)blockIdx.x * B2R_MAX_RANKS * sizeof(unsigned int);

  // ---- flow control: the slots of this parity were last used by call seq-2; its consumers must be done with them
  if (threadIdx.x == 0) fault_delay(c);
  if (seq > 2 && threadIdx.x < (unsigned)P && (int)threadIdx.x != me && (root < 0 || (int)threadIdx.x == root)) {
    const unsigned int* ack = (const unsigned int*)(c.heap[me] + my_acks) + threadIdx.x;
    wait_flag_ge(ack, seq - 2, c);  // on time-out the error word is set; the slots are always valid memory, so carry on
  }
  __syncthreads();

  // ---- my contribution: read once, keep in registers, push to every consumer
  V16 own[kLLVecPerThread];
#pragma unroll
  for (int u = 0; u < kLLVecPerThread; ++u) {
    const unsigned int v = v0 + threadIdx.x + u * blockDim.x;
    if (v < vend) {
      if (v < nfull) own[u] = ld_peer(A.send_ptr + (size_t)v * 16);
      else own[u] = ll_load_tail<E>(A.send_ptr + (size_t)nfull * 16, tail_bytes);
    }
  }
  if (!c.fault_skip_signal) {
    for (int d = 0; d < P; ++d) {
      if (d == me || (root >= 0 && d != root)) continue;
      char* base = c.heap[d] + parity_off + (size_t)me * slot;
#pragma unroll
      for (int u = 0; u < kLLVecPerThread; ++u) {
        const unsigned int v = v0 + threadIdx.x + u * blockDim.x;
        if (v < vend) ll_store_vec(base + (size_t)v * 32, own[u], seq);
      }
    }
  }

  // ---- consumers: wait for each peer's packets (four sources polled at a time), combine in rank order, store
  if (consumer) {
    unsigned int* err = (unsigned int*)(c.heap[me] + kOffError);
    const char* mine = c.heap[me] + parity_off;
#pragma unroll
    for (int u = 0; u < kLLVecPerThread; ++u) {
      const unsigned int v = v0 + threadIdx.x + u * blockDim.x;
      if (v >= vend) continue;
      typename V::Acc acc;
      bool ok = true;
      for (int p0 = 0; p0 < P && ok; p0 += 4) {
        unsigned long long pk[4][4];
        unsigned long long t0 = 0;
        unsigned int spins = 0;
        while (true) {
          bool all = true;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (p0 + k < P && p0 + k != me) ll_load_vec(mine + (size_t)(p0 + k) * slot + (size_t)v * 32, pk[k]);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (p0 + k < P && p0 + k != me) all = all && ll_valid(pk[k], seq);
          if (all) break;
          if (++spins >= 64) {
            spins = 0;
            const unsigned long long now = coll_globaltimer();
            if (t0 == 0) t0 = now;
            const bool late = c.timeout_ns && now - t0 > c.timeout_ns;
            if (late) atomicExch(err, 1u);
            if (late || ld_volatile_u32(err)) {  // give up as soon as anybody on this rank has
              ok = false;
              break;
            }
            __nanosleep(32);
          }
        }
        if (!ok) break;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int p = p0 + k;
          if (p >= P) continue;
          V16 x;
          if (p == me) x = own[u];
          else {
#pragma unroll
            for (int j = 0; j < 4; ++j) x.w[j] = (uint32_t)pk[k][j];
          }
          if (p == 0) acc = V::unpack(x);
          else V::combine(acc, x);
        }
      }
      if (ok) {
        const V16 r = V::pack(acc);
        if (v < nfull) st_peer(A.recv_ptr + (size_t)v * 16, r);
        else ll_store_tail<E>(A.recv_ptr + (size_t)nfull * 16, r, tail_bytes);
      }
    }
  }

  // ---- acknowledge the call to every peer. A consumer releases (its reads of the slots are complete before the ack is
  // visible); a rank that consumed nothing has nothing to order, and a release would only make it wait for its own packets
  // to be acknowledged by the link.
  __syncthreads();
  if (threadIdx.x < (unsigned)P && (int)threadIdx.x != me && !c.fault_skip_signal) {
    unsigned int* ack = (unsigned int*)(c.heap[threadIdx.x] + my_acks) + me;
    if (consumer) st_release_sys(ack, seq);
    else st_relaxed_sys_u32(ack, seq);
  }
}

}  // namespace b2r
