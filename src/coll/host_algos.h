// host_algos.h — the collective algorithms restated for CPU "ranks" (threads of one process, or processes sharing
// POSIX shared memory). This is the host loopback transport of SURVEY §4: the same slice / tree / ordering logic as the
// device kernels (coll_kernels.cuh) runs without a GPU, serves as their bit-exact oracle (same combination order), and is
// the engine behind the bundled mini-MPI's MPI_Reduce and the hand-written tree reduce of BASELINE.json config 1.
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>

#include "b2r.h"

namespace b2r {

// dst[i] = op(src[0][i], src[1][i], ..., src[n-1][i]) combined left to right; bf16 combines in fp32 and rounds once.
void combine_many_host(void* dst, const void* const* src, int nsrc, uint64_t first, uint64_t count, b2r_dtype_t dtype, b2r_op_t op);

struct HostRankView {
  int rank, nranks;
  const void* const* send;  // send[p]: rank p's send buffer (readable by everybody)
  void* const* recv;        // recv[p]: rank p's recv buffer (writable by everybody; may be null where unused)
  void* const* work;        // work[p]: rank p's scratch of `count` elements (tree only)
};

// Every rank calls this with its own view; `barrier()` must synchronise all ranks (and make their writes visible).
template <class Barrier>
void host_collective_rank(const HostRankView& v, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int root /* -1 = allreduce */,
                          b2r_algo_t algo, Barrier&& barrier) {
  const int P = v.nranks, r = v.rank;
  barrier();  // everybody's send buffer is ready
  if (algo == B2R_ALGO_ONESHOT || algo == B2R_ALGO_AUTO || algo == B2R_ALGO_NVLS || algo == B2R_ALGO_LL ||
      algo == B2R_ALGO_POSTED) {  // LL, posted: same rank-order fold
    if (root < 0 || root == r) combine_many_host(v.recv[r], v.send, P, 0, count, dtype, op);
    barrier();
  } else if (algo == B2R_ALGO_TWOSHOT || algo == B2R_ALGO_PUSH) {  // same slices, same rank-order combine
    int nworkers = root < 0 ? P : P - 1;
    int me = root < 0 ? r : (r == root ? -1 : (r < root ? r : r - 1));
    if (P == 1) { nworkers = 1; me = 0; }
    if (me >= 0) {
      const uint64_t lo = count * (uint64_t)me / nworkers, hi = count * (uint64_t)(me + 1) / nworkers;
      if (root < 0) {
        combine_many_host(v.recv[r], v.send, P, lo, hi - lo, dtype, op);
        const size_t esz = b2r_dtype_size(dtype);
        for (int p = 0; p < P; ++p)
          if (p != r) memcpy((char*)v.recv[p] + lo * esz, (const char*)v.recv[r] + lo * esz, (hi - lo) * esz);
      } else {
        combine_many_host(v.recv[root], v.send, P, lo, hi - lo, dtype, op);
      }
    }
    barrier();
  } else if (algo == B2R_ALGO_RING) {
    // Chain ending at a root: left fold from the farthest position down to the root (same association order as ring_body).
    // Hop by hop: the running partial travels in the buffer's own dtype, so bf16 is rounded at EVERY hop (unlike the
    // one-shot / two-shot / push / NVLS paths, which combine all P contributions in fp32 and round once).
    const size_t esz = b2r_dtype_size(dtype);
    auto chain_fold = [&](int rt, uint64_t first, uint64_t n) {  // recv[rt][first, first+n) = fold over the chain that ends at rt
      if (n == 0) return;
      memcpy((char*)v.recv[rt] + first * esz, (const char*)v.send[(rt + P - 1) % P] + first * esz, n * esz);
      for (int q = 1; q < P; ++q) {
        const void* pair[2] = {v.recv[rt], v.send[(rt + (P - 1 - q)) % P]};
        combine_many_host(v.recv[rt], pair, 2, first, n, dtype, op);
      }
    };
    if (root >= 0) {
      if (r == root) chain_fold(root, 0, count);
    } else {
      // allreduce = P chains at once: chain c ends at rank c and reduces slice c (cut at 16-byte vectors like the device
      // kernel; the sub-vector tail rides with the last chain), then every chain root copies its slice to every rank
      const uint64_t per_vec = 16 / esz, nvec = count / per_vec;
      const uint64_t lo = nvec * (uint64_t)r / P * per_vec;
      const uint64_t hi = r == P - 1 ? count : nvec * (uint64_t)(r + 1) / P * per_vec;
      chain_fold(r, lo, hi - lo);
      for (int p = 0; p < P; ++p)
        if (p != r && hi > lo) memcpy((char*)v.recv[p] + lo * esz, (const char*)v.recv[r] + lo * esz, (hi - lo) * esz);
    }
    barrier();
  } else {  // B2R_ALGO_TREE
    const int rt = root < 0 ? 0 : root;
    const int pos = (r - rt + P) % P;
    int rounds = 0;
    while ((1 << rounds) < P) ++rounds;
    if (rounds == 0) {
      const void* s[1] = {v.send[r]};
      combine_many_host(v.recv[r], s, 1, 0, count, dtype, op);
    }
    for (int k = 0; k < rounds; ++k) {
      const int bit = 1 << k;
      if (pos % (2 * bit) == 0 && pos + bit < P) {
        const int partner = (pos + bit + rt) % P;
        const bool last = k == rounds - 1;
        const bool mine_in_work = k >= 1 && pos + 1 < P;
        const bool peer_in_work = k >= 1 && (pos + bit) + 1 < P;
        const void* s[2] = {mine_in_work ? v.work[r] : v.send[r], peer_in_work ? v.work[partner] : v.send[partner]};
        void* d = last ? v.recv[r] : v.work[r];
        combine_many_host(d, s, 2, 0, count, dtype, op);
        if (last && root < 0) {
          const size_t esz = b2r_dtype_size(dtype);
          for (int p = 0; p < P; ++p)
            if (p != r) memcpy(v.recv[p], v.recv[r], count * esz);
        }
      }
      barrier();
    }
  }
}

}  // namespace b2r
