// host_coll.cpp — CPU execution of the collective algorithms (threads as ranks). See host_algos.h.
#include <condition_variable>
#include <mutex>
#include <thread>

#include "../core/rng.h"
#include "../util/check.h"
#include "host_algos.h"

namespace b2r {

namespace {
template <typename T>
T hmin_nan(T a, T b) { return (a != a) ? a : ((b != b) ? b : (b < a ? b : a)); }
template <typename T>
T hmax_nan(T a, T b) { return (a != a) ? a : ((b != b) ? b : (b > a ? b : a)); }

template <typename T>
void combine_fp(T* dst, const void* const* src, int nsrc, uint64_t first, uint64_t count, b2r_op_t op) {
  for (uint64_t i = first; i < first + count; ++i) {
    T acc = ((const T*)src[0])[i];
    for (int s = 1; s < nsrc; ++s) {
      const T x = ((const T*)src[s])[i];
      acc = op == B2R_SUM ? acc + x : (op == B2R_MIN ? hmin_nan(acc, x) : hmax_nan(acc, x));
    }
    dst[i] = acc;
  }
}

class ThreadBarrier {
 public:
  explicit ThreadBarrier(int n) : n_(n) {}
  void wait() {
    std::unique_lock<std::mutex> lk(mu_);
    const unsigned gen = gen_;
    if (++arrived_ == n_) {
      arrived_ = 0;
      ++gen_;
      cv_.notify_all();
    } else {
      cv_.wait(lk, [&] { return gen_ != gen; });
    }
  }

 private:
  std::mutex mu_;
  std::condition_variable cv_;
  int n_, arrived_ = 0;
  unsigned gen_ = 0;
};
}  // namespace

void combine_many_host(void* dst, const void* const* src, int nsrc, uint64_t first, uint64_t count, b2r_dtype_t dtype, b2r_op_t op) {
  switch (dtype) {
    case B2R_INT32: {
      int32_t* d = (int32_t*)dst;
      for (uint64_t i = first; i < first + count; ++i) {
        int32_t acc = ((const int32_t*)src[0])[i];
        for (int s = 1; s < nsrc; ++s) {
          const int32_t x = ((const int32_t*)src[s])[i];
          acc = op == B2R_SUM ? (int32_t)((uint32_t)acc + (uint32_t)x) : (op == B2R_MIN ? (x < acc ? x : acc) : (x > acc ? x : acc));
        }
        d[i] = acc;
      }
      return;
    }
    case B2R_FLOAT32: combine_fp((float*)dst, src, nsrc, first, count, op); return;
    case B2R_FLOAT64: combine_fp((double*)dst, src, nsrc, first, count, op); return;
    case B2R_BFLOAT16: {
      uint16_t* d = (uint16_t*)dst;
      for (uint64_t i = first; i < first + count; ++i) {
        float acc = bf16_bits_to_f32(((const uint16_t*)src[0])[i]);
        for (int s = 1; s < nsrc; ++s) {
          const float x = bf16_bits_to_f32(((const uint16_t*)src[s])[i]);
          acc = op == B2R_SUM ? acc + x : (op == B2R_MIN ? hmin_nan(acc, x) : hmax_nan(acc, x));
        }
        d[i] = f32_to_bf16_bits(acc);
      }
      return;
    }
  }
}

}  // namespace b2r

using namespace b2r;

extern "C" b2r_status_t b2r_host_reduce_threads(const void* const* sendbufs, void* recvbuf, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                                                int nranks, int root, b2r_algo_t algo) {
  B2R_REQUIRE(sendbufs && recvbuf, "b2r_host_reduce_threads: NULL buffer");
  B2R_REQUIRE(nranks >= 1 && nranks <= 1024, "b2r_host_reduce_threads: nranks=%d out of range", nranks);
  B2R_REQUIRE(root >= -1 && root < nranks, "b2r_host_reduce_threads: bad root %d", root);
  B2R_REQUIRE(dtype >= B2R_INT32 && dtype <= B2R_BFLOAT16 && op >= B2R_SUM && op <= B2R_MAX, "b2r_host_reduce_threads: bad dtype/op");
  const size_t esz = b2r_dtype_size(dtype);
  // root >= 0: recvbuf is the root's buffer. root == -1 (allreduce): recvbuf holds nranks consecutive result buffers.
  std::vector<std::vector<unsigned char>> work(nranks), scratch_recv(nranks);
  std::vector<void*> recv(nranks, nullptr), workp(nranks, nullptr);
  for (int r = 0; r < nranks; ++r) {
    if (algo == B2R_ALGO_TREE) {
      work[r].resize(count * esz + 16);
      workp[r] = work[r].data();
    }
    if (root < 0) recv[r] = (char*)recvbuf + (size_t)r * count * esz;
    else if (r == root) recv[r] = recvbuf;
  }
  ThreadBarrier bar(nranks);
  std::vector<std::thread> th;
  for (int r = 0; r < nranks; ++r) {
    th.emplace_back([&, r] {
      HostRankView v{r, nranks, sendbufs, recv.data(), workp.data()};
      host_collective_rank(v, count, dtype, op, root, algo, [&] { bar.wait(); });
    });
  }
  for (auto& t : th) t.join();
  return B2R_OK;
}
