// misc_kernels.cu — device-side data generation and the transform-reduce (dot product) hook.
//  * b2r_fill_device: the device twin of b2r_fill_host (rng.h); replaces the reference's host-side rand()/MT19937
//    fill + H2D copy (reduction.cpp:698-726, mpi/reduce.c:51-57) when the benchmark wants data born on the GPU.
//  * b2r_dot_local: vendored scalarProd analogue (scalarProd_kernel.cu:33-80) on top of the same block/ticket
//    machinery as the reduction family.
#include <cuda_runtime.h>

#include "../util/check.h"
#include "b2r.h"
#include "reduce_kernels.cuh"
#include "rng.h"
#include "workspace.h"

namespace b2r {
namespace {

template <int D>
__global__ void fill_kernel(void* dst, uint64_t n, uint64_t seed, uint64_t first, int mode) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (D == B2R_INT32) ((int32_t*)dst)[i] = gen_i32(seed, first + i, mode);
    else if (D == B2R_FLOAT32) ((float*)dst)[i] = gen_f32(seed, first + i, mode);
    else if (D == B2R_FLOAT64) ((double*)dst)[i] = gen_f64(seed, first + i, mode);
    else ((uint16_t*)dst)[i] = gen_bf16(seed, first + i, mode);
  }
}

// dot product: grid-stride, products accumulated per thread, then the family's finish_block (ticket combine).
template <int D>
__global__ void __launch_bounds__(512) dot_kernel(const void* __restrict__ a, const void* __restrict__ b,
                                                  const __grid_constant__ LocalParams p) {
  using R = Reducer<D, B2R_SUM>;
  using acc_t = typename R::acc_t;
  acc_t s0 = 0, s1 = 0;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + stride < p.n; i += 2 * stride) {
    if (D == B2R_INT32) {
      s0 = (acc_t)((uint32_t)s0 + (uint32_t)((const int32_t*)a)[i] * (uint32_t)((const int32_t*)b)[i]);
      s1 = (acc_t)((uint32_t)s1 + (uint32_t)((const int32_t*)a)[i + stride] * (uint32_t)((const int32_t*)b)[i + stride]);
    } else if (D == B2R_FLOAT32) {
      s0 = fmaf(((const float*)a)[i], ((const float*)b)[i], s0);
      s1 = fmaf(((const float*)a)[i + stride], ((const float*)b)[i + stride], s1);
    } else if (D == B2R_FLOAT64) {
      s0 = fma(((const double*)a)[i], ((const double*)b)[i], s0);
      s1 = fma(((const double*)a)[i + stride], ((const double*)b)[i + stride], s1);
    } else {
      s0 = fmaf(bf16_bits_to_f32(((const uint16_t*)a)[i]), bf16_bits_to_f32(((const uint16_t*)b)[i]), s0);
      s1 = fmaf(bf16_bits_to_f32(((const uint16_t*)a)[i + stride]), bf16_bits_to_f32(((const uint16_t*)b)[i + stride]), s1);
    }
  }
  if (i < p.n) {
    if (D == B2R_INT32) s0 = (acc_t)((uint32_t)s0 + (uint32_t)((const int32_t*)a)[i] * (uint32_t)((const int32_t*)b)[i]);
    else if (D == B2R_FLOAT32) s0 = fmaf(((const float*)a)[i], ((const float*)b)[i], s0);
    else if (D == B2R_FLOAT64) s0 = fma(((const double*)a)[i], ((const double*)b)[i], s0);
    else s0 = fmaf(bf16_bits_to_f32(((const uint16_t*)a)[i]), bf16_bits_to_f32(((const uint16_t*)b)[i]), s0);
  }
  acc_t v = D == B2R_INT32 ? (acc_t)((uint32_t)s0 + (uint32_t)s1) : (acc_t)(s0 + s1);
  finish_block<R, false>(v, p);
}

}  // namespace
}  // namespace b2r

using namespace b2r;

extern "C" {

b2r_status_t b2r_fill_device(void* dst, uint64_t count, b2r_dtype_t dtype, uint64_t seed, uint64_t first, int mode,
                             void* stream_) {
  B2R_REQUIRE(dst != nullptr || count == 0, "b2r_fill_device: dst is NULL");
  B2R_REQUIRE(mode >= 0 && mode <= 2, "b2r_fill_device: bad mode %d", mode);
  if (count == 0) return B2R_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  const int threads = 256;
  const int blocks = (int)std::min<uint64_t>((count + threads - 1) / threads, 148 * 16);
  switch (dtype) {
    case B2R_INT32: fill_kernel<B2R_INT32><<<blocks, threads, 0, stream>>>(dst, count, seed, first, mode); break;
    case B2R_FLOAT32: fill_kernel<B2R_FLOAT32><<<blocks, threads, 0, stream>>>(dst, count, seed, first, mode); break;
    case B2R_FLOAT64: fill_kernel<B2R_FLOAT64><<<blocks, threads, 0, stream>>>(dst, count, seed, first, mode); break;
    case B2R_BFLOAT16: fill_kernel<B2R_BFLOAT16><<<blocks, threads, 0, stream>>>(dst, count, seed, first, mode); break;
    default: return fail(B2R_ERR_INVALID, "b2r_fill_device: bad dtype %d", (int)dtype);
  }
  B2R_CUDA_CHECK(cudaGetLastError());
  return B2R_OK;
}

b2r_status_t b2r_dot_local(const void* a, const void* b, void* out, uint64_t count, b2r_dtype_t dtype,
                           b2r_workspace_t ws, void* stream_) {
  // same contract as b2r_reduce_local: validated up front, launched on the workspace's device whatever the caller's current
  // device is, and recorded as the workspace's last output so that a following reduction of `out` never overlaps with it
  B2R_REQUIRE(ws != nullptr && out != nullptr, "b2r_dot_local: NULL argument");
  B2R_REQUIRE((a && b) || count == 0, "b2r_dot_local: NULL input");
  B2R_REQUIRE(dtype >= B2R_INT32 && dtype <= B2R_BFLOAT16, "b2r_dot_local: bad dtype %d", (int)dtype);
  const size_t esz = b2r_dtype_size(dtype);
  B2R_REQUIRE(((uintptr_t)a % esz) == 0 && ((uintptr_t)b % esz) == 0 && ((uintptr_t)out % esz) == 0,
              "b2r_dot_local: pointers must be aligned to the element size %zu", esz);
  cudaStream_t stream = (cudaStream_t)stream_;
  int cur = -1;
  B2R_CUDA_CHECK(cudaGetDevice(&cur));
  struct DeviceGuard {
    int prev;
    bool active;
    ~DeviceGuard() { if (active) cudaSetDevice(prev); }
  } guard{cur, cur != ws->device};
  if (guard.active) B2R_CUDA_CHECK(cudaSetDevice(ws->device));
  LocalParams p;
  memset(&p, 0, sizeof(p));
  p.n = count;
  p.out = out;
  p.partials = ws->d_partials;
  p.ticket = ws->d_ticket;
  const int threads = 256;
  const uint64_t want = (count + (uint64_t)threads * 8 - 1) / ((uint64_t)threads * 8);
  const int blocks = (int)std::max<uint64_t>(1, std::min<uint64_t>(want, (uint64_t)ws->sm_count * 8));
  switch (dtype) {
    case B2R_INT32: dot_kernel<B2R_INT32><<<blocks, threads, 0, stream>>>(a, b, p); break;
    case B2R_FLOAT32: dot_kernel<B2R_FLOAT32><<<blocks, threads, 0, stream>>>(a, b, p); break;
    case B2R_FLOAT64: dot_kernel<B2R_FLOAT64><<<blocks, threads, 0, stream>>>(a, b, p); break;
    case B2R_BFLOAT16: dot_kernel<B2R_BFLOAT16><<<blocks, threads, 0, stream>>>(a, b, p); break;
  }
  B2R_CUDA_CHECK(cudaGetLastError());
  ws->last_out = out;
  ws->last_ctas = blocks;
  return B2R_OK;
}

}  // extern "C"
