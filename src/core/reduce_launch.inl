// reduce_launch.inl — included by one translation unit per dtype (parallel builds); instantiates the kernel
// matrix {op} x {LDG128, LDG256} x {unroll 4, 8} x {cluster, no cluster} + {TMA} x {cluster, no cluster} and the
// switch that picks one. This is the counterpart of the reference's switch(threads) template dispatch
// (reduction_kernel.cu:263-346), keyed on what matters on sm_100a instead of on a power-of-two block size.
#include <mutex>
#include <set>

#include "../util/check.h"
#include "reduce_kernels.cuh"

namespace b2r {
namespace {

using KernelFn = void (*)(const LocalParams);
// 227 KiB is the per-CTA opt-in limit on sm_100; the kernels also use <1 KiB of static shared memory.
constexpr int kMaxDynSmem = 226 * 1024;

template <int D, int O, bool CL>
KernelFn pick_kernel(const LaunchGeom& g) {
  switch (g.variant) {
    case B2R_VARIANT_LDG128:
      return g.unroll == 8 ? (KernelFn)reduce_ldg_kernel<D, O, 4, 8, CL> : (KernelFn)reduce_ldg_kernel<D, O, 4, 4, CL>;
    case B2R_VARIANT_LDG256:
      // (8 x 256-bit loads per thread makes ptxas serialise load->use; 4 in flight = 128 B/thread is the useful maximum)
      return (KernelFn)reduce_ldg_kernel<D, O, 8, 4, CL>;
    case B2R_VARIANT_TMA:
      return (KernelFn)reduce_tma_kernel<D, O, CL>;
  }
  return nullptr;
}

// the dynamically scheduled tail exists for the register-streaming variants without clusters
template <int D, int O>
KernelFn pick_dyn_kernel(const LaunchGeom& g) {
  switch (g.variant) {
    case B2R_VARIANT_LDG128:
      return g.unroll == 8 ? (KernelFn)reduce_ldg_dyn_kernel<D, O, 4, 8> : (KernelFn)reduce_ldg_dyn_kernel<D, O, 4, 4>;
    case B2R_VARIANT_LDG256:
      return (KernelFn)reduce_ldg_dyn_kernel<D, O, 8, 4>;
  }
  return nullptr;
}

template <int D, int O>
b2r_status_t launch_op(const LaunchGeom& g, const LocalParams& p, cudaStream_t stream) {
  KernelFn k = g.dyn_pct > 0 ? pick_dyn_kernel<D, O>(g) : (g.cluster > 1 ? pick_kernel<D, O, true>(g) : pick_kernel<D, O, false>(g));
  if (!k) return fail(B2R_ERR_INVALID, "launch_local: bad variant %d", g.variant);
  if (g.smem_bytes > 48 * 1024) {
    // opt in to large dynamic shared memory once per (kernel, device)
    static std::mutex mu;
    static std::set<std::pair<const void*, int>> done;
    int dev = 0;
    B2R_CUDA_CHECK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    if (!done.count({(const void*)k, dev})) {
      B2R_CUDA_CHECK(cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
      done.insert({(const void*)k, dev});
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)g.ctas);
  cfg.blockDim = dim3((unsigned)g.threads);
  cfg.dynamicSmemBytes = (size_t)g.smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  unsigned na = 0;
  if (g.cluster > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = (unsigned)g.cluster;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (g.pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = na ? attr : nullptr;
  cfg.numAttrs = na;
  B2R_CUDA_CHECK(cudaLaunchKernelEx(&cfg, k, p));
  return B2R_OK;
}

template <int D, int O>
b2r_status_t launch_multi_op(int nranks, int ctas, int threads, const MultiLocalParams& mp, cudaStream_t stream) {
  // cooperative launch: all gridDim.x * nranks CTAs must be co-resident because ranks wait for one another
  void* args[] = {(void*)&mp};
  B2R_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)reduce_ldg_multi_kernel<D, O>, dim3((unsigned)ctas, (unsigned)nranks), dim3((unsigned)threads),
                                             args, 0, stream));
  return B2R_OK;
}

}  // namespace

template <>
b2r_status_t launch_local_multi<B2R_LAUNCH_DTYPE>(int op, int nranks, int ctas, int threads, const MultiLocalParams& mp, cudaStream_t stream) {
  switch (op) {
    case B2R_SUM: return launch_multi_op<B2R_LAUNCH_DTYPE, B2R_SUM>(nranks, ctas, threads, mp, stream);
    case B2R_MIN: return launch_multi_op<B2R_LAUNCH_DTYPE, B2R_MIN>(nranks, ctas, threads, mp, stream);
    case B2R_MAX: return launch_multi_op<B2R_LAUNCH_DTYPE, B2R_MAX>(nranks, ctas, threads, mp, stream);
  }
  return fail(B2R_ERR_INVALID, "launch_local_multi: bad op %d", op);
}

template <>
b2r_status_t launch_local<B2R_LAUNCH_DTYPE>(int op, const LaunchGeom& g, const LocalParams& p, cudaStream_t stream) {
  switch (op) {
    case B2R_SUM: return launch_op<B2R_LAUNCH_DTYPE, B2R_SUM>(g, p, stream);
    case B2R_MIN: return launch_op<B2R_LAUNCH_DTYPE, B2R_MIN>(g, p, stream);
    case B2R_MAX: return launch_op<B2R_LAUNCH_DTYPE, B2R_MAX>(g, p, stream);
  }
  return fail(B2R_ERR_INVALID, "launch_local: bad op %d", op);
}

}  // namespace b2r
