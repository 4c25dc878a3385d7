// reduce_internal.h — entry points of the single-GPU reduction shared with the collectives (scalar-sharded reduce).
#pragma once
#include <cuda_runtime.h>

#include "b2r.h"
#include "reduce_kernels.cuh"
#include "workspace.h"

namespace b2r {

struct XgCall {  // cross-GPU tail of a reduction (see xgpu_finish in reduce_kernels.cuh)
  const CollCtx* ctx;  // DEVICE pointer to this rank's context copy
  int root, all;
  unsigned seq;
};

b2r_status_t plan_local(uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int sm_count, const b2r_launch_cfg_t* hint, LaunchGeom* g,
                        int* stages, int* tile_vecs);
b2r_status_t prepare_local(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, b2r_workspace* ws,
                           const b2r_launch_cfg_t* cfg, int out_as_acc, LaunchGeom* g, LocalParams* p);
b2r_status_t reduce_local_impl(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, b2r_workspace* ws,
                               cudaStream_t stream, const b2r_launch_cfg_t* cfg, int out_as_acc, const XgCall* xg);
b2r_status_t stream_chunks_from_host(const void* host_in, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, b2r_workspace* ws, cudaStream_t stream,
                                     uint64_t* nchunks);
void acc_to_dtype(const void* acc, b2r_dtype_t dtype, void* out);
b2r_status_t launch_any(b2r_dtype_t dtype, int op, const LaunchGeom& g, const LocalParams& p, cudaStream_t stream);
b2r_status_t launch_multi_any(b2r_dtype_t dtype, int op, int nranks, int ctas, int threads, const MultiLocalParams& mp, cudaStream_t stream);

}  // namespace b2r
