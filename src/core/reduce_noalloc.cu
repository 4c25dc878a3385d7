// reduce_noalloc.cu — b2r_reduce_local_noalloc: the single-GPU reduction for callers that must not allocate.
//
// b2r_reduce_local needs a b2r_workspace (block partials + an arrival ticket), and creating one calls cudaMalloc. Some
// callers cannot afford that call at all:
//   * processes that have handed every byte of device memory to another allocator (a framework's caching pool, a
//     fixed arena sized at start-up), where a late cudaMalloc can fail or fragment the pool;
//   * latency-critical loops that may not execute anything that synchronises the device — cudaMalloc and cudaFree do;
//   * restricted contexts such as library hooks and callbacks whose contract forbids allocation calls.
// For them the scratch lives in the library image itself: a file-scope __device__ array below. Every GPU that loads
// libb2r.so carries these 8448 bytes until the process exits, whether or not the entry point is ever used.
//
// Rules that follow from sharing ONE array per device:
//   * Limit: the array holds B2R_NOALLOC_MAX_PARTIALS (1024) block partials of 8 bytes plus the ticket. A launch
//     geometry that needs more partials (cfg->ctas / cluster > 1024) is rejected with B2R_ERR_INVALID; when the caller leaves
//     cfg->ctas at 0 the planner's choice is capped at 1024, so every element count is served.
//   * The array keeps whatever the previous call left in it. The only word that must start at zero is the ticket; it is
//     cleared with a 4-byte cudaMemsetAsync on the caller's stream before every launch (partials are fully overwritten
//     before they are read).
//   * Two calls that overlap on the same GPU (different streams, or different host threads on different streams) race on
//     the partials and the ticket and produce garbage. Serialise them on one stream, or use b2r_reduce_local with one
//     workspace per stream.
//   * The first call on each device performs a one-time set-up with that device current: it asks the runtime for the
//     array's device address (cudaGetSymbolAddress) and for the SM count. Make that first call before any stream
//     capture is begun; later calls only enqueue the memset and the kernel.
// Only the register-streaming kernel variants (ldg128, ldg256) are offered here.
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>

#include "../util/check.h"
#include "b2r.h"
#include "reduce_internal.h"

namespace {

constexpr size_t kPartialBytes = (size_t)B2R_NOALLOC_MAX_PARTIALS * 8;  // acc_t is at most 8 bytes (fp64)
constexpr size_t kTicketOffset = kPartialBytes;                        // 256-byte aligned: 8192
constexpr size_t kScratchBytes = kPartialBytes + 256;

// The scratch itself. In HOST code this name is only a stand-in handle for the runtime — it is not a device address and
// must never be passed to a kernel; the address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
struct PerDevice {
  bool ready = false;
  unsigned char* scratch = nullptr;  // device address of g_noalloc_scratch on this device
  int sm_count = 0;
};
PerDevice g_dev[kMaxDevices];
std::mutex g_mu;

// One-time set-up for the CURRENT device.
b2r_status_t setup_current_device(int dev, PerDevice** out) {
  std::lock_guard<std::mutex> lk(g_mu);
  PerDevice& d = g_dev[dev];
  if (!d.ready) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr)
      return b2r::fail(B2R_ERR_CUDA, "b2r_reduce_local_noalloc: cudaGetSymbolAddress on device %d failed: %s", dev, cudaGetErrorString(e));
    int sm = 0;
    e = cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return b2r::fail(B2R_ERR_CUDA, "b2r_reduce_local_noalloc: SM count of device %d: %s", dev, cudaGetErrorString(e));
    d.scratch = (unsigned char*)ptr;
    d.sm_count = sm;
    d.ready = true;
  }
  *out = &d;
  return B2R_OK;
}

}  // namespace

extern "C" b2r_status_t b2r_reduce_local_noalloc(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, void* stream_,
                                                 const b2r_launch_cfg_t* cfg) {
  using namespace b2r;
  cudaStream_t stream = (cudaStream_t)stream_;
  int dev = -1;
  B2R_CUDA_CHECK(cudaGetDevice(&dev));
  B2R_REQUIRE(dev >= 0 && dev < kMaxDevices, "b2r_reduce_local_noalloc: device ordinal %d not supported (max %d)", dev, kMaxDevices - 1);
  PerDevice* d = nullptr;
  B2R_CHECK(setup_current_device(dev, &d));

  b2r_launch_cfg_t c;
  memset(&c, 0, sizeof(c));
  if (cfg) c = *cfg;
  if (c.variant == B2R_VARIANT_AUTO) c.variant = B2R_VARIANT_LDG128;
  B2R_REQUIRE(c.variant == B2R_VARIANT_LDG128 || c.variant == B2R_VARIANT_LDG256,
              "b2r_reduce_local_noalloc: only the ldg128 / ldg256 variants are available without a workspace (got variant %d)", c.variant);
  // Consecutive calls share the scratch, so they must not overlap: programmatic dependent launch is forced OFF here
  // (a negative value is the one the planner honours as "off"; 0 would mean "whatever B2R_PDL says"), and so is the
  // dynamically scheduled tail (its per-tile partials could outgrow the 1024 entries of the static array).
  c.pdl = B2R_PDL_OFF;
  c.dyn_pct = -1;

  // a workspace view over the static array: nothing is allocated, nothing outlives this call
  b2r_workspace view;
  view.device = dev;
  view.sm_count = d->sm_count;
  view.d_partials = d->scratch;
  view.d_ticket = (unsigned*)(d->scratch + kTicketOffset);
  view.d_tile_ctr = nullptr;
  view.d_acc64 = nullptr;
  view.max_partials = B2R_NOALLOC_MAX_PARTIALS;
  view.d_trace = nullptr;
  view.trace_enabled = false;

  LaunchGeom g;
  LocalParams p;
  if (c.ctas == 0) {
    // the planner's own choice may exceed what the static array holds (up to 8 CTAs per SM on a B200 = 1184 partials):
    // cap it here; only a geometry the CALLER asked for is rejected below
    int stages = 0, tile_vecs = 0;
    B2R_CHECK(plan_local(count, dtype, op, view.sm_count, &c, &g, &stages, &tile_vecs));
    const int cl = g.cluster > 0 ? g.cluster : 1;
    if (g.ctas / cl > B2R_NOALLOC_MAX_PARTIALS) c.ctas = B2R_NOALLOC_MAX_PARTIALS * cl;
  }
  B2R_CHECK(prepare_local(in, out, count, dtype, op, &view, &c, 0, &g, &p));
  const int partials = g.ctas / (g.cluster > 0 ? g.cluster : 1);
  if (partials > B2R_NOALLOC_MAX_PARTIALS)
    return fail(B2R_ERR_INVALID,
                "b2r_reduce_local_noalloc: this launch needs %d block partials but the static scratch holds %d (%zu bytes); lower cfg->ctas "
                "or use b2r_reduce_local with a workspace",
                partials, B2R_NOALLOC_MAX_PARTIALS, kScratchBytes);
  if (partials > 1) B2R_CUDA_CHECK(cudaMemsetAsync(view.d_ticket, 0, sizeof(unsigned), stream));  // the one word that must start at zero
  return launch_any(dtype, (int)op, g, p, stream);
}
