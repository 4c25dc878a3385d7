// reduce_local.cu — host side of the single-GPU reduction: workspace, launch planning, public entry points.
// Counterpart of the reference's host wrappers sumreduce/minreduce/maxreduce (reduction_kernel.cu:263-524) and of
// getNumBlocksAndThreads (reduction.cpp:272-291). Differences by design: no power-of-two restriction, 64-bit
// counts, one launch instead of a multi-pass loop, geometry sized to the machine (k CTAs per SM, persistent).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../util/check.h"
#include "../util/env.h"
#include "../util/nvtx.h"
#include "b2r.h"
#include "reduce_kernels.cuh"
#include "reduce_internal.h"
#include "workspace.h"

namespace b2r {

namespace {

struct Tuned {
  int variant, threads, ctas_per_sm, cluster, unroll, stages, tile_bytes, dyn_pct;
};

// Defaults measured on B200 (profiles/local_sweep.md); every field can be overridden per call through
// b2r_launch_cfg_t or globally through B2R_VARIANT / B2R_THREADS / B2R_CTAS_PER_SM / B2R_CLUSTER / B2R_UNROLL /
// B2R_STAGES / B2R_PDL.
//  * >= 256 MiB: the bulk-copy (TMA) ring wins by ~4% (7.0 vs 6.76 TB/s at 1 GiB): 4 consumer warps, 3 CTAs/SM,
//    4 stages of 8 KiB -> 96 KiB in flight per SM without spending registers on it.
//  * smaller inputs are dominated by launch/ramp/tail (~7 us fixed): plain stream order likes many resident
//    threads (512 x 4 per SM); with programmatic dependent launch ONE CTA per SM is best, because the next
//    kernel's CTAs can become resident while this one drains (9.25 us vs 16.4 us per 64 MiB launch).
//  * `early` = programmatic launch with early streaming (cfg.pdl = 2); the safe flavour (wait before the first load) does
//    not overlap the streaming of consecutive kernels, so it takes the plain-order geometry.
//  * plain order at 8 MiB - 256 MiB: the last quarter of the input is claimed tile by tile (reduce_ldg_body<DYN>), which
//    removes the tail skew between CTAs (profiles/single_gpu.md, "dynamic tail").
Tuned tuned_defaults(b2r_dtype_t dtype, b2r_op_t op, uint64_t bytes, int early) {
  (void)op;
  (void)dtype;
  const int pdl = early;
  Tuned t;
  t.cluster = 1;
  t.unroll = 4;
  t.tile_bytes = 0;
  t.dyn_pct = 0;
  if (bytes >= (256ull << 20)) {
    t.variant = B2R_VARIANT_TMA;
    t.threads = 160;
    t.ctas_per_sm = 3;
    t.stages = 4;
  } else if (pdl) {
    t.variant = B2R_VARIANT_TMA;
    t.threads = 288;
    t.ctas_per_sm = 1;
    t.stages = 4;
  } else {
    // plain order / safe programmatic launch. 256 threads x 8 CTAs per SM + one prefetched line per thread measured best at
    // 64 MiB (10.8 us per launch; 512 x 4 with two lines: 11.4; no prefetch: 12.2; no programmatic launch at all: 16.4)
    t.variant = B2R_VARIANT_LDG128;
    t.threads = 256;
    t.ctas_per_sm = 8;
    t.stages = 4;
    if (bytes >= (8ull << 20)) t.dyn_pct = 25;
  }
  return t;
}

}  // namespace

b2r_status_t plan_local(uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int sm_count, const b2r_launch_cfg_t* hint,
                        LaunchGeom* g, int* stages, int* tile_vecs) {
  const uint64_t bytes = count * b2r_dtype_size(dtype);
  // PDL flavour: the environment can only ask for the safe one (wait before the first load); early streaming is a promise
  // about the input that only the caller can make, per call (b2r_launch_cfg_t.pdl = B2R_PDL_INPUT_READY)
  // The safe flavour is the DEFAULT (B2R_PDL=0 switches it off): it is correct after any producer, and back-to-back launches
  // stop paying the full kernel boundary (64 MiB int32 SUM: 16.4 us per launch without it, 13.8 with it).
  int pdl = B2R_ENV_INT("B2R_PDL", 1) > 0 ? B2R_PDL_SAFE : 0;  // cached: see util/env.h
  if (hint && hint->pdl) pdl = hint->pdl < 0 ? 0 : (hint->pdl >= B2R_PDL_INPUT_READY ? B2R_PDL_INPUT_READY : B2R_PDL_SAFE);
  const int early = pdl == B2R_PDL_INPUT_READY;
  Tuned t = tuned_defaults(dtype, op, bytes, early);
  if (hint && hint->variant && hint->variant != t.variant) {
    // a forced variant keeps its own sensible geometry
    if (hint->variant == B2R_VARIANT_TMA) { t.threads = early ? 288 : 160; t.ctas_per_sm = early ? 1 : 3; t.stages = 4; t.dyn_pct = 0; }
    else { t.threads = early ? 512 : 256; t.ctas_per_sm = early ? 1 : 8; }
  }
  t.variant = B2R_ENV_INT("B2R_VARIANT", t.variant);
  t.threads = B2R_ENV_INT("B2R_THREADS", t.threads);
  t.ctas_per_sm = B2R_ENV_INT("B2R_CTAS_PER_SM", t.ctas_per_sm);
  t.cluster = B2R_ENV_INT("B2R_CLUSTER", t.cluster);
  t.unroll = B2R_ENV_INT("B2R_UNROLL", t.unroll);
  t.stages = B2R_ENV_INT("B2R_STAGES", t.stages);
  t.dyn_pct = B2R_ENV_INT("B2R_DYN_PCT", t.dyn_pct);

  int variant = t.variant, threads = t.threads, cluster = t.cluster, unroll = t.unroll;
  int st = t.stages, tile_bytes = t.tile_bytes, dyn_pct = t.dyn_pct;
  int ctas = 0;
  if (hint) {
    if (hint->dyn_pct) dyn_pct = hint->dyn_pct;
    if (hint->variant) variant = hint->variant;
    if (hint->threads) threads = hint->threads;
    if (hint->cluster) cluster = hint->cluster;
    if (hint->unroll) unroll = hint->unroll;
    if (hint->stages) st = hint->stages;
    if (hint->tile_bytes) tile_bytes = hint->tile_bytes;
    ctas = hint->ctas;
  }
  B2R_REQUIRE(variant >= B2R_VARIANT_LDG128 && variant <= B2R_VARIANT_TMA, "plan: bad variant %d", variant);
  B2R_REQUIRE(threads >= 32 && threads <= 512 && threads % 32 == 0, "plan: threads=%d must be a multiple of 32 in [32,512]", threads);
  B2R_REQUIRE(cluster == 1 || cluster == 2 || cluster == 4 || cluster == 8, "plan: cluster=%d must be 1,2,4 or 8", cluster);
  B2R_REQUIRE(unroll == 4 || unroll == 8, "plan: unroll=%d must be 4 or 8", unroll);
  if (variant == B2R_VARIANT_LDG256) unroll = 4;
  if (variant == B2R_VARIANT_TMA) {
    B2R_REQUIRE(threads >= 64, "plan: the TMA variant needs a producer warp and at least one consumer warp");
    B2R_REQUIRE(st >= 1 && st <= 32, "plan: stages=%d out of range", st);
    tile_bytes = (threads - 32) * 16 * kTmaVecsPerThread;  // derived: every consumer thread owns 4 vectors of a tile
    B2R_REQUIRE((uint64_t)st * tile_bytes + 16ull * st <= 220 * 1024, "plan: ring of %d x %d bytes exceeds shared memory", st, tile_bytes);
  }

  const bool explicit_ctas = ctas > 0;
  if (!explicit_ctas) {
    // Small inputs: one CTA per ~32 KiB so that the ticket stage does not dominate; a single CTA (no ticket, no
    // partials) below 64 KiB. Large inputs: persistent grid, ctas_per_sm CTAs on every SM.
    const uint64_t min_bytes_per_cta = 32 * 1024;
    uint64_t want = (bytes + min_bytes_per_cta - 1) / min_bytes_per_cta;
    if (bytes <= 64 * 1024) want = 1;
    int cps = std::max(1, t.ctas_per_sm);
    if (variant == B2R_VARIANT_TMA) cps = std::max(1, std::min(cps, (220 * 1024) / (st * tile_bytes + 16 * st + 1024)));
    cps = std::max(1, std::min(cps, 2048 / threads));
    uint64_t cap = (uint64_t)sm_count * (uint64_t)cps;
    ctas = (int)std::max<uint64_t>(1, std::min<uint64_t>(want, cap));
    if (ctas == 1) {
      cluster = 1;
      if (bytes <= 8 * 1024) threads = std::min(threads, 128);
      else threads = std::min(threads, 256);
      if (variant == B2R_VARIANT_TMA) threads = std::max(threads, 64);
    }
  }
  B2R_REQUIRE(ctas >= 1 && ctas <= kMaxCtas, "plan: ctas=%d out of range [1,%d]", ctas, kMaxCtas);
  if (cluster > 1) {
    if (ctas < cluster) cluster = 1;
    else ctas -= ctas % cluster;
  }
  g->variant = variant;
  g->threads = threads;
  g->ctas = ctas;
  g->cluster = cluster;
  g->unroll = unroll;
  g->pdl = pdl;
  // the dynamic tail: register-streaming variants, no clusters, several CTAs, and never under early streaming (the tile
  // counter lives in the workspace, which an overlapped launch may not touch before its griddepcontrol.wait)
  if (dyn_pct < 0 || variant == B2R_VARIANT_TMA || cluster > 1 || ctas < 2 || early) dyn_pct = 0;
  g->dyn_pct = std::min(dyn_pct, 90);
  if (variant == B2R_VARIANT_TMA) {
    // the single-CTA path above may have lowered `threads`: the tile MUST follow (every consumer thread owns exactly
    // kTmaVecsPerThread vectors of a full tile; a stale, larger tile would leave part of each full tile unread)
    tile_bytes = (threads - 32) * 16 * kTmaVecsPerThread;
  }
  g->smem_bytes = variant == B2R_VARIANT_TMA ? st * tile_bytes + 16 * st : 0;
  *stages = st;
  *tile_vecs = tile_bytes / 16;
  return B2R_OK;
}

b2r_status_t launch_any(b2r_dtype_t dtype, int op, const LaunchGeom& g, const LocalParams& p, cudaStream_t stream) {
  switch (dtype) {
    case B2R_INT32: return launch_local<B2R_INT32>(op, g, p, stream);
    case B2R_FLOAT32: return launch_local<B2R_FLOAT32>(op, g, p, stream);
    case B2R_FLOAT64: return launch_local<B2R_FLOAT64>(op, g, p, stream);
    case B2R_BFLOAT16: return launch_local<B2R_BFLOAT16>(op, g, p, stream);
  }
  return fail(B2R_ERR_INVALID, "bad dtype %d", (int)dtype);
}

// Fills head/nvec/tail for a given vector width.
void split_edges(const void* in, uint64_t count, size_t esz, unsigned vec_bytes, LocalParams* p) {
  const uintptr_t addr = (uintptr_t)in;
  uint64_t head = ((vec_bytes - (addr % vec_bytes)) % vec_bytes) / esz;
  if (head > count) head = count;
  const uint64_t body_bytes = (count - head) * esz;
  p->head = head;
  p->nvec = body_bytes / vec_bytes;
  p->tail = count - head - p->nvec * (vec_bytes / esz);
}

// Validates the arguments, plans the launch and fills the kernel parameters (no launch).
b2r_status_t prepare_local(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, b2r_workspace* ws,
                           const b2r_launch_cfg_t* cfg, int out_as_acc, LaunchGeom* g, LocalParams* pp) {
  B2R_REQUIRE(ws != nullptr, "b2r_reduce_local: workspace is NULL");
  B2R_REQUIRE(out != nullptr, "b2r_reduce_local: out is NULL");
  B2R_REQUIRE(in != nullptr || count == 0, "b2r_reduce_local: in is NULL");
  B2R_REQUIRE(dtype >= B2R_INT32 && dtype <= B2R_BFLOAT16, "b2r_reduce_local: bad dtype %d", (int)dtype);
  B2R_REQUIRE(op >= B2R_SUM && op <= B2R_MAX, "b2r_reduce_local: bad op %d", (int)op);
  const size_t esz = b2r_dtype_size(dtype);
  B2R_REQUIRE(((uintptr_t)in % esz) == 0, "b2r_reduce_local: input pointer %p is not aligned to the element size %zu", in, esz);
  LocalParams& p = *pp;
  memset(&p, 0, sizeof(p));
  int stages = 0, tile_vecs = 0;
  B2R_CHECK(plan_local(count, dtype, op, ws->sm_count, cfg, g, &stages, &tile_vecs));
  p.in = in;
  p.out = out;
  p.n = count;
  split_edges(in, count, esz, g->variant == B2R_VARIANT_LDG256 ? 32u : 16u, &p);
  p.partials = ws->d_partials;
  p.ticket = ws->d_ticket;
  p.trace = nullptr;
  if (ws->trace_enabled && ws->d_trace) {  // the next slot of the trace ring
    const unsigned slot = ws->trace_seq++ % kTraceRing;
    p.trace = ws->d_trace + (size_t)slot * kMaxCtas * 4;
    ws->trace_ctas[slot] = g->ctas;
  }
  p.stages = stages;
  p.tile_vecs = tile_vecs;
  p.out_as_acc = out_as_acc;
  p.acc64 = B2R_ENV_INT("B2R_PACKED_SUM", 1) ? ws->d_acc64 : nullptr;  // int32 SUM: grid stage in one atomic (finish_grid_packed)
  // safe programmatic launch: lines per thread pulled into L2 while the previous kernel finishes (one line x 256 threads = 32 KiB
  // per CTA; more than that competes with the streaming loop of the kernel that is still running and costs more than it saves)
  p.prefetch_lines = g->pdl == B2R_PDL_SAFE ? (unsigned)std::max(0, std::min(64, B2R_ENV_INT("B2R_PREFETCH_LINES", 1))) : 0u;
  p.dyn_first = p.nvec;
  if (g->dyn_pct > 0) {
    // tiles of one load round of the CTA (threads x unroll vectors: 32 KiB for 512 x 4 x 16 B), grown when the per-tile partials
    // of a floating-point launch would not fit the workspace; a body too small for one tile per CTA stays static
    const uint64_t round_vecs = (uint64_t)g->threads * g->unroll;
    const uint64_t want = p.nvec / 100 * g->dyn_pct + (p.nvec % 100) * g->dyn_pct / 100;
    const uint64_t max_tiles = (uint64_t)std::max(1, ws->max_partials - g->ctas);
    uint64_t tile_vecs = std::max<uint64_t>(1, (uint64_t)std::max(1, B2R_ENV_INT("B2R_DYN_TILE_ROUNDS", 1))) * round_vecs;
    while ((want + tile_vecs - 1) / tile_vecs > max_tiles) tile_vecs += round_vecs;
    const uint64_t ntiles = want / tile_vecs;
    if (ntiles >= (uint64_t)g->ctas && ws->d_tile_ctr != nullptr && tile_vecs <= 0xffffffffull) {
      p.tile_ctr = ws->d_tile_ctr;
      p.dyn_tile_vecs = (unsigned)tile_vecs;
      p.dyn_ntiles = (unsigned)ntiles;
      p.dyn_first = p.nvec - ntiles * tile_vecs;
      p.dyn_first -= p.dyn_first % (128 / (g->variant == B2R_VARIANT_LDG256 ? 32u : 16u));  // 128-byte aligned tiles; the last one absorbs the slack
    } else {
      g->dyn_pct = 0;
    }
  }
  // never overlap with the previous launch if this input could be that launch's output scalar
  if (g->pdl && ws->last_out) {
    const char* lo = (const char*)in;
    const char* hi = lo + count * esz;
    const char* o = (const char*)ws->last_out;
    if (o + 8 > lo && o < hi) g->pdl = 0;
  }
  p.pdl = g->pdl;
  ws->last_out = out;
  ws->last_ctas = g->ctas;
  ws->last_geom[0] = g->variant;
  ws->last_geom[1] = g->threads;
  ws->last_geom[2] = g->ctas;
  ws->last_geom[3] = g->cluster;
  ws->last_dyn_tiles = (int)p.dyn_ntiles;
  return B2R_OK;
}

b2r_status_t reduce_local_impl(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                               b2r_workspace* ws, cudaStream_t stream, const b2r_launch_cfg_t* cfg, int out_as_acc,
                               const XgCall* xg) {
  NvtxRange nvtx(xg ? "b2r_reduce_scalar_sharded" : "b2r_reduce_local");
  B2R_REQUIRE(ws != nullptr, "b2r_reduce_local: workspace is NULL");
  // launch on the workspace's device whatever the caller's current device is (cheap when they already agree)
  int cur = -1;
  B2R_CUDA_CHECK(cudaGetDevice(&cur));
  struct DeviceGuard {
    int prev;
    bool active;
    ~DeviceGuard() { if (active) cudaSetDevice(prev); }
  } guard{cur, cur != ws->device};
  if (guard.active) B2R_CUDA_CHECK(cudaSetDevice(ws->device));
  LaunchGeom g;
  LocalParams p;
  B2R_CHECK(prepare_local(in, out, count, dtype, op, ws, cfg, out_as_acc, &g, &p));
  if (xg) {
    p.xg_ctx = xg->ctx;
    p.xg_root = xg->root;
    p.xg_all = xg->all;
    p.xg_seq = xg->seq;  // PDL stays as planned: the mailbox is only touched after griddepcontrol.wait
  }
  return launch_any(dtype, (int)op, g, p, stream);
}

b2r_status_t launch_multi_any(b2r_dtype_t dtype, int op, int nranks, int ctas, int threads, const MultiLocalParams& mp, cudaStream_t stream) {
  switch (dtype) {
    case B2R_INT32: return launch_local_multi<B2R_INT32>(op, nranks, ctas, threads, mp, stream);
    case B2R_FLOAT32: return launch_local_multi<B2R_FLOAT32>(op, nranks, ctas, threads, mp, stream);
    case B2R_FLOAT64: return launch_local_multi<B2R_FLOAT64>(op, nranks, ctas, threads, mp, stream);
    case B2R_BFLOAT16: return launch_local_multi<B2R_BFLOAT16>(op, nranks, ctas, threads, mp, stream);
  }
  return fail(B2R_ERR_INVALID, "bad dtype %d", (int)dtype);
}

// Accumulator-precision scalar (fp32 for bf16) -> one element of `dtype` (bf16: round to nearest even, NaN kept quiet).
void acc_to_dtype(const void* acc, b2r_dtype_t dtype, void* out) {
  if (dtype == B2R_BFLOAT16) {
    uint32_t u;
    memcpy(&u, acc, 4);
    uint16_t b;
    if ((u & 0x7fffffffu) > 0x7f800000u) b = (uint16_t)((u >> 16) | 0x40u);
    else b = (uint16_t)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);
    memcpy(out, &b, 2);
  } else {
    memcpy(out, acc, b2r_dtype_size(dtype));
  }
}

// The streaming half of the host -> device pipelines (b2r_reduce_from_host, b2r_reduce_scalar_sharded_from_host): on return
// the per-chunk scalars of the host array, in accumulator precision, are being written to ws->d_chunk_out[0, nchunks) by
// work queued on `stream`; nothing has been synchronised. The caller's current device must be ws->device.
b2r_status_t stream_chunks_from_host(const void* host_in, uint64_t count, b2r_dtype_t dtype, b2r_op_t op, b2r_workspace* ws,
                                     cudaStream_t stream, uint64_t* nchunks_out) {
  B2R_REQUIRE(host_in != nullptr || count == 0, "reduce from host: host_in is NULL");
  B2R_REQUIRE(dtype >= B2R_INT32 && dtype <= B2R_BFLOAT16, "bad dtype %d", (int)dtype);
  B2R_REQUIRE(op >= B2R_SUM && op <= B2R_MAX, "bad op %d", (int)op);
  const size_t esz = b2r_dtype_size(dtype);
  const size_t acc_sz = dtype == B2R_FLOAT64 ? 8 : 4;
  const uint64_t chunk_bytes = (uint64_t)std::max(1, B2R_ENV_INT("B2R_H2D_CHUNK_MB", 8)) << 20;
  const uint64_t chunk_elems = chunk_bytes / esz;
  const uint64_t nchunks = count == 0 ? 0 : (count + chunk_elems - 1) / chunk_elems;
  B2R_REQUIRE(nchunks <= kMaxChunks, "reduce from host: %llu chunks > %d; raise B2R_H2D_CHUNK_MB", (unsigned long long)nchunks, kMaxChunks);

  if (!ws->copy_stream) {
    B2R_CUDA_CHECK(cudaStreamCreateWithFlags(&ws->copy_stream, cudaStreamNonBlocking));
    B2R_CUDA_CHECK(cudaMalloc(&ws->d_chunk_out, (size_t)(kMaxChunks + 2) * 8));  // + the final scalar of the sharded flavour
    B2R_CUDA_CHECK(cudaMallocHost(&ws->h_chunk_out, (size_t)(kMaxChunks + 2) * 8));
    for (int i = 0; i < kStageBufs; ++i) {
      B2R_CUDA_CHECK(cudaEventCreateWithFlags(&ws->ev_copied[i], cudaEventDisableTiming));
      B2R_CUDA_CHECK(cudaEventCreateWithFlags(&ws->ev_consumed[i], cudaEventDisableTiming));
    }
  }
  if (ws->stage_bytes < chunk_bytes) {
    for (int i = 0; i < kStageBufs; ++i) {
      if (ws->d_stage[i]) B2R_CUDA_CHECK(cudaFree(ws->d_stage[i]));
      ws->d_stage[i] = nullptr;
      B2R_CUDA_CHECK(cudaMalloc(&ws->d_stage[i], chunk_bytes));
    }
    ws->stage_bytes = chunk_bytes;
  }

  // the copy stream must not run ahead of work already queued on `stream` that may still use the staging buffers
  B2R_CUDA_CHECK(cudaEventRecord(ws->ev_consumed[0], stream));
  B2R_CUDA_CHECK(cudaStreamWaitEvent(ws->copy_stream, ws->ev_consumed[0], 0));
  for (uint64_t c = 0; c < nchunks; ++c) {
    const int b = (int)(c % kStageBufs);
    const uint64_t first = c * chunk_elems;
    const uint64_t n = std::min(chunk_elems, count - first);
    if (c >= (uint64_t)kStageBufs) B2R_CUDA_CHECK(cudaStreamWaitEvent(ws->copy_stream, ws->ev_consumed[b], 0));
    B2R_CUDA_CHECK(cudaMemcpyAsync(ws->d_stage[b], (const char*)host_in + first * esz, n * esz, cudaMemcpyHostToDevice,
                                   ws->copy_stream));
    B2R_CUDA_CHECK(cudaEventRecord(ws->ev_copied[b], ws->copy_stream));
    B2R_CUDA_CHECK(cudaStreamWaitEvent(stream, ws->ev_copied[b], 0));
    B2R_CHECK(reduce_local_impl(ws->d_stage[b], (char*)ws->d_chunk_out + c * acc_sz, n, dtype, op, ws, stream, nullptr, 1, nullptr));
    B2R_CUDA_CHECK(cudaEventRecord(ws->ev_consumed[b], stream));
  }
  *nchunks_out = nchunks;
  return B2R_OK;
}

}  // namespace b2r

using namespace b2r;

extern "C" {

b2r_status_t b2r_workspace_create(b2r_workspace_t* out, int device) {
  B2R_REQUIRE(out != nullptr, "b2r_workspace_create: out is NULL");
  int prev = 0;
  B2R_CUDA_CHECK(cudaGetDevice(&prev));
  B2R_CUDA_CHECK(cudaSetDevice(device));
  b2r_workspace* ws = new b2r_workspace();
  ws->device = device;
  cudaError_t e = cudaDeviceGetAttribute(&ws->sm_count, cudaDevAttrMultiProcessorCount, device);
  if (e == cudaSuccess) e = cudaMalloc(&ws->d_partials, (size_t)kMaxCtas * 8);
  if (e == cudaSuccess) e = cudaMalloc((void**)&ws->d_ticket, 256);
  if (e == cudaSuccess) e = cudaMemset(ws->d_ticket, 0, 256);
  if (e == cudaSuccess) ws->d_tile_ctr = ws->d_ticket + 32;  // its own 128-byte line of the same allocation
  if (e == cudaSuccess) ws->d_acc64 = (unsigned long long*)(ws->d_ticket + 16);  // int32 SUM: packed arrivals + sum
  ws->max_partials = kMaxCtas;
  if (e == cudaSuccess) e = cudaMalloc((void**)&ws->d_trace, (size_t)kTraceRing * kMaxCtas * 4 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMemset(ws->d_trace, 0, (size_t)kTraceRing * kMaxCtas * 4 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaSetDevice(prev);
  if (e != cudaSuccess) {
    b2r_workspace_destroy(ws);
    return fail(B2R_ERR_CUDA, "b2r_workspace_create(device %d): %s", device, cudaGetErrorString(e));
  }
  *out = ws;
  return B2R_OK;
}

b2r_status_t b2r_workspace_destroy(b2r_workspace_t ws) {
  if (!ws) return B2R_OK;
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(ws->device);
  if (ws->d_partials) cudaFree(ws->d_partials);
  if (ws->d_ticket) cudaFree(ws->d_ticket);
  if (ws->d_trace) cudaFree(ws->d_trace);
  for (int i = 0; i < kStageBufs; ++i) {
    if (ws->d_stage[i]) cudaFree(ws->d_stage[i]);
    if (ws->ev_copied[i]) cudaEventDestroy(ws->ev_copied[i]);
    if (ws->ev_consumed[i]) cudaEventDestroy(ws->ev_consumed[i]);
  }
  if (ws->d_chunk_out) cudaFree(ws->d_chunk_out);
  if (ws->h_chunk_out) cudaFreeHost(ws->h_chunk_out);
  if (ws->copy_stream) cudaStreamDestroy(ws->copy_stream);
  cudaSetDevice(prev);
  delete ws;
  return B2R_OK;
}

b2r_status_t b2r_reduce_local(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                              b2r_workspace_t ws, void* stream) {
  return reduce_local_impl(in, out, count, dtype, op, ws, (cudaStream_t)stream, nullptr, 0, nullptr);
}

b2r_status_t b2r_reduce_local_ex(const void* in, void* out, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                                 b2r_workspace_t ws, void* stream, const b2r_launch_cfg_t* cfg) {
  return reduce_local_impl(in, out, count, dtype, op, ws, (cudaStream_t)stream, cfg, 0, nullptr);
}

void b2r_refresh_env(void) { env_refresh(); }

b2r_status_t b2r_plan_local(uint64_t count, b2r_dtype_t dtype, b2r_op_t op, int device, const b2r_launch_cfg_t* hint,
                            b2r_launch_cfg_t* chosen) {
  B2R_REQUIRE(chosen != nullptr, "b2r_plan_local: chosen is NULL");
  B2R_REQUIRE(dtype >= B2R_INT32 && dtype <= B2R_BFLOAT16, "b2r_plan_local: bad dtype %d", (int)dtype);
  env_refresh();  // the introspection call always looks at the environment as it is NOW (and the launch path follows)
  int sm = 148;  // B200; used when no device is present (planning is a pure function of the SM count)
  if (device >= 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && v > 0) sm = v;
    else cudaGetLastError();
  }
  LaunchGeom g;
  int stages = 0, tile_vecs = 0;
  B2R_CHECK(plan_local(count, dtype, op, sm, hint, &g, &stages, &tile_vecs));
  memset(chosen, 0, sizeof(*chosen));
  chosen->variant = g.variant;
  chosen->threads = g.threads;
  chosen->ctas = g.ctas;
  chosen->cluster = g.cluster;
  chosen->unroll = g.unroll;
  chosen->stages = stages;
  chosen->tile_bytes = tile_vecs * 16;
  chosen->pdl = g.pdl;
  chosen->dyn_pct = g.dyn_pct;
  return B2R_OK;
}

b2r_status_t b2r_workspace_trace_enable(b2r_workspace_t ws, int enable) {
  B2R_REQUIRE(ws != nullptr, "workspace is NULL");
  ws->trace_enabled = enable != 0;
  return B2R_OK;
}

b2r_status_t b2r_workspace_trace_read_at(b2r_workspace_t ws, int launches_back, uint64_t* stamps, int max_ctas, int* n_ctas) {
  B2R_REQUIRE(ws != nullptr && stamps != nullptr && n_ctas != nullptr, "b2r_workspace_trace_read: NULL argument");
  B2R_REQUIRE(launches_back >= 0 && launches_back < kTraceRing && (unsigned)launches_back < ws->trace_seq,
              "b2r_workspace_trace_read: launch %d back is not in the ring (%u traced launches, ring of %d)", launches_back, ws->trace_seq, kTraceRing);
  const unsigned slot = (ws->trace_seq - 1u - (unsigned)launches_back) % kTraceRing;
  int n = std::min(ws->trace_ctas[slot], max_ctas);
  int prev = 0;
  B2R_CUDA_CHECK(cudaGetDevice(&prev));
  B2R_CUDA_CHECK(cudaSetDevice(ws->device));
  cudaError_t e = cudaMemcpy(stamps, ws->d_trace + (size_t)slot * kMaxCtas * 4, (size_t)n * 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost);
  cudaSetDevice(prev);
  if (e != cudaSuccess) return fail(B2R_ERR_CUDA, "trace_read: %s", cudaGetErrorString(e));
  *n_ctas = n;
  return B2R_OK;
}

b2r_status_t b2r_workspace_trace_read(b2r_workspace_t ws, uint64_t* stamps, int max_ctas, int* n_ctas) {
  return b2r_workspace_trace_read_at(ws, 0, stamps, max_ctas, n_ctas);
}

// Pipelined host -> device + reduce. The host array is cut into chunks; chunk i is copied by the copy engine on a
// private stream while chunk i-1 is being reduced on `stream`; per-chunk scalars (in accumulator precision) are
// combined on the host in chunk order, so bf16 is still rounded only once.
b2r_status_t b2r_reduce_from_host(const void* host_in, void* out_host, uint64_t count, b2r_dtype_t dtype, b2r_op_t op,
                                  b2r_workspace_t ws, void* stream_) {
  NvtxRange nvtx("b2r_reduce_from_host");
  B2R_REQUIRE(ws != nullptr && out_host != nullptr, "b2r_reduce_from_host: NULL argument");
  cudaStream_t stream = (cudaStream_t)stream_;
  int prev = 0;
  B2R_CUDA_CHECK(cudaGetDevice(&prev));
  B2R_CUDA_CHECK(cudaSetDevice(ws->device));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  uint64_t nchunks = 0;
  B2R_CHECK(stream_chunks_from_host(host_in, count, dtype, op, ws, stream, &nchunks));
  const size_t acc_sz = dtype == B2R_FLOAT64 ? 8 : 4;
  const b2r_dtype_t acc_dtype = dtype == B2R_BFLOAT16 ? B2R_FLOAT32 : dtype;
  if (nchunks) {
    B2R_CUDA_CHECK(cudaMemcpyAsync(ws->h_chunk_out, ws->d_chunk_out, nchunks * acc_sz, cudaMemcpyDeviceToHost, stream));
  }
  B2R_CUDA_CHECK(cudaStreamSynchronize(stream));
  // combine chunk scalars in chunk order (accumulator precision), then round once
  unsigned char acc[8];
  b2r_reduce_host(ws->h_chunk_out, acc, nchunks, acc_dtype, op);
  acc_to_dtype(acc, dtype, out_host);
  return B2R_OK;
}

}  // extern "C"
