// comm.cu — communicators and the symmetric heap.
//
// Every rank owns one heap (control region | send staging | recv staging | user allocations) that is mapped into every
// other rank's address space, so kernels address peer memory with plain loads/stores over NVLink and, when the driver
// exposes multicast, through one extra "multicast" mapping whose loads reduce in the NVSwitch (multimem.*).
//   * kModeRank     one process per GPU (the reference's mpirun model, mpi/reduce.c:32-34): cuMemCreate + POSIX-FD export,
//                   descriptors exchanged over a Unix socket (bootstrap.cpp), cuMemImport + cuMemMap on every rank.
//   * kModeAll      one process, several GPUs (vendored simpleMultiGPU / MonteCarloMultiGPU single-thread mode): the
//                   same VMM mapping without any export.
//   * kModeLoopback several virtual ranks on ONE GPU with cudaMalloc'ed heaps, for protocol tests on a one-GPU box.
// Replaces the vendored simpleP2P recipe (cudaDeviceEnablePeerAccess + UVA pointers, simpleP2P.cu:164-176,250-275).
#include "comm.h"

#include <unistd.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../util/check.h"
#include "driver_api.h"

using namespace b2r;

namespace b2r {

namespace {

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// B2R_FAULT_INJECT="rank:1,delay_us:200000" or "rank:1,skip_signal:1" 
void parse_fault_injection(CollCtx* ctx, int rank) {
  ctx->fault_delay_us = 0;
  ctx->fault_skip_signal = 0;
  const char* s = getenv("B2R_FAULT_INJECT");
  if (!s || !*s) return;
  int frank = -1;
  unsigned delay = 0;
  int skip = 0;
  std::string str(s);
  size_t pos = 0;
  while (pos < str.size()) {
    size_t end = str.find(',', pos);
    if (end == std::string::npos) end = str.size();
    std::string kv = str.substr(pos, end - pos);
    size_t c = kv.find(':');
    if (c != std::string::npos) {
      std::string k = kv.substr(0, c), v = kv.substr(c + 1);
      if (k == "rank") frank = atoi(v.c_str());
      else if (k == "delay_us") delay = (unsigned)strtoul(v.c_str(), nullptr, 10);
      else if (k == "skip_signal") skip = atoi(v.c_str());
    }
    pos = end + 1;
  }
  if (frank == rank) {
    ctx->fault_delay_us = delay;
    ctx->fault_skip_signal = skip;
  }
}

struct Sizes {
  size_t stage_bytes, work_bytes, user_bytes, heap_bytes, ll_big_bytes;
};

Sizes compute_sizes(const b2r_comm_config_t* cfg, size_t granularity) {
  Sizes s;
  s.stage_bytes = (size_t)std::max(1, env_int("B2R_STAGE_MB", 64)) << 20;
  // the work region (tree partials, ring slots, the push algorithm's whole-slice staging: 8/7 of a message per launch at 8
  // GPUs) is sized separately: 320 MiB lets a 256 MiB reduce run as one launch. Same on every rank, like B2R_STAGE_MB.
  s.work_bytes = std::max(s.stage_bytes, (size_t)std::max(1, env_int("B2R_WORK_MB", 320)) << 20);
  s.user_bytes = cfg && cfg->heap_bytes ? (size_t)cfg->heap_bytes : ((size_t)256 << 20);
  s.user_bytes = round_up(s.user_bytes, 256);
  // the large LL region (64 CTAs x 4 KiB per rank and parity; 16 MiB covers 256 KiB messages on up to 16 ranks; 0 = none).
  // Like B2R_STAGE_MB it must be the same on every rank.
  s.ll_big_bytes = (size_t)std::max(0, env_int("B2R_LL_BIG_MB", 16)) << 20;
  s.heap_bytes = round_up(kCtrlBytes + 2 * s.stage_bytes + s.work_bytes + s.user_bytes + s.ll_big_bytes, granularity);
  return s;
}

void fill_common(b2r_comm* c, const b2r_comm_config_t* cfg, const Sizes& s) {
  c->heap_bytes = s.heap_bytes;
  c->stage_bytes = s.stage_bytes;
  c->work_bytes = s.work_bytes;
  c->user_off = kCtrlBytes + 2 * s.stage_bytes + s.work_bytes;
  c->ll_big_bytes = s.ll_big_bytes;
  c->ll_big_off = s.ll_big_bytes ? c->user_off + s.user_bytes : 0;
  c->user_end = s.ll_big_bytes ? c->ll_big_off : s.heap_bytes;
  c->symmetric_only = cfg && cfg->symmetric_only;
  c->bump = c->user_off;
  c->timeout_ms = cfg && cfg->timeout_ms ? cfg->timeout_ms : env_int("B2R_TIMEOUT_MS", 10000);
  c->ctx.timeout_ns = c->timeout_ms < 0 ? 0ull : (unsigned long long)c->timeout_ms * 1000000ull;
  c->max_ctas = cfg && cfg->max_ctas ? cfg->max_ctas : env_int("B2R_COLL_CTAS", 0);
  parse_fault_injection(&c->ctx, c->rank);
}

// zero my control region, publish the device copy of the context, create the scalar workspace
b2r_status_t finish_init(b2r_comm* c) {
  int prev = 0;
  B2R_CUDA_CHECK(cudaGetDevice(&prev));
  B2R_CUDA_CHECK(cudaSetDevice(c->device));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  B2R_CUDA_CHECK(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, c->device));
  B2R_CUDA_CHECK(cudaMemset(c->my_heap(), 0, kCtrlBytes));
  if (c->ll_big_bytes) B2R_CUDA_CHECK(cudaMemset(c->my_heap() + c->ll_big_off, 0, c->ll_big_bytes));  // LL flags must start at zero
  B2R_CUDA_CHECK(cudaMemcpy(c->my_heap() + kOffCtx, &c->ctx, sizeof(CollCtx), cudaMemcpyHostToDevice));
  B2R_CUDA_CHECK(cudaDeviceSynchronize());
  B2R_CHECK(b2r_workspace_create(&c->ws, c->device));
  return B2R_OK;
}

b2r_status_t vmm_prop(int device, bool exportable, CUmemAllocationProp* prop) {
  memset(prop, 0, sizeof(*prop));
  prop->type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop->location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop->location.id = device;
  prop->requestedHandleTypes = exportable ? CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR : CU_MEM_HANDLE_TYPE_NONE;
  return B2R_OK;
}

// error path of the init functions: free the members created so far (the heaps and mappings go with the group)
void discard_members(CommGroup* g) {
  for (b2r_comm* m : g->members) {
    if (m->ws) b2r_workspace_destroy(m->ws);
    delete m;
  }
  g->members.clear();
}

bool device_supports_multicast(const DriverApi* d, int device) {
  if (!d->MulticastCreate) return false;
  CUdevice cudev;
  int sup = 0;
  if (d->DeviceGet(&cudev, device) != CUDA_SUCCESS) return false;
  if (d->DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev) != CUDA_SUCCESS) return false;
  return sup != 0;
}

}  // namespace

CommGroup::~CommGroup() {
  const DriverApi* d = (mode == kModeLoopback) ? nullptr : driver_api();
  if (mode == kModeLoopback) {
    for (void* p : loop_heaps)
      if (p) cudaFree(p);
    return;
  }
  if (!d) return;
  if (mc_va) {
    d->MemUnmap(mc_va, va_size / std::max<size_t>(1, handles.size()));
    d->MemAddressFree(mc_va, va_size / std::max<size_t>(1, handles.size()));
  }
  if (mc_handle) d->MemRelease(mc_handle);
  if (va_base) {
    d->MemUnmap(va_base, va_size);
    d->MemAddressFree(va_base, va_size);
  }
  for (auto h : handles)
    if (h) d->MemRelease(h);
}

}  // namespace b2r

extern "C" {

// ------------------------------------------------------------------------------------------------ loopback
b2r_status_t b2r_comm_init_loopback(b2r_comm_t* comms, int nranks, int device, const b2r_comm_config_t* cfg) {
  B2R_REQUIRE(comms != nullptr, "b2r_comm_init_loopback: comms is NULL");
  B2R_REQUIRE(nranks >= 1 && nranks <= 8, "b2r_comm_init_loopback: nranks=%d must be in [1,8]", nranks);
  int prev = 0;
  B2R_CUDA_CHECK(cudaGetDevice(&prev));
  B2R_CUDA_CHECK(cudaSetDevice(device));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  Sizes s = compute_sizes(cfg, 256);
  auto group = std::make_shared<CommGroup>();
  group->nranks = nranks;
  group->mode = kModeLoopback;
  group->loop_heaps.assign(nranks, nullptr);
  for (int r = 0; r < nranks; ++r) B2R_CUDA_CHECK(cudaMalloc(&group->loop_heaps[r], s.heap_bytes));
  group->pending.assign(nranks, nullptr);
  for (int r = 0; r < nranks; ++r) {
    b2r_comm* c = new b2r_comm();
    c->mode = kModeLoopback;
    c->rank = r;
    c->nranks = nranks;
    c->device = device;
    memset(&c->ctx, 0, sizeof(c->ctx));
    for (int p = 0; p < nranks; ++p) c->ctx.heap[p] = (char*)group->loop_heaps[p];
    c->ctx.rank = r;
    c->ctx.nranks = nranks;
    fill_common(c, cfg, s);
    c->group = group;
    group->members.push_back(c);
    b2r_status_t st = finish_init(c);
    if (st != B2R_OK) {
      discard_members(group.get());
      for (int k = 0; k < r; ++k) comms[k] = nullptr;  // nothing usable is handed back
      return st;
    }
    comms[r] = c;
  }
  group->alive = nranks;
  return B2R_OK;
}

// ------------------------------------------------------------------------------------------------ one process, many devices
b2r_status_t b2r_comm_init_all(b2r_comm_t* comms, int ndev, const int* devices, const b2r_comm_config_t* cfg) {
  B2R_REQUIRE(comms != nullptr, "b2r_comm_init_all: comms is NULL");
  B2R_REQUIRE(ndev >= 1 && ndev <= B2R_MAX_RANKS, "b2r_comm_init_all: ndev=%d out of range", ndev);
  const DriverApi* d = driver_api();
  if (!d) return B2R_ERR_UNSUPPORTED;
  int prev = 0;
  B2R_CUDA_CHECK(cudaGetDevice(&prev));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  std::vector<int> devs(ndev);
  for (int i = 0; i < ndev; ++i) devs[i] = devices ? devices[i] : i;
  for (int i = 0; i < ndev; ++i)
    for (int j = 0; j < ndev; ++j) {
      if (i == j) continue;
      B2R_REQUIRE(devs[i] != devs[j], "b2r_comm_init_all: device %d listed twice (use b2r_comm_init_loopback for virtual ranks)", devs[i]);
      int can = 0;
      B2R_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, devs[i], devs[j]));
      if (!can) return fail(B2R_ERR_UNSUPPORTED, "device %d cannot access device %d (no NVLink/P2P path)", devs[i], devs[j]);
    }
  for (int i = 0; i < ndev; ++i) {  // make sure every primary context exists
    B2R_CUDA_CHECK(cudaSetDevice(devs[i]));
    B2R_CUDA_CHECK(cudaFree(0));
  }

  CUmemAllocationProp prop;
  size_t gran = 0;
  vmm_prop(devs[0], false, &prop);
  B2R_CU_CHECK(d->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  bool want_mc = !(cfg && cfg->use_multicast == 0) && ndev > 1;
  for (int i = 0; i < ndev && want_mc; ++i) want_mc = device_supports_multicast(d, devs[i]);
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  if (want_mc) {
    mp.numDevices = (unsigned)ndev;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_NONE;
    mp.size = compute_sizes(cfg, gran).heap_bytes;
    size_t mg = 0;
    if (d->MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg) gran = std::max(gran, mg);
    else want_mc = false;
  }
  if (cfg && cfg->use_multicast == 1 && !want_mc) return fail(B2R_ERR_UNSUPPORTED, "multicast (NVLS) required but not supported here");
  Sizes s = compute_sizes(cfg, gran);

  auto group = std::make_shared<CommGroup>();
  group->nranks = ndev;
  group->mode = kModeAll;
  group->devices = devs;
  group->handles.assign(ndev, 0);
  for (int i = 0; i < ndev; ++i) {
    vmm_prop(devs[i], false, &prop);
    B2R_CU_CHECK(d->MemCreate(&group->handles[i], s.heap_bytes, &prop, 0));
  }
  group->va_size = s.heap_bytes * ndev;
  B2R_CU_CHECK(d->MemAddressReserve(&group->va_base, group->va_size, gran, 0, 0));
  std::vector<CUmemAccessDesc> acc(ndev);
  for (int i = 0; i < ndev; ++i) {
    acc[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc[i].location.id = devs[i];
    acc[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  for (int i = 0; i < ndev; ++i) B2R_CU_CHECK(d->MemMap(group->va_base + (size_t)i * s.heap_bytes, s.heap_bytes, 0, group->handles[i], 0));
  B2R_CU_CHECK(d->MemSetAccess(group->va_base, group->va_size, acc.data(), (size_t)ndev));

  bool has_mc = false;
  if (want_mc) {
    mp.size = s.heap_bytes;
    CUresult r = d->MulticastCreate(&group->mc_handle, &mp);
    for (int i = 0; i < ndev && r == CUDA_SUCCESS; ++i) {
      CUdevice cudev;
      r = d->DeviceGet(&cudev, devs[i]);
      if (r == CUDA_SUCCESS) r = d->MulticastAddDevice(group->mc_handle, cudev);
    }
    for (int i = 0; i < ndev && r == CUDA_SUCCESS; ++i) r = d->MulticastBindMem(group->mc_handle, 0, group->handles[i], 0, s.heap_bytes, 0);
    if (r == CUDA_SUCCESS) r = d->MemAddressReserve(&group->mc_va, s.heap_bytes, gran, 0, 0);
    if (r == CUDA_SUCCESS) r = d->MemMap(group->mc_va, s.heap_bytes, 0, group->mc_handle, 0);
    if (r == CUDA_SUCCESS) r = d->MemSetAccess(group->mc_va, s.heap_bytes, acc.data(), (size_t)ndev);
    has_mc = (r == CUDA_SUCCESS);
    if (!has_mc) {
      if (cfg && cfg->use_multicast == 1) return fail(B2R_ERR_UNSUPPORTED, "multicast set-up failed with CUresult %d", (int)r);
      group->mc_va = 0;  // P2P paths still work
    }
  }

  for (int i = 0; i < ndev; ++i) {
    b2r_comm* c = new b2r_comm();
    c->mode = kModeAll;
    c->rank = i;
    c->nranks = ndev;
    c->device = devs[i];
    memset(&c->ctx, 0, sizeof(c->ctx));
    for (int p = 0; p < ndev; ++p) c->ctx.heap[p] = (char*)(group->va_base + (size_t)p * s.heap_bytes);
    c->ctx.mc = has_mc ? (char*)group->mc_va : nullptr;
    c->ctx.rank = i;
    c->ctx.nranks = ndev;
    c->has_mc = has_mc;
    fill_common(c, cfg, s);
    c->group = group;
    group->members.push_back(c);
    b2r_status_t st = finish_init(c);
    if (st != B2R_OK) {
      discard_members(group.get());
      for (int k = 0; k < i; ++k) comms[k] = nullptr;  // nothing usable is handed back
      return st;
    }
    comms[i] = c;
  }
  group->alive = ndev;
  return B2R_OK;
}

// ------------------------------------------------------------------------------------------------ one process per device
b2r_status_t b2r_comm_init_rank(b2r_comm_t* comm, int nranks, int rank, int device, const char* bootstrap_id,
                                const b2r_comm_config_t* cfg) {
  B2R_REQUIRE(comm != nullptr && bootstrap_id != nullptr, "b2r_comm_init_rank: NULL argument");
  B2R_REQUIRE(nranks >= 1 && nranks <= B2R_MAX_RANKS && rank >= 0 && rank < nranks, "b2r_comm_init_rank: bad rank %d / %d", rank, nranks);
  const DriverApi* d = driver_api();
  if (!d) return B2R_ERR_UNSUPPORTED;
  B2R_CUDA_CHECK(cudaSetDevice(device));
  B2R_CUDA_CHECK(cudaFree(0));

  std::unique_ptr<Bootstrap> boot(new Bootstrap());
  B2R_CHECK(boot->init(bootstrap_id, rank, nranks, 60000));

  CUmemAllocationProp prop;
  vmm_prop(device, true, &prop);
  size_t gran = 0;
  B2R_CU_CHECK(d->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));

  // multicast: every rank must support it, and every rank must take the same decision at every step
  int my_mc = (!(cfg && cfg->use_multicast == 0) && nranks > 1 && device_supports_multicast(d, device)) ? 1 : 0;
  std::vector<int> all_int(nranks);
  B2R_CHECK(boot->allgather(&my_mc, all_int.data(), sizeof(int)));
  bool want_mc = true;
  for (int v : all_int) want_mc = want_mc && v;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  if (want_mc) {
    mp.numDevices = (unsigned)nranks;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mp.size = compute_sizes(cfg, gran).heap_bytes;
    size_t mg = 0;
    int ok = (d->MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg) ? 1 : 0;
    B2R_CHECK(boot->allgather(&ok, all_int.data(), sizeof(int)));
    for (int v : all_int) want_mc = want_mc && v;
    if (want_mc) gran = std::max(gran, mg);
  }
  {  // everybody must agree on the granularity (same GPUs -> same value; checked, not assumed)
    std::vector<unsigned long long> g(nranks);
    unsigned long long mine = gran;
    B2R_CHECK(boot->allgather(&mine, g.data(), sizeof(mine)));
    for (auto v : g) gran = std::max<size_t>(gran, (size_t)v);
  }
  if (cfg && cfg->use_multicast == 1 && !want_mc) return fail(B2R_ERR_UNSUPPORTED, "multicast (NVLS) required but not supported here");
  Sizes s = compute_sizes(cfg, gran);

  auto group = std::make_shared<CommGroup>();
  group->nranks = nranks;
  group->mode = kModeRank;
  group->handles.assign(nranks, 0);
  B2R_CU_CHECK(d->MemCreate(&group->handles[rank], s.heap_bytes, &prop, 0));
  int my_fd = -1;
  B2R_CU_CHECK(d->MemExportToShareableHandle(&my_fd, group->handles[rank], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  std::vector<int> fds;
  B2R_CHECK(boot->allgather_fd(my_fd, &fds));
  for (int p = 0; p < nranks; ++p) {
    if (p == rank) continue;
    B2R_CU_CHECK(d->MemImportFromShareableHandle(&group->handles[p], (void*)(uintptr_t)fds[p], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  }
  for (int fd : fds)
    if (fd >= 0) ::close(fd);

  group->va_size = s.heap_bytes * nranks;
  B2R_CU_CHECK(d->MemAddressReserve(&group->va_base, group->va_size, gran, 0, 0));
  for (int p = 0; p < nranks; ++p) B2R_CU_CHECK(d->MemMap(group->va_base + (size_t)p * s.heap_bytes, s.heap_bytes, 0, group->handles[p], 0));
  CUmemAccessDesc acc;
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  {
    CUresult r = d->MemSetAccess(group->va_base, group->va_size, &acc, 1);
    if (r != CUDA_SUCCESS) return fail(B2R_ERR_UNSUPPORTED, "cuMemSetAccess on peer heaps failed (CUresult %d): no P2P path between the GPUs?", (int)r);
  }

  bool has_mc = false;
  if (want_mc) {
    mp.size = s.heap_bytes;
    int ok = 1;
    int mc_fd = -1;
    if (rank == 0) {
      ok = d->MulticastCreate(&group->mc_handle, &mp) == CUDA_SUCCESS;
      if (ok) ok = d->MemExportToShareableHandle(&mc_fd, group->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS;
    }
    B2R_CHECK(boot->allgather(&ok, all_int.data(), sizeof(int)));
    bool go = all_int[0] != 0;
    if (go) {
      B2R_CHECK(boot->bcast_fd(&mc_fd));
      if (rank != 0) ok = d->MemImportFromShareableHandle(&group->mc_handle, (void*)(uintptr_t)mc_fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) == CUDA_SUCCESS;
      if (mc_fd >= 0) ::close(mc_fd);
      CUdevice cudev;
      if (ok) ok = d->DeviceGet(&cudev, device) == CUDA_SUCCESS;
      if (ok) ok = d->MulticastAddDevice(group->mc_handle, cudev) == CUDA_SUCCESS;
      B2R_CHECK(boot->allgather(&ok, all_int.data(), sizeof(int)));  // all devices added before anybody binds
      for (int v : all_int) go = go && v;
      if (go) {
        ok = d->MulticastBindMem(group->mc_handle, 0, group->handles[rank], 0, s.heap_bytes, 0) == CUDA_SUCCESS;
        if (ok) ok = d->MemAddressReserve(&group->mc_va, s.heap_bytes, gran, 0, 0) == CUDA_SUCCESS;
        if (ok) ok = d->MemMap(group->mc_va, s.heap_bytes, 0, group->mc_handle, 0) == CUDA_SUCCESS;
        if (ok) ok = d->MemSetAccess(group->mc_va, s.heap_bytes, &acc, 1) == CUDA_SUCCESS;
        B2R_CHECK(boot->allgather(&ok, all_int.data(), sizeof(int)));
        for (int v : all_int) go = go && v;
      }
    }
    has_mc = go;
    if (!has_mc) {
      if (cfg && cfg->use_multicast == 1) return fail(B2R_ERR_UNSUPPORTED, "multicast set-up failed on at least one rank");
      group->mc_va = 0;
    }
  }

  b2r_comm* c = new b2r_comm();
  c->mode = kModeRank;
  c->rank = rank;
  c->nranks = nranks;
  c->device = device;
  memset(&c->ctx, 0, sizeof(c->ctx));
  for (int p = 0; p < nranks; ++p) c->ctx.heap[p] = (char*)(group->va_base + (size_t)p * s.heap_bytes);
  c->ctx.mc = has_mc ? (char*)group->mc_va : nullptr;
  c->ctx.rank = rank;
  c->ctx.nranks = nranks;
  c->has_mc = has_mc;
  fill_common(c, cfg, s);
  c->group = group;
  group->members.push_back(c);
  group->alive = 1;
  b2r_status_t st = finish_init(c);
  if (st == B2R_OK) st = boot->barrier();  // nobody signals into a control region that is not zeroed yet
  if (st != B2R_OK) {
    discard_members(group.get());
    return st;
  }
  c->boot = boot.release();
  *comm = c;
  return B2R_OK;
}

b2r_status_t b2r_comm_destroy(b2r_comm_t c) {
  if (!c) return B2R_OK;
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  if (c->boot) c->boot->barrier();  // peers may still be reading my heap
  if (c->ws) b2r_workspace_destroy(c->ws);
  delete c->boot;
  c->group.reset();  // the last member releases the mappings
  cudaSetDevice(prev);
  delete c;
  return B2R_OK;
}

b2r_status_t b2r_comm_rank(b2r_comm_t c, int* rank) {
  B2R_REQUIRE(c && rank, "NULL argument");
  *rank = c->rank;
  return B2R_OK;
}
b2r_status_t b2r_comm_size(b2r_comm_t c, int* n) {
  B2R_REQUIRE(c && n, "NULL argument");
  *n = c->nranks;
  return B2R_OK;
}
b2r_status_t b2r_comm_device(b2r_comm_t c, int* dev) {
  B2R_REQUIRE(c && dev, "NULL argument");
  *dev = c->device;
  return B2R_OK;
}
b2r_status_t b2r_comm_has_multicast(b2r_comm_t c, int* yes) {
  B2R_REQUIRE(c && yes, "NULL argument");
  *yes = c->has_mc ? 1 : 0;
  return B2R_OK;
}

b2r_status_t b2r_comm_workspace(b2r_comm_t c, b2r_workspace_t* ws) {
  B2R_REQUIRE(c && ws, "NULL argument");
  *ws = c->ws;
  return B2R_OK;
}

b2r_status_t b2r_symm_alloc(b2r_comm_t c, uint64_t bytes, void** ptr) {
  B2R_REQUIRE(c && ptr, "b2r_symm_alloc: NULL argument");
  const size_t need = round_up((size_t)std::max<uint64_t>(bytes, 1), 256);
  if (c->bump + need > c->user_end)
    return fail(B2R_ERR_NOMEM, "b2r_symm_alloc: %llu bytes requested, %zu left in the symmetric heap (raise b2r_comm_config_t.heap_bytes)",
                (unsigned long long)bytes, c->user_end - c->bump);
  *ptr = c->my_heap() + c->bump;
  c->bump += need;
  return B2R_OK;
}

b2r_status_t b2r_symm_free_all(b2r_comm_t c) {
  B2R_REQUIRE(c, "NULL communicator");
  c->bump = c->user_off;
  return B2R_OK;
}

b2r_status_t b2r_symm_peer_ptr(b2r_comm_t c, const void* local_ptr, int peer, void** peer_ptr) {
  B2R_REQUIRE(c && peer_ptr, "NULL argument");
  B2R_REQUIRE(peer >= 0 && peer < c->nranks, "bad peer %d", peer);
  const char* p = (const char*)local_ptr;
  B2R_REQUIRE(p >= c->my_heap() && p < c->my_heap() + c->heap_bytes, "pointer %p is not inside the symmetric heap", local_ptr);
  *peer_ptr = c->ctx.heap[peer] + (p - c->my_heap());
  return B2R_OK;
}

b2r_status_t b2r_comm_check(b2r_comm_t c, void* stream) {
  B2R_REQUIRE(c, "NULL communicator");
  int prev = 0;
  B2R_CUDA_CHECK(cudaGetDevice(&prev));
  B2R_CUDA_CHECK(cudaSetDevice(c->device));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  unsigned err = 0;
  B2R_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
  B2R_CUDA_CHECK(cudaMemcpy(&err, c->my_heap() + kOffError, sizeof(err), cudaMemcpyDeviceToHost));
  if (err) {
    B2R_CUDA_CHECK(cudaMemset(c->my_heap() + kOffError, 0, sizeof(err)));
    if (err == 2) return fail(B2R_ERR_INVALID, "rank %d: NVLS needs send/recv buffers at the SAME symmetric-heap offset on every rank", c->rank);
    return fail(B2R_ERR_TIMEOUT, "rank %d: a cross-GPU wait exceeded %d ms (peer missing, dead, or fault injected)", c->rank, c->timeout_ms);
  }
  return B2R_OK;
}

}  // extern "C"
