// guard_alloc.cpp — device buffers that END (or begin) at an unmapped page: the library's own out-of-bounds detector.
//
// The reference's only memory checker is cutil's emulation-mode bank checker (cuda/C/common/inc/bank_checker.h:25-60),
// dead since CUDA 3.1; the modern replacement, compute-sanitizer, is closed on the GPU pool this library is developed on
// ("runs under it have left GPUs needing a reset"). What remains is the hardware MMU: with the virtual-memory-management
// API a buffer can be placed so that the first byte past its end (or before its start) is reserved-but-unmapped address
// space. A kernel that reads or writes one vector too far then takes an illegal-address fault instead of silently touching
// a neighbouring allocation, which is exactly what memcheck would report. tests/test_guard_gpu.py runs the single-GPU
// kernel family over such buffers (in a child process, because a fault poisons the CUDA context for good).
//
// Layout of one allocation:   [ unmapped granule | mapped: round_up(bytes, granule) | unmapped granule ]
//   at_end = 1 : ptr = mapped_end - bytes   (the buffer's last byte is the last mapped byte)
//   at_end = 0 : ptr = mapped_begin         (the byte before the buffer is unmapped)
#include <cuda_runtime.h>

#include <map>
#include <mutex>

#include "../util/check.h"
#include "b2r.h"
#include "driver_api.h"

namespace {

struct GuardBlock {
  CUdeviceptr va;
  size_t va_size, mapped;
  CUmemGenericAllocationHandle handle;
};
std::mutex g_mu;
std::map<void*, GuardBlock> g_blocks;

}  // namespace

extern "C" b2r_status_t b2r_guard_alloc(uint64_t bytes, int device, int at_end, void** ptr) {
  using namespace b2r;
  B2R_REQUIRE(ptr != nullptr && bytes > 0, "b2r_guard_alloc: bad argument");
  const DriverApi* api = driver_api();
  if (!api) return B2R_ERR_UNSUPPORTED;
  int prev = 0;
  B2R_CUDA_CHECK(cudaGetDevice(&prev));
  B2R_CUDA_CHECK(cudaSetDevice(device));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  size_t gran = 0;
  B2R_CU_CHECK(api->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM));
  GuardBlock b = {};
  b.mapped = (size_t)((bytes + gran - 1) / gran * gran);
  b.va_size = b.mapped + 2 * gran;
  B2R_CU_CHECK(api->MemAddressReserve(&b.va, b.va_size, gran, 0, 0));
  CUresult r = api->MemCreate(&b.handle, b.mapped, &prop, 0);
  if (r == CUDA_SUCCESS) {
    r = api->MemMap(b.va + gran, b.mapped, 0, b.handle, 0);
    if (r == CUDA_SUCCESS) {
      CUmemAccessDesc acc = {};
      acc.location = prop.location;
      acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
      r = api->MemSetAccess(b.va + gran, b.mapped, &acc, 1);
      if (r != CUDA_SUCCESS) api->MemUnmap(b.va + gran, b.mapped);
    }
    if (r != CUDA_SUCCESS) api->MemRelease(b.handle);
  }
  if (r != CUDA_SUCCESS) {
    api->MemAddressFree(b.va, b.va_size);
    const char* msg = nullptr;
    api->GetErrorString(r, &msg);
    return fail(B2R_ERR_CUDA, "b2r_guard_alloc(%llu bytes): %s", (unsigned long long)bytes, msg ? msg : "driver error");
  }
  void* p = (void*)(at_end ? b.va + gran + b.mapped - bytes : b.va + gran);
  std::lock_guard<std::mutex> lk(g_mu);
  g_blocks[p] = b;
  *ptr = p;
  return B2R_OK;
}

extern "C" b2r_status_t b2r_guard_free(void* ptr) {
  using namespace b2r;
  if (!ptr) return B2R_OK;
  const DriverApi* api = driver_api();
  if (!api) return B2R_ERR_UNSUPPORTED;
  GuardBlock b;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_blocks.find(ptr);
    B2R_REQUIRE(it != g_blocks.end(), "b2r_guard_free: %p was not returned by b2r_guard_alloc", ptr);
    b = it->second;
    g_blocks.erase(it);
  }
  size_t gran = (b.va_size - b.mapped) / 2;
  api->MemUnmap(b.va + gran, b.mapped);
  api->MemRelease(b.handle);
  api->MemAddressFree(b.va, b.va_size);
  return B2R_OK;
}
