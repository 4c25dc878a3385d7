#include "driver_api.h"

#include <cuda_runtime.h>

#include <mutex>

#include "../util/check.h"

namespace b2r {

namespace {
DriverApi g_api;
bool g_ok = false;
std::once_flag g_once;

template <typename F>
bool resolve(const char* name, F* out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
    cudaGetLastError();
    fail(B2R_ERR_UNSUPPORTED, "driver entry point %s not available (%s)", name, e != cudaSuccess ? cudaGetErrorString(e) : "not found");
    return false;
  }
  *out = reinterpret_cast<F>(fn);
  return true;
}
}  // namespace

const DriverApi* driver_api() {
  std::call_once(g_once, [] {
    if (cudaFree(0) != cudaSuccess) {  // forces runtime + driver initialisation
      cudaGetLastError();
      fail(B2R_ERR_UNSUPPORTED, "CUDA driver not available");
      return;
    }
    bool ok = true;
    ok = ok && resolve("cuMemCreate", &g_api.MemCreate);
    ok = ok && resolve("cuMemRelease", &g_api.MemRelease);
    ok = ok && resolve("cuMemAddressReserve", &g_api.MemAddressReserve);
    ok = ok && resolve("cuMemAddressFree", &g_api.MemAddressFree);
    ok = ok && resolve("cuMemMap", &g_api.MemMap);
    ok = ok && resolve("cuMemUnmap", &g_api.MemUnmap);
    ok = ok && resolve("cuMemSetAccess", &g_api.MemSetAccess);
    ok = ok && resolve("cuMemGetAllocationGranularity", &g_api.MemGetAllocationGranularity);
    ok = ok && resolve("cuMemExportToShareableHandle", &g_api.MemExportToShareableHandle);
    ok = ok && resolve("cuMemImportFromShareableHandle", &g_api.MemImportFromShareableHandle);
    ok = ok && resolve("cuDeviceGet", &g_api.DeviceGet);
    ok = ok && resolve("cuDeviceGetAttribute", &g_api.DeviceGetAttribute);
    ok = ok && resolve("cuGetErrorString", &g_api.GetErrorString);
    // multicast is optional (CUDA >= 12.1 drivers); absence only disables the NVLS algorithms
    bool mc = resolve("cuMulticastCreate", &g_api.MulticastCreate) && resolve("cuMulticastAddDevice", &g_api.MulticastAddDevice) &&
              resolve("cuMulticastBindMem", &g_api.MulticastBindMem) && resolve("cuMulticastUnbind", &g_api.MulticastUnbind) &&
              resolve("cuMulticastGetGranularity", &g_api.MulticastGetGranularity);
    if (!mc) g_api.MulticastCreate = nullptr;
    g_ok = ok;
  });
  return g_ok ? &g_api : nullptr;
}

}  // namespace b2r
