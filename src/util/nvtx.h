// nvtx.h — optional NVTX ranges around the public entry points. Header-only NVTX v3: no link dependency;
// ranges are emitted only when B2R_NVTX=1 is set (checked once), so the default launch path pays one predictable branch.
#pragma once
#include <nvtx3/nvToolsExt.h>

#include <cstdlib>

namespace b2r {

inline bool nvtx_enabled() {
  static const bool on = [] {
    const char* v = getenv("B2R_NVTX");
    return v && *v && *v != '0';
  }();
  return on;
}

struct NvtxRange {
  bool active;
  explicit NvtxRange(const char* name) : active(nvtx_enabled()) {
    if (active) nvtxRangePushA(name);
  }
  ~NvtxRange() {
    if (active) nvtxRangePop();
  }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

}  // namespace b2r
