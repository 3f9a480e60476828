// scan_fused.cuh -- single-pass fused scan kernel (placeholder until the fused kernel lands).
#pragma once
#include "api_internal.cuh"
#include "scan_common.cuh"

namespace aha {
static int32_t build_hot_table(const aha_dfa*, DeviceDfa* dd) {
  dd->hot = HotTable();
  return AHA_OK;
}
}  // namespace aha
