// api_internal.cuh -- library-internal objects behind the opaque C handles.
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/aha_b200.h"
#include "dfa.hpp"
#include "host_cedar.hpp"

namespace aha {

extern thread_local std::string g_err;

inline int32_t set_err(int32_t code, const std::string& m) {
  g_err = m;
  return code;
}

#define AHA_CUDA(expr)                                                                               \
  do {                                                                                               \
    cudaError_t _e = (expr);                                                                         \
    if (_e != cudaSuccess) {                                                                         \
      return ::aha::set_err(AHA_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));       \
    }                                                                                                \
  } while (0)

// hot table placed in shared memory by the fused kernel (built at upload time from the blob)
struct HotTable {
  uint32_t n_hot = 0;         // states [0, n_hot) have rows in shared memory
  uint32_t row_stride = 0;    // u16 entries per row
  uint16_t* dev_rows = nullptr;  // [n_hot][row_stride]
};

struct DeviceDfa {
  int device = -1;
  uint8_t* blob = nullptr;  // device copy of the whole blob
  DfaView view;             // device pointers
  HotTable hot;
};

// workspace of one scan call: everything is stream-ordered (cudaMallocAsync) so repeated calls
// reuse the pool without synchronising frees.
struct HitsOwner {
  int device = 0;
  cudaStream_t stream = nullptr;
  void* dev_hits = nullptr;
  void* dev_doc_hit_off = nullptr;
  void* host_hits = nullptr;         // pinned
  void* host_doc_hit_off = nullptr;  // pinned
  void* dev_text = nullptr;          // owned copy made by aha_b200_scan
  void* dev_doc_off = nullptr;
};

}  // namespace aha

struct aha_dfa {
  std::vector<uint8_t> blob;  // host copy
  aha::BlobHeader header;
  std::vector<aha::DeviceDfa> dev;
  mutable double hits_per_byte = -1.0;  // largest hit density seen so far: sizes the hit buffer of the fused scan
  const aha::DeviceDfa* on(int device) const {
    for (const auto& d : dev)
      if (d.device == device) return &d;
    return nullptr;
  }
};

struct aha_builder {
  aha::HostCedar cedar;
  aha::HostAC ac;
};
