// api_internal.cuh -- library-internal objects behind the opaque C handles.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/aha_b200.h"
#include "dfa.hpp"
#include "host_cedar.hpp"

namespace aha {

extern thread_local std::string g_err;
extern std::atomic<long long> g_kernel_launches;  // kernels launched by this library (aha_b200_kernel_launches)

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

// character automaton placed in shared memory by scan_char_kernel (built at upload time from the blob)
struct HotTable {
  bool ok = false;            // dictionary is whole-UTF-8-unit and fits the shared-memory budget
  uint32_t n_hot = 0;         // character-level states resident in shared memory
  uint32_t n_root = 0, m = 0, m_pad = 0, n_mwords = 0;
  uint32_t lo[3] = {0, 0, 0}, n[3] = {0, 0, 0}, off[3] = {0, 0, 0};
  uint8_t* dev = nullptr;     // one allocation: root | val | check | gfail | ref_rec
  size_t off_root = 0, off_val = 0, off_check = 0, off_gfail = 0, off_ref = 0, off_mbits = 0;
  size_t smem_bytes = 0;
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
  void* host_hits = nullptr;         // pinned (recycled through PinnedPool)
  void* host_doc_hit_off = nullptr;  // pinned
  size_t host_hits_bytes = 0, host_doc_hit_off_bytes = 0;
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
