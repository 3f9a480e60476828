// api_internal.cuh -- library-internal objects behind the opaque C handles.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdio>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/aha_b200.h"
#include "dfa.hpp"
#include "folded_trie.hpp"
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

// restores the caller's current device when a function that switches devices returns (any path)
struct DeviceGuard {
  int prev = -1;
  DeviceGuard() { cudaGetDevice(&prev); }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// character automaton placed in shared memory by char_walk_kernel (built at upload time from the blob)
struct HotTable {
  bool ok = false;            // dictionary is whole-UTF-8-unit and fits the shared-memory budget
  uint32_t n_hot = 0;         // character-level states resident in shared memory
  uint32_t n_root = 0, m = 0, m_pad = 0, n_mwords = 0, n_gfail_s = 0;
  uint32_t lo[3] = {0, 0, 0}, n[3] = {0, 0, 0}, off[3] = {0, 0, 0};
  uint8_t* dev = nullptr;     // one allocation: root | val | check | gfail | ref_rec | mbits
  size_t off_root = 0, off_val = 0, off_check = 0, off_gfail = 0, off_ref = 0, off_mbits = 0, off_ref8 = 0, off_out8 = 0;
  uint32_t n_ref8 = 0;        // entries of the 8-byte ref table
  size_t smem_bytes = 0;
};

// ---- device memory that outlives a call --------------------------------------------------------------------
// One grow-only device allocation.  A warm scan call performs no CUDA allocation at all: every scratch array of
// the pipeline is a slot of the Workspace it runs in, every result buffer comes from the per-device DevCache.
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);  // growth only: synchronises the device, never on a warm call
    p = nullptr;
    cap = 0;
    const size_t want = ((bytes + bytes / 8 + 255) & ~size_t(255));  // head room: slightly larger inputs do not grow again
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
      e = cudaMalloc(&p, (bytes + 255) & ~size_t(255));
      if (e != cudaSuccess) return e;
      cap = (bytes + 255) & ~size_t(255);
      return cudaSuccess;
    }
    cap = want;
    return cudaSuccess;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <class T> T* as() const { return static_cast<T*>(p); }
};

// what the device reports at the end of a scan (one small D2H copy per call)
struct ScanStatus {
  unsigned long long total_hits;
  unsigned int n_direct;   // tiles delivered by the direct pass (diagnostics)
  unsigned int bad_docs;   // doc_off violations seen by validate_docs_kernel
};

enum WsSlot {
  kWsStore = 0,      // character path: event store
  kWsTileMeta,
  kWsTileHits,
  kWsTileDoc,
  kWsDirectList,
  kWsTilePrefix,
  kWsScanTmp,
  kWsLeadCounts,
  kWsTileLead,
  kWsDocLead,
  kWsStatus,         // ScanStatus + counters (device)
  kWsTileState,      // byte path: look-back state
  kWsScratch,
  kWsEvScratch,
  kWsMarks,
  kWsCounts,         // per-document machines: counts / offsets / scratch
  kWsOffsets,
  kWsDocScratch,
  kWsAligned,        // aligned copy of a text whose device pointer is not 16-byte aligned
  kWsLongA,          // match_longest in chunks: exit states (two rounds) and the entry state of every chunk's last walk
  kWsLongB,
  kWsLongEntry,
  kWsSlots
};

// Scratch of one scan call in flight on one device.  Acquired from the automaton's pool for the duration of a call
// (calls from several host threads get one each) and handed back when the call's last kernel has finished.
struct Workspace {
  int device = -1;
  DevBuf slot[kWsSlots];
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, evw0 = nullptr, evw1 = nullptr, evx0 = nullptr, evx1 = nullptr;
  ScanStatus* h_status = nullptr;  // pinned
  cudaError_t init(int dev) {
    device = dev;
    cudaError_t e;
    for (cudaEvent_t* ev : {&ev0, &ev1, &evw0, &evw1, &evx0, &evx1})
      if ((e = cudaEventCreate(ev)) != cudaSuccess) return e;
    return cudaMallocHost(&h_status, sizeof(ScanStatus));
  }
  ~Workspace() {
    if (device < 0) return;
    DeviceGuard g;
    cudaSetDevice(device);
    for (auto& s : slot) s.release();
    for (cudaEvent_t ev : {ev0, ev1, evw0, evw1, evx0, evx1})
      if (ev) cudaEventDestroy(ev);
    if (h_status) cudaFreeHost(h_status);
  }
  template <class T> cudaError_t get(WsSlot s, size_t count, T** out) {
    cudaError_t e = slot[s].ensure((count ? count : 1) * sizeof(T));
    *out = slot[s].as<T>();
    return e;
  }
};

struct DeviceDfa {
  int device = -1;
  uint8_t* blob = nullptr;  // device copy of the whole blob
  DfaView view;             // device pointers
  HotTable hot;
  uint2* out8 = nullptr;    // device: the output lists in 8 bytes per entry (byte-path expansion); NULL when a key does not fit
  // scan workspaces that are not in use (guarded by pool_mu of the owning aha_dfa)
  std::vector<std::unique_ptr<Workspace>> idle;
  double hits_per_byte = -1.0;  // largest hit density seen so far on this device: sizes the hit buffer up front
  // case-folded trie of the dictionary (folded_trie.hpp): built and uploaded at the first ignore_case scan
  int folded_state = 0;         // 0 = not tried yet, 1 = on the device, -1 = the dictionary does not qualify
  uint8_t* folded_dev = nullptr;  // one allocation: fdelta | fcand_off | fcand | fcmap
  size_t off_fcand_off = 0, off_fcand = 0, off_fcmap = 0, off_frec = 0;
  uint32_t folded_classes = 0, folded_max_chars = 0;
};

// Result buffers (hits, doc_hit_off) are handed to the caller and come back through aha_b200_hits_free: they are
// recycled through a per-device cache, so a steady stream of equally sized scans allocates nothing.
struct DevCache {
  struct Entry { void* p; size_t bytes; cudaEvent_t ev; };
  std::mutex mu;
  std::vector<Entry> free_list, in_use;
  size_t cached = 0;
  static constexpr size_t kMaxEntries = 16;
  void* acquire(size_t bytes, cudaStream_t st, size_t* got);
  void release(void* p, size_t bytes, cudaStream_t st);
  void trim();
};
DevCache& dev_cache(int device);

struct HitsOwner {
  int device = 0;
  cudaStream_t stream = nullptr;
  void* dev_hits = nullptr;          // from DevCache
  size_t dev_hits_bytes = 0;
  void* dev_doc_hit_off = nullptr;   // from DevCache
  size_t dev_doc_hit_off_bytes = 0;
  void* host_hits = nullptr;         // pinned (recycled through PinnedPool)
  void* host_doc_hit_off = nullptr;  // pinned
  size_t host_hits_bytes = 0, host_doc_hit_off_bytes = 0;
};

// test / diagnostics knobs, read from the environment once (aha_b200_init or first use) and settable through
// aha_b200_debug_set; none of them selects an alternate implementation of the default path
struct Knobs {
  int no_char = 0;          // AHA_B200_NO_CHAR=1: byte automaton even when the character tables fit
  int force_v1 = 0;         // AHA_B200_IMPL=0: two-pass checker kernels (the fallback for keys > 1024 bytes)
  int byte_chunk = 0;       // AHA_B200_BYTE_CHUNK: bytes per lane of the byte kernel (0 = default)
  int e2e_chunk_mb = 64;    // AHA_B200_E2E_CHUNK_MB: run size of the host pipeline
  int byte_span = 0;        // AHA_B200_BYTE_SPAN: bytes per lane of the byte kernel's span mode (0 = staged tiles)
};
Knobs& knobs();

}  // namespace aha

struct aha_dfa {
  std::vector<uint8_t> blob;  // host copy
  aha::BlobHeader header;
  std::vector<aha::DeviceDfa> dev;
  mutable std::mutex pool_mu;
  aha::DeviceDfa* on(int device) const {
    for (auto& d : const_cast<aha_dfa*>(this)->dev)
      if (d.device == device) return &d;
    return nullptr;
  }
};

struct aha_builder {
  aha::HostCedar cedar;
  aha::HostAC ac;
};
