// scan_common.cuh -- shared device-side definitions for the scan kernels.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/aha_b200.h"
#include "dfa.hpp"

namespace aha {

struct ScanParams {
  const uint8_t* text;       // concatenated documents
  int64_t total;             // bytes
  const int64_t* doc_off;    // n_docs + 1
  int64_t n_docs;
  // automaton (device pointers into the blob)
  const uint8_t* classmap;
  const uint32_t* delta;     // [S][C] u32: state | kOut | kEnd | kGoto
  const uint32_t* out_off;
  const int32_t* out_val;
  const uint32_t* key_len;
  const uint32_t* key_tail;
  uint32_t C;
  uint32_t warm;             // max_key_len - 1 bytes of warm-up before an owned region
  uint32_t mode;
  uint32_t sep_size;
  uint32_t sep_bits[8];
  // chunking of the concatenated text
  int32_t chunk;
  int64_t n_chunks;
  uint32_t* counts;              // [n_chunks + 1] hits per chunk (count pass)
  const uint64_t* offsets;       // [n_chunks + 1] exclusive prefix (emit pass)
  aha_hit* hits;                 // may be null (AHA_COUNT_ONLY)
  int64_t* doc_hit_off;          // [n_docs + 1]
  const uint64_t* lead_prefix;   // [n_chunks + 1] non-continuation bytes before each chunk (AHA_UNITS_CHAR)
  const uint64_t* doc_lead_base; // [n_docs + 1]  non-continuation bytes before each document
  aha_hit* doc_scratch;          // trie path: hits of document d at [doc_off[d] / 4, doc_off[d + 1] / 4) during the count pass
};

// "chr < sep.size && !sep[chr]" (ac.cr:326,333): neighbour byte vetoes the hit
__device__ __forceinline__ bool sep_rejects(const ScanParams& p, uint32_t chr) {
  return chr < p.sep_size && !((p.sep_bits[chr >> 5] >> (chr & 31)) & 1u);
}
__device__ __forceinline__ bool is_sep(const ScanParams& p, uint32_t chr) {
  return chr < p.sep_size && ((p.sep_bits[chr >> 5] >> (chr & 31)) & 1u);
}

// last index d in [0, n] with a[d] <= x  (a is non-decreasing, a[0] <= x)
__device__ __forceinline__ int64_t last_le(const int64_t* a, int64_t n, int64_t x) {
  int64_t lo = 0, hi = n;  // invariant a[lo] <= x
  while (lo < hi) {
    int64_t mid = (lo + hi + 1) >> 1;
    if (a[mid] <= x) lo = mid;
    else hi = mid - 1;
  }
  return lo;
}
// first index d in [0, n] with a[d] >= x (n if none is impossible here because a[n] = total >= x)
__device__ __forceinline__ int64_t first_ge(const int64_t* a, int64_t n, int64_t x) {
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (a[mid] >= x) hi = mid;
    else lo = mid + 1;
  }
  return lo;
}

}  // namespace aha
