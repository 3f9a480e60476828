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
  const uint32_t* delta;     // [S][C] u32: state | kOut | kEnd | kGoto; NULL when the automaton is too large for a dense table
  const uint4* rows;         // [S] sparse rows (dfa.hpp SparseRow): x = fail << 8 | flags, y z w = edges
  const uint32_t* wide;      // [n_wide][C] dense rows of the states with more than three children; row 0 = the root's
  const uint32_t* wide_fail; // [n_wide] fail links of the wide states, and
  uint32_t n_wide_first;     // ... n_wide when states are numbered wide first (id < n_wide_first <=> wide), else 0
  const uint32_t* out_off;
  const int32_t* out_val;
  const uint32_t* key_len;
  const uint32_t* key_tail;
  const uint2* out8;         // [n_out] output-list entries as {key id, key length << 16 | characters after the first}; NULL:
                             // a key too long for the packing (the lists are then read through out_val / key_len / key_tail)
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
  // case-folded trie (folded_trie.hpp), ignore_case scans of the trie path; fdelta == NULL: not available
  const uint32_t* fdelta;        // [n_fstates][fC]: target | kFoldCand, 0 = no edge
  const uint8_t* fcmap;          // [256] classes of the folded bytes
  const uint32_t* fcand_off;     // [n_fstates + 1]
  const uint4* fcand;            // {value, characters, mask low, mask high}
  const uint4* frec;             // [n_fstates][2]: {n, v0, v1, v2}, {m0, m1, m2, 0} (folded_trie.hpp)
  uint32_t fC;
  uint4* doc_scratch16;          // folded path, keys of at most 32 characters: {first character, characters, folded node, window} per hit,
                                 // the key behind the node is settled by trie_compact_folded_kernel (coalesced, off the walk)
};

// "chr < sep.size && !sep[chr]" (ac.cr:326,333): neighbour byte vetoes the hit
__device__ __forceinline__ bool sep_rejects(const ScanParams& p, uint32_t chr) {
  return chr < p.sep_size && !((p.sep_bits[chr >> 5] >> (chr & 31)) & 1u);
}
__device__ __forceinline__ bool is_sep(const ScanParams& p, uint32_t chr) {
  return chr < p.sep_size && ((p.sep_bits[chr >> 5] >> (chr & 31)) & 1u);
}

// The two neighbour tests of ACX#match(seq, sep).  Bytes / String forms (ac.cr:324-336): the neighbouring BYTE.
// Array(Char) / Slice(Char) forms (ac.cr:346-360, AHA_SEP_CODEPOINT): the code point of the neighbouring CHARACTER --
// only code points below sep.size <= 256 can veto, i.e. an ASCII byte or a two-byte unit led by 0xC2 / 0xC3.
// `next` = absolute index of the first byte after the match; `start` = absolute index of the key's first byte.
__device__ __forceinline__ bool sep_veto_after(const ScanParams& p, int64_t next, int64_t doc_end) {
  if (next >= doc_end) return false;
  uint32_t chr = p.text[next];
  if ((p.mode & AHA_SEP_CODEPOINT) && chr >= 0x80u) {
    if ((chr == 0xC2u || chr == 0xC3u) && next + 1 < doc_end) chr = ((chr & 0x1Fu) << 6) | (p.text[next + 1] & 0x3Fu);
    else return false;
  }
  return sep_rejects(p, chr);
}
__device__ __forceinline__ bool sep_veto_before(const ScanParams& p, int64_t start, int64_t doc_start) {
  if (start <= doc_start) return false;
  uint32_t chr = p.text[start - 1];
  if ((p.mode & AHA_SEP_CODEPOINT) && chr >= 0x80u) {
    if ((chr & 0xC0u) != 0x80u || start - 2 < doc_start) return false;
    const uint32_t b0 = p.text[start - 2];
    if (b0 != 0xC2u && b0 != 0xC3u) return false;
    chr = ((b0 & 0x1Fu) << 6) | (chr & 0x3Fu);
  }
  return sep_rejects(p, chr);
}

// ---- the automaton through its sparse rows -----------------------------------------------------------------------
// goto edge of state s on `byte` (edge format: target << 9 | kEdgeOut | label), 0 when there is none; *fail = fail link
__device__ __forceinline__ uint32_t sparse_goto(const ScanParams& p, const uint8_t* cmap, uint32_t s, uint32_t byte, uint32_t* fail) {
  if (s < p.n_wide_first) {  // wide-first numbering: the state's id is its row in wide[]
    *fail = __ldg(p.wide_fail + s);
    const uint32_t e = __ldg(p.wide + (uint64_t)s * p.C + cmap[byte]);
    return (e & 0xFFu) == 1u ? e : 0u;  // label field 2: a closure entry, not a goto edge
  }
  const uint4 r = __ldg(p.rows + s);
  *fail = r.x >> 8;
  if (r.x & kRowWide) {
    const uint32_t e = __ldg(p.wide + (uint64_t)r.y * p.C + cmap[byte]);
    return (e & 0xFFu) == 1u ? e : 0u;
  }
  return (r.y & 0xFFu) == byte ? r.y : ((r.z & 0xFFu) == byte ? r.z : ((r.w & 0xFFu) == byte ? r.w : 0u));
}
// entry of the goto/fail closure for (s, byte) in the dense table's format: state | kOut | kEnd | kGoto.  Read from the
// dense table when the automaton has one, else computed by the loop the table is the closure of (ac.cr:179-190).
__device__ __forceinline__ uint32_t delta_entry(const ScanParams& p, const uint8_t* cmap, uint32_t s, uint32_t byte) {
  if (p.delta) return p.delta[(uint64_t)s * p.C + cmap[byte]];
  uint32_t went = kGoto;
  for (;;) {
    uint32_t fail;
    const uint32_t e = byte ? sparse_goto(p, cmap, s, byte, &fail) : 0u;
    if (e) {
      const uint32_t t = e >> 9;
      const uint32_t head = __ldg(&p.rows[t].x);
      return t | ((e & kEdgeOut) ? kOut : 0u) | ((head & kRowEnd) ? kEnd : 0u) | went;
    }
    if (s == 0 || !byte) return 0u;
    s = fail;
    went = 0u;
  }
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
