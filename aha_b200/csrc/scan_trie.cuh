// scan_trie.cuh -- CedarX#match_longest(String, ignore_case, sep) (cedar.cr:1075-1121) on the GPU.
//
// One thread per document (the BASELINE config runs it over millions of 4 KiB documents).  The trie
// is the goto part of the flattened DFA (entries with kGoto); a state's own key id is the head of its
// output list.  The reference explores case variants with a breadth-first frontier and keeps the FIRST
// candidate of strictly greatest length (cedar.cr:993-1019, 1102-1104).  Frontier order at a given
// length is lexicographic in the per-character choice (exact before case-swapped), which is exactly
// depth-first preorder, so a bounded-memory DFS with the same strict '>' rule picks the same key.
// Offsets are character indices.  Characters are decoded structurally (lead byte -> length), mapped,
// re-encoded and walked byte by byte like CedarX#jump(Char) (cedar.cr:946-952).
#pragma once
#include "folded_trie.hpp"
#include "scan_common.cuh"

namespace aha {

constexpr int kTrieMaxChars = 96;  // DFS depth bound = longest key in characters

// simple 1:1 case pairs: ASCII, Latin-1, Greek, Cyrillic basic blocks (Char#uppercase?/upcase/downcase)
__host__ __device__ __forceinline__ bool cp_upper(uint32_t c) {
  return (c >= 'A' && c <= 'Z') || (c >= 0xC0 && c <= 0xDE && c != 0xD7) || (c >= 0x391 && c <= 0x3A9 && c != 0x3A2) ||
         (c >= 0x410 && c <= 0x42F);
}
__host__ __device__ __forceinline__ bool cp_lower(uint32_t c) {
  return (c >= 'a' && c <= 'z') || (c >= 0xE0 && c <= 0xFE && c != 0xF7) || (c >= 0x3B1 && c <= 0x3C9 && c != 0x3C2) ||
         (c >= 0x430 && c <= 0x44F);
}
// chr.uppercase? ? chr.downcase : chr.upcase   (cedar.cr:995-996)
__host__ __device__ __forceinline__ uint32_t cp_other(uint32_t c) {
  if (c < 0x80) return (c - 'A' < 26u) ? c + 32 : ((c - 'a' < 26u) ? c - 32 : c);  // ASCII: the common case, two compares
  return cp_upper(c) ? c + 32 : (cp_lower(c) ? c - 32 : c);
}

__device__ __forceinline__ int utf8_decode(const uint8_t* s, int64_t i, int64_t n, uint32_t* cp) {
  const uint32_t b = s[i];
  if (b < 0x80) { *cp = b; return 1; }
  if ((b & 0xE0) == 0xC0 && i + 1 < n) { *cp = ((b & 0x1F) << 6) | (s[i + 1] & 0x3F); return 2; }
  if ((b & 0xF0) == 0xE0 && i + 2 < n) { *cp = ((b & 0x0F) << 12) | ((s[i + 1] & 0x3F) << 6) | (s[i + 2] & 0x3F); return 3; }
  if ((b & 0xF8) == 0xF0 && i + 3 < n) {
    *cp = ((b & 0x07) << 18) | ((s[i + 1] & 0x3F) << 12) | ((s[i + 2] & 0x3F) << 6) | (s[i + 3] & 0x3F);
    return 4;
  }
  *cp = b;
  return 1;
}
// the same with 32-bit offsets (a document is < 2^31 bytes)
__device__ __forceinline__ int utf8_decode32(const uint8_t* s, int32_t i, int32_t n, uint32_t* cp) {
  const uint32_t b = s[i];
  if (b < 0x80) { *cp = b; return 1; }
  if ((b & 0xE0) == 0xC0 && i + 1 < n) { *cp = ((b & 0x1F) << 6) | (s[i + 1] & 0x3F); return 2; }
  if ((b & 0xF0) == 0xE0 && i + 2 < n) { *cp = ((b & 0x0F) << 12) | ((s[i + 1] & 0x3F) << 6) | (s[i + 2] & 0x3F); return 3; }
  if ((b & 0xF8) == 0xF0 && i + 3 < n) {
    *cp = ((b & 0x07) << 18) | ((s[i + 1] & 0x3F) << 12) | ((s[i + 2] & 0x3F) << 6) | (s[i + 3] & 0x3F);
    return 4;
  }
  *cp = b;
  return 1;
}
__device__ __forceinline__ int utf8_encode(uint32_t c, uint8_t* o) {
  if (c < 0x80) { o[0] = (uint8_t)c; return 1; }
  if (c < 0x800) { o[0] = 0xC0 | (c >> 6); o[1] = 0x80 | (c & 0x3F); return 2; }
  if (c < 0x10000) { o[0] = 0xE0 | (c >> 12); o[1] = 0x80 | ((c >> 6) & 0x3F); o[2] = 0x80 | (c & 0x3F); return 3; }
  o[0] = 0xF0 | (c >> 18); o[1] = 0x80 | ((c >> 12) & 0x3F); o[2] = 0x80 | ((c >> 6) & 0x3F); o[3] = 0x80 | (c & 0x3F);
  return 4;
}

// goto edge of `from` on `byte` in the dense table's format (kGoto set iff the edge exists, kOut: the target has an output
// list); the trie walk never follows a fail link, so without a dense table one sparse row answers
__device__ __forceinline__ uint32_t trie_goto(const ScanParams& p, const uint8_t* cmap, uint32_t from, uint32_t byte) {
  if (p.delta) return p.delta[(uint64_t)from * p.C + cmap[byte]];
  uint32_t fail;
  const uint32_t e = byte ? sparse_goto(p, cmap, from, byte, &fail) : 0u;
  return e ? ((e >> 9) | ((e & kEdgeOut) ? kOut : 0u) | kGoto) : 0u;
}

template <bool EMIT>
__global__ void __launch_bounds__(128) trie_longest_doc_kernel(const ScanParams p) {
  __shared__ uint8_t cmap[256];
  __shared__ uint32_t s_sep[8];  // is_sep? as one bit test: the BitArray with the bits at or above sep.size cleared
  for (int i = threadIdx.x; i < 256; i += blockDim.x) cmap[i] = p.classmap[i];
  if (threadIdx.x < 8) {
    uint32_t w = p.sep_bits[threadIdx.x];
    const uint32_t lo = threadIdx.x * 32u;
    if (p.sep_size <= lo) w = 0u;
    else if (p.sep_size < lo + 32u) w &= (1u << (p.sep_size - lo)) - 1u;
    s_sep[threadIdx.x] = w;
  }
  __syncthreads();
  const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= p.n_docs) return;
  auto sep_at = [&](uint32_t chr) -> bool { return chr < 256u && ((s_sep[chr >> 5] >> (chr & 31u)) & 1u); };  // cedar.cr:1067-1069
  const uint32_t* const delta = p.delta;  // dense goto table when the automaton has one (32-bit index: it is <= 128 MB)
  const uint8_t* s = p.text + p.doc_off[d];
  const int64_t n = p.doc_off[d + 1] - p.doc_off[d];
  // The count pass also records: a document keeps one scratch slot per four bytes of text, which holds every hit
  // unless the matches are shorter than that on average; trie_compact_kernel then moves the slice to its final place
  // and only documents that overflowed are walked again (EMIT).
  const int64_t sc_base = p.doc_off[d] >> 2, sc_cap = p.doc_scratch ? (p.doc_off[d + 1] >> 2) - sc_base : 0;
  if (EMIT && p.doc_scratch && (int64_t)p.counts[d] <= sc_cap) return;
  const bool icase = (p.mode & AHA_IGNORE_CASE) != 0;
  const bool use_sep = (p.mode & AHA_SEP) != 0;
  const uint32_t C = p.C;
  uint64_t run = EMIT ? p.offsets[d] : 0;
  // jump(chr, from): walk the character's UTF-8 bytes over goto edges only; -1 == no path, else the table entry of the
  // last edge (target state | flags: kOut tells whether the target has an output list, so out_off[] is only read then)
  auto jump = [&](uint32_t cp, uint32_t from) -> int64_t {
    if (cp < 0x80) {  // one byte, one edge
      const uint32_t e = delta ? __ldg(delta + (from * C + cmap[cp])) : trie_goto(p, cmap, from, cp);
      return (!(e & kGoto) || cp == 0) ? -1 : (int64_t)e;
    }
    uint8_t buf[4];
    const int k = utf8_encode(cp, buf);
    uint32_t e = from;
    for (int j = 0; j < k; ++j) {
      e = trie_goto(p, cmap, e & kStateMask, buf[j]);
      if (!(e & kGoto) || buf[j] == 0) return -1;
    }
    return (int64_t)e;
  };
  // ONE loop, one micro-step per iteration, so that the 32 documents of a warp stay in step (the nested loops of the
  // reference -- skip separators / depth-first walk / skip the rest of the token -- ran with 6.7 of 32 lanes active):
  //   phase 0  top of the reference loop: skip separators; the first other character starts a walk in the same iteration
  //   phase 1  one step of the depth-first walk: the character at c_bpos, its exact edge and its case-swapped edge
  //   phase 2  no key started here: on to the next separator
  // The walk keeps no per-level arrays: when the swapped edge of a level exists it is pushed as a pending branch (entry,
  // position, depth) and popped when the exact branch is exhausted -- last in, first out, i.e. depth-first preorder, the
  // order of the reference's frontier (cedar.cr:993-1019); with mixed-case keys a pending branch is rare, so the
  // stack is almost never touched.
  uint2 stk[kTrieMaxChars];  // x = table entry of the branch's first edge, y = (byte offset behind its character - pos) | depth << 16
  int sp = 0;
  int64_t pos = 0;
  int32_t cidx = 0;
  int phase = 0, depth = 0;
  uint32_t c_node = 0;
  int64_t c_bpos = 0;
  int32_t best_n = 0, best_v = 0, best_bytes = 0;
  while (phase != 3) {
    const int64_t cur = phase == 1 ? c_bpos : pos;
    bool have = cur < n;
    uint32_t cp = 0;
    int l = 1;
    if (have) l = utf8_decode(s, cur, n, &cp);
    if (phase == 2) {
      // up to eight characters per iteration: a walk step of the neighbouring lanes waits for two table lookups, the
      // skipping costs a byte load from L1 each
      for (int k = 0;;) {
        if (!have) { phase = 3; break; }
        const bool at_sep = sep_at(cp);
        pos += l;
        ++cidx;
        if (at_sep) { phase = 0; break; }  // ... consumed like any separator at the top of the loop
        if (++k == 8) break;
        have = pos < n;
        if (have) l = utf8_decode(s, pos, n, &cp);
      }
      continue;
    }
    if (phase == 0) {
      if (!have) { phase = 3; continue; }
      if (use_sep && sep_at(cp)) {
        pos += l;
        ++cidx;
        continue;
      }
      phase = 1;  // a walk starts at this character (this iteration is its first step)
      depth = 0;
      c_node = 0;
      c_bpos = pos;
      sp = 0;
      best_n = 0;
    }
    int64_t e0 = -1, e1 = -1;
    if (have && depth < kTrieMaxChars) {
      e0 = jump(cp, c_node);
      if (icase) {
        const uint32_t cc = cp_other(cp);  // chr.uppercase? ? chr.downcase : chr.upcase (cedar.cr:995-996)
        if (cc != cp) e1 = jump(cc, c_node);
      }
    }
    if (e1 >= 0) stk[sp++] = make_uint2((uint32_t)e1, (uint32_t)(c_bpos + l - pos) | ((uint32_t)(depth + 1) << 16));
    uint32_t arrived;
    if (e0 >= 0) {
      arrived = (uint32_t)e0;
      c_bpos += l;
      ++depth;
    } else if (sp > 0) {
      const uint2 t = stk[--sp];
      arrived = t.x;
      c_bpos = pos + (int64_t)(t.y & 0xFFFFu);
      depth = (int)(t.y >> 16);
    } else {
      // every branch from this start is exhausted: the strictly longest candidate, first in preorder, wins
      if (best_n == 0) {
        if (pos < n) {
          uint32_t skip;
          pos += utf8_decode(s, pos, n, &skip);
          ++cidx;
        }
        phase = use_sep ? 2 : 0;
      } else {
        if (EMIT) {
          if (p.hits) p.hits[run] = aha_hit{cidx, cidx + best_n, best_v};
        } else if ((int64_t)run < sc_cap) {
          p.doc_scratch[sc_base + (int64_t)run] = aha_hit{cidx, cidx + best_n, best_v};
        }
        ++run;
        pos += best_bytes;
        cidx += best_n;
        phase = 0;
      }
      continue;
    }
    // arrival at a node `depth` characters deep, the text cursor behind its last character
    c_node = arrived & kStateMask;
    if ((arrived & kOut) && depth > best_n) {
      bool ok = true;
      if (use_sep && c_bpos < n) {
        uint32_t ncp;
        utf8_decode(s, c_bpos, n, &ncp);
        ok = sep_at(ncp);
      }
      if (ok) {
        best_n = depth;
        best_v = p.out_val[p.out_off[c_node]];
        best_bytes = (int32_t)(c_bpos - pos);
      }
    }
  }
  if (!EMIT) p.counts[d] = (uint32_t)run;
}

// The candidate of folded node `nd` that the reference's frontier reaches first: smallest (key mask XOR text window) with
// character 0 most significant (folded_trie.hpp).  One 32-byte record for the common case, the node's list otherwise.
__device__ __forceinline__ int32_t folded_pick(const ScanParams& p, uint32_t nd, uint64_t window) {
  const uint4 ra = __ldg(p.frec + 2 * (size_t)nd), rb = __ldg(p.frec + 2 * (size_t)nd + 1);  // one sector
  int32_t best_v = (int32_t)ra.y;
  if (ra.x != 0xFFFFFFFFu) {  // up to three candidates, 32-bit masks
    const uint32_t w = (uint32_t)window;
    uint32_t bx = rb.x ^ w;
    if (ra.x > 1u && (rb.y ^ w) < bx) { bx = rb.y ^ w; best_v = (int32_t)ra.z; }
    if (ra.x > 2u && (rb.z ^ w) < bx) { best_v = (int32_t)ra.w; }
  } else {
    const uint32_t j0 = __ldg(p.fcand_off + nd), j1 = __ldg(p.fcand_off + nd + 1);
    uint64_t best_x = ~0ull;
    for (uint32_t j = j0; j < j1; ++j) {
      const uint4 c = __ldg(p.fcand + j);
      const uint64_t x = (((uint64_t)c.w << 32) | c.z) ^ window;
      if (x < best_x || j == j0) {
        best_x = x;
        best_v = (int32_t)c.x;
      }
    }
  }
  return best_v;
}

// ---- ignore_case over the case-folded trie (folded_trie.hpp): one lookup per character, no frontier ------------------
// Same flat state machine and the same count / record / compact / emit scheme as trie_longest_doc_kernel.  A walk step
// folds the text character, follows ONE edge of the folded trie and shifts the character's case into the window mask;
// a node with candidates settles which original key the reference's frontier would have reached first.
#ifndef AHA_TRIE_SKIP
#define AHA_TRIE_SKIP 1     // characters per iteration of the skip phase (measured on C4: 1 -> 55.4, 2 -> 52.9, 4 -> 50.6, 8 -> 48.7 GB/s)
#endif
// (measured: asking the compiler for a minimum number of resident blocks -- 1, 12 or 16 -- is slower than its own choice of
// 54 registers: 29 / 40 / 33 GB/s against 50)
#ifdef AHA_TRIE_BLOCKS
#define AHA_TRIE_BOUNDS __launch_bounds__(128, AHA_TRIE_BLOCKS)
#else
#define AHA_TRIE_BOUNDS __launch_bounds__(128)
#endif
template <bool EMIT>
__global__ void AHA_TRIE_BOUNDS trie_folded_doc_kernel(const ScanParams p) {
  __shared__ uint8_t fcmap[256];
  __shared__ uint32_t s_sep[8];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) fcmap[i] = p.fcmap[i];
  if (threadIdx.x < 8) {
    uint32_t w = p.sep_bits[threadIdx.x];
    const uint32_t lo = threadIdx.x * 32u;
    if (p.sep_size <= lo) w = 0u;
    else if (p.sep_size < lo + 32u) w &= (1u << (p.sep_size - lo)) - 1u;
    s_sep[threadIdx.x] = w;
  }
  __syncthreads();
  const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= p.n_docs) return;
  auto sep_at = [&](uint32_t chr) -> bool { return chr < 256u && ((s_sep[chr >> 5] >> (chr & 31u)) & 1u); };
  const uint8_t* s = p.text + p.doc_off[d];
  const int32_t n = (int32_t)(p.doc_off[d + 1] - p.doc_off[d]);  // documents are < 2^31 bytes (validated)
  const int64_t sc_base = p.doc_off[d] >> 2, sc_cap = p.doc_scratch ? (p.doc_off[d + 1] >> 2) - sc_base : 0;
  if (EMIT && p.doc_scratch && (int64_t)p.counts[d] <= sc_cap) return;
  const bool use_sep = (p.mode & AHA_SEP) != 0;
  const uint32_t* const fdelta = p.fdelta;
  const uint32_t fC = p.fC;
  // where this document's hits go: its final place (EMIT) or its scratch slice (count pass, up to out_cap records)
  aha_hit* const out = EMIT ? (p.hits ? p.hits + p.offsets[d] : nullptr) : (p.doc_scratch ? p.doc_scratch + sc_base : nullptr);
  const uint32_t out_cap = EMIT ? 0xFFFFFFFFu : (uint32_t)min(sc_cap, (int64_t)0x7FFFFFFF);
  uint4* const out16 = EMIT ? nullptr : (p.doc_scratch16 ? p.doc_scratch16 + sc_base : nullptr);
  uint32_t run = 0;
  int32_t pos = 0, c_bpos = 0;
  int32_t cidx = 0;
  int phase = 0, depth = 0, first_l = 1;
  uint32_t node = 0;
  uint64_t tmask = 0;  // bit j = the character j places behind the cursor is upper case
  bool pending = false;  // the node just reached has candidates: they stand if what follows is a separator (or the end)
  // the deepest node with candidates that passed the separator test; WHICH candidate is settled once, when the walk is over
  int32_t best_n = 0, best_bytes = 0;
  uint32_t best_node = 0;
  uint64_t best_window = 0;
  // Every iteration decodes ONE character -- at the walk's cursor, or at `pos` in the other two phases -- and every
  // decision of the iteration is taken from it: the separator test serves the skip phases AND the candidate reached by
  // the previous step (the character behind a candidate is the one the next step decodes anyway), and the length of the
  // walk's first character is kept for the miss.  Nothing is decoded twice and the phases share one instruction stream.
  while (phase != 3) {
    const bool walking = phase == 1;
    const int32_t cur = walking ? c_bpos : pos;
    const bool have = cur < n;
    uint32_t cp = 0;
    int l = 1;
    if (have) l = utf8_decode32(s, cur, n, &cp);
    const bool sep = have && sep_at(cp);
    if (pending && (!use_sep || !have || sep)) {
      best_node = node;
      best_window = depth >= 64 ? tmask : (tmask & ((1ull << depth) - 1ull));
      best_n = depth;
      best_bytes = c_bpos - pos;
    }
    pending = false;
    if (!walking && !have) {  // the document is finished (once per document)
      phase = 3;
      continue;
    }
    // Straight-line from here to the end of a step: lanes that skip, start and walk run the same instructions.
    // skip: phase 2 runs up to (and over) the next separator, phase 0 over all of them
    const bool skip = !walking && (phase == 2 || (use_sep && sep));
    const bool start = !walking && !skip;  // a walk starts at this character (this iteration is its first step)
    pos += skip ? l : 0;
    cidx += skip ? 1 : 0;
    phase = (skip && sep) ? 0 : (start ? 1 : phase);
    depth = start ? 0 : depth;
    node = start ? 0u : node;
    c_bpos = start ? pos : c_bpos;
    tmask = start ? 0ull : tmask;
    best_n = start ? 0 : best_n;
    first_l = start ? l : first_l;
    const bool step = walking || start;
    // one step: the folded character's edge (a multi-byte character walks its bytes)
    uint32_t e = 0;
    bool upper = false;
    if (step && have && depth < 64) {
      upper = cp < 0x80u ? (cp - 'A' < 26u) : cp_upper(cp);
      const uint32_t fc = upper ? cp + 32u : cp;
      if (fc < 0x80u) {
        e = fc ? __ldg(fdelta + (node * fC + fcmap[fc])) : 0u;
      } else {
        uint8_t buf[4];
        const int k = utf8_encode(fc, buf);
        e = node;
        for (int j = 0; j < k; ++j) {
          e = __ldg(fdelta + ((e & ~kFoldCand) * fC + fcmap[buf[j]]));
          if (e == 0u) break;
        }
      }
    }
    const bool adv = e != 0u;
    node = adv ? (e & ~kFoldCand) : node;
    c_bpos += adv ? l : 0;
    depth += adv ? 1 : 0;
    tmask = adv ? ((tmask << 1) | (upper ? 1ull : 0ull)) : tmask;
    pending = adv && (e & kFoldCand) != 0u;
    if (!step || adv) continue;
    // the walk from this start is over: the deepest candidate that passed the separator test wins
    if (best_n == 0) {
      pos += first_l;
      ++cidx;
      phase = use_sep ? 2 : 0;
      continue;
    }
    {
      if (!EMIT && out16) {
        // the count pass only notes the node: the compaction settles the key, coalesced and off the walk
        if (run < out_cap) out16[run] = make_uint4((uint32_t)cidx, (uint32_t)best_n, best_node, (uint32_t)best_window);
      } else {
        const int32_t best_v = folded_pick(p, best_node, best_window);
        if (out && run < out_cap) out[run] = aha_hit{cidx, cidx + best_n, best_v};
      }
      ++run;
      pos += best_bytes;
      cidx += best_n;
      phase = 0;
    }
  }
  if (!EMIT) p.counts[d] = (uint32_t)run;
}

// deferred records of the folded path -> final hits: one warp per document, the key behind each node settled here
__global__ void __launch_bounds__(256) trie_compact_folded_kernel(const ScanParams p) {
  const int64_t d = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (d >= p.n_docs || !p.hits) return;
  const int64_t sc_base = p.doc_off[d] >> 2, sc_cap = (p.doc_off[d + 1] >> 2) - sc_base;
  const int64_t cnt = p.counts[d];
  if (cnt > sc_cap) return;  // walked again by the EMIT pass
  const uint4* src = p.doc_scratch16 + sc_base;
  aha_hit* dst = p.hits + p.offsets[d];
  for (int64_t i = lane; i < cnt; i += 32) {
    const uint4 r = src[i];
    dst[i] = aha_hit{(int32_t)r.x, (int32_t)(r.x + r.y), folded_pick(p, r.z, (uint64_t)r.w)};
  }
}

// scratch slices -> final positions: one warp per document, 12-byte records copied as words
__global__ void __launch_bounds__(256) trie_compact_kernel(const ScanParams p) {
  const int64_t d = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (d >= p.n_docs || !p.hits) return;
  const int64_t sc_base = p.doc_off[d] >> 2, sc_cap = (p.doc_off[d + 1] >> 2) - sc_base;
  const int64_t cnt = p.counts[d];
  if (cnt > sc_cap) return;  // walked again by the EMIT pass
  const uint32_t* src = reinterpret_cast<const uint32_t*>(p.doc_scratch + sc_base);
  uint32_t* dst = reinterpret_cast<uint32_t*>(p.hits + p.offsets[d]);
  for (int64_t i = lane; i < cnt * 3; i += 32) dst[i] = src[i];
}

}  // namespace aha
