// scan_char.cuh -- the walk over UTF-8 units with the shared-memory CHARACTER automaton (char_table.hpp).
//
// One automaton step consumes a whole UTF-8 unit with shared-memory lookups only (root entry of the character, edge
// probes from the state and from its fail target); nothing on the walk touches global memory except fire-and-forget
// event stores.  An event is ONE 32-bit word:
//     bit 31 = 0 :  rel:15 << 16 | ref:16     rel = bytes (or, with AHA_UNITS_CHAR, unit starts) in [tile start, last
//                                             byte of the match], 1..4032; ref = firing edge slot, or m + cid for a
//                                             one-character state (index into ref_rec[])
//     bit 31 = 1 :  document index            a document starts here in the stream (marker)
// The document-relative end offset is rebuilt by the expansion kernel (scan_char_split.cuh) from the tile's base and
// the base of the document the event lies in.  Tiles whose text is not structurally valid UTF-8 are re-walked with
// the byte automaton (ByteGlobal).
#pragma once
#include "char_table.hpp"
#include "char_walk.cuh"
#include "scan_fused.cuh"

#include <type_traits>

namespace aha {

constexpr uint32_t kEvMark = 0x80000000u;  // marker event: low 31 bits = document index

// build-time variants of the walk (tools/ab.py compares them on one box)
#ifndef AHA_W_FAST
#define AHA_W_FAST 0     // 12-byte steps that lie wholly inside every sub-chunk's owned region skip the range and slot-bound tests
                         // (measured on C2: 9 % fewer instructions per character, walk 2.55 -> 2.69 ms: the second copy of the step costs more)
#endif
#ifndef AHA_W_COUNT
#define AHA_W_COUNT 0    // hits counted inside the walk from a 2-bit table in shared memory (measured on C2: walk 2.45 ms against
                         // 2.05 ms without: one more shared-memory lookup per character; char_count_kernel does it instead)
#endif
#ifndef AHA_W_PRED2
#define AHA_W_PRED2 1    // the fail-link probe only in lanes that have a fail link (C2: walk 2.008 -> 1.981 ms)
#endif
#ifndef AHA_W_FMA
#define AHA_W_FMA 0      // table addresses as multiply-adds (multiply pipe) instead of three-input adds / LEA (integer pipe): 2.55 -> 2.58 ms
#endif

struct CharDeviceTables {
  const uint16_t* root_g;   // global copies, loaded into shared memory at kernel start
  const uint16_t* check_g;
  const uint32_t* val_g;
  const uint16_t* gfail;    // [m]: the first n_gfail_s entries are copied to shared memory
  uint32_t n_gfail_s;
  const RefRec* ref_rec;    // stays in global memory
  uint32_t n_root, m;
  uint32_t m_pad;           // entries of check[] (m + n_cid + 1, padding = 0xFFFF)
  uint32_t n_mwords;        // words of the hit-count table
  const uint32_t* mbits_g;  // 2 bits per ref: keys that fire there (1..3), 0 = see ref_rec
  uint32_t lo[3], n[3], off[3];
};

struct CharParams {
  FusedParams f;            // f.scratch is unused here
  CharDeviceTables t;
  const uint8_t* cmap_g;    // byte class map for the fallback walk
};

// counts only (fallback / overflow paths)
struct CountSink {
  static constexpr bool kEvents = false;
  uint32_t n = 0;
  __device__ __forceinline__ void hit(int32_t, int32_t, int32_t) { ++n; }
  __device__ __forceinline__ void doc_start(int64_t) {}
};

// writes the hits of one event at out[at ...]
__device__ __forceinline__ void expand_event(const ScanParams& p, const RefRec& rr, int32_t end, bool chars, aha_hit* out, int64_t at,
                                             int64_t cap) {
  if (rr.n == 1) {
    if (at < cap) out[at] = chars ? aha_hit{end - 1 - (int32_t)rr.c, end, (int32_t)rr.a} : aha_hit{end - (int32_t)rr.b, end, (int32_t)rr.a};
    return;
  }
  for (uint32_t j = rr.a; j < rr.b; ++j, ++at) {
    const int32_t v = p.out_val[j];
    if (at < cap) out[at] = chars ? aha_hit{end - 1 - (int32_t)p.key_tail[v], end, v} : aha_hit{end - (int32_t)p.key_len[v], end, v};
  }
}

// ---- fast tile walk -----------------------------------------------------------------------------------
// Tables as seen by one warp (all in shared memory except gfail).
// 32-bit shared-window addresses: plain pointers make the compiler rebuild the generic->shared base (S2R of the
// CTA id in the cluster + LEA) in front of every predicated table load.
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
  uint16_t v;
  asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
  return v;
}
// makes a value opaque to the compiler: it stays in a register instead of being recomputed at every use
__device__ __forceinline__ uint32_t pin_u32(uint32_t x) {
  asm volatile("" : "+r"(x));
  return x;
}
template <class T>
__device__ __forceinline__ T* pin_ptr(T* p) {
  asm volatile("" : "+l"(p));
  return p;
}
__device__ __forceinline__ uint32_t mad_addr(uint32_t i, uint32_t scale, uint32_t base) {
#if AHA_W_FMA
  uint32_t a;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(a) : "r"(i), "r"(scale), "r"(base));
  return a;
#else
  return base + i * scale;
#endif
}
struct SmemU32 {
  uint32_t base;
  __device__ __forceinline__ uint32_t operator[](uint32_t i) const { return lds_u32(mad_addr(i, 4u, base)); }
};
struct SmemU16 {
  uint32_t base;
  __device__ __forceinline__ uint32_t operator[](uint32_t i) const { return lds_u16(mad_addr(i, 2u, base)); }
};
struct CharSmem {
  SmemU16 root;            // [n_root]
  SmemU16 root3;           // root + off[1]: entries of the 3-byte units
  SmemU32 val;             // [m]
  SmemU16 check;           // [m_pad]
  SmemU16 gfail_s;         // [n_gfail_s] effective fail of the states a chain can pass through (they have the smallest ids)
  uint32_t n_gfail_s;
  const uint16_t* gfail;   // global: ids >= n_gfail_s (only when the prefix did not fit)
  SmemU32 mbits;           // [n_mwords] 2 bits per ref: ref_rec[ref].n if 1..3, else 0
  uint32_t m;
  uint32_t lo2, n2, off2, lo3, n3, off3, lo4, n4, off4;  // root index ranges of 2-, 3-, 4-byte units
  uint32_t neg_lo3_64;     // -(lo3 / 64): lo is a multiple of 64 (char_table.hpp)
};

// effective fail of state t: shared memory for the ids a chain passes through, global memory beyond the prefix
__device__ __forceinline__ uint32_t gfail_of(const CharSmem& S, uint32_t t) { return t < S.n_gfail_s ? S.gfail_s[t] : (uint32_t)S.gfail[t]; }

// One character.  r = root entry of the character (0: it starts no key), col = its column (= its root index).
// s / f1 = current state and its effective fail link, deep = that fail link has a non-root fail link of its own.
// Returns v = next:15 | fires:1 | fail of next:15 | deep of next:1 and the firing slot.
// Two probes (the state, then its fail link) cover everything except "deep fail link missed as well", which walks
// the gfail chain; kVote = the whole warp is converged here.
template <bool kVote>
__device__ __forceinline__ uint32_t char_advance(const CharSmem& S, uint32_t r, uint32_t col, uint32_t s, uint32_t f1, bool deep, uint32_t& ref) {
  const uint32_t slot1 = s + col, slot2 = f1 + col;
  const uint32_t c1 = S.check[slot1], c2 = S.check[slot2];
  const bool hit1 = c1 == s, hit2 = c2 == f1;
  bool hit = hit1 || hit2;
  uint32_t slotv = hit1 ? slot1 : slot2;
  uint32_t v = r;  // the root entry in val[] format (fail = 0)
  if (hit) v = S.val[slotv];
  const bool rare = !hit && deep;
  if (kVote ? __any_sync(0xffffffffu, rare) : rare) {
    if (rare) {
      uint32_t t = gfail_of(S, f1);
      while (t != 0) {
        const uint32_t sl = t + col;
        if (S.check[sl] == t) {
          v = S.val[sl];
          slotv = sl;
          hit = true;
          break;
        }
        t = gfail_of(S, t);
      }
    }
  }
  ref = hit ? slotv : S.m + col;
  return v;
}

// The same transition split in two for the interleaved walk: char_probe() is straight-line code (so that the
// probes of several sub-chunks share one basic block and overlap their shared-memory latencies),
// char_rare() is the out-of-line continuation along gfail[].
struct ProbeResult {
  uint32_t v, slotv, cid;  // cid: the character's column
  bool hit, rare;
};
__device__ __forceinline__ ProbeResult char_probe(const CharSmem& S, uint32_t vroot, uint32_t col, uint32_t s, uint32_t f1, bool deep) {
  ProbeResult q;
  q.cid = col;
  const uint32_t slot1 = s + col, slot2 = f1 + col;
  const uint32_t c1 = S.check[slot1];
#if AHA_W_PRED2
  uint32_t c2 = 0xFFFFu;  // the second probe only where there is a fail link to probe: fewer lanes, fewer bank conflicts
  if (f1 != 0u) c2 = S.check[slot2];
#else
  const uint32_t c2 = S.check[slot2];
#endif
  const bool hit1 = c1 == s, hit2 = c2 == f1;
  q.hit = hit1 || hit2;
  const uint32_t slot12 = hit1 ? slot1 : slot2;
  q.v = vroot;  // the root entry in val[] format (fail = 0)
  if (q.hit) q.v = S.val[slot12];
  q.slotv = q.hit ? slot12 : S.m + col;  // what fires, as a ref_rec[] index
  q.rare = !q.hit && deep;
  return q;
}
// returns {val entry, slot} of the first fail-chain state behind f1 that has an edge in column col; slot 0 = none
// (out of line; the tables by value: a reference would force the caller's copy into local memory)
__device__ __noinline__ uint2 char_rare(SmemU16 check, SmemU32 val, SmemU16 gfail_s, uint32_t n_gfail_s, const uint16_t* gfail, uint32_t f1,
                                        uint32_t col) {
  uint32_t t = f1 < n_gfail_s ? gfail_s[f1] : (uint32_t)gfail[f1];
  while (t != 0) {
    const uint32_t sl = t + col;
    if (check[sl] == t) return make_uint2(val[sl], sl);
    t = t < n_gfail_s ? gfail_s[t] : (uint32_t)gfail[t];
  }
  return make_uint2(0u, 0u);
}

constexpr int kChains = 2;              // independent automata per lane: the lane's chunk is split in kChains sub-chunks

// Walks one tile.  Every lane owns [a, b) of the concatenated text, split into kChains sub-chunks of `sub`
// bytes that the lane walks INTERLEAVED (independent dependency chains hide the shared-memory latency of the
// root -> check -> val lookups).  Each sub-chunk runs the per-document segment logic of char_walk()
// (warm-up, owned region, document boundaries with markers) as a state machine driven by a warp-uniform loop:
// whenever every busy sub-chunk has 12 well-formed bytes of one kind in front of it (four 3-byte units, or
// twelve ASCII characters) the warp takes a branch-free unrolled step; near a boundary (region end, document
// end, mixed unit lengths) the warp falls back to one unit per iteration.
// one_doc: the whole tile lies in document cd = [cdoc_start, cdoc_end).
// The walk stores events only -- no hit counting -- slot-major: event i of sub-chunk c at my_ev[i * 64 + c], so the
// i-th events of all 64 sub-chunks of a tile share sectors and the store is written and read densely.  n_ev[c]
// counts in steps of 64 (that element stride), on_doc included; the caller divides when the walk returns.  Events
// beyond kStride slots are dropped but counted (the caller then sends the tile through the direct pass).
// rel_base / rel0: an event's `rel` is (offset of the last byte in sbuf) + rel_base in byte units, or
// rel0[c] + (units that start in the sub-chunk up to the match end) with kChars; both are relative to the tile start
// and never reset at a document boundary -- the expansion subtracts the document's base.
// Returns false (warp-uniform) when the text is not structurally valid UTF-8: the caller rescans the tile with the
// byte automaton.
// kSep (byte offsets only): ACX#match(seq, sep)'s veto of the byte AFTER a match (ac.cr:326) is applied here, where the text
// is in shared memory: an event whose next byte rejects (and lies in the same document) is never stored.  With token
// separators that is nearly every event, so the store, the count and the expansion shrink to what can still become a hit.
template <bool kChars, int kStride, bool kSep, class OnDoc>
__device__ __forceinline__ bool char_tile_walk(const CharSmem& S, const ScanParams& p, const uint8_t* sbuf, int64_t origin, int64_t a,
                                               int64_t b, int32_t sub, int32_t stage_end, int32_t warmc, bool one_doc, int64_t cd,
                                               int64_t cdoc_start, int64_t cdoc_end, int32_t rel_base, const uint32_t (&rel0)[kChains],
                                               uint32_t* my_ev, uint32_t (&n_ev)[kChains], uint32_t& n_hit, bool& lookup, OnDoc on_doc) {
  constexpr unsigned kFull = 0xffffffffu;
  const uint32_t sbuf_s = pin_u32((uint32_t)__cvta_generic_to_shared(sbuf));
  bool ok = true;
  // per sub-chunk state (all offsets relative to sbuf)
  int32_t rb[kChains], a_seg[kChains], pos[kChains], de[kChains], seg_end[kChains];
  uint32_t s[kChains], f1[kChains], k[kChains];
  bool deep[kChains], done[kChains];
  // (the document index and its 64-bit bounds are not carried: the boundary code below re-derives them from de)

  auto enter_segment = [&](int c, int64_t doc_start, int64_t doc_end) {
    de[c] = (int32_t)min(doc_end - origin, (int64_t)1 << 30);
    seg_end[c] = min(rb[c], de[c]);
    int64_t w = origin + a_seg[c] - warmc;
    const bool at_start = w < doc_start || origin + a_seg[c] == doc_start;
    if (at_start) w = doc_start;
    pos[c] = (int32_t)(w - origin);
    s[c] = 0; f1[c] = 0; deep[c] = false;
    if (!at_start) {  // the warm-up may begin inside a unit
      int skipped = 0;
      while (pos[c] < seg_end[c] && (sbuf[pos[c]] & 0xC0) == 0x80) {
        ++pos[c];
        if (++skipped > 3) { ok = false; break; }
      }
    }
  };
#pragma unroll
  for (int c = 0; c < kChains; ++c) {
    const int64_t ac = min(a + (int64_t)c * sub, b);
    const int64_t bc = c == kChains - 1 ? b : min(ac + sub, b);
    rb[c] = (int32_t)(bc - origin);
    a_seg[c] = (int32_t)(ac - origin);
    done[c] = ac >= bc;
    n_ev[c] = 0;
    pos[c] = rb[c]; de[c] = 0; seg_end[c] = 0; s[c] = 0; f1[c] = 0; k[c] = kChars ? rel0[c] : 0u; deep[c] = false;
    if (!done[c]) {
      int64_t d = cd, doc_start = cdoc_start, doc_end = cdoc_end;
      if (!one_doc) {
        d = last_le(p.doc_off, p.n_docs, ac);
        doc_start = p.doc_off[d];
        doc_end = p.doc_off[d + 1];
      }
      if (doc_start == ac) {
        for (int64_t e = first_ge(p.doc_off, p.n_docs, ac); e <= d; ++e) on_doc(c, e, n_ev[c]);
      }
      enter_segment(c, doc_start, doc_end);
    }
  }

  // consume one character of sub-chunk c whose root entry is r; cs / last = offsets of its first / last byte
  auto step = [&](int c, uint32_t r, uint32_t col, int32_t cs, int32_t last, auto vote) {
    uint32_t ref;
    const uint32_t v = char_advance<decltype(vote)::value>(S, r, col, s[c], f1[c], deep[c], ref);
    s[c] = v & 0x7FFFu;
    f1[c] = (v >> 16) & 0x7FFFu;
    deep[c] = (v >> 31) != 0u;
    if (kChars) k[c] += (cs >= a_seg[c]) ? 1u : 0u;
    bool fire = (v & 0x8000u) && (uint32_t)(last - a_seg[c]) < (uint32_t)(rb[c] - a_seg[c]);
    if (kSep && fire && last + 1 < de[c] && sep_rejects(p, sbuf[last + 1])) fire = false;
    if (fire) {  // n_ev[c] counts in steps of 64 = the slot-major element stride (see multi_step)
      if (n_ev[c] < (uint32_t)kStride * 64u) my_ev[n_ev[c] + c] = (kChars ? k[c] : (uint32_t)(last + rel_base)) * 65536u + ref;
      n_ev[c] += 64u;
#if AHA_W_COUNT
      const uint32_t cnt = (S.mbits[ref >> 4] >> ((ref & 15u) * 2u)) & 3u;
      n_hit += cnt;
      lookup = lookup || cnt == 0u;
#endif
    }
  };

  // one character of EVERY sub-chunk (warp converged): probes first, then the rare continuation behind a single
  // vote, then the state update, then the events (predicated, no branch).
  // evw[c] = the event word of this character without its ref (byte units: rel << 16); rl[c] = offset of the
  // character's last byte relative to a_seg[c], span[c] = rb[c] - a_seg[c] (hoisted by the caller).
  // fast: the caller has established that every sub-chunk of the warp owns this character and has slots left, so
  // the range test and the slot-bound test are not made.
  auto multi_step = [&](auto fast, const uint32_t (&vr)[kChains], const uint32_t (&ci)[kChains], const uint32_t (&evw)[kChains], int32_t len,
                        const int32_t (&rl)[kChains], const uint32_t (&span)[kChains]) {
    constexpr bool kFast = decltype(fast)::value;
    ProbeResult q[kChains];
    bool any_rare = false;
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
      q[c] = char_probe(S, vr[c], ci[c], s[c], f1[c], deep[c]);
      any_rare = any_rare || q[c].rare;
    }
    if (__any_sync(kFull, any_rare)) {
#pragma unroll
      for (int c = 0; c < kChains; ++c)
        if (q[c].rare) {
          const uint2 x = char_rare(S.check, S.val, S.gfail_s, S.n_gfail_s, S.gfail, f1[c], q[c].cid);
          if (x.y) { q[c].v = x.x; q[c].slotv = x.y; q[c].hit = true; }
        }
    }
    bool fire[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
      const uint32_t v = q[c].v;
      s[c] = v & 0x7FFFu;
      f1[c] = (v >> 16) & 0x7FFFu;
      deep[c] = (v >> 31) != 0u;
      if (kChars) k[c] += kFast ? 1u : ((rl[c] >= len - 1) ? 1u : 0u);
      // the fires bit as a sign test of v << 16: the shift can run on the multiply pipe, the test folds into the range compare
      int32_t vs;
      asm("shl.b32 %0, %1, 16;" : "=r"(vs) : "r"(v));
      fire[c] = kFast ? (vs < 0) : ((uint32_t)rl[c] < span[c] && vs < 0);
      if (kSep) {
        const int32_t nx = a_seg[c] + rl[c] + 1;  // the byte after the character (staged: the tile's 16 bytes of slack)
        const uint32_t chr = lds_u8(sbuf_s + (uint32_t)max(nx, 0));
        fire[c] = fire[c] && !(nx < de[c] && sep_rejects(p, chr));
      }
    }
    // branch-free: a handful of predicated instructions per sub-chunk cost less than the divergent branches around them
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
      // n_ev[c] is kept pre-multiplied by the slot-major element stride: the address is one multiply-add away
      if (fire[c] && (kFast || n_ev[c] < (uint32_t)kStride * 64u)) my_ev[n_ev[c] + c] = (kChars ? k[c] * 65536u : evw[c]) + q[c].slotv;
      if (fire[c]) n_ev[c] += 64u;
#if AHA_W_COUNT
      // hits of the event from the 2-bit table (0: more than three keys fire there, the caller looks it up)
      const uint32_t ref = q[c].slotv;
      const uint32_t cnt = fire[c] ? (S.mbits[ref >> 4] >> ((ref & 15u) * 2u)) & 3u : 4u;
      n_hit += cnt & 3u;
      lookup = lookup || cnt == 0u;
#endif
    }
  };

  if (__any_sync(kFull, !ok)) return false;
  {
    bool all_done = true;
#pragma unroll
    for (int c = 0; c < kChains; ++c) all_done = all_done && done[c];
    if (__all_sync(kFull, all_done)) return true;
  }
  // (ok and done[] only change in the one-unit branch below: that is where they are voted on)
  for (;;) {
    uint32_t A[kChains], B[kChains], C[kChains];
    // straight-line integer logic (no short-circuit branches): bad3 / bad1 = 0 iff the sub-chunk is done, or has twelve
    // bytes of room and they are four well-formed 3-byte units / twelve ASCII bytes
    uint32_t bad3 = 0, bad1 = 0;
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
      const uint32_t wa = sbuf_s + ((uint32_t)pos[c] & ~3u);
      const uint32_t w0 = lds_u32(wa), w1 = lds_u32(wa + 4), w2 = lds_u32(wa + 8), w3 = lds_u32(wa + 12);
      const uint32_t sh = ((uint32_t)pos[c] & 3u) * 8u;
      A[c] = __funnelshift_r(w0, w1, sh);
      B[c] = __funnelshift_r(w1, w2, sh);
      C[c] = __funnelshift_r(w2, w3, sh);
      // sign bit set when there is no room: pos >= seg_end, or pos + 12 > min(de, stage_end)
      const uint32_t noroom = (uint32_t)((seg_end[c] - 1 - pos[c]) | (min(de[c], stage_end) - 12 - pos[c])) & 0x80000000u;
      const uint32_t live = done[c] ? 0u : 0xFFFFFFFFu;
      const uint32_t x3 =
          ((A[c] ^ 0xE08080E0u) & 0xF0C0C0F0u) | ((B[c] ^ 0x80E08080u) & 0xC0F0C0C0u) | ((C[c] ^ 0x8080E080u) & 0xC0C0F0C0u);
      bad3 |= live & (noroom | x3);
      bad1 |= live & noroom;  // the ASCII test itself is only made when the 3-byte step is not taken
    }
    const bool can3 = bad3 == 0u;
    if (__all_sync(kFull, can3)) {
      // four 3-byte units per sub-chunk: payload = (b0 & 0xF) << 12 | (b1 & 0x3F) << 6 | (b2 & 0x3F), root index = payload - lo
      auto root3 = [&](uint32_t w) {  // index of the character inside its range, clamped to the zero entry behind it
        const uint32_t wm = w & 0x3F3F0Fu;
        const uint32_t h = __dp4a(wm, 0x00000140u, S.neg_lo3_64);
        const uint32_t rel = h * 64u + (wm >> 16);
        return min(rel, S.n3);  // root3[n3] = 0 (char_table.hpp)
      };
      // twelve bytes = three words: character u sits at a fixed place in A:B:C, one funnel shift each
      int32_t rl0[kChains];
      uint32_t span[kChains], pe[kChains];
      bool plain = AHA_W_FAST != 0;  // every sub-chunk of the lane owns all four characters and has four slots left
#pragma unroll
      for (int c = 0; c < kChains; ++c) {
        rl0[c] = pos[c] + 2 - a_seg[c];
        span[c] = (uint32_t)(rb[c] - a_seg[c]);
        pe[c] = (uint32_t)(pos[c] + rel_base + 2) << 16;  // event word of the first character, ref still to be added
        plain = plain && !done[c] && pos[c] >= a_seg[c] && pos[c] + 12 <= rb[c] && n_ev[c] <= (uint32_t)(kStride - 4) * 64u;
      }
      auto four = [&](auto fast) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          uint32_t vr[kChains], ci[kChains], evw[kChains];
          int32_t rl[kChains];
#pragma unroll
          for (int c = 0; c < kChains; ++c) {
            const uint32_t w = u == 0 ? A[c] : (u == 1 ? __funnelshift_r(A[c], B[c], 24) : (u == 2 ? __funnelshift_r(B[c], C[c], 16) : C[c] >> 8));
            const uint32_t ri = root3(w);
            vr[c] = S.root3[ri];      // state:15 | fires:1 of the character's own state
            ci[c] = ri + S.off3;      // its column = its root index: nothing to look up
            evw[c] = pe[c] + ((uint32_t)(3 * u) << 16);
            rl[c] = rl0[c] + 3 * u;
          }
          multi_step(fast, vr, ci, evw, 3, rl, span);
        }
      };
      if (AHA_W_FAST && __all_sync(kFull, plain)) four(std::true_type());
      else four(std::false_type());
#pragma unroll
      for (int c = 0; c < kChains; ++c)
        if (!done[c]) pos[c] += 12;
    } else if (__all_sync(kFull, [&] {
#pragma unroll
                 for (int c = 0; c < kChains; ++c) bad1 |= (done[c] ? 0u : 0xFFFFFFFFu) & ((A[c] | B[c] | C[c]) & 0x80808080u);
                 return bad1 == 0u;
               }())) {
      // twelve ASCII characters per sub-chunk
      uint32_t span[kChains];
#pragma unroll
      for (int c = 0; c < kChains; ++c) span[c] = (uint32_t)(rb[c] - a_seg[c]);
#pragma unroll 1
      for (int i = 0; i < 12; ++i) {
        uint32_t r[kChains], evw[kChains];
        int32_t rl[kChains];
#pragma unroll
        for (int c = 0; c < kChains; ++c) {
          r[c] = A[c] & 0xFFu;
          rl[c] = pos[c] + i - a_seg[c];
          evw[c] = (uint32_t)(pos[c] + i + rel_base) << 16;
          A[c] = __funnelshift_r(A[c], B[c], 8);
          B[c] = __funnelshift_r(B[c], C[c], 8);
          C[c] >>= 8;
        }
        uint32_t vr[kChains], ci[kChains];
#pragma unroll
        for (int c = 0; c < kChains; ++c) {
          vr[c] = S.root[r[c]];
          ci[c] = r[c];
        }
        multi_step(std::false_type(), vr, ci, evw, 1, rl, span);
      }
#pragma unroll
      for (int c = 0; c < kChains; ++c)
        if (!done[c]) pos[c] += 12;
    } else {
#pragma unroll
      for (int c = 0; c < kChains; ++c) {
        if (done[c]) continue;
        if (pos[c] >= seg_end[c]) {
          // end of the segment: the region is finished, or the next document starts here
          const int64_t bc = origin + rb[c];
          const int64_t at = origin + de[c];  // end of the current document (a clamped de only ever makes it larger than bc)
          const bool more = at <= bc && (at < bc || bc == p.total);
          if (!more) {
            done[c] = true;
          } else {
            int64_t d = last_le(p.doc_off, p.n_docs, at - 1) + 1;  // the current document is not empty
            while (d < p.n_docs) {
              on_doc(c, d, n_ev[c]);
              if (p.doc_off[d + 1] > at) break;
              ++d;
            }
            if (d >= p.n_docs || at >= bc) {
              done[c] = true;
            } else {
              a_seg[c] = (int32_t)(p.doc_off[d] - origin);
              enter_segment(c, p.doc_off[d], p.doc_off[d + 1]);
            }
          }
        } else {
          // one unit of any length
          const uint32_t b0 = A[c] & 0xFFu;
          int32_t len = 1;
          uint32_t r = 0, col = b0;
          if (b0 < 0x80u) {
            r = S.root[b0];
          } else if (b0 < 0xC0u || b0 >= 0xF8u) {
            ok = false;
          } else {
            len = b0 < 0xE0u ? 2 : (b0 < 0xF0u ? 3 : 4);
            const uint32_t b1 = (A[c] >> 8) & 0xFFu, b2 = (A[c] >> 16) & 0xFFu, b3 = A[c] >> 24;
            if (pos[c] + len > de[c]) ok = false;  // cut by the end of the document
            if ((b1 & 0xC0u) != 0x80u) ok = false;
            if (len >= 3 && (b2 & 0xC0u) != 0x80u) ok = false;
            if (len == 4 && (b3 & 0xC0u) != 0x80u) ok = false;
            uint32_t code, lo, n, off;
            if (len == 2) {
              code = ((b0 & 0x1Fu) << 6) | (b1 & 0x3Fu);
              lo = S.lo2; n = S.n2; off = S.off2;
            } else if (len == 3) {
              code = ((b0 & 0x0Fu) << 12) | ((b1 & 0x3Fu) << 6) | (b2 & 0x3Fu);
              lo = S.lo3; n = S.n3; off = S.off3;
            } else {
              code = ((b0 & 0x07u) << 18) | ((b1 & 0x3Fu) << 12) | ((b2 & 0x3Fu) << 6) | (b3 & 0x3Fu);
              lo = S.lo4; n = S.n4; off = S.off4;
            }
            col = off + min(code - lo, n);  // out of range: the zero entry behind the range (a column no state has)
            if (ok) r = S.root[col];
          }
          if (ok) {
            if (pos[c] + len - 1 >= rb[c]) {
              done[c] = true;  // the unit ends in the next region (the document goes on beyond it)
            } else {
              step(c, r, col, pos[c], pos[c] + len - 1, std::false_type());
              pos[c] += len;
            }
          }
        }
        if (done[c]) pos[c] = rb[c];
      }
      if (__any_sync(kFull, !ok)) return false;
      bool all_done = true;
#pragma unroll
      for (int c = 0; c < kChains; ++c) all_done = all_done && done[c];
      if (__all_sync(kFull, all_done)) break;
    }
  }
  return true;
}

}  // namespace aha
