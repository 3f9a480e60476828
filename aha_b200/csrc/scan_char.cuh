// scan_char.cuh -- fused single-pass scan with the shared-memory CHARACTER automaton (char_table.hpp).
//
// Same warp-per-tile pipeline as scan_fused.cuh (TMA bulk staging, owned regions, decoupled look-back),
// but one automaton step consumes a whole UTF-8 unit with two shared-memory lookups (root entry of the
// character, one edge probe) and nothing on the walk touches global memory except fire-and-forget event
// stores: an event is {end offset, firing edge}.  After the look-back the warp expands its events
// cooperatively (one event per lane per round) through ref_rec[] into final Aha::Hit records, so the
// L2 latency of the output tables is paid event-parallel, not on the dependent chain of the walk.
// Tiles whose text is not structurally valid UTF-8 are re-walked with the byte automaton (ByteGlobal).
#pragma once
#include "char_table.hpp"
#include "char_walk.cuh"
#include "scan_fused.cuh"

#include <type_traits>

namespace aha {

constexpr int kEvCap = 32;     // events per lane per tile the deferred delivery can handle (more: direct pass)
constexpr int kEvSlots = 144;  // ring slots per lane: one per byte of the lane's chunk, so the walk never checks a bound

struct alignas(8) CharEvent {
  int32_t end;   // document-relative end offset (bytes, or characters with AHA_UNITS_CHAR)
  uint32_t ref;  // firing edge slot, or m + cid for a one-character state: index into ref_rec[]
};

struct CharDeviceTables {
  const uint32_t* root_g;   // global copies, loaded into shared memory at kernel start
  const uint16_t* check_g;
  const uint32_t* val_g;
  const uint16_t* gfail;    // stays in global memory
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
  CharEvent* events;        // [n_warps_total][2][32][kEvSlots]  (two tiles in flight per warp)
  const uint8_t* cmap_g;    // byte class map for the fallback walk
};

// counts only (fallback / overflow paths)
struct CountSink {
  static constexpr bool kEvents = false;
  uint32_t n = 0;
  __device__ __forceinline__ void hit(int32_t, int32_t, int32_t) { ++n; }
  __device__ __forceinline__ void doc_start(int64_t) {}
};

// One lane's pass over its chunk [a, b) with the character automaton.  on_event(end, ref_with_multi_bit),
// on_doc(d).  Returns false when malformed UTF-8 was met.
template <class OnEvent, class OnDoc>
__device__ __forceinline__ bool char_lane_walk(const CharParams& P, const CharTablesView& T, const uint8_t* sbuf, int64_t origin,
                                               int64_t a, int64_t b, uint32_t lead0, OnEvent on_event, OnDoc on_doc) {
  const ScanParams& p = P.f.sp;
  if (a >= b) return true;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  int64_t d = last_le(p.doc_off, p.n_docs, a);
  int64_t doc_start = p.doc_off[d], doc_end = p.doc_off[d + 1];
  if (doc_start == a) {
    for (int64_t e = first_ge(p.doc_off, p.n_docs, a); e <= d; ++e) on_doc(e);
  }
  const int32_t warmc = (int32_t)p.warm + 3;
  int64_t seg_a = a;           // owned region start inside the current document
  uint32_t lead = lead0;       // characters of the document before seg_a
  bool ok = true;
  for (;;) {
    int64_t w = seg_a - warmc;
    if (w < doc_start || seg_a == doc_start) w = doc_start;
    const int64_t seg_end = min(b, doc_end);
    const int st = char_walk(T, sbuf, (int32_t)(w - origin), (int32_t)(seg_a - origin), (int32_t)(b - origin),
                             (int32_t)(seg_end - origin), (int32_t)(doc_end - origin), w == doc_start,
                             [&](uint32_t ref, int32_t last, uint32_t k) {
                               const int32_t end = chars ? (int32_t)(lead + k) : (int32_t)(last + origin - doc_start + 1);
                               on_event(end, ref);
                             });
    if (st != kWalkOk) { ok = false; break; }
    if (!(doc_end <= b && (doc_end < b || b == p.total))) break;  // chunk ends inside this document
    // step into the next non-empty document
    const int64_t pos = doc_end;
    ++d;
    while (d < p.n_docs) {
      on_doc(d);
      if (p.doc_off[d + 1] > pos) break;
      ++d;
    }
    if (d >= p.n_docs || pos >= b) break;
    doc_start = p.doc_off[d];
    doc_end = p.doc_off[d + 1];
    seg_a = doc_start;
    lead = 0;
  }
  return ok;
}

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
struct SmemU32 {
  uint32_t base;
  __device__ __forceinline__ uint32_t operator[](uint32_t i) const { return lds_u32(base + i * 4u); }
};
struct SmemU16 {
  uint32_t base;
  __device__ __forceinline__ uint32_t operator[](uint32_t i) const { return lds_u16(base + i * 2u); }
};
struct CharSmem {
  SmemU32 root;            // [n_root]
  SmemU32 root3;           // root + off[1]: entries of the 3-byte units
  SmemU32 val;             // [m]
  SmemU16 check;           // [m_pad]
  const uint16_t* gfail;   // global
  SmemU32 mbits;           // [n_mwords] 2 bits per ref: ref_rec[ref].n if 1..3, else 0
  uint32_t m;
  uint32_t lo2, n2, off2, lo3, n3, off3, lo4, n4, off4;  // root index ranges of 2-, 3-, 4-byte units
  uint32_t neg_lo3_64;     // -(lo3 / 64): lo is a multiple of 64 (char_table.hpp)
};

// One character.  r = root entry of the character (0: not in the dictionary).  s / f1 = current state and
// its effective fail link, deep = that fail link has a non-root fail link of its own.
// Returns v = next:15 | fires:1 | fail of next:15 | deep of next:1 and the firing slot.
// Two unconditional probes (the state, then its fail link) cover everything except "deep fail link
// missed as well", which walks gfail[] in global memory; kVote = the whole warp is converged here.
template <bool kVote>
__device__ __forceinline__ uint32_t char_advance(const CharSmem& S, uint32_t r, uint32_t s, uint32_t f1, bool deep, uint32_t& ref) {
  const uint32_t cid = r >> 16;
  const uint32_t slot1 = s + cid, slot2 = f1 + cid;
  const uint32_t c1 = S.check[slot1], c2 = S.check[slot2];
  const bool hit1 = c1 == s, hit2 = c2 == f1;
  bool hit = hit1 || hit2;
  uint32_t slotv = hit1 ? slot1 : slot2;
  uint32_t v = r & 0xFFFFu;  // the root entry in val[] format (fail = 0)
  if (hit) v = S.val[slotv];
  const bool rare = !hit && deep && cid != 0;
  if (kVote ? __any_sync(0xffffffffu, rare) : rare) {
    if (rare) {
      uint32_t t = S.gfail[f1];
      while (t != 0) {
        const uint32_t sl = t + cid;
        if (S.check[sl] == t) {
          v = S.val[sl];
          slotv = sl;
          hit = true;
          break;
        }
        t = S.gfail[t];
      }
    }
  }
  ref = hit ? slotv : S.m + cid;
  return v;
}

// The same transition split in two for the interleaved walk: char_probe() is straight-line code (so that the
// probes of several sub-chunks share one basic block and overlap their shared-memory latencies),
// char_rare() is the out-of-line continuation along gfail[].
struct ProbeResult {
  uint32_t v, slotv, cid;
  bool hit, rare;
};
__device__ __forceinline__ ProbeResult char_probe(const CharSmem& S, uint32_t vroot, uint32_t cid, uint32_t s, uint32_t f1, bool deep) {
  ProbeResult q;
  q.cid = cid;
  const uint32_t slot1 = s + q.cid, slot2 = f1 + q.cid;
  const uint32_t c1 = S.check[slot1], c2 = S.check[slot2];
  const bool hit1 = c1 == s, hit2 = c2 == f1;
  q.hit = hit1 || hit2;
  const uint32_t slot12 = hit1 ? slot1 : slot2;
  q.v = vroot;  // the root entry in val[] format (fail = 0)
  if (q.hit) q.v = S.val[slot12];
  q.slotv = q.hit ? slot12 : S.m + q.cid;  // what fires, as a ref_rec[] index
  q.rare = !q.hit && deep && q.cid != 0;
  return q;
}
// returns {val entry, slot} of the first fail-chain state behind f1 that has an edge for cid; slot 0 = none
__device__ __noinline__ uint2 char_rare(SmemU16 check, SmemU32 val, const uint16_t* gfail, uint32_t f1, uint32_t cid) {
  uint32_t t = gfail[f1];
  while (t != 0) {
    const uint32_t sl = t + cid;
    if (check[sl] == t) return make_uint2(val[sl], sl);
    t = gfail[t];
  }
  return make_uint2(0u, 0u);
}

constexpr int kChains = 2;              // independent automata per lane: the lane's chunk is split in kChains sub-chunks
constexpr int kEvCapC = kEvCap / kChains;      // events per sub-chunk the deferred delivery can handle
constexpr int kEvSlotsC = kEvSlots / kChains;  // ring slots per sub-chunk (>= bytes of a sub-chunk)

// Walks one tile.  Every lane owns [a, b) of the concatenated text, split into kChains sub-chunks of `sub`
// bytes that the lane walks INTERLEAVED (independent dependency chains hide the shared-memory latency of the
// root -> check -> val lookups).  Each sub-chunk runs the same per-document segment logic as char_lane_walk()
// (warm-up, owned region, document boundaries with markers) as a state machine driven by a warp-uniform loop:
// whenever every busy sub-chunk has 12 well-formed bytes of one kind in front of it (four 3-byte units, or
// twelve ASCII characters) the warp takes a branch-free unrolled step; near a boundary (region end, document
// end, mixed unit lengths) the warp falls back to one unit per iteration.
// one_doc: the whole tile lies in document cd = [cdoc_start, cdoc_end).
// Events of sub-chunk c go to my_ev[c * kStride ...) (slot-major when kBounded, see below); n_ev[c] counts all of them,
// n_hit[c] the hits they expand to as far as the 2-bit table knows (lookup_out = some event needs ref_rec[] for its
// count).  Returns false (warp-uniform) when the text is not structurally valid UTF-8: the caller rescans the tile
// with the byte automaton.
// kLean (the split pipeline): events only -- no hit counting in the walk -- and n_ev[] counts in steps of 64 (the
// slot-major element stride), on_doc included; the caller divides when the walk returns.
// kStride = ring slots per sub-chunk; kBounded: drop (but count) events beyond kStride instead of relying on
// kStride >= bytes of a sub-chunk, and lay the slots out slot-major (event i of sub-chunk c at my_ev[i * 64 + c]: the
// i-th events of all 64 sub-chunks of a tile share sectors, so the store is written and read densely).
template <bool kChars, int kStride, bool kBounded, bool kLean = false, class OnDoc>
__device__ __forceinline__ bool char_tile_walk(const CharSmem& S, const ScanParams& p, const uint8_t* sbuf, int64_t origin, int64_t a,
                                               int64_t b, int32_t sub, int32_t stage_end, int32_t warmc, bool one_doc, int64_t cd,
                                               int64_t cdoc_start, int64_t cdoc_end, const uint32_t (&lead0)[kChains], CharEvent* my_ev,
                                               uint32_t (&n_ev)[kChains], uint32_t (&n_hit)[kChains], bool& lookup_out, OnDoc on_doc) {
  constexpr unsigned kFull = 0xffffffffu;
  const uint32_t sbuf_s = pin_u32((uint32_t)__cvta_generic_to_shared(sbuf));
  bool ok = true;
  bool lookup = false;  // an event reports more than 3 keys: n_hit is a lower bound, the caller counts through ref_rec
  // per sub-chunk state (all offsets relative to sbuf)
  int32_t rb[kChains], a_seg[kChains], pos[kChains], de[kChains], seg_end[kChains], ebase[kChains];
  uint32_t s[kChains], f1[kChains], k[kChains], lead[kChains];
  bool deep[kChains], done[kChains];
  // (the document index and its 64-bit bounds are not carried: the boundary code below re-derives them from
  // ebase / de, which keeps 16 registers out of the hot loop)

  auto enter_segment = [&](int c, int64_t doc_start, int64_t doc_end) {
    de[c] = (int32_t)min(doc_end - origin, (int64_t)1 << 30);
    seg_end[c] = min(rb[c], de[c]);
    ebase[c] = (int32_t)(origin - doc_start) + 1;
    int64_t w = origin + a_seg[c] - warmc;
    const bool at_start = w < doc_start || origin + a_seg[c] == doc_start;
    if (at_start) w = doc_start;
    pos[c] = (int32_t)(w - origin);
    s[c] = 0; f1[c] = 0; k[c] = 0; deep[c] = false;
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
    n_hit[c] = 0;
    lead[c] = lead0[c];
    pos[c] = rb[c]; de[c] = 0; seg_end[c] = 0; ebase[c] = 0; s[c] = 0; f1[c] = 0; k[c] = 0; deep[c] = false;
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
  auto step = [&](int c, uint32_t r, int32_t cs, int32_t last, auto vote) {
    uint32_t ref;
    const uint32_t v = char_advance<decltype(vote)::value>(S, r, s[c], f1[c], deep[c], ref);
    s[c] = v & 0x7FFFu;
    f1[c] = (v >> 16) & 0x7FFFu;
    deep[c] = (v >> 31) != 0u;
    if (kChars) k[c] += (cs >= a_seg[c]) ? 1u : 0u;
    const bool fire = (v & 0x8000u) && (uint32_t)(last - a_seg[c]) < (uint32_t)(rb[c] - a_seg[c]);
    if (fire) {
      if (kLean) {  // n_ev[c] counts in steps of 64 = the slot-major element stride (see multi_step)
        if (n_ev[c] < (uint32_t)kStride * 64u) my_ev[n_ev[c] + c] = CharEvent{kChars ? (int32_t)(lead[c] + k[c]) : last + ebase[c], ref};
        n_ev[c] += 64u;
      } else {
        if (!kBounded || n_ev[c] < (uint32_t)kStride)
          my_ev[kBounded ? n_ev[c] * 64 + c : c * kStride + n_ev[c]] = CharEvent{kChars ? (int32_t)(lead[c] + k[c]) : last + ebase[c], ref};
        const uint32_t cnt = (S.mbits[ref >> 4] >> ((ref & 15u) * 2u)) & 3u;
        n_hit[c] += cnt;
        lookup = lookup || cnt == 0u;
        ++n_ev[c];
      }
    }
  };

  // one character of EVERY sub-chunk (warp converged): probes first, then the rare continuation behind a single
  // vote, then the state update, then the events behind a single branch
  // rl[c] = offset of the character's last byte relative to a_seg[c], span[c] = rb[c] - a_seg[c] (hoisted by the caller)
  auto multi_step = [&](const uint32_t (&vr)[kChains], const uint32_t (&ci)[kChains], const int32_t (&cs)[kChains], int32_t len,
                        const int32_t (&rl)[kChains], const uint32_t (&span)[kChains]) {
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
          const uint2 x = char_rare(S.check, S.val, S.gfail, f1[c], q[c].cid);
          if (x.y) { q[c].v = x.x; q[c].slotv = x.y; q[c].hit = true; }
        }
    }
    bool fire[kChains], any_fire = false;
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
      const uint32_t v = q[c].v;
      s[c] = v & 0x7FFFu;
      f1[c] = (v >> 16) & 0x7FFFu;
      deep[c] = (v >> 31) != 0u;
      if (kChars) k[c] += (rl[c] >= len - 1) ? 1u : 0u;
      // the fires bit as a sign test of v << 16: the shift can run on the multiply pipe, the test folds into the range compare
      int32_t vs;
      asm("shl.b32 %0, %1, 16;" : "=r"(vs) : "r"(v));
      fire[c] = (uint32_t)rl[c] < span[c] && vs < 0;
      any_fire = any_fire || fire[c];
    }
    if (kLean) {
      // branch-free: a handful of predicated instructions per sub-chunk cost less than the divergent branches around them
#pragma unroll
      for (int c = 0; c < kChains; ++c) {
        // n_ev[c] is kept pre-multiplied by the slot-major element stride: the address is one multiply-add away
        if (fire[c] && n_ev[c] < (uint32_t)kStride * 64u)
          my_ev[n_ev[c] + c] = CharEvent{kChars ? (int32_t)(lead[c] + k[c]) : cs[c] + len - 1 + ebase[c], q[c].slotv};
        if (fire[c]) n_ev[c] += 64u;
      }
    } else if (any_fire) {
#pragma unroll
      for (int c = 0; c < kChains; ++c) {
        if (fire[c]) {
          const uint32_t ref = q[c].slotv;
          if (!kBounded || n_ev[c] < (uint32_t)kStride)
            my_ev[kBounded ? n_ev[c] * 64 + c : c * kStride + n_ev[c]] =
                CharEvent{kChars ? (int32_t)(lead[c] + k[c]) : cs[c] + len - 1 + ebase[c], ref};
          if (!kLean) {
            const uint32_t cnt = (S.mbits[ref >> 4] >> ((ref & 15u) * 2u)) & 3u;
            n_hit[c] += cnt;
            lookup = lookup || cnt == 0u;
          }
          ++n_ev[c];
        }
      }
    }
  };

  if (__any_sync(kFull, !ok)) return false;
  {
    bool all_done = true;
#pragma unroll
    for (int c = 0; c < kChains; ++c) all_done = all_done && done[c];
    if (__all_sync(kFull, all_done)) { lookup_out = lookup; return true; }
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
      auto root3 = [&](uint32_t w) {
        const uint32_t wm = w & 0x3F3F0Fu;
        const uint32_t h = __dp4a(wm, 0x00000140u, S.neg_lo3_64);
        const uint32_t rel = h * 64u + (wm >> 16);
        return S.root3.base + min(rel, S.n3) * 4u;  // address of the entry; root3[n3] = 0 (char_table.hpp)
      };
      // twelve bytes = three words: character u sits at a fixed place in A:B:C, one funnel shift each
      int32_t rl0[kChains];
      uint32_t span[kChains];
#pragma unroll
      for (int c = 0; c < kChains; ++c) {
        rl0[c] = pos[c] + 2 - a_seg[c];
        span[c] = (uint32_t)(rb[c] - a_seg[c]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        uint32_t vr[kChains], ci[kChains];
        int32_t cs[kChains], rl[kChains];
#pragma unroll
        for (int c = 0; c < kChains; ++c) {
          const uint32_t w = u == 0 ? A[c] : (u == 1 ? __funnelshift_r(A[c], B[c], 24) : (u == 2 ? __funnelshift_r(B[c], C[c], 16) : C[c] >> 8));
          const uint32_t ra = root3(w);
          vr[c] = lds_u16(ra);      // the two halves of the entry as two loads: the load pipe has room, the integer pipe has not
          ci[c] = lds_u16(ra + 2);
          cs[c] = pos[c] + 3 * u;
          rl[c] = rl0[c] + 3 * u;
        }
        multi_step(vr, ci, cs, 3, rl, span);
      }
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
        uint32_t r[kChains];
        int32_t cs[kChains], rl[kChains];
#pragma unroll
        for (int c = 0; c < kChains; ++c) {
          r[c] = S.root[A[c] & 0xFFu];
          cs[c] = pos[c] + i;
          rl[c] = cs[c] - a_seg[c];
          A[c] = __funnelshift_r(A[c], B[c], 8);
          B[c] = __funnelshift_r(B[c], C[c], 8);
          C[c] >>= 8;
        }
        uint32_t vr[kChains], ci[kChains];
#pragma unroll
        for (int c = 0; c < kChains; ++c) {
          vr[c] = r[c] & 0xFFFFu;
          ci[c] = r[c] >> 16;
        }
        multi_step(vr, ci, cs, 1, rl, span);
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
              lead[c] = 0;
              enter_segment(c, p.doc_off[d], p.doc_off[d + 1]);
            }
          }
        } else {
          // one unit of any length
          const uint32_t b0 = A[c] & 0xFFu;
          int32_t len = 1;
          uint32_t r = 0;
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
            const uint32_t rel = code - lo;
            if (ok && rel < n) r = S.root[off + rel];
          }
          if (ok) {
            if (pos[c] + len - 1 >= rb[c]) {
              done[c] = true;  // the unit ends in the next region (the document goes on beyond it)
            } else {
              step(c, r, pos[c], pos[c] + len - 1, std::false_type());
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
  lookup_out = lookup;
  return true;
}

// ---- delivery of a walked tile --------------------------------------------------------------------------
// Document-start markers, then cooperative expansion of the warp's events (lane, sub-chunk, sequence order =
// output order) into final Aha::Hit records: one event per lane per round, kDeliverRounds rounds in flight so
// that the L2 latency of the event ring and of ref_rec[] is paid once per batch.
constexpr int kDeliverRounds = 4;

struct WalkedTile {
  uint32_t excl_lane;   // hits of lower lanes in the tile
  uint32_t hits0;       // hits of this lane's sub-chunk 0 (base of sub-chunk 1)
  uint32_t n_ev0, n_ev1;
  uint32_t n_mark0, n_mark1;
};

__device__ __noinline__ void deliver_events(const CharParams& P, unsigned int t, uint32_t Ttile, const WalkedTile w,
                                               const CharEvent* warp_ev, const int2* my_mark, int lane) {
  const ScanParams& p = P.f.sp;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const CharEvent* my_ev = warp_ev + lane * kEvSlots;
  const uint32_t n_ev = w.n_ev0 + w.n_ev1;
  uint32_t ev_incl = n_ev;
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t v = __shfl_up_sync(0xffffffffu, ev_incl, o);
    if (lane >= o) ev_incl += v;
  }
  const uint32_t ev_excl = ev_incl - n_ev;
  const uint32_t E = __shfl_sync(0xffffffffu, ev_incl, 31);
  // which event does this lane expand in round r of a batch; fetch it and what fires there
  auto fetch = [&](uint32_t base, CharEvent (&ev)[kDeliverRounds], RefRec (&rr)[kDeliverRounds]) {
#pragma unroll
    for (int r = 0; r < kDeliverRounds; ++r) {
      const uint32_t i = base + r * 32 + lane;
      ev[r] = CharEvent{0, 0};
      if (base + r * 32 < E) {  // warp-uniform
        uint32_t lo = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
          const uint32_t cand = lo + step;
          const uint32_t ex = __shfl_sync(0xffffffffu, ev_excl, cand & 31);
          if (cand < 32 && ex <= i) lo = cand;
        }
        const uint32_t ex = __shfl_sync(0xffffffffu, ev_excl, lo);
        const uint32_t n0 = __shfl_sync(0xffffffffu, w.n_ev0, lo);
        if (i < E) {
          const uint32_t idx = i - ex;
          ev[r] = warp_ev[lo * kEvSlots + (idx < n0 ? idx : kEvSlotsC + (idx - n0))];
        }
      }
    }
#pragma unroll
    for (int r = 0; r < kDeliverRounds; ++r) {
      rr[r] = RefRec{0, 0, 0, 0};
      if (base + r * 32 + lane < E) rr[r] = P.t.ref_rec[ev[r].ref];
    }
  };
  CharEvent ev[kDeliverRounds];
  RefRec rr[kDeliverRounds];
  const bool want = p.hits != nullptr && E > 0;
  if (want) fetch(0, ev, rr);  // in flight while the look-back runs
  const unsigned long long prefix = resolve_prefix(P.f, t, lane);
  if ((int64_t)t == P.f.n_tiles - 1 && lane == 0) *P.f.total_hits = prefix + Ttile;
#pragma unroll
  for (int c = 0; c < kChains; ++c) {
    const uint32_t n_mark = c ? w.n_mark1 : w.n_mark0;
    for (uint32_t k = 0; k < n_mark; ++k) {
      const int2 mk = my_mark[c * kMarkCapC + k];
      const int y = mk.y;  // events of this sub-chunk before the document start
      uint32_t before = 0;
      for (int i = 0; i < y; ++i) before += P.t.ref_rec[my_ev[c * kEvSlotsC + i].ref].n;
      p.doc_hit_off[mk.x] = (int64_t)(prefix + w.excl_lane + (c ? w.hits0 : 0u) + before);
    }
  }
  if (!want) return;
  uint32_t run = 0;
  for (uint32_t base = 0; base < E; base += 32 * kDeliverRounds) {
    if (base != 0) fetch(base, ev, rr);
#pragma unroll
    for (int r = 0; r < kDeliverRounds; ++r) {
      if (base + r * 32 < E) {  // warp-uniform
        uint32_t len_incl = rr[r].n;
        for (int o = 1; o < 32; o <<= 1) {
          uint32_t v = __shfl_up_sync(0xffffffffu, len_incl, o);
          if (lane >= o) len_incl += v;
        }
        if (base + r * 32 + lane < E)
          expand_event(p, rr[r], ev[r].end, chars, p.hits, (int64_t)prefix + run + (len_incl - rr[r].n), P.f.hit_cap);
        run += __shfl_sync(0xffffffffu, len_incl, 31);
      }
    }
  }
}

__global__ void __launch_bounds__(kFusedWarps * 32, 1) scan_char_kernel(const __grid_constant__ CharParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  const FusedParams& F = P.f;
  const ScanParams& p = F.sp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // shared layout: root u32[n_root] | val u32[m] | multi bits u32[n_mwords] | check u16[m_pad] | cmap u8[256] | per-warp {mbarrier, text}
  uint32_t* s_root = reinterpret_cast<uint32_t*>(smem);
  uint32_t* s_val = s_root + P.t.n_root;
  uint32_t* s_mbits = s_val + P.t.m;
  uint16_t* s_check = reinterpret_cast<uint16_t*>(s_mbits + P.t.n_mwords);
  for (uint32_t i = threadIdx.x; i < P.t.n_mwords; i += blockDim.x) s_mbits[i] = P.t.mbits_g[i];
  uint8_t* s_cmap = reinterpret_cast<uint8_t*>(s_check + P.t.m_pad);
  for (uint32_t i = threadIdx.x; i < P.t.n_root; i += blockDim.x) s_root[i] = P.t.root_g[i];
  for (uint32_t i = threadIdx.x; i < P.t.m; i += blockDim.x) s_val[i] = P.t.val_g[i];
  for (uint32_t i = threadIdx.x; i < P.t.m_pad; i += blockDim.x) s_check[i] = P.t.check_g[i];
  for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) s_cmap[i] = P.cmap_g[i];
  const uint32_t table_bytes = P.t.n_root * 4 + P.t.m * 4 + P.t.n_mwords * 4 + P.t.m_pad * 2 + 256;
  uint8_t* wbase = smem + ((table_bytes + 127) & ~127u) + (size_t)warp * (size_t)(F.buf_bytes + 128);
  uint64_t* bar = reinterpret_cast<uint64_t*>(wbase);
  uint8_t* sbuf = wbase + 128;
  if (lane == 0) mbar_init(bar, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();

  CharTablesView T;  // generic per-lane walk (tiles with document boundaries)
  T.root = s_root; T.check = s_check; T.val = s_val; T.gfail = P.t.gfail; T.m = P.t.m;
#pragma unroll
  for (int c = 0; c < 3; ++c) { T.lo[c] = P.t.lo[c]; T.n[c] = P.t.n[c]; T.off[c] = P.t.off[c]; }
  CharSmem S;        // fast walk
  S.root.base = pin_u32(smem_u32(s_root)); S.root3.base = pin_u32(smem_u32(s_root + P.t.off[1])); S.val.base = pin_u32(smem_u32(s_val));
  S.check.base = pin_u32(smem_u32(s_check)); S.gfail = P.t.gfail; S.mbits.base = pin_u32(smem_u32(s_mbits));
  S.m = P.t.m;
  S.lo2 = P.t.lo[0]; S.n2 = P.t.n[0]; S.off2 = P.t.off[0];
  S.lo3 = P.t.lo[1]; S.n3 = P.t.n[1]; S.off3 = P.t.off[1];
  S.lo4 = P.t.lo[2]; S.n4 = P.t.n[2]; S.off4 = P.t.off[2];
  S.neg_lo3_64 = 0u - (P.t.lo[1] >> 6);
  ByteGlobal fallback;
  fallback.cmap_g = P.cmap_g;
  fallback.cmap = s_cmap;

  const int64_t gwarp = (int64_t)blockIdx.x * kFusedWarps + warp;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  // warm-up: the state at a position is a suffix of at most max_key_len - 1 bytes, and it starts at a unit boundary,
  // so starting with the root at the first boundary inside that window reproduces it
  const int32_t warmc = (int32_t)p.warm;
  const int32_t sub = F.chunk / kChains;
  uint32_t parity = 0;
  int buf = 0;
  // the tile walked in the previous round, delivered after the next walk so that its look-back never waits
  bool pend = false;
  unsigned int pend_t = 0;
  uint32_t pend_T = 0;
  int pend_buf = 0;
  WalkedTile pw = {0, 0, 0, 0, 0, 0};
  // document that contained the previous tile (tiles come in increasing order)
  int64_t cd = 0, cdoc_start = 0, cdoc_end = -1;

  // takes the next ticket and starts staging its text: one TMA bulk copy (16-byte granules) plus a byte tail.
  // Called as soon as the previous tile no longer needs the buffer, so the copy overlaps that tile's delivery.
  unsigned int ticket_ahead = 0;  // lane 0: a ticket requested one tile early, so the atomic's latency is never waited for
  if (lane == 0) ticket_ahead = atomicAdd(F.ticket, 1u);
  auto next_tile = [&]() -> unsigned int {
    unsigned int nt = __shfl_sync(0xffffffffu, ticket_ahead, 0);
    if (lane == 0 && (int64_t)nt < F.n_tiles) ticket_ahead = atomicAdd(F.ticket, 1u);
    if ((int64_t)nt < F.n_tiles) {
      const int64_t tile_start = (int64_t)nt * F.tile;
      const int64_t tile_end = min(tile_start + (int64_t)F.tile, p.total);
      const int64_t org = max((int64_t)0, tile_start - (int64_t)F.w16);
      const int64_t want_end = min(p.total, tile_end + 16);
      const int64_t bulk_end = org + ((want_end - org) & ~(int64_t)15);
      if (lane == 0 && bulk_end > org) {
        mbar_expect_tx(bar, (uint32_t)(bulk_end - org));
        bulk_g2s(sbuf, p.text + org, (uint32_t)(bulk_end - org), bar);
      }
      for (int64_t i = bulk_end + lane; i < want_end; i += 32) sbuf[i - org] = p.text[i];
    }
    return nt;
  };
  unsigned int t = next_tile();
#ifdef AHA_B200_PHASES
  unsigned long long* phase_cyc = reinterpret_cast<unsigned long long*>(F.ticket + 4);
  long long ph_t0 = clock64();
#define AHA_PHASE(i) do { const long long _n = clock64(); if (lane == 0) atomicAdd(&phase_cyc[i], (unsigned long long)(_n - ph_t0)); ph_t0 = _n; } while (0)
#else
#define AHA_PHASE(i) do { } while (0)
#endif

  for (;;) {
    const bool have = (int64_t)t < F.n_tiles;
    bool bad = false, over = false;
    WalkedTile w = {0, 0, 0, 0, 0, 0};
    uint32_t Ttile = 0, lead0[kChains] = {0, 0};
    int64_t origin = 0, a = 0, b = 0;
    CharEvent* warp_ev = P.events + ((gwarp * 2 + buf) * 32) * kEvSlots;
    CharEvent* my_ev = pin_ptr(warp_ev + lane * kEvSlots);
    int2* my_mark = F.marks + ((gwarp * 2 + buf) * 32 + lane) * kMarkCap;
    if (have) {
      const int64_t tile_start = (int64_t)t * F.tile;
      const int64_t tile_end = min(tile_start + (int64_t)F.tile, p.total);
      origin = max((int64_t)0, tile_start - (int64_t)F.w16);
      const int64_t want_end = min(p.total, tile_end + 16);
      const int64_t bulk_end = origin + ((want_end - origin) & ~(int64_t)15);
      // which document is this?
      if (!(tile_start >= cdoc_start && tile_start < cdoc_end)) {
        cd = last_le(p.doc_off, p.n_docs, tile_start);
        cdoc_start = p.doc_off[cd];
        cdoc_end = p.doc_off[cd + 1];
      }
      const bool one_doc = cdoc_start <= tile_start && cdoc_end >= tile_end;  // the tile lies inside one document
      uint64_t tl = 0, dlb = 0;
      if (chars) {
        tl = F.tile_lead[t];
        if (one_doc) dlb = p.doc_lead_base[cd];
      }
      AHA_PHASE(1);
      if (bulk_end > origin) {
        mbar_wait(bar, parity);
        parity ^= 1u;
      }
      __syncwarp();
      AHA_PHASE(0);

      a = min(tile_start + (int64_t)lane * F.chunk, tile_end);
      b = min(a + (int64_t)F.chunk, tile_end);
      if (chars) {
        const int64_t a1 = min(a + sub, b);
        const uint32_t mine0 = count_leads_smem(sbuf, origin, a, a1);
        const uint32_t mine = mine0 + count_leads_smem(sbuf, origin, a1, b);
        uint32_t incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
          uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += v;
        }
        const uint64_t before_a = tl + (incl - mine);
        if (a < b) lead0[0] = (uint32_t)(before_a - (one_doc ? dlb : p.doc_lead_base[last_le(p.doc_off, p.n_docs, a)]));
        if (a1 < b) lead0[1] = (uint32_t)(before_a + mine0 - (one_doc ? dlb : p.doc_lead_base[last_le(p.doc_off, p.n_docs, a1)]));
      }

      AHA_PHASE(1);
      // ---- walk: events to the lane's ring slots, no global loads on the chain
      uint32_t n_ev[kChains] = {0, 0}, n_mark[kChains] = {0, 0}, n_hit[kChains] = {0, 0};
      bool lookup = false;
      auto on_doc = [&](int c, int64_t d, uint32_t nev) {
        if (n_mark[c] < (uint32_t)kMarkCapC) my_mark[c * kMarkCapC + n_mark[c]] = make_int2((int)d, (int)nev);
        ++n_mark[c];
      };
      const int32_t stage_end = (int32_t)(want_end - origin);
      const bool ok = chars ? char_tile_walk<true, kEvSlotsC, false>(S, p, sbuf, origin, a, b, sub, stage_end, warmc, one_doc, cd, cdoc_start, cdoc_end, lead0,
                                                   my_ev, n_ev, n_hit, lookup, on_doc)
                            : char_tile_walk<false, kEvSlotsC, false>(S, p, sbuf, origin, a, b, sub, stage_end, warmc, one_doc, cd, cdoc_start, cdoc_end, lead0,
                                                    my_ev, n_ev, n_hit, lookup, on_doc);
      AHA_PHASE(2);
      bad = __any_sync(0xffffffffu, !ok);
      over = __any_sync(0xffffffffu, n_ev[0] > (uint32_t)kEvCapC || n_ev[1] > (uint32_t)kEvCapC || n_mark[0] > (uint32_t)kMarkCapC ||
                                         n_mark[1] > (uint32_t)kMarkCapC);
      w.n_ev0 = n_ev[0]; w.n_ev1 = n_ev[1]; w.n_mark0 = n_mark[0]; w.n_mark1 = n_mark[1];
      __syncwarp();

      // ---- hits per lane, tile aggregate, published at once (nobody ever waits for a walk to be delivered)
      uint32_t n_hits;
      if (bad) {
        CountSink cs;
        fallback.walk(F, sbuf, origin, a, b, lead0[0], cs);
        n_hits = cs.n;
      } else if (over) {
        uint32_t c = 0;
        char_lane_walk(P, T, sbuf, origin, a, b, lead0[0], [&](int32_t, uint32_t ref) { c += P.t.ref_rec[ref].n; }, [&](int64_t) {});
        n_hits = c;
      } else {
        if (lookup) {  // some output list is longer than the 2-bit table can say
#pragma unroll
          for (int c = 0; c < kChains; ++c) {
            n_hit[c] = 0;
            for (uint32_t i = 0; i < n_ev[c]; ++i) n_hit[c] += P.t.ref_rec[my_ev[c * kEvSlotsC + i].ref].n;
          }
        }
        w.hits0 = n_hit[0];
        n_hits = n_hit[0] + n_hit[1];
      }
      uint32_t incl = n_hits;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      w.excl_lane = incl - n_hits;
      Ttile = __shfl_sync(0xffffffffu, incl, 31);
      publish_aggregate(F, t, Ttile, lane);
      AHA_PHASE(3);
    }

    if (have && (bad || over)) {
      // needs the staged text again: the previous tile goes out first, then this one, resolved right away
      if (pend) {
        const CharEvent* pev = P.events + ((gwarp * 2 + pend_buf) * 32) * kEvSlots;
        const int2* pmark = F.marks + ((gwarp * 2 + pend_buf) * 32 + lane) * kMarkCap;
        deliver_events(P, pend_t, pend_T, pw, pev, pmark, lane);
        pend = false;
      }
      const unsigned long long prefix = resolve_prefix(F, t, lane);
      if ((int64_t)t == F.n_tiles - 1 && lane == 0) *F.total_hits = prefix + Ttile;
      if (lane == 0) atomicAdd(F.n_direct, 1u);
      if (bad) {
        DirectSink ds;
        ds.out = p.hits; ds.doc_hit_off = p.doc_hit_off; ds.idx = (int64_t)(prefix + w.excl_lane); ds.cap = F.hit_cap;
        fallback.walk(F, sbuf, origin, a, b, lead0[0], ds);
      } else {
        int64_t idx = (int64_t)(prefix + w.excl_lane);
        char_lane_walk(
            P, T, sbuf, origin, a, b, lead0[0],
            [&](int32_t end, uint32_t ref) {
              const RefRec rr = P.t.ref_rec[ref];
              if (p.hits) expand_event(p, rr, end, chars, p.hits, idx, F.hit_cap);
              idx += rr.n;
            },
            [&](int64_t d) { p.doc_hit_off[d] = idx; });
      }
    }
    AHA_PHASE(6);
    // the text buffer is free: take the next ticket and start its copy before the (latency-bound) delivery
    unsigned int t_next = t;
    if (have) {
      __syncwarp();
      fence_proxy_async();  // reads of sbuf are done before the next bulk copy overwrites it
      t_next = next_tile();
    }
    AHA_PHASE(4);
    // ---- deliver the previous tile: its predecessors have had a whole walk of time to publish
    if (pend) {
      const CharEvent* pev = P.events + ((gwarp * 2 + pend_buf) * 32) * kEvSlots;
      const int2* pmark = F.marks + ((gwarp * 2 + pend_buf) * 32 + lane) * kMarkCap;
      deliver_events(P, pend_t, pend_T, pw, pev, pmark, lane);
      pend = false;
    }
    AHA_PHASE(5);
    if (!have) break;
    if (!(bad || over)) {
      pend = true;
      pend_t = t; pend_T = Ttile; pend_buf = buf;
      pw = w;
      buf ^= 1;
    }
    t = t_next;
  }
}

}  // namespace aha
