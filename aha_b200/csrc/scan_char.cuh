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

constexpr int kEvCap = 32;  // events per lane per tile kept in the scratch ring

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
  uint32_t n_mwords;        // words of the multi bitmap
  const uint32_t* mbits_g;  // bit ref: more than one key fires there
  uint32_t lo[3], n[3], off[3];
};

struct CharParams {
  FusedParams f;            // f.scratch is unused here
  CharDeviceTables t;
  CharEvent* events;        // [n_warps_total][2][32][kEvCap]  (two tiles in flight per warp)
  const uint8_t* cmap_g;    // byte class map for the fallback walk
};

// counts only (fallback / overflow paths)
struct CountSink {
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
struct CharSmem {
  const uint32_t* root;    // [n_root]
  const uint32_t* root3;   // root + off[1]: entries of the 3-byte units
  const uint32_t* val;     // [m]
  const uint16_t* check;   // [m_pad]
  const uint16_t* gfail;   // global
  const uint32_t* mbits;   // [n_mwords] bit ref: ref_rec[ref].n > 1
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
  const uint32_t cid = r & 0x7FFFu;
  const uint32_t slot1 = s + cid, slot2 = f1 + cid;
  const uint32_t c1 = S.check[slot1], c2 = S.check[slot2];
  const bool hit1 = c1 == s, hit2 = c2 == f1;
  bool hit = hit1 || hit2;
  uint32_t slotv = hit1 ? slot1 : slot2;
  uint32_t v = ((r >> 16) & 0x7FFFu) | (r & 0x8000u);  // the root entry in val[] format (fail = 0)
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

// Walks one tile.  Every lane owns [a, b) of the concatenated text and runs the same per-document segment
// logic as char_lane_walk() (warm-up, owned region, document boundaries with markers), but as a state machine
// driven by a warp-uniform loop, so that whenever every busy lane has 12 well-formed bytes of one kind in
// front of it (four 3-byte units, or twelve ASCII characters) the warp takes a branch-free unrolled step.
// Lanes near a boundary (region end, document end, mixed unit lengths) make the warp fall back to one
// unit per iteration.  one_doc: the whole tile lies in document cd = [cdoc_start, cdoc_end).
// Events of the lane go to my_ev[0 .. kEvCap); n_ev counts all of them.  Returns false (warp-uniform) when
// the text is not structurally valid UTF-8: the caller rescans the tile with the byte automaton.
template <bool kChars, class OnDoc>
__device__ __forceinline__ bool char_tile_walk(const CharSmem& S, const ScanParams& p, const uint8_t* sbuf, int64_t origin, int64_t a,
                                               int64_t b, int32_t stage_end, int32_t warmc, bool one_doc, int64_t cd,
                                               int64_t cdoc_start, int64_t cdoc_end, uint32_t lead0, CharEvent* my_ev,
                                               uint32_t& n_ev_out, uint32_t& multi_mask_out, OnDoc on_doc) {
  constexpr unsigned kFull = 0xffffffffu;
  const int32_t rb = (int32_t)(b - origin);
  bool done = a >= b, ok = true;
  int64_t d = cd, doc_start = cdoc_start, doc_end = cdoc_end;
  if (!done && !one_doc) {
    d = last_le(p.doc_off, p.n_docs, a);
    doc_start = p.doc_off[d];
    doc_end = p.doc_off[d + 1];
  }
  uint32_t n_ev = 0, multi_mask = 0;  // bit i: event i of this lane reports more than one key
  if (!done && doc_start == a) {
    for (int64_t e = first_ge(p.doc_off, p.n_docs, a); e <= d; ++e) on_doc(e, n_ev);
  }
  // current segment: owned bytes [a_seg, seg_end) of the document that ends at de (all relative to sbuf)
  int32_t a_seg = (int32_t)(a - origin), pos = rb, de = 0, seg_end = 0, ebase = 0;
  uint32_t s = 0, f1 = 0, k = 0, lead = lead0;
  bool deep = false;
  auto enter_segment = [&]() {
    de = (int32_t)min(doc_end - origin, (int64_t)1 << 30);
    seg_end = min(rb, de);
    ebase = (int32_t)(origin - doc_start) + 1;
    int64_t w = origin + a_seg - warmc;
    const bool at_start = w < doc_start || origin + a_seg == doc_start;
    if (at_start) w = doc_start;
    pos = (int32_t)(w - origin);
    s = 0; f1 = 0; k = 0; deep = false;
    if (!at_start) {  // the warm-up may begin inside a unit
      int skipped = 0;
      while (pos < seg_end && (sbuf[pos] & 0xC0) == 0x80) {
        ++pos;
        if (++skipped > 3) { ok = false; break; }
      }
    }
  };
  if (!done) enter_segment();

  // consume one character whose root entry is r; cs / last = offsets of its first / last byte
  auto step = [&](uint32_t r, int32_t cs, int32_t last, auto vote) {
    uint32_t ref;
    const uint32_t v = char_advance<decltype(vote)::value>(S, r, s, f1, deep, ref);
    s = v & 0x7FFFu;
    f1 = (v >> 16) & 0x7FFFu;
    deep = (v >> 31) != 0u;
    if (kChars) k += (cs >= a_seg) ? 1u : 0u;
    const bool fire = (v & 0x8000u) && (uint32_t)(last - a_seg) < (uint32_t)(rb - a_seg);
    if (fire) {
      if (n_ev < (uint32_t)kEvCap) {
        my_ev[n_ev] = CharEvent{kChars ? (int32_t)(lead + k) : last + ebase, ref};
        multi_mask |= ((S.mbits[ref >> 5] >> (ref & 31u)) & 1u) << n_ev;
      }
      ++n_ev;
    }
  };

  for (;;) {
    if (__any_sync(kFull, !ok)) return false;
    if (__all_sync(kFull, done)) break;
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(sbuf + (pos & ~3));
    const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
    const uint32_t sh = ((uint32_t)pos & 3u) * 8u;
    const uint32_t A = __funnelshift_r(w0, w1, sh), B = __funnelshift_r(w1, w2, sh), C = __funnelshift_r(w2, w3, sh);
    const bool room = pos < seg_end && pos + 12 <= min(de, stage_end);
    const uint32_t bad3 = ((A ^ 0xE08080E0u) & 0xF0C0C0F0u) | ((B ^ 0x80E08080u) & 0xC0F0C0C0u) | ((C ^ 0x8080E080u) & 0xC0C0F0C0u);
    if (__all_sync(kFull, done || (room && bad3 == 0u))) {
      // four 3-byte units: payload = (b0 & 0xF) << 12 | (b1 & 0x3F) << 6 | (b2 & 0x3F), root index = payload - lo
      auto step3 = [&](uint32_t w, int32_t cs) {
        const uint32_t wm = w & 0x3F3F0Fu;
        const uint32_t h = __dp4a(wm, 0x00000140u, S.neg_lo3_64);
        const uint32_t rel = h * 64u + (wm >> 16);
        const uint32_t r = rel < S.n3 ? S.root3[rel] : 0u;
        step(r, cs, cs + 2, std::true_type());
      };
      step3(A, pos);
      step3(__byte_perm(A, B, 0x0543), pos + 3);
      step3(__byte_perm(B, C, 0x0432), pos + 6);
      step3(C >> 8, pos + 9);
      if (!done) pos += 12;
    } else if (__all_sync(kFull, done || (room && ((A | B | C) & 0x80808080u) == 0u))) {
      // twelve ASCII characters
      const uint32_t w[3] = {A, B, C};
#pragma unroll
      for (int i = 0; i < 12; ++i) {
        const uint32_t byte = (w[i >> 2] >> ((i & 3) * 8)) & 0xFFu;
        step(S.root[byte], pos + i, pos + i, std::true_type());
      }
      if (!done) pos += 12;
    } else if (!done) {
      if (pos >= seg_end) {
        // end of the segment: the region is finished, or the next document starts here
        const bool more = doc_end <= b && (doc_end < b || b == p.total);
        if (!more) {
          done = true;
        } else {
          const int64_t at = doc_end;
          ++d;
          while (d < p.n_docs) {
            on_doc(d, n_ev);
            if (p.doc_off[d + 1] > at) break;
            ++d;
          }
          if (d >= p.n_docs || at >= b) {
            done = true;
          } else {
            doc_start = p.doc_off[d];
            doc_end = p.doc_off[d + 1];
            a_seg = (int32_t)(doc_start - origin);
            lead = 0;
            enter_segment();
          }
        }
      } else {
        // one unit of any length
        const uint32_t b0 = A & 0xFFu;
        int32_t len = 1;
        uint32_t r = 0;
        if (b0 < 0x80u) {
          r = S.root[b0];
        } else if (b0 < 0xC0u || b0 >= 0xF8u) {
          ok = false;
        } else {
          len = b0 < 0xE0u ? 2 : (b0 < 0xF0u ? 3 : 4);
          const uint32_t b1 = (A >> 8) & 0xFFu, b2 = (A >> 16) & 0xFFu, b3 = A >> 24;
          if (pos + len > de) ok = false;  // cut by the end of the document
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
          if (pos + len - 1 >= rb) {
            done = true;  // the unit ends in the next lane's region (the document goes on beyond b)
          } else {
            step(r, pos, pos + len - 1, std::false_type());
            pos += len;
          }
        }
      }
      if (done) pos = rb;
    }
  }
  n_ev_out = n_ev;
  multi_mask_out = multi_mask;
  return true;
}

// ---- delivery of a walked tile --------------------------------------------------------------------------
// Document-start markers, then cooperative expansion of the warp's events (lane, sequence order = output
// order) into final Aha::Hit records: one event per lane per round, output tables read event-parallel.
__device__ __forceinline__ void deliver_events(const CharParams& P, unsigned long long prefix, uint32_t excl_lane, uint32_t n_ev,
                                               uint32_t n_mark, uint32_t multi_mask, const CharEvent* warp_ev, const int2* my_mark,
                                               int lane) {
  const ScanParams& p = P.f.sp;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const CharEvent* my_ev = warp_ev + lane * kEvCap;
  for (uint32_t k = 0; k < n_mark; ++k) {
    const int y = my_mark[k].y;  // events of this lane before the document start
    uint32_t before = (uint32_t)y;
    for (uint32_t mm = multi_mask & (y >= 32 ? 0xffffffffu : ((1u << y) - 1u)); mm; mm &= mm - 1)
      before += P.t.ref_rec[my_ev[__ffs(mm) - 1].ref].n - 1u;
    p.doc_hit_off[my_mark[k].x] = (int64_t)(prefix + excl_lane + before);
  }
  uint32_t ev_incl = n_ev;
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t v = __shfl_up_sync(0xffffffffu, ev_incl, o);
    if (lane >= o) ev_incl += v;
  }
  const uint32_t ev_excl = ev_incl - n_ev;
  const uint32_t E = __shfl_sync(0xffffffffu, ev_incl, 31);
  if (!p.hits || E == 0) return;
  uint32_t run = 0;
  for (uint32_t base = 0; base < E; base += 32) {
    const uint32_t i = base + lane;
    uint32_t lo = 0;
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
      const uint32_t cand = lo + step;
      const uint32_t ex = __shfl_sync(0xffffffffu, ev_excl, cand & 31);
      if (cand < 32 && ex <= i) lo = cand;
    }
    const uint32_t ex = __shfl_sync(0xffffffffu, ev_excl, lo);
    CharEvent ev{0, 0};
    RefRec rr{0, 0, 0, 0};
    if (i < E) {
      ev = warp_ev[lo * kEvCap + (i - ex)];
      rr = P.t.ref_rec[ev.ref];
    }
    uint32_t len_incl = rr.n;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, len_incl, o);
      if (lane >= o) len_incl += v;
    }
    if (i < E) expand_event(p, rr, ev.end, chars, p.hits, (int64_t)prefix + run + (len_incl - rr.n), P.f.hit_cap);
    run += __shfl_sync(0xffffffffu, len_incl, 31);
  }
}

__global__ void __launch_bounds__(kFusedWarps * 32, 1) scan_char_kernel(const CharParams P) {
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
  S.root = s_root; S.root3 = s_root + P.t.off[1]; S.val = s_val; S.check = s_check; S.gfail = P.t.gfail; S.mbits = s_mbits;
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
  const int32_t warmc = (int32_t)p.warm + 3;
  uint32_t parity = 0;
  int buf = 0;
  // the tile walked in the previous round, delivered after the next walk so that its look-back never waits
  bool pend = false;
  unsigned int pend_t = 0;
  uint32_t pend_excl = 0, pend_T = 0, pend_n_ev = 0, pend_n_mark = 0, pend_multi = 0;
  int pend_buf = 0;
  // document that contained the previous tile (tiles come in increasing order)
  int64_t cd = 0, cdoc_start = 0, cdoc_end = -1;

  for (;;) {
    unsigned int t = 0;
    if (lane == 0) t = atomicAdd(F.ticket, 1u);
    t = __shfl_sync(0xffffffffu, t, 0);
    const bool have = (int64_t)t < F.n_tiles;
    bool bad = false, over = false;
    uint32_t n_ev = 0, n_mark = 0, multi_mask = 0, excl_lane = 0, Ttile = 0, lead0 = 0;
    int64_t origin = 0, a = 0, b = 0;
    CharEvent* warp_ev = P.events + ((gwarp * 2 + buf) * 32) * kEvCap;
    CharEvent* my_ev = warp_ev + lane * kEvCap;
    int2* my_mark = F.marks + ((gwarp * 2 + buf) * 32 + lane) * kMarkCap;
    if (have) {
      const int64_t tile_start = (int64_t)t * F.tile;
      const int64_t tile_end = min(tile_start + (int64_t)F.tile, p.total);
      origin = max((int64_t)0, tile_start - (int64_t)F.w16);
      const int64_t want_end = min(p.total, tile_end + 16);
      const int64_t bulk_end = origin + ((want_end - origin) & ~(int64_t)15);
      if (lane == 0 && bulk_end > origin) {
        mbar_expect_tx(bar, (uint32_t)(bulk_end - origin));
        bulk_g2s(sbuf, p.text + origin, (uint32_t)(bulk_end - origin), bar);
      }
      for (int64_t i = bulk_end + lane; i < want_end; i += 32) sbuf[i - origin] = p.text[i];
      // which document is this? (overlaps the copy)
      if (!(tile_start >= cdoc_start && tile_start < cdoc_end)) {
        cd = last_le(p.doc_off, p.n_docs, tile_start);
        cdoc_start = p.doc_off[cd];
        cdoc_end = p.doc_off[cd + 1];
      }
      const bool fast = cdoc_start <= tile_start && cdoc_end >= tile_end;  // the tile lies inside one document
      uint64_t tl = 0, dlb = 0;
      if (chars) {
        tl = F.tile_lead[t];
        if (fast) dlb = p.doc_lead_base[cd];
      }
      if (bulk_end > origin) {
        mbar_wait(bar, parity);
        parity ^= 1u;
      }
      __syncwarp();

      a = min(tile_start + (int64_t)lane * F.chunk, tile_end);
      b = min(a + (int64_t)F.chunk, tile_end);
      if (chars) {
        uint32_t mine = count_leads_smem(sbuf, origin, a, b);
        uint32_t incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
          uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += v;
        }
        const uint64_t before_a = tl + (incl - mine);
        if (a < b) lead0 = (uint32_t)(before_a - (fast ? dlb : p.doc_lead_base[last_le(p.doc_off, p.n_docs, a)]));
      }

      // ---- walk: events to the lane's ring slot, no global loads on the chain
      auto on_doc = [&](int64_t d, uint32_t nev) {
        if (n_mark < (uint32_t)kMarkCap) my_mark[n_mark] = make_int2((int)d, (int)nev);
        ++n_mark;
      };
      const int32_t stage_end = (int32_t)(want_end - origin);
      const bool ok = chars ? char_tile_walk<true>(S, p, sbuf, origin, a, b, stage_end, warmc, fast, cd, cdoc_start, cdoc_end, lead0, my_ev,
                                                   n_ev, multi_mask, on_doc)
                            : char_tile_walk<false>(S, p, sbuf, origin, a, b, stage_end, warmc, fast, cd, cdoc_start, cdoc_end, lead0, my_ev,
                                                    n_ev, multi_mask, on_doc);
      bad = __any_sync(0xffffffffu, !ok);
      over = __any_sync(0xffffffffu, n_ev > (uint32_t)kEvCap || n_mark > (uint32_t)kMarkCap);
      __syncwarp();

      // ---- hits per lane, tile aggregate, published at once (nobody ever waits for a walk to be delivered)
      uint32_t n_hits;
      if (bad) {
        CountSink cs;
        fallback.walk(F, sbuf, origin, a, b, lead0, cs);
        n_hits = cs.n;
      } else if (over) {
        uint32_t c = 0;
        char_lane_walk(P, T, sbuf, origin, a, b, lead0, [&](int32_t, uint32_t ref) { c += P.t.ref_rec[ref].n; },
                       [&](int64_t) {});
        n_hits = c;
      } else {
        n_hits = n_ev;
        for (uint32_t mm = multi_mask; mm; mm &= mm - 1) n_hits += P.t.ref_rec[my_ev[__ffs(mm) - 1].ref].n - 1u;
      }
      uint32_t incl = n_hits;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      excl_lane = incl - n_hits;
      Ttile = __shfl_sync(0xffffffffu, incl, 31);
      publish_aggregate(F, t, Ttile, lane);
    }

    // ---- deliver the previous tile: its predecessors have had a whole walk of time to publish
    if (pend) {
      const unsigned long long prefix = resolve_prefix(F, pend_t, lane);
      if ((int64_t)pend_t == F.n_tiles - 1 && lane == 0) *F.total_hits = prefix + pend_T;
      const CharEvent* pev = P.events + ((gwarp * 2 + pend_buf) * 32) * kEvCap;
      const int2* pmark = F.marks + ((gwarp * 2 + pend_buf) * 32 + lane) * kMarkCap;
      deliver_events(P, prefix, pend_excl, pend_n_ev, pend_n_mark, pend_multi, pev, pmark, lane);
      pend = false;
    }
    if (!have) break;

    if (bad || over) {
      // needs the staged text again: resolved and delivered right away
      const unsigned long long prefix = resolve_prefix(F, t, lane);
      if ((int64_t)t == F.n_tiles - 1 && lane == 0) *F.total_hits = prefix + Ttile;
      if (lane == 0) atomicAdd(F.n_direct, 1u);
      if (bad) {
        DirectSink ds;
        ds.out = p.hits; ds.doc_hit_off = p.doc_hit_off; ds.idx = (int64_t)(prefix + excl_lane); ds.cap = F.hit_cap;
        fallback.walk(F, sbuf, origin, a, b, lead0, ds);
      } else {
        int64_t idx = (int64_t)(prefix + excl_lane);
        char_lane_walk(
            P, T, sbuf, origin, a, b, lead0,
            [&](int32_t end, uint32_t ref) {
              const RefRec rr = P.t.ref_rec[ref];
              if (p.hits) expand_event(p, rr, end, chars, p.hits, idx, F.hit_cap);
              idx += rr.n;
            },
            [&](int64_t d) { p.doc_hit_off[d] = idx; });
      }
    } else {
      pend = true;
      pend_t = t; pend_excl = excl_lane; pend_T = Ttile; pend_n_ev = n_ev; pend_n_mark = n_mark; pend_multi = multi_mask;
      pend_buf = buf;
      buf ^= 1;
    }
    __syncwarp();
    fence_proxy_async();  // reads of sbuf are done before the next bulk copy overwrites it
  }
}

}  // namespace aha
