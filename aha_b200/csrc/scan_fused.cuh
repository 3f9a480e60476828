// scan_fused.cuh -- single-pass fused scan kernel for ACX#match / match(sep) (ac.cr:176-192, 265-286, 321-340).
//
// Persistent grid, one CTA per SM, every WARP is an independent worker:
//   * takes the next tile (32 lanes x `chunk` bytes of the concatenated text) from a global ticket,
//   * stages it (plus the warm-up bytes in front of it) into its private shared-memory buffer with one
//     TMA bulk copy (cp.async.bulk + mbarrier), no block-wide barrier in steady state,
//   * each lane walks its own chunk: warm-up over the preceding max_key_len-1 bytes of the document,
//     then reports the hits whose last byte falls in [a, b) ("owned region": boundary hits are neither
//     missed nor duplicated),
//   * hits go to a lane-private slot of an L2-resident scratch ring; lane counts are prefix-summed in the
//     warp, the tile's global offset comes from a decoupled look-back over tile_state[], and the warp
//     copies its hits to their final position -- the stream is already in reference order
//     (document, end ascending, output-chain order), no sort needed,
//   * if a lane's scratch slot (or its document-start markers) overflow, the warp re-walks the tile from
//     shared memory with the now known offsets and writes directly ("direct" pass).
// The automaton policy decides how one symbol is consumed: ByteGlobal (dense byte-class table in
// global/L2, any dictionary) or CharTable (UTF-8 character automaton resident in shared memory).
#pragma once
#include "api_internal.cuh"
#include "scan_common.cuh"

namespace aha {

constexpr int kLaneCap = 32;      // scratch records per lane per tile (record mode: sep scans)
#ifndef AHA_BYTE_EVCAP
#define AHA_BYTE_EVCAP 64
#endif
constexpr int kByteEvCap = AHA_BYTE_EVCAP;  // events (end position, state) per lane per tile (event mode: everything else)
constexpr int kMarkCap = 4;       // document-start markers per lane per tile
constexpr int kMarkCapC = 2;      // ... per sub-chunk (scan_char.cuh walks two sub-chunks per lane)
#ifndef AHA_B200_CHAR_WARPS
#define AHA_B200_CHAR_WARPS 16
#endif
#ifndef AHA_B200_CHAR_CHUNK
#define AHA_B200_CHAR_CHUNK 126
#endif
constexpr int kFusedWarps = AHA_B200_CHAR_WARPS;   // warps per CTA of scan_char_kernel (register- and shared-memory-bound)
constexpr int kByteSpanDefault = 512;   // bytes per lane of the byte kernel's span mode (AHA_B200_BYTE_SPAN; 0 = staged tiles only); C3 GB/s: 256 -> 84.1, 512 -> 84.9, 768 -> 80.9, staged 73.3
constexpr int kByteChunkDefault = 132;  // bytes per lane of the byte kernel (AHA_B200_BYTE_CHUNK overrides; see fused_geometry)
#ifndef AHA_BYTE_WARPS
#define AHA_BYTE_WARPS 32
#endif
#ifndef AHA_SPAN_WARPS
#define AHA_SPAN_WARPS 32
#endif
constexpr int kSpanWarps = AHA_SPAN_WARPS;  // ... of its span mode (no shared memory: bounded by registers only)
constexpr int kByteWarps = AHA_BYTE_WARPS;  // warps per CTA of scan_fused_kernel<ByteGlobal>: latency-bound on L2, so as many as fit

struct FusedParams {
  ScanParams sp;
  int32_t chunk;         // bytes per lane; chunk/4 is odd so lanes hit distinct banks
  int32_t tile;          // 32 * chunk
  int32_t w16;           // staged bytes in front of the tile (multiple of 16, > warm)
  int32_t buf_bytes;     // per-warp text buffer = w16 + tile + 16
  int64_t n_tiles;
  unsigned long long* tile_state;  // [n_tiles] (hits of the tile << 2) | 1 once the tile has been walked
  // groups of 32 consecutive tiles (two-level look-back: 32 tiles, then 32 groups = 1024 tiles per step)
  unsigned long long* group_excl;  // [n_groups] (hits before the group << 2) | 2, written by the first tile of the group that resolves
  unsigned int* ticket;
  unsigned long long* total_hits;  // out: total number of hits
  int64_t hit_cap;                 // capacity of sp.hits (records); writes beyond are dropped
  aha_hit* scratch;                // [n_warps_total][32][kLaneCap]
  uint2* ev_scratch;               // [n_warps_total][32][kByteEvCap] {state, end} (event mode)
  int2* marks;                     // [n_warps_total][32][kMarkCap] (doc index low/high packed, seq)
  unsigned int* n_direct;          // diagnostics: tiles that took the direct pass
  const uint64_t* tile_lead;       // [n_tiles + 1] non-continuation bytes before each tile (AHA_UNITS_CHAR)
  int32_t span;                    // != 0: span mode of the byte kernel -- no staging, every lane reads its own `chunk` bytes from global memory
  int32_t ev_cap;                  // events per lane slice of ev_scratch
};

// ---- PTX helpers ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}\n" ::"r"(
          smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// TMA bulk copy global -> shared, completion signalled on an mbarrier (UBLKCP in SASS)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- sinks ---------------------------------------------------------------------------------------
// ScratchSink: first pass; lane-private records + markers, counts everything.
struct ScratchSink {
  static constexpr bool kEvents = false;
  aha_hit* rec;   // lane slot [kLaneCap]
  int2* mark;     // lane slot [kMarkCap]
  uint32_t n = 0, n_mark = 0;
  __device__ __forceinline__ void hit(int32_t start, int32_t end, int32_t value) {
    if (n < (uint32_t)kLaneCap) rec[n] = aha_hit{start, end, value};
    ++n;
  }
  __device__ __forceinline__ void doc_start(int64_t d) {
    if (n_mark < (uint32_t)kMarkCap) mark[n_mark] = make_int2((int)d, (int)n);
    ++n_mark;
  }
  __device__ __forceinline__ bool want_records() const { return true; }
};
// DirectSink: offsets known; writes final positions.
struct DirectSink {
  static constexpr bool kEvents = false;
  aha_hit* out;
  int64_t* doc_hit_off;
  int64_t idx, cap;
  __device__ __forceinline__ void hit(int32_t start, int32_t end, int32_t value) {
    if (out && idx < cap) out[idx] = aha_hit{start, end, value};
    ++idx;
  }
  __device__ __forceinline__ void doc_start(int64_t d) { doc_hit_off[d] = idx; }
  __device__ __forceinline__ bool want_records() const { return true; }
};

// EventSink: the walk only notes where an output state was entered ({state, end}); the lists are expanded after the
// walk by all lanes at once.  With the expansion inside the walk a lane that fires holds its 31 neighbours for
// four dependent loads, and at one hit per 5-10 bytes some lane fires at nearly every step.
struct EventSink {
  static constexpr bool kEvents = true;
  uint2* ev;      // lane slot [cap]
  int2* mark;     // lane slot [kMarkCap]: {document, events before it}
  uint32_t cap = kByteEvCap;
  uint32_t n = 0, n_mark = 0;
  __device__ __forceinline__ void event(uint32_t state, int32_t end) {
    if (n < cap) ev[n] = make_uint2(state, (uint32_t)end);
    ++n;
  }
  __device__ __forceinline__ void hit(int32_t, int32_t, int32_t) {}
  __device__ __forceinline__ void doc_start(int64_t d) {
    if (n_mark < (uint32_t)kMarkCap) mark[n_mark] = make_int2((int)d, (int)n);
    ++n_mark;
  }
};
// TallySink: counts hits only (a tile whose events did not fit is counted, then delivered by the direct pass)
struct TallySink {
  static constexpr bool kEvents = false;
  uint32_t n = 0;
  __device__ __forceinline__ void hit(int32_t, int32_t, int32_t) { ++n; }
  __device__ __forceinline__ void doc_start(int64_t) {}
};

// expands the output list of `state` for a match whose last byte is absolute position `pos`
template <class Sink>
__device__ __forceinline__ void emit_list(const ScanParams& p, Sink& sink, uint32_t state, int64_t pos, int64_t doc_start,
                                          int64_t doc_end, uint32_t end_char) {
  if ((p.mode & AHA_SEP) && sep_veto_after(p, pos + 1, doc_end)) return;
  const uint32_t j1 = p.out_off[state + 1];
  const int32_t end = (int32_t)(pos - doc_start + 1);
  for (uint32_t j = p.out_off[state]; j < j1; ++j) {
    const int32_t v = p.out_val[j];
    const int32_t start = end - (int32_t)p.key_len[v];
    if ((p.mode & AHA_SEP) && sep_veto_before(p, doc_start + start, doc_start)) continue;
    if (p.mode & AHA_UNITS_CHAR) sink.hit((int32_t)end_char - 1 - (int32_t)p.key_tail[v], (int32_t)end_char, v);
    else sink.hit(start, end, v);
  }
}

// ---- policy: the byte automaton in global memory / L2 -------------------------------------------------------------
// Small dictionaries (dense table kept, dfa.hpp): one 4-byte load per byte from the closure table.
// Large ones: the sparse rows.  A narrow state (at most three children: 99 % of the states) is one 16-byte row -- the goto
// edge if the row has it, else on to the fail link's row: the loop of ACX#match_ (ac.cr:179-190) itself.  A wide state
// (numbered first: id < n_wide) owns a dense closure row and always answers with one 4-byte entry.  The dense table of
// the 1 M-pattern dictionary is 3.1 GB (every lookup an HBM sector); its rows are 100 MB and stay in L2.
// Where the walk reads its text.  StagedText: the warp's tile in shared memory (bulk copy).  SpanText: the lane's own span
// straight from global memory, eight (or sixteen) bytes at a time into registers (positions only ever grow): no shared memory, so
// a lane's span can be a kilobyte long -- the warm-up in front of it is then 3 % of the walk instead of 23 % -- and the
// SM's memory is L1 for the table lookups.
struct StagedText {
  const uint8_t* sbuf;
  __device__ __forceinline__ uint32_t get(int32_t pos) { return sbuf[pos]; }
};
#ifndef AHA_SPAN_WIN8
#define AHA_SPAN_WIN8 1   // 1: an 8-byte window (two registers less, twice the text loads): C3 86.0 against 84.8 GB/s with 16 bytes
#endif
#if AHA_SPAN_WIN8
struct SpanText {
  const uint8_t* base;
  int32_t limit;
  uint2 w;
  int32_t wpos = -8;
  __device__ __forceinline__ uint32_t get(int32_t pos) {
    const int32_t blk = pos & ~7;
    if (blk != wpos) {
      wpos = blk;
      if (blk + 8 <= limit) {
        w = __ldg(reinterpret_cast<const uint2*>(base + blk));
      } else {
        uint32_t t[2] = {0u, 0u};
        for (int i = 0; i < 8 && blk + i < limit; ++i) t[i >> 2] |= (uint32_t)base[blk + i] << ((i & 3) * 8);
        w = make_uint2(t[0], t[1]);
      }
    }
    const uint32_t k = (uint32_t)pos & 7u;
    const uint32_t x = (k & 4u) ? w.y : w.x;
    return (x >> ((k & 3u) * 8u)) & 0xFFu;
  }
};
#else
struct SpanText {
  const uint8_t* base;   // text + buf_origin (16-byte aligned)
  int32_t limit;         // bytes of text from base on
  uint4 w;
  int32_t wpos = -16;
  __device__ __forceinline__ uint32_t get(int32_t pos) {
    const int32_t blk = pos & ~15;
    if (blk != wpos) {
      wpos = blk;
      if (blk + 16 <= limit) {
        w = __ldg(reinterpret_cast<const uint4*>(base + blk));
      } else {  // the last bytes of the text: never read past them
        uint32_t t[4] = {0u, 0u, 0u, 0u};
        for (int i = 0; i < 16 && blk + i < limit; ++i) t[i >> 2] |= (uint32_t)base[blk + i] << ((i & 3) * 8);
        w = make_uint4(t[0], t[1], t[2], t[3]);
      }
    }
    const uint32_t k = (uint32_t)pos & 15u;
    const uint32_t x = (k & 8u) ? ((k & 4u) ? w.w : w.z) : ((k & 4u) ? w.y : w.x);
    return (x >> ((k & 3u) * 8u)) & 0xFFu;
  }
};
#endif

struct ByteGlobal {
  static constexpr int kExtraWarm = 0;
  const uint8_t* cmap_g;  // class map in global memory
  const uint8_t* cmap;    // shared memory copy (set by load_tables)
  __host__ __device__ static uint32_t table_bytes(const ByteGlobal&) { return 256; }
  __device__ __forceinline__ void load_tables(uint8_t* tables, int tid, int nthreads, const ScanParams&) {
    for (int i = tid; i < 256; i += nthreads) tables[i] = cmap_g[i];
    cmap = tables;
  }
  // Walks lane chunk [a, b) of the staged text.  sbuf[i] is absolute byte (buf_origin + i).
  // ONE loop, ONE load per iteration, no branch inside it: an iteration either consumes the byte (goto edge found,
  // or the state is a wide one -- its row is a closure row and always answers) or moves to the fail link and looks at
  // the same byte again.  The lanes of a warp are at different bytes of their chunks: with a loop per byte inside a
  // loop over bytes every byte would cost the warp the LONGEST fail chain of its 32 lanes, and with a branch per kind of
  // row the load latencies of the kinds would add up instead of overlapping.
  //   kDense: the automaton has a dense closure table (small dictionaries): every iteration consumes.
  template <bool kDense, class Text, class Sink>
  __device__ __forceinline__ void walk_impl(const FusedParams& P, Text text, int64_t buf_origin, int64_t a, int64_t b,
                                            uint32_t lead0, Sink& sink) const {
    const ScanParams& p = P.sp;
    int64_t d = last_le(p.doc_off, p.n_docs, a);
    int64_t doc_start = p.doc_off[d], doc_end = p.doc_off[d + 1];
    const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
    uint32_t lead = lead0;  // characters of the current document up to (excluding) position a
    if (doc_start == a) {
      for (int64_t e = first_ge(p.doc_off, p.n_docs, a); e <= d; ++e) sink.doc_start(e);
    }
    const uint32_t C = p.C, nw = p.n_wide_first;
    const uint4* wide4 = reinterpret_cast<const uint4*>(p.wide);
    uint32_t s = 0;
    int64_t pos64 = a - (int64_t)p.warm;  // warm-up over the bytes in front of the owned region: states only
    if (pos64 < doc_start) pos64 = doc_start;
    // positions relative to the staged buffer from here on (32-bit)
    int32_t pos = (int32_t)(pos64 - buf_origin);
    const int32_t a_rel = (int32_t)(a - buf_origin), b_rel = (int32_t)(b - buf_origin);
    for (;;) {
      const int32_t seg_end = (int32_t)(min(b, doc_end) - buf_origin);
      const int32_t end_base = (int32_t)(buf_origin - doc_start) + 1;  // end offset in the document = pos + end_base
      while (pos < seg_end) {
        const uint32_t byte = text.get(pos);
        const uint32_t cls = cmap[byte];
        uint32_t e, fail = 0;
        bool consume;
        if (kDense) {
          const uint32_t t = __ldg(p.delta + (uint64_t)s * C + cls);
          e = ((t & kStateMask) << 9) | ((t & kOut) ? kEdgeOut : 0u);
          consume = true;
        } else {
          const bool is_wide = s < nw;
          const uint32_t widx = s * C + cls;  // (< 2^32: a wide-first automaton has well under 2^32 / C wide states)
          const uint4 v = __ldg(is_wide ? wide4 + (widx >> 2) : p.rows + s);
          const uint32_t sel = widx & 3u;
          const uint32_t ew = sel == 0 ? v.x : (sel == 1 ? v.y : (sel == 2 ? v.z : v.w));
          const uint32_t en = (v.y & 0xFFu) == byte ? v.y : ((v.z & 0xFFu) == byte ? v.z : ((v.w & 0xFFu) == byte ? v.w : 0u));
          e = is_wide ? ew : en;
          fail = v.x >> 8;
          consume = is_wide || e != 0u;
        }
        const bool own = consume && pos >= a_rel;
        if (chars) lead += (own && (byte & 0xC0u) != 0x80u) ? 1u : 0u;
        if (own && (e & kEdgeOut)) {
          if constexpr (Sink::kEvents) sink.event(e >> 9, chars ? (int32_t)lead : pos + end_base);
          else emit_list(p, sink, e >> 9, buf_origin + pos, doc_start, doc_end, lead);
        }
        s = consume ? (e >> 9) : fail;
        pos += consume ? 1 : 0;
      }
      if (!(pos == (int32_t)(doc_end - buf_origin) && (pos < b_rel || b == p.total))) break;
      // the next document starts here: reset the automaton
      const int64_t at = doc_end;
      s = 0;
      lead = 0;
      ++d;
      while (d < p.n_docs) {
        sink.doc_start(d);
        if (p.doc_off[d + 1] > at) break;
        ++d;
      }
      if (d >= p.n_docs) break;
      doc_start = p.doc_off[d];
      doc_end = p.doc_off[d + 1];
    }
  }
  template <bool kSpan = false, class Sink>
  __device__ __forceinline__ void walk(const FusedParams& P, const uint8_t* sbuf, int64_t buf_origin, int64_t a, int64_t b,
                                       uint32_t lead0, Sink& sink) const {
    if (a >= b) return;
    if (kSpan) {  // sbuf is not used: the lane reads its span from global memory
      SpanText t;
      t.base = P.sp.text + buf_origin;
      t.limit = (int32_t)min(P.sp.total - buf_origin, (int64_t)0x7FFFFFF0);
      if (P.sp.delta) walk_impl<true>(P, t, buf_origin, a, b, lead0, sink);
      else walk_impl<false>(P, t, buf_origin, a, b, lead0, sink);
      return;
    }
    StagedText t{sbuf};
    if (P.sp.delta) walk_impl<true>(P, t, buf_origin, a, b, lead0, sink);
    else walk_impl<false>(P, t, buf_origin, a, b, lead0, sink);
  }
};

// characters (non-continuation bytes) of the document containing `a` that lie in [doc_start, a): only the
// part inside this tile is counted here; the caller adds the per-tile prefix.
__device__ __forceinline__ uint32_t count_leads_smem(const uint8_t* sbuf, int64_t buf_origin, int64_t from, int64_t to) {
  uint32_t c = 0;
  int32_t i = (int32_t)(from - buf_origin);
  const int32_t e = (int32_t)(to - buf_origin);
  for (; i < e && (i & 3); ++i) c += ((sbuf[i] & 0xC0) != 0x80);
  for (; i + 4 <= e; i += 4) {  // a continuation byte is 10xxxxxx
    const uint32_t w = *reinterpret_cast<const uint32_t*>(sbuf + i);
    c += 4u - (uint32_t)__popc(w & 0x80808080u & ~(w << 1));
  }
  for (; i < e; ++i) c += ((sbuf[i] & 0xC0) != 0x80);
  return c;
}

// unit starts (non-continuation bytes) of text[from, to): byte head, aligned words, byte tail
__device__ __forceinline__ uint32_t count_leads_global(const uint8_t* text, int64_t from, int64_t to) {
  uint32_t c = 0;
  int64_t i = from;
  for (; i < to && (i & 3); ++i) c += ((__ldg(text + i) & 0xC0) != 0x80);
  for (; i + 4 <= to; i += 4) {
    const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(text + i));
    c += 4u - (uint32_t)__popc(w & 0x80808080u & ~(w << 1));
  }
  for (; i < to; ++i) c += ((__ldg(text + i) & 0xC0) != 0x80);
  return c;
}

// ---- two-level decoupled look-back -------------------------------------------------------------------
// A tile needs the number of hits in all earlier tiles.  Tickets are handed out in order, so every earlier
// tile is already running on some warp and publishes its count without waiting for anybody.
// publish_aggregate: called once per tile right after its walk: one store, nothing to wait for.
__device__ __forceinline__ void publish_aggregate(const FusedParams& P, unsigned int t, uint32_t T, int lane) {
  if (lane == 0) st_relaxed_u64(&P.tile_state[t], ((unsigned long long)T << 2) | 1ull);
}
// resolve_prefix: hits in all tiles before t.  Whole warp; waits (with back-off) only for earlier tiles that
// have not finished their WALK yet (never for another warp's delivery).  The caller normally defers this call
// by one tile so that it does not wait at all.  The group's exclusive prefix is computed by the first tile of the
// group that gets here (summing whole groups of 32 tile counts backwards until a published prefix is met) and
// published for the other 31: all warps polling the same few lines would serialise in L2.
__device__ __forceinline__ unsigned long long resolve_prefix(const FusedParams& P, unsigned int t, int lane) {
  const unsigned int g = t >> 5, j = t & 31u;
  // tiles of the same group in front of t (lanes < j); lane 31 peeks at the group's exclusive prefix
  unsigned long long v = 1ull;
  for (;;) {
    if ((unsigned)lane < j) v = ld_relaxed_u64(&P.tile_state[((size_t)g << 5) + lane]);
    else if (lane == 31) v = ld_relaxed_u64(&P.group_excl[g]);
    if (__all_sync(0xffffffffu, lane == 31 || (v & 3ull) != 0ull)) break;
    __nanosleep(64);
  }
  unsigned long long in_group = (unsigned)lane < j ? (v >> 2) : 0ull;
  for (int o = 16; o > 0; o >>= 1) in_group += __shfl_xor_sync(0xffffffffu, in_group, o);
  const unsigned long long e31 = __shfl_sync(0xffffffffu, v, 31);
  if ((e31 & 3ull) == 2ull) return (e31 >> 2) + in_group;
  unsigned long long G = 0;
  for (long long h = (long long)g - 1; h >= 0; --h) {
    unsigned long long e, a;
    for (;;) {
      a = ld_relaxed_u64(&P.tile_state[((size_t)h << 5) + lane]);  // groups in front of g are complete groups
      e = ld_relaxed_u64(&P.group_excl[h]);
      if (__all_sync(0xffffffffu, (a & 3ull) != 0ull)) break;
      __nanosleep(64);
    }
    unsigned long long part = a >> 2;
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    G += part;
    e = __shfl_sync(0xffffffffu, e, 0);
    if ((e & 3ull) == 2ull) {
      G += e >> 2;
      break;
    }
  }
  if (lane == 0) st_relaxed_u64(&P.group_excl[g], (G << 2) | 2ull);
  return G + in_group;
}
// walk-then-resolve in one go (kernels that do not defer)
__device__ __forceinline__ unsigned long long tile_exclusive_prefix(const FusedParams& P, unsigned int t, uint32_t T, int lane) {
  publish_aggregate(P, t, T, lane);
  return resolve_prefix(P, t, lane);
}

// -DAHA_B200_PHASES: cycles per phase of scan_fused_kernel summed over warps (printed by run_fused to stderr)
#ifdef AHA_B200_PHASES
__device__ unsigned long long g_phase_cycles[8];
#define AHA_PHASE_DECL long long ph_t = clock64(); unsigned long long ph_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define AHA_PHASE(i) { const long long ph_n = clock64(); ph_acc[i] += (unsigned long long)(ph_n - ph_t); ph_t = ph_n; }
#define AHA_PHASE_FLUSH if (lane == 0) { for (int ph_i = 0; ph_i < 8; ++ph_i) atomicAdd(&g_phase_cycles[ph_i], ph_acc[ph_i]); }
#else
#define AHA_PHASE_DECL
#define AHA_PHASE(i)
#define AHA_PHASE_FLUSH
#endif

template <class Policy, bool EV, bool kSpan = false>
__global__ void __launch_bounds__((kSpan ? kSpanWarps : kByteWarps) * 32, 1) scan_fused_kernel(const FusedParams P, const Policy policy_in) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ScanParams& p = P.sp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // shared layout: [policy tables][per-warp: mbarrier(16 B) + text buffer]
  uint8_t* tables = smem;
  const uint32_t table_bytes = Policy::table_bytes(policy_in);
  Policy policy = policy_in;
  policy.load_tables(tables, threadIdx.x, blockDim.x, p);
  uint8_t* wbase = smem + ((table_bytes + 127) & ~127u) + (size_t)warp * (size_t)(P.buf_bytes + 128);
  uint64_t* bar = reinterpret_cast<uint64_t*>(wbase);
  uint8_t* sbuf = wbase + 128;
  if (lane == 0) mbar_init(bar, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();

  const int64_t gwarp = (int64_t)blockIdx.x * (kSpan ? kSpanWarps : kByteWarps) + warp;
  aha_hit* my_rec = P.scratch + (gwarp * 32 + lane) * kLaneCap;
  int2* my_mark = P.marks + (gwarp * 32 + lane) * kMarkCap;
  uint32_t parity = 0;
  AHA_PHASE_DECL

  for (;;) {
    unsigned int t = 0;
    if (lane == 0) t = atomicAdd(P.ticket, 1u);
    t = __shfl_sync(0xffffffffu, t, 0);
    if ((int64_t)t >= P.n_tiles) break;
    const int64_t tile_start = (int64_t)t * P.tile;
    const int64_t tile_end = min(tile_start + (int64_t)P.tile, p.total);
    // ---- stage [origin, copy_end) : 16-byte aligned bulk copy + byte tail
    const int64_t origin = max((int64_t)0, tile_start - (int64_t)P.w16);
    const int64_t want_end = min(p.total, tile_end + 1);              // +1: right neighbour for the sep test
    const int64_t bulk_end = origin + ((want_end - origin) & ~(int64_t)15);
    if (!kSpan) {
      if (lane == 0) {
        if (bulk_end > origin) {
          mbar_expect_tx(bar, (uint32_t)(bulk_end - origin));
          bulk_g2s(sbuf, p.text + origin, (uint32_t)(bulk_end - origin), bar);
        }
      }
      for (int64_t i = bulk_end + lane; i < want_end; i += 32) sbuf[i - origin] = p.text[i];
      if (bulk_end > origin) {
        mbar_wait(bar, parity);
        parity ^= 1u;
      }
    }
    __syncwarp();
    AHA_PHASE(0)  // ticket + staging

    const int64_t a = min(tile_start + (int64_t)lane * P.chunk, tile_end);
    const int64_t b = min(a + (int64_t)P.chunk, tile_end);
    // characters before `a` inside the current document (AHA_UNITS_CHAR)
    uint32_t lead0 = 0;
    if (p.mode & AHA_UNITS_CHAR) {
      // per-lane count of its own chunk, exclusive warp scan, plus the tile prefix, minus the document base
      uint32_t mine = kSpan ? count_leads_global(p.text, a, b) : count_leads_smem(sbuf, origin, a, b);
      uint32_t incl = mine;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const uint64_t before_a = P.tile_lead[t] + (incl - mine);
      if (a < b) {
        const int64_t d = last_le(p.doc_off, p.n_docs, a);
        lead0 = (uint32_t)(before_a - p.doc_lead_base[d]);
      }
    }

    if constexpr (EV) {
      // ---- event mode: walk (one dependent table load per byte, nothing else), count, look-back, expand in place
      const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
      uint2* my_ev = P.ev_scratch + (gwarp * 32 + lane) * (int64_t)P.ev_cap;
      EventSink es;
      es.ev = my_ev;
      es.mark = my_mark;
      es.cap = (uint32_t)P.ev_cap;
      policy.template walk<kSpan>(P, sbuf, origin, a, b, lead0, es);
      __syncwarp();
      AHA_PHASE(1)  // walk
      const uint32_t ne = es.n, nm = es.n_mark;
      const bool any_over = __any_sync(0xffffffffu, ne > (uint32_t)P.ev_cap || nm > (uint32_t)kMarkCap);
      uint32_t n_lane = 0;
      if (!any_over) {
        for (uint32_t k0 = 0; k0 < ne; k0 += 4) {  // list lengths of four events at a time: independent loads
          uint32_t c[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            c[u] = 0;
            if (k0 + u < ne) {
              const uint32_t st = my_ev[k0 + u].x;
              c[u] = __ldg(p.out_off + st + 1) - __ldg(p.out_off + st);
            }
          }
          n_lane += (c[0] + c[1]) + (c[2] + c[3]);
        }
      } else {
        TallySink ts;
        policy.template walk<kSpan>(P, sbuf, origin, a, b, lead0, ts);
        n_lane = ts.n;
      }
      uint32_t incl = n_lane;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const uint32_t excl_lane = incl - n_lane;
      const uint32_t T = __shfl_sync(0xffffffffu, incl, 31);
      AHA_PHASE(2)  // count
      const unsigned long long prefix = tile_exclusive_prefix(P, t, T, lane);
      AHA_PHASE(3)  // look-back
      if ((int64_t)t == P.n_tiles - 1 && lane == 0) *P.total_hits = prefix + T;
      if (!any_over) {
        // every lane expands its own events to their final positions (a lane's hits are consecutive in the stream)
        int64_t idx = (int64_t)(prefix + excl_lane);
        uint32_t m = 0;
        aha_hit* const out = p.hits;
        for (uint32_t k0 = 0; k0 < ne; k0 += 4) {
          uint2 e[4];
          uint32_t j0[4], cnt[4], kl[4];
          int32_t v0[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) e[u] = (k0 + u < ne) ? my_ev[k0 + u] : make_uint2(0u, 0u);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            j0[u] = 0;
            cnt[u] = 0;
            if (k0 + u < ne) {
              j0[u] = __ldg(p.out_off + e[u].x);
              cnt[u] = __ldg(p.out_off + e[u].x + 1) - j0[u];
            }
          }
          if (out && p.out8) {  // one 8-byte record per list entry: key id and both lengths
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const uint2 o = cnt[u] ? __ldg(p.out8 + j0[u]) : make_uint2(0u, 0u);
              v0[u] = (int32_t)o.x;
              kl[u] = chars ? (o.y & 0xFFFFu) : (o.y >> 16);
            }
          } else if (out) {
#pragma unroll
            for (int u = 0; u < 4; ++u) v0[u] = cnt[u] ? __ldg(p.out_val + j0[u]) : 0;
#pragma unroll
            for (int u = 0; u < 4; ++u) kl[u] = cnt[u] ? (chars ? __ldg(p.key_tail + v0[u]) : __ldg(p.key_len + v0[u])) : 0u;
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (k0 + u < ne) {
              while (m < nm && (uint32_t)my_mark[m].y == k0 + u) p.doc_hit_off[my_mark[m++].x] = idx;
              if (out && cnt[u]) {
                const int32_t end = (int32_t)e[u].y;
                if (idx < P.hit_cap) out[idx] = chars ? aha_hit{end - 1 - (int32_t)kl[u], end, v0[u]} : aha_hit{end - (int32_t)kl[u], end, v0[u]};
                for (uint32_t j = 1; j < cnt[u]; ++j) {
                  int32_t v, l;
                  if (p.out8) {
                    const uint2 o = __ldg(p.out8 + j0[u] + j);
                    v = (int32_t)o.x;
                    l = (int32_t)(chars ? (o.y & 0xFFFFu) : (o.y >> 16));
                  } else {
                    v = __ldg(p.out_val + j0[u] + j);
                    l = (int32_t)(chars ? __ldg(p.key_tail + v) : __ldg(p.key_len + v));
                  }
                  if (idx + j < P.hit_cap) out[idx + j] = chars ? aha_hit{end - 1 - l, end, v} : aha_hit{end - l, end, v};
                }
              }
              idx += cnt[u];
            }
          }
        }
        while (m < nm) p.doc_hit_off[my_mark[m++].x] = idx;
      } else {
        if (lane == 0) atomicAdd(P.n_direct, 1u);
        DirectSink ds;
        ds.out = p.hits;
        ds.doc_hit_off = p.doc_hit_off;
        ds.idx = (int64_t)(prefix + excl_lane);
        ds.cap = P.hit_cap;
        policy.template walk<kSpan>(P, sbuf, origin, a, b, lead0, ds);
      }
    } else {
      ScratchSink ss;
      ss.rec = my_rec;
      ss.mark = my_mark;
      policy.template walk<kSpan>(P, sbuf, origin, a, b, lead0, ss);

      // ---- warp prefix of lane counts, tile aggregate
      uint32_t incl = ss.n;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const uint32_t excl_lane = incl - ss.n;
      const uint32_t T = __shfl_sync(0xffffffffu, incl, 31);
      // ---- decoupled look-back
      const unsigned long long prefix = tile_exclusive_prefix(P, t, T, lane);
      if ((int64_t)t == P.n_tiles - 1 && lane == 0) *P.total_hits = prefix + T;

      // ---- deliver
      const bool over = (ss.n > (uint32_t)kLaneCap) || (ss.n_mark > (uint32_t)kMarkCap);
      const bool any_over = __any_sync(0xffffffffu, over);
      if (!any_over) {
        // markers first (they only need the lane base)
        for (uint32_t k = 0; k < ss.n_mark; ++k) p.doc_hit_off[my_mark[k].x] = (int64_t)(prefix + excl_lane + (uint32_t)my_mark[k].y);
        if (p.hits && T > 0) {
          // cooperative copy: output index i -> (lane l, record k) by binary search over the lane prefix
          __syncwarp();
          for (uint32_t i = lane; i < ((T + 31u) & ~31u); i += 32) {
            uint32_t lo = 0;
  #pragma unroll
            for (int step = 16; step > 0; step >>= 1) {
              const uint32_t cand = lo + step;
              const uint32_t ex = __shfl_sync(0xffffffffu, excl_lane, cand & 31);
              if (cand < 32 && ex <= i) lo = cand;
            }
            const uint32_t ex = __shfl_sync(0xffffffffu, excl_lane, lo);
            if (i < T) {
              const int64_t dst = (int64_t)prefix + i;
              if (dst < P.hit_cap) p.hits[dst] = P.scratch[(gwarp * 32 + lo) * kLaneCap + (i - ex)];
            }
          }
        }
      } else {
        if (lane == 0) atomicAdd(P.n_direct, 1u);
        DirectSink ds;
        ds.out = p.hits;
        ds.doc_hit_off = p.doc_hit_off;
        ds.idx = (int64_t)(prefix + excl_lane);
        ds.cap = P.hit_cap;
        policy.template walk<kSpan>(P, sbuf, origin, a, b, lead0, ds);
      }
    }
    __syncwarp();
    fence_proxy_async();  // reads of sbuf are done before the next bulk copy overwrites it
    AHA_PHASE(4)  // expansion / delivery
  }
  AHA_PHASE_FLUSH
}

// documents that start at the very end of the text (trailing empty documents and the n_docs sentinel)
__global__ void finalize_doc_off_kernel(const int64_t* doc_off, int64_t n_docs, int64_t total,
                                        const unsigned long long* total_hits, int64_t* doc_hit_off) {
  const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d <= n_docs && doc_off[d] >= total) doc_hit_off[d] = (int64_t)*total_hits;
}

}  // namespace aha
