// scan_char_split.cuh -- the character-automaton scan as a pipeline of kernels, all launched back to back with no
// host round trip between them:
//
//   0. tile_doc_kernel     which document every tile starts in (one binary search per tile, off the walkers' path).
//   1. char_walk_kernel    persistent walkers (tables in shared memory, TMA-staged text, two interleaved sub-chunks per
//                          lane).  A tile's events go to its own slice of a global event store (4 bytes per event), its
//                          per-sub-chunk event counts to tile_meta[].  The walk reads nothing back from global memory:
//                          nothing is counted, nothing is ordered, nobody waits for anybody.
//   2. char_count_kernel   one warp per tile at full occupancy: hits of the tile's events (2-bit table, L1-resident)
//                          -> tile_hits[]; cub exclusive sum -> tile_prefix[] (position of every tile's first hit).
//                          (A decoupled look-back inside the expansion was measured instead of these two launches: with
//                          half a million small tiles and ten thousand warps in flight the look-back chains were three
//                          orders of magnitude slower than the scan.)
//   3. char_expand_kernel  one warp per tile at FULL occupancy: events -> ref_rec[] -> Aha::Hit records at their final
//                          position.  The L2 round trips that would stall a 16-warp walker CTA are hidden by thousands
//                          of warps.  Writes beyond the hit buffer's capacity are dropped; the total is read by the host
//                          once, at the end of the call, and only an overflow re-runs this kernel with a larger buffer.
//   4. char_direct_kernel  the few tiles the walk could not finish (malformed UTF-8, more events than a slice holds):
//                          byte automaton over the staged tile, now with known offsets.  Always launched; reads the
//                          length of its list from device memory.
//
// Document starts travel inside the event stream as marker events: no separate marker arrays, no marker overflow.
// Output order = (tile, lane, sub-chunk, sequence) = reference order.
#pragma once
#include "scan_char.cuh"

namespace aha {

#ifndef AHA_W_PREFETCH
#define AHA_W_PREFETCH 0
#endif
constexpr int kStoreSlotsC = 16;                      // event slots per sub-chunk in the store (more: direct pass)
constexpr int kStoreSlots = kStoreSlotsC * kChains;   // per lane
constexpr uint32_t kTileDirect = 0x80000000u;         // tile_hits[] flag: the tile is delivered by char_direct_kernel
static_assert(32 * AHA_B200_CHAR_CHUNK < (1 << 15), "an event carries a 15-bit offset relative to its tile");

// which document a tile starts in: looked up once per tile by tile_doc_kernel, so that the walkers (whose
// consecutive tiles are megabytes apart) load one record a tile ahead instead of running a binary search
struct alignas(16) TileDoc {
  int64_t doc_start;
  int32_t doc_len;   // documents are < 2^31 bytes
  int32_t doc;
};

struct SplitParams {
  CharParams c;
  const TileDoc* tile_doc;   // [n_tiles]
  uint32_t* store;           // [n_tiles][kStoreSlotsC][32 lanes][kChains]: slot-major, event i of every sub-chunk side by side
  uint32_t* tile_meta;       // [n_tiles][32]  events of sub-chunk 0 | events of sub-chunk 1 << 16; direct tiles: hits of the lane
  uint32_t* tile_hits;       // [n_tiles + 1] hits of the tile (char_count_kernel), | kTileDirect for a tile the walk counted itself
  const uint64_t* tile_prefix;  // [n_tiles + 1] exclusive prefix of the hits
  const uint32_t* mbits_g;   // 2 bits per ref: hits it expands to when 1..3, 0 = see ref_rec[ref].n
  const uint2* ref8_g;       // [n_ref8] ref_rec in 8 bytes (see char_expand_kernel)
  uint32_t n_ref8;
  const uint2* out8_g;       // [n_out] every output-list entry as {key id, key length << 16 | characters after the first}
  uint32_t* direct_list;     // [n_tiles] tiles for char_direct_kernel
  ScanStatus* status;        // total_hits, n_direct (= length of direct_list), bad_docs
};

// tile_hits[] as the scan's input: the direct flag stripped, widened to 64 bits
struct TileHitsToU64 {
  __host__ __device__ __forceinline__ uint64_t operator()(const uint32_t& x) const { return (uint64_t)(x & ~kTileDirect); }
};

__global__ void tile_doc_kernel(const int64_t* doc_off, int64_t n_docs, int64_t tile, int64_t n_tiles, TileDoc* out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  const int64_t d = last_le(doc_off, n_docs, t * tile);
  const int64_t ds = doc_off[d];
  out[t] = TileDoc{ds, (int32_t)(doc_off[d + 1] - ds), (int32_t)d};
}

// ---- 1. walk ---------------------------------------------------------------------------------------------
// kSep: byte offsets with a separator set; the walk drops the events the byte after the match vetoes (char_tile_walk).
template <bool kSep>
__global__ void __launch_bounds__(kFusedWarps * 32, 1) char_walk_kernel(const __grid_constant__ SplitParams Q) {
  extern __shared__ __align__(128) uint8_t smem[];
  const CharParams& P = Q.c;
  const FusedParams& F = P.f;
  const ScanParams& p = F.sp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // shared layout: val u32[m] | hit-count bits (AHA_W_COUNT) | root u16[n_root] | check u16[m_pad] | gfail prefix u16 |
  //                class map of the fallback walk | per-warp {mbarrier, text}
  uint32_t* s_val = reinterpret_cast<uint32_t*>(smem);
  const uint32_t n_mw = AHA_W_COUNT ? P.t.n_mwords : 0u;
  uint32_t* s_mbits = s_val + P.t.m;
  uint16_t* s_root = reinterpret_cast<uint16_t*>(s_mbits + n_mw);
  uint16_t* s_check = s_root + P.t.n_root;
  uint16_t* s_gfail = s_check + P.t.m_pad;
  uint8_t* s_cmap = reinterpret_cast<uint8_t*>(s_gfail + P.t.n_gfail_s);
  for (uint32_t i = threadIdx.x; i < n_mw; i += blockDim.x) s_mbits[i] = P.t.mbits_g[i];
  for (uint32_t i = threadIdx.x; i < P.t.n_root; i += blockDim.x) s_root[i] = P.t.root_g[i];
  for (uint32_t i = threadIdx.x; i < P.t.m; i += blockDim.x) s_val[i] = P.t.val_g[i];
  for (uint32_t i = threadIdx.x; i < P.t.m_pad; i += blockDim.x) s_check[i] = P.t.check_g[i];
  for (uint32_t i = threadIdx.x; i < P.t.n_gfail_s; i += blockDim.x) s_gfail[i] = P.t.gfail[i];
  for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) s_cmap[i] = P.cmap_g[i];
  const uint32_t table_bytes = P.t.m * 4 + n_mw * 4 + (P.t.n_root + P.t.m_pad + P.t.n_gfail_s) * 2 + 256;
  uint8_t* wbase = smem + ((table_bytes + 127) & ~127u) + (size_t)warp * (size_t)(F.buf_bytes + 128);
  uint64_t* bar = reinterpret_cast<uint64_t*>(wbase);
  uint8_t* sbuf = wbase + 128;
  if (lane == 0) mbar_init(bar, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();

  CharSmem S;
  S.root.base = pin_u32(smem_u32(s_root)); S.root3.base = pin_u32(smem_u32(s_root + P.t.off[1])); S.val.base = pin_u32(smem_u32(s_val));
  S.check.base = pin_u32(smem_u32(s_check)); S.gfail = P.t.gfail; S.mbits.base = pin_u32(smem_u32(s_mbits));
  S.gfail_s.base = pin_u32(smem_u32(s_gfail)); S.n_gfail_s = P.t.n_gfail_s;
  S.m = P.t.m;
  S.lo2 = P.t.lo[0]; S.n2 = P.t.n[0]; S.off2 = P.t.off[0];
  S.lo3 = P.t.lo[1]; S.n3 = P.t.n[1]; S.off3 = P.t.off[1];
  S.lo4 = P.t.lo[2]; S.n4 = P.t.n[2]; S.off4 = P.t.off[2];
  S.neg_lo3_64 = 0u - (P.t.lo[1] >> 6);
  ByteGlobal fallback;
  fallback.cmap_g = P.cmap_g;
  fallback.cmap = s_cmap;

  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const int32_t warmc = (int32_t)p.warm;
  const int32_t sub = F.chunk / kChains;
  uint32_t parity = 0;

  // Tiles are dealt round-robin (tile t -> warp t mod W): every tile costs about the same, nothing orders them, and
  // a warp knows its next tile early enough to have that tile's document record in registers before it needs it.
  const int64_t W = (int64_t)gridDim.x * kFusedWarps;
  // the tile after the next one is pulled into L2 while this one is walked, so that the bulk copy that stages it
  // finds it there
  auto prefetch_tile = [&](int64_t nt) {
#if AHA_W_PREFETCH
    if (nt < F.n_tiles && lane == 0) {
      const int64_t tile_start = nt * F.tile;
      const int64_t org = max((int64_t)0, tile_start - (int64_t)F.w16);
      const int64_t end = min(p.total, tile_start + (int64_t)F.tile + 16) & ~(int64_t)15;
      if (end > org) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.text + org), "r"((uint32_t)(end - org)) : "memory");
    }
#endif
  };
  auto stage_tile = [&](int64_t nt) {
    if (nt < F.n_tiles) {
      const int64_t tile_start = nt * F.tile;
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
  };
  int64_t t = (int64_t)blockIdx.x * kFusedWarps + warp;
  TileDoc td_next = TileDoc{0, 0, 0};
  if (t < F.n_tiles) td_next = Q.tile_doc[t];
  stage_tile(t);

  while (t < F.n_tiles) {
    const int64_t tile_start = t * F.tile;
    const int64_t tile_end = min(tile_start + (int64_t)F.tile, p.total);
    const int64_t origin = max((int64_t)0, tile_start - (int64_t)F.w16);
    const int64_t want_end = min(p.total, tile_end + 16);
    const int64_t bulk_end = origin + ((want_end - origin) & ~(int64_t)15);
    const int64_t cd = td_next.doc, cdoc_start = td_next.doc_start, cdoc_end = td_next.doc_start + td_next.doc_len;
    if (t + W < F.n_tiles) td_next = Q.tile_doc[t + W];  // used a whole tile later
    prefetch_tile(t + W);
    const bool one_doc = cdoc_start <= tile_start && cdoc_end >= tile_end;
    if (bulk_end > origin) {
      mbar_wait(bar, parity);
      parity ^= 1u;
    }
    __syncwarp();

    const int64_t a = min(tile_start + (int64_t)lane * F.chunk, tile_end);
    const int64_t b = min(a + (int64_t)F.chunk, tile_end);
    // character units: unit starts of the tile in front of each sub-chunk (the event offsets count from the tile start)
    uint32_t rel0[kChains] = {0, 0};
    uint32_t lead_doc = 0;  // direct pass only: characters of the lane's first document in front of the lane
    if (chars) {
      const int64_t a1 = min(a + sub, b);
      const uint32_t mine0 = count_leads_smem(sbuf, origin, a, a1);
      const uint32_t mine = mine0 + count_leads_smem(sbuf, origin, a1, b);
      uint32_t incl = mine;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      rel0[0] = incl - mine;
      rel0[1] = rel0[0] + mine0;
    }

    uint32_t* my_ev = pin_ptr(Q.store + (size_t)t * 32 * kStoreSlots + lane * kChains);
    uint32_t n_ev[kChains] = {0, 0};
    uint32_t n_hits = 0;
    bool lookup = false;
    auto on_doc = [&](int c, int64_t d, uint32_t nev) {  // marker event in the stream of sub-chunk c
      if (nev < (uint32_t)kStoreSlotsC * 64u) my_ev[nev + c] = kEvMark | (uint32_t)d;  // nev in steps of 64
      n_ev[c] += 64u;
    };
    const int32_t stage_end = (int32_t)(want_end - origin);
    const int32_t rel_base = 1 - (int32_t)(tile_start - origin);
    bool ok;
    if constexpr (kSep)
      ok = char_tile_walk<false, kStoreSlotsC, true>(S, p, sbuf, origin, a, b, sub, stage_end, warmc, one_doc, cd, cdoc_start, cdoc_end,
                                                     rel_base, rel0, my_ev, n_ev, n_hits, lookup, on_doc);
    else
      ok = chars ? char_tile_walk<true, kStoreSlotsC, false>(S, p, sbuf, origin, a, b, sub, stage_end, warmc, one_doc, cd, cdoc_start,
                                                             cdoc_end, rel_base, rel0, my_ev, n_ev, n_hits, lookup, on_doc)
                 : char_tile_walk<false, kStoreSlotsC, false>(S, p, sbuf, origin, a, b, sub, stage_end, warmc, one_doc, cd, cdoc_start,
                                                              cdoc_end, rel_base, rel0, my_ev, n_ev, n_hits, lookup, on_doc);
    n_ev[0] >>= 6;
    n_ev[1] >>= 6;
    const bool bad = __any_sync(0xffffffffu, !ok);
    const bool over = __any_sync(0xffffffffu, n_ev[0] > (uint32_t)kStoreSlotsC || n_ev[1] > (uint32_t)kStoreSlotsC);
    if (bad || over) {
      // the expansion cannot use this tile's events: count with the byte automaton, deliver in char_direct_kernel
      if (chars && a < b) {
        const int64_t d = one_doc ? cd : last_le(p.doc_off, p.n_docs, a);
        lead_doc = (uint32_t)(F.tile_lead[t] + rel0[0] - p.doc_lead_base[d]);
      }
      CountSink cs;
      fallback.walk(F, sbuf, origin, a, b, lead_doc, cs);
      n_hits = cs.n;
    }
    // the staged text is not needed any more: the next tile's bulk copy flies while this tile's counts are settled
    __syncwarp();
    fence_proxy_async();  // reads of sbuf are done before the bulk copy overwrites it
    stage_tile(t + W);
    if (AHA_W_COUNT && !(bad || over) && __any_sync(0xffffffffu, lookup)) {
      // some output list is longer than the 2-bit table can say (rare): this lane's events once more, through ref_rec
      n_hits = 0;
#pragma unroll
      for (int c = 0; c < kChains; ++c)
        for (uint32_t i = 0; i < n_ev[c]; ++i) {
          const uint32_t e = my_ev[i * 64 + c];
          if (!(e & kEvMark)) n_hits += P.t.ref_rec[e & 0xFFFFu].n;
        }
    }
    Q.tile_meta[(size_t)t * 32 + lane] = (bad || over) ? n_hits : (n_ev[0] | (n_ev[1] << 16));
    uint32_t T = n_hits;  // (without AHA_W_COUNT: zero unless the tile goes to the direct pass; char_count_kernel counts)
    for (int o = 16; o > 0; o >>= 1) T += __shfl_xor_sync(0xffffffffu, T, o);
    if (lane == 0) {
      Q.tile_hits[t] = T | ((bad || over) ? kTileDirect : 0u);
      if (bad || over) Q.direct_list[atomicAdd(&Q.status->n_direct, 1u)] = (uint32_t)t;
    }
    t += W;
  }
}

// ---- 2. / 3. count and expansion: one warp per tile ---------------------------------------------------------------
// An event's end offset = its tile-relative `rel` + the tile's base - the base of the document it lies in (bytes:
// tile start / document start; characters: unit starts in front of the tile / in front of the document).  The
// document of an event is the one of the last marker in front of it in the tile's stream, else the tile's first.
__device__ __forceinline__ uint32_t event_hits(const SplitParams& Q, uint32_t ref) {
  const uint32_t n = (__ldg(Q.mbits_g + (ref >> 4)) >> ((ref & 15u) * 2u)) & 3u;
  return n ? n : Q.c.t.ref_rec[ref].n;
}

// the tile's events in stream order (lane, sub-chunk, sequence), 32 per round: which slot of the store holds event i
struct TileEvents {
  const uint32_t* tile_ev;
  uint32_t n_ev0, ev_excl, E;
  __device__ __forceinline__ void init(const SplitParams& Q, int64_t t, int lane) {
    const uint32_t meta = Q.tile_meta[(size_t)t * 32 + lane];
    n_ev0 = meta & 0xFFFFu;
    const uint32_t n_ev = n_ev0 + (meta >> 16);
    uint32_t ev_incl = n_ev;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, ev_incl, o);
      if (lane >= o) ev_incl += v;
    }
    ev_excl = ev_incl - n_ev;
    E = __shfl_sync(0xffffffffu, ev_incl, 31);
    tile_ev = Q.store + (size_t)t * 32 * kStoreSlots;
  }
  // event i (all lanes call it together); a lane beyond the end gets a marker-like word: no hits, nothing to expand
  __device__ __forceinline__ uint32_t fetch(uint32_t i) const {
    uint32_t lo = 0;
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
      const uint32_t cand = lo + step;
      const uint32_t ex = __shfl_sync(0xffffffffu, ev_excl, cand & 31);
      if (cand < 32 && ex <= i) lo = cand;
    }
    const uint32_t ex = __shfl_sync(0xffffffffu, ev_excl, lo);
    const uint32_t n0 = __shfl_sync(0xffffffffu, n_ev0, lo);
    if (i >= E) return kEvMark;
    const uint32_t idx = i - ex;
    return idx < n0 ? tile_ev[idx * 64 + lo * kChains] : tile_ev[(idx - n0) * 64 + lo * kChains + 1];
  }
};

// Counting needs no order: every lane adds up the events of its own two sub-chunks (slot k of both is one aligned
// 8-byte word, and the 32 lanes' words are one 256-byte row of the store), with the 2-bit table in shared memory.
// The rows of a tile are independent loads: eight are issued before the first is looked at.
__global__ void __launch_bounds__(256) char_count_kernel(const __grid_constant__ SplitParams Q) {
  extern __shared__ uint32_t s_mbits[];
  const uint32_t n_mwords = Q.c.t.n_mwords;
  // the 2-bit table spread to one byte per ref (a lookup is then a byte load, not a word load, a shift and a mask)
  for (uint32_t i = threadIdx.x; i < n_mwords * 4; i += blockDim.x) {
    const uint32_t x = (Q.mbits_g[i >> 2] >> ((i & 3u) * 8u)) & 0xFFu;
    s_mbits[i] = (x & 3u) | ((x & 0xCu) << 6) | ((x & 0x30u) << 12) | ((x & 0xC0u) << 18);
  }
  __syncthreads();
  const uint8_t* const s_cnt = reinterpret_cast<const uint8_t*>(s_mbits);
  const int lane = threadIdx.x & 31;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  // hits of an event; 0 with a live event = more than three keys: looked up afterwards (rare)
  auto bits_of = [&](uint32_t ev) -> uint32_t { return s_cnt[ev & 0xFFFFu]; };
  // (Measured and dropped: the first rows of the next tile in flight across the tile loop, 0.35 against 0.30 ms.)
  int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  uint32_t meta = t < Q.c.f.n_tiles ? Q.tile_meta[(size_t)t * 32 + lane] : 0u;
  uint32_t th = t < Q.c.f.n_tiles ? Q.tile_hits[t] : 0u;
  for (; t < Q.c.f.n_tiles; t += n_warps) {
    const uint32_t n0 = meta & 0xFFFFu, n1 = meta >> 16;
    const bool direct = (th & kTileDirect) != 0;  // counted by the walk with the byte automaton
    const int64_t tn = t + n_warps;  // the next tile's counts are on their way while this one's rows are read
    if (tn < Q.c.f.n_tiles) {
      meta = Q.tile_meta[(size_t)tn * 32 + lane];
      th = Q.tile_hits[tn];
    }
    if (direct) continue;
    const uint2* mine = reinterpret_cast<const uint2*>(Q.store + (size_t)t * 32 * kStoreSlots) + lane;
    uint32_t part = 0;
    bool more = false;
    const uint32_t mx = max(n0, n1);
    for (uint32_t k0 = 0; k0 < mx; k0 += 4) {
      uint2 e[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) e[j] = (k0 + j < mx) ? mine[(size_t)(k0 + j) * 32] : make_uint2(kEvMark, kEvMark);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool l0 = k0 + j < n0 && !(e[j].x & kEvMark), l1 = k0 + j < n1 && !(e[j].y & kEvMark);
        const uint32_t c0 = l0 ? bits_of(e[j].x) : 4u, c1 = l1 ? bits_of(e[j].y) : 4u;
        part += (c0 & 3u) + (c1 & 3u);
        more = more || c0 == 0u || c1 == 0u;
      }
    }
    if (__any_sync(0xffffffffu, more)) {  // some output list is longer than the table can say: this lane's events through ref_rec
      part = 0;
      for (uint32_t k = 0; k < mx; ++k) {
        const uint2 ev = mine[(size_t)k * 32];
        if (k < n0 && !(ev.x & kEvMark)) part += Q.c.t.ref_rec[ev.x & 0xFFFFu].n;
        if (k < n1 && !(ev.y & kEvMark)) part += Q.c.t.ref_rec[ev.y & 0xFFFFu].n;
      }
    }
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) Q.tile_hits[t] = part;
  }
}

// ref_rec[] in 8 bytes: x = key id, y = key length << 16 | characters after the first   (one key fires: y != 0, bit 31 clear)
//                       x = first index into out_val, y = 0x80000000 | n                (n >= 2 keys fire)
// One warp per tile.  The lookup goes through L1 (the 8-byte table of C2 is 200 KB; this kernel uses no shared memory).
// Measured and dropped: the table in shared memory -- 8-byte entries, one CTA of 32 warps per SM: 0.83 ms; single-key
// entries packed into 4 bytes, two CTAs of 24 warps: 0.74 ms -- against 0.63 ms from L1 at 64 warps per SM: the kernel
// lives on occupancy.  (16-byte ref_rec entries through L1: 0.73 ms.)
//
// kSep: ACX#match(seq, sep) (ac.cr:321-340) with byte offsets.  An event's absolute end is tile start + rel, so both
// neighbour bytes are one global load away: the byte after the match vetoes the whole event (ac.cr:326), the byte in
// front of a key's start vetoes that key (ac.cr:333); neither test looks across a document boundary.  The hits of a
// tile then depend on the text, not only on the events: kCount = true is the same loop without the writes and takes the
// place of char_count_kernel (tile_hits[] for the prefix sum).
template <bool kSep, bool kCount>
__global__ void __launch_bounds__(256, 4) char_expand_kernel(const __grid_constant__ SplitParams Q) {
  const CharParams& P = Q.c;
  const FusedParams& F = P.f;
  const ScanParams& p = F.sp;
  const bool chars = !kSep && (p.mode & AHA_UNITS_CHAR) != 0;
  // kSep with character offsets (String overloads): the events carry BYTE positions (the vetoes need them); the hit's
  // character offsets are counted here -- unit starts of the tile in front of the event, from the text itself
  const bool sep_chars = kSep && !kCount && (p.mode & AHA_UNITS_CHAR) != 0;
  const int lane = threadIdx.x & 31;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < F.n_tiles; t += n_warps) {
    if (Q.tile_hits[t] & kTileDirect) continue;
    TileEvents te;
    te.init(Q, t, lane);
    if (te.E == 0) {
      if (kCount && lane == 0) Q.tile_hits[t] = 0u;
      continue;
    }
    const unsigned long long prefix = kCount ? 0ull : Q.tile_prefix[t];
    const TileDoc td = Q.tile_doc[t];
    int64_t doc_lo = td.doc_start, doc_hi = td.doc_start + td.doc_len;  // kSep: bounds of the current document (carried like delta)
    // bases as 64-bit differences: end = rel + (tile base - document base)
    const int64_t tile_base = chars ? (int64_t)F.tile_lead[t] : t * (int64_t)F.tile;
    int64_t delta = tile_base - (chars ? (int64_t)p.doc_lead_base[td.doc] : td.doc_start);  // warp-uniform, updated at markers
    // sep_chars: unit starts of the tile in front of each lane's chunk, and of the text in front of the current document
    uint32_t lead_excl = 0;
    int64_t doc_lead = 0;
    if (sep_chars) {
      const int64_t ca = min(t * (int64_t)F.tile + (int64_t)lane * F.chunk, p.total), cb = min(ca + (int64_t)F.chunk, p.total);
      const uint32_t mine = count_leads_global(p.text, ca, cb);
      uint32_t incl = mine;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      lead_excl = incl - mine;
      doc_lead = (int64_t)p.doc_lead_base[td.doc];
    }
    uint32_t run = 0;
    for (uint32_t base = 0; base < te.E; base += 32) {
      const uint32_t ev = te.fetch(base + lane);
      const bool live = base + lane < te.E;
      const bool mark = live && (ev & kEvMark);
      uint2 r8 = make_uint2(0u, 0u);
      if (live && !mark) r8 = Q.ref8_g[ev & 0xFFFFu];
      uint32_t n = (r8.y & 0x80000000u) ? (r8.y & 0x7FFFFFFFu) : (r8.y != 0u ? 1u : 0u);
      int64_t my_delta = delta, my_lo = doc_lo, my_hi = doc_hi, my_dlead = doc_lead;
      if (__any_sync(0xffffffffu, mark)) {
        // document of every event of this round: the last marker at or in front of it, else the carried one
        int64_t d_here = mark ? (int64_t)(ev & ~kEvMark) : -1;
        for (int o = 1; o < 32; o <<= 1) {
          const int64_t v = __shfl_up_sync(0xffffffffu, d_here, o);
          if (lane >= o && d_here < 0) d_here = v;
        }
        if (d_here >= 0) {
          my_delta = tile_base - (chars ? (int64_t)p.doc_lead_base[d_here] : p.doc_off[d_here]);
          if (kSep) {
            my_lo = p.doc_off[d_here];
            my_hi = p.doc_off[d_here + 1];
            if (sep_chars) my_dlead = (int64_t)p.doc_lead_base[d_here];
          }
        }
        delta = __shfl_sync(0xffffffffu, my_delta, 31);
        if (kSep) {
          doc_lo = __shfl_sync(0xffffffffu, my_lo, 31);
          doc_hi = __shfl_sync(0xffffffffu, my_hi, 31);
          if (sep_chars) doc_lead = __shfl_sync(0xffffffffu, my_dlead, 31);
        }
      }
      const int64_t abs_end = t * (int64_t)F.tile + (int64_t)((ev >> 16) & 0x7FFFu);  // kSep: first byte after the match
      if (kSep && n != 0) {
        if (abs_end < my_hi && sep_rejects(p, __ldg(p.text + abs_end))) {
          n = 0;
        } else if (!(r8.y & 0x80000000u)) {
          const int64_t start = abs_end - (int64_t)(r8.y >> 16);
          if (start > my_lo && sep_rejects(p, __ldg(p.text + start - 1))) n = 0;
        } else {
          uint32_t keep = 0;
          for (uint32_t j = r8.x; j < r8.x + n; ++j) {
            const int64_t start = abs_end - (int64_t)(Q.out8_g[j].y >> 16);
            keep += (start > my_lo && sep_rejects(p, __ldg(p.text + start - 1))) ? 0u : 1u;
          }
          n = keep;
        }
      }
      int32_t end_char = 0;
      if (sep_chars) {
        const uint32_t rel = (ev >> 16) & 0x7FFFu;                       // the match ends rel bytes behind the tile start
        const uint32_t j = rel ? min((rel - 1u) / (uint32_t)F.chunk, 31u) : 0u;  // the lane chunk that holds its last byte
        const uint32_t before = __shfl_sync(0xffffffffu, lead_excl, j);
        if (n != 0) {
          const int64_t cs = t * (int64_t)F.tile + (int64_t)j * F.chunk;
          end_char = (int32_t)((int64_t)F.tile_lead[t] - my_dlead + before + count_leads_global(p.text, cs, abs_end));
        }
      }
      uint32_t len_incl = n;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, len_incl, o);
        if (lane >= o) len_incl += v;
      }
      int64_t at = (int64_t)prefix + run + (len_incl - n);
      if (kCount) {
        // nothing is written: the tile's total is all that is wanted
      } else if (mark) {
        p.doc_hit_off[ev & ~kEvMark] = at;
      } else if (n != 0 && p.hits) {
        const int32_t end = (int32_t)((int64_t)((ev >> 16) & 0x7FFFu) + my_delta);
        if (!(r8.y & 0x80000000u)) {
          if (at < F.hit_cap)
            p.hits[at] = sep_chars ? aha_hit{end_char - 1 - (int32_t)(r8.y & 0xFFFFu), end_char, (int32_t)r8.x}
                                   : (chars ? aha_hit{end - 1 - (int32_t)(r8.y & 0xFFFFu), end, (int32_t)r8.x}
                                            : aha_hit{end - (int32_t)(r8.y >> 16), end, (int32_t)r8.x});
        } else if (kSep) {
          const uint32_t n_all = r8.y & 0x7FFFFFFFu;
          for (uint32_t j = r8.x; j < r8.x + n_all; ++j) {
            const uint2 o = Q.out8_g[j];
            const int32_t v = (int32_t)o.x, len = (int32_t)(o.y >> 16);
            const int64_t start = abs_end - len;
            if (start > my_lo && sep_rejects(p, __ldg(p.text + start - 1))) continue;
            if (at < F.hit_cap) p.hits[at] = sep_chars ? aha_hit{end_char - 1 - (int32_t)(o.y & 0xFFFFu), end_char, v} : aha_hit{end - len, end, v};
            ++at;
          }
        } else {
          for (uint32_t j = r8.x; j < r8.x + n; ++j, ++at) {
            const uint2 o = Q.out8_g[j];
            if (at < F.hit_cap)
              p.hits[at] = chars ? aha_hit{end - 1 - (int32_t)(o.y & 0xFFFFu), end, (int32_t)o.x} : aha_hit{end - (int32_t)(o.y >> 16), end, (int32_t)o.x};
          }
        }
      }
      run += __shfl_sync(0xffffffffu, len_incl, 31);
    }
    if (kCount && lane == 0) Q.tile_hits[t] = run;
  }
}

// ---- 4. tiles the walk could not finish: byte automaton over the re-staged tile, one warp per tile ------------------
__global__ void __launch_bounds__(128) char_direct_kernel(const __grid_constant__ SplitParams Q) {
  extern __shared__ __align__(128) uint8_t smem[];
  const CharParams& P = Q.c;
  const FusedParams& F = P.f;
  const ScanParams& p = F.sp;
  const unsigned int n_list = Q.status->n_direct;
  if (n_list == 0) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* s_cmap = smem;
  for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) s_cmap[i] = P.cmap_g[i];
  uint8_t* sbuf = smem + 256 + (size_t)warp * (size_t)F.buf_bytes;
  __syncthreads();
  ByteGlobal fallback;
  fallback.cmap_g = P.cmap_g;
  fallback.cmap = s_cmap;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const unsigned int n_warps = gridDim.x * (blockDim.x >> 5);
  for (unsigned int k = blockIdx.x * (blockDim.x >> 5) + warp; k < n_list; k += n_warps) {
    const unsigned int t = Q.direct_list[k];
    const int64_t tile_start = (int64_t)t * F.tile;
    const int64_t tile_end = min(tile_start + (int64_t)F.tile, p.total);
    const int64_t origin = max((int64_t)0, tile_start - (int64_t)F.w16);
    const int64_t want_end = min(p.total, tile_end + 16);
    __syncwarp();
    for (int64_t i = origin + lane; i < want_end; i += 32) sbuf[i - origin] = p.text[i];
    __syncwarp();
    const int64_t a = min(tile_start + (int64_t)lane * F.chunk, tile_end);
    const int64_t b = min(a + (int64_t)F.chunk, tile_end);
    uint32_t lead0 = 0;
    if (chars) {
      uint32_t mine = count_leads_smem(sbuf, origin, a, b);
      uint32_t incl = mine;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const uint64_t before_a = F.tile_lead[t] + (incl - mine);
      if (a < b) lead0 = (uint32_t)(before_a - p.doc_lead_base[last_le(p.doc_off, p.n_docs, a)]);
    }
    // hits of the lanes in front of this one (counted by the walk kernel with the same automaton)
    const uint32_t mine_hits = Q.tile_meta[(size_t)t * 32 + lane];
    uint32_t incl = mine_hits;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    DirectSink ds;
    ds.out = p.hits; ds.doc_hit_off = p.doc_hit_off; ds.idx = (int64_t)(Q.tile_prefix[t] + (incl - mine_hits)); ds.cap = F.hit_cap;
    fallback.walk(F, sbuf, origin, a, b, lead0, ds);
  }
}

// ---- small helpers of the pipeline ---------------------------------------------------------------------------------
// doc_off must start at 0, be non-decreasing, end at `total`, and no document may reach 2^31 bytes (Aha::Hit offsets
// are Int32, matcher.cr:3-5): counted into status->bad_docs, reported by the host at the end of the call
__global__ void validate_docs_kernel(const int64_t* doc_off, int64_t n_docs, int64_t total, ScanStatus* status) {
  const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d > n_docs) return;
  bool bad = false;
  if (d == 0 && doc_off[0] != 0) bad = true;
  if (d == n_docs && doc_off[d] != total) bad = true;
  if (d < n_docs) {
    const int64_t len = doc_off[d + 1] - doc_off[d];
    if (len < 0 || len > (int64_t)INT32_MAX) bad = true;
  }
  if (bad) atomicAdd(&status->bad_docs, 1u);
}

// documents that start at the very end of the text (trailing empty documents and the n_docs sentinel); publishes the
// total for the host
__global__ void finalize_split_kernel(const int64_t* doc_off, int64_t n_docs, int64_t total, const uint64_t* total_hits,
                                      int64_t* doc_hit_off, ScanStatus* status) {
  const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d == 0) status->total_hits = *total_hits;
  if (d <= n_docs && doc_off[d] >= total) doc_hit_off[d] = (int64_t)*total_hits;
}

}  // namespace aha
