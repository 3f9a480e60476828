// scan_char_split.cuh -- the character-automaton scan as THREE kernels instead of one fused kernel:
//
//   1. char_walk_kernel    persistent walkers (same tables, staging and interleaved walk as scan_char.cuh).  A tile's
//                          events go to its own slice of a global event store, its per-sub-chunk counts to tile_meta[],
//                          its hit total to tile_hits[].  Nothing is ordered, nobody waits for anybody.
//   2. cub exclusive scan  over tile_hits[] -> tile_prefix[] (global position of every tile's first hit).
//   3. char_expand_kernel  one warp per tile at FULL occupancy: events -> ref_rec[] -> Aha::Hit records at their final
//                          position.  The L2 round trips that stall a 16-warp walker CTA are hidden by thousands of warps.
//   4. char_direct_kernel  the few tiles the walk could not finish (malformed UTF-8, more events than a slice holds):
//                          byte automaton over the staged tile, now with known offsets.
//
// Document starts travel inside the event stream as marker events (ref = kMarkRef, end = document index): no separate
// marker arrays, no marker overflow.  Output order = (tile, lane, sub-chunk, sequence) = reference order.
#pragma once
#include "scan_char.cuh"

namespace aha {

constexpr int kStoreSlotsC = 16;                      // event slots per sub-chunk in the store (more: direct pass)
constexpr int kStoreSlots = kStoreSlotsC * kChains;   // per lane
#ifndef AHA_B200_EXPAND_ROUNDS
#define AHA_B200_EXPAND_ROUNDS 1
#endif
constexpr int kExpandRounds = AHA_B200_EXPAND_ROUNDS;  // rounds of 32 events a warp of char_expand_kernel keeps in flight
                                                       // (measured on C2: 1 -> 0.91 ms, 4 -> 1.03 ms: occupancy wins)
constexpr uint32_t kMarkRef = 0xFFFFFFFFu;            // marker event: `end` holds a document index
constexpr uint32_t kTileDirect = 0x80000000u;         // tile_hits[] flag: the tile is delivered by char_direct_kernel

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
  CharEvent* store;          // [n_tiles][kStoreSlotsC][32 lanes][kChains]: slot-major, event i of every sub-chunk side by side
  uint32_t* tile_meta;       // [n_tiles][32][kChains]  hits:16 | events:16 of a sub-chunk
  uint32_t* tile_hits;       // [n_tiles + 1] hits of the tile | kTileDirect
  const uint64_t* tile_prefix;  // [n_tiles + 1] exclusive prefix of the hits (kernels 3 and 4)
  uint32_t* direct_list;     // [n_tiles] tiles for char_direct_kernel
  unsigned int* n_direct_list;
};

__global__ void tile_doc_kernel(const int64_t* doc_off, int64_t n_docs, int64_t tile, int64_t n_tiles, TileDoc* out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  const int64_t d = last_le(doc_off, n_docs, t * tile);
  const int64_t ds = doc_off[d];
  out[t] = TileDoc{ds, (int32_t)(doc_off[d + 1] - ds), (int32_t)d};
}

__global__ void tile_hits_to_u64_kernel(const uint32_t* tile_hits, int64_t n, uint64_t* out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = tile_hits[i] & ~kTileDirect;
}

// ---- 1. walk ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFusedWarps * 32, 1) char_walk_kernel(const __grid_constant__ SplitParams Q) {
  extern __shared__ __align__(128) uint8_t smem[];
  const CharParams& P = Q.c;
  const FusedParams& F = P.f;
  const ScanParams& p = F.sp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t* s_root = reinterpret_cast<uint32_t*>(smem);
  uint32_t* s_val = s_root + P.t.n_root;
  uint32_t* s_mbits = s_val + P.t.m;
  uint16_t* s_check = reinterpret_cast<uint16_t*>(s_mbits + P.t.n_mwords);
  uint8_t* s_cmap = reinterpret_cast<uint8_t*>(s_check + P.t.m_pad);
  for (uint32_t i = threadIdx.x; i < P.t.n_mwords; i += blockDim.x) s_mbits[i] = P.t.mbits_g[i];
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

  CharSmem S;
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

  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const int32_t warmc = (int32_t)p.warm;
  const int32_t sub = F.chunk / kChains;
  uint32_t parity = 0;

  // Tiles are dealt round-robin (tile t -> warp t mod W): every tile costs about the same, nothing orders them, and
  // a warp knows its next tile early enough to have that tile's document record in registers before it needs it.
  const int64_t W = (int64_t)gridDim.x * kFusedWarps;
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
    const bool one_doc = cdoc_start <= tile_start && cdoc_end >= tile_end;
    uint64_t tl = 0, dlb = 0;
    if (chars) {
      tl = F.tile_lead[t];
      if (one_doc) dlb = p.doc_lead_base[cd];
    }
    if (bulk_end > origin) {
      mbar_wait(bar, parity);
      parity ^= 1u;
    }
    __syncwarp();

    const int64_t a = min(tile_start + (int64_t)lane * F.chunk, tile_end);
    const int64_t b = min(a + (int64_t)F.chunk, tile_end);
    uint32_t lead0[kChains] = {0, 0};
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

    CharEvent* my_ev = pin_ptr(Q.store + (size_t)t * 32 * kStoreSlots + lane * kChains);
    uint32_t n_ev[kChains] = {0, 0}, n_hit[kChains] = {0, 0};
    bool lookup = false;
    auto on_doc = [&](int c, int64_t d, uint32_t nev) {  // marker event in the stream of sub-chunk c
      if (nev < (uint32_t)kStoreSlotsC * 64u) my_ev[nev + c] = CharEvent{(int32_t)d, kMarkRef};  // nev in steps of 64 (kLean)
      n_ev[c] += 64u;
    };
    const int32_t stage_end = (int32_t)(want_end - origin);
    const bool ok = chars ? char_tile_walk<true, kStoreSlotsC, true, true>(S, p, sbuf, origin, a, b, sub, stage_end, warmc, one_doc, cd, cdoc_start,
                                                                     cdoc_end, lead0, my_ev, n_ev, n_hit, lookup, on_doc)
                          : char_tile_walk<false, kStoreSlotsC, true, true>(S, p, sbuf, origin, a, b, sub, stage_end, warmc, one_doc, cd, cdoc_start,
                                                                      cdoc_end, lead0, my_ev, n_ev, n_hit, lookup, on_doc);
    n_ev[0] >>= 6;
    n_ev[1] >>= 6;
    const bool bad = __any_sync(0xffffffffu, !ok);
    const bool over = __any_sync(0xffffffffu, n_ev[0] > (uint32_t)kStoreSlotsC || n_ev[1] > (uint32_t)kStoreSlotsC);
    uint32_t n_hits = 0;
    if (bad || over) {
      // the expansion cannot use this tile's events: count with the byte automaton, deliver in char_direct_kernel
      CountSink cs;
      fallback.walk(F, sbuf, origin, a, b, lead0[0], cs);
      n_hits = cs.n;
      n_hit[0] = cs.n;
      n_hit[1] = 0;
    }
    // the staged text is not needed any more: the next tile's bulk copy flies while this tile's counts are settled
    __syncwarp();
    fence_proxy_async();  // reads of sbuf are done before the bulk copy overwrites it
    stage_tile(t + W);
    if (!(bad || over)) {
      // hits per sub-chunk: the walk only stored events; the lane reads its own back (both sub-chunks of a slot share
      // 16 bytes) and looks the hit counts up in the 2-bit table -- a few batched loads per tile instead of a table
      // lookup inside the walk at every event
      const uint32_t mx = max(n_ev[0], n_ev[1]);
      for (uint32_t i0 = 0; i0 < mx; i0 += 4) {
        uint4 e[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (i0 + j < mx) e[j] = *reinterpret_cast<const uint4*>(my_ev + (size_t)(i0 + j) * 64);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (i0 + j < n_ev[0] && e[j].y != kMarkRef) {
            const uint32_t cnt = (S.mbits[e[j].y >> 4] >> ((e[j].y & 15u) * 2u)) & 3u;
            n_hit[0] += cnt;
            lookup = lookup || cnt == 0u;
          }
          if (i0 + j < n_ev[1] && e[j].w != kMarkRef) {
            const uint32_t cnt = (S.mbits[e[j].w >> 4] >> ((e[j].w & 15u) * 2u)) & 3u;
            n_hit[1] += cnt;
            lookup = lookup || cnt == 0u;
          }
        }
      }
      if (lookup) {  // some output list is longer than the 2-bit table can say
#pragma unroll
        for (int c = 0; c < kChains; ++c) {
          n_hit[c] = 0;
          for (uint32_t i = 0; i < n_ev[c]; ++i) {
            const uint32_t ref = my_ev[i * 64 + c].ref;
            if (ref != kMarkRef) n_hit[c] += P.t.ref_rec[ref].n;
          }
        }
      }
      n_hits = n_hit[0] + n_hit[1];
    }
    uint2 meta;
    meta.x = (n_hit[0] << 16) | min(n_ev[0], 0xFFFFu);
    meta.y = (n_hit[1] << 16) | min(n_ev[1], 0xFFFFu);
    *reinterpret_cast<uint2*>(Q.tile_meta + ((size_t)t * 32 + lane) * kChains) = meta;
    uint32_t T = n_hits;
    for (int o = 16; o > 0; o >>= 1) T += __shfl_xor_sync(0xffffffffu, T, o);
    if (lane == 0) {
      Q.tile_hits[t] = T | ((bad || over) ? kTileDirect : 0u);
      if (bad || over) Q.direct_list[atomicAdd(Q.n_direct_list, 1u)] = t;
    }
    t += W;
  }
}

// ---- 3. expansion: one warp per tile ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) char_expand_kernel(const __grid_constant__ SplitParams Q) {
  const CharParams& P = Q.c;
  const ScanParams& p = P.f.sp;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const int lane = threadIdx.x & 31;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < P.f.n_tiles; t += n_warps) {
    if (Q.tile_hits[t] & kTileDirect) continue;
    const uint2 meta = *reinterpret_cast<const uint2*>(Q.tile_meta + ((size_t)t * 32 + lane) * kChains);
    const uint32_t n_ev0 = meta.x & 0xFFFFu, n_ev1 = meta.y & 0xFFFFu;
    const uint32_t n_ev = n_ev0 + n_ev1;
    uint32_t ev_incl = n_ev;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, ev_incl, o);
      if (lane >= o) ev_incl += v;
    }
    const uint32_t ev_excl = ev_incl - n_ev;
    const uint32_t E = __shfl_sync(0xffffffffu, ev_incl, 31);
    if (E == 0) continue;
    const unsigned long long prefix = Q.tile_prefix[t];
    const CharEvent* tile_ev = Q.store + (size_t)t * 32 * kStoreSlots;
    uint32_t run = 0;
    // kExpandRounds rounds of 32 events in flight: all event loads first, then all ref_rec loads, then the ordered stores
    for (uint32_t base = 0; base < E; base += 32 * kExpandRounds) {
      CharEvent ev[kExpandRounds];
      RefRec rr[kExpandRounds];
#pragma unroll
      for (int r = 0; r < kExpandRounds; ++r) {
        ev[r] = CharEvent{0, kMarkRef};
        if (base + r * 32 < E) {  // warp-uniform
          const uint32_t i = base + r * 32 + lane;
          uint32_t lo = 0;
#pragma unroll
          for (int step = 16; step > 0; step >>= 1) {
            const uint32_t cand = lo + step;
            const uint32_t ex = __shfl_sync(0xffffffffu, ev_excl, cand & 31);
            if (cand < 32 && ex <= i) lo = cand;
          }
          const uint32_t ex = __shfl_sync(0xffffffffu, ev_excl, lo);
          const uint32_t n0 = __shfl_sync(0xffffffffu, n_ev0, lo);
          if (i < E) {
            const uint32_t idx = i - ex;
            ev[r] = idx < n0 ? tile_ev[idx * 64 + lo * kChains] : tile_ev[(idx - n0) * 64 + lo * kChains + 1];
          }
        }
      }
#pragma unroll
      for (int r = 0; r < kExpandRounds; ++r) {
        rr[r] = RefRec{0, 0, 0, 0};
        if (ev[r].ref != kMarkRef) rr[r] = P.t.ref_rec[ev[r].ref];
      }
#pragma unroll
      for (int r = 0; r < kExpandRounds; ++r) {
        if (base + r * 32 < E) {  // warp-uniform
          uint32_t len_incl = rr[r].n;
          for (int o = 1; o < 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, len_incl, o);
            if (lane >= o) len_incl += v;
          }
          const int64_t at = (int64_t)prefix + run + (len_incl - rr[r].n);
          if (base + r * 32 + lane < E) {
            if (ev[r].ref == kMarkRef) p.doc_hit_off[ev[r].end] = at;
            else if (p.hits) expand_event(p, rr[r], ev[r].end, chars, p.hits, at, P.f.hit_cap);
          }
          run += __shfl_sync(0xffffffffu, len_incl, 31);
        }
      }
    }
  }
}

// ---- 4. tiles the walk could not finish: byte automaton over the re-staged tile, one warp per tile ------------------
__global__ void __launch_bounds__(128) char_direct_kernel(const __grid_constant__ SplitParams Q) {
  extern __shared__ __align__(128) uint8_t smem[];
  const CharParams& P = Q.c;
  const FusedParams& F = P.f;
  const ScanParams& p = F.sp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* s_cmap = smem;
  for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) s_cmap[i] = P.cmap_g[i];
  uint8_t* sbuf = smem + 256 + (size_t)warp * (size_t)F.buf_bytes;
  __syncthreads();
  ByteGlobal fallback;
  fallback.cmap_g = P.cmap_g;
  fallback.cmap = s_cmap;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const unsigned int n_list = *Q.n_direct_list;
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
    const uint32_t mine_hits = Q.tile_meta[((size_t)t * 32 + lane) * kChains] >> 16;
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

}  // namespace aha
