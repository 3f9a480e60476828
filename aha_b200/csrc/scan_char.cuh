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

namespace aha {

constexpr int kEvCap = 32;  // events per lane per tile kept in the scratch ring

struct CharEvent {
  int32_t end;   // document-relative end offset (bytes, or characters with AHA_UNITS_CHAR)
  uint32_t ref;  // firing edge slot / m + cid ; bit 31 = more than one key fires
};

struct CharDeviceTables {
  const uint32_t* root_g;   // global copies, loaded into shared memory at kernel start
  const uint16_t* check_g;
  const uint32_t* val_g;
  const uint16_t* gfail;    // stays in global memory
  const RefRec* ref_rec;    // stays in global memory
  uint32_t n_root, m;
  uint32_t lo[3], n[3], off[3];
};

struct CharParams {
  FusedParams f;            // f.scratch is unused here
  CharDeviceTables t;
  CharEvent* events;        // [n_warps_total][32][kEvCap]
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
                             [&](uint32_t ref, uint32_t multi, int32_t last, uint32_t k) {
                               const int32_t end = chars ? (int32_t)(lead + k) : (int32_t)(last + origin - doc_start + 1);
                               on_event(end, ref | (multi << 31));
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

__global__ void __launch_bounds__(kFusedWarps * 32, 1) scan_char_kernel(const CharParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  const FusedParams& F = P.f;
  const ScanParams& p = F.sp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // shared layout: root u32[n_root] | val u32[m] | check u16[m] | cmap u8[256] | per-warp {mbarrier, text}
  uint32_t* s_root = reinterpret_cast<uint32_t*>(smem);
  uint32_t* s_val = s_root + P.t.n_root;
  uint16_t* s_check = reinterpret_cast<uint16_t*>(s_val + P.t.m);
  uint8_t* s_cmap = reinterpret_cast<uint8_t*>(s_check + P.t.m);
  for (uint32_t i = threadIdx.x; i < P.t.n_root; i += blockDim.x) s_root[i] = P.t.root_g[i];
  for (uint32_t i = threadIdx.x; i < P.t.m; i += blockDim.x) {
    s_val[i] = P.t.val_g[i];
    s_check[i] = P.t.check_g[i];
  }
  for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) s_cmap[i] = P.cmap_g[i];
  const uint32_t table_bytes = P.t.n_root * 4 + P.t.m * 6 + 256;
  uint8_t* wbase = smem + ((table_bytes + 127) & ~127u) + (size_t)warp * (size_t)(F.buf_bytes + 128);
  uint64_t* bar = reinterpret_cast<uint64_t*>(wbase);
  uint8_t* sbuf = wbase + 128;
  if (lane == 0) mbar_init(bar, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();

  CharTablesView T;
  T.root = s_root; T.check = s_check; T.val = s_val; T.gfail = P.t.gfail; T.m = P.t.m;
#pragma unroll
  for (int c = 0; c < 3; ++c) { T.lo[c] = P.t.lo[c]; T.n[c] = P.t.n[c]; T.off[c] = P.t.off[c]; }
  ByteGlobal fallback;
  fallback.cmap_g = P.cmap_g;
  fallback.cmap = s_cmap;

  const int64_t gwarp = (int64_t)blockIdx.x * kFusedWarps + warp;
  CharEvent* my_ev = P.events + (gwarp * 32 + lane) * kEvCap;
  int2* my_mark = F.marks + (gwarp * 32 + lane) * kMarkCap;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  uint32_t parity = 0;

  for (;;) {
    unsigned int t = 0;
    if (lane == 0) t = atomicAdd(F.ticket, 1u);
    t = __shfl_sync(0xffffffffu, t, 0);
    if ((int64_t)t >= F.n_tiles) break;
    const int64_t tile_start = (int64_t)t * F.tile;
    const int64_t tile_end = min(tile_start + (int64_t)F.tile, p.total);
    const int64_t origin = max((int64_t)0, tile_start - (int64_t)F.w16);
    const int64_t want_end = min(p.total, tile_end + 1);
    const int64_t bulk_end = origin + ((want_end - origin) & ~(int64_t)15);
    if (lane == 0 && bulk_end > origin) {
      mbar_expect_tx(bar, (uint32_t)(bulk_end - origin));
      bulk_g2s(sbuf, p.text + origin, (uint32_t)(bulk_end - origin), bar);
    }
    for (int64_t i = bulk_end + lane; i < want_end; i += 32) sbuf[i - origin] = p.text[i];
    if (bulk_end > origin) {
      mbar_wait(bar, parity);
      parity ^= 1u;
    }
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

    // ---- walk: events to the lane's scratch slot, no global loads on the chain
    uint32_t n_ev = 0, n_single = 0, n_mark = 0, multi_mask = 0;
    bool ok = char_lane_walk(
        P, T, sbuf, origin, a, b, lead0,
        [&](int32_t end, uint32_t ref) {
          if (n_ev < (uint32_t)kEvCap) {
            my_ev[n_ev] = CharEvent{end, ref};
            multi_mask |= (ref >> 31) << n_ev;
          }
          ++n_ev;
          n_single += (ref >> 31) ^ 1u;
        },
        [&](int64_t d) {
          if (n_mark < (uint32_t)kMarkCap) my_mark[n_mark] = make_int2((int)d, (int)n_ev);
          ++n_mark;
        });
    const bool bad = __any_sync(0xffffffffu, !ok);
    const bool over = __any_sync(0xffffffffu, n_ev > (uint32_t)kEvCap || n_mark > (uint32_t)kMarkCap);
    __syncwarp();

    // ---- hits per lane
    uint32_t n_hits;
    if (bad) {
      CountSink cs;
      fallback.walk(F, sbuf, origin, a, b, lead0, cs);
      n_hits = cs.n;
    } else if (over) {
      uint32_t c = 0;
      char_lane_walk(P, T, sbuf, origin, a, b, lead0, [&](int32_t, uint32_t ref) { c += (ref >> 31) ? P.t.ref_rec[ref & 0x7FFFFFFFu].n : 1u; },
                     [&](int64_t) {});
      n_hits = c;
    } else {
      n_hits = n_single;
      for (uint32_t mm = multi_mask; mm; mm &= mm - 1) {
        const uint32_t ref = my_ev[__ffs(mm) - 1].ref;
        n_hits += P.t.ref_rec[ref & 0x7FFFFFFFu].n;
      }
    }
    uint32_t incl = n_hits;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    const uint32_t excl_lane = incl - n_hits;
    const uint32_t Ttile = __shfl_sync(0xffffffffu, incl, 31);
    const unsigned long long prefix = tile_exclusive_prefix(F, t, Ttile, lane);
    if ((int64_t)t == F.n_tiles - 1 && lane == 0) *F.total_hits = prefix + Ttile;

    // ---- deliver
    if (bad) {
      if (lane == 0) atomicAdd(F.n_direct, 1u);
      DirectSink ds;
      ds.out = p.hits; ds.doc_hit_off = p.doc_hit_off; ds.idx = (int64_t)(prefix + excl_lane); ds.cap = F.hit_cap;
      fallback.walk(F, sbuf, origin, a, b, lead0, ds);
    } else if (over) {
      if (lane == 0) atomicAdd(F.n_direct, 1u);
      int64_t idx = (int64_t)(prefix + excl_lane);
      char_lane_walk(
          P, T, sbuf, origin, a, b, lead0,
          [&](int32_t end, uint32_t ref) {
            const RefRec rr = P.t.ref_rec[ref & 0x7FFFFFFFu];
            if (p.hits) expand_event(p, rr, end, chars, p.hits, idx, F.hit_cap);
            idx += rr.n;
          },
          [&](int64_t d) { p.doc_hit_off[d] = idx; });
    } else {
      // markers: hits of this lane before the document start = sum over its first y events
      for (uint32_t k = 0; k < n_mark; ++k) {
        const int y = my_mark[k].y;  // events of this lane before the document start
        uint32_t before = (uint32_t)y;
        for (uint32_t mm = multi_mask & (y >= 32 ? 0xffffffffu : ((1u << y) - 1u)); mm; mm &= mm - 1) {
          const uint32_t ref = my_ev[__ffs(mm) - 1].ref;
          before += P.t.ref_rec[ref & 0x7FFFFFFFu].n - 1u;
        }
        p.doc_hit_off[my_mark[k].x] = (int64_t)(prefix + excl_lane + before);
      }
      // cooperative expansion: events in (lane, seq) order, 32 per round
      uint32_t ev_incl = n_ev;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, ev_incl, o);
        if (lane >= o) ev_incl += v;
      }
      const uint32_t ev_excl = ev_incl - n_ev;
      const uint32_t E = __shfl_sync(0xffffffffu, ev_incl, 31);
      if (p.hits && E > 0) {
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
            ev = P.events[(gwarp * 32 + lo) * kEvCap + (i - ex)];
            rr = P.t.ref_rec[ev.ref & 0x7FFFFFFFu];
          }
          uint32_t len_incl = rr.n;
          for (int o = 1; o < 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, len_incl, o);
            if (lane >= o) len_incl += v;
          }
          if (i < E) expand_event(p, rr, ev.end, chars, p.hits, (int64_t)prefix + run + (len_incl - rr.n), F.hit_cap);
          run += __shfl_sync(0xffffffffu, len_incl, 31);
        }
      }
    }
    __syncwarp();
    fence_proxy_async();
  }
}

}  // namespace aha
