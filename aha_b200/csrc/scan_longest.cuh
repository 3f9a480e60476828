// scan_longest.cuh -- ACX#match_longest (ac.cr:118-143, 249-263, 297-319) in parallel INSIDE a document.
//
// The machine is sequential: its state at a byte is {nid, prev_i, prev_nid} -- where the automaton stands and which end
// state is pending -- and without `intersectable` a yield resets nid, so the state depends on the whole history, not on a
// window of the text.  It is run speculatively:
//
//   1. guess   One thread per 512-byte chunk of the concatenated text.  The thread starts kWarm bytes in front of its
//              chunk (never in front of the document's start) in the state "root, nothing pending", and takes what it has
//              when it reaches the chunk as its entry state; then it walks the chunk: hits counted, exit state stored.
//              In practice the machine forgets its start within a few keys, so nearly every guessed entry is the true one.
//   2. settle  Chunk k's true entry is chunk k-1's exit.  Every round, a chunk whose last walk started from something
//              else walks again from it (hits recounted, new exit state).  Exits are double-buffered, so a round reads
//              only the previous round's; rounds repeat until nothing changes -- by induction from each document's first
//              chunk (whose entry is exact) every chunk has then been walked from its true entry.  Usually the first
//              round already changes nothing that matters and the second confirms it; the worst case (a text on which
//              the machine never re-synchronises) degenerates to one chunk per round, still exact.
//   3. emit    Prefix sum of the counts, then every chunk walks once more from its entry and writes its hits in place.
//
// A hit belongs to the chunk in which it is YIELDED (the failing byte, or the document's end), which may lie after the
// chunk of its end position: yields happen in position order, so chunk order is stream order.
#pragma once
#include "scan_common.cuh"

namespace aha {

#ifndef AHA_LONG_CHUNK
#define AHA_LONG_CHUNK 512
#endif
#ifndef AHA_LONG_WARM
#define AHA_LONG_WARM 96
#endif
constexpr int kLongChunk = AHA_LONG_CHUNK;  // bytes per thread
constexpr int kLongWarm = AHA_LONG_WARM;    // bytes of speculative run-up in front of a chunk

struct alignas(16) LongState {
  uint32_t s;       // automaton state
  uint32_t prev_s;  // pending end state (valid when prev_i >= 0)
  int64_t prev_i;   // absolute position of the pending end state's last byte, -1 = nothing pending
};
__device__ __forceinline__ bool same_state(const LongState& a, const LongState& b) {
  return a.s == b.s && a.prev_i == b.prev_i && (a.prev_i < 0 || a.prev_s == b.prev_s);
}
__device__ __forceinline__ LongState load_state(const LongState* p) {
  const uint4 v = *reinterpret_cast<const uint4*>(p);  // one 16-byte access: never seen half written
  LongState s;
  s.s = v.x; s.prev_s = v.y; s.prev_i = (int64_t)(((uint64_t)v.w << 32) | v.z);
  return s;
}
__device__ __forceinline__ void store_state(LongState* p, const LongState& s) {
  *reinterpret_cast<uint4*>(p) = make_uint4(s.s, s.prev_s, (uint32_t)(uint64_t)s.prev_i, (uint32_t)((uint64_t)s.prev_i >> 32));
}

struct LongParams {
  ScanParams sp;             // counts[] = hits yielded per chunk, offsets[] = their exclusive prefix (emit)
  LongState* exit_prev;      // [n_chunks] exit states of the previous round
  LongState* exit_next;      // [n_chunks] exit states of this round
  LongState* entry_used;     // [n_chunks] entry state of every chunk's last walk
  unsigned int* changed;     // settle: chunks that walked again in this round
};

// Walks chunk g from `st` (its state at the chunk's first byte).  EMIT = false: counts the hits it yields; EMIT = true:
// writes them from p.offsets[g] on and the doc_hit_off of the documents that start in the chunk.  Returns the exit state.
template <bool EMIT>
__device__ __forceinline__ LongState long_walk_chunk(const ScanParams& p, const uint8_t* cmap, int64_t g, LongState st, uint32_t* count_out) {
  const int64_t a = g * kLongChunk;
  const int64_t b = min(a + (int64_t)kLongChunk, p.total);
  const bool inter = (p.mode & AHA_INTERSECTABLE) != 0;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  int64_t d = last_le(p.doc_off, p.n_docs, a);
  if (d >= p.n_docs) d = p.n_docs - 1;
  int64_t doc_start = p.doc_off[d], doc_end = p.doc_off[d + 1];
  unsigned long long run = EMIT ? p.offsets[g] : 0ull;
  if (EMIT && doc_start == a)
    for (int64_t e = first_ge(p.doc_off, p.n_docs, a); e <= d; ++e) p.doc_hit_off[e] = (int64_t)run;
  // characters of the document up to and including the current byte (String overloads); a pending end carried in from
  // an earlier chunk has its count worked out when (if) it is written
  uint32_t lead = 0, prev_lead = 0;
  bool carried = st.prev_i >= 0;
  if (chars) lead = (uint32_t)(p.lead_prefix[g] - p.doc_lead_base[d]);
  auto flush = [&]() {  // fetch_one: the first live value of the state's chain is the state's own key (ac.cr:249-263)
    const uint32_t j = p.out_off[st.prev_s];
    if (j < p.out_off[st.prev_s + 1]) {
      if (EMIT && p.hits) {
        const int32_t v = p.out_val[j];
        aha_hit h;
        if (chars) {
          uint32_t pl = prev_lead;
          if (carried) {  // unit starts of the document up to the pending end: its chunk's prefix + the bytes of that chunk
            const int64_t cg = st.prev_i / kLongChunk;
            pl = (uint32_t)(p.lead_prefix[cg] - p.doc_lead_base[d]);
            for (int64_t i = cg * kLongChunk; i <= st.prev_i; ++i) pl += ((p.text[i] & 0xC0) != 0x80);
          }
          h.end = (int32_t)pl;
          h.start = h.end - 1 - (int32_t)p.key_tail[v];
        } else {
          h.end = (int32_t)(st.prev_i - doc_start + 1);
          h.start = h.end - (int32_t)p.key_len[v];
        }
        h.value = v;
        p.hits[run] = h;
      }
      ++run;
    }
    st.prev_i = -1;
    carried = false;
  };
  int64_t pos = a;
  while (pos < b) {
    const int64_t seg_end = min(b, doc_end);
    for (; pos < seg_end; ++pos) {
      const uint8_t byte = p.text[pos];
      lead += ((byte & 0xC0) != 0x80);
      const uint32_t e = delta_entry(p, cmap, st.s, byte);
      if (e & kGoto) {
        st.s = e & kStateMask;
        if (e & kEnd) { st.prev_i = pos; st.prev_s = st.s; prev_lead = lead; carried = false; }
      } else {
        if (st.prev_i >= 0) {
          flush();
          if (!inter) { st.s = 0; continue; }  // the failing byte is dropped (Q5)
        }
        st.s = e & kStateMask;  // fail walk, then the goto that finally succeeded (or the root)
        if (st.s != 0 && (e & kEnd)) { st.prev_i = pos; st.prev_s = st.s; prev_lead = lead; carried = false; }
      }
    }
    if (pos != doc_end) break;
    // the document ends here (possibly exactly at the end of the chunk): its pending end is yielded by this chunk
    if (st.prev_i >= 0) flush();
    st.s = 0;
    if (!(pos < b || b == p.total)) break;  // the next document starts with the next chunk
    lead = 0;
    ++d;
    while (d < p.n_docs) {
      if (EMIT) p.doc_hit_off[d] = (int64_t)run;
      if (p.doc_off[d + 1] > pos) break;
      ++d;
    }
    if (d >= p.n_docs) {
      if (EMIT) p.doc_hit_off[p.n_docs] = (int64_t)run;
      break;
    }
    doc_start = p.doc_off[d];
    doc_end = p.doc_off[d + 1];
  }
  if (!EMIT) *count_out = (uint32_t)run;
  return st;
}

// the state machine alone over [from, to) of one document (no hits, no boundaries): the speculative run-up
__device__ __forceinline__ LongState long_run_up(const ScanParams& p, const uint8_t* cmap, int64_t from, int64_t to) {
  const bool inter = (p.mode & AHA_INTERSECTABLE) != 0;
  LongState st{0u, 0u, -1};
  for (int64_t pos = from; pos < to; ++pos) {
    const uint32_t e = delta_entry(p, cmap, st.s, p.text[pos]);
    if (e & kGoto) {
      st.s = e & kStateMask;
      if (e & kEnd) { st.prev_i = pos; st.prev_s = st.s; }
    } else {
      if (st.prev_i >= 0) {
        st.prev_i = -1;
        if (!inter) { st.s = 0; continue; }
      }
      st.s = e & kStateMask;
      if (st.s != 0 && (e & kEnd)) { st.prev_i = pos; st.prev_s = st.s; }
    }
  }
  return st;
}

// kPhase 0 = guess, 1 = settle, 2 = emit
template <int kPhase>
__global__ void __launch_bounds__(128) longest_chunk_kernel(const __grid_constant__ LongParams q) {
  const ScanParams& p = q.sp;
  __shared__ uint8_t cmap[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) cmap[i] = p.classmap[i];
  __syncthreads();
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= p.n_chunks) return;
  const int64_t a = g * kLongChunk;
  if (kPhase == 2) {
    long_walk_chunk<true>(p, cmap, g, load_state(q.entry_used + g), nullptr);
    return;
  }
  // a chunk that begins where a document begins starts from the root with nothing pending: exact
  const int64_t d = last_le(p.doc_off, p.n_docs, a);
  const int64_t doc_start = p.doc_off[min(d, p.n_docs - 1)];
  const bool fresh = doc_start == a || g == 0;
  LongState entry{0u, 0u, -1};
  if (kPhase == 0) {
    if (!fresh) entry = long_run_up(p, cmap, max(doc_start, a - (int64_t)kLongWarm), a);
  } else {
    if (!fresh) entry = load_state(q.exit_prev + g - 1);
    if (same_state(entry, load_state(q.entry_used + g))) {
      store_state(q.exit_next + g, load_state(q.exit_prev + g));
      return;
    }
    atomicAdd(q.changed, 1u);
  }
  uint32_t count = 0;
  const LongState out = long_walk_chunk<false>(p, cmap, g, entry, &count);
  p.counts[g] = count;
  store_state(q.entry_used + g, entry);
  store_state(q.exit_next + g, out);
}

}  // namespace aha
