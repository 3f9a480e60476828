// scan_v1.cuh -- baseline two-pass scan (count, prefix, emit): one thread per fixed-size chunk of the
// concatenated text, DFA read from global memory.  Kept as the simple, obviously-correct device path
// (and as the checker for the fused kernel); the production kernel is scan_fused.cuh.
#pragma once
#include "scan_common.cuh"

namespace aha {

// Emits the output list of `entry`'s target state for a match ending at absolute byte `pos`
// (ACX#fetch, ac.cr:265-278 + the sep filter ac.cr:321-340 + the char remap matcher.cr:34-46).
// Returns the number of hits produced; writes them when EMIT.
template <bool EMIT>
__device__ __forceinline__ uint32_t emit_state(const ScanParams& p, uint32_t state, int64_t pos, int64_t doc_start,
                                                int64_t doc_end, uint64_t lead_incl, uint64_t doc_lead, uint64_t out_at) {
  if ((p.mode & AHA_SEP) && sep_veto_after(p, pos + 1, doc_end)) return 0;
  uint32_t n = 0;
  const uint32_t j1 = p.out_off[state + 1];
  const int32_t end = (int32_t)(pos - doc_start + 1);
  for (uint32_t j = p.out_off[state]; j < j1; ++j) {
    const int32_t v = p.out_val[j];
    const int32_t start = end - (int32_t)p.key_len[v];
    if ((p.mode & AHA_SEP) && sep_veto_before(p, doc_start + start, doc_start)) continue;
    if (EMIT && p.hits) {
      aha_hit h;
      if (p.mode & AHA_UNITS_CHAR) {
        h.end = (int32_t)(lead_incl - doc_lead);
        h.start = h.end - 1 - (int32_t)p.key_tail[v];
      } else {
        h.start = start;
        h.end = end;
      }
      h.value = v;
      p.hits[out_at + n] = h;
    }
    ++n;
  }
  return n;
}

template <bool EMIT>
__global__ void __launch_bounds__(128) scan_v1_kernel(const ScanParams p) {
  __shared__ uint8_t cmap[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) cmap[i] = p.classmap[i];
  __syncthreads();
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= p.n_chunks) return;
  const int64_t a = g * p.chunk;
  const int64_t b = min(a + (int64_t)p.chunk, p.total);
  int64_t d = last_le(p.doc_off, p.n_docs, a);
  if (d >= p.n_docs) d = p.n_docs - 1;
  int64_t doc_start = p.doc_off[d], doc_end = p.doc_off[d + 1];
  uint64_t run = EMIT ? p.offsets[g] : 0;
  uint64_t lead = (EMIT && (p.mode & AHA_UNITS_CHAR)) ? p.lead_prefix[g] : 0;
  uint64_t doc_lead = (EMIT && (p.mode & AHA_UNITS_CHAR)) ? p.doc_lead_base[d] : 0;
  const uint32_t C = p.C;
  // documents starting exactly at `a` (possibly several empty ones) are owned by this chunk
  if (EMIT && doc_start == a) {
    for (int64_t e = first_ge(p.doc_off, p.n_docs, a); e <= d; ++e) p.doc_hit_off[e] = (int64_t)run;
  }
  // warm-up: the DFA state at `a` depends only on the previous max_key_len-1 bytes of the document
  uint32_t s = 0;
  {
    int64_t w = a - (int64_t)p.warm;
    if (w < doc_start) w = doc_start;
    for (int64_t i = w; i < a; ++i) s = delta_entry(p, cmap, s, p.text[i]) & kStateMask;
  }
  int64_t pos = a;
  while (pos < b) {
    const int64_t seg_end = min(b, doc_end);
    for (; pos < seg_end; ++pos) {
      const uint8_t byte = p.text[pos];
      lead += ((byte & 0xC0) != 0x80);
      const uint32_t e = delta_entry(p, cmap, s, byte);
      s = e & kStateMask;
      if (e & kOut) run += emit_state<EMIT>(p, s, pos, doc_start, doc_end, lead, doc_lead, run);
    }
    if (pos == doc_end && (pos < b || b == p.total)) {
      // step into the next non-empty document (or finish): reset the automaton
      s = 0;
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
      if (EMIT && (p.mode & AHA_UNITS_CHAR)) doc_lead = p.doc_lead_base[d];
    }
  }
  if (!EMIT) p.counts[g] = (uint32_t)run;
}

// non-continuation bytes per chunk (matcher.cr:14-22: one character per lead byte)
// One warp per chunk, 16-byte loads when the chunk is aligned (HBM-bound: 1 byte read per text byte).
__global__ void count_leads_kernel(const uint8_t* text, int64_t total, int32_t chunk, int64_t n_chunks, uint32_t* counts) {
  const int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (g > n_chunks) return;
  uint32_t c = 0;
  if (g < n_chunks) {
    const int64_t a = g * chunk, b = min(a + (int64_t)chunk, total);
    int64_t i = a;
    if (((reinterpret_cast<uintptr_t>(text) + (uintptr_t)a) & 15) == 0) {
      const uint4* v = reinterpret_cast<const uint4*>(text + a);
      const int64_t n16 = (b - a) >> 4;
      for (int64_t k = lane; k < n16; k += 32) {
        const uint4 q = __ldg(&v[k]);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) c += 4u - (uint32_t)__popc(w[j] & 0x80808080u & ~(w[j] << 1));
      }
      i = a + (n16 << 4);
    }
    for (int64_t k = i + lane; k < b; k += 32) c += ((text[k] & 0xC0) != 0x80);
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  }
  if (lane == 0) counts[g] = c;
}

__global__ void doc_lead_base_kernel(const uint8_t* text, int64_t total, const int64_t* doc_off, int64_t n_docs,
                                     int32_t chunk, const uint64_t* lead_prefix, uint64_t* doc_lead_base) {
  // one warp per document boundary: the bytes between the tile start and the boundary, 32 at a time
  const int64_t d = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (d > n_docs) return;
  const int64_t pos = doc_off[d];
  const int64_t g = pos / chunk;
  uint32_t c = 0;
  for (int64_t i = g * chunk + lane; i < pos; i += 32) c += ((text[i] & 0xC0) != 0x80);
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) doc_lead_base[d] = lead_prefix[g] + c;
}

// ---- ACX#match_longest (ac.cr:118-143, 249-263): sequential machine, one thread per document ----
template <bool EMIT>
__global__ void __launch_bounds__(128) longest_doc_kernel(const ScanParams p) {
  __shared__ uint8_t cmap[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) cmap[i] = p.classmap[i];
  __syncthreads();
  const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= p.n_docs) return;
  const int64_t doc_start = p.doc_off[d], doc_end = p.doc_off[d + 1];
  const bool inter = (p.mode & AHA_INTERSECTABLE) != 0;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const uint32_t C = p.C;
  uint64_t run = EMIT ? p.offsets[d] : 0;
  uint32_t s = 0, prev_s = 0;
  int64_t prev_i = -1;
  uint32_t lead = 0, prev_lead = 0;
  // fetch_one: the first live value of the state's chain is the state's own key (ac.cr:249-263)
  auto flush = [&]() {
    const uint32_t j = p.out_off[prev_s];
    if (j < p.out_off[prev_s + 1]) {
      if (EMIT && p.hits) {
        const int32_t v = p.out_val[j];
        aha_hit h;
        if (chars) {
          h.end = (int32_t)prev_lead;
          h.start = h.end - 1 - (int32_t)p.key_tail[v];
        } else {
          h.end = (int32_t)(prev_i - doc_start + 1);
          h.start = h.end - (int32_t)p.key_len[v];
        }
        h.value = v;
        p.hits[run] = h;
      }
      ++run;
    }
    prev_i = -1;
  };
  for (int64_t pos = doc_start; pos < doc_end; ++pos) {
    const uint8_t byte = p.text[pos];
    lead += ((byte & 0xC0) != 0x80);
    const uint32_t e = delta_entry(p, cmap, s, byte);
    if (e & kGoto) {
      s = e & kStateMask;
      if (e & kEnd) { prev_i = pos; prev_s = s; prev_lead = lead; }
    } else {
      if (prev_i >= 0) {
        flush();
        if (!inter) { s = 0; continue; }  // the failing byte is dropped (Q5)
      }
      s = e & kStateMask;  // fail walk, then the goto that finally succeeded (or the root)
      if (s != 0 && (e & kEnd)) { prev_i = pos; prev_s = s; prev_lead = lead; }
    }
  }
  if (prev_i >= 0) flush();
  if (!EMIT) p.counts[d] = (uint32_t)run;
}

__global__ void offsets_to_doc_hit_off(const uint64_t* offsets, int64_t n, int64_t* doc_hit_off) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= n) doc_hit_off[i] = (int64_t)offsets[i];
}

}  // namespace aha
