// api_scan.cuh -- scan / hits / generator entry points (included at the end of api.cu).
//
// A warm scan call allocates nothing and never waits for the device in the middle of its pipeline: scratch lives in a
// Workspace that the automaton keeps per device, result buffers are recycled through a per-device cache, the hit
// buffer is sized from the densest batch seen so far, and the only host wait is the one at the end of the call that
// reads {total hits, status} back.  A batch denser than anything seen before overflows the buffer (writes beyond its
// capacity are dropped on the device) and re-runs the expansion alone with the exact size.
#pragma once
#include <chrono>
#include <condition_variable>
#include <thread>

namespace aha {

struct U32ToU64 {
  __host__ __device__ __forceinline__ uint64_t operator()(const uint32_t& x) const { return (uint64_t)x; }
};

// ---- result buffer cache -------------------------------------------------------------------------------------------
static std::mutex g_cache_mu;
static std::vector<std::unique_ptr<DevCache>> g_caches;  // index = device
DevCache& dev_cache(int device) {
  std::lock_guard<std::mutex> lk(g_cache_mu);
  if ((int)g_caches.size() <= device) g_caches.resize(device + 1);
  if (!g_caches[device]) g_caches[device].reset(new DevCache());
  return *g_caches[device];
}

// Every block carries an event recorded on the stream it was released on.  A block whose event has completed is free
// for anybody; when all fitting blocks are still busy (the D2H copy of the previous run is reading one) a new block is
// allocated instead of waiting, unless the block is large -- then the new owner's stream waits for it on the device.
void* DevCache::acquire(size_t bytes, cudaStream_t st, size_t* got) {
  bytes = (bytes + 255) & ~size_t(255);
  Entry pick{nullptr, 0, nullptr};
  bool wait = false;
  {
    std::lock_guard<std::mutex> lk(mu);
    size_t best = free_list.size(), best_busy = free_list.size();
    for (size_t i = 0; i < free_list.size(); ++i) {
      const Entry& e = free_list[i];
      if (e.bytes < bytes || e.bytes > 2 * bytes + (1 << 20)) continue;
      const bool idle = cudaEventQuery(e.ev) == cudaSuccess;
      size_t& slot = idle ? best : best_busy;
      if (slot == free_list.size() || e.bytes < free_list[slot].bytes) slot = i;
    }
    cudaGetLastError();  // cudaErrorNotReady of a busy block is not an error
    if (best == free_list.size() && best_busy != free_list.size() && bytes >= ((size_t)1 << 30)) {
      best = best_busy;
      wait = true;
    }
    if (best != free_list.size()) {
      pick = free_list[best];
      cached -= pick.bytes;
      free_list.erase(free_list.begin() + best);
      in_use.push_back(pick);
    }
  }
  if (pick.p) {
    if (wait) cudaStreamWaitEvent(st, pick.ev, 0);
    *got = pick.bytes;
    return pick.p;
  }
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) {
    cudaGetLastError();
    trim();  // give the cached blocks back and try once more
    if (cudaMalloc(&p, bytes) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
  }
  cudaEvent_t ev = nullptr;
  if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) {
    cudaFree(p);
    return nullptr;
  }
  {
    std::lock_guard<std::mutex> lk(mu);
    in_use.push_back(Entry{p, bytes, ev});
  }
  *got = bytes;
  return p;
}
// the caller's current device is the block's device
void DevCache::release(void* p, size_t /*bytes*/, cudaStream_t st) {
  if (!p) return;
  Entry e{nullptr, 0, nullptr}, drop{nullptr, 0, nullptr};
  {
    std::lock_guard<std::mutex> lk(mu);
    for (size_t i = 0; i < in_use.size(); ++i)
      if (in_use[i].p == p) {
        e = in_use[i];
        in_use.erase(in_use.begin() + i);
        break;
      }
    if (!e.p) return;  // not ours
    cudaEventRecord(e.ev, st);
    free_list.push_back(e);
    cached += e.bytes;
    if (free_list.size() > kMaxEntries) {  // drop the smallest block: the large ones are the expensive ones to get again
      size_t worst = 0;
      for (size_t i = 1; i < free_list.size(); ++i)
        if (free_list[i].bytes < free_list[worst].bytes) worst = i;
      drop = free_list[worst];
      cached -= drop.bytes;
      free_list.erase(free_list.begin() + worst);
    }
  }
  if (drop.p) {
    cudaEventSynchronize(drop.ev);
    cudaEventDestroy(drop.ev);
    cudaFree(drop.p);
  }
}
void DevCache::trim() {
  std::vector<Entry> all;
  {
    std::lock_guard<std::mutex> lk(mu);
    all.swap(free_list);
    cached = 0;
  }
  for (auto& e : all) {
    cudaEventSynchronize(e.ev);
    cudaEventDestroy(e.ev);
    cudaFree(e.p);
  }
}

// ---- workspaces ----------------------------------------------------------------------------------------------------
static int32_t ws_acquire(const aha_dfa* d, DeviceDfa* dd, std::unique_ptr<Workspace>* out) {
  {
    std::lock_guard<std::mutex> lk(d->pool_mu);
    if (!dd->idle.empty()) {
      *out = std::move(dd->idle.back());
      dd->idle.pop_back();
      return AHA_OK;
    }
  }
  std::unique_ptr<Workspace> w(new Workspace());
  AHA_CUDA(w->init(dd->device));
  *out = std::move(w);
  return AHA_OK;
}
static void ws_release(const aha_dfa* d, DeviceDfa* dd, std::unique_ptr<Workspace> w) {
  std::lock_guard<std::mutex> lk(d->pool_mu);
  dd->idle.push_back(std::move(w));
}

static inline void count_launch(int n = 1) { g_kernel_launches.fetch_add(n, std::memory_order_relaxed); }

template <class InIt>
static int32_t exclusive_sum_u64(Workspace* ws, InIt in, uint64_t* out, int64_t n, cudaStream_t st) {
  size_t tmp_bytes = 0;
  AHA_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, in, out, n, st));
  uint8_t* tmp = nullptr;
  AHA_CUDA(ws->get(kWsScanTmp, tmp_bytes, &tmp));
  AHA_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, in, out, n, st));
  count_launch(2);
  return AHA_OK;
}
static int32_t exclusive_sum_u32_to_u64(Workspace* ws, const uint32_t* in, uint64_t* out, int64_t n, cudaStream_t st) {
  cub::TransformInputIterator<uint64_t, U32ToU64, const uint32_t*> it(in, U32ToU64());
  return exclusive_sum_u64(ws, it, out, n, st);
}

static void fill_params(ScanParams* p, const DeviceDfa* dd, const void* dev_text, const void* dev_doc_off, int64_t n_docs,
                        int64_t total, uint32_t mode, const uint8_t* sep_bits, int32_t sep_size) {
  memset(p, 0, sizeof(*p));
  p->text = static_cast<const uint8_t*>(dev_text);
  p->total = total;
  p->doc_off = static_cast<const int64_t*>(dev_doc_off);
  p->n_docs = n_docs;
  p->classmap = dd->view.classmap;
  p->delta = dd->view.has_delta ? dd->view.delta : nullptr;
  p->rows = reinterpret_cast<const uint4*>(dd->view.rows);
  p->wide = dd->view.wide;
  p->wide_fail = dd->view.wide_fail;
  p->n_wide_first = dd->view.wide_first ? (uint32_t)dd->view.n_wide : 0u;
  p->out_off = dd->view.out_off;
  p->out_val = dd->view.out_val;
  p->key_len = dd->view.key_len;
  p->key_tail = dd->view.key_tail;
  p->out8 = dd->out8;
  p->C = dd->view.n_classes;
  p->warm = dd->view.max_key_len > 0 ? dd->view.max_key_len - 1 : 0;
  p->mode = mode;
  p->sep_size = (uint32_t)sep_size;
  if (sep_bits) memcpy(p->sep_bits, sep_bits, 32);
}

// one scan call in flight: everything the run_* pipelines share
struct ScanCall {
  const aha_dfa* dfa;
  DeviceDfa* dd;
  Workspace* ws;
  cudaStream_t st;
  aha_hits* res;
  HitsOwner* own;
  int n_sm;
  bool validated;  // doc_off was checked on the host (aha_b200_scan)
};

static int32_t alloc_doc_hit_off(ScanCall& c, int64_t n_docs, int64_t** out) {
  size_t got = 0;
  c.own->dev_doc_hit_off = dev_cache(c.dd->device).acquire((size_t)(n_docs + 1) * sizeof(int64_t), c.st, &got);
  if (!c.own->dev_doc_hit_off) return set_err(AHA_ERR_NOMEM, "device allocation of doc_hit_off failed");
  c.own->dev_doc_hit_off_bytes = got;
  *out = static_cast<int64_t*>(c.own->dev_doc_hit_off);
  return AHA_OK;
}
static int32_t alloc_hits(ScanCall& c, int64_t cap, aha_hit** out) {
  if (c.own->dev_hits) dev_cache(c.dd->device).release(c.own->dev_hits, c.own->dev_hits_bytes, c.st);
  c.own->dev_hits = nullptr;
  size_t got = 0;
  c.own->dev_hits = dev_cache(c.dd->device).acquire((size_t)(cap > 0 ? cap : 1) * sizeof(aha_hit), c.st, &got);
  if (!c.own->dev_hits) return set_err(AHA_ERR_NOMEM, "device allocation of the hit buffer failed");
  c.own->dev_hits_bytes = got;
  *out = static_cast<aha_hit*>(c.own->dev_hits);
  return AHA_OK;
}
// hit buffer capacity before anything has been counted: the densest batch seen so far on this device plus a quarter
static int64_t estimate_cap(const ScanCall& c, int64_t total) {
  double dens;
  {
    std::lock_guard<std::mutex> lk(c.dfa->pool_mu);
    dens = c.dd->hits_per_byte;
  }
  if (dens < 0) dens = 1.0 / 16;
  return (int64_t)((double)total * dens * 1.25) + 65536;
}
static void note_density(const ScanCall& c, int64_t total, unsigned long long hits) {
  if (total <= 0) return;
  const double dens = (double)hits / (double)total;
  std::lock_guard<std::mutex> lk(c.dfa->pool_mu);
  if (dens > c.dd->hits_per_byte) c.dd->hits_per_byte = dens;
}
// the one host wait of a call: status block home, then the end event
static int32_t read_status(ScanCall& c, const ScanStatus* dev_status, cudaEvent_t done) {
  AHA_CUDA(cudaMemcpyAsync(c.ws->h_status, dev_status, sizeof(ScanStatus), cudaMemcpyDeviceToHost, c.st));
  AHA_CUDA(cudaEventRecord(done, c.st));
  AHA_CUDA(cudaEventSynchronize(done));
  if (c.ws->h_status->bad_docs)
    return set_err(AHA_ERR_ARG, "doc_off must start at 0, be non-decreasing, end at total_bytes, and every document must be < 2^31 bytes");
  return AHA_OK;
}

// character units (String overloads): unit starts in front of every tile / every document
static int32_t lead_prepass(ScanCall& c, ScanParams& p, int32_t tile, int64_t n_tiles, const uint64_t** tile_lead_out) {
  uint32_t* lead_counts = nullptr;
  uint64_t* tile_lead = nullptr;
  uint64_t* doc_lead = nullptr;
  AHA_CUDA(c.ws->get(kWsLeadCounts, (size_t)n_tiles + 1, &lead_counts));
  AHA_CUDA(c.ws->get(kWsTileLead, (size_t)n_tiles + 1, &tile_lead));
  AHA_CUDA(c.ws->get(kWsDocLead, (size_t)p.n_docs + 1, &doc_lead));
  count_leads_kernel<<<(unsigned)((n_tiles + 1 + 7) / 8), 256, 0, c.st>>>(p.text, p.total, tile, n_tiles, lead_counts);
  count_launch();
  int32_t rc = exclusive_sum_u32_to_u64(c.ws, lead_counts, tile_lead, n_tiles + 1, c.st);
  if (rc) return rc;
  doc_lead_base_kernel<<<(unsigned)((p.n_docs + 1 + 7) / 8), 256, 0, c.st>>>(p.text, p.total, p.doc_off, p.n_docs, tile, tile_lead, doc_lead);
  count_launch();
  *tile_lead_out = tile_lead;
  p.doc_lead_base = doc_lead;
  return AHA_OK;
}

// case-folded trie of the dictionary on this device (folded_trie.hpp): built from the host blob and uploaded at the first
// ignore_case scan; a dictionary that does not qualify keeps the depth-first kernel
static bool ensure_folded(const aha_dfa* d, DeviceDfa* dd) {
  std::lock_guard<std::mutex> lk(d->pool_mu);
  if (dd->folded_state != 0) return dd->folded_state > 0;
  dd->folded_state = -1;
  FoldedTrieHost F = build_folded_trie(d->header, d->blob.data());
  if (!F.ok) return false;
  auto place = [](size_t& cursor, size_t bytes) { size_t at = cursor; cursor = (cursor + bytes + 255) & ~size_t(255); return at; };
  size_t cursor = 0;
  const size_t off_delta = place(cursor, F.delta.size() * 4);
  dd->off_fcand_off = place(cursor, F.cand_off.size() * 4);
  dd->off_fcand = place(cursor, (F.cand.size() + 1) * 16);
  dd->off_fcmap = place(cursor, 256);
  dd->off_frec = place(cursor, F.rec.size() * 4);
  std::vector<uint8_t> img(cursor, 0);
  memcpy(img.data() + off_delta, F.delta.data(), F.delta.size() * 4);
  memcpy(img.data() + dd->off_fcand_off, F.cand_off.data(), F.cand_off.size() * 4);
  uint32_t* c = reinterpret_cast<uint32_t*>(img.data() + dd->off_fcand);
  for (size_t i = 0; i < F.cand.size(); ++i) {
    c[4 * i] = (uint32_t)F.cand[i].value;
    c[4 * i + 1] = F.cand[i].n_chars;
    c[4 * i + 2] = (uint32_t)F.cand[i].mask;
    c[4 * i + 3] = (uint32_t)(F.cand[i].mask >> 32);
  }
  memcpy(img.data() + dd->off_fcmap, F.cmap, 256);
  memcpy(img.data() + dd->off_frec, F.rec.data(), F.rec.size() * 4);
  if (cudaMalloc(&dd->folded_dev, img.size()) != cudaSuccess) {
    cudaGetLastError();
    dd->folded_dev = nullptr;
    return false;
  }
  if (cudaMemcpy(dd->folded_dev, img.data(), img.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
    cudaGetLastError();
    cudaFree(dd->folded_dev);
    dd->folded_dev = nullptr;
    return false;
  }
  dd->folded_classes = F.n_classes;
  dd->folded_max_chars = F.max_chars;
  dd->folded_state = 1;
  return true;
}

// per-document machines (match_longest, trie longest) and the two-pass checker: count -> exclusive sum -> emit
static int32_t run_v1(ScanParams p, ScanCall& c) {
  cudaStream_t st = c.st;
  Workspace* ws = c.ws;
  const bool trie = (p.mode & AHA_TRIE_LONGEST) != 0;
  const bool longest = trie || (p.mode & AHA_LONGEST) != 0;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  p.chunk = 512;
  p.n_chunks = (p.total + p.chunk - 1) / p.chunk;
  const int64_t units = longest ? p.n_docs : p.n_chunks;
  uint32_t* counts = nullptr;
  uint64_t* offsets = nullptr;
  ScanStatus* status = nullptr;
  AHA_CUDA(ws->get(kWsCounts, (size_t)units + 1, &counts));
  AHA_CUDA(ws->get(kWsOffsets, (size_t)units + 1, &offsets));
  AHA_CUDA(ws->get(kWsStatus, 4, &status));
  int32_t rc = alloc_doc_hit_off(c, p.n_docs, &p.doc_hit_off);
  if (rc) return rc;
  AHA_CUDA(cudaMemsetAsync(counts, 0, (units + 1) * sizeof(uint32_t), st));
  AHA_CUDA(cudaMemsetAsync(p.doc_hit_off, 0, (p.n_docs + 1) * sizeof(int64_t), st));
  AHA_CUDA(cudaMemsetAsync(status, 0, sizeof(ScanStatus), st));
  p.counts = counts;
  p.offsets = offsets;
  const int tpb = 128;
  // trie path: the count pass keeps its hits in a scratch slice per document (3 bytes of scratch per byte of text), so
  // that the walk runs once; not for inputs whose scratch would not be reasonable to ask for
  // ignore_case: the case-folded trie (one lookup per character) when the dictionary qualifies; knob 1 keeps the
  // depth-first walk with the frontier
  const bool folded = trie && (p.mode & AHA_IGNORE_CASE) && !knobs().force_v1 && ensure_folded(c.dfa, c.dd);
  aha_hit* doc_scratch = nullptr;
  if (trie && !(p.mode & AHA_COUNT_ONLY) && p.total > 0 && p.total <= ((int64_t)16 << 30)) {
    if (folded && c.dd->folded_max_chars <= 32) {
      // 16-byte records {first character, characters, folded node, case window}: which key stands behind the node is
      // settled by the compaction kernel, not inside the walk
      uint4* s16 = nullptr;
      AHA_CUDA(ws->get(kWsDocScratch, (size_t)((p.total >> 2) + 1), &s16));
      p.doc_scratch16 = s16;
      doc_scratch = reinterpret_cast<aha_hit*>(s16);  // (non-null: "the count pass records")
    } else {
      AHA_CUDA(ws->get(kWsDocScratch, (size_t)((p.total >> 2) + 1), &doc_scratch));
    }
  }
  p.doc_scratch = doc_scratch;
  AHA_CUDA(cudaEventRecord(ws->ev0, st));
  if (!c.validated) {
    validate_docs_kernel<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, p.total, status);
    count_launch();
  }
  if (chars && !longest && p.total > 0) {
    const uint64_t* lead_prefix = nullptr;
    rc = lead_prepass(c, p, p.chunk, p.n_chunks, &lead_prefix);
    if (rc) return rc;
    p.lead_prefix = lead_prefix;
  }
  if (folded) {
    p.fdelta = reinterpret_cast<const uint32_t*>(c.dd->folded_dev);
    p.fcand_off = reinterpret_cast<const uint32_t*>(c.dd->folded_dev + c.dd->off_fcand_off);
    p.fcand = reinterpret_cast<const uint4*>(c.dd->folded_dev + c.dd->off_fcand);
    p.fcmap = c.dd->folded_dev + c.dd->off_fcmap;
    p.frec = reinterpret_cast<const uint4*>(c.dd->folded_dev + c.dd->off_frec);
    p.fC = c.dd->folded_classes;
  }
  const unsigned grid = (unsigned)((units + tpb - 1) / tpb);
  if (units > 0) {
    if (folded) trie_folded_doc_kernel<false><<<grid, tpb, 0, st>>>(p);
    else if (trie) trie_longest_doc_kernel<false><<<grid, tpb, 0, st>>>(p);
    else if (longest) longest_doc_kernel<false><<<grid, tpb, 0, st>>>(p);
    else scan_v1_kernel<false><<<grid, tpb, 0, st>>>(p);
    count_launch();
  }
  rc = exclusive_sum_u32_to_u64(ws, counts, offsets, units + 1, st);
  if (rc) return rc;
  // these machines emit at most one hit per failure event: the count pass decides the size of the buffer
  unsigned long long total_hits = 0;
  AHA_CUDA(cudaMemcpyAsync(&ws->h_status->total_hits, offsets + units, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  AHA_CUDA(cudaStreamSynchronize(st));
  total_hits = ws->h_status->total_hits;
  c.res->n_hits = (int64_t)total_hits;
  if (!(p.mode & AHA_COUNT_ONLY)) {
    rc = alloc_hits(c, (int64_t)total_hits, &p.hits);
    if (rc) return rc;
  }
  if (units > 0) {
    if (longest) {
      if (trie && p.doc_scratch && p.hits) {
        if (p.doc_scratch16) trie_compact_folded_kernel<<<(unsigned)((p.n_docs + 7) / 8), 256, 0, st>>>(p);
        else trie_compact_kernel<<<(unsigned)((p.n_docs + 7) / 8), 256, 0, st>>>(p);
        count_launch();
      }
      if (folded) trie_folded_doc_kernel<true><<<grid, tpb, 0, st>>>(p);
      else if (trie) trie_longest_doc_kernel<true><<<grid, tpb, 0, st>>>(p);  // with a scratch: only documents that overflowed it
      else longest_doc_kernel<true><<<grid, tpb, 0, st>>>(p);
      count_launch();
      offsets_to_doc_hit_off<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(offsets, p.n_docs, p.doc_hit_off);
      count_launch();
    } else {
      scan_v1_kernel<true><<<grid, tpb, 0, st>>>(p);
      count_launch();
    }
  }
  AHA_CUDA(cudaEventRecord(ws->ev1, st));
  AHA_CUDA(cudaGetLastError());
  return read_status(c, status, ws->evw1);
}

// ACX#match_longest in parallel inside documents: speculative chunks (scan_longest.cuh)
static int32_t run_longest_chunks(ScanParams p, ScanCall& c) {
  cudaStream_t st = c.st;
  Workspace* ws = c.ws;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  p.chunk = kLongChunk;
  p.n_chunks = (p.total + kLongChunk - 1) / kLongChunk;
  const int64_t n = p.n_chunks;
  uint32_t* counts = nullptr;
  uint64_t* offsets = nullptr;
  ScanStatus* status = nullptr;
  LongParams q;
  memset(&q, 0, sizeof(q));
  LongState *xa = nullptr, *xb = nullptr;
  AHA_CUDA(ws->get(kWsCounts, (size_t)n + 1, &counts));
  AHA_CUDA(ws->get(kWsOffsets, (size_t)n + 1, &offsets));
  AHA_CUDA(ws->get(kWsStatus, 4, &status));
  AHA_CUDA(ws->get(kWsLongA, (size_t)n + 1, &xa));
  AHA_CUDA(ws->get(kWsLongB, (size_t)n + 1, &xb));
  AHA_CUDA(ws->get(kWsLongEntry, (size_t)n + 1, &q.entry_used));
  int32_t rc = alloc_doc_hit_off(c, p.n_docs, &p.doc_hit_off);
  if (rc) return rc;
  AHA_CUDA(cudaMemsetAsync(counts, 0, (n + 1) * sizeof(uint32_t), st));
  AHA_CUDA(cudaMemsetAsync(p.doc_hit_off, 0, (p.n_docs + 1) * sizeof(int64_t), st));
  AHA_CUDA(cudaMemsetAsync(status, 0, 2 * sizeof(ScanStatus), st));
  p.counts = counts;
  p.offsets = offsets;
  q.changed = &status[1].n_direct;
  AHA_CUDA(cudaEventRecord(ws->ev0, st));
  if (!c.validated) {
    validate_docs_kernel<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, p.total, status);
    count_launch();
  }
  if (chars && p.total > 0) {
    const uint64_t* lead_prefix = nullptr;
    rc = lead_prepass(c, p, kLongChunk, n, &lead_prefix);
    if (rc) return rc;
    p.lead_prefix = lead_prefix;
  }
  const unsigned grid = (unsigned)((n + 127) / 128);
  if (n > 0) {
    q.sp = p;
    q.exit_prev = xb;
    q.exit_next = xa;
    longest_chunk_kernel<0><<<grid, 128, 0, st>>>(q);
    count_launch();
    // settle: until a round finds every chunk walked from its predecessor's exit (one small read-back per round)
    for (int64_t round = 0; round <= n; ++round) {
      std::swap(q.exit_prev, q.exit_next);
      AHA_CUDA(cudaMemsetAsync(q.changed, 0, sizeof(unsigned int), st));
      longest_chunk_kernel<1><<<grid, 128, 0, st>>>(q);
      count_launch();
      AHA_CUDA(cudaMemcpyAsync(&ws->h_status->n_direct, q.changed, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
      AHA_CUDA(cudaStreamSynchronize(st));
      if (ws->h_status->n_direct == 0) break;
    }
  }
  rc = exclusive_sum_u32_to_u64(ws, counts, offsets, n + 1, st);
  if (rc) return rc;
  AHA_CUDA(cudaMemcpyAsync(&ws->h_status->total_hits, offsets + n, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  AHA_CUDA(cudaStreamSynchronize(st));
  const unsigned long long total_hits = ws->h_status->total_hits;
  c.res->n_hits = (int64_t)total_hits;
  if (!(p.mode & AHA_COUNT_ONLY)) {
    rc = alloc_hits(c, (int64_t)total_hits, &p.hits);
    if (rc) return rc;
  }
  if (n > 0) {
    q.sp = p;
    longest_chunk_kernel<2><<<grid, 128, 0, st>>>(q);
    count_launch();
  }
  // documents that start at the very end of the text (and the sentinel) when the last chunk did not reach them
  finalize_doc_off_kernel<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, p.total, reinterpret_cast<const unsigned long long*>(offsets + n), p.doc_hit_off);
  count_launch();
  AHA_CUDA(cudaEventRecord(ws->ev1, st));
  AHA_CUDA(cudaGetLastError());
  return read_status(c, status, ws->evw1);
}

static void fill_char_tables(CharDeviceTables* t, const HotTable& h) {
  t->root_g = reinterpret_cast<const uint16_t*>(h.dev + h.off_root);
  t->val_g = reinterpret_cast<const uint32_t*>(h.dev + h.off_val);
  t->check_g = reinterpret_cast<const uint16_t*>(h.dev + h.off_check);
  t->gfail = reinterpret_cast<const uint16_t*>(h.dev + h.off_gfail);
  t->ref_rec = reinterpret_cast<const RefRec*>(h.dev + h.off_ref);
  t->n_root = h.n_root;
  t->m = h.m;
  t->m_pad = h.m_pad;
  t->n_mwords = h.n_mwords;
  t->n_gfail_s = h.n_gfail_s;
  t->mbits_g = reinterpret_cast<const uint32_t*>(h.dev + h.off_mbits);
  for (int k = 0; k < 3; ++k) { t->lo[k] = h.lo[k]; t->n[k] = h.n[k]; t->off[k] = h.off[k]; }
}

// character automaton: walk -> prefix over tiles -> expansion at full occupancy (scan_char_split.cuh)
static int32_t run_char_split(ScanParams p, ScanCall& c) {
  cudaStream_t st = c.st;
  Workspace* ws = c.ws;
  const int n_sm = c.n_sm;
  SplitParams sp;
  memset(&sp, 0, sizeof(sp));
  CharParams& cp = sp.c;
  FusedParams& fp = cp.f;
  fused_geometry(p.warm + 1, true, &fp.chunk, &fp.w16, &fp.buf_bytes);
  fp.tile = 32 * fp.chunk;
  fp.n_tiles = (p.total + fp.tile - 1) / fp.tile;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const bool count_only = (p.mode & AHA_COUNT_ONLY) != 0;
  const bool sep = (p.mode & AHA_SEP) != 0;  // events carry byte positions; character offsets are counted by the expansion
  const HotTable& h = c.dd->hot;
  fill_char_tables(&cp.t, h);
  cp.cmap_g = p.classmap;
  const int64_t nt = fp.n_tiles;
  const size_t nt1 = (size_t)(nt ? nt : 1);
  TileDoc* tile_doc = nullptr;
  uint64_t* tile_prefix = nullptr;
  AHA_CUDA(ws->get(kWsStore, nt1 * 32 * kStoreSlots, &sp.store));
  AHA_CUDA(ws->get(kWsTileMeta, nt1 * 32, &sp.tile_meta));
  AHA_CUDA(ws->get(kWsTileHits, (size_t)nt + 1, &sp.tile_hits));
  AHA_CUDA(ws->get(kWsTileDoc, nt1, &tile_doc));
  AHA_CUDA(ws->get(kWsDirectList, nt1, &sp.direct_list));
  AHA_CUDA(ws->get(kWsTilePrefix, (size_t)nt + 1, &tile_prefix));
  AHA_CUDA(ws->get(kWsStatus, 4, &sp.status));
  sp.tile_doc = tile_doc;
  sp.tile_prefix = tile_prefix;
  sp.mbits_g = reinterpret_cast<const uint32_t*>(h.dev + h.off_mbits);
  sp.ref8_g = reinterpret_cast<const uint2*>(h.dev + h.off_ref8);
  sp.n_ref8 = h.n_ref8;
  sp.out8_g = reinterpret_cast<const uint2*>(h.dev + h.off_out8);
  int32_t rc = alloc_doc_hit_off(c, p.n_docs, &p.doc_hit_off);
  if (rc) return rc;
  int64_t cap = 0;
  if (!count_only) {
    cap = estimate_cap(c, p.total);
    rc = alloc_hits(c, cap, &p.hits);
    if (rc) return rc;
  }
  fp.hit_cap = cap;
  const size_t smem = ((h.smem_bytes + 127) & ~size_t(127)) + (size_t)kFusedWarps * (fp.buf_bytes + 128);
  static std::once_flag attr_once[64];
  cudaError_t attr_err = cudaSuccess;
  std::call_once(attr_once[c.dd->device & 63], [&] {
    attr_err = cudaFuncSetAttribute(char_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(char_walk_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(char_walk_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  });
  AHA_CUDA(attr_err);

  AHA_CUDA(cudaEventRecord(ws->ev0, st));
  AHA_CUDA(cudaMemsetAsync(sp.status, 0, sizeof(ScanStatus), st));
  AHA_CUDA(cudaMemsetAsync(sp.tile_hits + nt, 0, sizeof(uint32_t), st));
  AHA_CUDA(cudaMemsetAsync(p.doc_hit_off, 0, (p.n_docs + 1) * sizeof(int64_t), st));
  if (!c.validated) {
    validate_docs_kernel<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, p.total, sp.status);
    count_launch();
  }
  if (chars && p.total > 0) {
    rc = lead_prepass(c, p, fp.tile, nt, &fp.tile_lead);
    if (rc) return rc;
  }
  fp.sp = p;
  if (nt > 0) {
    tile_doc_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, fp.tile, nt, tile_doc);
    count_launch();
    AHA_CUDA(cudaEventRecord(ws->evw0, st));
    if (sep) char_walk_kernel<true><<<n_sm, kFusedWarps * 32, smem, st>>>(sp);
    else char_walk_kernel<false><<<n_sm, kFusedWarps * 32, smem, st>>>(sp);
    count_launch();
    AHA_CUDA(cudaEventRecord(ws->evw1, st));
    AHA_CUDA(cudaGetLastError());
  }
#if !AHA_W_COUNT
  if (nt > 0) {
    if (sep) {  // the vetoes decide how many hits an event has: counted by the expansion's own loop
      const unsigned grid = (unsigned)std::min<int64_t>((nt + 7) / 8, (int64_t)n_sm * 64);
      char_expand_kernel<true, true><<<grid, 256, 0, st>>>(sp);
    } else {
      const unsigned grid = (unsigned)std::min<int64_t>((nt + 7) / 8, (int64_t)n_sm * 8);
      char_count_kernel<<<grid, 256, (size_t)h.n_mwords * 16, st>>>(sp);  // one byte per ref (<= 64 KB)
    }
    count_launch();
  }
#endif
  {
    cub::TransformInputIterator<uint64_t, TileHitsToU64, const uint32_t*> it(sp.tile_hits, TileHitsToU64());
    rc = exclusive_sum_u64(ws, it, tile_prefix, nt + 1, st);
    if (rc) return rc;
  }
  auto deliver = [&](bool timed) -> int32_t {
    if (nt > 0) {
      const unsigned grid = (unsigned)std::min<int64_t>((nt + 7) / 8, (int64_t)n_sm * 64);
      if (timed) AHA_CUDA(cudaEventRecord(ws->evx0, st));
      if (sep) char_expand_kernel<true, false><<<grid, 256, 0, st>>>(sp);
      else char_expand_kernel<false, false><<<grid, 256, 0, st>>>(sp);
      count_launch();
      if (timed) AHA_CUDA(cudaEventRecord(ws->evx1, st));
      // tiles the walk handed over (normally none: the kernel reads the list length on the device and leaves)
      const unsigned dgrid = (unsigned)std::min<int64_t>((nt + 3) / 4, (int64_t)n_sm * 8);
      char_direct_kernel<<<dgrid, 128, 256 + 4 * (size_t)fp.buf_bytes, st>>>(sp);
      count_launch();
    }
    finalize_split_kernel<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, p.total, tile_prefix + nt,
                                                                                   p.doc_hit_off, sp.status);
    count_launch();
    AHA_CUDA(cudaEventRecord(ws->ev1, st));
    AHA_CUDA(cudaGetLastError());
    return AHA_OK;
  };
  rc = deliver(true);
  if (rc) return rc;
  rc = read_status(c, sp.status, ws->ev1);
  if (rc) return rc;
  unsigned long long total_hits = ws->h_status->total_hits;
  if (!count_only && (int64_t)total_hits > cap) {
    // denser than any batch before it: the events are still in the store, only the expansion runs again
    cap = (int64_t)total_hits;
    rc = alloc_hits(c, cap, &p.hits);
    if (rc) return rc;
    sp.c.f.sp.hits = p.hits;
    sp.c.f.hit_cap = cap;
    rc = deliver(false);
    if (rc) return rc;
    rc = read_status(c, sp.status, ws->ev1);
    if (rc) return rc;
    total_hits = ws->h_status->total_hits;
  }
  note_density(c, p.total, total_hits);
  c.res->n_hits = (int64_t)total_hits;
  if (nt > 0) {
    float ms = 0;
    AHA_CUDA(cudaEventElapsedTime(&ms, ws->evw0, ws->evw1));
    c.res->walk_ms = ms;
    AHA_CUDA(cudaEventElapsedTime(&ms, ws->evx0, ws->evx1));
    c.res->expand_ms = ms;
  }
  return AHA_OK;
}

// single-pass fused kernel with the byte automaton in global memory / L2: see scan_fused.cuh
template <class Policy>
static int32_t launch_fused(const FusedParams& fp, const Policy& pol, int n_sm, size_t smem, cudaStream_t st) {
  // event mode (lists expanded after the walk) for everything but sep scans, whose neighbour tests sit in emit_list
  if ((fp.sp.mode & AHA_SEP) || !fp.ev_scratch) {
    AHA_CUDA(cudaFuncSetAttribute(scan_fused_kernel<Policy, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    scan_fused_kernel<Policy, false><<<n_sm, kByteWarps * 32, smem, st>>>(fp, pol);
  } else if (fp.span) {
    scan_fused_kernel<Policy, true, true><<<n_sm, kSpanWarps * 32, smem, st>>>(fp, pol);
  } else {
    AHA_CUDA(cudaFuncSetAttribute(scan_fused_kernel<Policy, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    scan_fused_kernel<Policy, true><<<n_sm, kByteWarps * 32, smem, st>>>(fp, pol);
  }
  count_launch();
  AHA_CUDA(cudaGetLastError());
  return AHA_OK;
}

static int32_t run_fused(ScanParams p, ScanCall& c) {
  cudaStream_t st = c.st;
  Workspace* ws = c.ws;
  const int n_sm = c.n_sm;
  FusedParams fp;
  memset(&fp, 0, sizeof(fp));
  fused_geometry(p.warm + 1, false, &fp.chunk, &fp.w16, &fp.buf_bytes);
  fp.ev_cap = kByteEvCap;
  // Span mode (event delivery, documents that are not tiny): no staging, every lane reads a span of kSpanBytes from
  // global memory; the warm-up in front of a span is then a few percent of the walk, and no shared memory is taken from L1.
  // Tiny documents keep the staged tiles (a span would hold more document starts than a lane has marker slots).
  if (!(p.mode & AHA_SEP) && knobs().byte_span && p.n_docs > 0 && p.total / p.n_docs >= 4 * knobs().byte_span) {
    fp.span = 1;
    fp.chunk = knobs().byte_span;
    fp.buf_bytes = 0;
    fp.ev_cap = std::max(kByteEvCap, fp.chunk / 4);
  }
  fp.tile = 32 * fp.chunk;
  fp.n_tiles = (p.total + fp.tile - 1) / fp.tile;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const bool count_only = (p.mode & AHA_COUNT_ONLY) != 0;
  const int64_t n_warps = (int64_t)n_sm * std::max(kByteWarps, kSpanWarps);
  unsigned long long* tile_state = nullptr;
  unsigned int* counters = nullptr;  // ScanStatus {total_hits, n_direct, bad_docs} | [4] ticket
  const int64_t n_groups = (fp.n_tiles + 31) / 32;
  const size_t state_words = (size_t)(fp.n_tiles + 1) + (size_t)(n_groups + 1);
  AHA_CUDA(ws->get(kWsTileState, state_words, &tile_state));
  AHA_CUDA(ws->get(kWsStatus, 8, &counters));
  if (p.mode & AHA_SEP) AHA_CUDA(ws->get(kWsScratch, (size_t)n_warps * 32 * kLaneCap, &fp.scratch));
  else AHA_CUDA(ws->get(kWsEvScratch, (size_t)n_warps * 32 * (size_t)fp.ev_cap, &fp.ev_scratch));
  AHA_CUDA(ws->get(kWsMarks, (size_t)n_warps * 2 * 32 * kMarkCap, &fp.marks));
  int32_t rc = alloc_doc_hit_off(c, p.n_docs, &p.doc_hit_off);
  if (rc) return rc;
  ScanStatus* status = reinterpret_cast<ScanStatus*>(counters);
  fp.tile_state = tile_state;
  fp.group_excl = tile_state + (fp.n_tiles + 1);
  fp.total_hits = &status->total_hits;
  fp.n_direct = &status->n_direct;
  fp.ticket = counters + 4;
  ByteGlobal pol;
  pol.cmap_g = p.classmap;
  pol.cmap = nullptr;
  const size_t smem = ((ByteGlobal::table_bytes(pol) + 127) & ~127u) + (size_t)kByteWarps * (fp.buf_bytes + 128);
  int64_t cap = 0;
  if (!count_only) {
    cap = estimate_cap(c, p.total);
    rc = alloc_hits(c, cap, &p.hits);
    if (rc) return rc;
  }
  AHA_CUDA(cudaEventRecord(ws->ev0, st));
  AHA_CUDA(cudaMemsetAsync(counters, 0, 8 * sizeof(unsigned int), st));
  if (!c.validated) {
    validate_docs_kernel<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, p.total, status);
    count_launch();
  }
  if (chars && p.total > 0) {
    rc = lead_prepass(c, p, fp.tile, fp.n_tiles, &fp.tile_lead);
    if (rc) return rc;
  }
  unsigned long long total_hits = 0;
  for (int attempt = 0; attempt < 2; ++attempt) {
    fp.hit_cap = cap;
    fp.sp = p;
    AHA_CUDA(cudaMemsetAsync(tile_state, 0, state_words * sizeof(unsigned long long), st));
    if (attempt) AHA_CUDA(cudaMemsetAsync(counters, 0, 8 * sizeof(unsigned int), st));
    AHA_CUDA(cudaMemsetAsync(p.doc_hit_off, 0, (p.n_docs + 1) * sizeof(int64_t), st));
    if (fp.n_tiles > 0) {
      rc = launch_fused(fp, pol, n_sm, smem, st);
      if (rc) return rc;
    }
    finalize_doc_off_kernel<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, p.total, fp.total_hits,
                                                                                 p.doc_hit_off);
    count_launch();
    AHA_CUDA(cudaEventRecord(ws->ev1, st));
    rc = read_status(c, status, ws->ev1);
    if (rc) return rc;
#ifdef AHA_B200_PHASES
    {
      unsigned long long ph[8], zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      cudaMemcpyFromSymbol(ph, g_phase_cycles, sizeof(ph));
      cudaMemcpyToSymbol(g_phase_cycles, zero, sizeof(zero));
      double sum = 0;
      for (int i = 0; i < 5; ++i) sum += (double)ph[i];
      fprintf(stderr, "phases (share of warp cycles): stage %.3f walk %.3f count %.3f look-back %.3f deliver %.3f  [%.0f Mcycles per warp]\n",
              ph[0] / sum, ph[1] / sum, ph[2] / sum, ph[3] / sum, ph[4] / sum, sum / ((double)n_warps * 1e6));
    }
#endif
    total_hits = ws->h_status->total_hits;
    if (count_only || (int64_t)total_hits <= cap) break;
    if (attempt == 1) return set_err(AHA_ERR_STATE, "hit buffer sizing did not converge");
    cap = (int64_t)total_hits;  // denser than any batch before it: once more with the exact size
    rc = alloc_hits(c, cap, &p.hits);
    if (rc) return rc;
  }
  note_density(c, p.total, total_hits);
  c.res->n_hits = (int64_t)total_hits;
  return AHA_OK;
}

static int32_t scan_on_device(const aha_dfa* d, const void* dev_text, const void* dev_doc_off, int64_t n_docs,
                              int64_t total, uint32_t mode, const uint8_t* sep_bits, int32_t sep_size, cudaStream_t st,
                              aha_hits* res, HitsOwner* own, bool validated) {
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  DeviceDfa* dd = d->on(dev);
  if (!dd) return set_err(AHA_ERR_STATE, "automaton was not uploaded to the current device (aha_b200_init)");
  if ((mode & AHA_SEP) && (sep_size > 256 || sep_size < 0 || !sep_bits))
    return set_err(AHA_ERR_ARG, "sep BitArray size > 256 is not supported");  // ac.cr:322
  if ((mode & AHA_TRIE_LONGEST) && dd->view.max_key_len > (uint32_t)kTrieMaxChars)
    return set_err(AHA_ERR_ARG, "AHA_TRIE_LONGEST supports keys of at most 96 bytes");
  own->device = dev;
  own->stream = st;
  std::unique_ptr<Workspace> ws;
  int32_t rc = ws_acquire(d, dd, &ws);
  if (rc) return rc;
  ScanCall c;
  c.dfa = d; c.dd = dd; c.ws = ws.get(); c.st = st; c.res = res; c.own = own; c.validated = validated;
  static int sm_count[64] = {0};
  if (!sm_count[dev & 63]) AHA_CUDA(cudaDeviceGetAttribute(&sm_count[dev & 63], cudaDevAttrMultiProcessorCount, dev));
  c.n_sm = sm_count[dev & 63];
  // the bulk copies of the tiled kernels need a 16-byte aligned text: a misaligned device pointer is copied once
  if ((reinterpret_cast<uintptr_t>(dev_text) & 15) != 0 && total > 0) {
    uint8_t* aligned = nullptr;
    cudaError_t e = ws->get(kWsAligned, (size_t)total + 64, &aligned);
    if (e == cudaSuccess) e = cudaMemcpyAsync(aligned, dev_text, (size_t)total, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) {
      ws_release(d, dd, std::move(ws));
      return set_err(AHA_ERR_CUDA, std::string("aligned text copy: ") + cudaGetErrorString(e));
    }
    dev_text = aligned;
  }
  ScanParams p;
  fill_params(&p, dd, dev_text, dev_doc_off, n_docs, total, mode, sep_bits, sep_size);
  const Knobs& k = knobs();
  const bool tiled_ok = !k.force_v1 && !(mode & (AHA_LONGEST | AHA_TRIE_LONGEST)) && dd->view.max_key_len <= 1024;
  // the code-point form of the neighbour test (Array(Char) overloads) stays on the byte kernel
  const bool sep_chars = (mode & AHA_SEP) && (mode & AHA_SEP_CODEPOINT);
  const bool split = tiled_ok && dd->hot.ok && !k.no_char && !sep_chars;
  // match_longest: speculative chunks inside documents (knob 1 keeps the one-thread-per-document machine)
  const bool long_chunks = (mode & AHA_LONGEST) && !(mode & AHA_TRIE_LONGEST) && !k.force_v1;
  if (split) rc = run_char_split(p, c);
  else if (tiled_ok) rc = run_fused(p, c);
  else if (long_chunks) rc = run_longest_chunks(p, c);
  else rc = run_v1(p, c);
  if (rc == AHA_OK) {
    float ms = 0;
    cudaError_t e = cudaEventElapsedTime(&ms, ws->ev0, ws->ev1);
    if (e != cudaSuccess) rc = set_err(AHA_ERR_CUDA, std::string("cudaEventElapsedTime: ") + cudaGetErrorString(e));
    res->scan_ms = ms;
  } else {
    cudaStreamSynchronize(st);  // nothing of this call may still be running on the workspace when it goes back
    cudaGetLastError();
  }
  ws_release(d, dd, std::move(ws));
  return rc;
}

static void free_owner(HitsOwner* own);

}  // namespace aha

extern "C" {

int32_t aha_b200_scan_device(const aha_dfa* d, const void* dev_text, const void* dev_doc_off, int64_t n_docs,
                             int64_t total_bytes, uint32_t mode, const uint8_t* sep_bits, int32_t sep_size,
                             void* stream, aha_hits** out) {
  if (!d || !out || n_docs < 0 || total_bytes < 0 || (!dev_doc_off)) return set_err(AHA_ERR_ARG, "null argument");
  std::unique_ptr<aha_hits> res(new aha_hits());
  memset(res.get(), 0, sizeof(aha_hits));
  HitsOwner* own = new HitsOwner();
  res->opaque = own;
  res->n_docs = n_docs;
  res->on_device = 1;
  int32_t rc = scan_on_device(d, dev_text, dev_doc_off, n_docs, total_bytes, mode, sep_bits, sep_size,
                              static_cast<cudaStream_t>(stream), res.get(), own, false);
  if (rc) {
    aha_b200_hits_free(res.release());
    return rc;
  }
  res->device = own->device;
  res->hits = static_cast<aha_hit*>(own->dev_hits);
  res->doc_hit_off = static_cast<int64_t*>(own->dev_doc_hit_off);
  *out = res.release();
  return AHA_OK;
}

}  // extern "C"

// ---- pinned host buffers: cudaMallocHost costs ~100 ms per GB, so result buffers are recycled --------------
namespace aha {
struct PinnedPool {
  std::mutex mu;
  std::vector<std::pair<void*, size_t>> free_list;
  size_t cached = 0;
  static constexpr size_t kMaxCached = (size_t)16 << 30;
  void* acquire(size_t bytes, size_t* got) {
    bytes = (bytes + 4095) & ~size_t(4095);
    {
      std::lock_guard<std::mutex> lk(mu);
      size_t best = free_list.size();
      for (size_t i = 0; i < free_list.size(); ++i)
        if (free_list[i].second >= bytes && (best == free_list.size() || free_list[i].second < free_list[best].second)) best = i;
      if (best != free_list.size() && free_list[best].second <= 2 * bytes + (1 << 20)) {
        void* p = free_list[best].first;
        *got = free_list[best].second;
        cached -= *got;
        free_list.erase(free_list.begin() + best);
        return p;
      }
    }
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) return nullptr;
    *got = bytes;
    return p;
  }
  void release(void* p, size_t bytes) {
    if (!p) return;
    std::lock_guard<std::mutex> lk(mu);
    if (cached + bytes > kMaxCached) {
      cudaFreeHost(p);
      return;
    }
    free_list.push_back({p, bytes});
    cached += bytes;
  }
};
static PinnedPool g_pinned;

static void free_owner(HitsOwner* own) {
  if (!own) return;
  DeviceGuard guard;
  cudaSetDevice(own->device);
  DevCache& cache = dev_cache(own->device);
  if (own->dev_hits) cache.release(own->dev_hits, own->dev_hits_bytes, own->stream);
  if (own->dev_doc_hit_off) cache.release(own->dev_doc_hit_off, own->dev_doc_hit_off_bytes, own->stream);
  if (own->host_hits) g_pinned.release(own->host_hits, own->host_hits_bytes);
  if (own->host_doc_hit_off) g_pinned.release(own->host_doc_hit_off, own->host_doc_hit_off_bytes);
  delete own;
}

// ---- host pipeline: one context per device, kept between calls ------------------------------------------------------
struct E2ECtx {
  int device = -1;
  cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
  cudaEvent_t e_in[2] = {nullptr, nullptr};
  DevBuf d_text[2], d_off[2];
  int64_t* h_off_stage = nullptr;  // pinned, rebased doc_off of the runs in flight
  size_t h_off_cap = 0;            // entries per buffer
  cudaError_t init(int dev) {
    device = dev;
    cudaError_t e;
    if ((e = cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking)) != cudaSuccess) return e;
    if ((e = cudaStreamCreateWithFlags(&s_run, cudaStreamNonBlocking)) != cudaSuccess) return e;
    if ((e = cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking)) != cudaSuccess) return e;
    for (int b = 0; b < 2; ++b)
      if ((e = cudaEventCreateWithFlags(&e_in[b], cudaEventDisableTiming)) != cudaSuccess) return e;
    return cudaSuccess;
  }
  cudaError_t ensure(int64_t max_bytes, int64_t max_docs) {
    cudaError_t e;
    for (int b = 0; b < 2; ++b) {
      if ((e = d_text[b].ensure((size_t)max_bytes + 64)) != cudaSuccess) return e;
      if ((e = d_off[b].ensure((size_t)(max_docs + 1) * sizeof(int64_t))) != cudaSuccess) return e;
    }
    if ((size_t)(max_docs + 1) > h_off_cap) {
      if (h_off_stage) cudaFreeHost(h_off_stage);
      h_off_stage = nullptr;
      h_off_cap = (size_t)(max_docs + 1) + (size_t)(max_docs + 1) / 4;
      if ((e = cudaMallocHost(&h_off_stage, 2 * h_off_cap * sizeof(int64_t))) != cudaSuccess) return e;
    }
    return cudaSuccess;
  }
  ~E2ECtx() {
    if (device < 0) return;
    DeviceGuard g;
    cudaSetDevice(device);
    for (int b = 0; b < 2; ++b) {
      d_text[b].release();
      d_off[b].release();
      if (e_in[b]) cudaEventDestroy(e_in[b]);
    }
    if (h_off_stage) cudaFreeHost(h_off_stage);
    for (cudaStream_t s : {s_in, s_run, s_out})
      if (s) cudaStreamDestroy(s);
  }
};
static std::mutex g_e2e_mu;
static std::vector<std::unique_ptr<E2ECtx>> g_e2e_idle;
static std::unique_ptr<E2ECtx> e2e_acquire(int device, cudaError_t* err) {
  {
    std::lock_guard<std::mutex> lk(g_e2e_mu);
    for (size_t i = 0; i < g_e2e_idle.size(); ++i)
      if (g_e2e_idle[i]->device == device) {
        std::unique_ptr<E2ECtx> c = std::move(g_e2e_idle[i]);
        g_e2e_idle.erase(g_e2e_idle.begin() + i);
        return c;
      }
  }
  std::unique_ptr<E2ECtx> c(new E2ECtx());
  *err = c->init(device);
  return c;
}
static void e2e_release(std::unique_ptr<E2ECtx> c) {
  std::lock_guard<std::mutex> lk(g_e2e_mu);
  g_e2e_idle.push_back(std::move(c));
}

// What the device workers of one aha_b200_scan call share.  The batch is cut into runs of whole documents; run k is
// scanned on device k mod G.  A run's hits go straight to their final place in ONE pinned buffer: the place is the
// sum of the hit counts of the runs before it, which their workers publish as soon as their scans have counted.
struct E2EShared {
  const aha_dfa* d;
  const uint8_t* text;
  const int64_t* doc_off;
  uint32_t mode;
  const uint8_t* sep_bits;
  int32_t sep_size;
  std::vector<int64_t> cut;            // run k = documents [cut[k], cut[k+1])
  std::vector<int> devices;
  int64_t* h_doc_hit_off = nullptr;    // pinned, n_docs + 1
  std::mutex mu;
  std::condition_variable cv;
  std::vector<int64_t> run_hits;       // -1 until the run has been counted
  std::vector<int64_t> run_base;       // first hit of the run in the stream
  size_t known = 0;                    // runs [0, known) have their base
  aha_hit* hits = nullptr;             // primary pinned buffer
  size_t hits_bytes = 0;
  int64_t hits_cap = -1;               // records; -1 until run 0 has been counted
  std::vector<std::pair<void*, size_t>> side;  // per run: pinned overflow block (hits that did not fit the primary buffer)
  int32_t rc = AHA_OK;                 // first error
  std::string err;
  double scan_ms = 0, walk_ms = 0, expand_ms = 0;
  bool failed() {
    std::lock_guard<std::mutex> lk(mu);
    return rc != AHA_OK;
  }
  void fail(int32_t code, const std::string& msg) {
    std::lock_guard<std::mutex> lk(mu);
    if (rc == AHA_OK) { rc = code; err = msg; }
    cv.notify_all();
  }
};

static void e2e_worker(E2EShared* S, int wi) {
  const int G = (int)S->devices.size();
  const int dev = S->devices[wi];
  const size_t n_runs = S->cut.size() - 1;
  if ((size_t)wi >= n_runs) return;
  cudaError_t ce = cudaSetDevice(dev);
  if (ce != cudaSuccess) return S->fail(AHA_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(ce));
  std::unique_ptr<E2ECtx> ctx = e2e_acquire(dev, &ce);
  auto bail = [&](int32_t code, const std::string& msg) {
    S->fail(code, msg);
    cudaStreamSynchronize(ctx->s_in);
    cudaStreamSynchronize(ctx->s_out);
    e2e_release(std::move(ctx));
  };
  if (ce != cudaSuccess) return bail(AHA_ERR_CUDA, std::string("e2e context: ") + cudaGetErrorString(ce));
  int64_t max_bytes = 16, max_docs = 1;
  for (size_t k = wi; k < n_runs; k += G) {
    max_bytes = std::max<int64_t>(max_bytes, S->doc_off[S->cut[k + 1]] - S->doc_off[S->cut[k]]);
    max_docs = std::max<int64_t>(max_docs, S->cut[k + 1] - S->cut[k]);
  }
  if ((ce = ctx->ensure(max_bytes, max_docs)) != cudaSuccess) return bail(AHA_ERR_CUDA, std::string("e2e buffers: ") + cudaGetErrorString(ce));
  const bool count_only = (S->mode & AHA_COUNT_ONLY) != 0;
  auto stage_in = [&](size_t k, int b) -> cudaError_t {
    const int64_t lo = S->cut[k], hi = S->cut[k + 1];
    const int64_t base = S->doc_off[lo], bytes = S->doc_off[hi] - base;
    int64_t* ho = ctx->h_off_stage + (size_t)b * ctx->h_off_cap;
    for (int64_t i = lo; i <= hi; ++i) ho[i - lo] = S->doc_off[i] - base;
    cudaError_t e = cudaSuccess;
    if (bytes > 0) e = cudaMemcpyAsync(ctx->d_text[b].p, S->text + base, (size_t)bytes, cudaMemcpyHostToDevice, ctx->s_in);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ctx->d_off[b].p, ho, (size_t)(hi - lo + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->s_in);
    if (e == cudaSuccess) e = cudaEventRecord(ctx->e_in[b], ctx->s_in);
    return e;
  };
  double scan_ms = 0, walk_ms = 0, expand_ms = 0;
  int b = 0;
  if ((ce = stage_in(wi, 0)) != cudaSuccess) return bail(AHA_ERR_CUDA, std::string("H2D: ") + cudaGetErrorString(ce));
  for (size_t k = wi; k < n_runs; k += G, b ^= 1) {
    if (S->failed()) break;
    const int64_t lo = S->cut[k], hi = S->cut[k + 1];
    const int64_t bytes = S->doc_off[hi] - S->doc_off[lo];
    if ((ce = cudaStreamWaitEvent(ctx->s_run, ctx->e_in[b], 0)) != cudaSuccess) return bail(AHA_ERR_CUDA, cudaGetErrorString(ce));
    // the other text buffer was released by this worker's previous scan (scan_on_device returns when its kernels are
    // done): the copy of the next run overlaps this scan, the hits of the previous run are still going home on s_out
    if (k + G < n_runs && (ce = stage_in(k + G, b ^ 1)) != cudaSuccess) return bail(AHA_ERR_CUDA, std::string("H2D: ") + cudaGetErrorString(ce));
    aha_hits part;
    memset(&part, 0, sizeof(part));
    HitsOwner pown;
    const int32_t rc = scan_on_device(S->d, ctx->d_text[b].p, ctx->d_off[b].p, hi - lo, bytes, S->mode, S->sep_bits, S->sep_size, ctx->s_run,
                                      &part, &pown, true);
    DevCache& cache = dev_cache(dev);
    auto drop = [&] {
      if (pown.dev_hits) cache.release(pown.dev_hits, pown.dev_hits_bytes, ctx->s_out);
      if (pown.dev_doc_hit_off) cache.release(pown.dev_doc_hit_off, pown.dev_doc_hit_off_bytes, ctx->s_out);
    };
    if (rc) {
      drop();
      return bail(rc, g_err);
    }
    scan_ms += part.scan_ms;
    walk_ms += part.walk_ms;
    expand_ms += part.expand_ms;
    // publish the count, then wait for the place: the counts of all earlier runs
    int64_t base = 0;
    aha_hit* dst = nullptr;
    {
      std::unique_lock<std::mutex> lk(S->mu);
      S->run_hits[k] = part.n_hits;
      while (S->known < n_runs && S->run_hits[S->known] >= 0) {
        S->run_base[S->known + 1] = S->run_base[S->known] + S->run_hits[S->known];
        ++S->known;
      }
      S->cv.notify_all();
      S->cv.wait(lk, [&] { return S->known > k || S->rc != AHA_OK; });
      if (S->rc != AHA_OK) {
        lk.unlock();
        drop();
        break;
      }
      base = S->run_base[k];
      if (!count_only && part.n_hits > 0) {
        if (S->hits_cap < 0) {
          // first sized when run 0 is known (every run waits for run 0's count anyway): the batch at that density
          const int64_t total = S->doc_off[S->cut[n_runs]];
          const double dens = (double)S->run_base[S->known] / (double)std::max<int64_t>(1, S->doc_off[S->cut[S->known]]);
          int64_t want = (int64_t)((double)total * dens * 1.2) + 4096;
          if (S->known == n_runs) want = S->run_base[n_runs];
          size_t got = 0;
          S->hits = static_cast<aha_hit*>(g_pinned.acquire((size_t)std::max<int64_t>(want, 1) * sizeof(aha_hit), &got));
          S->hits_bytes = got;
          S->hits_cap = S->hits ? (int64_t)(got / sizeof(aha_hit)) : 0;
        }
        if (S->hits && base + part.n_hits <= S->hits_cap) {
          dst = S->hits + base;
        } else {
          size_t got = 0;
          void* blk = g_pinned.acquire((size_t)part.n_hits * sizeof(aha_hit), &got);
          S->side[k] = {blk, got};
          dst = static_cast<aha_hit*>(blk);
        }
      }
    }
    if (!count_only && part.n_hits > 0) {
      if (!dst) {
        drop();
        return bail(AHA_ERR_NOMEM, "pinned allocation failed");
      }
      ce = cudaMemcpyAsync(dst, pown.dev_hits, (size_t)part.n_hits * sizeof(aha_hit), cudaMemcpyDeviceToHost, ctx->s_out);
      if (ce != cudaSuccess) { drop(); return bail(AHA_ERR_CUDA, std::string("D2H: ") + cudaGetErrorString(ce)); }
    }
    ce = cudaMemcpyAsync(S->h_doc_hit_off + lo, pown.dev_doc_hit_off, (size_t)(hi - lo) * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->s_out);
    if (ce != cudaSuccess) { drop(); return bail(AHA_ERR_CUDA, std::string("D2H: ") + cudaGetErrorString(ce)); }
    drop();  // ordered behind the copies on s_out
  }
  cudaStreamSynchronize(ctx->s_in);
  ce = cudaStreamSynchronize(ctx->s_out);
  if (ce != cudaSuccess) S->fail(AHA_ERR_CUDA, std::string("D2H: ") + cudaGetErrorString(ce));
  {
    std::lock_guard<std::mutex> lk(S->mu);
    S->scan_ms = std::max(S->scan_ms, scan_ms);
    S->walk_ms = std::max(S->walk_ms, walk_ms);
    S->expand_ms = std::max(S->expand_ms, expand_ms);
  }
  e2e_release(std::move(ctx));
}
}  // namespace aha

extern "C" {

// Host buffers in, host buffers out.  The batch is cut into runs of whole documents of ~64 MB, dealt round-robin to
// the devices of aha_b200_init (one host thread and three streams per device): the H2D copy of a worker's next run
// and the D2H copy of its previous run's hits overlap the scan of the current one, so that on a PCIe-attached GPU the
// call is bound by the slower of the two copy directions, not by their sum.
int32_t aha_b200_scan(const aha_dfa* d, const uint8_t* text, const int64_t* doc_off, int64_t n_docs, uint32_t mode,
                      const uint8_t* sep_bits, int32_t sep_size, aha_hits** out) {
  if (!d || !out || n_docs < 0 || !doc_off || (!text && n_docs > 0 && doc_off[n_docs] > 0))
    return set_err(AHA_ERR_ARG, "null argument");
  if (doc_off[0] != 0) return set_err(AHA_ERR_ARG, "doc_off[0] must be 0");
  for (int64_t i = 0; i < n_docs; ++i) {
    const int64_t len = doc_off[i + 1] - doc_off[i];
    if (len < 0) return set_err(AHA_ERR_ARG, "doc_off must be non-decreasing");
    if (len > INT32_MAX) return set_err(AHA_ERR_ARG, "document >= 2^31 bytes: Aha::Hit offsets are Int32");
  }
  if ((mode & AHA_SEP) && (sep_size > 256 || sep_size < 0 || !sep_bits))
    return set_err(AHA_ERR_ARG, "sep BitArray size > 256 is not supported");  // ac.cr:322
  int32_t rc = ensure_devices();
  if (rc) return rc;
  DeviceGuard guard;
  E2EShared S;
  S.d = d; S.text = text; S.doc_off = doc_off; S.mode = mode; S.sep_bits = sep_bits; S.sep_size = sep_size;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    for (int dev : g_devices)
      if (d->on(dev)) S.devices.push_back(dev);
  }
  if (S.devices.empty()) return set_err(AHA_ERR_STATE, "automaton was not uploaded to any selected device (aha_b200_init)");
  // runs of whole documents
  S.cut.assign(1, 0);
  {
    int64_t target = (int64_t)knobs().e2e_chunk_mb << 20;
    // the trie-longest machine is one thread per document: a run must hold enough documents to fill the device
    // (a 64 MB run of 4 KiB documents is 16 k threads on a GPU that holds 300 k)
    if ((mode & AHA_TRIE_LONGEST) && n_docs > 0 && knobs().e2e_chunk_mb >= 16) {
      const int64_t avg = std::max<int64_t>(1, doc_off[n_docs] / n_docs);
      static const int64_t run_docs = [] {  // AHA_B200_TRIE_RUN_DOCS: documents per run of the per-document machine
        const char* e = getenv("AHA_B200_TRIE_RUN_DOCS");
        const long long v = e ? atoll(e) : 0;
        return (int64_t)(v > 0 ? v : ((long long)1 << 16));  // measured on C4 (2 GiB, e2e GB/s): 32 k 26.4, 64 k 34.9, 128 k 33.5, 256 k 27.8
      }();
      target = std::min<int64_t>(std::max<int64_t>(target, avg * run_docs), (int64_t)2 << 30);
    }
    int64_t start = 0;
    for (int64_t i = 0; i < n_docs; ++i) {
      if (doc_off[i + 1] - doc_off[start] >= target) {
        S.cut.push_back(i + 1);
        start = i + 1;
      }
    }
    if (S.cut.back() != n_docs || S.cut.size() == 1) S.cut.push_back(n_docs);
  }
  const size_t n_runs = S.cut.size() - 1;
  S.run_hits.assign(n_runs, -1);
  S.run_base.assign(n_runs + 1, 0);
  S.side.assign(n_runs, {nullptr, 0});
  std::unique_ptr<aha_hits> res(new aha_hits());
  memset(res.get(), 0, sizeof(aha_hits));
  HitsOwner* own = new HitsOwner();
  res->opaque = own;
  res->n_docs = n_docs;
  own->device = S.devices[0];
  own->host_doc_hit_off = g_pinned.acquire((size_t)(n_docs + 1) * sizeof(int64_t), &own->host_doc_hit_off_bytes);
  if (!own->host_doc_hit_off) {
    aha_b200_hits_free(res.release());
    return set_err(AHA_ERR_NOMEM, "pinned allocation failed");
  }
  S.h_doc_hit_off = static_cast<int64_t*>(own->host_doc_hit_off);

  const int G = (int)std::min<size_t>(S.devices.size(), n_runs);
  S.devices.resize(G);
  if (G == 1) {
    e2e_worker(&S, 0);
  } else {
    std::vector<std::thread> th;
    for (int w = 0; w < G; ++w) th.emplace_back(e2e_worker, &S, w);
    for (auto& t : th) t.join();
  }
  auto release_blocks = [&] {
    for (auto& s : S.side)
      if (s.first) g_pinned.release(s.first, s.second);
    if (S.hits) g_pinned.release(S.hits, S.hits_bytes);
  };
  if (S.rc != AHA_OK) {
    release_blocks();
    aha_b200_hits_free(res.release());
    return set_err(S.rc, S.err);
  }
  const int64_t n_hits = S.run_base[n_runs];
  // hits that did not fit the primary buffer (a batch whose density grows after its first run): gathered once
  bool any_side = false;
  for (auto& s : S.side) any_side = any_side || s.first;
  if (any_side) {
    size_t got = 0;
    aha_hit* all = static_cast<aha_hit*>(g_pinned.acquire((size_t)std::max<int64_t>(n_hits, 1) * sizeof(aha_hit), &got));
    if (!all) {
      release_blocks();
      aha_b200_hits_free(res.release());
      return set_err(AHA_ERR_NOMEM, "pinned allocation failed");
    }
    for (size_t k = 0; k < n_runs; ++k) {
      const int64_t n = S.run_hits[k];
      if (n <= 0) continue;
      const aha_hit* src = S.side[k].first ? static_cast<const aha_hit*>(S.side[k].first) : S.hits + S.run_base[k];
      memcpy(all + S.run_base[k], src, (size_t)n * sizeof(aha_hit));
    }
    release_blocks();
    S.hits = all;
    S.hits_bytes = got;
  }
  if (!S.hits) S.hits = static_cast<aha_hit*>(g_pinned.acquire(sizeof(aha_hit), &S.hits_bytes));
  own->host_hits = S.hits;
  own->host_hits_bytes = S.hits_bytes;
  // every run's offsets start at 0: rebase by the run's first hit
  int64_t* h_doc_hit_off = S.h_doc_hit_off;
  for (size_t k = 0; k < n_runs; ++k) {
    const int64_t base = S.run_base[k];
    if (base != 0)
      for (int64_t i = S.cut[k]; i < S.cut[k + 1]; ++i) h_doc_hit_off[i] += base;
  }
  h_doc_hit_off[n_docs] = n_hits;
  res->n_hits = n_hits;
  res->scan_ms = S.scan_ms;
  res->walk_ms = S.walk_ms;
  res->expand_ms = S.expand_ms;
  res->device = own->device;
  res->on_device = 0;
  res->hits = static_cast<aha_hit*>(own->host_hits);
  res->doc_hit_off = h_doc_hit_off;
  *out = res.release();
  return AHA_OK;
}

void aha_b200_hits_free(aha_hits* h) {
  if (!h) return;
  aha::free_owner(static_cast<HitsOwner*>(h->opaque));
  delete h;
}

// ---- synthetic corpora --------------------------------------------------------------------
int32_t aha_b200_gen_text_host(int32_t corpus, uint64_t seed, int64_t first, int64_t bytes, uint8_t* out) {
  if (corpus < 0 || corpus > 2 || first < 0 || bytes < 0 || (!out && bytes > 0)) return set_err(AHA_ERR_ARG, "bad argument");
  uint8_t buf[kGenBlock];
  int64_t pos = first;
  const int64_t end = first + bytes;
  while (pos < end) {
    const int64_t k = pos / kGenBlock;
    gen_block(corpus, seed, (uint64_t)k, buf);
    const int64_t b0 = k * kGenBlock;
    const int64_t lo = pos - b0;
    const int64_t hi = std::min<int64_t>(kGenBlock, end - b0);
    memcpy(out + (pos - first), buf + lo, (size_t)(hi - lo));
    pos = b0 + hi;
  }
  return AHA_OK;
}

}  // extern "C"

namespace aha {
__global__ void gen_text_kernel(int corpus, uint64_t seed, int64_t first_block, int64_t n_blocks, int64_t bytes, uint8_t* out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_blocks) return;
  uint8_t buf[kGenBlock];
  gen_block(corpus, seed, (uint64_t)(first_block + i), buf);
  const int64_t o = i * kGenBlock;
  const int64_t n = min((int64_t)kGenBlock, bytes - o);
  for (int64_t j = 0; j < n; ++j) out[o + j] = buf[j];
}
}  // namespace aha

extern "C" int32_t aha_b200_gen_text_device(int32_t corpus, uint64_t seed, int64_t first, int64_t bytes, void* dev_out,
                                            void* stream) {
  if (corpus < 0 || corpus > 2 || first < 0 || bytes < 0 || (!dev_out && bytes > 0)) return set_err(AHA_ERR_ARG, "bad argument");
  if (first % kGenBlock != 0) return set_err(AHA_ERR_ARG, "first must be a multiple of 192");
  if (bytes == 0) return AHA_OK;
  const int64_t n_blocks = (bytes + kGenBlock - 1) / kGenBlock;
  gen_text_kernel<<<(unsigned)((n_blocks + 127) / 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      corpus, seed, first / kGenBlock, n_blocks, bytes, static_cast<uint8_t*>(dev_out)); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
  AHA_CUDA(cudaGetLastError());
  return AHA_OK;
}
