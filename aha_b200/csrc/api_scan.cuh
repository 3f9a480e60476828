// api_scan.cuh -- scan / hits / generator entry points (included at the end of api.cu).
#pragma once
#include <chrono>

namespace aha {

struct U32ToU64 {
  __host__ __device__ __forceinline__ uint64_t operator()(const uint32_t& x) const { return (uint64_t)x; }
};

static int32_t pool_setup(int device) {
  static std::mutex mu;
  static std::vector<int> done;
  std::lock_guard<std::mutex> lk(mu);
  if (std::find(done.begin(), done.end(), device) != done.end()) return AHA_OK;
  cudaMemPool_t pool;
  AHA_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
  uint64_t thr = UINT64_MAX;
  AHA_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
  done.push_back(device);
  return AHA_OK;
}

static int32_t exclusive_sum_u32_to_u64(const uint32_t* in, uint64_t* out, int64_t n, cudaStream_t st) {
  cub::TransformInputIterator<uint64_t, U32ToU64, const uint32_t*> it(in, U32ToU64());
  size_t tmp_bytes = 0;
  AHA_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, it, out, n, st));
  void* tmp = nullptr;
  AHA_CUDA(cudaMallocAsync(&tmp, tmp_bytes ? tmp_bytes : 16, st));
  AHA_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, it, out, n, st));
  AHA_CUDA(cudaFreeAsync(tmp, st));
  return AHA_OK;
}

static void fill_params(ScanParams* p, const DeviceDfa* dd, const void* dev_text, const void* dev_doc_off, int64_t n_docs,
                        int64_t total, uint32_t mode, const uint8_t* sep_bits, int32_t sep_size) {
  memset(p, 0, sizeof(*p));
  p->text = static_cast<const uint8_t*>(dev_text);
  p->total = total;
  p->doc_off = static_cast<const int64_t*>(dev_doc_off);
  p->n_docs = n_docs;
  p->classmap = dd->view.classmap;
  p->delta = dd->view.delta;
  p->out_off = dd->view.out_off;
  p->out_val = dd->view.out_val;
  p->key_len = dd->view.key_len;
  p->key_tail = dd->view.key_tail;
  p->C = dd->view.n_classes;
  p->warm = dd->view.max_key_len > 0 ? dd->view.max_key_len - 1 : 0;
  p->mode = mode;
  p->sep_size = (uint32_t)sep_size;
  if (sep_bits) memcpy(p->sep_bits, sep_bits, 32);
}

// two-pass baseline: count -> exclusive sum -> emit
static int32_t run_v1(ScanParams p, cudaStream_t st, aha_hits* res, HitsOwner* own, cudaEvent_t ev0, cudaEvent_t ev1) {
  const bool trie = (p.mode & AHA_TRIE_LONGEST) != 0;
  const bool longest = trie || (p.mode & AHA_LONGEST) != 0;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  p.chunk = 512;
  p.n_chunks = (p.total + p.chunk - 1) / p.chunk;
  const int64_t units = longest ? p.n_docs : p.n_chunks;
  uint32_t* counts = nullptr;
  uint64_t* offsets = nullptr;
  uint32_t* lead_counts = nullptr;
  uint64_t* lead_prefix = nullptr;
  uint64_t* doc_lead = nullptr;
  AHA_CUDA(cudaMallocAsync(&counts, (units + 1) * sizeof(uint32_t), st));
  AHA_CUDA(cudaMallocAsync(&offsets, (units + 1) * sizeof(uint64_t), st));
  AHA_CUDA(cudaMemsetAsync(counts, 0, (units + 1) * sizeof(uint32_t), st));
  AHA_CUDA(cudaMallocAsync(&own->dev_doc_hit_off, (p.n_docs + 1) * sizeof(int64_t), st));
  AHA_CUDA(cudaMemsetAsync(own->dev_doc_hit_off, 0, (p.n_docs + 1) * sizeof(int64_t), st));
  p.counts = counts;
  p.offsets = offsets;
  p.doc_hit_off = static_cast<int64_t*>(own->dev_doc_hit_off);
  const int tpb = 128;
  // trie path: the count pass keeps its hits in a scratch slice per document (3 bytes of scratch per byte of text), so
  // that the walk runs once; not for inputs whose scratch would not be reasonable to ask for
  aha_hit* doc_scratch = nullptr;
  if (trie && !(p.mode & AHA_COUNT_ONLY) && p.total > 0 && p.total <= ((int64_t)16 << 30))
    AHA_CUDA(cudaMallocAsync(&doc_scratch, (size_t)((p.total >> 2) + 1) * sizeof(aha_hit), st));
  p.doc_scratch = doc_scratch;
  AHA_CUDA(cudaEventRecord(ev0, st));
  if (chars && !longest && p.total > 0) {
    AHA_CUDA(cudaMallocAsync(&lead_counts, (p.n_chunks + 1) * sizeof(uint32_t), st));
    AHA_CUDA(cudaMallocAsync(&lead_prefix, (p.n_chunks + 1) * sizeof(uint64_t), st));
    AHA_CUDA(cudaMallocAsync(&doc_lead, (p.n_docs + 1) * sizeof(uint64_t), st));
    count_leads_kernel<<<(unsigned)((p.n_chunks + 1 + 7) / 8), 256, 0, st>>>(p.text, p.total, p.chunk, p.n_chunks, lead_counts); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    int32_t rc = exclusive_sum_u32_to_u64(lead_counts, lead_prefix, p.n_chunks + 1, st);
    if (rc) return rc;
    doc_lead_base_kernel<<<(unsigned)((p.n_docs + 1 + 7) / 8), 256, 0, st>>>(p.text, p.total, p.doc_off, p.n_docs, p.chunk,
                                                                              lead_prefix, doc_lead); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    p.lead_prefix = lead_prefix;
    p.doc_lead_base = doc_lead;
  }
  const unsigned grid = (unsigned)((units + tpb - 1) / tpb);
  if (units > 0) {
    if (trie) trie_longest_doc_kernel<false><<<grid, tpb, 0, st>>>(p);
    else if (longest) longest_doc_kernel<false><<<grid, tpb, 0, st>>>(p);
    else scan_v1_kernel<false><<<grid, tpb, 0, st>>>(p);
    aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
  }
  int32_t rc = exclusive_sum_u32_to_u64(counts, offsets, units + 1, st);
  if (rc) return rc;
  uint64_t total_hits = 0;
  AHA_CUDA(cudaMemcpyAsync(&total_hits, offsets + units, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  AHA_CUDA(cudaStreamSynchronize(st));
  res->n_hits = (int64_t)total_hits;
  if (!(p.mode & AHA_COUNT_ONLY)) {
    AHA_CUDA(cudaMallocAsync(&own->dev_hits, (total_hits ? total_hits : 1) * sizeof(aha_hit), st));
    p.hits = static_cast<aha_hit*>(own->dev_hits);
  }
  if (units > 0) {
    if (longest) {
      if (trie && p.doc_scratch && p.hits) {
        trie_compact_kernel<<<(unsigned)((p.n_docs + 7) / 8), 256, 0, st>>>(p);
        aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
      }
      if (trie) trie_longest_doc_kernel<true><<<grid, tpb, 0, st>>>(p);  // with a scratch: only documents that overflowed it
      else longest_doc_kernel<true><<<grid, tpb, 0, st>>>(p);
      aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
      offsets_to_doc_hit_off<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(offsets, p.n_docs, p.doc_hit_off); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    } else {
      scan_v1_kernel<true><<<grid, tpb, 0, st>>>(p); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    }
  }
  AHA_CUDA(cudaEventRecord(ev1, st));
  AHA_CUDA(cudaGetLastError());
  AHA_CUDA(cudaFreeAsync(counts, st));
  AHA_CUDA(cudaFreeAsync(offsets, st));
  if (doc_scratch) AHA_CUDA(cudaFreeAsync(doc_scratch, st));
  if (lead_counts) {
    AHA_CUDA(cudaFreeAsync(lead_counts, st));
    AHA_CUDA(cudaFreeAsync(lead_prefix, st));
    AHA_CUDA(cudaFreeAsync(doc_lead, st));
  }
  return AHA_OK;
}

// single-pass fused kernel: see scan_fused.cuh
template <class Policy>
static int32_t launch_fused(const FusedParams& fp, const Policy& pol, int n_sm, size_t smem, cudaStream_t st) {
  // event mode (lists expanded after the walk) for everything but sep scans, whose neighbour tests sit in emit_list
  const char* fr = getenv("AHA_B200_BYTE_RECORDS");
  const bool force_records = fr && fr[0] == '1';  // in-product cross-check of the older delivery
  if ((fp.sp.mode & AHA_SEP) || force_records || !fp.ev_scratch) {
    AHA_CUDA(cudaFuncSetAttribute(scan_fused_kernel<Policy, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    scan_fused_kernel<Policy, false><<<n_sm, kByteWarps * 32, smem, st>>>(fp, pol); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
  } else {
    AHA_CUDA(cudaFuncSetAttribute(scan_fused_kernel<Policy, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    scan_fused_kernel<Policy, true><<<n_sm, kByteWarps * 32, smem, st>>>(fp, pol); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
  }
  AHA_CUDA(cudaGetLastError());
  return AHA_OK;
}

// character automaton, three kernels (scan_char_split.cuh): walk -> prefix over tiles -> expansion at full occupancy
static int32_t run_char_split(ScanParams p, const DeviceDfa* dd, cudaStream_t st, aha_hits* res, HitsOwner* own, cudaEvent_t ev0,
                              cudaEvent_t ev1) {
  int dev = 0, n_sm = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  AHA_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  SplitParams sp;
  memset(&sp, 0, sizeof(sp));
  CharParams& cp = sp.c;
  FusedParams& fp = cp.f;
  fused_geometry(p.warm + 1, true, &fp.chunk, &fp.w16, &fp.buf_bytes);
  fp.tile = 32 * fp.chunk;
  fp.n_tiles = (p.total + fp.tile - 1) / fp.tile;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const HotTable& h = dd->hot;
  cp.t.root_g = reinterpret_cast<const uint32_t*>(h.dev + h.off_root);
  cp.t.val_g = reinterpret_cast<const uint32_t*>(h.dev + h.off_val);
  cp.t.check_g = reinterpret_cast<const uint16_t*>(h.dev + h.off_check);
  cp.t.gfail = reinterpret_cast<const uint16_t*>(h.dev + h.off_gfail);
  cp.t.ref_rec = reinterpret_cast<const RefRec*>(h.dev + h.off_ref);
  cp.t.n_root = h.n_root;
  cp.t.m = h.m;
  cp.t.m_pad = h.m_pad;
  cp.t.n_mwords = h.n_mwords;
  cp.t.mbits_g = reinterpret_cast<const uint32_t*>(h.dev + h.off_mbits);
  for (int c = 0; c < 3; ++c) { cp.t.lo[c] = h.lo[c]; cp.t.n[c] = h.n[c]; cp.t.off[c] = h.off[c]; }
  cp.cmap_g = p.classmap;
  const int64_t nt = fp.n_tiles;
  unsigned int* counters = nullptr;  // [0] ticket, [1] n_direct_list
  uint64_t* tile_prefix = nullptr;
  uint32_t* lead_counts = nullptr;
  uint64_t* tile_lead = nullptr;
  uint64_t* doc_lead = nullptr;
  void* scan_tmp = nullptr;
  size_t scan_tmp_bytes = 0;
  AHA_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_tmp_bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, nt + 1, st));
  // every allocation whose size is known happens before the timed region
  AHA_CUDA(cudaMallocAsync(&counters, 16, st));
  AHA_CUDA(cudaMallocAsync(&sp.store, (size_t)(nt ? nt : 1) * 32 * kStoreSlots * sizeof(CharEvent), st));
  AHA_CUDA(cudaMallocAsync(&sp.tile_meta, (size_t)(nt ? nt : 1) * 32 * kChains * sizeof(uint32_t), st));
  AHA_CUDA(cudaMallocAsync(&sp.tile_hits, (size_t)(nt + 1) * sizeof(uint32_t), st));
  TileDoc* tile_doc = nullptr;
  AHA_CUDA(cudaMallocAsync(&tile_doc, (size_t)(nt ? nt : 1) * sizeof(TileDoc), st));
  sp.tile_doc = tile_doc;
  AHA_CUDA(cudaMallocAsync(&sp.direct_list, (size_t)(nt ? nt : 1) * sizeof(uint32_t), st));
  AHA_CUDA(cudaMallocAsync(&tile_prefix, (size_t)(2 * (nt + 1)) * sizeof(uint64_t), st));
  AHA_CUDA(cudaMallocAsync(&scan_tmp, scan_tmp_bytes ? scan_tmp_bytes : 16, st));
  AHA_CUDA(cudaMallocAsync(&own->dev_doc_hit_off, (p.n_docs + 1) * sizeof(int64_t), st));
  p.doc_hit_off = static_cast<int64_t*>(own->dev_doc_hit_off);
  if (chars && p.total > 0) {
    AHA_CUDA(cudaMallocAsync(&lead_counts, (nt + 1) * sizeof(uint32_t), st));
    AHA_CUDA(cudaMallocAsync(&tile_lead, (nt + 1) * sizeof(uint64_t), st));
    AHA_CUDA(cudaMallocAsync(&doc_lead, (p.n_docs + 1) * sizeof(uint64_t), st));
  }
  sp.n_direct_list = counters + 1;
  fp.ticket = counters;
  fp.n_direct = counters + 1;
  sp.tile_prefix = tile_prefix;
  uint64_t* hits_u64 = tile_prefix + (nt + 1);
  const size_t smem = ((h.smem_bytes + 127) & ~size_t(127)) + (size_t)kFusedWarps * (fp.buf_bytes + 128);
  AHA_CUDA(cudaFuncSetAttribute(char_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));

  AHA_CUDA(cudaEventRecord(ev0, st));
  if (chars && p.total > 0) {
    count_leads_kernel<<<(unsigned)((nt + 1 + 7) / 8), 256, 0, st>>>(p.text, p.total, fp.tile, nt, lead_counts);
    g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    int32_t rc = exclusive_sum_u32_to_u64(lead_counts, tile_lead, nt + 1, st);
    if (rc) return rc;
    doc_lead_base_kernel<<<(unsigned)((p.n_docs + 1 + 7) / 8), 256, 0, st>>>(p.text, p.total, p.doc_off, p.n_docs, fp.tile, tile_lead,
                                                                              doc_lead);
    g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    fp.tile_lead = tile_lead;
    p.doc_lead_base = doc_lead;
  }
  AHA_CUDA(cudaMemsetAsync(counters, 0, 16, st));
  AHA_CUDA(cudaMemsetAsync(sp.tile_hits + nt, 0, sizeof(uint32_t), st));
  AHA_CUDA(cudaMemsetAsync(own->dev_doc_hit_off, 0, (p.n_docs + 1) * sizeof(int64_t), st));
  p.hits = nullptr;
  fp.hit_cap = 0;
  fp.sp = p;
  cudaEvent_t evw0 = nullptr, evw1 = nullptr;
  if (nt > 0) {
    tile_doc_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, fp.tile, nt, tile_doc);
    g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    AHA_CUDA(cudaEventCreate(&evw0));
    AHA_CUDA(cudaEventCreate(&evw1));
    AHA_CUDA(cudaEventRecord(evw0, st));
    char_walk_kernel<<<n_sm, kFusedWarps * 32, smem, st>>>(sp);
    g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    AHA_CUDA(cudaEventRecord(evw1, st));
    AHA_CUDA(cudaGetLastError());
  }
  tile_hits_to_u64_kernel<<<(unsigned)((nt + 1 + 255) / 256), 256, 0, st>>>(sp.tile_hits, nt + 1, hits_u64);
  g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
  AHA_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_tmp_bytes, hits_u64, tile_prefix, nt + 1, st));
  g_kernel_launches.fetch_add(2, std::memory_order_relaxed);
  unsigned long long total_hits = 0;
  unsigned int n_direct = 0;
  AHA_CUDA(cudaMemcpyAsync(&total_hits, tile_prefix + nt, sizeof(total_hits), cudaMemcpyDeviceToHost, st));
  AHA_CUDA(cudaMemcpyAsync(&n_direct, sp.n_direct_list, sizeof(n_direct), cudaMemcpyDeviceToHost, st));
  AHA_CUDA(cudaStreamSynchronize(st));
  if (evw0) {
    float wms = 0;
    AHA_CUDA(cudaEventElapsedTime(&wms, evw0, evw1));
    res->walk_ms = wms;
    cudaEventDestroy(evw0);
    cudaEventDestroy(evw1);
  }
  // the hit buffer is sized exactly: the walk has already counted
  if (!(p.mode & AHA_COUNT_ONLY)) {
    AHA_CUDA(cudaMallocAsync(&own->dev_hits, (size_t)(total_hits ? total_hits : 1) * sizeof(aha_hit), st));
    fp.sp.hits = static_cast<aha_hit*>(own->dev_hits);
    fp.hit_cap = (int64_t)total_hits;
  }
  if (nt > 0) {
    const unsigned grid = (unsigned)std::min<int64_t>((nt + 7) / 8, (int64_t)n_sm * 64);
    char_expand_kernel<<<grid, 256, 0, st>>>(sp);
    g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    if (n_direct > 0) {
      const unsigned dgrid = (unsigned)std::min<int64_t>(((int64_t)n_direct + 3) / 4, (int64_t)n_sm * 8);
      char_direct_kernel<<<dgrid, 128, 256 + 4 * (size_t)fp.buf_bytes, st>>>(sp);
      g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    }
    AHA_CUDA(cudaGetLastError());
  }
  finalize_doc_off_kernel<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, p.total, reinterpret_cast<const unsigned long long*>(tile_prefix + nt), p.doc_hit_off);
  g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
  AHA_CUDA(cudaEventRecord(ev1, st));
  AHA_CUDA(cudaGetLastError());
  res->n_hits = (int64_t)total_hits;
  AHA_CUDA(cudaFreeAsync(counters, st));
  AHA_CUDA(cudaFreeAsync(sp.store, st));
  AHA_CUDA(cudaFreeAsync(sp.tile_meta, st));
  AHA_CUDA(cudaFreeAsync(sp.tile_hits, st));
  AHA_CUDA(cudaFreeAsync(tile_doc, st));
  AHA_CUDA(cudaFreeAsync(sp.direct_list, st));
  AHA_CUDA(cudaFreeAsync(tile_prefix, st));
  AHA_CUDA(cudaFreeAsync(scan_tmp, st));
  if (lead_counts) {
    AHA_CUDA(cudaFreeAsync(lead_counts, st));
    AHA_CUDA(cudaFreeAsync(tile_lead, st));
    AHA_CUDA(cudaFreeAsync(doc_lead, st));
  }
  return AHA_OK;
}

static int32_t run_fused(ScanParams p, const DeviceDfa* dd, const aha_dfa* dfa, cudaStream_t st, aha_hits* res, HitsOwner* own,
                         cudaEvent_t ev0, cudaEvent_t ev1) {
  int dev = 0, n_sm = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  AHA_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  FusedParams fp;
  memset(&fp, 0, sizeof(fp));
  // character automaton in shared memory when the dictionary allows it (plain match, with or without char units)
  const bool use_char = dd->hot.ok && !(p.mode & AHA_SEP);
  fused_geometry(p.warm + 1, use_char, &fp.chunk, &fp.w16, &fp.buf_bytes);
  fp.tile = 32 * fp.chunk;
  fp.n_tiles = (p.total + fp.tile - 1) / fp.tile;
  const bool chars = (p.mode & AHA_UNITS_CHAR) != 0;
  const int64_t n_warps = (int64_t)n_sm * (use_char ? kFusedWarps : kByteWarps);
  unsigned long long* tile_state = nullptr;
  unsigned int* counters = nullptr;  // [0] ticket, [1] n_direct, [2..3] total_hits (u64)
  uint32_t* lead_counts = nullptr;
  uint64_t* tile_lead = nullptr;
  uint64_t* doc_lead = nullptr;
  const int64_t n_groups = (fp.n_tiles + 31) / 32;
  const size_t state_words = (size_t)(fp.n_tiles + 1) + (size_t)(n_groups + 1);
  AHA_CUDA(cudaMallocAsync(&tile_state, state_words * sizeof(unsigned long long), st));
  AHA_CUDA(cudaMallocAsync(&counters, 16 + 8 * 8, st));  // + 8 phase-cycle counters (AHA_B200_PHASES)
  CharParams cp;
  memset(&cp, 0, sizeof(cp));
  if (use_char) {
    const HotTable& h = dd->hot;
    cp.t.root_g = reinterpret_cast<const uint32_t*>(h.dev + h.off_root);
    cp.t.val_g = reinterpret_cast<const uint32_t*>(h.dev + h.off_val);
    cp.t.check_g = reinterpret_cast<const uint16_t*>(h.dev + h.off_check);
    cp.t.gfail = reinterpret_cast<const uint16_t*>(h.dev + h.off_gfail);
    cp.t.ref_rec = reinterpret_cast<const RefRec*>(h.dev + h.off_ref);
    cp.t.n_root = h.n_root;
    cp.t.m = h.m;
    cp.t.m_pad = h.m_pad;
    cp.t.n_mwords = h.n_mwords;
    cp.t.mbits_g = reinterpret_cast<const uint32_t*>(h.dev + h.off_mbits);
    for (int c = 0; c < 3; ++c) { cp.t.lo[c] = h.lo[c]; cp.t.n[c] = h.n[c]; cp.t.off[c] = h.off[c]; }
    cp.cmap_g = p.classmap;
    AHA_CUDA(cudaMallocAsync(&cp.events, (size_t)n_warps * 2 * 32 * kEvSlots * sizeof(CharEvent), st));
  } else {
    AHA_CUDA(cudaMallocAsync(&fp.scratch, (size_t)n_warps * 32 * kLaneCap * sizeof(aha_hit), st));
    if (!(p.mode & AHA_SEP)) AHA_CUDA(cudaMallocAsync(&fp.ev_scratch, (size_t)n_warps * 32 * kByteEvCap * sizeof(uint2), st));
  }
  AHA_CUDA(cudaMallocAsync(&fp.marks, (size_t)n_warps * 2 * 32 * kMarkCap * sizeof(int2), st));
  AHA_CUDA(cudaMallocAsync(&own->dev_doc_hit_off, (p.n_docs + 1) * sizeof(int64_t), st));
  p.doc_hit_off = static_cast<int64_t*>(own->dev_doc_hit_off);
  fp.tile_state = tile_state;
  fp.group_excl = tile_state + (fp.n_tiles + 1);
  fp.ticket = counters;
  fp.n_direct = counters + 1;
  fp.total_hits = reinterpret_cast<unsigned long long*>(counters + 2);
  ByteGlobal pol;
  pol.cmap_g = p.classmap;
  pol.cmap = nullptr;
  size_t smem = ((ByteGlobal::table_bytes(pol) + 127) & ~127u) + (size_t)kByteWarps * (fp.buf_bytes + 128);
  if (use_char) {
    smem = ((dd->hot.smem_bytes + 127) & ~size_t(127)) + (size_t)kFusedWarps * (fp.buf_bytes + 128);
    AHA_CUDA(cudaFuncSetAttribute(scan_char_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  }

#ifdef AHA_B200_PHASES
  auto h_now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double h0 = h_now();
#define AHA_HSTAMP(what) fprintf(stderr, "[host] %-14s +%.3f ms\n", what, h_now() - h0)
#else
#define AHA_HSTAMP(what) do { } while (0)
#endif
  // every allocation happens before the timed region: growing the pool synchronises with the stream
  const bool count_only = (p.mode & AHA_COUNT_ONLY) != 0;
  int64_t cap = 0;
  if (!count_only && dfa->hits_per_byte >= 0) {
    cap = (int64_t)((double)p.total * dfa->hits_per_byte * 1.25) + 4096;
    AHA_CUDA(cudaMallocAsync(&own->dev_hits, (size_t)cap * sizeof(aha_hit), st));
  }
  if (chars && p.total > 0) {
    AHA_CUDA(cudaMallocAsync(&lead_counts, (fp.n_tiles + 1) * sizeof(uint32_t), st));
    AHA_CUDA(cudaMallocAsync(&tile_lead, (fp.n_tiles + 1) * sizeof(uint64_t), st));
    AHA_CUDA(cudaMallocAsync(&doc_lead, (p.n_docs + 1) * sizeof(uint64_t), st));
  }
  AHA_CUDA(cudaEventRecord(ev0, st));
  if (chars && p.total > 0) {
    count_leads_kernel<<<(unsigned)((fp.n_tiles + 1 + 7) / 8), 256, 0, st>>>(p.text, p.total, fp.tile, fp.n_tiles, lead_counts); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    int32_t rc = exclusive_sum_u32_to_u64(lead_counts, tile_lead, fp.n_tiles + 1, st);
    if (rc) return rc;
    doc_lead_base_kernel<<<(unsigned)((p.n_docs + 1 + 7) / 8), 256, 0, st>>>(p.text, p.total, p.doc_off, p.n_docs, fp.tile,
                                                                              tile_lead, doc_lead); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    fp.tile_lead = tile_lead;
    p.doc_lead_base = doc_lead;
    AHA_HSTAMP("lead prepass");
  }
  // capacity of the hit buffer: exact after a counting run on first use, density history afterwards
  unsigned long long total_hits = 0;
  bool have_result = false;
  for (int attempt = 0; attempt < 3 && !have_result; ++attempt) {
    bool counting = count_only || (attempt == 0 && cap == 0);
    if (!counting && attempt > 0) {  // the first estimate was missing or too small: cap holds the exact size now
      if (own->dev_hits) AHA_CUDA(cudaFreeAsync(own->dev_hits, st));
      own->dev_hits = nullptr;
      AHA_CUDA(cudaMallocAsync(&own->dev_hits, (size_t)cap * sizeof(aha_hit), st));
    }
    AHA_HSTAMP("hits alloc");
    p.hits = counting ? nullptr : static_cast<aha_hit*>(own->dev_hits);
    fp.hit_cap = counting ? 0 : cap;
    fp.sp = p;
    AHA_CUDA(cudaMemsetAsync(tile_state, 0, state_words * sizeof(unsigned long long), st));
    AHA_CUDA(cudaMemsetAsync(counters, 0, 16 + 8 * 8, st));
    AHA_CUDA(cudaMemsetAsync(own->dev_doc_hit_off, 0, (p.n_docs + 1) * sizeof(int64_t), st));
    if (fp.n_tiles > 0) {
      if (use_char) {
        cp.f = fp;
        scan_char_kernel<<<n_sm, kFusedWarps * 32, smem, st>>>(cp); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
        AHA_CUDA(cudaGetLastError());
      } else {
        int32_t rc = launch_fused(fp, pol, n_sm, smem, st);
        if (rc) return rc;
      }
    }
    finalize_doc_off_kernel<<<(unsigned)((p.n_docs + 1 + 255) / 256), 256, 0, st>>>(p.doc_off, p.n_docs, p.total, fp.total_hits,
                                                                                 p.doc_hit_off); aha::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    AHA_CUDA(cudaEventRecord(ev1, st));
    AHA_HSTAMP("launched");
    AHA_CUDA(cudaMemcpyAsync(&total_hits, fp.total_hits, sizeof(total_hits), cudaMemcpyDeviceToHost, st));
    AHA_CUDA(cudaStreamSynchronize(st));
#ifdef AHA_B200_PHASES
    {
      unsigned long long ph[8];
      cudaMemcpy(ph, counters + 4, sizeof(ph), cudaMemcpyDeviceToHost);
      unsigned int nd = 0;
      cudaMemcpy(&nd, counters + 1, 4, cudaMemcpyDeviceToHost);
      fprintf(stderr, "[phases] tiles=%lld direct=%u  mega-cycles: stage_wait=%.1f prep=%.1f walk=%.1f count=%.1f next=%.1f deliver=%.1f direct=%.1f\n",
              (long long)fp.n_tiles, nd, ph[0] / 1e6, ph[1] / 1e6, ph[2] / 1e6, ph[3] / 1e6, ph[4] / 1e6, ph[5] / 1e6, ph[6] / 1e6);
    }
#endif
    const double dens = p.total > 0 ? (double)total_hits / (double)p.total : 0.0;
    if (dens > dfa->hits_per_byte) dfa->hits_per_byte = dens;
    if (counting) {
      if (count_only) have_result = true;
      else cap = (int64_t)total_hits + 1;  // exact capacity for the real run
    } else if ((int64_t)total_hits <= cap) {
      have_result = true;
    } else {
      cap = (int64_t)total_hits + 1;  // estimate was too small: rerun with the exact size
    }
    if (!have_result) AHA_CUDA(cudaEventRecord(ev0, st));  // time only the run that delivers
  }
  if (!have_result) return set_err(AHA_ERR_STATE, "hit buffer sizing did not converge");
  res->n_hits = (int64_t)total_hits;
  AHA_CUDA(cudaFreeAsync(tile_state, st));
  AHA_CUDA(cudaFreeAsync(counters, st));
  if (fp.scratch) AHA_CUDA(cudaFreeAsync(fp.scratch, st));
  if (fp.ev_scratch) AHA_CUDA(cudaFreeAsync(fp.ev_scratch, st));
  if (cp.events) AHA_CUDA(cudaFreeAsync(cp.events, st));
  AHA_CUDA(cudaFreeAsync(fp.marks, st));
  if (lead_counts) {
    AHA_CUDA(cudaFreeAsync(lead_counts, st));
    AHA_CUDA(cudaFreeAsync(tile_lead, st));
    AHA_CUDA(cudaFreeAsync(doc_lead, st));
  }
  return AHA_OK;
}

static int32_t scan_on_device(const aha_dfa* d, const void* dev_text, const void* dev_doc_off, int64_t n_docs,
                              int64_t total, uint32_t mode, const uint8_t* sep_bits, int32_t sep_size, cudaStream_t st,
                              aha_hits* res, HitsOwner* own) {
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  const DeviceDfa* dd = d->on(dev);
  if (!dd) return set_err(AHA_ERR_STATE, "automaton was not uploaded to the current device (aha_b200_init)");
  if ((mode & AHA_SEP) && (sep_size > 256 || sep_size < 0 || !sep_bits))
    return set_err(AHA_ERR_ARG, "sep BitArray size > 256 is not supported");  // ac.cr:322
  if ((mode & AHA_TRIE_LONGEST) && dd->view.max_key_len > (uint32_t)kTrieMaxChars)
    return set_err(AHA_ERR_ARG, "AHA_TRIE_LONGEST supports keys of at most 96 bytes");
  int32_t rc = pool_setup(dev);
  if (rc) return rc;
  own->device = dev;
  own->stream = st;
  ScanParams p;
  fill_params(&p, dd, dev_text, dev_doc_off, n_docs, total, mode, sep_bits, sep_size);
  cudaEvent_t ev0, ev1;
  AHA_CUDA(cudaEventCreate(&ev0));
  AHA_CUDA(cudaEventCreate(&ev1));
  const bool fused_ok = scan_impl() == 1 && !(mode & (AHA_LONGEST | AHA_TRIE_LONGEST)) &&
                        (reinterpret_cast<uintptr_t>(dev_text) & 15) == 0 && dd->view.max_key_len <= 1024;
  const char* cimpl = getenv("AHA_B200_CHAR_IMPL");  // "fused": the single-kernel variant of scan_char.cuh (kept for A/B runs)
  const bool split = fused_ok && dd->hot.ok && !(mode & AHA_SEP) && !(cimpl && cimpl[0] == 'f');
  if (split) rc = run_char_split(p, dd, st, res, own, ev0, ev1);
  else if (fused_ok) rc = run_fused(p, dd, d, st, res, own, ev0, ev1);
  else rc = run_v1(p, st, res, own, ev0, ev1);
  if (rc == AHA_OK) {
    AHA_CUDA(cudaEventSynchronize(ev1));
    float ms = 0;
    AHA_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
    res->scan_ms = ms;
  }
  cudaEventDestroy(ev0);
  cudaEventDestroy(ev1);
  return rc;
}

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
                              static_cast<cudaStream_t>(stream), res.get(), own);
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

// bytes per pipeline stage of aha_b200_scan (documents are never split)
static int64_t e2e_chunk_bytes() {
  const char* e = getenv("AHA_B200_E2E_CHUNK_MB");
  const int64_t mb = e ? atoll(e) : 64;
  return (mb > 0 ? mb : 64) << 20;
}
}  // namespace aha

// Host buffers in, host buffers out.  The batch is cut into runs of whole documents of ~64 MB; the H2D copy of
// run k+1 and the D2H copy of the hits of run k-1 overlap the scan of run k (three streams), so that on a
// PCIe-attached GPU the call is bound by the slower of the two copy directions, not by their sum.
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
  const int64_t total = doc_off[n_docs];
  int32_t rc = ensure_devices();
  if (rc) return rc;
  std::unique_ptr<aha_hits> res(new aha_hits());
  memset(res.get(), 0, sizeof(aha_hits));
  HitsOwner* own = new HitsOwner();
  res->opaque = own;
  res->n_docs = n_docs;
  const bool count_only = (mode & AHA_COUNT_ONLY) != 0;

  // runs of whole documents
  std::vector<int64_t> cut(1, 0);
  {
    const int64_t target = e2e_chunk_bytes();
    int64_t start = 0;
    for (int64_t i = 0; i < n_docs; ++i) {
      if (doc_off[i + 1] - doc_off[start] >= target) {
        cut.push_back(i + 1);
        start = i + 1;
      }
    }
    if (cut.back() != n_docs || cut.size() == 1) cut.push_back(n_docs);
  }
  const size_t n_runs = cut.size() - 1;
  int64_t max_bytes = 16, max_docs = 1;
  for (size_t k = 0; k < n_runs; ++k) {
    max_bytes = std::max<int64_t>(max_bytes, doc_off[cut[k + 1]] - doc_off[cut[k]]);
    max_docs = std::max<int64_t>(max_docs, cut[k + 1] - cut[k]);
  }

  cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
  cudaEvent_t e_in[2] = {nullptr, nullptr}, e_out = nullptr;
  uint8_t* d_text[2] = {nullptr, nullptr};
  int64_t* d_off[2] = {nullptr, nullptr};
  int64_t* h_off_stage = nullptr;  // pinned, rebased doc_off of the runs in flight
  size_t h_off_stage_bytes = 0;
  std::vector<void*> late_free;    // device hit blocks whose D2H copy may still be running
  size_t hits_cap = 0, hits_bytes = 0, off_bytes = 0;
  auto cleanup = [&]() {
    if (s_out) cudaStreamSynchronize(s_out);
    if (s_in) cudaStreamSynchronize(s_in);
    for (void* p : late_free) cudaFree(p);
    for (int b = 0; b < 2; ++b) {
      if (d_text[b]) cudaFree(d_text[b]);
      if (d_off[b]) cudaFree(d_off[b]);
      if (e_in[b]) cudaEventDestroy(e_in[b]);
    }
    if (e_out) cudaEventDestroy(e_out);
    if (h_off_stage) g_pinned.release(h_off_stage, h_off_stage_bytes);
    if (s_in) cudaStreamDestroy(s_in);
    if (s_run) cudaStreamDestroy(s_run);
    if (s_out) cudaStreamDestroy(s_out);
  };
  auto fail = [&](int32_t code) {
    cleanup();
    aha_b200_hits_free(res.release());
    return code;
  };
#define AHA_E2E(expr)                                                                                      \
  do {                                                                                                     \
    cudaError_t _e = (expr);                                                                               \
    if (_e != cudaSuccess) return fail(set_err(AHA_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e))); \
  } while (0)
  AHA_E2E(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
  AHA_E2E(cudaStreamCreateWithFlags(&s_run, cudaStreamNonBlocking));
  AHA_E2E(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
  AHA_E2E(cudaEventCreateWithFlags(&e_out, cudaEventDisableTiming));
  const int n_buf = n_runs > 1 ? 2 : 1;
  for (int b = 0; b < n_buf; ++b) {
    AHA_E2E(cudaEventCreateWithFlags(&e_in[b], cudaEventDisableTiming));
    AHA_E2E(cudaMalloc(&d_text[b], (size_t)max_bytes + 64));
    AHA_E2E(cudaMalloc(&d_off[b], (size_t)(max_docs + 1) * sizeof(int64_t)));
  }
  h_off_stage = static_cast<int64_t*>(g_pinned.acquire((size_t)n_buf * (max_docs + 1) * sizeof(int64_t), &h_off_stage_bytes));
  own->host_doc_hit_off = g_pinned.acquire((size_t)(n_docs + 1) * sizeof(int64_t), &off_bytes);
  if (!h_off_stage || !own->host_doc_hit_off) return fail(set_err(AHA_ERR_NOMEM, "pinned allocation failed"));
  own->host_doc_hit_off_bytes = off_bytes;
  int64_t* h_doc_hit_off = static_cast<int64_t*>(own->host_doc_hit_off);

  auto stage_in = [&](size_t k) -> cudaError_t {
    const int b = (int)(k & 1);
    const int64_t lo = cut[k], hi = cut[k + 1];
    const int64_t base = doc_off[lo], bytes = doc_off[hi] - base;
    int64_t* ho = h_off_stage + (size_t)b * (max_docs + 1);
    for (int64_t i = lo; i <= hi; ++i) ho[i - lo] = doc_off[i] - base;
    cudaError_t e = cudaSuccess;
    if (bytes > 0) e = cudaMemcpyAsync(d_text[b], text + base, (size_t)bytes, cudaMemcpyHostToDevice, s_in);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_off[b], ho, (size_t)(hi - lo + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s_in);
    if (e == cudaSuccess) e = cudaEventRecord(e_in[b], s_in);
    return e;
  };

  int64_t hit_base = 0;
  double scan_ms = 0, walk_ms = 0;
  std::vector<int64_t> run_hit_base(n_runs + 1, 0);
  AHA_E2E(stage_in(0));
  for (size_t k = 0; k < n_runs; ++k) {
    const int b = (int)(k & 1);
    const int64_t lo = cut[k], hi = cut[k + 1];
    const int64_t bytes = doc_off[hi] - doc_off[lo];
    AHA_E2E(cudaStreamWaitEvent(s_run, e_in[b], 0));
    aha_hits part;
    memset(&part, 0, sizeof(part));
    HitsOwner pown;
    // the other buffer was released by the scan of run k-1 (scan_on_device synchronises its stream): the copy of
    // run k+1 overlaps this scan, the hits of run k-1 are still going home on s_out
    if (k + 1 < n_runs) AHA_E2E(stage_in(k + 1));
    rc = scan_on_device(d, d_text[b], d_off[b], hi - lo, bytes, mode, sep_bits, sep_size, s_run, &part, &pown);
    if (rc) return fail(rc);
    scan_ms += part.scan_ms;
    walk_ms += part.walk_ms;
    run_hit_base[k] = hit_base;
    if (!count_only && part.n_hits > 0) {
      const size_t need = (size_t)(hit_base + part.n_hits);
      if (need > hits_cap) {  // grow: estimate the whole batch from the density seen so far
        const double dens = (double)need / (double)std::max<int64_t>(1, doc_off[hi]);
        size_t want = (size_t)((double)total * dens * 1.15) + 4096;
        if (k + 1 == n_runs || want < need) want = need;
        size_t got = 0;
        void* nb = g_pinned.acquire(want * sizeof(aha_hit), &got);
        if (!nb) return fail(set_err(AHA_ERR_NOMEM, "pinned allocation failed"));
        if (own->host_hits) {
          AHA_E2E(cudaStreamSynchronize(s_out));
          memcpy(nb, own->host_hits, (size_t)hit_base * sizeof(aha_hit));
          g_pinned.release(own->host_hits, hits_bytes);
        }
        own->host_hits = nb;
        hits_bytes = got;
        own->host_hits_bytes = got;
        hits_cap = got / sizeof(aha_hit);
      }
      AHA_E2E(cudaMemcpyAsync(static_cast<aha_hit*>(own->host_hits) + hit_base, pown.dev_hits, (size_t)part.n_hits * sizeof(aha_hit),
                              cudaMemcpyDeviceToHost, s_out));
    }
    AHA_E2E(cudaMemcpyAsync(h_doc_hit_off + lo, pown.dev_doc_hit_off, (size_t)(hi - lo + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost,
                            s_out));
    if (pown.dev_hits) AHA_E2E(cudaFreeAsync(pown.dev_hits, s_out));
    if (pown.dev_doc_hit_off) AHA_E2E(cudaFreeAsync(pown.dev_doc_hit_off, s_out));
    // per-run offsets start at 0: rebase once the copy has landed (next iteration's host work overlaps it)
    if (k > 0) {
      // run k-1's offsets are complete after the previous D2H; the copy just issued is behind it on s_out
    }
    hit_base += part.n_hits;
    own->device = pown.device;
  }
  run_hit_base[n_runs] = hit_base;
  AHA_E2E(cudaStreamSynchronize(s_out));
  // runs overlap in one slot of doc_hit_off (the last offset of run k is the first of run k+1): rebase back to front
  for (size_t k = n_runs; k-- > 0;) {
    const int64_t lo = cut[k], hi = cut[k + 1], base = run_hit_base[k];
    if (k + 1 < n_runs) h_doc_hit_off[hi] = run_hit_base[k + 1];
    if (base != 0)
      for (int64_t i = lo; i < hi; ++i) h_doc_hit_off[i] += base;
  }
  h_doc_hit_off[n_docs] = hit_base;
  if (n_docs == 0) h_doc_hit_off[0] = 0;
  if (!own->host_hits) {
    size_t got = 0;
    own->host_hits = g_pinned.acquire(sizeof(aha_hit), &got);
    own->host_hits_bytes = got;
  }
  cleanup();
#undef AHA_E2E
  res->n_hits = hit_base;
  res->scan_ms = scan_ms;
  res->walk_ms = walk_ms;
  res->device = own->device;
  res->on_device = 0;
  res->hits = static_cast<aha_hit*>(own->host_hits);
  res->doc_hit_off = h_doc_hit_off;
  *out = res.release();
  return AHA_OK;
}

void aha_b200_hits_free(aha_hits* h) {
  if (!h) return;
  HitsOwner* own = static_cast<HitsOwner*>(h->opaque);
  if (own) {
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(own->device);
    if (own->dev_hits) cudaFreeAsync(own->dev_hits, own->stream);
    if (own->dev_doc_hit_off) cudaFreeAsync(own->dev_doc_hit_off, own->stream);
    if (own->dev_text) cudaFreeAsync(own->dev_text, own->stream);
    if (own->dev_doc_off) cudaFreeAsync(own->dev_doc_off, own->stream);
    if (own->host_hits) aha::g_pinned.release(own->host_hits, own->host_hits_bytes);
    if (own->host_doc_hit_off) aha::g_pinned.release(own->host_doc_hit_off, own->host_doc_hit_off_bytes);
    cudaSetDevice(prev);
    delete own;
  }
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
