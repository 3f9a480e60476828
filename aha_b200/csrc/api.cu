// api.cu -- extern "C" entry points of libaha_b200.so (see include/aha_b200.h).
#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "api_internal.cuh"
#include "../../include/aha_synth.h"
#include "char_table.hpp"
#include "char_walk.cuh"
#include "scan_common.cuh"
#include "scan_v1.cuh"
#include "scan_trie.cuh"
#include "scan_longest.cuh"
#include "scan_fused.cuh"
#include "scan_char.cuh"
#include "scan_char_split.cuh"

namespace aha {

thread_local std::string g_err;
std::atomic<long long> g_kernel_launches{0};
static std::mutex g_mu;
static std::vector<int> g_devices;  // devices selected by aha_b200_init

// the environment is read once, when the library is first used; aha_b200_debug_set changes a knob afterwards
Knobs& knobs() {
  static Knobs k = [] {
    Knobs v;
    auto num = [](const char* name, int dflt) {
      const char* e = getenv(name);
      return e && *e ? atoi(e) : dflt;
    };
    v.no_char = num("AHA_B200_NO_CHAR", 0) == 1;
    const char* impl = getenv("AHA_B200_IMPL");
    v.force_v1 = impl && impl[0] == '0';
    v.byte_chunk = num("AHA_B200_BYTE_CHUNK", 0);
    v.e2e_chunk_mb = num("AHA_B200_E2E_CHUNK_MB", 64);
    if (v.e2e_chunk_mb <= 0) v.e2e_chunk_mb = 64;
    v.byte_span = num("AHA_B200_BYTE_SPAN", kByteSpanDefault);
    if (v.byte_span < 0 || v.byte_span > 8192 || (v.byte_span & 15)) v.byte_span = 0;
    return v;
  }();
  return k;
}

static int32_t ensure_devices() {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_devices.empty()) return AHA_OK;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    return set_err(AHA_ERR_CUDA, std::string("no CUDA device: libaha_b200 has no CPU fallback (") +
                                     cudaGetErrorString(e) + ")");
  }
  int cur = 0;
  AHA_CUDA(cudaGetDevice(&cur));
  g_devices.push_back(cur);
  return AHA_OK;
}

// per-warp text buffer geometry of the fused kernels (shared by the table budget and the launcher)
// byte kernel: 132 = 4 * 33 bytes per lane, the lanes' chunks start in distinct banks.
// character kernels: 126 = 2 sub-chunks of 63 bytes; with the 9 warm-up bytes that a run of 3-byte units needs, a
// sub-chunk is exactly six 12-byte steps (132 would need a seventh, three quarters empty)
static void fused_geometry(uint32_t max_key_len, bool char_kernel, int32_t* chunk, int32_t* w16, int32_t* buf_bytes) {
  const uint32_t warm = max_key_len > 0 ? max_key_len - 1 : 0;
  *w16 = (int32_t)((warm + 1 + 3 + 15) & ~15u);
  if (*w16 < 16) *w16 = 16;
  auto buf_for = [&](int32_t c) { return (*w16 + 32 * c + 16 + 32 + 127) & ~127; };  // 16 staged bytes past the tile + read slack
  *chunk = char_kernel ? AHA_B200_CHAR_CHUNK : 132;
  if (!char_kernel) {
    // byte kernel: bound by table lookups per second, and every lane repeats max_key_len - 1 of them as warm-up, so the
    // longest chunk (chunk / 4 odd: lanes in distinct banks) whose 32 per-warp buffers still fit next to the class map
    int32_t want = knobs().byte_chunk ? knobs().byte_chunk : kByteChunkDefault;
    if (want >= 36 && want <= 1024 && want % 4 == 0 && (want / 4) % 2 == 1 &&
        (size_t)kByteWarps * (size_t)(buf_for(want) + 128) + 512 <= (size_t)227 * 1024)
      *chunk = want;
  }
  *buf_bytes = buf_for(*chunk);
}

static int32_t build_hot_table(const aha_dfa* d, DeviceDfa* dd) {
  dd->hot = HotTable();
  if (!d->header.has_delta) return AHA_OK;  // a dictionary too large for a dense table is far too large for shared memory
  int32_t chunk, w16, buf;
  fused_geometry(d->header.max_key_len, true, &chunk, &w16, &buf);
  const size_t total_smem = 227 * 1024;
  const size_t per_warps = (size_t)kFusedWarps * (buf + 128) + 512;
  if (per_warps + 4096 > total_smem) return AHA_OK;
  CharTableHost T = build_char_table(d->header, d->blob.data(), total_smem - per_warps);
  if (!T.ok) return AHA_OK;  // not eligible: the byte automaton in global memory serves this dictionary
  HotTable& h = dd->hot;
  auto place = [](size_t& cursor, size_t bytes) { size_t at = cursor; cursor = (cursor + bytes + 255) & ~size_t(255); return at; };
  size_t cursor = 0;
  h.off_root = place(cursor, T.root.size() * 2);
  h.off_val = place(cursor, T.val.size() * 4);
  h.off_check = place(cursor, T.check.size() * 2);
  h.off_gfail = place(cursor, T.gfail.size() * 2);
  h.off_ref = place(cursor, T.ref_rec.size() * sizeof(RefRec));
  h.off_mbits = place(cursor, T.multi_bits.size() * 4);
  // ref_rec in 8 bytes for the expansion's shared-memory table (scan_char_split.cuh)
  std::vector<uint2> ref8(T.ref_rec.size());
  bool ref8_ok = true;
  for (size_t i = 0; i < T.ref_rec.size(); ++i) {
    const RefRec& r = T.ref_rec[i];
    if (r.n == 1) {
      if (r.b >= 0x8000u || r.c >= 0x10000u || r.b == 0) ref8_ok = false;
      ref8[i] = make_uint2(r.a, (r.b << 16) | r.c);
    } else {
      ref8[i] = make_uint2(r.a, r.n ? (0x80000000u | r.n) : 0u);
    }
  }
  h.off_ref8 = place(cursor, ref8.size() * sizeof(uint2));
  // every entry of the output lists in the same 8 bytes {key id, key length << 16 | characters after the first}: a list of
  // several keys is then expanded with one load per hit instead of out_val[] -> key_len[]
  const DfaView hv = view_of(d->header, d->blob.data());
  std::vector<uint2> out8(hv.n_out);
  for (uint64_t j = 0; j < hv.n_out; ++j) {
    const int32_t v = hv.out_val[j];
    if (v < 0 || (uint64_t)v >= hv.n_keys || hv.key_len[v] >= 0x8000u || hv.key_tail[v] >= 0x10000u) {
      ref8_ok = false;
      break;
    }
    out8[j] = make_uint2((uint32_t)v, (hv.key_len[v] << 16) | hv.key_tail[v]);
  }
  h.off_out8 = place(cursor, (out8.size() + 1) * sizeof(uint2));
  std::vector<uint8_t> stage(cursor, 0);
  memcpy(stage.data() + h.off_ref8, ref8.data(), ref8.size() * sizeof(uint2));
  if (!out8.empty()) memcpy(stage.data() + h.off_out8, out8.data(), out8.size() * sizeof(uint2));
  if (!ref8_ok) return AHA_OK;  // a key longer than 32 KB: the byte automaton serves this dictionary (h.ok stays false)
  h.n_ref8 = (uint32_t)ref8.size();
  memcpy(stage.data() + h.off_root, T.root.data(), T.root.size() * 2);
  memcpy(stage.data() + h.off_val, T.val.data(), T.val.size() * 4);
  memcpy(stage.data() + h.off_check, T.check.data(), T.check.size() * 2);
  memcpy(stage.data() + h.off_gfail, T.gfail.data(), T.gfail.size() * 2);
  memcpy(stage.data() + h.off_ref, T.ref_rec.data(), T.ref_rec.size() * sizeof(RefRec));
  memcpy(stage.data() + h.off_mbits, T.multi_bits.data(), T.multi_bits.size() * 4);
  AHA_CUDA(cudaMalloc(&h.dev, cursor));
  AHA_CUDA(cudaMemcpy(h.dev, stage.data(), cursor, cudaMemcpyHostToDevice));
  h.ok = true;
  h.n_hot = T.n_states;
  h.n_root = T.n_root;
  h.m = T.m;
  h.m_pad = (uint32_t)T.check.size();
  h.n_mwords = (uint32_t)T.multi_bits.size();
  h.n_gfail_s = T.n_gfail_s;
  for (int c = 0; c < 3; ++c) { h.lo[c] = T.lo[c]; h.n[c] = T.n[c]; h.off[c] = T.off[c]; }
  h.smem_bytes = T.smem_bytes();
  return AHA_OK;
}

static int32_t upload(aha_dfa* d) {
  int32_t rc = ensure_devices();
  if (rc) return rc;
  DeviceGuard guard;  // the caller's current device is restored on every path
  std::vector<int> devices;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    devices = g_devices;
  }
  for (int dev : devices) {
    AHA_CUDA(cudaSetDevice(dev));
    d->dev.emplace_back();
    DeviceDfa& dd = d->dev.back();
    dd.device = dev;
    AHA_CUDA(cudaMalloc(&dd.blob, d->blob.size()));
    AHA_CUDA(cudaMemcpy(dd.blob, d->blob.data(), d->blob.size(), cudaMemcpyHostToDevice));
    dd.view = view_of(d->header, dd.blob);
    {
      // the output lists as {key id, length << 16 | tail characters}: one load per hit in the byte path's expansion
      const DfaView hv = view_of(d->header, d->blob.data());
      std::vector<uint2> out8(hv.n_out + 1);
      bool fits = true;
      for (uint64_t j = 0; j < hv.n_out && fits; ++j) {
        const int32_t v = hv.out_val[j];
        fits = v >= 0 && (uint64_t)v < hv.n_keys && hv.key_len[v] < 0x10000u && hv.key_tail[v] < 0x10000u;
        if (fits) out8[j] = make_uint2((uint32_t)v, (hv.key_len[v] << 16) | hv.key_tail[v]);
      }
      if (fits) {
        AHA_CUDA(cudaMalloc(&dd.out8, out8.size() * sizeof(uint2)));
        AHA_CUDA(cudaMemcpy(dd.out8, out8.data(), out8.size() * sizeof(uint2), cudaMemcpyHostToDevice));
      }
    }
    rc = build_hot_table(d, &dd);
    if (rc) return rc;
  }
  return AHA_OK;
}

static void free_dfa(aha_dfa* d) {
  if (!d) return;
  DeviceGuard guard;
  for (auto& dd : d->dev) {
    cudaSetDevice(dd.device);
    dd.idle.clear();  // workspaces (their destructor frees on dd.device)
    if (dd.blob) cudaFree(dd.blob);
    if (dd.hot.dev) cudaFree(dd.hot.dev);
    if (dd.folded_dev) cudaFree(dd.folded_dev);
    if (dd.out8) cudaFree(dd.out8);
  }
  delete d;
}

static int32_t finish_dfa(std::unique_ptr<aha_dfa> d, aha_dfa** out) {
  std::string err;
  if (!parse_blob(d->blob.data(), (int64_t)d->blob.size(), &d->header, &err)) return set_err(AHA_ERR_ARG, err);
  int32_t rc = upload(d.get());
  if (rc) {
    free_dfa(d.release());  // device copies made so far
    return rc;
  }
  *out = d.release();
  return AHA_OK;
}

}  // namespace aha

using namespace aha;

extern "C" {

int32_t aha_b200_version(void) { return 200; }
int64_t aha_b200_kernel_launches(void) { return (int64_t)aha::g_kernel_launches.load(); }
const char* aha_b200_last_error(void) { return g_err.c_str(); }

int32_t aha_b200_init(const int32_t* devices, int32_t n) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return set_err(AHA_ERR_CUDA, std::string("no CUDA device: libaha_b200 has no CPU fallback (") + cudaGetErrorString(e) + ")");
  knobs();
  std::lock_guard<std::mutex> lk(g_mu);
  g_devices.clear();
  if (devices == nullptr || n <= 0) {
    int cur = 0;
    AHA_CUDA(cudaGetDevice(&cur));
    g_devices.push_back(cur);
  } else {
    for (int i = 0; i < n; ++i) {
      if (devices[i] < 0 || devices[i] >= count) return set_err(AHA_ERR_ARG, "device index out of range");
      g_devices.push_back(devices[i]);
    }
  }
  return AHA_OK;
}

// ---- builder ------------------------------------------------------------------------------
int32_t aha_b200_builder_new(aha_builder** out) {
  if (!out) return set_err(AHA_ERR_ARG, "out is null");
  *out = new aha_builder();
  return AHA_OK;
}
int32_t aha_b200_builder_insert(aha_builder* b, const uint8_t* key, int32_t len, int32_t* id_out) {
  if (!b || (!key && len > 0) || len < 0) return set_err(AHA_ERR_ARG, "bad argument");
  int32_t id = b->cedar.insert(key, len);
  if (id == -2) return set_err(AHA_ERR_KEY, b->cedar.error());
  if (id_out) *id_out = id;
  return AHA_OK;
}
int32_t aha_b200_builder_lookup(const aha_builder* b, const uint8_t* key, int32_t len, int32_t* id_out) {
  if (!b || !id_out || len < 0) return set_err(AHA_ERR_ARG, "bad argument");
  *id_out = b->cedar.find(key, len);
  return AHA_OK;
}
int32_t aha_b200_builder_key(const aha_builder* b, int32_t id, uint8_t* out, int32_t cap, int32_t* len_out) {
  if (!b || !len_out) return set_err(AHA_ERR_ARG, "bad argument");
  std::string k;
  if (!b->cedar.key_of(id, &k)) return set_err(AHA_ERR_ARG, "Index out of bounds");  // IndexError, cedar.cr:832
  *len_out = (int32_t)k.size();
  if ((int32_t)k.size() > cap) return set_err(AHA_ERR_ARG, "buffer too small");
  memcpy(out, k.data(), k.size());
  return AHA_OK;
}
int32_t aha_b200_builder_compile(aha_builder* b, const void** nodes, int64_t* array_size, const int32_t** fails,
                                 const int32_t** output, const uint32_t** key_lens, int64_t* n_keys) {
  if (!b) return set_err(AHA_ERR_ARG, "builder is null");
  b->ac.compile(b->cedar);
  if (nodes) *nodes = b->cedar.nodes();
  if (array_size) *array_size = b->cedar.array_size();
  if (fails) *fails = b->ac.fails.data();
  if (output) *output = reinterpret_cast<const int32_t*>(b->ac.output.data());
  if (key_lens) *key_lens = b->ac.key_lens.data();
  if (n_keys) *n_keys = b->cedar.leaf_size();
  return AHA_OK;
}
void aha_b200_builder_free(aha_builder* b) { delete b; }

// ---- automaton ----------------------------------------------------------------------------
int32_t aha_b200_dfa_from_cedar(const void* nodes, int64_t array_size, const int32_t* fails, const int32_t* output,
                                const uint32_t* key_lens, int64_t n_keys, aha_dfa** out) {
  if (!nodes || !fails || !output || (!key_lens && n_keys > 0) || !out) return set_err(AHA_ERR_ARG, "null argument");
  std::unique_ptr<aha_dfa> d(new aha_dfa());
  std::string err;
  if (!flatten(static_cast<const CedarNode*>(nodes), array_size, fails, reinterpret_cast<const CedarOut*>(output),
               key_lens, n_keys, &d->blob, &err))
    return set_err(AHA_ERR_ARG, err);
  return finish_dfa(std::move(d), out);
}

int32_t aha_b200_flatten_host(const void* nodes, int64_t array_size, const int32_t* fails, const int32_t* output,
                              const uint32_t* key_lens, int64_t n_keys, uint8_t** blob, int64_t* bytes) {
  if (!nodes || !fails || !output || (!key_lens && n_keys > 0) || !blob || !bytes) return set_err(AHA_ERR_ARG, "null argument");
  std::vector<uint8_t> v;
  std::string err;
  if (!flatten(static_cast<const CedarNode*>(nodes), array_size, fails, reinterpret_cast<const CedarOut*>(output),
               key_lens, n_keys, &v, &err))
    return set_err(AHA_ERR_ARG, err);
  *blob = static_cast<uint8_t*>(malloc(v.size()));
  if (!*blob) return set_err(AHA_ERR_NOMEM, "malloc failed");
  memcpy(*blob, v.data(), v.size());
  *bytes = (int64_t)v.size();
  return AHA_OK;
}
void aha_b200_blob_free(uint8_t* blob) { free(blob); }

int32_t aha_b200_dfa_compile(const uint8_t* keys, const int64_t* key_off, int64_t n_keys, aha_dfa** out) {
  if ((!keys && n_keys > 0) || !key_off || !out || n_keys < 0) return set_err(AHA_ERR_ARG, "null argument");
  aha_builder b;
  for (int64_t i = 0; i < n_keys; ++i) {
    const int64_t len = key_off[i + 1] - key_off[i];
    if (len < 0 || len > INT32_MAX) return set_err(AHA_ERR_ARG, "bad key_off");
    int32_t id = b.cedar.insert(keys + key_off[i], (int32_t)len);
    if (id == -2) return set_err(AHA_ERR_KEY, b.cedar.error());
    if (id != i)  // ac.cr:66
      return set_err(AHA_ERR_KEY, "key:" + std::string(reinterpret_cast<const char*>(keys + key_off[i]), (size_t)len) + " appear twice.");
  }
  b.ac.compile(b.cedar);
  return aha_b200_dfa_from_cedar(b.cedar.nodes(), b.cedar.array_size(), b.ac.fails.data(),
                                 reinterpret_cast<const int32_t*>(b.ac.output.data()), b.ac.key_lens.data(),
                                 b.cedar.leaf_size(), out);
}

int32_t aha_b200_dfa_blob(const aha_dfa* d, const uint8_t** blob, int64_t* bytes) {
  if (!d || !blob || !bytes) return set_err(AHA_ERR_ARG, "null argument");
  *blob = d->blob.data();
  *bytes = (int64_t)d->blob.size();
  return AHA_OK;
}

int32_t aha_b200_dfa_from_blob(const uint8_t* blob, int64_t bytes, aha_dfa** out) {
  if (!blob || bytes <= 0 || !out) return set_err(AHA_ERR_ARG, "null argument");
  std::unique_ptr<aha_dfa> d(new aha_dfa());
  d->blob.assign(blob, blob + bytes);
  return finish_dfa(std::move(d), out);
}

int32_t aha_b200_dfa_from_device_blob(const void* dev_blob, int64_t bytes, aha_dfa** out) {
  if (!dev_blob || bytes <= 0 || !out) return set_err(AHA_ERR_ARG, "null argument");
  std::unique_ptr<aha_dfa> d(new aha_dfa());
  d->blob.resize((size_t)bytes);
  AHA_CUDA(cudaMemcpy(d->blob.data(), dev_blob, (size_t)bytes, cudaMemcpyDeviceToHost));
  return finish_dfa(std::move(d), out);
}

int64_t aha_b200_dfa_info(const aha_dfa* d, int32_t what) {
  if (!d) return -1;
  switch (what) {
    case 0: return (int64_t)d->header.n_states;
    case 1: return d->header.n_classes;
    case 2: return (int64_t)d->header.n_keys;
    case 3: return d->header.max_key_len;
    case 4: return (int64_t)d->header.n_out;
    case 5: return (int64_t)d->header.total_bytes;
    case 6: return d->dev.empty() ? 0 : d->dev[0].hot.n_hot;
    case 7: return 4;
    default: return -1;
  }
}

void aha_b200_dfa_free(aha_dfa* d) { free_dfa(d); }

int32_t aha_b200_debug_set(int32_t knob, int32_t value) {
  Knobs& k = knobs();
  switch (knob) {
    case 0: k.no_char = value != 0; return AHA_OK;
    case 1: k.force_v1 = value != 0; return AHA_OK;
    case 2: k.byte_chunk = value; return AHA_OK;
    case 3: k.e2e_chunk_mb = value > 0 ? value : 64; return AHA_OK;
    case 4: k.byte_span = value < 0 ? kByteSpanDefault : ((value <= 8192 && !(value & 15)) ? value : 0); return AHA_OK;  // -1: default
    default: return set_err(AHA_ERR_ARG, "unknown knob");
  }
}

int32_t aha_b200_debug_char_scan(const uint8_t* blob, int64_t blob_bytes, const uint8_t* text, int64_t n, int32_t chunk,
                                 int64_t smem_budget, aha_hit* hits, int64_t cap, int64_t* n_hits, int64_t* info) {
  if (!blob || !n_hits || !info || chunk <= 0 || n < 0 || n > INT32_MAX - 64) return set_err(AHA_ERR_ARG, "bad argument");
  BlobHeader h;
  std::string err;
  if (!parse_blob(blob, blob_bytes, &h, &err)) return set_err(AHA_ERR_ARG, err);
  const DfaView v = view_of(h, blob);
  CharTableHost T;
  if (h.has_delta) T = build_char_table(h, blob, (size_t)smem_budget);
  else T.why = "automaton too large";
  memset(info, 0, 8 * sizeof(int64_t));
  *n_hits = 0;
  if (!T.ok) return set_err(AHA_OK, T.why);
  info[0] = 1;
  info[2] = T.n_cid; info[3] = T.m; info[4] = T.n_states; info[5] = (int64_t)T.root.size(); info[6] = T.n_edges;
  info[7] = (int64_t)T.smem_bytes();
  CharTablesView tv;
  tv.root = T.root.data(); tv.check = T.check.data(); tv.val = T.val.data(); tv.gfail = T.gfail.data();
  for (int c = 0; c < 3; ++c) { tv.lo[c] = T.lo[c]; tv.n[c] = T.n[c]; tv.off[c] = T.off[c]; }
  tv.m = T.m;
  const int32_t warmc = (int32_t)(v.max_key_len > 0 ? v.max_key_len - 1 : 0) + 3;
  int64_t cnt = 0;
  bool valid = true;
  for (int64_t a = 0; a < n && valid; a += chunk) {
    const int32_t b = (int32_t)std::min<int64_t>(a + chunk, n);
    const int32_t w = (int32_t)std::max<int64_t>(0, a - warmc);
    const int st = char_walk(tv, text, w, (int32_t)a, b, b, (int32_t)n, w == 0, [&](uint32_t ref, int32_t last, uint32_t) {
      const RefRec rr = T.ref_rec[ref];
      if (rr.n == 0) { valid = false; return; }
      if (rr.n == 1) {
        if (hits && cnt < cap) hits[cnt] = aha_hit{last + 1 - (int32_t)rr.b, last + 1, (int32_t)rr.a};
        ++cnt;
        return;
      }
      for (uint32_t j = rr.a; j < rr.b; ++j) {
        const int32_t val = v.out_val[j];
        if (hits && cnt < cap) hits[cnt] = aha_hit{last + 1 - (int32_t)v.key_len[val], last + 1, val};
        ++cnt;
      }
    });
    if (st != kWalkOk) valid = false;
  }
  info[1] = valid ? 1 : 0;
  *n_hits = cnt;
  return AHA_OK;
}

// Host replay of trie_folded_doc_kernel (scan_trie.cuh) over one document: the same folded trie, the same flat state
// machine, the same candidate pick -- so that the builder's claim (one folded node = the whole frontier of
// cedar.cr:993-1019, first-of-longest = smallest mask distance) is checked against the oracle without a device.
int32_t aha_b200_debug_folded_longest(const uint8_t* blob, int64_t blob_bytes, const uint8_t* text, int64_t n64, uint32_t mode,
                                      const uint8_t* sep_bits, int32_t sep_size, aha_hit* hits, int64_t cap, int64_t* n_hits,
                                      int64_t* info) {
  if (!blob || !n_hits || !info || n64 < 0 || n64 > INT32_MAX - 64 || (n64 && !text)) return set_err(AHA_ERR_ARG, "bad argument");
  if ((mode & AHA_SEP) && (!sep_bits || sep_size < 0 || sep_size > 256)) return set_err(AHA_ERR_ARG, "bad separator set");
  BlobHeader h;
  std::string err;
  if (!parse_blob(blob, blob_bytes, &h, &err)) return set_err(AHA_ERR_ARG, err);
  memset(info, 0, 8 * sizeof(int64_t));
  *n_hits = 0;
  const FoldedTrieHost F = build_folded_trie(h, blob);
  if (!F.ok) return set_err(AHA_OK, "no folded trie for this dictionary");
  info[0] = 1; info[1] = F.n_states; info[2] = F.n_classes; info[3] = F.max_chars; info[4] = (int64_t)F.cand.size();
  for (uint32_t i = 0; i < F.n_states; ++i) info[5] += F.rec[(size_t)i * 8] == 0xFFFFFFFFu;
  const bool use_sep = (mode & AHA_SEP) != 0;
  auto sep_at = [&](uint32_t c) { return c < 256u && (int32_t)c < sep_size && ((sep_bits[c >> 3] >> (c & 7u)) & 1u); };
  const int32_t n = (int32_t)n64;
  const uint8_t* s = text;
  auto decode = [&](int32_t i, uint32_t* cp) -> int {  // utf8_decode32
    const uint32_t b = s[i];
    if (b < 0x80) { *cp = b; return 1; }
    if ((b & 0xE0) == 0xC0 && i + 1 < n) { *cp = ((b & 0x1F) << 6) | (s[i + 1] & 0x3F); return 2; }
    if ((b & 0xF0) == 0xE0 && i + 2 < n) { *cp = ((b & 0x0F) << 12) | ((s[i + 1] & 0x3F) << 6) | (s[i + 2] & 0x3F); return 3; }
    if ((b & 0xF8) == 0xF0 && i + 3 < n) {
      *cp = ((b & 0x07) << 18) | ((s[i + 1] & 0x3F) << 12) | ((s[i + 2] & 0x3F) << 6) | (s[i + 3] & 0x3F);
      return 4;
    }
    *cp = b;
    return 1;
  };
  auto pick = [&](uint32_t nd, uint64_t window) -> int32_t {  // folded_pick
    const uint32_t* r = &F.rec[(size_t)nd * 8];
    int32_t best_v = (int32_t)r[1];
    if (r[0] != 0xFFFFFFFFu) {
      const uint32_t w = (uint32_t)window;
      uint32_t bx = r[4] ^ w;
      if (r[0] > 1u && (r[5] ^ w) < bx) { bx = r[5] ^ w; best_v = (int32_t)r[2]; }
      if (r[0] > 2u && (r[6] ^ w) < bx) best_v = (int32_t)r[3];
    } else {
      uint64_t best_x = ~0ull;
      for (uint32_t j = F.cand_off[nd]; j < F.cand_off[nd + 1]; ++j) {
        const uint64_t x = F.cand[j].mask ^ window;
        if (x < best_x || j == F.cand_off[nd]) { best_x = x; best_v = F.cand[j].value; }
      }
    }
    return best_v;
  };
  int64_t run = 0;
  int32_t pos = 0, c_bpos = 0, cidx = 0, best_n = 0, best_bytes = 0;
  int phase = 0, depth = 0, first_l = 1;
  uint32_t node = 0, best_node = 0;
  uint64_t tmask = 0, best_window = 0;
  bool pending = false;
  while (phase != 3) {
    const bool walking = phase == 1;
    const int32_t cur = walking ? c_bpos : pos;
    const bool have = cur < n;
    uint32_t cp = 0;
    int l = 1;
    if (have) l = decode(cur, &cp);
    const bool sep = have && sep_at(cp);
    if (pending && (!use_sep || !have || sep)) {
      best_node = node;
      best_window = depth >= 64 ? tmask : (tmask & ((1ull << depth) - 1ull));
      best_n = depth;
      best_bytes = c_bpos - pos;
    }
    pending = false;
    if (!walking && !have) break;
    const bool skip = !walking && (phase == 2 || (use_sep && sep));
    const bool start = !walking && !skip;
    if (skip) { pos += l; ++cidx; if (sep) phase = 0; }
    if (start) { phase = 1; depth = 0; node = 0; c_bpos = pos; tmask = 0; best_n = 0; first_l = l; }
    const bool step = walking || start;
    uint32_t e = 0;
    bool upper = false;
    if (step && have && depth < 64) {
      upper = fold_is_upper(cp);
      const uint32_t fc = upper ? cp + 32u : cp;
      std::string enc;
      fold_encode(fc, &enc);
      e = fc ? node : 0u;
      for (size_t j = 0; j < enc.size() && fc; ++j) {
        e = F.delta[(size_t)(e & ~kFoldCand) * F.n_classes + F.cmap[(uint8_t)enc[j]]];
        if (e == 0u) break;
      }
    }
    const bool adv = e != 0u;
    if (adv) {
      node = e & ~kFoldCand;
      c_bpos += l;
      ++depth;
      tmask = (tmask << 1) | (upper ? 1ull : 0ull);
      pending = (e & kFoldCand) != 0u;
    }
    if (!step || adv) continue;
    if (best_n == 0) {
      pos += first_l;
      ++cidx;
      phase = use_sep ? 2 : 0;
      continue;
    }
    if (hits && run < cap) hits[run] = aha_hit{cidx, cidx + best_n, pick(best_node, best_window)};
    ++run;
    pos += best_bytes;
    cidx += best_n;
    phase = 0;
  }
  *n_hits = run;
  return AHA_OK;
}

}  // extern "C"

#include "api_scan.cuh"
