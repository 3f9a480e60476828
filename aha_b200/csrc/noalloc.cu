// noalloc.cu -- aha_b200_scan_device_noalloc: ACX#match over a batch already on the device, for callers that must not
// allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// A translation unit of its own: nothing here is on any other path.
//
// The two-pass scan (count per chunk -> prefix -> emit) needs per-chunk scratch.  Here that scratch is a file-scope
// __device__ array of a fixed size (kNaScratchBytes, about 1.5 MiB, held by every GPU that loads the library until the
// process exits), so the call makes no cudaMalloc / cudaFree / cudaMallocHost and uses no allocating library routine;
// it creates no stream or event and never waits for the device: three kernels go to the caller's stream and the call
// returns.  Limit: kNaMaxChunks chunks of kNaChunk bytes = 64 MiB of text per call; larger batches get AHA_ERR_ARG.
//
// All calls on one device share the array.  It keeps the previous call's contents: every word a call reads is written
// by that call first (the count kernel writes counts[0 .. n_chunks], the prefix kernel offsets[0 .. n_chunks]),
// so nothing needs a memset.  Two calls that overlap on the GPU (different streams, or graphs launched
// concurrently) race on it: callers serialise them.
#include <cstring>
#include <mutex>

#include "api_internal.cuh"
#include "scan_common.cuh"

namespace aha {

constexpr int kNaChunk = 512;                       // bytes of text per thread, as in the two-pass checker (scan_v1.cuh)
constexpr int64_t kNaMaxChunks = (int64_t)1 << 17;  // 64 MiB of text per call
constexpr int64_t kNaMaxBytes = kNaMaxChunks * kNaChunk;
constexpr int kNaMaxDevices = 64;

struct NaScratch {
  unsigned long long offsets[kNaMaxChunks + 1];  // exclusive prefix of counts[]
  uint32_t counts[kNaMaxChunks + 1];             // hits of every chunk's owned region
  uint32_t pad[3];
};
constexpr size_t kNaScratchBytes = sizeof(NaScratch);

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress: the kernels get the device
// address that call returns for the current device.
static __device__ __align__(256) unsigned char g_na_scratch[kNaScratchBytes];

struct NaParams {
  ScanParams sp;       // text, doc_off, automaton (counts / offsets point into the scratch)
  uint32_t* counts;
  unsigned long long* offsets;
  int64_t hit_cap;     // records the caller's hit buffer holds
  int64_t* n_hits;     // device: total number of hits (may exceed hit_cap: the excess is not written)
};

// One thread per chunk: warm-up over the max_key_len - 1 bytes before the chunk (never across a document start), then
// the automaton over the chunk's bytes; EMIT = false counts the hits of the chunk, EMIT = true writes them at the
// chunk's offset and the per-document offsets as it crosses document starts.
template <bool EMIT>
__global__ void __launch_bounds__(128) na_walk_kernel(const __grid_constant__ NaParams q) {
  const ScanParams& p = q.sp;
  __shared__ uint8_t cmap[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) cmap[i] = p.classmap[i];
  __syncthreads();
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= p.n_chunks) return;
  const int64_t a = g * kNaChunk;
  const int64_t b = min(a + (int64_t)kNaChunk, p.total);
  int64_t d = last_le(p.doc_off, p.n_docs, a);
  if (d >= p.n_docs) d = p.n_docs - 1;
  int64_t doc_start = p.doc_off[d], doc_end = p.doc_off[d + 1];
  unsigned long long run = EMIT ? q.offsets[g] : 0ull;
  if (EMIT && doc_start == a)  // documents starting exactly here (possibly several empty ones) belong to this chunk
    for (int64_t e = first_ge(p.doc_off, p.n_docs, a); e <= d; ++e) p.doc_hit_off[e] = (int64_t)run;
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
      const uint32_t e = delta_entry(p, cmap, s, p.text[pos]);
      s = e & kStateMask;
      if (e & kOut) {  // ACX#fetch (ac.cr:265-278): every key ending here, longest first
        const uint32_t j0 = p.out_off[s], j1 = p.out_off[s + 1];
        if (EMIT && p.hits) {
          const int32_t end = (int32_t)(pos - doc_start + 1);
          for (uint32_t j = j0; j < j1; ++j) {
            const unsigned long long at = run + (j - j0);
            if (at >= (unsigned long long)q.hit_cap) break;
            const int32_t v = p.out_val[j];
            p.hits[at] = aha_hit{end - (int32_t)p.key_len[v], end, v};
          }
        }
        run += j1 - j0;
      }
    }
    if (pos == doc_end && (pos < b || b == p.total)) {  // into the next non-empty document: the automaton restarts
      s = 0;
      ++d;
      while (d < p.n_docs) {
        if (EMIT) p.doc_hit_off[d] = (int64_t)run;
        if (p.doc_off[d + 1] > pos) break;
        ++d;
      }
      if (d >= p.n_docs) break;
      doc_start = p.doc_off[d];
      doc_end = p.doc_off[d + 1];
    }
  }
  if (!EMIT) q.counts[g] = (uint32_t)run;
}

// exclusive prefix of counts[0 .. n) in one block (n <= 2^17: a few microseconds); offsets[n] = total
__global__ void __launch_bounds__(1024) na_prefix_kernel(const __grid_constant__ NaParams q) {
  __shared__ unsigned long long warp_sum[32];
  __shared__ unsigned long long carry;
  const int64_t n = q.sp.n_chunks;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < n; base += blockDim.x) {
    const int64_t i = base + threadIdx.x;
    const unsigned long long x = i < n ? q.counts[i] : 0u;
    unsigned long long incl = x;
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      unsigned long long w = warp_sum[lane], wi = w;
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += y;
      }
      warp_sum[lane] = wi - w;  // exclusive
    }
    __syncthreads();
    const unsigned long long c = carry;
    if (i < n) q.offsets[i] = c + warp_sum[warp] + incl - x;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = c + warp_sum[warp] + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    q.offsets[n] = carry;
    *q.n_hits = (int64_t)carry;
    q.sp.doc_hit_off[q.sp.n_docs] = (int64_t)carry;
    if (n == 0)
      for (int64_t d = 0; d < q.sp.n_docs; ++d) q.sp.doc_hit_off[d] = 0;  // only empty documents
  }
}

// device address of the scratch, resolved once per device with that device current
static std::mutex g_na_mu;
static NaScratch* g_na_ptr[kNaMaxDevices];

static int32_t na_scratch_on(int device, NaScratch** out) {
  if (device < 0 || device >= kNaMaxDevices) return set_err(AHA_ERR_ARG, "scan_device_noalloc: device index out of range");
  std::lock_guard<std::mutex> lk(g_na_mu);
  if (!g_na_ptr[device]) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_na_scratch);  // the current device is `device` (checked by the caller)
    if (e != cudaSuccess || !ptr) {
      cudaGetLastError();
      return set_err(AHA_ERR_CUDA, std::string("cudaGetSymbolAddress(g_na_scratch): ") + cudaGetErrorString(e));
    }
    g_na_ptr[device] = static_cast<NaScratch*>(ptr);
  }
  *out = g_na_ptr[device];
  return AHA_OK;
}

}  // namespace aha

extern "C" int32_t aha_b200_scan_device_noalloc(const aha_dfa* d, const void* dev_text, const void* dev_doc_off,
                                                int64_t n_docs, int64_t total_bytes, void* dev_hits, int64_t hit_cap,
                                                void* dev_doc_hit_off, void* dev_n_hits, void* stream) {
  using namespace aha;
  if (!d || !dev_doc_off || !dev_doc_hit_off || !dev_n_hits || n_docs < 0 || total_bytes < 0 || hit_cap < 0 ||
      (total_bytes > 0 && !dev_text) || (hit_cap > 0 && !dev_hits))
    return set_err(AHA_ERR_ARG, "scan_device_noalloc: null or negative argument");
  if (n_docs == 0 && total_bytes != 0) return set_err(AHA_ERR_ARG, "scan_device_noalloc: text without documents");
  if (total_bytes > kNaMaxBytes)
    return set_err(AHA_ERR_ARG, "scan_device_noalloc: " + std::to_string(total_bytes) + " bytes of text exceed the " +
                                    std::to_string(kNaMaxBytes) + " bytes (64 MiB) the static scratch is sized for; "
                                    "split the batch or use aha_b200_scan_device");
  int device = -1;
  AHA_CUDA(cudaGetDevice(&device));
  DeviceDfa* dd = d->on(device);
  if (!dd) return set_err(AHA_ERR_STATE, "dfa not uploaded to the current device");
  NaScratch* scratch = nullptr;
  int32_t rc = na_scratch_on(device, &scratch);
  if (rc) return rc;

  NaParams q;
  memset(&q, 0, sizeof(q));
  ScanParams& p = q.sp;
  p.text = static_cast<const uint8_t*>(dev_text);
  p.total = total_bytes;
  p.doc_off = static_cast<const int64_t*>(dev_doc_off);
  p.n_docs = n_docs;
  p.classmap = dd->view.classmap;
  p.delta = dd->view.has_delta ? dd->view.delta : nullptr;
  p.rows = reinterpret_cast<const uint4*>(dd->view.rows);
  p.wide = dd->view.wide;
  p.wide_fail = dd->view.wide_fail;
  p.n_wide_first = dd->view.wide_first ? (uint32_t)dd->view.n_wide : 0u;
  p.out_off = dd->view.out_off;
  p.out_val = dd->view.out_val;
  p.key_len = dd->view.key_len;
  p.key_tail = dd->view.key_tail;
  p.C = dd->view.n_classes;
  p.warm = dd->view.max_key_len > 0 ? dd->view.max_key_len - 1 : 0;
  p.chunk = kNaChunk;
  p.n_chunks = (total_bytes + kNaChunk - 1) / kNaChunk;
  p.hits = static_cast<aha_hit*>(dev_hits);
  p.doc_hit_off = static_cast<int64_t*>(dev_doc_hit_off);
  q.counts = scratch->counts;
  q.offsets = scratch->offsets;
  q.hit_cap = hit_cap;
  q.n_hits = static_cast<int64_t*>(dev_n_hits);

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const unsigned grid = (unsigned)((p.n_chunks + 127) / 128);
  int launches = 1;
  if (p.n_chunks > 0) na_walk_kernel<false><<<grid, 128, 0, st>>>(q);
  na_prefix_kernel<<<1, 1024, 0, st>>>(q);
  if (p.n_chunks > 0) {
    na_walk_kernel<true><<<grid, 128, 0, st>>>(q);
    launches = 3;
  }
  AHA_CUDA(cudaGetLastError());
  g_kernel_launches.fetch_add(launches, std::memory_order_relaxed);
  return AHA_OK;
}
