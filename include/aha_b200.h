/* aha_b200.h -- C ABI of libaha_b200.so, the B200 drop-in for the Aha::AC scan path.
 *
 * The reference (chenkovsky/aha) is pure Crystal and has no FFI boundary; this header IS the
 * boundary a Crystal `lib LibAhaB200` block binds (see src/aha_gpu.cr and INTEGRATION.md).
 * Every entry point names the reference code it replaces.  All functions return 0 on success,
 * a negative aha_status otherwise; aha_b200_last_error() returns a thread-local message that
 * the host wrapper turns into `raise` (the reference raises strings: ac.cr:66,322; cedar.cr:235,756).
 *
 * Ownership: inputs are borrowed for the duration of the call.  aha_dfa / aha_hits are owned by
 * the library and released only by the *_free calls.  Offsets in aha_hit are per document and
 * Int32 like Aha::Hit (matcher.cr:2-11): a document must be < 2^31 bytes.
 */
#ifndef AHA_B200_H
#define AHA_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct aha_dfa aha_dfa;         /* flattened automaton, host copy + per-device copies */
typedef struct aha_builder aha_builder; /* host cedar trie under construction (Aha::Cedar)   */

/* == Aha::Hit (matcher.cr:2-11): [start, end) half open, value = key id */
typedef struct aha_hit {
  int32_t start, end, value;
} aha_hit;

/* result of one scan call */
typedef struct aha_hits {
  aha_hit* hits;          /* n_hits records, document order, reference emission order        */
  int64_t n_hits;
  int64_t* doc_hit_off;   /* n_docs + 1 offsets into hits[] (exclusive prefix of per-doc counts) */
  int64_t n_docs;
  int32_t on_device;      /* 1: hits/doc_hit_off are device pointers (aha_b200_scan_device)  */
  int32_t device;
  double scan_ms;         /* device time of all launches of the scan (CUDA events)           */
  double walk_ms;         /* of which char_walk_kernel; 0 on other paths                     */
  double expand_ms;       /* of which char_expand_kernel; 0 on other paths                   */
  void* opaque;
} aha_hits;

enum aha_status {
  AHA_OK = 0,
  AHA_ERR_ARG = -1,       /* bad argument / contract violation (doc >= 2^31 bytes, sep_size > 256 ...) */
  AHA_ERR_CUDA = -2,      /* CUDA runtime error */
  AHA_ERR_KEY = -3,       /* reference raise: duplicate key, empty key, NUL in key */
  AHA_ERR_NOMEM = -4,
  AHA_ERR_STATE = -5      /* library not initialised / dfa not on this device */
};

/* scan mode bits */
enum aha_mode {
  AHA_UNITS_CHAR = 1,     /* String overloads: offsets are character indices (matcher.cr:34-46, ac.cr:305-310) */
  AHA_SEP = 2,            /* ACX#match(seq, sep : BitArray)  (ac.cr:321-340)                           */
  AHA_LONGEST = 4,        /* ACX#match_longest               (ac.cr:118-143, 297-303)                  */
  AHA_INTERSECTABLE = 8,  /* match_longest(seq, true)        (ac.cr:131-137)                           */
  AHA_TRIE_LONGEST = 16,  /* CedarX#match_longest(String, ignore_case, sep) (cedar.cr:1075-1121)       */
  AHA_IGNORE_CASE = 32,   /* with AHA_TRIE_LONGEST           (cedar.cr:995-996)                        */
  AHA_COUNT_ONLY = 64,    /* no hit records, only doc_hit_off (bench / sizing)                         */
  AHA_SEP_CODEPOINT = 128 /* with AHA_SEP: the Array(Char) / Slice(Char) forms of match(seq, sep) (ac.cr:342-364):
                             the neighbour test is on the CODE POINT of the neighbouring character, not on
                             its byte, so a neighbour >= U+0100 (>= sep.size) never vetoes; used with
                             AHA_UNITS_CHAR (those forms yield character offsets)                        */
};

/* ---- library / devices ------------------------------------------------------------------- */
/* Selects the CUDA devices this process scans on.  One per process under torchrun; with n > 1 a
 * single process (a Crystal host has one thread) drives several GPUs: automata built afterwards are
 * uploaded to every selected device, and aha_b200_scan deals the batch to them in runs of whole
 * documents, one host thread and three streams per device, hits landing in document order in one
 * pinned buffer.  devices == NULL, n == 0 -> current device only.  Also reads the AHA_B200_*
 * environment knobs (once).                                                                   */
int32_t aha_b200_init(const int32_t* devices, int32_t n);
const char* aha_b200_last_error(void);
int32_t aha_b200_version(void);
/* number of CUDA kernels this library has launched so far in the process (diagnostics / bench accounting) */
int64_t aha_b200_kernel_launches(void);

/* ---- host builder: Aha::Cedar#insert + Aha::AC.compile where Crystal is absent ----------- */
/* Aha::Cedar.new (cedar.cr:198-221) */
int32_t aha_b200_builder_new(aha_builder** out);
/* Aha::Cedar#insert(Bytes) (cedar.cr:755-778); *id_out = key id.  AHA_ERR_KEY on empty key / NUL byte. */
int32_t aha_b200_builder_insert(aha_builder* b, const uint8_t* key, int32_t len, int32_t* id_out);
/* Aha::Cedar#[]?(Bytes) (cedar.cr:822-828); *id_out = -1 when absent */
int32_t aha_b200_builder_lookup(const aha_builder* b, const uint8_t* key, int32_t len, int32_t* id_out);
/* Aha::Cedar#[](Int) (cedar.cr:747-749): key bytes of id into out[cap]; *len_out = length */
int32_t aha_b200_builder_key(const aha_builder* b, int32_t id, uint8_t* out, int32_t cap, int32_t* len_out);
/* raw views of the arrays the reference holds after AC.compile (ac.cr:35-39); valid until the
 * next insert / free.  nodes: 12-byte Node(Int32)[array_size]; fails: Int32[array_size];
 * output: {next,value}[array_size]; key_lens: UInt32[n_keys].  Runs ACX.compile(da) (ac.cr:71-112). */
int32_t aha_b200_builder_compile(aha_builder* b, const void** nodes, int64_t* array_size, const int32_t** fails,
                                 const int32_t** output, const uint32_t** key_lens, int64_t* n_keys);
void aha_b200_builder_free(aha_builder* b);

/* ---- automaton --------------------------------------------------------------------------- */
/* Flattens the arrays of a compiled Aha::AC (@da.array, @fails, @output, @key_lens: ac.cr:35-39,
 * cedar.cr:29-36,147-148) into the automaton blob (sparse 16-byte rows, closure rows for the wide
 * states, a dense table only while it stays small) and uploads it to every selected device.        */
int32_t aha_b200_dfa_from_cedar(const void* nodes, int64_t array_size, const int32_t* fails, const int32_t* output,
                                const uint32_t* key_lens, int64_t n_keys, aha_dfa** out);
/* Aha::AC.compile(keys) (ac.cr:62-69) in one call: keys concatenated, key_off has n_keys+1 entries.
 * AHA_ERR_KEY with "key:... appear twice." on a duplicate, like the reference.                */
int32_t aha_b200_dfa_compile(const uint8_t* keys, const int64_t* key_off, int64_t n_keys, aha_dfa** out);
/* host-only flatten (no device needed): malloc'ed blob, release with aha_b200_blob_free */
int32_t aha_b200_flatten_host(const void* nodes, int64_t array_size, const int32_t* fails, const int32_t* output,
                              const uint32_t* key_lens, int64_t n_keys, uint8_t** blob, int64_t* bytes);
void aha_b200_blob_free(uint8_t* blob);
/* serialised DFA (host memory, library-owned until dfa_free) for broadcast / caching */
int32_t aha_b200_dfa_blob(const aha_dfa* d, const uint8_t** blob, int64_t* bytes);
/* rebuild from a blob in host memory (e.g. received on another rank) and upload */
int32_t aha_b200_dfa_from_blob(const uint8_t* blob, int64_t bytes, aha_dfa** out);
/* rebuild from a blob already resident on the current device (after ncclBroadcast) */
int32_t aha_b200_dfa_from_device_blob(const void* dev_blob, int64_t bytes, aha_dfa** out);
/* info: 0 n_states, 1 n_classes, 2 n_keys, 3 max_key_len, 4 n_out_entries, 5 blob bytes,
 *       6 states of the character automaton resident in shared memory (0: byte path),
 *       7 entry width of the global table in bytes                                           */
int64_t aha_b200_dfa_info(const aha_dfa* d, int32_t what);
void aha_b200_dfa_free(aha_dfa* d);

/* ---- scan -------------------------------------------------------------------------------- */
/* Batch form of ACX#match / match(sep) / match_longest (ac.cr:280-340) and
 * CedarX#match_longest (cedar.cr:1075-1121) over n_docs documents stored back to back in HOST
 * memory: document i = text[doc_off[i] .. doc_off[i+1]).  Copies the text to the device(s),
 * scans, copies the hits back (pinned).  sep_bits: 256-bit set, bit b of byte b>>3; sep_size =
 * BitArray size (bytes >= sep_size always pass: ac.cr:326,333).                               */
int32_t aha_b200_scan(const aha_dfa* d, const uint8_t* text, const int64_t* doc_off, int64_t n_docs, uint32_t mode,
                      const uint8_t* sep_bits, int32_t sep_size, aha_hits** out);
/* Same, text and doc_off already in the current device's memory; results stay on the device.
 * stream: a cudaStream_t (0 = default stream).  doc_off is validated on the device (must start at
 * 0, be non-decreasing, end at total_bytes, no document >= 2^31 bytes): AHA_ERR_ARG otherwise.
 * The call returns when the scan has finished (the reference call is synchronous, ac.cr:280-286);
 * inside it nothing waits for the host: a warm call makes no allocation and one host wait.   */
int32_t aha_b200_scan_device(const aha_dfa* d, const void* dev_text, const void* dev_doc_off, int64_t n_docs,
                             int64_t total_bytes, uint32_t mode, const uint8_t* sep_bits, int32_t sep_size,
                             void* stream, aha_hits** out);
void aha_b200_hits_free(aha_hits* h);
/* ACX#match (byte offsets, mode 0) over a batch on the current device WITHOUT any allocation.  No counterpart in
 * the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  The caller passes every buffer:
 *   dev_hits         aha_hit[hit_cap] (device); hits beyond hit_cap are counted but not written
 *   dev_doc_hit_off  int64[n_docs + 1] (device), as aha_hits.doc_hit_off
 *   dev_n_hits       int64[1] (device): total hits; > hit_cap means the buffer was too small
 * The per-chunk scratch of the two-pass scan is a static __device__ array of about 1.5 MiB that every GPU which
 * loads the library holds until exit: at most 64 MiB of text per call (AHA_ERR_ARG beyond that).  The call
 * enqueues three kernels on `stream` and returns without waiting (results are complete when the stream reaches
 * that point); it makes no cudaMalloc / cudaFree / cudaMallocHost, creates no stream or event, and calls no
 * allocating library routine.  doc_off must satisfy the contract of aha_b200_scan_device (it is NOT validated
 * here: there is no host wait to report it through).  All calls on one device share the scratch: calls that
 * overlap on the GPU (different streams, concurrently launched graphs) race; serialise them.  The first call
 * on a device resolves the scratch's address (cudaGetSymbolAddress) and should be made outside capture.       */
int32_t aha_b200_scan_device_noalloc(const aha_dfa* d, const void* dev_text, const void* dev_doc_off, int64_t n_docs,
                                     int64_t total_bytes, void* dev_hits, int64_t hit_cap, void* dev_doc_hit_off,
                                     void* dev_n_hits, void* stream);

/* ---- diagnostics ------------------------------------------------------------------------- */
/* Test knobs (also read once from the environment): 0 = AHA_B200_NO_CHAR (byte automaton even when
 * the character tables fit shared memory), 1 = two-pass checker kernels (AHA_B200_IMPL=0; otherwise
 * only the path of keys > 1024 bytes), 2 = AHA_B200_BYTE_CHUNK, 3 = AHA_B200_E2E_CHUNK_MB,
 * 4 = AHA_B200_BYTE_SPAN (bytes per lane of the byte kernel's span mode, 0 = off, -1 = default). */
int32_t aha_b200_debug_set(int32_t knob, int32_t value);
/* Host-side replay of the shared-memory character automaton the fused kernel uses (same table builder,
 * same per-lane walk, lanes emulated sequentially with `chunk`-byte owned regions) over ONE document.
 * info[0] = 1 if the dictionary is eligible (whole UTF-8 units, fits `smem_budget`), info[1] = 1 if the
 * text was structurally valid UTF-8, info[2..7] = n_cid, slots, b1, root entries, edges, table bytes.
 * Writes at most cap hits (byte offsets); *n_hits = number found.  No device needed.            */
int32_t aha_b200_debug_char_scan(const uint8_t* blob, int64_t blob_bytes, const uint8_t* text, int64_t n, int32_t chunk,
                                 int64_t smem_budget, aha_hit* hits, int64_t cap, int64_t* n_hits, int64_t* info);
/* Host-side replay of the case-folded trie walk behind CedarX#match_longest(ignore_case: true) (cedar.cr:984-1027,
 * 1075-1121; aha_b200/csrc/folded_trie.hpp): the same table builder, state machine and candidate pick as the kernel,
 * over ONE document; mode = AHA_TRIE_LONGEST | AHA_IGNORE_CASE [| AHA_SEP], sep_bits / sep_size as in aha_b200_scan.
 * info[0] = 1 if the dictionary has a folded trie (keys <= 64 characters, valid UTF-8, table <= 256 MB),
 * info[1..5] = folded nodes, byte classes, longest key in characters, candidates, nodes with an overflow record.
 * Writes at most cap hits (character offsets); *n_hits = number found.  No device needed.       */
int32_t aha_b200_debug_folded_longest(const uint8_t* blob, int64_t blob_bytes, const uint8_t* text, int64_t n, uint32_t mode,
                                      const uint8_t* sep_bits, int32_t sep_size, aha_hit* hits, int64_t cap, int64_t* n_hits,
                                      int64_t* info);

/* ---- synthetic corpora: counter-based, identical bytes on host and device -- */
enum aha_corpus { AHA_CORPUS_EN = 0, AHA_CORPUS_ZH = 1, AHA_CORPUS_MIX = 2 };
/* fills out[0..bytes) with corpus text starting at absolute byte `first` of the stream */
int32_t aha_b200_gen_text_host(int32_t corpus, uint64_t seed, int64_t first, int64_t bytes, uint8_t* out);
int32_t aha_b200_gen_text_device(int32_t corpus, uint64_t seed, int64_t first, int64_t bytes, void* dev_out,
                                 void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AHA_B200_H */
