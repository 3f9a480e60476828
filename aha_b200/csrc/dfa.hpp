// dfa.hpp -- flattening of a compiled Aha::AC (cedar double array + @fails + @output + @key_lens)
// into the automaton the GPU scans, and its serialised ("blob") form.
//
// Two forms of the same automaton live in the blob:
//   * SPARSE ROWS (always): 16 bytes per state = {fail link, flags, up to three goto edges {label, target}}; the one
//     state in a hundred that has more than three children (and the root) owns a dense row of `wide[]` instead.  Of the
//     states of a 100 k-pattern dictionary 83 % have one child and 13 % none; the rows of the 1 M-pattern dictionary
//     are 100 MB where its dense table is 3.1 GB.  The match kernel of large dictionaries walks these: one 16-byte
//     load per state visited (goto, else the fail link's row), exactly the loop of ACX#match_ (ac.cr:176-192).
//   * DENSE TABLE delta[s][c] (only while S * C * 4 <= 128 MB): the closure of that loop, one lookup per byte; the
//     shared-memory character automaton is derived from it, the per-document machines and small dictionaries use it.
//
// delta[s][c] is the closure of the goto/fail loop of ACX#match_ (ac.cr:176-192):
//     delta(s,b) = child(s,b)                      if the goto edge exists (cedar.cr:441-447)
//                = 0                               if s is the root
//                = delta(fails[s], b)              otherwise
// States are the label != 0 trie nodes, renumbered in BFS order (root = 0) so that shallow (hot)
// states have small ids.  Every entry carries three bits about its TARGET state / edge:
//     kOut  target has a non-empty output list (ACX#fetch, ac.cr:265-278, rule Q1)
//     kEnd  target is_end? (raw flag bits, cedar.cr:657-660, Q2/Q4) -- drives match_longest
//     kGoto the edge is a real goto edge (not derived through fail links)
#pragma once
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "host_cedar.hpp"

namespace aha {

constexpr uint32_t kOut = 0x80000000u;
constexpr uint32_t kEnd = 0x40000000u;
constexpr uint32_t kGoto = 0x20000000u;
constexpr uint32_t kStateMask = 0x1FFFFFFFu;
constexpr uint64_t kBlobMagic = 0x3130414644414841ull;  // "AHADFA01"
constexpr uint64_t kDenseMaxBytes = (uint64_t)128 << 20;  // the dense table is kept only up to this size
constexpr uint64_t kSparseMaxStates = (uint64_t)1 << 23;  // an edge carries its target in 23 bits

// sparse row of one state.  head = fail << 8 | flags; edge = target << 9 | target-has-output << 8 | label (0 = no edge)
// wide rows: e[0] = index of the state's dense row in wide[] (entries in edge format, label field = 1 when present)
struct SparseRow {
  uint32_t head;
  uint32_t e[3];
};
constexpr uint32_t kRowOut = 1u, kRowWide = 2u, kRowEnd = 4u;
constexpr uint32_t kEdgeOut = 0x100u;

struct BlobHeader {
  uint64_t magic;
  uint32_t version;
  uint32_t n_classes;
  uint64_t n_states;
  uint64_t n_keys;
  uint64_t n_out;
  uint32_t max_key_len;
  uint32_t reserved;
  uint64_t off_classmap;  // u8[256]
  uint64_t off_delta;     // u32[n_states * n_classes]
  uint64_t off_out_off;   // u32[n_states + 1]
  uint64_t off_out_val;   // i32[n_out]
  uint64_t off_key_len;   // u32[n_keys]
  uint64_t off_key_tail;  // u32[n_keys]: non-continuation bytes in key[1..]: char-unit start remap
  uint64_t off_depth;     // u16[n_states]: depth in bytes (diagnostics, hot-set selection)
  uint64_t off_fail;      // u32[n_states]: failure link (ac.cr:96-105) in state numbering
  uint64_t total_bytes;
  uint64_t off_rows;      // SparseRow[n_states]
  uint64_t off_wide;      // u32[n_wide * n_classes]
  uint64_t n_wide;
  uint64_t has_delta;     // 1: off_delta holds the dense table; 0: too large, the section is empty
  uint64_t wide_first;    // 1: states are numbered wide states first (state id < n_wide <=> wide, wide row = id);
                          //    set when there is no dense table: the walk then reads a wide state's edge with ONE load
  uint64_t off_wide_fail; // u32[n_wide]: fail link of the wide states (wide_first only)
};

inline uint64_t align16(uint64_t x) { return (x + 15) & ~uint64_t(15); }

// View over a blob (host or device pointer arithmetic only)
struct DfaView {
  const uint8_t* classmap;
  const uint32_t* delta;
  const uint32_t* out_off;
  const int32_t* out_val;
  const uint32_t* key_len;
  const uint32_t* key_tail;
  const uint16_t* depth;
  const uint32_t* fail;
  const SparseRow* rows;
  const uint32_t* wide;
  const uint32_t* wide_fail;
  uint64_t n_wide;
  bool has_delta, wide_first;
  uint32_t n_classes;
  uint32_t max_key_len;
  uint64_t n_states, n_keys, n_out;
};

inline DfaView view_of(const BlobHeader& h, const uint8_t* base) {
  DfaView v;
  v.classmap = base + h.off_classmap;
  v.delta = reinterpret_cast<const uint32_t*>(base + h.off_delta);
  v.out_off = reinterpret_cast<const uint32_t*>(base + h.off_out_off);
  v.out_val = reinterpret_cast<const int32_t*>(base + h.off_out_val);
  v.key_len = reinterpret_cast<const uint32_t*>(base + h.off_key_len);
  v.key_tail = reinterpret_cast<const uint32_t*>(base + h.off_key_tail);
  v.depth = reinterpret_cast<const uint16_t*>(base + h.off_depth);
  v.fail = reinterpret_cast<const uint32_t*>(base + h.off_fail);
  v.rows = reinterpret_cast<const SparseRow*>(base + h.off_rows);
  v.wide = reinterpret_cast<const uint32_t*>(base + h.off_wide);
  v.n_wide = h.n_wide;
  v.has_delta = h.has_delta != 0;
  v.wide_first = h.wide_first != 0;
  v.wide_fail = reinterpret_cast<const uint32_t*>(base + h.off_wide_fail);
  v.n_classes = h.n_classes;
  v.max_key_len = h.max_key_len;
  v.n_states = h.n_states;
  v.n_keys = h.n_keys;
  v.n_out = h.n_out;
  return v;
}

// Static chunks of [0, n) on up to `threads` host threads: fn(begin, end, thread index).  The flatten of a large
// dictionary is a handful of passes over millions of states whose cost is the random reads of the cedar arrays.
template <class F>
inline void parallel_chunks(uint64_t n, int threads, F fn) {
  if (threads <= 1 || n < 8192) {
    fn((uint64_t)0, n, 0);
    return;
  }
  std::vector<std::thread> pool;
  const uint64_t per = (n + (uint64_t)threads - 1) / (uint64_t)threads;
  for (int t = 0; t < threads; ++t) {
    const uint64_t b = (uint64_t)t * per, e = std::min<uint64_t>(n, b + per);
    if (b >= e) break;
    pool.emplace_back([=] { fn(b, e, t); });
  }
  for (auto& th : pool) th.join();
}
// AHA_B200_FLATTEN_THREADS (1 = the sequential passes); default: the host's hardware threads, at most 16
inline int flatten_threads() {
  if (const char* e = getenv("AHA_B200_FLATTEN_THREADS")) return std::max(1, atoi(e));
  const unsigned hc = std::thread::hardware_concurrency();
  return (int)std::min<unsigned>(hc ? hc : 1u, 16u);
}

// Flatten.  Returns false + err on malformed input.
inline bool flatten(const CedarNode* nodes, int64_t array_size, const int32_t* fails, const CedarOut* output,
                    const uint32_t* key_lens, int64_t n_keys, std::vector<uint8_t>* blob, std::string* err) {
  auto base_of = [&](int32_t id) { return -(nodes[id].value + 1); };
  auto is_end = [&](int32_t id) { return (nodes[id].flags & 0x200) != 0 || nodes[id].child == 0; };
  if (array_size < 256) {
    *err = "array_size < 256";
    return false;
  }
#ifdef AHA_FLATTEN_TIMES  // -DAHA_FLATTEN_TIMES: seconds per phase on stderr (tools/blob_stats.py)
  auto t_last = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "flatten %-28s %.3f s\n", what, std::chrono::duration<double>(now - t_last).count());
    t_last = now;
  };
#else
  auto lap = [](const char*) {};
#endif
  // 1. BFS renumbering over the sibling lists (cedar.cr:450-463), same order ACX.compile walks
  std::vector<int32_t> slot_of;              // state -> cedar slot
  std::vector<int32_t> parent;               // state -> parent state
  std::vector<uint8_t> label;                // state -> incoming label
  std::vector<int32_t> state_of(array_size, -1);
  slot_of.push_back(0);
  parent.push_back(0);
  label.push_back(0);
  state_of[0] = 0;
  bool used[256] = {false};
  const int T = array_size >= (1 << 18) ? flatten_threads() : 1;
  if (T > 1) {
    // the same breadth-first order, one level at a time: count the children of every state of the level, prefix, place
    std::atomic<int> bad(0);
    std::vector<uint8_t> used_t((size_t)T * 256, 0);
    std::vector<uint32_t> first;  // first[i] = children of the level's states in front of state i
    size_t lo = 0;
    auto first_sibling = [&](int32_t id, int32_t* b_out) -> int {  // -1: corrupt
      const int32_t b = base_of(id);
      *b_out = b;
      uint8_t sib = nodes[id].child;
      if (sib == 0 && b > 0) {
        if (b >= array_size) return -1;
        sib = nodes[b].sibling;
      }
      return sib;
    };
    while (lo < slot_of.size() && !bad.load()) {
      const size_t hi = slot_of.size(), n_level = hi - lo;
      first.assign(n_level + 1, 0);
      parallel_chunks(n_level, T, [&](uint64_t b0, uint64_t e0, int) {
        for (uint64_t i = b0; i < e0; ++i) {
          int32_t b;
          int sib = first_sibling(slot_of[lo + i], &b);
          if (sib < 0) { bad.store(1); return; }
          uint32_t n = 0;
          while (sib != 0) {
            const int32_t to = b ^ sib;
            if (to < 0) break;
            if (to >= array_size || ++n > 256) { bad.store(2); return; }
            sib = nodes[to].sibling;
          }
          first[i + 1] = n;
        }
      });
      if (bad.load()) break;
      for (size_t i = 0; i < n_level; ++i) first[i + 1] += first[i];
      const size_t total = first[n_level];
      if (total == 0) break;
      slot_of.resize(hi + total);
      parent.resize(hi + total);
      label.resize(hi + total);
      parallel_chunks(n_level, T, [&](uint64_t b0, uint64_t e0, int t) {
        for (uint64_t i = b0; i < e0; ++i) {
          int32_t b;
          int sib = first_sibling(slot_of[lo + i], &b);
          size_t at = hi + first[i];
          while (sib != 0) {
            const int32_t to = b ^ sib;
            if (to < 0) break;
            if (__atomic_exchange_n(&state_of[to], (int32_t)at, __ATOMIC_RELAXED) >= 0) { bad.store(2); return; }
            slot_of[at] = to;
            parent[at] = (int32_t)(lo + i);
            label[at] = (uint8_t)sib;
            used_t[(size_t)t * 256 + sib] = 1;
            ++at;
            sib = nodes[to].sibling;
          }
        }
      });
      lo = hi;
    }
    if (bad.load()) {
      *err = bad.load() == 1 ? "corrupt node: base out of range" : "corrupt sibling list";
      return false;
    }
    for (int t = 0; t < T; ++t)
      for (int b = 0; b < 256; ++b) used[b] = used[b] || used_t[(size_t)t * 256 + b];
  }
  for (size_t s = 0; T <= 1 && s < slot_of.size(); ++s) {
    const int32_t id = slot_of[s];
    const int32_t b = base_of(id);
    uint8_t sib = nodes[id].child;
    if (sib == 0 && b > 0) {
      if (b >= array_size) {  // the arrays may come from a file (AC.load): never index past them
        *err = "corrupt node: base out of range";
        return false;
      }
      sib = nodes[b].sibling;
    }
    while (sib != 0) {
      const int32_t to = b ^ sib;
      if (to < 0) break;
      if (to >= array_size || state_of[to] >= 0) {
        *err = "corrupt sibling list";
        return false;
      }
      state_of[to] = static_cast<int32_t>(slot_of.size());
      slot_of.push_back(to);
      parent.push_back(static_cast<int32_t>(s));
      label.push_back(sib);
      used[sib] = true;
      sib = nodes[to].sibling;
    }
  }
  const uint64_t S = slot_of.size();
  if (S > kStateMask || S > kSparseMaxStates) {
    *err = "too many states";
    return false;
  }
  lap("1 bfs");
  // 2. byte classes: every byte used as a label is its own class; all others (incl. NUL, Q6) share class 0
  uint8_t classmap[256];
  uint32_t C = 1;
  for (int b = 0; b < 256; ++b) classmap[b] = (b != 0 && used[b]) ? static_cast<uint8_t>(C++) : 0;
  // 3. per-state facts
  std::vector<uint16_t> depth(S, 0);
  std::vector<uint32_t> tail(S, 0);  // non-continuation bytes among path bytes at index >= 1
  std::vector<uint32_t> sflags(S, 0);
  std::vector<uint32_t> out_off(S + 1, 0);
  std::vector<int32_t> out_val;
  uint32_t max_depth = 0;
  for (uint64_t s = 1; s < S; ++s) {
    const uint32_t d = depth[parent[s]] + 1u;
    depth[s] = static_cast<uint16_t>(d > 65535u ? 65535u : d);
    if (d > max_depth) max_depth = d;
    tail[s] = tail[parent[s]] + ((d >= 2 && (label[s] & 0xC0) != 0x80) ? 1u : 0u);
  }
  std::vector<uint32_t> key_tail(n_keys > 0 ? n_keys : 1, 0);
  // ACX#fetch (ac.cr:265-278): walk while value >= 0, follow next while >= 0.  Two passes (count, prefix, fill) so that the
  // states can be handled independently; returns 0 or an error code, emit(v) for every value that is kept.
  auto walk_outputs = [&](uint64_t st, auto emit) -> int {
    int32_t e = slot_of[st];
    int guard = 0;
    while (output[e].value >= 0) {
      const int32_t v = output[e].value;
      if (v >= n_keys) return 1;
      if (key_lens[v] < 0x7FFFFFFFu) emit(v);
      if (!(output[e].next >= 0)) break;
      e = output[e].next;
      if (e >= array_size) return 2;
      if (++guard > 1 << 20) return 3;
    }
    return 0;
  };
  {
    std::atomic<int> bad(0);
    parallel_chunks(S, T, [&](uint64_t b0, uint64_t e0, int) {
      for (uint64_t st = std::max<uint64_t>(b0, 1); st < e0; ++st) {
        uint32_t n = 0;
        const int rc = walk_outputs(st, [&](int32_t) { ++n; });
        if (rc) { bad.store(rc); return; }
        out_off[st + 1] = n;
      }
    });
    if (bad.load()) {
      *err = bad.load() == 1 ? "output value out of range" : (bad.load() == 2 ? "output chain leaves the array" : "output chain loop");
      return false;
    }
    out_off[0] = 0;
    out_off[1] = 0;  // the root has no output list
    for (uint64_t st = 1; st < S; ++st) out_off[st + 1] += out_off[st];
    out_val.resize(out_off[S]);
    parallel_chunks(S, T, [&](uint64_t b0, uint64_t e0, int) {
      for (uint64_t st = std::max<uint64_t>(b0, 1); st < e0; ++st) {
        const int32_t id = slot_of[st];
        if (is_end(id)) sflags[st] |= kEnd;
        uint32_t at = out_off[st];
        walk_outputs(st, [&](int32_t v) { out_val[at++] = v; });
        if (at > out_off[st]) sflags[st] |= kOut;
        if (output[id].value >= 0) key_tail[output[id].value] = tail[st];  // a key id belongs to one state
      }
    });
  }
  out_off[S] = static_cast<uint32_t>(out_val.size());
  lap("3 facts + output lists");
  // 4. sparse rows: the children of a state are contiguous in BFS numbering
  std::vector<uint64_t> kid_begin(S + 1, S);  // smallest child id, S when the state has no child
  parallel_chunks(S, T, [&](uint64_t b0, uint64_t e0, int) {  // a state's children are consecutive: the first one of each run
    for (uint64_t st = std::max<uint64_t>(b0, 1); st < e0; ++st)
      if (st == 1 || parent[st] != parent[st - 1]) kid_begin[parent[st]] = st;
  });
  auto kids_of = [&](uint64_t s, uint64_t* lo, uint64_t* hi) {
    *lo = kid_begin[s];
    *hi = *lo;
    while (*hi < S && static_cast<uint64_t>(parent[*hi]) == s) ++*hi;
  };
  std::vector<uint32_t> fail_state(S, 0);
  {
    std::atomic<int> bad(0);
    parallel_chunks(S, T, [&](uint64_t b0, uint64_t e0, int) {
      for (uint64_t st = std::max<uint64_t>(b0, 1); st < e0; ++st) {
        const int32_t f = fails[slot_of[st]];
        if (f < 0 || f >= array_size || state_of[f] < 0 || static_cast<uint64_t>(state_of[f]) >= st) { bad.store(1); return; }
        fail_state[st] = static_cast<uint32_t>(state_of[f]);
      }
    });
    if (bad.load()) {
      *err = "fail link does not point to an earlier state";
      return false;
    }
  }
  std::vector<SparseRow> rows(S);
  std::vector<uint32_t> wide;
  uint64_t n_wide = 0;
  auto edge_to = [&](uint64_t t, uint32_t lab) {
    return (static_cast<uint32_t>(t) << 9) | ((sflags[t] & kOut) ? kEdgeOut : 0u) | lab;
  };
  // which states are wide (the root, and more than three children), their rows in wide[] in state order
  std::vector<uint32_t> wide_row(S, 0xFFFFFFFFu);
  {
    std::vector<uint8_t> is_wide(S, 0);
    parallel_chunks(S, T, [&](uint64_t b0, uint64_t e0, int) {
      for (uint64_t st = b0; st < e0; ++st) {
        uint64_t lo, hi;
        kids_of(st, &lo, &hi);
        is_wide[st] = (st == 0 || hi - lo > 3) ? 1 : 0;
      }
    });
    for (uint64_t st = 0; st < S; ++st)
      if (is_wide[st]) wide_row[st] = static_cast<uint32_t>(n_wide++);
  }
  wide.assign(n_wide * C, 0u);
  parallel_chunks(S, T, [&](uint64_t b0, uint64_t e0, int) {
    for (uint64_t st = b0; st < e0; ++st) {
      uint64_t lo, hi;
      kids_of(st, &lo, &hi);
      SparseRow r;
      r.head = (fail_state[st] << 8) | ((sflags[st] & kOut) ? kRowOut : 0u) | ((sflags[st] & kEnd) ? kRowEnd : 0u);
      r.e[0] = r.e[1] = r.e[2] = 0;
      if (wide_row[st] != 0xFFFFFFFFu) {
        r.head |= kRowWide;
        r.e[0] = wide_row[st];
        for (uint64_t t = lo; t < hi; ++t) wide[static_cast<uint64_t>(wide_row[st]) * C + classmap[label[t]]] = edge_to(t, 1u);
      } else {
        for (uint64_t t = lo; t < hi; ++t) r.e[t - lo] = edge_to(t, label[t]);
      }
      rows[st] = r;
    }
  });
  lap("4 rows");
  const bool dense = S * static_cast<uint64_t>(C) * 4 <= kDenseMaxBytes;
  // 4a. no dense table: the dense rows of the wide states are completed to closure rows -- an entry without a goto edge
  // takes what the fail chain would find (label field 2 instead of 1; 0 = back to the root).  A wide state then never
  // misses: the walk consumes a byte with ONE load there, and the wide states are where fail chains usually end.
  if (!dense) {
    auto goto_of = [&](uint32_t t, uint32_t c, uint32_t byte) -> uint32_t {  // real goto edge of state t, 0 = none
      const SparseRow& r = rows[t];
      if (r.head & kRowWide) {
        const uint32_t e = wide[static_cast<uint64_t>(r.e[0]) * C + c];
        return (e & 0xFFu) == 1u ? e : 0u;
      }
      for (int k = 0; k < 3; ++k)
        if (r.e[k] && (r.e[k] & 0xFFu) == byte) return r.e[k];
      return 0u;
    };
    uint8_t byte_of_class[256] = {0};
    for (int b = 1; b < 256; ++b)
      if (classmap[b]) byte_of_class[classmap[b]] = static_cast<uint8_t>(b);
    for (uint64_t s = 1; s < S; ++s) {  // BFS order: the wide states on the fail chain (shallower) are complete already
      if (!(rows[s].head & kRowWide)) continue;
      uint32_t* row = &wide[static_cast<uint64_t>(rows[s].e[0]) * C];
      for (uint32_t c = 1; c < C; ++c) {
        if (row[c]) continue;
        uint32_t t = fail_state[s], e = 0;
        for (;;) {
          if (rows[t].head & kRowWide) {  // complete row (or the root's): its entry is the answer
            e = wide[static_cast<uint64_t>(rows[t].e[0]) * C + c];
            break;
          }
          e = goto_of(t, c, byte_of_class[c]);
          if (e || t == 0) break;
          t = fail_state[t];
        }
        row[c] = e ? ((e & ~0xFFu) | 2u) : 0u;
      }
    }
  }
  lap("4a wide closure");
  // 4b. no dense table: renumber, wide states first in BFS order, the others behind them in BFS order.  A wide state's
  // id is then its row in wide[], and the walk reads its edge with one load instead of two (row, then wide entry):
  // the wide states are the shallow ones the walk is in most of the time.
  std::vector<uint32_t> wide_fail;
  if (!dense) {
    std::vector<uint32_t> nid(S);
    uint32_t next_wide = 0, next_narrow = static_cast<uint32_t>(n_wide);
    for (uint64_t s = 0; s < S; ++s) nid[s] = (rows[s].head & kRowWide) ? next_wide++ : next_narrow++;
    auto re_edge = [&](uint32_t e) { return e ? ((nid[e >> 9] << 9) | (e & 0x1FFu)) : 0u; };
    std::vector<SparseRow> rows2(S);
    std::vector<uint32_t> wide2(wide.size());
    std::vector<uint32_t> out_off2(S + 1, 0), fail2(S), cnt(S, 0);
    std::vector<uint16_t> depth2(S);
    wide_fail.assign(n_wide, 0);
    parallel_chunks(S, T, [&](uint64_t b0, uint64_t e0, int) {
      for (uint64_t s = b0; s < e0; ++s) {
        const uint32_t t = nid[s];
        SparseRow r = rows[s];
        r.head = (nid[r.head >> 8] << 8) | (r.head & 0xFFu);
        if (r.head & kRowWide) {
          for (uint32_t c = 0; c < C; ++c) wide2[static_cast<uint64_t>(t) * C + c] = re_edge(wide[static_cast<uint64_t>(r.e[0]) * C + c]);
          r.e[0] = t;
          wide_fail[t] = r.head >> 8;
        } else {
          for (int k = 0; k < 3; ++k) r.e[k] = re_edge(r.e[k]);
        }
        rows2[t] = r;
        fail2[t] = nid[fail_state[s]];
        depth2[t] = depth[s];
        cnt[t] = out_off[s + 1] - out_off[s];
      }
    });
    for (uint64_t t = 0; t < S; ++t) out_off2[t + 1] = out_off2[t] + cnt[t];
    std::vector<int32_t> out_val2(out_val.size());
    parallel_chunks(S, T, [&](uint64_t b0, uint64_t e0, int) {
      for (uint64_t s = b0; s < e0; ++s)
        for (uint32_t j = 0; j < out_off[s + 1] - out_off[s]; ++j) out_val2[out_off2[nid[s]] + j] = out_val[out_off[s] + j];
    });
    rows.swap(rows2);
    wide.swap(wide2);
    out_off.swap(out_off2);
    out_val.swap(out_val2);
    fail_state.swap(fail2);
    depth.swap(depth2);
  }
  lap("4b renumber");
  // 5. layout
  BlobHeader h;
  memset(&h, 0, sizeof(h));
  h.magic = kBlobMagic;
  h.version = 3;
  h.n_classes = C;
  h.n_states = S;
  h.n_keys = static_cast<uint64_t>(n_keys);
  h.n_out = out_val.size();
  h.max_key_len = max_depth;
  h.has_delta = dense ? 1 : 0;
  h.wide_first = dense ? 0 : 1;
  h.n_wide = n_wide;
  uint64_t off = align16(sizeof(BlobHeader));
  h.off_classmap = off; off = align16(off + 256);
  h.off_delta = off; off = align16(off + (dense ? S * C * 4 : 0));
  h.off_out_off = off; off = align16(off + (S + 1) * 4);
  h.off_out_val = off; off = align16(off + out_val.size() * 4);
  h.off_key_len = off; off = align16(off + static_cast<uint64_t>(n_keys) * 4);
  h.off_key_tail = off; off = align16(off + static_cast<uint64_t>(n_keys) * 4);
  h.off_depth = off; off = align16(off + S * 2);
  h.off_fail = off; off = align16(off + S * 4);
  h.off_rows = off; off = align16(off + S * sizeof(SparseRow));
  h.off_wide = off; off = align16(off + wide.size() * 4);
  h.off_wide_fail = off; off = align16(off + wide_fail.size() * 4);
  h.total_bytes = off;
  blob->assign(off, 0);
  uint8_t* p = blob->data();
  memcpy(p, &h, sizeof(h));
  memcpy(p + h.off_classmap, classmap, 256);
  memcpy(p + h.off_out_off, out_off.data(), (S + 1) * 4);
  if (!out_val.empty()) memcpy(p + h.off_out_val, out_val.data(), out_val.size() * 4);
  if (n_keys > 0) {
    memcpy(p + h.off_key_len, key_lens, static_cast<uint64_t>(n_keys) * 4);
    memcpy(p + h.off_key_tail, key_tail.data(), static_cast<uint64_t>(n_keys) * 4);
  }
  memcpy(p + h.off_depth, depth.data(), S * 2);
  memcpy(p + h.off_fail, fail_state.data(), S * 4);
  memcpy(p + h.off_rows, rows.data(), S * sizeof(SparseRow));
  if (!wide.empty()) memcpy(p + h.off_wide, wide.data(), wide.size() * 4);
  if (!wide_fail.empty()) memcpy(p + h.off_wide_fail, wide_fail.data(), wide_fail.size() * 4);
  lap("5 layout + copies");
  // 6. dense closure in BFS order (small automata only): fail targets are shallower, hence already complete
  if (dense) {
    uint32_t* delta = reinterpret_cast<uint32_t*>(p + h.off_delta);
    for (uint64_t s = 0; s < S; ++s) {
      uint32_t* row = delta + s * C;
      if (s == 0) {
        for (uint32_t c = 0; c < C; ++c) row[c] = 0;
      } else {
        const uint32_t* frow = delta + static_cast<uint64_t>(fail_state[s]) * C;
        for (uint32_t c = 0; c < C; ++c) row[c] = frow[c] & ~kGoto;
      }
      uint64_t lo, hi;
      kids_of(s, &lo, &hi);
      for (uint64_t t = lo; t < hi; ++t) row[classmap[label[t]]] = static_cast<uint32_t>(t) | sflags[t] | kGoto;
    }
  }
  return true;
}

inline bool parse_blob(const uint8_t* blob, int64_t bytes, BlobHeader* h, std::string* err) {
  if (bytes < static_cast<int64_t>(sizeof(BlobHeader))) {
    *err = "blob too small";
    return false;
  }
  memcpy(h, blob, sizeof(BlobHeader));
  if (h->magic != kBlobMagic || h->version != 3 || h->total_bytes != static_cast<uint64_t>(bytes)) {
    *err = "bad blob header";
    return false;
  }
  // a blob may arrive from a file or from another rank: every section must lie inside it and every index it holds
  // must stay inside the section it points into, or the kernels would read out of bounds
  const uint64_t S = h->n_states, C = h->n_classes, K = h->n_keys, O = h->n_out, T = h->total_bytes;
  if (S == 0 || S > kStateMask || C == 0 || C > 256 || K > 0x7FFFFFFFull || O > 0xFFFFFFFFull) {
    *err = "bad blob header: counts out of range";
    return false;
  }
  auto inside = [&](uint64_t off, uint64_t n, uint64_t width) {
    return (off & 3) == 0 && off <= T && n <= (T - off) / width;
  };
  if (S > kSparseMaxStates || h->has_delta > 1 || h->n_wide == 0 || h->n_wide > S) {
    *err = "bad blob header: counts out of range";
    return false;
  }
  if (h->wide_first > 1 || h->wide_first == h->has_delta) {
    *err = "bad blob header: flags";
    return false;
  }
  if (!inside(h->off_rows, S, sizeof(SparseRow)) || !inside(h->off_wide, h->n_wide * C, 4) ||
      !inside(h->off_wide_fail, h->wide_first ? h->n_wide : 0, 4)) {
    *err = "bad blob header: section out of bounds";
    return false;
  }
  if (!inside(h->off_classmap, 256, 1) || !inside(h->off_delta, h->has_delta ? S * C : 0, 4) || !inside(h->off_out_off, S + 1, 4) ||
      !inside(h->off_out_val, O, 4) || !inside(h->off_key_len, K, 4) || !inside(h->off_key_tail, K, 4) ||
      !inside(h->off_depth, S, 2) || !inside(h->off_fail, S, 4)) {
    *err = "bad blob header: section out of bounds";
    return false;
  }
  const DfaView v = view_of(*h, blob);
  for (int b = 0; b < 256; ++b)
    if (v.classmap[b] >= C) {
      *err = "corrupt blob: class map";
      return false;
    }
  uint32_t bad = 0;
  for (uint64_t s = 0; s < S; ++s) bad |= (v.out_off[s] > v.out_off[s + 1]) | (v.fail[s] >= S);
  bad |= v.out_off[S] != O;
  for (uint64_t j = 0; j < O; ++j) bad |= (static_cast<uint64_t>(static_cast<uint32_t>(v.out_val[j])) >= K);
  for (uint64_t s = 0; s < S; ++s) {  // sparse rows: fail links, inline targets, wide row indices
    const SparseRow& r = v.rows[s];
    bad |= (r.head >> 8) >= S;
    if (r.head & kRowWide) bad |= r.e[0] >= h->n_wide;
    else
      for (int k = 0; k < 3; ++k) bad |= (r.e[k] >> 9) >= S;
  }
  bad |= !(v.rows[0].head & kRowWide);
  if (h->wide_first)
    for (uint64_t s = 0; s < S; ++s) {  // the numbering the walk relies on: id < n_wide <=> wide, and a wide row is its id
      const SparseRow& r = v.rows[s];
      const bool w = (r.head & kRowWide) != 0;
      bad |= w != (s < h->n_wide);
      if (w) bad |= (r.e[0] != s) | (s < h->n_wide && v.wide_fail[s] != (r.head >> 8));
    }
  for (uint64_t i = 0; i < h->n_wide * C; ++i) bad |= (v.wide[i] >> 9) >= S;
  const uint64_t n_delta = h->has_delta ? S * C : 0;
  uint32_t worst = 0;
  for (uint64_t i = 0; i < n_delta; ++i) {
    const uint32_t t = v.delta[i] & kStateMask;
    worst = t > worst ? t : worst;
  }
  if (bad || worst >= S) {
    *err = "corrupt blob: an index leaves its section";
    return false;
  }
  return true;
}

}  // namespace aha
