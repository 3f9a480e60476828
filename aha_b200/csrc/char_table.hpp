// char_table.hpp -- derives the shared-memory resident CHARACTER automaton from the byte-level DFA.
//
// When every key is a sequence of whole UTF-8 units (lead byte + its continuation bytes) and the text
// is structurally valid UTF-8, a key can only match at unit boundaries (a lead/ASCII byte never equals
// a continuation byte), so the byte automaton of ACX#match_ (ac.cr:176-192) visited at unit boundaries
// IS the automaton over characters: same states, same failure links, same output lists, and no output
// can fire in the middle of a unit.  One transition then consumes 1-4 text bytes.
//
// Representation (all in shared memory, 6 bytes per edge slot):
//   root[idx(char)]      -> id of the one-character state:15 | fires:1                          (direct table, u16)
//   column of a char     =  idx(char), its index into root[]: the walk has it in a register already, no lookup
//   state id == base     -> edge slot = base + column ; check[slot] == base means the goto edge exists
//   val[slot]            -> next state id:15 | fires:1 | effective fail of next:15 | deep:1
//                           deep = the fail link of next has a non-root fail link itself (a third probe may be needed)
//   gfail[id]            -> effective fail of state id; the states a chain can pass through are packed first, so that
//                           they get the smallest ids and the part of gfail[] the walk ever reads is a short prefix
//                           that fits in shared memory too (the rest stays in global memory)
// States without children are never "current": an edge into one stores the nearest fail-ancestor that
// has children (its transitions are identical by the closure of the goto/fail loop) but still reports
// the target's own output list (ref_rec[slot], global memory, read only when hits are written out).
// One-character states carry fail = 0, so "follow fail until 0, then take the root entry" needs no depth test.
#pragma once
#include <algorithm>
#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "dfa.hpp"

namespace aha {

constexpr uint32_t kCidOut = 0x8000u;
constexpr uint32_t kNextOut = 0x8000u;

// what an edge / a one-character state reports: n == 1 -> a = key id, b = key length in bytes, c = characters
// after the first (char-unit start); n > 1 -> [a, b) indexes out_val
struct RefRec {
  uint32_t a, b, c, n;
};

struct CharTableHost {
  bool ok = false;
  std::string why;                 // reason when !ok
  // multi-byte root index, one contiguous range per unit length L = 2, 3, 4:
  //   idx = off[L-2] + (payload - lo[L-2]) when payload - lo < n ; ASCII: idx = byte
  uint32_t lo[3] = {0, 0, 0}, n[3] = {0, 0, 0}, off[3] = {256, 256, 256};  // entries 128..255 stay 0: idle lanes of the ASCII step may index them
  uint32_t n_root = 256;
  uint32_t m = 0;                  // slots
  std::vector<uint16_t> root;      // [n_root]
  std::vector<uint16_t> check;     // [m_pad = m + n_root + 1 rounded up]: unused / padding = 0xFFFF, so base + column never needs a bound test
  std::vector<uint32_t> val;       // [m]
  std::vector<uint32_t> multi_bits; // 2 bits per ref: ref_rec[ref].n when it is 1..3, 0 = look it up (shared memory: the walk
                                    // counts hits without touching ref_rec)
  std::vector<RefRec> ref_rec;     // [m + n_root] what fires at edge slot / at the one-character state of column c: m + c (global)
  std::vector<uint16_t> gfail;     // [m] effective fail of the state whose id is the index (global; prefix in shared memory)
  uint32_t n_gfail_s = 0;          // gfail[0 .. n_gfail_s) covers every state a fail chain can pass through
  uint32_t n_cid = 0, n_edges = 0, n_states = 0;
  // shared memory of the walk: val | root | check | gfail prefix | class map of the fallback walk
  size_t smem_bytes() const { return (size_t)m * 4 + root.size() * 2 + check.size() * 2 + (size_t)n_gfail_s * 2 + 256 + 16; }
  uint32_t root_index(uint32_t code) const {
    const uint32_t cls = code >> 30;
    return cls ? off[cls - 1] + ((code & 0x3FFFFFFFu) - lo[cls - 1]) : code;
  }
};

// tagged code of a UTF-8 unit: injective over (length, payload bits)
// bits 30-31: length - 1, bits 0-20: payload
inline uint32_t unit_code(const uint8_t* b, int len) {
  if (len == 1) return b[0];
  if (len == 2) return (1u << 30) | ((b[0] & 0x1Fu) << 6) | (b[1] & 0x3Fu);
  if (len == 3) return (2u << 30) | ((b[0] & 0x0Fu) << 12) | ((b[1] & 0x3Fu) << 6) | (b[2] & 0x3Fu);
  return (3u << 30) | ((b[0] & 0x07u) << 18) | ((b[1] & 0x3Fu) << 12) | ((b[2] & 0x3Fu) << 6) | (b[3] & 0x3Fu);
}
// bytes that must follow a lead byte, -1 if the byte cannot start a unit
inline int unit_follow(uint8_t b) {
  if (b < 0x80) return 0;
  if (b >= 0xC0 && b < 0xE0) return 1;
  if (b >= 0xE0 && b < 0xF0) return 2;
  if (b >= 0xF0 && b < 0xF8) return 3;
  return -1;
}

struct CharEdge { uint32_t from, to, code; };
struct CharRow { uint32_t st; uint32_t first, count; bool depth1; bool chain; };

inline bool parent_is_root_unit(uint32_t st, const std::vector<int8_t>& phase, const std::vector<uint32_t>& parent) {
  uint32_t p = parent[st];
  while (phase[p] != 0) p = parent[p];
  return p == 0;
}

// First-fit row displacement: state id = base, edge slot = base + column (column = root index of the character).
inline CharTableHost pack_rows(CharTableHost T, const DfaView& v, const std::vector<CharEdge>& edges,
                               const std::vector<CharRow>& rows, const std::vector<uint32_t>& eff, size_t smem_budget) {
  T.ok = false;
  const size_t root_bytes = (size_t)T.n_root * 2;
  // check[] is padded by n_root + 1 entries; 16 KB at most for the gfail prefix (sized below)
  const size_t pad_bytes = ((size_t)T.n_root + 1 + 16) * 2 + 8;
  if (smem_budget < root_bytes + 256 + 16 + pad_bytes + 64) { T.why = "shared memory budget too small"; return T; }
  size_t m_max = (smem_budget - root_bytes - 256 - 16 - pad_bytes) / 6;  // 6 bytes per slot
  if (m_max + T.n_root > 0xFFF0) m_max = 0xFFF0 - T.n_root;  // an event carries its ref (slot, or m + column) in 16 bits
  const size_t base_max = 0x7FFF;                             // a state id has 15 bits
  std::vector<uint8_t> used(m_max + 1, 0), base_used(base_max + 1, 0);
  used[0] = 1;
  std::vector<uint32_t> id_of(v.n_states, 0);
  std::vector<uint32_t> cols;
  uint32_t max_slot = 0, max_base = 0, max_chain_base = 0;
  uint32_t first_free = 1;  // lowest free slot
  for (const CharRow& r : rows) {
    cols.clear();
    for (uint32_t i = r.first; i < r.first + r.count; ++i) cols.push_back(T.root_index(edges[i].code));
    std::sort(cols.begin(), cols.end());
    const uint32_t c0 = cols[0], c_last = cols.back();
    while (first_free <= m_max && used[first_free]) ++first_free;
    bool placed = false;
    uint32_t b = 0;
    for (uint32_t f = std::max(first_free, c0 + 1); f <= m_max; ++f) {  // candidate slot for the smallest column
      if (used[f]) continue;
      b = f - c0;
      if ((size_t)b + c_last > m_max || b > base_max) break;
      if (base_used[b]) continue;
      bool fit = true;
      for (size_t i = 1; i < cols.size(); ++i)
        if (used[b + cols[i]]) { fit = false; break; }
      if (fit) { placed = true; break; }
    }
    if (!placed) { T.why = "character automaton does not fit in shared memory"; return T; }
    base_used[b] = 1;
    for (uint32_t c : cols) used[b + c] = 1;
    id_of[r.st] = b;
    if (b + c_last > max_slot) max_slot = b + c_last;
    if (b > max_base) max_base = b;
    if (r.chain && b > max_chain_base) max_chain_base = b;
  }
  T.m = (max_slot + 1 + 15) & ~15u;
  T.check.assign(((size_t)max_base + T.n_root + 1 + 15) & ~size_t(15), 0xFFFFu);
  if (T.check.size() < T.m) T.check.resize(T.m, 0xFFFFu);
  T.val.assign(T.m, 0);
  T.gfail.assign(T.m, 0);
  T.root.assign(T.n_root, 0);
  T.ref_rec.assign((size_t)T.m + T.n_root, RefRec{0, 0, 0, 0});
  // reference record of a firing state: single output -> {value, key_len, key_tail, 1}; else list bounds
  auto rec_of = [&](uint32_t st) -> RefRec {
    const uint32_t j0 = v.out_off[st], j1 = v.out_off[st + 1];
    if (j1 - j0 == 1) {
      const int32_t val = v.out_val[j0];
      return RefRec{(uint32_t)val, v.key_len[val], v.key_tail[val], 1};
    }
    return RefRec{j0, j1, 0, j1 - j0};
  };
  for (const CharEdge& e : edges) {
    const uint32_t col = T.root_index(e.code);
    const uint32_t cont = eff[e.to];  // state we continue in (byte-level id, 0 = root)
    const uint32_t next_id = cont ? id_of[cont] : 0;
    const uint32_t n_out = v.out_off[e.to + 1] - v.out_off[e.to];
    const uint32_t flags = (n_out ? 0x8000u : 0u);
    if (e.from == 0) {
      T.root[col] = (uint16_t)(next_id | flags);
      if (n_out) T.ref_rec[(size_t)T.m + col] = rec_of(e.to);
      continue;
    }
    const uint32_t base = id_of[e.from];
    const uint32_t slot = base + col;
    const uint32_t cf = cont ? eff[v.fail[cont]] : 0;
    const uint32_t cont_fail = cf ? id_of[cf] : 0;
    T.check[slot] = (uint16_t)base;
    const uint32_t deep = (cf && eff[v.fail[cf]]) ? 0x80000000u : 0u;
    T.val[slot] = next_id | flags | (cont_fail << 16) | deep;
    if (n_out) T.ref_rec[slot] = rec_of(e.to);
  }
  for (const CharRow& r : rows) {
    const uint32_t f = eff[v.fail[r.st]];
    T.gfail[id_of[r.st]] = (uint16_t)(f ? id_of[f] : 0);
  }
  // the chain states were packed first: gfail[0 .. max_chain_base] is all a chain ever reads.  What does not fit
  // the budget stays in global memory (the walk tests the id against n_gfail_s).
  {
    const size_t fixed = (size_t)T.m * 4 + T.root.size() * 2 + T.check.size() * 2 + 256 + 16;
    size_t room = smem_budget > fixed ? (smem_budget - fixed) / 2 : 0;
    size_t want = max_chain_base ? ((size_t)max_chain_base + 1 + 7) & ~size_t(7) : 0;
    T.n_gfail_s = (uint32_t)std::min(std::min(want, room), (size_t)T.m);
  }
  T.multi_bits.assign((T.ref_rec.size() + 15) / 16, 0u);
  for (size_t i = 0; i < T.ref_rec.size(); ++i)
    if (T.ref_rec[i].n >= 1 && T.ref_rec[i].n <= 3) T.multi_bits[i >> 4] |= T.ref_rec[i].n << ((i & 15) * 2);
  T.ok = true;
  return T;
}

inline CharTableHost build_char_table(const BlobHeader& h, const uint8_t* blob, size_t smem_budget) {
  CharTableHost T;
  const DfaView v = view_of(h, blob);
  const uint64_t S = v.n_states;
  const uint32_t C = v.n_classes;
  if (S < 2) { T.why = "empty dictionary"; return T; }
  uint8_t byte_of_class[256] = {0};
  for (int b = 1; b < 256; ++b)
    if (v.classmap[b]) byte_of_class[v.classmap[b]] = (uint8_t)b;
  // 1. phase of every byte state (bytes still owed to the current unit); children are goto edges
  std::vector<int8_t> phase(S, -1);
  std::vector<uint32_t> parent(S, 0);
  std::vector<uint8_t> label(S, 0);
  std::vector<uint32_t> kid_begin(S + 1, 0);  // children are contiguous in BFS numbering
  phase[0] = 0;
  {
    uint64_t cursor = 1;
    for (uint64_t s = 0; s < S; ++s) {
      kid_begin[s] = (uint32_t)cursor;
      const uint32_t* row = v.delta + s * C;
      // children of s, in BFS numbering order == increasing id; find them by scanning the row
      std::vector<std::pair<uint32_t, uint8_t>> kids;
      for (uint32_t c = 1; c < C; ++c)
        if (row[c] & kGoto) kids.push_back({row[c] & kStateMask, byte_of_class[c]});
      std::sort(kids.begin(), kids.end());
      for (auto& k : kids) {
        if (k.first != cursor) { T.why = "children not contiguous"; return T; }
        parent[cursor] = (uint32_t)s;
        label[cursor] = k.second;
        const int ph = phase[s];
        if (ph == 0) {
          const int f = unit_follow(k.second);
          if (f < 0) { T.why = "a key does not start a UTF-8 unit at a unit boundary"; return T; }
          phase[cursor] = (int8_t)f;
        } else {
          if ((k.second & 0xC0) != 0x80) { T.why = "a key breaks a UTF-8 unit"; return T; }
          phase[cursor] = (int8_t)(ph - 1);
        }
        ++cursor;
      }
    }
    kid_begin[S] = (uint32_t)S;
  }
  for (uint64_t s = 1; s < S; ++s)
    if (v.out_off[s] < v.out_off[s + 1] && phase[s] != 0) { T.why = "a key ends inside a UTF-8 unit"; return T; }
  // 2. character edges: from every phase-0 state, follow goto edges to the next phase-0 states
  typedef CharEdge Edge;
  std::vector<Edge> edges;
  std::vector<uint8_t> has_kids(S, 0);
  {
    struct Item { uint32_t st; uint8_t bytes[4]; int n; };
    std::vector<Item> stack;
    for (uint64_t s = 0; s < S; ++s) {
      if (phase[s] != 0) continue;
      stack.clear();
      stack.push_back(Item{(uint32_t)s, {0, 0, 0, 0}, 0});
      while (!stack.empty()) {
        Item it = stack.back();
        stack.pop_back();
        for (uint32_t t = kid_begin[it.st]; t < kid_begin[it.st + 1]; ++t) {
          Item nx = it;
          nx.st = t;
          nx.bytes[nx.n++] = label[t];
          if (phase[t] == 0) {
            edges.push_back(Edge{(uint32_t)s, t, unit_code(nx.bytes, nx.n)});
            has_kids[s] = 1;
          } else {
            stack.push_back(nx);
          }
        }
      }
    }
  }
  // 3. the dictionary's characters and the root index ranges (the root index of a character is also its column)
  std::map<uint32_t, uint32_t> seen;  // unit code -> 1
  uint32_t lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0, 0, 0};
  for (auto& e : edges) {
    const uint32_t cls = e.code >> 30;
    if (seen.emplace(e.code, 1).second && cls) {
      lo[cls - 1] = std::min(lo[cls - 1], e.code & 0x3FFFFFFFu);
      hi[cls - 1] = std::max(hi[cls - 1], e.code & 0x3FFFFFFFu);
    }
  }
  T.n_cid = (uint32_t)seen.size();
  T.n_root = 256;
  for (int c = 0; c < 3; ++c) {
    T.off[c] = T.n_root;
    if (lo[c] != 0xFFFFFFFFu) {
      T.lo[c] = lo[c] & ~63u;  // multiple of 64: the kernel folds it into the decode (scan_char.cuh)
      T.n[c] = hi[c] - T.lo[c] + 1;
      if (T.n[c] > (1u << 15)) { T.why = "character range too wide for the root table"; return T; }
    }
    T.n_root += T.n[c] + 1;  // one zero entry behind every range: the kernel clamps an out-of-range index to it
  }
  if ((size_t)T.n_root * 2 > smem_budget / 2 || T.n_root > 0xC000) { T.why = "character range too wide for the root table"; return T; }
  // 4. effective state: nearest fail-ancestor (self included) that has children; 0 = root
  std::vector<uint32_t> eff(S, 0);
  for (uint64_t s = 1; s < S; ++s) {
    if (phase[s] != 0) continue;
    eff[s] = has_kids[s] ? (uint32_t)s : eff[v.fail[s]];  // fail[s] < s and is a phase-0 state
    if (phase[v.fail[s]] != 0) { T.why = "failure link leaves unit boundaries"; return T; }
  }
  // 5. rows: edges grouped by source state (root excluded: the direct table serves it)
  std::sort(edges.begin(), edges.end(), [](const Edge& a, const Edge& b) { return a.from != b.from ? a.from < b.from : a.code < b.code; });
  typedef CharRow Row;
  std::vector<Row> rows;
  for (uint32_t i = 0; i < edges.size();) {
    uint32_t j = i;
    while (j < edges.size() && edges[j].from == edges[i].from) ++j;
    if (edges[i].from != 0) rows.push_back(Row{edges[i].from, i, j - i, parent_is_root_unit(edges[i].from, phase, parent), false});
    i = j;
  }
  // states a fail chain can pass through: the effective fail of a state with children (or of an edge's continuation)
  // that has a non-root effective fail of its own.  gfail[] is read for exactly these.
  {
    std::vector<uint8_t> is_target(S, 0);
    for (uint64_t s = 1; s < S; ++s)
      if (phase[s] == 0) is_target[eff[v.fail[eff[s] ? eff[s] : 0]]] = 1;
    for (Row& r : rows) r.chain = is_target[r.st] && eff[v.fail[r.st]] != 0;
  }
  // first-fit: the widest rows first (they are the hardest to place), then the chain states (smallest ids: their part
  // of gfail[] goes to shared memory), then everything else by width
  std::stable_sort(rows.begin(), rows.end(), [](const Row& a, const Row& b) {
    const int ka = a.count >= 16 ? 0 : (a.chain ? 1 : 2), kb = b.count >= 16 ? 0 : (b.chain ? 1 : 2);
    return ka != kb ? ka < kb : a.count > b.count;
  });
  T.n_edges = (uint32_t)edges.size();
  T.n_states = (uint32_t)rows.size() + 1;
  T.ok = true;  // provisional; packing below may still fail
  return pack_rows(T, v, edges, rows, eff, smem_budget);
}

}  // namespace aha
