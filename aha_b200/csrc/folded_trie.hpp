// folded_trie.hpp -- the dictionary's trie with case folded away, for CedarX#match_longest(ignore_case: true)
// (cedar.cr:984-1027, 1075-1121).
//
// The reference explores the case variants of the text with a frontier: at every character it tries the character as it
// stands and, if it is cased, its counterpart (cedar.cr:993-1019), and keeps the FIRST candidate of strictly greatest
// length.  Every node that frontier can reach after L characters spells a string whose case-folded form is the folded
// form of the L characters of text -- so all of them are ONE node of the trie of the folded keys, and a walk over that
// trie needs one lookup per character, no frontier and no backtracking.  What the frontier order decides is which of the
// original nodes merged into a folded node wins when several are key ends: the frontier is ordered, at every length,
// lexicographically by "took the character as it stands (first) / took its counterpart", i.e. depth-first preorder.  With
// bit i of a candidate's case mask = "character i of the key is upper case", and the same mask of the text window, the
// winner among the candidates of a node is the one whose (mask XOR text mask), read with character 0 as the most
// significant bit, is smallest.  Masks are 64 bits: dictionaries with keys of more than 64 characters, keys that are not
// valid UTF-8, or a folded table above 256 MB keep the depth-first kernel.
//
// Built on the host from the blob (the goto edges of the sparse rows), lazily at the first ignore_case scan.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "dfa.hpp"

namespace aha {

// simple 1:1 case pairs: ASCII, Latin-1, Greek, Cyrillic basic blocks (the same table as scan_trie.cuh and the oracle)
inline bool fold_is_upper(uint32_t c) {
  return (c >= 'A' && c <= 'Z') || (c >= 0xC0 && c <= 0xDE && c != 0xD7) || (c >= 0x391 && c <= 0x3A9 && c != 0x3A2) ||
         (c >= 0x410 && c <= 0x42F);
}

struct FoldedCand {
  int32_t value;      // key id (head of the original node's output list)
  uint32_t n_chars;   // length of the key in characters
  uint64_t mask;      // bit (n_chars - 1 - i) = character i is upper case
};

struct FoldedTrieHost {
  bool ok = false;
  uint32_t n_states = 0, n_classes = 0, max_chars = 0;
  uint8_t cmap[256];
  std::vector<uint32_t> delta;     // [n_states][n_classes]: target | kFoldCand (target has candidates); 0 = no edge
  std::vector<uint32_t> cand_off;  // [n_states + 1]
  std::vector<FoldedCand> cand;
  // the common case in one 32-byte sector per node: up to three candidates whose masks fit 32 bits,
  // {n, value 0, value 1, value 2, mask 0, mask 1, mask 2, 0}; n = 0xFFFFFFFF: look at cand[] (more candidates, longer keys)
  std::vector<uint32_t> rec;       // [n_states][8]
};
constexpr uint32_t kFoldCand = 0x80000000u;
constexpr uint64_t kFoldedMaxBytes = (uint64_t)256 << 20;

inline bool fold_decode(const std::string& s, std::vector<uint32_t>* cps) {
  cps->clear();
  size_t i = 0;
  while (i < s.size()) {
    const uint8_t b = (uint8_t)s[i];
    uint32_t cp;
    int l;
    if (b < 0x80) { cp = b; l = 1; }
    else if ((b & 0xE0) == 0xC0) { cp = b & 0x1F; l = 2; }
    else if ((b & 0xF0) == 0xE0) { cp = b & 0x0F; l = 3; }
    else if ((b & 0xF8) == 0xF0) { cp = b & 0x07; l = 4; }
    else return false;
    if (i + l > s.size()) return false;
    for (int k = 1; k < l; ++k) {
      const uint8_t c = (uint8_t)s[i + k];
      if ((c & 0xC0) != 0x80) return false;
      cp = (cp << 6) | (c & 0x3F);
    }
    cps->push_back(cp);
    i += l;
  }
  return true;
}
inline void fold_encode(uint32_t c, std::string* out) {
  if (c < 0x80) out->push_back((char)c);
  else if (c < 0x800) { out->push_back((char)(0xC0 | (c >> 6))); out->push_back((char)(0x80 | (c & 0x3F))); }
  else if (c < 0x10000) {
    out->push_back((char)(0xE0 | (c >> 12))); out->push_back((char)(0x80 | ((c >> 6) & 0x3F))); out->push_back((char)(0x80 | (c & 0x3F)));
  } else {
    out->push_back((char)(0xF0 | (c >> 18))); out->push_back((char)(0x80 | ((c >> 12) & 0x3F)));
    out->push_back((char)(0x80 | ((c >> 6) & 0x3F))); out->push_back((char)(0x80 | (c & 0x3F)));
  }
}

inline FoldedTrieHost build_folded_trie(const BlobHeader& h, const uint8_t* blob) {
  FoldedTrieHost F;
  const DfaView v = view_of(h, blob);
  const uint64_t S = h.n_states;
  const uint32_t C = h.n_classes;
  if (S > ((uint64_t)1 << 21)) return F;  // a folded table of that many states would not pass the size test below: do not build it
  uint8_t byte_of_class[256] = {0};
  for (int b = 1; b < 256; ++b)
    if (v.classmap[b]) byte_of_class[v.classmap[b]] = (uint8_t)b;
  // goto edges of an original state: (label, target, target has an output list)
  struct Edge { uint8_t label; uint32_t to; bool out; };
  auto edges_of = [&](uint32_t s, std::vector<Edge>* out) {
    out->clear();
    const SparseRow& r = v.rows[s];
    const bool wide = v.wide_first ? s < v.n_wide : (r.head & kRowWide) != 0;
    if (wide) {
      const uint64_t row = v.wide_first ? s : r.e[0];
      for (uint32_t c = 1; c < C; ++c) {
        const uint32_t e = v.wide[row * C + c];
        if ((e & 0xFFu) == 1u) out->push_back(Edge{byte_of_class[c], e >> 9, (e & kEdgeOut) != 0});
      }
    } else {
      for (int k = 0; k < 3; ++k)
        if (r.e[k]) out->push_back(Edge{(uint8_t)(r.e[k] & 0xFFu), r.e[k] >> 9, (r.e[k] & kEdgeOut) != 0});
    }
  };
  // folded trie under construction: children in a map per node
  struct Node { std::map<uint8_t, uint32_t> kid; std::vector<FoldedCand> cand; };
  std::vector<Node> nodes(1);
  // depth-first over the original trie; every pending state carries the bytes of its path
  std::vector<Edge> edges;
  std::vector<uint32_t> cps;
  std::string folded;
  struct Todo { uint32_t state; std::string path; bool out; };
  std::vector<Todo> todo;
  todo.push_back(Todo{0u, std::string(), false});
  uint64_t visited = 0;
  while (!todo.empty()) {
    Todo t = std::move(todo.back());
    todo.pop_back();
    if (++visited > S) return F;  // not a tree: corrupt blob
    if (t.out) {
      if (!fold_decode(t.path, &cps) || cps.empty() || cps.size() > 64) return F;
      const uint32_t j = v.out_off[t.state];
      if (j >= v.out_off[t.state + 1]) return F;
      FoldedCand c;
      c.value = v.out_val[j];
      c.n_chars = (uint32_t)cps.size();
      c.mask = 0;
      folded.clear();
      for (size_t i = 0; i < cps.size(); ++i) {
        const bool up = fold_is_upper(cps[i]);
        if (up) c.mask |= (uint64_t)1 << (cps.size() - 1 - i);
        fold_encode(up ? cps[i] + 32 : cps[i], &folded);
      }
      uint32_t at = 0;
      for (char ch : folded) {
        const uint8_t b = (uint8_t)ch;
        auto it = nodes[at].kid.find(b);
        if (it == nodes[at].kid.end()) {
          const uint32_t id = (uint32_t)nodes.size();
          nodes[at].kid.emplace(b, id);
          nodes.emplace_back();
          at = id;
        } else {
          at = it->second;
        }
      }
      nodes[at].cand.push_back(c);
      if (c.n_chars > F.max_chars) F.max_chars = c.n_chars;
    }
    edges_of(t.state, &edges);
    for (const Edge& e : edges) {
      if (e.to >= S) return F;
      std::string p = t.path;
      p.push_back((char)e.label);
      todo.push_back(Todo{e.to, std::move(p), e.out});
    }
  }
  // classes over the folded labels, breadth-first numbering, dense table
  bool used[256] = {false};
  for (const Node& n : nodes)
    for (const auto& kv : n.kid) used[kv.first] = true;
  uint32_t Cf = 1;
  for (int b = 0; b < 256; ++b) F.cmap[b] = (b != 0 && used[b]) ? (uint8_t)Cf++ : 0;
  const uint64_t Sf = nodes.size();
  if (Sf * Cf * 4 > kFoldedMaxBytes || Sf >= 0x7FFFFFFFu) return F;
  std::vector<uint32_t> order, nid(Sf, 0);
  order.reserve(Sf);
  order.push_back(0);
  for (size_t i = 0; i < order.size(); ++i)
    for (const auto& kv : nodes[order[i]].kid) {
      nid[kv.second] = (uint32_t)order.size();
      order.push_back(kv.second);
    }
  F.n_states = (uint32_t)Sf;
  F.n_classes = Cf;
  F.delta.assign(Sf * Cf, 0u);
  F.cand_off.assign(Sf + 1, 0u);
  for (size_t i = 0; i < Sf; ++i) {
    const Node& n = nodes[order[i]];
    F.cand_off[i + 1] = F.cand_off[i] + (uint32_t)n.cand.size();
    for (const FoldedCand& c : n.cand) F.cand.push_back(c);
    for (const auto& kv : n.kid) {
      const uint32_t to = nid[kv.second];
      F.delta[i * Cf + F.cmap[kv.first]] = to | (nodes[kv.second].cand.empty() ? 0u : kFoldCand);
    }
  }
  F.rec.assign(Sf * 8, 0u);
  for (size_t i = 0; i < Sf; ++i) {
    const uint32_t j0 = F.cand_off[i], nc = F.cand_off[i + 1] - j0;
    uint32_t* r = &F.rec[i * 8];
    bool small = nc <= 3;
    for (uint32_t j = 0; j < nc && small; ++j) small = F.cand[j0 + j].n_chars <= 32;
    if (!small) {
      r[0] = 0xFFFFFFFFu;
      continue;
    }
    r[0] = nc;
    for (uint32_t j = 0; j < nc; ++j) {
      r[1 + j] = (uint32_t)F.cand[j0 + j].value;
      r[4 + j] = (uint32_t)F.cand[j0 + j].mask;
    }
  }
  F.ok = true;
  return F;
}

}  // namespace aha
