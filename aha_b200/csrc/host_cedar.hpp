// host_cedar.hpp -- host-side builder mirroring Aha::Cedar / Aha::AC.compile.
//
// In a Crystal deployment the reference builder (src/aha/cedar.cr, src/aha/ac.cr:62-112) runs
// unchanged and hands its arrays to aha_b200_dfa_from_cedar().  Where Crystal is absent (this
// image) the same arrays are produced here, slot for slot, so that is_end?/value read the same
// raw bits and the DFA derived from them is the one the reference would scan.
//
// The double array is a reduced trie: node.value < 0 encodes base = -(value+1); value >= 0 is a
// key id stored in the base field of a leaf; free slots form a per-256-block ring through
// (value, check) = (-prev, -next).  Blocks live on three rings: Full, Closed (1 free slot or
// given up on) and Open.  All arithmetic is 32-bit two's complement with wrap (compile -fwrapv).
#pragma once
#include <cstdint>
#include <cstring>
#include <deque>
#include <string>
#include <vector>

namespace aha {

#pragma pack(push, 1)
struct CedarNode {  // == Aha::CedarX::Node(Int32), cedar.cr:29-36
  int32_t value;
  int32_t check;
  uint8_t sibling;
  uint8_t child;
  uint16_t flags;  // bits 0-8 child count, bit 9 END
};
#pragma pack(pop)
static_assert(sizeof(CedarNode) == 12, "node layout");

struct CedarOut {  // == Aha::ACX::OutNode(Int32), ac.cr:15-33
  int32_t next;
  int32_t value;
};

class HostCedar {
 public:
  static constexpr int32_t kLimit = INT32_MAX;
  static constexpr int32_t kNoFreeLeaf = INT32_MIN;
  static constexpr uint16_t kCountMask = 0x1FF;
  static constexpr uint16_t kEnd = 0x200;

  HostCedar() {  // cedar.cr:198-221
    cap_ = 256;
    size_ = 256;
    nodes_.resize(cap_);
    ring_.resize(1);
    leaf_of_.resize(256);
    leaf_cap_ = 256;
    nodes_[0] = fresh(-1, -1);
    for (int i = 1; i < 256; ++i) nodes_[i] = fresh(-(i - 1), -(i + 1));
    nodes_[1].value = -255;
    nodes_[255].check = -1;
    ring_[0] = BlockInfo{0, 0, 256, 257, 0, 1};
    for (int i = 0; i <= 256; ++i) reject_[i] = i + 1;
  }

  // Aha::Cedar#insert (cedar.cr:755-778).  Returns the key id; -2 + error() when the reference raises.
  int32_t insert(const uint8_t* key, int32_t len) {
    if (len == 0) return fail_with("Cannot insert empty key");
    int32_t known = find(key, len);
    if (known >= 0) return known;
    int32_t at = descend_creating(key, len);
    if (at < 0) return -2;
    int32_t id;
    if (free_leaf_ != kNoFreeLeaf) {
      id = -free_leaf_ - 1;
      free_leaf_ = leaf_of_[id];
    } else {
      if (leaf_count_ == leaf_cap_) {
        leaf_cap_ *= 2;
        leaf_of_.resize(leaf_cap_);
      }
      id = leaf_count_++;
    }
    nodes_[at].value = id;
    nodes_[at].flags |= kEnd;
    leaf_of_[id] = at;
    ++keys_;
    return id;
  }

  // Aha::Cedar#[]? (cedar.cr:822-828): -1 == nil
  int32_t find(const uint8_t* key, int32_t len) const {
    int32_t at = 0;
    for (int32_t i = 0; i < len; ++i) {
      at = step(at, key[i]);
      if (at < 0) return -1;
    }
    return value(at);
  }

  // cedar.cr:678-686 (private jump, one byte)
  int32_t step(int32_t from, uint8_t byte) const {
    if (nodes_[from].value >= 0) return -1;
    int32_t to = base(from) ^ byte;
    return nodes_[to].check == from ? to : -1;
  }
  // cedar.cr:441-447
  int32_t child(int32_t id, uint8_t label) const {
    int32_t c = base(id) ^ label;
    if (c < 0 || c >= size_ || nodes_[c].check != id) return -1;
    return c;
  }
  // cedar.cr:657-660
  bool is_end(int32_t id) const { return (nodes_[id].flags & kEnd) != 0 || nodes_[id].child == 0; }
  // cedar.cr:726-734
  int32_t value(int32_t id) const {
    int32_t v = nodes_[id].value;
    if (v >= 0) return v;
    const CedarNode& t = nodes_[base(id)];
    return (t.check == id && t.value >= 0 && t.value != kLimit) ? t.value : -1;
  }
  // cedar.cr:450-463: labelled children in sibling-list order, label 0 skipped
  template <class Fn> void each_child(int32_t id, Fn fn) const {
    int32_t b = base(id);
    uint8_t s = nodes_[id].child;
    if (s == 0 && b > 0) s = nodes_[b].sibling;
    while (s != 0) {
      int32_t to = b ^ s;
      if (to < 0) break;
      fn(to, s);
      s = nodes_[to].sibling;
    }
  }
  // Aha::Cedar#[](Int) (cedar.cr:707-722, 747-749): key bytes of id, false if none
  bool key_of(int32_t id, std::string* out) const {
    if (id < 0 || id >= leaf_count_) return false;
    int32_t at = leaf_of_[id];
    std::string rev;
    while (at > 0) {
      int32_t up = nodes_[at].check;
      if (up < 0) return false;
      int32_t ch = base(up) ^ at;
      if (ch != 0) rev.push_back(static_cast<char>(ch));
      at = up;
    }
    if (at != 0 || rev.empty()) return false;
    out->assign(rev.rbegin(), rev.rend());
    return true;
  }

  const CedarNode* nodes() const { return nodes_.data(); }
  int32_t array_size() const { return size_; }
  int32_t leaf_size() const { return leaf_count_; }
  int32_t key_num() const { return keys_; }
  const std::string& error() const { return err_; }

 private:
  struct BlockInfo {  // cedar.cr:98-129
    int32_t prev, next, num, reject, trial, ehead;
  };
  enum Ring { kFull, kClosed, kOpen };

  static CedarNode fresh(int32_t v, int32_t c) { return CedarNode{v, c, 0, 0, 0}; }
  int32_t base(int32_t id) const { return -(nodes_[id].value + 1); }
  int32_t fail_with(const std::string& m) {
    err_ = m;
    return -2;
  }
  int32_t& head(Ring r) { return heads_[r]; }
  uint16_t kids(int32_t id) const { return nodes_[id].flags & kCountMask; }
  void bump_kids(int32_t id) {  // cedar.cr:69-71, 416: un-masked OR kept bit-for-bit
    uint16_t f = nodes_[id].flags;
    nodes_[id].flags = static_cast<uint16_t>((f & static_cast<uint16_t>(~kCountMask)) | static_cast<uint16_t>((f & kCountMask) + 1));
  }

  // ---- block rings (cedar.cr:266-323) ----
  void unlink(int32_t bi, Ring r, bool only_one) {
    if (only_one) {
      head(r) = 0;
      return;
    }
    BlockInfo& b = ring_[bi];
    ring_[b.prev].next = b.next;
    ring_[b.next].prev = b.prev;
    if (head(r) == bi) head(r) = b.next;
  }
  void link(int32_t bi, Ring r, bool ring_empty) {
    BlockInfo& b = ring_[bi];
    if (ring_empty) {
      head(r) = b.prev = b.next = bi;
      return;
    }
    int32_t h = head(r);
    b.prev = ring_[h].prev;
    b.next = h;
    ring_[ring_[h].prev].next = bi;
    head(r) = bi;
    ring_[h].prev = bi;
  }
  void move_block(int32_t bi, Ring from, Ring to) {
    unlink(bi, from, ring_[bi].next == bi);
    link(bi, to, head(to) == 0 && ring_[bi].num != 0);
  }
  int32_t grow() {  // cedar.cr:299-313
    if (size_ == cap_) {
      cap_ *= 2;
      ring_.resize(cap_ >> 8);
      nodes_.resize(cap_);
    }
    int32_t bi = size_ >> 8;
    ring_[bi] = BlockInfo{0, 0, 256, 257, 0, size_};
    for (int i = 0; i < 256; ++i) nodes_[size_ + i] = fresh(-(((i + 255) & 255) + size_), -(((i + 1) & 255) + size_));
    link(bi, kOpen, head(kOpen) == 0);
    size_ += 256;
    return bi;
  }

  // ---- free slots (cedar.cr:329-397, 526-580) ----
  int32_t any_free_slot() {
    if (head(kClosed) != 0) return ring_[head(kClosed)].ehead;
    if (head(kOpen) != 0) return ring_[head(kOpen)].ehead;
    return grow() << 8;
  }
  int32_t take_slot(int32_t b, uint8_t label, int32_t parent) {
    int32_t e = b < 0 ? any_free_slot() : (b ^ label);
    int32_t bi = e >> 8;
    BlockInfo& blk = ring_[bi];
    if (--blk.num == 0) {
      if (bi != 0) move_block(bi, kClosed, kFull);
    } else {
      int32_t pv = nodes_[e].value, nx = nodes_[e].check;
      nodes_[-pv].check = nx;
      nodes_[-nx].value = pv;
      if (blk.ehead == e) blk.ehead = -nx;
      if (bi != 0 && blk.num == 1 && blk.trial != max_trial_) move_block(bi, kOpen, kClosed);
    }
    nodes_[e].value = kLimit;
    nodes_[e].check = parent;
    if (b < 0) nodes_[parent].value = -(e ^ label) - 1;
    return e;
  }
  void give_back(int32_t e) {
    int32_t bi = e >> 8;
    BlockInfo& blk = ring_[bi];
    if (++blk.num == 1) {
      blk.ehead = e;
      nodes_[e].value = nodes_[e].check = -e;
      if (bi != 0) move_block(bi, kFull, kClosed);
    } else {
      int32_t pv = blk.ehead;
      int32_t nx = -nodes_[pv].check;
      nodes_[e].value = -pv;
      nodes_[e].check = -nx;
      nodes_[pv].check = -e;
      nodes_[nx].value = -e;
      if ((blk.num == 2 || blk.trial == max_trial_) && bi != 0) move_block(bi, kClosed, kOpen);
      blk.trial = 0;
    }
    if (blk.reject < reject_[blk.num]) blk.reject = reject_[blk.num];
    nodes_[e].child = nodes_[e].sibling = 0;
    nodes_[e].flags = 0;
  }
  int32_t free_run_for(const uint8_t* labels, int n) {
    int32_t bi = head(kOpen);
    if (bi != 0) {
      const int32_t last = ring_[head(kOpen)].prev;
      for (;;) {
        BlockInfo& blk = ring_[bi];
        if (blk.num >= n && n < blk.reject) {
          int32_t e = blk.ehead;
          do {
            int32_t b = e ^ labels[0];
            int i = 0;
            while (i < n && nodes_[b ^ labels[i]].check < 0) ++i;
            if (i == n) {
              blk.ehead = e;
              return e;
            }
            e = -nodes_[e].check;
          } while (e != blk.ehead);
        }
        blk.reject = n;
        if (blk.reject < reject_[blk.num]) reject_[blk.num] = blk.reject;
        int32_t nxt = blk.next;
        if (++blk.trial == max_trial_) move_block(bi, kOpen, kClosed);
        if (bi == last) break;
        bi = nxt;
      }
    }
    return grow() << 8;
  }

  // ---- sibling lists (cedar.cr:403-417, 498-523; unordered mode) ----
  void add_sibling(int32_t parent, int32_t b, uint8_t label, bool has_child) {
    uint8_t* link_at = &nodes_[parent].child;
    if (has_child && *link_at == 0) link_at = &nodes_[b ^ *link_at].sibling;
    nodes_[b ^ label].sibling = *link_at;
    *link_at = label;
    bump_kids(parent);
  }
  int collect_labels(int32_t b, uint8_t first, uint8_t extra, bool with_extra, uint8_t* out) const {
    int n = 0;
    uint8_t c = first;
    if (c == 0) {
      out[n++] = 0;
      c = nodes_[b].sibling;
    }
    if (with_extra) out[n++] = extra;
    for (; c != 0; c = nodes_[b ^ c].sibling) out[n++] = c;
    return n;
  }

  // cedar.cr:582-655: slot base_n^label_n is taken by another family; move the smaller family
  int32_t relocate(int32_t from_n, int32_t base_n, uint8_t label_n) {
    const int32_t contested = base_n ^ label_n;
    const int32_t from_p = nodes_[contested].check;
    const int32_t base_p = base(from_p);
    const bool move_self = kids(from_n) < kids(from_p);
    uint8_t labels[257] = {0};
    const int n = move_self ? collect_labels(base_n, nodes_[from_n].child, label_n, true, labels)
                            : collect_labels(base_p, nodes_[from_p].child, 255, false, labels);
    int32_t nb = (n == 1 ? any_free_slot() : free_run_for(labels, n)) ^ labels[0];
    const int32_t owner = move_self ? from_n : from_p;
    const int32_t old_base = move_self ? base_n : base_p;
    if (move_self && labels[0] == label_n) nodes_[owner].child = label_n;
    nodes_[owner].value = -nb - 1;
    for (int i = 0; i < n; ++i) {
      const uint8_t lab = labels[i];
      const int32_t dst = take_slot(nb, lab, owner);
      const int32_t src = old_base ^ lab;
      nodes_[dst].sibling = (i + 1 == n) ? 0 : labels[i + 1];
      if (move_self && src == contested) continue;  // the brand-new child: nothing to carry over
      nodes_[dst].value = nodes_[src].value;
      if (nodes_[src].value >= 0 && nodes_[src].value != kLimit) leaf_of_[nodes_[src].value] = dst;
      nodes_[dst].flags = nodes_[src].flags;
      if (nodes_[dst].value < 0 && lab != 0) {  // re-parent grandchildren
        uint8_t c = nodes_[src].child;
        nodes_[dst].child = c;
        const int32_t gb = base(dst);
        do {
          CedarNode& g = nodes_[gb ^ c];
          g.check = dst;
          c = g.sibling;
        } while (c != 0);
      }
      if (!move_self && src == from_n) from_n = dst;
      if (!move_self && src == contested) {
        // the evicted slot is handed straight to the new child; flags are NOT cleared (Q4)
        add_sibling(from_n, contested ^ label_n, label_n, true);
        nodes_[src].child = 0;
        nodes_[src].value = kLimit;
        nodes_[src].check = from_n;
      } else {
        give_back(src);
      }
    }
    return move_self ? (nb ^ label_n) : contested;
  }

  // cedar.cr:244-259
  int32_t child_creating(int32_t from, uint8_t label) {
    const int32_t b = base(from);
    int32_t to = b ^ label;
    if (b < 0 || nodes_[to].check < 0) {
      const bool has_child = b >= 0 && nodes_[b ^ nodes_[from].child].check == from;
      to = take_slot(b, label, from);
      add_sibling(from, to ^ label, label, has_child);
    } else if (nodes_[to].check != from) {
      to = relocate(from, b, label);
    }
    return to;
  }
  // cedar.cr:224-241
  int32_t descend_creating(const uint8_t* key, int32_t len) {
    int32_t at = 0;
    for (int32_t pos = 0; pos < len; ++pos) {
      const int32_t v = nodes_[at].value;
      if (v >= 0 && v != kLimit) {  // leaf turns interior: push its id down to a label-0 child
        const int32_t z = child_creating(at, 0);
        nodes_[z].value = v;
        leaf_of_[v] = z;
      }
      if (key[pos] == 0) return fail_with("key[" + std::to_string(pos) + "] is zero");
      at = child_creating(at, key[pos]);
    }
    return nodes_[at].value < 0 ? child_creating(at, 0) : at;
  }

  std::vector<CedarNode> nodes_;
  std::vector<BlockInfo> ring_;
  std::vector<int32_t> leaf_of_;
  int32_t reject_[257];
  int32_t heads_[3] = {0, 0, 0};
  int32_t cap_ = 0, size_ = 0, keys_ = 0, leaf_count_ = 0, leaf_cap_ = 0;
  int32_t free_leaf_ = kNoFreeLeaf;
  int32_t max_trial_ = 1;
  std::string err_;
};

// Aha::AC.compile(da) (ac.cr:71-112): failure links, one-deep output chain, key lengths.
struct HostAC {
  std::vector<int32_t> fails;
  std::vector<CedarOut> output;
  std::vector<uint32_t> key_lens;

  void compile(const HostCedar& da) {
    const int32_t n = da.array_size();
    fails.assign(n, -1);
    output.assign(n, CedarOut{-1, -1});
    key_lens.assign(da.leaf_size() > 0 ? da.leaf_size() : 1, 0u);
    struct Item {
      int32_t id;
      int32_t depth;
    };
    std::deque<Item> bfs;
    fails[0] = 0;
    da.each_child(0, [&](int32_t c, uint8_t) {
      fails[c] = 0;
      bfs.push_back(Item{c, 1});
    });
    while (!bfs.empty()) {
      const Item it = bfs.front();
      bfs.pop_front();
      if (da.is_end(it.id)) {
        const int32_t vk = da.value(it.id);
        if (vk >= 0) key_lens[vk] = static_cast<uint32_t>(it.depth);  // reference writes [-1] when vk<0 (Q4): dropped
        output[it.id].value = vk;
      }
      da.each_child(it.id, [&](int32_t c, uint8_t label) {
        bfs.push_back(Item{c, it.depth + 1});
        int32_t f = it.id;
        while (f != 0) {
          const int32_t via = da.child(fails[f], label);
          if (via >= 0) {
            f = via;
            break;
          }
          f = fails[f];
        }
        fails[c] = f;
        if (da.is_end(f)) output[c].next = f;
      });
    }
  }
};

}  // namespace aha
