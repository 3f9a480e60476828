// oracle/aha_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the chenkovsky/aha `Aha::AC` hot path (Crystal, cannot be
// built in this image: no crystal/shards/LLVM).  Every function cites the
// reference file:line it follows.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load this library; the
// product (aha_b200/) never links, imports or executes anything under oracle/.
//
// Parity pinning: the known-answer vectors of spec/ac_spec.cr:5-54,
// spec/ac_longest_match_spec.cr:5-63, spec/cedar_match_spec.cr:4-63,
// spec/cedar_spec.cr:4-13,28-293 and README.md:29-38 are replayed against this
// file by tests/test_oracle_kat.py.  On-disk (save/load) byte parity is
// UNPINNED (format lives in the un-vendored shard chenkovsky/super_io).
//
// Arithmetic is int32 with wrap-around (Crystal 0.23 integers wrap silently;
// cedar.cr:81-83 evaluates -(value+1) on value == Int32::MAX): build with -fwrapv.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <string>
#include <thread>
#include <atomic>
#include <vector>
#include <algorithm>

#include "../include/aha_synth.h"  // corpus generator only (bench input), not part of the restatement

namespace {

// cedar.cr:29-36  Node(Int32): value, check, sibling, child, flags  (12 bytes)
struct Node {
  int32_t value;
  int32_t check;
  uint8_t sibling;
  uint8_t child;
  uint16_t flags;
};
static_assert(sizeof(Node) == 12, "Node(Int32) is 12 bytes");

// cedar.cr:98-129
struct Block {
  int32_t prev, next, num, reject, trial, ehead;
};

const int32_t VALUE_LIMIT = INT32_MAX;  // cedar.cr:10-12
const int32_t VALUE_MIN = INT32_MIN;    // cedar.cr:14-16
const uint16_t CHILD_NUM_MASK = (1 << 9) - 1;  // cedar.cr:62
const uint16_t END_MASK = 1 << 9;              // cedar.cr:63

struct Cedar {
  std::vector<Node> array;
  std::vector<Block> blocks;
  std::vector<int32_t> leafs;
  int32_t reject[257];
  int32_t key_num, key_capacity, bheadF, bheadC, bheadO, array_size, capacity;
  int32_t max_trial, leaf_size, free_leaf_slot;
  std::string err;

  // cedar.cr:81-83
  int32_t base_of(int32_t id) const { return -(array[id].value + 1); }
  // cedar.cr:65-71 (the OR is un-masked on purpose: bit-for-bit)
  uint16_t child_num(int32_t id) const { return array[id].flags & CHILD_NUM_MASK; }
  void set_child_num(int32_t id, uint16_t v) {
    array[id].flags = (uint16_t)((array[id].flags & (uint16_t)~CHILD_NUM_MASK) | v);
  }

  static Node mk(int32_t value, int32_t check) {  // cedar.cr:56-60
    Node n; n.value = value; n.check = check; n.sibling = 0; n.child = 0; n.flags = 0; return n;
  }
  static Block mkb(int32_t prev, int32_t next, int32_t trial, int32_t ehead) {  // cedar.cr:127
    Block b; b.prev = prev; b.next = next; b.trial = trial; b.ehead = ehead; b.num = 256; b.reject = 257; return b;
  }

  // cedar.cr:198-221
  Cedar() {
    capacity = 256; key_num = 0; leaf_size = 0; key_capacity = capacity;
    leafs.assign(key_capacity, 0);
    array.assign(capacity, mk(0, 0));
    array_size = capacity;
    blocks.assign(capacity >> 8, mkb(0, 0, 0, 0));
    max_trial = 1;
    array[0] = mk(-1, -1);
    for (int i = 1; i < 256; i++) array[i] = mk(-(i - 1), -(i + 1));
    array[1].value = -255;
    array[255].check = -1;
    blocks[0] = mkb(0, 0, 0, 1);
    for (int i = 0; i < 257; i++) reject[i] = i + 1;
    bheadF = bheadC = bheadO = 0;
    free_leaf_slot = VALUE_MIN;
  }

  int32_t* head(int which) { return which == 0 ? &bheadF : which == 1 ? &bheadC : &bheadO; }

  // cedar.cr:266-278
  void pop_block(int32_t bi, int32_t* head_in, bool last) {
    if (last) { *head_in = 0; return; }
    Block& b = blocks[bi];
    blocks[b.prev].next = b.next;
    blocks[b.next].prev = b.prev;
    if (bi == *head_in) *head_in = b.next;
  }
  // cedar.cr:285-296
  void push_block(int32_t bi, int32_t* head_out, bool empty) {
    Block& b = blocks[bi];
    if (empty) { *head_out = bi; b.prev = bi; b.next = bi; return; }
    int32_t tail_idx = *head_out;
    b.prev = blocks[tail_idx].prev;
    b.next = *head_out;
    blocks[blocks[tail_idx].prev].next = bi;
    *head_out = bi;
    blocks[tail_idx].prev = bi;
  }
  // cedar.cr:299-313
  int32_t add_block() {
    if (array_size == capacity) {
      capacity *= 2;
      blocks.resize(capacity >> 8, mkb(0, 0, 0, 0));
      array.resize(capacity, mk(0, 0));
    }
    blocks[array_size >> 8] = mkb(0, 0, 0, array_size);
    for (int i = 0; i < 256; i++)
      array[array_size + i] = mk(-(((i + 255) & 255) + array_size), -(((i + 1) & 255) + array_size));
    push_block(array_size >> 8, &bheadO, bheadO == 0);
    array_size += 256;
    return (array_size >> 8) - 1;
  }
  // cedar.cr:320-323
  void transfer_block(int32_t bi, int32_t* head_in, int32_t* head_out) {
    pop_block(bi, head_in, bi == blocks[bi].next);
    push_block(bi, head_out, *head_out == 0 && blocks[bi].num != 0);
  }
  // cedar.cr:526-530
  int32_t find_place() {
    if (bheadC != 0) return blocks[bheadC].ehead;
    if (bheadO != 0) return blocks[bheadO].ehead;
    return add_block() << 8;
  }
  // cedar.cr:329-359
  int32_t pop_enode(int32_t base, uint8_t label, int32_t from) {
    int32_t e = base < 0 ? find_place() : (base ^ label);
    int32_t bi = e >> 8;
    Block& b = blocks[bi];
    b.num -= 1;
    if (b.num == 0) {
      if (bi != 0) transfer_block(bi, &bheadC, &bheadF);
    } else {
      Node n = array[e];
      array[-n.value].check = n.check;
      array[-n.check].value = n.value;
      if (e == b.ehead) b.ehead = -n.check;
      if (bi != 0 && b.num == 1 && b.trial != max_trial) transfer_block(bi, &bheadO, &bheadC);
    }
    array[e].value = VALUE_LIMIT;
    array[e].check = from;
    if (base < 0) array[from].value = -(e ^ label) - 1;
    return e;
  }
  // cedar.cr:361-397
  void push_enode(int32_t e) {
    int32_t bi = e >> 8;
    Block& b = blocks[bi];
    b.num += 1;
    if (b.num == 1) {
      b.ehead = e;
      array[e].value = -e;
      array[e].check = -e;
      if (bi != 0) transfer_block(bi, &bheadF, &bheadC);
    } else {
      int32_t prev = b.ehead;
      int32_t next_ = -array[prev].check;
      array[e].value = -prev;
      array[e].check = -next_;
      array[prev].check = -e;
      array[next_].value = -e;
      if (b.num == 2 || b.trial == max_trial) {
        if (bi != 0) transfer_block(bi, &bheadC, &bheadO);
      }
      b.trial = 0;
    }
    if (b.reject < reject[b.num]) b.reject = reject[b.num];
    array[e].child = 0;
    array[e].sibling = 0;
    array[e].flags = 0;
  }
  // cedar.cr:403-417 (unordered mode only: @ordered == false, cedar.cr:198)
  void push_sibling(int32_t from, int32_t base, uint8_t label, bool has_child) {
    uint8_t* child_ptr = &array[from].child;
    bool keep_order = (*child_ptr == 0);
    if (has_child && keep_order) child_ptr = &array[base ^ *child_ptr].sibling;
    array[base ^ label].sibling = *child_ptr;
    *child_ptr = label;
    set_child_num(from, (uint16_t)(child_num(from) + 1));
  }
  // cedar.cr:498-523
  int set_child(int32_t base, uint8_t c, uint8_t label, bool append_label, uint8_t* children) {
    int idx = 0;
    if (c == 0) {
      children[idx++] = c;
      c = array[base ^ c].sibling;
    }
    if (append_label) children[idx++] = label;
    while (c != 0) {
      children[idx++] = c;
      c = array[base ^ c].sibling;
    }
    return idx;
  }
  // cedar.cr:533-580
  int32_t find_places(const uint8_t* ch, int n) {
    int32_t bi = bheadO;
    if (bi != 0) {
      int32_t bz = blocks[bheadO].prev;
      int32_t nc = n;
      while (true) {
        Block& b = blocks[bi];
        if (b.num >= nc && nc < b.reject) {
          int32_t e = b.ehead;
          while (true) {
            int32_t base = e ^ ch[0];
            for (int i = 0; i < n; i++) {
              if (!(array[base ^ ch[i]].check < 0)) break;
              if (i == n - 1) { b.ehead = e; return e; }
            }
            e = -array[e].check;
            if (e == b.ehead) break;
          }
        }
        b.reject = nc;
        if (b.reject < reject[b.num]) reject[b.num] = b.reject;
        int32_t bin = b.next;
        b.trial += 1;
        if (b.trial == max_trial) transfer_block(bi, &bheadO, &bheadC);
        if (bi == bz) break;
        bi = bin;
      }
    }
    return add_block() << 8;
  }
  // cedar.cr:582-655
  int32_t resolve(int32_t from_n, int32_t base_n, uint8_t label_n) {
    int32_t to_pn = base_n ^ label_n;
    int32_t from_p = array[to_pn].check;
    int32_t base_p = base_of(from_p);
    bool flag = child_num(from_n) < child_num(from_p);  // consult, cedar.cr:432-434
    uint8_t children_[257];
    memset(children_, 0, sizeof(children_));
    int n;
    if (flag) n = set_child(base_n, array[from_n].child, label_n, true, children_);
    else n = set_child(base_p, array[from_p].child, 255, false, children_);
    int32_t base = (n == 1) ? find_place() : find_places(children_, n);
    base ^= children_[0];
    int32_t from, base_;
    if (flag) { from = from_n; base_ = base_n; } else { from = from_p; base_ = base_p; }
    if (flag && children_[0] == label_n) array[from].child = label_n;
    array[from].value = -base - 1;
    for (int i = 0; i < n; i++) {
      uint8_t chl = children_[i];
      int32_t to = pop_enode(base, chl, from);
      int32_t to_ = base_ ^ chl;
      array[to].sibling = (i == n - 1) ? 0 : children_[i + 1];
      if (flag && to_ == to_pn) continue;
      array[to].value = array[to_].value;
      if (array[to_].value >= 0 && array[to_].value != VALUE_LIMIT) leafs[array[to_].value] = to;
      array[to].flags = array[to_].flags;
      if (array[to].value < 0 && chl != 0) {
        uint8_t c = array[to_].child;
        array[to].child = c;
        int32_t nb = base_of(to);
        int32_t p = nb ^ c;
        array[p].check = to;
        c = array[p].sibling;
        while (c != 0) {
          p = nb ^ c;
          array[p].check = to;
          c = array[p].sibling;
        }
      }
      if (!flag && to_ == from_n) from_n = to;
      if (!flag && to_ == to_pn) {
        push_sibling(from_n, to_pn ^ label_n, label_n, true);
        array[to_].child = 0;
        array[to_].value = VALUE_LIMIT;
        array[to_].check = from_n;
      } else {
        push_enode(to_);
      }
    }
    if (flag) return base ^ label_n;
    return to_pn;
  }
  // cedar.cr:244-259
  int32_t follow(int32_t from, uint8_t label) {
    int32_t base = base_of(from);
    int32_t to = base ^ label;
    if (base < 0 || array[to].check < 0) {
      bool has_child = base >= 0 && (array[base ^ array[from].child].check == from);
      to = pop_enode(base, label, from);
      push_sibling(from, to ^ label, label, has_child);
    } else if (array[to].check != from) {
      to = resolve(from, base, label);
    }
    return to;
  }
  // cedar.cr:224-241 ; returns -1 and sets err on a NUL byte (cedar.cr:235)
  int32_t get(const uint8_t* key, int32_t len) {
    int32_t from = 0;
    for (int32_t pos = 0; pos < len; pos++) {
      int32_t value = array[from].value;
      if (value >= 0 && value != VALUE_LIMIT) {
        int32_t to = follow(from, 0);
        array[to].value = value;
        leafs[value] = to;
      }
      if (key[pos] == 0) { err = "key[" + std::to_string(pos) + "] is zero"; return -1; }
      from = follow(from, key[pos]);
    }
    return array[from].value < 0 ? follow(from, 0) : from;
  }
  // cedar.cr:678-694
  int32_t jump1(uint8_t byte, int32_t from) const {
    if (array[from].value >= 0) return -1;
    int32_t to = base_of(from) ^ byte;
    if (array[to].check != from) return -1;
    return to;
  }
  int32_t jump(const uint8_t* p, int32_t len, int32_t from) const {
    for (int32_t i = 0; i < len; i++) {
      from = jump1(p[i], from);
      if (from < 0) return -1;
    }
    return from;
  }
  // cedar.cr:726-734
  int32_t value(int32_t id) const {
    int32_t val = array[id].value;
    if (val >= 0) return val;
    int32_t to = base_of(id);
    const Node& t = array[to];
    if (t.check == id && t.value >= 0 && t.value != VALUE_LIMIT) return t.value;
    return -1;
  }
  // cedar.cr:822-828 ; -1 = nil
  int32_t lookup(const uint8_t* key, int32_t len) const {
    int32_t to = jump(key, len, 0);
    if (to < 0) return -1;
    return value(to);
  }
  // cedar.cr:755-778 ; -2 = raised (err set)
  int32_t insert(const uint8_t* key, int32_t len) {
    if (len == 0) { err = "Cannot insert empty key"; return -2; }
    int32_t id = lookup(key, len);
    if (id >= 0) return id;
    int32_t p = get(key, len);
    if (p < 0) return -2;
    if (free_leaf_slot != VALUE_MIN) {
      int32_t cur = -free_leaf_slot - 1;
      free_leaf_slot = leafs[cur];
      id = cur;
    } else {
      if (leaf_size == key_capacity) {
        key_capacity *= 2;
        leafs.resize(key_capacity, 0);
      }
      leaf_size += 1;
      id = leaf_size - 1;
    }
    array[p].value = id;
    array[p].flags |= END_MASK;  // end!, cedar.cr:77-79
    leafs[id] = p;
    key_num += 1;
    return id;
  }
  // cedar.cr:441-447
  int32_t child(int32_t id, uint8_t label) const {
    int32_t cid = base_of(id) ^ label;
    if (cid < 0 || cid >= array_size || array[cid].check != id) return -1;
    return cid;
  }
  // cedar.cr:657-660
  bool is_end(int32_t id) const {
    if (array[id].flags & END_MASK) return true;
    return array[id].child == 0;
  }
  // cedar.cr:450-463
  template <class F> void children(int32_t id, F f) const {
    int32_t base = base_of(id);
    uint8_t s = array[id].child;
    if (s == 0 && base > 0) s = array[base].sibling;
    while (s != 0) {
      int32_t to = base ^ s;
      if (to < 0) break;
      f(to, s);
      s = array[to].sibling;
    }
  }
  // cedar.cr:707-722 without the trailing NUL; false if "no path"/"invalid key"
  bool key_of(int32_t id, std::string& out) const {
    std::string rev;
    while (id > 0) {
      int32_t from = array[id].check;
      if (from < 0) return false;
      int32_t chr = base_of(from) ^ id;
      if (chr != 0) rev.push_back((char)(uint8_t)chr);
      id = from;
    }
    if (id != 0 || rev.empty()) return false;
    out.assign(rev.rbegin(), rev.rend());
    return true;
  }
};

// matcher.cr:2-11
struct Hit {
  int32_t start, end, value;
};

struct OutNode {  // ac.cr:15-33
  int32_t next, value;
};

struct AC {
  Cedar* da;
  std::vector<OutNode> output;
  std::vector<int32_t> fails;
  std::vector<uint32_t> key_lens;
  std::vector<Hit> hits;  // result buffer of the last call
};

// ac.cr:71-112
AC* ac_compile(Cedar* da) {
  AC* ac = new AC();
  ac->da = da;
  int32_t nlen = da->array_size;
  ac->fails.assign(nlen, -1);
  OutNode on; on.next = -1; on.value = -1;
  ac->output.assign(nlen, on);
  ac->key_lens.assign(da->leaf_size > 0 ? da->leaf_size : 1, 0);
  struct QE { int32_t id; uint8_t label; int32_t len; };
  std::deque<QE> q;
  const int32_t ro = 0;
  ac->fails[ro] = ro;
  da->children(ro, [&](int32_t id, uint8_t label) {
    ac->fails[id] = ro;
    q.push_back(QE{id, label, 1});
  });
  while (!q.empty()) {
    QE n = q.front();
    q.pop_front();
    int32_t nid = n.id, l = n.len;
    if (da->is_end(nid)) {
      int32_t vk = da->value(nid);
      // reference writes key_lens[-1] out of bounds when vk == -1 (stale END flag,
      // SURVEY Q4); the write is unobservable, so it is skipped here.
      if (vk >= 0) ac->key_lens[vk] = (uint32_t)l;
      ac->output[nid].value = vk;
    }
    da->children(nid, [&](int32_t cid, uint8_t clabel) {
      q.push_back(QE{cid, clabel, l + 1});
      int32_t fid = nid;
      while (fid != ro) {
        int32_t fs = ac->fails[fid];
        if (da->child(fs, clabel) >= 0) {
          fid = da->child(fs, clabel);
          break;
        }
        fid = ac->fails[fid];
      }
      ac->fails[cid] = fid;
      if (da->is_end(fid)) ac->output[cid].next = fid;
    });
  }
  return ac;
}

const uint32_t KEY_LEN_MASK = 0x7FFFFFFFu;  // ac.cr:237

// ac.cr:265-278
template <class F> inline void fetch(const AC* ac, int32_t idx, int32_t nid, F f) {
  const OutNode* e = &ac->output[nid];
  while (e->value >= 0) {
    int32_t val = e->value;
    if (ac->key_lens[val] < KEY_LEN_MASK) {
      int32_t len = (int32_t)(ac->key_lens[val] & KEY_LEN_MASK);
      f(Hit{idx - len + 1, idx + 1, val});
    }
    if (!(e->next >= 0)) break;
    e = &ac->output[e->next];
  }
}
// ac.cr:249-263
template <class F> inline void fetch_one(const AC* ac, int32_t idx, int32_t nid, F f) {
  const OutNode* e = &ac->output[nid];
  while (e->value >= 0) {
    int32_t val = e->value;
    if (ac->key_lens[val] < KEY_LEN_MASK) {
      int32_t len = (int32_t)(ac->key_lens[val] & KEY_LEN_MASK);
      f(Hit{idx - len + 1, idx + 1, val});
      break;
    }
    if (!(e->next >= 0)) break;
    e = &ac->output[e->next];
  }
}

// ac.cr:176-192
template <class F> inline void match_(const AC* ac, const uint8_t* seq, int64_t n, F yield) {
  const Cedar* da = ac->da;
  int32_t nid = 0;
  for (int64_t i = 0; i < n; i++) {
    uint8_t b = seq[i];
    while (true) {
      int32_t nid_ = da->child(nid, b);
      if (nid_ >= 0) {
        nid = nid_;
        if (da->is_end(nid)) yield((int32_t)i, nid);
        break;
      }
      if (nid == 0) break;
      nid = ac->fails[nid];
    }
  }
}

// ac.cr:118-143
template <class F>
inline void match_longest_(const AC* ac, const uint8_t* seq, int64_t n, bool intersectable, F yield) {
  const Cedar* da = ac->da;
  int32_t nid = 0;
  int32_t prev_i = -1, prev_nid = -1;
  for (int64_t i = 0; i < n; i++) {
    uint8_t b = seq[i];
    while (true) {
      int32_t nid_ = da->child(nid, b);
      if (nid_ >= 0) {
        nid = nid_;
        if (da->is_end(nid)) { prev_i = (int32_t)i; prev_nid = nid; }
        break;
      }
      if (prev_i != -1) {
        yield(prev_i, prev_nid);
        prev_i = -1;
        if (!intersectable) nid = 0;
      }
      if (nid == 0) break;
      nid = ac->fails[nid];
    }
  }
  if (prev_i != -1) yield(prev_i, prev_nid);
}

// 256-bit separator set + its BitArray size (ac.cr:326,333: bytes >= sep.size pass)
struct Sep {
  const uint8_t* bits;
  int32_t size;
  bool rejects(uint32_t chr) const {  // "chr < sep.size && !sep[chr]"
    return chr < (uint32_t)size && !((bits[chr >> 3] >> (chr & 7)) & 1);
  }
  bool is_sep(uint32_t chr) const {  // cedar.cr:1067-1069
    return chr < (uint32_t)size && ((bits[chr >> 3] >> (chr & 7)) & 1);
  }
};

// matcher.cr:14-22 on valid UTF-8: char index of byte j = (#non-continuation bytes in [0..j]) - 1
void char_map(const uint8_t* s, int64_t n, std::vector<int32_t>& cob) {
  cob.resize(n);
  int32_t c = -1;
  for (int64_t i = 0; i < n; i++) {
    if ((s[i] & 0xC0) != 0x80) c++;
    cob[i] = c;
  }
}

// --- CedarX#match_longest (cedar.cr:1089-1121) on a decoded char slice -------------
// Crystal Char#uppercase?/downcase/upcase restricted to ASCII + Latin-1 + Greek + Cyrillic
// basic ranges (simple 1:1 mappings).  Anything else is treated as caseless.
inline bool cp_is_upper(uint32_t c) {
  if (c >= 'A' && c <= 'Z') return true;
  if (c >= 0xC0 && c <= 0xDE && c != 0xD7) return true;
  if (c >= 0x391 && c <= 0x3A9 && c != 0x3A2) return true;
  if (c >= 0x410 && c <= 0x42F) return true;
  return false;
}
inline uint32_t cp_down(uint32_t c) { return cp_is_upper(c) ? c + 32 : c; }
inline uint32_t cp_up(uint32_t c) {
  if (c >= 'a' && c <= 'z') return c - 32;
  if (c >= 0xE0 && c <= 0xFE && c != 0xF7) return c - 32;
  if (c >= 0x3B1 && c <= 0x3C9 && c != 0x3C2) return c - 32;
  if (c >= 0x430 && c <= 0x44F) return c - 32;
  return c;
}
inline int cp_encode(uint32_t c, uint8_t* o) {
  if (c < 0x80) { o[0] = (uint8_t)c; return 1; }
  if (c < 0x800) { o[0] = 0xC0 | (c >> 6); o[1] = 0x80 | (c & 0x3F); return 2; }
  if (c < 0x10000) { o[0] = 0xE0 | (c >> 12); o[1] = 0x80 | ((c >> 6) & 0x3F); o[2] = 0x80 | (c & 0x3F); return 3; }
  o[0] = 0xF0 | (c >> 18); o[1] = 0x80 | ((c >> 12) & 0x3F); o[2] = 0x80 | ((c >> 6) & 0x3F); o[3] = 0x80 | (c & 0x3F);
  return 4;
}
// valid UTF-8 only (contract); malformed bytes decode as single Latin-1-ish units
void decode_utf8(const uint8_t* s, int64_t n, std::vector<uint32_t>& out) {
  out.clear();
  int64_t i = 0;
  while (i < n) {
    uint8_t b = s[i];
    if (b < 0x80) { out.push_back(b); i += 1; }
    else if ((b & 0xE0) == 0xC0 && i + 1 < n) { out.push_back(((b & 0x1F) << 6) | (s[i + 1] & 0x3F)); i += 2; }
    else if ((b & 0xF0) == 0xE0 && i + 2 < n) { out.push_back(((b & 0x0F) << 12) | ((s[i + 1] & 0x3F) << 6) | (s[i + 2] & 0x3F)); i += 3; }
    else if ((b & 0xF8) == 0xF0 && i + 3 < n) { out.push_back(((b & 0x07) << 18) | ((s[i + 1] & 0x3F) << 12) | ((s[i + 2] & 0x3F) << 6) | (s[i + 3] & 0x3F)); i += 4; }
    else { out.push_back(b); i += 1; }
  }
}
// cedar.cr:946-952
inline int32_t jump_char(const Cedar* da, uint32_t cp, int32_t from) {
  uint8_t buf[4];
  int k = cp_encode(cp, buf);
  return da->jump(buf, k, from);
}

// cedar.cr:984-1027 (prefix, Slice(Char), limit = -1) fused with the candidate
// selection of cedar.cr:1097-1105.  early_exit=false walks to the end of the slice
// like the reference (Q9: O(n^2)); true stops when the frontier is empty (same result).
void trie_longest(const Cedar* da, const uint32_t* s, int64_t size, bool ignore_case, const Sep* sep,
                  bool early_exit, std::vector<Hit>& out) {
  std::vector<int32_t> queue, new_queue;
  int64_t offset = 0;
  while (offset < size) {
    if (sep) {
      while (offset < size && sep->is_sep(s[offset])) offset++;
    }
    int32_t last_vk = 0;
    int64_t last_char_num = 0;
    queue.clear();
    queue.push_back(0);
    new_queue.clear();
    int64_t char_num = 0;
    for (int64_t p = offset; p < size; p++) {
      uint32_t chr = s[p];
      char_num += 1;
      uint32_t other = chr;
      if (ignore_case) other = cp_is_upper(chr) ? cp_down(chr) : cp_up(chr);
      for (size_t qi = 0; qi < queue.size(); qi++) {
        int32_t node = queue[qi];
        for (int pass = 0; pass < 2; pass++) {
          if (pass == 1 && !(ignore_case && other != chr)) break;
          int32_t nn = jump_char(da, pass == 0 ? chr : other, node);
          if (nn >= 0) {
            new_queue.push_back(nn);
            int32_t vk = da->value(nn);
            if (vk >= 0) {
              bool ok = true;
              if (sep) ok = (char_num + offset == size) || sep->is_sep(s[char_num + offset]);
              if (ok && char_num > last_char_num) { last_vk = vk; last_char_num = char_num; }
            }
          }
        }
      }
      queue.swap(new_queue);
      new_queue.clear();
      if (early_exit && queue.empty()) break;
    }
    if (last_char_num == 0) {
      offset += 1;
      if (sep) {
        while (offset < size && !sep->is_sep(s[offset])) offset++;
      }
    } else {
      out.push_back(Hit{(int32_t)offset, (int32_t)(offset + last_char_num), last_vk});
      offset += last_char_num;
    }
  }
}

}  // namespace

// ------------------------------- C ABI for ctypes ---------------------------------
// ACX#match(seq : Array(Char) | Slice(Char), sep) (ac.cr:342-364): the walk is over the characters' bytes, the two
// neighbour tests look at the neighbouring CHARACTER's code point (`chr.ord < sep.size && !sep[chr.ord]`).  Hits come out
// in byte offsets here; the caller remaps them (the reference yields character offsets from this overload).
template <class F>
inline void match_sep_chars_(const AC* ac, const uint8_t* seq, int64_t n, const Sep& sep, F yield) {
  std::vector<uint32_t> chars;   // seq as Array(Char)
  std::vector<int32_t> cob;      // char_of_byte (matcher.cr:24-32)
  decode_utf8(seq, n, chars);
  char_map(seq, n, cob);
  const int64_t n_chars = (int64_t)chars.size();
  match_(ac, seq, n, [&](int32_t idx, int32_t nid) {
    const int64_t chr_idx = cob[idx];
    if (chr_idx + 1 < n_chars && sep.rejects(chars[chr_idx + 1])) return;
    fetch(ac, idx, nid, [&](Hit h) {
      if (h.start > 0 && cob[h.start] > 0 && sep.rejects(chars[cob[h.start] - 1])) return;  // (keys are whole characters)
      yield(h);
    });
  });
}

extern "C" {

void* orc_cedar_new() { return new Cedar(); }
void orc_cedar_free(void* c) { delete (Cedar*)c; }
// returns key id, or -2 when the reference would raise (message via orc_cedar_error)
int32_t orc_cedar_insert(void* c, const uint8_t* key, int32_t len) { return ((Cedar*)c)->insert(key, len); }
int32_t orc_cedar_lookup(void* c, const uint8_t* key, int32_t len) { return ((Cedar*)c)->lookup(key, len); }
const char* orc_cedar_error(void* c) { return ((Cedar*)c)->err.c_str(); }
int32_t orc_cedar_array_size(void* c) { return ((Cedar*)c)->array_size; }
int32_t orc_cedar_leaf_size(void* c) { return ((Cedar*)c)->leaf_size; }
int32_t orc_cedar_key_num(void* c) { return ((Cedar*)c)->key_num; }
const void* orc_cedar_nodes(void* c) { return ((Cedar*)c)->array.data(); }
// trie[i] (cedar.cr:747-749): writes the key bytes, returns length or -1
int32_t orc_cedar_key(void* c, int32_t id, uint8_t* out, int32_t cap) {
  Cedar* d = (Cedar*)c;
  if (id < 0 || id >= d->leaf_size) return -1;
  std::string k;
  if (!d->key_of(d->leafs[id], k)) return -1;
  if ((int32_t)k.size() > cap) return -1;
  memcpy(out, k.data(), k.size());
  return (int32_t)k.size();
}

// ACX.compile(da) (ac.cr:71-112); the AC borrows the cedar
void* orc_ac_compile(void* c) { return ac_compile((Cedar*)c); }
// ACX.compile(keys) (ac.cr:62-69): keys concatenated, offs has n+1 entries.
// Returns NULL when the reference raises; *err_key = offending index.
void* orc_ac_compile_keys(const uint8_t* keys, const int64_t* offs, int64_t n, void** cedar_out, int64_t* err_key) {
  Cedar* da = new Cedar();
  for (int64_t i = 0; i < n; i++) {
    int32_t kid = da->insert(keys + offs[i], (int32_t)(offs[i + 1] - offs[i]));
    if (kid != (int32_t)i) {
      if (err_key) *err_key = i;
      delete da;
      return nullptr;
    }
  }
  *cedar_out = da;
  return ac_compile(da);
}
void orc_ac_free(void* a) { delete (AC*)a; }
const int32_t* orc_ac_fails(void* a) { return ((AC*)a)->fails.data(); }
const int32_t* orc_ac_output(void* a) { return (const int32_t*)((AC*)a)->output.data(); }
const uint32_t* orc_ac_key_lens(void* a) { return ((AC*)a)->key_lens.data(); }
const Hit* orc_ac_hits(void* a) { return ((AC*)a)->hits.data(); }

enum { ORC_UNITS_CHAR = 1, ORC_SEP = 2, ORC_LONGEST = 4, ORC_INTERSECTABLE = 8, ORC_SEP_CODEPOINT = 128 };

// ACX#match / match(sep) / match_longest on Bytes, optionally remapped to character
// offsets like the String overloads (ac.cr:280-340, 297-310; matcher.cr:34-46).
// Result in orc_ac_hits(); returns the hit count.
int64_t orc_ac_scan(void* a, const uint8_t* seq, int64_t n, uint32_t mode, const uint8_t* sep_bits, int32_t sep_size) {
  AC* ac = (AC*)a;
  std::vector<Hit>& out = ac->hits;
  out.clear();
  Sep sep{sep_bits, sep_size};
  if (mode & ORC_LONGEST) {
    match_longest_(ac, seq, n, (mode & ORC_INTERSECTABLE) != 0, [&](int32_t idx, int32_t nid) {
      fetch_one(ac, idx, nid, [&](Hit h) { out.push_back(h); });
    });
  } else if ((mode & ORC_SEP) && (mode & ORC_SEP_CODEPOINT)) {
    match_sep_chars_(ac, seq, n, sep, [&](Hit h) { out.push_back(h); });
  } else if (mode & ORC_SEP) {
    match_(ac, seq, n, [&](int32_t idx, int32_t nid) {
      if ((int64_t)idx + 1 < n && sep.rejects(seq[idx + 1])) return;
      fetch(ac, idx, nid, [&](Hit h) {
        if (h.start > 0 && sep.rejects(seq[h.start - 1])) return;
        out.push_back(h);
      });
    });
  } else {
    match_(ac, seq, n, [&](int32_t idx, int32_t nid) { fetch(ac, idx, nid, [&](Hit h) { out.push_back(h); }); });
  }
  if (mode & ORC_UNITS_CHAR) {
    std::vector<int32_t> cob;
    char_map(seq, n, cob);
    for (Hit& h : out) h = Hit{cob[h.start], cob[h.end - 1] + 1, h.value};
  }
  return (int64_t)out.size();
}

// Count-only variant for timing (no result vector growth): returns hits, xors a checksum.
int64_t orc_ac_scan_count(void* a, const uint8_t* seq, int64_t n, uint64_t* checksum) {
  AC* ac = (AC*)a;
  int64_t cnt = 0;
  uint64_t cs = 0;
  match_(ac, seq, n, [&](int32_t idx, int32_t nid) {
    fetch(ac, idx, nid, [&](Hit h) {
      cnt++;
      cs = cs * 0x9E3779B97F4A7C15ull + ((uint64_t)(uint32_t)h.end << 32 | (uint32_t)h.value);
    });
  });
  if (checksum) *checksum = cs;
  return cnt;
}

// Document-parallel timing harness: `threads` workers pull documents from a shared counter and run
// ACX#match over each (same loop as orc_ac_scan_count); the automaton is shared read-only.
int64_t orc_ac_scan_count_mt(void* a, const uint8_t* text, const int64_t* doc_off, int64_t n_docs, int32_t threads) {
  AC* ac = (AC*)a;
  std::atomic<int64_t> next(0), total(0);
  auto work = [&]() {
    int64_t local = 0;
    for (;;) {
      int64_t d = next.fetch_add(1);
      if (d >= n_docs) break;
      match_(ac, text + doc_off[d], doc_off[d + 1] - doc_off[d], [&](int32_t idx, int32_t nid) {
        fetch(ac, idx, nid, [&](Hit) { local++; });
      });
    }
    total += local;
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; t++) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
  return total.load();
}

// Order-sensitive checksum of the hit stream of a batch, document-parallel: sum over documents d and over the j-th hit
// of d of mix(d, j, start, end, value) mod 2^64 (bench.py computes the same on the GPU's stream with torch).
// *hash_out = checksum; returns the hit count.
static inline uint64_t mix_hit(int64_t d, int64_t j, const Hit& h) {
  uint64_t x = (uint64_t)(int64_t)h.start * 0x9E3779B97F4A7C15ull + (uint64_t)(int64_t)h.end * 0xC2B2AE3D27D4EB4Full +
               (uint64_t)(int64_t)h.value * 0x165667B19E3779F9ull + (uint64_t)j * 0x85EBCA77C2B2AE63ull +
               (uint64_t)d * 0x27D4EB2F165667C5ull;
  x = (x ^ (uint64_t)((int64_t)x >> 29)) * 0xBF58476D1CE4E5B9ull;  // arithmetic shift: what torch's int64 >> does
  return x;
}
int64_t orc_ac_scan_hash_mt(void* a, const uint8_t* text, const int64_t* doc_off, int64_t n_docs, int32_t threads, uint64_t* hash_out) {
  AC* ac = (AC*)a;
  std::atomic<int64_t> next(0), total(0);
  std::atomic<uint64_t> hash(0);
  auto work = [&]() {
    int64_t local = 0;
    uint64_t hsum = 0;
    for (;;) {
      int64_t d = next.fetch_add(1);
      if (d >= n_docs) break;
      int64_t j = 0;
      match_(ac, text + doc_off[d], doc_off[d + 1] - doc_off[d], [&](int32_t idx, int32_t nid) {
        fetch(ac, idx, nid, [&](Hit h) { hsum += mix_hit(d, j++, h); });
      });
      local += j;
    }
    total += local;
    hash += hsum;
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; t++) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
  *hash_out = hash.load();
  return total.load();
}

// Any mode of orc_ac_scan over a batch, document-parallel: ACX#match / match(sep) / match_longest (+ intersectable), byte
// or character offsets.  Returns the hit count; *hash_out = the order-sensitive checksum of orc_ac_scan_hash_mt.
int64_t orc_ac_scan_mode_mt(void* a, const uint8_t* text, const int64_t* doc_off, int64_t n_docs, uint32_t mode,
                            const uint8_t* sep_bits, int32_t sep_size, int32_t threads, uint64_t* hash_out) {
  AC* ac = (AC*)a;
  Sep sep{sep_bits, sep_size};
  std::atomic<int64_t> next(0), total(0);
  std::atomic<uint64_t> hash(0);
  auto work = [&]() {
    int64_t local = 0;
    uint64_t hsum = 0;
    std::vector<Hit> out;
    std::vector<int32_t> cob;
    for (;;) {
      const int64_t d = next.fetch_add(1);
      if (d >= n_docs) break;
      const uint8_t* seq = text + doc_off[d];
      const int64_t n = doc_off[d + 1] - doc_off[d];
      out.clear();
      if (mode & ORC_LONGEST) {
        match_longest_(ac, seq, n, (mode & ORC_INTERSECTABLE) != 0, [&](int32_t idx, int32_t nid) {
          fetch_one(ac, idx, nid, [&](Hit h) { out.push_back(h); });
        });
      } else if ((mode & ORC_SEP) && (mode & ORC_SEP_CODEPOINT)) {
        match_sep_chars_(ac, seq, n, sep, [&](Hit h) { out.push_back(h); });
      } else if (mode & ORC_SEP) {
        match_(ac, seq, n, [&](int32_t idx, int32_t nid) {
          if ((int64_t)idx + 1 < n && sep.rejects(seq[idx + 1])) return;
          fetch(ac, idx, nid, [&](Hit h) {
            if (h.start > 0 && sep.rejects(seq[h.start - 1])) return;
            out.push_back(h);
          });
        });
      } else {
        match_(ac, seq, n, [&](int32_t idx, int32_t nid) { fetch(ac, idx, nid, [&](Hit h) { out.push_back(h); }); });
      }
      if (mode & ORC_UNITS_CHAR) {
        char_map(seq, n, cob);
        for (Hit& h : out) h = Hit{cob[h.start], cob[h.end - 1] + 1, h.value};
      }
      int64_t j = 0;
      for (const Hit& h : out) hsum += mix_hit(d, j++, h);
      local += j;
    }
    total += local;
    hash += hsum;
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; t++) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
  if (hash_out) *hash_out = hash.load();
  return total.load();
}

// synthetic corpus text (include/aha_synth.h): the reference arm of bench.py generates its own input
int32_t orc_gen_text(int32_t corpus, uint64_t seed, int64_t first, int64_t bytes, uint8_t* out, int32_t threads) {
  if (corpus < 0 || corpus > 2 || first < 0 || bytes < 0) return -1;
  const int64_t end = first + bytes;
  const int64_t k0 = first / aha::kGenBlock, k1 = (end + aha::kGenBlock - 1) / aha::kGenBlock;
  std::atomic<int64_t> next(k0);
  auto work = [&]() {
    uint8_t buf[aha::kGenBlock];
    for (;;) {
      const int64_t ka = next.fetch_add(1024);
      if (ka >= k1) break;
      for (int64_t k = ka; k < std::min<int64_t>(ka + 1024, k1); ++k) {
        aha::gen_block(corpus, seed, (uint64_t)k, buf);
        const int64_t b0 = k * aha::kGenBlock;
        const int64_t lo = std::max<int64_t>(first, b0), hi = std::min<int64_t>(end, b0 + aha::kGenBlock);
        if (hi > lo) memcpy(out + (lo - first), buf + (lo - b0), (size_t)(hi - lo));
      }
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; t++) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
  return 0;
}

// CedarX#match_longest(String, ignore_case, sep) (cedar.cr:1075-1121); character offsets.
// Result buffer is owned by the cedar-side holder passed as `hits_holder` (an AC or NULL ->
// static thread-local buffer).
static thread_local std::vector<Hit> g_trie_hits;
int64_t orc_cedar_match_longest(void* c, const uint8_t* s, int64_t n, int32_t ignore_case, int32_t use_sep,
                                const uint8_t* sep_bits, int32_t sep_size, int32_t early_exit) {
  Cedar* da = (Cedar*)c;
  std::vector<uint32_t> chars;
  decode_utf8(s, n, chars);
  g_trie_hits.clear();
  Sep sep{sep_bits, sep_size};
  trie_longest(da, chars.data(), (int64_t)chars.size(), ignore_case != 0, use_sep ? &sep : nullptr, early_exit != 0,
               g_trie_hits);
  return (int64_t)g_trie_hits.size();
}
const Hit* orc_trie_hits() { return g_trie_hits.data(); }

// the same over a batch, document-parallel (early-exit form, identical results: Q9): hit count, and in *hash_out the
// order-sensitive checksum of the per-document streams (mix_hit) when hash_out is not null
int64_t orc_cedar_match_longest_mt(void* c, const uint8_t* text, const int64_t* doc_off, int64_t n_docs, int32_t ignore_case,
                                   int32_t use_sep, const uint8_t* sep_bits, int32_t sep_size, int32_t threads, uint64_t* hash_out) {
  Cedar* da = (Cedar*)c;
  std::atomic<int64_t> next(0), total(0);
  std::atomic<uint64_t> hash(0);
  Sep sep{sep_bits, sep_size};
  auto work = [&]() {
    std::vector<uint32_t> chars;
    std::vector<Hit> hits;
    int64_t local = 0;
    uint64_t hsum = 0;
    for (;;) {
      const int64_t d0 = next.fetch_add(16);
      if (d0 >= n_docs) break;
      for (int64_t d = d0; d < std::min<int64_t>(d0 + 16, n_docs); ++d) {
        decode_utf8(text + doc_off[d], doc_off[d + 1] - doc_off[d], chars);
        hits.clear();
        trie_longest(da, chars.data(), (int64_t)chars.size(), ignore_case != 0, use_sep ? &sep : nullptr, true, hits);
        if (hash_out)
          for (size_t j = 0; j < hits.size(); ++j) hsum += mix_hit(d, (int64_t)j, hits[j]);
        local += (int64_t)hits.size();
      }
    }
    total += local;
    hash += hsum;
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; t++) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
  if (hash_out) *hash_out = hash.load();
  return total.load();
}

}  // extern "C"
