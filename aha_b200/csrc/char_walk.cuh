// char_walk.cuh -- one lane's walk over UTF-8 units with the character automaton of char_table.hpp.
// Compiled for the device (tables in shared memory) and for the host (the CPU self-check entry
// aha_b200_debug_char_scan, used by the no-GPU tests to validate the derived tables).
#pragma once
#include <cstdint>

#include "../../include/aha_synth.h"  // AHA_HD

namespace aha {

struct CharTablesView {
  const uint16_t* root;       // [n_root] one-character state:15 | fires:1; the index of a character is also its column
  const uint16_t* check;      // [m_pad]
  const uint32_t* val;        // [m]
  const uint16_t* gfail;      // [m]   (global memory on the device)
  uint32_t lo[3], n[3], off[3];  // root index ranges for 2-, 3-, 4-byte units
  uint32_t m;
};

enum CharWalkStatus { kWalkOk = 0, kWalkBadUtf8 = 1 };

// Walks buf[w .. ) of one document segment.  All positions are offsets into `buf`.
//   w        first byte to read (document start, or a few bytes before `a` for the warm-up)
//   a, b     owned region: units whose LAST byte lies in [a, b) are reported
//   seg_end  min(b, end of the document): no unit may start at or beyond it
//   doc_end  end of the document: a unit reaching past it is malformed
//   at_doc_start  w is the first byte of the document (otherwise leading continuation bytes are skipped)
// emit(ref, last, k): ref = firing edge slot, or m + cid for a one-character state;
// k = number of units that START in [a, last] (character offsets).
// Returns kWalkBadUtf8 when the text is not
// structurally valid UTF-8 inside the walked range (the caller then rescans with the byte automaton).
template <class Emit>
AHA_HD int char_walk(const CharTablesView& T, const uint8_t* buf, int32_t w, int32_t a, int32_t b, int32_t seg_end,
                     int32_t doc_end, bool at_doc_start, Emit emit) {
  int32_t cs = w;
  if (!at_doc_start) {
    int skipped = 0;
    while (cs < seg_end && (buf[cs] & 0xC0) == 0x80) {
      ++cs;
      if (++skipped > 3) return kWalkBadUtf8;
    }
  }
  uint32_t s = 0, f1 = 0;
  uint32_t k = 0;
  while (cs < seg_end) {
    const uint32_t b0 = buf[cs];
    int32_t len;
    uint32_t idx;
    if (b0 < 0x80) {
      len = 1;
      idx = b0;
    } else {
      uint32_t code;
      if (b0 < 0xC0 || b0 >= 0xF8) return kWalkBadUtf8;
      len = b0 < 0xE0 ? 2 : (b0 < 0xF0 ? 3 : 4);
      if (cs + len > doc_end) return kWalkBadUtf8;
      const uint32_t b1 = buf[cs + 1];
      if ((b1 & 0xC0) != 0x80) return kWalkBadUtf8;
      if (len == 2) {
        code = ((b0 & 0x1Fu) << 6) | (b1 & 0x3Fu);
      } else {
        const uint32_t b2 = buf[cs + 2];
        if ((b2 & 0xC0) != 0x80) return kWalkBadUtf8;
        if (len == 3) {
          code = ((b0 & 0x0Fu) << 12) | ((b1 & 0x3Fu) << 6) | (b2 & 0x3Fu);
        } else {
          const uint32_t b3 = buf[cs + 3];
          if ((b3 & 0xC0) != 0x80) return kWalkBadUtf8;
          code = ((b0 & 0x07u) << 18) | ((b1 & 0x3Fu) << 12) | ((b2 & 0x3Fu) << 6) | (b3 & 0x3Fu);
        }
      }
      const uint32_t rel = code - T.lo[len - 2];
      idx = T.off[len - 2] + (rel < T.n[len - 2] ? rel : T.n[len - 2]);  // out of range: the zero entry behind the range
    }
    const int32_t last = cs + len - 1;
    if (last >= b) break;  // the unit ends in the next lane's region
    if (cs >= a) ++k;
    cs += len;
    const uint32_t r = T.root[idx];
    const uint32_t col = idx;  // no state has an edge in the column of a character outside the dictionary: every probe misses
    uint32_t t = s;
    bool first = true;
    for (;;) {
      if (t == 0) {  // the root: direct table entry of this character
        if ((r & 0x8000u) && last >= a) emit(T.m + col, last, k);
        s = r & 0x7FFFu;
        f1 = 0;
        break;
      }
      const uint32_t slot = t + col;
      if (slot < T.m && T.check[slot] == t) {
        const uint32_t v = T.val[slot];
        if ((v & 0x8000u) && last >= a) emit(slot, last, k);
        s = v & 0x7FFFu;
        f1 = (v >> 16) & 0x7FFFu;
        break;
      }
      t = first ? f1 : (uint32_t)T.gfail[t];  // effective failure link (0 for one-character states)
      first = false;
    }
  }
  return kWalkOk;
}

}  // namespace aha
