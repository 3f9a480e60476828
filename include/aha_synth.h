// aha_synth.h -- counter-based synthetic corpora, identical bytes on host and device.
// Bench / test infrastructure shared by the product library (device generator for corpora that do not fit the host,
// aha_b200_gen_text_*) and by the CPU oracle (orc_gen_text: the reference arm of bench.py generates its own text and
// never loads the product).  Nothing of the scan path depends on it except the AHA_HD macro.
//
// The stream is a sequence of independent 192-byte blocks (192 = 64 * 3, so CJK blocks end on a
// character boundary).  Block k depends only on (corpus, seed, k); integer arithmetic only, so
// the host and device versions agree bit for bit.  No block ever contains a NUL byte (Q6).
//   EN : pseudo-words (2-10 lowercase letters, 1/8 capitalised) over a 50 000-word vocabulary,
//        ranks drawn log-uniformly (a discrete Zipf(s=1)), joined by ' ' ('\n' about every 80th).
//   ZH : 3-byte CJK code points U+4E00 + r, r < 6000 log-uniform; '，' / '。' every 8-24 chars.
//   MIX: paragraphs of 8 blocks alternating EN / ZH.
#pragma once
#include <cstdint>

#ifdef __CUDACC__
#define AHA_HD __host__ __device__ __forceinline__
#else
#define AHA_HD inline
#endif

namespace aha {

constexpr int kGenBlock = 192;
constexpr uint32_t kEnVocab = 50000;
constexpr uint32_t kZhVocab = 6000;

AHA_HD uint64_t splitmix64(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// log-uniform rank in [0, vocab): octave uniform in [0, octaves), position uniform inside it
AHA_HD uint32_t zipf_rank(uint64_t r, uint32_t octaves, uint32_t vocab) {
  uint32_t oct = (uint32_t)(r % octaves);
  uint32_t in = (uint32_t)(r >> 16) & ((1u << oct) - 1u);
  return ((1u << oct) - 1u + in) % vocab;
}

// word of a rank: length 2..10, letters a-z; returns length
AHA_HD int en_word(uint64_t seed, uint32_t rank, uint8_t* out) {
  uint64_t s = seed * 0xD1342543DE82EF95ull + rank;
  uint64_t h = splitmix64(s);
  int len = 2 + (int)(h % 9);
  bool cap = ((h >> 8) & 7) == 0;
  uint64_t g = splitmix64(s);
  for (int i = 0; i < len; ++i) {
    out[i] = (uint8_t)('a' + (g % 26));
    g /= 26;
  }
  if (cap) out[0] = (uint8_t)(out[0] - 32);
  return len;
}

AHA_HD void gen_block_en(uint64_t seed, uint64_t k, uint8_t* out) {
  uint64_t s = seed ^ (k * 0xA24BAED4963EE407ull);
  int pos = 0;
  uint8_t w[12];
  while (pos < kGenBlock) {
    uint64_t r = splitmix64(s);
    int len = en_word(seed, zipf_rank(r, 16, kEnVocab), w);
    if (pos + len + 1 > kGenBlock) {  // does not fit: pad the block with spaces
      while (pos < kGenBlock) out[pos++] = ' ';
      break;
    }
    for (int i = 0; i < len; ++i) out[pos++] = w[i];
    out[pos++] = ((r >> 40) % 80 == 0) ? '\n' : ' ';
  }
}

AHA_HD void gen_block_zh(uint64_t seed, uint64_t k, uint8_t* out) {
  uint64_t s = seed ^ (k * 0x9FB21C651E98DF25ull);
  int until_punct = 8 + (int)(splitmix64(s) % 17);
  for (int pos = 0; pos < kGenBlock; pos += 3) {
    uint64_t r = splitmix64(s);
    uint32_t cp;
    if (--until_punct == 0) {
      cp = (r & 1) ? 0xFF0Cu : 0x3002u;
      until_punct = 8 + (int)((r >> 8) % 17);
    } else {
      cp = 0x4E00u + zipf_rank(r, 13, kZhVocab);
    }
    out[pos] = (uint8_t)(0xE0 | (cp >> 12));
    out[pos + 1] = (uint8_t)(0x80 | ((cp >> 6) & 0x3F));
    out[pos + 2] = (uint8_t)(0x80 | (cp & 0x3F));
  }
}

AHA_HD void gen_block(int corpus, uint64_t seed, uint64_t k, uint8_t* out) {
  bool zh = corpus == 1 || (corpus == 2 && ((k >> 3) & 1));
  if (zh) gen_block_zh(seed, k, out);
  else gen_block_en(seed, k, out);
}

}  // namespace aha
