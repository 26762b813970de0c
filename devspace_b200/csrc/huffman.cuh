// huffman.cuh -- per-block dynamic Huffman tables for deflate (RFC 1951 3.2.5-3.2.7).
//
// Stands where compress/flate's huffmanBitWriter builds its literal/length, distance and code-length
// encoders for every block of the level-6 stream that gzip.NewWriter (tar.go:156) produces
// [Go stdlib, not vendored]. Length-limited code lengths are derived the classic way (Huffman tree by
// two-queue merge over frequency-sorted leaves, overflow repair on the per-length counts as in
// zlib's gen_bitlen), canonical codes are assigned per RFC 1951 3.2.2 and bit-reversed for the
// LSB-first bit writer. All functions are serial __host__ __device__ code working on caller-provided
// arrays (shared memory on the device): a block has at most 286 + 30 + 19 symbols.
#pragma once

#include "dsg_common.cuh"

namespace dsg {

constexpr int kNumLitLen = 286;   // 0..255 literals, 256 end-of-block, 257..285 lengths
constexpr int kNumDist = 30;
constexpr int kNumCodeLen = 19;
constexpr int kMaxBits = 15;
constexpr int kMaxClBits = 7;
constexpr int kEob = 256;

// RFC 1951 3.2.5 tables
DSG_HD uint32_t len_code(uint32_t len /*3..258*/) {
  // code index 0..28 (symbol 257 + idx)
  if (len <= 10) return len - 3;
  if (len == 258) return 28;
  uint32_t l = len - 3;
  uint32_t hb = 31 - (uint32_t)
#if defined(__CUDA_ARCH__)
      __clz(l);
#else
      __builtin_clz(l);
#endif
  // l in [8,255): groups of 4 codes per power of two starting at l=8 (hb=3)
  return (hb - 1) * 4 + ((l >> (hb - 2)) & 3);
}
DSG_HD uint32_t len_extra_bits(uint32_t code /*0..28*/) { return (code < 8 || code == 28) ? 0 : (code - 4) >> 2; }
DSG_HD uint32_t len_base(uint32_t code) {
  if (code < 8) return code + 3;
  if (code == 28) return 258;
  uint32_t eb = (code - 4) >> 2;
  return 3 + ((4 + (code & 3)) << eb);
}
DSG_HD uint32_t dist_code(uint32_t dist /*1..32768*/) {
  if (dist <= 4) return dist - 1;
  uint32_t d = dist - 1;
  uint32_t hb = 31 - (uint32_t)
#if defined(__CUDA_ARCH__)
      __clz(d);
#else
      __builtin_clz(d);
#endif
  return hb * 2 + ((d >> (hb - 1)) & 1);
}
DSG_HD uint32_t dist_extra_bits(uint32_t code /*0..29*/) { return code < 4 ? 0 : (code - 2) >> 1; }
DSG_HD uint32_t dist_base(uint32_t code) {
  if (code < 4) return code + 1;
  uint32_t eb = (code - 2) >> 1;
  return 1 + ((2 + (code & 1)) << eb);
}

DSG_HD uint32_t bit_reverse(uint32_t v, int n) {
  uint32_t r = 0;
  for (int i = 0; i < n; i++) {
    r = (r << 1) | (v & 1);
    v >>= 1;
  }
  return r;
}

// Scratch for huff_build_lengths: 2*n nodes.
struct HuffScratch {
  uint32_t weight[2 * kNumLitLen];
  uint16_t parent[2 * kNumLitLen];
  uint16_t order[kNumLitLen];   // leaves sorted by (freq, symbol)
  uint16_t bl_count[kMaxBits + 2];
};

// Two-queue Huffman merge over m >= 2 leaves whose weights S.weight[0..m) are sorted ascending. Internal nodes
// m..2m-2 are created in non-decreasing weight; S.parent[] receives the tree; returns the root index.
// The heads of both queues are kept in registers so that an iteration costs one or two shared-memory loads.
DSG_HD int huff_merge_sorted(HuffScratch &S, int m) {
  const uint32_t kInf = 0xffffffffu;
  int li = 0, ii = m, next = m;
  uint32_t wl = S.weight[0], wi = kInf;
  for (int k = 0; k < m - 1; k++) {
    int a, b;
    uint32_t wa, wb;
    if (wl <= wi) {  // ties prefer the leaf, as zlib's heap does for equal frequency / lower depth
      a = li++;
      wa = wl;
      wl = li < m ? S.weight[li] : kInf;
    } else {
      a = ii++;
      wa = wi;
      wi = ii < next ? S.weight[ii] : kInf;
    }
    if (wl <= wi) {
      b = li++;
      wb = wl;
      wl = li < m ? S.weight[li] : kInf;
    } else {
      b = ii++;
      wb = wi;
      wi = ii < next ? S.weight[ii] : kInf;
    }
    const uint32_t w = wa + wb;
    S.weight[next] = w;
    S.parent[a] = (uint16_t)next;
    S.parent[b] = (uint16_t)next;
    if (ii == next) wi = w;  // the internal queue was empty: the new node is its head
    next++;
  }
  return next - 1;
}

// Code lengths from the tree in S.parent[] (leaves 0..m-1 in ascending weight order = S.order[]), limited to
// max_bits the way zlib's gen_bitlen does it: depths are propagated top-down, a node deeper than max_bits is
// clamped and counted as an overflow (internal nodes included), then leaves are moved down level by level
// until the Kraft sum is exactly one again, and the longest codes go to the least frequent symbols.
DSG_HD void huff_lengths_from_tree(HuffScratch &S, int m, int root, int max_bits, uint8_t *lens) {
  int overflow = 0;
  S.weight[root] = 0;  // weight[] is reused as depth
  for (int i = root - 1; i >= m; i--) {
    int d = (int)S.weight[S.parent[i]] + 1;
    if (d > max_bits) {
      d = max_bits;
      overflow++;
    }
    S.weight[i] = (uint32_t)d;
  }
  for (int b = 0; b <= max_bits + 1; b++) S.bl_count[b] = 0;
  for (int i = 0; i < m; i++) {
    int d = (int)S.weight[S.parent[i]] + 1;
    if (d > max_bits) {
      d = max_bits;
      overflow++;
    }
    S.weight[i] = (uint32_t)d;
    S.bl_count[d]++;
  }
  if (overflow > 0) {
    do {
      int bits = max_bits - 1;
      while (S.bl_count[bits] == 0) bits--;
      S.bl_count[bits]--;
      S.bl_count[bits + 1] += 2;
      S.bl_count[max_bits]--;
      overflow -= 2;
    } while (overflow > 0);
    int i = 0;
    for (int bits = max_bits; bits >= 1; bits--)
      for (int c = S.bl_count[bits]; c > 0; c--) S.weight[i++] = (uint32_t)bits;
  }
  for (int i = 0; i < m; i++) lens[S.order[i]] = (uint8_t)S.weight[i];
}

// Code lengths (0 = unused symbol) for n symbols with the given frequencies, limited to max_bits.
// At least two codes get a non-zero length so that decoders never see an incomplete one-symbol tree
// (same convention as zlib's build_tree). Returns the highest used symbol index, or -1 when none.
DSG_HD int huff_build_lengths(const uint32_t *freq, int n, int max_bits, uint8_t *lens, HuffScratch &S) {
  int m = 0;
  int max_sym = -1;
  for (int i = 0; i < n; i++) {
    lens[i] = 0;
    if (freq[i]) {
      S.order[m++] = (uint16_t)i;
      max_sym = i;
    }
  }
  if (m == 0) {  // force two symbols: 0 and 1
    lens[0] = 1;
    lens[1] = 1;
    return 1;
  }
  if (m == 1) {
    int s = S.order[0];
    int other = s == 0 ? 1 : 0;
    lens[s] = 1;
    lens[other] = 1;
    return s > other ? s : other;
  }
  // insertion sort by (freq, symbol): m <= 286, frequencies mostly small; stable and branch-light
  for (int i = 1; i < m; i++) {
    uint16_t s = S.order[i];
    uint32_t f = freq[s];
    int j = i - 1;
    while (j >= 0 && freq[S.order[j]] > f) {
      S.order[j + 1] = S.order[j];
      j--;
    }
    S.order[j + 1] = s;
  }
  for (int i = 0; i < m; i++) S.weight[i] = freq[S.order[i]];
  const int root = huff_merge_sorted(S, m);
  huff_lengths_from_tree(S, m, root, max_bits, lens);
  return max_sym;
}

// Canonical codes (RFC 1951 3.2.2), bit-reversed so that they can be OR-ed LSB first.
DSG_HD void huff_make_codes(const uint8_t *lens, int n, uint16_t *codes) {
  uint16_t bl_count[kMaxBits + 1];
  uint16_t next_code[kMaxBits + 2];
  for (int b = 0; b <= kMaxBits; b++) bl_count[b] = 0;
  for (int i = 0; i < n; i++) bl_count[lens[i]]++;
  bl_count[0] = 0;
  uint32_t code = 0;
  for (int b = 1; b <= kMaxBits; b++) {
    code = (code + bl_count[b - 1]) << 1;
    next_code[b] = (uint16_t)code;
  }
  for (int i = 0; i < n; i++) {
    int l = lens[i];
    codes[i] = l ? (uint16_t)bit_reverse(next_code[l]++, l) : 0;
  }
}

// Fixed Huffman code lengths (RFC 1951 3.2.6)
DSG_HD uint32_t fixed_litlen_bits(uint32_t sym) { return sym < 144 ? 8 : (sym < 256 ? 9 : (sym < 280 ? 7 : 8)); }

// Order in which code-length-code lengths are transmitted
DSG_HD int cl_order(int i) {
  const uint8_t o[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
  return o[i];
}

// Run-length encoding of the concatenated literal/length + distance code lengths (RFC 1951 3.2.7):
// symbols 0..15 literal lengths, 16 = repeat previous 3..6 (2 extra bits), 17 = zeros 3..10 (3 bits),
// 18 = zeros 11..138 (7 bits). Emits (symbol | extra << 8) entries; returns how many. cl_freq[19] is
// incremented. Same greedy rule as zlib's scan_tree/send_tree.
DSG_HD int rle_code_lengths(const uint8_t *lens, int n, uint16_t *out, uint32_t *cl_freq) {
  int w = 0;
  int i = 0;
  while (i < n) {
    uint8_t v = lens[i];
    int run = 1;
    while (i + run < n && lens[i + run] == v) run++;
    i += run;
    if (v == 0) {
      while (run >= 11) {
        int r = run > 138 ? 138 : run;
        out[w++] = (uint16_t)(18 | ((r - 11) << 8));
        cl_freq[18]++;
        run -= r;
      }
      if (run >= 3) {
        out[w++] = (uint16_t)(17 | ((run - 3) << 8));
        cl_freq[17]++;
        run = 0;
      }
      while (run-- > 0) {
        out[w++] = 0;
        cl_freq[0]++;
      }
    } else {
      out[w++] = v;
      cl_freq[v]++;
      run--;
      while (run >= 3) {
        int r = run > 6 ? 6 : run;
        out[w++] = (uint16_t)(16 | ((r - 3) << 8));
        cl_freq[16]++;
        run -= r;
      }
      while (run-- > 0) {
        out[w++] = v;
        cl_freq[v]++;
      }
    }
  }
  return w;
}

DSG_HD uint32_t cl_extra_bits(uint32_t sym) { return sym == 16 ? 2 : (sym == 17 ? 3 : (sym == 18 ? 7 : 0)); }

}  // namespace dsg
