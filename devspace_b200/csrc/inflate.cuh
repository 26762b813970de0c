// inflate.cuh -- chunk-parallel gunzip of the streams THIS LIBRARY writes.
//
// Stands where untarAll reads its archive, /root/reference/pkg/devspace/sync/tar.go:16-33 (gzip.NewReader(reader) ->
// tar.NewReader): the decompression half of the downstream direction. A foreign gzip stream (GNU `tar -czf` on the
// remote side, downstream.go:394) is one chain of bits and cannot be inflated in parallel; a stream written by
// deflate_lz.cuh / deflate_stored.cuh can, because of how it is built:
//   * every chunk (16 .. 256 KiB of input) is ONE fixed/dynamic block that starts on a byte boundary and is followed
//     by an empty stored block (000, pad, 00 00 ff ff), or a few stored blocks (00 LEN NLEN data);
//   * the body ends with an empty final fixed block (03 00), or, at level 0, with a final stored block.
// So the byte offsets where a "unit" (a compressed chunk, a stored block, the final block) may start can be found
// without decoding -- behind every 00 00 ff ff, behind every plausible stored-block header, behind every gzip header
// -- and every candidate can be decoded on its own thread. Candidates that only look like unit starts (the patterns
// occur by chance once per ~4 GB of compressed data) decode to garbage or fail; the serial walk over the unit chain,
// which needs only each unit's END offset, never visits them.
//
// Kernels (one WARP per unit: a unit is a serial bit stream, so all 32 lanes decode it in lockstep -- the same loads,
// the same branches, no divergence -- and share only the work that is parallel, the copies of a match; the
// parallelism is the number of units, thousands to tens of thousands per GB):
//   gz_scan_count / gz_scan_write   candidate unit starts, sorted by offset
//   gz_measure_kernel               dry decode of every candidate: kind, output bytes, end offset
//   gz_chain_kernel                 serial walk from the first member header: the true units, their output offsets,
//                                   member trailers (also callable on the host: gz_chain_walk)
//   gz_decode_kernel                decode of the true units. A chunk may copy from the 32 KiB of output BEFORE it
//                                   (the compressor's dictionary), which another thread is producing at the same time:
//                                   the decode writes 16-bit symbols -- a byte, or 256 + index into that window --
//                                   and copies symbols, so that window references propagate through copies of copies;
//   gz_tails_plain / gz_tails_round the last 32 KiB of every unit: plain bytes at once, window references in parallel
//                                   rounds (a reference whose source in the tails before it is a byte by now);
//   gz_tails_kernel                 one CTA, units in order: whatever the rounds left open (long chains of copies);
//   gz_bodies_kernel                all units in parallel: resolves the rest against the finished tails;
//   gz_stored_copy_kernel           stored blocks straight from the input.
// CRC32 and ISIZE of every member are checked by the caller with the CRC kernels of crc32.cuh.
// Bound: thread-serial Huffman decoding (instruction issue), not HBM; algorithmic bytes: read C, write T (+ 2 x 2T for
// the symbol scratch).
#pragma once

#include "copy_util.cuh"
#include "huffman.cuh"

namespace dsg {

constexpr uint32_t kGzWindow = 32768;
constexpr uint32_t kGzMaxUnitOut = 4u * 65536u;          // kMaxChunk of the compressor
constexpr uint32_t kGzMaxUnitIn = kGzMaxUnitOut + 8192;  // a chunk never exceeds its stored encoding
constexpr int kGzScanSpan = 64;                          // positions per thread in the candidate scan
constexpr int kGzScanThreads = 256;
constexpr int kGzScanCta = kGzScanSpan * kGzScanThreads;  // 16384 positions per CTA

enum GzKind : uint32_t { kGzInvalid = 0, kGzCompressed = 1, kGzStored = 2, kGzStoredFinal = 3, kGzFinal = 4 };

struct GzRec {          // what the dry decode found at a candidate offset
  uint64_t end;         // offset behind the unit (behind the sync marker of a compressed chunk)
  uint32_t out_len;
  uint32_t kind;
};

struct GzUnit {         // a unit of the true chain
  uint64_t in_off;      // compressed chunk: its first byte; stored block: first payload byte
  uint64_t out_off;     // offset in the decompressed stream
  uint32_t out_len;
  uint32_t kind;
  uint32_t win_avail;   // decompressed bytes of the same member before the unit (<= 32768)
  uint32_t member;
};

struct GzMember {
  uint64_t out_off, out_len;
  uint32_t crc, isize;
};

struct GzSummary {      // written by the chain walk
  uint64_t total_out;
  uint32_t nunits, nmembers;
  int32_t status;       // DSG_OK, DSG_E_ARG (not a stream of this library), DSG_E_CAPACITY (unit / member table full)
  uint32_t pad;
  uint64_t bad_offset;
};

DSG_HD bool gz_is_header(const uint8_t *p) {
  return p[0] == 0x1f && p[1] == 0x8b && p[2] == 0x08 && p[3] == 0 && p[4] == 0 && p[5] == 0 && p[6] == 0 && p[7] == 0 &&
         p[8] == 0 && p[9] == 0xff;
}

// Candidate unit starts implied by the bytes at position p: behind a sync marker, behind a stored block whose header
// starts at p, behind a gzip header. Up to two (a marker can also read as a stored header). Returns how many.
DSG_HD int gz_candidates_at(const uint8_t *in, uint64_t n, uint64_t p, uint64_t *c) {
  int k = 0;
  if (p + 4 <= n && in[p] == 0 && in[p + 1] == 0 && in[p + 2] == 0xff && in[p + 3] == 0xff && p + 4 < n) c[k++] = p + 4;
  if (p + 5 <= n && (in[p] & 0xfe) == 0) {
    const uint32_t len = in[p + 1] | ((uint32_t)in[p + 2] << 8), nlen = in[p + 3] | ((uint32_t)in[p + 4] << 8);
    if ((len ^ nlen) == 0xffffu && len != 0 && p + 5 + len < n) c[k++] = p + 5 + len;
  }
  if (k < 2 && p + 10 < n && in[p] == 0x1f && gz_is_header(in + p)) c[k++] = p + 10;
  return k;
}

__global__ void __launch_bounds__(kGzScanThreads)
gz_scan_count(const uint8_t *__restrict__ in, uint64_t n, uint32_t *__restrict__ cta_counts) {
  __shared__ uint32_t total;
  if (threadIdx.x == 0) total = 0;
  __syncthreads();
  const uint64_t p0 = (uint64_t)blockIdx.x * kGzScanCta + (uint64_t)threadIdx.x * kGzScanSpan;
  uint32_t c = 0;
  uint64_t tmp[2];
  for (int k = 0; k < kGzScanSpan; k++) {
    const uint64_t p = p0 + k;
    if (p >= n) break;
    c += (uint32_t)gz_candidates_at(in, n, p, tmp);
  }
  if (c) atomicAdd(&total, c);
  __syncthreads();
  if (threadIdx.x == 0) cta_counts[blockIdx.x] = total;
}

// cta_offs: exclusive prefix of cta_counts. The list is NOT sorted by offset (a stored header points far ahead): the
// host sorts the (short) list afterwards.
__global__ void __launch_bounds__(kGzScanThreads)
gz_scan_write(const uint8_t *__restrict__ in, uint64_t n, const uint32_t *__restrict__ cta_offs,
              uint64_t *__restrict__ cand) {
  __shared__ uint32_t warp_sum[kGzScanThreads / 32];
  const uint64_t p0 = (uint64_t)blockIdx.x * kGzScanCta + (uint64_t)threadIdx.x * kGzScanSpan;
  uint32_t c = 0;
  uint64_t tmp[2];
  for (int k = 0; k < kGzScanSpan; k++) {
    const uint64_t p = p0 + k;
    if (p >= n) break;
    c += (uint32_t)gz_candidates_at(in, n, p, tmp);
  }
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t x = c;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (uint32_t)o) x += y;
  }
  if (lane == 31) warp_sum[warp] = x;
  __syncthreads();
  uint32_t before = 0;
  for (uint32_t w = 0; w < warp; w++) before += warp_sum[w];
  uint32_t at = cta_offs[blockIdx.x] + before + (x - c);
  for (int k = 0; k < kGzScanSpan && c; k++) {
    const uint64_t p = p0 + k;
    if (p >= n) break;
    const int m = gz_candidates_at(in, n, p, tmp);
    for (int j = 0; j < m; j++) cand[at++] = tmp[j];
  }
}

// ---- the decoder of one unit ------------------------------------------------------------------------------------

struct GzBits {
  const uint8_t *in;
  uint64_t pos, end;  // next byte to load, end of the input
  uint64_t buf;
  uint32_t cnt;
  bool bad;
  DSG_D void refill() {
    while (cnt <= 56 && pos < end) {
      buf |= (uint64_t)in[pos++] << cnt;
      cnt += 8;
    }
  }
  DSG_D uint32_t bits(uint32_t n) {  // n <= 16
    if (cnt < n) {
      refill();
      if (cnt < n) {
        bad = true;
        return 0;
      }
    }
    const uint32_t v = (uint32_t)buf & ((1u << n) - 1u);
    buf >>= n;
    cnt -= n;
    return v;
  }
  DSG_D uint64_t byte_pos() const { return pos - cnt / 8; }  // first byte that has not been consumed entirely
};

// Canonical Huffman decoding table in its smallest form: codes per length and the symbols in code order. A symbol
// is decoded one bit at a time (deflate sends codes most significant bit first).
template <int kSyms>
struct GzTable {
  uint16_t count[16];
  uint16_t sym[kSyms];
  // false when the lengths over-subscribe the code space
  DSG_D bool build(const uint8_t *lens, int n) {
    for (int l = 0; l < 16; l++) count[l] = 0;
    for (int i = 0; i < n; i++) count[lens[i]]++;
    count[0] = 0;
    int left = 1;
    for (int l = 1; l < 16; l++) {
      left <<= 1;
      left -= count[l];
      if (left < 0) return false;
    }
    uint16_t offs[16];
    offs[1] = 0;
    for (int l = 1; l < 15; l++) offs[l + 1] = (uint16_t)(offs[l] + count[l]);
    for (int i = 0; i < n; i++)
      if (lens[i]) sym[offs[lens[i]]++] = (uint16_t)i;
    return true;
  }
  DSG_D int decode(GzBits &b) const {
    int code = 0, first = 0, index = 0;
    for (int l = 1; l < 16; l++) {
      code |= (int)b.bits(1);
      const int c = count[l];
      if (code - c < first) return sym[index + (code - first)];
      index += c;
      first += c;
      first <<= 1;
      code <<= 1;
    }
    return -1;
  }
};

// Direct lookup table of one code (per warp, shared memory): the next kBits bits of the stream index an entry
// symbol | length << 12; 0 = the code is longer than kBits (rare: the bit-by-bit walk takes over). Built by the 32
// lanes together from the canonical form. The bit-by-bit walk alone cost ~900 cycles per symbol.
constexpr int kGzLitBits = 10, kGzDistBits = 9;

template <int kSyms>
DSG_D void gz_fast_table(const GzTable<kSyms> &t, const uint8_t *lens, uint16_t *tab, int bits, uint32_t lane) {
  uint32_t first[16], offs[16];
  uint32_t code = 0, at = 0;
  first[0] = offs[0] = 0;
  for (int l = 1; l < 16; l++) {
    code = (code + (l > 1 ? t.count[l - 1] : 0u)) << 1;
    first[l] = code;
    offs[l] = at;
    at += t.count[l];
  }
  for (uint32_t k = lane; k < (1u << bits); k += 32) tab[k] = 0;
  __syncwarp();
  for (uint32_t idx = lane; idx < at; idx += 32) {
    const uint32_t sy = t.sym[idx], l = lens[sy];
    if (l <= (uint32_t)bits) {
      const uint32_t c = first[l] + (idx - offs[l]);
      const uint32_t rev = __brev(c) >> (32 - l);
      for (uint32_t k = rev; k < (1u << bits); k += 1u << l) tab[k] = (uint16_t)(sy | (l << 12));
    }
  }
  __syncwarp();
}

template <int kSyms>
DSG_D int gz_decode_sym(GzBits &b, const uint16_t *tab, int bits, const GzTable<kSyms> &slow) {
  if (b.cnt < 32) b.refill();
  const uint32_t e = tab[(uint32_t)b.buf & ((1u << bits) - 1u)];
  const uint32_t l = e >> 12;
  if (e != 0 && l <= b.cnt) {
    b.buf >>= l;
    b.cnt -= l;
    return (int)(e & 0xfffu);
  }
  return slow.decode(b);
}

// Decodes the unit that starts at byte `start`; called by all 32 lanes of a warp with the same arguments (every lane
// runs the whole bit-serial decode, so control flow is uniform; measured against one thread per unit, where the copy
// loop of the longest match in the warp stalled 31 other decoders on a dependent global load per byte: 2 GB/s).
// kWrite: 16-bit symbols to out16[0 .. out_len) (a byte, or 256 + index into the 32 KiB window before the unit): lane
// 0 writes the literals, a match is copied by all lanes (element j of a match comes from j mod dist of its source,
// which lies entirely before the match: no dependency inside a copy). win_avail: how far back a copy may reach
// before the unit's first byte. Returns the kind; *out_len and *end are valid unless the kind is kGzInvalid.
template <bool kWrite>
DSG_D uint32_t gz_decode_unit(const uint8_t *in, uint64_t n, uint64_t start, uint32_t win_avail, uint16_t *out16,
                              uint32_t *out_len, uint64_t *end, uint32_t lane, uint16_t *s_lt, uint16_t *s_dt) {
  *out_len = 0;
  *end = start;
  if (start >= n) return kGzInvalid;
  const uint32_t b0 = in[start];
  if ((b0 & 0xfe) == 0) {  // stored block (final at level 0)
    if (start + 5 > n) return kGzInvalid;
    const uint32_t len = in[start + 1] | ((uint32_t)in[start + 2] << 8), nlen = in[start + 3] | ((uint32_t)in[start + 4] << 8);
    if ((len ^ nlen) != 0xffffu || start + 5 + len > n) return kGzInvalid;
    *out_len = len;
    *end = start + 5 + len;
    return b0 ? kGzStoredFinal : kGzStored;
  }
  if (b0 == 0x03) {  // empty final fixed block
    if (start + 2 > n || in[start + 1] != 0) return kGzInvalid;
    *end = start + 2;
    return kGzFinal;
  }
  const uint32_t btype = (b0 >> 1) & 3u;
  if ((b0 & 1u) || (btype != 1 && btype != 2)) return kGzInvalid;
  GzBits b{in, start, n < start + kGzMaxUnitIn ? n : start + kGzMaxUnitIn, 0, 0, false};
  b.bits(3);
  GzTable<288> lt;
  GzTable<32> dt;
  uint8_t lens[320];
  if (btype == 1) {
    for (int i = 0; i < 288; i++) lens[i] = (uint8_t)fixed_litlen_bits((uint32_t)i);
    lt.build(lens, 288);
    gz_fast_table(lt, lens, s_lt, kGzLitBits, lane);
    for (int i = 0; i < 30; i++) lens[i] = 5;
    dt.build(lens, 30);
    gz_fast_table(dt, lens, s_dt, kGzDistBits, lane);
  } else {
    const uint32_t hlit = b.bits(5) + 257, hdist = b.bits(5) + 1, hclen = b.bits(4) + 4;
    if (b.bad || hlit > 286 || hdist > 30) return kGzInvalid;
    uint8_t cl[19];
    for (int i = 0; i < 19; i++) cl[i] = 0;
    for (uint32_t i = 0; i < hclen; i++) cl[cl_order((int)i)] = (uint8_t)b.bits(3);
    GzTable<32> ct;
    if (b.bad || !ct.build(cl, 19)) return kGzInvalid;
    uint32_t i = 0;
    while (i < hlit + hdist) {
      const int s = ct.decode(b);
      if (s < 0 || b.bad) return kGzInvalid;
      if (s < 16) {
        lens[i++] = (uint8_t)s;
      } else {
        uint32_t rep, v = 0;
        if (s == 16) {
          if (i == 0) return kGzInvalid;
          v = lens[i - 1];
          rep = 3 + b.bits(2);
        } else if (s == 17) {
          rep = 3 + b.bits(3);
        } else {
          rep = 11 + b.bits(7);
        }
        if (i + rep > hlit + hdist) return kGzInvalid;
        while (rep--) lens[i++] = (uint8_t)v;
      }
    }
    if (b.bad || lens[256] == 0) return kGzInvalid;
    if (!lt.build(lens, (int)hlit) || !dt.build(lens + hlit, (int)hdist)) return kGzInvalid;
    gz_fast_table(lt, lens, s_lt, kGzLitBits, lane);
    gz_fast_table(dt, lens + hlit, s_dt, kGzDistBits, lane);
  }
  uint32_t q = 0;
  for (;;) {
    const int s = gz_decode_sym(b, s_lt, kGzLitBits, lt);
    if (s < 0 || b.bad) return kGzInvalid;
    if (s < 256) {
      if (q >= kGzMaxUnitOut) return kGzInvalid;
      if (kWrite && lane == 0) out16[q] = (uint16_t)s;
      q++;
    } else if (s == 256) {
      break;
    } else {
      const uint32_t lc = (uint32_t)s - 257;
      if (lc > 28) return kGzInvalid;
      const uint32_t len = len_base(lc) + b.bits(len_extra_bits(lc));
      const int ds = gz_decode_sym(b, s_dt, kGzDistBits, dt);
      if (ds < 0 || ds > 29 || b.bad) return kGzInvalid;
      const uint32_t dist = dist_base((uint32_t)ds) + b.bits(dist_extra_bits((uint32_t)ds));
      if (b.bad || dist > q + win_avail || q + len > kGzMaxUnitOut) return kGzInvalid;
      if (kWrite) {
        __syncwarp();  // the literals and copies before this match are visible to every lane
        for (uint32_t j = lane; j < len; j += 32) {
          const int32_t src = (int32_t)(q + (dist >= len ? j : j % dist)) - (int32_t)dist;
          out16[q + j] = src >= 0 ? out16[src] : (uint16_t)(256 + ((int32_t)kGzWindow + src));
        }
        __syncwarp();
      }
      q += len;
    }
  }
  // the chunk's closing sync marker: an empty stored block
  if (b.bits(3) != 0 || b.bad) return kGzInvalid;
  // after the three header bits the writer pads to a byte: the unconsumed bits of the current byte must be zero
  const uint32_t partial = b.cnt & 7;
  if (partial && b.bits(partial) != 0) return kGzInvalid;
  const uint64_t at = b.byte_pos();
  if (at + 4 > n || in[at] != 0 || in[at + 1] != 0 || in[at + 2] != 0xff || in[at + 3] != 0xff) return kGzInvalid;
  *out_len = q;
  *end = at + 4;
  return kGzCompressed;
}

constexpr int kGzWarpsPerCta = 4;

__global__ void __launch_bounds__(32 * kGzWarpsPerCta)
gz_measure_kernel(const uint8_t *__restrict__ in, uint64_t n, const uint64_t *__restrict__ cand, uint32_t ncand,
                  GzRec *__restrict__ rec) {
  __shared__ uint16_t s_tab[kGzWarpsPerCta][(1 << kGzLitBits) + (1 << kGzDistBits)];
  const uint32_t i = blockIdx.x * kGzWarpsPerCta + (threadIdx.x >> 5);  // one warp per candidate
  if (i >= ncand) return;
  uint16_t *tab = s_tab[threadIdx.x >> 5];
  GzRec r;
  r.kind = gz_decode_unit<false>(in, n, cand[i], kGzWindow, nullptr, &r.out_len, &r.end, threadIdx.x & 31, tab,
                                 tab + (1 << kGzLitBits));
  if ((threadIdx.x & 31) == 0) rec[i] = r;
}

// ---- the chain -----------------------------------------------------------------------------------------------------
// Serial walk from offset 0 (a gzip header) over the true units. cand[] sorted ascending (duplicates allowed), rec[]
// aligned with it. Host and device.
DSG_HD void gz_chain_walk(const uint8_t *in, uint64_t n, const uint64_t *cand, const GzRec *rec, uint32_t ncand,
                          GzUnit *units, uint32_t max_units, GzMember *members, uint32_t max_members, GzSummary *S) {
  S->total_out = 0;
  S->nunits = 0;
  S->nmembers = 0;
  S->status = DSG_OK;
  S->bad_offset = 0;
  uint64_t p = 0, out = 0;
  uint32_t at = 0;  // the chain only moves forward: the candidate list is walked, not searched
  while (p < n) {
    if (p + 18 > n || !gz_is_header(in + p)) {
      S->status = DSG_E_ARG;
      S->bad_offset = p;
      return;
    }
    p += 10;
    const uint64_t member_out = out;
    bool open = true;
    while (open) {
      // the record of the candidate at offset p
      while (at < ncand && cand[at] < p) at++;
      const uint32_t lo = at;
      if (lo >= ncand || cand[lo] != p || rec[lo].kind == kGzInvalid) {
        S->status = DSG_E_ARG;
        S->bad_offset = p;
        return;
      }
      const GzRec r = rec[lo];
      if (r.kind == kGzFinal) {
        open = false;
      } else {
        if (S->nunits >= max_units) {
          S->status = DSG_E_CAPACITY;
          return;
        }
        GzUnit u;
        u.in_off = r.kind == kGzCompressed ? p : p + 5;
        u.out_off = out;
        u.out_len = r.out_len;
        u.kind = r.kind;
        u.win_avail = (uint32_t)(out - member_out < kGzWindow ? out - member_out : kGzWindow);
        u.member = S->nmembers;
        if (r.out_len) units[S->nunits++] = u;
        out += r.out_len;
        if (r.kind == kGzStoredFinal) open = false;
      }
      p = r.end;
    }
    if (p + 8 > n || S->nmembers >= max_members) {
      S->status = p + 8 > n ? DSG_E_ARG : DSG_E_CAPACITY;
      S->bad_offset = p;
      return;
    }
    GzMember m;
    m.out_off = member_out;
    m.out_len = out - member_out;
    m.crc = in[p] | ((uint32_t)in[p + 1] << 8) | ((uint32_t)in[p + 2] << 16) | ((uint32_t)in[p + 3] << 24);
    m.isize = in[p + 4] | ((uint32_t)in[p + 5] << 8) | ((uint32_t)in[p + 6] << 16) | ((uint32_t)in[p + 7] << 24);
    if (m.isize != (uint32_t)m.out_len) {
      S->status = DSG_E_ARG;
      S->bad_offset = p;
      return;
    }
    members[S->nmembers++] = m;
    p += 8;
  }
  S->total_out = out;
}

__global__ void gz_chain_kernel(const uint8_t *in, uint64_t n, const uint64_t *cand, const GzRec *rec, uint32_t ncand,
                                GzUnit *units, uint32_t max_units, GzMember *members, uint32_t max_members,
                                GzSummary *S) {
  if (blockIdx.x == 0 && threadIdx.x == 0) gz_chain_walk(in, n, cand, rec, ncand, units, max_units, members, max_members, S);
}

// ---- decode, resolve ---------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(32 * kGzWarpsPerCta)
gz_decode_kernel(const uint8_t *__restrict__ in, uint64_t n, const GzUnit *__restrict__ units, uint32_t nunits,
                 uint16_t *__restrict__ out16, uint64_t sym_base, uint32_t *__restrict__ error) {
  __shared__ uint16_t s_tab[kGzWarpsPerCta][(1 << kGzLitBits) + (1 << kGzDistBits)];
  const uint32_t i = blockIdx.x * kGzWarpsPerCta + (threadIdx.x >> 5);  // one warp per unit
  if (i >= nunits) return;
  const GzUnit u = units[i];
  if (u.kind != kGzCompressed) return;
  uint16_t *tab = s_tab[threadIdx.x >> 5];
  uint32_t len;
  uint64_t end;
  const uint32_t k = gz_decode_unit<true>(in, n, u.in_off, u.win_avail, out16 + (u.out_off - sym_base), &len, &end, threadIdx.x & 31,
                                          tab, tab + (1 << kGzLitBits));
  if ((k != kGzCompressed || len != u.out_len) && (threadIdx.x & 31) == 0) atomicExch(error, i + 1);
}

__global__ void __launch_bounds__(256)
gz_stored_copy_kernel(const uint8_t *__restrict__ in, const GzUnit *__restrict__ units, uint32_t nunits,
                      uint8_t *__restrict__ out) {
  const uint32_t i = blockIdx.x;
  if (i >= nunits) return;
  const GzUnit u = units[i];
  if (u.kind == kGzCompressed) return;
  block_copy_unaligned(out + u.out_off, in + u.in_off, u.out_len, threadIdx.x, 256);
}

// The last 32 KiB (or all) of every compressed unit, in two steps.
// gz_tails_plain (one CTA per unit, all at once): the symbols that are plain bytes, and a flag per unit whose tail holds
// window references at all (few do: a reference in a tail is a copy of a copy that started in the unit's first 32 KiB).
// gz_tails_kernel (one CTA, the flagged units in stream order): their window references, which point into the 32 KiB
// before the unit -- the tails of the units before it, final by the time this unit's turn comes. (As one sequential
// kernel over every unit and every tail symbol this step was 118 of the 150 ms of a 476 MB stream.)
__global__ void __launch_bounds__(256)
gz_tails_plain(const uint8_t *__restrict__ in, const GzUnit *__restrict__ units, uint32_t nunits,
               uint16_t *__restrict__ out16, uint64_t sym_base, uint8_t *__restrict__ out,
               uint32_t *__restrict__ needs_fix) {
  const uint32_t i = blockIdx.x;
  if (i >= nunits) return;
  const GzUnit u = units[i];
  if (threadIdx.x == 0) needs_fix[i] = 0;
  __syncthreads();
  const uint32_t t0 = u.out_len > kGzWindow ? u.out_len - kGzWindow : 0;
  if (u.kind != kGzCompressed) {  // a stored block's tail as symbols too: the rounds below look sources up in out16
    for (uint32_t j = t0 + threadIdx.x; j < u.out_len; j += 256) out16[u.out_off - sym_base + j] = in[u.in_off + j];
    return;
  }
  bool any = false;
  for (uint32_t j = t0 + threadIdx.x; j < u.out_len; j += 256) {
    const uint32_t v = out16[u.out_off - sym_base + j];
    if (v < 256) out[u.out_off + j] = (uint8_t)v;
    else any = true;
  }
  if (any) needs_fix[i] = 1;
}

// One round over the flagged tails, all at once: a window reference whose source symbol (in the tails in front of the
// unit) is a plain byte by now becomes that byte; one whose source is still a reference waits for the next round. The
// number of rounds is the longest chain of copies ACROSS units (a handful; a phrase handed down through thousands of
// chunks would need thousands: the caller stops after a fixed number of rounds and leaves the rest to the serial
// kernel below). A 16-bit symbol is read and written whole, and it only ever changes from reference to byte, so a
// racing reader sees one of two valid states.
__global__ void __launch_bounds__(256)
gz_tails_round(const GzUnit *__restrict__ units, uint32_t nunits, uint16_t *out16, uint64_t sym_base, uint8_t *out,
               uint32_t *__restrict__ needs_fix, uint32_t *__restrict__ open_units) {
  const uint32_t i = blockIdx.x;
  if (i >= nunits || needs_fix[i] == 0) return;
  const GzUnit u = units[i];
  const uint32_t t0 = u.out_len > kGzWindow ? u.out_len - kGzWindow : 0;
  __shared__ uint32_t left;
  if (threadIdx.x == 0) left = 0;
  __syncthreads();
  uint32_t mine = 0;
  for (uint32_t j = t0 + threadIdx.x; j < u.out_len; j += 256) {
    const uint32_t v = out16[u.out_off - sym_base + j];
    if (v < 256) continue;
    const uint64_t src = u.out_off - kGzWindow + (v - 256);  // a source in front of this batch of units is final
    const uint32_t sv = src < sym_base ? (uint32_t)out[src] : *reinterpret_cast<volatile uint16_t *>(out16 + (src - sym_base));
    if (sv < 256) {
      out16[u.out_off - sym_base + j] = (uint16_t)sv;
      out[u.out_off + j] = (uint8_t)sv;
    } else {
      mine++;
    }
  }
  if (mine) atomicAdd(&left, mine);
  __syncthreads();
  if (threadIdx.x == 0) {
    if (left == 0) needs_fix[i] = 0;
    else atomicAdd(open_units, 1u);
  }
}

__global__ void __launch_bounds__(1024)
gz_tails_kernel(const GzUnit *__restrict__ units, uint32_t nunits, const uint16_t *__restrict__ out16,
                uint64_t sym_base, uint8_t *__restrict__ out, const uint32_t *__restrict__ needs_fix) {
  // unit descriptors and flags come in batches through shared memory: per unit only its symbols and the window bytes
  // are fetched from global memory (a descriptor read per unit is one more dependent L2 round trip in the serial chain)
  __shared__ uint64_t s_off[1024];
  __shared__ uint32_t s_len[1024];  // 0 = nothing to do for this unit
  for (uint32_t base = 0; base < nunits; base += 1024) {
    __syncthreads();
    {
      const uint32_t i = base + threadIdx.x;
      uint32_t len = 0;
      uint64_t off = 0;
      if (i < nunits && needs_fix[i]) {
        const GzUnit u = units[i];
        len = u.out_len;
        off = u.out_off;
      }
      s_off[threadIdx.x] = off;
      s_len[threadIdx.x] = len;
    }
    __syncthreads();
    const uint32_t cnt = nunits - base < 1024 ? nunits - base : 1024;
    for (uint32_t k0 = 0; k0 < cnt; k0++) {
      const uint32_t out_len = s_len[k0];
      if (out_len == 0) continue;  // uniform
      const uint64_t out_off = s_off[k0];
      const uint32_t t0 = out_len > kGzWindow ? out_len - kGzWindow : 0;
      // Every reference of this tail points in front of the unit, never into the tail itself: all loads first (32
      // independent symbol loads and gathers per thread in flight), then the stores. A load-store loop would serialise
      // on the possible aliasing of `out` with itself: one L2 round trip per symbol.
      uint32_t v[32];
#pragma unroll
      for (int k = 0; k < 32; k++) {
        const uint32_t j = t0 + (uint32_t)k * 1024u + threadIdx.x;
        v[k] = j < out_len ? out16[out_off - sym_base + j] : 0u;
      }
#pragma unroll
      for (int k = 0; k < 32; k++)
        if (v[k] >= 256) v[k] = 256u + out[out_off - kGzWindow + (v[k] - 256)];
#pragma unroll
      for (int k = 0; k < 32; k++)
        if (v[k] >= 256) out[out_off + t0 + (uint32_t)k * 1024u + threadIdx.x] = (uint8_t)(v[k] - 256);
      __syncthreads();  // the tail is visible to the whole CTA before the next unit reads its window
    }
  }
}

// One CTA per unit: everything in front of the tail.
__global__ void __launch_bounds__(256)
gz_bodies_kernel(const GzUnit *__restrict__ units, uint32_t nunits, const uint16_t *__restrict__ out16,
                 uint64_t sym_base, uint8_t *__restrict__ out) {
  const uint32_t i = blockIdx.x;
  if (i >= nunits) return;
  const GzUnit u = units[i];
  if (u.kind != kGzCompressed || u.out_len <= kGzWindow) return;
  const uint32_t body = u.out_len - kGzWindow;
  for (uint32_t j = threadIdx.x; j < body; j += 256) {
    const uint32_t v = out16[u.out_off - sym_base + j];
    out[u.out_off + j] = v < 256 ? (uint8_t)v : out[u.out_off - kGzWindow + (v - 256)];
  }
}

}  // namespace dsg
