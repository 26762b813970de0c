// crc32.cuh -- hash/crc32 ChecksumIEEE of the tar stream, as compress/gzip's Writer keeps it
// (z.digest = crc32.Update(z.digest, crc32.IEEETable, p) per Write; trailer at Close).
// Reached from /root/reference/pkg/devspace/sync/tar.go:156-157 (gzip.NewWriter ... gw.Close()).
//
// The CRC is linear over GF(2), so the stream is cut into 128 KiB tiles that are reduced independently
// and joined by multiplying each tile remainder by x^(8 * bytes that follow) mod P ("parallel combine",
// BASELINE.json north_star). Inside a tile the 16-byte words are dealt round-robin to the threads
// (coalesced 128-bit loads, no staging); a thread folds its words Horner style:
//     acc = acc * x^(8*stride) xor raw16(word)
// with 4 + 16 shared-memory table lookups per 16 bytes. HBM traffic: read T once; roofline: HBM
// (the kernel itself is shared-memory-lookup bound at about one lookup per byte).
#pragma once

#include "dsg_common.cuh"

namespace dsg {

constexpr uint32_t kCrcPoly = 0xEDB88320u;   // reflected IEEE polynomial
constexpr int kCrcThreads = 256;
constexpr int kCrcTileBytes = 128 * 1024;     // 256 threads x 32 words x 16 B
constexpr int kCrcTables = 20;                // 16 slice tables + 4 stride-multiply tables
constexpr int kCrcShiftEntries = 256;         // x^(8*16*k), k = 0..255

// a(x) * b(x) mod P, reflected bit order (bit 31 = x^0); same recurrence as zlib's multmodp
DSG_HD uint32_t crc_mul(uint32_t a, uint32_t b) {
  uint32_t p = 0;
  for (int i = 0; i < 32; i++) {
    if (a & 0x80000000u) p ^= b;
    a <<= 1;
    b = (b & 1u) ? (b >> 1) ^ kCrcPoly : (b >> 1);
  }
  return p;
}

// x^(8*nbytes) mod P by square and multiply
DSG_HD uint32_t crc_xpow8(uint64_t nbytes) {
  uint32_t r = 0x80000000u;   // x^0
  uint32_t sq = 0x00800000u;  // x^8
  while (nbytes) {
    if (nbytes & 1) r = crc_mul(r, sq);
    sq = crc_mul(sq, sq);
    nbytes >>= 1;
  }
  return r;
}

// Standard (finalised) CRC of a || b from the finalised CRCs of a and b (zlib crc32_combine identity).
DSG_HD uint32_t crc_combine(uint32_t crc_a, uint32_t crc_b, uint64_t len_b) {
  return crc_mul(crc_xpow8(len_b), crc_a) ^ crc_b;
}

// Finalised CRC from the raw remainder (init 0, no xorout) of an n-byte message:
// init 0xffffffff contributes 0xffffffff * x^(8n); xorout 0xffffffff.
DSG_HD uint32_t crc_finalise(uint32_t raw, uint64_t n) { return raw ^ crc_mul(0xffffffffu, crc_xpow8(n)) ^ 0xffffffffu; }

// Host-side table construction (ctx creation). Layout: tab[m][b], m = 0..15: raw remainder of byte b
// followed by m zero bytes; tab[16+k][b]: (b placed in byte k of the accumulator) * x^(8*stride_bytes).
// shift[k] = x^(8*16*k).
inline void crc_build_tables(uint32_t *tab /* 20*256 */, uint32_t *shift /* 256 */, uint64_t stride_bytes) {
  for (uint32_t b = 0; b < 256; b++) {
    uint32_t c = b;
    for (int k = 0; k < 8; k++) c = (c & 1u) ? (c >> 1) ^ kCrcPoly : (c >> 1);
    tab[b] = c;
  }
  for (int m = 1; m < 16; m++)
    for (uint32_t b = 0; b < 256; b++) {
      uint32_t c = tab[(m - 1) * 256 + b];
      tab[m * 256 + b] = tab[c & 0xff] ^ (c >> 8);  // one more zero byte
    }
  const uint32_t K = crc_xpow8(stride_bytes);
  for (int k = 0; k < 4; k++)
    for (uint32_t b = 0; b < 256; b++) tab[(16 + k) * 256 + b] = crc_mul(b << (8 * k), K);
  for (int k = 0; k < kCrcShiftEntries; k++) shift[k] = crc_xpow8((uint64_t)16 * k);
}

// raw remainder (init 0) of one 16-byte word, memory order x,y,z,w little endian
DSG_D uint32_t crc_raw16(const uint32_t *__restrict__ T, uint4 v) {
  uint32_t r = 0;
  r ^= T[15 * 256 + (v.x & 0xff)] ^ T[14 * 256 + ((v.x >> 8) & 0xff)] ^ T[13 * 256 + ((v.x >> 16) & 0xff)] ^
       T[12 * 256 + (v.x >> 24)];
  r ^= T[11 * 256 + (v.y & 0xff)] ^ T[10 * 256 + ((v.y >> 8) & 0xff)] ^ T[9 * 256 + ((v.y >> 16) & 0xff)] ^
       T[8 * 256 + (v.y >> 24)];
  r ^= T[7 * 256 + (v.z & 0xff)] ^ T[6 * 256 + ((v.z >> 8) & 0xff)] ^ T[5 * 256 + ((v.z >> 16) & 0xff)] ^
       T[4 * 256 + (v.z >> 24)];
  r ^= T[3 * 256 + (v.w & 0xff)] ^ T[2 * 256 + ((v.w >> 8) & 0xff)] ^ T[1 * 256 + ((v.w >> 16) & 0xff)] ^
       T[0 * 256 + (v.w >> 24)];
  return r;
}

// One CTA per tile of `words16` 16-byte words starting at data (16-byte aligned).
// raw_out[tile] = raw remainder of the tile's bytes.
__global__ void __launch_bounds__(kCrcThreads)
crc_tiles_kernel(const uint4 *__restrict__ data, uint64_t words16, const uint32_t *__restrict__ g_tab,
                 const uint32_t *__restrict__ g_shift, uint32_t *__restrict__ raw_out) {
  __shared__ uint32_t T[kCrcTables * 256];
  __shared__ uint32_t S[kCrcShiftEntries];
  __shared__ uint32_t red[kCrcThreads / 32];
  for (int i = threadIdx.x; i < kCrcTables * 256; i += kCrcThreads) T[i] = g_tab[i];
  for (int i = threadIdx.x; i < kCrcShiftEntries; i += kCrcThreads) S[i] = g_shift[i];
  __syncthreads();
  constexpr uint64_t kTileWords = kCrcTileBytes / 16;
  const uint64_t w0 = (uint64_t)blockIdx.x * kTileWords;
  const uint64_t nw = words16 - w0 < kTileWords ? words16 - w0 : kTileWords;
  const uint4 *base = data + w0;
  uint32_t acc = 0;
  uint64_t last = 0;
  bool any = false;
  const uint32_t *MK = T + 16 * 256;
  for (uint64_t w = threadIdx.x; w < nw; w += kCrcThreads) {
    uint4 v = base[w];
    // acc * x^(8 * 16 * 256): four lookups, one per accumulator byte
    acc = MK[acc & 0xff] ^ MK[256 + ((acc >> 8) & 0xff)] ^ MK[512 + ((acc >> 16) & 0xff)] ^ MK[768 + (acc >> 24)];
    acc ^= crc_raw16(T, v);
    last = w;
    any = true;
  }
  if (any) {
    // bring the thread's remainder to the end of the tile: (nw - 1 - last) words follow it
    uint64_t after = nw - 1 - last;
    uint32_t sh = after < (uint64_t)kCrcShiftEntries ? S[after] : crc_xpow8(after * 16);
    acc = crc_mul(acc, sh);
  }
  for (int o = 16; o > 0; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t r = 0;
    for (int i = 0; i < kCrcThreads / 32; i++) r ^= red[i];
    raw_out[blockIdx.x] = r;
  }
}

// Joins the tile remainders, the unaligned head/tail bytes, and a running CRC from earlier segments.
//   message = [head bytes][ntiles tiles, last one possibly short][tail bytes]
//   state[0] (in/out) = finalised CRC of everything before this message; state[1] = its length (low),
//   state[2] = length (high). *state is updated to cover the message as well.
__global__ void __launch_bounds__(256)
crc_join_kernel(const uint32_t *__restrict__ raw, uint32_t ntiles, uint64_t body_bytes, const uint8_t *__restrict__ head,
                uint32_t head_len, const uint8_t *__restrict__ tail, uint32_t tail_len, uint32_t *__restrict__ state) {
  __shared__ uint32_t red[8];
  uint32_t acc = 0;
  for (uint32_t t = threadIdx.x; t < ntiles; t += 256) {
    uint64_t end = (uint64_t)(t + 1) * kCrcTileBytes;
    if (end > body_bytes) end = body_bytes;
    uint64_t after = (body_bytes - end) + tail_len;
    acc ^= crc_mul(raw[t], crc_xpow8(after));
  }
  for (int o = 16; o > 0; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t r = 0;
    for (int i = 0; i < 8; i++) r ^= red[i];
    // bytewise raw remainder of the few unaligned bytes
    auto raw_bytes = [](const uint8_t *p, uint32_t n) {
      uint32_t c = 0;
      for (uint32_t i = 0; i < n; i++) {
        c ^= p[i];
        for (int k = 0; k < 8; k++) c = (c & 1u) ? (c >> 1) ^ kCrcPoly : (c >> 1);
      }
      return c;
    };
    if (head_len) r ^= crc_mul(raw_bytes(head, head_len), crc_xpow8(body_bytes + tail_len));
    if (tail_len) r ^= raw_bytes(tail, tail_len);
    const uint64_t n = (uint64_t)head_len + body_bytes + tail_len;
    const uint32_t crc_msg = crc_finalise(r, n);
    const uint64_t prev_len = (uint64_t)state[1] | ((uint64_t)state[2] << 32);
    const uint32_t crc_all = prev_len ? crc_combine(state[0], crc_msg, n) : crc_msg;
    state[0] = n ? crc_all : state[0];
    const uint64_t tot = prev_len + n;
    state[1] = (uint32_t)tot;
    state[2] = (uint32_t)(tot >> 32);
  }
}

}  // namespace dsg
