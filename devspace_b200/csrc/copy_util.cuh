// copy_util.cuh -- CTA-cooperative byte copy between arbitrarily aligned global addresses.
// Used wherever the deflate bit stream forces odd byte offsets (stored blocks behind 5-byte headers,
// compaction of per-chunk outputs): destination words are written as aligned 128-bit stores, the source
// is read as aligned 32-bit words and realigned with funnel shifts.
#pragma once

#include "dsg_common.cuh"

namespace dsg {

// May read up to 3 bytes beyond src+n inside the last aligned 32-bit word (never across it).
// kStreamStore: the 128-bit stores carry the evict-first hint (data that is written once and read much later by
// another kernel should not push the working set out of L2).
template <bool kStreamStore = false>
DSG_D void block_copy_unaligned(uint8_t *__restrict__ dst, const uint8_t *__restrict__ src, uint64_t n, uint32_t tid,
                                uint32_t nthr) {
  if (n == 0) return;
  uint64_t head = (16 - ((uintptr_t)dst & 15)) & 15;
  if (head > n) head = n;
  for (uint64_t i = tid; i < head; i += nthr) dst[i] = src[i];
  const uint64_t body = (n - head) / 16;
  uint8_t *d = dst + head;
  const uint8_t *s = src + head;
  const uint32_t m = (uint32_t)((uintptr_t)s & 3);
  const uint32_t *s32 = reinterpret_cast<const uint32_t *>(s - m);
  const uint32_t sh = 8 * m;
  for (uint64_t w = tid; w < body; w += nthr) {
    const uint32_t *p = s32 + w * 4;
    uint32_t a0 = p[0], a1 = p[1], a2 = p[2], a3 = p[3];
    uint4 v;
    if (m == 0) {
      v = make_uint4(a0, a1, a2, a3);
    } else {
      uint32_t a4 = p[4];
      v = make_uint4(__funnelshift_r(a0, a1, sh), __funnelshift_r(a1, a2, sh), __funnelshift_r(a2, a3, sh),
                     __funnelshift_r(a3, a4, sh));
    }
#if defined(DSG_EMU)
    *reinterpret_cast<uint4 *>(d + w * 16) = v;
#else
    if (kStreamStore) __stcs(reinterpret_cast<uint4 *>(d + w * 16), v);
    else *reinterpret_cast<uint4 *>(d + w * 16) = v;
#endif
  }
  for (uint64_t i = head + body * 16 + tid; i < n; i += nthr) dst[i] = src[i];
}

}  // namespace dsg
