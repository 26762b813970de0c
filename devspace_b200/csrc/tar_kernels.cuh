// tar_kernels.cuh -- the archive/tar Writer as three small plan kernels and one layout kernel.
//
// Replaces, for a whole manifest at once, the per-entry sequence of
//   tw.WriteHeader(hdr) ; io.Copy(tw, f) ; (implicit) tw.Flush()      tar.go:253,294,298
// and the final tarWriter.Close() (tar.go:160).
//
//   tar_classify_kernel : one thread per entry -> 512-byte blocks the entry occupies
//                         (Header.allowedFormats decides USTAR vs PAX), per-CTA partial sums
//   tar_scan_partials   : exclusive scan of the CTA sums (single CTA)
//   tar_offsets_kernel  : per-entry stream offset ent_off[i] (bytes), ent_off[n] = end of entries
//   tar_layout_kernel   : one CTA per 32 KiB tile of the OUTPUT stream; every byte of the tile is
//                         written exactly once (header blocks, payload, zero padding, trailer), so the
//                         stream needs no memset and HBM traffic is read P + manifest, write T.
//                         Headers are formatted one thread per entry into bank-skewed shared-memory
//                         rows and leave the SM as coalesced 128-byte stores.
#pragma once

#include "tar_format.cuh"

namespace dsg {

constexpr int kPlanThreads = 256;          // entries per CTA in the plan kernels
constexpr int kTileBlocks = 64;            // 512-byte blocks per layout tile
constexpr int kTileBytes = kTileBlocks * kBlock;  // 32 KiB
constexpr int kLayoutThreads = 256;
constexpr int kRowWords = 129;             // 512 B + 4 B skew: row r, word w -> bank (r + w) % 32

DSG_D void report_error(DevStatus *st, int code, unsigned int entry) {
  // first error = lowest entry index. Entry (high half) and code (low half) travel in ONE 64-bit atomicMin, so the
  // code that the host reads always belongs to the entry it reads.
  const unsigned long long w = ((unsigned long long)entry << 32) | (unsigned long long)(unsigned int)code;
  atomicMin(reinterpret_cast<unsigned long long *>(st), w);
}

// ---- plan -------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kPlanThreads)
tar_classify_kernel(const dsg_entry *__restrict__ entries, uint32_t n, const char *__restrict__ strtab,
                    uint64_t strtab_len, uint64_t arena_len, uint64_t *__restrict__ ent_blocks,
                    uint64_t *__restrict__ cta_sums, unsigned long long *__restrict__ payload_total,
                    DevStatus *status) {
  __shared__ uint64_t warp_sums[kPlanThreads / 32];
  __shared__ uint64_t warp_pay[kPlanThreads / 32];
  uint32_t i = blockIdx.x * kPlanThreads + threadIdx.x;
  uint64_t blocks = 0, pay = 0;
  if (i < n) {
    dsg_entry e = entries[i];
    HdrPlan p;
    int rc = classify_entry(e, strtab, strtab_len, p);
    if (rc == DSG_OK && e.typeflag == '0' && e.size > 0 &&
        (e.data_off > arena_len || (uint64_t)e.size > arena_len - e.data_off))
      rc = DSG_E_RANGE;
    if (rc != DSG_OK) {
      report_error(status, rc, i);
    } else {
      blocks = (p.hdr_bytes + entry_payload_padded(e)) / kBlock;
      pay = e.typeflag == '0' ? (uint64_t)e.size : 0;
    }
    ent_blocks[i] = blocks;
  }
  // CTA reduction (warp shuffles, then 8 partials)
  for (int o = 16; o > 0; o >>= 1) {
    blocks += __shfl_down_sync(0xffffffffu, blocks, o);
    pay += __shfl_down_sync(0xffffffffu, pay, o);
  }
  if ((threadIdx.x & 31) == 0) {
    warp_sums[threadIdx.x >> 5] = blocks;
    warp_pay[threadIdx.x >> 5] = pay;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint64_t s = 0, q = 0;
    for (int w = 0; w < kPlanThreads / 32; w++) {
      s += warp_sums[w];
      q += warp_pay[w];
    }
    cta_sums[blockIdx.x] = s;
    if (q) atomicAdd(payload_total, (unsigned long long)q);
  }
}

// single CTA: cta_sums[0..m) -> exclusive prefix in place, total blocks -> *total_blocks
__global__ void __launch_bounds__(1024) tar_scan_partials(uint64_t *__restrict__ cta_sums, uint32_t m,
                                                           uint64_t *__restrict__ total_blocks) {
  __shared__ uint64_t warp_tot[32];
  __shared__ uint64_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t base = 0; base < m; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint64_t v = i < m ? cta_sums[i] : 0;
    uint64_t x = v;  // inclusive scan inside the warp
    for (int o = 1; o < 32; o <<= 1) {
      uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (uint32_t)o) x += y;
    }
    if (lane == 31) warp_tot[warp] = x;
    __syncthreads();
    if (warp == 0) {  // exclusive scan of the 32 warp totals
      const uint64_t t = warp_tot[lane];
      uint64_t s = t;
      for (int o = 1; o < 32; o <<= 1) {
        uint64_t y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane >= (uint32_t)o) s += y;
      }
      warp_tot[lane] = s - t;
    }
    __syncthreads();
    const uint64_t carry = carry_s;
    const uint64_t excl = carry + warp_tot[warp] + (x - v);
    if (i < m) cta_sums[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl + v;  // total through this chunk of 1024
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_blocks = carry_s;
}

__global__ void __launch_bounds__(kPlanThreads)
tar_offsets_kernel(const uint64_t *__restrict__ ent_blocks, uint32_t n, const uint64_t *__restrict__ cta_base,
                   uint64_t *__restrict__ ent_off) {
  __shared__ uint64_t warp_tot[kPlanThreads / 32];
  uint32_t i = blockIdx.x * kPlanThreads + threadIdx.x;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint64_t v = i < n ? ent_blocks[i] : 0;
  uint64_t x = v;
  for (int o = 1; o < 32; o <<= 1) {
    uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (uint32_t)o) x += y;
  }
  if (lane == 31) warp_tot[warp] = x;
  __syncthreads();
  uint64_t wbase = 0;
  for (uint32_t w = 0; w < warp; w++) wbase += warp_tot[w];
  uint64_t excl = cta_base[blockIdx.x] + wbase + (x - v);
  if (i < n) ent_off[i] = excl * kBlock;
  if (i == n - 1) ent_off[n] = (excl + v) * kBlock;
}

// Arena byte range each pipeline segment of the stream needs: seg_lo[s] = min, seg_hi[s] = max over the
// payload pieces falling into stream bytes [s*seg_bytes, (s+1)*seg_bytes). One thread per entry. The host
// uploads exactly [seg_lo, seg_hi) before the segment's layout kernel (any arena order is correct; an
// arena laid out in entry order gives minimal, non-overlapping uploads). seg_lo starts at ~0, seg_hi at 0.
__global__ void __launch_bounds__(kPlanThreads)
tar_segment_ranges_kernel(const dsg_entry *__restrict__ entries, uint32_t n, const uint64_t *__restrict__ ent_off,
                          uint64_t seg_bytes, unsigned long long *__restrict__ seg_lo,
                          unsigned long long *__restrict__ seg_hi) {
  const uint32_t i = blockIdx.x * kPlanThreads + threadIdx.x;
  if (i >= n) return;
  const dsg_entry e = entries[i];
  if (e.typeflag != '0' || e.size <= 0) return;
  const uint64_t ps = ent_off[i + 1] - pad512((uint64_t)e.size);  // first payload byte in the stream
  const uint64_t pe = ps + (uint64_t)e.size;
  for (uint64_t s = ps / seg_bytes; s * seg_bytes < pe; s++) {
    const uint64_t lo = ps > s * seg_bytes ? ps : s * seg_bytes;
    const uint64_t hi = pe < (s + 1) * seg_bytes ? pe : (s + 1) * seg_bytes;
    atomicMin(&seg_lo[s], (unsigned long long)(e.data_off + (lo - ps)));
    atomicMax(&seg_hi[s], (unsigned long long)(e.data_off + (hi - ps)));
  }
}

// ---- layout -----------------------------------------------------------------------------------------

// Largest index e in [0, n] with ent_off[e] <= off  (ent_off is non-decreasing, ent_off[0] = 0).
// Warp-cooperative 32-ary search: 4 rounds for a million entries instead of 20 dependent loads.
DSG_D uint32_t warp_find_entry(const uint64_t *__restrict__ ent_off, uint32_t n, uint64_t off) {
  const uint32_t lane = threadIdx.x & 31;
  uint32_t lo = 0, hi = n + 1;  // answer in [lo, hi)
  while (hi - lo > 1) {
    uint32_t span = hi - lo;
    uint32_t step = (span + 31) / 32;
    uint32_t idx = lo + lane * step;  // lane 0 probes lo (always <= off)
    bool le = idx < hi && ent_off[idx] <= off;
    uint32_t m = __ballot_sync(0xffffffffu, le);
    uint32_t k = 31 - __clz(m);  // highest lane whose probe is <= off (m != 0: lane 0)
    uint32_t nlo = lo + k * step;
    uint32_t nhi = nlo + step < hi ? nlo + step : hi;
    lo = nlo;
    hi = nhi;
  }
  return lo;
}

DSG_D void copy_block_16(uint8_t *__restrict__ dst, const uint8_t *__restrict__ src, uint32_t valid, uint32_t lane) {
  // one warp moves one 512-byte block: lane l owns bytes [16l, 16l+16); bytes >= valid are zero padding
  uint32_t b0 = lane * 16;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (b0 + 16 <= valid && ((uintptr_t)(src + b0) & 15) == 0) {
    v = *reinterpret_cast<const uint4 *>(src + b0);
  } else if (b0 < valid) {
    uint32_t w[4] = {0, 0, 0, 0};
    uint32_t lim = valid - b0 < 16 ? valid - b0 : 16;
    if (((uintptr_t)(src + b0) & 3) == 0) {
      uint32_t k = 0;
      for (; k + 4 <= lim; k += 4) w[k >> 2] = *reinterpret_cast<const uint32_t *>(src + b0 + k);
      for (; k < lim; k++) w[k >> 2] |= (uint32_t)src[b0 + k] << (8 * (k & 3));
    } else {
      for (uint32_t k = 0; k < lim; k++) w[k >> 2] |= (uint32_t)src[b0 + k] << (8 * (k & 3));
    }
    v = make_uint4(w[0], w[1], w[2], w[3]);
  }
  *reinterpret_cast<uint4 *>(dst + b0) = v;
}

// One CTA per 32 KiB tile of the stream [seg_lo + tile*32K, ...) clipped to [seg_lo, seg_hi).
// `out` points at stream offset seg_lo (segments are multiples of 512, SURVEY.md 8e / engine).
// `arena` points at arena offset arena_base (the engine uploads only the slice a segment needs).
__global__ void __launch_bounds__(kLayoutThreads)
tar_layout_kernel(const dsg_entry *__restrict__ entries, uint32_t n, const char *__restrict__ strtab,
                  const uint64_t *__restrict__ ent_off, const uint8_t *__restrict__ arena, uint64_t arena_base,
                  uint8_t *__restrict__ out, uint64_t seg_lo, uint64_t seg_hi, DevStatus *status) {
  __shared__ uint32_t rows[kTileBlocks * kRowWords];  // header staging, bank-skewed rows
  __shared__ uint64_t s_off[kTileBlocks + 2];         // ent_off[e0 .. e0+65]
  __shared__ uint32_t s_e0, s_cnt;
  __shared__ uint32_t blk_ent[kTileBlocks];           // entry owning each block (n = trailer / beyond)
  __shared__ uint8_t blk_kind[kTileBlocks];           // 0 payload/zero, 1 staged USTAR header, 2 skip (PAX area)

  const uint64_t t0 = seg_lo + (uint64_t)blockIdx.x * kTileBytes;
  if (t0 >= seg_hi) return;
  const uint64_t t1 = t0 + kTileBytes < seg_hi ? t0 + kTileBytes : seg_hi;
  const uint32_t nblk = (uint32_t)((t1 - t0) / kBlock);
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  if (warp == 0) {
    uint32_t e0 = warp_find_entry(ent_off, n, t0);
    if (lane == 0) s_e0 = e0;
    // entries overlapping the tile: at most 64 start inside it, plus the one straddling t0
    uint32_t cnt = 0;
    for (uint32_t k = lane; k < kTileBlocks + 2; k += 32) {
      uint32_t idx = e0 + k;
      uint64_t v = idx <= n ? ent_off[idx] : ~uint64_t(0);
      s_off[k] = v;
    }
    __syncwarp();
    (void)cnt;
  }
  __syncthreads();
  const uint32_t e0 = s_e0;

  // phase A: thread b < nblk resolves block b and, if it is the first header block of an entry,
  // formats that entry's header(s).
  if (threadIdx.x < nblk) {
    const uint32_t b = threadIdx.x;
    const uint64_t off = t0 + (uint64_t)b * kBlock;
    // largest k with s_off[k] <= off
    uint32_t lo = 0, hi = kTileBlocks + 1;
    while (lo < hi) {
      uint32_t mid = (lo + hi + 1) >> 1;
      if (s_off[mid] <= off) lo = mid; else hi = mid - 1;
    }
    const uint32_t ei = e0 + lo;
    blk_ent[b] = ei;
    uint8_t kind = 0;
    if (ei < n) {
      const dsg_entry e = entries[ei];
      const uint64_t eoff = s_off[lo];
      const uint64_t hdr_bytes = (s_off[lo + 1] - eoff) - entry_payload_padded(e);
      if (off < eoff + hdr_bytes) {
        kind = 2;
        // the thread on the entry's first header block formats it; if a pipeline segment starts
        // inside a (multi-block, PAX) header area the thread on the segment's first block takes over
        if (off == eoff || off == seg_lo) {
          HdrPlan p;
          classify_entry(e, strtab, ~uint64_t(0), p);  // validated by tar_classify_kernel
          if (p.flags & kNeedPax) {
            // rare path: extended header + records + main header straight to global memory,
            // clipped to this segment
            const uint64_t lo_abs = eoff > seg_lo ? eoff : seg_lo;
            const uint64_t hi_abs = eoff + hdr_bytes < seg_hi ? eoff + hdr_bytes : seg_hi;
            // dst index 0 corresponds to stream offset eoff; only the window is dereferenced
            uint8_t *area = out + ((int64_t)eoff - (int64_t)seg_lo);
            int rc = format_pax_headers(area, (uint32_t)(lo_abs - eoff), (uint32_t)(hi_abs - eoff), e, strtab, p);
            if (rc != DSG_OK) report_error(status, rc, ei);
          } else {
            kind = 1;
            uint8_t *row = reinterpret_cast<uint8_t *>(rows + b * kRowWords);
            for (int w = 0; w < 128; w++) rows[b * kRowWords + w] = 0;
            format_main_header(row, e, strtab, p);
          }
        }
      }
    }
    blk_kind[b] = kind;
  }
  __syncthreads();

  // phase B: each warp moves blocks warp, warp+8, ... as 32 x 16-byte lanes
  for (uint32_t b = warp; b < nblk; b += kLayoutThreads / 32) {
    const uint64_t off = t0 + (uint64_t)b * kBlock;
    uint8_t *dst = out + (off - seg_lo);
    const uint8_t kind = blk_kind[b];
    if (kind == 2) continue;
    if (kind == 1) {
      const uint32_t *row = rows + b * kRowWords;
      uint32_t *d32 = reinterpret_cast<uint32_t *>(dst);
#pragma unroll
      for (int j = 0; j < 4; j++) d32[lane + 32 * j] = row[lane + 32 * j];
      continue;
    }
    const uint32_t ei = blk_ent[b];
    uint32_t valid = 0;
    const uint8_t *src = arena;
    if (ei < n) {
      const dsg_entry e = entries[ei];
      const uint32_t k = ei - e0;
      const uint64_t pstart = s_off[k + 1] - entry_payload_padded(e);  // first payload byte in the stream
      const uint64_t rel = off - pstart;                               // multiple of 512
      if (e.typeflag == '0' && rel < (uint64_t)e.size) {
        uint64_t left = (uint64_t)e.size - rel;
        valid = left < kBlock ? (uint32_t)left : kBlock;
        src = arena + (e.data_off - arena_base) + rel;
      }
    }
    copy_block_16(dst, src, valid, lane);  // valid == 0: trailer / padding block -> zeros
  }
}

}  // namespace dsg
