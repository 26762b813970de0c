// file_crc.cuh -- CRC32-IEEE of every file of a manifest in one pass over the arena: the per-file checksum of
// DevSpace's build-context hash, /root/reference/pkg/util/hash/hash.go:165-197 (hashFileCRC32: crc32.New over
// the file bytes, polynomial 0xedb88320), called for every context file at hash.go:146 before every image build
// decision (pkg/devspace/image/build.go:213). The same bytes are what writeTar packs, so the arena that is
// resident for dsg_pack_device also yields the checksums.
//
// A file is cut into [head < 16 B][body: units of kFileCrcUnit bytes, 16-byte aligned in device memory][tail < 16 B].
// One CTA reduces one unit to its raw remainder with the tile fold of crc32.cuh; one warp per file joins the
// units (remainder * x^(8 * bytes that follow) mod P), adds head and tail bytewise and finalises.
// HBM traffic: the payload once. Bound: shared-memory table lookups (as crc_tiles_kernel), not HBM.
#pragma once

#include "crc32.cuh"

namespace dsg {

constexpr int kFileCrcUnit = 64 * 1024;

struct FileCrcEntry {  // the fields of dsg_entry this path reads (copied out by the caller's kernel)
  uint64_t off;
  uint64_t size;
};

DSG_HD uint32_t fcrc_head_len(uint64_t addr, uint64_t size) {
  const uint64_t h = (16 - (addr & 15)) & 15;
  return (uint32_t)(h < size ? h : size);
}

// units[i] = number of body units of file i (thread per file)
__global__ void fcrc_count_kernel(const FileCrcEntry *__restrict__ files, uint32_t n, uint64_t arena_addr,
                                  uint32_t *__restrict__ units) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t size = files[i].size;
  const uint32_t head = fcrc_head_len(arena_addr + files[i].off, size);
  const uint64_t body = (size - head) & ~15ull;
  units[i] = (uint32_t)((body + kFileCrcUnit - 1) / kFileCrcUnit);
}

// Exclusive scan of n counts in three launches: per-block (1024 values) scan + block totals, scan of the totals
// (one block, any count), add. first[n] = total.
constexpr int kScanBlock = 1024;

__device__ inline uint32_t fcrc_block_scan(uint32_t v, uint32_t *s_warp, uint32_t *total) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t x = v;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (uint32_t)o) x += y;
  }
  if (lane == 31) s_warp[warp] = x;
  __syncthreads();
  if (warp == 0) {
    const uint32_t t = s_warp[lane];
    uint32_t s = t;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= (uint32_t)o) s += y;
    }
    s_warp[lane] = s - t;
    if (lane == 31) s_warp[32] = s;
  }
  __syncthreads();
  *total = s_warp[32];
  return s_warp[warp] + (x - v);
}

__global__ void __launch_bounds__(kScanBlock)
fcrc_scan_blocks_kernel(const uint32_t *__restrict__ in, uint32_t n, uint32_t *__restrict__ out,
                        uint32_t *__restrict__ block_totals) {
  __shared__ uint32_t s_warp[33];
  const uint32_t i = blockIdx.x * kScanBlock + threadIdx.x;
  uint32_t total;
  const uint32_t ex = fcrc_block_scan(i < n ? in[i] : 0u, s_warp, &total);
  if (i < n) out[i] = ex;
  if (threadIdx.x == 0) block_totals[blockIdx.x] = total;
}

// one block: exclusive scan of the block totals in place; grand total to *grand
__global__ void __launch_bounds__(kScanBlock)
fcrc_scan_totals_kernel(uint32_t *__restrict__ totals, uint32_t nblocks, uint32_t *__restrict__ grand) {
  __shared__ uint32_t s_warp[33];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < nblocks; base += kScanBlock) {
    const uint32_t i = base + threadIdx.x;
    uint32_t total;
    const uint32_t ex = fcrc_block_scan(i < nblocks ? totals[i] : 0u, s_warp, &total);
    const uint32_t c = carry;
    __syncthreads();
    if (i < nblocks) totals[i] = c + ex;
    if (threadIdx.x == 0) carry = c + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *grand = carry;
}

__global__ void __launch_bounds__(kScanBlock)
fcrc_scan_add_kernel(uint32_t *__restrict__ out, uint32_t n, const uint32_t *__restrict__ block_offsets,
                     const uint32_t *__restrict__ grand) {
  const uint32_t i = blockIdx.x * kScanBlock + threadIdx.x;
  if (i < n) out[i] += block_offsets[blockIdx.x];
  if (i == 0) out[n] = *grand;
}

// unit_file[u] = the file unit u belongs to (thread per file; a binary search per unit in the units kernel cost 17
// dependent global loads per unit, most of the time of a tree of small files)
__global__ void fcrc_map_kernel(const uint32_t *__restrict__ first, uint32_t n, uint32_t *__restrict__ unit_file) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (uint32_t u = first[i]; u < first[i + 1]; u++) unit_file[u] = i;
}

// One CTA per kFileCrcUnitsPerCta consecutive units (the 21 KB of lookup tables are loaded once per CTA: with one
// unit per CTA a tree of 100 000 small files spent its time loading tables): raw remainder of each unit's bytes.
// first[] = exclusive scan of the unit counts (n + 1 entries).
constexpr int kFileCrcUnitsPerCta = 8;

__global__ void __launch_bounds__(kCrcThreads)
fcrc_units_kernel(const FileCrcEntry *__restrict__ files, uint32_t n, const uint8_t *__restrict__ arena,
                  const uint32_t *__restrict__ first, const uint32_t *__restrict__ unit_file, uint32_t nunits,
                  const uint32_t *__restrict__ g_tab,
                  const uint32_t *__restrict__ g_shift, uint32_t *__restrict__ raw_out) {
  __shared__ uint32_t T[kCrcTables * 256];
  __shared__ uint32_t S[kCrcShiftEntries];
  __shared__ uint32_t red[kCrcThreads / 32];
  for (int i = threadIdx.x; i < kCrcTables * 256; i += kCrcThreads) T[i] = g_tab[i];
  for (int i = threadIdx.x; i < kCrcShiftEntries; i += kCrcThreads) S[i] = g_shift[i];
  const uint32_t *MK = T + 16 * 256;
  const uint32_t u0 = blockIdx.x * kFileCrcUnitsPerCta;
  for (uint32_t u = u0; u < u0 + kFileCrcUnitsPerCta && u < nunits; u++) {
    __syncthreads();  // tables loaded; red of the previous unit consumed
    const uint32_t f = unit_file[u];
    const uint64_t size = files[f].size;
    const uint8_t *p = arena + files[f].off;
    const uint32_t head = fcrc_head_len((uint64_t)(uintptr_t)p, size);
    const uint64_t body = (size - head) & ~15ull;
    const uint64_t b0 = (uint64_t)(u - first[f]) * kFileCrcUnit;
    const uint64_t nbytes = body - b0 < (uint64_t)kFileCrcUnit ? body - b0 : (uint64_t)kFileCrcUnit;
    const uint4 *base = reinterpret_cast<const uint4 *>(p + head + b0);
    const uint32_t nw = (uint32_t)(nbytes / 16);
    uint32_t acc = 0, last = 0;
    bool any = false;
    for (uint32_t w = threadIdx.x; w < nw; w += kCrcThreads) {
      const uint4 v = base[w];
      acc = MK[acc & 0xff] ^ MK[256 + ((acc >> 8) & 0xff)] ^ MK[512 + ((acc >> 16) & 0xff)] ^ MK[768 + (acc >> 24)];
      acc ^= crc_raw16(T, v);
      last = w;
      any = true;
    }
    if (any) {
      const uint32_t after = nw - 1 - last;
      acc = crc_mul(acc, after < (uint32_t)kCrcShiftEntries ? S[after] : crc_xpow8((uint64_t)after * 16));
    }
    for (int o = 16; o > 0; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t r = 0;
      for (int i = 0; i < kCrcThreads / 32; i++) r ^= red[i];
      raw_out[u] = r;
    }
  }
}

// One warp per file: joins the unit remainders, adds the unaligned head and tail, finalises. crc of an empty file = 0.
__global__ void __launch_bounds__(256)
fcrc_join_kernel(const FileCrcEntry *__restrict__ files, uint32_t n, const uint8_t *__restrict__ arena,
                 const uint32_t *__restrict__ first, const uint32_t *__restrict__ raw, uint32_t *__restrict__ crcs) {
  const uint32_t f = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  if (f >= n) return;
  const uint64_t size = files[f].size;
  const uint8_t *p = arena + files[f].off;
  const uint32_t head = fcrc_head_len((uint64_t)(uintptr_t)p, size);
  const uint64_t body = (size - head) & ~15ull;
  const uint32_t tail = (uint32_t)(size - head - body);
  const uint32_t u0 = first[f], nu = first[f + 1] - u0;
  // Each lane folds a contiguous stretch of units Horner style (one multiplication per unit: by x^(8 * length of
  // the unit it takes in), then moves its remainder to the end of the file with one x^(8n).
  uint32_t acc = 0;
  if (nu) {
    const uint32_t per = (nu + 31) / 32;
    const uint32_t lo = lane * per, hi = lo + per < nu ? lo + per : nu;
    if (lo < hi) {
      const uint32_t k_full = hi - lo > 1 ? crc_xpow8(kFileCrcUnit) : 0u;  // (a power costs ~40 multiplications)
      const uint64_t last_len = body - (uint64_t)(nu - 1) * kFileCrcUnit;
      acc = raw[u0 + lo];
      for (uint32_t k = lo + 1; k < hi; k++) {
        const uint32_t mult = (k == nu - 1) ? crc_xpow8(last_len) : k_full;
        acc = crc_mul(acc, mult) ^ raw[u0 + k];
      }
      uint64_t end = (uint64_t)hi * kFileCrcUnit;
      if (end > body) end = body;
      acc = crc_mul(acc, crc_xpow8((body - end) + tail));
    }
  }
  for (int o = 16; o > 0; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
  // The two remaining powers are computed side by side on lanes 1 and 2: for a tree of small files these powers,
  // on one lane of an otherwise idle warp, were most of the kernel's time.
  uint32_t pw = 0;
  if (lane == 1 && head) pw = crc_xpow8(body + tail);
  if (lane == 2 && size) pw = crc_xpow8(size);
  const uint32_t pw_head = __shfl_sync(0xffffffffu, pw, 1), pw_size = __shfl_sync(0xffffffffu, pw, 2);
  if (lane == 0) {
    auto raw_bytes = [](const uint8_t *q, uint32_t m) {
      uint32_t c = 0;
      for (uint32_t i = 0; i < m; i++) {
        c ^= q[i];
        for (int b = 0; b < 8; b++) c = (c & 1u) ? (c >> 1) ^ kCrcPoly : (c >> 1);
      }
      return c;
    };
    if (head) acc ^= crc_mul(raw_bytes(p, head), pw_head);
    if (tail) acc ^= raw_bytes(p + head + body, tail);
    crcs[f] = size ? (acc ^ crc_mul(0xffffffffu, pw_size) ^ 0xffffffffu) : 0u;  // crc_finalise with the power given
  }
}

// dsg_entry (80-byte manifest record) -> FileCrcEntry; directories and empty files get size 0
__global__ void fcrc_extract_kernel(const uint8_t *__restrict__ entries, uint32_t n, uint32_t stride, uint32_t off_size,
                                    uint32_t off_data, uint32_t off_type, uint64_t arena_len,
                                    FileCrcEntry *__restrict__ out, uint32_t *__restrict__ error) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint8_t *e = entries + (size_t)i * stride;
  const int64_t size = *reinterpret_cast<const int64_t *>(e + off_size);
  const uint64_t off = *reinterpret_cast<const uint64_t *>(e + off_data);
  const uint8_t type = e[off_type];
  FileCrcEntry r{0, 0};
  if (type != '5' && size > 0) {
    if (off > arena_len || (uint64_t)size > arena_len - off) atomicExch(error, 1u + i);
    else r = FileCrcEntry{off, (uint64_t)size};
  } else if (size < 0) {
    atomicExch(error, 1u + i);
  }
  out[i] = r;
}

}  // namespace dsg
