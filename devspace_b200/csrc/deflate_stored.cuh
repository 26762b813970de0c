// deflate_stored.cuh -- compress/flate "stored" blocks (BTYPE=00), RFC 1951 3.2.4.
// The level-0 body of the gzip member and the per-chunk escape hatch for incompressible input
// (Go's writeBlock picks stored when it is the smallest encoding, compress/flate huffman_bit_writer.go
// [stdlib, not vendored]; call site /root/reference/pkg/devspace/sync/tar.go:156).
// Pure data movement: HBM bound, read n + write n + 5 bytes per 65535.
#pragma once

#include "copy_util.cuh"

namespace dsg {

constexpr uint32_t kStoredMax = 65535;  // LEN is 16 bits
constexpr int kStoredThreads = 256;

DSG_HD uint64_t stored_deflate_size(uint64_t n) { return n + 5 * ((n + kStoredMax - 1) / kStoredMax) + 2; }

// One CTA per stored block. The stream ends with an empty final fixed-Huffman block (03 00), the same
// terminator the LZ path uses, so every data block carries BFINAL=0 and blocks from different CTAs /
// segments concatenate by byte copy.
__global__ void __launch_bounds__(kStoredThreads)
deflate_stored_kernel(const uint8_t *__restrict__ in, uint64_t n, uint8_t *__restrict__ out, int write_terminator) {
  const uint64_t nblocks = (n + kStoredMax - 1) / kStoredMax;
  const uint64_t k = blockIdx.x;
  if (k < nblocks) {
    const uint64_t s0 = k * kStoredMax;
    const uint32_t len = (uint32_t)(n - s0 < kStoredMax ? n - s0 : kStoredMax);
    uint8_t *dst = out + k * (kStoredMax + 5);
    if (threadIdx.x == 0) {
      dst[0] = 0;  // BFINAL=0, BTYPE=00, byte aligned
      dst[1] = (uint8_t)(len & 0xff);
      dst[2] = (uint8_t)(len >> 8);
      dst[3] = (uint8_t)(~len & 0xff);
      dst[4] = (uint8_t)((~len >> 8) & 0xff);
    }
    block_copy_unaligned(dst + 5, in + s0, len, threadIdx.x, kStoredThreads);
  }
  if (write_terminator && k == (nblocks ? nblocks - 1 : 0) && threadIdx.x == 0) {
    uint8_t *end = out + n + 5 * nblocks;
    end[0] = 0x03;  // BFINAL=1, BTYPE=01, end-of-block code 0000000
    end[1] = 0x00;
  }
}

}  // namespace dsg
