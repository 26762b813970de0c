// deflate_lz.cuh -- LZ77 + dynamic Huffman deflate kernels (compress/flate level 6 equivalent).
#pragma once

#include "dsg_common.cuh"
#include "deflate_stored.cuh"

namespace dsg {

// Upper bound of the raw deflate output for n input bytes (any level): a chunk never exceeds its
// stored encoding, plus the 2-byte terminator.
DSG_HD uint64_t lz_deflate_bound(uint64_t n) { return stored_deflate_size(n) + 64; }

}  // namespace dsg
