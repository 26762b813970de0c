// dsg_common.cuh -- shared definitions for the libdsgpack kernels.
//
// The kernels live in .cuh headers so that the same source is compiled twice:
//   * by nvcc for sm_100a into libdsgpack.so (the product; devspace_b200/csrc/dsgpack.cu), and
//   * by g++ with -DDSG_EMU against tests/simt_emu/cuda_emu.h, a fiber-based SIMT emulator used by
//     the CPU test-suite to check kernel LOGIC without a GPU (test infrastructure, never shipped and
//     never reachable from the C ABI).
#pragma once

#include <stddef.h>
#include <stdint.h>

#include "../../include/dsgpack.h"

#if defined(DSG_EMU)
#include "cuda_emu.h"
#define DSG_DYN_SMEM(name) unsigned char *name = ::emu::dyn_smem()
#else
#include <cuda_runtime.h>
#define DSG_DYN_SMEM(name) extern __shared__ __align__(128) unsigned char name[]
#endif

#if defined(__CUDACC__) || defined(DSG_EMU)
#define DSG_HD __host__ __device__ __forceinline__
#define DSG_D __device__ __forceinline__
#else
#define DSG_HD inline
#define DSG_D inline
#endif

namespace dsg {

constexpr int kBlock = 512;  // tar block size (archive/tar format.go: blockSize)

DSG_HD uint64_t pad512(uint64_t n) { return (n + 511u) & ~uint64_t(511); }

// Device-side status word: first error wins (lowest entry index), see dsg_status.
// One 64-bit word (little endian: code in the low half, entry in the high half) updated with a single atomicMin.
struct alignas(8) DevStatus {
  int code;            // dsg_status, 0 = ok
  unsigned int entry;  // index of the offending entry
};
static_assert(sizeof(DevStatus) == 8, "DevStatus is one 64-bit word");

}  // namespace dsg
