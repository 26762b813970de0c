// tests/emu_harness.cc -- TEST INFRASTRUCTURE ONLY.
// Compiles the product's kernel headers (devspace_b200/csrc/*.cuh) with -DDSG_EMU against the SIMT
// emulator and exposes, via a C ABI for ctypes, the same launch sequences the CUDA host code performs,
// so the CPU test-suite can check kernel logic against the oracle before any GPU time is spent.
// Nothing in the product links, loads or calls this file.
#include <stdio.h>
#include <string.h>

#include <vector>

#include "tar_kernels.cuh"

dim3 threadIdx, blockIdx, blockDim, gridDim;

using namespace dsg;

extern "C" {

// Mirrors Engine::tar_build in dsgpack.cu: classify -> scan -> offsets -> layout (segmented).
int emu_tar_build(const dsg_entry *entries, uint32_t n, const char *strtab, uint64_t strtab_len, const uint8_t *arena,
                  uint64_t arena_len, uint32_t flags, uint64_t segment_bytes, uint8_t *out, uint64_t cap,
                  uint64_t *tar_len, uint32_t *err_entry) {
  const uint32_t ncta = (n + kPlanThreads - 1) / kPlanThreads;
  std::vector<uint64_t> ent_blocks(n + 1), cta_sums(ncta + 1), ent_off(n + 1, 0);
  unsigned long long payload_total = 0;
  uint64_t total_blocks = 0;
  DevStatus st{0, 0xffffffffu};
  if (n) {
    emu::launch(dim3(ncta), dim3(kPlanThreads), 0, [&] {
      tar_classify_kernel(entries, n, strtab, strtab_len, arena_len, ent_blocks.data(), cta_sums.data(),
                          &payload_total, &st);
    });
    if (st.code) {
      if (err_entry) *err_entry = st.entry;
      return st.code;
    }
    emu::launch(dim3(1), dim3(1024), 0, [&] { tar_scan_partials(cta_sums.data(), ncta, &total_blocks); });
    emu::launch(dim3(ncta), dim3(kPlanThreads), 0,
                [&] { tar_offsets_kernel(ent_blocks.data(), n, cta_sums.data(), ent_off.data()); });
  }
  const uint64_t total = ent_off[n] + ((flags & DSG_F_NO_TRAILER) ? 0 : 2 * kBlock);
  *tar_len = total;
  if (total > cap) return DSG_E_CAPACITY;
  if (segment_bytes == 0) segment_bytes = total ? total : 512;
  segment_bytes = (segment_bytes + kTileBytes - 1) / kTileBytes * kTileBytes;
  for (uint64_t lo = 0; lo < total; lo += segment_bytes) {
    uint64_t hi = lo + segment_bytes < total ? lo + segment_bytes : total;
    uint32_t tiles = (uint32_t)((hi - lo + kTileBytes - 1) / kTileBytes);
    emu::launch(dim3(tiles), dim3(kLayoutThreads), 0, [&] {
      tar_layout_kernel(entries, n, strtab, ent_off.data(), arena, 0, out + lo, lo, hi, &st);
    });
  }
  if (st.code) {
    if (err_entry) *err_entry = st.entry;
    return st.code;
  }
  return 0;
}

}  // extern "C"
