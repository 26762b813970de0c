// tests/emu_harness.cc -- TEST INFRASTRUCTURE ONLY.
// Compiles the product's kernel headers (devspace_b200/csrc/*.cuh) with -DDSG_EMU against the SIMT
// emulator and exposes, via a C ABI for ctypes, the same launch sequences the CUDA host code performs,
// so the CPU test-suite can check kernel logic against the oracle before any GPU time is spent.
// Nothing in the product links, loads or calls this file.
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "crc32.cuh"
#include "deflate_lz.cuh"
#include "file_crc.cuh"
#include "inflate.cuh"
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

// Mirrors lz_deflate_run in dsgpack.cu: one launch of the persistent LZ77+Huffman kernel.
int emu_lz_deflate(const uint8_t *in, uint64_t len, const uint8_t *dict, uint32_t dict_len, uint8_t *out,
                   uint64_t cap, uint64_t *out_len, uint32_t grid, uint32_t run_chunks) {
  if (len == 0) {
    out[0] = 3;
    out[1] = 0;
    *out_len = 2;
    return 0;
  }
  // bits 16..23 of run_chunks: chunk size in units of 64 KiB (0 = 1), as the host picks it for large inputs;
  // bits 24.. : chunk size in units of 16 KiB (1 or 2), as the host picks it for small inputs (run_chunks = 1 then)
  uint32_t units = (run_chunks >> 16) & 0xffu;
  const uint32_t quarters = run_chunks >> 24;
  run_chunks &= 0xffffu;
  if (units != 2 && units != 4) units = 1;
  uint32_t chunk_bytes = units * (uint32_t)kChunk;
  if (quarters == 1 || quarters == 2) {
    chunk_bytes = quarters * (uint32_t)kMinChunk;
    run_chunks = 1;
  }
  if (run_chunks == 0) run_chunks = 1;
  // as many whole runs of big chunks as fit, the remainder in 64 KiB chunks one by one (the host's DSG_LZ_TAIL=0)
  const uint32_t big_runs = (uint32_t)(len / ((uint64_t)run_chunks * chunk_bytes));
  const uint32_t big_chunks = big_runs * run_chunks;
  const uint32_t nchunks = big_chunks + (uint32_t)((len - (uint64_t)big_chunks * chunk_bytes + kChunk - 1) / kChunk);
  const uint32_t nruns = big_runs + (nchunks - big_chunks);
  if (grid == 0 || grid > nruns) grid = nruns;
  std::vector<uint8_t> slots((size_t)nchunks * lz_slot_bytes(chunk_bytes), 0xAB);
  std::vector<uint32_t> meta(nchunks, 0);
  std::vector<uint64_t> offs(nchunks, 0);
  std::vector<uint32_t> tok((size_t)grid * kTokSlotWords, 0xDEADBEEF);
  uint32_t ticket = 0, error = 0;
  uint64_t total = 0;
  LzParams P;
  P.in = in;
  P.len = len;
  P.dict = dict;
  P.dict_len = dict_len;
  P.chunk_bytes = chunk_bytes;
  P.slots = slots.data();
  P.chunk_meta = meta.data();
  P.nchunks = nchunks;
  P.run_chunks = run_chunks;
  P.big_runs = big_runs;
  P.ticket = &ticket;
  P.tok_scratch = tok.data();
  P.prof = nullptr;
  P.no_bulk = 0;
  P.depth = kDepth;
  P.deep_depth = kDeepDepth;
  P.refine_rounds = kRefineRounds;
  emu::launch(dim3(grid), dim3(kLzThreads), LzLayout::total, [&] { lz_deflate_kernel(P); });
  emu::launch(dim3(1), dim3(1024), 0, [&] { lz_scan_kernel(meta.data(), nchunks, offs.data(), cap, 1, &total, &error); });
  emu::launch(dim3(nchunks), dim3(256), 0, [&] {
    lz_gather_kernel(in, len, slots.data(), meta.data(), offs.data(), nchunks, chunk_bytes, big_chunks, out, 1, &error);
  });
  *out_len = total;
  return error ? DSG_E_CAPACITY : 0;
}

// Mirrors crc_update in dsgpack.cu.
uint32_t emu_crc32(const uint8_t *buf, uint64_t len) {
  std::vector<uint32_t> tab(kCrcTables * 256), shift(kCrcShiftEntries);
  crc_build_tables(tab.data(), shift.data(), (uint64_t)kCrcThreads * 16);
  uint32_t state[4] = {0, 0, 0, 0};
  if (len == 0) return 0;
  uint32_t head = (uint32_t)((16 - ((uintptr_t)buf & 15)) & 15);
  if (head > len) head = (uint32_t)len;
  const uint64_t body = (len - head) & ~uint64_t(15);
  const uint32_t tail = (uint32_t)(len - head - body);
  const uint32_t ntiles = (uint32_t)((body + kCrcTileBytes - 1) / kCrcTileBytes);
  std::vector<uint32_t> raw(ntiles + 1);
  if (ntiles)
    emu::launch(dim3(ntiles), dim3(kCrcThreads), 0, [&] {
      crc_tiles_kernel(reinterpret_cast<const uint4 *>(buf + head), body / 16, tab.data(), shift.data(), raw.data());
    });
  emu::launch(dim3(1), dim3(256), 0,
              [&] { crc_join_kernel(raw.data(), ntiles, body, buf, head, buf + head + body, tail, state); });
  return state[0];
}

// Mirrors file_crc_run in dsgpack.cu: entries = n records of `stride` bytes (dsg_entry layout offsets passed in).
int emu_file_crc32(const uint8_t *entries, uint32_t n, uint32_t stride, uint32_t off_size, uint32_t off_data,
                   uint32_t off_type, const uint8_t *arena, uint64_t arena_len, uint32_t *crcs) {
  if (n == 0) return 0;
  std::vector<uint32_t> tab(kCrcTables * 256), shift(kCrcShiftEntries);
  crc_build_tables(tab.data(), shift.data(), (uint64_t)kCrcThreads * 16);
  const uint32_t nblk = (n + kScanBlock - 1) / kScanBlock;
  std::vector<FileCrcEntry> files(n);
  std::vector<uint32_t> units(n), first(n + 1), totals(nblk + 2, 0);
  uint32_t *grand = totals.data() + nblk, *error = totals.data() + nblk + 1;
  emu::launch(dim3((n + 255) / 256), dim3(256), 0, [&] {
    fcrc_extract_kernel(entries, n, stride, off_size, off_data, off_type, arena_len, files.data(), error);
  });
  emu::launch(dim3((n + 255) / 256), dim3(256), 0,
              [&] { fcrc_count_kernel(files.data(), n, (uint64_t)(uintptr_t)arena, units.data()); });
  emu::launch(dim3(nblk), dim3(kScanBlock), 0,
              [&] { fcrc_scan_blocks_kernel(units.data(), n, first.data(), totals.data()); });
  emu::launch(dim3(1), dim3(kScanBlock), 0, [&] { fcrc_scan_totals_kernel(totals.data(), nblk, grand); });
  emu::launch(dim3(nblk), dim3(kScanBlock), 0, [&] { fcrc_scan_add_kernel(first.data(), n, totals.data(), grand); });
  if (*error) return -(int)*error;
  const uint32_t nunits = *grand;
  std::vector<uint32_t> raw(nunits + 1), unit_file(nunits + 1);
  if (nunits)
    emu::launch(dim3((n + 255) / 256), dim3(256), 0, [&] { fcrc_map_kernel(first.data(), n, unit_file.data()); });
  if (nunits)
    emu::launch(dim3((nunits + kFileCrcUnitsPerCta - 1) / kFileCrcUnitsPerCta), dim3(kCrcThreads), 0, [&] {
      fcrc_units_kernel(files.data(), n, arena, first.data(), unit_file.data(), nunits, tab.data(), shift.data(), raw.data());
    });
  emu::launch(dim3((n + 7) / 8), dim3(256), 0,
              [&] { fcrc_join_kernel(files.data(), n, arena, first.data(), raw.data(), crcs); });
  return 0;
}

// Serial Huffman length construction (huffman.cuh: two-queue merge + zlib gen_bitlen repair), for the Kraft tests.
int emu_huff_lengths(const uint32_t *freq, int n, int max_bits, uint8_t *lens) {
  static HuffScratch S;
  return huff_build_lengths(freq, n, max_bits, lens, S);
}


// Mirrors gunzip_device in dsgpack.cu: candidate scan -> dry decode -> chain -> decode -> stored copy -> tails ->
// bodies. Returns 0, DSG_E_ARG (not a stream of this library), DSG_E_CAPACITY or DSG_E_CORRUPT (second pass differs);
// crcs_ok = every member's CRC32 matches (checked here with the serial table-free recurrence, test infrastructure).
int emu_gunzip(const uint8_t *in, uint64_t n, uint8_t *out, uint64_t cap, uint64_t *out_len, uint32_t *nunits_out,
               uint32_t *nmembers_out, uint32_t *ncand_out, uint32_t rounds, uint64_t batch_bytes) {
  *out_len = 0;
  if (n < 20) return DSG_E_ARG;
  const uint32_t ncta = (uint32_t)((n + kGzScanCta - 1) / kGzScanCta);
  std::vector<uint32_t> counts(ncta, 0), offs(ncta, 0);
  emu::launch(dim3(ncta), dim3(kGzScanThreads), 0, [&] { gz_scan_count(in, n, counts.data()); });
  uint64_t total = 0;
  for (uint32_t i = 0; i < ncta; i++) {
    offs[i] = (uint32_t)total;
    total += counts[i];
  }
  if (total == 0) return DSG_E_ARG;
  std::vector<uint64_t> cand(total);
  emu::launch(dim3(ncta), dim3(kGzScanThreads), 0, [&] { gz_scan_write(in, n, offs.data(), cand.data()); });
  std::sort(cand.begin(), cand.end());
  cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
  const uint32_t ncand = (uint32_t)cand.size();
  if (ncand_out) *ncand_out = ncand;
  std::vector<GzRec> rec(ncand);
  emu::launch(dim3((ncand + kGzWarpsPerCta - 1) / kGzWarpsPerCta), dim3(32 * kGzWarpsPerCta), 0, [&] { gz_measure_kernel(in, n, cand.data(), ncand, rec.data()); });
  std::vector<GzUnit> units(ncand + 16);
  std::vector<GzMember> members(4096);
  GzSummary S;
  emu::launch(dim3(1), dim3(32), 0, [&] {
    gz_chain_kernel(in, n, cand.data(), rec.data(), ncand, units.data(), (uint32_t)units.size(), members.data(),
                    (uint32_t)members.size(), &S);
  });
  if (S.status != DSG_OK) return DSG_E_ARG;
  *out_len = S.total_out;
  if (nunits_out) *nunits_out = S.nunits;
  if (nmembers_out) *nmembers_out = S.nmembers;
  if (S.total_out > cap) return DSG_E_CAPACITY;
  if (S.nunits == 0) return 0;
  // batches of consecutive units, as the host does (batch_bytes == 0: one batch)
  const uint64_t batch_out = batch_bytes ? batch_bytes : S.total_out;
  std::vector<uint16_t> sym(std::min<uint64_t>(S.total_out, batch_out) + kGzMaxUnitOut + 8, 0xEEEE);
  emu::launch(dim3(S.nunits), dim3(256), 0, [&] { gz_stored_copy_kernel(in, units.data(), S.nunits, out); });
  uint32_t u0 = 0;
  while (u0 < S.nunits) {
    const uint64_t base = units[u0].out_off;
    uint32_t u1 = u0 + 1;
    while (u1 < S.nunits && units[u1].out_off + units[u1].out_len - base <= batch_out) u1++;
    const uint32_t nb = u1 - u0;
    const GzUnit *bu = units.data() + u0;
    std::fill(sym.begin(), sym.end(), (uint16_t)0xEEEE);
    uint32_t err = 0;
    emu::launch(dim3((nb + kGzWarpsPerCta - 1) / kGzWarpsPerCta), dim3(32 * kGzWarpsPerCta), 0,
                [&] { gz_decode_kernel(in, n, bu, nb, sym.data(), base, &err); });
    if (err) return DSG_E_CORRUPT;
    std::vector<uint32_t> fix(nb, 7);
    emu::launch(dim3(nb), dim3(256), 0, [&] { gz_tails_plain(in, bu, nb, sym.data(), base, out, fix.data()); });
    // `rounds` parallel rounds (the host's limit is 48; the tests also run with few, so that the serial kernel is reached)
    uint32_t open_units = 1;
    for (uint32_t r = 0; r < rounds && open_units; r++) {
      open_units = 0;
      emu::launch(dim3(nb), dim3(256), 0, [&] { gz_tails_round(bu, nb, sym.data(), base, out, fix.data(), &open_units); });
    }
    if (open_units)
      emu::launch(dim3(1), dim3(1024), 0, [&] { gz_tails_kernel(bu, nb, sym.data(), base, out, fix.data()); });
    emu::launch(dim3(nb), dim3(256), 0, [&] { gz_bodies_kernel(bu, nb, sym.data(), base, out); });
    u0 = u1;
  }
  return 0;
}

}  // extern "C"
