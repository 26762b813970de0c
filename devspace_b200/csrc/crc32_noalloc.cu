// crc32_noalloc.cu -- CRC32 (IEEE) of a device buffer without a single allocation.
//
// Who needs this: a caller that has to checksum bytes already sitting in GPU memory but is not in a position to
// ask the CUDA allocator for anything at that moment -- for example a sync daemon embedded in a process whose
// device memory was handed out completely at start-up (another runtime owns the pool, so a cudaMalloc here
// would fail or evict), or a hot loop that cannot afford the device-wide synchronisation cudaMalloc / cudaFree
// imply or an out-of-memory error in the middle of an upload. The regular entry point (dsg_dev_crc32) keeps its
// per-tile partial remainders in a buffer it grows with cudaMalloc on demand; this one keeps them in a
// file-scope device array that is part of the library image:
//
//     every GPU that loads libdsgpack.so holds these 55 KB (33 KB scratch + 22 KB lookup tables) until the
//     process exits, whether or not this entry point is ever called;
//     the scratch bounds the input to 1 GiB per call (8192 tiles of 128 KiB); longer inputs are rejected;
//     all calls on a device share the one array: it still holds the previous call's contents when the next
//     call starts (the words that must start at zero are cleared at the top of each call), and two calls whose
//     GPU work overlaps on the same device -- from two host threads, say -- race on it. Serialise them.
//
// The first call on each device does a one-time set-up with that device current (resolves the device addresses
// of the arrays, uploads the lookup tables). Make that first call before starting any stream capture.
// The call itself launches on the default stream and returns after the 4-byte result has been copied back.
// It calls no cudaMalloc / cudaFree / cudaHostAlloc and no library routine that allocates.
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "dsg_common.cuh"

// The CRC kernels and their helper functions, compiled a second time under a private namespace so that this
// translation unit can launch them against its own static scratch (dsgpack.cu owns the public instantiation).
namespace dsg_noalloc {
#include "crc32.cuh"
}

namespace {

using namespace dsg_noalloc::dsg;

constexpr uint64_t kMaxBytes = 1ull << 30;                            // 1 GiB per call
constexpr uint32_t kMaxTiles = (uint32_t)(kMaxBytes / kCrcTileBytes); // 8192 per-tile remainders
constexpr uint32_t kStateWords = 64;                                  // [0..2] running CRC state, rest padding

// Scratch: words [0, 64) = running state read and updated by the join kernel (MUST be zero when a call starts),
// words [64, 64 + kMaxTiles) = one raw remainder per 128 KiB tile (every used entry is overwritten before it is
// read). In host code the names below are host-side stand-ins, not device addresses.
static __device__ __align__(256) uint32_t g_scratch[kStateWords + kMaxTiles];
static __device__ __align__(256) uint32_t g_tab[kCrcTables * 256];
static __device__ __align__(256) uint32_t g_shift[kCrcShiftEntries];

struct DeviceState {
  bool ready = false;
  uint32_t *scratch = nullptr, *tab = nullptr, *shift = nullptr;
};
DeviceState g_dev[64];
std::mutex g_mu;

int fail(char *err, size_t err_len, int code, const char *what, cudaError_t e) {
  if (err && err_len) {
    if (e != cudaSuccess) snprintf(err, err_len, "%s: %s", what, cudaGetErrorString(e));
    else snprintf(err, err_len, "%s", what);
  }
  if (e != cudaSuccess) cudaGetLastError();
  return code;
}

// Once per device, with that device current: device addresses of the arrays + table upload.
int setup_device(int dev, DeviceState **out, char *err, size_t err_len) {
  std::lock_guard<std::mutex> g(g_mu);
  DeviceState &d = g_dev[dev];
  if (!d.ready) {
    cudaError_t e;
    if ((e = cudaGetSymbolAddress((void **)&d.scratch, g_scratch)) != cudaSuccess)
      return fail(err, err_len, DSG_E_CUDA, "cudaGetSymbolAddress(scratch)", e);
    if ((e = cudaGetSymbolAddress((void **)&d.tab, g_tab)) != cudaSuccess)
      return fail(err, err_len, DSG_E_CUDA, "cudaGetSymbolAddress(tables)", e);
    if ((e = cudaGetSymbolAddress((void **)&d.shift, g_shift)) != cudaSuccess)
      return fail(err, err_len, DSG_E_CUDA, "cudaGetSymbolAddress(shift table)", e);
    static uint32_t h_tab[kCrcTables * 256], h_shift[kCrcShiftEntries];  // no heap either
    crc_build_tables(h_tab, h_shift, (uint64_t)kCrcThreads * 16);
    if ((e = cudaMemcpy(d.tab, h_tab, sizeof h_tab, cudaMemcpyHostToDevice)) != cudaSuccess)
      return fail(err, err_len, DSG_E_CUDA, "table upload", e);
    if ((e = cudaMemcpy(d.shift, h_shift, sizeof h_shift, cudaMemcpyHostToDevice)) != cudaSuccess)
      return fail(err, err_len, DSG_E_CUDA, "shift table upload", e);
    d.ready = true;
  }
  *out = &d;
  return DSG_OK;
}

}  // namespace

extern "C" int dsg_crc32_noalloc(const uint8_t *d_buf, size_t len, uint32_t *crc, char *err, size_t err_len) {
  if (!crc || (len && !d_buf)) return fail(err, err_len, DSG_E_ARG, "dsg_crc32_noalloc: null pointer", cudaSuccess);
  if (len > kMaxBytes) {
    if (err && err_len)
      snprintf(err, err_len, "dsg_crc32_noalloc: %zu bytes exceed the %llu-byte limit of the static scratch "
               "(use dsg_dev_crc32, which allocates, or split the buffer)", len, (unsigned long long)kMaxBytes);
    return DSG_E_CAPACITY;
  }
  if (len == 0) {
    *crc = 0;
    return DSG_OK;
  }
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(err, err_len, DSG_E_CUDA, "cudaGetDevice", e);
  if (dev < 0 || dev >= 64) return fail(err, err_len, DSG_E_ARG, "dsg_crc32_noalloc: device index out of range", cudaSuccess);
  DeviceState *d = nullptr;
  int rc = setup_device(dev, &d, err, err_len);  // first call on this device: before any stream capture
  if (rc != DSG_OK) return rc;

  // The running state must start at zero; the per-tile words need no clearing (written before they are read).
  if ((e = cudaMemsetAsync(d->scratch, 0, kStateWords * sizeof(uint32_t), 0)) != cudaSuccess)
    return fail(err, err_len, DSG_E_CUDA, "cudaMemsetAsync(state)", e);
  uint32_t head = (uint32_t)((16 - ((uintptr_t)d_buf & 15)) & 15);
  if (head > len) head = (uint32_t)len;
  const uint64_t body = (len - head) & ~uint64_t(15);
  const uint32_t tail = (uint32_t)(len - head - body);
  const uint32_t ntiles = (uint32_t)((body + kCrcTileBytes - 1) / kCrcTileBytes);
  uint32_t *raw = d->scratch + kStateWords;
  if (ntiles) {
    crc_tiles_kernel<<<ntiles, kCrcThreads, 0, 0>>>(reinterpret_cast<const uint4 *>(d_buf + head), body / 16, d->tab,
                                                    d->shift, raw);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, err_len, DSG_E_CUDA, "crc_tiles_kernel launch", e);
  }
  crc_join_kernel<<<1, 256, 0, 0>>>(raw, ntiles, body, d_buf, head, d_buf + head + body, tail, d->scratch);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, err_len, DSG_E_CUDA, "crc_join_kernel launch", e);
  if ((e = cudaMemcpy(crc, d->scratch, sizeof(uint32_t), cudaMemcpyDeviceToHost)) != cudaSuccess)
    return fail(err, err_len, DSG_E_CUDA, "result copy", e);
  return DSG_OK;
}
