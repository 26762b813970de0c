// dsgpack.cu -- host engine and C ABI of libdsgpack.so (include/dsgpack.h).
//
// Replaces the body of writeTar below the directory walk
// (/root/reference/pkg/devspace/sync/tar.go:146-182: gzip.NewWriter -> tar.NewWriter -> per entry
// WriteHeader + io.Copy -> tarWriter.Close -> gw.Close) with CUDA kernels for sm_100a.
// No CPU fallback: every compute entry point needs a CUDA device and fails with DSG_E_CUDA otherwise.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <errno.h>
#include <unistd.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "crc32.cuh"
#include "deflate_lz.cuh"
#include "file_crc.cuh"
#include "inflate.cuh"
#include "deflate_stored.cuh"
#include "tar_kernels.cuh"

using namespace dsg;

namespace {

struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
      cudaGetLastError();
      e = cudaMalloc(&p, bytes + 256);
      want = bytes + 256;
    }
    if (e == cudaSuccess) cap = want - 256;  // 256 bytes of slack behind every buffer (aligned over-reads)
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct Timer {
  cudaEvent_t a = nullptr, b = nullptr;
  bool used = false;
};

enum { kTimPlan, kTimTar, kTimDeflate, kTimCrc, kTimGather, kTimTotal, kTimCount };

}  // namespace

struct dsg_ctx {
  int device = 0;
  dsg_opts opts{};
  cudaStream_t stream = nullptr;      // compute (and everything in the single-stream entry points)
  cudaStream_t stream_in = nullptr;   // host -> device copies of the pipelined path
  cudaStream_t stream_out = nullptr;  // device -> host copies of the pipelined path
  std::string err;
  dsg_stats stats{};
  std::mutex mu;
  // constant tables
  DevBuf crc_tab, crc_shift;
  // plan
  DevBuf ent_blocks, cta_sums, ent_off, scalars;  // scalars: [0] total_blocks u64, [1] payload u64, [2..] status, crc state
  // staging of host inputs / outputs
  DevBuf d_entries, d_strtab, d_arena, d_tar, d_out;
  // crc
  DevBuf crc_raw;
  // per-file crc (context hash)
  DevBuf fc_files, fc_units, fc_first, fc_totals, fc_raw, fc_crcs;
  // deflate
  DevBuf lz_scratch, chunk_sizes, chunk_slots, chunk_offs, lz_prof;
  // pipelined path: two slots each for the arena slice, the tar segment and the compressed segment
  DevBuf slot_arena[2], slot_tar[2], slot_out[2], seg_ranges;
  // gunzip (inflate.cuh)
  DevBuf gz_counts, gz_cand, gz_rec, gz_units, gz_members, gz_sym, gz_misc, gz_in, gz_out;
  Timer tim[kTimCount];
  uint64_t *h_scalars = nullptr;  // pinned mirror of `scalars`
  uint8_t *h_stage[2] = {nullptr, nullptr};  // pinned staging of the file-descriptor sink (dsg_pack_fd)
  size_t h_stage_cap = 0;
  uint32_t launches = 0;
  int sm_count = 148;
};

namespace {

constexpr size_t kScalarWords = 16;   // u64 slots
constexpr size_t kMaxSegSizes = 64;   // pinned ring of per-segment output sizes behind the scalar mirror
// slot layout inside `scalars` (u64 units)
enum { kSTotalBlocks = 0, kSPayload = 1, kSStatus = 2 /* DevStatus, 8 bytes */, kSCrc = 3 /* 3 x u32 in 2 slots */,
       kSDeflateBytes = 5, kSSpare = 6 };

int fail(dsg_ctx *c, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (c) c->err = buf;
  return code;
}

#define CU(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess) {                                                                             \
      cudaGetLastError();                                                                                \
      return fail(ctx, e_ == cudaErrorMemoryAllocation ? DSG_E_NOMEM : DSG_E_CUDA, "%s: %s (%s:%d)", #call, \
                  cudaGetErrorString(e_), __FILE__, __LINE__);                                           \
    }                                                                                                    \
  } while (0)

#define LAUNCH_CHECK()                                                      \
  do {                                                                      \
    ctx->launches++;                                                        \
    CU(cudaGetLastError());                                                 \
  } while (0)

int lz_ctx_init(dsg_ctx *ctx);
int lz_deflate_run(dsg_ctx *ctx, const uint8_t *d_in, uint64_t len, uint8_t *d_out, uint64_t out_cap, uint64_t *out_len);

void tic(dsg_ctx *ctx, int which) {
  Timer &t = ctx->tim[which];
  cudaEventRecord(t.a, ctx->stream);
}
void toc(dsg_ctx *ctx, int which) {
  Timer &t = ctx->tim[which];
  cudaEventRecord(t.b, ctx->stream);
  t.used = true;
}
float elapsed(dsg_ctx *ctx, int which) {
  Timer &t = ctx->tim[which];
  if (!t.used) return 0.f;
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, t.a, t.b) != cudaSuccess) {
    cudaGetLastError();
    return 0.f;
  }
  return ms;
}

int begin_call(dsg_ctx *ctx) {
  CU(cudaSetDevice(ctx->device));
  ctx->launches = 0;
  memset(&ctx->stats, 0, sizeof ctx->stats);
  for (auto &t : ctx->tim) t.used = false;
  ctx->err.clear();
  return DSG_OK;
}

// ---- stage: archive/tar Writer ---------------------------------------------------------------------------

// Enqueues the plan kernels (no synchronisation): per-entry stream offsets in ent_off, totals and the status word in
// `scalars` on the device.
int tar_plan_enqueue(dsg_ctx *ctx, const dsg_entry *d_entries, uint32_t n, const char *d_strtab, uint64_t strtab_len,
                     uint64_t arena_len) {
  const uint32_t ncta = (n + kPlanThreads - 1) / kPlanThreads;
  CU(ctx->ent_blocks.ensure(sizeof(uint64_t) * (size_t)(n + 1)));
  CU(ctx->cta_sums.ensure(sizeof(uint64_t) * (size_t)(ncta + 1)));
  CU(ctx->ent_off.ensure(sizeof(uint64_t) * (size_t)(n + 1)));
  uint64_t *sc = ctx->scalars.as<uint64_t>();
  memset(ctx->h_scalars, 0, sizeof(uint64_t) * kScalarWords);
  DevStatus st0{0, 0xffffffffu};
  memcpy(&ctx->h_scalars[kSStatus], &st0, sizeof st0);
  CU(cudaMemcpyAsync(sc, ctx->h_scalars, sizeof(uint64_t) * kScalarWords, cudaMemcpyHostToDevice, ctx->stream));
  tic(ctx, kTimPlan);
  if (n) {
    tar_classify_kernel<<<ncta, kPlanThreads, 0, ctx->stream>>>(
        d_entries, n, d_strtab, strtab_len, arena_len, ctx->ent_blocks.as<uint64_t>(), ctx->cta_sums.as<uint64_t>(),
        reinterpret_cast<unsigned long long *>(sc + kSPayload), reinterpret_cast<DevStatus *>(sc + kSStatus));
    LAUNCH_CHECK();
    tar_scan_partials<<<1, 1024, 0, ctx->stream>>>(ctx->cta_sums.as<uint64_t>(), ncta, sc + kSTotalBlocks);
    LAUNCH_CHECK();
    tar_offsets_kernel<<<ncta, kPlanThreads, 0, ctx->stream>>>(ctx->ent_blocks.as<uint64_t>(), n,
                                                              ctx->cta_sums.as<uint64_t>(), ctx->ent_off.as<uint64_t>());
    LAUNCH_CHECK();
  } else {
    CU(cudaMemsetAsync(ctx->ent_off.p, 0, sizeof(uint64_t), ctx->stream));
  }
  toc(ctx, kTimPlan);
  return DSG_OK;
}

// Runs the plan kernels and returns T (synchronises once to read it and the status word).
int tar_plan(dsg_ctx *ctx, const dsg_entry *d_entries, uint32_t n, const char *d_strtab, uint64_t strtab_len,
             uint64_t arena_len, uint32_t flags, uint64_t *tar_len, uint64_t *payload) {
  int rc = tar_plan_enqueue(ctx, d_entries, n, d_strtab, strtab_len, arena_len);
  if (rc) return rc;
  uint64_t *sc = ctx->scalars.as<uint64_t>();
  CU(cudaMemcpyAsync(ctx->h_scalars, sc, sizeof(uint64_t) * kScalarWords, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  DevStatus st;
  memcpy(&st, &ctx->h_scalars[kSStatus], sizeof st);
  if (st.code != 0) return fail(ctx, st.code, "entry %u: %s", st.entry, dsg_strerror(st.code));
  *tar_len = ctx->h_scalars[kSTotalBlocks] * kBlock + ((flags & DSG_F_NO_TRAILER) ? 0 : 2 * kBlock);
  if (payload) *payload = ctx->h_scalars[kSPayload];
  return DSG_OK;
}

// Writes stream bytes [lo, hi) (multiples of 512) to d_dst; d_arena holds arena bytes from arena_base on.
int tar_layout(dsg_ctx *ctx, const dsg_entry *d_entries, uint32_t n, const char *d_strtab, const uint8_t *d_arena,
               uint64_t arena_base, uint8_t *d_dst, uint64_t lo, uint64_t hi) {
  if (hi <= lo) return DSG_OK;
  const uint64_t tiles = (hi - lo + kTileBytes - 1) / kTileBytes;
  uint64_t *sc = ctx->scalars.as<uint64_t>();
  tar_layout_kernel<<<(unsigned)tiles, kLayoutThreads, 0, ctx->stream>>>(
      d_entries, n, d_strtab, ctx->ent_off.as<uint64_t>(), d_arena, arena_base, d_dst, lo, hi,
      reinterpret_cast<DevStatus *>(sc + kSStatus));
  LAUNCH_CHECK();
  return DSG_OK;
}

int check_status(dsg_ctx *ctx) {
  uint64_t *sc = ctx->scalars.as<uint64_t>();
  CU(cudaMemcpyAsync(ctx->h_scalars, sc, sizeof(uint64_t) * kScalarWords, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  DevStatus st;
  memcpy(&st, &ctx->h_scalars[kSStatus], sizeof st);
  if (st.code != 0) return fail(ctx, st.code, "entry %u: %s", st.entry, dsg_strerror(st.code));
  return DSG_OK;
}

// ---- stage: CRC32 -------------------------------------------------------------------------------------------

// Folds d_buf[0,len) into the running CRC state kept in scalars[kSCrc..] (device resident, no sync).
int crc_update(dsg_ctx *ctx, const uint8_t *d_buf, uint64_t len) {
  if (len == 0) return DSG_OK;
  uint32_t head = (uint32_t)((16 - ((uintptr_t)d_buf & 15)) & 15);
  if (head > len) head = (uint32_t)len;
  const uint64_t body = (len - head) & ~uint64_t(15);
  const uint32_t tail = (uint32_t)(len - head - body);
  const uint32_t ntiles = (uint32_t)((body + kCrcTileBytes - 1) / kCrcTileBytes);
  CU(ctx->crc_raw.ensure(sizeof(uint32_t) * (size_t)(ntiles + 1)));
  uint32_t *state = reinterpret_cast<uint32_t *>(ctx->scalars.as<uint64_t>() + kSCrc);
  if (ntiles) {
    crc_tiles_kernel<<<ntiles, kCrcThreads, 0, ctx->stream>>>(reinterpret_cast<const uint4 *>(d_buf + head), body / 16,
                                                            ctx->crc_tab.as<uint32_t>(), ctx->crc_shift.as<uint32_t>(),
                                                            ctx->crc_raw.as<uint32_t>());
    LAUNCH_CHECK();
  }
  crc_join_kernel<<<1, 256, 0, ctx->stream>>>(ctx->crc_raw.as<uint32_t>(), ntiles, body, d_buf, head,
                                              d_buf + head + body, tail, state);
  LAUNCH_CHECK();
  return DSG_OK;
}

int crc_reset(dsg_ctx *ctx) {
  CU(cudaMemsetAsync(ctx->scalars.as<uint64_t>() + kSCrc, 0, 2 * sizeof(uint64_t), ctx->stream));
  return DSG_OK;
}

// ---- stage: deflate -----------------------------------------------------------------------------------------

// out_cap >= 18 (checked by the caller). *deflate_bytes is the NEEDED size when the deflate stage found the output
// too small (*error != 0, nothing was written): the trailer then lies beyond the buffer and must not be written.
__global__ void gzip_frame_kernel(uint8_t *out, uint64_t out_cap, const uint64_t *deflate_bytes,
                                  const uint32_t *crc_state, const uint32_t *error) {
  // compress/gzip Writer: 1f 8b 08 00 | mtime 0 | XFL 0 | OS ff ... CRC32 LE, ISIZE LE
  if (threadIdx.x == 0) {
    const uint8_t hdr[10] = {0x1f, 0x8b, 0x08, 0, 0, 0, 0, 0, 0, 0xff};
    for (int i = 0; i < 10; i++) out[i] = hdr[i];
    if (*error != 0 || *deflate_bytes > out_cap - 18) return;
    uint8_t *t = out + 10 + *deflate_bytes;
    const uint32_t crc = crc_state[0], isize = crc_state[1];
    for (int i = 0; i < 4; i++) t[i] = (uint8_t)(crc >> (8 * i));
    for (int i = 0; i < 4; i++) t[4 + i] = (uint8_t)(isize >> (8 * i));
  }
}

__global__ void set_u64_kernel(uint64_t *p, uint64_t v) { *p = v; }

// Raw deflate of d_in[0,len) to d_out (capacity out_cap). Size lands in scalars[kSDeflateBytes] on the
// device and, via *out_len (after a sync), on the host.
int deflate_run(dsg_ctx *ctx, const uint8_t *d_in, uint64_t len, int level, uint8_t *d_out, uint64_t out_cap,
                uint64_t *out_len) {
  uint64_t *sc = ctx->scalars.as<uint64_t>();
  tic(ctx, kTimDeflate);
  if (level == DSG_LEVEL_STORED) {
    const uint64_t need = stored_deflate_size(len);
    if (need > out_cap) {
      *out_len = need;
      return fail(ctx, DSG_E_CAPACITY, "deflate output needs %llu bytes, capacity %llu", (unsigned long long)need,
                  (unsigned long long)out_cap);
    }
    const uint64_t nblocks = (len + kStoredMax - 1) / kStoredMax;
    deflate_stored_kernel<<<(unsigned)std::max<uint64_t>(nblocks, 1), kStoredThreads, 0, ctx->stream>>>(d_in, len, d_out, 1);
    LAUNCH_CHECK();
    set_u64_kernel<<<1, 1, 0, ctx->stream>>>(sc + kSDeflateBytes, need);
    LAUNCH_CHECK();
    toc(ctx, kTimDeflate);
    *out_len = need;
    return DSG_OK;
  }
  int rc = lz_deflate_run(ctx, d_in, len, d_out, out_cap, out_len);
  toc(ctx, kTimDeflate);
  return rc;
}

// ---- LZ77 + Huffman path (level 6) -------------------------------------------------------------------------

int lz_ctx_init(dsg_ctx *ctx) {
  cudaError_t e = cudaFuncSetAttribute(lz_deflate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LzLayout::total);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(ctx, DSG_E_CUDA, "cudaFuncSetAttribute(lz_deflate_kernel, %d bytes of shared memory): %s",
                (int)LzLayout::total, cudaGetErrorString(e));
  }
  return DSG_OK;
}

__global__ void lz_finish_empty_kernel(uint8_t *out, uint64_t *total_out, int write_terminator) {
  if (threadIdx.x == 0) {
    if (write_terminator) {
      out[0] = 0x03;
      out[1] = 0x00;
    }
    *total_out = write_terminator ? 2 : 0;
  }
}

// Launch geometry of one level-6 launch over `len` bytes: the one place that decides it (the launch itself and the
// code that pre-sizes the scratch buffers of the pipelined path both ask here).
struct LzGeometry {
  uint32_t chunk_bytes, run_chunks, nchunks, big_runs, big_chunks, grid;
  size_t slot_bytes;  // scratch for the per-chunk output slots
};
bool lz_geometry(const dsg_ctx *ctx, uint64_t len, LzGeometry &G) {
  const dsg_opts &o = ctx->opts;
  // Chunk size: the largest of 64 / 128 / 256 KiB that still leaves every SM eight chunks (see kMaxChunk).
  uint32_t chunk_bytes = kChunk;
  while (chunk_bytes < (uint32_t)kMaxChunk && len / (2ull * chunk_bytes) >= 8ull * (uint64_t)ctx->sm_count) chunk_bytes *= 2;
  // Small inputs: a call lasts as long as its slowest SM, i.e. a whole number of chunk times (1.2 ms per 64 KiB chunk,
  // 0.6 ms per 32 KiB one). While the input has no more 32 KiB chunks than there are SMs, halving the chunks halves the
  // call (measured on the B200: a 4 MB tree 73 x -> 110 x the CPU path, +0.4 % output on token-like text: each chunk
  // carries its own Huffman tables and sync marker); beyond that it only adds a second wave of chunks and dictionary
  // builds (16 MB: slower), so 64 KiB chunks stay. Below 128 KiB -- a change set of a few files, where the call is
  // latency and 2 % of the output is a few hundred bytes -- the chunks are 16 KiB.
  if (len <= (uint64_t)(2 * kMinChunk) * (uint64_t)ctx->sm_count) chunk_bytes = len <= (128u << 10) ? kMinChunk : 2 * kMinChunk;
  if (o.lz_chunk_units == 1 || o.lz_chunk_units == 2 || o.lz_chunk_units == 4) chunk_bytes = o.lz_chunk_units * kChunk;
  // Runs of run_chunks big chunks first (long runs amortise hashing the dictionary), short enough that every SM takes
  // five or more of them: the launch ends when the slowest SM does, and with two or three runs per SM the ends lie a
  // good part of a run apart whatever follows. The last stretch of the input -- half a run per SM, a quarter for short
  // runs -- is cut into 64 KiB chunks and handed out one by one, so that the CTAs finish within a millisecond of each
  // other. Measured on the bench generator (tools/sweep_geometry.py, profiles/r02_geometry_sweep.jsonl): one 512 MiB
  // pipeline segment (50.1 ms at the rate of one long launch) in runs of 3 x 256 KiB + 58 MB of small chunks (the rule
  // until then) 52.6 ms, runs of 2 + 19 MB 50.0 ms, runs of 1 + 9 MB 50.3 ms, runs of 4 or 6 with less than 74 MB of
  // small chunks 56 .. 63 ms; 96 MiB in 64 KiB chunks: runs of 1 12.05 ms, of 2 11.6, of 3 11.5, of 4 12.3.
  const uint64_t big_total = (len + chunk_bytes - 1) / chunk_bytes;
  uint32_t run_chunks = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(8, big_total / (5 * (uint64_t)ctx->sm_count)));
  if (o.lz_run_chunks > 0 && o.lz_run_chunks <= 64) run_chunks = o.lz_run_chunks;
  uint64_t tail_bytes = run_chunks > 1 || chunk_bytes > (uint32_t)kChunk
                            ? std::min<uint64_t>(len, (uint64_t)ctx->sm_count * run_chunks * chunk_bytes / (run_chunks <= 2 ? 4 : 2))
                            : len;
  if (chunk_bytes < (uint32_t)kChunk) {  // every full chunk is a "big" one; the remainder is one (short) tail chunk
    run_chunks = 1;
    tail_bytes = 0;
  }
  if (o.lz_flags & DSG_LZ_TAIL_SET) tail_bytes = std::min<uint64_t>(len, o.lz_tail_bytes);
  const uint64_t big_runs64 = (len - tail_bytes) / ((uint64_t)run_chunks * chunk_bytes);
  const uint64_t big_chunks64 = big_runs64 * run_chunks;
  const uint64_t nchunks64 = big_chunks64 + (len - big_chunks64 * chunk_bytes + kChunk - 1) / kChunk;
  if (nchunks64 > 0xfffffff0ull) return false;
  G.chunk_bytes = chunk_bytes;
  G.run_chunks = run_chunks;
  G.nchunks = (uint32_t)nchunks64;
  G.big_runs = (uint32_t)big_runs64;
  G.big_chunks = (uint32_t)big_chunks64;
  const uint32_t nruns = G.big_runs + (G.nchunks - G.big_chunks);
  G.grid = std::min<uint32_t>(std::max<uint32_t>(nruns, 1), (uint32_t)ctx->sm_count);
  if (o.lz_grid > 0 && o.lz_grid < G.grid) G.grid = o.lz_grid;
  G.slot_bytes = (size_t)lz_slot_bytes(chunk_bytes) * G.nchunks;  // every chunk, also a 64 KiB one, has a big slot
  return true;
}

// Level-6 deflate of d_in[0,len) on `stream`: one persistent kernel. d_dict/dict_len: up to 32 KiB of the
// member that precede d_in (previous pipeline segment). The produced size is written to *d_total (device),
// a too-small d_out sets *d_error (device) to 1.
int lz_launch(dsg_ctx *ctx, cudaStream_t stream, const uint8_t *d_in, uint64_t len, const uint8_t *d_dict,
              uint32_t dict_len, uint8_t *d_out, uint64_t out_cap, uint64_t *d_total, uint32_t *d_error,
              int write_terminator) {
  DevBuf &b_sizes = ctx->chunk_sizes;
  DevBuf &b_slots = ctx->chunk_slots;
  DevBuf &b_tok = ctx->lz_scratch;
  if (len == 0) {
    if (out_cap < 2 && write_terminator) return fail(ctx, DSG_E_CAPACITY, "deflate output needs 2 bytes");
    lz_finish_empty_kernel<<<1, 32, 0, stream>>>(d_out, d_total, write_terminator);
    LAUNCH_CHECK();
    return DSG_OK;
  }
  LzGeometry G;
  if (!lz_geometry(ctx, len, G)) return fail(ctx, DSG_E_ARG, "input too large for one call");
  const uint32_t chunk_bytes = G.chunk_bytes, run_chunks = G.run_chunks, nchunks = G.nchunks, big_runs = G.big_runs,
                 big_chunks = G.big_chunks, grid = G.grid;
  // scratch: per-chunk output slots + sizes + offsets, the ticket word, token slots (one per resident CTA)
  const size_t meta_bytes = (sizeof(uint64_t) + sizeof(uint32_t)) * (size_t)nchunks + 64;
  CU(b_sizes.ensure(meta_bytes));
  CU(b_slots.ensure(G.slot_bytes));
  CU(b_tok.ensure(sizeof(uint32_t) * (size_t)kTokSlotWords * ctx->sm_count));
  uint64_t *offs = b_sizes.as<uint64_t>();
  uint32_t *meta = reinterpret_cast<uint32_t *>(offs + nchunks);
  uint32_t *ticket = meta + nchunks;
  CU(cudaMemsetAsync(ticket, 0, sizeof(uint32_t), stream));
  LzParams P;
  P.in = d_in;
  P.len = len;
  P.dict = dict_len ? d_dict : nullptr;
  P.dict_len = dict_len;
  P.chunk_bytes = chunk_bytes;
  P.slots = b_slots.as<uint8_t>();
  P.chunk_meta = meta;
  P.nchunks = nchunks;
  P.run_chunks = run_chunks;
  P.big_runs = big_runs;
  P.ticket = ticket;
  P.tok_scratch = b_tok.as<uint32_t>();
  P.prof = nullptr;
  P.no_bulk = (ctx->opts.lz_flags & DSG_LZ_NO_BULK) ? 1u : 0u;
  {  // search effort: the defaults, unless a profiling tool passes others in the upper bits of lz_flags
    const uint32_t f = ctx->opts.lz_flags;
    P.deep_depth = ((f >> 8) & 0xffu) ? 4u * ((f >> 8) & 0xffu) : (uint32_t)kDeepDepth;
    P.refine_rounds = ((f >> 16) & 3u) == 0 ? (uint32_t)kRefineRounds : (((f >> 16) & 3u) == 3u ? 0u : ((f >> 16) & 3u));
    P.depth = ((f >> 18) & 15u) ? ((f >> 18) & 15u) : (uint32_t)kDepth;
  }
  const bool prof_on = (ctx->opts.lz_flags & DSG_LZ_PROFILE) != 0;  // development aid: per-phase cycle counts on stderr
  if (prof_on) {
    CU(ctx->lz_prof.ensure(sizeof(unsigned long long) * kProfSlots));
    CU(cudaMemsetAsync(ctx->lz_prof.p, 0, sizeof(unsigned long long) * kProfSlots, stream));
    P.prof = ctx->lz_prof.as<unsigned long long>();
  }
  lz_deflate_kernel<<<grid, kLzThreads, LzLayout::total, stream>>>(P);
  LAUNCH_CHECK();
  lz_scan_kernel<<<1, 1024, 0, stream>>>(meta, nchunks, offs, out_cap, write_terminator, d_total, d_error);
  LAUNCH_CHECK();
  lz_gather_kernel<<<nchunks, 256, 0, stream>>>(d_in, len, P.slots, meta, offs, nchunks, chunk_bytes, big_chunks,
                                                d_out, write_terminator, d_error);
  LAUNCH_CHECK();
  if (prof_on) {
    unsigned long long h[kProfSlots];
    CU(cudaMemcpyAsync(h, P.prof, sizeof h, cudaMemcpyDeviceToHost, stream));
    CU(cudaStreamSynchronize(stream));
    static const char *names[kProfCount] = {"load", "insert", "search", "parse", "emit", "huffman", "sizes", "pack",
                                              "ref_parse", "ref_list", "ref_deep", "wait_a", "wait_b", "join", "link", "flat"};
    unsigned long long tot = 0;
    for (int i = 0; i < kProfCount; i++) tot += h[i];
    fprintf(stderr, "[lz prof] chunks=%u grid=%u cycles/chunk:", nchunks, grid);
    for (int i = 0; i < kProfCount; i++) fprintf(stderr, " %s=%llu", names[i], h[i] / nchunks);
    fprintf(stderr, " total=%llu (mean warp finish: search=%llu ref_deep=%llu)\n", tot / nchunks,
            h[kProfWarpSearch] / nchunks, h[kProfWarpDeep] / nchunks);
  }
  return DSG_OK;
}

// Whole-buffer variant on the context's main stream: size -> scalars[kSDeflateBytes], error -> scalars[kSSpare]
// (both read back with the other scalars after the caller's synchronisation).
int lz_deflate_run(dsg_ctx *ctx, const uint8_t *d_in, uint64_t len, uint8_t *d_out, uint64_t out_cap,
                   uint64_t *out_len) {
  uint64_t *sc = ctx->scalars.as<uint64_t>();
  *out_len = lz_deflate_bound(len);
  CU(cudaMemsetAsync(sc + kSSpare, 0, sizeof(uint64_t), ctx->stream));
  return lz_launch(ctx, ctx->stream, d_in, len, nullptr, 0, d_out, out_cap, sc + kSDeflateBytes,
                   reinterpret_cast<uint32_t *>(sc + kSSpare), 1);
}

}  // namespace

// =================================================================================================================
// C ABI
// =================================================================================================================

extern "C" {

int dsg_abi_version(void) { return DSG_ABI_VERSION; }

const char *dsg_strerror(int status) {
  switch (status) {
    case DSG_OK: return "ok";
    case DSG_E_ARG: return "invalid argument";
    case DSG_E_CUDA: return "CUDA error";
    case DSG_E_HEADER: return "archive/tar: header field cannot be encoded";
    case DSG_E_TYPEFLAG: return "typeflag is none of '0' (file), '5' (directory), '1' (hard link), '2' (symlink)";
    case DSG_E_CAPACITY: return "output buffer too small";
    case DSG_E_UNCLEAN_NAME: return "PAX entry name is not path.Clean()ed";
    case DSG_E_NOMEM: return "out of device or pinned memory";
    case DSG_E_RANGE: return "payload range outside the arena";
    case DSG_E_CORRUPT: return "gzip stream fails its structure or CRC32 check";
    default: return "unknown status";
  }
}

int dsg_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

const char *dsg_last_error(const dsg_ctx *ctx) { return ctx ? ctx->err.c_str() : "no context"; }

int dsg_ctx_create(int device, const dsg_opts *opts, dsg_ctx **out) {
  if (!out) return DSG_E_ARG;
  *out = nullptr;
  int ndev = dsg_device_count();
  if (ndev <= 0 || device < 0 || device >= ndev) return DSG_E_CUDA;  // no CPU fallback
  dsg_ctx *ctx = new dsg_ctx();
  ctx->device = device;
  ctx->opts.struct_size = sizeof(dsg_opts);
  ctx->opts.level = DSG_LEVEL_DEFAULT;
  ctx->opts.segment_bytes = 0;
  if (opts) {
    size_t k = std::min<size_t>(opts->struct_size, sizeof(dsg_opts));
    memcpy(&ctx->opts, opts, k);
    ctx->opts.struct_size = sizeof(dsg_opts);
  }
  auto bail = [&](cudaError_t e, const char *what) {
    fprintf(stderr, "dsg_ctx_create: %s: %s\n", what, cudaGetErrorString(e));
    cudaGetLastError();
    dsg_ctx_destroy(ctx);
    return DSG_E_CUDA;
  };
  cudaError_t e;
  if ((e = cudaSetDevice(device)) != cudaSuccess) return bail(e, "cudaSetDevice");
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail(e, "cudaGetDeviceProperties");
  ctx->sm_count = prop.multiProcessorCount;
  if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "stream");
  for (auto &t : ctx->tim) {
    if ((e = cudaEventCreate(&t.a)) != cudaSuccess) return bail(e, "event");
    if ((e = cudaEventCreate(&t.b)) != cudaSuccess) return bail(e, "event");
  }
  if ((e = cudaStreamCreateWithFlags(&ctx->stream_in, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "stream");
  if ((e = cudaStreamCreateWithFlags(&ctx->stream_out, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "stream");
  if ((e = cudaMallocHost(reinterpret_cast<void **>(&ctx->h_scalars),
                          sizeof(uint64_t) * (kScalarWords + kMaxSegSizes))) != cudaSuccess)
    return bail(e, "cudaMallocHost");
  if ((e = ctx->scalars.ensure(sizeof(uint64_t) * kScalarWords)) != cudaSuccess) return bail(e, "scalars");
  // CRC tables: built once on the host (2 x 5120 table entries), resident for the context's lifetime
  std::vector<uint32_t> tab(kCrcTables * 256), shift(kCrcShiftEntries);
  crc_build_tables(tab.data(), shift.data(), (uint64_t)kCrcThreads * 16);
  if ((e = ctx->crc_tab.ensure(tab.size() * 4)) != cudaSuccess) return bail(e, "crc_tab");
  if ((e = ctx->crc_shift.ensure(shift.size() * 4)) != cudaSuccess) return bail(e, "crc_shift");
  if ((e = cudaMemcpy(ctx->crc_tab.p, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess)
    return bail(e, "crc_tab upload");
  if ((e = cudaMemcpy(ctx->crc_shift.p, shift.data(), shift.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess)
    return bail(e, "crc_shift upload");
  int rc = lz_ctx_init(ctx);
  if (rc != DSG_OK) {
    fprintf(stderr, "dsg_ctx_create: %s\n", ctx->err.c_str());
    dsg_ctx_destroy(ctx);
    return rc;
  }
  *out = ctx;
  return DSG_OK;
}

int dsg_ctx_set_opts(dsg_ctx *ctx, const dsg_opts *opts) {
  if (!ctx || !opts) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  dsg_opts o{};
  memcpy(&o, opts, std::min<size_t>(opts->struct_size, sizeof(dsg_opts)));
  if (o.level != DSG_LEVEL_STORED && o.level != DSG_LEVEL_DEFAULT) return fail(ctx, DSG_E_ARG, "level must be 0 or 6");
  o.struct_size = sizeof(dsg_opts);
  ctx->opts = o;
  return DSG_OK;
}

void dsg_ctx_destroy(dsg_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  for (DevBuf *b : {&ctx->crc_tab, &ctx->crc_shift, &ctx->ent_blocks, &ctx->cta_sums, &ctx->ent_off, &ctx->scalars,
                    &ctx->d_entries, &ctx->d_strtab, &ctx->d_arena, &ctx->d_tar, &ctx->d_out, &ctx->crc_raw,
                    &ctx->fc_files, &ctx->fc_units, &ctx->fc_first, &ctx->fc_totals, &ctx->fc_raw, &ctx->fc_crcs,
                    &ctx->gz_counts, &ctx->gz_cand, &ctx->gz_rec, &ctx->gz_units, &ctx->gz_members, &ctx->gz_sym,
                    &ctx->gz_misc, &ctx->gz_in, &ctx->gz_out, &ctx->lz_scratch, &ctx->chunk_sizes, &ctx->chunk_slots, &ctx->chunk_offs, &ctx->lz_prof, &ctx->seg_ranges, &ctx->slot_arena[0],
                    &ctx->slot_arena[1], &ctx->slot_tar[0], &ctx->slot_tar[1], &ctx->slot_out[0], &ctx->slot_out[1]})
    b->release();
  for (auto &t : ctx->tim) {
    if (t.a) cudaEventDestroy(t.a);
    if (t.b) cudaEventDestroy(t.b);
  }
  if (ctx->h_scalars) cudaFreeHost(ctx->h_scalars);
  for (uint8_t *b : ctx->h_stage)
    if (b) cudaFreeHost(b);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  if (ctx->stream_in) cudaStreamDestroy(ctx->stream_in);
  if (ctx->stream_out) cudaStreamDestroy(ctx->stream_out);
  cudaGetLastError();
  delete ctx;
}

void *dsg_pinned_alloc(dsg_ctx *ctx, size_t bytes) {
  if (!ctx) return nullptr;
  cudaSetDevice(ctx->device);
  void *p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return p;
}
void dsg_pinned_free(dsg_ctx *ctx, void *p) {
  if (!ctx || !p) return;
  cudaSetDevice(ctx->device);
  cudaFreeHost(p);
}

void *dsg_dev_alloc(dsg_ctx *ctx, size_t bytes) {
  if (!ctx) return nullptr;
  cudaSetDevice(ctx->device);
  void *p = nullptr;
  if (cudaMalloc(&p, bytes + 256) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return p;
}
void dsg_dev_free(dsg_ctx *ctx, void *p) {
  if (!ctx || !p) return;
  cudaSetDevice(ctx->device);
  cudaFree(p);
}
int dsg_dev_upload(dsg_ctx *ctx, void *d_dst, const void *src, size_t bytes) {
  if (!ctx) return DSG_E_ARG;
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemcpyAsync(d_dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DSG_OK;
}
int dsg_dev_download(dsg_ctx *ctx, void *dst, const void *d_src, size_t bytes) {
  if (!ctx) return DSG_E_ARG;
  CU(cudaSetDevice(ctx->device));
  CU(cudaMemcpyAsync(dst, d_src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return DSG_OK;
}

// ---- host-side layout helpers ---------------------------------------------------------------------------------

int dsg_entry_header_bytes(const dsg_entry *e, const char *strtab, size_t strtab_len, uint64_t *out) {
  if (!e || !strtab || !out) return DSG_E_ARG;
  HdrPlan p;
  int rc = classify_entry(*e, strtab, strtab_len, p);
  if (rc != DSG_OK) return rc;
  *out = p.hdr_bytes;
  return DSG_OK;
}

int dsg_tar_size(const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len, uint32_t flags,
                 uint64_t *out) {
  if ((!entries && n) || !strtab || !out) return DSG_E_ARG;
  uint64_t t = 0;
  for (size_t i = 0; i < n; i++) {
    HdrPlan p;
    int rc = classify_entry(entries[i], strtab, strtab_len, p);
    if (rc != DSG_OK) return rc;
    t += p.hdr_bytes + entry_payload_padded(entries[i]);
  }
  *out = t + ((flags & DSG_F_NO_TRAILER) ? 0 : 2 * kBlock);
  return DSG_OK;
}

uint64_t dsg_gzip_bound(uint64_t tar_bytes) { return 18 + lz_deflate_bound(tar_bytes); }

// The shards' gzip members as ONE member ("optional host-side stitch" for decoders that stop after
// the first member): every deflate body this library writes consists of non-final blocks that end on a byte
// boundary plus the 2-byte terminator 03 00 (an empty final block), so the bodies of all members but the last lose
// their terminator and are concatenated; the checksum is folded with x^(8n) mod P like the tile remainders on the
// device, ISIZE is the sum mod 2^32. Host integer work over O(number of shards) headers; the bodies are copied once.
int dsg_stitch_members(const uint8_t *const *members, const size_t *member_lens, const uint64_t *tar_lens, size_t n,
                       uint8_t *out, size_t out_cap, size_t *out_len) {
  static const uint8_t hdr[10] = {0x1f, 0x8b, 0x08, 0, 0, 0, 0, 0, 0, 0xff};
  if (!members || !member_lens || !tar_lens || !out_len || n == 0) return DSG_E_ARG;
  size_t need = 10 + 8;
  for (size_t i = 0; i < n; i++) {
    const uint8_t *m = members[i];
    const size_t len = member_lens[i];
    if (!m || len < 20 || memcmp(m, hdr, 10) != 0) return DSG_E_ARG;              // not one of our members
    if (m[len - 10] != 0x03 || m[len - 9] != 0x00) return DSG_E_ARG;              // body must end with the terminator
    uint32_t isize = 0;
    for (int b = 0; b < 4; b++) isize |= (uint32_t)m[len - 4 + b] << (8 * b);
    if (isize != (uint32_t)tar_lens[i]) return DSG_E_ARG;                          // tar_lens does not belong to it
    need += len - 18 - (i + 1 < n ? 2 : 0);
  }
  *out_len = need;
  if (!out || out_cap < need) return DSG_E_CAPACITY;
  memcpy(out, hdr, 10);
  size_t w = 10;
  uint32_t crc = 0;
  uint64_t total = 0;
  for (size_t i = 0; i < n; i++) {
    const uint8_t *m = members[i];
    const size_t len = member_lens[i];
    const size_t body = len - 18 - (i + 1 < n ? 2 : 0);
    memcpy(out + w, m + 10, body);
    w += body;
    uint32_t c = 0;
    for (int b = 0; b < 4; b++) c |= (uint32_t)m[len - 8 + b] << (8 * b);
    crc = total ? crc_combine(crc, c, tar_lens[i]) : c;
    total += tar_lens[i];
  }
  for (int b = 0; b < 4; b++) out[w + b] = (uint8_t)(crc >> (8 * b));
  for (int b = 0; b < 4; b++) out[w + 4 + b] = (uint8_t)((uint32_t)total >> (8 * b));
  return DSG_OK;
}

uint32_t dsg_crc32_combine(uint32_t crc_a, uint32_t crc_b, uint64_t len_b) { return crc_combine(crc_a, crc_b, len_b); }

int dsg_shard_bounds(const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len, int nshards,
                     uint64_t *bounds) {
  if ((!entries && n) || !strtab || !bounds || nshards <= 0) return DSG_E_ARG;
  std::vector<uint64_t> pre(n + 1, 0);
  for (size_t i = 0; i < n; i++) {
    HdrPlan p;
    int rc = classify_entry(entries[i], strtab, strtab_len, p);
    if (rc != DSG_OK) return rc;
    pre[i + 1] = pre[i] + p.hdr_bytes + entry_payload_padded(entries[i]);
  }
  const uint64_t total = pre[n];
  bounds[0] = 0;
  for (int s = 1; s < nshards; s++) {
    // first entry whose start offset reaches s/nshards of the stream; files are never split
    const uint64_t target = (uint64_t)(((unsigned __int128)total * (unsigned)s) / (unsigned)nshards);
    size_t k = std::lower_bound(pre.begin(), pre.end(), target) - pre.begin();
    if (k > n) k = n;
    if (k < bounds[s - 1]) k = bounds[s - 1];
    bounds[s] = k;
  }
  bounds[nshards] = n;
  return DSG_OK;
}

// ---- single stages ------------------------------------------------------------------------------------------

int dsg_dev_tar_build(dsg_ctx *ctx, const dsg_entry *d_entries, size_t n, const char *d_strtab, size_t strtab_len,
                      const uint8_t *d_arena, size_t arena_len, uint32_t flags, uint8_t *d_tar, size_t tar_cap,
                      uint64_t *tar_len) {
  if (!ctx || !tar_len || (n && (!d_entries || !d_strtab)) || n > 0xfffffff0ull) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  uint64_t T = 0, P = 0;
  rc = tar_plan(ctx, d_entries, (uint32_t)n, d_strtab, strtab_len, arena_len, flags, &T, &P);
  if (rc) return rc;
  *tar_len = T;
  if (T > tar_cap) return fail(ctx, DSG_E_CAPACITY, "tar stream needs %llu bytes", (unsigned long long)T);
  tic(ctx, kTimTar);
  rc = tar_layout(ctx, d_entries, (uint32_t)n, d_strtab, d_arena, 0, d_tar, 0, T);
  if (rc) return rc;
  toc(ctx, kTimTar);
  rc = check_status(ctx);
  ctx->stats.entries = n;
  ctx->stats.payload_bytes = P;
  ctx->stats.tar_bytes = T;
  ctx->stats.ms_plan = elapsed(ctx, kTimPlan);
  ctx->stats.ms_tar = elapsed(ctx, kTimTar);
  ctx->stats.kernel_launches = ctx->launches;
  return rc;
}

int dsg_dev_crc32(dsg_ctx *ctx, const uint8_t *d_buf, size_t len, uint32_t *crc) {
  if (!ctx || !crc || (len && !d_buf)) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  if ((rc = crc_reset(ctx))) return rc;
  tic(ctx, kTimCrc);
  if ((rc = crc_update(ctx, d_buf, len))) return rc;
  toc(ctx, kTimCrc);
  CU(cudaMemcpyAsync(ctx->h_scalars, ctx->scalars.p, sizeof(uint64_t) * kScalarWords, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  *crc = (uint32_t)ctx->h_scalars[kSCrc];
  ctx->stats.ms_crc = elapsed(ctx, kTimCrc);
  ctx->stats.crc32 = *crc;
  ctx->stats.kernel_launches = ctx->launches;
  return DSG_OK;
}

// ---- per-file CRC32 (build-context hash, pkg/util/hash/hash.go:165-197) ----------------------------------------

namespace {
// d_crcs[i] = CRC32-IEEE of entry i's payload (0 for directories and empty files). Asynchronous on ctx->stream
// except for one synchronisation that reads the number of units.
int file_crc_run(dsg_ctx *ctx, const dsg_entry *d_entries, uint32_t n, const uint8_t *d_arena, uint64_t arena_len,
                 uint32_t *d_crcs) {
  if (n == 0) return DSG_OK;
  cudaStream_t st = ctx->stream;
  const uint32_t nblk = (n + kScanBlock - 1) / kScanBlock;
  CU(ctx->fc_files.ensure(sizeof(FileCrcEntry) * (size_t)n));
  CU(ctx->fc_units.ensure(sizeof(uint32_t) * (size_t)n));
  CU(ctx->fc_first.ensure(sizeof(uint32_t) * ((size_t)n + 1)));
  CU(ctx->fc_totals.ensure(sizeof(uint32_t) * ((size_t)nblk + 2)));
  FileCrcEntry *files = ctx->fc_files.as<FileCrcEntry>();
  uint32_t *units = ctx->fc_units.as<uint32_t>(), *first = ctx->fc_first.as<uint32_t>();
  uint32_t *totals = ctx->fc_totals.as<uint32_t>();
  uint32_t *grand = totals + nblk, *error = totals + nblk + 1;
  CU(cudaMemsetAsync(error, 0, sizeof(uint32_t), st));
  fcrc_extract_kernel<<<(n + 255) / 256, 256, 0, st>>>(reinterpret_cast<const uint8_t *>(d_entries), n,
                                                      (uint32_t)sizeof(dsg_entry), (uint32_t)offsetof(dsg_entry, size),
                                                      (uint32_t)offsetof(dsg_entry, data_off),
                                                      (uint32_t)offsetof(dsg_entry, typeflag), arena_len, files, error);
  LAUNCH_CHECK();
  fcrc_count_kernel<<<(n + 255) / 256, 256, 0, st>>>(files, n, (uint64_t)(uintptr_t)d_arena, units);
  LAUNCH_CHECK();
  fcrc_scan_blocks_kernel<<<nblk, kScanBlock, 0, st>>>(units, n, first, totals);
  LAUNCH_CHECK();
  fcrc_scan_totals_kernel<<<1, kScanBlock, 0, st>>>(totals, nblk, grand);
  LAUNCH_CHECK();
  fcrc_scan_add_kernel<<<nblk, kScanBlock, 0, st>>>(first, n, totals, grand);
  LAUNCH_CHECK();
  uint32_t h[2] = {0, 0};  // grand total, error
  CU(cudaMemcpyAsync(h, grand, sizeof h, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (h[1]) return fail(ctx, DSG_E_RANGE, "entry %u: payload range outside the arena", h[1] - 1);
  const uint32_t nunits = h[0];
  CU(ctx->fc_raw.ensure(sizeof(uint32_t) * (2 * (size_t)nunits + 2)));  // remainders, then the unit -> file map
  uint32_t *unit_file = ctx->fc_raw.as<uint32_t>() + nunits + 1;
  if (nunits) {
    fcrc_map_kernel<<<(n + 255) / 256, 256, 0, st>>>(first, n, unit_file);
    LAUNCH_CHECK();
    fcrc_units_kernel<<<(nunits + kFileCrcUnitsPerCta - 1) / kFileCrcUnitsPerCta, kCrcThreads, 0, st>>>(
        files, n, d_arena, first, unit_file, nunits, ctx->crc_tab.as<uint32_t>(),
                                                       ctx->crc_shift.as<uint32_t>(), ctx->fc_raw.as<uint32_t>());
    LAUNCH_CHECK();
  }
  fcrc_join_kernel<<<(n + 7) / 8, 256, 0, st>>>(files, n, d_arena, first, ctx->fc_raw.as<uint32_t>(), d_crcs);
  LAUNCH_CHECK();
  return DSG_OK;
}
}  // namespace

extern "C" int dsg_dev_file_crc32(dsg_ctx *ctx, const dsg_entry *d_entries, size_t n, const uint8_t *d_arena,
                                  size_t arena_len, uint32_t *d_crcs) {
  if (!ctx || (n && (!d_entries || !d_crcs)) || (arena_len && !d_arena) || n > 0x7fffffffu) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  tic(ctx, kTimCrc);
  if ((rc = file_crc_run(ctx, d_entries, (uint32_t)n, d_arena, arena_len, d_crcs))) return rc;
  toc(ctx, kTimCrc);
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->stats.entries = n;
  ctx->stats.ms_crc = elapsed(ctx, kTimCrc);
  ctx->stats.ms_total = ctx->stats.ms_crc;
  ctx->stats.kernel_launches = ctx->launches;
  return DSG_OK;
}

extern "C" int dsg_file_crc32(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const uint8_t *arena, size_t arena_len,
                              uint32_t *crcs) {
  if (!ctx || (n && (!entries || !crcs)) || (arena_len && !arena) || n > 0x7fffffffu) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  if (n == 0) return DSG_OK;
  CU(ctx->d_entries.ensure(sizeof(dsg_entry) * n));
  CU(ctx->d_arena.ensure(arena_len + 16));
  CU(ctx->fc_crcs.ensure(sizeof(uint32_t) * n));
  Timer &tt = ctx->tim[kTimTotal];
  cudaEventRecord(tt.a, ctx->stream);
  CU(cudaMemcpyAsync(ctx->d_entries.p, entries, sizeof(dsg_entry) * n, cudaMemcpyHostToDevice, ctx->stream));
  if (arena_len) CU(cudaMemcpyAsync(ctx->d_arena.p, arena, arena_len, cudaMemcpyHostToDevice, ctx->stream));
  tic(ctx, kTimCrc);
  if ((rc = file_crc_run(ctx, ctx->d_entries.as<dsg_entry>(), (uint32_t)n, ctx->d_arena.as<uint8_t>(), arena_len,
                         ctx->fc_crcs.as<uint32_t>())))
    return rc;
  toc(ctx, kTimCrc);
  CU(cudaMemcpyAsync(crcs, ctx->fc_crcs.p, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, ctx->stream));
  cudaEventRecord(tt.b, ctx->stream);
  tt.used = true;
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->stats.entries = n;
  ctx->stats.h2d_bytes = sizeof(dsg_entry) * n + arena_len;
  ctx->stats.d2h_bytes = sizeof(uint32_t) * n;
  ctx->stats.ms_crc = elapsed(ctx, kTimCrc);
  ctx->stats.ms_total = elapsed(ctx, kTimTotal);
  ctx->stats.kernel_launches = ctx->launches;
  return DSG_OK;
}

int dsg_dev_deflate(dsg_ctx *ctx, const uint8_t *d_in, size_t len, int level, uint8_t *d_out, size_t out_cap,
                    size_t *out_len) {
  if (!ctx || !out_len || !d_out || (len && !d_in)) return DSG_E_ARG;
  if (level != DSG_LEVEL_STORED && level != DSG_LEVEL_DEFAULT) return fail(ctx, DSG_E_ARG, "level must be 0 or 6");
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  uint64_t n = 0;
  rc = deflate_run(ctx, d_in, len, level, d_out, out_cap, &n);
  *out_len = (size_t)n;
  if (rc) return rc;
  CU(cudaStreamSynchronize(ctx->stream));
  if (level != DSG_LEVEL_STORED) {
    CU(cudaMemcpy(ctx->h_scalars, ctx->scalars.p, sizeof(uint64_t) * kScalarWords, cudaMemcpyDeviceToHost));
    n = ctx->h_scalars[kSDeflateBytes];
    *out_len = (size_t)n;
    if (ctx->h_scalars[kSSpare] != 0)
      return fail(ctx, DSG_E_CAPACITY, "deflate output needs %llu bytes, capacity %llu", (unsigned long long)n,
                  (unsigned long long)out_cap);
  }
  ctx->stats.tar_bytes = len;
  ctx->stats.out_bytes = n;
  ctx->stats.ms_deflate = elapsed(ctx, kTimDeflate);
  ctx->stats.ms_gather = elapsed(ctx, kTimGather);
  ctx->stats.kernel_launches = ctx->launches;
  return DSG_OK;
}

// gzip member around an existing device buffer: header | deflate | crc32 | isize
static int gzip_device(dsg_ctx *ctx, const uint8_t *d_in, uint64_t len, int level, uint8_t *d_out, uint64_t out_cap,
                       uint64_t *out_len) {
  if (out_cap < 18) {
    *out_len = dsg_gzip_bound(len);
    return fail(ctx, DSG_E_CAPACITY, "gzip output needs at least 18 bytes");
  }
  int rc;
  uint64_t *sc = ctx->scalars.as<uint64_t>();
  if ((rc = crc_reset(ctx))) return rc;
  CU(cudaMemsetAsync(sc + kSSpare, 0, sizeof(uint64_t), ctx->stream));  // the deflate stage's "too small" flag
  tic(ctx, kTimCrc);
  if ((rc = crc_update(ctx, d_in, len))) return rc;
  toc(ctx, kTimCrc);
  uint64_t dn = 0;
  rc = deflate_run(ctx, d_in, len, level, d_out + 10, out_cap - 18, &dn);
  if (rc) {
    *out_len = dn + 18;
    return rc;
  }
  gzip_frame_kernel<<<1, 32, 0, ctx->stream>>>(d_out, out_cap, sc + kSDeflateBytes,
                                               reinterpret_cast<uint32_t *>(sc + kSCrc),
                                               reinterpret_cast<uint32_t *>(sc + kSSpare));
  LAUNCH_CHECK();
  CU(cudaMemcpyAsync(ctx->h_scalars, sc, sizeof(uint64_t) * kScalarWords, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  *out_len = ctx->h_scalars[kSDeflateBytes] + 18;
  ctx->stats.crc32 = (uint32_t)ctx->h_scalars[kSCrc];
  if (level != DSG_LEVEL_STORED && ctx->h_scalars[kSSpare] != 0)
    return fail(ctx, DSG_E_CAPACITY, "gzip output needs %llu bytes, capacity %llu", (unsigned long long)*out_len,
                (unsigned long long)out_cap);
  return DSG_OK;
}

int dsg_dev_gzip(dsg_ctx *ctx, const uint8_t *d_in, size_t len, int level, uint8_t *d_out, size_t out_cap,
                 size_t *out_len) {
  if (!ctx || !out_len || !d_out || (len && !d_in)) return DSG_E_ARG;
  if (level != DSG_LEVEL_STORED && level != DSG_LEVEL_DEFAULT) return fail(ctx, DSG_E_ARG, "level must be 0 or 6");
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  uint64_t n = 0;
  rc = gzip_device(ctx, d_in, len, level, d_out, out_cap, &n);
  *out_len = (size_t)n;
  if (rc) return rc;
  ctx->stats.tar_bytes = len;
  ctx->stats.out_bytes = n;
  ctx->stats.ms_deflate = elapsed(ctx, kTimDeflate);
  ctx->stats.ms_crc = elapsed(ctx, kTimCrc);
  ctx->stats.ms_gather = elapsed(ctx, kTimGather);
  ctx->stats.kernel_launches = ctx->launches;
  return DSG_OK;
}

// ---- the hot path -----------------------------------------------------------------------------------------------

static int pack_resident(dsg_ctx *ctx, const dsg_entry *d_entries, size_t n, const char *d_strtab, size_t strtab_len,
                         const uint8_t *d_arena, size_t arena_len, bool gzip, uint8_t *d_out, size_t out_cap,
                         uint64_t *out_len) {
  uint64_t T = 0, P = 0;
  int rc = tar_plan(ctx, d_entries, (uint32_t)n, d_strtab, strtab_len, arena_len, ctx->opts.flags, &T, &P);
  if (rc) return rc;
  ctx->stats.entries = n;
  ctx->stats.payload_bytes = P;
  ctx->stats.tar_bytes = T;
  if (!gzip) {
    *out_len = T;
    if (T > out_cap) return fail(ctx, DSG_E_CAPACITY, "tar stream needs %llu bytes", (unsigned long long)T);
    tic(ctx, kTimTar);
    if ((rc = tar_layout(ctx, d_entries, (uint32_t)n, d_strtab, d_arena, 0, d_out, 0, T))) return rc;
    toc(ctx, kTimTar);
    return check_status(ctx);
  }
  CU(ctx->d_tar.ensure(T + 64));
  tic(ctx, kTimTar);
  if ((rc = tar_layout(ctx, d_entries, (uint32_t)n, d_strtab, d_arena, 0, ctx->d_tar.as<uint8_t>(), 0, T))) return rc;
  toc(ctx, kTimTar);
  rc = gzip_device(ctx, ctx->d_tar.as<uint8_t>(), T, ctx->opts.level, d_out, out_cap, out_len);
  if (rc) return rc;
  DevStatus st;
  memcpy(&st, &ctx->h_scalars[kSStatus], sizeof st);
  if (st.code != 0) return fail(ctx, st.code, "entry %u: %s", st.entry, dsg_strerror(st.code));
  return DSG_OK;
}

static void finish_stats(dsg_ctx *ctx, uint64_t out_bytes) {
  ctx->stats.out_bytes = out_bytes;
  ctx->stats.ms_plan = elapsed(ctx, kTimPlan);
  ctx->stats.ms_tar = elapsed(ctx, kTimTar);
  ctx->stats.ms_deflate = elapsed(ctx, kTimDeflate);
  ctx->stats.ms_crc = elapsed(ctx, kTimCrc);
  ctx->stats.ms_gather = elapsed(ctx, kTimGather);
  ctx->stats.ms_total = elapsed(ctx, kTimTotal);
  ctx->stats.kernel_launches = ctx->launches;
}

int dsg_pack_device(dsg_ctx *ctx, const dsg_entry *d_entries, size_t n, const char *d_strtab, size_t strtab_len,
                    const uint8_t *d_arena, size_t arena_len, uint8_t *d_out, size_t out_cap, size_t *out_len) {
  if (!ctx || !out_len || !d_out || (n && (!d_entries || !d_strtab)) || n > 0xfffffff0ull) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  tic(ctx, kTimTotal);
  uint64_t produced = 0;
  rc = pack_resident(ctx, d_entries, n, d_strtab, strtab_len, d_arena, arena_len, true, d_out, out_cap, &produced);
  *out_len = (size_t)produced;
  if (rc) return rc;
  toc(ctx, kTimTotal);
  CU(cudaStreamSynchronize(ctx->stream));
  finish_stats(ctx, produced);
  return DSG_OK;
}

// Host buffers in, host buffer out, pipelined per segment of the tar stream:
//   copy stream   : arena slice of segment i+1  host -> device
//   compute stream: layout -> CRC32 -> deflate of segment i        (dictionary = tail of segment i-1)
//   copy-back     : compressed bytes of segment i-1  device -> host, at the running output offset
// Two device slots per buffer kind; the host only blocks on the per-segment size read-back.
static int pack_host_pipelined(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                               const uint8_t *arena, size_t arena_len, bool gzip, uint8_t *out, size_t out_cap,
                               size_t *out_len, int sink_fd = -1) {
  // sink_fd >= 0: the stream goes to that file descriptor instead of `out`: every segment is copied into one of two
  // pinned staging buffers and written while the GPU works on the following segments.
  constexpr int NS = 2;
  const bool sink = sink_fd >= 0;
  auto write_all = [&](const void *buf, size_t nbytes) -> bool {
    const uint8_t *q = static_cast<const uint8_t *>(buf);
    while (nbytes) {
      const ssize_t k = write(sink_fd, q, nbytes);
      if (k <= 0) return false;
      q += k;
      nbytes -= (size_t)k;
    }
    return true;
  };
  cudaStream_t s_cmp = ctx->stream, s_in = ctx->stream_in, s_out = ctx->stream_out;
  CU(cudaEventRecord(ctx->tim[kTimTotal].a, s_cmp));
  // ---- manifest + plan ---------------------------------------------------------------------------------------
  CU(ctx->d_entries.ensure(sizeof(dsg_entry) * std::max<size_t>(n, 1)));
  CU(ctx->d_strtab.ensure(std::max<size_t>(strtab_len, 1)));
  if (n) CU(cudaMemcpyAsync(ctx->d_entries.p, entries, sizeof(dsg_entry) * n, cudaMemcpyHostToDevice, s_cmp));
  if (strtab_len) CU(cudaMemcpyAsync(ctx->d_strtab.p, strtab, strtab_len, cudaMemcpyHostToDevice, s_cmp));
  uint64_t h2d = sizeof(dsg_entry) * n + strtab_len, d2h = 0;
  const dsg_entry *d_entries = ctx->d_entries.as<dsg_entry>();
  const char *d_strtab = ctx->d_strtab.as<char>();
  uint64_t T = 0, P = 0;
  int rc = tar_plan(ctx, d_entries, (uint32_t)n, d_strtab, strtab_len, arena_len, ctx->opts.flags, &T, &P);
  if (rc) return rc;
  ctx->stats.entries = n;
  ctx->stats.payload_bytes = P;
  ctx->stats.tar_bytes = T;
  // Default segment: a sixth of the stream between 64 and 512 MiB. Every segment is one deflate launch, and a launch
  // ends with up to one chunk time (1.2 ms) of idle SMs and starts with every CTA hashing a dictionary: measured
  // 2.9 ms per 256 MiB segment, 1.8 ms per 512 MiB one; small streams still want a few segments for the overlap.
  uint64_t S = ctx->opts.segment_bytes;
  if (S == 0) S = std::min<uint64_t>(512ull << 20, std::max<uint64_t>(64ull << 20, T / 6));
  S = std::max<uint64_t>((S + kChunk - 1) / kChunk * kChunk, kChunk);
  const uint32_t nseg = (uint32_t)std::max<uint64_t>((T + S - 1) / S, 1);
  // per-segment arena ranges (device -> pinned host)
  CU(ctx->seg_ranges.ensure(sizeof(uint64_t) * 3 * (size_t)nseg));
  std::vector<uint64_t> h_rng(3 * (size_t)nseg);
  unsigned long long *d_lo = ctx->seg_ranges.as<unsigned long long>(), *d_hi = d_lo + nseg;
  uint64_t *d_segout = reinterpret_cast<uint64_t *>(d_hi + nseg);
  CU(cudaMemsetAsync(d_lo, 0xff, sizeof(uint64_t) * nseg, s_cmp));
  CU(cudaMemsetAsync(d_hi, 0, sizeof(uint64_t) * 2 * nseg, s_cmp));
  if (n) {
    tar_segment_ranges_kernel<<<(unsigned)((n + kPlanThreads - 1) / kPlanThreads), kPlanThreads, 0, s_cmp>>>(
        d_entries, (uint32_t)n, ctx->ent_off.as<uint64_t>(), S, d_lo, d_hi);
    LAUNCH_CHECK();
  }
  CU(cudaMemcpyAsync(h_rng.data(), d_lo, sizeof(uint64_t) * 2 * nseg, cudaMemcpyDeviceToHost, s_cmp));
  CU(cudaStreamSynchronize(s_cmp));
  uint64_t max_arena = 16;
  for (uint32_t i = 0; i < nseg; i++) {
    uint64_t lo = h_rng[i], hi = h_rng[nseg + i];
    if (hi <= lo || lo == ~0ull) {
      h_rng[i] = h_rng[nseg + i] = 0;
      continue;
    }
    lo &= ~uint64_t(15);  // keeps the 16-byte alignment of payload starts relative to the slot base
    h_rng[i] = lo;
    max_arena = std::max(max_arena, hi - lo);
  }
  // ---- output framing (host memory: written directly) ----------------------------------------------------------
  const uint64_t seg_out_cap = gzip ? lz_deflate_bound(std::min(S, std::max<uint64_t>(T, 1))) : 0;
  size_t produced = 0;
  if (gzip) {
    static const uint8_t hdr[10] = {0x1f, 0x8b, 0x08, 0, 0, 0, 0, 0, 0, 0xff};  // compress/gzip Writer defaults
    if (sink) {
      if (!write_all(hdr, 10)) return fail(ctx, DSG_E_ARG, "write: %s", strerror(errno));
    } else {
      if (out_cap < 18) {
        *out_len = (size_t)dsg_gzip_bound(T);
        return fail(ctx, DSG_E_CAPACITY, "gzip output needs at least 18 bytes");
      }
      memcpy(out, hdr, 10);
    }
    produced = 10;
  }
  if (sink) {
    const size_t want = (size_t)(gzip ? seg_out_cap : std::min(S, std::max<uint64_t>(T, 1))) + 64;
    if (want > ctx->h_stage_cap) {
      for (uint8_t *&b : ctx->h_stage) {
        if (b) cudaFreeHost(b);
        b = nullptr;
        CU(cudaHostAlloc(reinterpret_cast<void **>(&b), want, cudaHostAllocDefault));
      }
      ctx->h_stage_cap = want;
    }
  }
  std::vector<uint64_t> seg_bytes(nseg, 0);
  // ---- slots ---------------------------------------------------------------------------------------------------
  for (int k = 0; k < NS; k++) {
    CU(ctx->slot_arena[k].ensure(max_arena + 64));
    if (gzip) {
      CU(ctx->slot_tar[k].ensure(std::min(S, T) + 64));
      CU(ctx->slot_out[k].ensure(seg_out_cap + 64));
    } else {
      CU(ctx->slot_tar[k].ensure(std::min(S, T) + 64));
    }
  }
  if (gzip && ctx->opts.level != DSG_LEVEL_STORED) {
    // Sized with the launch's own arithmetic for the largest segment, so that no launch of the pipeline has to
    // re-allocate (cudaFree + cudaMalloc synchronise the device in the middle of the overlap).
    LzGeometry G;
    if (!lz_geometry(ctx, std::min(S, T), G)) return fail(ctx, DSG_E_ARG, "segment too large for one launch");
    CU(ctx->chunk_sizes.ensure((sizeof(uint64_t) + sizeof(uint32_t)) * (size_t)G.nchunks + 64));
    CU(ctx->chunk_slots.ensure(G.slot_bytes));
    CU(ctx->lz_scratch.ensure(sizeof(uint32_t) * (size_t)kTokSlotWords * ctx->sm_count));
  }
  if (gzip) {
    if ((rc = crc_reset(ctx))) return rc;
  }
  uint64_t *sc = ctx->scalars.as<uint64_t>();
  CU(cudaMemsetAsync(sc + kSSpare, 0, sizeof(uint64_t), s_cmp));
  std::vector<cudaEvent_t> ev_in(nseg), ev_cmp(nseg), ev_out(nseg), ev_t(2 * (size_t)nseg);
  auto destroy_events = [&]() {
    for (auto *v : {&ev_in, &ev_cmp, &ev_out, &ev_t})
      for (cudaEvent_t e : *v)
        if (e) cudaEventDestroy(e);
  };
  for (auto *v : {&ev_in, &ev_cmp, &ev_out})
    for (auto &e : *v) {
      e = nullptr;
      if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
        destroy_events();
        return fail(ctx, DSG_E_CUDA, "cudaEventCreate failed");
      }
    }
  for (auto &e : ev_t) {
    e = nullptr;
    if (cudaEventCreate(&e) != cudaSuccess) {
      destroy_events();
      return fail(ctx, DSG_E_CUDA, "cudaEventCreate failed");
    }
  }
#define CUP(call)                                                                                          \
  do {                                                                                                     \
    cudaError_t e_ = (call);                                                                               \
    if (e_ != cudaSuccess) {                                                                               \
      cudaGetLastError();                                                                                  \
      cudaDeviceSynchronize();                                                                             \
      destroy_events();                                                                                    \
      return fail(ctx, DSG_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);   \
    }                                                                                                      \
  } while (0)

  uint64_t *h_sizes = ctx->h_scalars + kScalarWords;  // pinned, kMaxSegSizes slots (ring)
  int status = DSG_OK;
  auto enqueue = [&](uint32_t i) -> int {
    const int k = (int)(i % NS);
    const uint64_t lo = (uint64_t)i * S, hi = std::min<uint64_t>(lo + S, T);
    const uint64_t a_lo = h_rng[i], a_hi = h_rng[nseg + i];
    // copy-in: slot k's previous user (segment i-NS) must have finished computing
    if (i >= NS) CUP(cudaStreamWaitEvent(s_in, ev_cmp[i - NS], 0));
    if (a_hi > a_lo) {
      CUP(cudaMemcpyAsync(ctx->slot_arena[k].p, arena + a_lo, a_hi - a_lo, cudaMemcpyHostToDevice, s_in));
      h2d += a_hi - a_lo;
    }
    CUP(cudaEventRecord(ev_in[i], s_in));
    // compute
    CUP(cudaStreamWaitEvent(s_cmp, ev_in[i], 0));
    if (i >= NS) CUP(cudaStreamWaitEvent(s_cmp, ev_out[i - NS], 0));  // slot k's output has been copied back
    CUP(cudaEventRecord(ev_t[2 * i], s_cmp));
    uint8_t *d_seg = ctx->slot_tar[k].as<uint8_t>();
    int r = tar_layout(ctx, d_entries, (uint32_t)n, d_strtab, ctx->slot_arena[k].as<uint8_t>(), a_lo, d_seg, lo, hi);
    if (r) return r;
    if (gzip) {
      if ((r = crc_update(ctx, d_seg, hi - lo))) return r;
      const bool last = i + 1 == nseg;
      if (ctx->opts.level == DSG_LEVEL_STORED) {
        const uint64_t nblocks = (hi - lo + kStoredMax - 1) / kStoredMax;
        deflate_stored_kernel<<<(unsigned)std::max<uint64_t>(nblocks, 1), kStoredThreads, 0, s_cmp>>>(
            d_seg, hi - lo, ctx->slot_out[k].as<uint8_t>(), last ? 1 : 0);
        ctx->launches++;
        set_u64_kernel<<<1, 1, 0, s_cmp>>>(d_segout + i, stored_deflate_size(hi - lo) - (last ? 0 : 2));
        ctx->launches++;
      } else {
        const uint8_t *dict = nullptr;
        uint32_t dict_len = 0;
        if (i > 0) {  // tail of the previous segment, still resident in the other slot
          dict_len = (uint32_t)std::min<uint64_t>(S, kDict);
          dict = ctx->slot_tar[(i - 1) % NS].as<uint8_t>() + (S - dict_len);
        }
        r = lz_launch(ctx, s_cmp, d_seg, hi - lo, dict, dict_len, ctx->slot_out[k].as<uint8_t>(), seg_out_cap,
                      d_segout + i, reinterpret_cast<uint32_t *>(sc + kSSpare), last ? 1 : 0);
        if (r) return r;
      }
      CUP(cudaMemcpyAsync(&h_sizes[i % kMaxSegSizes], d_segout + i, sizeof(uint64_t), cudaMemcpyDeviceToHost, s_cmp));
    }
    CUP(cudaEventRecord(ev_t[2 * i + 1], s_cmp));
    CUP(cudaEventRecord(ev_cmp[i], s_cmp));
    return DSG_OK;
  };
  auto drain = [&](uint32_t i) -> int {
    const int k = (int)(i % NS);
    const uint64_t lo = (uint64_t)i * S, hi = std::min<uint64_t>(lo + S, T);
    CUP(cudaEventSynchronize(ev_cmp[i]));
    const uint64_t nbytes = gzip ? h_sizes[i % kMaxSegSizes] : hi - lo;
    const uint64_t tail = gzip && i + 1 == nseg ? 8 : 0;
    seg_bytes[i] = nbytes;
    if (!sink && produced + nbytes + tail > out_cap) {
      produced += nbytes + tail;
      status = DSG_E_CAPACITY;
      CUP(cudaEventRecord(ev_out[i], s_out));
      return DSG_OK;
    }
    CUP(cudaStreamWaitEvent(s_out, ev_cmp[i], 0));
    const void *src = gzip ? ctx->slot_out[k].p : ctx->slot_tar[k].p;
    if (nbytes && status == DSG_OK) {
      CUP(cudaMemcpyAsync(sink ? ctx->h_stage[i % 2] : out + produced, src, nbytes, cudaMemcpyDeviceToHost, s_out));
      d2h += nbytes;
    }
    produced += nbytes;
    CUP(cudaEventRecord(ev_out[i], s_out));
    return DSG_OK;
  };
  auto flush = [&](uint32_t j) -> int {  // file-descriptor sink: segment j has arrived in its staging buffer
    if (!sink) return DSG_OK;
    CUP(cudaEventSynchronize(ev_out[j]));
    if (!write_all(ctx->h_stage[j % 2], (size_t)seg_bytes[j])) {
      cudaDeviceSynchronize();
      destroy_events();
      return fail(ctx, DSG_E_ARG, "write: %s", strerror(errno));
    }
    return DSG_OK;
  };
  for (uint32_t i = 0; i < nseg && rc == DSG_OK; i++) {
    rc = enqueue(i);
    if (rc == DSG_OK && i >= 2) rc = flush(i - 2);  // written while segments i-1 and i are on the GPU
    if (rc == DSG_OK && i >= 1) rc = drain(i - 1);
  }
  if (rc == DSG_OK && nseg >= 2) rc = flush(nseg - 2);
  if (rc == DSG_OK) rc = drain(nseg - 1);
  if (rc == DSG_OK) rc = flush(nseg - 1);
  if (rc != DSG_OK) {
    cudaDeviceSynchronize();
    destroy_events();
    return rc;
  }
  // trailer: CRC32 + ISIZE from the device-resident running state
  CUP(cudaMemcpyAsync(ctx->h_scalars, sc, sizeof(uint64_t) * kScalarWords, cudaMemcpyDeviceToHost, s_cmp));
  CUP(cudaStreamSynchronize(s_cmp));
  CUP(cudaStreamSynchronize(s_out));
  CUP(cudaEventRecord(ctx->tim[kTimTotal].b, s_out));
  CUP(cudaEventSynchronize(ctx->tim[kTimTotal].b));
  ctx->tim[kTimTotal].used = true;
  float ms_cmp = 0.f;
  for (uint32_t i = 0; i < nseg; i++) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ev_t[2 * i], ev_t[2 * i + 1]) == cudaSuccess) ms_cmp += ms;
  }
  destroy_events();
#undef CUP
  DevStatus st;
  memcpy(&st, &ctx->h_scalars[kSStatus], sizeof st);
  if (st.code != 0) return fail(ctx, st.code, "entry %u: %s", st.entry, dsg_strerror(st.code));
  if (gzip) {
    const uint32_t *cs = reinterpret_cast<const uint32_t *>(&ctx->h_scalars[kSCrc]);
    const uint32_t crc = cs[0], isize = cs[1];
    ctx->stats.crc32 = crc;
    if (status == DSG_OK) {
      uint8_t tr[8];
      for (int b = 0; b < 4; b++) tr[b] = (uint8_t)(crc >> (8 * b));
      for (int b = 0; b < 4; b++) tr[4 + b] = (uint8_t)(isize >> (8 * b));
      if (sink) {
        if (!write_all(tr, 8)) return fail(ctx, DSG_E_ARG, "write: %s", strerror(errno));
      } else {
        memcpy(out + produced, tr, 8);
      }
    }
    produced += 8;
    if (ctx->h_scalars[kSSpare] != 0 && status == DSG_OK)
      return fail(ctx, DSG_E_CUDA, "internal: segment output exceeded its bound");
  }
  *out_len = produced;
  ctx->stats.h2d_bytes = h2d;
  ctx->stats.d2h_bytes = d2h;
  ctx->stats.segments = nseg;
  finish_stats(ctx, produced);
  ctx->stats.ms_deflate = ms_cmp;  // pipelined mode: layout + CRC + deflate per segment, summed
  if (status != DSG_OK) return fail(ctx, status, "output needs %llu bytes, capacity %llu", (unsigned long long)produced,
                                    (unsigned long long)out_cap);
  return DSG_OK;
}

// Small archives -- the change sets of the dev loop (upstream.go:136: batches of up to 1000 files every 600 ms;
// BASELINE config 1 is 5 files / 93 KB) -- are bound by launch and synchronisation latency, not by bandwidth: the
// pipelined path synchronises four times (plan read-back, segment ranges, per-segment sizes, trailer). Here the stream
// length comes from the host-side plan (pure integer work on the manifest), everything is enqueued on one stream --
// copy in, plan, layout, CRC, deflate, frame, copy out of the bound -- and the host synchronises ONCE.
constexpr uint64_t kSmallPathBytes = 4ull << 20;

static int pack_host_small(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                           const uint8_t *arena, size_t arena_len, bool gzip, uint8_t *out, size_t out_cap,
                           size_t *out_len, uint64_t T) {
  cudaStream_t st = ctx->stream;
  CU(cudaEventRecord(ctx->tim[kTimTotal].a, st));
  // the arena range this manifest touches (a shard of a larger arena copies only its own bytes)
  uint64_t a_lo = ~0ull, a_hi = 0, P = 0;
  for (size_t i = 0; i < n; i++) {
    if (entries[i].typeflag != '0' || entries[i].size <= 0) continue;
    const uint64_t lo = entries[i].data_off, sz = (uint64_t)entries[i].size;
    if (lo > arena_len || sz > arena_len - lo) return fail(ctx, DSG_E_RANGE, "entry %zu: %s", i, dsg_strerror(DSG_E_RANGE));
    a_lo = std::min(a_lo, lo);
    a_hi = std::max(a_hi, lo + sz);
    P += sz;
  }
  if (a_hi <= a_lo || a_lo == ~0ull) a_lo = a_hi = 0;
  a_lo &= ~uint64_t(15);
  const uint64_t bound = gzip ? 18 + lz_deflate_bound(T) : T;
  CU(ctx->d_entries.ensure(sizeof(dsg_entry) * std::max<size_t>(n, 1)));
  CU(ctx->d_strtab.ensure(std::max<size_t>(strtab_len, 1)));
  CU(ctx->d_arena.ensure(a_hi - a_lo + 64));
  CU(ctx->d_tar.ensure(T + 64));
  if (gzip) CU(ctx->d_out.ensure(bound + 64));
  if (n) CU(cudaMemcpyAsync(ctx->d_entries.p, entries, sizeof(dsg_entry) * n, cudaMemcpyHostToDevice, st));
  if (strtab_len) CU(cudaMemcpyAsync(ctx->d_strtab.p, strtab, strtab_len, cudaMemcpyHostToDevice, st));
  if (a_hi > a_lo) CU(cudaMemcpyAsync(ctx->d_arena.p, arena + a_lo, a_hi - a_lo, cudaMemcpyHostToDevice, st));
  const dsg_entry *d_entries = ctx->d_entries.as<dsg_entry>();
  const char *d_strtab = ctx->d_strtab.as<char>();
  int rc = tar_plan_enqueue(ctx, d_entries, (uint32_t)n, d_strtab, strtab_len, arena_len);
  if (rc) return rc;
  uint8_t *d_tar = ctx->d_tar.as<uint8_t>();
  tic(ctx, kTimTar);
  if ((rc = tar_layout(ctx, d_entries, (uint32_t)n, d_strtab, ctx->d_arena.as<uint8_t>(), a_lo, d_tar, 0, T))) return rc;
  toc(ctx, kTimTar);
  uint64_t *sc = ctx->scalars.as<uint64_t>();
  uint64_t copied = T;
  if (gzip) {
    uint8_t *d_out = ctx->d_out.as<uint8_t>();
    if ((rc = crc_reset(ctx))) return rc;
    CU(cudaMemsetAsync(sc + kSSpare, 0, sizeof(uint64_t), st));
    tic(ctx, kTimCrc);
    if ((rc = crc_update(ctx, d_tar, T))) return rc;
    toc(ctx, kTimCrc);
    uint64_t dn = 0;
    if ((rc = deflate_run(ctx, d_tar, T, ctx->opts.level, d_out + 10, bound - 18, &dn))) return rc;
    gzip_frame_kernel<<<1, 32, 0, st>>>(d_out, bound, sc + kSDeflateBytes, reinterpret_cast<uint32_t *>(sc + kSCrc),
                                        reinterpret_cast<uint32_t *>(sc + kSSpare));
    LAUNCH_CHECK();
    copied = std::min<uint64_t>(out_cap, bound);
    if (copied) CU(cudaMemcpyAsync(out, d_out, copied, cudaMemcpyDeviceToHost, st));
  } else {
    copied = std::min<uint64_t>(out_cap, T);
    if (copied) CU(cudaMemcpyAsync(out, d_tar, copied, cudaMemcpyDeviceToHost, st));
  }
  CU(cudaMemcpyAsync(ctx->h_scalars, sc, sizeof(uint64_t) * kScalarWords, cudaMemcpyDeviceToHost, st));
  CU(cudaEventRecord(ctx->tim[kTimTotal].b, st));
  ctx->tim[kTimTotal].used = true;
  CU(cudaStreamSynchronize(st));  // the one synchronisation of the call
  DevStatus dst;
  memcpy(&dst, &ctx->h_scalars[kSStatus], sizeof dst);
  if (dst.code != 0) return fail(ctx, dst.code, "entry %u: %s", dst.entry, dsg_strerror(dst.code));
  const uint64_t produced = gzip ? ctx->h_scalars[kSDeflateBytes] + 18 : T;
  *out_len = (size_t)produced;
  ctx->stats.entries = n;
  ctx->stats.payload_bytes = P;
  ctx->stats.tar_bytes = T;
  ctx->stats.h2d_bytes = sizeof(dsg_entry) * n + strtab_len + (a_hi - a_lo);
  ctx->stats.d2h_bytes = copied;
  ctx->stats.segments = 1;
  if (gzip) ctx->stats.crc32 = (uint32_t)ctx->h_scalars[kSCrc];
  finish_stats(ctx, produced);
  if (produced > out_cap)
    return fail(ctx, DSG_E_CAPACITY, "output needs %llu bytes, capacity %llu", (unsigned long long)produced,
                (unsigned long long)out_cap);
  return DSG_OK;
}

static int pack_host(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                     const uint8_t *arena, size_t arena_len, bool gzip, uint8_t *out, size_t out_cap, size_t *out_len) {
  if (!ctx || !out_len || !out || (n && (!entries || !strtab)) || (arena_len && !arena) || n > 0xfffffff0ull)
    return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  if (ctx->opts.segment_bytes == 0 && n <= 4096) {  // (a forced segment size asks for the pipelined path)
    uint64_t T = 0;
    if (dsg_tar_size(entries, n, strtab, strtab_len, ctx->opts.flags, &T) == DSG_OK && T <= kSmallPathBytes)
      return pack_host_small(ctx, entries, n, strtab, strtab_len, arena, arena_len, gzip, out, out_cap, out_len, T);
  }
  return pack_host_pipelined(ctx, entries, n, strtab, strtab_len, arena, arena_len, gzip, out, out_cap, out_len);
}

int dsg_pack(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
             const uint8_t *arena, size_t arena_len, uint8_t *out, size_t out_cap, size_t *out_len) {
  return pack_host(ctx, entries, n, strtab, strtab_len, arena, arena_len, true, out, out_cap, out_len);
}

int dsg_pack_tar_only(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                      const uint8_t *arena, size_t arena_len, uint8_t *out, size_t out_cap, size_t *out_len) {
  return pack_host(ctx, entries, n, strtab, strtab_len, arena, arena_len, false, out, out_cap, out_len);
}

int dsg_pack_fd(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                const uint8_t *arena, size_t arena_len, int fd, int gzip, uint64_t *out_len) {
  if (!ctx || fd < 0 || (n && (!entries || !strtab)) || (arena_len && !arena) || n > 0xfffffff0ull) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  size_t produced = 0;
  rc = pack_host_pipelined(ctx, entries, n, strtab, strtab_len, arena, arena_len, gzip != 0, nullptr, 0, &produced, fd);
  if (out_len) *out_len = produced;
  return rc;
}

int dsg_get_stats(const dsg_ctx *ctx, dsg_stats *out) {
  if (!ctx || !out) return DSG_E_ARG;
  *out = ctx->stats;
  return DSG_OK;
}

}  // extern "C"

// ---- gunzip of this library's own streams (downstream direction, tar.go:16-33) ----------------------------------------

namespace {
int gunzip_device(dsg_ctx *ctx, const uint8_t *d_in, uint64_t n, uint8_t *d_out, uint64_t out_cap, uint64_t *out_len) {
  cudaStream_t st = ctx->stream;
  *out_len = 0;
  if (n < 20) return fail(ctx, DSG_E_ARG, "not a gzip stream of this library: %llu bytes", (unsigned long long)n);
  if (n > 0xffffffffull * (uint64_t)kGzScanCta / 2) return fail(ctx, DSG_E_ARG, "input too large for one call");
  // 1. candidate unit starts
  const uint32_t ncta = (uint32_t)((n + kGzScanCta - 1) / kGzScanCta);
  CU(ctx->gz_counts.ensure(sizeof(uint32_t) * 2 * (size_t)ncta + 64));
  uint32_t *d_counts = ctx->gz_counts.as<uint32_t>(), *d_offs = d_counts + ncta;
  gz_scan_count<<<ncta, kGzScanThreads, 0, st>>>(d_in, n, d_counts);
  LAUNCH_CHECK();
  std::vector<uint32_t> counts(ncta);
  CU(cudaMemcpyAsync(counts.data(), d_counts, sizeof(uint32_t) * ncta, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  uint64_t total = 0;
  for (uint32_t i = 0; i < ncta; i++) {
    const uint32_t c = counts[i];
    counts[i] = (uint32_t)total;
    total += c;
  }
  if (total == 0 || total > 0x7fffffffull) return fail(ctx, DSG_E_ARG, "not a gzip stream of this library (no unit starts found)");
  CU(cudaMemcpyAsync(d_offs, counts.data(), sizeof(uint32_t) * ncta, cudaMemcpyHostToDevice, st));
  CU(ctx->gz_cand.ensure(sizeof(uint64_t) * (size_t)total + 64));
  uint64_t *d_cand = ctx->gz_cand.as<uint64_t>();
  gz_scan_write<<<ncta, kGzScanThreads, 0, st>>>(d_in, n, d_offs, d_cand);
  LAUNCH_CHECK();
  std::vector<uint64_t> cand((size_t)total);
  CU(cudaMemcpyAsync(cand.data(), d_cand, sizeof(uint64_t) * total, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  std::sort(cand.begin(), cand.end());
  cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
  const uint32_t ncand = (uint32_t)cand.size();
  CU(cudaMemcpyAsync(d_cand, cand.data(), sizeof(uint64_t) * ncand, cudaMemcpyHostToDevice, st));
  // 2. dry decode of every candidate, 3. the chain
  CU(ctx->gz_rec.ensure(sizeof(GzRec) * (size_t)ncand + 64));
  GzRec *d_rec = ctx->gz_rec.as<GzRec>();
  gz_measure_kernel<<<(ncand + kGzWarpsPerCta - 1) / kGzWarpsPerCta, 32 * kGzWarpsPerCta, 0, st>>>(d_in, n, d_cand, ncand, d_rec);
  LAUNCH_CHECK();
  const uint32_t max_units = ncand + 16, max_members = 65536;
  CU(ctx->gz_units.ensure(sizeof(GzUnit) * (size_t)max_units));
  CU(ctx->gz_members.ensure(sizeof(GzMember) * (size_t)max_members));
  CU(ctx->gz_misc.ensure(sizeof(GzSummary) + 128));
  GzUnit *d_units = ctx->gz_units.as<GzUnit>();
  GzMember *d_members = ctx->gz_members.as<GzMember>();
  GzSummary *d_sum = ctx->gz_misc.as<GzSummary>();
  uint32_t *d_err = reinterpret_cast<uint32_t *>(d_sum + 1);
  CU(cudaMemsetAsync(d_err, 0, sizeof(uint32_t), st));
  gz_chain_kernel<<<1, 32, 0, st>>>(d_in, n, d_cand, d_rec, ncand, d_units, max_units, d_members, max_members, d_sum);
  LAUNCH_CHECK();
  GzSummary S;
  CU(cudaMemcpyAsync(&S, d_sum, sizeof S, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (S.status == DSG_E_CAPACITY) return fail(ctx, DSG_E_ARG, "gzip stream has more members than one call takes");
  if (S.status != DSG_OK)
    return fail(ctx, DSG_E_ARG, "not a gzip stream written by this library (structure breaks at byte %llu): only "
                "streams of dsg_pack* can be inflated in parallel", (unsigned long long)S.bad_offset);
  *out_len = S.total_out;
  ctx->stats.tar_bytes = S.total_out;
  ctx->stats.segments = S.nmembers;
  if (S.total_out > out_cap)
    return fail(ctx, DSG_E_CAPACITY, "output needs %llu bytes, capacity %llu", (unsigned long long)S.total_out,
                (unsigned long long)out_cap);
  if (S.nunits == 0) return DSG_OK;
  // 4. decode into 16-bit symbols, stored blocks straight to the output, then resolve the window references -- in
  // batches of consecutive units, so that the symbol scratch (2 bytes per output byte) stays bounded for streams of
  // tens of GB: a batch's references into the bytes in front of it find them final in `out`
  std::vector<GzUnit> h_units(S.nunits);
  CU(cudaMemcpyAsync(h_units.data(), d_units, sizeof(GzUnit) * S.nunits, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  const uint64_t batch_out = (ctx->opts.gunzip_batch_mib ? (uint64_t)ctx->opts.gunzip_batch_mib : 2048ull) << 20;
  CU(ctx->gz_sym.ensure(sizeof(uint16_t) * (size_t)(std::min<uint64_t>(S.total_out, batch_out) + kGzMaxUnitOut) + 64));
  uint16_t *d_sym = ctx->gz_sym.as<uint16_t>();
  tic(ctx, kTimDeflate);
  gz_stored_copy_kernel<<<S.nunits, 256, 0, st>>>(d_in, d_units, S.nunits, d_out);
  LAUNCH_CHECK();
  uint32_t u0 = 0, nbatches = 0;
  while (u0 < S.nunits) {
    const uint64_t base = h_units[u0].out_off;
    uint32_t u1 = u0 + 1;
    while (u1 < S.nunits && h_units[u1].out_off + h_units[u1].out_len - base <= batch_out) u1++;
    const uint32_t nb = u1 - u0;
    const GzUnit *bu = d_units + u0;
    nbatches++;
    gz_decode_kernel<<<(nb + kGzWarpsPerCta - 1) / kGzWarpsPerCta, 32 * kGzWarpsPerCta, 0, st>>>(d_in, n, bu, nb, d_sym, base, d_err);
    LAUNCH_CHECK();
    CU(ctx->gz_counts.ensure(sizeof(uint32_t) * (size_t)nb + 64));  // (the scan counters are done with)
    uint32_t *d_fix = ctx->gz_counts.as<uint32_t>();
    gz_tails_plain<<<nb, 256, 0, st>>>(d_in, bu, nb, d_sym, base, d_out, d_fix);
    LAUNCH_CHECK();
    // parallel rounds, four at a time, until no tail holds an open reference; a stream that still has some after 48
    // rounds, or whose open tails stop halving, finishes in the serial kernel
    uint32_t *d_open = d_err + 1;  // 4 counters behind the error word
    uint32_t open_units = 1, before = 0xffffffffu;
    for (int batch = 0; batch < 12 && open_units; batch++) {
      CU(cudaMemsetAsync(d_open, 0, sizeof(uint32_t) * 4, st));
      for (int r = 0; r < 4; r++) {
        gz_tails_round<<<nb, 256, 0, st>>>(bu, nb, d_sym, base, d_out, d_fix, d_open + r);
        LAUNCH_CHECK();
      }
      CU(cudaMemcpyAsync(&open_units, d_open + 3, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      // rounds that do not at least halve the open tails are a long chain working its way down the stream one unit
      // per round (template-like text: a phrase handed down through every chunk): the serial kernel is faster there
      if (open_units > before / 2) break;
      before = open_units;
    }
    if (open_units) {
      gz_tails_kernel<<<1, 1024, 0, st>>>(bu, nb, d_sym, base, d_out, d_fix);
      LAUNCH_CHECK();
    }
    gz_bodies_kernel<<<nb, 256, 0, st>>>(bu, nb, d_sym, base, d_out);
    LAUNCH_CHECK();
    u0 = u1;
  }
  toc(ctx, kTimDeflate);
  uint32_t h_err = 0;
  CU(cudaMemcpyAsync(&h_err, d_err, sizeof h_err, cudaMemcpyDeviceToHost, st));
  std::vector<GzMember> members(S.nmembers);
  CU(cudaMemcpyAsync(members.data(), d_members, sizeof(GzMember) * S.nmembers, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (h_err) return fail(ctx, DSG_E_CORRUPT, "unit %u decodes differently in the second pass", h_err - 1);
  // 5. every member's CRC32 (ISIZE was checked by the chain walk)
  uint32_t *state = reinterpret_cast<uint32_t *>(ctx->scalars.as<uint64_t>() + kSCrc);
  tic(ctx, kTimCrc);
  for (uint32_t m = 0; m < S.nmembers; m++) {
    int rc = crc_reset(ctx);
    if (rc) return rc;
    if ((rc = crc_update(ctx, d_out + members[m].out_off, members[m].out_len))) return rc;
    uint32_t got[2] = {0, 0};
    CU(cudaMemcpyAsync(got, state, sizeof got, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (got[0] != members[m].crc)
      return fail(ctx, DSG_E_CORRUPT, "member %u: CRC32 %08x, trailer says %08x", m, got[0], members[m].crc);
    ctx->stats.crc32 = got[0];
  }
  toc(ctx, kTimCrc);
  return DSG_OK;
}
}  // namespace

extern "C" int dsg_dev_gunzip(dsg_ctx *ctx, const uint8_t *d_in, size_t in_len, uint8_t *d_out, size_t out_cap,
                              size_t *out_len) {
  if (!ctx || !d_in || !out_len || (out_cap && !d_out)) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  tic(ctx, kTimTotal);
  uint64_t produced = 0;
  rc = gunzip_device(ctx, d_in, in_len, d_out, out_cap, &produced);
  *out_len = (size_t)produced;
  if (rc) return rc;
  toc(ctx, kTimTotal);
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->stats.payload_bytes = in_len;
  finish_stats(ctx, produced);
  return DSG_OK;
}

extern "C" int dsg_gunzip(dsg_ctx *ctx, const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, size_t *out_len) {
  if (!ctx || !in || !out_len || (out_cap && !out)) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(ctx->mu);
  int rc = begin_call(ctx);
  if (rc) return rc;
  CU(ctx->gz_in.ensure(in_len + 64));
  // ISIZE of the last member bounds nothing (mod 2^32, several members): the output buffer is sized from out_cap
  CU(ctx->gz_out.ensure(out_cap + 64));
  tic(ctx, kTimTotal);
  CU(cudaMemcpyAsync(ctx->gz_in.p, in, in_len, cudaMemcpyHostToDevice, ctx->stream));
  uint64_t produced = 0;
  rc = gunzip_device(ctx, ctx->gz_in.as<uint8_t>(), in_len, ctx->gz_out.as<uint8_t>(), out_cap, &produced);
  *out_len = (size_t)produced;
  if (rc) return rc;
  if (produced) CU(cudaMemcpyAsync(out, ctx->gz_out.p, produced, cudaMemcpyDeviceToHost, ctx->stream));
  toc(ctx, kTimTotal);
  CU(cudaStreamSynchronize(ctx->stream));
  ctx->stats.payload_bytes = in_len;
  ctx->stats.h2d_bytes = in_len;
  ctx->stats.d2h_bytes = produced;
  finish_stats(ctx, produced);
  return DSG_OK;
}
