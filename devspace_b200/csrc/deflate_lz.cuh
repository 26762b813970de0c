// deflate_lz.cuh -- block-parallel DEFLATE for sm_100a: LZ77 match finding in shared memory, per-chunk dynamic
// Huffman tables, warp-scan bit packing; the compress/flate level-6 writer behind gzip.NewWriter
// (/root/reference/pkg/devspace/sync/tar.go:156) re-designed for the GPU [Go stdlib compress/flate is not
// vendored; the published algorithm family and level-6 parameters are restated in oracle/gogzip.c].
//
// Stream structure (pigz style, BASELINE.json north_star): the input is cut into 64 KiB chunks. A chunk is
// compressed by one CTA into one deflate block (dynamic, fixed or stored, whichever is smallest) followed by an
// empty stored block (sync flush marker 00 00 ff ff) so that every chunk ends on a byte boundary and chunks
// concatenate by byte copy. The previous 32 KiB of INPUT serve as the chunk's dictionary (same member, so
// back-references into it are legal), which keeps the size within ~1 % of a serial level-6 stream.
//
// One persistent kernel (grid = #SMs, 1024 threads, 64 registers, 224 KB shared memory, 1 CTA/SM):
//   per run     ticket = atomicAdd(counter): a run of up to 8 consecutive chunks with ONE sliding window (chunks of
//               16 / 32 KiB for small inputs, 64 / 128 / 256 KiB for large ones; the last chunks of a launch one by one)
//   per chunk   a sampled byte histogram of every 8 KiB: flat chunks (random bytes) whose first tiles find no match
//               leave through a stored exit; the window is rebuilt from the stream behind them
//   per 1 KiB tile, by one of EIGHT teams of 4 warps (a team takes the next free tile when it has emitted its own;
//   linking and emission stay in tile order through one-shot mbarrier signals; at most eight consecutive tiles in flight):
//     load      cp.async.bulk (the TMA engine's 1-D copy) + mbarrier into the 40 KiB ring (32 KiB window + one tile
//               per team + look-ahead), issued before the wait for the previous tile's hashing
//     insert    13-bit hash of 4 bytes; EXACT-order chains in two parts: a stable partition of the tile's positions by
//               bucket class (not ordered), then one warp per class links its positions 32 per step (in tile order;
//               lap-stamped head words, no CAS); run interiors are not inserted
//     search    shallow pass over EVERY position (depth 1, nice 64, good 8): per-lane state machine over runs of 8
//               positions handed out through a shared counter; a position inherits (len-1, dist) from its predecessor
//     refine    stateless lazy parse from the tile start -> work list of the token starts and lazy successors whose
//               shallow walk hit the depth limit -> deep search of only those (depth 160 / 40, links chased four
//               ahead); a match that grows puts the position behind it on a second list, searched at depth 53 right
//               behind the first (what a second parse + list + search round did, at 1/3 of its cost)
//     parse     stateless lazy rule next(p) = p + (len[p] >= 4 && len[p+1] <= len[p] ? len[p] : 1) from the carried
//               position: one bitmap word per lane, exits travel by shuffle until nothing changes
//     emit      token starts -> team-wide rank (scan) -> 32-bit tokens to a per-CTA global scratch (L2 resident) +
//               shared-memory histograms
//   per chunk   Huffman code lengths (rank sort, two-queue merge, zlib-style overflow repair), code-length RLE, exact
//               sizes of the three encodings -> chunk size known BEFORE any bit is written
//               bit packing: per-token bit lengths -> block scan -> atomicOr into a shared-memory window ->
//               128-bit stores into the chunk's slot of a scratch buffer (stored chunks write nothing)
// followed by lz_scan_kernel (exclusive prefix of the chunk sizes) and lz_gather_kernel (slots -> final byte
// offsets, stored chunks straight from the input). A first version placed the output with a decoupled
// look-back inside the main kernel; CTAs then spent 20 % of their time waiting for slower predecessors
// (profiles/r01_lz_phases.md), and the wait chain serialises completely once a CTA owns a run of chunks.
// Roofline: the kernel is bound by shared-memory bandwidth / LSU issue (random 1-2 byte reads), not by HBM:
// algorithmic HBM bytes per chunk are 64 KiB read (+32 KiB dictionary re-read from L2) + C written.
#pragma once

#include "copy_util.cuh"
#include "deflate_stored.cuh"
#include "huffman.cuh"

namespace dsg {

constexpr int kLzThreads = 1024;
constexpr int kLzWarps = kLzThreads / 32;
constexpr int kChunk = 65536;
#if !defined(DSG_LZ_TEAMS)
#define DSG_LZ_TEAMS 8
#endif
constexpr int kTeams = DSG_LZ_TEAMS;     // the CTA works as eight teams of 4 warps (4: four teams of 8), each on a tile of its own (see the kernel)
constexpr int kTeamThreads = kLzThreads / kTeams;
constexpr int kTeamWarps = kTeamThreads / 32;
constexpr int kTile = 8 * kTeamThreads;  // 2048 (1024 with eight teams): eight positions per thread of the team
constexpr int kDict = 32768;
constexpr int kRing = kDict + kTeams * kTile;  // 40960 ring slots (bytes): the window + one tile per team
constexpr int kLook = 272;               // look-ahead beyond a tile (>= 258 + 3, multiple of 16)
constexpr int kMaxDist = kDict - kLook;  // 32496
constexpr int kHashBits = 13;
constexpr int kHashSize = 1 << kHashBits;
constexpr int kMinMatch = 4;
constexpr int kMaxMatch = 258;
#if !defined(DSG_LZ_DEPTH)
#define DSG_LZ_DEPTH 1
#endif
constexpr int kDepth = DSG_LZ_DEPTH;          // first pass: every position, shallow
// The search constants below can be overridden on the compiler's command line (the ratio experiments behind
// profiles/r02_lz_phases.md were run that way, under the emulator for the size and on the GPU for the speed).
#if !defined(DSG_LZ_DEEP)
#define DSG_LZ_DEEP 160
#endif
constexpr int kDeepDepth = DSG_LZ_DEEP;    // second pass: token starts and their lazy successors only
constexpr int kRefineRounds = 1;   // parse + list + deep pass over the list (two until the second list below replaced the second round)
#if !defined(DSG_LZ_CHAIN_DIV)
#define DSG_LZ_CHAIN_DIV 3
#endif
constexpr int kChainDepthDiv = DSG_LZ_CHAIN_DIV;  // depth of the second list: the positions behind the matches the deep pass made longer
#if !defined(DSG_LZ_CMPS)
#define DSG_LZ_CMPS 16
#endif
constexpr int kMaxCompares = DSG_LZ_CMPS;   // full compares per position
#if !defined(DSG_LZ_NICE)
#define DSG_LZ_NICE 64
#endif
constexpr int kNice = DSG_LZ_NICE;  // (zlib level 6: 128; 64 gives the same sizes -- text 1.0185, binary mix 1.0020 against 1.0019 -- 1.2 % faster)
constexpr int kGood = 8;
constexpr int kRun = kTile / kTeamThreads;  // 8 consecutive positions per thread (list building, token emission)
constexpr uint16_t kNil = 0xffff;
constexpr int kStageBytes = 16384;        // bit-packing window
constexpr int kBatchTokens = 2048;        // tokens per bit-packing batch (<= 48 bits each = 12 KiB)
constexpr int kTokPerThread = kBatchTokens / kLzThreads;
constexpr uint32_t kTokMatch = 0x80000000u;
// kChunk is the unit; a launch may use chunks of 1, 2 or 4 units (LzParams::chunk_bytes): a chunk is one deflate
// block with one pair of Huffman tables, and its epilogue (tree construction, sizes, bit packing, the drain of the
// tile pipeline) is the part of the kernel in which most threads wait -- large inputs amortise it over more bytes.
constexpr int kMaxChunk = 4 * kChunk;
constexpr int kMinChunk = kChunk / 4;  // small inputs: 16 / 32 KiB chunks (the host decides, dsgpack.cu lz_geometry)
constexpr int kTokSlotWords = kMaxChunk + 64;  // per-CTA token scratch (u32), worst case all literals + EOB
DSG_HD uint32_t lz_slot_bytes(uint32_t chunk_bytes) { return chunk_bytes + 128; }  // per-chunk output slot (16-byte aligned)
// The first big_chunks chunks of a launch have chunk_bytes bytes (they are handed out in runs), the chunks behind
// them kChunk bytes (handed out one by one at the end of the launch, where a big chunk would leave SMs idle for long).
DSG_HD uint64_t lz_chunk_offset(uint32_t c, uint32_t chunk_bytes, uint32_t big_chunks) {
  return c < big_chunks ? (uint64_t)c * chunk_bytes
                        : (uint64_t)big_chunks * chunk_bytes + (uint64_t)(c - big_chunks) * (uint64_t)kChunk;
}
DSG_HD uint32_t lz_chunk_size(uint32_t c, uint32_t chunk_bytes, uint32_t big_chunks) {
  return c < big_chunks ? chunk_bytes : (uint32_t)kChunk;
}

// Upper bound of the raw deflate output for n input bytes: a chunk never exceeds its stored encoding (one stored
// block for a 16 KiB chunk, two for a 64 KiB one: at most 10 bytes per 16 KiB), plus the 2-byte terminator.
DSG_HD uint64_t lz_deflate_bound(uint64_t n) { return n + 10 * ((n + kMinChunk - 1) / kMinChunk) + 2 + 64; }

// Shared-memory carve-up (dynamic shared memory; byte offsets, 16-byte aligned)
struct LzLayout {
  static constexpr int data = 0;                                  // kRing + kLook + 48 bytes (mirror + slack)
  static constexpr int prev = data + kRing + kLook + 48;          // u16[kRing]
  static constexpr int head = prev + 2 * kRing;                   // u32[kHashSize]
  static constexpr int mlen = head + 4 * kHashSize;               // u8[kTeams][kTile]   } reused as the bit-packing
  static constexpr int mdist = mlen + kTeams * kTile;             // u16[kTeams][kTile]  } window after the last tile
  static constexpr int bitmap = mdist + 2 * kTeams * kTile;       // u32[kTeams][kTile/32 + 16] token-start flags + exits
  static constexpr int lfreq = bitmap + kTeams * (kTile / 8 + 64); // u32[288]
  static constexpr int dfreq = lfreq + 4 * 288;                   // u32[32]
  static constexpr int llen = dfreq + 4 * 32;                     // u8[288] literal/length code lengths
  static constexpr int dlen = llen + 288;                         // u8[32]
  static constexpr int lcode = dlen + 32;                         // u16[288]
  static constexpr int dcode = lcode + 2 * 288;                   // u16[32]
  static constexpr int cllen = dcode + 2 * 32;                    // u8[32]
  static constexpr int clcode = cllen + 32;                       // u16[32]
  static constexpr int rle = clcode + 2 * 32;                     // u16[320]
  static constexpr int misc = rle + 2 * 320;                      // u32[288] scalars / scan scratch / mbarriers
  static constexpr int huff = misc + 4 * 288 + 16;                 // 2 x HuffScratch + comb[512] + ctrl[80]
  static constexpr int work = huff + 2 * 4064 + 512 + 4 * 80 + 16; // u16[kTeams][kTile] positions for the deep pass
  static constexpr int deep = work + 2 * kTeams * kTile;           // u32[kTeams][kTile/32] positions searched deeply
  static constexpr int chain = deep + kTeams * kTile / 8 + 16;     // u16[kTeams][kTile] second list of the deep pass
  static constexpr int total = chain + 2 * kTeams * kTile;
};
static_assert(sizeof(HuffScratch) <= 4064, "HuffScratch grew");
static_assert(LzLayout::prev % 16 == 0 && LzLayout::head % 16 == 0 && LzLayout::mlen % 16 == 0, "alignment");
static_assert(kTeams * 3 * kTile >= kStageBytes, "bit-packing window must fit in the match arrays");
static_assert(LzLayout::total <= 227 * 1024, "shared memory budget");
static_assert(kMaxDist % 16 == 0 && kLook % 16 == 0, "ring loads move 16-byte granules");

// misc[] slots
enum { kMTicket = 0, kMStart, kMNTok, kMMode, kMOutLo, kMOutHi, kMHdrBits, kMChunkBytes, kMHlit, kMHdist, kMHclen,
       kMNRle, kMBodyDyn, kMBodyFix, kMError, kMRunCtr /* kTeams slots */, kMRunStart = kMRunCtr + kTeams,
       kMFlat /* every 8 KiB of the chunk has a flat byte histogram */, kMProbe /* matches found in the first tile of
       every team */, kMStale /* the window has to be rebuilt before the next chunk */,
       kMScan = 32 /* CTA scan: 32 warp partials + carry; team scans: 20 words each */, kMTeamScan = kMScan,
       kMTeamScanStride = 20, kMSig = 192 /* kSigCount hand-off signals, 8 bytes each */,
       kMBulkBar = 256 /* kTeams mbarriers, 8 bytes each */, kMTileCtr = 272 /* next tile of the chunk */,
       kMTeamTile = 273 /* kTeams slots: the tile a team has just taken */, kMiscWords = 288 };
static_assert(kMBulkBar % 2 == 0 && kMBulkBar + 2 * kTeams <= kMTileCtr && kMSig % 2 == 0 && kMSig + 2 * 32 <= kMBulkBar &&
                  kMTeamTile + kTeams <= kMiscWords, "mbarrier words in misc[]");
static_assert(kMStale < kMScan && kRun == 8 && kMTeamScan + kTeams * kMTeamScanStride <= kMSig, "misc[] slots");
static_assert(kTeams == 4 || kTeams == 8, "teams");

// Kernel parameter block
struct LzParams {
  const uint8_t *in;          // input stream (this call's bytes)
  uint64_t len;               // bytes
  const uint8_t *dict;        // up to 32 KiB that precede `in` in the member (previous segment), or null
  uint32_t dict_len;
  uint32_t chunk_bytes;       // kChunk, 2 * kChunk or kMaxChunk
  uint8_t *slots;             // per-chunk output slots, lz_slot_bytes(chunk_bytes) apart (scratch)
  uint32_t *chunk_meta;       // per chunk: (bytes << 1) | stored flag (scratch)
  uint32_t nchunks;
  uint32_t run_chunks;        // consecutive chunks one ticket hands out (the LZ window slides across them)
  uint32_t big_runs;          // the first big_runs tickets are runs of run_chunks chunks, the rest single chunks
  uint32_t *ticket;           // zeroed before launch
  uint32_t *tok_scratch;      // gridDim.x * kTokSlotWords
  unsigned long long *prof;   // optional: cycles per phase, summed over CTAs (thread 0's clock64 deltas)
  uint32_t no_bulk;           // tuning knob: stage the tiles with plain loads instead of cp.async.bulk
  uint32_t depth, deep_depth, refine_rounds;  // search effort (kDepth, kDeepDepth, kRefineRounds unless the tools say otherwise)
};

enum { kProfLoad = 0, kProfInsert, kProfSearch, kProfParse, kProfEmit, kProfHuff, kProfSizes, kProfPack, kProfRefParse,
       kProfRefList, kProfRefDeep, kProfWaitA, kProfWaitB, kProfJoin, kProfLink, kProfFlat, kProfCount,
       // not phases: sum over warps of (own finish - phase start) / 32, to tell the tail of a phase from its bulk
       kProfWarpSearch = kProfCount, kProfWarpDeep, kProfSlots };

// named barriers: 0 = the CTA
enum { kBarTeam = 1 /* +team */, kBarEnd = kBarTeam + kTeams };
static_assert(kBarEnd <= 16, "an SM has 16 named barriers");

DSG_D uint32_t ring_slot(int rp /* run-relative position, >= -kDict */) {
  return (uint32_t)(rp + kDict) % (uint32_t)kRing;
}

// A thread group inside the CTA: all threads (barrier 0 = __syncthreads) or warps 1.. (named barrier 1), used to
// overlap the single-warp parse of one tile with loading and hashing the next tile.
struct LzGroup {
  int tid;   // index inside the group
  int n;     // threads in the group (multiple of 32)
  int bar;   // 0 = the whole CTA
};
DSG_D void group_sync(const LzGroup &g) {
  if (g.bar == 0) {
    __syncthreads();
  } else {
#if defined(DSG_EMU)
    emu::named_barrier_sync(g.bar, g.n);
#else
    asm volatile("bar.sync %0, %1;" ::"r"(g.bar), "r"(g.n) : "memory");
#endif
  }
}

// Tile hand-offs between teams: "tile t is in the hash chains" and "tile t has emitted its tokens", each a one-shot
// signal. A signal lives in a ring of kSigRing slots (slot t % kSigRing, use t / kSigRing) -- at most kTeams consecutive
// tiles are in flight, so a slot is never armed again before its waiters have passed. On the GPU a slot is an mbarrier
// with arrival count 1: ONE thread of the producing team arrives (release) after the team's own barrier, the
// threads of the consuming team wait on the phase parity (try_wait sleeps in hardware instead of holding a named
// barrier with 512 participants). Under the emulator a slot is a counter and the wait yields the fiber. The slots are
// initialised once per CTA; tiles are numbered through all chunks the CTA compresses (seq_base), and EVERY tile that is
// taken arrives exactly once on both of its signals, so slot and phase follow from the tile's number alone.
constexpr int kSigRing = 16;
enum { kSigIns = 0, kSigEmit = kSigRing, kSigCount = 2 * kSigRing };
static_assert(kSigRing >= 2 * kTeams, "signal ring");

#if defined(DSG_EMU)
DSG_D void sig_init(uint32_t *sig, int i) { sig[2 * i] = 0; }
DSG_D void sig_arrive(uint32_t *sig, int i) {
  __threadfence_block();
  atomicAdd(&sig[2 * i], 1u);
}
DSG_D void sig_wait(uint32_t *sig, int i, uint32_t use) {
  while (*reinterpret_cast<volatile uint32_t *>(&sig[2 * i]) < use + 1u) emu::yield();
}
#else
DSG_D void sig_init(uint32_t *sig, int i) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(sig + 2 * i);
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(1) : "memory");
}
DSG_D void sig_arrive(uint32_t *sig, int i) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(sig + 2 * i);
  asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(a) : "memory");
}
DSG_D void sig_wait(uint32_t *sig, int i, uint32_t use) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(sig + 2 * i);
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "SWAIT_%=:\n"
      "mbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%0], %1;\n"
      "@p bra SDONE_%=;\n"
      "bra SWAIT_%=;\n"
      "SDONE_%=:\n"
      "}\n" ::"r"(a),
      "r"(use & 1u)
      : "memory");
}
#endif

// Shared-memory loads through 32-bit shared-window addresses. Inside the chain-walk loops the compiler otherwise
// keeps the arrays as generic pointers and rebuilds the window base (S2UR/ULEA) in every iteration.
#if defined(DSG_EMU)
struct SmemBase { const uint8_t *p; };
DSG_D SmemBase smem_base(const void *p) { return SmemBase{static_cast<const uint8_t *>(p)}; }
DSG_D uint32_t lds_u8(SmemBase b, uint32_t off) { return b.p[off]; }
DSG_D uint32_t lds_u16(SmemBase b, uint32_t off) {
  return (uint32_t)b.p[off] | ((uint32_t)b.p[off + 1] << 8);
}
#else
struct SmemBase { uint32_t a; };
DSG_D SmemBase smem_base(const void *p) { return SmemBase{(uint32_t)__cvta_generic_to_shared(p)}; }
DSG_D uint32_t lds_u8(SmemBase b, uint32_t off) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(b.a + off));
  return v;
}
DSG_D uint32_t lds_u16(SmemBase b, uint32_t off) {
  uint32_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(b.a + off));
  return v;
}
#endif

DSG_D uint32_t hash4(uint32_t v) { return (v * 2654435761u) >> (32 - kHashBits); }

// 4 bytes at an arbitrary shared-memory byte offset (little endian)
DSG_D uint32_t load4(const uint8_t *p) {
  const uint32_t m = (uint32_t)((uintptr_t)p & 3);
  const uint32_t *a = reinterpret_cast<const uint32_t *>(p - m);
  uint32_t lo = a[0];
  if (m == 0) return lo;
  uint32_t hi = a[1];
  return __funnelshift_r(lo, hi, 8 * m);
}

// Block-wide exclusive scan of one value per thread (all kLzThreads threads). Returns the exclusive prefix and the
// block total through *total. `scratch` has >= 17 words. Two barriers.
DSG_D uint32_t block_exclusive_scan(uint32_t v, uint32_t *scratch, uint32_t *total) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (uint32_t)o) x += y;
  }
  if (lane == 31) scratch[warp] = x;
  __syncthreads();
  if (warp == 0) {
    uint32_t t = lane < (uint32_t)kLzWarps ? scratch[lane] : 0;
    uint32_t s = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= (uint32_t)o) s += y;
    }
    if (lane < (uint32_t)kLzWarps) scratch[lane] = s - t;
    if (lane == kLzWarps - 1) scratch[kLzWarps] = s;
  }
  __syncthreads();
  uint32_t r = scratch[warp] + (x - v);
  *total = scratch[kLzWarps];
  return r;
}

// The same scan over the threads of a group (a team: kTeamWarps warps, its own named barrier). `scratch` >= 17 words.
DSG_D uint32_t group_exclusive_scan(uint32_t v, uint32_t *scratch, uint32_t *total, const LzGroup &g) {
  const uint32_t lane = (uint32_t)g.tid & 31, warp = (uint32_t)g.tid >> 5, nw = (uint32_t)g.n >> 5;
  uint32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (uint32_t)o) x += y;
  }
  if (lane == 31) scratch[warp] = x;
  group_sync(g);
  if (warp == 0) {
    uint32_t t = lane < nw ? scratch[lane] : 0;
    uint32_t s = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= (uint32_t)o) s += y;
    }
    if (lane < nw) scratch[lane] = s - t;
    if (lane == nw - 1) scratch[nw] = s;
  }
  group_sync(g);
  uint32_t r = scratch[warp] + (x - v);
  *total = scratch[nw];
  return r;
}

// Exclusive running maximum over the threads of a group (identity 0). `scratch` >= 17 words (33 for the CTA).
DSG_D uint32_t group_exclusive_max(uint32_t v, uint32_t *scratch, uint32_t *total, const LzGroup &g) {
  const uint32_t lane = (uint32_t)g.tid & 31, warp = (uint32_t)g.tid >> 5, nw = (uint32_t)g.n >> 5;
  uint32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (uint32_t)o) x = x > y ? x : y;
  }
  uint32_t ex = __shfl_up_sync(0xffffffffu, x, 1);
  if (lane == 0) ex = 0;
  if (lane == 31) scratch[warp] = x;
  group_sync(g);
  if (warp == 0) {
    const uint32_t t = lane < nw ? scratch[lane] : 0;
    uint32_t m = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, m, o);
      if (lane >= (uint32_t)o) m = m > y ? m : y;
    }
    uint32_t before = __shfl_up_sync(0xffffffffu, m, 1);
    if (lane == 0) before = 0;
    if (lane < nw) scratch[lane] = before;
    if (lane == nw - 1) scratch[nw] = m;
  }
  group_sync(g);
  const uint32_t w = scratch[warp];
  *total = scratch[nw];
  return ex > w ? ex : w;
}

// Sequential LSB-first bit writer into a zeroed shared-memory window (single thread).
struct SmemBitWriter {
  uint32_t *w;
  uint32_t pos;  // bit position
  DSG_D void put(uint32_t v, uint32_t n) {
    if (n == 0) return;
    uint32_t i = pos >> 5, s = pos & 31;
    w[i] |= v << s;
    if (s + n > 32) w[i + 1] |= v >> (32 - s);
    pos += n;
  }
};

// ---------------------------------------------------------------------------------------------------------
// incompressible chunks
// ---------------------------------------------------------------------------------------------------------
// Random bytes (compressed or encrypted payloads: archives, images, .gz files inside the tree) used to cost the full
// search, parse and Huffman work -- 1.04 M cycles per 64 KiB, a third of a text chunk -- only to end as a stored
// block. A chunk now takes the stored exit early when BOTH hold:
//   (1) every 8 KiB of it has a flat byte histogram: 2 KiB sampled (16 of every 64 bytes), chi-square against the
//       uniform distribution below kFlatChi2 (uniform data: 255 +- 23; base64, text, machine code, tables: thousands).
//       Integer arithmetic only, so the GPU and the emulator decide alike;
//   (2) the first tile of every team (the first 8 KiB, searched normally against the 32 KiB window) found at most
//       kFlatMaxMatches matches -- a copy of earlier incompressible bytes (a duplicated image) shows up here.
// What this gives up: a repeat that begins later than 8 KiB into an all-flat chunk is not used for that chunk.
// The window is rebuilt from the stream at the next chunk (kMStale), which therefore sees its full dictionary.
constexpr int kFlatBlock = 8192;
constexpr int kFlatSampleStride = 64;   // one 16-byte granule of every 64 bytes
constexpr int kFlatMinSamples = 1024;
constexpr int kFlatChi2 = 480;          // ten standard deviations above the mean of uniform data
constexpr int kFlatMaxMatches = 8;

// All threads of the CTA. hist: 32 x 128 words (one u16 histogram per warp). Writes 1 / 0 to *flat. Ends with a barrier.
DSG_D void chunk_flatness(const uint8_t *in, uint64_t len, int64_t cs, int chunk_len, uint32_t *hist, uint32_t *flat) {
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) *flat = ((((uintptr_t)(in + cs)) & 15) == 0 && chunk_len >= kFlatBlock) ? 1u : 0u;
  __syncthreads();
  if (*flat == 0u) return;  // uniform
  uint32_t *h = hist + warp * 128;
  const int nblocks = (chunk_len + kFlatBlock - 1) / kFlatBlock;
  for (int b = (int)warp; b < nblocks; b += kLzWarps) {
    for (uint32_t i = lane; i < 128; i += 32) h[i] = 0;
    __syncwarp();
    const int blen = chunk_len - b * kFlatBlock < kFlatBlock ? chunk_len - b * kFlatBlock : kFlatBlock;
    uint32_t n = 0;
    // the block's samples are four granules per lane: all loads are issued before the first histogram update (the
    // bytes come from DRAM, nobody has touched them yet; one after the other they cost four round trips per chunk
    // with the whole CTA waiting at the barrier behind this function)
    constexpr int kPer = kFlatBlock / (32 * kFlatSampleStride);  // 4
    uint4 v[kPer];
    bool have[kPer];
#pragma unroll
    for (int k = 0; k < kPer; k++) {
      const int o = ((int)lane + 32 * k) * kFlatSampleStride;
      const int64_t g = cs + (int64_t)b * kFlatBlock + o;
      have[k] = o + 16 <= blen && g + 16 <= (int64_t)len;
      v[k] = make_uint4(0, 0, 0, 0);
      if (have[k]) {
#if defined(DSG_EMU)
        v[k] = *reinterpret_cast<const uint4 *>(in + g);
#else
        v[k] = __ldg(reinterpret_cast<const uint4 *>(in + g));
#endif
      }
    }
#pragma unroll
    for (int k = 0; k < kPer; k++) {
      if (!have[k]) continue;
      const uint32_t w[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
      for (int q = 0; q < 4; q++) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const uint32_t byte = (w[q] >> (8 * j)) & 0xffu;
          atomicAdd(&h[byte >> 1], (byte & 1u) ? 0x10000u : 1u);
        }
      }
      n += 16;
    }
    __syncwarp();
    // sum of squared counts, 8 bins per lane
    uint32_t sq = 0;
    for (uint32_t i = lane; i < 128; i += 32) {
      const uint32_t c0 = h[i] & 0xffffu, c1 = h[i] >> 16;
      sq += c0 * c0 + c1 * c1;
    }
    for (int o = 16; o > 0; o >>= 1) {
      sq += __shfl_xor_sync(0xffffffffu, sq, o);
      n += __shfl_xor_sync(0xffffffffu, n, o);
    }
    // chi2 = 256 / n * sum(c^2) - n < kFlatChi2
    const bool ok = n >= (uint32_t)kFlatMinSamples && 256ull * sq < (uint64_t)n * (n + (uint32_t)kFlatChi2);
    if (lane == 0 && !ok) atomicAnd(flat, 0u);
    __syncwarp();
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------------------------------------
// tile phases
// ---------------------------------------------------------------------------------------------------------

// Copies stream bytes [g0, g1) (absolute offsets in the call's input, may be negative = dictionary from the
// previous segment, may exceed len = zero fill) to the ring slots of chunk-relative positions rp0...
DSG_D void load_ring(uint8_t *s_data, const LzParams &P, int64_t cs, int rp0, int rp1, const LzGroup &g) {
  // 16-byte granules: rp0 and rp1 are multiples of 16 and the ring size is a multiple of 16
  for (int rp = rp0 + 16 * g.tid; rp < rp1; rp += 16 * g.n) {
    const int64_t g = cs + rp;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (g >= 0 && g + 16 <= (int64_t)P.len && (((uintptr_t)(P.in + g)) & 15) == 0) {
      // streamed: read once; the per-CTA token lists are what should stay in L2
#if defined(DSG_EMU)
      v = *reinterpret_cast<const uint4 *>(P.in + g);
#else
      v = __ldcs(reinterpret_cast<const uint4 *>(P.in + g));
#endif
    } else {
      uint32_t w[4] = {0, 0, 0, 0};
      for (int k = 0; k < 16; k++) {
        const int64_t gg = g + k;
        uint32_t b = 0;
        if (gg >= 0) {
          if (gg < (int64_t)P.len) b = P.in[gg];
        } else if (P.dict && -gg <= (int64_t)P.dict_len) {
          b = P.dict[(int64_t)P.dict_len + gg];
        }
        w[k >> 2] |= b << (8 * (k & 3));
      }
      v = make_uint4(w[0], w[1], w[2], w[3]);
    }
    const uint32_t slot = ring_slot(rp);
    *reinterpret_cast<uint4 *>(s_data + slot) = v;
    if (slot < (uint32_t)(kLook + 48)) {  // mirror of the ring start behind its end: reads never wrap
      if (slot + 16 <= (uint32_t)(kLook + 48)) *reinterpret_cast<uint4 *>(s_data + kRing + slot) = v;
    }
  }
}

// ---- bulk staging of a tile: cp.async.bulk (the TMA engine's 1-D copy) + mbarrier ------------------------------------
// A tile's bytes travel global -> shared memory without passing through registers: ONE thread of the team arms the
// team's mbarrier with the byte count and issues the copies (a tile is one to two ring pieces plus the mirror of
// the ring start), the team waits on the mbarrier phase right before it hashes the tile. Source: this GPU's own
// global memory, no tensor map, no multicast. Needs 16-byte aligned source, destination and sizes: tiles at the
// edges of the stream (dictionary of the previous segment, zero fill behind the end) take load_ring instead.
#if !defined(DSG_EMU)
DSG_D void mbar_init(uint32_t mbar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
DSG_D void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
DSG_D void mbar_wait(uint32_t mbar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(mbar),
      "r"(parity)
      : "memory");
}
DSG_D void bulk_g2s(uint32_t dst_smem, const void *src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src), "r"(bytes), "r"(mbar)
               : "memory");
}
#endif

// True when the positions [rp0, rp1) of the run can be staged by bulk copies.
DSG_D bool bulk_ok(const LzParams &P, int64_t cs, int rp0, int rp1) {
#if defined(DSG_EMU)
  (void)P; (void)cs; (void)rp0; (void)rp1;
  return false;
#else
  if (P.no_bulk) return false;
  const int64_t g0 = cs + rp0, g1 = cs + rp1;
  return g0 >= 0 && g1 <= (int64_t)P.len && ((((uintptr_t)P.in) | (uint64_t)g0 | (uint64_t)g1) & 15) == 0;
#endif
}

// One thread: arms `mbar` and issues the copies of stream bytes [cs + rp0, cs + rp1) to their ring slots (and to the
// mirror behind the ring end). rp0, rp1 multiples of 16, rp1 - rp0 <= kRing.
DSG_D void bulk_load_ring(uint8_t *s_data, const LzParams &P, int64_t cs, int rp0, int rp1, uint32_t mbar) {
#if !defined(DSG_EMU)
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(s_data);
  const uint8_t *src = P.in + cs + rp0;
  const uint32_t n = (uint32_t)(rp1 - rp0);
  const uint32_t s0 = ring_slot(rp0);
  const uint32_t first = s0 + n <= (uint32_t)kRing ? n : (uint32_t)kRing - s0;  // piece up to the ring end
  const uint32_t mir = (uint32_t)(kLook + 48);
  // mirror parts: slots [0, mir) of either piece
  const uint32_t m1 = s0 < mir ? (s0 + first < mir ? first : mir - s0) : 0u;    // of the first piece (starts at s0)
  const uint32_t m2 = n - first < mir ? n - first : mir;                        // of the wrapped piece (starts at 0)
  // generic-proxy accesses of these slots (older laps) are ordered before the async-proxy writes
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  mbar_expect_tx(mbar, n + m1 + m2);
  bulk_g2s(base + s0, src, first, mbar);
  if (m1) bulk_g2s(base + (uint32_t)kRing + s0, src, m1, mbar);
  if (n > first) {
    bulk_g2s(base, src + first, n - first, mbar);
    bulk_g2s(base + (uint32_t)kRing, src + first, m2, mbar);
  }
#else
  (void)s_data; (void)P; (void)cs; (void)rp0; (void)rp1; (void)mbar;
#endif
}

// Inserts positions [rp0, rp1) into the hash chains in EXACT stream order (nearest candidate first). Exact order
// matters: chains that are only ordered per round cost 1-3 % of compressed size (tools/lzmodel.cc, SCRAMBLE=512).
// Two steps per batch of positions:
//   pre-hash  all threads of the group, in parallel: the bucket of every position (or "not inserted") into `hpre`;
//   link      ONE warp, 32 consecutive positions per step: MATCH.ANY groups the lanes of a bucket, a lane links to
//             the nearest lower lane of its group (same step: the intervening positions are in other buckets) or,
//             if it is the first of its group, to the bucket's head; the last lane of a group becomes the head.
//             No compare-and-swap, no retries, no barrier between steps (a warp's shared-memory accesses are
//             performed in program order; __syncwarp orders them between lanes).
// The link step is the one strictly ordered step of the tile pipeline (tile t after tile t-1), so its LATENCY
// bounds the pipeline: 64 short steps per tile instead of 8 rounds of racing sorted insertions (a CAS loop per
// position with a barrier per round, profiles/r02_lz_phases.md: 33 k cycles per tile, and the teams queued for it).
// Positions in the interior of a byte run (the five bytes p-1..p+3 are equal) are not inserted: every one of
// them would land in the same bucket (useless candidates); the search gives them the distance-1 candidate
// directly. Their chain start is the start of their run (second pass of insert_positions). Nothing here depends
// on how the threads interleave -- the output of the kernel is the same on every run (tools/determinism.py).
constexpr int kInsRound = kLzThreads;  // slots this close to a lap behind the current position count as gone (head_entry)
constexpr uint32_t kNoInsert = 0xffffu;

// interior of a byte run: the five bytes p-1..p+3 are equal
DSG_D bool run_interior(const uint8_t *s_data, uint32_t slot, uint32_t v, int rp, int lo_rp) {
  return rp > lo_rp && v == (v & 0xffu) * 0x01010101u && s_data[slot ? slot - 1 : kRing - 1] == (uint8_t)v;
}

// A head word holds the slot of the newest position of its bucket in the low half and that position's ring lap in
// the high half. The lap tells a live entry from one that is a whole lap (or more) old: its slot belongs to a newer
// position by now, and following it would walk another bucket's chain. Entries in the last kInsRound positions of
// the lap window are treated as gone too; they lie beyond kMaxDist anyway.
DSG_D uint32_t head_word(int rp) { return ring_slot(rp) | ((uint32_t)((rp + kDict) / kRing) << 16); }

// Slot of the entry a head word names, or kNil when the entry is not live as seen from position rp.
DSG_D uint32_t head_entry(uint32_t word, int rp) {
  const uint32_t s = word & 0xffffu;
  if (s == kNil) return kNil;
  const int cur = rp + kDict;                                   // ring position index of rp
  const int ent = (int)(word >> 16) * kRing + (int)s;           // ... of the entry
  const int delta = cur - ent;                                  // > 0: the entry is older
  if (delta >= kRing - kInsRound || delta <= 0) return kNil;
  return s;
}

// What a thread keeps between the two parts of an insertion.
struct InsPrep {
  uint32_t list_start, list_count;  // the class list of this thread's WARP inside l_pos / l_hv (uniform in the warp)
  uint32_t lead;                    // run interiors at the start of this thread's positions whose run starts before
                                    // the range: their chain start is the carried run start (ordered part)
  uint32_t last;                    // range index + 1 of the last position of the range that is not an interior, 0 = none
};

constexpr int kInsClasses = kTeamWarps;  // bucket classes = warps of a team
static_assert((kInsClasses == 8 || kInsClasses == 4) && kRun == 8 && kTile <= 65535, "packed class counters: 2 x 4 fields of 16 bits");

// Part one, NOT ordered (needs only the bytes of the range): buckets of positions [rp0, rp0 + n), n <= kTile, eight
// consecutive positions per thread; a stable partition of the inserted positions by bucket class (the low bits of the bucket) into
// l_pos (range index) / l_hv (bucket) -- class c occupies one contiguous stretch, in stream order; chain starts of
// the run interiors whose run starts inside the range. `scratch`: 40 words, 8-byte aligned, not in use by the group
// since its last barrier. One team barrier inside; the lists are complete behind the NEXT barrier of the group.
DSG_D InsPrep insert_prepare(const uint8_t *s_data, unsigned short *prev16, uint16_t *l_pos, uint16_t *l_hv,
                             uint32_t *scratch, int rp0, int n, int lo_rp, const LzGroup &g) {
  const uint32_t lane = (uint32_t)g.tid & 31u, w = (uint32_t)g.tid >> 5;
  const int i0 = g.tid * kRun;
  uint32_t hv[kRun];
  uint64_t ca = 0, cb = 0;  // positions per class: classes 0-3 and 4-7, 16 bits each
  uint32_t last = 0;
  const uint32_t slot0 = ring_slot(rp0 + i0);
  if (i0 < n) {
    // the thread's 8 + 3 bytes from four aligned words (reads run on into the mirror behind the ring end), the byte
    // before them on its own
    const uint32_t m8 = 8u * (slot0 & 3u);
    const uint32_t *wp = reinterpret_cast<const uint32_t *>(s_data + (slot0 & ~3u));
    const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
    const uint32_t x0 = __funnelshift_r(w0, w1, m8), x1 = __funnelshift_r(w1, w2, m8), x2 = __funnelshift_r(w2, w3, m8);
    uint32_t before = s_data[slot0 ? slot0 - 1 : kRing - 1];
#pragma unroll
    for (int j = 0; j < kRun; j++) {
      const int i = i0 + j;
      const uint32_t v = j < 4 ? __funnelshift_r(x0, x1, 8 * (j & 3)) : __funnelshift_r(x1, x2, 8 * (j & 3));
      // interior of a byte run (run_interior): the five bytes p-1..p+3 are equal
      const bool interior = rp0 + i > lo_rp && v == (v & 0xffu) * 0x01010101u && before == (v & 0xffu);
      before = v & 0xffu;
      hv[j] = kNoInsert;
      if (i < n && !interior) {
        hv[j] = hash4(v);
        last = (uint32_t)i + 1u;
        const uint32_t cls = hv[j] & (uint32_t)(kInsClasses - 1);
        const uint64_t one = 1ull << (16 * (cls & 3u));
        if (cls & 4u) cb += one;
        else ca += one;
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < kRun; j++) hv[j] = kNoInsert;
  }
  // inclusive scan inside the warp: class counts (sum) and the last non-interior position (max)
  uint64_t sa = ca, sb = cb;
  uint32_t sl = last;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint64_t ta = __shfl_up_sync(0xffffffffu, sa, o);
    const uint64_t tb = kInsClasses > 4 ? __shfl_up_sync(0xffffffffu, sb, o) : 0ull;  // four classes: one word of counts
    const uint32_t tl = __shfl_up_sync(0xffffffffu, sl, o);
    if (lane >= (uint32_t)o) {
      sa += ta;
      sb += tb;
      sl = sl > tl ? sl : tl;
    }
  }
  uint32_t el = __shfl_up_sync(0xffffffffu, sl, 1);  // exclusive: the threads before this one
  if (lane == 0) el = 0;
  uint64_t *scr64 = reinterpret_cast<uint64_t *>(scratch);
  // (the scratch is free: a barrier of the group lies between its previous use and this call)
  if (lane == 31) {
    scr64[2 * w] = sa;
    scr64[2 * w + 1] = sb;
    scratch[32 + w] = sl;
  }
  group_sync(g);
  uint64_t ta = 0, tb = 0, pa = 0, pb = 0;
  uint32_t tl = 0, pl = 0;
#pragma unroll
  for (uint32_t k = 0; k < (uint32_t)kInsClasses; k++) {
    const uint64_t xa = scr64[2 * k], xb = scr64[2 * k + 1];
    const uint32_t xl = scratch[32 + k];
    if (k < w) {
      pa += xa;
      pb += xb;
      pl = pl > xl ? pl : xl;
    }
    ta += xa;
    tb += xb;
    tl = tl > xl ? tl : xl;
  }
  if (pl > el) el = pl;
  // class stretches: exclusive prefix over the eight class totals (a multiplication sums the 16-bit fields below)
  const uint64_t kOnes = 0x0001000100010001ull;
  const uint64_t ia = ta * kOnes, ib = tb * kOnes;      // inclusive inside each half
  const uint64_t ea = ia << 16, eb = (ib << 16) + (ia >> 48) * kOnes;
  uint64_t oa = ea + pa + (sa - ca), ob = eb + pb + (sb - cb);  // where this thread's positions of every class go
#pragma unroll
  for (int j = 0; j < kRun; j++) {
    if (hv[j] != kNoInsert) {
      const uint32_t cls = hv[j] & (uint32_t)(kInsClasses - 1);
      const uint32_t sh = 16 * (cls & 3u);
      uint32_t at;
      if (cls & 4u) {
        at = (uint32_t)(ob >> sh) & 0xffffu;
        ob += 1ull << sh;
      } else {
        at = (uint32_t)(oa >> sh) & 0xffffu;
        oa += 1ull << sh;
      }
      l_pos[at] = (uint16_t)(i0 + j);
      l_hv[at] = (uint16_t)hv[j];
    }
  }
  // Run interiors get the START of their byte run as their chain start -- the nearest position before them that is
  // not an interior; that position sits in the same bucket (its four bytes are the run byte too), so its own link
  // leads on to the earlier runs.
  InsPrep r;
  r.lead = 0;
  uint32_t cur = el;
#pragma unroll
  for (int j = 0; j < kRun; j++) {
    const int i = i0 + j;
    if (i < n) {
      if (hv[j] != kNoInsert) {
        cur = (uint32_t)i + 1u;
      } else if (cur == 0) {
        r.lead++;
      } else {
        const int dist = i - ((int)cur - 1);
        const uint32_t sl = slot0 + (uint32_t)j;
        prev16[sl < (uint32_t)kRing ? sl : sl - (uint32_t)kRing] =
            dist <= kMaxDist ? (unsigned short)ring_slot(rp0 + (int)cur - 1) : (unsigned short)kNil;
      }
    }
  }
  const uint32_t sh = 16 * (w & 3u);
  r.list_start = (uint32_t)(((w & 4u) ? eb : ea) >> sh) & 0xffffu;
  r.list_count = (uint32_t)(((w & 4u) ? tb : ta) >> sh) & 0xffffu;
  r.last = tl;
  return r;
}

// Part two, ORDERED (range after range along the stream): every warp links the positions of its bucket class, 32
// per step, in stream order; the classes are disjoint sets of buckets, so the warps do not meet. In a step every
// lane reads the head of its bucket, then writes itself there and reads the head back: a lane that finds itself is
// alone in its bucket (the common case: 32 positions, 1024 buckets per class) and links to the old head. Lanes that
// share a bucket find each other with a second write (now the first winner sees a foreign word too) and a MATCH.ANY
// in which all the other lanes pass one common value -- the instruction costs ~12 cycles per DISTINCT value and
// runs one warp at a time per SM (tools/micro_match.cu: 377 cycles for 32 distinct buckets, 512 with eight warps
// at it, 18 for four values): a lane of a shared bucket links to the nearest lower lane of its group (the positions between
// them are in other buckets), the first of a group to the old head, and the last lane of a group becomes the head.
// `run_start` (shared memory, one word per run of chunks): position + kDict + 1 of the last position before rp0 that
// is not a run interior, 0 = none; updated to the same for the end of the range. Ends with a team barrier.
DSG_D void insert_link(unsigned short *prev16, uint32_t *s_head, const uint16_t *l_pos, const uint16_t *l_hv,
                       const InsPrep &prep, int rp0, const LzGroup &g, uint32_t *run_start) {
  const uint32_t lane = (uint32_t)g.tid & 31u;
  uint32_t pos_n = 0, hv_n = 0;  // the next step's entry, fetched a step ahead
  if (lane < prep.list_count) {
    pos_n = l_pos[prep.list_start + lane];
    hv_n = l_hv[prep.list_start + lane];
  }
  for (uint32_t k0 = 0; k0 < prep.list_count; k0 += 32) {
    const bool active = k0 + lane < prep.list_count;
    const uint32_t pos = pos_n, hv = hv_n;
    if (k0 + 32 + lane < prep.list_count) {
      pos_n = l_pos[prep.list_start + k0 + 32 + lane];
      hv_n = l_hv[prep.list_start + k0 + 32 + lane];
    }
    const int rp = rp0 + (int)pos;
    const uint32_t mine = head_word(rp);
    const uint32_t old = active ? s_head[hv] : 0u;
    __syncwarp();  // the old heads are read before any lane of this step replaces one
    if (active) s_head[hv] = mine;
    __syncwarp();
    const bool lost = active && s_head[hv] != mine;
    uint32_t link = old;      // what this position links to: a head word ...
    bool link_is_word = true;  // ... or the range index of a position of this step
    if (__any_sync(0xffffffffu, lost)) {
      __syncwarp();
      if (lost) s_head[hv] = mine;  // the first winner of a shared bucket now reads a foreign word as well
      __syncwarp();
      const bool shared = active && (lost || s_head[hv] != mine);
      __syncwarp();
      // all lanes take part; the lanes that are alone in their bucket pass ONE common value (cost per distinct value)
      const uint32_t grp = __match_any_sync(0xffffffffu, shared ? hv : 0xffffffffu);
      const uint32_t lower = shared ? grp & ((1u << lane) - 1u) : 0u;
      const uint32_t below = __shfl_sync(0xffffffffu, pos, lower ? 31 - __clz((int)lower) : (int)lane);
      if (lower) {
        link = below;
        link_is_word = false;
      }
      if (shared && (grp >> lane) == 1u) s_head[hv] = mine;  // the last position of the bucket in this step
    }
    if (active)
      prev16[ring_slot(rp)] = (unsigned short)(link_is_word ? head_entry(link, rp) : ring_slot(rp0 + (int)link));
    __syncwarp();  // the heads are written before the next step reads them
  }
  if (prep.lead) {
    const uint32_t carry = *run_start;
    const int start_rp = (int)carry - kDict - 1;
    const unsigned short start_slot = carry ? (unsigned short)ring_slot(start_rp) : (unsigned short)kNil;
    for (uint32_t j = 0; j < prep.lead; j++) {
      const int r = rp0 + g.tid * kRun + (int)j;
      prev16[ring_slot(r)] = (carry && r - start_rp <= kMaxDist) ? start_slot : (unsigned short)kNil;
    }
  }
  group_sync(g);  // everybody has read the carry; the range is in the chains
  if (g.tid == 0 && prep.last) *run_start = (uint32_t)(rp0 + (int)prep.last - 1 + kDict + 1);
}

// Longest match for every position of the tile. tile_rp = run-relative position of the tile start, chunk_end =
// run-relative end of the chunk (matches never cross it), lo_rp = lowest position holding real data.
// A thread owns kRun consecutive positions, split into kWays independent sub-runs that it advances in an
// interleaved fashion (two dependent shared-memory chains in flight per thread: the walk is latency bound, ncu
// shows 16 warps cannot hide it, profiles/r01_lz_ncu.md). Inside a sub-run a position inherits (len-1, dist)
// from its predecessor. Each sub-run is a small state machine -- an iteration performs ONE chain step, ONE
// 4-byte compare step, or finishes a position and sets up the next -- so the lanes of a warp never wait for
// the lane with the longest chain or the longest compare.
constexpr int kWays = 1;
constexpr int kSearchRun = 8;  // positions per dynamically assigned run
constexpr int kSubRun = kRun / kWays;

struct SearchWay {
  int j, st, pos0;  // position index inside the sub-run, state, tile-relative first position
  int inh_len, inh_dist;
  int maxlen, best, bdist, depth, max_back, last_d, cmps;
  int cl, cd;  // compare state: bytes matched so far, distance of the candidate
  uint32_t slot, q, qn, aw, v_prev;
  const uint8_t *a, *cb;
};

DSG_D void search_way_init(SearchWay &w, const uint8_t *s_data, int tile_rp, int pos0, int lo_rp) {
  w.j = 0;
  w.st = 0;
  w.pos0 = pos0;
  w.inh_len = w.inh_dist = 0;
  w.maxlen = w.best = w.bdist = w.depth = w.max_back = w.last_d = w.cmps = w.cl = w.cd = 0;
  w.slot = ring_slot(tile_rp + pos0);
  w.q = w.qn = kNil;
  w.aw = 0;
  w.a = s_data + w.slot;
  w.cb = s_data;
  if (tile_rp + pos0 > lo_rp) w.v_prev = load4(s_data + ring_slot(tile_rp + pos0 - 1));
  else w.v_prev = ~load4(w.a);
}

// Sets up position w.j of the run (inheritance, window, chain start, byte-run candidate).
DSG_D void search_way_setup(SearchWay &w, const uint8_t *s_data, const uint16_t *s_prev, int tile_rp, int chunk_end,
                            int lo_rp, int depth_cfg) {
  enum { kWalk = 1, kCmp = 2 };
  const int rp = tile_rp + w.pos0 + w.j;
  w.maxlen = chunk_end - rp < kMaxMatch ? chunk_end - rp : kMaxMatch;
  w.best = 0;
  w.bdist = 0;
  w.depth = 0;
  w.q = kNil;
  w.last_d = 0;
  w.cmps = 0;
  w.st = kWalk;
  const uint32_t v = load4(w.a);
  if (w.maxlen >= kMinMatch) {
    if (w.inh_len > kMinMatch) {
      w.best = w.inh_len - 1 < w.maxlen ? w.inh_len - 1 : w.maxlen;
      w.bdist = w.inh_dist;
    }
    w.max_back = rp - lo_rp < kMaxDist ? rp - lo_rp : kMaxDist;
    w.depth = w.best >= kNice ? 0 : (w.best >= kGood ? depth_cfg / 4 : depth_cfg);
    w.q = s_prev[w.slot];
    w.aw = w.a[w.best];  // the byte a longer match has to reproduce (cheap first filter)
    const bool in_run = v == w.v_prev && w.max_back > 0;  // interior of a byte run
    if (w.inh_len >= kMaxMatch && w.best >= kMinMatch && w.best < w.maxlen) {
      // The predecessor's match was cut by the length limit, so the inherited one (a byte shorter) very likely goes
      // on: without this the lengths inside a long run or a long periodic stretch decay 258, 257, 256 ... and the
      // parse pays extra bits for the odd lengths (a megabyte of zeros: 3.7 KB instead of 1.3 KB). Extend it from
      // where it ends, at the same distance: one or two compare steps.
      int cslot = (int)w.slot - w.bdist;
      if (cslot < 0) cslot += kRing;
      w.cb = s_data + cslot;
      w.cd = w.bdist;
      w.cl = w.best & ~3;
      w.qn = w.q;
      w.st = kCmp;
    } else if (in_run && w.best < kNice && !(w.bdist == 1 && w.best >= kMinMatch)) {
      // try the distance-1 candidate (never in the chains, see insert_one)
      w.cb = w.a - 1 + (w.slot == 0 ? kRing : 0);
      w.cd = 1;
      w.cl = 4;  // v == v_prev: the first four bytes are known to match
      w.qn = w.q;
      w.st = kCmp;
    }
  }
  w.v_prev = v;
}

// The three actions of the per-lane state machine.
DSG_D void search_cmp_step(SearchWay &w) {  // one 4-byte step of the candidate at distance cd
  enum { kWalk = 1 };
  const uint32_t x = w.cl < w.maxlen ? (load4(w.a + w.cl) ^ load4(w.cb + w.cl)) : 1u;
  if (x == 0) {
    w.cl += 4;
    return;
  }
  int l = w.cl < w.maxlen ? w.cl + ((__ffs((int)x) - 1) >> 3) : w.maxlen;
  if (l > w.maxlen) l = w.maxlen;
  if (l >= kMinMatch && (l > w.best || (l == w.best && w.cd < w.bdist))) {
    w.best = l;
    w.bdist = w.cd;
    if (l >= kNice || l >= w.maxlen) w.depth = 0;
    else w.aw = w.a[l];
  }
  w.q = w.qn;
  w.st = kWalk;
}

DSG_D void search_walk_step(SearchWay &w, const uint8_t *s_data, const uint16_t *s_prev) {  // one chain step
  enum { kCmp = 2 };
  w.depth--;
  int d = (int)w.slot - (int)w.q;
  if (d <= 0) d += kRing;
  if (d > w.max_back || d <= w.last_d) {
    w.q = kNil;  // left the window, or a stale link (chains are in exact stream order: distances grow)
    return;
  }
  w.last_d = d;
  const uint8_t *b = s_data + w.q;
  const uint32_t qn = s_prev[w.q];  // next link, fetched before the compare so that its latency overlaps
  if (b[w.best] == (uint8_t)w.aw) {
    w.qn = qn;
    w.cb = b;
    w.cd = d;
    w.cl = 0;
    w.st = kCmp;
    // Periodic data (records, tables) makes every candidate pass the one-byte filter and cost a full compare
    // without ever getting longer: after kMaxCompares of them the position is done.
    if (++w.cmps >= kMaxCompares) w.depth = 0;
  } else {
    w.q = qn;
  }
}

// Chain exhausted: record the position and set up the next one of the run (the caller hands out a new run).
DSG_D void search_finish_position(SearchWay &w, const uint8_t *s_data, const uint16_t *s_prev, uint8_t *s_mlen,
                                  uint16_t *s_mdist, int tile_rp, int chunk_end, int lo_rp, int depth_cfg) {
  if (w.best < kMinMatch) w.best = 0;
  s_mlen[w.pos0 + w.j] = (uint8_t)(w.best ? w.best - 3 : 0);  // 0 = no match, else len-3 in 1..255
  // bit 15: the walk stopped at the depth limit with the chain unfinished -> worth a deep second look
  const bool limited = w.q != kNil && w.best < kNice && w.best < w.maxlen && w.maxlen >= kMinMatch;
  s_mdist[w.pos0 + w.j] = (uint16_t)(w.bdist | (limited ? 0x8000 : 0));
  w.inh_len = w.best;
  w.inh_dist = w.bdist;
  w.j++;
  w.slot++;
  w.a++;
  if (w.slot == (uint32_t)kRing) {
    w.slot = 0;
    w.a = s_data;
  }
  if (w.j < kSearchRun) search_way_setup(w, s_data, s_prev, tile_rp, chunk_end, lo_rp, depth_cfg);
}

// Positions are handed out in runs of kSearchRun through a shared-memory counter: a lane that finishes its run
// takes the next free one while the other lanes of its warp keep walking, so neither lanes nor warps wait for the
// slowest run at the barrier behind the search. Every iteration performs ONE action per lane (chain step,
// 4-byte compare step, or finishing a position and setting up the next).
// Measured alternatives that were slower on B200 (profiles/r01_lz_phases.md): fixed 16-position runs per
// thread, two interleaved runs per thread, and warp-voted scheduling of the three actions.
// `ctr` must be zero on entry.
DSG_D void search_tile(const uint8_t *s_data, const uint16_t *s_prev, uint8_t *s_mlen, uint16_t *s_mdist, int tile_rp,
                       int tile_len, int chunk_end, int lo_rp, uint32_t *ctr, int depth_cfg) {
  enum { kWalk = 1, kCmp = 2 };
  const uint32_t nruns = (uint32_t)((tile_len + kSearchRun - 1) / kSearchRun);
  SearchWay w;
  uint32_t run = atomicAdd(ctr, 1u);
  if (run >= nruns) return;
  search_way_init(w, s_data, tile_rp, (int)run * kSearchRun, lo_rp);
  search_way_setup(w, s_data, s_prev, tile_rp, chunk_end, lo_rp, depth_cfg);
  for (;;) {
    if (w.st == kCmp) {
      search_cmp_step(w);
    } else if (w.depth > 0 && w.q != kNil) {
      search_walk_step(w, s_data, s_prev);
    } else {
      search_finish_position(w, s_data, s_prev, s_mlen, s_mdist, tile_rp, chunk_end, lo_rp, depth_cfg);
      if (w.j >= kSearchRun) {
        run = atomicAdd(ctr, 1u);
        if (run >= nruns) break;
        search_way_init(w, s_data, tile_rp, (int)run * kSearchRun, lo_rp);
        search_way_setup(w, s_data, s_prev, tile_rp, chunk_end, lo_rp, depth_cfg);
      }
    }
  }
}

// ---- second pass: deep search of the positions the parse actually uses ---------------------------------------
// zlib spends its 128 chain steps only where a token starts (and at the byte behind a match start, for the lazy
// decision); spending them at every position is what made the first version of this kernel slow, and a uniform
// depth of 32 left 5-7 % of compression on the table for text. So: shallow pass everywhere, parse, then a deep
// pass over the token starts + their successors (a compacted, well balanced work list), parse again.

// Work list: the token starts of `bitmap` and, behind every start that carries a match, the next position --
// each at most once per tile (s_deep). Returns the list length (all threads).
DSG_D uint32_t refine_build_list(const uint8_t *s_mlen, const uint16_t *s_mdist, const uint32_t *s_bitmap,
                                 uint32_t *s_deep, uint16_t *s_work, uint32_t *scan_scratch, int tile_len,
                                 const LzGroup &g) {
  const int t0 = g.tid * kRun;
  uint16_t mine[2 * kRun];
  uint32_t n = 0, nfull = 0;
  const uint32_t bits = (s_bitmap[t0 >> 5] >> (t0 & 31)) & ((1u << kRun) - 1u);
  // zlib spends a quarter of the effort at the byte behind an already good match. The class is decided here, from
  // the state before the round (not while other lanes update their neighbours), and travels in bit 15.
  auto quarter = [&](int tp) -> uint32_t {
    return (tp > 0 && s_mlen[tp - 1] != 0 && s_mlen[tp - 1] + 3u >= (uint32_t)kGood) ? 0x8000u : 0u;
  };
  uint32_t m = bits;
  while (m) {
    const int j = __ffs((int)m) - 1;
    m &= m - 1;
    const int tp = t0 + j;
    // only positions whose shallow walk ran into the depth limit (bit 15 of mdist) can gain anything
    if ((s_mdist[tp] & 0x8000) && !(s_deep[tp >> 5] & (1u << (tp & 31)))) {
      const uint32_t q = quarter(tp);
      mine[n++] = (uint16_t)(tp | q);
      nfull += q ? 0 : 1;
    }
    const int nx = tp + 1;
    if (s_mlen[tp] != 0 && nx < tile_len && (s_mdist[nx] & 0x8000) && !(s_bitmap[nx >> 5] & (1u << (nx & 31))) &&
        !(s_deep[nx >> 5] & (1u << (nx & 31)))) {
      const uint32_t q = quarter(nx);
      mine[n++] = (uint16_t)(nx | q);
      nfull += q ? 0 : 1;
    }
  }
  // (every thread reads s_deep above and updates it below the scan, whose barriers lie between the two)
  // The full-depth positions go to the front of the list and the quarter-depth ones behind them: the lanes of a warp
  // take 32 consecutive entries and wait for the longest walk among them, so walks of one length belong together,
  // and the long ones start first. One scan carries both counts.
  uint32_t total;
  const uint32_t at = group_exclusive_scan(nfull | ((n - nfull) << 16), scan_scratch, &total, g);
  uint32_t at_full = at & 0xffffu, at_quarter = (total & 0xffffu) + (at >> 16);
  for (uint32_t k = 0; k < n; k++) {
    const uint32_t tp = mine[k] & 0x7fffu;
    if (mine[k] & 0x8000u) s_work[at_quarter++] = mine[k];
    else s_work[at_full++] = mine[k];
    atomicOr(&s_deep[tp >> 5], 1u << (tp & 31));
  }
  group_sync(g);
  return (total & 0xffffu) + (total >> 16);
}

// Deep search of the listed positions, handed out through a shared counter (`ctr` zero on entry).
// The walk is bound by the latency of one lane's chain of dependent shared-memory loads (the stall samples are
// fixed-latency waits and scoreboard waits, not issue contention), so the links are chased kChase ahead: the loads
// of prev[] depend only on each other, the filter bytes and the window checks of the candidates then overlap.
// Speculative links past the end of a chain are kNil or stale slots: both are valid shared-memory indices.
constexpr int kChase = 4;

DSG_D void refine_search(const uint8_t *s_data, const uint16_t *s_prev, uint8_t *s_mlen, uint16_t *s_mdist,
                         const uint16_t *s_work, uint32_t nwork, int tile_rp, int tile_len, int chunk_end, int lo_rp,
                         uint32_t *ctr, int deep_cfg, uint32_t ctr_base = 0, uint32_t *s_deep = nullptr,
                         uint16_t *s_chain = nullptr, uint32_t *n_chain = nullptr) {
  const SmemBase sd = smem_base(s_data), sp = smem_base(s_prev);
  for (;;) {
    const uint32_t k = atomicAdd(ctr, 1u) - ctr_base;
    if (k >= nwork) break;
    const uint32_t item = s_work[k];
    const int tp = (int)(item & 0x7fffu);
    const int rp = tile_rp + tp;
    const int slot = (int)ring_slot(rp);
    const uint8_t *a = s_data + slot;
    const int maxlen = chunk_end - rp < kMaxMatch ? chunk_end - rp : kMaxMatch;
    const uint32_t cur = s_mlen[tp];
    int best = cur ? (int)cur + 3 : 0;
    const int best0 = best;
    int bdist = cur ? (int)(s_mdist[tp] & 0x7fff) : 0;
    if (maxlen < kMinMatch || best >= kNice || best >= maxlen) continue;
    const int max_back = rp - lo_rp < kMaxDist ? rp - lo_rp : kMaxDist;
    int depth = (item & 0x8000u) ? deep_cfg / 4 : deep_cfg;  // see refine_build_list
    int cmps = kMaxCompares;
    int last_d = 0;
    uint32_t q = s_prev[slot];
    uint32_t want = a[best];
    while (depth > 0 && q != kNil) {
      uint32_t c[kChase + 1];
      c[0] = q;
#pragma unroll
      for (int i = 0; i < kChase; i++) c[i + 1] = c[i] == kNil ? (uint32_t)kNil : lds_u16(sp, 2u * c[i]);
      uint32_t f[kChase];
#pragma unroll
      for (int i = 0; i < kChase; i++) f[i] = lds_u8(sd, (c[i] == kNil ? 0u : c[i]) + (uint32_t)best);
      bool stop = false;
      uint32_t nq = c[kChase];
#pragma unroll
      for (int i = 0; i < kChase; i++) {
        if (!stop) {
          const uint32_t ci = c[i];
          if (ci == kNil || depth <= 0) {
            stop = true;
            nq = kNil;
          } else {
            int d = slot - (int)ci;
            if (d <= 0) d += kRing;
            depth--;
            if (d > max_back || d <= last_d) {
              stop = true;  // left the window, or a stale link (chains are in exact stream order: distances grow)
              nq = kNil;
            } else {
              last_d = d;
              if (f[i] == want) {
                const uint8_t *b = s_data + ci;
                int l = 0;
                while (l < maxlen) {
                  const uint32_t x = load4(a + l) ^ load4(b + l);
                  if (x) {
                    l += (__ffs((int)x) - 1) >> 3;
                    break;
                  }
                  l += 4;
                }
                if (l > maxlen) l = maxlen;
                if (l > best) {
                  best = l;
                  bdist = d;
                  stop = true;  // the filter bytes of the remaining candidates were read for the old length
                  if (l >= kNice || l >= maxlen) {
                    nq = kNil;
                  } else {
                    want = a[best];
                    nq = c[i + 1];
                  }
                }
                if (--cmps == 0) {  // periodic data: every candidate passes the filter and none gets longer
                  stop = true;
                  nq = kNil;
                }
              }
            }
          }
        }
      }
      q = nq;
    }
    if (best >= kMinMatch) {
      s_mlen[tp] = (uint8_t)(best - 3);
      s_mdist[tp] = (uint16_t)bdist;
    }
    // The match grew: the token behind it starts somewhere else now. That position joins a second list (once: the
    // deep-search marks arbitrate), searched behind this one without another parse.
    if (s_chain && best != best0) {
      const int q = tp + best;
      if (q < tile_len && (s_mdist[q] & 0x8000)) {
        const uint32_t bit = 1u << (q & 31);
        if (!(atomicOr(&s_deep[q >> 5], bit) & bit)) s_chain[atomicAdd(n_chain, 1u)] = (uint16_t)q;
      }
    }
  }
}

// next(p) for the stateless lazy parse; tp is tile-relative, tile_len = positions of this tile inside the chunk
DSG_D int parse_step(const uint8_t *s_mlen, int tp, int tile_len) {
  const uint32_t l = s_mlen[tp];
  if (l == 0) return 1;
  if (tp + 1 < tile_len && s_mlen[tp + 1] > l) return 1;  // a longer match starts at the next byte: defer
  return (int)l + 3;
}

// ---------------------------------------------------------------------------------------------------------
// Parse of one tile: marks the token start positions in s_bitmap, computed by kParseWarps warps. `start` = first
// token start (tile-relative, may lie beyond the tile). Every lane owns ONE bitmap word
// (32 positions) and keeps it in a register; a warp owns a segment of 1024 positions.
//   1. every lane walks speculatively from the start of its word (the very first lane: from `start`);
//   2. inside a warp the exits travel by shuffle: a lane whose true entry (its left neighbour's exit) differs from
//      the entry its marks were made for re-walks until it meets a marked position (next() is stateless, so from
//      there on the old marks and the old exit stand); repeated until no exit of the warp changes;
//   3. the segment exits travel through shared memory the same way, until no segment exit changes.
// Called by all threads of a team (its first kParseThreads threads do the work; the synchronisation is the team's
// barrier). Returns the first token start of the following tile, relative to that tile. Ends with a barrier.
// ---------------------------------------------------------------------------------------------------------
constexpr int kParseWarps = kTile / 1024;
constexpr int kParseThreads = kParseWarps * 32;

// Re-walk of one word from entry e (>= sub_lo). Returns the new exit, or -1 when the walk met the marked path.
DSG_D int parse_lane_walk(const uint8_t *s_mlen, uint32_t &cur, int e, int sub_lo, int sub_hi, int tile_len) {
  if (e >= sub_hi) {  // the whole word lies inside an earlier match (or beyond the tile)
    cur = 0;
    return e;
  }
  cur &= ~((1u << (e - sub_lo)) - 1u);
  int q = e;
  while (q < sub_hi) {
    const int b = q - sub_lo;
    if ((cur >> b) & 1u) return -1;
    if ((q & 3) == 0 && q + 4 <= sub_hi && ((cur >> b) & 0xfu) == 0u &&
        *reinterpret_cast<const uint32_t *>(s_mlen + q) == 0u) {  // four literals in a row
      cur |= 0xfu << b;
      q += 4;
      continue;
    }
    cur |= 1u << b;
    const int nq = q + parse_step(s_mlen, q, tile_len);
    const int hi = (nq < sub_hi ? nq : sub_hi) - sub_lo;  // clear the marks the token covers: bits (b, hi)
    const uint32_t upto_hi = hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u);
    const uint32_t upto_b = (2u << b) - 1u;               // bits [0, b]
    cur &= ~(upto_hi & ~upto_b);
    q = nq;
  }
  return q;
}

// In-warp fixed point; seg_entry = the entry of the warp's first lane. Returns this lane's exit.
DSG_D int parse_warp_settle(const uint8_t *s_mlen, uint32_t &cur, int &used, int ex, int seg_entry, int sub_lo,
                            int sub_hi, int tile_len) {
  const int lane = (int)(threadIdx.x & 31);
  for (int pass = 0; pass < 34; pass++) {
    const int left = __shfl_up_sync(0xffffffffu, ex, 1);
    const int e = lane == 0 ? seg_entry : left;
    bool changed = false;
    if (e != used) {
      const int r = parse_lane_walk(s_mlen, cur, e, sub_lo, sub_hi, tile_len);
      if (r >= 0) {
        changed = r != ex;
        ex = r;
      }
      used = e;
    }
    if (!__any_sync(0xffffffffu, changed)) break;
  }
  return ex;
}

DSG_D int parse_tile_wide(const uint8_t *s_mlen, uint32_t *s_bitmap, int start, int tile_len, const LzGroup &team) {
  const int g = team.tid;
  const bool active = g < kParseThreads;  // the other threads of the team only keep the barriers company
  const int lane = g & 31, w = g >> 5;
  const int sub_lo = g * 32;
  const int sub_hi = sub_lo + 32 < tile_len ? sub_lo + 32 : tile_len;
  uint32_t *s_x = s_bitmap + kTile / 32;  // [0, kParseWarps): segment exits; [kParseWarps, +2): change flags
  uint32_t cur = 0;
  int seg_entry = w == 0 ? start : w * 1024;
  int used = lane == 0 ? seg_entry : sub_lo;
  int ex = 0;
  if (active) {
    ex = parse_lane_walk(s_mlen, cur, used, sub_lo, sub_hi, tile_len);
    ex = parse_warp_settle(s_mlen, cur, used, ex, seg_entry, sub_lo, sub_hi, tile_len);
    if (lane == 31) s_x[w] = (uint32_t)ex;
  }
  if (kParseWarps == 1) {  // one warp holds the whole tile: no segment exits to exchange, one barrier
    if (active) s_bitmap[g] = cur;
    group_sync(team);
    return (int)s_x[0] - tile_len;
  }
  if (g == 0) {
    s_x[kParseWarps] = 0;
    s_x[kParseWarps + 1] = 0;
  }
  for (int round = 0; round < kParseWarps + 1; round++) {
    uint32_t *flag = s_x + kParseWarps + (round & 1);
    group_sync(team);  // exits and the cleared flag are visible
    if (active) {
      const int e = w == 0 ? start : (int)s_x[w - 1];
      bool changed = false;
      if (e != seg_entry) {
        seg_entry = e;
        const int old = ex;
        ex = parse_warp_settle(s_mlen, cur, used, ex, seg_entry, sub_lo, sub_hi, tile_len);
        changed = lane == 31 && ex != old;
      }
      if (changed) *flag = 1u;
    }
    group_sync(team);  // every exit has been read, every flag written
    if (active && lane == 31) s_x[w] = (uint32_t)ex;
    if (g == 0) s_x[kParseWarps + ((round + 1) & 1)] = 0;
    if (*flag == 0u) break;
  }
  if (active) s_bitmap[g] = cur;
  group_sync(team);
  return (int)s_x[kParseWarps - 1] - tile_len;  // >= 0: first token start of the following tile, relative to it
}

// ---------------------------------------------------------------------------------------------------------
// CTA-cooperative Huffman construction (literal/length and distance trees side by side)
// ---------------------------------------------------------------------------------------------------------
// Thread roles: the literal/length tree uses threads [0, 286), the distance tree threads [288, 318); the two
// serial merges run on thread 0 and thread 288 (different warps). Every step is followed by a barrier in the
// caller-visible function below, so all threads of the CTA must call it.
struct HuffJob {
  const uint32_t *freq;  // n frequencies
  uint8_t *lens;         // n code lengths out
  uint16_t *codes;       // n bit-reversed canonical codes out
  HuffScratch *S;
  uint32_t *ctrl;        // [0] m (used symbols), [1] max depth, [2] max used symbol, [3..19] bl_count, [20..36] next_code
  int n;
  int t0;                // first thread of this job
};

DSG_D void huff_jobs_build(const HuffJob &A, const HuffJob &B) {
  const int tid = (int)threadIdx.x;
  const bool inA = tid >= A.t0 && tid < A.t0 + A.n;
  const bool inB = tid >= B.t0 && tid < B.t0 + B.n;
  const HuffJob &J = inB ? B : A;
  const bool active = inA || inB;
  const int i = tid - J.t0;
  // step 0: clear control words and lengths
  if (active) {
    J.lens[i] = 0;
    for (int k = i; k < 40; k += J.n) J.ctrl[k] = 0;
  }
  __syncthreads();
  // step 1: rank sort by (frequency, symbol) among the used symbols
  uint32_t f = 0;
  if (active) {
    f = J.freq[i];
    if (f) {
      int rank = 0;
      for (int j = 0; j < J.n; j++) {
        const uint32_t g = J.freq[j];
        rank += (g != 0 && (g < f || (g == f && j < i))) ? 1 : 0;
      }
      J.S->order[rank] = (uint16_t)i;
      J.S->weight[rank] = f;
      atomicAdd(&J.ctrl[0], 1u);
      atomicMax(&J.ctrl[2], (uint32_t)i);
    }
  }
  __syncthreads();
  // step 2: serial two-queue merge (one thread per tree)
  if (active && i == 0) {
    const int m = (int)J.ctrl[0];
    if (m >= 2) J.ctrl[37] = (uint32_t)huff_merge_sorted(*J.S, m);
  }
  __syncthreads();
  // step 3: leaf depths by walking up the parent links
  if (active) {
    const int m = (int)J.ctrl[0];
    if (m >= 2 && i < m) {
      const int root = (int)J.ctrl[37];
      int node = i, d = 0;
      while (node != root) {
        node = J.S->parent[node];
        d++;
      }
      atomicMax(&J.ctrl[1], (uint32_t)d);
      J.lens[J.S->order[i]] = (uint8_t)d;  // final when no code exceeds the limit (checked next)
    }
  }
  __syncthreads();
  // step 4: degenerate trees and the (rare) length-limit repair, serial
  if (active && i == 0) {
    const int m = (int)J.ctrl[0];
    if (m == 0) {
      J.lens[0] = 1;
      J.lens[1] = 1;
      J.ctrl[2] = 1;
    } else if (m == 1) {
      const int sy = J.S->order[0];
      const int other = sy == 0 ? 1 : 0;
      J.lens[sy] = 1;
      J.lens[other] = 1;
      J.ctrl[2] = (uint32_t)(sy > other ? sy : other);
    } else if (J.ctrl[1] > (uint32_t)kMaxBits) {
      huff_lengths_from_tree(*J.S, m, (int)J.ctrl[37], kMaxBits, J.lens);
    }
  }
  __syncthreads();
  // step 5: canonical codes: per-length counts, first code of each length, then rank inside the length
  if (active && J.lens[i]) atomicAdd(&J.ctrl[3 + J.lens[i]], 1u);
  __syncthreads();
  if (active && i == 0) {
    uint32_t code = 0;
    J.ctrl[3 + 0] = 0;
    for (int b = 1; b <= kMaxBits; b++) {
      code = (code + J.ctrl[3 + b - 1]) << 1;
      J.ctrl[20 + b] = code;
    }
  }
  __syncthreads();
  if (active) {
    const int l = J.lens[i];
    uint32_t c = 0;
    if (l) {
      uint32_t before = 0;
      for (int j = 0; j < i; j++) before += (J.lens[j] == l) ? 1u : 0u;
      c = bit_reverse(J.ctrl[20 + l] + before, l);
    }
    J.codes[i] = (uint16_t)c;
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------------------

// Search effort: compile-time constants (1 % faster than reading them from the parameter block); a library built
// with -DDSG_LZ_RUNTIME_EFFORT takes them from LzParams instead (tools/sweep_effort.py: speed against size).
#if defined(DSG_LZ_RUNTIME_EFFORT)
#define LZ_EFFORT(field, constant) ((int)P.field)
#else
#define LZ_EFFORT(field, constant) (constant)
#endif
__global__ void __launch_bounds__(kLzThreads, 1) lz_deflate_kernel(LzParams P) {
  DSG_DYN_SMEM(smem);
  uint8_t *s_data = smem + LzLayout::data;
  uint16_t *s_prev = reinterpret_cast<uint16_t *>(smem + LzLayout::prev);
  uint32_t *s_head = reinterpret_cast<uint32_t *>(smem + LzLayout::head);
  uint8_t *s_mlen = smem + LzLayout::mlen;
  uint16_t *s_mdist = reinterpret_cast<uint16_t *>(smem + LzLayout::mdist);
  uint32_t *s_bitmap = reinterpret_cast<uint32_t *>(smem + LzLayout::bitmap);
  uint32_t *s_lfreq = reinterpret_cast<uint32_t *>(smem + LzLayout::lfreq);
  uint32_t *s_dfreq = reinterpret_cast<uint32_t *>(smem + LzLayout::dfreq);
  uint8_t *s_llen = smem + LzLayout::llen;
  uint8_t *s_dlen = smem + LzLayout::dlen;
  uint16_t *s_lcode = reinterpret_cast<uint16_t *>(smem + LzLayout::lcode);
  uint16_t *s_dcode = reinterpret_cast<uint16_t *>(smem + LzLayout::dcode);
  uint8_t *s_cllen = smem + LzLayout::cllen;
  uint16_t *s_clcode = reinterpret_cast<uint16_t *>(smem + LzLayout::clcode);
  uint16_t *s_rle = reinterpret_cast<uint16_t *>(smem + LzLayout::rle);
  uint32_t *s_misc = reinterpret_cast<uint32_t *>(smem + LzLayout::misc);
  uint16_t *s_work = reinterpret_cast<uint16_t *>(smem + LzLayout::work);
  uint32_t *s_deep = reinterpret_cast<uint32_t *>(smem + LzLayout::deep);
  uint32_t *s_stage = reinterpret_cast<uint32_t *>(smem + LzLayout::mlen);  // bit-packing window (after the tiles)
  uint32_t *tokens = P.tok_scratch + (size_t)blockIdx.x * kTokSlotWords;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // the team's private arrays
  const uint32_t team = warp / kTeamWarps;
  const LzGroup g_team{(int)(tid - team * kTeamThreads), kTeamThreads, kBarTeam + (int)team};
  uint8_t *t_mlen = s_mlen + team * kTile;
  uint16_t *t_mdist = s_mdist + team * kTile;
  uint32_t *t_bitmap = s_bitmap + team * (kTile / 32 + 16);
  uint16_t *t_work = s_work + team * kTile;
  uint32_t *t_deep = s_deep + team * (kTile / 32);
  uint16_t *t_chain = reinterpret_cast<uint16_t *>(smem + LzLayout::chain) + team * kTile;
  uint32_t *t_nchain = s_misc + kMTeamScan + team * kMTeamScanStride + 19;  // (the team scans use the first 9 words)
  uint32_t *t_scan = s_misc + kMTeamScan + team * kMTeamScanStride;
  uint32_t *t_ctr = s_misc + kMRunCtr + team;
  unsigned short *prev16 = reinterpret_cast<unsigned short *>(s_prev);

  // one mbarrier per team for the bulk staging of its tiles (8-byte words behind the scan scratch of s_misc)
  uint32_t bulk_phase = 0;  // parity of the team's next mbarrier phase (uniform inside the team)
#if !defined(DSG_EMU)
  const uint32_t t_mbar = (uint32_t)__cvta_generic_to_shared(s_misc + kMBulkBar + 2 * team);
  if (tid == 0) {
    for (int i = 0; i < kTeams; i++) mbar_init((uint32_t)__cvta_generic_to_shared(s_misc + kMBulkBar + 2 * i), 1);
    for (int i = 0; i < kSigCount; i++) sig_init(s_misc + kMSig, i);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
#else
  const uint32_t t_mbar = 0;
#endif
  uint32_t *s_sig = s_misc + kMSig;
#if defined(DSG_EMU)
  if (tid < (uint32_t)kSigCount) sig_init(s_sig, (int)tid);
  __syncthreads();
#endif
  uint32_t seq_base = 0;  // number of the chunk's first tile in the CTA's running tile count (uniform)

  long long t_last = 0;
#define LZ_PROF(k)                                                    \
  do {                                                                \
    if (P.prof && tid == 0) {                                         \
      const long long now_ = clock64();                               \
      atomicAdd(&P.prof[k], (unsigned long long)(now_ - t_last));     \
      t_last = now_;                                                  \
    }                                                                 \
  } while (0)
  // Tickets: first the long runs, then the last stretch of the input chunk by chunk, so that the CTAs finish within
  // one chunk of each other (with runs of 6-8 chunks the last run of a 256 MiB pipeline segment cost 10 % idle time).
  const uint32_t R = P.run_chunks;
  const uint32_t nbig = P.big_runs;
  const uint32_t nruns = nbig + (P.nchunks - nbig * R);
  for (;;) {
    __syncthreads();  // previous run fully done with shared memory
    if (P.prof && tid == 0) t_last = clock64();
    if (tid == 0) s_misc[kMTicket] = atomicAdd(P.ticket, 1u);
    __syncthreads();
    const uint32_t run = s_misc[kMTicket];
    if (run >= nruns) break;
    // A run is a stretch of R consecutive chunks compressed by this CTA with ONE sliding LZ window: the hash
    // chains and the ring carry over from chunk to chunk, only the first chunk has to hash its dictionary.
    const uint32_t c0 = run < nbig ? run * R : nbig * R + (run - nbig);
    const uint32_t rlen = run < nbig ? R : 1u;
    const int chunk = (int)lz_chunk_size(c0, P.chunk_bytes, nbig * R);  // a run has chunks of one size
    const int64_t rs = (int64_t)lz_chunk_offset(c0, P.chunk_bytes, nbig * R);  // stream offset of the run
    const int64_t avail = rs + (P.dict ? (int64_t)P.dict_len : 0);
    int lo_rp = -(int)(avail < (int64_t)kMaxDist ? avail : (int64_t)kMaxDist);
    const LzGroup g_all{(int)tid, kLzThreads, 0};
    if (tid == 0) s_misc[kMStale] = 1;  // a run starts by building its window

    for (uint32_t k = 0; k < rlen; k++) {
    const uint32_t c = c0 + k;
    if (c >= P.nchunks) break;
    const int64_t cs = rs + (int64_t)k * chunk;
    const int crp = (int)k * chunk;  // run-relative position of the chunk start
    const int chunk_len = (int)((int64_t)P.len - cs < chunk ? (int64_t)P.len - cs : chunk);
    __syncthreads();
    if (s_misc[kMStale]) {
      // The window: at the start of a run, and again behind a chunk that took the stored exit (its tiles were not
      // loaded). Dictionary + first tile + look-ahead are resident before the dictionary is hashed (the last three
      // dictionary positions hash over the first bytes of the chunk). kMaxDist and crp are multiples of 16.
      if (crp > 0) lo_rp = crp - kMaxDist;
      for (uint32_t i = tid; i < (uint32_t)kHashSize; i += kLzThreads) s_head[i] = kNil;
      load_ring(s_data, P, rs, crp - (((crp - lo_rp) + 15) & ~15), crp + kTile + kLook, g_all);
      __syncthreads();
      if (tid == 0) {
        s_misc[kMRunStart] = 0;
        s_misc[kMStale] = 0;
      }
      __syncthreads();
      // the window goes into the chains like tiles do: every team prepares a stretch of kTile positions, then the
      // teams link their stretches one after the other
      for (int base = lo_rp; base < crp; base += kTeams * kTile) {
        const int b0 = base + (int)team * kTile;
        const int nb = crp - b0 < 0 ? 0 : (crp - b0 < kTile ? crp - b0 : kTile);
        const InsPrep prep = insert_prepare(s_data, prev16, t_work, t_mdist, t_bitmap, b0, nb, lo_rp, g_team);
        for (uint32_t j = 0; j < (uint32_t)kTeams; j++) {
          __syncthreads();
          if (team == j && nb > 0) insert_link(prev16, s_head, t_work, t_mdist, prep, b0, g_team, &s_misc[kMRunStart]);
        }
      }
      __syncthreads();
      LZ_PROF(kProfLoad);
    }
    for (uint32_t i = tid; i < 288; i += kLzThreads) s_lfreq[i] = 0;
    if (tid < 32) s_dfreq[tid] = 0;
    if (tid == 0) {
      s_misc[kMStart] = 0;
      s_misc[kMNTok] = 0;
      s_misc[kMProbe] = 0;
      s_misc[kMTileCtr] = (uint32_t)kTeams;  // every team's first tile is its own number, the later ones are taken
    }
    chunk_flatness(P.in, P.len, cs, chunk_len, reinterpret_cast<uint32_t *>(s_work), &s_misc[kMFlat]);  // barriers inside
    __syncthreads();
    LZ_PROF(kProfFlat);

    // ---- tiles: a pipeline over kTeams teams -------------------------------------------------------------------
    // A team (kTeamWarps warps) runs, for the tile it holds, stage A (load, hash, shallow search), stage B (parse / deep search
    // rounds) and stage C (final parse, token emission). Hashing must follow the hashing of tile t-1 and stage C the
    // emission of tile t-1 (token order, the carried parse position), each done by whichever team holds that tile: one
    // signal per tile and direction. So while one team is in stage B of tile t the others are in other stages of the
    // tiles around it, and the long tails of the search phases (the last few chains of a tile: 40 % of the phase with
    // one tile in flight) are filled with their work. The ring holds the 32 KiB window plus one tile per team:
    // loading tile t overwrites only bytes that lie beyond the window of every tile that is still in flight.
    const int ntiles = (chunk_len + kTile - 1) / kTile;
    // every team has a first tile, so the CTA-wide barrier behind it is reached by all threads
    const bool flat_candidate = s_misc[kMFlat] != 0u && ntiles >= kTeams;
    bool stored_exit = false;
    // Tiles are TAKEN, not dealt: a team that has emitted its tile takes the next free one (shared counter), so a team
    // with a cheap tile (random or zero page) does not sit out the turn of a team with an expensive one (token-like
    // text: three times the deep-search work). Hashing and emission stay in tile order through the signals; at most
    // kTeams consecutive tiles are in flight (a team cannot emit, hence not take a tile, before its predecessor has).
    int t = (int)team;
    while (t < ntiles) {
      const int tile_rp = crp + t * kTile;
      const int tile_len = chunk_len - t * kTile < kTile ? chunk_len - t * kTile : kTile;
      const bool has_next = t + 1 < ntiles;
      // -- stage A
      // The tile's bytes + look-ahead replace the oldest ring bytes: those lie beyond the window of every tile that is
      // still being searched (the nearest one, tile t-kTeams+1, reaches back exactly to the first byte that is kept), so the
      // copy does not have to wait for the hashing of tile t-1 and overlaps that wait. (The first tile of a window came
      // with the dictionary; it is loaded again, the same bytes.)
      const bool staged = tile_rp > 0;
      // 16 bytes more at the front (the byte before the tile decides whether its first position is a run interior)
      // and the first kLook bytes, which the tile before this one has loaded or is loading as its look-ahead -- the
      // same bytes to the same slots: with them the team can prepare the insertion without waiting for that tile.
      const int load_lo = tile_rp - 16, load_hi = tile_rp + kTile + kLook;
      const bool bulk = staged && bulk_ok(P, rs, load_lo, load_hi);
      if (bulk) {
        if (g_team.tid == 0) bulk_load_ring(s_data, P, rs, load_lo, load_hi, t_mbar);
      } else if (staged) {
        load_ring(s_data, P, rs, load_lo, load_hi, g_team);
      }
      const uint32_t seq = seq_base + (uint32_t)t;  // t > 0: seq - 1 is the previous tile of this chunk
      // One warp of the team does the waiting (a waiting warp re-issues its try_wait; eight of them take issue slots
      // from the teams that are working), the others park at the team's barrier.
#if !defined(DSG_EMU)
      if (bulk && g_team.tid < 32) mbar_wait(t_mbar, bulk_phase);
#endif
      if (bulk) bulk_phase ^= 1u;
      if (staged) group_sync(g_team);
      LZ_PROF(kProfLoad);
      const InsPrep prep = insert_prepare(s_data, prev16, t_work, t_mdist, t_bitmap, tile_rp, tile_len, lo_rp, g_team);
      LZ_PROF(kProfInsert);
      // tile t-1 is in the chains
      if (t > 0 && g_team.tid < 32) sig_wait(s_sig, kSigIns + (int)((seq - 1) & (kSigRing - 1)), (seq - 1) / kSigRing);
      group_sync(g_team);  // ... and the team's class lists are complete
      LZ_PROF(kProfWaitA);
      // (the search counter and the deep-search marks of the tile: nobody uses them now, the barrier inside
      // insert_link publishes them)
      if (g_team.tid == 0) *t_ctr = 0;
      for (uint32_t i = (uint32_t)g_team.tid; i < (uint32_t)kTile / 32; i += kTeamThreads) t_deep[i] = 0;
      insert_link(prev16, s_head, t_work, t_mdist, prep, tile_rp, g_team, &s_misc[kMRunStart]);  // ends with a barrier
      if (g_team.tid == 0) sig_arrive(s_sig, kSigIns + (int)(seq & (kSigRing - 1)));  // (behind the team barrier)
      LZ_PROF(kProfLink);
      const long long t_ws = P.prof ? clock64() : 0;
      search_tile(s_data, s_prev, t_mlen, t_mdist, tile_rp, tile_len, crp + chunk_len, lo_rp, t_ctr, LZ_EFFORT(depth, kDepth));
      if (P.prof && team == 0 && lane == 0)
        atomicAdd(&P.prof[kProfWarpSearch], (unsigned long long)(clock64() - t_ws) / kTeamWarps);
      group_sync(g_team);
      LZ_PROF(kProfSearch);
      if (flat_candidate && t == (int)team) {
        // condition (2) of the stored exit: what the shallow search found in the first tile of every team
        const uint2 m8 = *reinterpret_cast<const uint2 *>(t_mlen + g_team.tid * kRun);  // kRun == 8 positions
        uint32_t found = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) found += (((m8.x >> (8 * j)) & 0xffu) ? 1u : 0u) + (((m8.y >> (8 * j)) & 0xffu) ? 1u : 0u);
        if (g_team.tid * kRun >= tile_len) found = 0;
        for (int o = 16; o > 0; o >>= 1) found += __shfl_xor_sync(0xffffffffu, found, o);
        if (lane == 0 && found) atomicAdd(&s_misc[kMProbe], found);
        __syncthreads();
        if (s_misc[kMProbe] <= (uint32_t)kFlatMaxMatches) {
          // The chunk is stored. Every tile that was taken completes both of its signals.
          if (g_team.tid == 0) sig_arrive(s_sig, kSigEmit + (int)(seq & (kSigRing - 1)));
          stored_exit = true;
          break;
        }
      }
      // -- stage B: the refine rounds parse from the tile start (the true first token start is not known before tile
      // t-1 has been parsed for good; a parse from a wrong start joins the true one within a few tokens, and this
      // way the choice of deep-searched positions does not depend on timing)
      for (int r = 0; r < LZ_EFFORT(refine_rounds, kRefineRounds); r++) {
        if (g_team.tid == 0) *t_ctr = 0;  // (idle since the barrier behind the last search; published by the parse's barrier)
        if (g_team.tid == 0) *t_nchain = 0;
        parse_tile_wide(t_mlen, t_bitmap, 0, tile_len, g_team);
        LZ_PROF(kProfRefParse);
        const uint32_t nwork = refine_build_list(t_mlen, t_mdist, t_bitmap, t_deep, t_work, t_scan, tile_len, g_team);
        LZ_PROF(kProfRefList);
        if (nwork == 0) break;  // nothing to look at more deeply (random, zero and record pages): a further round would
                                // parse the same lengths again and find the same empty list (uniform: nwork is a scan total)
        const long long t_wd = P.prof ? clock64() : 0;
        refine_search(s_data, s_prev, t_mlen, t_mdist, t_work, nwork, tile_rp, tile_len, crp + chunk_len, lo_rp, t_ctr,
                      LZ_EFFORT(deep_depth, kDeepDepth), 0u, t_deep, t_chain, t_nchain);
        group_sync(g_team);
        // the second list: the positions behind the matches that grew, at half the depth, no further chaining
        const uint32_t nchain = *t_nchain;  // uniform
        if (nchain)  // (the counter stands at nwork + one failed take per thread)
          refine_search(s_data, s_prev, t_mlen, t_mdist, t_chain, nchain, tile_rp, tile_len, crp + chunk_len, lo_rp, t_ctr,
                        LZ_EFFORT(deep_depth, kDeepDepth) / kChainDepthDiv, nwork + (uint32_t)kTeamThreads);
        if (P.prof && team == 0 && lane == 0)
          atomicAdd(&P.prof[kProfWarpDeep], (unsigned long long)(clock64() - t_wd) / kTeamWarps);
        group_sync(g_team);
        LZ_PROF(kProfRefDeep);
      }
      // -- stage C: final parse and token emission, in tile order
      if (t > 0) {  // tile t-1 has emitted: kMStart and kMNTok are ours
        if (g_team.tid < 32) sig_wait(s_sig, kSigEmit + (int)((seq - 1) & (kSigRing - 1)), (seq - 1) / kSigRing);
        group_sync(g_team);
      }
      LZ_PROF(kProfWaitB);
      const int nxt = parse_tile_wide(t_mlen, t_bitmap, (int)s_misc[kMStart], tile_len, g_team);
      LZ_PROF(kProfParse);
      // emit tokens: rank of every start position, then token words + histograms
      {
        const int t0 = g_team.tid * kRun;
        const uint32_t bits = (t_bitmap[t0 >> 5] >> (t0 & 31)) & ((1u << kRun) - 1u);
        uint32_t total;
        uint32_t rank = group_exclusive_scan((uint32_t)__popc(bits), t_scan, &total, g_team);
        const uint32_t base = s_misc[kMNTok];
        uint32_t m = bits;
        while (m) {
          const int j = __ffs((int)m) - 1;
          m &= m - 1;
          const int tp = t0 + j;
          const uint32_t l = t_mlen[tp];
          uint32_t tok;
          // the lazy rule may have deferred a match: a start with a longer match right behind it is a literal
          const bool lit = l == 0 || (tp + 1 < tile_len && t_mlen[tp + 1] > l);
          if (lit) {
            tok = s_data[ring_slot(tile_rp + tp)];
            atomicAdd(&s_lfreq[tok], 1u);
          } else {
            const uint32_t d = t_mdist[tp] & 0x7fffu;
            tok = kTokMatch | (l << 16) | (d - 1);  // l = len-3
            atomicAdd(&s_lfreq[257 + len_code(l + 3)], 1u);
            atomicAdd(&s_dfreq[dist_code(d)], 1u);
          }
          tokens[base + rank++] = tok;
        }
        group_sync(g_team);  // everybody has read kMNTok and kMStart
        if (g_team.tid == 0) {
          s_misc[kMNTok] = base + total;
          s_misc[kMStart] = (uint32_t)nxt;
          sig_arrive(s_sig, kSigEmit + (int)(seq & (kSigRing - 1)));  // release: the two words above are visible
          s_misc[kMTeamTile + team] = atomicAdd(&s_misc[kMTileCtr], 1u);    // the team's next tile
        }
        group_sync(g_team);
        t = (int)s_misc[kMTeamTile + team];
      }
      LZ_PROF(kProfEmit);
    }
    __syncthreads();
    LZ_PROF(kProfJoin);
    seq_base += stored_exit ? (uint32_t)kTeams : (uint32_t)ntiles;  // the tiles that were taken (and have signalled)
    if (stored_exit) {  // uniform: decided from shared memory behind a CTA-wide barrier
      if (tid == 0) {
        const uint32_t n_stored_blocks = ((uint32_t)chunk_len + kStoredMax - 1) / kStoredMax;
        P.chunk_meta[c] = (((uint32_t)chunk_len + 5 * n_stored_blocks) << 1) | 1u;  // the gather kernel copies the input
        s_misc[kMStale] = 1;
      }
      continue;  // the next chunk starts with a barrier
    }

    // ---- Huffman tables and exact sizes ------------------------------------------------------------------
    if (tid == 0) {
      tokens[s_misc[kMNTok]] = kEob;  // end-of-block travels as the last "literal" symbol
      s_misc[kMNTok] += 1;
      s_lfreq[kEob] += 1;
    }
    __syncthreads();
    HuffScratch *hs = reinterpret_cast<HuffScratch *>(smem + LzLayout::huff);  // two scratches, 4064 bytes apart
    {
      uint32_t *ctrl = reinterpret_cast<uint32_t *>(hs + 2) + 128;  // behind `comb` (512 bytes) in the head area
      HuffJob jl{s_lfreq, s_llen, s_lcode, &hs[0], ctrl, kNumLitLen, 0};
      HuffJob jd{s_dfreq, s_dlen, s_dcode, &hs[1], ctrl + 40, kNumDist, 288};
      huff_jobs_build(jl, jd);
      if (tid == 0) {
        const uint32_t max_l = ctrl[2], max_d = ctrl[40 + 2];
        s_misc[kMHlit] = (max_l < 256 ? 256 : max_l) + 1;
        s_misc[kMHdist] = max_d + 1;
      }
      __syncthreads();
    }
    LZ_PROF(kProfHuff);
    // body sizes: sum freq * len (+ extra bits) for the dynamic and the fixed code
    {
      uint32_t dyn = 0, fix = 0;
      for (uint32_t s = tid; s < (uint32_t)(kNumLitLen + kNumDist); s += kLzThreads) {
        if (s < (uint32_t)kNumLitLen) {
          const uint32_t f = s_lfreq[s];
          const uint32_t eb = s > 256 ? len_extra_bits(s - 257) : 0;
          dyn += f * (s_llen[s] + eb);
          fix += f * (fixed_litlen_bits(s) + eb);
        } else {
          const uint32_t dsy = s - kNumLitLen;
          const uint32_t f = s_dfreq[dsy];
          const uint32_t eb = dist_extra_bits(dsy);
          dyn += f * (s_dlen[dsy] + eb);
          fix += f * (5 + eb);
        }
      }
      for (int o = 16; o > 0; o >>= 1) {
        dyn += __shfl_xor_sync(0xffffffffu, dyn, o);
        fix += __shfl_xor_sync(0xffffffffu, fix, o);
      }
      if (lane == 0) {
        s_misc[72 + warp] = dyn;
        s_misc[112 + warp] = fix;
      }
    }
    __syncthreads();
    if (tid == 0) {
      uint32_t body_dyn = 0, body_fix = 0;
      for (int w = 0; w < kLzWarps; w++) {
        body_dyn += s_misc[72 + w];
        body_fix += s_misc[112 + w];
      }
      // code-length code
      const uint32_t hlit = s_misc[kMHlit], hdist = s_misc[kMHdist];
      uint8_t *comb = reinterpret_cast<uint8_t *>(hs + 2);  // 316 bytes behind the two tree scratches
      for (uint32_t i = 0; i < hlit; i++) comb[i] = s_llen[i];
      for (uint32_t i = 0; i < hdist; i++) comb[hlit + i] = s_dlen[i];
      uint32_t cf[kNumCodeLen];
      for (int i = 0; i < kNumCodeLen; i++) cf[i] = 0;
      const int nr = rle_code_lengths(comb, (int)(hlit + hdist), s_rle, cf);
      huff_build_lengths(cf, kNumCodeLen, kMaxClBits, s_cllen, hs[0]);
      huff_make_codes(s_cllen, kNumCodeLen, s_clcode);
      int hclen = kNumCodeLen;
      while (hclen > 4 && s_cllen[cl_order(hclen - 1)] == 0) hclen--;
      uint32_t hdr = 3 + 5 + 5 + 4 + 3 * (uint32_t)hclen;
      for (int i = 0; i < nr; i++) hdr += s_cllen[s_rle[i] & 0xff] + cl_extra_bits(s_rle[i] & 0xff);
      const uint32_t dyn_bits = hdr + body_dyn;
      const uint32_t fix_bits = 3 + body_fix;
      const uint32_t n_stored_blocks = ((uint32_t)chunk_len + kStoredMax - 1) / kStoredMax;
      const uint32_t stored_bytes = (uint32_t)chunk_len + 5 * n_stored_blocks;
      // compressed encodings end with the sync marker: 3 header bits, pad to a byte, 00 00 ff ff
      const uint32_t dyn_bytes = (dyn_bits + 3 + 7) / 8 + 4;
      const uint32_t fix_bytes = (fix_bits + 3 + 7) / 8 + 4;
      uint32_t mode, bytes, hbits;
      if (stored_bytes <= dyn_bytes && stored_bytes <= fix_bytes) {
        mode = 0; bytes = stored_bytes; hbits = 0;
      } else if (fix_bytes <= dyn_bytes) {
        mode = 1; bytes = fix_bytes; hbits = 3;
      } else {
        mode = 2; bytes = dyn_bytes; hbits = hdr;
      }
      s_misc[kMMode] = mode;
      s_misc[kMChunkBytes] = bytes;
      s_misc[kMHdrBits] = hbits;
      s_misc[kMHclen] = (uint32_t)hclen;
      s_misc[kMNRle] = (uint32_t)nr;
      P.chunk_meta[c] = (bytes << 1) | (mode == 0 ? 1u : 0u);
    }
    __syncthreads();
    LZ_PROF(kProfSizes);
    const uint32_t mode = s_misc[kMMode];
    const uint32_t chunk_bytes = s_misc[kMChunkBytes];
    uint8_t *dst = P.slots + (size_t)c * lz_slot_bytes(P.chunk_bytes);

    if (mode == 0) {
      // stored: the gather kernel copies the bytes straight from the input
    } else {
      // ---- bit packing ----------------------------------------------------------------------------------
      const uint32_t ntok = s_misc[kMNTok];
      for (uint32_t i = tid; i < (uint32_t)kStageBytes / 4; i += kLzThreads) s_stage[i] = 0;
      __syncthreads();
      if (tid == 0) {
        SmemBitWriter bw{s_stage, 0};
        if (mode == 1) {
          bw.put(2, 3);  // BFINAL=0, BTYPE=01
        } else {
          bw.put(4, 3);  // BFINAL=0, BTYPE=10
          bw.put(s_misc[kMHlit] - 257, 5);
          bw.put(s_misc[kMHdist] - 1, 5);
          bw.put(s_misc[kMHclen] - 4, 4);
          for (uint32_t i = 0; i < s_misc[kMHclen]; i++) bw.put(s_cllen[cl_order((int)i)], 3);
          for (uint32_t i = 0; i < s_misc[kMNRle]; i++) {
            const uint32_t sym = s_rle[i] & 0xff, ex = s_rle[i] >> 8;
            bw.put(s_clcode[sym], s_cllen[sym]);
            bw.put(ex, cl_extra_bits(sym));
          }
        }
      }
      __syncthreads();
      uint32_t bitpos = s_misc[kMHdrBits];  // absolute bit position inside the chunk's output
      uint32_t flushed = 0;                 // bytes already written to global memory
      for (uint32_t b0 = 0; b0 < ntok; b0 += kBatchTokens) {
        // each thread encodes kTokPerThread consecutive tokens
        uint64_t val[kTokPerThread];
        uint32_t nb[kTokPerThread];
        uint32_t mine = 0;
#pragma unroll
        for (int k = 0; k < kTokPerThread; k++) {
          const uint32_t ti = b0 + tid * kTokPerThread + k;
          uint64_t v = 0;
          uint32_t n = 0;
          if (ti < ntok) {
            const uint32_t tok = tokens[ti];
            if (tok & kTokMatch) {
              const uint32_t len = ((tok >> 16) & 0xff) + 3, dist = (tok & 0x7fff) + 1;
              const uint32_t lc = len_code(len), dc = dist_code(dist);
              const uint32_t leb = len_extra_bits(lc), deb = dist_extra_bits(dc);
              uint32_t lcd, lcl, dcd, dcl;
              if (mode == 2) {
                lcd = s_lcode[257 + lc]; lcl = s_llen[257 + lc]; dcd = s_dcode[dc]; dcl = s_dlen[dc];
              } else {
                const uint32_t sym = 257 + lc;
                lcl = fixed_litlen_bits(sym);
                lcd = bit_reverse(sym < 280 ? sym - 256 : 0xc0 + (sym - 280), (int)lcl);
                dcl = 5;
                dcd = bit_reverse(dc, 5);
              }
              v = lcd;
              n = lcl;
              v |= (uint64_t)(len - len_base(lc)) << n;
              n += leb;
              v |= (uint64_t)dcd << n;
              n += dcl;
              v |= (uint64_t)(dist - dist_base(dc)) << n;
              n += deb;
            } else {
              if (mode == 2) {
                v = s_lcode[tok];
                n = s_llen[tok];
              } else {
                n = fixed_litlen_bits(tok);
                const uint32_t code = tok < 144 ? 0x30 + tok : (tok < 256 ? 0x190 + (tok - 144) : tok - 256);
                v = bit_reverse(code, (int)n);
              }
            }
          }
          val[k] = v;
          nb[k] = n;
          mine += n;
        }
        uint32_t batch_bits;
        uint32_t off = block_exclusive_scan(mine, s_misc + kMScan, &batch_bits);
        uint32_t o = bitpos - flushed * 8 + off;  // bit offset inside the window
#pragma unroll
        for (int k = 0; k < kTokPerThread; k++) {
          if (nb[k]) {
            const uint32_t w = o >> 5, s = o & 31;
            const uint64_t v = val[k];
            const uint32_t a0 = (uint32_t)(v << s);
            const uint32_t a1 = s ? (uint32_t)(v >> (32 - s)) : (uint32_t)(v >> 32);
            const uint32_t a2 = s ? (uint32_t)(v >> (64 - s)) : 0u;
            if (a0) atomicOr(&s_stage[w], a0);
            if (a1) atomicOr(&s_stage[w + 1], a1);
            if (a2) atomicOr(&s_stage[w + 2], a2);
            o += nb[k];
          }
        }
        bitpos += batch_bits;
        __syncthreads();
        // flush whole bytes; carry the partial byte to the front of the window
        const uint32_t whole = bitpos / 8 - flushed;
        const bool last = b0 + kBatchTokens >= ntok;
        if (!last) {
          // every thread reads the partial byte itself; only the words this batch touched are cleared again
          const uint32_t carry = (bitpos & 7) ? reinterpret_cast<const uint8_t *>(s_stage)[whole] : 0u;
          block_copy_unaligned<true>(dst + flushed, reinterpret_cast<const uint8_t *>(s_stage), whole, tid, kLzThreads);
          __syncthreads();
          for (uint32_t i = tid; i < whole / 4 + 2; i += kLzThreads) s_stage[i] = i == 0 ? carry : 0u;
          flushed += whole;
          __syncthreads();
        } else {
          // sync flush marker: stored block header (000), pad to byte, LEN=0000 NLEN=ffff
          if (tid == 0) {
            uint32_t p = bitpos - flushed * 8 + 3;  // three zero bits
            p = (p + 7) & ~7u;
            uint8_t *sb = reinterpret_cast<uint8_t *>(s_stage) + p / 8;
            sb[0] = 0;
            sb[1] = 0;
            sb[2] = 0xff;
            sb[3] = 0xff;
          }
          __syncthreads();
          const uint32_t remain = chunk_bytes - flushed;
          block_copy_unaligned<true>(dst + flushed, reinterpret_cast<const uint8_t *>(s_stage), remain, tid, kLzThreads);
        }
      }
    }
    __syncthreads();
#if !defined(DSG_EMU)
    {
      // The chunk's token list is dead now. Dropping its lines from L2 keeps them from being written back to DRAM
      // later (with 256 KiB chunks the lists of all CTAs are 70 MB: measured 7 GB of dead write-backs per 8 GiB).
      const uint32_t lines = s_misc[kMNTok] * 4u / 128u;
      const char *tb = reinterpret_cast<const char *>(tokens);
      for (uint32_t i = tid; i < lines; i += kLzThreads)
        asm volatile("discard.global.L2 [%0], 128;" ::"l"(tb + (size_t)i * 128) : "memory");
    }
#endif
    LZ_PROF(kProfPack);
    }  // chunks of the run
  }
#undef LZ_PROF
}

// Exclusive prefix of the chunk sizes (single CTA): offs[c] = byte offset of chunk c in the output,
// *total_out = all chunks (+2 for the terminator); *error = 1 when the output buffer is too small.
__global__ void __launch_bounds__(1024) lz_scan_kernel(const uint32_t *__restrict__ chunk_meta, uint32_t nchunks,
                                                        uint64_t *__restrict__ offs, uint64_t out_cap,
                                                        int write_terminator, uint64_t *__restrict__ total_out,
                                                        uint32_t *__restrict__ error) {
  __shared__ uint64_t warp_tot[32];
  __shared__ uint64_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t base = 0; base < nchunks; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint64_t v = i < nchunks ? (uint64_t)(chunk_meta[i] >> 1) : 0;
    uint64_t x = v;
    for (int o = 1; o < 32; o <<= 1) {
      uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (uint32_t)o) x += y;
    }
    if (lane == 31) warp_tot[warp] = x;
    __syncthreads();
    if (warp == 0) {
      const uint64_t t = warp_tot[lane];
      uint64_t sc = t;
      for (int o = 1; o < 32; o <<= 1) {
        uint64_t y = __shfl_up_sync(0xffffffffu, sc, o);
        if (lane >= (uint32_t)o) sc += y;
      }
      warp_tot[lane] = sc - t;
    }
    __syncthreads();
    const uint64_t excl = carry_s + warp_tot[warp] + (x - v);
    if (i < nchunks) offs[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const uint64_t tot = carry_s + (write_terminator ? 2 : 0);
    *total_out = tot;
    if (tot > out_cap) atomicExch(error, 1u);
  }
}

// One CTA per chunk: moves the chunk's bytes from its slot (or, for stored chunks, from the input behind
// 5-byte stored-block headers) to its final offset. The last chunk appends the terminator.
__global__ void __launch_bounds__(256)
lz_gather_kernel(const uint8_t *__restrict__ in, uint64_t len, const uint8_t *__restrict__ slots,
                 const uint32_t *__restrict__ chunk_meta, const uint64_t *__restrict__ offs, uint32_t nchunks,
                 uint32_t chunk_bytes, uint32_t big_chunks, uint8_t *__restrict__ out, int write_terminator,
                 const uint32_t *__restrict__ error) {
  if (*error) return;  // output too small: nothing is written
  const uint32_t c = blockIdx.x;
  const uint32_t meta = chunk_meta[c];
  const uint32_t bytes = meta >> 1;
  uint8_t *dst = out + offs[c];
  if (meta & 1u) {
    const uint64_t cs = lz_chunk_offset(c, chunk_bytes, big_chunks);
    const uint64_t csz = lz_chunk_size(c, chunk_bytes, big_chunks);
    const uint32_t chunk_len = (uint32_t)(len - cs < csz ? len - cs : csz);
    uint32_t done = 0, o = 0;
    while (done < chunk_len) {
      const uint32_t n = chunk_len - done < kStoredMax ? chunk_len - done : kStoredMax;
      if (threadIdx.x == 0) {
        dst[o + 0] = 0;  // BFINAL=0, BTYPE=00
        dst[o + 1] = (uint8_t)(n & 0xff);
        dst[o + 2] = (uint8_t)(n >> 8);
        dst[o + 3] = (uint8_t)(~n & 0xff);
        dst[o + 4] = (uint8_t)((~n >> 8) & 0xff);
      }
      block_copy_unaligned(dst + o + 5, in + cs + done, n, threadIdx.x, 256);
      done += n;
      o += n + 5;
    }
  } else {
    block_copy_unaligned(dst, slots + (size_t)c * lz_slot_bytes(chunk_bytes), bytes, threadIdx.x, 256);
  }
  if (write_terminator && c == nchunks - 1 && threadIdx.x == 0) {
    dst[bytes] = 0x03;  // BFINAL=1, BTYPE=01, end-of-block
    dst[bytes + 1] = 0x00;
  }
}

}  // namespace dsg
