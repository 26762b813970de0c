/*
 * dsgpack.h -- C ABI of libdsgpack.so: B200-native tar+gzip packing for DevSpace's upstream sync /
 * kaniko build-context path.
 *
 * Drop-in boundary: the Go seam that stays is
 *     func writeTar(files []*fileInformation, config *SyncConfig) (string, map[string]*fileInformation, error)
 *                                                     /root/reference/pkg/devspace/sync/tar.go:146
 * The Go side keeps the walk, exclude matching, os.Stat, the fileMap lookups and reading file bytes;
 * what crosses this ABI is the list of tar headers it would have handed to tw.WriteHeader
 * (tar.go:253,294) as POD `dsg_entry` records, plus the payload bytes it would have io.Copy'd
 * (tar.go:298) in one arena. What comes back is the byte stream that gzip.NewWriter(f) (tar.go:156)
 * would have received: a standard gzip file whose decompressed content is bit-identical to the
 * archive/tar stream. INTEGRATION.md shows the cgo stub.
 *
 * Conventions: every function returns 0 (DSG_OK) or a negative dsg_status; no torch / C++ types;
 * all pointers are plain host pointers unless the parameter name starts with `d_` (device pointer on
 * the context's device). A context serves one caller at a time (SURVEY.md 8b "Threading").
 * There is NO CPU fallback: without a usable CUDA device every compute entry point fails with
 * DSG_E_CUDA.
 */
#ifndef DSGPACK_H
#define DSGPACK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DSG_ABI_VERSION 2

typedef enum dsg_status {
  DSG_OK = 0,
  DSG_E_ARG = -1,        /* NULL / inconsistent arguments                                          */
  DSG_E_CUDA = -2,       /* CUDA runtime error; text via dsg_last_error()                           */
  DSG_E_HEADER = -3,     /* archive/tar would return ErrHeader / ErrFieldTooLong for some entry     */
  DSG_E_TYPEFLAG = -4,   /* typeflag other than '0' / '5' (tar.go only emits files and empty dirs)  */
  DSG_E_CAPACITY = -5,   /* caller's output buffer too small; required size in *out_len             */
  DSG_E_UNCLEAN_NAME = -6, /* a PAX entry's name is not path.Clean()ed (tar.go names always are)    */
  DSG_E_NOMEM = -7,      /* device / pinned allocation failed                                       */
  DSG_E_RANGE = -8,      /* payload range outside the arena                                         */
  DSG_E_CORRUPT = -9     /* dsg_gunzip: the stream fails its structure or CRC32 check                */
} dsg_status;

/*
 * One tar.Header as tarFile / tarFolder build it before tw.WriteHeader:
 *   tar.FileInfoHeader(stat) (tar.go:242,280) -> hdr.Name = relative path (tar.go:243,284)
 *   -> Mode/Uid/Gid <- RemoteMode/RemoteUID/RemoteGID when the path is indexed (tar.go:246-250,287-291).
 * 80 bytes, little endian, no pointers: the array is copied to the device as is.
 * Strings live in a caller-provided string table (`strtab`); several entries may share one string.
 */
typedef struct dsg_entry {
  uint64_t data_off;   /* payload offset in the arena (bytes); ignored when size == 0 or typeflag '5'.
                        * typeflag '1' / '2': hdr.Linkname instead, as strtab offset | (length << 32)  */
  int64_t size;        /* hdr.Size: stat.Size() for files, 0 for directories                         */
  int64_t mtime_sec;   /* stat.ModTime() seconds; rounding to the nearest second (WriteHeader,        */
  int64_t mode;        /*   Format unset) is done on the device. mode: perm|04000|02000|01000         */
  int64_t uid;         /* hdr.Uid / hdr.Gid (Go int)                                                  */
  int64_t gid;
  uint32_t mtime_nsec; /* stat.ModTime() nanoseconds, 0..999999999                                    */
  uint32_t name_off;   /* hdr.Name: bytes [name_off, name_off+name_len) of strtab                     */
  uint32_t name_len;
  uint32_t uname_off;  /* hdr.Uname (passwd lookup done by tar.FileInfoHeader on the host)            */
  uint32_t gname_off;  /* hdr.Gname                                                                   */
  uint16_t uname_len;
  uint16_t gname_len;
  uint8_t typeflag;    /* '0' regular file, '5' directory; '1' hard link, '2' symlink (docker dialect) */
  uint8_t reserved[7]; /* must be zero                                                                */
} dsg_entry;

typedef struct dsg_ctx dsg_ctx; /* opaque */

/* Compression level semantics follow compress/gzip: the reference hard-wires DefaultCompression
 * (gzip.NewWriter, tar.go:156). DSG_LEVEL_STORED exists for incompressible inputs and tests. */
#define DSG_LEVEL_STORED 0
#define DSG_LEVEL_DEFAULT 6

typedef struct dsg_opts {
  uint32_t struct_size;    /* sizeof(dsg_opts), for forward compatibility                           */
  int32_t level;           /* DSG_LEVEL_DEFAULT or DSG_LEVEL_STORED                                  */
  uint64_t segment_bytes;  /* tar-stream bytes per pipeline segment (0 = a sixth of the stream,
                            * between 64 and 512 MiB)                                            */
  uint32_t flags;          /* DSG_F_*                                                                */
  uint32_t gunzip_batch_mib; /* dsg_gunzip: decompressed MiB per batch of chunks (0 = 2048); bounds the scratch */
  /* Launch geometry of the level-6 kernel. All zero (the default) = chosen from the input size. The parity
   * tests and the profiling tools set these to reach every geometry with small inputs; a caller never has to. */
  uint32_t lz_chunk_units; /* deflate chunk = 1, 2 or 4 units of 64 KiB (0 = automatic)                    */
  uint32_t lz_run_chunks;  /* consecutive chunks one CTA compresses with one sliding window, 1..64 (0 = auto) */
  uint32_t lz_grid;        /* CTAs of the launch (0 = one per SM)                                           */
  uint32_t lz_flags;       /* DSG_LZ_*                                                                      */
  uint64_t lz_tail_bytes;  /* with DSG_LZ_TAIL_SET: bytes at the end of a launch handed out in 64 KiB chunks */
} dsg_opts;

#define DSG_F_NO_TRAILER 1u /* omit the 1024-byte tar trailer: for every shard but the last */
#define DSG_LZ_TAIL_SET 1u  /* lz_tail_bytes is meaningful (0 is a valid value)                             */
#define DSG_LZ_PROFILE 2u   /* per-phase cycle counts of the level-6 kernel on stderr (development aid)     */
#define DSG_LZ_NO_BULK 4u   /* stage the tiles with plain loads instead of cp.async.bulk (measurement aid)  */
/* bits 8..29 of lz_flags: search effort for tools/sweep_effort.py; read only by a library built with
 * -DDSG_LZ_RUNTIME_EFFORT (the shipped one compiles the effort in and ignores them) */

/* Per-call timing and size counters (bench.py reads them; times are CUDA-event milliseconds). */
typedef struct dsg_stats {
  uint64_t entries;        /* N                                                                      */
  uint64_t payload_bytes;  /* P = sum of file sizes                                                  */
  uint64_t tar_bytes;      /* T = tar stream length                                                  */
  uint64_t out_bytes;      /* C = gzip bytes produced                                                */
  uint64_t h2d_bytes;      /* bytes copied host -> device (manifest + arena)                         */
  uint64_t d2h_bytes;      /* bytes copied device -> host                                            */
  float ms_total;          /* first enqueue -> last byte landed / produced                           */
  float ms_plan;           /* tar plan kernels (classify + scan)                                     */
  float ms_tar;            /* header + payload layout kernels                                        */
  float ms_deflate;        /* LZ77 + Huffman + bit-pack kernel(s)                                    */
  float ms_crc;            /* CRC32 kernel + combine                                                 */
  float ms_gather;         /* compaction of per-chunk outputs                                        */
  uint32_t kernel_launches;/* kernels launched by this call                                          */
  uint32_t segments;       /* pipeline segments processed                                            */
  uint32_t crc32;          /* CRC32-IEEE of the tar stream (gzip trailer)                            */
  uint32_t reserved;
} dsg_stats;

/* ---- context ------------------------------------------------------------------------------------ */

/* Replaces: nothing in Go (gzip.NewWriter / tar.NewWriter carry the per-archive state, tar.go:156,159);
 * here the context owns the device, streams and reusable device/pinned buffers. */
int dsg_ctx_create(int device, const dsg_opts *opts, dsg_ctx **out);
void dsg_ctx_destroy(dsg_ctx *ctx);
/* Replaces the options of a live context (level, segment size, flags, launch geometry) between calls. */
int dsg_ctx_set_opts(dsg_ctx *ctx, const dsg_opts *opts);
const char *dsg_last_error(const dsg_ctx *ctx);   /* text of the last failure (CUDA error string etc.) */
const char *dsg_strerror(int status);
int dsg_abi_version(void);
int dsg_device_count(void);                       /* 0 when no CUDA device is usable                   */

/* Page-locked staging memory for the arena / output (the Go shim wraps it with unsafe.Slice; it stands
 * where tarFile's io.Copy source buffer and the temp file's write buffer were, tar.go:147,298). */
void *dsg_pinned_alloc(dsg_ctx *ctx, size_t bytes);
void dsg_pinned_free(dsg_ctx *ctx, void *p);

/* ---- host-side layout helpers (pure integer work on the manifest; no GPU) ------------------------- */

/* Bytes tw.WriteHeader emits for one entry: 512, or 512 + pad512(PAX records) + 512.
 * Same decision procedure as Header.allowedFormats in go1.11 archive/tar. */
int dsg_entry_header_bytes(const dsg_entry *e, const char *strtab, size_t strtab_len, uint64_t *out);

/* T: total tar stream length for the manifest, including the 1024-byte trailer (Writer.Close) unless
 * flags has DSG_F_NO_TRAILER. */
int dsg_tar_size(const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len, uint32_t flags,
                 uint64_t *out);

/* Upper bound of the gzip output for a tar stream of `tar_bytes` bytes (for sizing `out`). */
uint64_t dsg_gzip_bound(uint64_t tar_bytes);

/* One gzip member out of the shards' members (for decoders that stop after the first member; GNU
 * gzip/tar and CPython read the plain concatenation as well). members[i] / member_lens[i]: the output of dsg_pack
 * for shard i, in shard order; tar_lens[i]: that shard's tar-stream length (dsg_tar_size with the shard's flags, or
 * dsg_stats.tar_bytes). Host work only, no context: strips the inner headers, trailers and block terminators and
 * folds the checksums (the crc32_combine identity). DSG_E_CAPACITY with *out_len = the size needed; DSG_E_ARG when
 * a piece is not a member this library wrote or tar_lens[i] does not match its ISIZE. No counterpart in the
 * reference (it writes one member with one writer). */
int dsg_stitch_members(const uint8_t *const *members, const size_t *member_lens, const uint64_t *tar_lens, size_t n,
                       uint8_t *out, size_t out_cap, size_t *out_len);

/* CRC32-IEEE of A || B from crc(A), crc(B) and len(B) (x^(8 len_b) mod P): what a caller that concatenates the
 * shards' members itself needs for a single trailer. Host integer work; hash/crc32 has no such function, the Go
 * side would otherwise have to re-read the stream. */
uint32_t dsg_crc32_combine(uint32_t crc_a, uint32_t crc_b, uint64_t len_b);

/* Contiguous entry ranges balanced by tar-stream bytes for `nshards` GPUs:
 * shard s owns entries [bounds[s], bounds[s+1]). `bounds` has nshards+1 slots. Files are never split. */
int dsg_shard_bounds(const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len, int nshards,
                     uint64_t *bounds);

/* ---- the hot path ------------------------------------------------------------------------------- */

/* Replaces the body of writeTar (tar.go:146-182) below the walk: tar.NewWriter(gzip.NewWriter(f)),
 * WriteHeader + io.Copy per entry, tarWriter.Close, gw.Close. Host buffers in, host buffer out;
 * copies are pipelined with the kernels per segment. `out` receives one complete gzip member.
 * On DSG_E_CAPACITY *out_len holds the size that would have been needed. */
int dsg_pack(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
             const uint8_t *arena, size_t arena_len, uint8_t *out, size_t out_cap, size_t *out_len);

/* Same, without gzip: the uncompressed archive/tar stream (what tar.NewWriter(w) writes to w).
 * Next-row use: the docker builder's uncompressed context (builder/docker/docker.go:94-97). */
int dsg_pack_tar_only(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                      const uint8_t *arena, size_t arena_len, uint8_t *out, size_t out_cap, size_t *out_len);

/* dsg_pack / dsg_pack_tar_only (gzip = 1 / 0) with the stream written to the file descriptor `fd` instead of a
 * caller buffer: writeTar's temp file (tar.go:148-157; SURVEY.md 8f rank 3, "overlap D2H with the temp-file
 * write"). Every pipeline segment lands in one of two pinned staging buffers of the context and is written while
 * the GPU works on the following segments; no output bound has to be allocated. *out_len = bytes written. On an
 * error part of the stream may have been written already (the caller removes the file, as tar.go:172 does). */
int dsg_pack_fd(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                const uint8_t *arena, size_t arena_len, int fd, int gzip, uint64_t *out_len);

/* Device-resident variant of dsg_pack for callers that already hold manifest + arena in HBM
 * (bench.py's kernel-only number): every `d_` pointer is device memory on the context's device.
 * *out_len is written on the host after the call's own synchronisation. */
int dsg_pack_device(dsg_ctx *ctx, const dsg_entry *d_entries, size_t n, const char *d_strtab, size_t strtab_len,
                    const uint8_t *d_arena, size_t arena_len, uint8_t *d_out, size_t out_cap, size_t *out_len);

/* Counters of the most recent dsg_pack* call on this context. */
int dsg_get_stats(const dsg_ctx *ctx, dsg_stats *out);

/* ---- one call, several GPUs, one ordered archive (devspace_b200/csrc/multipack.cu) ------------------------------ *
 * Replaces the same lines as dsg_pack -- ONE writeTar call per archive from one goroutine (upstream.go:342; kaniko's
 * whole tree, util.go:68-73) -- on a box with several GPUs: the header list is cut into ndev
 * contiguous ranges balanced by tar-stream bytes (dsg_shard_bounds), device devices[s] packs range s on a host thread
 * of its own, and the calling thread writes the members in shard order to `fd` (or `out`) while later shards are
 * still being packed. mode 0: the members back to back, a multi-member .tar.gz (GNU gzip/tar, CPython read it as one
 * stream); DSG_MULTI_ONE_MEMBER: stitched into a single member on the fly (for decoders that stop after the first
 * member). The same device may be listed more than once (several contexts on one GPU). Every shard's member is
 * staged in a pinned buffer of its context until its turn to be written. No CPU fallback: dsg_multi_create fails
 * with DSG_E_CUDA when a listed device is missing. */
typedef struct dsg_multi dsg_multi; /* opaque */
#define DSG_MULTI_ONE_MEMBER 1u
int dsg_multi_create(const int *devices, int ndev, const dsg_opts *opts, dsg_multi **out);
void dsg_multi_destroy(dsg_multi *m);
const char *dsg_multi_last_error(const dsg_multi *m);
int dsg_multi_shards(const dsg_multi *m);
int dsg_multi_get_stats(const dsg_multi *m, int shard, dsg_stats *out);   /* of the most recent call */
int dsg_multi_pack_fd(dsg_multi *m, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                      const uint8_t *arena, size_t arena_len, int fd, uint32_t mode, uint64_t *out_len);
int dsg_multi_pack(dsg_multi *m, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                   const uint8_t *arena, size_t arena_len, uint32_t mode, uint8_t *out, size_t out_cap, size_t *out_len);

/* ---- host mirror of the Go call site (devspace_b200/csrc/hostpack.cc; host code only: walk, stat, read) ------ */

/* fileInformation (pkg/devspace/sync/file_information.go:21-32) */
typedef struct dsg_file_info {
  const char *name;     /* Name: path relative to WatchPath ("" = the whole tree)                              */
  int64_t size;         /* Size                                                                                */
  int64_t mtime;        /* Mtime: seconds, rounded (roundMtime, util.go:87-89)                                 */
  int32_t is_directory; /* IsDirectory                                                                         */
  int32_t is_symlink;   /* IsSymbolicLink                                                                      */
  int64_t remote_mode;  /* RemoteMode / RemoteUID / RemoteGID: replace Mode/Uid/Gid when the path is indexed   */
  int64_t remote_uid;
  int64_t remote_gid;
} dsg_file_info;

/* ignoreMatcher.MatchesPath / uploadIgnoreMatcher.MatchesPath (tar.go:196-206): non-zero = excluded */
typedef int (*dsg_match_fn)(const char *relative_path, void *user);
typedef void (*dsg_log_fn)(const char *line, void *user);

/* The fields of SyncConfig that writeTar reads (sync_config.go): WatchPath, the two matchers, fileIndex.fileMap,
 * Logf. retry_sleep_ms: the 4 s of tar.go:174 (0 = 4000); max_retries: 0 = retry forever like the reference. */
typedef struct dsg_sync_config {
  const char *watch_path;
  dsg_match_fn ignore;
  void *ignore_user;
  dsg_match_fn upload_ignore;
  void *upload_ignore_user;
  const dsg_file_info *file_map;
  size_t file_map_len;
  dsg_log_fn log;
  void *log_user;
  uint32_t retry_sleep_ms;
  uint32_t max_retries;
} dsg_sync_config;

typedef struct dsg_written dsg_written; /* writtenFiles + the manifest that was packed */

/* recursiveTar / tarFolder / tarFile (tar.go:185-304) without the tar.Writer: the ordered header list and
 * writtenFiles for a change list. Host only (no GPU). */
int dsg_walk(const dsg_file_info *files, size_t nfiles, const dsg_sync_config *cfg, dsg_written **out);
size_t dsg_written_count(const dsg_written *w);
const dsg_file_info *dsg_written_get(const dsg_written *w, size_t i);
const dsg_entry *dsg_written_entries(const dsg_written *w, size_t *n, const char **strtab, size_t *strtab_len,
                                     uint64_t *arena_bytes);
void dsg_written_free(dsg_written *w);

/* writeTar (tar.go:146-182): temp file path in out_path, writtenFiles in *written; a failing walk / read is
 * logged ("[Upstream] Tar failed: ..."), the temp file removed, and the whole call retried after the sleep. */
int dsg_write_tar(dsg_ctx *ctx, const dsg_file_info *files, size_t nfiles, const dsg_sync_config *cfg, char *out_path,
                  size_t out_path_cap, dsg_written **written);

/* ---- single stages, device pointers (parity tests drive each kernel alone) ----------------------- */

/* archive/tar Writer: headers + payload layout + padding + trailer into d_tar[tar_cap]. */
int dsg_dev_tar_build(dsg_ctx *ctx, const dsg_entry *d_entries, size_t n, const char *d_strtab, size_t strtab_len,
                      const uint8_t *d_arena, size_t arena_len, uint32_t flags, uint8_t *d_tar, size_t tar_cap,
                      uint64_t *tar_len);

/* hash/crc32 ChecksumIEEE over d_buf[0,len) (gzip.Writer keeps it as z.digest). */
int dsg_dev_crc32(dsg_ctx *ctx, const uint8_t *d_buf, size_t len, uint32_t *crc);

/* Build-context hash, per-file part: hashFileCRC32, pkg/util/hash/hash.go:165-197
 * (crc32.New(crc32.MakeTable(0xedb88320)) over the file's bytes), which DirectoryExcludes calls for every context
 * file (hash.go:146) before each build decision (pkg/devspace/image/build.go:213). crcs[i] = CRC32-IEEE of entry
 * i's payload arena[data_off, data_off+size); 0 for directories and empty files (Go: the checksum of no bytes).
 * The manifest and arena are the ones dsg_pack / dsg_pack_device take, so an arena that is resident for packing
 * yields the checksums without another host->device pass. dsg_file_crc32: host buffers in, host array out
 * (arena ideally pinned); dsg_dev_file_crc32: everything device resident. DSG_E_RANGE when a payload range lies
 * outside the arena. The SHA-256 over the short "path;crc" list stays on the host (dsg_directory_hash below). */
int dsg_file_crc32(dsg_ctx *ctx, const dsg_entry *entries, size_t n, const uint8_t *arena, size_t arena_len,
                   uint32_t *crcs);
int dsg_dev_file_crc32(dsg_ctx *ctx, const dsg_entry *d_entries, size_t n, const uint8_t *d_arena, size_t arena_len,
                       uint32_t *d_crcs);

/* compress/flate Writer at `level`: raw deflate stream (byte aligned, BFINAL set on the last block)
 * of d_in[0,len) into d_out[out_cap]. */
int dsg_dev_deflate(dsg_ctx *ctx, const uint8_t *d_in, size_t len, int level, uint8_t *d_out, size_t out_cap,
                    size_t *out_len);

/* compress/gzip Writer: header + deflate + CRC32 + ISIZE for an existing device buffer. */
int dsg_dev_gzip(dsg_ctx *ctx, const uint8_t *d_in, size_t len, int level, uint8_t *d_out, size_t out_cap,
                 size_t *out_len);

/* No counterpart in the reference. CRC32-IEEE of d_buf[0,len) (device memory on the CURRENT device) for callers
 * that cannot go to the CUDA allocator when they need the checksum: a process whose device memory was handed
 * out completely at start-up, or a loop that cannot take cudaMalloc/cudaFree's device-wide synchronisation or an
 * out-of-memory failure mid-upload. Needs no dsg_ctx and allocates nothing: the per-tile partial results live
 * in a static device array inside the library (55 KB with the lookup tables, resident on every GPU that loads
 * the library until the process exits), which limits len to 1 GiB -- longer inputs return DSG_E_CAPACITY with a
 * message in err. The first call on a device does a one-time set-up (device current; make it before any stream
 * capture). All calls on a device share the array: do not let two calls overlap on the GPU. Runs on the default
 * stream and returns once *crc (host memory) holds the result. err/err_len: optional buffer for error text. */
int dsg_crc32_noalloc(const uint8_t *d_buf, size_t len, uint32_t *crc, char *err, size_t err_len);

/* ---- downstream direction ------------------------------------------------------------------ *
 * Replaces the decompression half of untarAll, tar.go:16-33 (gzip.NewReader(reader) in front of tar.NewReader), for
 * archives that THIS LIBRARY packed -- the case SURVEY.md names as the one worth accelerating: both ends run the
 * packer (a GPU pod packs the downstream archive instead of `tar -czf`, downstream.go:394). Such a stream inflates
 * chunk-parallel: every chunk starts on a byte boundary behind a sync marker or a stored block, the unit starts are
 * found without decoding, every unit is decoded on its own thread, and the references a chunk makes into the 32 KiB
 * before it are resolved afterwards (devspace_b200/csrc/inflate.cuh). Multi-member streams (dsg_multi_*, shard
 * members back to back) are taken as they are. Every member's ISIZE and CRC32 are checked. `out` receives the tar
 * stream; tar parsing and the file writes of untarNext (tar.go:36-144) stay on the host.
 * A gzip stream of any other writer (GNU gzip, Go) is one chain of bits and is refused with DSG_E_ARG -- there is no
 * serial fallback on the GPU and none on the CPU; the caller keeps gzip.NewReader for those. DSG_E_CAPACITY with
 * *out_len = the size needed; DSG_E_CORRUPT when a member fails its CRC32. */
int dsg_gunzip(dsg_ctx *ctx, const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, size_t *out_len);
int dsg_dev_gunzip(dsg_ctx *ctx, const uint8_t *d_in, size_t in_len, uint8_t *d_out, size_t out_cap, size_t *out_len);

/* Plain device memory helpers so that ctypes / cgo callers need no CUDA binding of their own. */
void *dsg_dev_alloc(dsg_ctx *ctx, size_t bytes);
void dsg_dev_free(dsg_ctx *ctx, void *d_ptr);
int dsg_dev_upload(dsg_ctx *ctx, void *d_dst, const void *src, size_t bytes);
int dsg_dev_download(dsg_ctx *ctx, void *dst, const void *d_src, size_t bytes);

/* ---- Build-context hash, host side -------------------------------------------------------
 * DirectoryExcludes(srcPath, excludePatterns), pkg/util/hash/hash.go:43-163: filepath.Walk order, the exclude
 * decisions, "filePath" for a directory and "filePath;crc32hex" for a file into SHA-256, lower-case hex out.
 * The pattern matcher of the reference (docker/pkg/fileutils) stays with the caller: `exclude` gets the relative
 * path as hash.go builds it ("./a/b") and returns 0 = keep, 1 = skip (a directory is not entered, hash.go:120,134),
 * 2 = leave this entry out but enter the directory (an exclusion pattern "!dir/x" lies below it, hash.go:122-131).
 * NULL keeps everything. File bytes are read into one pinned arena and checksummed by dsg_file_crc32 on the GPU.
 * out_hex: 64 hex digits + NUL. err/err_len: optional buffer for the error text. */
typedef int (*dsg_exclude_fn)(const char *rel_path, int is_dir, void *user);
int dsg_directory_hash(dsg_ctx *ctx, const char *src_path, dsg_exclude_fn exclude, void *user, char out_hex[65],
                       char *err, size_t err_len);

/* ---- Docker build-context tar, host side -------------------------------------------------
 * archive.TarWithOptions(contextDir, {ExcludePatterns, ChownOpts{0,0}}) as pkg/devspace/builder/docker/docker.go:94-97
 * calls it (github.com/docker/docker/pkg/archive, go.mod:22; not under /root/reference: restated, parity unpinned):
 * filepath.Walk order, no entry for the context directory itself, every directory an entry named "dir/", symlinks
 * ('2') and further names of a multiply linked inode ('1') as link entries, mode with the file-type bits, mtime
 * truncated to seconds, uid = gid = 0 with the owner's names kept, no compression. `exclude` as for
 * dsg_directory_hash, except that rel_path is "a/b" (TarWithOptions does not prefix "./"). Device nodes, FIFOs and
 * sockets are left out. The stream (what the reference pipes to the daemon) is written to out_path. */
int dsg_docker_context_tar(dsg_ctx *ctx, const char *src_path, dsg_exclude_fn exclude, void *user, const char *out_path,
                           uint64_t *tar_bytes, uint64_t *n_entries, char *err, size_t err_len);

#ifdef __cplusplus
}
#endif
#endif /* DSGPACK_H */
