/*
 * oracle/gotar.h -- TEST INFRASTRUCTURE ONLY (CPU oracle).
 *
 * CPU restatement of what DevSpace's upstream pack path emits at the tar level:
 *   pkg/devspace/sync/tar.go:146-304 (writeTar -> recursiveTar -> tarFolder / tarFile)
 * which hands one tar.Header per entry to Go's archive/tar Writer (go1.11.4, .travis.yml:12;
 * call sites tar.go:159,242,253,280,294,298) and wraps it in compress/gzip (tar.go:156).
 *
 * The arithmetic lives in the Go standard library, which is NOT under /root/reference and for which
 * no toolchain exists in this image. This file restates the published go1.11 archive/tar writer
 * algorithm (Writer.WriteHeader, Header.allowedFormats, splitUSTARPath, writeUSTARHeader,
 * writePAXHeader, writeRawFile, formatPAXRecord, formatter.formatOctal/formatString, block.SetFormat,
 * Writer.Close). PARITY UNPINNED against Go itself: upstream holds no golden bytes for this path
 * (SURVEY.md 8c). It is cross-checked against two independent implementations of the format
 * (GNU tar 1.35 and CPython tarfile as decoders, CPython's USTAR encoder on the fields the two
 * writers agree on) in tests/test_oracle_*.py.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use
 * anything under oracle/. The product (devspace_b200/, libdsgpack.so) never links or calls it.
 */
#ifndef ORACLE_GOTAR_H
#define ORACLE_GOTAR_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* One tar.Header as tarFile/tarFolder hand it to tw.WriteHeader (tar.go:253,294). */
typedef struct gotar_entry {
  const char *name;    size_t name_len;   /* hdr.Name = fileInformation.Name (tar.go:243,284) */
  const char *uname;   size_t uname_len;  /* tar.FileInfoHeader: passwd lookup of stat uid      */
  const char *gname;   size_t gname_len;
  int64_t size;                            /* stat.Size() for regular files, 0 for directories   */
  int64_t mtime_sec;                       /* stat.ModTime(): seconds (floor) + nanoseconds      */
  uint32_t mtime_nsec;
  int64_t mode;                            /* perm | setuid 04000 | setgid 02000 | sticky 01000,
                                              or RemoteMode when indexed (tar.go:246-250,287-291) */
  int64_t uid, gid;                        /* stat uid/gid or RemoteUID/RemoteGID                 */
  uint8_t typeflag;                        /* '0' regular file, '5' (empty) directory; docker's
                                              build-context dialect also '1' hard link, '2' symlink */
  const uint8_t *data;                     /* size bytes of payload (io.Copy, tar.go:298)        */
  const char *linkname; size_t linkname_len; /* hdr.Linkname for '1' / '2' (docker pkg/archive
                                              addTarFile: os.Readlink / the first name of the inode) */
} gotar_entry;

enum {
  GOTAR_OK = 0,
  GOTAR_E_HEADER = -1,   /* archive/tar: header field cannot be encoded (ErrHeader / ErrFieldTooLong) */
  GOTAR_E_TYPE = -2,     /* typeflag outside what tar.go can produce */
  GOTAR_E_NOMEM = -3
};

/* Bytes the tar.Writer emits for one entry's header(s): 512, or 512 + pad512(records) + 512 for PAX. */
int gotar_header_bytes(const gotar_entry *e, size_t *out);

/* Total stream size: sum(header bytes + pad512(size)) + 1024 trailer (Writer.Close). */
int gotar_stream_size(const gotar_entry *es, size_t n, size_t *out);

/* Emit the whole tar stream (headers, payloads, padding, trailer) into out[cap]. */
int gotar_write(const gotar_entry *es, size_t n, uint8_t *out, size_t cap, size_t *written);

/* path.Clean + path.Join(dir, "PaxHeaders.0", file) helpers, exported for unit tests. */
size_t gotar_path_clean(const char *s, size_t n, char *out /* cap >= n+2 */);

/* Go compress/gzip framing (header 1f 8b 08 00 00000000 00 ff, CRC32-IEEE, ISIZE) around a raw
 * deflate body produced by zlib at `level` (proxy for Go's compress/flate level 6, see gogzip.c). */
size_t gogzip_bound(size_t n);
int gogzip_pack(const uint8_t *in, size_t n, int level, uint8_t *out, size_t cap, size_t *written);
uint32_t gogzip_crc32(const uint8_t *in, size_t n);


/* One archive (tar stream + gzip at `level`) per shard [lo[s], hi[s]) of the header list, on a pool of `nthreads`
 * pthreads: the all-host-cores number of bench.py --impl reference. Per-shard sizes and worker seconds come back. */
int gooracle_pack_shards_mt(const gotar_entry *es, const size_t *lo, const size_t *hi, size_t nshards, int level,
                            int nthreads, size_t *tar_bytes, size_t *gz_bytes, double *seconds);

#ifdef __cplusplus
}
#endif
#endif
