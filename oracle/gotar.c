/*
 * oracle/gotar.c -- TEST INFRASTRUCTURE ONLY (CPU oracle); see gotar.h for scope and pinning status
 * ("parity unpinned" against Go itself; cross-checked against GNU tar and CPython tarfile).
 *
 * Restates, for the headers pkg/devspace/sync/tar.go:242-255,280-298 can produce, the go1.11
 * archive/tar writer. Function names below name the Go stdlib function each one follows.
 */
#include "gotar.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define BLOCK 512
#define NAME_SIZE 100   /* archive/tar format.go: nameSize   */
#define PREFIX_SIZE 155 /* archive/tar format.go: prefixSize */

/* ---- strfunc.go ---------------------------------------------------------------------------- */

/* isASCII: every byte < 0x80 and != NUL */
static int is_ascii(const char *s, size_t n) {
  for (size_t i = 0; i < n; i++) {
    unsigned char c = (unsigned char)s[i];
    if (c >= 0x80 || c == 0) return 0;
  }
  return 1;
}

static int has_nul(const char *s, size_t n) { return memchr(s, 0, n) != NULL; }

/* toASCII: drop every non-ASCII rune and NUL. Ranging over runes drops all bytes >= 0x80. */
static size_t to_ascii(const char *s, size_t n, char *out) {
  size_t w = 0;
  for (size_t i = 0; i < n; i++) {
    unsigned char c = (unsigned char)s[i];
    if (c < 0x80 && c != 0) out[w++] = (char)c;
  }
  return w;
}

/* fitsInOctal(n, x) */
static int fits_in_octal(int n, int64_t x) {
  unsigned oct_bits = (unsigned)(n - 1) * 3;
  return x >= 0 && (n >= 22 || x < ((int64_t)1 << oct_bits));
}

/* formatter.formatString, including the trailing-slash quirk for over-long paths */
static void format_string(uint8_t *b, size_t blen, const char *s, size_t slen) {
  size_t n = slen < blen ? slen : blen;
  memcpy(b, s, n);
  if (slen < blen) b[slen] = 0;
  if (slen > blen && b[blen - 1] == '/') {
    size_t t = blen;
    while (t > 0 && s[t - 1] == '/') t--;
    b[t] = 0;
  }
}

/* formatter.formatOctal; returns 0 if the value did not fit (Go then writes zero). */
static int format_octal(uint8_t *b, int blen, int64_t x) {
  int ok = 1;
  if (!fits_in_octal(blen, x)) {
    x = 0;
    ok = 0;
  }
  char digits[32];
  int nd = 0;
  uint64_t u = (uint64_t)x;
  do {
    digits[nd++] = (char)('0' + (u & 7));
    u >>= 3;
  } while (u);
  char s[64];
  int w = 0;
  int pad = blen - nd - 1;
  for (int i = 0; i < pad; i++) s[w++] = '0';
  while (nd) s[w++] = digits[--nd];
  format_string(b, (size_t)blen, s, (size_t)w);
  return ok;
}

/* ---- common.go: splitUSTARPath -------------------------------------------------------------- */
static int split_ustar_path(const char *name, size_t len, size_t *plen_out) {
  size_t length = len;
  if (length <= NAME_SIZE || !is_ascii(name, len)) return 0;
  if (length > PREFIX_SIZE + 1)
    length = PREFIX_SIZE + 1;
  else if (name[length - 1] == '/')
    length--;
  long i = -1;
  for (long k = (long)length - 1; k >= 0; k--)
    if (name[k] == '/') {
      i = k;
      break;
    }
  long nlen = (long)len - i - 1;
  long plen = i;
  if (i <= 0 || nlen > NAME_SIZE || nlen == 0 || plen > PREFIX_SIZE) return 0;
  *plen_out = (size_t)plen;
  return 1;
}

/* ---- path.Clean (Go "path" package), used by path.Join in writePAXHeader -------------------- */
size_t gotar_path_clean(const char *p, size_t n, char *out) {
  if (n == 0) {
    out[0] = '.';
    return 1;
  }
  int rooted = p[0] == '/';
  size_t w = 0, r = 0, dotdot = 0;
  if (rooted) {
    out[w++] = '/';
    r = 1;
    dotdot = 1;
  }
  while (r < n) {
    if (p[r] == '/') {
      r++;
    } else if (p[r] == '.' && (r + 1 == n || p[r + 1] == '/')) {
      r++;
    } else if (p[r] == '.' && p[r + 1] == '.' && (r + 2 == n || p[r + 2] == '/')) {
      r += 2;
      if (w > dotdot) {
        w--;
        while (w > dotdot && out[w] != '/') w--;
      } else if (!rooted) {
        if (w > 0) out[w++] = '/';
        out[w++] = '.';
        out[w++] = '.';
        dotdot = w;
      }
    } else {
      if ((rooted && w != 1) || (!rooted && w != 0)) out[w++] = '/';
      for (; r < n && p[r] != '/'; r++) out[w++] = p[r];
    }
  }
  if (w == 0) {
    out[0] = '.';
    return 1;
  }
  return w;
}

/* ---- format.go: block.SetFormat for USTAR and PAX (both: magic "ustar\0", version "00") ----- */
static void set_format_ustar(uint8_t *blk) {
  memcpy(blk + 257, "ustar\0", 6);
  memcpy(blk + 263, "00", 2);
  /* ComputeChecksum: unsigned sum with the chksum field read as spaces */
  uint32_t sum = 0;
  for (int i = 0; i < BLOCK; i++) sum += (i >= 148 && i < 156) ? (uint32_t)' ' : blk[i];
  format_octal(blk + 148, 7, (int64_t)sum);
  blk[155] = ' ';
}

/* ---- PAX record helpers ---------------------------------------------------------------------- */
static int dec_digits(uint64_t v) {
  int d = 1;
  while (v >= 10) {
    v /= 10;
    d++;
  }
  return d;
}

/* formatPAXRecord: "%d %s=%s\n" with the length counting itself */
static size_t pax_record(const char *k, const char *v, size_t vlen, char *out) {
  size_t klen = strlen(k);
  size_t size = klen + vlen + 3;
  size += (size_t)dec_digits(size);
  size_t hd = (size_t)dec_digits(size);
  size_t total = hd + 1 + klen + 1 + vlen + 1;
  if (total != size) size = total; /* "final adjustment if adding size field increased the record size" */
  size_t actual = (size_t)dec_digits(size) + 1 + klen + 1 + vlen + 1; /* len(record) as Go builds it */
  if (out) {
    int w = sprintf(out, "%zu %s=", size, k);
    memcpy(out + w, v, vlen);
    out[w + vlen] = '\n';
  }
  return actual;
}

typedef struct {
  int need_pax;     /* USTAR cannot hold it */
  int has_path, has_uname, has_gname, has_uid, has_gid, has_size, has_mtime, has_linkpath;
  char uid_s[24], gid_s[24], size_s[24], mtime_s[24];
  size_t ustar_prefix_len; /* >0 when splitUSTARPath succeeded */
  int64_t mtime_rounded;
} plan_t;

/* Header.allowedFormats restricted to what tar.go fills in (no Linkname, Dev*, Xattrs, PAXRecords,
 * AccessTime/ChangeTime are zeroed by WriteHeader because Format is FormatUnknown). */
static int plan_entry(const gotar_entry *e, plan_t *p) {
  memset(p, 0, sizeof *p);
  const int link = e->typeflag == '1' || e->typeflag == '2';
  if (e->typeflag != '0' && e->typeflag != '5' && !link) return GOTAR_E_TYPE;
  /* WriteHeader: ModTime.Round(time.Second): half rounds up */
  p->mtime_rounded = e->mtime_sec + (e->mtime_nsec >= 500000000u ? 1 : 0);

  int ustar_ok = 1, pax_ok = 1;
  /* verifyString(h.Name, 100, paxPath) */
  if (!is_ascii(e->name, e->name_len) || e->name_len > NAME_SIZE) {
    size_t pl = 0;
    if (!split_ustar_path(e->name, e->name_len, &pl)) ustar_ok = 0;
    else p->ustar_prefix_len = pl;
    p->has_path = 1;
  }
  /* verifyString(h.Linkname, 100, paxLinkpath): only paxPath may be split for USTAR */
  if (link && (!is_ascii(e->linkname, e->linkname_len) || e->linkname_len > NAME_SIZE)) {
    ustar_ok = 0;
    p->has_linkpath = 1;
  }
  /* verifyString(h.Uname / h.Gname, 32, paxUname / paxGname) */
  if (!is_ascii(e->uname, e->uname_len) || e->uname_len > 32) {
    ustar_ok = 0;
    p->has_uname = 1;
  }
  if (!is_ascii(e->gname, e->gname_len) || e->gname_len > 32) {
    ustar_ok = 0;
    p->has_gname = 1;
  }
  /* verifyNumeric(h.Mode, 8, paxNone): neither USTAR nor PAX; GNU base-256 is unreachable from
   * tar.go (mode <= 07777), so it is reported as ErrHeader here. */
  if (!fits_in_octal(8, e->mode)) return GOTAR_E_HEADER;
  if (!fits_in_octal(8, e->uid)) {
    ustar_ok = 0;
    p->has_uid = 1;
    snprintf(p->uid_s, sizeof p->uid_s, "%lld", (long long)e->uid);
  }
  if (!fits_in_octal(8, e->gid)) {
    ustar_ok = 0;
    p->has_gid = 1;
    snprintf(p->gid_s, sizeof p->gid_s, "%lld", (long long)e->gid);
  }
  if (!fits_in_octal(12, e->size)) {
    ustar_ok = 0;
    p->has_size = 1;
    snprintf(p->size_s, sizeof p->size_s, "%lld", (long long)e->size);
  }
  /* verifyTime(h.ModTime, 12, paxMtime): nanoseconds are zero after rounding */
  if (!fits_in_octal(12, p->mtime_rounded)) {
    ustar_ok = 0;
    p->has_mtime = 1;
    snprintf(p->mtime_s, sizeof p->mtime_s, "%lld", (long long)p->mtime_rounded); /* formatPAXTime, nsecs==0 */
  }
  /* header-only / trailing-slash / negative-size checks */
  if (e->typeflag == '0' && e->name_len > 0 && e->name[e->name_len - 1] == '/') return GOTAR_E_HEADER;
  if (e->size < 0) return GOTAR_E_HEADER; /* negative size: ErrHeader for files; tar.go never sets it for dirs */
  if ((e->typeflag == '5' || link) && e->size != 0) return GOTAR_E_HEADER; /* FileInfoHeader leaves Size 0 for dirs and links */
  /* validPAXRecord: NUL not allowed in path/uname/gname values */
  if ((p->has_path && has_nul(e->name, e->name_len)) || (p->has_uname && has_nul(e->uname, e->uname_len)) ||
      (p->has_gname && has_nul(e->gname, e->gname_len)) || (p->has_linkpath && has_nul(e->linkname, e->linkname_len)))
    pax_ok = 0;
  if (ustar_ok) {
    p->need_pax = 0;
    return GOTAR_OK;
  }
  if (!pax_ok) return GOTAR_E_HEADER; /* GNU is excluded as well: hasNUL => mustNotBe(FormatGNU) */
  p->need_pax = 1;
  return GOTAR_OK;
}

/* sorted keys: gid, gname, linkpath, mtime, path, size, uid, uname (sort.Strings in writePAXHeader) */
static size_t pax_records(const gotar_entry *e, const plan_t *p, char *out) {
  size_t w = 0;
#define REC(cond, key, val, vlen)                           \
  if (cond) w += pax_record(key, val, vlen, out ? out + w : NULL)
  REC(p->has_gid, "gid", p->gid_s, strlen(p->gid_s));
  REC(p->has_gname, "gname", e->gname, e->gname_len);
  REC(p->has_linkpath, "linkpath", e->linkname, e->linkname_len);
  REC(p->has_mtime, "mtime", p->mtime_s, strlen(p->mtime_s));
  REC(p->has_path, "path", e->name, e->name_len);
  REC(p->has_size, "size", p->size_s, strlen(p->size_s));
  REC(p->has_uid, "uid", p->uid_s, strlen(p->uid_s));
  REC(p->has_uname, "uname", e->uname, e->uname_len);
#undef REC
  return w;
}

static size_t pad512(size_t n) { return (n + 511) & ~(size_t)511; }

int gotar_header_bytes(const gotar_entry *e, size_t *out) {
  plan_t p;
  int rc = plan_entry(e, &p);
  if (rc) return rc;
  if (!p.need_pax) {
    *out = BLOCK;
    return GOTAR_OK;
  }
  size_t rec = pax_records(e, &p, NULL);
  *out = BLOCK + pad512(rec) + BLOCK;
  return GOTAR_OK;
}

int gotar_stream_size(const gotar_entry *es, size_t n, size_t *out) {
  size_t t = 0;
  for (size_t i = 0; i < n; i++) {
    size_t h;
    int rc = gotar_header_bytes(&es[i], &h);
    if (rc) return rc;
    t += h + (es[i].typeflag == '0' ? pad512((size_t)es[i].size) : 0);
  }
  *out = t + 2 * BLOCK;
  return GOTAR_OK;
}

/* Writer.templateV7Plus; strict=1 is the USTAR path, strict=0 the PAX path (toASCII, errors ignored) */
static void template_v7plus(uint8_t *blk, const gotar_entry *e, const plan_t *p, const char *name, size_t name_len,
                            int to_ascii_strings) {
  memset(blk, 0, BLOCK);
  char *tmp = (char *)malloc(name_len + e->uname_len + e->gname_len + e->linkname_len + 8);
  blk[156] = e->typeflag;
  if (e->typeflag == '1' || e->typeflag == '2') {
    if (to_ascii_strings) {
      size_t n = to_ascii(e->linkname, e->linkname_len, tmp);
      format_string(blk + 157, 100, tmp, n);
    } else {
      format_string(blk + 157, 100, e->linkname, e->linkname_len);
    }
  }
  if (to_ascii_strings) {
    size_t n = to_ascii(name, name_len, tmp);
    format_string(blk + 0, 100, tmp, n);
  } else {
    format_string(blk + 0, 100, name, name_len);
  }
  /* Linkname "" : formatString writes a NUL at [157] (already zero) */
  format_octal(blk + 100, 8, e->mode);
  format_octal(blk + 108, 8, e->uid);
  format_octal(blk + 116, 8, e->gid);
  format_octal(blk + 124, 12, e->size);
  format_octal(blk + 136, 12, p->mtime_rounded);
  if (to_ascii_strings) {
    size_t n = to_ascii(e->uname, e->uname_len, tmp);
    format_string(blk + 265, 32, tmp, n);
    n = to_ascii(e->gname, e->gname_len, tmp);
    format_string(blk + 297, 32, tmp, n);
  } else {
    format_string(blk + 265, 32, e->uname, e->uname_len);
    format_string(blk + 297, 32, e->gname, e->gname_len);
  }
  format_octal(blk + 329, 8, 0); /* Devmajor */
  format_octal(blk + 337, 8, 0); /* Devminor */
  free(tmp);
}

static int write_entry_headers(const gotar_entry *e, uint8_t *out, size_t *hbytes) {
  plan_t p;
  int rc = plan_entry(e, &p);
  if (rc) return rc;
  if (!p.need_pax) {
    /* writeUSTARHeader */
    if (p.ustar_prefix_len) {
      size_t pl = p.ustar_prefix_len;
      template_v7plus(out, e, &p, e->name + pl + 1, e->name_len - pl - 1, 0);
      format_string(out + 345, PREFIX_SIZE, e->name, pl);
    } else {
      template_v7plus(out, e, &p, e->name, e->name_len, 0);
    }
    set_format_ustar(out);
    *hbytes = BLOCK;
    return GOTAR_OK;
  }
  /* writePAXHeader: extended header file first (writeRawFile) */
  size_t rec = pax_records(e, &p, NULL);
  char *recbuf = (char *)malloc(rec + 64);
  pax_records(e, &p, recbuf);
  /* dir, file := path.Split(realName); name = path.Join(dir, "PaxHeaders.0", file) */
  size_t slash = e->name_len;
  while (slash > 0 && e->name[slash - 1] != '/') slash--;
  size_t dir_len = slash, file_len = e->name_len - slash;
  char *joined = (char *)malloc(e->name_len + 32);
  size_t j = 0;
  if (dir_len) {
    memcpy(joined, e->name, dir_len);
    j = dir_len;
    joined[j++] = '/';
  }
  memcpy(joined + j, "PaxHeaders.0", 12);
  j += 12;
  if (file_len) {
    joined[j++] = '/';
    memcpy(joined + j, e->name + slash, file_len);
    j += file_len;
  }
  char *clean = (char *)malloc(j + 4);
  size_t cl = gotar_path_clean(joined, j, clean);
  /* writeRawFile: toASCII, truncate to nameSize, TrimRight "/" */
  char *xa = (char *)malloc(cl + 4);
  size_t xl = to_ascii(clean, cl, xa);
  if (xl > NAME_SIZE) xl = NAME_SIZE;
  while (xl > 0 && xa[xl - 1] == '/') xl--;
  uint8_t *blk = out;
  memset(blk, 0, BLOCK);
  blk[156] = 'x';
  format_string(blk + 0, 100, xa, xl);
  format_octal(blk + 100, 8, 0);
  format_octal(blk + 108, 8, 0);
  format_octal(blk + 116, 8, 0);
  format_octal(blk + 124, 12, (int64_t)rec);
  format_octal(blk + 136, 12, 0);
  set_format_ustar(blk); /* FormatPAX shares magic/version with USTAR */
  size_t recpad = pad512(rec);
  memset(out + BLOCK, 0, recpad);
  memcpy(out + BLOCK, recbuf, rec);
  /* main header: strings through toASCII, numeric overflow silently written as zero */
  uint8_t *mh = out + BLOCK + recpad;
  template_v7plus(mh, e, &p, e->name, e->name_len, 1);
  set_format_ustar(mh);
  *hbytes = BLOCK + recpad + BLOCK;
  free(recbuf);
  free(joined);
  free(clean);
  free(xa);
  return GOTAR_OK;
}

int gotar_write(const gotar_entry *es, size_t n, uint8_t *out, size_t cap, size_t *written) {
  size_t total;
  int rc = gotar_stream_size(es, n, &total);
  if (rc) return rc;
  if (total > cap) return GOTAR_E_NOMEM;
  size_t off = 0;
  for (size_t i = 0; i < n; i++) {
    size_t hb = 0;
    rc = write_entry_headers(&es[i], out + off, &hb);
    if (rc) return rc;
    off += hb;
    if (es[i].typeflag == '0') { /* writeRawHeader: header-only types carry no payload */
      size_t sz = (size_t)es[i].size;
      if (sz) memcpy(out + off, es[i].data, sz);
      size_t padded = pad512(sz);
      memset(out + off + sz, 0, padded - sz); /* Writer.Flush: zero padding */
      off += padded;
    }
  }
  memset(out + off, 0, 2 * BLOCK); /* Writer.Close: two zero blocks, no record padding */
  off += 2 * BLOCK;
  *written = off;
  return GOTAR_OK;
}
