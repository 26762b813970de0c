// tar_format.cuh -- one tar.Header -> header block(s), as go1.11 archive/tar's Writer encodes it.
//
// Follows what /root/reference/pkg/devspace/sync/tar.go:253,294 (tw.WriteHeader) reaches in the Go
// standard library (go1.11.4): Writer.WriteHeader -> Header.allowedFormats -> writeUSTARHeader /
// writePAXHeader -> templateV7Plus -> formatter.formatString / formatOctal -> block.SetFormat.
// All functions are __host__ __device__: the device kernels (tar_kernels.cuh) call them one thread
// per entry; the host-side layout helpers of the C ABI (dsg_entry_header_bytes, dsg_tar_size,
// dsg_shard_bounds) call the classification half on the CPU (integer work on the manifest only).
#pragma once

#include "dsg_common.cuh"

namespace dsg {

constexpr int kNameSize = 100;    // format.go nameSize
constexpr int kPrefixSize = 155;  // format.go prefixSize

enum : uint32_t {
  kNeedPax = 1u,
  kPaxPath = 2u,
  kPaxUname = 4u,
  kPaxGname = 8u,
  kPaxUid = 16u,
  kPaxGid = 32u,
  kPaxSize = 64u,
  kPaxMtime = 128u,
  kPaxLinkpath = 256u,
};

// Link entries (typeflag '1' hard link, '2' symbolic link; docker's build-context dialect, not tar.go) carry no
// payload: their data_off field holds the link target instead, as a strtab reference.
DSG_HD bool is_link(const dsg_entry &e) { return e.typeflag == '1' || e.typeflag == '2'; }
DSG_HD uint32_t link_off(const dsg_entry &e) { return (uint32_t)e.data_off; }
DSG_HD uint32_t link_len(const dsg_entry &e) { return (uint32_t)(e.data_off >> 32); }

struct HdrPlan {
  int64_t mtime;        // ModTime.Round(time.Second).Unix()
  uint32_t hdr_bytes;   // 512, or 512 + pad512(rec_bytes) + 512
  uint32_t rec_bytes;   // PAX records length (0 for USTAR)
  uint32_t flags;       // kNeedPax | kPax*
  uint32_t prefix_len;  // >0: splitUSTARPath succeeded, name = [prefix]/[suffix]
};

// strfunc.go isASCII: every byte < 0x80 and != NUL
DSG_HD bool is_ascii(const char *s, uint32_t n) {
  for (uint32_t i = 0; i < n; i++) {
    unsigned char c = (unsigned char)s[i];
    if (c >= 0x80 || c == 0) return false;
  }
  return true;
}

DSG_HD bool has_nul(const char *s, uint32_t n) {
  for (uint32_t i = 0; i < n; i++)
    if (s[i] == 0) return true;
  return false;
}

// strfunc.go fitsInOctal
DSG_HD bool fits_in_octal(int n, int64_t x) {
  unsigned oct_bits = (unsigned)(n - 1) * 3;
  return x >= 0 && (n >= 22 || x < (int64_t(1) << oct_bits));
}

DSG_HD int dec_digits_u64(uint64_t v) {
  int d = 1;
  while (v >= 10) {
    v /= 10;
    d++;
  }
  return d;
}

// strconv.FormatInt(v, 10) length
DSG_HD int dec_len_i64(int64_t v) {
  if (v < 0) return 1 + dec_digits_u64(uint64_t(0) - uint64_t(v));
  return dec_digits_u64(uint64_t(v));
}

// writes strconv.FormatInt(v, 10) to out, returns length
DSG_HD int dec_write_i64(int64_t v, char *out) {
  int n = dec_len_i64(v);
  uint64_t u = v < 0 ? uint64_t(0) - uint64_t(v) : uint64_t(v);
  int i = n;
  do {
    out[--i] = char('0' + (u % 10));
    u /= 10;
  } while (u);
  if (v < 0) out[0] = '-';
  return n;
}

// formatPAXRecord length: "%d %s=%s\n", the length prefix counts itself
DSG_HD uint32_t pax_record_len(uint32_t klen, uint32_t vlen, uint32_t *size_field) {
  uint32_t size = klen + vlen + 3;
  size += (uint32_t)dec_digits_u64(size);
  uint32_t total = (uint32_t)dec_digits_u64(size) + klen + vlen + 3;
  if (total != size) size = total;  // "final adjustment if adding size field increased the record size"
  if (size_field) *size_field = size;
  return (uint32_t)dec_digits_u64(size) + klen + vlen + 3;
}

// common.go splitUSTARPath: returns prefix length (index of the splitting '/'), 0 = cannot split
DSG_HD uint32_t split_ustar_path(const char *name, uint32_t len, bool ascii) {
  uint32_t length = len;
  if (length <= (uint32_t)kNameSize || !ascii) return 0;
  if (length > (uint32_t)kPrefixSize + 1)
    length = kPrefixSize + 1;
  else if (name[length - 1] == '/')
    length--;
  int i = -1;
  for (int k = (int)length - 1; k >= 0; k--)
    if (name[k] == '/') {
      i = k;
      break;
    }
  int nlen = (int)len - i - 1;
  int plen = i;
  if (i <= 0 || nlen > kNameSize || nlen == 0 || plen > kPrefixSize) return 0;
  return (uint32_t)plen;
}

// Header.allowedFormats, restricted to the fields tar.go and docker's pkg/archive fill (no Dev* / Xattrs /
// PAXRecords; AccessTime and ChangeTime are zero: cleared by WriteHeader when Format is unset, by docker's
// FileInfoHeader otherwise). Format = FormatPAX (docker) decides nothing here: "PAX implies USTAR allowed too"
// (allowedFormats), so USTAR is still written whenever it can hold the header.
// Returns a dsg_status.
DSG_HD int classify_entry(const dsg_entry &e, const char *strtab, uint64_t strtab_len, HdrPlan &p) {
  p.mtime = 0;
  p.hdr_bytes = 0;
  p.rec_bytes = 0;
  p.flags = 0;
  p.prefix_len = 0;
  if (e.typeflag != '0' && e.typeflag != '5' && !is_link(e)) return DSG_E_TYPEFLAG;
  if (is_link(e) && (uint64_t)link_off(e) + link_len(e) > strtab_len) return DSG_E_ARG;
  if ((uint64_t)e.name_off + e.name_len > strtab_len || (uint64_t)e.uname_off + e.uname_len > strtab_len ||
      (uint64_t)e.gname_off + e.gname_len > strtab_len || e.mtime_nsec > 999999999u)
    return DSG_E_ARG;
  const char *name = strtab + e.name_off;
  const char *uname = strtab + e.uname_off;
  const char *gname = strtab + e.gname_off;
  // WriteHeader: ModTime.Round(time.Second), halves round up
  p.mtime = e.mtime_sec + (e.mtime_nsec >= 500000000u ? 1 : 0);

  bool ustar_ok = true, pax_ok = true;
  uint32_t flags = 0;
  bool name_ascii = is_ascii(name, e.name_len);
  if (!name_ascii || e.name_len > (uint32_t)kNameSize) {  // verifyString(h.Name, paxPath)
    uint32_t pl = split_ustar_path(name, e.name_len, name_ascii);
    if (!pl) ustar_ok = false;
    p.prefix_len = pl;
    flags |= kPaxPath;
  }
  const char *lname = is_link(e) ? strtab + link_off(e) : name;
  const uint32_t llen = is_link(e) ? link_len(e) : 0;
  if (!is_ascii(lname, llen) || llen > (uint32_t)kNameSize) {  // verifyString(h.Linkname, paxLinkpath): no split
    ustar_ok = false;
    flags |= kPaxLinkpath;
  }
  if (!is_ascii(uname, e.uname_len) || e.uname_len > 32) {  // verifyString(h.Uname, paxUname)
    ustar_ok = false;
    flags |= kPaxUname;
  }
  if (!is_ascii(gname, e.gname_len) || e.gname_len > 32) {
    ustar_ok = false;
    flags |= kPaxGname;
  }
  // verifyNumeric(h.Mode, paxNone): neither USTAR nor PAX can hold it; GNU base-256 is unreachable
  // from tar.go (mode <= 07777) and is reported as ErrHeader.
  if (!fits_in_octal(8, e.mode)) return DSG_E_HEADER;
  if (!fits_in_octal(8, e.uid)) {
    ustar_ok = false;
    flags |= kPaxUid;
  }
  if (!fits_in_octal(8, e.gid)) {
    ustar_ok = false;
    flags |= kPaxGid;
  }
  if (!fits_in_octal(12, e.size)) {
    ustar_ok = false;
    flags |= kPaxSize;
  }
  if (!fits_in_octal(12, p.mtime)) {  // verifyTime(h.ModTime, paxMtime); nanoseconds are zero after rounding
    ustar_ok = false;
    flags |= kPaxMtime;
  }
  if (e.typeflag == '0' && e.name_len > 0 && name[e.name_len - 1] == '/') return DSG_E_HEADER;
  if (e.size < 0) return DSG_E_HEADER;
  if ((e.typeflag == '5' || is_link(e)) && e.size != 0) return DSG_E_HEADER;
  // validPAXRecord: no NUL inside path / uname / gname values
  if (((flags & kPaxPath) && has_nul(name, e.name_len)) || ((flags & kPaxUname) && has_nul(uname, e.uname_len)) ||
      ((flags & kPaxGname) && has_nul(gname, e.gname_len)) || ((flags & kPaxLinkpath) && has_nul(lname, llen)))
    pax_ok = false;
  if (ustar_ok) {
    p.flags = 0;
    p.hdr_bytes = kBlock;
    return DSG_OK;
  }
  if (!pax_ok) return DSG_E_HEADER;
  p.prefix_len = 0;
  p.flags = flags | kNeedPax;
  uint32_t rec = 0;  // sorted keys: gid gname linkpath mtime path size uid uname
  if (flags & kPaxGid) rec += pax_record_len(3, (uint32_t)dec_len_i64(e.gid), nullptr);
  if (flags & kPaxGname) rec += pax_record_len(5, e.gname_len, nullptr);
  if (flags & kPaxLinkpath) rec += pax_record_len(8, llen, nullptr);
  if (flags & kPaxMtime) rec += pax_record_len(5, (uint32_t)dec_len_i64(p.mtime), nullptr);
  if (flags & kPaxPath) rec += pax_record_len(4, e.name_len, nullptr);
  if (flags & kPaxSize) rec += pax_record_len(4, (uint32_t)dec_len_i64(e.size), nullptr);
  if (flags & kPaxUid) rec += pax_record_len(3, (uint32_t)dec_len_i64(e.uid), nullptr);
  if (flags & kPaxUname) rec += pax_record_len(5, e.uname_len, nullptr);
  p.rec_bytes = rec;
  p.hdr_bytes = kBlock + (uint32_t)pad512(rec) + kBlock;
  return DSG_OK;
}

// Payload bytes (padded to 512) following the header(s): writeRawHeader gives header-only types none.
DSG_HD uint64_t entry_payload_padded(const dsg_entry &e) { return e.typeflag == '0' ? pad512((uint64_t)e.size) : 0; }

// ---- field formatters (blk points at a zero-filled 512-byte block) --------------------------------

// formatter.formatString for a string that fits or is truncated; slen_total is len(s) in Go,
// `s` yields the first min(slen_total, blen) bytes.
DSG_HD void format_string(uint8_t *b, uint32_t blen, const char *s, uint32_t slen_total) {
  uint32_t n = slen_total < blen ? slen_total : blen;
  for (uint32_t i = 0; i < n; i++) b[i] = (uint8_t)s[i];
  // slen < blen: b[slen] = 0 (already zero)
  if (slen_total > blen && b[blen - 1] == '/') {  // over-long path ending in '/': cut at the trimmed length
    uint32_t t = blen;
    while (t > 0 && b[t - 1] == '/') t--;
    b[t] = 0;
  }
}

// formatString(b, toASCII(s)): non-ASCII bytes and NULs are dropped before copying
DSG_HD void format_string_to_ascii(uint8_t *b, uint32_t blen, const char *s, uint32_t slen) {
  uint32_t w = 0, total = 0;
  for (uint32_t i = 0; i < slen; i++) {
    unsigned char c = (unsigned char)s[i];
    if (c < 0x80 && c != 0) {
      if (w < blen) b[w++] = c;
      total++;
    }
  }
  if (total > blen && b[blen - 1] == '/') {
    uint32_t t = blen;
    while (t > 0 && b[t - 1] == '/') t--;
    b[t] = 0;
  }
}

// formatter.formatOctal: zero-padded octal, width blen-1, NUL terminated; out-of-range writes zero
DSG_HD void format_octal(uint8_t *b, int blen, int64_t x) {
  if (!fits_in_octal(blen, x)) x = 0;
  uint64_t u = (uint64_t)x;
  // fits => at most blen-1 digits (blen >= 22 never occurs for tar fields)
  for (int i = blen - 2; i >= 0; i--) {
    b[i] = (uint8_t)('0' + (u & 7));
    u >>= 3;
  }
  b[blen - 1] = 0;
}

// block.SetFormat(FormatUSTAR | FormatPAX): magic "ustar\0", version "00", then the checksum
// (unsigned byte sum with the field read as spaces) as 6 octal digits, NUL, space.
DSG_HD void set_format_and_checksum(uint8_t *blk) {
  blk[257] = 'u';
  blk[258] = 's';
  blk[259] = 't';
  blk[260] = 'a';
  blk[261] = 'r';
  blk[262] = 0;
  blk[263] = '0';
  blk[264] = '0';
  uint32_t sum = 8 * 32;
  for (int i = 0; i < 148; i++) sum += blk[i];
  for (int i = 156; i < kBlock; i++) sum += blk[i];
  format_octal(blk + 148, 7, (int64_t)sum);
  blk[155] = ' ';
}

// Writer.templateV7Plus + SetFormat for the entry's main header.
// USTAR path: strings verbatim (they were verified ASCII and fitting), optional prefix split.
// PAX path:   strings through toASCII, numeric overflow silently written as zero.
DSG_HD void format_main_header(uint8_t *blk, const dsg_entry &e, const char *strtab, const HdrPlan &p) {
  const char *name = strtab + e.name_off;
  const char *uname = strtab + e.uname_off;
  const char *gname = strtab + e.gname_off;
  const bool pax = (p.flags & kNeedPax) != 0;
  blk[156] = e.typeflag;
  if (is_link(e)) {
    if (pax) format_string_to_ascii(blk + 157, 100, strtab + link_off(e), link_len(e));
    else format_string(blk + 157, 100, strtab + link_off(e), link_len(e));
  }
  if (pax) {
    format_string_to_ascii(blk, 100, name, e.name_len);
    format_string_to_ascii(blk + 265, 32, uname, e.uname_len);
    format_string_to_ascii(blk + 297, 32, gname, e.gname_len);
  } else {
    if (p.prefix_len) {
      uint32_t pl = p.prefix_len;
      format_string(blk, 100, name + pl + 1, e.name_len - pl - 1);
      format_string(blk + 345, kPrefixSize, name, pl);
    } else {
      format_string(blk, 100, name, e.name_len);
    }
    format_string(blk + 265, 32, uname, e.uname_len);
    format_string(blk + 297, 32, gname, e.gname_len);
  }
  format_octal(blk + 100, 8, e.mode);
  format_octal(blk + 108, 8, e.uid);
  format_octal(blk + 116, 8, e.gid);
  format_octal(blk + 124, 12, e.size);
  format_octal(blk + 136, 12, p.mtime);
  format_octal(blk + 329, 8, 0);  // Devmajor
  format_octal(blk + 337, 8, 0);  // Devminor
  set_format_and_checksum(blk);
}

// True when path.Clean(name) == name can be relied on for building the PaxHeaders.0 name without
// running Clean (names produced by tar.go always are: path.Join at tar.go:186,261). A directory may end in one
// '/' (docker's canonicalTarName): path.Split then yields an empty file part and Join drops the slash again.
DSG_HD bool name_is_clean_for_pax(const char *s, uint32_t n, bool dir_slash) {
  if (n == 0) return false;
  if (n > 1 && s[n - 1] == '/') {
    if (!dir_slash) return false;
    n--;
    if (n > 1 && s[n - 1] == '/') return false;
  }
  uint32_t i = 0;
  while (i < n) {
    if (s[i] == '/') {
      if (i + 1 < n && s[i + 1] == '/') return false;
      i++;
      continue;
    }
    uint32_t j = i;
    while (j < n && s[j] != '/') j++;
    uint32_t cl = j - i;
    if ((cl == 1 && s[i] == '.') || (cl == 2 && s[i] == '.' && s[i + 1] == '.')) return false;
    i = j;
  }
  return true;
}

// writePAXHeader + writeRawFile: the 'x' extended header block, its records (zero padded to 512),
// then the main header: hdr_bytes in all. Only bytes whose offset r inside that area satisfies
// win_lo <= r < win_hi are stored (dst[r]); the layout kernel uses the window to clip an area that
// straddles a pipeline-segment boundary. Every byte of the window is written.
DSG_HD int format_pax_headers(uint8_t *dst, uint32_t win_lo, uint32_t win_hi, const dsg_entry &e,
                              const char *strtab, const HdrPlan &p) {
  const char *name = strtab + e.name_off;
  if (!name_is_clean_for_pax(name, e.name_len, e.typeflag == '5')) return DSG_E_UNCLEAN_NAME;
  if (win_hi > p.hdr_bytes) win_hi = p.hdr_bytes;
  uint8_t tmp[kBlock];
  auto emit_block = [&](uint32_t at) {
    for (uint32_t i = 0; i < (uint32_t)kBlock; i++) {
      uint32_t r = at + i;
      if (r >= win_lo && r < win_hi) dst[r] = tmp[i];
    }
  };
  // --- extended header block: name = dir + "PaxHeaders.0/" + file, toASCII, cut to 100, TrimRight "/"
  for (int i = 0; i < kBlock; i++) tmp[i] = 0;
  uint32_t slash = e.name_len;
  while (slash > 0 && name[slash - 1] != '/') slash--;
  const char mid[13] = {'P', 'a', 'x', 'H', 'e', 'a', 'd', 'e', 'r', 's', '.', '0', '/'};
  uint32_t w = 0;
  const uint32_t total = e.name_len + 13;
  for (uint32_t i = 0; i < total && w < 100; i++) {
    unsigned char c = i < slash ? (unsigned char)name[i]
                                : (i < slash + 13 ? (unsigned char)mid[i - slash] : (unsigned char)name[i - 13]);
    if (c < 0x80 && c != 0) tmp[w++] = c;
  }
  while (w > 0 && tmp[w - 1] == '/') tmp[--w] = 0;
  tmp[156] = 'x';
  format_octal(tmp + 100, 8, 0);
  format_octal(tmp + 108, 8, 0);
  format_octal(tmp + 116, 8, 0);
  format_octal(tmp + 124, 12, (int64_t)p.rec_bytes);
  format_octal(tmp + 136, 12, 0);
  set_format_and_checksum(tmp);
  emit_block(0);
  // --- records, sorted by key, then zero padding up to the next block
  uint32_t r = kBlock;
  auto out1 = [&](uint8_t c) {
    if (r >= win_lo && r < win_hi) dst[r] = c;
    r++;
  };
  char num[24];
  auto put = [&](const char *k, uint32_t klen, const char *v, uint32_t vlen) {
    uint32_t size_field;
    pax_record_len(klen, vlen, &size_field);
    char sz[24];
    int sl = dec_write_i64((int64_t)size_field, sz);
    for (int i = 0; i < sl; i++) out1((uint8_t)sz[i]);
    out1(' ');
    for (uint32_t i = 0; i < klen; i++) out1((uint8_t)k[i]);
    out1('=');
    for (uint32_t i = 0; i < vlen; i++) out1((uint8_t)v[i]);
    out1('\n');
  };
  uint32_t nl;
  if (p.flags & kPaxGid) { nl = (uint32_t)dec_write_i64(e.gid, num); put("gid", 3, num, nl); }
  if (p.flags & kPaxGname) put("gname", 5, strtab + e.gname_off, e.gname_len);
  if (p.flags & kPaxLinkpath) put("linkpath", 8, strtab + link_off(e), link_len(e));
  if (p.flags & kPaxMtime) { nl = (uint32_t)dec_write_i64(p.mtime, num); put("mtime", 5, num, nl); }
  if (p.flags & kPaxPath) put("path", 4, name, e.name_len);
  if (p.flags & kPaxSize) { nl = (uint32_t)dec_write_i64(e.size, num); put("size", 4, num, nl); }
  if (p.flags & kPaxUid) { nl = (uint32_t)dec_write_i64(e.uid, num); put("uid", 3, num, nl); }
  if (p.flags & kPaxUname) put("uname", 5, strtab + e.uname_off, e.uname_len);
  const uint32_t main_at = kBlock + (uint32_t)pad512(p.rec_bytes);
  while (r < main_at) out1(0);
  // --- main header
  for (int i = 0; i < kBlock; i++) tmp[i] = 0;
  format_main_header(tmp, e, strtab, p);
  emit_block(main_at);
  return DSG_OK;
}

}  // namespace dsg
