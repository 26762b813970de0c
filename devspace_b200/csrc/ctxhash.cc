// ctxhash.cc -- host side of DevSpace's build-context hash, DirectoryExcludes in
// /root/reference/pkg/util/hash/hash.go:43-163 (called from pkg/devspace/image/build.go:213): the walk, the
// exclude decisions (through a callback: the docker/pkg/fileutils pattern matcher stays with the caller), reading
// the files into one pinned arena, and the SHA-256 over the short "path;crc" list. The per-file CRC32
// (hashFileCRC32, hash.go:165-197) is computed on the GPU by dsg_file_crc32 -- there is no CPU checksum in here.
#include <dirent.h>
#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <set>
#include <string>
#include <vector>

#include "../../include/dsgpack.h"

namespace {

// ---- SHA-256 (FIPS 180-4), as crypto/sha256 computes it -------------------------------------------------------
struct Sha256 {
  uint32_t h[8] = {0x6a09e667u, 0xbb67ae85u, 0x3c6ef372u, 0xa54ff53au, 0x510e527fu, 0x9b05688cu, 0x1f83d9abu, 0x5be0cd19u};
  uint8_t buf[64];
  size_t fill = 0;
  uint64_t total = 0;

  static uint32_t rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

  void block(const uint8_t *p) {
    static const uint32_t K[64] = {
        0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u,
        0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u,
        0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau,
        0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u,
        0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u,
        0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u,
        0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u,
        0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};
    uint32_t w[64];
    for (int i = 0; i < 16; i++)
      w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 | (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
    for (int i = 16; i < 64; i++) {
      const uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3);
      const uint32_t s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10);
      w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
    for (int i = 0; i < 64; i++) {
      const uint32_t t1 = hh + (rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25)) + ((e & f) ^ (~e & g)) + K[i] + w[i];
      const uint32_t t2 = (rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
      hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
  }

  void write(const void *data, size_t n) {
    const uint8_t *p = static_cast<const uint8_t *>(data);
    total += n;
    while (n) {
      const size_t take = std::min(n, sizeof buf - fill);
      memcpy(buf + fill, p, take);
      fill += take; p += take; n -= take;
      if (fill == sizeof buf) {
        block(buf);
        fill = 0;
      }
    }
  }

  void hex(char out[65]) {
    const uint64_t bits = total * 8;
    const uint8_t one = 0x80, zero = 0;
    write(&one, 1);
    while (fill != 56) write(&zero, 1);
    uint8_t len[8];
    for (int i = 0; i < 8; i++) len[i] = (uint8_t)(bits >> (56 - 8 * i));
    write(len, 8);
    for (int i = 0; i < 8; i++) snprintf(out + 8 * i, 9, "%08x", h[i]);
    out[64] = 0;
  }
};

// ---- the walk (filepath.Walk: Lstat, names of a directory in byte order, pre-order) -----------------------------
struct Rec {
  std::string path;  // filePath as Go sees it
  bool is_dir;
  int file = -1;     // index into the entry array (files that could be opened)
};

struct Walk {
  std::string src;
  dsg_exclude_fn exclude;
  void *user;
  std::vector<Rec> recs;
  std::vector<dsg_entry> entries;
  std::vector<std::string> file_paths;
  std::set<std::string> seen;
  uint64_t arena_bytes = 0;
  std::string err;
};

std::string clean_path(const std::string &p) {  // filepath.Clean, enough of it for filepath.Join(srcPath, ".")
  if (p.empty()) return ".";
  const bool rooted = p[0] == '/';
  std::vector<std::string> parts;
  size_t i = 0;
  while (i <= p.size()) {
    size_t j = p.find('/', i);
    if (j == std::string::npos) j = p.size();
    const std::string s = p.substr(i, j - i);
    if (s == "..") {
      if (!parts.empty() && parts.back() != "..") parts.pop_back();
      else if (!rooted) parts.push_back("..");
    } else if (!s.empty() && s != ".") {
      parts.push_back(s);
    }
    i = j + 1;
  }
  std::string out = rooted ? "/" : "";
  for (size_t k = 0; k < parts.size(); k++) out += (k ? "/" : "") + parts[k];
  return out.empty() ? "." : out;
}

int walk_rec(Walk &w, const std::string &path, const std::string &rel, const struct stat &st) {
  const bool is_dir = S_ISDIR(st.st_mode);
  const std::string rel_file = rel.empty() ? "." : "./" + rel;  // hash.go:87-89
  int verdict = 0;
  if (rel_file != ".") {
    verdict = w.exclude ? w.exclude(rel_file.c_str(), is_dir ? 1 : 0, w.user) : 0;  // hash.go:99-104
    if (verdict == 1) return 0;                  // a file: return nil; a directory: filepath.SkipDir (hash.go:113-134)
    if (verdict == 2 && !is_dir) return 0;
  }
  if (verdict == 0 && w.seen.insert(rel_file).second) {  // hash.go:136-139
    Rec r{path, is_dir, -1};
    if (!is_dir) {
      // hashFileCRC32 opens (and so follows links); what cannot be opened or read is left out (hash.go:147-149)
      const int fd = open(path.c_str(), O_RDONLY | O_NONBLOCK);
      struct stat fs;
      if (fd >= 0 && fstat(fd, &fs) == 0 && S_ISREG(fs.st_mode)) {
        dsg_entry e;
        memset(&e, 0, sizeof e);
        e.typeflag = '0';
        e.size = (int64_t)fs.st_size;
        e.data_off = w.arena_bytes;
        w.arena_bytes += ((uint64_t)fs.st_size + 15) & ~15ull;
        r.file = (int)w.entries.size();
        w.entries.push_back(e);
        w.file_paths.push_back(path);
      }
      if (fd >= 0) close(fd);
      if (r.file >= 0) w.recs.push_back(r);
    } else {
      w.recs.push_back(r);                       // "Path is enough" (hash.go:141-143)
    }
  }
  if (!is_dir) return 0;
  DIR *d = opendir(path.c_str());
  if (!d) {
    w.err = "Hash: Can't stat file " + w.src + " to hash: " + strerror(errno);
    return DSG_E_ARG;
  }
  std::vector<std::string> names;
  while (struct dirent *de = readdir(d)) {
    if (!strcmp(de->d_name, ".") || !strcmp(de->d_name, "..")) continue;
    names.push_back(de->d_name);
  }
  closedir(d);
  std::sort(names.begin(), names.end());
  for (const std::string &nm : names) {
    const std::string child = path == "/" ? "/" + nm : path + "/" + nm;
    struct stat cst;
    if (lstat(child.c_str(), &cst) != 0) {
      w.err = "Hash: Can't stat file " + w.src + " to hash: " + strerror(errno);
      return DSG_E_ARG;
    }
    const int rc = walk_rec(w, child, rel.empty() ? nm : rel + "/" + nm, cst);
    if (rc) return rc;
  }
  return 0;
}

}  // namespace

extern "C" int dsg_directory_hash(dsg_ctx *ctx, const char *src_path, dsg_exclude_fn exclude, void *user,
                                  char out_hex[65], char *err, size_t err_len) {
  auto fail = [&](int rc, const std::string &msg) {
    if (err && err_len) snprintf(err, err_len, "%s", msg.c_str());
    return rc;
  };
  if (!ctx || !src_path || !out_hex) return DSG_E_ARG;
  struct stat st;
  if (lstat(src_path, &st) != 0) return fail(DSG_E_ARG, std::string("lstat ") + src_path + ": " + strerror(errno));
  if (!S_ISDIR(st.st_mode)) return fail(DSG_E_ARG, std::string("Path ") + src_path + " is not a directory");
  Walk w;
  w.src = src_path;
  w.exclude = exclude;
  w.user = user;
  const std::string root = clean_path(src_path);
  struct stat rst;
  if (lstat(root.c_str(), &rst) != 0) return fail(DSG_E_ARG, "lstat " + root + ": " + strerror(errno));
  int rc = walk_rec(w, root, "", rst);
  if (rc) return fail(rc, "Error hashing " + w.src + ": " + w.err);

  // read every file into one pinned arena; a file that shrank is hashed as far as it reads (io.Copy reads to EOF)
  std::vector<uint32_t> crcs(w.entries.size() + 1, 0);
  if (!w.entries.empty()) {
    uint8_t *arena = (uint8_t *)dsg_pinned_alloc(ctx, (size_t)w.arena_bytes + 16);
    if (!arena) return fail(DSG_E_NOMEM, "pinned arena allocation failed");
    for (size_t i = 0; i < w.entries.size(); i++) {
      dsg_entry &e = w.entries[i];
      uint64_t got = 0;
      const int fd = open(w.file_paths[i].c_str(), O_RDONLY);
      if (fd >= 0) {
        while (got < (uint64_t)e.size) {
          const ssize_t n = read(fd, arena + e.data_off + got, (size_t)((uint64_t)e.size - got));
          if (n <= 0) break;
          got += (uint64_t)n;
        }
        close(fd);
      }
      e.size = (int64_t)got;
    }
    rc = dsg_file_crc32(ctx, w.entries.data(), w.entries.size(), arena, (size_t)w.arena_bytes, crcs.data());
    dsg_pinned_free(ctx, arena);
    if (rc) return fail(rc, dsg_last_error(ctx));
  }

  Sha256 sha;
  char hex8[16];
  for (const Rec &r : w.recs) {
    sha.write(r.path.data(), r.path.size());
    if (!r.is_dir) {
      snprintf(hex8, sizeof hex8, ";%08x", crcs[(size_t)r.file]);  // hex.EncodeToString(hash.Sum(nil)): big endian
      sha.write(hex8, 9);
    }
  }
  sha.hex(out_hex);  // fmt.Sprintf("%x", hash.Sum(nil))
  return DSG_OK;
}
