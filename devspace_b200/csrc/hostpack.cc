// hostpack.cc -- C++ host mirror of the reference's Go call site above the C ABI.
//
// Mirrors /root/reference/pkg/devspace/sync/tar.go:146-324 function by function (same names, same argument
// meaning, same error behaviour) for environments without a Go toolchain:
//   writeTar (146-182)        -> dsg_write_tar      temp file, walk, pack, writtenFiles; on a walk error: log,
//                                                   remove the temp file, sleep, retry
//   recursiveTar (185-229)    -> recursive_tar      de-dup, exclude / uploadExclude match, os.Stat (follows
//                                                   symlinks), vanished files skipped silently
//   tarFolder (231-267)       -> tar_folder         sorted ReadDir; only EMPTY, non-root directories get a header
//   tarFile (269-304)         -> tar_file           header from stat, Name = relative path, Remote* override when
//                                                   indexed, payload read into the pinned arena
//   createFileInformationFromStat (306-324), roundMtime (util.go:87-89)
// The tar.Writer / gzip.Writer work itself happens on the GPU through dsg_pack (dsgpack.cu). This file is host
// code only (walk, stat, read): what north_star keeps on the host.
#include <dirent.h>
#include <errno.h>
#include <fcntl.h>
#include <grp.h>
#include <pwd.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <map>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/dsgpack.h"

namespace {

struct Walked {
  std::vector<dsg_entry> entries;
  std::string strtab;
  std::vector<std::string> paths;        // absolute path per entry (files only; "" for directories)
  std::vector<dsg_file_info> written;    // writtenFiles, insertion order
  std::vector<std::string> written_names;
  std::map<std::string, size_t> written_idx;
  std::map<uint32_t, uint32_t> uname_off, gname_off;  // uid/gid -> strtab offset (tar.FileInfoHeader's userMap/groupMap)
  std::map<uint32_t, uint16_t> uname_len, gname_len;
  uint64_t arena_bytes = 0;
  std::string error;
};

// path.Join for two elements: skip empties, join with '/', Clean (inputs here are clean already)
std::string join(const std::string &a, const std::string &b) {
  if (a.empty()) return b;
  if (b.empty()) return a;
  if (a.back() == '/') return a + b;
  return a + "/" + b;
}

int64_t round_mtime(const struct timespec &ts) { return (int64_t)ts.tv_sec + (ts.tv_nsec >= 500000000L ? 1 : 0); }

// fileIndex.fileMap (file_index.go:8-11) is a Go map: the index handed in as an array is hashed once per walk (a
// 100 k-file tree with a 100 k-entry index was 10^10 string compares as a linear scan). The first entry of a name wins,
// as in the scan it replaces.
struct FileMapIndex {
  const dsg_sync_config *cfg = nullptr;
  std::unordered_map<std::string, const dsg_file_info *> by_name;
};
thread_local FileMapIndex g_index;

void index_file_map(const dsg_sync_config *cfg) {
  g_index.cfg = cfg;
  g_index.by_name.clear();
  g_index.by_name.reserve(cfg->file_map_len * 2 + 1);
  for (size_t i = 0; i < cfg->file_map_len; i++)
    if (cfg->file_map[i].name) g_index.by_name.emplace(cfg->file_map[i].name, &cfg->file_map[i]);
}

const dsg_file_info *lookup(const dsg_sync_config *cfg, const std::string &name) {
  if (g_index.cfg != cfg) index_file_map(cfg);
  auto it = g_index.by_name.find(name);
  return it == g_index.by_name.end() ? nullptr : it->second;
}

uint32_t intern(Walked &w, const std::string &s) {
  uint32_t off = (uint32_t)w.strtab.size();
  w.strtab += s;
  return off;
}

void names_of(Walked &w, uint32_t uid, uint32_t gid, dsg_entry &e) {
  if (!w.uname_off.count(uid)) {
    struct passwd *pw = getpwuid(uid);
    std::string n = pw ? pw->pw_name : "";
    w.uname_off[uid] = intern(w, n);
    w.uname_len[uid] = (uint16_t)n.size();
  }
  if (!w.gname_off.count(gid)) {
    struct group *gr = getgrgid(gid);
    std::string n = gr ? gr->gr_name : "";
    w.gname_off[gid] = intern(w, n);
    w.gname_len[gid] = (uint16_t)n.size();
  }
  e.uname_off = w.uname_off[uid];
  e.uname_len = w.uname_len[uid];
  e.gname_off = w.gname_off[gid];
  e.gname_len = w.gname_len[gid];
}

// createFileInformationFromStat (tar.go:306-324)
dsg_file_info info_from_stat(const dsg_sync_config *cfg, const std::string &rel, const struct stat &st) {
  dsg_file_info fi;
  memset(&fi, 0, sizeof fi);
  fi.size = st.st_size;
  fi.mtime = round_mtime(st.st_mtim);
  fi.is_directory = S_ISDIR(st.st_mode) ? 1 : 0;
  if (const dsg_file_info *known = lookup(cfg, rel)) {
    fi.remote_mode = known->remote_mode;
    fi.remote_uid = known->remote_uid;
    fi.remote_gid = known->remote_gid;
  }
  return fi;
}

void add_header(Walked &w, const dsg_sync_config *cfg, const std::string &rel, const struct stat &st, char typeflag,
                const dsg_file_info &fi, const std::string &abs) {
  dsg_entry e;
  memset(&e, 0, sizeof e);
  e.name_off = intern(w, rel);
  e.name_len = (uint32_t)rel.size();
  // tar.FileInfoHeader: Mode = perm | setuid | setgid | sticky (no file-type bits)
  int64_t mode = st.st_mode & 0777;
  if (st.st_mode & S_ISUID) mode |= 04000;
  if (st.st_mode & S_ISGID) mode |= 02000;
  if (st.st_mode & S_ISVTX) mode |= 01000;
  e.mode = mode;
  e.uid = st.st_uid;
  e.gid = st.st_gid;
  e.mtime_sec = st.st_mtim.tv_sec;
  e.mtime_nsec = (uint32_t)st.st_mtim.tv_nsec;
  e.typeflag = (uint8_t)typeflag;
  e.size = typeflag == '0' ? st.st_size : 0;
  names_of(w, st.st_uid, st.st_gid, e);
  if (lookup(cfg, rel)) {  // tar.go:246-250,287-291: Uname/Gname are NOT overridden
    e.mode = fi.remote_mode;
    e.uid = fi.remote_uid;
    e.gid = fi.remote_gid;
  }
  if (typeflag == '0' && st.st_size > 0) {
    w.arena_bytes = (w.arena_bytes + 15) & ~uint64_t(15);
    e.data_off = w.arena_bytes;
    w.arena_bytes += (uint64_t)st.st_size;
  }
  w.entries.push_back(e);
  w.paths.push_back(typeflag == '0' ? abs : std::string());
  w.written_idx[rel] = w.written.size();
  w.written_names.push_back(rel);
  w.written.push_back(fi);
}

int recursive_tar(Walked &w, const dsg_sync_config *cfg, const std::string &rel);

// tarFolder (tar.go:231-267)
int tar_folder(Walked &w, const dsg_sync_config *cfg, const std::string &rel, const struct stat &st,
               const dsg_file_info &fi) {
  const std::string dirpath = join(cfg->watch_path, rel);
  DIR *d = opendir(dirpath.c_str());
  if (!d) return 0;  // "Couldn't read dir": logged and skipped
  std::vector<std::string> names;
  while (struct dirent *de = readdir(d)) {
    if (!strcmp(de->d_name, ".") || !strcmp(de->d_name, "..")) continue;
    names.push_back(de->d_name);
  }
  closedir(d);
  std::sort(names.begin(), names.end());  // ioutil.ReadDir sorts by file name
  if (names.empty() && !rel.empty()) add_header(w, cfg, rel, st, '5', fi, dirpath);
  for (const auto &n : names) {
    int rc = recursive_tar(w, cfg, join(rel, n));
    if (rc) return rc;
  }
  return 0;
}

// tarFile (tar.go:269-304)
int tar_file(Walked &w, const dsg_sync_config *cfg, const std::string &rel, const struct stat &st,
             const dsg_file_info &fi) {
  const std::string abs = join(cfg->watch_path, rel);
  if (!S_ISREG(st.st_mode)) {  // checked before the open: opening a FIFO would block
    w.error = "archive/tar: unsupported file type: " + abs;
    return DSG_E_TYPEFLAG;
  }
  int fd = open(abs.c_str(), O_RDONLY);
  if (fd < 0) {  // os.Open failed: errors.Trace(err) -> writeTar retries
    w.error = "open " + abs + ": " + strerror(errno);
    return DSG_E_ARG;
  }
  close(fd);
  add_header(w, cfg, rel, st, '0', fi, abs);
  return 0;
}

// recursiveTar (tar.go:185-229)
int recursive_tar(Walked &w, const dsg_sync_config *cfg, const std::string &rel) {
  if (w.written_idx.count(rel)) return 0;
  bool excluded = false;
  if (cfg->ignore && cfg->ignore(rel.c_str(), cfg->ignore_user)) excluded = true;
  if (cfg->upload_ignore && cfg->upload_ignore(rel.c_str(), cfg->upload_ignore_user)) excluded = true;
  if (excluded) return 0;
  const std::string abs = join(cfg->watch_path, rel);
  struct stat st;
  if (stat(abs.c_str(), &st) != 0) return 0;  // "We skip files that are suddenly not there anymore"
  dsg_file_info fi = info_from_stat(cfg, rel, st);
  if (S_ISDIR(st.st_mode)) return tar_folder(w, cfg, rel, st, fi);
  return tar_file(w, cfg, rel, st, fi);
}

int walk_all(Walked &w, const dsg_file_info *files, size_t nfiles, const dsg_sync_config *cfg) {
  for (size_t i = 0; i < nfiles; i++) {
    const std::string rel = files[i].name ? files[i].name : "";
    if (!w.written_idx.count(rel)) {
      int rc = recursive_tar(w, cfg, rel);
      if (rc) return rc;
    }
  }
  for (size_t i = 0; i < w.written.size(); i++) w.written[i].name = w.written_names[i].c_str();
  return 0;
}

int read_all(const Walked &w, uint8_t *arena, std::string *err) {
  for (size_t i = 0; i < w.entries.size(); i++) {
    const dsg_entry &e = w.entries[i];
    if (e.typeflag != '0' || e.size == 0) continue;
    int fd = open(w.paths[i].c_str(), O_RDONLY);
    if (fd < 0) {
      *err = "open " + w.paths[i] + ": " + strerror(errno);
      return DSG_E_ARG;
    }
    uint64_t got = 0;
    while (got < (uint64_t)e.size) {
      ssize_t n = read(fd, arena + e.data_off + got, (size_t)((uint64_t)e.size - got));
      if (n <= 0) break;
      got += (uint64_t)n;
    }
    // one more byte: archive/tar fails with ErrWriteTooLong when the file grew after stat
    char extra;
    const bool grew = got == (uint64_t)e.size && read(fd, &extra, 1) == 1;
    close(fd);
    if (got != (uint64_t)e.size || grew) {
      *err = "archive/tar: " + w.paths[i] + (grew ? ": write too long" : ": missed writing bytes") + " (file changed size)";
      return DSG_E_HEADER;
    }
  }
  return 0;
}

}  // namespace

struct dsg_written {
  Walked w;
};

extern "C" {

int dsg_walk(const dsg_file_info *files, size_t nfiles, const dsg_sync_config *cfg, dsg_written **out) {
  if (!cfg || !cfg->watch_path || !out || (nfiles && !files)) return DSG_E_ARG;
  dsg_written *h = new dsg_written();
  index_file_map(cfg);  // once per walk: the caller may have changed the index since the last one
  int rc = walk_all(h->w, files, nfiles, cfg);
  g_index.cfg = nullptr;
  g_index.by_name.clear();
  if (rc) {
    delete h;
    return rc;
  }
  *out = h;
  return DSG_OK;
}

size_t dsg_written_count(const dsg_written *h) { return h ? h->w.written.size() : 0; }
const dsg_file_info *dsg_written_get(const dsg_written *h, size_t i) {
  return (h && i < h->w.written.size()) ? &h->w.written[i] : nullptr;
}
const dsg_entry *dsg_written_entries(const dsg_written *h, size_t *n, const char **strtab, size_t *strtab_len,
                                     uint64_t *arena_bytes) {
  if (!h) return nullptr;
  if (n) *n = h->w.entries.size();
  if (strtab) *strtab = h->w.strtab.data();
  if (strtab_len) *strtab_len = h->w.strtab.size();
  if (arena_bytes) *arena_bytes = h->w.arena_bytes;
  return h->w.entries.data();
}
void dsg_written_free(dsg_written *h) { delete h; }

int dsg_write_tar(dsg_ctx *ctx, const dsg_file_info *files, size_t nfiles, const dsg_sync_config *cfg, char *out_path,
                  size_t out_path_cap, dsg_written **written) {
  if (!ctx || !cfg || !out_path || out_path_cap < 32 || !written) return DSG_E_ARG;
  const unsigned sleep_ms = cfg->retry_sleep_ms ? cfg->retry_sleep_ms : 4000;  // tar.go:174: 4 seconds
  for (unsigned attempt = 0;; attempt++) {
    char tmpl[] = "/tmp/devspace-gpack-XXXXXX";  // ioutil.TempFile("", "")
    int fd = mkstemp(tmpl);
    if (fd < 0) return DSG_E_ARG;
    dsg_written *h = nullptr;
    std::string err;
    int rc = dsg_walk(files, nfiles, cfg, &h);
    uint8_t *arena = nullptr;
    if (rc == DSG_OK) {
      Walked &w = h->w;
      const char *strtab = w.strtab.empty() ? "" : w.strtab.data();
      arena = (uint8_t *)dsg_pinned_alloc(ctx, (size_t)w.arena_bytes + 16);
      rc = arena ? read_all(w, arena, &err) : DSG_E_NOMEM;
      // gzip.NewWriter(file) + tar.NewWriter: the stream goes to the temp file segment by segment while the GPU works
      // on the next segments (dsg_pack_fd)
      uint64_t nbytes = 0;
      if (rc == DSG_OK) {
        rc = dsg_pack_fd(ctx, w.entries.data(), w.entries.size(), strtab, w.strtab.size(), arena,
                         (size_t)w.arena_bytes, fd, 1, &nbytes);
        if (rc != DSG_OK) err = dsg_last_error(ctx);
      }
    }
    if (arena) dsg_pinned_free(ctx, arena);
    close(fd);
    if (rc == DSG_OK) {
      snprintf(out_path, out_path_cap, "%s", tmpl);
      *written = h;
      return DSG_OK;
    }
    // "[Upstream] Tar failed: %s. Will retry in 4 seconds..." then writeTar(files, config) again (tar.go:170-177)
    if (cfg->log) {
      std::string msg = "[Upstream] Tar failed: " + (err.empty() ? std::string(dsg_strerror(rc)) : err) +
                        ". Will retry in " + std::to_string(sleep_ms / 1000.0).substr(0, 4) + " seconds...\n";
      cfg->log(msg.c_str(), cfg->log_user);
    }
    if (h) dsg_written_free(h);
    unlink(tmpl);
    if (cfg->max_retries && attempt + 1 >= cfg->max_retries) return rc;  // the reference retries forever (0)
    std::this_thread::sleep_for(std::chrono::milliseconds(sleep_ms));
  }
}

}  // extern "C"
