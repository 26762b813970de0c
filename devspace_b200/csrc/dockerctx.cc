// dockerctx.cc -- host mirror of the docker build-context tar DevSpace sends to the daemon
// (/root/reference/pkg/devspace/builder/docker/docker.go:94-97: archive.TarWithOptions(contextDir,
// {ExcludePatterns, ChownOpts{0,0}}), github.com/docker/docker/pkg/archive pinned in go.mod:22). The walk, the
// header construction (docker's FileInfoHeader / addTarFile) and the file reads happen here; the archive bytes are
// produced on the GPU by dsg_pack_tar_only (same header + layout kernels as writeTar, link entries added).
// Restated from the published algorithm, PARITY UNPINNED -- see oracle/dockerctx.py for what is and is not covered.
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
#include <map>
#include <set>
#include <string>
#include <vector>

#include "../../include/dsgpack.h"

namespace {

struct Ctx {
  dsg_exclude_fn exclude;
  void *user;
  std::vector<dsg_entry> entries;
  std::vector<std::string> paths;  // per entry: file to read ("" when there is no payload)
  std::string strtab;
  std::set<std::string> seen;
  std::map<std::pair<uint64_t, uint64_t>, std::string> inodes;  // (dev, ino) -> first tar name (SeenFiles)
  std::map<uint32_t, std::pair<uint32_t, uint16_t>> unames, gnames;
  uint64_t arena_bytes = 0;
  std::string err;
};

uint32_t intern(Ctx &c, const std::string &s) {
  const uint32_t off = (uint32_t)c.strtab.size();
  c.strtab += s;
  return off;
}

void owner_names(Ctx &c, uint32_t uid, uint32_t gid, dsg_entry &e) {  // tar.FileInfoHeader: os/user lookups, cached
  if (!c.unames.count(uid)) {
    struct passwd *pw = getpwuid(uid);
    const std::string n = pw ? pw->pw_name : "";
    c.unames[uid] = {intern(c, n), (uint16_t)n.size()};
  }
  if (!c.gnames.count(gid)) {
    struct group *gr = getgrgid(gid);
    const std::string n = gr ? gr->gr_name : "";
    c.gnames[gid] = {intern(c, n), (uint16_t)n.size()};
  }
  e.uname_off = c.unames[uid].first;
  e.uname_len = c.unames[uid].second;
  e.gname_off = c.gnames[gid].first;
  e.gname_len = c.gnames[gid].second;
}

// addTarFile(path, name)
void add_tar_file(Ctx &c, const std::string &path, const std::string &rel, const struct stat &st) {
  const bool is_dir = S_ISDIR(st.st_mode), is_link = S_ISLNK(st.st_mode);
  if (!S_ISREG(st.st_mode) && !is_dir && !is_link) return;  // devices, FIFOs, sockets: not covered, left out
  dsg_entry e;
  memset(&e, 0, sizeof e);
  int64_t mode = st.st_mode & 0777;
  if (st.st_mode & S_ISUID) mode |= 04000;
  if (st.st_mode & S_ISGID) mode |= 02000;
  if (st.st_mode & S_ISVTX) mode |= 01000;
  mode |= is_dir ? 040000 : is_link ? 0120000 : 0100000;  // fillGo18FileTypeBits
  e.mode = mode;
  e.mtime_sec = st.st_mtim.tv_sec;  // ModTime.Truncate(time.Second)
  e.mtime_nsec = 0;
  owner_names(c, st.st_uid, st.st_gid, e);
  e.uid = 0;  // ChownOpts{0, 0}: the names stay
  e.gid = 0;
  const std::string name = is_dir ? rel + "/" : rel;  // canonicalTarName
  std::string link, file;
  if (is_dir) {
    e.typeflag = '5';
  } else if (is_link) {
    std::vector<char> buf(4096);
    ssize_t n;
    while ((n = readlink(path.c_str(), buf.data(), buf.size())) == (ssize_t)buf.size()) buf.resize(buf.size() * 2);
    if (n < 0) return;
    link.assign(buf.data(), (size_t)n);
    e.typeflag = '2';
  } else {
    e.typeflag = '0';
    e.size = st.st_size;
  }
  if (!is_dir && st.st_nlink > 1) {  // hasHardlinks: the second name of an inode becomes a link to the first
    const auto key = std::make_pair((uint64_t)st.st_dev, (uint64_t)st.st_ino);
    auto it = c.inodes.find(key);
    if (it != c.inodes.end()) {
      e.typeflag = '1';
      link = it->second;
      e.size = 0;
    } else {
      c.inodes[key] = name;
    }
  }
  if (e.typeflag == '0' && e.size > 0) {
    const int fd = open(path.c_str(), O_RDONLY);  // "Can't add file %s to tar": logged and left out by docker
    if (fd < 0) return;
    close(fd);
    c.arena_bytes = (c.arena_bytes + 15) & ~uint64_t(15);
    e.data_off = c.arena_bytes;
    c.arena_bytes += (uint64_t)e.size;
    file = path;
  } else if (e.typeflag == '1' || e.typeflag == '2') {
    e.data_off = (uint64_t)intern(c, link) | ((uint64_t)link.size() << 32);
  }
  e.name_off = intern(c, name);
  e.name_len = (uint32_t)name.size();
  c.entries.push_back(e);
  c.paths.push_back(file);
}

void walk(Ctx &c, const std::string &path, const std::string &rel, const struct stat &st) {
  const bool is_dir = S_ISDIR(st.st_mode);
  if (rel != ".") {
    const int verdict = c.exclude ? c.exclude(rel.c_str(), is_dir ? 1 : 0, c.user) : 0;
    if (verdict == 1 || (verdict == 2 && !is_dir)) return;
    if (verdict == 0 && c.seen.insert(rel).second) add_tar_file(c, path, rel, st);
  }
  if (!is_dir) return;
  DIR *d = opendir(path.c_str());
  if (!d) return;
  std::vector<std::string> names;
  while (struct dirent *de = readdir(d)) {
    if (!strcmp(de->d_name, ".") || !strcmp(de->d_name, "..")) continue;
    names.push_back(de->d_name);
  }
  closedir(d);
  std::sort(names.begin(), names.end());
  for (const std::string &nm : names) {
    const std::string child = path + "/" + nm;
    struct stat cst;
    if (lstat(child.c_str(), &cst) != 0) continue;
    walk(c, child, rel == "." ? nm : rel + "/" + nm, cst);
  }
}

}  // namespace

extern "C" int dsg_docker_context_tar(dsg_ctx *ctx, const char *src_path, dsg_exclude_fn exclude, void *user,
                                      const char *out_path, uint64_t *tar_bytes, uint64_t *n_entries, char *err,
                                      size_t err_len) {
  auto fail = [&](int rc, const std::string &msg) {
    if (err && err_len) snprintf(err, err_len, "%s", msg.c_str());
    return rc;
  };
  if (!ctx || !src_path || !out_path) return DSG_E_ARG;
  std::string root = src_path;
  while (root.size() > 1 && root.back() == '/') root.pop_back();
  struct stat st;
  if (lstat(root.c_str(), &st) != 0) return fail(DSG_E_ARG, "lstat " + root + ": " + strerror(errno));
  if (!S_ISDIR(st.st_mode)) return fail(DSG_E_ARG, root + " is not a directory");
  Ctx c;
  c.exclude = exclude;
  c.user = user;
  walk(c, root, ".", st);

  const char *strtab = c.strtab.empty() ? "" : c.strtab.data();
  uint64_t T = 0;
  int rc = dsg_tar_size(c.entries.data(), c.entries.size(), strtab, c.strtab.size(), 0, &T);
  if (rc) return fail(rc, std::string("archive/tar: ") + dsg_strerror(rc));
  uint8_t *arena = (uint8_t *)dsg_pinned_alloc(ctx, (size_t)c.arena_bytes + 16);
  if (!arena) return fail(DSG_E_NOMEM, "pinned allocation failed");
  for (size_t i = 0; i < c.entries.size() && rc == DSG_OK; i++) {
    if (c.paths[i].empty()) continue;
    const dsg_entry &e = c.entries[i];
    const int fd = open(c.paths[i].c_str(), O_RDONLY);
    uint64_t got = 0;
    if (fd >= 0) {
      while (got < (uint64_t)e.size) {
        const ssize_t n = read(fd, arena + e.data_off + got, (size_t)((uint64_t)e.size - got));
        if (n <= 0) break;
        got += (uint64_t)n;
      }
      close(fd);
    }
    if (got != (uint64_t)e.size) rc = fail(DSG_E_HEADER, "archive/tar: " + c.paths[i] + ": file changed size");
  }
  uint64_t n = 0;
  if (rc == DSG_OK) {
    const int fd = open(out_path, O_WRONLY | O_CREAT | O_TRUNC, 0600);
    if (fd < 0) {
      rc = fail(DSG_E_ARG, std::string("open ") + out_path + ": " + strerror(errno));
    } else {
      rc = dsg_pack_fd(ctx, c.entries.data(), c.entries.size(), strtab, c.strtab.size(), arena, (size_t)c.arena_bytes,
                       fd, 0, &n);  // uncompressed: what TarWithOptions pipes to the daemon
      if (rc) fail(rc, dsg_last_error(ctx));
      close(fd);
      if (rc == DSG_OK && n != T) rc = fail(DSG_E_CUDA, "internal: stream length differs from the plan");
    }
  }
  dsg_pinned_free(ctx, arena);
  if (rc == DSG_OK) {
    if (tar_bytes) *tar_bytes = n;
    if (n_entries) *n_entries = c.entries.size();
  }
  return rc;
}
