// multipack.cc -- ONE call, N GPUs, one ordered archive (SURVEY.md 8b "dsg_ctx_create(const int* devices, int ndev)",
// 8e "Multi-GPU").
//
// The reference packs an archive with one writeTar call from one goroutine (/root/reference/pkg/devspace/sync/
// upstream.go:342; kaniko's whole tree: util.go:68-73). dsg_multi_* keeps that shape for a box with several GPUs:
// the header list is cut into contiguous ranges balanced by tar-stream bytes (dsg_shard_bounds, files are never
// split), one context per device packs its range on its own host thread (every shard a complete gzip member, only
// the last one carries the tar trailer), and the calling thread writes the members IN SHARD ORDER to one file
// descriptor (or one buffer) while the later shards are still being packed. Either the members go out back to back
// (a multi-member .tar.gz: GNU gzip/tar and CPython read it as one stream) or, with DSG_MULTI_ONE_MEMBER, stitched
// into a single member on the fly (inner headers, trailers and 03 00 terminators dropped, checksums folded with the
// crc32_combine identity) for decoders that stop after the first member (busybox).
// No data-path collective exists: the GPUs never exchange payload. Host code + the C ABI of dsgpack.cu only.
#include <stdio.h>
#include <string.h>
#include <errno.h>
#include <unistd.h>

#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/dsgpack.h"

struct dsg_multi {
  std::vector<int> devices;
  std::vector<dsg_ctx *> ctx;
  std::vector<uint8_t *> out;      // pinned member buffer per shard (grown on demand)
  std::vector<size_t> out_cap;
  std::vector<dsg_stats> stats;
  dsg_opts opts{};
  std::string err;
  std::mutex mu;                   // one caller at a time
};

namespace {

int mfail(dsg_multi *m, int code, const std::string &text) {
  if (m) m->err = text;
  return code;
}

struct Sink {
  int fd = -1;
  uint8_t *buf = nullptr;
  size_t cap = 0, pos = 0;
  bool overflow = false, io_error = false;
  void put(const void *p, size_t n) {
    if (fd >= 0) {
      const uint8_t *q = static_cast<const uint8_t *>(p);
      size_t left = n;
      while (left && !io_error) {
        const ssize_t k = write(fd, q, left);
        if (k <= 0) {
          io_error = true;
          break;
        }
        q += k;
        left -= (size_t)k;
      }
    } else if (pos + n <= cap) {
      memcpy(buf + pos, p, n);
    } else {
      overflow = true;
    }
    pos += n;
  }
};

int multi_pack(dsg_multi *m, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
               const uint8_t *arena, size_t arena_len, uint32_t mode, Sink &sink) {
  const int S = (int)m->ctx.size();
  std::vector<uint64_t> bounds((size_t)S + 1, 0);
  int rc = dsg_shard_bounds(entries, n, strtab, strtab_len, S, bounds.data());
  if (rc) return mfail(m, rc, std::string("dsg_shard_bounds: ") + dsg_strerror(rc));
  // member buffers (host integer work on the manifest: the size of every shard's stream is known before any GPU runs)
  std::vector<uint64_t> tar_len((size_t)S, 0);
  for (int s = 0; s < S; s++) {
    const uint32_t flags = (m->opts.flags & ~DSG_F_NO_TRAILER) | (s + 1 < S ? DSG_F_NO_TRAILER : 0u);
    rc = dsg_tar_size(entries + bounds[s], (size_t)(bounds[s + 1] - bounds[s]), strtab, strtab_len, flags, &tar_len[s]);
    if (rc) return mfail(m, rc, std::string("dsg_tar_size: ") + dsg_strerror(rc));
    const size_t want = (size_t)dsg_gzip_bound(tar_len[s]) + 64;
    if (want > m->out_cap[s]) {
      if (m->out[s]) dsg_pinned_free(m->ctx[s], m->out[s]);
      m->out[s] = static_cast<uint8_t *>(dsg_pinned_alloc(m->ctx[s], want));
      m->out_cap[s] = m->out[s] ? want : 0;
      if (!m->out[s]) return mfail(m, DSG_E_NOMEM, "pinned member buffer of shard " + std::to_string(s));
    }
  }
  // one host thread per shard; the caller's thread drains them in shard order
  std::vector<int> status((size_t)S, 1);  // 1 = running, <= 0 = the shard's return code
  std::vector<size_t> produced((size_t)S, 0);
  std::vector<std::string> errs((size_t)S);
  std::mutex mu;
  std::condition_variable cv;
  std::vector<std::thread> th;
  for (int s = 0; s < S; s++) {
    th.emplace_back([&, s]() {
      dsg_opts o = m->opts;
      o.struct_size = sizeof(dsg_opts);
      o.flags = (m->opts.flags & ~DSG_F_NO_TRAILER) | (s + 1 < S ? DSG_F_NO_TRAILER : 0u);
      int r = dsg_ctx_set_opts(m->ctx[s], &o);
      size_t got = 0;
      if (r == DSG_OK)
        r = dsg_pack(m->ctx[s], entries + bounds[s], (size_t)(bounds[s + 1] - bounds[s]), strtab, strtab_len, arena,
                     arena_len, m->out[s], m->out_cap[s], &got);
      if (r == DSG_OK) dsg_get_stats(m->ctx[s], &m->stats[s]);
      std::lock_guard<std::mutex> g(mu);
      if (r != DSG_OK) errs[s] = dsg_last_error(m->ctx[s]);
      produced[s] = got;
      status[s] = r;
      cv.notify_all();
    });
  }
  static const uint8_t hdr[10] = {0x1f, 0x8b, 0x08, 0, 0, 0, 0, 0, 0, 0xff};
  const bool one = (mode & DSG_MULTI_ONE_MEMBER) != 0;
  uint32_t crc = 0;
  uint64_t total = 0;
  int first_err = DSG_OK;
  std::string first_text;
  if (one) sink.put(hdr, 10);
  for (int s = 0; s < S; s++) {
    {
      std::unique_lock<std::mutex> g(mu);
      cv.wait(g, [&] { return status[s] <= 0; });
    }
    if (status[s] != DSG_OK) {
      if (first_err == DSG_OK) {
        first_err = status[s];
        first_text = "shard " + std::to_string(s) + " (device " + std::to_string(m->devices[s]) + "): " + errs[s];
      }
      continue;
    }
    if (first_err != DSG_OK) continue;
    const uint8_t *mem = m->out[s];
    const size_t len = produced[s];
    if (!one) {
      sink.put(mem, len);
      continue;
    }
    // body without the member's own header, trailer and (except for the last shard) the 03 00 terminator
    const size_t body = len - 18 - (s + 1 < S ? 2 : 0);
    sink.put(mem + 10, body);
    uint32_t c = 0;
    for (int b = 0; b < 4; b++) c |= (uint32_t)mem[len - 8 + b] << (8 * b);
    crc = total ? dsg_crc32_combine(crc, c, tar_len[s]) : c;
    total += tar_len[s];
  }
  for (auto &t : th) t.join();
  if (first_err != DSG_OK) return mfail(m, first_err, first_text);
  if (one) {
    uint8_t tr[8];
    for (int b = 0; b < 4; b++) tr[b] = (uint8_t)(crc >> (8 * b));
    for (int b = 0; b < 4; b++) tr[4 + b] = (uint8_t)((uint32_t)total >> (8 * b));
    sink.put(tr, 8);
  }
  if (sink.io_error) return mfail(m, DSG_E_ARG, std::string("write: ") + strerror(errno));
  if (sink.overflow) return mfail(m, DSG_E_CAPACITY, "output needs " + std::to_string(sink.pos) + " bytes");
  return DSG_OK;
}

}  // namespace

extern "C" {

int dsg_multi_create(const int *devices, int ndev, const dsg_opts *opts, dsg_multi **out) {
  if (!out || !devices || ndev <= 0 || ndev > 64) return DSG_E_ARG;
  *out = nullptr;
  dsg_multi *m = new dsg_multi();
  m->opts.struct_size = sizeof(dsg_opts);
  m->opts.level = DSG_LEVEL_DEFAULT;
  if (opts) memcpy(&m->opts, opts, opts->struct_size < sizeof(dsg_opts) ? opts->struct_size : sizeof(dsg_opts));
  m->opts.struct_size = sizeof(dsg_opts);
  for (int i = 0; i < ndev; i++) {
    dsg_ctx *c = nullptr;
    const int rc = dsg_ctx_create(devices[i], &m->opts, &c);
    if (rc != DSG_OK) {
      dsg_multi_destroy(m);
      return rc;  // no CPU fallback: a missing device is an error
    }
    m->devices.push_back(devices[i]);
    m->ctx.push_back(c);
  }
  m->out.assign((size_t)ndev, nullptr);
  m->out_cap.assign((size_t)ndev, 0);
  m->stats.assign((size_t)ndev, dsg_stats{});
  *out = m;
  return DSG_OK;
}

void dsg_multi_destroy(dsg_multi *m) {
  if (!m) return;
  for (size_t s = 0; s < m->ctx.size(); s++) {
    if (s < m->out.size() && m->out[s]) dsg_pinned_free(m->ctx[s], m->out[s]);
    dsg_ctx_destroy(m->ctx[s]);
  }
  delete m;
}

const char *dsg_multi_last_error(const dsg_multi *m) { return m ? m->err.c_str() : "no context"; }

int dsg_multi_shards(const dsg_multi *m) { return m ? (int)m->ctx.size() : 0; }

int dsg_multi_get_stats(const dsg_multi *m, int shard, dsg_stats *out) {
  if (!m || !out || shard < 0 || shard >= (int)m->stats.size()) return DSG_E_ARG;
  *out = m->stats[(size_t)shard];
  return DSG_OK;
}

int dsg_multi_pack_fd(dsg_multi *m, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                      const uint8_t *arena, size_t arena_len, int fd, uint32_t mode, uint64_t *out_len) {
  if (!m || fd < 0 || (n && (!entries || !strtab)) || (arena_len && !arena)) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(m->mu);
  m->err.clear();
  Sink sink;
  sink.fd = fd;
  const int rc = multi_pack(m, entries, n, strtab, strtab_len, arena, arena_len, mode, sink);
  if (out_len) *out_len = sink.pos;
  return rc;
}

int dsg_multi_pack(dsg_multi *m, const dsg_entry *entries, size_t n, const char *strtab, size_t strtab_len,
                   const uint8_t *arena, size_t arena_len, uint32_t mode, uint8_t *out, size_t out_cap, size_t *out_len) {
  if (!m || !out || !out_len || (n && (!entries || !strtab)) || (arena_len && !arena)) return DSG_E_ARG;
  std::lock_guard<std::mutex> g(m->mu);
  m->err.clear();
  Sink sink;
  sink.buf = out;
  sink.cap = out_cap;
  const int rc = multi_pack(m, entries, n, strtab, strtab_len, arena, arena_len, mode, sink);
  *out_len = sink.pos;
  return rc;
}

}  // extern "C"
