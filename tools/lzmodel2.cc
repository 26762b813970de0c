// tools/lzmodel2.cc -- DESIGN-EXPLORATION TOOL (not built by build(), not linked into the product, not a
// fallback): serial model of the round-2 match finder -- per-hash ROWS of the K most recent positions (two tables:
// a short hash of 4 bytes and a long hash of L bytes) probed at every position, inheritance of (len-1, dist), the
// stateless lazy parse, per-chunk dynamic Huffman sizes -- to predict the compressed size against zlib level 6 and
// the amount of work per position before GPU time is spent.
//   g++ -O2 -std=c++17 -DDSG_EMU -Itests/simt_emu -Idevspace_b200/csrc tools/lzmodel2.cc -lz -o /tmp/lzmodel2
//   /tmp/lzmodel2 file   (parameters through the environment, see P(...) below)
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include <algorithm>
#include <vector>

#include "huffman.cuh"

dim3 threadIdx, blockIdx, blockDim, gridDim;
using namespace dsg;

static int P(const char *name, int def) { const char *e = getenv(name); return e ? atoi(e) : def; }

static const int MAXDIST = 32768 - 272;

static size_t zlib6(const uint8_t *d, size_t n) {
  z_stream zs;
  memset(&zs, 0, sizeof zs);
  deflateInit2(&zs, 6, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY);
  std::vector<uint8_t> out(n + n / 100 + 1024);
  zs.next_in = (Bytef *)d;
  zs.avail_in = n;
  zs.next_out = out.data();
  zs.avail_out = out.size();
  deflate(&zs, Z_FINISH);
  size_t r = zs.total_out;
  deflateEnd(&zs);
  return r;
}

static inline uint64_t load8(const uint8_t *p) { uint64_t v; memcpy(&v, p, 8); return v; }

struct Rows {
  int bits, K, L;
  std::vector<int32_t> e;   // [1<<bits][K] ring per row
  std::vector<uint32_t> cnt;
  void init(int b, int k, int l) { bits = b; K = k; L = l; e.assign((size_t)K << bits, -1); cnt.assign(1u << bits, 0); }
  uint32_t hash(const uint8_t *p) const {
    uint64_t v = load8(p);
    if (L < 8) v &= (1ull << (8 * L)) - 1;
    uint64_t h = v * 0x9E3779B185EBCA87ull;
    for (int o = 8; o < L; o += 8) {
      uint64_t w = load8(p + o);
      if (L - o < 8) w &= (1ull << (8 * (L - o))) - 1;
      h = (h ^ (h >> 29)) * 0xC2B2AE3D27D4EB4Full + w * 0x9E3779B185EBCA87ull;
    }
    return (uint32_t)(h >> (64 - bits));
  }
  void insert(const uint8_t *d, size_t p) {
    uint32_t h = hash(d + p);
    e[(size_t)h * K + cnt[h] % K] = (int32_t)p;
    cnt[h]++;
  }
  // candidates of position p, newest first
  int get(const uint8_t *d, size_t p, int32_t *out) const {
    uint32_t h = hash(d + p);
    int n = 0;
    uint32_t c = cnt[h];
    for (int k = 0; k < K && (uint32_t)k < c; k++) out[n++] = e[(size_t)h * K + (c - 1 - k) % K];
    return n;
  }
};

int main(int argc, char **argv) {
  if (argc < 2) return 1;
  const int CHUNK = P("CHUNK", 65536), B4 = P("B4", 12), K4 = P("K4", 8), B8 = P("B8", 12), K8 = P("K8", 8), L8 = P("L8", 8),
            L4 = P("L4", 4), MINM = P("MINM", 4), NICE = P("NICE", 128), SKIP = P("SKIP", 258), RUN = P("RUN", 1 << 30),
            LAZY2 = P("LAZY2", 0), TILE = P("TILE", 0), REP = P("REP", 1), MAXCMP = P("MAXCMP", 64), FAR4 = P("FAR4", 0);
  FILE *f = fopen(argv[1], "rb");
  fseek(f, 0, SEEK_END);
  size_t n = ftell(f);
  fseek(f, 0, SEEK_SET);
  std::vector<uint8_t> d(n + 80);
  if (fread(d.data(), 1, n, f) != n) return 2;
  fclose(f);

  Rows r4, r8, r16, r32;
  const int K16 = P("K16", 0), L16 = P("L16", 16), B16 = P("B16", 12), K32 = P("K32", 0), L32 = P("L32", 32), B32 = P("B32", 12);
  r4.init(B4, K4 ? K4 : 1, L4);
  if (K8) r8.init(B8, K8, L8);
  if (K16) r16.init(B16, K16, L16);
  if (K32) r32.init(B32, K32, L32);
  std::vector<uint16_t> mlen(n), mdist(n);
  const int NEARD = P("NEARD", 0), NEARW = P("NEARW", 4096), NEARB = P("NEARB", 13);
  std::vector<int32_t> nhead(1u << NEARB, -1), nprev(n + 1, -1);
  auto nhash = [&](const uint8_t *q) { uint32_t v; memcpy(&v, q, 4); return (v * 2654435761u) >> (32 - NEARB); };
  uint64_t n_cand = 0, n_true = 0, n_cmpbytes = 0, n_search = 0, n_probe = 0, n_ins = 0;
  int casc_depth = 4;
  const int CASCSLACK = P("CASCSLACK", 0);
  // the tables slide over the whole input (a run of chunks shares one window); matches never cross a chunk end
  int inh_len = 0, inh_dist = 0;
  std::vector<int32_t> cand(K4 + K8 + K16 + K32 + 4 + 256);
  const int KS = P("SORTED", 0), UNIT = P("UNIT", 32768), KF = P("KFWD", 0);
  const int F8 = P("F8", 8), F12 = P("F12", 8), F20 = P("F20", 8), H4B = P("H4B", 16), E8 = P("E8", 8), E12 = P("E12", 12), E20 = P("E20", 20);
  std::vector<uint64_t> skey;      // sorted (key << 20 | index in unit window)
  std::vector<uint32_t> srank;     // rank of every position of the current unit window
  size_t sbase = 0, s_unit_end = 0;
  auto fieldhash = [&](const uint8_t *q, int from, int to, int bits) -> uint64_t {
    if (bits == 0) return 0;
    uint64_t h = 0x9E3779B185EBCA87ull;
    for (int i = from; i < to; i++) h = (h ^ q[i]) * 0x100000001B3ull;
    h ^= h >> 29; h *= 0xC2B2AE3D27D4EB4Full;
    return (h >> 40) & ((1ull << bits) - 1);
  };
  auto build_unit = [&](size_t u0) {
    sbase = u0 > 32768 ? u0 - 32768 : 0;
    s_unit_end = u0 + UNIT < n ? u0 + UNIT : n;
    skey.clear();
    for (size_t q = sbase; q < s_unit_end; q++) {
      if (q + 24 > n + 60) break;
      const uint8_t *a = &d[q];
      bool interior = REP && q > 0 && memcmp(a, a - 1, 4) == 0 && a[3] == a[4];
      if (interior) continue;
      uint64_t k = fieldhash(a, 0, 4, H4B);
      k = (k << F8) | fieldhash(a, 4, E8, F8);
      k = (k << F12) | fieldhash(a, E8, E12, F12);
      k = (k << F20) | fieldhash(a, E12, E20, F20);
      skey.push_back((k << 20) | (uint64_t)(q - sbase));
    }
    std::sort(skey.begin(), skey.end());
    srank.assign(s_unit_end - sbase, 0xffffffffu);
    for (size_t i = 0; i < skey.size(); i++) srank[skey[i] & 0xfffff] = (uint32_t)i;
  };
  for (size_t p = 0; p < n; p++) {
    const size_t cs = p / CHUNK * CHUNK, ce = cs + CHUNK < n ? cs + CHUNK : n;
    if ((p - cs) % RUN == 0) inh_len = 0;
    const int maxlen = (int)(ce - p < 258 ? ce - p : 258);
    int best = 0, bdist = 0;
    if (inh_len > MINM) { best = inh_len - 1 < maxlen ? inh_len - 1 : maxlen; bdist = inh_dist; }
    if (maxlen >= MINM && best < SKIP && best < NICE) {
      n_search++;
      int nc = 0;
      if (REP && p > 0) cand[nc++] = (int32_t)(p - 1);
      if (KS) {
        if (p >= s_unit_end || skey.empty() && p == 0) build_unit(p);
        // own key (the position may be a skipped interior: then locate by lower bound)
        const uint8_t *a = &d[p];
        uint64_t k = fieldhash(a, 0, 4, H4B);
        k = (k << F8) | fieldhash(a, 4, E8, F8);
        k = (k << F12) | fieldhash(a, E8, E12, F12);
        k = (k << F20) | fieldhash(a, E12, E20, F20);
        const uint64_t me = (k << 20) | (uint64_t)(p - sbase);
        size_t r = std::lower_bound(skey.begin(), skey.end(), me) - skey.begin();
        const int shift = 20 + F8 + F12 + F20;
        int taken = 0;
        for (size_t i = r; i-- > 0 && taken < KS;) {
          if ((skey[i] >> shift) != (me >> shift)) break;
          const size_t q = sbase + (skey[i] & 0xfffff);
          taken++;
          cand[nc++] = (int32_t)q;
        }
        taken = 0;
        for (size_t i = r + 1; i < skey.size() && taken < KF; i++) {
          if ((skey[i] >> shift) != (me >> shift)) break;
          const size_t q = sbase + (skey[i] & 0xfffff);
          taken++;
          if (q < p) cand[nc++] = (int32_t)q;
        }
      }
      static const int CASC = P("CASC", 0);
      if (!CASC) {
      if (K32) nc += r32.get(d.data(), p, &cand[nc]);
      if (K16) nc += r16.get(d.data(), p, &cand[nc]);
      if (K8) nc += r8.get(d.data(), p, &cand[nc]);
      if (K4) nc += r4.get(d.data(), p, &cand[nc]);
      } else {
        // cascade: a deeper level is probed (and later inserted into) only when the level above is saturated with
        // true matches of its gram
        Rows *lv[4] = {&r4, &r8, &r16, &r32};
        const int KK[4] = {K4, K8, K16, K32};
        casc_depth = 1;
        for (int L = 0; L < 4 && KK[L]; L++) {
          n_probe++;
          const int n0 = nc;
          nc += lv[L]->get(d.data(), p, &cand[nc]);
          int ntrue = 0;
          for (int i = n0; i < nc; i++) {
            const int32_t q = cand[i];
            if (q >= 0 && (size_t)q < p && p - q <= (size_t)MAXDIST && memcmp(&d[p], &d[q], CASC == 2 ? 4 : lv[L]->L) == 0) ntrue++;
          }
          if (ntrue < KK[L] - CASCSLACK) break;
          casc_depth = L + 2;
        }
        // deepest first for the compare order
        std::reverse(cand.begin() + (REP && p > 0 ? 1 : 0), cand.begin() + nc);
      }
      const int n_far = nc;
      if (NEARD) {
        int32_t q = nhead[nhash(&d[p])];
        for (int k = 0; k < NEARD && q >= 0 && p - q <= (size_t)NEARW; k++, q = nprev[q]) cand[nc++] = q;
        // near candidates first (they are the most recent)
        std::rotate(cand.begin() + (REP && p > 0 ? 1 : 0), cand.begin() + n_far, cand.begin() + nc);
      }
      const int n_near = nc - n_far;
      int cmps = 0;
      const uint8_t *a = &d[p];
      for (int i = 0; i < nc; i++) {
        const int32_t q = cand[i];
        if (q < 0 || (size_t)q >= p) continue;
        if (TILE && (size_t)q / TILE == p / TILE && q != (int32_t)(p - 1) && !(i < n_near + 1)) continue;
        const size_t dist = p - q;
        if (dist > (size_t)MAXDIST) continue;
        n_cand++;
        const uint8_t *b = &d[q];
        if (memcmp(a, b, 4) != 0) continue;
        n_true++;
        if (best >= MINM && a[best] != b[best]) continue;
        if (++cmps > MAXCMP) break;
        int l = 0;
        while (l < maxlen && a[l] == b[l]) l++;
        n_cmpbytes += l;
        if (l >= MINM && (l > best || (l == best && (int)dist < bdist))) {
          if (FAR4 && l == 4 && dist > (size_t)FAR4) continue;
          best = l;
          bdist = (int)dist;
          if (l >= NICE || l >= maxlen) break;
        }
      }
    }
    else if (P("CASC", 0)) casc_depth = 1;
    if (best < MINM) best = 0;
    mlen[p] = (uint16_t)best;
    mdist[p] = (uint16_t)bdist;
    inh_len = best;
    inh_dist = bdist;
    if (NEARD && p + 4 <= n) { const uint32_t h = nhash(&d[p]); nprev[p] = nhead[h]; nhead[h] = (int32_t)p; }
    // insert p (run interiors are skipped: the distance-1 probe covers them)
    if (p + 8 <= n) {
      const bool interior = p > 0 && memcmp(&d[p], &d[p - 1], 4) == 0 && d[p + 3] == d[p + 4 < n ? p + 4 : p + 3];
      if (!interior || !REP) {
        r4.insert(d.data(), p);
        static const int S8 = P("S8", 1), S16 = P("S16", 1), S32 = P("S32", 1);
        static const int CASC2 = P("CASC", 0);
        const int cdp = CASC2 ? casc_depth : 4;
        if (K8 && p % S8 == 0 && cdp >= 2) { r8.insert(d.data(), p); n_ins++; }
        if (K16 && p % S16 == 0 && cdp >= 3) { r16.insert(d.data(), p); n_ins++; }
        if (K32 && p % S32 == 0 && cdp >= 4) { r32.insert(d.data(), p); n_ins++; }
      }
    }
  }

  // parse + sizes per chunk
  uint64_t total_bits = 0, n_tok = 0, n_match = 0, n_stored = 0, n_fixed = 0, n_dyn = 0;
  HuffScratch S;
  for (size_t cs = 0; cs < n; cs += CHUNK) {
    const size_t ce = cs + CHUNK < n ? cs + CHUNK : n;
    uint32_t lf[kNumLitLen] = {0}, df[kNumDist] = {0};
    uint64_t extra = 0;
    size_t p = cs;
    while (p < ce) {
      int l = mlen[p];
      bool take = l >= MINM;
      if (take && p + 1 < ce && mlen[p + 1] > l) take = false;
      if (take && LAZY2 && p + 2 < ce && mlen[p + 2] > l + 1) take = false;
      if (take) {
        uint32_t lc = len_code(l), dc = dist_code(mdist[p]);
        lf[257 + lc]++;
        df[dc]++;
        extra += len_extra_bits(lc) + dist_extra_bits(dc);
        p += l;
        n_match++;
      } else {
        lf[d[p]]++;
        p++;
      }
      n_tok++;
    }
    lf[kEob]++;
    uint8_t ll[kNumLitLen], dl[kNumDist], cl[kNumCodeLen];
    int max_l = huff_build_lengths(lf, kNumLitLen, kMaxBits, ll, S);
    int max_d = huff_build_lengths(df, kNumDist, kMaxBits, dl, S);
    int hlit = (max_l < 256 ? 256 : max_l) + 1, hdist = max_d + 1;
    uint8_t comb[kNumLitLen + kNumDist];
    memcpy(comb, ll, hlit);
    memcpy(comb + hlit, dl, hdist);
    uint16_t rle[kNumLitLen + kNumDist];
    uint32_t cf[kNumCodeLen] = {0};
    int nr = rle_code_lengths(comb, hlit + hdist, rle, cf);
    huff_build_lengths(cf, kNumCodeLen, kMaxClBits, cl, S);
    int hclen = 19;
    while (hclen > 4 && cl[cl_order(hclen - 1)] == 0) hclen--;
    uint64_t dyn = 3 + 5 + 5 + 4 + 3 * hclen;
    for (int i = 0; i < nr; i++) dyn += cl[rle[i] & 0xff] + cl_extra_bits(rle[i] & 0xff);
    uint64_t body = extra, fixed = 3 + extra;
    for (int s = 0; s < kNumLitLen; s++) { body += (uint64_t)lf[s] * ll[s]; fixed += (uint64_t)lf[s] * fixed_litlen_bits(s); }
    for (int s = 0; s < kNumDist; s++) { body += (uint64_t)df[s] * dl[s]; fixed += (uint64_t)df[s] * 5; }
    dyn += body;
    size_t clen = ce - cs;
    uint64_t stored = 8 * (clen + 5 * ((clen + 65534) / 65535));
    uint64_t bits;
    if (stored <= dyn && stored <= fixed) { bits = stored; n_stored++; }
    else if (fixed <= dyn) { bits = (fixed + 3 + 7) / 8 * 8 + 32; n_fixed++; }
    else { bits = (dyn + 3 + 7) / 8 * 8 + 32; n_dyn++; }
    total_bits += bits;
  }
  size_t ours = (total_bits + 7) / 8 + 2;
  size_t z = zlib6(d.data(), n);
  printf("%s n=%zu ours=%zu zlib6=%zu ratio=%.4f | stored/fixed/dyn=%llu/%llu/%llu tokens=%llu matches=%llu | "
         "search/pos=%.3f cand/pos=%.2f true4/pos=%.2f cmpbytes/pos=%.2f probes/pos=%.2f deepins/pos=%.2f\n",
         argv[1], n, ours, z, (double)ours / z, (unsigned long long)n_stored, (unsigned long long)n_fixed,
         (unsigned long long)n_dyn, (unsigned long long)n_tok, (unsigned long long)n_match, (double)n_search / n,
         (double)n_cand / n, (double)n_true / n, (double)n_cmpbytes / n, (double)n_probe / n, (double)n_ins / n);
  return 0;
}
