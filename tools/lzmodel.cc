// tools/lzmodel.cc -- DESIGN-EXPLORATION TOOL (not built by build(), not linked into the product, not a
// fallback): a serial model of the block-parallel deflate algorithm the CUDA kernel implements, used in the
// build container (no GPU) to predict the compressed size against zlib level 6 and the amount of
// match-search work for a given set of parameters before GPU time is spent.
//   g++ -O2 -std=c++17 -DDSG_EMU -Itests/simt_emu -Idevspace_b200/csrc tools/lzmodel.cc -lz -o /tmp/lzmodel
//   /tmp/lzmodel file [chunk=65536] [hashbits=14] [depth=64] [minmatch=4] [nice=128] [policy=1]
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include <vector>

#include "huffman.cuh"

dim3 threadIdx, blockIdx, blockDim, gridDim;
using namespace dsg;

static int CHUNK = 65536, HB = 14, DEPTH = 64, MINM = 4, NICE = 128, POLICY = 1, GOOD = 8, LAZY = 258, RUN = 16, INSIDE = 0;
static const int WIN = 32768, MAXDIST = 32768 - 512;

static inline uint32_t hash4(const uint8_t *p) {
  uint32_t v;
  memcpy(&v, p, 4);
  if (MINM == 3) v &= 0xffffff;
  return (v * 2654435761u) >> (32 - HB);
}

struct Tok {
  uint16_t len;  // 0 = literal
  uint16_t dist; // or literal byte
};

static uint64_t g_steps = 0, g_searches = 0, g_cmpbytes = 0, g_work = 0, g_fut = 0;

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

int main(int argc, char **argv) {
  if (argc < 2) return 1;
  if (argc > 2) CHUNK = atoi(argv[2]);
  if (argc > 3) HB = atoi(argv[3]);
  if (argc > 4) DEPTH = atoi(argv[4]);
  if (argc > 5) MINM = atoi(argv[5]);
  if (argc > 6) NICE = atoi(argv[6]);
  if (argc > 7) POLICY = atoi(argv[7]);
  if (argc > 8) LAZY = atoi(argv[8]);
  if (argc > 9) RUN = atoi(argv[9]);
  if (argc > 10) INSIDE = atoi(argv[10]);
  FILE *f = fopen(argv[1], "rb");
  fseek(f, 0, SEEK_END);
  size_t n = ftell(f);
  fseek(f, 0, SEEK_SET);
  std::vector<uint8_t> d(n + 8);
  if (fread(d.data(), 1, n, f) != n) return 2;
  fclose(f);

  std::vector<int32_t> head(1 << HB), prev(n + 1), start_of(n + 1, -1);
  std::vector<uint16_t> mlen(CHUNK), mdist(CHUNK);
  std::vector<Tok> toks;
  uint64_t total_bits = 0, n_stored = 0, n_fixed = 0, n_dyn = 0, n_tok = 0, n_match = 0, hdr_bits = 0;
  HuffScratch S;
  for (size_t cs = 0; cs < n; cs += CHUNK) {
    size_t ce = cs + CHUNK < n ? cs + CHUNK : n;
    size_t ds = cs > (size_t)WIN ? cs - WIN : 0;
    std::fill(head.begin(), head.end(), -1);
    {
      static int SCR = getenv("SCRAMBLE") ? atoi(getenv("SCRAMBLE")) : 0;
      std::vector<size_t> ord;
      for (size_t p = ds; p < ce; p++) ord.push_back(p);
      if (SCR > 1) {
        // rounds of SCR consecutive positions (aligned to the chunk start like the kernel's tiles), random order inside
        size_t i0 = 0;
        while (i0 < ord.size()) {
          size_t i1 = i0 + SCR - ((ord[i0] - ds) % SCR);
          if (i1 > ord.size()) i1 = ord.size();
          for (size_t k = i1 - 1; k > i0; k--) { size_t j = i0 + rand() % (k - i0 + 1); std::swap(ord[k], ord[j]); }
          i0 = i1;
        }
      }
      static int SNAP = getenv("SNAP") ? atoi(getenv("SNAP")) : 0;
      size_t k0 = 0;
      for (size_t k = 0; k < ord.size(); k++) {
        size_t p = ord[k];
        static int RUNX = getenv("RUNX") ? atoi(getenv("RUNX")) : 0;
        if (p + 4 > n) { prev[p] = -1; }
        else {
          uint32_t h = hash4(&d[p]);
          prev[p] = head[h];
          bool interior = RUNX && p > ds && memcmp(&d[p], &d[p - 1], 4) == 0;
          if (!interior) head[h] = (int32_t)p;
        }
        // end of a round: snapshot the head for every position of the round
        bool round_end = (k + 1 == ord.size()) || (SCR > 1 && ((k + 1 + (ord.size() ? 0 : 0)) , ((ord[k0] - ds) / SCR != (ord[k + 1] - ds) / SCR)));
        if (SNAP && round_end) {
          for (size_t j = k0; j <= k; j++) { size_t pp = ord[j]; start_of[pp] = (pp + 4 > n) ? -1 : head[hash4(&d[pp])]; }
          k0 = k + 1;
        }
      }
    }
    // per-position best match
    std::vector<uint8_t> deep(CHUNK, 0);
    int inh_len = 0, inh_dist = 0, rep_dist = 0;
    size_t cover_end = cs;   // policy 3: end of the match the thread-local speculative parse is inside of
    size_t lazy_pos = (size_t)-1;
    int lazy_len = 0;
    for (size_t p = cs; p < ce; p++) {
      if ((p - cs) % RUN == 0) { inh_len = 0; cover_end = p; lazy_pos = (size_t)-1; rep_dist = 0; }
      int maxlen = (int)(ce - p < 258 ? ce - p : 258);
      int best = 0, bdist = 0;
      // inherited candidate from p-1
      if (inh_len > MINM) { best = inh_len - 1; bdist = inh_dist; if (best > maxlen) best = maxlen; }
      int depth = DEPTH;
      if (POLICY == 1 || POLICY == 2) {
        if (best >= NICE) depth = 0;
        else if (best >= 32 && POLICY >= 2) depth = 0;
        else if (best >= GOOD) depth = DEPTH / 4;
      } else if (POLICY == 4 || POLICY == 5) {
        depth = INSIDE;
        if (best >= NICE) depth = 0;
      } else if (POLICY == 3) {
        if (p == lazy_pos) depth = lazy_len >= GOOD ? DEPTH / 4 : DEPTH;   // zlib: lazy evaluation of p+1
        else if (p < cover_end) depth = INSIDE;                             // inside a taken match
        else depth = DEPTH;                                                 // token start: full effort
        if (best >= NICE) depth = 0;
      }
      static int REP = getenv("REP") ? atoi(getenv("REP")) : 0;
      static int RUNX2 = getenv("RUNX") ? atoi(getenv("RUNX")) : 0;
      if (RUNX2 && maxlen >= MINM && best < NICE && p > ds && memcmp(&d[p], &d[p - 1], 4) == 0) {
        const uint8_t *a = &d[p], *b = &d[p - 1];
        int l = 0;
        while (l < maxlen && a[l] == b[l]) l++;
        if (l > best || (l == best && 1 < bdist)) { best = l; bdist = 1; }
        if (best >= GOOD && depth > DEPTH / 4) depth = DEPTH / 4;
        if (best >= NICE) depth = 0;
      }
      if (REP && maxlen >= MINM && best < NICE) {   // explicit near probes: distance 1 and the run's last distance
        int cand[2] = {1, rep_dist};
        for (int ci = 0; ci < 2; ci++) {
          int dd = cand[ci];
          if (dd <= 0 || (size_t)dd > p - ds) continue;
          const uint8_t *a = &d[p], *b = &d[p - dd];
          int l = 0;
          while (l < maxlen && a[l] == b[l]) l++;
          if (l >= MINM && (l > best || (l == best && dd < bdist))) { best = l; bdist = dd; }
        }
      }
      if (maxlen >= MINM && depth > 0) {
        g_searches++;
        static int SNAP2 = getenv("SNAP") ? atoi(getenv("SNAP")) : 0;
        int32_t q = SNAP2 ? start_of[p] : prev[p];
        static int FF = getenv("FUTURE_FREE") ? atoi(getenv("FUTURE_FREE")) : 0;
        while (q >= 0 && depth-- > 0) {
          if ((size_t)q >= p) { if (FF) depth++; g_steps++; g_fut++; q = prev[q]; continue; }   // same-round entry later in the stream
          size_t dist = p - q;
          if (dist > (size_t)MAXDIST) break;
          g_steps++;
          const uint8_t *a = &d[p], *b = &d[q];
          if (best < MINM || (a[best] == b[best] && a[best - 1] == b[best - 1])) {
            int l = 0;
            while (l < maxlen && a[l] == b[l]) l++;
            g_cmpbytes += l;
            if (l >= MINM && (l > best || (l == best && (int)dist < bdist))) {
              best = l;
              bdist = (int)dist;
              if (l >= NICE || l >= maxlen) break;
            }
          }
          q = prev[q];
        }
      }
      if (best < MINM) best = 0;
      if (MINM == 3 && best == 3 && bdist > 4096) best = 0;  // zlib TOO_FAR
      mlen[p - cs] = (uint16_t)best;
      mdist[p - cs] = (uint16_t)bdist;
      if (POLICY == 3) {
        if (p == lazy_pos) {
          if (best > lazy_len) { lazy_pos = p + 1; lazy_len = best; cover_end = p + 1; }   // defer again
          else { cover_end = (p - 1) + lazy_len; lazy_pos = (size_t)-1; }                 // take the match at p-1
        } else if (p >= cover_end) {
          if (best >= MINM) { lazy_pos = p + 1; lazy_len = best; cover_end = p + 1; }
          else cover_end = p + 1;
        }
      }
      inh_len = best;
      inh_dist = bdist;
      if (best >= MINM) rep_dist = bdist;
    }
    if (POLICY == 5) {
      // refine without a first parse: deep-search every position that is not a pure inheritor of its predecessor
      std::vector<size_t> work;
      for (size_t p = cs; p < ce; p++) {
        bool inh = p > cs && (p - cs) % RUN != 0 && mlen[p - cs] >= (LAZY > 100 ? MINM : LAZY) && mlen[p - 1 - cs] == mlen[p - cs] + 1 &&
                   mdist[p - 1 - cs] == mdist[p - cs];
        if (!inh) work.push_back(p);
      }
      g_work += work.size();
      std::vector<uint16_t> nl(mlen), nd(mdist);
      for (size_t p : work) {
        int maxlen = (int)(ce - p < 258 ? ce - p : 258);
        int best = mlen[p - cs], bdist = mdist[p - cs];
        int depth = DEPTH;
        if (p > cs && mlen[p - 1 - cs] >= GOOD) depth = DEPTH / 4;
        if (best >= NICE || maxlen < MINM) continue;
        g_searches++;
        int32_t c = prev[p];
        while (c >= 0 && depth-- > 0) {
          size_t dist = p - c;
          if (dist > (size_t)MAXDIST) break;
          g_steps++;
          const uint8_t *a = &d[p], *b = &d[c];
          if (best < MINM || (a[best] == b[best] && a[best - 1] == b[best - 1])) {
            int l = 0;
            while (l < maxlen && a[l] == b[l]) l++;
            g_cmpbytes += l;
            if (l > best && l >= MINM) { best = l; bdist = (int)dist; if (l >= NICE || l >= maxlen) break; }
          }
          c = prev[c];
        }
        nl[p - cs] = (uint16_t)best;
        nd[p - cs] = (uint16_t)bdist;
      }
      mlen = nl; mdist = nd;
    }
    if (POLICY == 4) {
      // two-phase refine: the loop above ran with the shallow depth (INSIDE); now parse, deep-search the token
      // starts and their lazy successors, and parse again. ROUNDS passes.
      for (int round = 0; round < (LAZY > 100 ? 1 : LAZY); round++) {
        std::vector<size_t> work;
        size_t q = cs;
        while (q < ce) {
          int l = mlen[q - cs];
          bool take = l >= MINM && !(q + 1 < ce && mlen[q + 1 - cs] > l);
          if (!deep[q - cs]) work.push_back(q);
          if (l >= MINM && q + 1 < ce && !deep[q + 1 - cs]) work.push_back(q + 1);
          q += take ? l : 1;
        }
        for (size_t p : work) {
          if (deep[p - cs]) continue;
          deep[p - cs] = 1;
          int maxlen = (int)(ce - p < 258 ? ce - p : 258);
          int best = mlen[p - cs], bdist = mdist[p - cs];
          int depth = DEPTH;
          // lazy successor of a good match: zlib quarters the effort
          if (p > cs && mlen[p - 1 - cs] >= GOOD) depth = DEPTH / 4;
          if (best >= NICE || maxlen < MINM) continue;
          g_searches++;
          int32_t c = prev[p];
          while (c >= 0 && depth-- > 0) {
            size_t dist = p - c;
            if (dist > (size_t)MAXDIST) break;
            g_steps++;
            const uint8_t *a = &d[p], *b = &d[c];
            if (best < MINM || (a[best] == b[best] && a[best - 1] == b[best - 1])) {
              int l = 0;
              while (l < maxlen && a[l] == b[l]) l++;
              g_cmpbytes += l;
              if (l > best && l >= MINM) { best = l; bdist = (int)dist; if (l >= NICE || l >= maxlen) break; }
            }
            c = prev[c];
          }
          mlen[p - cs] = (uint16_t)best;
          mdist[p - cs] = (uint16_t)bdist;
        }
      }
    }
    // stateless lazy parse
    toks.clear();
    uint32_t lf[kNumLitLen] = {0}, df[kNumDist] = {0};
    uint64_t extra = 0;
    size_t p = cs;
    while (p < ce) {
      int l = mlen[p - cs];
      bool take = l >= MINM;
      if (take && (POLICY >= 4 || l < LAZY) && p + 1 < ce && mlen[p + 1 - cs] > l) take = false;
      if (take) {
        toks.push_back({(uint16_t)l, mdist[p - cs]});
        uint32_t lc = len_code(l), dc = dist_code(mdist[p - cs]);
        lf[257 + lc]++;
        df[dc]++;
        extra += len_extra_bits(lc) + dist_extra_bits(dc);
        p += l;
        n_match++;
      } else {
        toks.push_back({0, d[p]});
        lf[d[p]]++;
        p++;
      }
    }
    lf[kEob]++;
    n_tok += toks.size();
    // sizes
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
    uint64_t hb = dyn;
    uint64_t body = extra, fixed = 3 + extra;
    for (int s = 0; s < kNumLitLen; s++) { body += (uint64_t)lf[s] * ll[s]; fixed += (uint64_t)lf[s] * fixed_litlen_bits(s); }
    for (int s = 0; s < kNumDist; s++) { body += (uint64_t)df[s] * dl[s]; fixed += (uint64_t)df[s] * 5; }
    dyn += body;
    size_t clen = ce - cs;
    uint64_t stored = 8 * (clen + 5 * ((clen + 65534) / 65535));
    uint64_t bits;
    if (stored <= dyn && stored <= fixed) { bits = stored; n_stored++; }
    else if (fixed <= dyn) { bits = fixed; n_fixed++; bits = (bits + 7) / 8 * 8 + 0; bits += 40; }
    else { bits = dyn; n_dyn++; hdr_bits += hb; bits += 3 + 7; bits = (bits + 7) / 8 * 8; bits += 32; }  // + sync marker: empty stored block
    total_bits += bits;
  }
  size_t ours = (total_bits + 7) / 8 + 2;
  size_t z = zlib6(d.data(), n);
  printf("%s n=%zu ours=%zu zlib6=%zu ratio=%.4f | chunks stored/fixed/dyn=%llu/%llu/%llu tokens=%llu matches=%llu "
         "hdr=%.2f%% | searches/pos=%.3f steps/pos=%.2f cmpbytes/pos=%.2f deepfrac=%.3f\n",
         argv[1], n, ours, z, (double)ours / z, (unsigned long long)n_stored, (unsigned long long)n_fixed,
         (unsigned long long)n_dyn, (unsigned long long)n_tok, (unsigned long long)n_match,
         100.0 * hdr_bits / 8 / ours, (double)g_searches / n, (double)g_steps / n, (double)g_cmpbytes / n, (double)g_work / n); printf("future skips/pos=%.2f\n", (double)g_fut / n);
  return 0;
}
