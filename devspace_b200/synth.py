"""Synthetic file trees of the shapes BASELINE.json names, generated with torch so that the same
code runs on the CPU (small trees for tests) and on the GPU (multi-GiB arenas for bench.py, then copied to
pinned host memory). Seeded and deterministic per (config, seed, device type). Data only: no compute of the
pack path happens here.

  C2  node_modules-like: many ~4 KB JS/JSON-like text files, nested paths, ~5 % names needing the USTAR split
  C3  mixed source tree: log-normal sizes 1 KB..8 MB; 70 % text, 20 % semi-structured binary, 10 % random
  C4  build context: few large files (256 MiB each at full scale) of an ELF-like page mix (~2x at level 6)
  C5a one million empty/tiny files (header bound)      C5b incompressible random bytes
"""
from __future__ import annotations

import math
from typing import List, Optional, Tuple

import numpy as np
import torch

from .manifest import ENTRY_DTYPE, Manifest, TYPE_REG

MTIME0 = 1_546_300_800
SEED0 = 0xD5C0DE


def _gen(seed: int, device) -> torch.Generator:
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return g


# ---- content generators: return uint8 tensors on `device` -----------------------------------------------------

_VOCAB_CACHE = {}


def _vocab(device, nwords: int = 6000, seed: int = 7):
    """Identifier-like words with Zipf weights + punctuation/keyword tokens. Returns (flat bytes, starts, lens, probs)."""
    key = (str(device), nwords, seed)
    if key in _VOCAB_CACHE:
        return _VOCAB_CACHE[key]
    rng = np.random.default_rng(seed)
    kw = [b"function ", b"return ", b"const ", b"let ", b"var ", b"if (", b") {\n", b"}\n", b"};\n", b" = ", b" => ", b"(",
          b")", b", ", b";\n", b".", b"require('", b"')", b"module.exports", b"\n", b"  ", b"    ", b"      ", b"{ ",
          b" }", b"[", b"]", b": ", b"\"", b"\",\n", b"null", b"true", b"false", b"this.", b"new ", b"typeof ",
          b" === ", b" !== ", b" && ", b" || ", b"// ", b"/* ", b" */\n", b"0", b"1", b"2", b"10", b"100", b"'use strict';\n",
          b"\"version\": \"", b"\"resolved\": \"https://registry.npmjs.org/", b"\"integrity\": \"sha512-", b"\"requires\": {\n",
          b"\"dependencies\": {\n", b".tgz\",\n", b"\"dev\": true,\n"]
    syll = [b"get", b"set", b"on", b"is", b"has", b"to", b"from", b"with", b"name", b"value", b"index", b"item", b"list",
            b"map", b"key", b"node", b"path", b"file", b"data", b"error", b"event", b"state", b"props", b"config", b"option",
            b"result", b"callback", b"handler", b"request", b"response", b"buffer", b"stream", b"length", b"type", b"id",
            b"user", b"module", b"export", b"parse", b"format", b"create", b"update", b"delete", b"render", b"load"]
    words: List[bytes] = list(kw)
    while len(words) < nwords:
        k = int(rng.integers(1, 4))
        parts = [syll[int(i)] for i in rng.integers(0, len(syll), k)]
        w = parts[0] + b"".join(p.capitalize() for p in parts[1:])
        if rng.random() < 0.3:
            w += bytes(rng.integers(97, 123, int(rng.integers(1, 5)), dtype=np.uint8))
        words.append(w)
    lens = np.array([len(w) for w in words], np.int64)
    starts = np.concatenate([[0], np.cumsum(lens)[:-1]])
    flat = np.frombuffer(b"".join(words), np.uint8).copy()
    ranks = np.arange(1, len(words) + 1, dtype=np.float64)
    probs = 1.0 / ranks ** 1.05
    # keywords/punctuation are the most frequent tokens
    probs = probs / probs.sum()
    out = (torch.from_numpy(flat).to(device), torch.from_numpy(starts).to(device), torch.from_numpy(lens).to(device),
           torch.from_numpy(probs).to(device=device, dtype=torch.float32))
    _VOCAB_CACHE[key] = out
    return out


def text_like(nbytes: int, seed: int, device="cpu") -> torch.Tensor:
    """JS/JSON-like source text (gzip -6 ratio ~3.5-4.5x): lines are drawn from a pool of line templates (token
    sequences over a Zipf vocabulary) and re-emitted with a few tokens swapped for fresh identifiers, the way
    real code repeats statements with different names. Most of the redundancy therefore sits in repeated
    phrases of 20-80 bytes at distances up to tens of KiB, as in real source trees, not in the unigram
    statistics of a token soup."""
    if nbytes <= 0:
        return torch.zeros(0, dtype=torch.uint8, device=device)
    flat, starts, lens, probs = _vocab(device)
    g = _gen(seed, device)
    mean_len = float((lens.double() * probs.double()).sum())
    TPL, LINE = 384, 12                                    # templates per stream, tokens per line
    ntok = int(nbytes / mean_len * 1.2) + 4 * LINE
    nlines = ntok // LINE + 1
    tpl = torch.multinomial(probs, TPL * LINE, replacement=True, generator=g).view(TPL, LINE)
    # template choice: mostly recent templates (locality), sometimes any
    pick = torch.randint(0, TPL, (nlines,), generator=g, device=device)
    near = (torch.arange(nlines, device=device) // 24 * 7 + torch.randint(0, 24, (nlines,), generator=g, device=device)) % TPL
    pick = torch.where(torch.rand(nlines, generator=g, device=device) < 0.8, near, pick)
    ids = tpl[pick]                                        # (nlines, LINE)
    # substitute ~25 % of the tokens by fresh draws (identifiers / literals differ from line to line)
    fresh = torch.multinomial(probs, nlines * LINE, replacement=True, generator=g).view(nlines, LINE)
    sub = torch.rand(nlines, LINE, generator=g, device=device) < 0.25
    ids = torch.where(sub, fresh, ids).reshape(-1)
    ntok = ids.numel()
    tl = lens[ids]
    # every line ends with a newline token's bytes already present in many templates; lengths as drawn
    ends = torch.cumsum(tl, 0)
    total = int(ends[-1])
    tok_of_byte = torch.repeat_interleave(torch.arange(ntok, device=device), tl)
    within = torch.arange(total, device=device) - (ends - tl)[tok_of_byte]
    out = flat[starts[ids][tok_of_byte] + within]
    if out.numel() < nbytes:  # extremely unlikely; pad by tiling
        out = out.repeat(nbytes // out.numel() + 1)
    return out[:nbytes].contiguous()


def random_bytes(nbytes: int, seed: int, device="cpu") -> torch.Tensor:
    g = _gen(seed, device)
    return torch.randint(0, 256, (nbytes,), dtype=torch.uint8, generator=g, device=device)


def binary_like(nbytes: int, seed: int, device="cpu") -> torch.Tensor:
    """ELF-like page mix: 4 KiB pages of four kinds (random code-like, zero-heavy, 16-byte records, string table)."""
    if nbytes <= 0:
        return torch.zeros(0, dtype=torch.uint8, device=device)
    g = _gen(seed, device)
    npages = (nbytes + 4095) // 4096
    kind = torch.randint(0, 4, (npages,), generator=g, device=device)
    out = torch.empty(npages * 4096, dtype=torch.uint8, device=device).view(npages, 4096)
    # kind 0: high entropy
    idx = (kind == 0).nonzero().flatten()
    if idx.numel():
        out[idx] = torch.randint(0, 256, (idx.numel(), 4096), dtype=torch.uint8, generator=g, device=device)
    # kind 1: zero heavy with sparse bytes
    idx = (kind == 1).nonzero().flatten()
    if idx.numel():
        r = torch.randint(0, 256, (idx.numel(), 4096), dtype=torch.uint8, generator=g, device=device)
        m = torch.rand(idx.numel(), 4096, generator=g, device=device) < 0.04
        out[idx] = torch.where(m, r, torch.zeros_like(r))
    # kind 2: 16-byte records: counter(4) | per-page constant(8) | random(4)
    idx = (kind == 2).nonzero().flatten()
    if idx.numel():
        n2 = idx.numel()
        rec = torch.empty(n2, 256, 16, dtype=torch.uint8, device=device)
        ctr = (torch.arange(256, device=device).view(1, 256) * 24 +
               torch.randint(0, 1 << 20, (n2, 1), generator=g, device=device))
        for b in range(4):
            rec[:, :, b] = ((ctr >> (8 * b)) & 0xff).to(torch.uint8)
        rec[:, :, 4:12] = torch.randint(0, 256, (n2, 1, 8), dtype=torch.uint8, generator=g, device=device)
        rec[:, :, 12:14] = torch.randint(0, 256, (n2, 256, 2), dtype=torch.uint8, generator=g, device=device)
        rec[:, :, 14:16] = 0
        out[idx] = rec.view(n2, 4096)
    # kind 3: string-table-like text
    idx = (kind == 3).nonzero().flatten()
    if idx.numel():
        t = text_like(idx.numel() * 4096, seed * 7919 + 13, device)
        out[idx] = t.view(idx.numel(), 4096)
    return out.view(-1)[:nbytes].contiguous()


# ---- trees ---------------------------------------------------------------------------------------------------------

def _lognormal_sizes(n: int, mu: float, sigma: float, lo: int, hi: int, total: int, seed: int) -> np.ndarray:
    rng = np.random.default_rng(seed)
    s = np.clip(rng.lognormal(mu, sigma, n), lo, hi)
    s = np.maximum(lo, (s * (total / s.sum())).astype(np.int64))
    return np.minimum(s, hi * 4).astype(np.int64)


def _names_tree(n: int, seed: int, prefix: str, depth_lo: int, depth_hi: int, long_frac: float, ext: List[str]) -> List[bytes]:
    rng = np.random.default_rng(seed)
    pk = ["lodash", "express", "react", "webpack", "babel-core", "eslint", "moment", "async", "chalk", "commander",
          "debug", "request", "uuid", "yargs", "glob", "rimraf", "semver", "minimist", "mkdirp", "body-parser"]
    sub = ["lib", "dist", "src", "test", "node_modules", "internal", "helpers", "utils", "build", "es", "cjs", "types"]
    names = []
    for i in range(n):
        d = int(rng.integers(depth_lo, depth_hi + 1))
        parts = [prefix, "%s-%d" % (pk[int(rng.integers(0, len(pk)))], i // 37)]
        parts += [sub[int(rng.integers(0, len(sub)))] for _ in range(max(0, d - 3))]
        base = "f%06d%s" % (i, ext[int(rng.integers(0, len(ext)))])
        if rng.random() < long_frac:  # total length 101..200: exercises splitUSTARPath
            parts.insert(2, "a-rather-long-scoped-package-directory-name-" + "x" * int(rng.integers(20, 60)))
        names.append(("/".join(parts) + "/" + base).encode())
    return names


def build_manifest(names: List[bytes], sizes: np.ndarray, arena: np.ndarray, offsets: np.ndarray) -> Manifest:
    n = len(names)
    ent = np.zeros(n, ENTRY_DTYPE)
    strtab = bytearray(b"dev")
    offs = np.empty(n, np.uint32)
    lens = np.empty(n, np.uint32)
    for i, nm in enumerate(names):
        offs[i] = len(strtab)
        lens[i] = len(nm)
        strtab += nm
    ent["name_off"], ent["name_len"] = offs, lens
    ent["uname_off"] = ent["gname_off"] = 0
    ent["uname_len"] = ent["gname_len"] = 3
    ent["typeflag"] = TYPE_REG
    ent["mode"] = np.where(np.arange(n) % 17 == 0, 0o755, 0o644)
    ent["uid"] = ent["gid"] = 1000
    ent["mtime_sec"] = MTIME0 + np.arange(n)
    ent["size"] = sizes
    ent["data_off"] = offsets
    return Manifest(ent, np.frombuffer(bytes(strtab), np.uint8).copy(), arena)


def _layout(sizes: np.ndarray, align: int = 16) -> Tuple[np.ndarray, int]:
    padded = (sizes + (align - 1)) // align * align
    offs = np.concatenate([[0], np.cumsum(padded)[:-1]]).astype(np.uint64)
    return offs, int(padded.sum())


def _fill(arena_t: torch.Tensor, offs: np.ndarray, sizes: np.ndarray, kinds: np.ndarray, seed: int, device) -> None:
    """Content is generated per kind as one long stream and cut into the files of that kind."""
    for k, fn in ((0, text_like), (1, binary_like), (2, random_bytes)):
        idx = np.nonzero(kinds == k)[0]
        if idx.size == 0:
            continue
        tot = int(sizes[idx].sum())
        pos = 0
        CH = 64 << 20
        # stream in <=64 MiB pieces to bound temporaries
        piece = None
        piece_lo = 0
        for i in idx:
            sz = int(sizes[i])
            if sz == 0:
                continue
            if piece is None or pos + sz > piece_lo + piece.numel():
                piece_lo = pos
                piece = fn(min(max(CH, sz), tot - pos), seed + 1000 * k + int(i), device)
            o = int(offs[i])
            arena_t[o:o + sz] = piece[pos - piece_lo: pos - piece_lo + sz]
            pos += sz


def tree_c2(nfiles: int = 100_000, total: int = 400_000_000, seed: int = SEED0 + 2, device="cpu") -> Manifest:
    sizes = _lognormal_sizes(nfiles, math.log(2900), 1.0, 1, 256 << 10, total, seed)
    names = _names_tree(nfiles, seed, "node_modules", 3, 8, 0.05, [".js", ".json", ".d.ts", ".md", ".map"])
    offs, alen = _layout(sizes)
    arena = torch.zeros(max(alen, 16), dtype=torch.uint8, device=device)
    _fill(arena, offs, sizes, np.zeros(nfiles, np.int64), seed, device)
    return build_manifest(names, sizes, arena.cpu().numpy(), offs)


def tree_c3(nfiles: int = 20_000, total: int = 4_000_000_000, seed: int = SEED0 + 3, device="cpu") -> Manifest:
    sizes = _lognormal_sizes(nfiles, math.log(60_000), 1.6, 1024, 8 << 20, total, seed)
    names = _names_tree(nfiles, seed, "src", 2, 6, 0.03, [".go", ".js", ".py", ".o", ".bin", ".png", ".yaml"])
    rng = np.random.default_rng(seed + 1)
    kinds = rng.choice(3, nfiles, p=[0.7, 0.2, 0.1])
    offs, alen = _layout(sizes)
    arena = torch.zeros(max(alen, 16), dtype=torch.uint8, device=device)
    _fill(arena, offs, sizes, kinds, seed, device)
    return build_manifest(names, sizes, arena.cpu().numpy(), offs)


def tree_c4_names(nfiles: int, first: int = 0) -> List[bytes]:
    return [b"context/bin/artifact-%04d.bin" % (first + i) for i in range(nfiles)]


def tree_c4(nfiles: int = 256, file_bytes: int = 256 << 20, seed: int = SEED0 + 4, device="cpu", first: int = 0) -> Manifest:
    sizes = np.full(nfiles, file_bytes, np.int64)
    offs, alen = _layout(sizes)
    arena = torch.empty(max(alen, 16), dtype=torch.uint8, device=device)
    for i in range(nfiles):
        arena[int(offs[i]): int(offs[i]) + file_bytes] = binary_like(file_bytes, seed + 31 * (first + i), device)
    return build_manifest(tree_c4_names(nfiles, first), sizes, arena.cpu().numpy(), offs)


def tree_c5a(nfiles: int = 1_000_000, seed: int = SEED0 + 5) -> Manifest:
    rng = np.random.default_rng(seed)
    sizes = np.where(rng.random(nfiles) < 0.9, 0, rng.integers(1, 65, nfiles)).astype(np.int64)
    names = [b"tiny/d%03d/f%07d" % (i // 1000, i) for i in range(nfiles)]
    offs, alen = _layout(sizes)
    arena = rng.integers(97, 123, max(alen, 16), dtype=np.uint8)
    return build_manifest(names, sizes, arena, offs)


def tree_c5b(nfiles: int = 32, total: int = 8_000_000_000, seed: int = SEED0 + 6, device="cpu") -> Manifest:
    sizes = np.full(nfiles, total // nfiles, np.int64)
    offs, alen = _layout(sizes)
    arena = random_bytes(max(alen, 16), seed, device)
    return build_manifest([b"blobs/random-%02d.bin" % i for i in range(nfiles)], sizes, arena.cpu().numpy(), offs)
