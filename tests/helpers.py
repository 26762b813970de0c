"""Shared test fixtures: deterministic header lists (the golden/edge suite of SURVEY.md 8c) and loaders.

Everything here is machine independent: uid/gid/uname/gname/mtime are fixed in the manifest, never read
from the host.
"""
from __future__ import annotations

import ctypes
import hashlib
import io
import os
import subprocess
import sys
import tarfile
from typing import Dict, List

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import oracle  # noqa: E402  (tests are allowed to use the oracle)
from oracle import Entry  # noqa: E402

MTIME0 = 1_546_300_800  # 2019-01-01T00:00:00Z


def rng_bytes(seed: int, n: int) -> bytes:
    return np.random.default_rng(seed).integers(0, 256, n, dtype=np.uint8).tobytes()


def text_bytes(seed: int, n: int) -> bytes:
    """Compressible pseudo-text (word soup)."""
    rng = np.random.default_rng(seed)
    words = [b"const", b"function", b"return", b"require", b"module", b"exports", b"value", b"index", b"error",
             b"callback", b"promise", b"string", b"number", b"object", b"length", b"=>", b"{", b"}", b"(", b")", b";",
             b"\n", b"  ", b"    ", b"if", b"else", b"for", b"var", b"let", b"this", b"null", b"true", b"false"]
    out = bytearray()
    idx = rng.integers(0, len(words), n // 3 + 8)
    for i in idx:
        out += words[int(i)]
        out += b" "
        if len(out) >= n:
            break
    return bytes(out[:n])


def F(name, data=b"", mode=0o644, uid=1000, gid=1000, uname="dev", gname="dev", mtime=MTIME0, nsec=0) -> Entry:
    return Entry(name=name, size=len(data), mtime_sec=mtime, mtime_nsec=nsec, mode=mode, uid=uid, gid=gid,
                 uname=uname, gname=gname, typeflag=ord("0"), data=data)


def D(name, mode=0o755, uid=1000, gid=1000, uname="dev", gname="dev", mtime=MTIME0, nsec=0) -> Entry:
    return Entry(name=name, size=0, mtime_sec=mtime, mtime_nsec=nsec, mode=mode, uid=uid, gid=gid, uname=uname,
                 gname=gname, typeflag=ord("5"), data=b"")


def util_test_tree() -> List[Entry]:
    """What writeTar emits for TestCopyToContainerTestable's tree (sync/util_test.go:25-40) after the
    excludes: DFS in sorted ReadDir order, files 0666 & ~umask(022) = 0644, empty dir testFolder2."""
    c = b"TestContents"
    return [
        F("testFile1", c), F("testFile2", c),
        F("testFolder/testFile1", c), F("testFolder/testFile2", c),
        D("testFolder2"),
    ]


def edge_suite() -> Dict[str, List[Entry]]:
    s: Dict[str, List[Entry]] = {}
    s["util_test_tree"] = util_test_tree()
    s["empty_archive"] = []
    s["single_empty_file"] = [F("empty")]
    s["sizes_511_512_513"] = [F("f511", rng_bytes(1, 511)), F("f512", rng_bytes(2, 512)), F("f513", rng_bytes(3, 513))]
    s["one_byte"] = [F("b", b"x")]
    # name lengths around the USTAR limits
    s["name_100"] = [F("n" * 100, b"a")]
    s["name_101_nosplit"] = [F("n" * 101, b"a")]                       # no '/', cannot split -> PAX
    s["name_101_split"] = [F("dir/" + "n" * 97, b"a")]                 # 101 bytes, splits at "dir"
    s["name_155_100"] = [F("p" * 155 + "/" + "s" * 100, b"a")]          # maximal USTAR split
    s["name_156_100"] = [F("p" * 156 + "/" + "s" * 100, b"a")]          # prefix too long -> PAX
    s["name_155_101"] = [F("p" * 155 + "/" + "s" * 101, b"a")]          # suffix too long -> PAX
    s["name_256_deep"] = [F("/".join(["seg%02d" % i + "x" * 10 for i in range(16)]) + "/" + "f" * 47, b"abc")]
    s["name_300_pax"] = [F("a/" + "b" * 140 + "/" + "c" * 156, b"abc")]
    s["name_leading_slash"] = [F("/sub/file", b"initial-sync style name")]
    s["name_leading_slash_long"] = [F("/" + "d" * 120 + "/" + "f" * 30, b"x")]
    s["name_pax_trailing_slash_cut"] = [F("q" * 99 + "/" + "r" * 160, b"x")]  # byte 100 of the name is '/'
    s["name_nonascii"] = [F("dü/ñandú.txt", b"utf8 name -> PAX path record")]
    s["name_nonascii_long"] = [F("é" * 70 + "/" + "z" * 20, b"x")]
    s["uname_nonascii"] = [F("f", b"x", uname="dév")]
    s["gname_33"] = [F("f", b"x", gname="g" * 33)]
    s["uname_32"] = [F("f", b"x", uname="u" * 32, gname="g" * 32)]
    # numeric limits
    s["uid_2097151"] = [F("f", b"x", uid=0o7777777, gid=0o7777777)]
    s["uid_2097152"] = [F("f", b"x", uid=0o10000000)]
    s["gid_2097152"] = [F("f", b"x", gid=0o10000000)]
    s["uid_gid_big"] = [F("f", b"x", uid=4_000_000_000, gid=3_000_000_000)]
    s["modes"] = [F("suid", b"x", mode=0o4755), F("sgid", b"x", mode=0o2755), F("sticky", b"x", mode=0o1777),
                  F("none", b"x", mode=0), F("all", b"x", mode=0o7777)]
    s["mtime_rounding"] = [F("down", b"x", mtime=MTIME0, nsec=499_999_999), F("up", b"x", mtime=MTIME0, nsec=500_000_000),
                           F("exact", b"x", mtime=MTIME0, nsec=0), F("zero", b"x", mtime=0)]
    s["mtime_negative"] = [F("old", b"x", mtime=-5)]
    s["mtime_huge"] = [F("far", b"x", mtime=0o100000000000)]           # 2^33: needs PAX mtime
    s["empty_dir"] = [D("testFolder2")]
    s["dir_long_name"] = [D("e" * 90 + "/" + "f" * 90)]
    s["remote_override"] = [F("indexed", b"x", mode=0o600, uid=0, gid=0, uname="dev", gname="dev")]  # tar.go:287-291
    s["pax_multi"] = [F("ü" * 60, b"x", uid=0o10000000, gid=0o10000000, uname="üser", gname="g" * 40,
                        mtime=-1)]
    s["pax_record_len_99_100"] = [F("k" * 101, b"x"), F("k" * 102 + "é"[:0], b"y"), F("m" * 90 + "é", b"z")]
    big_name = "L" * 5000
    s["name_5000"] = [F("d/" + big_name, b"x")]
    # mixed archive: many small files, PAX entries in the middle, a large file crossing several layout tiles
    mixed: List[Entry] = []
    for i in range(300):
        sz = [0, 1, 17, 511, 512, 513, 4096, 5000][i % 8]
        mixed.append(F("node_modules/pkg%03d/lib/file%03d.js" % (i // 7, i), text_bytes(100 + i, sz), mtime=MTIME0 + i))
        if i % 50 == 25:
            mixed.append(F("node_modules/pkg%03d/" % i + "é" * 30 + ".js", text_bytes(7, 700), mtime=MTIME0 + i))
        if i % 60 == 30:
            mixed.append(D("node_modules/pkg%03d/emptydir" % i, mtime=MTIME0 + i))
    mixed.append(F("big/blob.bin", rng_bytes(42, 200_000)))
    mixed.append(F("big/text.txt", text_bytes(43, 150_001)))
    s["mixed_tree"] = mixed
    return s


ERROR_CASES: Dict[str, List[Entry]] = {
    "trailing_slash_file": [F("dir/", b"")],
    "name_with_nul": [F("a\0b" + "c" * 120, b"x")],
    "mode_too_big": [F("f", b"x", mode=0o10000000)],
    "bad_typeflag": [Entry(name=b"l", typeflag=ord("3"))],   # character device: neither dialect produces it
}


def sha256(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def decode_with_tarfile(stream: bytes):
    tf = tarfile.open(fileobj=io.BytesIO(stream), mode="r:")
    out = []
    for m in tf.getmembers():
        data = tf.extractfile(m).read() if m.isreg() else b""
        out.append((m, data))
    return out


# ---- large manifests: oracle stream and gunzip as running SHA-256 (nothing is held twice) --------------------

def manifest_entries(man, lo: int, hi: int) -> List[Entry]:
    """oracle.Entry views of manifest rows [lo, hi) (payloads are numpy slices of the arena, not copies)."""
    st = man.strtab
    out = []
    for i in range(lo, hi):
        e = man.entries[i]
        tf = int(e["typeflag"])
        data = man.arena[int(e["data_off"]): int(e["data_off"]) + int(e["size"])] if tf == ord("0") else b""
        out.append(Entry(name=st[int(e["name_off"]): int(e["name_off"]) + int(e["name_len"])].tobytes(),
                         size=int(e["size"]), mtime_sec=int(e["mtime_sec"]), mtime_nsec=int(e["mtime_nsec"]),
                         mode=int(e["mode"]), uid=int(e["uid"]), gid=int(e["gid"]),
                         uname=st[int(e["uname_off"]): int(e["uname_off"]) + int(e["uname_len"])].tobytes(),
                         gname=st[int(e["gname_off"]): int(e["gname_off"]) + int(e["gname_len"])].tobytes(),
                         typeflag=tf, data=data))
    return out


def oracle_tar_sha256(man, group_bytes: int = 256 << 20, group_entries: int = 20000):
    """(sha256 hex, length) of the oracle's tar stream for a whole manifest, built group by group."""
    h = hashlib.sha256()
    total = 0
    sizes = man.entries["size"]
    lo = 0
    while lo < man.n:
        hi, acc = lo, 0
        while hi < man.n and hi - lo < group_entries and (hi == lo or acc + int(sizes[hi]) <= group_bytes):
            acc += int(sizes[hi])
            hi += 1
        total += oracle.tar_stream_update(manifest_entries(man, lo, hi), h.update, trailer=False)
        lo = hi
    h.update(bytes(1024))
    return h.hexdigest(), total + 1024


def gunzip_sha256(buf, piece: int = 16 << 20):
    """(sha256 hex, length, members) of the decompressed content of a (multi-member) gzip buffer, streamed."""
    import zlib
    h = hashlib.sha256()
    mv = memoryview(buf)
    n, pos, members = 0, 0, 0
    while pos < len(mv):
        d = zlib.decompressobj(31)
        members += 1
        while not d.eof:
            if pos >= len(mv):
                raise ValueError("gzip stream ends inside a member")
            out = d.decompress(mv[pos:pos + piece])
            pos += min(piece, len(mv) - pos)
            h.update(out)
            n += len(out)
        pos -= len(d.unused_data)
    return h.hexdigest(), n, members


# ---- SIMT emulator harness (kernel logic on CPU) ------------------------------------------------------------

_EMU = None


def emu_lib():
    global _EMU
    if _EMU is None:
        so = os.path.join(ROOT, "tests", "simt_emu", "libdsgemu.so")
        src = [os.path.join(ROOT, "tests", "emu_harness.cc"), os.path.join(ROOT, "tests", "simt_emu", "cuda_emu.h")]
        csrc = os.path.join(ROOT, "devspace_b200", "csrc")
        src += [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".cuh")]
        src.append(os.path.join(ROOT, "include", "dsgpack.h"))
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
            subprocess.run(["g++", "-O1", "-std=c++17", "-DDSG_EMU", "-fPIC", "-shared",
                            "-I" + os.path.join(ROOT, "tests", "simt_emu"), "-I" + csrc, "-o", so,
                            os.path.join(ROOT, "tests", "emu_harness.cc")], check=True)
        _EMU = ctypes.CDLL(so)
    return _EMU


def emu_tar_build(man, flags: int = 0, segment_bytes: int = 0):
    """Run the tar plan + layout kernels under the emulator. Returns (status, bytes, err_entry)."""
    lib = emu_lib()
    ent = np.ascontiguousarray(man.entries)
    strtab = np.ascontiguousarray(man.strtab)
    arena = np.ascontiguousarray(man.arena) if man.arena.size else np.zeros(16, np.uint8)
    cap = int(oracle_bound(man)) + 4096
    out = np.full(cap, 0xEE, np.uint8)
    tar_len = ctypes.c_uint64()
    err = ctypes.c_uint32(0xFFFFFFFF)
    rc = lib.emu_tar_build(ctypes.c_void_p(ent.ctypes.data), ctypes.c_uint32(man.n), ctypes.c_void_p(strtab.ctypes.data),
                           ctypes.c_uint64(strtab.size), ctypes.c_void_p(arena.ctypes.data),
                           ctypes.c_uint64(man.arena.size), ctypes.c_uint32(flags), ctypes.c_uint64(segment_bytes),
                           ctypes.c_void_p(out.ctypes.data), ctypes.c_uint64(cap), ctypes.byref(tar_len),
                           ctypes.byref(err))
    return rc, out[: tar_len.value].tobytes() if rc == 0 else b"", err.value


def oracle_bound(man) -> int:
    # generous: 3 header blocks + name-sized PAX records per entry
    n = man.n
    return int(man.payload_bytes + 512 * n + (1536 + 2048) * n + int(man.entries["name_len"].sum()) * 2 + 4096)
