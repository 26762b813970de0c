"""CPU: the chunk-parallel gunzip kernels (devspace_b200/csrc/inflate.cuh) under the SIMT emulator, on streams made by
the emulated deflate kernel and on hand-built streams of the same structure; the real parity tests are the -m gpu ones
(tests/test_gpu_gunzip.py). Reference behaviour: gzip.NewReader in untarAll, tar.go:16-33."""
import ctypes
import gzip
import struct
import zlib

import numpy as np
import pytest

import helpers
from test_emu_deflate import emu_deflate, realistic_text

HDR = bytes.fromhex("1f8b08000000000000ff")


def member(raw_deflate: bytes, data: bytes) -> bytes:
    return HDR + raw_deflate + struct.pack("<II", zlib.crc32(data), len(data) & 0xFFFFFFFF)


def emu_gunzip(blob: bytes, cap=None, rounds=48, batch=0):
    L = helpers.emu_lib()
    a = np.frombuffer(blob, np.uint8)
    cap = len(blob) * 60 + 1024 if cap is None else cap
    out = np.full(max(cap, 1), 0xAA, np.uint8)
    n, nu, nm, nc = ctypes.c_uint64(), ctypes.c_uint32(), ctypes.c_uint32(), ctypes.c_uint32()
    rc = L.emu_gunzip(ctypes.c_void_p(a.ctypes.data), ctypes.c_uint64(a.size), ctypes.c_void_p(out.ctypes.data),
                      ctypes.c_uint64(cap), ctypes.byref(n), ctypes.byref(nu), ctypes.byref(nm), ctypes.byref(nc),
                      ctypes.c_uint32(rounds), ctypes.c_uint64(batch))
    return rc, out[: n.value].tobytes() if rc == 0 else b"", n.value, nu.value, nm.value, nc.value


CASES = {
    "text_150k": realistic_text(150_000, 41),                                      # three chunks, window references
    "mixed": realistic_text(40_000, 42) + helpers.rng_bytes(43, 70_000) + b"\0" * 30_000 + realistic_text(30_000, 44),
    "rand_140k": helpers.rng_bytes(45, 140_000),                                    # stored chunks (several blocks each)
    "tiny": b"abc",
    "empty": b"",
    "repeat_across_chunks": (realistic_text(20_000, 46) * 9)[:170_000],             # copies of copies of the window
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_emulated_gunzip_of_emulated_deflate(name):
    data = CASES[name]
    for kw in ({}, {"small_chunk_kib": 16}, {"run_chunks": 2}):
        if kw and len(data) < 50_000:
            continue
        blob = member(emu_deflate(data, **kw), data)
        assert gzip.decompress(blob) == data
        rc, out, n, nunits, nmembers, ncand = emu_gunzip(blob)
        assert rc == 0 and out == data and nmembers == 1, (name, kw, rc, n, len(data))
        for rounds in (0, 1):                     # the serial kernel takes what the parallel rounds leave open
            assert emu_gunzip(blob, rounds=rounds)[1] == data, (name, kw, rounds)
        assert emu_gunzip(blob, batch=70_000)[1] == data     # batches of one or two units: references across batches
        if len(data) > 70_000:
            assert nunits >= 2


def test_emulated_gunzip_multi_member_and_level0_structure():
    a, b = realistic_text(70_000, 51), helpers.rng_bytes(52, 30_000)
    # second member in the level-0 structure: stored blocks of up to 65535 bytes + the 03 00 terminator
    stored = b"".join(b"\0" + struct.pack("<HH", len(p), len(p) ^ 0xFFFF) + p for p in (b[:20_000], b[20_000:])) + b"\x03\0"
    blob = member(emu_deflate(a), a) + member(stored, b) + member(b"\x03\0", b"")
    assert gzip.decompress(blob) == a + b
    rc, out, n, nunits, nmembers, _ = emu_gunzip(blob)
    assert rc == 0 and out == a + b and nmembers == 3


def test_emulated_gunzip_refuses_foreign_and_damaged_streams():
    data = realistic_text(90_000, 61)
    assert emu_gunzip(gzip.compress(data, 6, mtime=0))[0] == -1          # zlib's stream: one chain of bits, refused
    good = member(emu_deflate(data), data)
    assert emu_gunzip(good[:-9])[0] == -1                                  # truncated trailer
    bad = bytearray(good)
    bad[len(bad) // 2] ^= 0x40                                             # a flipped bit inside a chunk
    rc, out, *_ = emu_gunzip(bytes(bad))
    assert rc != 0 or out != data                                          # structure error, or a CRC the caller catches
    rc, _, n, *_ = emu_gunzip(good, cap=1000)
    assert rc == -5 and n == len(data)                                     # DSG_E_CAPACITY with the size needed


def test_marker_bytes_inside_payload_do_not_derail_the_chain():
    # stored payload full of byte patterns that look like unit starts: sync markers, stored headers, gzip headers
    rnd = bytearray(helpers.rng_bytes(70, 140_000))                       # incompressible: ends up in stored blocks
    rng = np.random.default_rng(7)
    pats = [b"\0\0\xff\xff", b"\0\x05\0\xfa\xff", HDR + b"\x03\0", b"\0\0\xff\xff\x04"]
    for o in rng.integers(0, len(rnd) - 16, 1500):
        pat = pats[int(o) % len(pats)]
        rnd[int(o): int(o) + len(pat)] = pat
    trap = bytes(rnd)
    # the level-0 structure: the payload sits in the stream byte for byte, behind a compressed member
    stored = b"".join(b"\0" + struct.pack("<HH", len(p), len(p) ^ 0xFFFF) + p
                      for p in (trap[o:o + 65535] for o in range(0, len(trap), 65535))) + b"\x03\0"
    text = realistic_text(30_000, 71)
    blob = member(emu_deflate(text), text) + member(stored, trap)
    data = text + trap
    assert gzip.decompress(blob) == data
    rc, out, n, nunits, nmembers, ncand = emu_gunzip(blob)
    assert rc == 0 and out == data and nmembers == 2
    assert ncand > nunits + 100                                            # many false candidates were decoded and ignored
