"""CPU: the LZ77 + Huffman deflate kernel and the CRC kernels (devspace_b200/csrc/deflate_lz.cuh, crc32.cuh)
compiled for the SIMT emulator. Checks kernel LOGIC without a GPU: the stream must inflate to the input with a
stock decoder (zlib) and stay close to zlib level 6 in size. Inputs are small: the emulator runs every CUDA
thread as a fiber. The real parity tests are the -m gpu ones."""
import ctypes
import zlib

import numpy as np
import pytest

import helpers

lib = None


def _lib():
    global lib
    if lib is None:
        lib = helpers.emu_lib()
        lib.emu_crc32.restype = ctypes.c_uint32
    return lib


def emu_deflate(data: bytes, grid: int = 0, dict_: bytes = b"", run_chunks: int = 1, chunk_units: int = 1,
                small_chunk_kib: int = 0):
    L = _lib()
    a = np.frombuffer(data, np.uint8)
    d = np.frombuffer(dict_, np.uint8)
    cap = len(data) + 10 * ((len(data) + 16383) // 16384) + 128
    out = np.zeros(cap, np.uint8)
    n = ctypes.c_uint64()
    rc = L.emu_lz_deflate(ctypes.c_void_p(a.ctypes.data if len(data) else None), ctypes.c_uint64(len(data)),
                          ctypes.c_void_p(d.ctypes.data if len(dict_) else None), ctypes.c_uint32(len(dict_)),
                          ctypes.c_void_p(out.ctypes.data), ctypes.c_uint64(cap), ctypes.byref(n), ctypes.c_uint32(grid),
                          ctypes.c_uint32(run_chunks | (chunk_units << 16) | ((small_chunk_kib // 16) << 24)))
    assert rc == 0
    return out[: n.value].tobytes()


def zlib6(data: bytes) -> int:
    c = zlib.compressobj(6, zlib.DEFLATED, -15, 8)
    return len(c.compress(data) + c.flush())


def realistic_text(n: int, seed: int) -> bytes:
    from devspace_b200 import synth
    return synth.text_like(n, seed).numpy().tobytes()


CASES = {
    "empty": b"",
    "3B": b"abc",
    "rep600": b"abcabcabcabc" * 50,
    "zeros_70k": b"\0" * 70_000,                      # two chunks, maximal matches, distance 1
    "rand_66k": helpers.rng_bytes(2, 66_000),          # stored chunks (two stored blocks for a full chunk)
    "text_20k": helpers.text_bytes(1, 20_000),
    "text_140k": realistic_text(140_000, 3),           # three chunks: dictionary carried across chunk boundaries
    "mixed_100k": realistic_text(30_000, 4) + helpers.rng_bytes(5, 40_000) + b"\0" * 30_000,
    "chunk_exact": realistic_text(65_536, 6),
    "chunk_plus1": realistic_text(65_537, 7),
    "tile_edges": (b"0123456789abcdef" * 600)[:8192 - 3] + helpers.rng_bytes(8, 7) + b"0123456789abcdef" * 700,
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_emulated_deflate_inflates_with_zlib(name):
    data = CASES[name]
    z = emu_deflate(data)
    assert zlib.decompress(z, -15) == data
    assert z[-2:] == b"\x03\x00"                                   # final empty fixed block
    if name.startswith("text") or name.startswith("chunk"):
        assert len(z) <= 1.12 * zlib6(data)                        # sanity only; the 3 % bar is judged on the GPU
    if name == "zeros_70k":
        assert len(z) < 400                                        # ~150 B per 64 KiB chunk: tables + sync marker


def test_emulated_deflate_persistent_grid_and_dictionary():
    data = realistic_text(200_000, 9)
    one = emu_deflate(data, grid=1)            # one CTA walks all four chunks through the ticket counter
    two = emu_deflate(data, grid=3)
    assert one == two                           # chunk outputs do not depend on which CTA ran them
    slid = emu_deflate(data, run_chunks=3)      # one sliding window across three chunks, then a run of one
    assert zlib.decompress(slid, -15) == data
    assert abs(len(slid) - len(one)) <= 0.005 * len(one)
    assert zlib.decompress(one, -15) == data
    # a segment that continues a member: the previous 32 KiB are handed in as the dictionary
    head, tail = data[:100_000], data[100_000:]
    z = emu_deflate(tail, dict_=head[-32768:])
    d = zlib.decompressobj(-15, zdict=head[-32768:])
    assert d.decompress(z) == tail
    assert len(z) < len(emu_deflate(tail))      # the dictionary is really used


def test_emulated_incompressible_is_stored_with_small_overhead():
    data = helpers.rng_bytes(10, 131_072)
    z = emu_deflate(data)
    assert zlib.decompress(z, -15) == data
    assert len(z) == len(data) + 2 * 10 + 2     # two stored blocks per 64 KiB chunk + terminator


@pytest.mark.parametrize("n", [0, 1, 15, 16, 17, 4095, 131072, 131073, 300_001])
def test_emulated_crc32(n):
    data = helpers.rng_bytes(n + 1, n)
    a = np.frombuffer(data, np.uint8)
    got = _lib().emu_crc32(ctypes.c_void_p(a.ctypes.data if n else None), ctypes.c_uint64(n))
    assert got == zlib.crc32(data)


def test_emulated_crc32_unaligned_pointer():
    data = np.frombuffer(helpers.rng_bytes(3, 200_000), np.uint8)
    for off in (1, 5, 9, 13):
        got = _lib().emu_crc32(ctypes.c_void_p(data.ctypes.data + off), ctypes.c_uint64(150_000))
        assert got == zlib.crc32(data[off:off + 150_000].tobytes())


@pytest.mark.parametrize("sched", ["reverse", "random"])
def test_emulated_deflate_under_other_thread_schedules(sched):
    """The emulator normally runs the threads of a block in thread-id order, which hides code that relies on that
    order (atomics "arriving" in stream order, a missing barrier). EMU_SCHED is read once per process: subprocess."""
    import os
    import subprocess
    import sys
    code = (
        "import sys, zlib; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import test_emu_deflate as t\n"
        "for name in ('mixed_100k', 'text_140k', 'tile_edges', 'zeros_70k'):\n"
        "    d = t.CASES[name]; z = t.emu_deflate(d, grid=2, run_chunks=2)\n"
        "    assert zlib.decompress(z, -15) == d, name\n"
        "    assert len(z) <= 1.05 * t.zlib6(d) + 512, (name, len(z))\n"
        "    assert z == t.emu_deflate(d, grid=1, run_chunks=2), name\n"
        "    import hashlib; print(name, hashlib.sha256(z).hexdigest())\n"
        "print('ok')\n" % (helpers.ROOT, os.path.join(helpers.ROOT, "tests")))
    env = dict(os.environ, EMU_SCHED=sched)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stderr[-2000:]
    # The output must not depend on the order in which the threads run (lap-stamped hash heads, chain starts of run
    # interiors taken after the insertion): the same bytes as under the default thread-id order.
    import hashlib
    for line in r.stdout.strip().split("\n")[:-1]:
        name, digest = line.split()
        assert hashlib.sha256(emu_deflate(CASES[name], grid=2, run_chunks=2)).hexdigest() == digest, (sched, name)


def test_emulated_stored_exit_keeps_the_window_and_sees_copies():
    """Chunks of flat bytes leave through the stored exit after their first tiles; the chunk behind one rebuilds its
    window from the stream, and a copy of earlier random bytes (visible in the first tiles) still compresses."""
    rnd = helpers.rng_bytes(21, 65_536 * 2)
    text = realistic_text(40_000, 22)
    # random chunk, random chunk, then text whose first 32 KiB repeat the text that precedes the random bytes
    data = text[:32_768] + rnd[: 65_536 * 2 - 32_768] + text[:32_768] + text[32_768:]
    for rc in (1, 3):
        z = emu_deflate(data, run_chunks=rc, grid=1)
        assert zlib.decompress(z, -15) == data
    # a duplicated random block: the second copy starts a chunk, so its first tiles match the window -> not stored
    blk = helpers.rng_bytes(23, 20_000)
    dup = helpers.rng_bytes(24, 65_536 - 20_000) + blk + blk + helpers.rng_bytes(25, 30_000)
    z = emu_deflate(dup, run_chunks=2, grid=1)
    assert zlib.decompress(z, -15) == dup
    assert len(z) < len(dup) - 15_000                 # the second copy costs (almost) nothing
    # text right behind a stored chunk in the same run still uses matches into that chunk's tail (rebuilt window)
    tail = text[:20_000]
    mix = helpers.rng_bytes(26, 65_536 - 20_000) + tail + tail + realistic_text(30_000, 27)
    z = emu_deflate(mix, run_chunks=2, grid=1)
    assert zlib.decompress(z, -15) == mix
    assert len(z) < 65_536 + 5 * 2 + 0.45 * 50_000


@pytest.mark.parametrize("kib", [16, 32])
def test_emulated_small_chunks(kib):
    """Small inputs are cut into 16 / 32 KiB chunks (one deflate block each, the 32 KiB dictionary unchanged)."""
    data = realistic_text(70_001, 31) + helpers.rng_bytes(32, 20_000) + b"\0" * 9_000
    z = emu_deflate(data, small_chunk_kib=kib, grid=2)
    assert zlib.decompress(z, -15) == data
    assert z == emu_deflate(data, small_chunk_kib=kib, grid=1)
    assert len(z) <= 1.04 * len(emu_deflate(data))            # the price of more block headers
    rnd = helpers.rng_bytes(33, 40_000)
    z = emu_deflate(rnd, small_chunk_kib=kib)
    assert zlib.decompress(z, -15) == rnd and len(z) <= len(rnd) + 5 * ((len(rnd) + kib * 1024 - 1) // (kib * 1024)) + 2


def test_emulated_sizes_stay_within_three_percent_of_zlib6():
    """The judged bound (1.03 x zlib level 6) on the generators of the bench and of the hardest size case, with the
    kernel's real search effort, on the CPU: the deep pass with its second work list (the positions behind matches
    that grew) has to keep token-like text where two full refine rounds had it (1.0206 on 1.5 MB, 1.0185 on the GPU's
    512 MiB) -- one round alone is 1.035."""
    from devspace_b200 import synth
    text = synth.text_like(600_000, 21).numpy().tobytes()
    mix = synth.binary_like(600_000, 22).numpy().tobytes()
    for name, data, bound in (("text", text, 1.03), ("binary mix", mix, 1.01)):
        z = emu_deflate(data, grid=4, run_chunks=2)
        assert zlib.decompress(z, -15) == data
        assert len(z) <= bound * zlib6(data), (name, len(z), zlib6(data))


def test_emulated_deflate_random_mixtures_round_trip():
    """The fuzz of tools/fuzz_deflate.py on the CPU, at sizes around the tile (1 KiB), window-batch (8 KiB), small-chunk
    (16 / 32 KiB) and chunk (64 KiB) edges: mixtures of text, page mix, random bytes, byte runs, short periods and 2-bit
    symbols, random run lengths and chunk sizes; every stream inflates with zlib and stays near zlib level 6."""
    from devspace_b200 import synth
    rng = np.random.default_rng(7)

    def piece(n, seed):
        kind = int(rng.integers(0, 6))
        if kind == 0:
            return synth.text_like(n, seed).numpy().tobytes()
        if kind == 1:
            return synth.binary_like(n, seed).numpy().tobytes()
        if kind == 2:
            return synth.random_bytes(n, seed).numpy().tobytes()
        if kind == 3:
            return bytes([int(rng.integers(0, 256))]) * n
        if kind == 4:
            unit = rng.integers(0, 256, int(rng.integers(1, 40)), dtype=np.uint8).tobytes()
            return (unit * (n // len(unit) + 1))[:n]
        return rng.integers(0, 4, n, dtype=np.uint8).tobytes()

    edges = [0, 1, 3, 100, 1023, 1024, 1025, 2048, 8191, 8192, 8193, 16384, 32768, 40961, 65536, 65537]
    for c in range(40):
        parts = []
        for k in range(int(rng.integers(1, 4))):
            n = max(0, int(rng.choice(edges)) + int(rng.integers(-2, 3)))
            parts.append(piece(n, 100 * c + k))
        data = b"".join(parts)
        kw = {}
        if rng.random() < 0.4:
            kw = dict(run_chunks=int(rng.choice([1, 2, 3])), grid=int(rng.choice([1, 2, 5])))
        elif rng.random() < 0.3 and len(data) > 16384:
            kw = dict(small_chunk_kib=int(rng.choice([16, 32])))
        z = emu_deflate(data, **kw)
        assert zlib.decompress(z, -15) == data, (c, len(data), kw)
        if len(data) > 50_000:
            assert len(z) <= 1.7 * zlib6(data) + 512, (c, len(data), len(z), zlib6(data))
