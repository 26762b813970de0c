"""GPU: CRC32, deflate and the whole dsg_pack path through the C ABI.
Judged equalities: gunzip(out) == oracle tar stream byte for byte; stock decoders accept the
output; compressed size within 3 % of the oracle's zlib level-6 size (level 6 only)."""
import ctypes
import gzip
import os
import shutil
import struct
import subprocess
import zlib

import numpy as np
import pytest

import helpers
import oracle
from devspace_b200.manifest import from_entries

pytestmark = pytest.mark.gpu
SUITE = helpers.edge_suite()


@pytest.fixture(scope="module")
def packer0():
    from devspace_b200.pack import Packer
    p = Packer(level=0)
    yield p
    p.close()


BUFS = {
    "empty": b"", "1B": b"a", "15B": b"0123456789abcde", "16B": b"0123456789abcdef", "17B": b"0123456789abcdefg",
    "rand_131071": helpers.rng_bytes(3, 131071), "rand_131072": helpers.rng_bytes(4, 131072),
    "rand_1M+5": helpers.rng_bytes(5, (1 << 20) + 5), "text_3M": helpers.text_bytes(6, 3_000_000),
    "zeros_300k": b"\0" * 300_000,
}


@pytest.mark.parametrize("name", sorted(BUFS))
def test_crc32_matches_zlib(packer0, name):
    data = BUFS[name]
    assert packer0.crc32(data) == zlib.crc32(data) == oracle.crc32(data)


def test_crc32_unaligned_device_pointer(packer0):
    from devspace_b200.pack import DeviceBuffer
    data = np.frombuffer(helpers.rng_bytes(8, 500_003), np.uint8)
    d = DeviceBuffer(packer0, data.size + 64).upload(data)
    try:
        for off in (1, 3, 7, 15, 16, 33):
            assert packer0.crc32_device(d, data.size - off - 5, offset=off) == zlib.crc32(data[off:data.size - 5].tobytes())
    finally:
        d.free()


@pytest.mark.parametrize("name", sorted(BUFS))
def test_stored_deflate_roundtrip(packer0, name):
    data = BUFS[name]
    raw = packer0.deflate(data, level=0)
    assert zlib.decompress(raw, -15) == data
    assert len(raw) == len(data) + 5 * ((len(data) + 65534) // 65535) + 2
    g = packer0.deflate(data, level=0, gzip=True)
    assert g[:10] == bytes.fromhex("1f8b08000000000000ff")
    assert struct.unpack("<II", g[-8:]) == (zlib.crc32(data), len(data) & 0xFFFFFFFF)
    assert gzip.decompress(g) == data


@pytest.mark.parametrize("name", ["util_test_tree", "empty_archive", "mixed_tree", "name_5000", "pax_multi"])
def test_pack_stored_gunzips_to_oracle_tar(packer0, name, tmp_path):
    es = SUITE[name]
    out = packer0.pack(from_entries(es))
    ref = oracle.tar_stream(es)
    assert gzip.decompress(out) == ref
    st = packer0.stats()
    assert st["tar_bytes"] == len(ref) and st["out_bytes"] == len(out) and st["crc32"] == zlib.crc32(ref)
    assert st["kernel_launches"] > 0
    if shutil.which("tar") and name != "name_5000":
        p = tmp_path / "a.tgz"
        p.write_bytes(out)
        r = subprocess.run(["tar", "tzf", str(p)], capture_output=True)
        assert r.returncode == 0 and r.stderr == b"" and len(r.stdout.split(b"\n")) - 1 == len(es)
        assert subprocess.run(["gzip", "-t", str(p)], capture_output=True).returncode == 0


def test_pack_device_resident_matches_host_path(packer0):
    from devspace_b200.pack import DeviceBuffer, DeviceManifest
    es = SUITE["mixed_tree"]
    man = from_entries(es)
    dm = DeviceManifest(packer0, man)
    d_out = DeviceBuffer(packer0, packer0.gzip_bound(packer0.tar_size(man)) + 64)
    try:
        n = packer0.pack_device(dm, d_out)
        assert gzip.decompress(d_out.download(n).tobytes()) == oracle.tar_stream(es)
    finally:
        dm.free()
        d_out.free()


# ---- level 6: LZ77 + dynamic Huffman kernels ---------------------------------------------------------------------

@pytest.fixture(scope="module")
def packer6():
    from devspace_b200.pack import Packer
    p = Packer(level=6)
    yield p
    p.close()


def _zlib6(data: bytes) -> int:
    c = zlib.compressobj(6, zlib.DEFLATED, -15, 8)
    return len(c.compress(data) + c.flush())


@pytest.mark.parametrize("name", sorted(BUFS))
def test_level6_deflate_roundtrip(packer6, name):
    data = BUFS[name]
    raw = packer6.deflate(data, level=6)
    assert zlib.decompress(raw, -15) == data
    assert raw[-2:] == b"\x03\x00"
    g = packer6.deflate(data, level=6, gzip=True)
    assert g[:10] == bytes.fromhex("1f8b08000000000000ff")
    assert struct.unpack("<II", g[-8:]) == (zlib.crc32(data), len(data) & 0xFFFFFFFF)
    assert gzip.decompress(g) == data
    assert len(raw) <= len(data) + 5 * ((len(data) + 16383) // 16384) + 2        # never worse than stored chunks


# Judged equality (3): compressed size <= 1.03 x the oracle's zlib level-6 size. Measured on the B200
# (tools/size_ratios.py, end of round 2): ELF-like binary 1.0001, BASELINE config 1 0.9991, JS-like token text 1.0212 (the
# hardest case: every 4-gram recurs hundreds of times per 32 KiB window; it needs the deep second search pass).
SIZE_CASES = {"elf_like_binary": 1.03, "c1_quickstart": 1.03, "js_like_text": 1.03}


@pytest.mark.parametrize("name", sorted(SIZE_CASES))
def test_level6_size_vs_zlib6(packer6, name):
    from devspace_b200 import synth
    if name == "js_like_text":
        data = synth.text_like(6_000_000, 21).numpy().tobytes()
    elif name == "elf_like_binary":
        data = synth.binary_like(6_000_000, 22).numpy().tobytes()
    else:
        data = gzip.decompress(open(os.path.join(os.path.dirname(__file__), "golden", "c1_quickstart.tgz"), "rb").read())
    raw = packer6.deflate(data, level=6)
    assert zlib.decompress(raw, -15) == data
    ref = _zlib6(data)
    assert len(raw) <= SIZE_CASES[name] * ref, (name, len(raw), ref, len(raw) / ref)


@pytest.mark.parametrize("name", ["util_test_tree", "empty_archive", "mixed_tree", "name_5000", "pax_multi",
                                  "single_empty_file"])
def test_level6_pack_gunzips_to_oracle_tar(packer6, name, tmp_path):
    es = SUITE[name]
    out = packer6.pack(from_entries(es))
    ref = oracle.tar_stream(es)
    assert gzip.decompress(out) == ref                      # judged equality (1): bit-exact tar stream
    st = packer6.stats()
    assert st["crc32"] == zlib.crc32(ref) and st["tar_bytes"] == len(ref) and st["kernel_launches"] > 0
    if shutil.which("tar") and name != "name_5000":        # judged equality (2): stock decoders accept it
        p = tmp_path / "a.tgz"
        p.write_bytes(out)
        r = subprocess.run(["tar", "tzf", str(p)], capture_output=True)
        assert r.returncode == 0 and r.stderr == b"" and len(r.stdout.split(b"\n")) - 1 == len(es)
        assert subprocess.run(["gzip", "-t", str(p)], capture_output=True).returncode == 0


def test_level6_pipeline_segments_share_one_member(tmp_path):
    """64 KiB segments: 15 segments for mixed_tree; the dictionary crosses segment boundaries, the CRC is
    folded across segments, and the result is still one gzip member equal to the oracle's tar stream."""
    from devspace_b200.pack import Packer
    es = SUITE["mixed_tree"]
    ref = oracle.tar_stream(es)
    with Packer(level=6, segment_bytes=65536) as p:
        out = p.pack(from_entries(es))
        st = p.stats()
    assert st["segments"] == (len(ref) + 65535) // 65536 and st["segments"] > 8
    assert gzip.decompress(out) == ref
    with Packer(level=6) as p:
        whole = p.pack(from_entries(es))
    assert abs(len(out) - len(whole)) <= 0.01 * len(whole)            # segmentation costs (almost) nothing
    with Packer(level=0, segment_bytes=131072) as p:
        assert gzip.decompress(p.pack(from_entries(es))) == ref
        assert p.pack(from_entries(es), tar_only=True) == ref


def test_level6_small_trees_size_vs_oracle(packer6):
    from devspace_b200 import synth
    for man, bound in ((synth.tree_c2(3000, 12_000_000), 1.03), (synth.tree_c3(200, 40_000_000), 1.03),
                       (synth.tree_c4(2, 6 << 20), 1.03)):
        out = packer6.pack(man)
        tar = gzip.decompress(out)
        # the tar stream itself is checked against the oracle through the generic path
        es = [oracle.Entry(name=man.name(i), size=int(e["size"]), mtime_sec=int(e["mtime_sec"]), mode=int(e["mode"]),
                           uid=1000, gid=1000, uname=b"dev", gname=b"dev", data=man.payload(i))
              for i, e in enumerate(man.entries)]
        assert tar == oracle.tar_stream(es)
        ref = len(oracle.gzip_pack(tar, 6))
        assert len(out) <= bound * ref, (len(out), ref, len(out) / ref)


def test_level6_pack_device_and_capacity_error(packer6):
    from devspace_b200._lib import DsgError
    from devspace_b200.pack import DeviceBuffer, DeviceManifest
    es = SUITE["mixed_tree"]
    man = from_entries(es)
    dm = DeviceManifest(packer6, man)
    d_out = DeviceBuffer(packer6, packer6.gzip_bound(packer6.tar_size(man)) + 64)
    small = DeviceBuffer(packer6, 4096)
    try:
        n = packer6.pack_device(dm, d_out)
        assert gzip.decompress(d_out.download(n).tobytes()) == oracle.tar_stream(es)
        with pytest.raises(DsgError) as ei:
            packer6.pack_device(dm, small)
        assert ei.value.status == -5
        out = np.empty(1000, np.uint8)
        with pytest.raises(DsgError) as ei:
            packer6.pack_into(man, out)
        assert ei.value.status == -5
    finally:
        dm.free()
        d_out.free()
        small.free()


@pytest.mark.parametrize("tar_only", [False, True])
def test_pack_to_fd_streams_the_same_bytes(tmp_path, tar_only):
    """dsg_pack_fd: the segments reach the file through the staging buffers in order; the file
    equals what dsg_pack returns, also with many small segments and with a single one."""
    from devspace_b200.pack import Packer
    es = SUITE["mixed_tree"]
    man = from_entries(es)
    ref = oracle.tar_stream(es)
    for seg in (65536, 1 << 20, 0):
        with Packer(level=6, segment_bytes=seg) as p:
            path = tmp_path / ("o%d" % seg)
            with open(path, "wb") as f:
                n = p.pack_to_fd(man, f.fileno(), tar_only=tar_only)
            st = p.stats()
            mem = p.pack(man, tar_only=tar_only)
        got = path.read_bytes()
        assert n == len(got) and st["segments"] >= 1
        assert (got if tar_only else gzip.decompress(got)) == ref
        if seg != 65536 or tar_only:
            assert got == mem
    with Packer(level=6) as p:                      # an unusable descriptor is an error, not a crash
        from devspace_b200._lib import DsgError
        r, w = os.pipe()
        os.close(r)
        import signal
        old = signal.signal(signal.SIGPIPE, signal.SIG_IGN)
        try:
            with pytest.raises(DsgError):
                p.pack_to_fd(man, w)
        finally:
            signal.signal(signal.SIGPIPE, old)
            os.close(w)
        assert gzip.decompress(p.pack(man)) == ref  # the context is still usable


def test_level6_output_is_deterministic_and_equals_the_emulator(packer6):
    """Same input, same bytes: run to run on the GPU, and GPU against the same kernel source under the CPU emulator
    (so the emulator tests of the CPU suite speak for the GPU build)."""
    from devspace_b200 import synth
    from test_emu_deflate import emu_deflate
    data = (synth.text_like(90_000, 31).numpy().tobytes() + b"\0" * 5000 + synth.binary_like(80_000, 32).numpy().tobytes()
            + synth.random_bytes(20_000, 33).numpy().tobytes())
    a = packer6.deflate(data, level=6)                      # 195 000 bytes: the host picks 32 KiB chunks
    assert a == packer6.deflate(data, level=6)
    assert zlib.decompress(a, -15) == data
    assert a == emu_deflate(data, grid=2, small_chunk_kib=32)
    packer6.set_lz_geometry(chunk_units=1)                  # 64 KiB chunks
    try:
        b = packer6.deflate(data, level=6)
    finally:
        packer6.set_lz_geometry()
    assert zlib.decompress(b, -15) == data and b == emu_deflate(data, grid=2, run_chunks=1)
    big = synth.binary_like(8_000_000, 34).numpy().tobytes()
    assert packer6.deflate(big, level=6) == packer6.deflate(big, level=6)


@pytest.mark.parametrize("units", [2, 4])
def test_level6_large_chunks(packer6, units):
    """Inputs of hundreds of MB are cut into 128 / 256 KiB chunks (one deflate block each) with 64 KiB chunks at the
    end of the launch; dsg_opts.lz_chunk_units / lz_tail_bytes force that on a small input: the stream must inflate,
    stay within the size bound, and equal the emulator's bytes."""
    from devspace_b200 import synth
    from test_emu_deflate import emu_deflate
    packer6.set_lz_geometry(chunk_units=units, tail_bytes=0)   # no 64 KiB tail chunks beyond the last big one's rest
    try:
        data = synth.text_like(3_000_000, 41).numpy().tobytes() + synth.binary_like(2_000_001, 42).numpy().tobytes()
        raw = packer6.deflate(data, level=6)
        assert zlib.decompress(raw, -15) == data
        assert len(raw) <= 1.03 * _zlib6(data)
        small = data[2_900_000:3_300_000]                  # text, then binary: 400 000 bytes, a short last chunk
        assert packer6.deflate(small, level=6) == emu_deflate(small, grid=2, run_chunks=1, chunk_units=units)
    finally:
        packer6.set_lz_geometry()


@pytest.mark.parametrize("level", [0, 6])
def test_gzip_capacity_probe_never_writes_past_the_buffer(level):
    """DSG_E_CAPACITY is the documented size probe: with a too-small device buffer dsg_dev_gzip / dsg_pack_device report
    the size they need and leave every byte behind the buffer alone (the gzip trailer used to be written at
    out + 10 + needed, far past a small buffer)."""
    from devspace_b200 import synth
    from devspace_b200._lib import DsgError
    from devspace_b200.pack import DeviceBuffer, DeviceManifest, Packer
    data = np.frombuffer(synth.text_like(400_000, 51).numpy().tobytes() + helpers.rng_bytes(52, 100_000), np.uint8)
    man = from_entries(SUITE["mixed_tree"])
    with Packer(level=level) as p:
        d_in = DeviceBuffer(p, data.size + 16).upload(data)
        dm = DeviceManifest(p, man)
        ARENA = 4 << 20                                         # the output buffer is carved out of this allocation
        d_area = DeviceBuffer(p, ARENA)
        canary = np.full(ARENA, 0xA5, np.uint8)
        try:
            whole = p.deflate(data, level=level, gzip=True)
            need = len(whole)
            whole_pack = p.pack(man)
            for fn, n_need, ref in ((lambda cap: p._lib.dsg_dev_gzip(p._ctx, d_in.ptr, data.size, level, d_area.ptr, cap,
                                                                   ctypes.byref(n_out)), need, whole),
                                    (lambda cap: p._lib.dsg_pack_device(p._ctx, dm.entries.ptr, dm.n, dm.strtab.ptr,
                                                                      dm.strtab_len, dm.arena.ptr, dm.arena_len,
                                                                      d_area.ptr, cap, ctypes.byref(n_out)),
                                     len(whole_pack), whole_pack)):
                for cap in (18, 1000, n_need - 1, n_need):
                    d_area.upload(canary)
                    n_out = ctypes.c_size_t()
                    rc = fn(cap)
                    got = d_area.download()
                    assert (got[cap:] == 0xA5).all(), (level, cap, "bytes behind the buffer were written")
                    if cap < n_need:
                        assert rc == -5 and n_out.value == n_need, (level, cap, rc, n_out.value, n_need)
                    else:
                        assert rc == 0 and n_out.value == n_need and got[:n_need].tobytes() == ref
        finally:
            for b in (d_in, d_area):
                b.free()
            dm.free()
