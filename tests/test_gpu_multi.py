"""GPU: dsg_multi_* -- one call, N shards, one ordered archive (the reference packs an archive with
one writeTar call, upstream.go:342). Checked against the oracle's tar stream. With one GPU the shards are several
contexts on that GPU (the same code path); with two or more GPUs every shard runs on a device of its own."""
import gzip
import os
import zlib

import pytest

import helpers
import oracle
from devspace_b200 import synth
from devspace_b200._lib import DsgError, load

pytestmark = pytest.mark.gpu


def _ngpu():
    return load().dsg_device_count()


def _one_member(blob: bytes) -> bytes:
    d = zlib.decompressobj(31)
    out = d.decompress(blob)
    assert d.eof and d.unused_data == b""
    return out


@pytest.mark.parametrize("level", [6, 0])
@pytest.mark.parametrize("devices", [[0], [0, 0, 0], "all"])
def test_multi_pack_matches_the_oracle(devices, level, tmp_path):
    from devspace_b200.multi import MultiPacker
    if devices == "all":
        if _ngpu() < 2:
            pytest.skip("needs two or more GPUs")
        devices = list(range(_ngpu()))
    man = synth.tree_c3(150, 30_000_000)
    ref = oracle.tar_stream(helpers.manifest_entries(man, 0, man.n))
    with MultiPacker(devices, level=level, segment_bytes=4 << 20) as mp:
        members = mp.pack(man)                                   # members back to back
        assert gzip.decompress(members) == ref
        sha, n, count = helpers.gunzip_sha256(members)
        assert count == len(devices) and n == len(ref)
        one = mp.pack(man, one_member=True)                      # stitched on the fly
        assert _one_member(one) == ref
        path = tmp_path / "a.tgz"
        with open(path, "wb") as f:
            written = mp.pack_to_fd(man, f.fileno(), one_member=True)
        assert written == len(one) and path.read_bytes() == one
        st = mp.stats()
        assert len(st) == len(devices) and sum(s["tar_bytes"] for s in st) == len(ref)
        assert all(s["kernel_launches"] > 0 for s in st)


def test_multi_pack_reports_errors_and_small_buffers():
    from devspace_b200.multi import MultiPacker
    import ctypes
    import numpy as np
    from devspace_b200.manifest import from_entries
    with pytest.raises(DsgError):
        MultiPacker([0, 99])                                      # a device that does not exist: no fallback
    man = synth.tree_c2(300, 1_500_000)
    with MultiPacker([0, 0]) as mp:
        out = np.empty(1000, np.uint8)
        n_out = ctypes.c_size_t()
        rc = mp._lib.dsg_multi_pack(mp._h, man.entries.ctypes.data, man.n, man.strtab.ctypes.data, man.strtab.size,
                                    man.arena.ctypes.data, man.arena.size, 0, out.ctypes.data, out.size,
                                    ctypes.byref(n_out))
        assert rc == -5 and n_out.value == len(mp.pack(man))     # DSG_E_CAPACITY with the size that is needed
        bad = from_entries(helpers.ERROR_CASES["name_with_nul"])
        with pytest.raises(DsgError):
            mp.pack(bad)
        assert gzip.decompress(mp.pack(man)) == oracle.tar_stream(helpers.manifest_entries(man, 0, man.n))


def test_multi_pack_more_shards_than_entries():
    """Four shards, two entries: the empty shards write empty members (header, empty final block, trailer); the stream
    still gunzips to the oracle's tar stream, with a stock decoder and with dsg_gunzip."""
    from devspace_b200.manifest import from_entries
    from devspace_b200.multi import MultiPacker
    from devspace_b200.pack import Packer
    es = [helpers.F("a.txt", b"TestContents" * 5000), helpers.F("b.bin", helpers.rng_bytes(3, 70_000))]
    ref = oracle.tar_stream(es)
    with MultiPacker([0, 0, 0, 0]) as mp, Packer() as p:
        for one in (False, True):
            blob = mp.pack(from_entries(es), one_member=one)
            assert gzip.decompress(blob) == ref
            assert p.gunzip(blob) == ref
    with MultiPacker([0, 0]) as mp:
        assert gzip.decompress(mp.pack(from_entries([]))) == oracle.tar_stream([])
