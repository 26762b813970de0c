"""GPU: dsg_gunzip / dsg_dev_gunzip -- chunk-parallel inflate of the streams this library writes (SURVEY.md 8f rank 4;
the decompression half of untarAll, /root/reference/pkg/devspace/sync/tar.go:16-33). Judged equality: the bytes are
what a stock inflate (zlib) returns for the same stream, and that is the oracle's tar stream."""
import gzip
import io
import tarfile
import zlib

import numpy as np
import pytest

import helpers
import oracle
from devspace_b200 import synth
from devspace_b200._lib import DsgError
from devspace_b200.manifest import from_entries

pytestmark = pytest.mark.gpu
SUITE = helpers.edge_suite()


@pytest.fixture(scope="module")
def packer6():
    from devspace_b200.pack import Packer
    p = Packer(level=6)
    yield p
    p.close()


def _entries(man):
    return helpers.manifest_entries(man, 0, man.n)


@pytest.mark.parametrize("name", ["util_test_tree", "empty_archive", "mixed_tree", "single_empty_file", "pax_multi"])
def test_gunzip_of_pack_is_the_oracle_tar_stream(packer6, name):
    es = SUITE[name]
    blob = packer6.pack(from_entries(es))
    ref = oracle.tar_stream(es)
    assert zlib.decompress(blob, 31) == ref
    got = packer6.gunzip(blob)
    assert got == ref
    st = packer6.stats()
    assert st["crc32"] == zlib.crc32(ref) and st["kernel_launches"] > 0
    names = [m.name for m in tarfile.open(fileobj=io.BytesIO(got), mode="r:").getmembers()]
    assert len(names) == len(es)                               # untarNext's side: a stock tar reader takes it


@pytest.mark.parametrize("tree", ["c2", "c3", "c4"])
def test_gunzip_small_trees_every_geometry(tree):
    from devspace_b200.pack import Packer
    man = {"c2": lambda: synth.tree_c2(3000, 12_000_000), "c3": lambda: synth.tree_c3(200, 40_000_000),
           "c4": lambda: synth.tree_c4(2, 6 << 20)}[tree]()
    ref = oracle.tar_stream(_entries(man))
    for kw in ({}, {"segment_bytes": 1 << 20}, {"chunk_units": 4, "tail_bytes": 0}, {"chunk_units": 1, "run_chunks": 3}):
        with Packer(level=6, **kw) as p:
            blob = p.pack(man)
            assert p.gunzip(blob) == ref, (tree, kw)
    with Packer(level=0) as p:                                  # stored blocks only
        assert p.gunzip(p.pack(man)) == ref
    with Packer(level=6, gunzip_batch_mib=1) as p:              # many batches: references across batch boundaries
        assert p.gunzip(p.pack(man)) == ref


def test_gunzip_members_of_several_shards_and_the_stitched_member():
    from devspace_b200.multi import MultiPacker
    from devspace_b200.pack import Packer
    man = synth.tree_c3(150, 30_000_000)
    ref = oracle.tar_stream(_entries(man))
    with MultiPacker([0, 0, 0], level=6, segment_bytes=4 << 20) as mp, Packer(level=6) as p:
        members = mp.pack(man)
        assert p.gunzip(members) == ref and p.stats()["segments"] == 3        # three members back to back
        assert p.gunzip(mp.pack(man, one_member=True)) == ref


def test_gunzip_device_resident_and_capacity(packer6):
    import ctypes
    from devspace_b200.pack import DeviceBuffer
    data = synth.text_like(3_000_000, 81).numpy().tobytes() + helpers.rng_bytes(82, 500_000) + b"\0" * 200_000
    blob = np.frombuffer(packer6.deflate(data, level=6, gzip=True), np.uint8)
    d_in = DeviceBuffer(packer6, blob.size + 64).upload(blob)
    d_out = DeviceBuffer(packer6, len(data) + 64)
    try:
        n = ctypes.c_size_t()
        rc = packer6._lib.dsg_dev_gunzip(packer6._ctx, d_in.ptr, blob.size, d_out.ptr, len(data), ctypes.byref(n))
        assert rc == 0 and n.value == len(data) and d_out.download(n.value).tobytes() == data
        rc = packer6._lib.dsg_dev_gunzip(packer6._ctx, d_in.ptr, blob.size, d_out.ptr, 1000, ctypes.byref(n))
        assert rc == -5 and n.value == len(data)                 # DSG_E_CAPACITY with the size needed
    finally:
        d_in.free()
        d_out.free()


def test_gunzip_refuses_foreign_streams_and_catches_damage(packer6):
    data = synth.text_like(400_000, 91).numpy().tobytes()
    with pytest.raises(DsgError) as ei:
        packer6.gunzip(gzip.compress(data, 6, mtime=0), out_cap=len(data))   # zlib's stream: one chain of bits
    assert ei.value.status == -1
    good = bytearray(packer6.deflate(data, level=6, gzip=True))
    assert packer6.gunzip(bytes(good)) == data
    good[-6] ^= 0x01                                             # the CRC32 in the trailer
    with pytest.raises(DsgError) as ei:
        packer6.gunzip(bytes(good), out_cap=len(data))
    assert ei.value.status == -9
    good[-6] ^= 0x01
    good[len(good) // 3] ^= 0x10                                 # a bit inside a chunk: structure or checksum fails
    with pytest.raises(DsgError) as ei:
        packer6.gunzip(bytes(good), out_cap=len(data))
    assert ei.value.status in (-1, -9)
    assert packer6.gunzip(bytes(packer6.deflate(data, level=6, gzip=True))) == data   # the context is still usable
