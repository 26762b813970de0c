"""Build-context hash on the GPU: dsg_file_crc32 / dsg_dev_file_crc32 / dsg_directory_hash through the C ABI against
zlib.crc32 and the oracle (oracle/ctxhash.py restating pkg/util/hash/hash.go), and the golden fixture."""
import gzip
import json
import os
import zlib

import numpy as np
import pytest

import ctxhash_trees
import oracle
from oracle import ctxhash as oracle_ctxhash

pytestmark = pytest.mark.gpu
GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "ctxhash.json")))


@pytest.fixture(scope="module")
def packer():
    from devspace_b200.pack import Packer
    with Packer(level=6) as p:
        yield p


@pytest.fixture(scope="module")
def trees(tmp_path_factory):
    d = tmp_path_factory.mktemp("ctx")
    for name in ctxhash_trees.TREES:
        ctxhash_trees.materialise(str(d), name)
    return str(d)


def _expect(man):
    return [zlib.crc32(man.payload(i).tobytes()) if chr(int(e["typeflag"])) == "0" else 0 for i, e in enumerate(man.entries)]


def test_file_crc32_edge_sizes(packer):
    from devspace_b200.manifest import ManifestBuilder
    rng = np.random.default_rng(3)
    for align in (16, 1):                         # 16-byte aligned payloads and byte-packed (unaligned) ones
        b = ManifestBuilder(arena_align=align)
        for i, s in enumerate([0, 1, 15, 16, 17, 33, 4095, 4096, 65535, 65536, 65537, 65536 * 3 + 21, 1 << 20]):
            b.add(b"f%02d" % i, s, data=rng.integers(0, 256, s, dtype=np.uint8))
        b.add(b"emptydir", typeflag=ord("5"))
        man = b.build()
        got = packer.file_crc32(man)
        assert [int(c) for c in got] == _expect(man)
        st = packer.stats()
        assert st["kernel_launches"] >= 6 and st["h2d_bytes"] >= man.arena.size


def test_file_crc32_device_resident_shares_the_arena_with_pack(packer):
    """One resident arena: the per-file checksums and the archive, no second host->device pass."""
    from devspace_b200 import synth
    from devspace_b200.pack import DeviceBuffer, DeviceManifest
    man = synth.tree_c2(3000, 12_000_000)
    dm = DeviceManifest(packer, man)
    d_crcs = DeviceBuffer(packer, 4 * man.n + 16)
    d_out = DeviceBuffer(packer, packer.gzip_bound(packer.tar_size(man)) + 64)
    try:
        packer.file_crc32_device(dm, d_crcs)
        got = d_crcs.download(4 * man.n).view(np.uint32)
        assert [int(c) for c in got] == _expect(man)
        n = packer.pack_device(dm, d_out)
        tar = gzip.decompress(d_out.download(n).tobytes())
        assert len(tar) == packer.tar_size(man)
    finally:
        dm.free(), d_crcs.free(), d_out.free()


def test_file_crc32_many_tiny_and_one_large(packer):
    from devspace_b200.manifest import ManifestBuilder
    rng = np.random.default_rng(4)
    b = ManifestBuilder()
    for i in range(5000):
        s = int(rng.integers(0, 65))
        b.add(b"t/%05d" % i, s, data=rng.integers(0, 256, s, dtype=np.uint8))
    b.add(b"big", 40_000_003, data=rng.integers(0, 256, 40_000_003, dtype=np.uint8))
    man = b.build()
    assert [int(c) for c in packer.file_crc32(man)] == _expect(man)


def test_file_crc32_range_error(packer):
    from devspace_b200._lib import DsgError
    from devspace_b200.manifest import ManifestBuilder
    b = ManifestBuilder()
    b.add(b"a", 100, data=np.full(100, 120, np.uint8))
    man = b.build()
    man.entries[0]["size"] = man.arena.size + 1
    with pytest.raises(DsgError) as ei:
        packer.file_crc32(man)
    assert ei.value.status == -8


@pytest.mark.parametrize("name", sorted(ctxhash_trees.TREES))
def test_directory_hash_matches_oracle_and_golden(packer, trees, name, monkeypatch):
    from devspace_b200.ctxhash import directory_excludes
    monkeypatch.chdir(trees)
    got = directory_excludes(packer, name)
    assert got == oracle_ctxhash.directory_excludes(name) == GOLDEN[name]["hash"]
    # an absolute path hashes differently (filePath is part of the digest) but still like the oracle
    absolute = os.path.join(trees, name)
    assert directory_excludes(packer, absolute) == oracle_ctxhash.directory_excludes(absolute) != got
    assert directory_excludes(packer, name + "/./") == got          # filepath.Join(srcPath, ".") cleans the root


def test_directory_hash_excludes(packer, trees, monkeypatch):
    from devspace_b200.ctxhash import directory_excludes
    monkeypatch.chdir(trees)
    got = directory_excludes(packer, "util_like", ctxhash_trees.util_like_exclude)
    assert got == GOLDEN["util_like/excluded"]["hash"]
    skip = lambda rel, d: 1 if rel == "./testFolder" else 0  # noqa: E731
    assert directory_excludes(packer, "util_like", skip) == oracle_ctxhash.directory_excludes("util_like", skip)


def test_directory_hash_errors(packer, tmp_path):
    from devspace_b200._lib import DsgError
    from devspace_b200.ctxhash import directory_excludes
    (tmp_path / "f").write_bytes(b"x")
    with pytest.raises(DsgError) as ei:
        directory_excludes(packer, str(tmp_path / "f"))
    assert "is not a directory" in str(ei.value)
    with pytest.raises(DsgError):
        directory_excludes(packer, str(tmp_path / "missing"))
    (tmp_path / "empty").mkdir()
    assert directory_excludes(packer, str(tmp_path / "empty")) == oracle_ctxhash.directory_excludes(str(tmp_path / "empty"))
