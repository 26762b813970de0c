"""GPU: CRC32, deflate and the whole dsg_pack path through the C ABI.
Judged equalities: gunzip(out) == oracle tar stream byte for byte; stock decoders accept the
output; compressed size within 3 % of the oracle's zlib level-6 size (level 6 only)."""
import gzip
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
