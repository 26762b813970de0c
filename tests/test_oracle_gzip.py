"""CPU: gzip framing of the oracle (compress/gzip Writer restatement) against CPython zlib/gzip and GNU gzip."""
import gzip
import shutil
import struct
import subprocess
import zlib

import pytest

import helpers
import oracle


@pytest.mark.parametrize("data", [b"", b"a", b"TestContents" * 100, helpers.rng_bytes(5, 70000),
                                  helpers.text_bytes(6, 300000)], ids=["empty", "1B", "rep", "rand70k", "text300k"])
def test_gzip_framing(data):
    g = oracle.gzip_pack(data)
    assert g[:10] == bytes.fromhex("1f8b08000000000000ff")        # flags 0, mtime 0, XFL 0, OS 255 (Go defaults)
    crc, isize = struct.unpack("<II", g[-8:])
    assert crc == zlib.crc32(data) == oracle.crc32(data)
    assert isize == len(data) & 0xFFFFFFFF
    assert zlib.decompress(g[10:-8], -15) == data                   # raw deflate body
    assert gzip.decompress(g) == data
    # body is zlib level 6: CPython's zlib (same library) gives the same size
    c = zlib.compressobj(6, zlib.DEFLATED, -15, 8)
    assert len(c.compress(data) + c.flush()) == len(g) - 18


@pytest.mark.skipif(shutil.which("gzip") is None, reason="gzip not installed")
def test_gnu_gzip_accepts(tmp_path):
    p = tmp_path / "x.gz"
    data = helpers.text_bytes(9, 100000)
    p.write_bytes(oracle.gzip_pack(data))
    assert subprocess.run(["gzip", "-t", str(p)], capture_output=True).returncode == 0
    assert subprocess.run(["gzip", "-dc", str(p)], capture_output=True).stdout == data
