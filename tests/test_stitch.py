"""dsg_stitch_members: one gzip member out of the shards' members. The CPU test builds members the way
the library does (non-final byte-aligned blocks + the 03 00 terminator) with zlib; the GPU test stitches real shards."""
import gzip
import struct
import zlib

import numpy as np
import pytest

from devspace_b200 import shard
from devspace_b200._lib import DsgError

HDR = bytes.fromhex("1f8b08000000000000ff")


def member(data: bytes) -> bytes:
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    body = c.compress(data) + c.flush(zlib.Z_SYNC_FLUSH) if data else b""
    return HDR + body + b"\x03\x00" + struct.pack("<II", zlib.crc32(data), len(data) & 0xFFFFFFFF)


def one_member(blob: bytes) -> bytes:
    d = zlib.decompressobj(31)
    out = d.decompress(blob)
    assert d.eof and d.unused_data == b""            # exactly one member, nothing behind it
    return out


def test_stitch_zlib_made_members():
    rng = np.random.default_rng(1)
    parts = [b"hello world " * 1000, rng.integers(0, 256, 70_001, dtype=np.uint8).tobytes(), b"", b"x", b"tail" * 33]
    ms = [member(p) for p in parts]
    for m, p in zip(ms, parts):
        assert gzip.decompress(m) == p               # each piece is a valid member by itself
    st = shard.stitch_members(ms, [len(p) for p in parts])
    assert one_member(st) == b"".join(parts)
    assert st[:10] == HDR and len(st) == sum(len(m) - 18 for m in ms) - 2 * (len(ms) - 1) + 18
    assert struct.unpack("<II", st[-8:]) == (zlib.crc32(b"".join(parts)), len(b"".join(parts)))
    assert shard.stitch_members(ms[:1], [len(parts[0])]) == ms[0]


def test_stitch_rejects_foreign_input():
    good = member(b"abc")
    with pytest.raises(DsgError):
        shard.stitch_members([gzip.compress(b"abc"), good], [3, 3])      # header with mtime / final block inside
    with pytest.raises(DsgError):
        shard.stitch_members([good, good], [3, 4])                       # tar_lens does not match ISIZE
    with pytest.raises(DsgError):
        shard.stitch_members([good[:15]], [3])


@pytest.mark.gpu
@pytest.mark.parametrize("level", [6, 0])
def test_stitch_real_shards(level):
    import oracle
    from devspace_b200 import synth
    from devspace_b200.pack import Packer
    man = synth.tree_c2(1200, 5_000_000)
    members, tar_lens = [], []
    world = 3
    for r in range(world):
        sub, flags = shard.rank_slice(man, r, world)
        with Packer(level=level, flags=flags, segment_bytes=1 << 20) as p:
            members.append(p.pack(sub))
            tar_lens.append(p.stats()["tar_bytes"])
    import helpers
    whole = oracle.tar_stream(helpers.manifest_entries(man, 0, man.n))     # the oracle, not the GPU's own whole pack
    st = shard.stitch_members(members, tar_lens)
    assert one_member(st) == whole == gzip.decompress(shard.concat_members(members))
