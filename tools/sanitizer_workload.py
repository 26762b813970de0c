"""Workload for compute-sanitizer (memcheck / racecheck / synccheck / initcheck): every kernel of libdsgpack.so once,
the level-6 kernel in each of its geometries (16 / 32 / 64 / 128 / 256 KiB chunks, runs of chunks, the stored exit with
the window rebuild behind it, dictionary from a previous segment), small inputs so that a tool that slows the GPU a
hundredfold still ends in minutes. Output summaries are kept under profiles/.
  compute-sanitizer --tool racecheck python tools/sanitizer_workload.py"""
import gzip
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
import oracle  # noqa: E402
from devspace_b200 import synth  # noqa: E402
from devspace_b200.manifest import ManifestBuilder, from_entries  # noqa: E402
from devspace_b200.multi import MultiPacker  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402

quick = "--quick" in sys.argv
es = helpers.edge_suite()["mixed_tree"]
ref = oracle.tar_stream(es)
with Packer(level=6, segment_bytes=65536) as p:          # pipelined path, 15 segments, dictionary across segments
    assert gzip.decompress(p.pack(from_entries(es))) == ref
with Packer(level=6) as p:
    assert gzip.decompress(p.pack(from_entries(es))) == ref                      # small-archive path (one sync)
    text = synth.text_like(300_000, 1).numpy().tobytes()
    mixed = (text[:70_000] + synth.random_bytes(140_000, 3).numpy().tobytes() + text[70_000:160_000] + b"\0" * 40_000
             + synth.binary_like(90_000, 2).numpy().tobytes())
    for geo in ({}, {"chunk_units": 1, "run_chunks": 3, "tail_bytes": 0}, {"chunk_units": 2, "tail_bytes": 0},
                {"chunk_units": 4, "tail_bytes": 65536}, {"chunk_units": 1, "grid": 2}):
        if quick and geo.get("chunk_units") in (2, 4):
            continue
        p.set_lz_geometry(**geo)
        for name, data in (("mixed", mixed), ("tiny", b"abc"), ("text40k", text[:40_000])):
            raw = p.deflate(data, level=6)
            assert zlib.decompress(raw, -15) == data, (geo, name)
    p.set_lz_geometry()
    rng = np.random.default_rng(1)
    b = ManifestBuilder(arena_align=1)
    for i, s in enumerate([0, 1, 17, 4095, 65536, 65537, 300_001]):
        b.add(b"f%d" % i, s, data=rng.integers(0, 256, s, dtype=np.uint8))
    man = b.build()
    got = p.file_crc32(man)
    assert [int(c) for c in got] == [zlib.crc32(man.payload(i).tobytes()) for i in range(man.n)]
    assert p.pack(from_entries(es), tar_only=True) == ref
    assert p.crc32(mixed) == zlib.crc32(mixed)
with Packer(level=0) as p:
    assert gzip.decompress(p.pack(from_entries(es))) == ref
with MultiPacker([0, 0], level=6, segment_bytes=131072) as mp:
    assert gzip.decompress(mp.pack(from_entries(es), one_member=True)) == ref
print("sanitizer workload ok")
