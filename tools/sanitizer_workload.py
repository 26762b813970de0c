import sys, zlib, gzip
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
import helpers, oracle
from devspace_b200 import synth
from devspace_b200.manifest import from_entries, ManifestBuilder
from devspace_b200.pack import Packer
es = helpers.edge_suite()["mixed_tree"]
ref = oracle.tar_stream(es)
with Packer(level=6, segment_bytes=65536) as p:
    assert gzip.decompress(p.pack(from_entries(es))) == ref
with Packer(level=6) as p:
    for name, data in (("text", synth.text_like(700_000, 1).numpy().tobytes()), ("bin", synth.binary_like(700_000, 2).numpy().tobytes()),
                       ("rand", synth.random_bytes(200_000, 3).numpy().tobytes()), ("zeros", b"\0" * 300_000), ("tiny", b"abc")):
        raw = p.deflate(data, level=6); assert zlib.decompress(raw, -15) == data, name
    rng = np.random.default_rng(1)
    b = ManifestBuilder(arena_align=1)
    for i, s in enumerate([0, 1, 17, 4095, 65536, 65537, 300_001]):
        b.add(b"f%d" % i, s, data=rng.integers(0, 256, s, dtype=np.uint8))
    man = b.build()
    got = p.file_crc32(man)
    assert [int(c) for c in got] == [zlib.crc32(man.payload(i).tobytes()) for i in range(man.n)]
    assert p.pack(from_entries(es), tar_only=True) == ref
print("sanitizer workload ok")
