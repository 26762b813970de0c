import sys, os, zlib
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np
import fuzz_deflate as fz
from devspace_b200.pack import Packer
rng = np.random.default_rng(3)
def z6(d):
    co = zlib.compressobj(6, zlib.DEFLATED, -15); return len(co.compress(d) + co.flush())
with Packer(level=6) as p:
    for name, data in (("const", b"\x07" * 1_000_000), ("period37", (bytes(range(37)) * 30000)[:1_000_000]),
                       ("period3", (b"abc" * 400000)[:1_000_000]), ("sym4", rng.integers(0, 4, 1_000_000, dtype=np.uint8).tobytes()),
                       ("sym16", rng.integers(0, 16, 1_000_000, dtype=np.uint8).tobytes()),
                       ("period300", (rng.integers(0,256,300,dtype=np.uint8).tobytes() * 4000)[:1_000_000])):
        raw = p.deflate(data, level=6); assert zlib.decompress(raw, -15) == data
        print(name, len(raw), z6(data), round(len(raw) / z6(data), 3), flush=True)
