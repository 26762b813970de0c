#!/usr/bin/env python
"""Randomised round trips through the level-6 kernel on the GPU: mixed inputs of random sizes, random chunk-size /
tail overrides, every output inflated with zlib and compared; sizes against zlib level 6. Usage: fuzz_deflate.py [n]"""
import os
import sys
import time
import zlib

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200 import synth  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402


def piece(rng, n, seed):
    kind = int(rng.integers(0, 6))
    if kind == 0:
        return synth.text_like(n, seed).numpy().tobytes()
    if kind == 1:
        return synth.binary_like(n, seed).numpy().tobytes()
    if kind == 2:
        return synth.random_bytes(n, seed).numpy().tobytes()
    if kind == 3:
        return bytes([int(rng.integers(0, 256))]) * n
    if kind == 4:
        unit = rng.integers(0, 256, int(rng.integers(1, 40)), dtype=np.uint8).tobytes()
        return (unit * (n // len(unit) + 1))[:n]
    return rng.integers(0, 4, n, dtype=np.uint8).tobytes()


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    rng = np.random.default_rng(int(os.environ.get("FUZZ_SEED", "1")))
    worst = 0.0
    t0 = time.time()
    with Packer(level=6) as p:
        for c in range(cases):
            parts = []
            for k in range(int(rng.integers(1, 5))):
                n = int(rng.choice([0, 1, 3, 100, 4095, 65536, 65537, 200_000, 1_000_000, 3_000_000]))
                n = max(0, n + int(rng.integers(-3, 4)))
                parts.append(piece(rng, n, 1000 * c + k))
            data = b"".join(parts)
            geo = {}
            if rng.random() < 0.5:
                geo = dict(chunk_units=int(rng.choice([1, 2, 4])), tail_bytes=int(rng.choice([0, 65536, 300_000, 10 ** 9])),
                           run_chunks=int(rng.choice([1, 2, 3, 8])))
            if rng.random() < 0.3:
                geo["grid"] = int(rng.choice([1, 2, 7, 100]))
            p.set_lz_geometry(**geo)
            raw = p.deflate(data, level=6)
            assert zlib.decompress(raw, -15) == data, ("round trip", c, len(data), geo)
            if c % 3 == 0:   # the downstream direction on the same stream, in small batches now and then
                g = p.deflate(data, level=6, gzip=True)
                p.set_lz_geometry(gunzip_batch_mib=int(rng.choice([0, 1])), **geo)
                assert p.gunzip(g, out_cap=len(data)) == data, ("gunzip", c, len(data), geo)
            co = zlib.compressobj(6, zlib.DEFLATED, -15)
            ref = len(co.compress(data) + co.flush())
            if len(data) > 100_000:
                worst = max(worst, len(raw) / ref)
            assert len(raw) <= len(data) + 5 * ((len(data) + 16383) // 16384) + 2, ("bound", c)
    print(f"fuzz ok: {cases} cases in {time.time() - t0:.0f} s, worst size ratio vs zlib-6 on inputs > 100 KB: {worst:.4f}")


if __name__ == "__main__":
    main()
