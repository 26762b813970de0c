#!/usr/bin/env python
"""Runs the level-6 kernel several times on the same inputs and reports whether the output bytes are identical."""
import hashlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200 import synth  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
with Packer(level=6) as p:
    for name, data in (("text", synth.text_like(n, 1).numpy().tobytes()), ("bin", synth.binary_like(n, 3).numpy().tobytes()),
                       ("zeros+text", b"\0" * 300_000 + synth.text_like(1_000_000, 5).numpy().tobytes() * 3)):
        hs = set()
        sizes = set()
        for _ in range(4):
            z = p.deflate(data, level=6)
            hs.add(hashlib.sha256(z).hexdigest())
            sizes.add(len(z))
        print(name, "identical" if len(hs) == 1 else f"{len(hs)} different outputs", sorted(sizes), flush=True)
