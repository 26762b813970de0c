#!/usr/bin/env python
"""Deflate throughput of one resident launch by input length (the pipelined host path launches per segment)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200 import synth  # noqa: E402
from devspace_b200.pack import DeviceBuffer, Packer  # noqa: E402

with Packer(level=6) as p:
    data = synth.binary_like(2 << 30, 7).numpy()
    d_in = DeviceBuffer(p, data.size + 64).upload(data)
    d_out = DeviceBuffer(p, p.gzip_bound(data.size) + 64)
    for n in (64 << 20, 128 << 20, 256 << 20, 512 << 20, 1 << 30, 2 << 30):
        best = 1e9
        for _ in range(3):
            p.deflate_device(d_in, n, d_out)
            best = min(best, p.stats()["ms_deflate"])
        print(f"{n >> 20} MiB: {best:.1f} ms, {n / 1e9 / (best / 1e3):.2f} GB/s", flush=True)
