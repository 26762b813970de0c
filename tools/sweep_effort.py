#!/usr/bin/env python
"""Development aid: speed and size of the level-6 kernel under other search efforts (deep depth, refine rounds,
shallow depth -- the upper bits of dsg_opts.lz_flags), device resident, on the generators of devspace_b200.synth.
The shipped library compiles the effort in; the knobs exist only in a library built with -DDSG_LZ_RUNTIME_EFFORT
(tools/final_batch.sh shows how one is built next to the tree and swapped in for this tool alone).
Usage: python tools/sweep_effort.py [MiB=512] [configs, e.g. r1d32,r3,s4: rounds / deep depth / shallow depth]"""
import json
import os
import sys
import zlib

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200 import synth  # noqa: E402
from devspace_b200.pack import DeviceBuffer, Packer  # noqa: E402


def effort(deep=0, rounds=0, depth=0):
    return (deep // 4) | (rounds << 8) | (depth << 10)


def main():
    mib = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    n = mib << 20
    gens = {"bin": synth.binary_like(n, 3), "text": synth.text_like(n, 1)}
    zl = {}
    for k, t in gens.items():
        sample = t[: 32 << 20].numpy().tobytes()
        c = zlib.compressobj(6, zlib.DEFLATED, -15, 8)
        zl[k] = len(c.compress(sample) + c.flush()) / len(sample)
    configs = [("default", {})]
    for spec in (sys.argv[2].split(",") if len(sys.argv) > 2 else []):  # e.g. r1d32 = 1 round, deep depth 32; r3 = no round
        kw = {}
        for part in spec.replace("r", " r").replace("d", " d").replace("s", " s").split():
            kw[{"r": "rounds", "d": "deep", "s": "depth"}[part[0]]] = int(part[1:])
        configs.append((spec, kw))
    with Packer(level=6) as p:
        bufs = {k: DeviceBuffer(p, n + 64).upload(t.numpy()) for k, t in gens.items()}
        d_out = DeviceBuffer(p, n + (n >> 8) + (1 << 20))
        for name, kw in configs:
            p.set_lz_geometry(effort=effort(**kw))
            row = {"config": name}
            for k in gens:
                best = 1e9
                for _ in range(3):
                    out = p.deflate_device(bufs[k], n, d_out)
                    best = min(best, p.stats()["ms_deflate"])
                row[k] = {"GBps": round(n / 1e6 / best, 2), "ratio_vs_zlib6": round(out / n / zl[k], 4)}
            print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
