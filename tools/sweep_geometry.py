#!/usr/bin/env python
"""Development aid: the level-6 kernel on ONE pipeline segment's worth of the bench generator under other launch
geometries (run length, tail bytes, chunk size): how much of a segment's time is the end of the launch.
Usage: python tools/sweep_geometry.py [MiB=512] [runs=1,2,3,4,6,8] [tail MiB=auto,0,9,19,37,74] [chunk units of 64 KiB=4]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200 import synth  # noqa: E402
from devspace_b200.pack import DeviceBuffer, Packer  # noqa: E402


def main():
    mib = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    n = mib << 20
    data = synth.binary_like(n, 3)
    with Packer(level=6) as p:
        d_in = DeviceBuffer(p, n + 64).upload(data.numpy())
        d_out = DeviceBuffer(p, n + (n >> 8) + (1 << 20))
        units = int(sys.argv[4]) if len(sys.argv) > 4 else 4
        rows = [("auto", {})]
        runs = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 2, 3, 4, 6, 8]
        tails = [None if x == "auto" else int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [None, 0, 9, 19, 37, 74]
        for rc in runs:
            for tail_mib in tails:
                kw = {"run_chunks": rc, "chunk_units": units}
                if tail_mib is not None:
                    kw["tail_bytes"] = tail_mib << 20
                rows.append((f"run{rc}_tail{'auto' if tail_mib is None else tail_mib}", kw))
        for name, kw in rows:
            p.set_lz_geometry(**kw)
            best = 1e9
            for _ in range(3):
                out = p.deflate_device(d_in, n, d_out)
                best = min(best, p.stats()["ms_deflate"])
            print(json.dumps({"geometry": name, "ms": round(best, 2), "GBps": round(n / 1e6 / best, 3), "out": out}), flush=True)


if __name__ == "__main__":
    main()
