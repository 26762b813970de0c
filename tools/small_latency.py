#!/usr/bin/env python
"""Latency of small archives (the dev-loop change sets, BASELINE config 1): wall clock of dsg_pack with pinned buffers
and the CUDA-event stage times of the same call, for the quickstart tree (5 files, 93 KB) and a few small sizes."""
import gzip
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import helpers  # noqa: E402
from devspace_b200 import synth  # noqa: E402
from devspace_b200.manifest import Manifest, from_entries  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402


def c1_manifest():
    t = gzip.decompress(open(os.path.join(helpers.ROOT, "tests", "golden", "c1_quickstart.tgz"), "rb").read())
    es = [helpers.F(m.name, d, mode=m.mode, uid=m.uid, gid=m.gid, uname=m.uname, gname=m.gname, mtime=m.mtime)
          for m, d in helpers.decode_with_tarfile(t)]
    return from_entries(es), t


def main():
    with Packer(level=6) as p:
        cases = [("C1 quickstart (5 files, 93 KB)", c1_manifest()[0])]
        for total in (16 << 10, 64 << 10, 256 << 10, 1 << 20, 4 << 20):
            cases.append((f"config-3 generator, {total >> 10} KB", synth.tree_c3(max(4, total // 40_000), total, seed=77)))
        for name, man in cases:
            h_in = p.pinned(max(16, man.arena.size))
            h_in.array[: man.arena.size] = man.arena
            m = Manifest(man.entries, man.strtab, h_in.array[: man.arena.size])
            h_out = p.pinned(p.gzip_bound(p.tar_size(man)) + 64)
            for _ in range(5):
                n = p.pack_into(m, h_out.array)
            ts = []
            for _ in range(30):
                t0 = time.perf_counter()
                n = p.pack_into(m, h_out.array)
                ts.append(time.perf_counter() - t0)
            st = p.stats()
            ok = gzip.decompress(h_out.array[:n].tobytes()) is not None
            print(json.dumps({"case": name, "payload_bytes": man.payload_bytes, "tar_bytes": st["tar_bytes"],
                              "gzip_bytes": n, "wall_ms_min": round(min(ts) * 1e3, 3),
                              "wall_ms_median": round(sorted(ts)[len(ts) // 2] * 1e3, 3),
                              "event_ms_total": round(st["ms_total"], 3), "ms_plan": round(st["ms_plan"], 3),
                              "ms_tar": round(st["ms_tar"], 3), "ms_crc": round(st["ms_crc"], 3),
                              "ms_deflate": round(st["ms_deflate"], 3), "launches": st["kernel_launches"],
                              "gunzip_ok": ok}), flush=True)
            h_in.free()
            h_out.free()


if __name__ == "__main__":
    main()
