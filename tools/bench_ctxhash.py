#!/usr/bin/env python
"""Per-file CRC32 of a build context (hash.go:165-197) on the GPU against zlib.crc32 on one host core.
One JSON line per workload: resident (dsg_dev_file_crc32, CUDA events in the library) and end to end
(dsg_file_crc32 from pinned host memory, wall clock), payload GB/s. Usage: python tools/bench_ctxhash.py [reps]"""
import json
import os
import sys
import time
import zlib

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200 import synth  # noqa: E402
from devspace_b200.pack import DeviceBuffer, DeviceManifest, Packer, PinnedArray  # noqa: E402
from devspace_b200.manifest import Manifest  # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    work = [("c2_node_modules_100k_files_400MB", lambda: synth.tree_c2()),
            ("c3_mixed_5k_files_1GB", lambda: synth.tree_c3(5000, 1_000_000_000)),
            ("c4_8_files_x_256MiB", lambda: synth.tree_c4(8, 256 << 20))]
    with Packer(level=6) as p:
        for name, make in work:
            man = make()
            P = man.payload_bytes
            # pinned copy of the arena for the end-to-end leg
            pin = PinnedArray(p, man.arena.size + 16)
            pin.array[:man.arena.size] = man.arena
            pman = Manifest(man.entries, man.strtab, pin.array[:man.arena.size])
            dm = DeviceManifest(p, man)
            d_crcs = DeviceBuffer(p, 4 * man.n + 16)
            res, e2e = [], []
            for _ in range(reps + 1):
                p.file_crc32_device(dm, d_crcs)
                res.append(p.stats()["ms_crc"])
                t0 = time.perf_counter()
                got = p.file_crc32(pman)
                e2e.append((time.perf_counter() - t0) * 1e3)
            launches = p.stats()["kernel_launches"]
            dev = d_crcs.download(4 * man.n).view(np.uint32)
            # CPU: zlib.crc32 over a bounded sample of the files (first ~256 MB), one core
            t0 = time.perf_counter()
            done, k = 0, 0
            while k < man.n and done < 256_000_000:
                pl = man.payload(k)
                c = zlib.crc32(pl) if pl.size else 0
                assert c == int(got[k]) == int(dev[k]), (name, k)
                done += pl.size
                k += 1
            cpu_s = time.perf_counter() - t0
            line = {"workload": name, "files": int(man.n), "payload_bytes": int(P),
                    "resident_GBps": P / 1e9 / (min(res[1:]) / 1e3), "resident_ms": min(res[1:]),
                    "e2e_GBps": P / 1e9 / (min(e2e[1:]) / 1e3), "e2e_ms": min(e2e[1:]), "kernel_launches": launches,
                    "cpu_zlib_crc32_GBps_1core": done / 1e9 / cpu_s, "cpu_sample_files": k, "parity_on_sample": True}
            print(json.dumps(line), flush=True)
            dm.free(), d_crcs.free(), pin.free()


if __name__ == "__main__":
    main()
