#!/usr/bin/env python
"""Docker build-context tar (archive.TarWithOptions at builder/docker/docker.go:94-97 produces an
UNCOMPRESSED tar): dsg_pack_tar_only end to end with pinned buffers (H2D + header/layout kernels + D2H, wall clock)
and device resident (dsg_dev_tar_build, CUDA events), against the CPU oracle's tar writer on one host core.
One JSON line per workload. Usage: python tools/bench_docker_tar.py [reps]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from devspace_b200 import synth  # noqa: E402
from devspace_b200.manifest import Manifest  # noqa: E402
from devspace_b200.pack import DeviceBuffer, DeviceManifest, Packer  # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    work = [("c2_node_modules_100k_files_400MB", lambda: synth.tree_c2(device="cuda")),
            ("c3_mixed_5k_files_1GB", lambda: synth.tree_c3(5000, 1_000_000_000, device="cuda")),
            ("c4_8_files_x_256MiB", lambda: synth.tree_c4(8, 256 << 20, device="cuda"))]
    with Packer(level=6) as p:
        for name, make in work:
            man = make()
            T = p.tar_size(man)
            pin = p.pinned(man.arena.size + 16)
            pin.array[: man.arena.size] = man.arena
            pman = Manifest(man.entries, man.strtab, pin.array[: man.arena.size])
            out = p.pinned(T + 64)
            e2e = []
            for _ in range(reps + 1):
                t0 = time.perf_counter()
                n = p.pack_into(pman, out.array, tar_only=True)
                e2e.append(time.perf_counter() - t0)
            dm = DeviceManifest(p, man)
            d_tar = DeviceBuffer(p, T + 64)
            res = []
            for _ in range(reps + 1):
                p.tar_build_device(dm, d_tar)
                res.append(p.stats()["ms_tar"] + p.stats()["ms_plan"])
            t0 = time.perf_counter()
            ref_sha, ref_len = helpers.oracle_tar_sha256(man)
            cpu_s = time.perf_counter() - t0
            import hashlib
            ok = n == ref_len and hashlib.sha256(out.array[:n]).hexdigest() == ref_sha
            print(json.dumps({"workload": name, "entries": man.n, "payload_bytes": man.payload_bytes, "tar_bytes": T,
                              "e2e_GBps_tar": T / 1e9 / min(e2e[1:]), "e2e_ms": min(e2e[1:]) * 1e3,
                              "resident_GBps_tar": T / 1e9 / (min(res[1:]) / 1e3), "resident_ms": min(res[1:]),
                              "cpu_oracle_tar_plus_sha256_GBps_1core": T / 1e9 / cpu_s,
                              "bit_exact_vs_oracle": bool(ok)}), flush=True)
            for b in (pin, out):
                b.free()
            dm.free()
            d_tar.free()


if __name__ == "__main__":
    main()
