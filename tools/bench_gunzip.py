#!/usr/bin/env python
"""dsg_gunzip: chunk-parallel inflate of this library's streams against zlib's inflate on one host
core. One JSON line per workload: end to end (pinned host in, pinned host out, wall clock), device resident (CUDA
events), bytes of tar stream per second. Usage: python tools/bench_gunzip.py"""
import ctypes
import json
import os
import sys
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from devspace_b200 import synth  # noqa: E402
from devspace_b200.manifest import Manifest  # noqa: E402
from devspace_b200.pack import DeviceBuffer, Packer  # noqa: E402


def main():
    only = os.environ.get("GUNZIP_ONLY")
    work = [("c2_node_modules_100k_files_400MB", lambda: synth.tree_c2(device="cuda")),
            ("c3_mixed_5k_files_1GB", lambda: synth.tree_c3(5000, 1_000_000_000, device="cuda")),
            ("c4_8_files_x_256MiB", lambda: synth.tree_c4(8, 256 << 20, device="cuda"))]
    with Packer(level=6) as p:
        for name, make in work:
            if only and only not in name:
                continue
            man = make()
            T = p.tar_size(man)
            blob_h = p.pinned(p.gzip_bound(T) + 64)
            n = p.pack_into(man, blob_h.array)
            out_h = p.pinned(T + 64)
            n_out = ctypes.c_size_t()
            e2e = []
            for _ in range(3):
                t0 = time.perf_counter()
                rc = p._lib.dsg_gunzip(p._ctx, ctypes.c_void_p(blob_h.array.ctypes.data), n,
                                       ctypes.c_void_p(out_h.array.ctypes.data), T, ctypes.byref(n_out))
                e2e.append(time.perf_counter() - t0)
                assert rc == 0 and n_out.value == T, (rc, n_out.value, T)
            d_in = DeviceBuffer(p, n + 64).upload(blob_h.array[:n])
            d_out = DeviceBuffer(p, T + 64)
            res = []
            for _ in range(3):
                rc = p._lib.dsg_dev_gunzip(p._ctx, d_in.ptr, n, d_out.ptr, T, ctypes.byref(n_out))
                assert rc == 0
                st = p.stats()
                res.append(st["ms_total"])
            k = min(n, 256 << 20)                                  # CPU: zlib inflate of (a prefix of) the same stream
            t0 = time.perf_counter()
            d = zlib.decompressobj(31)
            got = 0
            mv = memoryview(blob_h.array)[:n]
            for o in range(0, n, 16 << 20):
                got += len(d.decompress(mv[o:o + (16 << 20)]))
                if time.perf_counter() - t0 > 20:
                    break
            cpu_s = time.perf_counter() - t0
            same = bool(np.array_equal(d_out.download(min(T, 64 << 20)), out_h.array[: min(T, 64 << 20)]))
            print(json.dumps({"workload": name, "gzip_bytes": int(n), "tar_bytes": int(T),
                              "e2e_GBps_tar": T / 1e9 / min(e2e[1:]), "e2e_ms": min(e2e[1:]) * 1e3,
                              "resident_GBps_tar": T / 1e9 / (min(res[1:]) / 1e3), "resident_ms": min(res[1:]),
                              "ms_decode_resolve": st["ms_deflate"], "ms_crc": st["ms_crc"], "members": st["segments"],
                              "cpu_zlib_inflate_GBps_1core": got / 1e9 / cpu_s, "device_equals_host_path": same}),
                  flush=True)
            for b in (blob_h, out_h):
                b.free()
            d_in.free()
            d_out.free()


if __name__ == "__main__":
    main()
