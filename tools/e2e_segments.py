#!/usr/bin/env python
"""Development aid: end-to-end pack (pinned host in, pinned host out) of a bench shard under several pipeline
segment sizes, with the library's stage times. Usage: python tools/e2e_segments.py [files=16] [seg MiB ...]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402


def main():
    files = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    segs = [int(x) for x in sys.argv[2:]] or [0, 64, 128, 256, 512, 1024]
    man, arena = bench.make_shard(0, files, 256 << 20, torch.device("cuda", 0))
    P = files * (256 << 20)
    # the same shard resident in HBM (dsg_pack_device), for the stage times without copies
    import ctypes
    from devspace_b200.pack import DeviceBuffer
    with Packer(level=6) as p:
        T = p.tar_size(man)
        d_e = DeviceBuffer(p, man.entries.nbytes).upload(man.entries.view(np.uint8).reshape(-1))
        d_s = DeviceBuffer(p, man.strtab.size).upload(man.strtab)
        d_o = DeviceBuffer(p, int(p.gzip_bound(T) * 0.62) + (1 << 20))
        n = ctypes.c_size_t()
        for _ in range(3):
            p._check(p._lib.dsg_pack_device(p._ctx, d_e.ptr, man.n, d_s.ptr, man.strtab.size, ctypes.c_void_p(arena.data_ptr()),
                                            arena.numel(), d_o.ptr, d_o.nbytes, ctypes.byref(n)))
            st = p.stats()
        print(json.dumps({"resident": True, "GBps": round(P / 1e6 / st["ms_total"], 3), **{k: round(v, 1) for k, v in st.items() if k.startswith("ms_")},
                          "segments": st["segments"], "launches": st["kernel_launches"]}), flush=True)
        d_o.free()
    for seg in segs:
        with Packer(level=6, segment_bytes=seg << 20) as p:
            T = p.tar_size(man)
            h_arena = p.pinned(arena.numel())
            torch.from_numpy(h_arena.array).copy_(arena)
            torch.cuda.synchronize()
            h_out = p.pinned(int(p.gzip_bound(T) * 0.62) + (1 << 20))
            mh = type(man)(man.entries, man.strtab, h_arena.array)
            best = None
            for _ in range(3):
                t0 = time.perf_counter()
                n = p.pack_into(mh, h_out.array)
                dt = time.perf_counter() - t0
                st = p.stats()
                if best is None or dt < best[0]:
                    best = (dt, st)
            dt, st = best
            print(json.dumps({"segment_MiB": seg, "e2e_GBps": round(P / 1e9 / dt, 3), "wall_ms": round(dt * 1e3, 1),
                              "ms_total": round(st["ms_total"], 1), "ms_deflate": round(st["ms_deflate"], 1),
                              "ms_tar": round(st["ms_tar"], 1), "ms_crc": round(st["ms_crc"], 1),
                              "ms_gather": round(st["ms_gather"], 1), "segments": st["segments"],
                              "launches": st["kernel_launches"]}), flush=True)
            del h_arena, h_out


if __name__ == "__main__":
    main()
