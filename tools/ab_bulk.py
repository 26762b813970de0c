#!/usr/bin/env python
"""A/B of the tile staging on the bench generator, device resident: cp.async.bulk + mbarrier against plain loads
(dsg_opts.lz_flags & DSG_LZ_NO_BULK). Usage: python tools/ab_bulk.py [files=8]"""
import ctypes
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from devspace_b200.pack import DeviceBuffer, Packer  # noqa: E402


def main():
    files = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    man, arena = bench.make_shard(0, files, 256 << 20, torch.device("cuda", 0))
    with Packer(level=6) as p:
        T = p.tar_size(man)
        d_e = DeviceBuffer(p, man.entries.nbytes).upload(man.entries.view(np.uint8).reshape(-1))
        d_s = DeviceBuffer(p, man.strtab.size).upload(man.strtab)
        d_o = DeviceBuffer(p, int(p.gzip_bound(T) * 0.62) + (1 << 20))
        res = {}
        for rnd in range(3):
            for name, kw in (("bulk", {}), ("plain_loads", {"no_bulk": True})):
                p.set_lz_geometry(**kw)
                n = ctypes.c_size_t()
                for _ in range(2):
                    rc = p._lib.dsg_pack_device(p._ctx, d_e.ptr, man.n, d_s.ptr, man.strtab.size,
                                                ctypes.c_void_p(arena.data_ptr()), arena.numel(), d_o.ptr, d_o.nbytes,
                                                ctypes.byref(n))
                    p._check(rc)
                res.setdefault(name, []).append(p.stats()["ms_total"])
        P = files * (256 << 20)
        print(json.dumps({k: {"ms": [round(x, 2) for x in v], "GBps_best": round(P / 1e6 / min(v), 3)} for k, v in res.items()}))


if __name__ == "__main__":
    main()
