#!/usr/bin/env python
"""Stage timings on one GPU (CUDA events inside libdsgpack, read through dsg_get_stats): tar plan/layout,
CRC32, deflate. Development aid for the HBM-bound kernels; bench.py is the contract benchmark."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200.manifest import ENTRY_DTYPE, Manifest  # noqa: E402
from devspace_b200.pack import DeviceBuffer, DeviceManifest, Packer  # noqa: E402


def synth(nfiles, size, seed=1):
    rng = np.random.default_rng(seed)
    ent = np.zeros(nfiles, ENTRY_DTYPE)
    names = [b"dir%04d/file%07d.bin" % (i // 100, i) for i in range(nfiles)]
    strtab = b"dev" + b"".join(names)
    offs = np.cumsum([3] + [len(x) for x in names[:-1]]).astype(np.uint32)
    ent["name_off"] = offs
    ent["name_len"] = [len(x) for x in names]
    ent["uname_off"] = 0
    ent["gname_off"] = 0
    ent["uname_len"] = 3
    ent["gname_len"] = 3
    ent["typeflag"] = ord("0")
    ent["mode"] = 0o644
    ent["uid"] = ent["gid"] = 1000
    ent["mtime_sec"] = 1546300800 + np.arange(nfiles)
    ent["size"] = size
    stride = (size + 15) & ~15
    ent["data_off"] = np.arange(nfiles, dtype=np.uint64) * np.uint64(stride)
    arena = rng.integers(0, 256, max(16, nfiles * stride), dtype=np.uint8)
    return Manifest(ent, np.frombuffer(strtab, np.uint8).copy(), arena)


def main():
    level = int(os.environ.get("DSG_LEVEL", "0"))
    with Packer(level=level) as p:
        for nfiles, size in ((2048, 512 * 1024), (200_000, 0), (100_000, 4096)):
            man = synth(nfiles, size)
            dm = DeviceManifest(p, man)
            T = p.tar_size(man)
            d_tar = DeviceBuffer(p, T + 64)
            d_out = DeviceBuffer(p, p.gzip_bound(T) + 64)
            res = {"nfiles": nfiles, "size": size, "tar_bytes": T, "level": level}
            for it in range(3):
                p.tar_build_device(dm, d_tar)
                st = p.stats()
            res["tar_plan_ms"], res["tar_layout_ms"] = st["ms_plan"], st["ms_tar"]
            res["tar_layout_GBps_rw"] = (man.payload_bytes + T) / 1e6 / max(st["ms_tar"], 1e-6)
            for it in range(3):
                p.crc32_device(d_tar, T)
                st = p.stats()
            res["crc_ms"] = st["ms_crc"]
            res["crc_GBps"] = T / 1e6 / max(st["ms_crc"], 1e-6)
            for it in range(3):
                n = p.deflate_device(d_tar, T, d_out, level)
                st = p.stats()
            res["deflate_ms"] = st["ms_deflate"]
            res["deflate_in_GBps"] = T / 1e6 / max(st["ms_deflate"], 1e-6)
            res["deflate_out_bytes"] = n
            t0 = time.time()
            for it in range(3):
                n = p.pack_device(dm, d_out)
                st = p.stats()
            res["pack_device_ms_total"] = st["ms_total"]
            res["pack_device_GBps_payload"] = man.payload_bytes / 1e6 / max(st["ms_total"], 1e-6)
            res["launches"] = st["kernel_launches"]
            print(json.dumps(res), flush=True)
            for b in (d_tar, d_out):
                b.free()
            dm.free()


if __name__ == "__main__":
    main()
