#!/usr/bin/env python
"""ONE call, N GPUs, one ordered archive (dsg_multi_pack_fd): wall clock of the whole call -- sharding, every device's
pipelined pack, and the concatenating (or stitching) write of the members to one file descriptor in shard order --
on the BASELINE config-4 generator, `files_per_gpu` files of 256 MiB per device (weak scaling). One JSON line per N.
Usage: python tools/multi_pack_bench.py [files_per_gpu=8] [N ...]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from devspace_b200 import synth  # noqa: E402
from devspace_b200._lib import load  # noqa: E402
from devspace_b200.multi import MultiPacker  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402

FILE_BYTES = 256 << 20


def main():
    fpg = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    ndev = load().dsg_device_count()
    ns = [int(a) for a in sys.argv[2:]] or [n for n in (1, 2, 4, 8) if n <= ndev]
    nmax = max(ns)
    files = nmax * fpg
    sizes = np.full(files, FILE_BYTES, np.int64)
    offs, alen = synth._layout(sizes)
    with Packer(level=6) as p0:
        arena = p0.pinned(alen)                       # cudaHostAllocPortable: every device copies from it
        host = torch.from_numpy(arena.array)
        for i in range(files):
            host[int(offs[i]): int(offs[i]) + FILE_BYTES].copy_(synth.binary_like(FILE_BYTES, synth.SEED0 + 4 + 31 * i, "cuda"))
        torch.cuda.synchronize()
        man_all = synth.build_manifest(synth.tree_c4_names(files), sizes, arena.array, offs)
        sink = os.open("/dev/null", os.O_WRONLY)
        for n in ns:
            man = man_all.slice(0, n * fpg)
            with MultiPacker(list(range(n)), level=6) as mp:
                mp.pack_to_fd(man.slice(0, n), sink)                 # warm-up: contexts, buffers, kernels
                mp.pack_to_fd(man, sink)
                res = {}
                for mode, one in (("members", False), ("one_member", True)):
                    t0 = time.perf_counter()
                    written = mp.pack_to_fd(man, sink, one_member=one)
                    res[mode] = (time.perf_counter() - t0, written)
                st = mp.stats()
            P = n * fpg * FILE_BYTES
            print(json.dumps({"what": "dsg_multi_pack_fd, one call, wall clock incl. the ordered write to one fd (/dev/null)",
                              "n_gpus": n, "files": n * fpg, "payload_bytes": P,
                              "members_GBps": P / 1e9 / res["members"][0], "members_s": round(res["members"][0], 3),
                              "one_member_GBps": P / 1e9 / res["one_member"][0],
                              "one_member_s": round(res["one_member"][0], 3), "gzip_bytes": res["members"][1],
                              "per_shard_ms_total": [round(s["ms_total"], 1) for s in st]}), flush=True)
        os.close(sink)
        arena.free()


if __name__ == "__main__":
    main()
