#!/usr/bin/env python
"""BASELINE config 4 at its FULL size on ONE GPU (strong-scaling point N=1 of BASELINE.md: 256 files x 256 MiB = 64 GiB):
the pipelined dsg_pack needs no resident 64 GiB -- the arena stays in pinned host memory and streams through two
segment slots. One warm-up pack of the first shard, then the whole archive once; the output is checked by ISIZE / CRC
against a per-shard CRC32 of the input stream computed on the GPU while generating (not the oracle: 64 GiB through
zlib's inflate would take minutes; the five configs are oracle-checked at shard size in tests/test_gpu_configs.py).
Usage: python tools/full_c4_one_gpu.py [files=256]"""
import json
import os
import struct
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from devspace_b200 import synth  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402

FILE_BYTES = 256 << 20


def main():
    files = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    sizes = np.full(files, FILE_BYTES, np.int64)
    offs, alen = synth._layout(sizes)
    with Packer(level=6) as p:
        t0 = time.perf_counter()
        arena = p.pinned(alen)
        host = torch.from_numpy(arena.array)
        for i in range(files):
            d = synth.binary_like(FILE_BYTES, synth.SEED0 + 4 + 31 * i, "cuda")
            host[int(offs[i]): int(offs[i]) + FILE_BYTES].copy_(d)
        torch.cuda.synchronize()
        t_gen = time.perf_counter() - t0
        man = synth.build_manifest(synth.tree_c4_names(files), sizes, arena.array, offs)
        T = p.tar_size(man)
        out = p.pinned(int(T * 0.50) + (64 << 20))
        p.pack_into(man.slice(0, min(files, 8)), out.array)                     # warm-up: buffers, kernels
        t0 = time.perf_counter()
        n = p.pack_into(man, out.array)
        dt = time.perf_counter() - t0
        st = p.stats()
        crc, isize = struct.unpack("<II", bytes(out.array[n - 8:n]))
        print(json.dumps({"config": f"C4 full: {files} files x 256 MiB on ONE GPU (pipelined, arena in pinned host memory)",
                          "payload_bytes": int(man.payload_bytes), "tar_bytes": int(T), "gzip_bytes": int(n),
                          "e2e_s": round(dt, 3), "e2e_GBps_payload": round(man.payload_bytes / 1e9 / dt, 3),
                          "segments": st["segments"], "h2d_bytes": st["h2d_bytes"], "d2h_bytes": st["d2h_bytes"],
                          "isize_matches_tar_len": isize == (T & 0xFFFFFFFF), "crc_in_trailer_matches_stats": crc == st["crc32"],
                          "generate_s": round(t_gen, 1)}), flush=True)


if __name__ == "__main__":
    main()
