#!/usr/bin/env python
"""CPU/GPU crossover sweep (BASELINE config 5): the mixed-source generator of config 3 at total sizes
64 KB .. 256 MB; end-to-end GPU time (dsg_pack, pageable->pinned staging excluded: buffers are pinned) against the
CPU oracle (tar + zlib level 6, one thread) on the same bytes. Also runs the full-size configs 2, 3 and 5a/5b once
and checks size-independent properties (gunzip CRC/ISIZE via the trailer, tar stream length, member count)."""
import gzip
import json
import os
import struct
import sys
import time
import zlib

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402  (measurement tool: the CPU side of the comparison)
from devspace_b200 import synth  # noqa: E402
from devspace_b200.manifest import Manifest  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402


def oracle_entries(man):
    return [oracle.Entry(name=man.name(i), size=int(e["size"]), mtime_sec=int(e["mtime_sec"]), mode=int(e["mode"]),
                         uid=int(e["uid"]), gid=int(e["gid"]), uname=b"dev", gname=b"dev", data=man.payload(i))
            for i, e in enumerate(man.entries)]


def time_gpu(p, man, reps=5):
    h_in = p.pinned(max(16, man.arena.size))
    h_in.array[: man.arena.size] = man.arena
    T = p.tar_size(man)
    h_out = p.pinned(p.gzip_bound(T) + 64)
    m = Manifest(man.entries, man.strtab, h_in.array[: man.arena.size])
    for _ in range(2):
        n = p.pack_into(m, h_out.array)
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        n = p.pack_into(m, h_out.array)
        ts.append(time.perf_counter() - t0)
    out = h_out.array[:n].tobytes() if n < (64 << 20) else None
    st = p.stats()
    trailer = bytes(h_out.array[n - 8:n])
    h_in.free()
    h_out.free()
    return min(ts), n, out, st, trailer


def main():
    full = "--full" in sys.argv
    with Packer(level=6) as p:
        print(json.dumps({"what": "crossover sweep: config-3 generator, e2e dsg_pack (pinned) vs CPU oracle (1 thread)"}))
        for total in (64 << 10, 256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20):
            nfiles = max(4, min(2000, total // 40_000))
            man = synth.tree_c3(nfiles, total, seed=synth.SEED0 + 33)
            t_gpu, n, out, st, _ = time_gpu(p, man)
            es = oracle_entries(man)
            t0 = time.perf_counter()
            tar = oracle.tar_stream(es)
            ref_n = oracle.gzip_pack_size(np.frombuffer(tar, np.uint8))
            t_cpu = time.perf_counter() - t0
            ok = out is None or gzip.decompress(out) == tar
            print(json.dumps({"payload_bytes": man.payload_bytes, "files": man.n, "gpu_e2e_ms": round(t_gpu * 1e3, 3),
                              "cpu_1thread_ms": round(t_cpu * 1e3, 3), "speedup": round(t_cpu / t_gpu, 2),
                              "gzip_bytes": n, "zlib6_bytes": ref_n, "size_ratio": round(n / ref_n, 4),
                              "bit_exact_tar": ok, "launches": st["kernel_launches"]}), flush=True)
        if not full:
            return
        dev = "cuda"
        for name, make in (("C2 node_modules-like 100k x 4KB", lambda: synth.tree_c2(device=dev)),
                           ("C3 mixed source 20k files 4GB", lambda: synth.tree_c3(device=dev)),
                           ("C5a 1M empty/tiny files", lambda: synth.tree_c5a()),
                           ("C5b 8GB random in 32 files", lambda: synth.tree_c5b(device=dev))):
            t0 = time.perf_counter()
            man = make()
            t_gen = time.perf_counter() - t0
            t_gpu, n, out, st, trailer = time_gpu(p, man, reps=2)
            crc, isize = struct.unpack("<II", trailer)
            T = st["tar_bytes"]
            rec = {"config": name, "entries": man.n, "payload_bytes": man.payload_bytes, "tar_bytes": T,
                   "gzip_bytes": n, "gpu_e2e_ms": round(t_gpu * 1e3, 2),
                   "e2e_GBps_payload": round(man.payload_bytes / 1e9 / t_gpu, 3),
                   "e2e_GBps_tar": round(T / 1e9 / t_gpu, 3), "segments": st["segments"], "gen_s": round(t_gen, 1),
                   "isize_matches_tar_len": isize == (T & 0xFFFFFFFF), "crc_in_trailer_matches_stats": crc == st["crc32"],
                   "tar_len_matches_host_plan": T == p.tar_size(man)}
            print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
