#!/usr/bin/env python
"""writeTar's temp file: pack into a pinned buffer and write it afterwards (before) against dsg_pack_fd, which writes
every segment while the GPU works on the next ones. One JSON line."""
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200 import synth  # noqa: E402
from devspace_b200.manifest import Manifest  # noqa: E402
from devspace_b200.pack import Packer, PinnedArray  # noqa: E402

files = int(sys.argv[1]) if len(sys.argv) > 1 else 16
man = synth.tree_c4(files, 256 << 20)
with Packer(level=6) as p, tempfile.TemporaryDirectory() as d:
    pin = PinnedArray(p, man.arena.size + 16)
    pin.array[:man.arena.size] = man.arena
    pman = Manifest(man.entries, man.strtab, pin.array[:man.arena.size])
    out = PinnedArray(p, p.gzip_bound(p.tar_size(man)) + 64)
    res = {"workload": f"{files} x 256 MiB ELF-like files", "payload_bytes": int(man.payload_bytes)}
    for rep in range(3):
        path = os.path.join(d, "a.tgz")
        t0 = time.perf_counter()
        n = p.pack_into(pman, out.array)
        t1 = time.perf_counter()
        with open(path, "wb") as f:
            f.write(out.array[:n].data)
        t2 = time.perf_counter()
        os.unlink(path)
        with open(path, "wb") as f:
            t3 = time.perf_counter()
            m = p.pack_to_fd(pman, f.fileno())
            t4 = time.perf_counter()
        assert os.path.getsize(path) == m and abs(m - n) < 1e-4 * n   # (chain insertion races: sizes differ by bytes)
        os.unlink(path)
        res.update(pack_ms=(t1 - t0) * 1e3, write_after_ms=(t2 - t1) * 1e3, pack_then_write_ms=(t2 - t0) * 1e3,
                   pack_fd_ms=(t4 - t3) * 1e3, out_bytes=int(n))
    res["GBps_pack_then_write"] = res["payload_bytes"] / 1e9 / (res["pack_then_write_ms"] / 1e3)
    res["GBps_pack_fd"] = res["payload_bytes"] / 1e9 / (res["pack_fd_ms"] / 1e3)
    print(json.dumps(res))
