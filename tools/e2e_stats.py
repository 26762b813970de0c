import sys, time, json
sys.path.insert(0, '/root/repo')
import numpy as np
from devspace_b200 import synth
from devspace_b200.pack import Packer, PinnedArray, DeviceManifest, DeviceBuffer
from devspace_b200.manifest import Manifest
man = synth.tree_c4(8, 256 << 20)
for seg in (256 << 20, 512 << 20, 1 << 30):
    with Packer(level=6, segment_bytes=seg) as p:
        pin = PinnedArray(p, man.arena.size + 16); pin.array[:man.arena.size] = man.arena
        pman = Manifest(man.entries, man.strtab, pin.array[:man.arena.size])
        out = PinnedArray(p, p.gzip_bound(p.tar_size(man)) + 64)
        for i in range(3):
            t0 = time.perf_counter(); n = p.pack_into(pman, out.array); dt = time.perf_counter() - t0
        st = p.stats()
        print(seg >> 20, "MiB segments: wall ms", round(dt*1e3,1), {k: round(v,1) if isinstance(v,float) else v for k,v in st.items() if k.startswith("ms_") or k in ("segments","kernel_launches")}, flush=True)
        dm = DeviceManifest(p, man); d_out = DeviceBuffer(p, p.gzip_bound(p.tar_size(man)) + 64)
        for i in range(2): p.pack_device(dm, d_out)
        st = p.stats(); print("   resident:", {k: round(v,1) if isinstance(v,float) else v for k,v in st.items() if k.startswith("ms_")}, flush=True)
        dm.free(); d_out.free(); pin.free(); out.free()
