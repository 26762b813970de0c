#!/usr/bin/env python
"""Development aid: deflate buffers on the GPU, inflate on the host, report the first chunk that breaks."""
import os
import sys
import zlib

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200 import synth  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402


def check(p, name, data):
    raw = p.deflate(data, level=6)
    d = zlib.decompressobj(-15)
    try:
        back = d.decompress(raw)
        ok = back == data
        pos = None
        if not ok:
            n = min(len(back), len(data))
            a, b = np.frombuffer(back[:n], np.uint8), np.frombuffer(data[:n], np.uint8)
            diff = np.nonzero(a != b)[0]
            pos = int(diff[0]) if diff.size else n
    except zlib.error as e:
        ok = False
        # find how far the stream decodes
        d = zlib.decompressobj(-15)
        got = 0
        try:
            for i in range(0, len(raw), 4096):
                got += len(d.decompress(raw[i:i + 4096]))
        except zlib.error:
            pass
        pos = got
        name += f" ({e})"
    exp_chunks = (len(data) + 65535) // 65536
    print(f"{name}: in={len(data)} out={len(raw)} chunks={exp_chunks} ok={ok}" +
          ("" if ok else f" first_bad_byte~{pos} (chunk {pos // 65536})"), flush=True)
    return ok


def main():
    sizes = [int(x) for x in (sys.argv[1:] or ["1000000", "20000000"])]
    # LZ_PROF=1: per-phase cycle counts of the kernel on stderr (dsg_opts.lz_flags & DSG_LZ_PROFILE)
    with Packer(level=6, profile=bool(os.environ.get("LZ_PROF")), chunk_units=int(os.environ.get("LZ_UNITS", "0"))) as p:
        kinds = os.environ.get("LZ_KINDS", "text,rand,bin").split(",")  # which generators
        for n in sizes:
            if "text" in kinds:
                check(p, f"text[{n}]", synth.text_like(n, 1).numpy().tobytes())
            if "rand" in kinds:
                check(p, f"rand[{n}]", synth.random_bytes(n, 2).numpy().tobytes())
            if "bin" in kinds:
                check(p, f"bin[{n}]", synth.binary_like(n, 3).numpy().tobytes())


if __name__ == "__main__":
    main()
