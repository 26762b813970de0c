#!/usr/bin/env python
"""Development aid: what the deep search gains per listed position on the page kinds of synth.binary_like (the
numbers behind the gate of the refine rounds). LZ profile lines go to stderr. Usage: python tools/gain_kinds.py [MiB=64]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from devspace_b200 import synth  # noqa: E402
from devspace_b200.pack import Packer  # noqa: E402


def pages(kind: int, n: int) -> bytes:
    """binary_like with one page kind only: generate 4x and keep the pages of that kind (same generator, same seed)."""
    g = synth._gen(3, "cpu")
    npages = (4 * n + 4095) // 4096
    k = torch.randint(0, 4, (npages,), generator=g)
    full = synth.binary_like(4 * n, 3).view(-1, 4096)[: npages]
    sel = full[(k == kind).nonzero().flatten()[: n // 4096]]
    return sel.contiguous().view(-1).numpy().tobytes()


def main():
    n = (int(sys.argv[1]) if len(sys.argv) > 1 else 64) << 20
    eff = int(os.environ.get("LZ_EFFORT", "0"))
    with Packer(level=6, profile=True) as p:
        if eff:
            p.set_lz_geometry(profile=True, effort=eff)
        for name, data in (("text", synth.text_like(n, 1).numpy().tobytes()), ("bin", synth.binary_like(n, 3).numpy().tobytes()),
                           ("zero_heavy", pages(1, n)), ("records", pages(2, n)), ("strings", pages(3, n))):
            sys.stderr.write(f"--- {name} {len(data)} bytes\n")
            sys.stderr.flush()
            raw = p.deflate(data, level=6)
            sys.stderr.write(f"    out={len(raw)}\n")


if __name__ == "__main__":
    main()
