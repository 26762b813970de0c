"""Builds libdsgpack.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdsgpack.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-O2", "-shared", "-diag-suppress", "550,177"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libdsgpack.so cannot be built (there is no CPU fallback)")


def sources():
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".cc", ".h", ".inl"))]
    srcs.append(os.path.join(os.path.dirname(HERE), "include", "dsgpack.h"))
    return srcs


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in sources())


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB
    cus = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cu") and not f.startswith("bench_")]
    ccs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cc")]
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", LIB, *cus, *ccs]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    r = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


def build_microbench() -> str:
    """Stand-alone calibration binary (shared-memory atomics / match.any / random LDS rates)."""
    out = os.path.join(HERE, "bench_micro")
    src = os.path.join(CSRC, "bench_micro.cu")
    if os.path.exists(src) and (not os.path.exists(out) or os.path.getmtime(src) > os.path.getmtime(out)):
        subprocess.run([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-o",
                        out, src], check=True, cwd=CSRC)
    return out


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
