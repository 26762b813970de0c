"""Code-length construction (devspace_b200/csrc/huffman.cuh): the lengths must form a complete prefix code within
the DEFLATE limits (15 bits for literal/length and distance codes, 7 for the code-length code) for any histogram --
an over-subscribed set is what zlib rejects with "invalid code lengths set". CPU only."""
import ctypes

import numpy as np
import pytest

from helpers import emu_lib


def lengths(freq, max_bits):
    lib = emu_lib()
    f = np.ascontiguousarray(freq, np.uint32)
    lens = np.zeros(len(f), np.uint8)
    lib.emu_huff_lengths(ctypes.c_void_p(f.ctypes.data), ctypes.c_int(len(f)), ctypes.c_int(max_bits),
                         ctypes.c_void_p(lens.ctypes.data))
    return lens


def kraft(lens, max_bits):
    return sum(1 << (max_bits - int(l)) for l in lens if l)


@pytest.mark.parametrize("n,max_bits", [(286, 15), (30, 15), (19, 7)])
def test_complete_code_for_skewed_histograms(n, max_bits):
    rng = np.random.default_rng(n * 100 + max_bits)
    for trial in range(400):
        kind = trial % 4
        if kind == 0:      # geometric: forces the depth limit (Fibonacci-like weights)
            f = np.floor(1.6 ** rng.permutation(n).astype(np.float64)).clip(0, 2 ** 31).astype(np.uint32)
        elif kind == 1:    # a few huge, many tiny
            f = rng.integers(0, 3, n).astype(np.uint32)
            f[rng.integers(0, n, 3)] = rng.integers(10 ** 5, 10 ** 7, 3)
        elif kind == 2:    # sparse
            f = np.where(rng.random(n) < 0.15, rng.integers(1, 1000, n), 0).astype(np.uint32)
        else:
            f = rng.integers(0, 10 ** 4, n).astype(np.uint32)
        used = int((f > 0).sum())
        lens = lengths(f, max_bits)
        assert ((lens > 0) == (f > 0)).all() or used < 2
        assert int(lens.max()) <= max_bits
        if used >= 2:
            assert kraft(lens, max_bits) == 1 << max_bits, (trial, f.tolist(), lens.tolist())
            # a more frequent symbol never gets a longer code
            order = np.argsort(f[f > 0], kind="stable")
            ls = lens[f > 0][order]
            fs = f[f > 0][order]
            assert all(ls[i] >= ls[j] for i, j in zip(range(len(ls) - 1), range(1, len(ls))) if fs[i] < fs[j])


def test_degenerate_histograms():
    # zlib's build_tree forces at least two codes so that the code is complete: the same here
    for f in ([0, 0, 0, 0], [0, 7, 0, 0], [9, 0, 0, 0], [0, 0, 0, 3]):
        lens = lengths(f, 15)
        assert kraft(lens, 15) == 1 << 15 and sorted(lens.tolist())[-2:] == [1, 1], (f, lens.tolist())
        assert all(lens[i] == 1 for i, v in enumerate(f) if v)
    assert lengths([5, 0, 9, 0], 15).tolist() == [1, 0, 1, 0]
