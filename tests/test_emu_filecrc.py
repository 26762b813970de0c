"""The per-file CRC32 kernels (devspace_b200/csrc/file_crc.cuh) under the SIMT emulator against zlib.crc32. CPU only."""
import ctypes
import zlib

import numpy as np
import pytest

from devspace_b200.manifest import ENTRY_DTYPE
from helpers import emu_lib


def emu_file_crc32(ent, arena):
    lib = emu_lib()
    f = ENTRY_DTYPE.fields
    crcs = np.zeros(max(1, len(ent)), np.uint32)
    rc = lib.emu_file_crc32(ctypes.c_void_p(ent.ctypes.data), ctypes.c_uint32(len(ent)), ctypes.c_uint32(ENTRY_DTYPE.itemsize),
                            ctypes.c_uint32(f["size"][1]), ctypes.c_uint32(f["data_off"][1]),
                            ctypes.c_uint32(f["typeflag"][1]), ctypes.c_void_p(arena.ctypes.data),
                            ctypes.c_uint64(arena.size), ctypes.c_void_p(crcs.ctypes.data))
    return rc, crcs[:len(ent)]


def make(sizes, seed, gap=20, first=3):
    rng = np.random.default_rng(seed)
    offs, pos = [], first
    for s in sizes:
        offs.append(pos)
        pos += s + int(rng.integers(0, gap + 1))
    arena = rng.integers(0, 256, pos + 64, dtype=np.uint8)
    ent = np.zeros(len(sizes), ENTRY_DTYPE)
    ent["data_off"], ent["size"], ent["typeflag"] = offs, sizes, ord("0")
    return ent, arena


def expect(ent, arena):
    return [0 if e["typeflag"] == ord("5") else zlib.crc32(arena[int(e["data_off"]):int(e["data_off"]) + int(e["size"])].tobytes())
            for e in ent]


def test_edge_sizes_and_alignments():
    sizes = [0, 1, 15, 16, 17, 31, 32, 33, 100, 4095, 4096, 65535, 65536, 65537, 65536 + 16 + 5, 200000, 0, 7, 131072 + 31]
    ent, arena = make(sizes, 5)
    ent = np.concatenate([ent, np.zeros(1, ENTRY_DTYPE)])
    ent[-1]["typeflag"] = ord("5")
    rc, crcs = emu_file_crc32(ent, arena)
    assert rc == 0 and [int(c) for c in crcs] == expect(ent, arena)


def test_check_value():
    arena = np.frombuffer(b"....123456789....", np.uint8).copy()
    ent = np.zeros(1, ENTRY_DTYPE)
    ent["data_off"], ent["size"], ent["typeflag"] = 4, 9, ord("0")
    rc, crcs = emu_file_crc32(ent, arena)
    assert rc == 0 and int(crcs[0]) == 0xCBF43926


def test_many_files_cross_scan_blocks():
    rng = np.random.default_rng(9)
    sizes = [int(s) for s in rng.integers(0, 600, 2500)]       # > 2 scan blocks of 1024 files
    sizes[1024] = 70000
    ent, arena = make(sizes, 11, gap=3, first=0)
    rc, crcs = emu_file_crc32(ent, arena)
    assert rc == 0 and [int(c) for c in crcs] == expect(ent, arena)


def test_range_error():
    ent, arena = make([10, 20], 1)
    ent[1]["data_off"] = arena.size - 5
    rc, _ = emu_file_crc32(ent, arena)
    assert rc == -2          # -(1 + index of the bad entry)


def test_more_units_than_lanes():
    ent, arena = make([2_300_007, 70_000, 0, 5 * 65536], 21)      # 36 units: two per lane in the join, a full-unit tail
    rc, crcs = emu_file_crc32(ent, arena)
    assert rc == 0 and [int(c) for c in crcs] == expect(ent, arena)
