"""File-range sharding across GPUs: contiguous entry ranges balanced by tar-stream bytes, one
complete gzip member per shard, the last shard carries the tar trailer, the host concatenates the members in
shard order. No data-path collective exists: ranks never exchange payload."""
from __future__ import annotations

import ctypes
from typing import List, Tuple

import numpy as np

from . import _lib
from .manifest import Manifest


def shard_bounds(man: Manifest, nshards: int) -> np.ndarray:
    """dsg_shard_bounds: shard s owns entries [b[s], b[s+1]). Pure host integer work (no GPU needed)."""
    lib = _lib.load()
    b = np.zeros(nshards + 1, np.uint64)
    rc = lib.dsg_shard_bounds(man.entries.ctypes.data if man.n else None, man.n, man.strtab.ctypes.data,
                              man.strtab.size, nshards, b.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)))
    if rc != 0:
        raise _lib.DsgError(rc, lib.dsg_strerror(rc).decode())
    return b


def rank_slice(man: Manifest, rank: int, world: int) -> Tuple[Manifest, int]:
    """The entries rank `rank` packs and the dsg_opts.flags to pack them with (no trailer except on the last)."""
    b = shard_bounds(man, world)
    sub = man.slice(int(b[rank]), int(b[rank + 1]))
    flags = 0 if rank == world - 1 else _lib.DSG_F_NO_TRAILER
    return sub, flags


def concat_members(members: List[bytes]) -> bytes:
    """A multi-member .tar.gz: GNU gzip/tar and CPython's gzip read the members back to back as one stream."""
    return b"".join(members)


def stitch_members(members: List[bytes], tar_lens: List[int]) -> bytes:
    """dsg_stitch_members: the shards' members as ONE gzip member (inner headers, trailers and block terminators
    dropped, checksums folded). Pure host work (no GPU needed)."""
    lib = _lib.load()
    n = len(members)
    bufs = [np.frombuffer(m, np.uint8) for m in members]
    ptrs = (ctypes.c_void_p * n)(*[b.ctypes.data for b in bufs])
    lens = (ctypes.c_size_t * n)(*[len(m) for m in members])
    tl = (ctypes.c_uint64 * n)(*[int(t) for t in tar_lens])
    need = ctypes.c_size_t()
    lib.dsg_stitch_members.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                                       ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
    rc = lib.dsg_stitch_members(ptrs, lens, tl, n, None, 0, ctypes.byref(need))
    if rc not in (0, _lib.DSG_E_CAPACITY):
        raise _lib.DsgError(rc, lib.dsg_strerror(rc).decode())
    out = np.zeros(need.value, np.uint8)
    rc = lib.dsg_stitch_members(ptrs, lens, tl, n, out.ctypes.data, out.size, ctypes.byref(need))
    if rc != 0:
        raise _lib.DsgError(rc, lib.dsg_strerror(rc).decode())
    return out[:need.value].tobytes()
