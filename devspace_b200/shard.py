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
