"""One call, several GPUs, one ordered archive: the host mirror of dsg_multi_* (include/dsgpack.h).

Stands where ONE ``writeTar`` call stood (``/root/reference/pkg/devspace/sync/upstream.go:342``, kaniko's whole tree
``util.go:68-73``) on a box with several GPUs: the header list is sharded by tar-stream bytes, every device packs its
range, the members reach the file descriptor in shard order. All compute happens in libdsgpack.so."""
from __future__ import annotations

import ctypes
from ctypes import byref, c_int, c_size_t, c_uint64, c_void_p
from typing import List, Sequence

import numpy as np

from . import _lib
from ._lib import DsgError, DsgOpts, DsgStats
from .manifest import Manifest
from .pack import _ptr


class MultiPacker:
    def __init__(self, devices: Sequence[int], level: int = _lib.DSG_LEVEL_DEFAULT, segment_bytes: int = 0):
        self._lib = _lib.load()
        opts = DsgOpts(ctypes.sizeof(DsgOpts), level, segment_bytes, 0, 0, 0, 0, 0, 0, 0)
        devs = (c_int * len(devices))(*devices)
        h = c_void_p()
        rc = self._lib.dsg_multi_create(devs, len(devices), byref(opts), byref(h))
        if rc != 0:
            raise DsgError(rc, "dsg_multi_create failed (a listed CUDA device is missing? there is no CPU fallback): "
                           + self._lib.dsg_strerror(rc).decode())
        self._h = h
        self.devices = list(devices)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.dsg_multi_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int):
        if rc != 0:
            raise DsgError(rc, self._lib.dsg_multi_last_error(self._h).decode() or self._lib.dsg_strerror(rc).decode())

    def pack_to_fd(self, man: Manifest, fd: int, one_member: bool = False) -> int:
        n_out = c_uint64()
        rc = self._lib.dsg_multi_pack_fd(self._h, _ptr(man.entries), man.n, _ptr(man.strtab), man.strtab.size,
                                         _ptr(man.arena), man.arena.size, int(fd),
                                         _lib.DSG_MULTI_ONE_MEMBER if one_member else 0, byref(n_out))
        self._check(rc)
        return n_out.value

    def pack(self, man: Manifest, one_member: bool = False) -> bytes:
        lib = self._lib
        T = c_uint64()
        rc = lib.dsg_tar_size(_ptr(man.entries), man.n, _ptr(man.strtab), man.strtab.size, 0, byref(T))
        if rc != 0:
            raise DsgError(rc, lib.dsg_strerror(rc).decode())
        cap = int(lib.dsg_gzip_bound(T.value)) + 32 * len(self.devices) + 64
        out = np.empty(cap, np.uint8)
        n_out = c_size_t()
        rc = lib.dsg_multi_pack(self._h, _ptr(man.entries), man.n, _ptr(man.strtab), man.strtab.size, _ptr(man.arena),
                                man.arena.size, _lib.DSG_MULTI_ONE_MEMBER if one_member else 0, _ptr(out), out.size,
                                byref(n_out))
        self._check(rc)
        return out[: n_out.value].tobytes()

    def stats(self) -> List[dict]:
        res = []
        for s in range(self._lib.dsg_multi_shards(self._h)):
            st = DsgStats()
            self._check(self._lib.dsg_multi_get_stats(self._h, s, byref(st)))
            res.append(st.as_dict())
        return res
