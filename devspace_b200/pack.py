"""Python host side above the C ABI: a thin mirror of the Go call site.

``Packer.pack(manifest)`` stands where the body of ``writeTar`` stood
(``/root/reference/pkg/devspace/sync/tar.go:146-182``): it takes the headers + payload bytes the walk
collected and returns the ``.tar.gz`` bytes that the Go code would have written to its temp file.
All compute happens in libdsgpack.so on the GPU; this module only moves pointers.
"""
from __future__ import annotations

import ctypes
from ctypes import byref, c_size_t, c_uint32, c_uint64, c_void_p
from typing import Optional

import numpy as np

from . import _lib
from ._lib import DsgError, DsgOpts, DsgStats
from .manifest import Manifest


def _ptr(a: Optional[np.ndarray]) -> c_void_p:
    if a is None or a.size == 0:
        return c_void_p(0)
    return c_void_p(a.ctypes.data)


class PinnedArray:
    """uint8 numpy view over cudaHostAlloc memory owned by a Packer (dsg_pinned_alloc)."""

    def __init__(self, packer: "Packer", nbytes: int):
        self._packer = packer
        self.nbytes = int(nbytes)
        self._p = packer._lib.dsg_pinned_alloc(packer._ctx, max(1, self.nbytes))
        if not self._p:
            raise DsgError(_lib.DSG_E_NOMEM, f"dsg_pinned_alloc({nbytes}) failed")
        buf = (ctypes.c_uint8 * max(1, self.nbytes)).from_address(self._p)
        self.array = np.frombuffer(buf, dtype=np.uint8)[: self.nbytes]

    def free(self):
        if self._p:
            self.array = None
            self._packer._lib.dsg_pinned_free(self._packer._ctx, self._p)
            self._p = None


class DeviceBuffer:
    def __init__(self, packer: "Packer", nbytes: int):
        self._packer = packer
        self.nbytes = int(nbytes)
        self.ptr = packer._lib.dsg_dev_alloc(packer._ctx, max(1, self.nbytes))
        if not self.ptr:
            raise DsgError(_lib.DSG_E_NOMEM, f"dsg_dev_alloc({nbytes}) failed")

    def upload(self, a: np.ndarray):
        a = np.ascontiguousarray(a)
        assert a.nbytes <= self.nbytes
        if a.nbytes:
            self._packer._check(self._packer._lib.dsg_dev_upload(self._packer._ctx, self.ptr, _ptr(a), a.nbytes))
        return self

    def download(self, nbytes: Optional[int] = None) -> np.ndarray:
        n = self.nbytes if nbytes is None else int(nbytes)
        out = np.empty(n, np.uint8)
        if n:
            self._packer._check(self._packer._lib.dsg_dev_download(self._packer._ctx, _ptr(out), self.ptr, n))
        return out

    def free(self):
        if self.ptr:
            self._packer._lib.dsg_dev_free(self._packer._ctx, self.ptr)
            self.ptr = None


class DeviceManifest:
    """Manifest + arena resident in HBM (for the kernel-only measurements and the stage tests)."""

    def __init__(self, packer: "Packer", man: Manifest):
        self.n = man.n
        self.strtab_len = int(man.strtab.size)
        self.arena_len = int(man.arena.size)
        self.payload_bytes = man.payload_bytes
        self.entries = DeviceBuffer(packer, max(1, man.entries.nbytes)).upload(man.entries.view(np.uint8).reshape(-1))
        self.strtab = DeviceBuffer(packer, max(1, man.strtab.size)).upload(man.strtab)
        self.arena = DeviceBuffer(packer, max(1, man.arena.size)).upload(man.arena)

    def free(self):
        for b in (self.entries, self.strtab, self.arena):
            b.free()


class Packer:
    """One dsg_ctx: one GPU, one caller at a time (SURVEY.md 8b "Threading")."""

    def __init__(self, device: int = 0, level: int = _lib.DSG_LEVEL_DEFAULT, segment_bytes: int = 0, flags: int = 0,
                 **lz):
        self._lib = _lib.load()
        self._base = (level, segment_bytes, flags)
        opts = self._opts(**lz)
        ctx = c_void_p()
        rc = self._lib.dsg_ctx_create(device, byref(opts), byref(ctx))
        if rc != 0:
            raise DsgError(rc, "dsg_ctx_create failed (no usable CUDA device? there is no CPU fallback): "
                           + self._lib.dsg_strerror(rc).decode())
        self._ctx = ctx
        self.device = device
        self.level = level

    # -- plumbing -------------------------------------------------------------------------------------
    def _opts(self, chunk_units: int = 0, run_chunks: int = 0, grid: int = 0, tail_bytes: Optional[int] = None,
              profile: bool = False, no_bulk: bool = False, gunzip_batch_mib: int = 0, effort: int = 0) -> DsgOpts:
        level, segment_bytes, flags = self._base
        # effort (profiling tools only): search effort in bits 8.. of lz_flags, see lz_launch in dsgpack.cu; 0 = default
        lz_flags = ((_lib.DSG_LZ_TAIL_SET if tail_bytes is not None else 0) | (_lib.DSG_LZ_PROFILE if profile else 0)
                    | (_lib.DSG_LZ_NO_BULK if no_bulk else 0) | ((effort & 0xffffff) << 8))
        return DsgOpts(ctypes.sizeof(DsgOpts), level, segment_bytes, flags, gunzip_batch_mib, chunk_units, run_chunks, grid,
                       lz_flags,
                       int(tail_bytes or 0))

    def set_lz_geometry(self, **lz) -> None:
        """dsg_ctx_set_opts: launch geometry of the level-6 kernel (chunk_units, run_chunks, grid, tail_bytes,
        profile); no arguments = automatic. Used by the parity tests and the profiling tools."""
        opts = self._opts(**lz)
        self._check(self._lib.dsg_ctx_set_opts(self._ctx, byref(opts)))

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.dsg_ctx_destroy(self._ctx)
            self._ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int):
        if rc != 0:
            raise DsgError(rc, self._lib.dsg_last_error(self._ctx).decode() or self._lib.dsg_strerror(rc).decode())

    def stats(self) -> dict:
        st = DsgStats()
        self._check(self._lib.dsg_get_stats(self._ctx, byref(st)))
        return st.as_dict()

    def pinned(self, nbytes: int) -> PinnedArray:
        return PinnedArray(self, nbytes)

    # -- manifest-only helpers (host integer work) ---------------------------------------------------------
    def tar_size(self, man: Manifest, flags: int = 0) -> int:
        out = c_uint64()
        self._check(self._lib.dsg_tar_size(_ptr(man.entries), man.n, _ptr(man.strtab), man.strtab.size, flags, byref(out)))
        return out.value

    def gzip_bound(self, tar_bytes: int) -> int:
        return int(self._lib.dsg_gzip_bound(tar_bytes))

    def shard_bounds(self, man: Manifest, nshards: int) -> np.ndarray:
        b = np.zeros(nshards + 1, np.uint64)
        self._check(self._lib.dsg_shard_bounds(_ptr(man.entries), man.n, _ptr(man.strtab), man.strtab.size, nshards,
                                               b.ctypes.data_as(ctypes.POINTER(c_uint64))))
        return b

    # -- the hot path ------------------------------------------------------------------------------------------
    def pack_into(self, man: Manifest, out: np.ndarray, tar_only: bool = False) -> int:
        """dsg_pack / dsg_pack_tar_only with caller-owned (ideally pinned) buffers. Returns bytes produced."""
        n_out = c_size_t()
        fn = self._lib.dsg_pack_tar_only if tar_only else self._lib.dsg_pack
        rc = fn(self._ctx, _ptr(man.entries), man.n, _ptr(man.strtab), man.strtab.size, _ptr(man.arena),
                man.arena.size, _ptr(out), out.size, byref(n_out))
        self._check(rc)
        return n_out.value

    def pack_to_fd(self, man: Manifest, fd: int, tar_only: bool = False) -> int:
        """dsg_pack_fd: the stream goes to the file descriptor segment by segment, overlapped with the GPU work."""
        n_out = ctypes.c_uint64()
        rc = self._lib.dsg_pack_fd(self._ctx, _ptr(man.entries), man.n, _ptr(man.strtab), man.strtab.size,
                                   _ptr(man.arena), man.arena.size, int(fd), 0 if tar_only else 1, byref(n_out))
        self._check(rc)
        return n_out.value

    def pack(self, man: Manifest, tar_only: bool = False) -> bytes:
        """writeTar's archive bytes for the manifest: gzip(tar stream) (or the bare tar stream)."""
        T = self.tar_size(man, 0)
        cap = T if tar_only else self.gzip_bound(T)
        out = np.empty(cap + 64, np.uint8)
        n = self.pack_into(man, out, tar_only)
        return out[:n].tobytes()

    def pack_device(self, dm: DeviceManifest, d_out: DeviceBuffer) -> int:
        n_out = c_size_t()
        rc = self._lib.dsg_pack_device(self._ctx, dm.entries.ptr, dm.n, dm.strtab.ptr, dm.strtab_len, dm.arena.ptr,
                                       dm.arena_len, d_out.ptr, d_out.nbytes, byref(n_out))
        self._check(rc)
        return n_out.value

    # -- single stages (device buffers) ----------------------------------------------------------------------
    def tar_build_device(self, dm: DeviceManifest, d_tar: DeviceBuffer, flags: int = 0) -> int:
        n = c_uint64()
        rc = self._lib.dsg_dev_tar_build(self._ctx, dm.entries.ptr, dm.n, dm.strtab.ptr, dm.strtab_len, dm.arena.ptr,
                                         dm.arena_len, flags, d_tar.ptr, d_tar.nbytes, byref(n))
        self._check(rc)
        return n.value

    def crc32_device(self, d_buf: DeviceBuffer, nbytes: int, offset: int = 0) -> int:
        crc = c_uint32()
        self._check(self._lib.dsg_dev_crc32(self._ctx, c_void_p(d_buf.ptr + offset), nbytes, byref(crc)))
        return crc.value

    def file_crc32_device(self, dm: "DeviceManifest", d_crcs: DeviceBuffer) -> None:
        """hashFileCRC32 (pkg/util/hash/hash.go:165) of every entry of a resident manifest, into d_crcs (u32[n])."""
        self._check(self._lib.dsg_dev_file_crc32(self._ctx, dm.entries.ptr, dm.n, dm.arena.ptr, dm.arena_len,
                                                 d_crcs.ptr))

    def file_crc32(self, man: Manifest) -> np.ndarray:
        """Per-file CRC32-IEEE of a host manifest (one host->device pass of the arena)."""
        crcs = np.zeros(max(1, man.n), np.uint32)
        ent = np.ascontiguousarray(man.entries)
        arena = np.ascontiguousarray(man.arena)
        self._check(self._lib.dsg_file_crc32(self._ctx, c_void_p(ent.ctypes.data), man.n,
                                             c_void_p(arena.ctypes.data), int(arena.size), c_void_p(crcs.ctypes.data)))
        return crcs[:man.n]

    def deflate_device(self, d_in: DeviceBuffer, nbytes: int, d_out: DeviceBuffer, level: Optional[int] = None) -> int:
        n = c_size_t()
        lvl = self.level if level is None else level
        self._check(self._lib.dsg_dev_deflate(self._ctx, d_in.ptr, nbytes, lvl, d_out.ptr, d_out.nbytes, byref(n)))
        return n.value

    def gzip_device(self, d_in: DeviceBuffer, nbytes: int, d_out: DeviceBuffer, level: Optional[int] = None) -> int:
        n = c_size_t()
        lvl = self.level if level is None else level
        self._check(self._lib.dsg_dev_gzip(self._ctx, d_in.ptr, nbytes, lvl, d_out.ptr, d_out.nbytes, byref(n)))
        return n.value

    # convenience wrappers used by tests: host bytes in, host bytes out, one stage on the GPU
    def tar_stream(self, man: Manifest, flags: int = 0) -> bytes:
        dm = DeviceManifest(self, man)
        T = self.tar_size(man, flags)
        d_tar = DeviceBuffer(self, T + 64)
        try:
            n = self.tar_build_device(dm, d_tar, flags)
            return d_tar.download(n).tobytes()
        finally:
            dm.free()
            d_tar.free()

    def crc32(self, data) -> int:
        a = np.frombuffer(bytes(data), np.uint8) if not isinstance(data, np.ndarray) else data
        d = DeviceBuffer(self, a.size + 16).upload(a)
        try:
            return self.crc32_device(d, a.size)
        finally:
            d.free()

    def gunzip(self, blob, out_cap: Optional[int] = None) -> bytes:
        """dsg_gunzip: chunk-parallel inflate of a (multi-member) gzip stream this library wrote; the tar stream back.
        out_cap defaults to the sum the member trailers promise (read on the host: ISIZE of every member is only
        known after the walk, so the default asks the library once with no buffer)."""
        a = np.frombuffer(bytes(blob), np.uint8) if not isinstance(blob, np.ndarray) else blob
        n = c_size_t()
        if out_cap is None:
            rc = self._lib.dsg_gunzip(self._ctx, _ptr(a), a.size, c_void_p(0), 0, byref(n))
            if rc not in (0, _lib.DSG_E_CAPACITY):
                self._check(rc)
            out_cap = n.value
        out = np.empty(max(out_cap, 1), np.uint8)
        self._check(self._lib.dsg_gunzip(self._ctx, _ptr(a), a.size, _ptr(out), out_cap, byref(n)))
        return out[: n.value].tobytes()

    def deflate(self, data, level: Optional[int] = None, gzip: bool = False) -> bytes:
        a = np.frombuffer(bytes(data), np.uint8) if not isinstance(data, np.ndarray) else data
        d_in = DeviceBuffer(self, a.size + 16).upload(a)
        d_out = DeviceBuffer(self, self.gzip_bound(a.size) + 64)
        try:
            fn = self.gzip_device if gzip else self.deflate_device
            n = fn(d_in, a.size, d_out, level)
            return d_out.download(n).tobytes()
        finally:
            d_in.free()
            d_out.free()
