"""ctypes binding of the C++ host mirror of writeTar (devspace_b200/csrc/hostpack.cc, include/dsgpack.h).

``write_tar(packer, files, config)`` is the reference's ``writeTar(files, config)``
(``/root/reference/pkg/devspace/sync/tar.go:146``): it returns the temp-file path of the ``.tar.gz`` and the
``writtenFiles`` map. ``walk`` returns what the walk alone produces (host only, no GPU)."""
from __future__ import annotations

import ctypes
from ctypes import POINTER, byref, c_char_p, c_int, c_int32, c_int64, c_size_t, c_uint32, c_uint64, c_void_p
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from .manifest import ENTRY_DTYPE, Manifest


class CFileInfo(ctypes.Structure):
    _fields_ = [("name", c_char_p), ("size", c_int64), ("mtime", c_int64), ("is_directory", c_int32),
                ("is_symlink", c_int32), ("remote_mode", c_int64), ("remote_uid", c_int64), ("remote_gid", c_int64)]


MATCH_FN = ctypes.CFUNCTYPE(c_int, c_char_p, c_void_p)
LOG_FN = ctypes.CFUNCTYPE(None, c_char_p, c_void_p)


class CSyncConfig(ctypes.Structure):
    _fields_ = [("watch_path", c_char_p), ("ignore", MATCH_FN), ("ignore_user", c_void_p), ("upload_ignore", MATCH_FN),
                ("upload_ignore_user", c_void_p), ("file_map", POINTER(CFileInfo)), ("file_map_len", c_size_t),
                ("log", LOG_FN), ("log_user", c_void_p), ("retry_sleep_ms", c_uint32), ("max_retries", c_uint32)]


@dataclass
class FileInformation:
    """fileInformation (file_information.go:21-32)."""
    Name: str
    Size: int = 0
    Mtime: int = 0
    IsDirectory: bool = False
    IsSymbolicLink: bool = False
    RemoteMode: int = 0
    RemoteUID: int = 0
    RemoteGID: int = 0


@dataclass
class SyncConfig:
    """The part of SyncConfig writeTar reads."""
    WatchPath: str
    ignoreMatcher: Optional[Callable[[str], bool]] = None
    uploadIgnoreMatcher: Optional[Callable[[str], bool]] = None
    fileMap: Dict[str, FileInformation] = field(default_factory=dict)
    Logf: Optional[Callable[[str], None]] = None
    retry_sleep_ms: int = 4000
    max_retries: int = 0


def _bind(lib):
    if getattr(lib, "_hostpack_bound", False):
        return
    lib.dsg_walk.argtypes = [POINTER(CFileInfo), c_size_t, POINTER(CSyncConfig), POINTER(c_void_p)]
    lib.dsg_written_count.argtypes = [c_void_p]
    lib.dsg_written_count.restype = c_size_t
    lib.dsg_written_get.argtypes = [c_void_p, c_size_t]
    lib.dsg_written_get.restype = POINTER(CFileInfo)
    lib.dsg_written_entries.argtypes = [c_void_p, POINTER(c_size_t), POINTER(c_void_p), POINTER(c_size_t), POINTER(c_uint64)]
    lib.dsg_written_entries.restype = c_void_p
    lib.dsg_written_free.argtypes = [c_void_p]
    lib.dsg_written_free.restype = None
    lib.dsg_write_tar.argtypes = [c_void_p, POINTER(CFileInfo), c_size_t, POINTER(CSyncConfig), ctypes.c_char_p, c_size_t,
                                  POINTER(c_void_p)]
    lib._hostpack_bound = True


def _enc(s: str) -> bytes:
    return s.encode("utf-8", "surrogateescape")


def _c_files(files: Sequence[FileInformation]):
    arr = (CFileInfo * max(1, len(files)))()
    keep = []
    for i, f in enumerate(files):
        nm = _enc(f.Name)
        keep.append(nm)
        arr[i] = CFileInfo(nm, f.Size, f.Mtime, int(f.IsDirectory), int(f.IsSymbolicLink), f.RemoteMode, f.RemoteUID,
                           f.RemoteGID)
    return arr, keep


def _c_config(cfg: SyncConfig):
    keep: list = []
    c = CSyncConfig()
    wp = _enc(cfg.WatchPath)
    keep.append(wp)
    c.watch_path = wp
    if cfg.ignoreMatcher:
        fn = MATCH_FN(lambda p, _u, m=cfg.ignoreMatcher: 1 if m(p.decode("utf-8", "surrogateescape")) else 0)
        keep.append(fn)
        c.ignore = fn
    if cfg.uploadIgnoreMatcher:
        fn = MATCH_FN(lambda p, _u, m=cfg.uploadIgnoreMatcher: 1 if m(p.decode("utf-8", "surrogateescape")) else 0)
        keep.append(fn)
        c.upload_ignore = fn
    if cfg.fileMap:
        fm, k2 = _c_files(list(cfg.fileMap.values()))
        keep.extend([fm, k2])
        c.file_map = fm
        c.file_map_len = len(cfg.fileMap)
    if cfg.Logf:
        fn = LOG_FN(lambda line, _u, f=cfg.Logf: f(line.decode("utf-8", "replace")))
        keep.append(fn)
        c.log = fn
    c.retry_sleep_ms = cfg.retry_sleep_ms
    c.max_retries = cfg.max_retries
    return c, keep


def _written(lib, h) -> Tuple[Dict[str, FileInformation], Manifest]:
    out: Dict[str, FileInformation] = {}
    for i in range(lib.dsg_written_count(h)):
        f = lib.dsg_written_get(h, i).contents
        nm = f.name.decode("utf-8", "surrogateescape")
        out[nm] = FileInformation(nm, f.size, f.mtime, bool(f.is_directory), bool(f.is_symlink), f.remote_mode,
                                  f.remote_uid, f.remote_gid)
    n, st, sl, ab = c_size_t(), c_void_p(), c_size_t(), c_uint64()
    p = lib.dsg_written_entries(h, byref(n), byref(st), byref(sl), byref(ab))
    ent = np.zeros(n.value, ENTRY_DTYPE)
    if n.value:
        ctypes.memmove(ent.ctypes.data, p, n.value * ENTRY_DTYPE.itemsize)
    strtab = np.frombuffer(ctypes.string_at(st.value, sl.value), np.uint8).copy() if sl.value else np.zeros(1, np.uint8)
    return out, Manifest(ent, strtab, np.zeros(0, np.uint8))


def walk(files: Sequence[FileInformation], config: SyncConfig) -> Tuple[Dict[str, FileInformation], Manifest]:
    """recursiveTar / tarFolder / tarFile without the writer: (writtenFiles, header list). No GPU needed."""
    lib = _lib.load()
    _bind(lib)
    cf, k1 = _c_files(files)
    cc, k2 = _c_config(config)
    h = c_void_p()
    rc = lib.dsg_walk(cf, len(files), byref(cc), byref(h))
    if rc != 0:
        raise _lib.DsgError(rc, lib.dsg_strerror(rc).decode())
    try:
        return _written(lib, h)
    finally:
        lib.dsg_written_free(h)


def write_tar(packer, files: Sequence[FileInformation], config: SyncConfig) -> Tuple[str, Dict[str, FileInformation]]:
    """writeTar(files, config) -> (temp file path, writtenFiles); the archive is produced on the GPU."""
    lib = packer._lib
    _bind(lib)
    cf, k1 = _c_files(files)
    cc, k2 = _c_config(config)
    h = c_void_p()
    path = ctypes.create_string_buffer(256)
    rc = lib.dsg_write_tar(packer._ctx, cf, len(files), byref(cc), path, len(path), byref(h))
    if rc != 0:
        raise _lib.DsgError(rc, lib.dsg_last_error(packer._ctx).decode() or lib.dsg_strerror(rc).decode())
    try:
        written, _ = _written(lib, h)
    finally:
        lib.dsg_written_free(h)
    return path.value.decode(), written
