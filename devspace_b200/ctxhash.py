"""ctypes binding of the host mirror of DirectoryExcludes (devspace_b200/csrc/ctxhash.cc, include/dsgpack.h).

``directory_excludes(packer, src_path, exclude)`` is the reference's ``hash.DirectoryExcludes(srcPath, patterns)``
(``/root/reference/pkg/util/hash/hash.go:43``) with the docker pattern matcher replaced by a callback
``exclude(rel_path, is_dir) -> 0 keep | 1 skip | 2 skip the entry, enter the directory``. The per-file CRC32 is
computed on the GPU; there is no CPU checksum in this package."""
from __future__ import annotations

import ctypes
from ctypes import c_char_p, c_int, c_size_t, c_void_p
from typing import Callable, Optional

from . import _lib

EXCLUDE_FN = ctypes.CFUNCTYPE(c_int, c_char_p, c_int, c_void_p)
KEEP, SKIP, SKIP_ENTER = 0, 1, 2


def directory_excludes(packer, src_path: str, exclude: Optional[Callable[[str, bool], int]] = None) -> str:
    lib = packer._lib
    lib.dsg_directory_hash.argtypes = [c_void_p, c_char_p, EXCLUDE_FN, c_void_p, c_char_p, c_char_p, c_size_t]
    cb = EXCLUDE_FN(lambda p, d, _u: int(exclude(p.decode("utf-8", "surrogateescape"), bool(d)))) if exclude \
        else ctypes.cast(None, EXCLUDE_FN)
    out = ctypes.create_string_buffer(65)
    err = ctypes.create_string_buffer(512)
    rc = lib.dsg_directory_hash(packer._ctx, src_path.encode("utf-8", "surrogateescape"), cb, None, out, err, len(err))
    if rc != 0:
        raise _lib.DsgError(rc, err.value.decode("utf-8", "replace") or lib.dsg_strerror(rc).decode())
    return out.value.decode()
