"""ctypes binding of the host mirror of docker's build-context tar (devspace_b200/csrc/dockerctx.cc).

``context_tar(packer, context_dir, out_path, exclude)`` is what ``archive.TarWithOptions(contextDir, ...)`` produces at
``/root/reference/pkg/devspace/builder/docker/docker.go:94-97``; the stream is written to ``out_path``. ``exclude`` is
the ``.dockerignore`` matcher as a callback ``(rel_path, is_dir) -> 0 keep | 1 skip | 2 skip the entry, enter the dir``."""
from __future__ import annotations

import ctypes
from ctypes import POINTER, byref, c_char_p, c_size_t, c_uint64, c_void_p
from typing import Callable, Optional, Tuple

from . import _lib
from .ctxhash import EXCLUDE_FN


def context_tar(packer, context_dir: str, out_path: str,
                exclude: Optional[Callable[[str, bool], int]] = None) -> Tuple[int, int]:
    """Returns (tar bytes, entries)."""
    lib = packer._lib
    lib.dsg_docker_context_tar.argtypes = [c_void_p, c_char_p, EXCLUDE_FN, c_void_p, c_char_p, POINTER(c_uint64),
                                           POINTER(c_uint64), c_char_p, c_size_t]
    cb = EXCLUDE_FN(lambda p, d, _u: int(exclude(p.decode("utf-8", "surrogateescape"), bool(d)))) if exclude \
        else ctypes.cast(None, EXCLUDE_FN)
    nbytes, nent = c_uint64(), c_uint64()
    err = ctypes.create_string_buffer(512)
    rc = lib.dsg_docker_context_tar(packer._ctx, context_dir.encode("utf-8", "surrogateescape"), cb, None,
                                    out_path.encode("utf-8", "surrogateescape"), byref(nbytes), byref(nent), err, len(err))
    if rc != 0:
        raise _lib.DsgError(rc, err.value.decode("utf-8", "replace") or lib.dsg_strerror(rc).decode())
    return nbytes.value, nent.value
