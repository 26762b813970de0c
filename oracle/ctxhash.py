"""oracle.ctxhash -- TEST INFRASTRUCTURE ONLY.

CPU restatement of DevSpace's build-context hash, ``/root/reference/pkg/util/hash/hash.go``:

* ``hash_file_crc32``     -- hashFileCRC32 (hash.go:165-197): CRC32 with the reflected IEEE polynomial 0xedb88320
  over the file's bytes, ``hex.EncodeToString(hash.Sum(nil))`` = 8 lower-case hex digits, big endian.
* ``directory_excludes``  -- DirectoryExcludes (hash.go:43-163): ``filepath.Walk`` order (root first, then the
  names of every directory in byte-wise sorted order, ``os.Lstat`` so links are not followed), exclude decisions,
  SHA-256 over ``filePath`` for directories and ``filePath + ";" + crc`` for everything else (a file that cannot
  be read is left out, hash.go:147-149), ``fmt.Sprintf("%x", sum)``.
* ``directory``           -- Directory (hash.go:20-40): SHA-256 over ``path;size;mtime_ns`` of every walked entry.

The pattern matching of the reference is ``github.com/docker/docker/pkg/fileutils`` (go.mod:22, pinned
v0.0.0-20181211214838-62d80835abe3), which is not in /root/reference: it stays on the host in the product too, so
here it is a callback ``exclude(rel_path, is_dir) -> 0 keep | 1 skip (a directory is not entered) | 2 skip this
entry but enter the directory`` (2 = the "!exception below an excluded directory" case, hash.go:112-131).

Parity status: CRC32-IEEE and SHA-256 are pinned by their standard check values (tests/test_oracle_ctxhash.py);
the walk order and string format are UNPINNED against Go itself (no toolchain, no golden values upstream).
"""
from __future__ import annotations

import hashlib
import os
import stat as _stat
import zlib
from typing import Callable, List, Optional, Tuple

KEEP, SKIP, SKIP_ENTER = 0, 1, 2


def hash_file_crc32(path: str) -> Optional[str]:
    """hashFileCRC32: None where Go returns an error (open or read failure)."""
    try:
        crc = 0
        with open(path, "rb") as f:
            while True:
                b = f.read(1 << 20)
                if not b:
                    break
                crc = zlib.crc32(b, crc)
        return "%08x" % (crc & 0xFFFFFFFF)
    except OSError:
        return None


def _clean(p: str) -> str:
    # filepath.Join(srcPath, ".") == filepath.Clean(srcPath)
    return os.path.normpath(p) if p else "."


def walk(root: str):
    """filepath.Walk(root): yields (path, lstat) pre-order, children in sorted byte order."""
    root_b = os.fsencode(root)

    def rec(path_b: bytes, st):
        yield os.fsdecode(path_b), st
        if not _stat.S_ISDIR(st.st_mode):
            return
        try:
            names = sorted(os.listdir(path_b))
        except OSError:
            return
        for nm in names:
            child = path_b + b"/" + nm if path_b != b"/" else b"/" + nm
            try:
                cst = os.lstat(child)
            except OSError:
                continue
            yield from rec(child, cst)

    yield from rec(root_b, os.lstat(root_b))


def records(src_path: str, exclude: Optional[Callable[[str, bool], int]] = None) -> List[Tuple[str, bool, Optional[str]]]:
    """The ordered (filePath, is_dir, crc_hex) list DirectoryExcludes feeds to SHA-256."""
    st = os.lstat(src_path)
    if not _stat.S_ISDIR(st.st_mode):
        raise NotADirectoryError("Path %s is not a directory" % src_path)
    root = _clean(src_path)
    out: List[Tuple[str, bool, Optional[str]]] = []
    seen = set()

    def rec(path: str, st) -> None:
        rel = os.path.relpath(path, src_path)
        is_dir = _stat.S_ISDIR(st.st_mode)
        rel_file = rel if rel == "." else "./" + rel          # hash.go:87-89
        if rel_file != ".":
            verdict = exclude(rel_file, is_dir) if exclude else KEEP
            if verdict == SKIP:
                return                                         # file: return nil; directory: filepath.SkipDir
            if verdict == SKIP_ENTER and not is_dir:
                return
        else:
            verdict = KEEP
        if verdict == KEEP and rel_file not in seen:
            seen.add(rel_file)
            if is_dir:
                out.append((path, True, None))                 # hash.go:141-143
            else:
                crc = hash_file_crc32(path)                    # hash.go:146-151
                if crc is not None:
                    out.append((path, False, crc))
        if is_dir:
            for nm in sorted(os.listdir(os.fsencode(path))):
                child = path + "/" + os.fsdecode(nm) if path != "/" else "/" + os.fsdecode(nm)
                try:
                    cst = os.lstat(child)
                except OSError:
                    continue
                rec(child, cst)

    rec(root, os.lstat(root))
    return out


def digest(recs) -> str:
    h = hashlib.sha256()
    for path, is_dir, crc in recs:
        h.update(os.fsencode(path) if is_dir else os.fsencode(path) + b";" + crc.encode())
    return h.hexdigest()


def directory_excludes(src_path: str, exclude: Optional[Callable[[str, bool], int]] = None) -> str:
    return digest(records(src_path, exclude))


def directory(path: str) -> str:
    """Directory (hash.go:20-40): path;size;mtime(ns) of every entry; walk errors are ignored."""
    h = hashlib.sha256()
    for p, st in walk(path):
        h.update(os.fsencode(p) + b";" + str(st.st_size).encode() + b";" + str(st.st_mtime_ns).encode())
    return h.hexdigest()
