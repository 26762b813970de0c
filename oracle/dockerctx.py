"""oracle.dockerctx -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the docker build-context tar DevSpace sends to the daemon
(``/root/reference/pkg/devspace/builder/docker/docker.go:94-97``: ``archive.TarWithOptions(contextDir,
&archive.TarOptions{ExcludePatterns: excludes, ChownOpts: &idtools.Identity{UID: 0, GID: 0}})``).

The arithmetic lives in ``github.com/docker/docker/pkg/archive`` (go.mod:22, pinned
v0.0.0-20181211214838-62d80835abe3) on top of go1.11 ``archive/tar``; neither is under /root/reference and no Go
toolchain exists here, so this restates the published algorithm. **PARITY UNPINNED** ("no
tests at all" upstream). What is restated:

* ``TarWithOptions``: ``filepath.Walk`` over the context (``os.Lstat``, names byte-sorted, pre-order); the source
  directory itself gets no entry (``IncludeSourceDir`` false); names are relative paths without ``./``; the exclude
  decisions are those of hash.go (which copied them from here), given by a callback like oracle.ctxhash; a path is
  added once (``seen``); an entry that cannot be added is left out (docker logs and continues).
* ``addTarFile`` / ``FileInfoHeader``: ``tar.FileInfoHeader`` (perm | setuid | setgid | sticky; Size only for
  regular files; Uname/Gname looked up from the file's owner), ``Format = PAX`` (which still writes plain USTAR
  whenever it can: "PAX implies USTAR allowed too"), ``ModTime`` truncated to seconds, the file-type bits put back
  into Mode (``fillGo18FileTypeBits``), directories named with a trailing ``/`` (``canonicalTarName``), symlinks
  as typeflag '2' with ``os.Readlink``'s target, a second name of an inode with several links as typeflag '1'
  pointing at the first, then ``Uid = Gid = 0`` (ChownOpts; the names stay). No compression.
* Not restated: device nodes, FIFOs (typeflags '3' '4' '6') and sockets are left out (sockets are an error in
  tar.FileInfoHeader as well); security.capability xattrs; Windows path handling.

The byte level (USTAR / PAX blocks, linkpath records) is oracle/gotar.c.
"""
from __future__ import annotations

import grp
import os
import pwd
import stat as _stat
from typing import Callable, Dict, List, Optional, Tuple

from . import Entry

KEEP, SKIP, SKIP_ENTER = 0, 1, 2
_TYPE_BITS = ((_stat.S_ISREG, 0o100000), (_stat.S_ISDIR, 0o40000), (_stat.S_ISLNK, 0o120000))


def _names(uid: int, gid: int, cache: Dict) -> Tuple[bytes, bytes]:
    if ("u", uid) not in cache:
        try:
            cache[("u", uid)] = pwd.getpwuid(uid).pw_name.encode()
        except KeyError:
            cache[("u", uid)] = b""
    if ("g", gid) not in cache:
        try:
            cache[("g", gid)] = grp.getgrgid(gid).gr_name.encode()
        except KeyError:
            cache[("g", gid)] = b""
    return cache[("u", uid)], cache[("g", gid)]


def entries(src_path: str, exclude: Optional[Callable[[str, bool], int]] = None) -> List[Entry]:
    """The tar.Header list (with payloads) TarWithOptions writes for the context directory."""
    out: List[Entry] = []
    seen = set()
    inodes: Dict[Tuple[int, int], bytes] = {}
    cache: Dict = {}
    root = os.path.normpath(src_path)

    def add(path: str, rel: str, st) -> None:
        mode = st.st_mode
        is_dir = _stat.S_ISDIR(mode)
        if not (_stat.S_ISREG(mode) or is_dir or _stat.S_ISLNK(mode)):
            return                                                    # devices, FIFOs, sockets: not restated
        m = _stat.S_IMODE(mode) & 0o777
        m |= 0o4000 if mode & _stat.S_ISUID else 0
        m |= 0o2000 if mode & _stat.S_ISGID else 0
        m |= 0o1000 if mode & _stat.S_ISVTX else 0
        for pred, bits in _TYPE_BITS:                                 # fillGo18FileTypeBits
            if pred(mode):
                m |= bits
        uname, gname = _names(st.st_uid, st.st_gid, cache)
        name = os.fsencode(rel) + (b"/" if is_dir else b"")           # canonicalTarName
        e = Entry(name=name, mtime_sec=int(st.st_mtime_ns // 1_000_000_000), mtime_nsec=0, mode=m, uid=0, gid=0,
                  uname=uname, gname=gname)
        if is_dir:
            e.typeflag = ord("5")
        elif _stat.S_ISLNK(mode):
            e.typeflag = ord("2")
            e.linkname = os.fsencode(os.readlink(path))
        else:
            e.typeflag = ord("0")
            e.size = st.st_size
        if not is_dir and st.st_nlink > 1:                            # hasHardlinks
            key = (st.st_dev, st.st_ino)
            if key in inodes:
                e.typeflag, e.linkname, e.size = ord("1"), inodes[key], 0
            else:
                inodes[key] = name
        if e.typeflag == ord("0") and e.size > 0:
            try:
                with open(path, "rb") as f:
                    e.data = f.read()
            except OSError:
                return                                                # "Can't add file ... to tar": left out
            if len(e.data) != e.size:
                return
        out.append(e)

    def rec(path: str, rel: str, st) -> None:
        is_dir = _stat.S_ISDIR(st.st_mode)
        verdict = KEEP
        if rel != ".":
            verdict = exclude(rel, is_dir) if exclude else KEEP
            if verdict == SKIP or (verdict == SKIP_ENTER and not is_dir):
                return
            if verdict == KEEP and rel not in seen:
                seen.add(rel)
                add(path, rel, st)
        if is_dir:
            for nm in sorted(os.listdir(os.fsencode(path))):
                child = path + "/" + os.fsdecode(nm)
                try:
                    cst = os.lstat(child)
                except OSError:
                    continue
                rec(child, os.fsdecode(nm) if rel == "." else rel + "/" + os.fsdecode(nm), cst)

    st = os.lstat(root)
    if not _stat.S_ISDIR(st.st_mode):
        raise NotADirectoryError(src_path)
    rec(root, ".", st)
    return out
