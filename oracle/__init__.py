"""oracle -- TEST INFRASTRUCTURE ONLY.

CPU oracle for DevSpace's tar+gzip pack path (``/root/reference/pkg/devspace/sync/tar.go:146-304``).
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this package; the product (``devspace_b200``) never does.

* ``gotar.c``  -- go1.11 ``archive/tar`` writer restatement (USTAR + PAX), byte level.
* ``gogzip.c`` -- go1.11 ``compress/gzip`` framing around a zlib level-6 raw deflate body.
* ``walk()``   -- restatement of the traversal in ``writeTar`` / ``recursiveTar`` / ``tarFolder`` /
  ``tarFile`` (tar.go:146-304): order, de-duplication, excludes, empty-directory rule, Remote* override.

Parity status: UNPINNED against Go itself (no Go toolchain, stdlib not vendored, upstream tests hold
no golden bytes). Cross-checked against GNU tar / CPython tarfile / CPython zlib.
"""
from __future__ import annotations

import ctypes
import os
import stat as _stat
import subprocess
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libgooracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile ``libgooracle.so`` from gotar.c + gogzip.c (gcc, -lz)."""
    srcs = [os.path.join(_HERE, f) for f in ("gotar.c", "gogzip.c", "gotar.h")]
    stale = force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs
    )
    if stale:
        subprocess.run(["make", "-C", _HERE, "-B" if force else "-s", "libgooracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _LIB_PATH


class _CEntry(ctypes.Structure):
    _fields_ = [
        ("name", ctypes.c_char_p), ("name_len", ctypes.c_size_t),
        ("uname", ctypes.c_char_p), ("uname_len", ctypes.c_size_t),
        ("gname", ctypes.c_char_p), ("gname_len", ctypes.c_size_t),
        ("size", ctypes.c_int64),
        ("mtime_sec", ctypes.c_int64),
        ("mtime_nsec", ctypes.c_uint32),
        ("mode", ctypes.c_int64),
        ("uid", ctypes.c_int64), ("gid", ctypes.c_int64),
        ("typeflag", ctypes.c_uint8),
        ("data", ctypes.c_void_p),
        ("linkname", ctypes.c_char_p), ("linkname_len", ctypes.c_size_t),
    ]


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(_LIB_PATH)
        lib.gotar_header_bytes.argtypes = [ctypes.POINTER(_CEntry), ctypes.POINTER(ctypes.c_size_t)]
        lib.gotar_stream_size.argtypes = [ctypes.POINTER(_CEntry), ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
        lib.gotar_write.argtypes = [ctypes.POINTER(_CEntry), ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t,
                                    ctypes.POINTER(ctypes.c_size_t)]
        lib.gotar_path_clean.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p]
        lib.gotar_path_clean.restype = ctypes.c_size_t
        lib.gogzip_bound.argtypes = [ctypes.c_size_t]
        lib.gogzip_bound.restype = ctypes.c_size_t
        lib.gogzip_pack.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t,
                                    ctypes.POINTER(ctypes.c_size_t)]
        lib.gogzip_crc32.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
        lib.gogzip_crc32.restype = ctypes.c_uint32
        lib.gooracle_pack_shards_mt.argtypes = [ctypes.POINTER(_CEntry), ctypes.POINTER(ctypes.c_size_t),
                                                ctypes.POINTER(ctypes.c_size_t), ctypes.c_size_t, ctypes.c_int,
                                                ctypes.c_int, ctypes.POINTER(ctypes.c_size_t),
                                                ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_double)]
        _lib = lib
    return _lib


class OracleError(Exception):
    """archive/tar refused the header (ErrHeader / ErrFieldTooLong) or the entry kind is impossible."""

    def __init__(self, code: int):
        super().__init__({-1: "archive/tar: header field cannot be encoded",
                          -2: "typeflag not producible by tar.go",
                          -3: "buffer too small"}.get(code, f"oracle error {code}"))
        self.code = code


@dataclass
class Entry:
    """One tar.Header as tar.go:242-255 / 280-294 builds it (plus the payload for io.Copy, tar.go:298)."""
    name: bytes
    size: int = 0
    mtime_sec: int = 0
    mtime_nsec: int = 0
    mode: int = 0o644
    uid: int = 0
    gid: int = 0
    uname: bytes = b""
    gname: bytes = b""
    typeflag: int = ord("0")
    data: object = b""   # bytes / bytearray / memoryview / numpy uint8 array of `size` bytes
    linkname: bytes = b""  # hdr.Linkname for typeflag '1' / '2' (docker's build-context dialect only)

    def __post_init__(self):
        if isinstance(self.name, str):
            self.name = self.name.encode("utf-8", "surrogateescape")
        if isinstance(self.uname, str):
            self.uname = self.uname.encode("utf-8", "surrogateescape")
        if isinstance(self.gname, str):
            self.gname = self.gname.encode("utf-8", "surrogateescape")
        if isinstance(self.linkname, str):
            self.linkname = self.linkname.encode("utf-8", "surrogateescape")


def _data_ptr(d, keep: list) -> Optional[int]:
    if d is None:
        return None
    try:
        import numpy as np
        if isinstance(d, np.ndarray):
            a = np.ascontiguousarray(d.view(np.uint8).reshape(-1))
            keep.append(a)
            return a.ctypes.data if a.size else None
    except ImportError:  # pragma: no cover
        pass
    b = bytes(d) if not isinstance(d, bytes) else d
    keep.append(b)
    if not b:
        return None
    return ctypes.cast(ctypes.c_char_p(b), ctypes.c_void_p).value


def _to_c(entries: Sequence[Entry]):
    keep: list = []
    arr = (_CEntry * max(1, len(entries)))()
    for i, e in enumerate(entries):
        c = arr[i]
        keep.extend((e.name, e.uname, e.gname, e.linkname))
        c.linkname, c.linkname_len = e.linkname, len(e.linkname)
        c.name, c.name_len = e.name, len(e.name)
        c.uname, c.uname_len = e.uname, len(e.uname)
        c.gname, c.gname_len = e.gname, len(e.gname)
        c.size, c.mtime_sec, c.mtime_nsec = e.size, e.mtime_sec, e.mtime_nsec
        c.mode, c.uid, c.gid, c.typeflag = e.mode, e.uid, e.gid, e.typeflag
        c.data = _data_ptr(e.data, keep) if e.typeflag == ord("0") and e.size > 0 else None
    return arr, keep


def header_bytes(e: Entry) -> int:
    lib = _load()
    arr, _keep = _to_c([e])
    out = ctypes.c_size_t()
    rc = lib.gotar_header_bytes(arr, ctypes.byref(out))
    if rc:
        raise OracleError(rc)
    return out.value


def stream_size(entries: Sequence[Entry]) -> int:
    lib = _load()
    arr, _keep = _to_c(entries)
    out = ctypes.c_size_t()
    rc = lib.gotar_stream_size(arr, len(entries), ctypes.byref(out))
    if rc:
        raise OracleError(rc)
    return out.value


def tar_stream(entries: Sequence[Entry]) -> bytes:
    """The byte stream archive/tar's Writer emits for these headers + payloads + Close()."""
    lib = _load()
    arr, _keep = _to_c(entries)
    total = ctypes.c_size_t()
    rc = lib.gotar_stream_size(arr, len(entries), ctypes.byref(total))
    if rc:
        raise OracleError(rc)
    buf = ctypes.create_string_buffer(total.value)
    written = ctypes.c_size_t()
    rc = lib.gotar_write(arr, len(entries), buf, total.value, ctypes.byref(written))
    if rc:
        raise OracleError(rc)
    assert written.value == total.value
    return buf.raw


def tar_stream_update(entries: Sequence[Entry], update: Callable[[memoryview], None], trailer: bool) -> int:
    """Feeds the stream of `entries` to `update` (e.g. a hashlib object's update) without materialising a bytes copy;
    the 1024-byte trailer of Writer.Close() only when `trailer`. For streams too large to hold twice: the caller
    walks the header list in groups. Returns the bytes fed."""
    lib = _load()
    arr, _keep = _to_c(entries)
    total = ctypes.c_size_t()
    rc = lib.gotar_stream_size(arr, len(entries), ctypes.byref(total))
    if rc:
        raise OracleError(rc)
    buf = ctypes.create_string_buffer(total.value)
    written = ctypes.c_size_t()
    rc = lib.gotar_write(arr, len(entries), buf, total.value, ctypes.byref(written))
    if rc:
        raise OracleError(rc)
    n = total.value if trailer else total.value - 1024
    update(memoryview(buf)[:n])
    return n


def pack_shards_mt(entries: Sequence[Entry], bounds: Sequence[int], threads: int, level: int = 6):
    """One archive per shard [bounds[s], bounds[s+1]) of `entries`, packed by a pool of `threads` pthreads inside the
    oracle (no interpreter lock in the loop). Returns (tar_bytes[], gz_bytes[], worker_seconds[]) per shard."""
    lib = _load()
    arr, _keep = _to_c(entries)
    ns = len(bounds) - 1
    SZ = ctypes.c_size_t * max(ns, 1)
    lo, hi = SZ(*bounds[:-1]), SZ(*bounds[1:])
    tb, gb = SZ(), SZ()
    sec = (ctypes.c_double * max(ns, 1))()
    rc = lib.gooracle_pack_shards_mt(arr, lo, hi, ns, level, threads, tb, gb, sec)
    if rc:
        raise OracleError(rc)
    return list(tb)[:ns], list(gb)[:ns], list(sec)[:ns]


def path_clean(p: bytes) -> bytes:
    lib = _load()
    out = ctypes.create_string_buffer(len(p) + 4)
    n = lib.gotar_path_clean(p, len(p), out)
    return out.raw[:n]


def crc32(data) -> int:
    lib = _load()
    keep: list = []
    ptr = _data_ptr(data, keep)
    n = len(keep[0]) if keep else 0
    return lib.gogzip_crc32(ptr, n)


def gzip_pack(data, level: int = 6) -> bytes:
    """gzip.NewWriter(f) (tar.go:156) + Write(data) + Close(), deflate body from zlib at `level`."""
    lib = _load()
    keep: list = []
    ptr = _data_ptr(data, keep)
    n = len(keep[0]) if keep else 0
    cap = lib.gogzip_bound(n)
    out = ctypes.create_string_buffer(cap)
    written = ctypes.c_size_t()
    rc = lib.gogzip_pack(ptr, n, level, out, cap, ctypes.byref(written))
    if rc:
        raise OracleError(rc)
    return out.raw[: written.value]


def gzip_pack_size(data, level: int = 6, scratch=None) -> int:
    """Size only (for large buffers: avoids the bytes copy)."""
    lib = _load()
    keep: list = []
    ptr = _data_ptr(data, keep)
    n = len(keep[0]) if keep else 0
    cap = lib.gogzip_bound(n)
    out = scratch if scratch is not None and len(scratch) >= cap else ctypes.create_string_buffer(cap)
    written = ctypes.c_size_t()
    rc = lib.gogzip_pack(ptr, n, level, out, cap, ctypes.byref(written))
    if rc:
        raise OracleError(rc)
    return written.value


def pack(entries: Sequence[Entry], level: int = 6) -> bytes:
    """writeTar's archive bytes for an already-walked header list: gzip(tar stream)."""
    return gzip_pack(tar_stream(entries), level)


# ------------------------------------------------------------------------------------------------
# Traversal restatement: writeTar / recursiveTar / tarFolder / tarFile  (tar.go:146-304)
# ------------------------------------------------------------------------------------------------

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


def round_mtime(sec: int, nsec: int) -> int:
    """roundMtime (util.go:87-89): time.Round(time.Second).Unix(), half rounds up."""
    return sec + (1 if nsec >= 500_000_000 else 0)


def go_path_join(a: str, b: str) -> str:
    """path.Join for two elements (tar.go:186,232,261,270): skip empties, join with '/', Clean."""
    parts = [p for p in (a, b) if p != ""]
    if not parts:
        return ""
    return path_clean("/".join(parts).encode("utf-8", "surrogateescape")).decode("utf-8", "surrogateescape")


def _go_file_mode(st: os.stat_result) -> int:
    """tar.FileInfoHeader: Mode = perm | c_ISUID | c_ISGID | c_ISVTX (no file-type bits)."""
    m = _stat.S_IMODE(st.st_mode) & 0o777
    if st.st_mode & _stat.S_ISUID:
        m |= 0o4000
    if st.st_mode & _stat.S_ISGID:
        m |= 0o2000
    if st.st_mode & _stat.S_ISVTX:
        m |= 0o1000
    return m


def walk(
    files: Iterable[FileInformation],
    watch_path: str,
    file_map: Optional[Dict[str, FileInformation]] = None,
    is_excluded: Optional[Callable[[str], bool]] = None,
    lookup_user: Optional[Callable[[int], str]] = None,
    lookup_group: Optional[Callable[[int], str]] = None,
    read_data: bool = True,
) -> Tuple[List[Entry], Dict[str, FileInformation]]:
    """Ordered header list + writtenFiles map that writeTar (tar.go:146-182) produces.

    ``is_excluded(relpath)`` stands for ignoreMatcher / uploadIgnoreMatcher (tar.go:192-207).
    """
    import grp
    import pwd

    file_map = file_map or {}
    written: Dict[str, FileInformation] = {}
    entries: List[Entry] = []

    def uname_of(uid: int) -> str:
        if lookup_user:
            return lookup_user(uid)
        try:
            return pwd.getpwuid(uid).pw_name
        except KeyError:
            return ""

    def gname_of(gid: int) -> str:
        if lookup_group:
            return lookup_group(gid)
        try:
            return grp.getgrgid(gid).gr_name
        except KeyError:
            return ""

    def info_from_stat(rel: str, st: os.stat_result) -> FileInformation:          # tar.go:306-324
        fi = FileInformation(Name=rel, Size=st.st_size,
                             Mtime=round_mtime(st.st_mtime_ns // 10**9, st.st_mtime_ns % 10**9),
                             IsDirectory=_stat.S_ISDIR(st.st_mode))
        known = file_map.get(rel)
        if known is not None:
            fi.RemoteMode, fi.RemoteGID, fi.RemoteUID = known.RemoteMode, known.RemoteGID, known.RemoteUID
        return fi

    def header(fi: FileInformation, st: os.stat_result, typeflag: str, data: bytes) -> Entry:
        mode, uid, gid = _go_file_mode(st), st.st_uid, st.st_gid
        e = Entry(name=fi.Name, size=len(data) if typeflag == "0" else 0,
                  mtime_sec=st.st_mtime_ns // 10**9, mtime_nsec=st.st_mtime_ns % 10**9,
                  mode=mode, uid=uid, gid=gid, uname=uname_of(uid), gname=gname_of(gid),
                  typeflag=ord(typeflag), data=data)
        if fi.Name in file_map:                                                   # tar.go:246-250,287-291
            e.mode, e.uid, e.gid = fi.RemoteMode, fi.RemoteUID, fi.RemoteGID      # Uname/Gname NOT overridden
        return e

    def recursive_tar(rel: str) -> None:                                           # tar.go:185-229
        if rel in written:
            return
        if is_excluded is not None and is_excluded(rel):
            return
        absp = go_path_join(watch_path, rel)
        try:
            st = os.stat(absp)                                                     # follows symlinks
        except OSError:
            return                                                                 # vanished: skipped
        fi = info_from_stat(rel, st)
        if _stat.S_ISDIR(st.st_mode):
            tar_folder(fi, st)
        else:
            tar_file(fi, st)

    def tar_folder(fi: FileInformation, st: os.stat_result) -> None:               # tar.go:231-267
        p = go_path_join(watch_path, fi.Name)
        try:
            names = sorted(os.listdir(p), key=lambda s: s.encode("utf-8", "surrogateescape"))  # ioutil.ReadDir sorts
        except OSError:
            return
        if not names and fi.Name != "":
            entries.append(header(fi, st, "5", b""))
            written[fi.Name] = fi
        for n in names:
            recursive_tar(go_path_join(fi.Name, n))

    def tar_file(fi: FileInformation, st: os.stat_result) -> None:                 # tar.go:269-304
        p = go_path_join(watch_path, fi.Name)
        with open(p, "rb") as f:
            data = f.read() if read_data else b"\0" * st.st_size
        if not _stat.S_ISREG(st.st_mode):
            raise OracleError(-2)
        entries.append(header(fi, st, "0", data))
        written[fi.Name] = fi

    for element in files:                                                          # tar.go:164-179
        if element.Name not in written:
            recursive_tar(element.Name)
    return entries, written
