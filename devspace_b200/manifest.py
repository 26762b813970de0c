"""Manifest: the list of tar headers + payload arena that crosses the C ABI (include/dsgpack.h).

Host-side mirror of what ``tarFile`` / ``tarFolder`` (``/root/reference/pkg/devspace/sync/tar.go:231-304``)
hand to ``tw.WriteHeader`` and ``io.Copy``: one ``dsg_entry`` per header, strings in a shared string
table, file bytes in one arena. The Go shim builds exactly this while it walks.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterable, List, Optional, Sequence

import numpy as np

# struct dsg_entry (include/dsgpack.h), 80 bytes, little endian
ENTRY_DTYPE = np.dtype(
    [
        ("data_off", "<u8"),
        ("size", "<i8"),
        ("mtime_sec", "<i8"),
        ("mode", "<i8"),
        ("uid", "<i8"),
        ("gid", "<i8"),
        ("mtime_nsec", "<u4"),
        ("name_off", "<u4"),
        ("name_len", "<u4"),
        ("uname_off", "<u4"),
        ("gname_off", "<u4"),
        ("uname_len", "<u2"),
        ("gname_len", "<u2"),
        ("typeflag", "u1"),
        ("reserved", "u1", (7,)),
    ],
    align=False,
)
assert ENTRY_DTYPE.itemsize == 80

TYPE_REG = ord("0")
TYPE_DIR = ord("5")
TYPE_LINK = ord("1")      # hard link   } docker build-context dialect only: no payload, the target
TYPE_SYMLINK = ord("2")   # symbolic link }   is a strtab reference in data_off (offset | length << 32)
ARENA_ALIGN = 16  # payload starts are 16-byte aligned so the layout kernel moves 128-bit words


def _b(s) -> bytes:
    return s.encode("utf-8", "surrogateescape") if isinstance(s, str) else bytes(s)


@dataclass
class Manifest:
    """Arrays ready to hand to ``dsg_pack``: entries (ENTRY_DTYPE), strtab (uint8), arena (uint8)."""

    entries: np.ndarray
    strtab: np.ndarray
    arena: np.ndarray

    @property
    def n(self) -> int:
        return int(self.entries.shape[0])

    @property
    def payload_bytes(self) -> int:
        reg = self.entries["typeflag"] == TYPE_REG
        return int(self.entries["size"][reg].sum())

    def name(self, i: int) -> bytes:
        e = self.entries[i]
        return self.strtab[int(e["name_off"]): int(e["name_off"]) + int(e["name_len"])].tobytes()

    def payload(self, i: int) -> np.ndarray:
        e = self.entries[i]
        if e["typeflag"] != TYPE_REG:
            return self.arena[:0]
        return self.arena[int(e["data_off"]): int(e["data_off"]) + int(e["size"])]

    def slice(self, lo: int, hi: int) -> "Manifest":
        """Entries [lo, hi) with the same strtab / arena (shards share the staging buffers)."""
        return Manifest(self.entries[lo:hi], self.strtab, self.arena)


class ManifestBuilder:
    """Accumulates headers the way the Go shim would while walking (one ``add`` per WriteHeader)."""

    def __init__(self, arena_align: int = ARENA_ALIGN):
        self._rows: List[tuple] = []
        self._strs: dict = {}
        self._strtab = bytearray()
        self._chunks: List[np.ndarray] = []
        self._arena_len = 0
        self._align = arena_align

    def _intern(self, s: bytes):
        hit = self._strs.get(s)
        if hit is None:
            hit = len(self._strtab)
            self._strtab += s
            self._strs[s] = hit
        return hit

    def _name(self, s: bytes) -> int:
        off = len(self._strtab)  # names are unique: no interning dictionary growth
        self._strtab += s
        return off

    def add(self, name, size: int = 0, mtime_sec: int = 0, mtime_nsec: int = 0, mode: int = 0o644, uid: int = 0,
            gid: int = 0, uname=b"", gname=b"", typeflag: int = TYPE_REG, data=None, linkname=b"") -> None:
        name, uname, gname = _b(name), _b(uname), _b(gname)
        data_off = 0
        if typeflag in (TYPE_LINK, TYPE_SYMLINK):  # the link target travels in data_off as a strtab reference
            linkname = _b(linkname)
            data_off = self._name(linkname) | (len(linkname) << 32)
        if typeflag == TYPE_REG and size > 0:
            pad = (-self._arena_len) % self._align
            if pad:
                self._chunks.append(np.zeros(pad, np.uint8))
                self._arena_len += pad
            a = np.frombuffer(data, np.uint8) if not isinstance(data, np.ndarray) else data.view(np.uint8).reshape(-1)
            if a.size != size:
                raise ValueError(f"payload of {name!r}: {a.size} bytes, header says {size}")
            data_off = self._arena_len
            self._chunks.append(a)
            self._arena_len += size
        self._rows.append((data_off, size, mtime_sec, mode, uid, gid, mtime_nsec, self._name(name), len(name),
                           self._intern(uname), self._intern(gname), len(uname), len(gname), typeflag))

    def add_entry(self, e) -> None:
        """Accepts any object with the oracle.Entry attribute names (duck typed; no import of oracle)."""
        self.add(e.name, e.size, e.mtime_sec, e.mtime_nsec, e.mode, e.uid, e.gid, e.uname, e.gname, e.typeflag,
                 e.data if e.typeflag == TYPE_REG and e.size > 0 else None, getattr(e, "linkname", b""))

    def build(self) -> Manifest:
        ent = np.zeros(len(self._rows), ENTRY_DTYPE)
        cols = ("data_off", "size", "mtime_sec", "mode", "uid", "gid", "mtime_nsec", "name_off", "name_len",
                "uname_off", "gname_off", "uname_len", "gname_len", "typeflag")
        if self._rows:
            for j, c in enumerate(cols):
                ent[c] = [r[j] for r in self._rows]
        strtab = np.frombuffer(bytes(self._strtab), np.uint8).copy() if self._strtab else np.zeros(1, np.uint8)
        arena = np.concatenate(self._chunks) if self._chunks else np.zeros(0, np.uint8)
        return Manifest(ent, strtab, arena)


def from_entries(entries: Iterable) -> Manifest:
    b = ManifestBuilder()
    for e in entries:
        b.add_entry(e)
    return b.build()
