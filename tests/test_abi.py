"""CPU: libdsgpack.so loads and exports every symbol include/dsgpack.h declares; the host-side layout helpers
(pure integer work on the manifest) agree with the oracle; and without a GPU the compute path refuses to run
(no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

import helpers
import oracle
from devspace_b200 import _lib
from devspace_b200.manifest import ENTRY_DTYPE, from_entries

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SUITE = helpers.edge_suite()


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "dsgpack.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(dsg_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    declared = _declared_symbols()
    assert len(declared) >= 20
    for s in declared:
        assert hasattr(lib, s), f"{s} declared in include/dsgpack.h but not exported"
    assert sorted(_lib.SYMBOLS) == declared
    assert lib.dsg_abi_version() == 2


def test_entry_struct_layout_is_80_bytes():
    assert ENTRY_DTYPE.itemsize == 80
    assert ENTRY_DTYPE.fields["mtime_nsec"][1] == 48 and ENTRY_DTYPE.fields["typeflag"][1] == 72


def _tar_size(lib, man, flags=0):
    out = ctypes.c_uint64()
    rc = lib.dsg_tar_size(man.entries.ctypes.data if man.n else None, man.n, man.strtab.ctypes.data, man.strtab.size,
                          flags, ctypes.byref(out))
    return rc, out.value


@pytest.mark.parametrize("name", sorted(SUITE))
def test_host_tar_size_and_header_bytes_match_oracle(name):
    lib = _lib.load()
    es = SUITE[name]
    man = from_entries(es)
    rc, T = _tar_size(lib, man)
    assert rc == 0 and T == oracle.stream_size(es)
    rc, T2 = _tar_size(lib, man, _lib.DSG_F_NO_TRAILER)
    assert rc == 0 and T2 == T - 1024
    for i, e in enumerate(es[:40]):
        out = ctypes.c_uint64()
        rc = lib.dsg_entry_header_bytes(man.entries[i:i + 1].ctypes.data, man.strtab.ctypes.data, man.strtab.size,
                                        ctypes.byref(out))
        assert rc == 0 and out.value == oracle.header_bytes(e)


@pytest.mark.parametrize("name,code", [("trailing_slash_file", -3), ("name_with_nul", -3), ("mode_too_big", -3),
                                       ("bad_typeflag", -4)])
def test_host_helpers_report_header_errors(name, code):
    lib = _lib.load()
    rc, _ = _tar_size(lib, from_entries(helpers.ERROR_CASES[name]))
    assert rc == code


def test_shard_bounds_are_contiguous_and_balanced():
    lib = _lib.load()
    es = SUITE["mixed_tree"]
    man = from_entries(es)
    for g in (1, 2, 4, 8):
        b = np.zeros(g + 1, np.uint64)
        rc = lib.dsg_shard_bounds(man.entries.ctypes.data, man.n, man.strtab.ctypes.data, man.strtab.size, g,
                                  b.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)))
        assert rc == 0
        assert b[0] == 0 and b[-1] == man.n and all(b[i] <= b[i + 1] for i in range(g))
        # concatenating the shards' tar streams (trailer only on the last) gives the whole stream
        parts = []
        for s in range(g):
            t = oracle.tar_stream(es[int(b[s]):int(b[s + 1])])
            parts.append(t if s == g - 1 else t[:-1024])
        assert b"".join(parts) == oracle.tar_stream(es)


def test_no_cpu_fallback_without_a_gpu():
    lib = _lib.load()
    if lib.dsg_device_count() > 0:
        pytest.skip("a CUDA device is present")
    ctx = ctypes.c_void_p()
    assert lib.dsg_ctx_create(0, None, ctypes.byref(ctx)) == _lib.DSG_E_CUDA
    assert not ctx.value
    from devspace_b200.pack import Packer
    with pytest.raises(_lib.DsgError):
        Packer()
