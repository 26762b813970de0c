"""ctypes binding of libdsgpack.so (include/dsgpack.h). Loading fails loudly when the CUDA library has
not been built: there is no CPU fallback anywhere in this package."""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, byref, c_char_p, c_float, c_int, c_int32, c_size_t, c_uint32, c_uint64, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdsgpack.so")

DSG_OK = 0
DSG_E_ARG, DSG_E_CUDA, DSG_E_HEADER, DSG_E_TYPEFLAG = -1, -2, -3, -4
DSG_E_CAPACITY, DSG_E_UNCLEAN_NAME, DSG_E_NOMEM, DSG_E_RANGE, DSG_E_CORRUPT = -5, -6, -7, -8, -9
DSG_LEVEL_STORED, DSG_LEVEL_DEFAULT = 0, 6
DSG_F_NO_TRAILER = 1
DSG_LZ_TAIL_SET, DSG_LZ_PROFILE, DSG_LZ_NO_BULK = 1, 2, 4

# every symbol include/dsgpack.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "dsg_abi_version", "dsg_strerror", "dsg_device_count", "dsg_last_error", "dsg_ctx_create", "dsg_ctx_destroy",
    "dsg_ctx_set_opts",
    "dsg_pinned_alloc", "dsg_pinned_free", "dsg_entry_header_bytes", "dsg_tar_size", "dsg_gzip_bound",
    "dsg_shard_bounds", "dsg_crc32_combine", "dsg_stitch_members", "dsg_pack", "dsg_pack_tar_only", "dsg_pack_fd", "dsg_pack_device", "dsg_get_stats", "dsg_dev_tar_build",
    "dsg_dev_crc32", "dsg_dev_deflate", "dsg_dev_gzip", "dsg_dev_alloc", "dsg_dev_free", "dsg_dev_upload",
    "dsg_dev_download", "dsg_crc32_noalloc", "dsg_file_crc32", "dsg_dev_file_crc32", "dsg_directory_hash", "dsg_docker_context_tar", "dsg_walk", "dsg_written_count", "dsg_written_get",
    "dsg_written_entries", "dsg_written_free", "dsg_write_tar",
    "dsg_multi_create", "dsg_multi_destroy", "dsg_multi_last_error", "dsg_multi_shards", "dsg_multi_get_stats",
    "dsg_multi_pack_fd", "dsg_multi_pack", "dsg_gunzip", "dsg_dev_gunzip",
]
DSG_MULTI_ONE_MEMBER = 1


class DsgOpts(ctypes.Structure):
    _fields_ = [("struct_size", c_uint32), ("level", c_int32), ("segment_bytes", c_uint64), ("flags", c_uint32),
                ("gunzip_batch_mib", c_uint32), ("lz_chunk_units", c_uint32), ("lz_run_chunks", c_uint32),
                ("lz_grid", c_uint32), ("lz_flags", c_uint32), ("lz_tail_bytes", c_uint64)]


class DsgStats(ctypes.Structure):
    _fields_ = [("entries", c_uint64), ("payload_bytes", c_uint64), ("tar_bytes", c_uint64), ("out_bytes", c_uint64),
                ("h2d_bytes", c_uint64), ("d2h_bytes", c_uint64), ("ms_total", c_float), ("ms_plan", c_float),
                ("ms_tar", c_float), ("ms_deflate", c_float), ("ms_crc", c_float), ("ms_gather", c_float),
                ("kernel_launches", c_uint32), ("segments", c_uint32), ("crc32", c_uint32), ("reserved", c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


class DsgError(RuntimeError):
    def __init__(self, status: int, text: str):
        super().__init__(f"dsgpack status {status}: {text}")
        self.status = status


_lib = None


def load(build_if_missing: bool = True):
    """Returns the loaded library; builds it with nvcc when missing/stale and a compiler is present."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_missing:
        from . import build as _build
        try:
            if _build.is_stale():
                _build.build()
        except RuntimeError:
            if not os.path.exists(LIB_PATH):
                raise
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m devspace_b200.build` "
                           "(the CUDA library is the only implementation; there is no CPU fallback)")
    lib = ctypes.CDLL(LIB_PATH)
    vp, sz, u64p = c_void_p, c_size_t, POINTER(c_uint64)
    lib.dsg_abi_version.restype = c_int
    lib.dsg_strerror.argtypes = [c_int]
    lib.dsg_strerror.restype = c_char_p
    lib.dsg_device_count.restype = c_int
    lib.dsg_last_error.argtypes = [vp]
    lib.dsg_last_error.restype = c_char_p
    lib.dsg_ctx_create.argtypes = [c_int, POINTER(DsgOpts), POINTER(vp)]
    lib.dsg_ctx_destroy.argtypes = [vp]
    lib.dsg_ctx_destroy.restype = None
    lib.dsg_ctx_set_opts.argtypes = [vp, POINTER(DsgOpts)]
    lib.dsg_pinned_alloc.argtypes = [vp, sz]
    lib.dsg_pinned_alloc.restype = vp
    lib.dsg_pinned_free.argtypes = [vp, vp]
    lib.dsg_pinned_free.restype = None
    lib.dsg_entry_header_bytes.argtypes = [vp, vp, sz, u64p]
    lib.dsg_tar_size.argtypes = [vp, sz, vp, sz, c_uint32, u64p]
    lib.dsg_gzip_bound.argtypes = [c_uint64]
    lib.dsg_gzip_bound.restype = c_uint64
    lib.dsg_shard_bounds.argtypes = [vp, sz, vp, sz, c_int, u64p]
    pack_args = [vp, vp, sz, vp, sz, vp, sz, vp, sz, POINTER(sz)]
    lib.dsg_pack.argtypes = pack_args
    lib.dsg_pack_tar_only.argtypes = pack_args
    lib.dsg_pack_device.argtypes = pack_args
    lib.dsg_pack_fd.argtypes = [vp, vp, sz, vp, sz, vp, sz, c_int, c_int, u64p]
    lib.dsg_get_stats.argtypes = [vp, POINTER(DsgStats)]
    lib.dsg_dev_tar_build.argtypes = [vp, vp, sz, vp, sz, vp, sz, c_uint32, vp, sz, u64p]
    lib.dsg_dev_crc32.argtypes = [vp, vp, sz, POINTER(c_uint32)]
    lib.dsg_dev_deflate.argtypes = [vp, vp, sz, c_int, vp, sz, POINTER(sz)]
    lib.dsg_dev_gzip.argtypes = [vp, vp, sz, c_int, vp, sz, POINTER(sz)]
    lib.dsg_dev_alloc.argtypes = [vp, sz]
    lib.dsg_dev_alloc.restype = vp
    lib.dsg_dev_free.argtypes = [vp, vp]
    lib.dsg_dev_free.restype = None
    lib.dsg_dev_upload.argtypes = [vp, vp, vp, sz]
    lib.dsg_dev_download.argtypes = [vp, vp, vp, sz]
    lib.dsg_crc32_noalloc.argtypes = [vp, sz, POINTER(c_uint32), c_char_p, sz]
    lib.dsg_file_crc32.argtypes = [vp, vp, sz, vp, sz, vp]
    lib.dsg_dev_file_crc32.argtypes = [vp, vp, sz, vp, sz, vp]
    lib.dsg_gunzip.argtypes = [vp, vp, sz, vp, sz, POINTER(sz)]
    lib.dsg_dev_gunzip.argtypes = [vp, vp, sz, vp, sz, POINTER(sz)]
    lib.dsg_crc32_combine.argtypes = [c_uint32, c_uint32, c_uint64]
    lib.dsg_crc32_combine.restype = c_uint32
    lib.dsg_multi_create.argtypes = [POINTER(c_int), c_int, POINTER(DsgOpts), POINTER(vp)]
    lib.dsg_multi_destroy.argtypes = [vp]
    lib.dsg_multi_destroy.restype = None
    lib.dsg_multi_last_error.argtypes = [vp]
    lib.dsg_multi_last_error.restype = c_char_p
    lib.dsg_multi_shards.argtypes = [vp]
    lib.dsg_multi_get_stats.argtypes = [vp, c_int, POINTER(DsgStats)]
    lib.dsg_multi_pack_fd.argtypes = [vp, vp, sz, vp, sz, vp, sz, c_int, c_uint32, u64p]
    lib.dsg_multi_pack.argtypes = [vp, vp, sz, vp, sz, vp, sz, c_uint32, vp, sz, POINTER(sz)]
    _lib = lib
    return lib
