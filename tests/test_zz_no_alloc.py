"""GPU: dsg_crc32_noalloc -- the CRC32 entry point for callers that cannot allocate device memory at call time
(a process whose GPU memory is fully handed out at start-up, or a loop that cannot take the allocator's
device-wide synchronisation). It keeps its per-tile partial results in a static device array inside the library
instead of a cudaMalloc'd buffer, so this test checks the thing that can go wrong with a shared, never-cleared
scratch: two calls in a row, the second on a SHORTER and different buffer, must both equal the regular,
allocating entry point (dsg_dev_crc32) bit for bit -- CRC work is integer work, the tolerance is zero.
The file sorts last on purpose: a run that stops at the first failure has then already run every other test.
It must pass both alone in a fresh process (first call = one-time set-up) and after the rest of the suite."""
import ctypes
import zlib

import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def packer():
    from devspace_b200.pack import Packer
    p = Packer(level=0)
    yield p
    p.close()


def _noalloc(lib, d_ptr, n):
    crc = ctypes.c_uint32(0xDEADBEEF)
    err = ctypes.create_string_buffer(256)
    rc = lib.dsg_crc32_noalloc(ctypes.c_void_p(d_ptr), n, ctypes.byref(crc), err, len(err))
    return rc, crc.value, err.value.decode()


def test_two_calls_in_a_row_match_the_allocating_entry_point(packer):
    from devspace_b200.pack import DeviceBuffer
    lib = packer._lib
    first = np.frombuffer(helpers.rng_bytes(71, 9_000_001), np.uint8)       # 69 tiles + an unaligned tail
    second = np.frombuffer(helpers.text_bytes(72, 300_007), np.uint8)       # shorter: stale tile words stay behind
    d1 = DeviceBuffer(packer, first.size + 64).upload(first)
    d2 = DeviceBuffer(packer, second.size + 64).upload(second)
    try:
        want1 = packer.crc32_device(d1, first.size)                         # existing entry point (allocates)
        want2 = packer.crc32_device(d2, second.size)
        assert want1 == zlib.crc32(first.tobytes()) and want2 == zlib.crc32(second.tobytes())
        rc, got1, err = _noalloc(lib, d1.ptr, first.size)
        assert rc == 0, err
        rc, got2, err = _noalloc(lib, d2.ptr, second.size)                  # second call straight after the first
        assert rc == 0, err
        assert got1 == want1
        assert got2 == want2
        # and once more in the other order, with an unaligned device pointer
        rc, got3, err = _noalloc(lib, d1.ptr + 3, first.size - 10)
        assert rc == 0, err
        assert got3 == packer.crc32_device(d1, first.size - 10, offset=3)
    finally:
        d1.free()
        d2.free()


def test_rejects_inputs_beyond_the_static_scratch(packer):
    lib = packer._lib
    rc, _, err = _noalloc(lib, 0x1000, (1 << 30) + 1)      # the size is checked before the pointer is touched
    assert rc == -5 and "limit" in err and "1073741824" in err
    rc, crc, _ = _noalloc(lib, 0, 0)                        # empty input: CRC 0, nothing launched
    assert rc == 0 and crc == 0
