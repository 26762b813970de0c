"""CPU: the product's tar kernels (devspace_b200/csrc/tar_kernels.cuh), compiled for the SIMT emulator, against
the oracle. This checks kernel LOGIC (plan, scan, layout tiles, PAX clipping across segments) without a GPU;
the real parity tests are the -m gpu ones that go through libdsgpack.so."""
import pytest

import helpers
import oracle
from devspace_b200.manifest import from_entries

SUITE = helpers.edge_suite()


@pytest.mark.parametrize("name", sorted(SUITE))
def test_emulated_tar_kernels_match_oracle(name):
    es = SUITE[name]
    ref = oracle.tar_stream(es)
    man = from_entries(es)
    rc, got, _ = helpers.emu_tar_build(man)
    assert rc == 0
    assert got == ref


@pytest.mark.parametrize("name", ["mixed_tree", "name_5000", "pax_multi", "util_test_tree"])
@pytest.mark.parametrize("seg", [32768, 65536])
def test_emulated_segmented_layout(name, seg):
    """Pipeline segments cut the stream at arbitrary 32 KiB multiples, also inside a PAX header area."""
    es = SUITE[name]
    if name == "name_5000":   # put the 7 KiB PAX area across the first segment boundary
        es = [helpers.F("pad%d" % i, helpers.rng_bytes(i, 4096)) for i in range(6)] + es + es
    ref = oracle.tar_stream(es)
    rc, got, _ = helpers.emu_tar_build(from_entries(es), segment_bytes=seg)
    assert rc == 0 and got == ref


def test_emulated_no_trailer_flag():
    es = SUITE["util_test_tree"]
    rc, got, _ = helpers.emu_tar_build(from_entries(es), flags=1)
    assert rc == 0 and got == oracle.tar_stream(es)[:-1024]


@pytest.mark.parametrize("name,code", [("trailing_slash_file", -3), ("name_with_nul", -3), ("mode_too_big", -3),
                                       ("bad_typeflag", -4)])
def test_emulated_error_codes(name, code):
    rc, _, err_entry = helpers.emu_tar_build(from_entries(helpers.ERROR_CASES[name]))
    assert rc == code and err_entry == 0


def test_emulated_first_error_is_lowest_entry():
    es = [helpers.F("ok%d" % i, b"x") for i in range(300)]
    es[270] = helpers.F("bad/", b"")
    es[40] = helpers.F("f", b"x", mode=0o10000000)
    rc, _, err_entry = helpers.emu_tar_build(from_entries(es))
    assert rc == -3 and err_entry == 40
