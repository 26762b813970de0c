"""GPU: the tar kernels through the C ABI (dsg_dev_tar_build / dsg_pack_tar_only) against the oracle and the
frozen golden vectors. Bit-exact: byte and integer work."""
import gzip
import json
import os

import numpy as np
import pytest

import helpers
import oracle
from devspace_b200.manifest import from_entries

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
VEC = json.load(open(os.path.join(GOLD, "vectors.json")))
SUITE = helpers.edge_suite()


@pytest.fixture(scope="module")
def packer():
    from devspace_b200.pack import Packer
    p = Packer(level=0)
    yield p
    p.close()


@pytest.mark.parametrize("name", sorted(SUITE))
def test_tar_stream_matches_oracle_and_golden(packer, name):
    es = SUITE[name]
    got = packer.tar_stream(from_entries(es))
    assert helpers.sha256(got) == VEC[name]["sha256"]          # frozen vector (no oracle needed on the box)
    assert got == oracle.tar_stream(es)                          # and the live oracle on the same inputs


def test_c1_quickstart_golden(packer):
    t = gzip.decompress(open(os.path.join(GOLD, "c1_quickstart.tgz"), "rb").read())
    members = helpers.decode_with_tarfile(t)
    es = [helpers.F(m.name, d, mode=m.mode, uid=m.uid, gid=m.gid, uname=m.uname, gname=m.gname, mtime=m.mtime)
          for m, d in members]
    assert packer.tar_stream(from_entries(es)) == t
    assert packer.pack(from_entries(es), tar_only=True) == t


@pytest.mark.parametrize("case", ["name_101_nosplit", "uid_2097152"])
def test_kernels_match_hand_derived_pax_bytes(packer, case):
    """The PAX path of the header kernels against bytes derived by hand from the go1.11 source, independent of the oracle
    (tests/golden/make_pax_hand.py)."""
    import json
    c = json.load(open(os.path.join(GOLD, "pax_hand_derived.json")))[case]
    m = c["manifest"]
    es = [helpers.F(m["name"], m["data"].encode(), mode=m["mode"], uid=m["uid"], gid=m["gid"], uname=m["uname"],
                    gname=m["gname"], mtime=m["mtime"])]
    assert packer.tar_stream(from_entries(es)) == bytes.fromhex(c["stream_hex"])


def test_no_trailer_flag(packer):
    es = SUITE["mixed_tree"]
    assert packer.tar_stream(from_entries(es), flags=1) == oracle.tar_stream(es)[:-1024]


@pytest.mark.parametrize("name,code", [("trailing_slash_file", -3), ("name_with_nul", -3), ("mode_too_big", -3),
                                       ("bad_typeflag", -4)])
def test_error_codes(packer, name, code):
    from devspace_b200._lib import DsgError
    with pytest.raises(DsgError) as ei:
        packer.tar_stream_checked = packer.pack(from_entries(helpers.ERROR_CASES[name]), tar_only=True)
    assert ei.value.status == code


def test_payload_out_of_arena_is_rejected(packer):
    from devspace_b200._lib import DsgError
    man = from_entries(SUITE["sizes_511_512_513"])
    man.entries["data_off"][2] = 10_000_000
    with pytest.raises(DsgError) as ei:
        packer.pack(man, tar_only=True)
    assert ei.value.status == -8


def test_many_small_entries_and_unaligned_arena(packer):
    """100k-entry style: lots of one-block entries; arena offsets deliberately NOT 16-byte aligned."""
    from devspace_b200.manifest import ManifestBuilder
    rng = np.random.default_rng(11)
    b = ManifestBuilder(arena_align=1)
    es = []
    for i in range(5000):
        sz = int(rng.integers(0, 1500))
        e = helpers.F("pkg%04d/f%05d.js" % (i // 40, i), helpers.text_bytes(i, sz), mtime=helpers.MTIME0 + i)
        es.append(e)
        b.add_entry(e)
    got = packer.pack(b.build(), tar_only=True)
    assert got == oracle.tar_stream(es)


def test_large_file_spans_many_tiles(packer):
    es = [helpers.F("a.bin", helpers.rng_bytes(1, 3_000_001)), helpers.F("b.bin", helpers.rng_bytes(2, 1_234_567))]
    assert packer.pack(from_entries(es), tar_only=True) == oracle.tar_stream(es)
