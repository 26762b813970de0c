"""CPU: the oracle (go1.11 archive/tar restatement) against the frozen golden vectors and against two
independent implementations of the format (CPython tarfile, GNU tar). Upstream has no golden bytes for this
path: parity with Go itself is UNPINNED; these checks pin everything that can be pinned here."""
import gzip
import io
import json
import os
import shutil
import subprocess
import tarfile

import pytest

import helpers
import oracle
from helpers import D, F, MTIME0

GOLD = os.path.join(os.path.dirname(__file__), "golden")
VEC = json.load(open(os.path.join(GOLD, "vectors.json")))
SUITE = helpers.edge_suite()


@pytest.mark.parametrize("name", sorted(SUITE))
def test_golden_vectors(name):
    t = oracle.tar_stream(SUITE[name])
    assert len(t) == VEC[name]["tar_bytes"]
    assert helpers.sha256(t) == VEC[name]["sha256"]
    assert len(t) == oracle.stream_size(SUITE[name])
    assert len(t) % 512 == 0 and t[-1024:] == b"\0" * 1024           # Writer.Close: exactly two zero blocks


def test_c1_quickstart_golden():
    t = gzip.decompress(open(os.path.join(GOLD, "c1_quickstart.tgz"), "rb").read())
    assert len(t) == 98304 == VEC["c1_quickstart"]["tar_bytes"]        # 5 files / 93 116 B -> 98 304 B
    assert helpers.sha256(t) == VEC["c1_quickstart"]["sha256"]
    members = helpers.decode_with_tarfile(t)
    assert [m.name for m, _ in members] == ["/.dockerignore", "/README.md", "/index.js", "/package-lock.json",
                                            "/package.json"]
    assert sum(len(d) for _, d in members) == 93116
    # re-encoding the decoded members with the oracle reproduces the stream
    es = [F(m.name, d, mode=m.mode, uid=m.uid, gid=m.gid, uname=m.uname, gname=m.gname, mtime=m.mtime)
          for m, d in members]
    assert oracle.tar_stream(es) == t


@pytest.mark.parametrize("name", sorted(SUITE))
def test_decodes_with_cpython_tarfile(name):
    es = SUITE[name]
    t = oracle.tar_stream(es)
    got = helpers.decode_with_tarfile(t)
    assert len(got) == len(es)
    for e, (m, data) in zip(es, got):
        want_name = e.name.decode("utf-8", "surrogateescape")
        assert m.name == want_name
        assert m.isdir() == (e.typeflag == ord("5"))
        assert data == (bytes(e.data) if e.typeflag == ord("0") else b"")
        assert m.mode == e.mode
        assert m.uid == e.uid and m.gid == e.gid
        assert m.mtime == e.mtime_sec + (1 if e.mtime_nsec >= 500_000_000 else 0)   # ModTime.Round(time.Second)
        assert m.uname == e.uname.decode("utf-8", "surrogateescape")
        assert m.gname == e.gname.decode("utf-8", "surrogateescape")


@pytest.mark.skipif(shutil.which("tar") is None, reason="GNU tar not installed")
@pytest.mark.parametrize("name", ["util_test_tree", "mixed_tree", "name_155_100", "name_300_pax", "name_nonascii",
                                  "uid_gid_big", "modes", "empty_dir", "name_5000"])
def test_gnu_tar_accepts_and_extracts(name, tmp_path):
    """The remote side runs `tar xzpf` (upstream.go:409): GNU tar must take the stream, gzip'd, without a word."""
    es = SUITE[name]
    arc = tmp_path / "a.tar.gz"
    arc.write_bytes(oracle.pack(es))
    r = subprocess.run(["tar", "tzf", str(arc)], capture_output=True)
    assert r.returncode == 0 and r.stderr == b"", r.stderr
    listed = r.stdout.split(b"\n")[:-1]
    assert len(listed) == len(es)
    if name == "name_5000":
        return                                   # a 5000-byte path component cannot exist on a file system
    dest = tmp_path / "x"
    dest.mkdir()
    r = subprocess.run(["tar", "xzpf", str(arc), "-C", str(dest), "--no-same-owner"], capture_output=True)
    assert r.returncode == 0, r.stderr
    for e in es:
        p = os.path.join(bytes(str(dest), "utf-8"), e.name.lstrip(b"/"))
        if e.typeflag == ord("5"):
            assert os.path.isdir(p)
        else:
            assert open(p, "rb").read() == bytes(e.data)


def _python_ustar_header(e) -> bytes:
    ti = tarfile.TarInfo(e.name.decode())
    ti.size, ti.mtime, ti.mode, ti.uid, ti.gid = e.size, e.mtime_sec, e.mode, e.uid, e.gid
    ti.uname, ti.gname = e.uname.decode(), e.gname.decode()
    ti.type = tarfile.DIRTYPE if e.typeflag == ord("5") else tarfile.REGTYPE
    return ti.tobuf(format=tarfile.USTAR_FORMAT)


@pytest.mark.parametrize("e", [
    F("testFile1", b"TestContents"), F("n" * 100, b"a"), F("dir/" + "n" * 97, b"a"), F("p" * 155 + "/" + "s" * 100),
    F("f", b"x", uid=0o7777777, gid=0o7777777), F("suid", b"x", mode=0o4755), F("zero", b"x", mtime=0),
    F("big", b"", mtime=0o77777777777),
], ids=lambda e: e.name[:12].decode())
def test_ustar_block_matches_cpython_encoder_where_formats_agree(e):
    """CPython's USTAR writer is an independent encoder. Known, documented differences from Go's writer:
    (1) devmajor/devminor of non-device entries: Go formats them as "0000000\\0", CPython leaves NULs;
    (2) hence a different checksum. Everything else in the 512 bytes must be identical."""
    ours = bytearray(oracle.tar_stream([e])[:512])
    theirs = bytearray(_python_ustar_header(e))
    assert bytes(ours[329:345]) == b"0000000\x000000000\x00" and bytes(theirs[329:345]) == b"\0" * 16
    for blk in (ours, theirs):
        blk[329:345] = b"\0" * 16
        blk[148:156] = b" " * 8
    assert ours == theirs
    # and the checksum field is the byte sum of our own block: 6 octal digits, NUL, space
    raw = bytearray(oracle.tar_stream([e])[:512])
    stored = bytes(raw[148:156])
    raw[148:156] = b" " * 8
    assert stored == b"%06o\0 " % sum(raw)


def test_directory_entry_has_no_trailing_slash_and_typeflag_5():
    """tar.go:243 overwrites hdr.Name, so the '/' FileInfoHeader appended to directory names is lost."""
    blk = oracle.tar_stream([D("testFolder2")])[:512]
    assert blk[:12] == b"testFolder2\0" and blk[156:157] == b"5" and blk[124:136] == b"00000000000\0"


def test_pax_extended_header_layout():
    e = F("a/" + "b" * 140 + "/" + "c" * 156, b"abc")
    t = oracle.tar_stream([e])
    assert oracle.header_bytes(e) == 1536
    x = t[:512]
    assert x[156:157] == b"x"
    assert x[:100].rstrip(b"\0") == (b"a/" + b"b" * 140 + b"/PaxHeaders.0/" + b"c" * 156)[:100]
    assert x[100:108] == b"0000000\0" and x[136:148] == b"00000000000\0"          # mode/uid/gid/mtime zero
    assert x[265:345] == b"\0" * 80                                                  # uname/gname/dev* untouched
    rec = t[512:1024].rstrip(b"\0")
    name = e.name
    assert rec == b"%d path=%s\n" % (len(name) + 10 + 0, name) or rec.endswith(b" path=" + name + b"\n")
    n = int(rec.split(b" ", 1)[0])
    assert n == len(rec)                                                             # the length counts itself
    main = t[1024:1536]
    assert main[156:157] == b"0" and main[:100] == name[:100]


def test_pax_record_length_self_reference_boundaries():
    # "%d path=%s\n": lengths that cross 99 -> 100 need the second adjustment step of formatPAXRecord
    for ln in range(80, 130):
        e = F("é" + "k" * ln, b"")
        rec = oracle.tar_stream([e])[512:1024].rstrip(b"\0")
        assert int(rec.split(b" ", 1)[0]) == len(rec)


def test_path_clean_matches_go_semantics():
    cases = {b"": b".", b"a/b": b"a/b", b"a//b": b"a/b", b"a/./b": b"a/b", b"a/b/..": b"a", b"a/b/../../..": b"..",
             b"/..": b"/", b"/a/../..": b"/", b"../a": b"../a", b"a/": b"a", b"/": b"/", b"//a//": b"/a",
             b"a/b//PaxHeaders.0/c": b"a/b/PaxHeaders.0/c"}
    for k, v in cases.items():
        assert oracle.path_clean(k) == v, k


@pytest.mark.parametrize("name", sorted(helpers.ERROR_CASES))
def test_header_errors(name):
    with pytest.raises(oracle.OracleError):
        oracle.tar_stream(helpers.ERROR_CASES[name])


def test_piecewise_oracle_hash_and_streamed_gunzip_agree_with_the_whole_stream():
    """The helpers the full-size config tests use (tests/test_gpu_configs.py): the oracle stream hashed group by group
    and a multi-member gzip inflated piece by piece give the SHA-256 of the whole stream."""
    import gzip
    import hashlib
    from devspace_b200 import synth
    man = synth.tree_c2(400, 1_500_000)
    sha, n = helpers.oracle_tar_sha256(man, group_bytes=200_000, group_entries=33)
    ref = oracle.tar_stream(helpers.manifest_entries(man, 0, man.n))
    assert (sha, n) == (hashlib.sha256(ref).hexdigest(), len(ref))
    two = gzip.compress(ref[:700_001], 1) + gzip.compress(ref[700_001:], 1)
    assert helpers.gunzip_sha256(two, piece=65_000) == (sha, n, 2)


def _hand_derived():
    import json
    return json.load(open(os.path.join(GOLD, "pax_hand_derived.json")))


@pytest.mark.parametrize("case", ["name_101_nosplit", "uid_2097152"])
def test_oracle_matches_hand_derived_pax_bytes(case):
    """tests/golden/make_pax_hand.py builds these streams field by field from the go1.11 archive/tar source without the
    oracle (PaxHeaders.0 entry name, self-inclusive record length, name field cut without terminator, zeroed overflow
    uid, NUL devmajor/devminor in the PAX block): a second derivation the oracle has to agree with."""
    c = _hand_derived()[case]
    m = c["manifest"]
    es = [helpers.F(m["name"], m["data"].encode(), mode=m["mode"], uid=m["uid"], gid=m["gid"], uname=m["uname"],
                    gname=m["gname"], mtime=m["mtime"])]
    assert oracle.tar_stream(es) == bytes.fromhex(c["stream_hex"])
