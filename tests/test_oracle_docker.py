"""The docker build-context dialect (oracle/dockerctx.py + oracle/gotar.c link entries): decoded by CPython tarfile
and GNU tar (independent implementations of the format), and the tar kernels under the emulator against the oracle.
CPU only."""
import io
import os
import shutil
import subprocess
import tarfile

import pytest

import docker_trees
import oracle
from oracle import dockerctx
from devspace_b200.manifest import from_entries
from helpers import emu_tar_build


@pytest.fixture(scope="module")
def ctx(tmp_path_factory):
    return docker_trees.materialise(str(tmp_path_factory.mktemp("dk")))


def test_walk_headers(ctx):
    es = dockerctx.entries(ctx)
    by = {e.name: e for e in es}
    names = [e.name for e in es]
    assert names == sorted(names, key=lambda n: n.rstrip(b"/").split(b"/"))           # filepath.Walk order
    assert b"src/" in by and b"src/pkg/" in by and b"emptydir/" in by                 # every directory, with "/"
    assert b"./" not in by and b"" not in by                                          # not the context dir itself
    assert by[b"src/"].typeflag == ord("5") and by[b"src/"].mode & 0o170000 == 0o040000
    assert by[b"emptydir/"].mode == 0o041777                                          # sticky + type bits
    f = by[b"src/main.go"]
    assert f.typeflag == ord("0") and f.mode == 0o100755 and f.uid == f.gid == 0
    assert f.mtime_sec == docker_trees.FIXED_MTIME and f.mtime_nsec == 0              # truncated, not rounded
    assert by[b"link_file"].typeflag == ord("2") and by[b"link_file"].linkname == b"src/main.go"
    assert by[b"link_file"].mode & 0o170000 == 0o120000 and by[b"link_file"].size == 0
    assert by[b"dangling"].linkname == b"nowhere/at/all"
    h = by[b"zz_hardlink.go"]
    assert h.typeflag == ord("1") and h.linkname == b"src/main.go" and h.size == 0    # second name of the inode
    assert by[b"src/empty.txt"].size == 0


def test_excludes(ctx):
    names = [e.name for e in dockerctx.entries(ctx, docker_trees.dockerignore)]
    assert b"build.log" not in names and b"node_modules/" not in names
    assert b"node_modules/lodash/" not in names and b"node_modules/lodash/index.js" not in names
    assert b"node_modules/keep/" in names and b"node_modules/keep/kept.js" in names


def test_stream_decodes_with_tarfile_and_gnu_tar(ctx, tmp_path):
    es = dockerctx.entries(ctx)
    stream = oracle.tar_stream(es)
    tf = tarfile.open(fileobj=io.BytesIO(stream))
    members = {m.name: m for m in tf.getmembers()}
    assert len(members) == len(es)
    assert members["link_file"].issym() and members["link_file"].linkname == "src/main.go"
    assert members["long_link"].issym() and members["long_link"].linkname == "t" * 150       # PAX linkpath record
    assert members["zz_hardlink.go"].islnk() and members["zz_hardlink.go"].linkname == "src/main.go"
    assert members["src"].isdir() and members["q" * 130].isdir()                             # PAX path of a directory
    assert members["p" * 120 + ".bin"].isreg()
    assert tf.extractfile("src/pkg/util/util.go").read() == open(os.path.join(ctx, "src/pkg/util/util.go"), "rb").read()
    assert members["café/menü.txt"].size == len(b"utf-8 names")
    if shutil.which("tar"):
        p = tmp_path / "ctx.tar"
        p.write_bytes(stream)
        r = subprocess.run(["tar", "tvf", str(p)], capture_output=True)
        assert r.returncode == 0 and r.stderr == b"", r.stderr
        listing = r.stdout.decode("utf-8", "replace")
        assert "link_file -> src/main.go" in listing and "zz_hardlink.go link to src/main.go" in listing
        assert len(listing.strip().split("\n")) == len(es)


def _rec(key: bytes, value: bytes) -> bytes:
    """formatPAXRecord: "%d %s=%s\\n" where the length counts itself."""
    body = b" " + key + b"=" + value + b"\n"
    n = len(body) + len(str(len(body)))
    if len(str(n)) + len(body) != n:
        n = len(str(n)) + len(body)
    return str(n).encode() + body


def test_header_level_details(ctx):
    es = {e.name: e for e in dockerctx.entries(ctx)}
    long_dir = b"q" * 130 + b"/"
    blocks = oracle.tar_stream([es[long_dir]])
    # path.Split("qq..q/") = ("qq..q/", ""): Join(dir, "PaxHeaders.0", "") = "qq..q/PaxHeaders.0", cut to 100 bytes
    assert blocks[156:157] == b"x" and blocks[:100] == b"q" * 100
    assert _rec(b"path", long_dir) in blocks[512:1024]
    assert blocks[1024 + 156:1024 + 157] == b"5" and blocks[1024:1024 + 100] == long_dir[:100]
    sym = oracle.tar_stream([es[b"link_file"]])
    assert sym[156:157] == b"2" and sym[157:257].rstrip(b"\0") == b"src/main.go" and sym[124:135] == b"0" * 11
    assert sym[100:107] == b"0120777"
    ll = oracle.tar_stream([es[b"long_link"]])
    assert ll[156:157] == b"x" and _rec(b"linkpath", b"t" * 150) in ll[512:1024]
    assert ll[1024 + 157:1024 + 257] == b"t" * 100                                               # truncated in the block
    split = oracle.tar_stream([es[b"d" * 60 + b"/" + b"e" * 60 + b"/file.txt"]])
    assert split[156:157] == b"0" and split[345:345 + 60] == b"d" * 60                            # USTAR prefix, no PAX


@pytest.mark.parametrize("excluded", [False, True])
def test_emulated_kernels_match_oracle(ctx, excluded):
    es = dockerctx.entries(ctx, docker_trees.dockerignore if excluded else None)
    rc, got, err = emu_tar_build(from_entries(es))
    assert rc == 0, (rc, err)
    assert got == oracle.tar_stream(es)
    rc, seg, _ = emu_tar_build(from_entries(es), segment_bytes=32768)                           # clipped PAX areas
    assert rc == 0 and seg == got


def test_link_entry_errors():
    from oracle import Entry, OracleError
    bad = Entry(name=b"l", typeflag=ord("2"), linkname=b"a\0b" + b"c" * 120)                    # NUL in a PAX value
    with pytest.raises(OracleError):
        oracle.tar_stream([bad])
    rc, _, err = emu_tar_build(from_entries([bad]))
    assert rc == -3 and err == 0
    sized = Entry(name=b"l", typeflag=ord("1"), linkname=b"x", size=5, data=b"12345")
    with pytest.raises(OracleError):
        oracle.tar_stream([sized])
    assert emu_tar_build(from_entries([sized]))[0] == -3
