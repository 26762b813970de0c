"""Docker build-context tar through the C ABI (dsg_docker_context_tar -> dsg_pack_tar_only on the GPU) against the oracle
(oracle/dockerctx.py + oracle/gotar.c), bit exact, and accepted by the stock decoders."""
import os
import shutil
import subprocess
import tarfile

import pytest

import docker_trees
import oracle
from oracle import dockerctx

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def packer():
    from devspace_b200.pack import Packer
    with Packer(level=0) as p:
        yield p


@pytest.fixture(scope="module")
def ctx(tmp_path_factory):
    return docker_trees.materialise(str(tmp_path_factory.mktemp("dk")))


@pytest.mark.parametrize("excluded", [False, True])
def test_context_tar_matches_oracle(packer, ctx, tmp_path, excluded):
    from devspace_b200.dockerctx import context_tar
    ex = docker_trees.dockerignore if excluded else None
    out = str(tmp_path / "ctx.tar")
    nbytes, nent = context_tar(packer, ctx, out, ex)
    es = dockerctx.entries(ctx, ex)
    ref = oracle.tar_stream(es)
    got = open(out, "rb").read()
    assert nent == len(es) and nbytes == len(ref) == len(got)
    assert got == ref                                              # bit exact against the oracle
    names = [m.name for m in tarfile.open(out).getmembers()]
    assert len(names) == len(es) and ("build.log" in names) == (not excluded)
    if shutil.which("tar"):
        r = subprocess.run(["tar", "tf", out], capture_output=True)
        assert r.returncode == 0 and r.stderr == b""


def test_context_tar_extracts_to_the_same_tree(packer, ctx, tmp_path):
    from devspace_b200.dockerctx import context_tar
    out = str(tmp_path / "ctx.tar")
    context_tar(packer, ctx, out)
    dst = tmp_path / "x"
    dst.mkdir()
    with tarfile.open(out) as tf:
        tf.extractall(dst, filter="fully_trusted")
    assert os.readlink(dst / "link_file") == "src/main.go" and os.readlink(dst / "long_link") == "t" * 150
    assert (dst / "src/main.go").read_bytes() == open(os.path.join(ctx, "src/main.go"), "rb").read()
    assert os.stat(dst / "zz_hardlink.go").st_ino == os.stat(dst / "src/main.go").st_ino    # hard link restored
    assert (dst / "emptydir").is_dir() and (dst / ("q" * 130)).is_dir()
    assert (dst / ("p" * 120 + ".bin")).read_bytes() == b"pax path"


def test_context_tar_errors(packer, tmp_path):
    from devspace_b200._lib import DsgError
    from devspace_b200.dockerctx import context_tar
    (tmp_path / "f").write_bytes(b"x")
    with pytest.raises(DsgError) as ei:
        context_tar(packer, str(tmp_path / "f"), str(tmp_path / "o.tar"))
    assert "not a directory" in str(ei.value)
    (tmp_path / "empty").mkdir()
    n, k = context_tar(packer, str(tmp_path / "empty"), str(tmp_path / "o.tar"))
    assert (n, k) == (1024, 0) and (tmp_path / "o.tar").read_bytes() == b"\0" * 1024        # just the trailer
