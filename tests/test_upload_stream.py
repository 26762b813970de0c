"""SURVEY.md 8f rank 3, second half: the archive is uploaded WHILE it is being packed. The size contract of the
reference's remote script (upstream.go:387-407 waits for an exact byte count) becomes a frame contract
(devspace_b200/upload.py). The remote side is the local ``sh`` the reference itself uses when testing
(upstream.go:67-95). CPU: the protocol with a zlib-made archive; GPU: the same with dsg_pack_fd as the producer."""
import gzip
import io
import os
import shutil
import subprocess
import tarfile

import pytest

import helpers
from devspace_b200 import upload

needs_sh = pytest.mark.skipif(not (shutil.which("sh") and shutil.which("tar")), reason="needs sh and tar")


def _tree(tmp_path):
    src = tmp_path / "src"
    (src / "sub" / "deep").mkdir(parents=True)
    (src / "emptydir").mkdir()
    files = {"a.txt": b"TestContents", "sub/b.bin": helpers.rng_bytes(1, 300_000),
             "sub/deep/c.js": helpers.text_bytes(2, 2_500_000), "zero": b""}
    for rel, data in files.items():
        (src / rel).write_bytes(data)
    return src, files


def _check_extracted(dest, files):
    for rel, data in files.items():
        assert (dest / rel).read_bytes() == data, rel
    assert (dest / "emptydir").is_dir()


def _sh():
    return subprocess.Popen(["sh"], stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.PIPE)


@needs_sh
def test_frames_reach_the_remote_tar_cpu(tmp_path):
    src, files = _tree(tmp_path)
    buf = io.BytesIO()
    with tarfile.open(fileobj=buf, mode="w:gz") as tf:
        tf.add(src, arcname=".")
    blob = buf.getvalue()
    dest = tmp_path / "dest"
    sh = _sh()

    def produce(fd):
        for o in range(0, len(blob), 70_001):      # the producer writes in pieces of its own choosing
            os.write(fd, blob[o:o + 70_001])

    sent = upload.upload_archive_streaming(sh.stdin, sh.stdout, str(dest), produce)
    sh.stdin.close()
    sh.wait(timeout=30)
    assert sent == len(blob)
    _check_extracted(dest, files)


@needs_sh
def test_producer_failure_and_bad_archive_are_errors(tmp_path):
    sh = _sh()

    def broken(fd):
        os.write(fd, b"not a gzip stream")

    with pytest.raises(IOError):
        upload.upload_archive_streaming(sh.stdin, sh.stdout, str(tmp_path / "d1"), broken)
    sh.stdin.close()
    sh.wait(timeout=30)
    sh = _sh()

    def raising(fd):
        raise RuntimeError("pack failed")

    with pytest.raises(RuntimeError):
        upload.upload_archive_streaming(sh.stdin, sh.stdout, str(tmp_path / "d2"), raising)
    sh.kill()


@needs_sh
@pytest.mark.gpu
def test_pack_fd_streams_into_the_upload(tmp_path):
    """writeTar + uploadArchive in one motion: the host mirror walks the tree, dsg_pack_fd writes the archive into the
    pipe segment by segment, the frames leave while later segments are still on the GPU, the remote tar extracts."""
    import numpy as np
    from devspace_b200 import hostpack
    from devspace_b200.manifest import Manifest
    from devspace_b200.pack import Packer
    src, files = _tree(tmp_path)
    dest = tmp_path / "dest"
    written, hdrs = hostpack.walk([hostpack.FileInformation(Name="")], hostpack.SyncConfig(WatchPath=str(src)))
    arena = np.zeros(int((hdrs.entries["data_off"] + hdrs.entries["size"].astype(np.uint64)).max()) + 16, np.uint8)
    for i, e in enumerate(hdrs.entries):                       # tarFile's io.Copy: the file bytes into the arena
        if e["typeflag"] == ord("0") and e["size"] > 0:
            data = (src / hdrs.name(i).decode().lstrip("/")).read_bytes()
            arena[int(e["data_off"]): int(e["data_off"]) + len(data)] = np.frombuffer(data, np.uint8)
    man = Manifest(hdrs.entries, hdrs.strtab, arena)
    with Packer(level=6, segment_bytes=1 << 20) as p:          # several segments: the stream arrives in pieces
        sh = _sh()
        sent = upload.upload_archive_streaming(sh.stdin, sh.stdout, str(dest), lambda fd: p.pack_to_fd(man, fd))
        sh.stdin.close()
        sh.wait(timeout=60)
        assert p.stats()["segments"] >= 3 and sent == p.stats()["out_bytes"]
    _check_extracted(dest, files)
    assert {k.lstrip("/") for k, v in written.items() if not v.IsDirectory} == set(files)
