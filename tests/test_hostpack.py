"""The C++ host mirror of writeTar (devspace_b200/csrc/hostpack.cc). The tree, the excludes and the checks are
those of TestCopyToContainerTestable (/root/reference/pkg/devspace/sync/util_test.go:16-104): the walk is compared
with the oracle's restatement of the traversal on CPU; on the GPU the archive goes through the system
`tar xzpf` like the reference's fake remote does (upstream.go:409) and the extracted tree is checked."""
import gzip
import os
import shutil
import subprocess

import pytest

import helpers
import oracle
from devspace_b200 import hostpack

FILE_CONTENTS = b"TestContents"     # util_test.go:147
EXCLUDES = ["ignoredFile", "ignoredFolder", "testFolder/ignoredFile"]   # util_test.go:28-38


def make_local_tree(local):
    for f in ("testFile1", "testFile2", "ignoredFile"):
        (local / f).write_bytes(FILE_CONTENTS)
    for d in ("testFolder", "testFolder2", "ignoredFolder"):
        (local / d).mkdir()
    for f in ("testFolder/testFile1", "testFolder/testFile2", "testFolder/ignoredFile", "ignoredFolder/testFile1"):
        (local / f).write_bytes(FILE_CONTENTS)
    os.chmod(local / "testFile2", 0o755)
    for root, dirs, fs in os.walk(local):          # fixed, past mtimes: a header mtime rounded UP (WriteHeader) must
        for n in dirs + fs:                        # not lie in the future when the remote tar extracts it
            os.utime(os.path.join(root, n), ns=(1546300900_000000000, 1546300900_000000000))
    os.utime(local / "testFile1", ns=(1546300800_600000000, 1546300800_600000000))   # .6 s: rounds up


def excluded(rel: str) -> bool:
    """Stand-in for the gitignore matcher: a path is excluded when it or one of its parents is listed."""
    return any(rel == e or rel.startswith(e + "/") for e in EXCLUDES)


def test_walk_matches_the_oracle_traversal(tmp_path):
    make_local_tree(tmp_path)
    files = [hostpack.FileInformation(Name="", IsDirectory=True)]            # CopyToContainer: util.go:68-73
    cfg = hostpack.SyncConfig(WatchPath=str(tmp_path), ignoreMatcher=excluded)
    written, man = hostpack.walk(files, cfg)
    want_entries, want_written = oracle.walk([oracle.FileInformation(Name="", IsDirectory=True)], str(tmp_path),
                                             is_excluded=excluded)
    assert [man.name(i) for i in range(man.n)] == [e.name for e in want_entries]
    assert [man.name(i).decode() for i in range(man.n)] == ["testFile1", "testFile2", "testFolder/testFile1",
                                                            "testFolder/testFile2", "testFolder2"]
    assert sorted(written) == sorted(want_written)
    for i, e in enumerate(want_entries):
        g = man.entries[i]
        assert (int(g["size"]), int(g["mode"]), int(g["uid"]), int(g["gid"]), int(g["typeflag"])) == \
               (e.size, e.mode, e.uid, e.gid, e.typeflag)
        assert (int(g["mtime_sec"]), int(g["mtime_nsec"])) == (e.mtime_sec, e.mtime_nsec)
        un = man.strtab[int(g["uname_off"]): int(g["uname_off"]) + int(g["uname_len"])].tobytes()
        assert un == e.uname
    assert written["testFile1"].Mtime == 1546300801                           # roundMtime
    assert written["testFolder2"].IsDirectory and not written["testFile2"].IsDirectory


def test_walk_dedup_remote_override_and_vanished(tmp_path):
    make_local_tree(tmp_path)
    files = [hostpack.FileInformation("testFolder"), hostpack.FileInformation("testFolder/testFile1"),
             hostpack.FileInformation("doesNotExist"), hostpack.FileInformation("testFile2")]
    fm = {"testFile2": hostpack.FileInformation("testFile2", RemoteMode=0o600, RemoteUID=7, RemoteGID=8)}
    written, man = hostpack.walk(files, hostpack.SyncConfig(WatchPath=str(tmp_path), fileMap=fm,
                                                            uploadIgnoreMatcher=lambda p: p.endswith("ignoredFile")))
    names = [man.name(i).decode() for i in range(man.n)]
    assert names == ["testFolder/testFile1", "testFolder/testFile2", "testFile2"]      # change-list order, de-duplicated
    e = man.entries[2]
    assert (int(e["mode"]), int(e["uid"]), int(e["gid"])) == (0o600, 7, 8)             # tar.go:287-291
    assert written["testFile2"].RemoteMode == 0o600


@pytest.mark.gpu
@pytest.mark.skipif(shutil.which("tar") is None, reason="tar not installed")
def test_copy_to_container_roundtrip(tmp_path):
    from devspace_b200.pack import Packer
    local, remote = tmp_path / "local", tmp_path / "remote"
    local.mkdir()
    remote.mkdir()
    make_local_tree(local)
    with Packer(level=6) as p:
        path, written = hostpack.write_tar(p, [hostpack.FileInformation(Name="", IsDirectory=True)],
                                           hostpack.SyncConfig(WatchPath=str(local), ignoreMatcher=excluded))
    try:
        r = subprocess.run(["tar", "xzpf", path, "-C", str(remote), "--no-same-owner"], capture_output=True)
        assert r.returncode == 0 and r.stderr == b""
        blob = open(path, "rb").read()
    finally:
        os.remove(path)                                                    # upstream.go:441
    # checkFilesAndFolders (util_test.go:149-297): existence + exact content on the remote side
    for f in ("testFile1", "testFile2", "testFolder/testFile1", "testFolder/testFile2"):
        assert (remote / f).read_bytes() == FILE_CONTENTS
    for f in ("ignoredFile", "testFolder/ignoredFile", "ignoredFolder/testFile1", "ignoredFolder"):
        assert not (remote / f).exists()
    assert (remote / "testFolder").is_dir() and (remote / "testFolder2").is_dir()
    assert os.stat(remote / "testFile2").st_mode & 0o777 == 0o755
    assert int(os.stat(remote / "testFile1").st_mtime) == 1546300801
    assert sorted(written) == ["testFile1", "testFile2", "testFolder/testFile1", "testFolder/testFile2", "testFolder2"]
    want, _ = oracle.walk([oracle.FileInformation(Name="", IsDirectory=True)], str(local), is_excluded=excluded)
    assert gzip.decompress(blob) == oracle.tar_stream(want)                # and the stream is the oracle's, bit for bit


@pytest.mark.gpu
def test_write_tar_retries_like_the_reference(tmp_path):
    """tar.go:170-177: a failing pack is logged, the temp file removed, and writeTar runs again after the sleep."""
    from devspace_b200._lib import DsgError
    from devspace_b200.pack import Packer
    (tmp_path / "secret").write_bytes(b"x")
    os.chmod(tmp_path / "secret", 0)
    if os.access(tmp_path / "secret", os.R_OK):
        (tmp_path / "secret").unlink()
        os.mkfifo(tmp_path / "secret")       # root reads anything: use a file type archive/tar cannot take via io.Copy
    logs = []
    with Packer(level=6) as p:
        with pytest.raises(DsgError):
            hostpack.write_tar(p, [hostpack.FileInformation("secret")],
                               hostpack.SyncConfig(WatchPath=str(tmp_path), Logf=logs.append, retry_sleep_ms=10,
                                                   max_retries=2))
    assert len(logs) == 2 and all(l.startswith("[Upstream] Tar failed: ") and "Will retry" in l for l in logs)
