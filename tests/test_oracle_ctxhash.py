"""The build-context hash oracle (oracle/ctxhash.py, restating pkg/util/hash/hash.go) against standard check values,
its own golden fixture, and the walk semantics of DirectoryExcludes. CPU only."""
import hashlib
import json
import os
import zlib

import pytest

import ctxhash_trees
from oracle import ctxhash

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "ctxhash.json")))


@pytest.fixture(scope="module")
def trees(tmp_path_factory):
    d = tmp_path_factory.mktemp("ctx")
    for name in ctxhash_trees.TREES:
        ctxhash_trees.materialise(str(d), name)
    return str(d)


def test_crc32_check_value(tmp_path):
    # the CRC-32/ISO-HDLC check value: crc("123456789") = cbf43926; Go prints hash.Sum(nil) big endian
    p = tmp_path / "f"
    p.write_bytes(b"123456789")
    assert ctxhash.hash_file_crc32(str(p)) == "cbf43926"
    (tmp_path / "e").write_bytes(b"")
    assert ctxhash.hash_file_crc32(str(tmp_path / "e")) == "00000000"
    assert ctxhash.hash_file_crc32(str(tmp_path / "missing")) is None
    assert ctxhash.hash_file_crc32(str(tmp_path)) is None          # a directory: Go's read fails, entry left out


def test_digest_is_sha256_of_the_concatenation():
    recs = [("a", True, None), ("a/b", False, "cbf43926")]
    assert ctxhash.digest(recs) == hashlib.sha256(b"a" + b"a/b;cbf43926").hexdigest()
    assert ctxhash.digest([]) == "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855"


@pytest.mark.parametrize("name", sorted(ctxhash_trees.TREES))
def test_golden(trees, name, monkeypatch):
    monkeypatch.chdir(trees)
    recs = ctxhash.records(name)
    assert len(recs) == GOLDEN[name]["records"]
    assert [list(r) for r in recs[:4]] == GOLDEN[name]["first"]
    assert ctxhash.digest(recs) == GOLDEN[name]["hash"]


def test_walk_order_and_links(trees, monkeypatch):
    monkeypatch.chdir(trees)
    recs = ctxhash.records("util_like")
    paths = [r[0] for r in recs]
    assert paths[0] == "util_like" and paths == sorted(paths, key=lambda p: p.split("/"))  # pre-order, names sorted
    by = {r[0]: r for r in recs}
    # a link to a file is hashed through the link (os.Open follows it); a link to a directory cannot be read
    assert by["util_like/link_to_file1"][2] == "%08x" % zlib.crc32(b"TestContents")
    assert "util_like/link_to_dir" not in by
    assert by["util_like/testFolder2"][1] is True                    # empty directory: its path is hashed
    assert by["util_like/testFile2"][2] == "00000000"


def test_excludes(trees, monkeypatch):
    monkeypatch.chdir(trees)
    recs = ctxhash.records("util_like", ctxhash_trees.util_like_exclude)
    paths = [r[0] for r in recs]
    assert "util_like/ignoreFolder" not in paths                      # excluded itself ...
    assert "util_like/ignoreFolder/kept/exception.txt" in paths       # ... but entered for the exception
    assert "util_like/ignoreFolder/ignoredFile" not in paths
    assert ctxhash.digest(recs) == GOLDEN["util_like/excluded"]["hash"]
    skipped = ctxhash.records("util_like", lambda rel, d: 1 if rel == "./testFolder" else 0)
    assert not any(p.startswith("util_like/testFolder/") or p == "util_like/testFolder" for p, _, _ in skipped)


def test_not_a_directory(tmp_path):
    (tmp_path / "f").write_bytes(b"x")
    with pytest.raises(NotADirectoryError):
        ctxhash.directory_excludes(str(tmp_path / "f"))


def test_directory_variant(trees, monkeypatch):
    monkeypatch.chdir(trees)
    h = hashlib.sha256()
    for p, st in ctxhash.walk("one_big"):
        h.update(("%s;%d;%d" % (p, st.st_size, st.st_mtime_ns)).encode())
    assert ctxhash.directory("one_big") == h.hexdigest()
    assert [p for p, _ in ctxhash.walk("one_big")] == ["one_big", "one_big/blob.bin", "one_big/z"]
