"""Small directory trees for the build-context hash tests (shared by the oracle, emulator and GPU tests and by
tests/golden/make_golden_ctxhash.py). Content is seeded, names are fixed: the hashes are machine independent as
long as the tree is hashed through the same relative path."""
from __future__ import annotations

import os

import numpy as np


def _bytes(n: int, seed: int) -> bytes:
    return np.random.default_rng(seed).integers(0, 256, n, dtype=np.uint8).tobytes()


TREES = {
    # the shape of the reference's sync test tree (sync/util_test.go:25-40) with some content sizes that matter to
    # the checksum kernel: empty, < 16 bytes, exactly one unit, just over one unit
    "util_like": {
        "testFile1": b"TestContents",
        "testFile2": b"",
        "testFolder/testFile1": b"123456789",
        "testFolder/testFile2": _bytes(65536, 1),
        "testFolder/sub/deep.bin": _bytes(65536 + 17, 2),
        "testFolder2/": None,
        "ignoreFolder/ignoredFile": b"x" * 100,
        "ignoreFolder/kept/exception.txt": b"kept",
        ".dockerignore": b"ignoreFolder\n!ignoreFolder/kept\n",
    },
    "many_small": {("d%02d/f%04d.js" % (i % 7, i)): _bytes(1 + (i * 37) % 300, 100 + i) for i in range(1500)},
    "one_big": {"blob.bin": _bytes(3 * 1024 * 1024 + 5, 7), "z/": None},
}


def materialise(root: str, name: str) -> str:
    """Writes TREES[name] under root/<name> and returns that path."""
    base = os.path.join(root, name)
    os.makedirs(base, exist_ok=True)
    for rel, data in TREES[name].items():
        p = os.path.join(base, rel)
        if data is None:
            os.makedirs(p, exist_ok=True)
        else:
            os.makedirs(os.path.dirname(p), exist_ok=True)
            with open(p, "wb") as f:
                f.write(data)
    if name == "util_like":
        link = os.path.join(base, "link_to_file1")
        if not os.path.lexists(link):
            os.symlink("testFile1", link)
        dlink = os.path.join(base, "link_to_dir")
        if not os.path.lexists(dlink):
            os.symlink("testFolder2", dlink)
    return base


def util_like_exclude(rel: str, is_dir: bool) -> int:
    """What docker's matcher answers for the .dockerignore of util_like: ignoreFolder is excluded, with an exception
    below it (so the directory is entered), everything in it but kept/ is excluded."""
    if rel == "./ignoreFolder":
        return 2
    if rel.startswith("./ignoreFolder/") and not rel.startswith("./ignoreFolder/kept"):
        return 1
    return 0
