"""A build-context tree for the docker-dialect tests: directories at several depths, an empty directory, symlinks (file,
directory, dangling, a target longer than 100 bytes), a second hard link, a name that needs the USTAR prefix split,
a name that needs a PAX path record, a directory whose name needs one, a non-ASCII name."""
from __future__ import annotations

import os

import numpy as np

FIXED_MTIME = 1_546_300_800


def materialise(root: str) -> str:
    base = os.path.join(root, "context")
    os.makedirs(base, exist_ok=True)
    rng = np.random.default_rng(11)

    def w(rel, data):
        p = os.path.join(base, rel)
        os.makedirs(os.path.dirname(p), exist_ok=True)
        with open(p, "wb") as f:
            f.write(data)

    w("Dockerfile", b"FROM scratch\nCOPY . /app\n")
    w(".dockerignore", b"node_modules\n!node_modules/keep\n*.log\n")
    w("src/main.go", b"package main\n" * 50)
    w("src/pkg/util/util.go", rng.integers(0, 256, 5000, dtype=np.uint8).tobytes())
    w("src/empty.txt", b"")
    w("build.log", b"noise")
    w("node_modules/lodash/index.js", b"module.exports = {}\n")
    w("node_modules/keep/kept.js", b"keep me\n")
    os.makedirs(os.path.join(base, "emptydir"), exist_ok=True)
    w("d" * 60 + "/" + "e" * 60 + "/file.txt", b"prefix split")            # 126 bytes: USTAR prefix/name split
    w("p" * 120 + ".bin", b"pax path")                                     # 124-byte component: PAX path record
    os.makedirs(os.path.join(base, "q" * 130), exist_ok=True)              # directory name + "/" needs PAX
    w("café/menü.txt", b"utf-8 names")
    for rel, target in (("link_file", "src/main.go"), ("link_dir", "src"), ("dangling", "nowhere/at/all"),
                        ("long_link", "t" * 150)):
        p = os.path.join(base, rel)
        if not os.path.lexists(p):
            os.symlink(target, p)
    hard = os.path.join(base, "zz_hardlink.go")
    if not os.path.lexists(hard):
        os.link(os.path.join(base, "src/main.go"), hard)
    os.chmod(os.path.join(base, "src/main.go"), 0o755)
    os.chmod(os.path.join(base, "emptydir"), 0o1777)
    for dp, dn, fn in os.walk(base):
        for n in dn + fn:
            p = os.path.join(dp, n)
            os.utime(p, (FIXED_MTIME, FIXED_MTIME + 0.75), follow_symlinks=False)
    os.utime(base, (FIXED_MTIME, FIXED_MTIME))
    return base


def dockerignore(rel: str, is_dir: bool) -> int:
    """What docker's matcher answers for the .dockerignore above."""
    if rel == "node_modules":
        return 2                                   # excluded, but "!node_modules/keep" lies below it
    if rel.startswith("node_modules/") and not (rel == "node_modules/keep" or rel.startswith("node_modules/keep/")):
        return 1
    if rel.endswith(".log") and "/" not in rel:
        return 1
    return 0
