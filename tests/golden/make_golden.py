#!/usr/bin/env python
"""Generates the golden fixtures under tests/golden/ (run in the build container, where /root/reference exists).

Upstream holds no golden bytes for this path, so the vectors are produced by the CPU oracle
(oracle/gotar.c, parity UNPINNED against Go itself) and then frozen: they pin the oracle against regressions,
travel to the GPU box (where /root/reference does not exist), and give the GPU tests fixed answers.

  vectors.json        -- per edge-suite case: tar stream length + SHA-256 (+ first header block, hex, for review)
  c1_quickstart.tgz   -- BASELINE config 1: the examples/quickstart upstream-sync set (5 files, 93 116 payload bytes)
                         as the oracle's tar stream (98 304 bytes), stored gzip -9 for size. Metadata is fixed
                         (uid=gid=1000, dev/dev, 0644, mtime 1546300800+i) so the vector is machine independent.
"""
import gzip
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import helpers  # noqa: E402
import oracle  # noqa: E402

QUICKSTART = "/root/reference/examples/quickstart"
# uploadExcludePaths of examples/quickstart/.devspace/config.yaml:25-29 remove Dockerfile, .devspace/, chart/, node_modules/
C1_FILES = [".dockerignore", "README.md", "index.js", "package-lock.json", "package.json"]


def c1_entries():
    es = []
    for i, f in enumerate(sorted(C1_FILES)):
        data = open(os.path.join(QUICKSTART, f), "rb").read()
        es.append(helpers.F("/" + f, data, mtime=helpers.MTIME0 + i))  # initial-sync names carry a leading slash
    return es


def main():
    vec = {}
    for name, es in helpers.edge_suite().items():
        t = oracle.tar_stream(es)
        vec[name] = {"entries": len(es), "tar_bytes": len(t), "sha256": hashlib.sha256(t).hexdigest(),
                     "first_block_hex": t[:512].hex() if len(es) == 1 else None,
                     "header_bytes": [oracle.header_bytes(e) for e in es] if len(es) <= 8 else None}
    t = oracle.tar_stream(c1_entries())
    vec["c1_quickstart"] = {"entries": 5, "tar_bytes": len(t), "sha256": hashlib.sha256(t).hexdigest(),
                            "payload_bytes": sum(e.size for e in c1_entries()),
                            "zlib6_gzip_bytes": len(oracle.gzip_pack(t, 6))}
    with open(os.path.join(HERE, "c1_quickstart.tgz"), "wb") as f:
        f.write(gzip.compress(t, 9, mtime=0))
    with open(os.path.join(HERE, "vectors.json"), "w") as f:
        json.dump(vec, f, indent=1, sort_keys=True)
    print("wrote", len(vec), "vectors; c1:", vec["c1_quickstart"])


if __name__ == "__main__":
    main()
