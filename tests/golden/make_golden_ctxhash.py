#!/usr/bin/env python
"""Writes tests/golden/ctxhash.json: the oracle's build-context hashes (oracle/ctxhash.py, a restatement of
pkg/util/hash/hash.go) of the trees in tests/ctxhash_trees.py, hashed through the relative path "<name>" from
inside a scratch directory. Run from the repo root: python tests/golden/make_golden_ctxhash.py"""
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import ctxhash_trees  # noqa: E402
from oracle import ctxhash  # noqa: E402

out = {}
with tempfile.TemporaryDirectory() as d:
    for name in ctxhash_trees.TREES:
        ctxhash_trees.materialise(d, name)
    cwd = os.getcwd()
    os.chdir(d)
    try:
        for name in ctxhash_trees.TREES:
            recs = ctxhash.records(name)
            out[name] = {"hash": ctxhash.digest(recs), "records": len(recs),
                         "first": [list(r) for r in recs[:4]]}
        out["util_like/excluded"] = {"hash": ctxhash.directory_excludes("util_like", ctxhash_trees.util_like_exclude)}
    finally:
        os.chdir(cwd)
json.dump(out, open(os.path.join(HERE, "ctxhash.json"), "w"), indent=1, sort_keys=True)
print(json.dumps(out, indent=1)[:600])
