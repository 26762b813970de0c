"""CPU, world_size 2 over gloo: the N>1 host logic -- shard bounds, per-rank packing with the trailer only on the
last shard, member concatenation on rank 0, and the max/sum reductions bench.py uses for its timing. The per-rank
"GPU" work is done by the product kernels under the SIMT emulator (no GPU in this container); the checker is the
oracle. There is no data-path collective to test: ranks never exchange payload."""
import gzip
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import helpers as h
    import test_emu_deflate as emu
    from devspace_b200 import shard
    from devspace_b200.manifest import from_entries

    es = h.edge_suite()["mixed_tree"]
    man = from_entries(es)
    sub, flags = shard.rank_slice(man, rank, world)
    rc, tar, _ = h.emu_tar_build(sub, flags=flags)
    assert rc == 0
    # one gzip member per rank: header | deflate (emulated LZ kernel) | crc32 | isize
    import struct
    import zlib
    member = bytes.fromhex("1f8b08000000000000ff") + emu.emu_deflate(tar) + struct.pack("<II", zlib.crc32(tar), len(tar))
    gathered = [None] * world
    dist.all_gather_object(gathered, member)
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)          # bench.py: time = max over ranks
    s = torch.tensor([float(sub.payload_bytes)], dtype=torch.float64)
    dist.all_reduce(s, op=dist.ReduceOp.SUM)          # bench.py: value = units of all ranks / that time
    if rank == 0:
        q.put((shard.concat_members(gathered), float(t.item()), float(s.item()), man.payload_bytes))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_shard_pack_and_concatenate():
    import oracle
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    blob, tmax, psum, ptotal = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert tmax == float(world) and psum == float(ptotal)
    es = helpers.edge_suite()["mixed_tree"]
    assert gzip.decompress(blob) == oracle.tar_stream(es)       # two members, one archive, bit-exact tar stream
    assert blob.count(bytes.fromhex("1f8b08000000000000ff")) >= 2


def test_shard_bounds_balance_large_files():
    from devspace_b200 import shard, synth
    man = synth.build_manifest(synth.tree_c4_names(256), np.full(256, 256 << 20, np.int64), np.zeros(16, np.uint8),
                               np.arange(256, dtype=np.uint64) * np.uint64(256 << 20))
    for g in (1, 2, 4, 8):
        b = shard.shard_bounds(man, g)
        assert list(np.diff(b.astype(np.int64))) == [256 // g] * g      # config 4: 32 files per GPU at N=8
