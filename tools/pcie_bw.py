#!/usr/bin/env python
"""PCIe roofline denominator: pinned cudaMemcpyAsync host->device and device->host
bandwidth of every GPU of the job AT THE SAME TIME. One process per GPU (torchrun for N > 1); all ranks copy between
barriers, every rank times its own copies with CUDA events, rank 0 prints one JSON line:
  python tools/pcie_bw.py                                   # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/pcie_bw.py
The line is kept in profiles/r02_pcie.jsonl; bench.py reads the per-N figure from there."""
import json
import os
import sys

import torch
import torch.distributed as dist


def main():
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nbytes = int(os.environ.get("PCIE_BYTES", 1 << 30))
    reps = 5
    h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h.fill_(7)
    d = torch.empty(nbytes, dtype=torch.uint8, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn):
        fn()
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        b.synchronize()
        barrier()
        return nbytes * reps / 1e9 / (a.elapsed_time(b) / 1e3)

    h2d = timed(lambda: d.copy_(h, non_blocking=True))
    d2h = timed(lambda: h.copy_(d, non_blocking=True))
    s2 = torch.cuda.Stream()
    h_b = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_b = torch.empty(nbytes, dtype=torch.uint8, device=dev)

    def both():
        d.copy_(h, non_blocking=True)
        with torch.cuda.stream(s2):
            h_b.copy_(d_b, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s2)

    bidir = timed(both)          # GB/s per direction while both directions run
    vals = torch.tensor([h2d, d2h, bidir], dtype=torch.float64, device=dev)
    if world > 1:
        lo = vals.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        tot = vals.clone()
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    else:
        lo, tot = vals, vals
    if rank == 0:
        print(json.dumps({"what": "pinned cudaMemcpyAsync bandwidth, all GPUs of the job copying at the same time",
                          "n_gpus": world, "bytes_per_copy": nbytes, "reps": reps,
                          "h2d_GBps_per_gpu_min": float(lo[0]), "d2h_GBps_per_gpu_min": float(lo[1]),
                          "bidir_GBps_per_direction_min": float(lo[2]),
                          "h2d_GBps_aggregate": float(tot[0]), "d2h_GBps_aggregate": float(tot[1]),
                          "gpu": torch.cuda.get_device_name(local), "host_cpus": os.cpu_count()}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
