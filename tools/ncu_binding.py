#!/usr/bin/env python
"""What binds lz_deflate_kernel (shared-memory bandwidth / instruction issue, with the HBM fraction
alongside): reads the raw page of an ncu report and writes the figures bench.py puts into `roofline.binding`.
Usage: python tools/ncu_binding.py report.ncu-rep profiles/r02_lz_binding.json "<what was captured>" """
import csv
import json
import subprocess
import sys


def main():
    rep, out, what = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[0]
    vals = rows[-1]          # the last (only) captured launch
    m = {}
    for k, v in zip(hdr, vals):
        try:
            m[k] = float(v.replace(",", ""))
        except ValueError:
            m[k] = v
    stall = {k.replace("smsp__pcsamp_warps_issue_stalled_", ""): v for k, v in m.items()
             if k.startswith("smsp__pcsamp_warps_issue_stalled_") and not k.endswith("_not_issued") and isinstance(v, float)}
    tot = sum(stall.values()) or 1.0
    top = sorted(stall.items(), key=lambda kv: -kv[1])[:6]
    wf = m.get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", 0.0)
    bc = m.get("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", 0.0)
    res = {
        "kernel": m.get("Kernel Name", "lz_deflate_kernel"),
        "captured": what,
        "duration_ms": m.get("gpu__time_duration.sum", 0.0) / (1e6 if m.get("gpu__time_duration.sum", 0) > 1e5 else 1.0),
        "smem_wavefronts_pct_of_peak": m.get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
        "smem_bank_conflict_share": (bc / wf) if wf else None,
        "issue_active_pct": m.get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "active_lanes_per_instruction": m.get("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "eligible_warps_per_cycle": m.get("smsp__warps_eligible.avg.per_cycle_active"),
        "warps_active_pct": m.get("sm__warps_active.avg.pct_of_peak_sustained_active"),
        "sm_throughput_pct": m.get("sm__throughput.avg.pct_of_peak_sustained_elapsed"),
        "dram_throughput_pct": m.get("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        "dram_bytes_read": m.get("dram__bytes_read.sum"), "dram_bytes_write": m.get("dram__bytes_write.sum"),
        "barrier_share_of_stall_samples": stall.get("barrier", 0.0) / tot,
        "top_stalls": {k: round(v / tot, 4) for k, v in top},
        "registers_per_thread": m.get("launch__registers_per_thread"),
        "shared_mem_per_block_dynamic": m.get("launch__shared_mem_per_block_dynamic"),
    }
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
