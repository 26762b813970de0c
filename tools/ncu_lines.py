#!/usr/bin/env python
"""Per-source-line summary of an ncu report (cuda,sass source page): share of stall samples, share of executed
instructions and average active lanes. Usage: python tools/ncu_lines.py report.ncu-rep [top]"""
import csv
import subprocess
import sys


def main():
    rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    cur, hdr, agg = None, None, {}
    for r in csv.reader(out.splitlines()):
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif r[0].isdigit() and hdr:
            try:
                s = int(r[6]) if r[6] not in ("-", "") else 0
                i = int(r[7]) if r[7] not in ("-", "") else 0
                ti = hdr.index("Thread Instructions Executed")
                t = int(r[ti]) if r[ti] not in ("-", "") else 0
            except (ValueError, IndexError):
                continue
            a = agg.setdefault((cur, int(r[0])), [0, 0, 0, r[1]])
            a[0] += s
            a[1] += i
            a[2] += t
    ts = sum(a[0] for a in agg.values()) or 1
    ti = sum(a[1] for a in agg.values()) or 1
    print(f"total samples {ts}, warp instructions {ti}, lanes/instr {sum(a[2] for a in agg.values()) / ti:.1f}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{k[0][:16]:16s}:{k[1]:4d} {100 * a[0] / ts:5.1f}%s {100 * a[1] / ti:5.1f}%i lanes={a[2] / max(1, a[1]):4.1f}  {a[3][:90]}")


if __name__ == "__main__":
    main()
