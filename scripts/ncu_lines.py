#!/usr/bin/env python3
"""Per-source-line instruction and stall-sample totals of one kernel in an .ncu-rep captured
with --import-source on (read here, no GPU):
    python scripts/ncu_lines.py gpurun_out/prof.ncu-rep [top]"""
import csv
import io
import subprocess
import sys

path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(r for r in rows if r and r[0] == "Line No")
i_inst, i_samp = hdr.index("Instructions Executed"), hdr.index("# Samples")
lines = {}
for r in rows:
    if len(r) <= i_inst or not r[0].isdigit() or r[2] != "-":  # per-line aggregate rows have "-" as address
        continue
    lines[int(r[0])] = (r[1], int(r[i_inst] or 0), int(r[i_samp] or 0))
total_i = sum(v[1] for v in lines.values())
total_s = sum(v[2] for v in lines.values())
print("total instructions %d, samples %d" % (total_i, total_s))
for no, (src, inst, samp) in sorted(lines.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%5d %6.2f%% inst %6.2f%% stall | %s" % (no, 100.0 * inst / total_i, 100.0 * samp / max(1, total_s), src.strip()[:110]))
print("-- by stall samples")
for no, (src, inst, samp) in sorted(lines.items(), key=lambda kv: -kv[1][2])[:top]:
    print("%5d %6.2f%% inst %6.2f%% stall | %s" % (no, 100.0 * inst / total_i, 100.0 * samp / max(1, total_s), src.strip()[:110]))
