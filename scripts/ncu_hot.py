#!/usr/bin/env python
"""Summarise the source page of one kernel of an .ncu-rep: per-instruction stall
samples, showing the instructions that carry the samples and every LDG/STG/BRA between."""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.004
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; data = rows[2:]
ia = hdr.index("Source"); isamp = hdr.index("# Samples"); iex = hdr.index("Instructions Executed")
tot = sum(int(r[isamp]) for r in data)
print(rows[0][1], "total samples", tot, "instructions", len(data))
for n, r in enumerate(data):
    s = int(r[isamp]); src = r[ia].strip()
    if s > tot * thr or any(t in src for t in ("LDG", "STG", "BAR", "LDS", "STS", "ATOM", "RED")):
        print(f"{n:5d} {s:7d} {100.0*s/tot:5.1f}% {r[iex]:>9s} {src[:80]}")
