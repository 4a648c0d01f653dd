#!/usr/bin/env python
"""Attribute executed instructions and stall samples of one kernel to CUDA source lines:
joins the SASS page of an .ncu-rep with nvdisasm's line info of the object file.
  python scripts/ncu_lines.py report.ncu-rep kernel object.o [top]"""
import csv, os, re, subprocess, sys, tempfile
rep, kern, obj = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines())); hdr = rows[1]; data = rows[2:]
ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
# walk the function of this kernel
lines, cur, infn = [], None, False
for ln in dis.splitlines():
    if ln.startswith(".text."):
        infn = kern in ln
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if infn and re.match(r"\s+/\*[0-9a-f]{4}\*/", ln): lines.append(cur)
if len(lines) != len(data):
    print("instruction count mismatch", len(lines), len(data), "(object rebuilt since the profile?)")
n = min(len(lines), len(data))
agg = {}
for k in range(n):
    a = agg.setdefault(lines[k], [0, 0, 0])
    a[0] += int(data[k][iex]); a[1] += int(data[k][isamp]); a[2] += 1
tex = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
src = {}
print(f"{kern}: {tex} warp-instructions, {ts} samples")
for key, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    f, l = key if key else ("?", 0)
    if f not in src:
        p = os.path.join(os.path.dirname(os.path.abspath(obj)), f)
        src[f] = open(p).read().splitlines() if os.path.exists(p) else []
    text = src[f][l - 1].strip()[:90] if 0 < l <= len(src[f]) else ""
    print(f"{f}:{l:<4d} inst {100.0*v[0]/tex:5.1f}%  samples {100.0*v[1]/ts:5.1f}%  | {text}")
