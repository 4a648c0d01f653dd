#!/usr/bin/env python
"""profiles/r1_traffic.json from an .ncu-rep: DRAM bytes per launch of every kernel (mean over the captured launches).
  python scripts/ncu_traffic.py report.ncu-rep > profiles/r1_traffic.json"""
import csv, json, re, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: n for n, h in enumerate(hdr)}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
acc = {}
for r in data:
    name = r[col["Kernel Name"]].split("(")[0]
    name = re.sub(r"^void |<unnamed>::|<\d>|_tma", "", name).strip()
    b = sum(float(r[col[k]].replace(",", "")) * scale[units[col[k]]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    acc.setdefault(name, []).append(b)
print(json.dumps({"workload": "C4", "n_gpus": 1,
                  "source": f"ncu --set full, {rep}, dram__bytes_read.sum + dram__bytes_write.sum per launch, mean over the captured launches "
                            "(staged kernels k_uv_tma/k_tq_tma/k_ppw_tma listed under the plain names)",
                  "bytes_per_launch": {k: int(sum(v) / len(v)) for k, v in sorted(acc.items())}}, indent=1))
