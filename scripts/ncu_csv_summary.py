#!/usr/bin/env python
"""Per-kernel summary of an `ncu -i report --page raw --csv` export (the report itself is too large to travel from
the GPU box): markdown table on stdout, DRAM traffic and FP64 instruction counts per launch as JSON.
  python scripts/ncu_csv_summary.py gpurun_out/prof_r2_raw.csv C4 profiles/r2_traffic.json profiles/r2_fp64.json > profiles/r2_ncu_table.md"""
import csv, json, re, sys
src, workload, jt, jf = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4]
rows = [r for r in csv.reader(open(src)) if len(r) > 10]
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: n for n, h in enumerate(hdr)}
def g(r, k, d=float("nan")):
    try: return float(r[col[k]].replace(",", ""))
    except Exception: return d
def unit(k): return units[col[k]] if k in col else ""
def to_bytes(v, u): return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1)
def to_ms(v, u): return v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "usecond": 1e-3, "msecond": 1.0, "nsecond": 1e-6, "second": 1e3}.get(u, 1)
def short(n):
    n = n.split("(")[0]
    return re.sub(r"^void |<unnamed>::", "", n).strip()
stall_keys = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
seen = {}
for r in data: seen.setdefault(short(r[col["Kernel Name"]]), []).append(r)
FP64_RATE = 18.57e12 / 32.0   # warp-instructions/s at the measured issue rate (profiles/r2_measurements.md, scripts/fp64_peak.cu)
print("| kernel | launches | ms | DRAM GB | DRAM GB/s | L2 hit % | regs | warps/SM | IPC/SM | fp64 pipe % | FP64 warp-inst (M) | FP64 floor ms | inst/warp | top stalls (warps per issue) |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
traffic, fp64 = {}, {}
for name, rs in seen.items():
    ms = sum(to_ms(g(r, "gpu__time_duration.sum"), unit("gpu__time_duration.sum")) for r in rs) / len(rs)
    byt = sum(to_bytes(g(r, "dram__bytes_read.sum"), unit("dram__bytes_read.sum")) + to_bytes(g(r, "dram__bytes_write.sum"), unit("dram__bytes_write.sum")) for r in rs) / len(rs)
    f64 = sum(g(r, "smsp__inst_executed_pipe_fp64.sum", 0.0) for r in rs) / len(rs)
    r = rs[len(rs) // 2]
    stalls = sorted(((g(r, k, 0.0), k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]) for k in stall_keys), reverse=True)[:4]
    inst = g(r, "smsp__inst_executed.sum"); nw = g(r, "launch__grid_size") * g(r, "launch__block_size") / 32.0
    key = re.sub(r"<\d>|_tma", "", name)
    traffic.setdefault(key, []).append(byt); fp64.setdefault(key, []).append(f64)
    print(f"| {name} | {len(rs)} | {ms:.3f} | {byt/1e9:.2f} | {byt/ms/1e6:.0f} | {g(r,'lts__t_sector_hit_rate.pct'):.0f} | "
          f"{g(r,'launch__registers_per_thread'):.0f} | {g(r,'sm__warps_active.avg.per_cycle_active'):.1f} | {g(r,'sm__inst_executed.avg.per_cycle_elapsed'):.2f} | "
          f"{g(r,'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active'):.0f} | {f64/1e6:.1f} | {f64/FP64_RATE*1e3:.3f} | {inst/nw:.0f} | "
          + ", ".join(f"{n} {v:.1f}" for v, n in stalls) + " |")
note = f"ncu --set full, {src}; mean over the captured launches of each kernel (template instances and _tma variants under the plain name: max)"
json.dump({"workload": workload, "n_gpus": 1, "source": note + "; dram__bytes_read.sum + dram__bytes_write.sum per launch",
           "bytes_per_launch": {k: int(max(v)) for k, v in sorted(traffic.items())}}, open(jt, "w"), indent=1)
json.dump({"workload": workload, "n_gpus": 1, "source": note + "; smsp__inst_executed_pipe_fp64.sum (warp instructions) per launch",
           "fp64_warp_inst_per_launch": {k: int(max(v)) for k, v in sorted(fp64.items())}}, open(jf, "w"), indent=1)
