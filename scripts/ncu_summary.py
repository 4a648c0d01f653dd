#!/usr/bin/env python
"""Per-kernel summary table of an .ncu-rep (ncu --set full): time, DRAM bytes and
throughput, occupancy, registers, IPC, FP64-pipe use and the top warp-stall reasons.
  python scripts/ncu_summary.py report.ncu-rep [> profiles/xxx.md]"""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: n for n, h in enumerate(hdr)}
def g(r, k, d=float("nan")):
    try: return float(r[col[k]].replace(",", ""))
    except Exception: return d
def unit(k): return units[col[k]] if k in col else ""
def to_bytes(v, u):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1)
def to_ms(v, u):
    return v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "usecond": 1e-3, "msecond": 1.0, "nsecond": 1e-6, "second": 1e3}.get(u, 1)
stall_keys = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
seen = {}
for r in data:
    name = r[col["Kernel Name"]].split("(")[0]
    seen.setdefault(name, []).append(r)
print("| kernel | launches | ms | DRAM rd GB | DRAM wr GB | DRAM GB/s | % of 6552 GB/s | L2 hit % | regs | warps/SM | IPC/SM | fp64 pipe % | inst/warp | top stalls (warps per issue) |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
for name, rs in seen.items():
    r = rs[len(rs) // 2]
    ms = to_ms(g(r, "gpu__time_duration.sum"), unit("gpu__time_duration.sum"))
    rd = to_bytes(g(r, "dram__bytes_read.sum"), unit("dram__bytes_read.sum"))
    wr = to_bytes(g(r, "dram__bytes_write.sum"), unit("dram__bytes_write.sum"))
    warps = g(r, "sm__warps_active.avg.per_cycle_active")
    ipc = g(r, "sm__inst_executed.avg.per_cycle_elapsed")
    stalls = sorted(((g(r, k, 0.0), k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]) for k in stall_keys), reverse=True)[:4]
    inst = g(r, "smsp__inst_executed.sum"); nw = g(r, "launch__grid_size") * g(r, "launch__block_size") / 32.0
    print(f"| {name} | {len(rs)} | {ms:.3f} | {rd/1e9:.2f} | {wr/1e9:.2f} | {(rd+wr)/ms/1e6:.0f} | "
          f"{(rd+wr)/ms/1e6/65.52:.0f} | {g(r,'lts__t_sector_hit_rate.pct'):.0f} | "
          f"{g(r,'launch__registers_per_thread'):.0f} | {warps:.1f} | {ipc:.2f} | "
          f"{g(r,'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active'):.0f} | {inst/nw:.0f} | "
          + ", ".join(f"{n} {v:.1f}" for v, n in stalls) + " |")
