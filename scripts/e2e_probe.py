"""Where does the end-to-end time go?  Wall-clock of the pieces of bench.py's e2e leg, alone and combined (C4)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from regcm_b200 import capi, synth

wl = sys.argv[1] if len(sys.argv) > 1 else "C4"
K = 20
lib = capi.cuda_lib()
base = synth.CASES[wl]
case = synth.make_case(wl)
core = capi.DynCore(lib, case.cfg)
synth.load_case(core, case, stage_next=True)
core.step(4)
rec = synth.make_case(wl, records=(capi.BDY_NEXT,)).bdy[capi.BDY_NEXT]
pin_in, pin_out = {}, {}
for fid, a in rec.items():
    t = torch.empty(a.shape[::-1], dtype=torch.float64).pin_memory(); v = t.numpy().T; v[...] = a; pin_in[fid] = (t, v)
for fid in range(7):
    shp = core.field_shape(fid)
    t = torch.empty(shp[::-1], dtype=torch.float64).pin_memory(); pin_out[fid] = (t, t.numpy().T, shp)

def export():
    for fid, (t, v, shp) in pin_out.items():
        core.get_state_async(fid, capi.LEVEL_A, v, (1, 1, 1), shp)
def stage():
    for fid, (t, v) in pin_in.items():
        core.stage_bdy_async(fid, v)
def steps():
    for _ in range(K):
        core.step_async(1); core.diag()
def wait():
    core.sync(); core.stage_wait(); core.get_state_wait(); torch.cuda.synchronize()

def timed(name, *fns):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for f in fns: f()
    wait()
    dt = time.perf_counter() - t0
    print(f"{name:28s} {dt*1e3:8.1f} ms   ({dt*1e3/K:6.2f} ms per step of {K})", flush=True)

export(); wait()
timed("steps only", steps)
timed("export only (D2H 7 fields)", export)
timed("stage only (H2D 6 fields)", stage)
timed("export + steps", export, steps)
timed("stage + steps", stage, steps)
timed("stage + export", stage, export)
timed("stage + export + steps", stage, export, steps)
timed("steps(K) one call", lambda: core.step_async(K))
print("device ms of last batch", core.last_step_ms())
