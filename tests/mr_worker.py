"""Worker of the multi-rank tests: one process = one rank = one tile.

Runs the dyn-core on a 2-D decomposition and writes this rank's tile of every
prognostic to <outdir>/rank<r>.npz.  Two transports:
  --lib emu   host harness (tests/emu) + torch.distributed/gloo callbacks  (CPU, no GPU)
  --lib cuda  the product library + NCCL                                   (one GPU per rank)
Launched by tests/test_multirank_emu.py / test_gpu_multirank.py with RANK,
WORLD_SIZE, MASTER_ADDR, MASTER_PORT in the environment.
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from regcm_b200 import capi, synth  # noqa: E402

FIELDS = (capi.F_U, capi.F_V, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PP, capi.F_W)
HYDRO_FIELDS = (capi.F_U, capi.F_V, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PS)
MOLOCH_FIELDS = (capi.F_U, capi.F_V, capi.F_W, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PAI, capi.F_PS)


def tile_window(tile, jx, iy, pad=3):
    j1, j2, i1, i2 = tile
    return (max(1, j1 - pad), min(jx, j2 + pad), max(1, i1 - pad), min(iy, i2 + pad))


def owned_bounds(core, field):
    """Global index range of the points of `field` this rank owns."""
    j1, j2, i1, i2 = core.tile()
    kz = core.kz
    if core.cfg.idynamic == 3:   # Arakawa C grid: u (jx, iy-1), v (jx-1, iy), ps 2-D
        if field == capi.F_U:
            return (j1, i1, 1), (j2, min(i2, core.iy - 1), kz)
        if field == capi.F_V:
            return (j1, i1, 1), (min(j2, core.jx - 1), i2, kz)
        if field == capi.F_PS:
            return (j1, i1, 1), (min(j2, core.jx - 1), min(i2, core.iy - 1), 1)
    if field in (capi.F_U, capi.F_V):
        return (j1, i1, 1), (j2, i2, kz)
    if field == capi.F_PS:
        return (j1, i1, 1), (min(j2, core.jx - 1), min(i2, core.iy - 1), 1)
    nk = kz + 1 if field in (capi.F_W, capi.F_TKE) else kz
    return (j1, i1, 1), (min(j2, core.jx - 1), min(i2, core.iy - 1), nk)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", choices=("emu", "cuda"), required=True)
    ap.add_argument("--case", default="tiny")
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--cpus-j", type=int, default=0)
    ap.add_argument("--cpus-i", type=int, default=0)
    ap.add_argument("--ibdyfrq", type=int, default=6)
    ap.add_argument("--iboudy", type=int, default=-1)
    ap.add_argument("--idynamic", type=int, default=0, help="override the case's core (1 = hydrostatic)")
    ap.add_argument("--cfg", action="append", default=[], help="name=value override of an rcm_cfg member")
    ap.add_argument("--restart-at", type=int, default=0, help="stop after this many steps, carry the OWNED points of the state into "
                    "a new handle through the restart entry points, continue there")
    ap.add_argument("--work", action="store_true", help="also dump the per-step work arrays")
    ap.add_argument("--spiky-tracers", action="store_true", help="tracers with sharp edges (drives their negative fix)")
    ap.add_argument("--sharp-q", action="store_true", help="moisture with sharp fronts (drives the negative-q repair)")
    ap.add_argument("--phys", action="store_true", help="every step through pre_physics / set_phys_tend / post_physics with the synthetic "
                    "tendencies of tests/test_physics_coupling.py")
    ap.add_argument("--uw", action="store_true", help="ibltyp = 2 inputs of tests/test_tke.py: TKE state, TKE physics source, kpbl, cloud water")
    ap.add_argument("--from-files", default="", help="directory with case_DOMAIN000.nc / case_ICBC.*.nc written by tests/test_ncio.py: every "
                    "rank reads the files itself (the reference's do_parallel_netcdf_in path) and hands the library global arrays")
    ap.add_argument("--out", required=True)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)

    if args.lib == "emu":
        lib = capi.CoreLib(os.path.join(ROOT, "tests", "emu", "libregcm_dyn_emu.so"), "rcm_dyn_")

        def as_tensor(ptr, n):
            arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(n,))
            return torch.from_numpy(arr)

        @C.CFUNCTYPE(C.c_int, C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong)
        def sendrecv(peer, sbuf, ns, rbuf, nr):
            try:
                reqs = [dist.isend(as_tensor(sbuf, ns), peer), dist.irecv(as_tensor(rbuf, nr), peer)]
                for r in reqs:
                    r.wait()
                return 0
            except Exception as e:  # pragma: no cover
                print("sendrecv failed:", e, flush=True)
                return 1

        @C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int)
        def allreduce(v, n, op):
            try:
                dist.all_reduce(as_tensor(v, n), op=dist.ReduceOp.MAX if op == 1 else dist.ReduceOp.SUM)
                return 0
            except Exception as e:  # pragma: no cover
                print("allreduce failed:", e, flush=True)
                return 1

        lib.lib.rcm_emu_set_transport(sendrecv, allreduce)
        uid = b"\0" * 128
    else:
        lib = capi.cuda_lib()
        box = [capi.nccl_unique_id(lib) if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]

    over = {}
    if args.idynamic:
        over["idynamic"] = args.idynamic
    for kv in args.cfg:
        k, v = kv.split("=")
        over[k] = float(v) if "." in v else int(v)
    base = synth.make_case(args.case, records=(), ibdyfrq=args.ibdyfrq, **over)  # cfg only
    cfg = base.cfg
    if args.iboudy >= 0:
        cfg.iboudy = args.iboudy
    cfg.nproc, cfg.myid, cfg.cpus_j, cfg.cpus_i = world, rank, args.cpus_j, args.cpus_i
    if args.lib == "cuda":
        cfg.device = int(os.environ.get("LOCAL_RANK", rank))
    core = capi.DynCore(lib, cfg)
    core.comm_init(uid)
    win = tile_window(core.tile(), cfg.jx, cfg.iy)
    case = synth.make_case(args.case, window=win, ibdyfrq=args.ibdyfrq, **over)
    case.cfg = cfg
    if args.from_files:
        from regcm_b200 import ncio
        from tests.test_ncio import START, _hours
        ncio.load_domain(core, ncio.read_domain_info(os.path.join(args.from_files, "case_DOMAIN000.nc"), cfg))
        rd = ncio.IcbcReader(args.from_files, "case", cfg, START)
        units = "hours since 1949-12-01 00:00:00 UTC"
        for slot, k in ((capi.BDY_B0, 0), (capi.BDY_B1, 1), (capi.BDY_NEXT, 2)):
            rd.feed(core, slot, ncio.timeval2date(_hours(START) + k * cfg.ibdyfrq, units))
        core.init()
        rd.close()
    else:
        synth.load_case(core, case, stage_next=True)
    if args.spiky_tracers:
        from tests.test_tracers import _spiky
        for n in range(cfg.ntr):
            q = _spiky(base, n)                    # global array; the library keeps this rank's window
            for lv in (capi.LEVEL_A, capi.LEVEL_B):
                core.set_state(capi.F_CHI + n, lv, q)
    if args.uw:
        from tests.test_tke import load_uw, uw_inputs
        load_uw(core, case, *uw_inputs(base))      # global arrays; the library keeps this rank's window
    if args.sharp_q:
        from tests.helpers import sharp_moisture
        qv, qc = sharp_moisture(base)          # global arrays; the library keeps this rank's window
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            core.set_state(capi.F_QV, lv, qv)
            core.set_state(capi.F_QC, lv, qc)
    if args.restart_at > 0:
        # what write_savefile / read_savefile would do on every rank (tests/test_restart.py), owned points only:
        # rcm_dyn_set_clock refreshes the ghost rows of the restored state
        from tests.test_restart import fields_of
        core.step(args.restart_at)
        flds, levels = fields_of(cfg)
        d = core.diag()
        sav = {}
        for f in flds:
            lb, ub = owned_bounds(core, f)
            for lv in levels:
                sav[(f, lv)] = (lb, core.get_state(f, lv, lb=lb, ub=ub))
        aux = {}
        j1, j2, i1, i2 = core.tile()
        clb = (j1, i1, 1)
        cub = lambda nk: (min(j2, cfg.jx - 1), min(i2, cfg.iy - 1), nk)
        if cfg.idynamic == 1:
            for rid in (capi.R_DSTOR, capi.R_HSTOR):
                a = np.zeros(tuple(cub(cfg.nsplit)[q] - clb[q] + 1 for q in range(3)), order="F")
                assert lib.get_restart(core.h, rid, a.ctypes.data_as(C.POINTER(C.c_double)), *capi._bounds(clb, cub(cfg.nsplit))) == 0
                aux[rid] = a
        if cfg.idynamic == 2:
            aux[capi.R_TMASK] = core.get_restart(capi.R_TMASK)
        core.close()
        if args.lib == "cuda":
            box = [capi.nccl_unique_id(lib) if rank == 0 else None]
            dist.broadcast_object_list(box, src=0)
            uid = box[0]
        core = capi.DynCore(lib, cfg)
        core.comm_init(uid)
        synth.load_case(core, case, stage_next=True)
        if args.uw:     # kpbl and the physics tendency belong to the caller (kpbl_io of the SAV file), the state is restored below
            load_uw(core, case, *uw_inputs(base))
        for (f, lv), (lb, a) in sav.items():
            core.set_state(f, lv, a, lb=lb)
        for rid, a in aux.items():
            core.set_restart(rid, a, lb=(1, 1, 1) if rid == capi.R_TMASK else clb)
        core.set_clock(d.lcount, d.xbctime)
        core.step(args.steps - args.restart_at)
    elif args.phys:
        from tests.test_physics_coupling import _mo_tendencies, _phys_tendencies
        ten = _mo_tendencies(base) if cfg.idynamic == 3 else _phys_tendencies(base, cfg.idynamic == 1)
        for n in range(args.steps):
            core.step_pre_physics()
            for f, a in ten.items():
                core.set_phys_tend(f, a * (1.0 + 0.25 * n))      # global arrays; the library keeps this rank's window
            core.step_post_physics()
    else:
        core.step(args.steps)
    out = {"tile": np.array(core.tile())}
    moloch = cfg.idynamic == 3
    tracers = tuple(capi.F_CHI + n for n in range(cfg.ntr)) + ((capi.F_TKE,) if cfg.ibltyp == 2 and not moloch else ())
    for fld in (MOLOCH_FIELDS if moloch else HYDRO_FIELDS if cfg.idynamic == 1 else FIELDS) + tracers:
        for lev, tag in (((capi.LEVEL_A, "a"),) if moloch else ((capi.LEVEL_A, "a"), (capi.LEVEL_B, "b"))):
            lb, ub = owned_bounds(core, fld)
            out[f"f{fld}{tag}"] = core.get_state(fld, lev, lb=lb, ub=ub)
            out[f"lb{fld}"] = np.array(lb)
    if args.work:
        jx, iy, kz = cfg.jx, cfg.iy, cfg.kz
        j1, j2, i1, i2 = core.tile()
        for wid in range(10):
            if wid in (capi.W_UTEN, capi.W_VTEN):
                glb, gub = (2, 2, 1), (jx - 1, iy - 1, kz)
            elif wid == capi.W_PPTEN:
                glb, gub = (2, 2, 1), (jx - 2, iy - 2, kz)
            elif wid == capi.W_WTEN:
                glb, gub = (2, 2, 1), (jx - 2, iy - 2, kz + 1)
            elif wid == capi.W_QDOT:
                glb, gub = (1, 1, 1), (jx - 1, iy - 1, kz + 1)
            else:
                glb, gub = (1, 1, 1), (jx - 1, iy - 1, kz)
            lb = (max(glb[0], j1), max(glb[1], i1), 1)
            ub = (min(gub[0], j2), min(gub[1], i2), gub[2])
            a = np.zeros(tuple(ub[q] - lb[q] + 1 for q in range(3)), order="F")
            rc = lib.get_work(core.h, wid, a.ctypes.data_as(C.POINTER(C.c_double)), *capi._bounds(lb, ub))
            assert rc == 0
            out[f"w{wid}"] = a
            out[f"wlb{wid}"] = np.array(lb)
    d = core.diag()
    out["max_cfl"] = np.array(d.max_cfl)
    out["neg_q"] = np.array(d.neg_q_points)
    out["neg_chi"] = np.array(d.neg_chi_points)
    out["ptn"] = np.array([d.ptntot, d.pt2tot])
    out["ps_extremes"] = np.array([d.ps_max, d.ps_min])
    np.savez(os.path.join(args.out, f"rank{rank}.npz"), **out)
    core.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
