"""Multi-GPU parity (needs >= 2 GPUs; skipped otherwise): the NCCL halo exchange of the
product library against the same library on one GPU, and against the CPU oracle."""
import numpy as np
import pytest

from regcm_b200 import capi, synth
from test_multirank_emu import assemble, compare, run_ranks

pytestmark = pytest.mark.gpu


def ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def single_gpu(cuda_lib, case_name, steps):
    case = synth.make_case(case_name)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    core.step(steps)
    ref = {(f, "ab"[lv]): core.get_state(f, lv) for f in range(7) for lv in (0, 1)}
    core.close()
    return case, ref


@pytest.mark.parametrize("world,layout,case_name", [(2, (0, 0), "small"), (2, (1, 2), "small"), (4, (2, 2), "C2"), (8, (0, 0), "C2")])
def test_nccl_halo_matches_single_gpu(cuda_lib, tmp_path, world, layout, case_name):
    if ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    steps = 6
    run_ranks(world, tmp_path, ["--case", case_name, "--steps", str(steps), "--cpus-j", str(layout[0]), "--cpus-i", str(layout[1])],
              lib="cuda")
    case, ref = single_gpu(cuda_lib, case_name, steps)
    glob = assemble(tmp_path, world, case.cfg)
    compare(glob, ref, 1e-12)
    # and against the CPU oracle (same tolerance as the single-GPU parity tests)
    from oracle import pyoracle
    orc = pyoracle.make_oracle(case.cfg)
    synth.load_case(orc, case)
    orc.step(steps)
    oref = {(f, "ab"[lv]): orc.get_state(f, lv) for f in range(7) for lv in (0, 1)}
    compare(glob, oref, 1e-10)


@pytest.mark.parametrize("world,layout,case_name", [(2, (0, 0), "smallm"), (2, (1, 2), "smallm"), (4, (2, 2), "smallm"), (8, (0, 0), "C5q")])
def test_moloch_halo_matches_single_gpu(cuda_lib, tmp_path, world, layout, case_name):
    """MOLOCH core (idynamic = 3) on a 2-D decomposition: equal to the one-GPU result of the same library
    (the exchange is exact) and within the single-GPU tolerance of the CPU oracle."""
    if ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    from test_multirank_emu import MOLOCH_FIELDS, assemble_moloch
    steps = 3
    run_ranks(world, tmp_path, ["--case", case_name, "--steps", str(steps), "--cpus-j", str(layout[0]), "--cpus-i", str(layout[1])],
              lib="cuda")
    case = synth.make_case(case_name)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    core.step(steps)
    glob, exts = assemble_moloch(tmp_path, world, case.cfg)
    for f in MOLOCH_FIELDS:
        a, b = glob[f], core.get_state(f)
        b = b if b.ndim == 3 else b[:, :, None]
        assert not np.isnan(a).any(), capi.FIELD_NAMES[f]
        assert float(np.abs(a - b).max()) <= 1e-12 * max(float(np.abs(b).max()), 1e-300), capi.FIELD_NAMES[f]
    d = core.diag()
    for e in exts:
        assert abs(e[0] - d.ps_max) <= 1e-12 * d.ps_max and abs(e[1] - d.ps_min) <= 1e-12 * d.ps_min
    core.close()
    if case_name == "smallm":
        from oracle import pyoracle
        orc = pyoracle.make_oracle(case.cfg)
        synth.load_case(orc, case)
        orc.step(steps)
        for f in MOLOCH_FIELDS:
            a, b = glob[f], orc.get_state(f)
            b = b if b.ndim == 3 else b[:, :, None]
            assert float(np.abs(a - b).max()) <= 1e-10 * max(float(np.abs(b).max()), 1e-300), capi.FIELD_NAMES[f]


@pytest.mark.parametrize("world,layout,case_name", [(2, (0, 0), "small"), (2, (1, 2), "small"), (4, (2, 2), "C1"), (8, (0, 0), "C3h")])
def test_hydro_halo_matches_single_gpu(cuda_lib, tmp_path, world, layout, case_name):
    """Hydrostatic core (idynamic = 1) on a 2-D decomposition: equal to the one-GPU result of the same library
    (lstand = 1: nothing is summed over ranks, the exchange is exact) and within the single-GPU tolerance of the
    CPU oracle (tests/test_gpu_hydro.py)."""
    if ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    from test_multirank_emu import HYDRO_FIELDS, assemble_hydro
    steps = 6
    run_ranks(world, tmp_path, ["--case", case_name, "--idynamic", "1", "--steps", str(steps), "--cpus-j", str(layout[0]),
                                "--cpus-i", str(layout[1])], lib="cuda")
    case = synth.make_case(case_name, idynamic=1)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    core.step(steps)
    glob, _ = assemble_hydro(tmp_path, world, case.cfg)
    for f in HYDRO_FIELDS:
        for lv, tag in ((0, "a"), (1, "b")):
            a, b = glob[(f, tag)], core.get_state(f, lv)
            b = b if b.ndim == 3 else b[:, :, None]
            assert not np.isnan(a).any(), capi.FIELD_NAMES[f]
            assert float(np.abs(a - b).max()) <= 1e-12 * max(float(np.abs(b).max()), 1e-300), capi.FIELD_NAMES[f] + tag
    core.close()
    if case_name != "C3h":
        from oracle import pyoracle
        orc = pyoracle.make_oracle(case.cfg)
        synth.load_case(orc, case)
        orc.step(steps)
        for f in HYDRO_FIELDS:
            a, b = glob[(f, "a")], orc.get_state(f, 0)
            b = b if b.ndim == 3 else b[:, :, None]
            assert float(np.abs(a - b).max()) <= 1e-11 * max(float(np.abs(b).max()), 1e-300), capi.FIELD_NAMES[f]


@pytest.mark.parametrize("idynamic", [2, 1])
def test_uw_idiffu2_halo_matches_single_gpu(cuda_lib, tmp_path, idynamic):
    """Round-2 options on two GPUs: TKE and kpbl (ibltyp = 2, iuwvadv = 1), one tracer, the nine-point diffusion (idiffu = 2,
    corner ghost points).  Same library, same arithmetic: the two-GPU result equals the one-GPU result."""
    import os
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    from tests.test_tke import load_uw, uw_inputs
    steps = 5
    opts = dict(ibltyp=2, iuwvadv=1, ntr=1, idiffu=2)
    args = ["--case", "small", "--idynamic", str(idynamic), "--steps", str(steps), "--cpus-j", "0", "--cpus-i", "0", "--uw"]
    for k, v in opts.items():
        args += ["--cfg", f"{k}={v}"]
    run_ranks(2, tmp_path, args, lib="cuda")
    case = synth.make_case("small", idynamic=idynamic, **opts)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    load_uw(core, case, *uw_inputs(case))
    core.step(steps)
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    for fld, nk in ((capi.F_TKE, kz + 1), (capi.F_CHI, kz), (capi.F_QC, kz), (capi.F_T, kz)):
        for lv, tag in ((0, "a"), (1, "b")):
            glob = np.full((jx - 1, iy - 1, nk), np.nan)
            for r in range(2):
                d = np.load(os.path.join(tmp_path, f"rank{r}.npz"))
                lb, a = d[f"lb{fld}"], d[f"f{fld}{tag}"]
                glob[lb[0] - 1:lb[0] - 1 + a.shape[0], lb[1] - 1:lb[1] - 1 + a.shape[1], :] = a
            want = core.get_state(fld, lv)
            assert not np.isnan(glob).any()
            assert np.abs(glob - want).max() <= 1e-12 * np.abs(want).max(), (fld, tag)
    core.close()


@pytest.mark.parametrize("idynamic", [2, 1])
def test_sladvec_halo_matches_single_gpu(cuda_lib, tmp_path, idynamic):
    """isladvec = 1 on two GPUs: tiles allocated with four ghost rows, the level-b species exchanged three rows wide through
    the peer-memory transport before k_sl (the staged kernels clamp their tiles against the wider rows).  Same library, same
    arithmetic: the two-GPU result equals the one-GPU result."""
    import os
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    steps = 5
    opts = dict(isladvec=1, ntr=2)
    args = ["--case", "small", "--idynamic", str(idynamic), "--steps", str(steps), "--cpus-j", "0", "--cpus-i", "0"]
    for k, v in opts.items():
        args += ["--cfg", f"{k}={v}"]
    run_ranks(2, tmp_path, args, lib="cuda")
    case = synth.make_case("small", idynamic=idynamic, **opts)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    core.step(steps)
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    for fld in (capi.F_CHI, capi.F_CHI + 1, capi.F_QV, capi.F_QC, capi.F_T):
        for lv, tag in ((0, "a"), (1, "b")):
            glob = np.full((jx - 1, iy - 1, kz), np.nan)
            for r in range(2):
                d = np.load(os.path.join(tmp_path, f"rank{r}.npz"))
                lb, a = d[f"lb{fld}"], d[f"f{fld}{tag}"]
                glob[lb[0] - 1:lb[0] - 1 + a.shape[0], lb[1] - 1:lb[1] - 1 + a.shape[1], :] = a
            want = core.get_state(fld, lv)
            assert not np.isnan(glob).any()
            assert np.abs(glob - want).max() <= 1e-12 * np.abs(want).max(), (fld, tag)
    core.close()
