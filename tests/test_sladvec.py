"""isladvec = 1: semi-Lagrangian horizontal advection of qv, qc and the tracers
(Main/mod_sladvection.F90:67-229,335-479,545-660 as `tend` calls it, Main/mod_tendency.F90:995-1000,1283-1285,1361-1363,
1378-1380,1397-1399), with the quasi-monotonic limiter (iqmsl).

CPU suite: the CUDA sources through the host harness against the oracle — bit for bit for idynamic = 2, eigen-solver
tolerance for idynamic = 1.  The oracle is pinned by a NumPy restatement written from the Fortran
(tests/numpy_ref.py::sl_tendency through tests/test_numpy_ref.py::test_tracer_step_numpy_vs_oracle[sl, sl-noclip, sl-hydro])."""
import numpy as np
import pytest

from oracle import pyoracle
from regcm_b200 import capi, synth
from tests.helpers import make_pair, rel_errors, smooth_qc

NTR = 2


def _run(lib, name, steps, tol, checked=False, **kw):
    case, core, ref = make_pair(lib, name, oracle_checked=checked, isladvec=1, **kw)
    qc = smooth_qc(case)
    for c in (core, ref):
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            c.set_state(capi.F_QC, lv, qc)
        c.step(steps)
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) <= tol, errs
    for n in range(kw.get("ntr", 0)):
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            a, r = core.get_state(capi.F_CHI + n, lv), ref.get_state(capi.F_CHI + n, lv)
            assert np.abs(a - r).max() <= tol * np.abs(r).max()
    return core, ref


@pytest.mark.parametrize("kw", [{}, {"iboudy": 1}, {"iboudy": 0}, {"iboudy": 3}, {"iboudy": 4}, {"iqmsl": 0}, {"ntr": NTR}, {"idynamic": 1},
                                {"idynamic": 1, "ntr": NTR, "iboudy": 1}],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_emu_sladvec(emu_lib, kw):
    _run(emu_lib, "tiny", 6, 1e-11 if kw.get("idynamic") == 1 else 0.0, checked=True, **kw)


def test_sladvec_switch_is_live():
    out = []
    for sl in (0, 1):
        case = synth.make_case("tiny", isladvec=sl, ntr=1)
        o = pyoracle.make_oracle(case.cfg); synth.load_case(o, case)
        o.step(6)
        out.append((o.get_state(capi.F_QV), o.get_state(capi.F_CHI)))
    for a, b in zip(*out):
        assert np.isfinite(a).all() and np.abs(a - b).max() > 1e-6 * np.abs(a).max()


@pytest.mark.parametrize("idynamic", [2, 1])
def test_departure_point_too_far_is_an_error(emu_lib, idynamic):
    """The reference calls fatal when the departure point lies more than one cell away (Main/mod_sladvection.F90:152-157)."""
    case, core, ref = make_pair(emu_lib, "tiny", isladvec=1, idynamic=idynamic, dt=9000.0)
    for c in (core, ref):
        with pytest.raises(capi.RcmError) as e:
            c.step(1)
        assert e.value.code == capi.ERR_CFL if hasattr(capi, "ERR_CFL") else "CFL" in str(e.value)


def test_create_rejects_sladvec_where_it_is_not_built(emu_lib):
    cfg = synth.make_case("tinym", records=()).cfg
    cfg.isladvec = 1                                   # the reference switches it off for the MOLOCH core (mod_params.F90:1158-1162)
    with pytest.raises(capi.RcmError):
        capi.DynCore(emu_lib, cfg)


@pytest.mark.gpu
@pytest.mark.parametrize("kw", [{}, {"iboudy": 4}, {"ntr": NTR}, {"idynamic": 1, "ntr": NTR}],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_gpu_sladvec_vs_oracle(cuda_lib, kw):
    core, ref = _run(cuda_lib, "small", 8, 1e-11, **kw)
    core.close()
