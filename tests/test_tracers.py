"""Tracer transport (ichem = 1 with ntr species, physics off, ichebdy = 1): hadvtr, vadv4d(ind = 2),
nudge_chi, diffu_x4d, forecast, the negative fix with factor 0.1 and the RAW filter
(Main/mod_tendency.F90:548-586,1393-1411,1502-1511,1539-1543; Main/mod_advection.F90:666-720,895-914;
Main/chemlib/mod_che_bdyco.F90:118-420,565-760,792-986,1177-1217).

CPU suite: the CUDA sources through the host harness (tests/emu) against the oracle — the tracer kernel evaluates
the reference's expressions in the reference's order, so the comparison is bit for bit for idynamic = 2 — plus pins of the oracle
itself (independence of the other prognostics; the NumPy restatement of the tracer step is in tests/test_numpy_ref.py)."""
import numpy as np
import pytest

from oracle import pyoracle
from regcm_b200 import capi, synth
from tests.helpers import make_pair

NTR = 3


def _tracers(c, ntr=NTR):
    return [(c.get_state(capi.F_CHI + n, capi.LEVEL_A), c.get_state(capi.F_CHI + n, capi.LEVEL_B)) for n in range(ntr)]


@pytest.mark.parametrize("kw", [{}, {"iboudy": 1}, {"iboudy": 0}, {"iboudy": 4}, {"upstream_mode": 0}, {"idynamic": 1},
                                {"idynamic": 1, "iboudy": 1}],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_emu_tracers_bit_exact(emu_lib, kw):
    case, core, ref = make_pair(emu_lib, "tiny", oracle_checked=True, ntr=NTR, **kw)
    for c in (core, ref):
        c.step(6)
    # idynamic = 1: the split-explicit tables of the product and of the oracle come from two independent eigen-solvers
    # (tests/test_hydro.py), so the winds agree to ~1e-13 and the tracers they carry are compared at the same 1e-11
    tol = 1e-11 if kw.get("idynamic") == 1 else 0.0
    for n, ((a, b), (ra, rb)) in enumerate(zip(_tracers(core), _tracers(ref))):
        assert ra.max() > 1e-6 and np.isfinite(ra).all()
        assert np.abs(a - ra).max() <= tol * ra.max(), f"chi{n} level a"
        assert np.abs(b - rb).max() <= tol * rb.max(), f"chi{n} level b"
    assert core.diag().neg_chi_points == ref.diag().neg_chi_points


def test_emu_tracers_record_change(emu_lib):
    """chem_bdyin: chib0 <- chib1, the staged record coupled with the boundary p*, chibt by division by dtbdys."""
    case, core, ref = make_pair(emu_lib, "tiny", ntr=2, ibdyfrq=1)
    core.step(34); ref.step(34)      # record change at step 30
    assert core.diag().xbctime == ref.diag().xbctime
    for (a, b), (ra, rb) in zip(_tracers(core, 2), _tracers(ref, 2)):
        sc = np.abs(ra).max()
        assert np.abs(a - ra).max() <= 1e-11 * sc and np.abs(b - rb).max() <= 1e-11 * sc


def _spiky(case, n):
    """Coupled tracer field with top-hat patches and single-cell spikes next to empty air: advection and the
    4th-order diffusion undershoot around them, so the forecast goes negative at isolated points and in groups."""
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    rng = np.random.default_rng(100 + n)
    q = np.zeros((jx - 1, iy - 1, kz))
    q[4:9, 3:8, :] = 96.0 * 1.0e-6
    q[11:13, 6:12, 1::2] = 96.0 * 3.0e-6
    for _ in range(12):
        q[rng.integers(3, jx - 4), rng.integers(3, iy - 4), rng.integers(0, kz)] = 96.0 * 5.0e-6
    return np.asfortranarray(q)


@pytest.mark.parametrize("idynamic", [2, 1])
def test_emu_tracer_negative_fix(emu_lib, idynamic):
    """The sequential 3x3 repair (factor 0.1) is triggered and reproduced exactly, adjacent negatives included."""
    case, core, ref = make_pair(emu_lib, "tiny", ntr=2, idynamic=idynamic)
    for c in (core, ref):
        for n in range(2):
            q = _spiky(case, n)
            c.set_state(capi.F_CHI + n, capi.LEVEL_A, q)
            c.set_state(capi.F_CHI + n, capi.LEVEL_B, q)
    total = 0
    for _ in range(4):
        core.step(1); ref.step(1)
        assert core.diag().neg_chi_points == ref.diag().neg_chi_points
        total += ref.diag().neg_chi_points
        tol = 1e-11 if idynamic == 1 else 0.0
        for (a, b), (ra, rb) in zip(_tracers(core, 2), _tracers(ref, 2)):
            assert np.abs(a - ra).max() <= tol * ra.max() and np.abs(b - rb).max() <= tol * rb.max()
            assert ra[1:-1, 1:-1].min() >= 0.0
    assert total > 50


def test_tracers_leave_the_other_prognostics_alone():
    for idyn in (2, 1):
        ca = synth.make_case("tiny", idynamic=idyn, ntr=NTR)
        cb = synth.make_case("tiny", idynamic=idyn)
        oa, ob = pyoracle.make_oracle(ca.cfg), pyoracle.make_oracle(cb.cfg)
        synth.load_case(oa, ca); synth.load_case(ob, cb)
        oa.step(5); ob.step(5)
        for f in (capi.F_U, capi.F_V, capi.F_T, capi.F_QV, capi.F_QC):
            assert np.array_equal(oa.get_state(f), ob.get_state(f))


def test_create_rejects_bad_tracer_counts(emu_lib):
    cfg = synth.make_case("tiny", records=()).cfg
    cfg.ntr = capi.MAXTR + 1
    with pytest.raises(capi.RcmError):
        capi.DynCore(emu_lib, cfg)
    cfg = synth.make_case("tinym", records=()).cfg
    cfg.ntr = 1                     # the MOLOCH core's tracers are not built
    with pytest.raises(capi.RcmError):
        capi.DynCore(emu_lib, cfg)
