"""Parity tests proper: the CUDA library (through the C ABI) against the CPU
oracle on the same seeded synthetic inputs, and against committed golden
fixtures.  Tolerances (double precision, no FMA contraction on either side, so
the only arithmetic difference is CUDA libm `pow` vs glibc `pow`, <= 2 ulp, in
theta and in the log-linear vertical qv flux):
  * work arrays after one step          <= 1e-13 relative L-inf
  * prognostic fields after N <= 12 steps <= 1e-11 relative L-inf
  * after ~200 steps (record change)    <= 1e-9  relative L-inf  
"""
import glob
import os

import numpy as np
import pytest

from regcm_b200 import capi, synth
from tests.golden_util import load_golden
from tests.helpers import make_pair, rel_errors, smooth_qc, work_errors

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worst(errs):
    return max(e[0] for e in errs.values())


def test_one_step_work_arrays(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "small")
    core.step(1); ref.step(1)
    w = work_errors(core, ref)
    assert max(w.values()) <= 1e-13, w
    assert _worst(rel_errors(core, ref)) <= 1e-13
    assert core.launch_count() > 20


@pytest.mark.parametrize("name,nsteps", [("tiny", 12), ("small", 8), ("C2", 10)])
def test_parity_vs_oracle(cuda_lib, name, nsteps):
    case, core, ref = make_pair(cuda_lib, name)
    core.step(nsteps); ref.step(nsteps)
    errs = rel_errors(core, ref)
    assert _worst(errs) <= 1e-11, errs
    dc, dr = core.diag(), ref.diag()
    assert dc.istep == dr.istep and dc.lcount == dr.lcount and dc.neg_q_points == 0
    assert abs(dc.max_cfl - dr.max_cfl) <= 1e-11 * dr.max_cfl


VARIANTS = [{"iboudy": 1}, {"iboudy": 0}, {"iboudy": 2}, {"iboudy": 3}, {"iboudy": 4}, {"ifupr": 0}, {"ifrayd": 0}, {"ipptls": 0}, {"diffu_hgtf": 1},
            {"upstream_mode": 0}, {"stability_enhance": 0}, {"flat": True, "unit_map": True}]


@pytest.mark.parametrize("kw", VARIANTS, ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()))
def test_parity_option_variants(cuda_lib, kw):
    case, core, ref = make_pair(cuda_lib, "tiny", **kw)
    core.step(6); ref.step(6)
    errs = rel_errors(core, ref)
    assert _worst(errs) <= 1e-11, errs


def test_boundary_record_change_and_long_run(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "tiny", ibdyfrq=2)
    core.step(110); ref.step(110)   # one boundary-record change at step 60
    assert core.diag().xbctime == ref.diag().xbctime
    errs = rel_errors(core, ref)
    assert _worst(errs) <= 1e-9, errs


@pytest.mark.parametrize("name", ["small", "C2"])
def test_staged_kernels_equal_plain_kernels(cuda_lib, name, monkeypatch):
    """The TMA/shared-memory staged variants of the slow-tendency kernels and of the forward sweep share their
    arithmetic with the plain kernels (k_*_level.inc): results must be bit-identical."""
    case = synth.make_case(name)
    for v in ("RCM_UV_TMA", "RCM_TQ_TMA", "RCM_PPW_TMA", "RCM_WFWD_TMA"):
        monkeypatch.setenv(v, "1")
    a = capi.DynCore(cuda_lib, case.cfg); synth.load_case(a, case)
    for v in ("RCM_UV_TMA", "RCM_TQ_TMA", "RCM_PPW_TMA", "RCM_WFWD_TMA"):
        monkeypatch.setenv(v, "0")
    b = capi.DynCore(cuda_lib, case.cfg); synth.load_case(b, case)
    a.step(5); b.step(5)
    for f in range(7):
        for lv in (0, 1):
            assert np.array_equal(a.get_state(f, lv), b.get_state(f, lv)), (f, lv)
    for w in range(10):
        assert np.array_equal(a.get_work(w), b.get_work(w)), w


def test_async_record_staging(cuda_lib):
    """rcm_dyn_stage_bdy_async while steps run == synchronous rcm_dyn_set_bdy before them."""
    case = synth.make_case("tiny", ibdyfrq=2)
    a = capi.DynCore(cuda_lib, case.cfg); synth.load_case(a, case, stage_next=True)
    b = capi.DynCore(cuda_lib, case.cfg); synth.load_case(b, case, stage_next=False)
    b.step(20)
    keep = {fid: np.asfortranarray(arr) for fid, arr in case.bdy[capi.BDY_NEXT].items()}
    for fid, arr in keep.items():
        b.stage_bdy_async(fid, arr)
    b.step(90)          # crosses the record change at step 60 without an explicit wait
    b.stage_wait()
    a.step(110)
    for f in range(7):
        for lv in (0, 1):
            assert np.array_equal(a.get_state(f, lv), b.get_state(f, lv)), (f, lv)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "nh_*.npz"))))
def test_golden_fixtures(cuda_lib, path):
    case, nsteps, fields = load_golden(path)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    core.step(nsteps)
    for (f, lv), ref in fields.items():
        got = core.get_state(f, lv)
        sc = max(float(np.abs(ref).max()), 1e-300)
        assert float(np.abs(got - ref).max()) / sc <= 1e-11, (capi.FIELD_NAMES[f], lv)


def test_c3_full_size_vs_oracle(cuda_lib):
    """CORDEX-like 250x200x23 domain (BASELINE config 3): 2 steps against the oracle."""
    case, core, ref = make_pair(cuda_lib, "C3")
    core.step(2); ref.step(2)
    assert _worst(rel_errors(core, ref)) <= 1e-12


def test_rest_state_full_size(cuda_lib):
    """Size-independent property at the C4-quarter size (500x500x41): a balanced
    atmosphere at rest stays at rest (cf. tests/test_oracle.py)."""
    case = synth.make_case("C4q", flat=True, unit_map=True, noise=False)
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    sig = case.statics[capi.S_SIGMA]
    hs = 0.5 * (sig[1:] + sig[:-1])
    p0, tlp, ts0, ptop = 101325.0, 47.70, 288.15, case.cfg.ptop * 1000.0
    ps0 = p0 - ptop
    t0 = np.maximum(ts0 + tlp * np.log((ps0 * hs + ptop) / p0), 216.65)
    for slot in case.bdy:
        b = case.bdy[slot]
        b[capi.F_U] = np.zeros_like(b[capi.F_U]); b[capi.F_V] = np.zeros_like(b[capi.F_V])
        b[capi.F_T] = np.asfortranarray(np.broadcast_to(t0, (jx - 1, iy - 1, kz)).copy())
        b[capi.F_QV] = np.full_like(b[capi.F_QV], 1.0e-8)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    core.step(10)
    d = core.diag()
    assert d.nan_flag == 0 and d.max_cfl < 1e-3
    assert np.abs(core.get_state(capi.F_U)).max() / 96.0 < 1e-4
    assert np.abs(core.get_state(capi.F_W)).max() / 96.0 < 1e-5


def test_error_codes(cuda_lib):
    case = synth.make_case("tiny", ibdyfrq=1)
    core = capi.DynCore(cuda_lib, case.cfg)
    with pytest.raises(capi.RcmError) as ei:
        core.step(1)                      # step before init
    assert ei.value.code == -7
    synth.load_case(core, case, stage_next=False)
    core.step(29)
    with pytest.raises(capi.RcmError) as ei:
        core.step(1)                      # record change without a staged record
    assert ei.value.code == -6
    # CFL violation is reported, not silently integrated (Main/mod_sound.F90:650-665)
    case2 = synth.make_case("tiny", dt=2000.0)
    core2 = capi.DynCore(cuda_lib, case2.cfg)
    synth.load_case(core2, case2)
    with pytest.raises(capi.RcmError) as ei:
        core2.step(20)
    assert ei.value.code in (-3, -4)


def test_set_state_roundtrip_and_qc_path(cuda_lib):
    """Non-zero cloud water exercises hadvqx / vadv4d / water loading on both sides."""
    case, core, ref = make_pair(cuda_lib, "tiny")
    qc = smooth_qc(case)
    for c in (core, ref):
        c.set_state(capi.F_QC, capi.LEVEL_A, qc)
        c.set_state(capi.F_QC, capi.LEVEL_B, qc)
    np.testing.assert_array_equal(core.get_state(capi.F_QC), qc)
    core.step(5); ref.step(5)
    errs = rel_errors(core, ref)
    assert _worst(errs) <= 1e-11, errs
    assert np.abs(ref.get_state(capi.F_QC)).max() > 0
