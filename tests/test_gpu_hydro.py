"""Parity tests of the hydrostatic core (idynamic = 1) on the GPU: the CUDA
library through the C ABI against the CPU oracle on the same seeded synthetic
inputs, and against the committed golden fixtures tests/golden/hy_*.npz.
Tolerances (double precision, no FMA contraction on either side; the arithmetic
differences are CUDA libm `pow`/`log` vs glibc, <= 2 ulp, in the vertical
T advection, the log-linear qv flux, the geopotential and the log-p pressure
gradient):
  * vertical-mode tables (two independent host solvers: oracle/linalg_ref.hpp vs csrc/linalg_small.hpp)  <= 1e-11 relative
  * work arrays after one step                       <= 1e-12 relative L-inf
  * prognostic fields after N <= 12 steps            <= 1e-10 relative L-inf
  * after 110 steps with a boundary-record change    <= 1e-8  relative L-inf
The measured errors are written to gpurun_out/hydro_parity.json when that
directory exists (copied into profiles/ by hand)."""
import glob
import json
import os

import numpy as np
import pytest

from regcm_b200 import capi, synth
from tests.golden_util import load_golden
from tests.helpers import make_pair, rel_errors, smooth_qc, work_errors

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LOG = {}


def _worst(errs):
    return max(e[0] for e in errs.values())


def _record(key, value):
    _LOG[key] = value
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "hydro_parity.json"), "w") as fh:
            json.dump(_LOG, fh, indent=1, sort_keys=True)


def test_hydro_tables_and_one_step(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "small", idynamic=1)
    for t in (capi.T_HBAR, capi.T_ZMATX, capi.T_ZMATXR, capi.T_TAU, capi.T_VARPA1, capi.T_AM, capi.T_AN):
        a, b = core.get_table(t), ref.get_table(t)
        assert np.abs(a - b).max() <= 1e-11 * np.abs(b).max(), t
    core.step(1); ref.step(1)
    w = work_errors(core, ref)
    _record("one_step_work", w)
    assert max(w.values()) <= 1e-12, w
    e = rel_errors(core, ref)
    _record("one_step_state", {k: v[0] for k, v in e.items()})
    assert _worst(e) <= 1e-12, e
    assert core.launch_count() >= 20


@pytest.mark.parametrize("name,nsteps", [("tiny", 12), ("small", 8), ("C1", 10)])
def test_hydro_parity_vs_oracle(cuda_lib, name, nsteps):
    case, core, ref = make_pair(cuda_lib, name, idynamic=1)
    core.step(nsteps); ref.step(nsteps)
    errs = rel_errors(core, ref)
    _record(f"parity_{name}_{nsteps}", {k: v[0] for k, v in errs.items()})
    assert _worst(errs) <= 1e-10, errs
    dc, dr = core.diag(), ref.diag()
    assert dc.lcount == dr.lcount and dc.neg_q_points == 0 and dc.nan_flag == 0
    assert abs(dc.ptntot - dr.ptntot) <= 1e-9 * dr.ptntot and abs(dc.pt2tot - dr.pt2tot) <= 1e-7 * dr.pt2tot


HYDRO_VARIANTS = [{"iboudy": 1}, {"iboudy": 0}, {"iboudy": 2}, {"iboudy": 3}, {"iboudy": 4}, {"ipptls": 0}, {"diffu_hgtf": 0}, {"upstream_mode": 0},
                  {"stability_enhance": 0}, {"flat": True, "unit_map": True}, {"lstand": 0}, {"nsplit": 1}, {"nsplit": 3}]


@pytest.mark.parametrize("kw", HYDRO_VARIANTS, ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()))
def test_hydro_option_variants(cuda_lib, kw):
    case, core, ref = make_pair(cuda_lib, "tiny", idynamic=1, **kw)
    core.step(6); ref.step(6)
    errs = rel_errors(core, ref)
    assert _worst(errs) <= 1e-10, errs


def test_hydro_boundary_record_change_and_long_run(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "tiny", idynamic=1, ibdyfrq=2)
    core.step(110); ref.step(110)   # one boundary-record change at step 60
    assert core.diag().xbctime == ref.diag().xbctime
    errs = rel_errors(core, ref)
    _record("long_run_tiny_110", {k: v[0] for k, v in errs.items()})
    assert _worst(errs) <= 1e-8, errs


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "hy_*.npz"))))
def test_hydro_golden_fixtures(cuda_lib, path):
    case, nsteps, fields = load_golden(path)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    core.step(nsteps)
    assert np.abs(core.get_table(capi.T_HBAR) - fields["hbar"]).max() <= 1e-12 * fields["hbar"].max()
    for key, ref in fields.items():
        if not isinstance(key, tuple):
            continue
        f, lv = key
        got = core.get_state(f, lv)
        sc = max(float(np.abs(ref).max()), 1e-300)
        assert float(np.abs(got - ref).max()) / sc <= 1e-10, (capi.FIELD_NAMES[f], lv)
    assert abs(core.diag().ptntot - fields["ptntot"]) <= 1e-9 * fields["ptntot"]


def test_hydro_state_roundtrip_and_qc_path(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "tiny", idynamic=1)
    qc = smooth_qc(case)
    for c in (core, ref):
        c.set_state(capi.F_QC, capi.LEVEL_A, qc)
        c.set_state(capi.F_QC, capi.LEVEL_B, qc)
    np.testing.assert_array_equal(core.get_state(capi.F_QC), qc)
    ps = ref.get_state(capi.F_PS)
    core.set_state(capi.F_PS, capi.LEVEL_A, ps)
    np.testing.assert_array_equal(core.get_state(capi.F_PS), ps)
    core.step(5); ref.step(5)
    errs = rel_errors(core, ref)
    assert _worst(errs) <= 1e-10, errs
    assert np.abs(ref.get_state(capi.F_QC)).max() > 0


def test_hydro_rest_state_full_size(cuda_lib):
    """Size-independent property at a size the oracle is not run at (500x400x23): a
    horizontally uniform atmosphere at rest over flat terrain stays at rest."""
    case = synth.make_case("C3", idynamic=1, jx=500, iy=400, flat=True, unit_map=True, noise=False, dt=100.0)
    tcol = case.bdy[capi.BDY_B0][capi.F_T][0:1, 0:1, :].copy()
    qcol = case.bdy[capi.BDY_B0][capi.F_QV][0:1, 0:1, :].copy()
    for slot in case.bdy:
        b = case.bdy[slot]
        b[capi.F_U][...] = 0.0
        b[capi.F_V][...] = 0.0
        b[capi.F_T][...] = tcol
        b[capi.F_QV][...] = qcol
        b[capi.F_PS][...] = 1000.0
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    t0, ps0 = core.get_state(capi.F_T), core.get_state(capi.F_PS)
    core.step(10)
    assert core.diag().nan_flag == 0
    assert np.abs(core.get_state(capi.F_U)).max() < 1e-8 and np.abs(core.get_state(capi.F_V)).max() < 1e-8
    assert np.abs(core.get_state(capi.F_PS) - ps0).max() < 1e-9
    assert np.abs(core.get_state(capi.F_T) - t0).max() / np.abs(t0).max() < 1e-11


def test_hydro_error_codes(cuda_lib):
    case = synth.make_case("tiny", idynamic=1, ibdyfrq=1)
    core = capi.DynCore(cuda_lib, case.cfg)
    with pytest.raises(capi.RcmError) as ei:
        core.step(1)                      # step before init
    assert ei.value.code == -7
    synth.load_case(core, case, stage_next=False)
    core.step(29)
    with pytest.raises(capi.RcmError) as ei:
        core.step(1)                      # record change without a staged record
    assert ei.value.code == -6
