"""Parity tests of the MOLOCH core (idynamic = 3) on the GPU: the CUDA library through the C ABI
against the CPU oracle on the same seeded synthetic inputs (all option variants of tests/test_moloch.py,
including the ROTLLR projection branches and the sponge boundary option), against the committed golden fixtures
tests/golden/mo_*.npz, and through the reference's own analytic cases at a size the oracle is not
run at.  Tolerances (double precision, no FMA contraction on either side; the arithmetic differences
are CUDA libm `pow`/`exp` vs glibc, <= 2 ulp, in paicompute and in the p / ps closure):
  * static fields and the state after one step       <= 1e-12 relative L-inf
  * prognostic fields after N <= 10 steps            <= 1e-10 relative L-inf
  * after 70 steps with a boundary-record change     <= 1e-8  relative L-inf
w is normalised by max|w| of the reference (O(0.5 m/s)).  The measured errors are written to
gpurun_out/moloch_parity.json when that directory exists."""
import glob
import json
import os

import numpy as np
import pytest

from regcm_b200 import capi, synth
from tests.golden_util import load_golden
from tests.helpers import make_pair, rel_errors, work_errors
from tests.test_moloch import MOLOCH_VARIANTS, _analytic_case

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LOG = {}


def _worst(errs):
    return max(e[0] for e in errs.values())


def _record(key, value):
    _LOG[key] = value
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "moloch_parity.json"), "w") as fh:
            json.dump(_LOG, fh, indent=1, sort_keys=True)


def test_moloch_statics_and_one_step(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "smallm")
    e0 = rel_errors(core, ref)
    _record("init_state", {k: v[0] for k, v in e0.items()})
    assert _worst(e0) <= 1e-13, e0
    core.step(1); ref.step(1)
    w = work_errors(core, ref)
    _record("one_step_work", w)
    assert max(w.values()) <= 1e-12, w
    e = rel_errors(core, ref)
    _record("one_step_state", {k: v[0] for k, v in e.items()})
    assert _worst(e) <= 1e-12, e
    # nadv * (nsound * 3 + 5) + pre + close + boundary kernels
    assert core.launch_count() >= 3 * (6 * 3 + 5) + 2


@pytest.mark.parametrize("name,nsteps", [("tinym", 10), ("smallm", 6)])
def test_moloch_parity_vs_oracle(cuda_lib, name, nsteps):
    case, core, ref = make_pair(cuda_lib, name)
    core.step(nsteps); ref.step(nsteps)
    errs = rel_errors(core, ref)
    _record(f"parity_{name}_{nsteps}", {k: v[0] for k, v in errs.items()})
    assert _worst(errs) <= 1e-10, errs
    dc, dr = core.diag(), ref.diag()
    assert dc.lcount == dr.lcount and dc.nan_flag == 0
    assert abs(dc.ps_max - dr.ps_max) <= 1e-10 * dr.ps_max and abs(dc.ps_min - dr.ps_min) <= 1e-10 * dr.ps_min


@pytest.mark.parametrize("kw", MOLOCH_VARIANTS[1:], ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()))
def test_moloch_option_variants(cuda_lib, kw):
    case, core, ref = make_pair(cuda_lib, "tinym", **kw)
    core.step(4); ref.step(4)
    errs = rel_errors(core, ref)
    assert _worst(errs) <= 1e-10, errs


def test_moloch_boundary_record_change_and_long_run(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "tinym", ibdyfrq=1)
    core.step(70); ref.step(70)     # one boundary-record change at step 40
    assert core.diag().xbctime == ref.diag().xbctime
    errs = rel_errors(core, ref)
    _record("long_run_tinym_70", {k: v[0] for k, v in errs.items()})
    assert _worst(errs) <= 1e-8, errs


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "mo_*.npz"))))
def test_moloch_golden_fixtures(cuda_lib, path):
    case, nsteps, fields = load_golden(path)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    core.step(nsteps)
    for key, ref in fields.items():
        if not isinstance(key, tuple):
            continue
        got = core.get_state(key[0])
        sc = max(float(np.abs(ref).max()), 1e-300)
        assert float(np.abs(got - ref).max()) / sc <= 1e-10, capi.FIELD_NAMES[key[0]]
    d = core.diag()
    assert abs(d.ps_max - fields["ps_max"]) <= 1e-10 * fields["ps_max"]


def test_moloch_cloud_water_and_state_roundtrip(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "tinym")
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    J, I, K = np.meshgrid(np.arange(jx - 1), np.arange(iy - 1), np.arange(kz), indexing="ij")
    qc = np.asfortranarray(1e-4 * (1.0 + 0.5 * np.sin(2 * np.pi * J / (jx - 1)) * np.cos(2 * np.pi * I / (iy - 1))) * np.sin(np.pi * (K + 0.5) / kz))
    for c in (core, ref):
        c.set_state(capi.F_QC, capi.LEVEL_A, qc)
    np.testing.assert_array_equal(core.get_state(capi.F_QC), qc)
    core.step(5); ref.step(5)
    errs = rel_errors(core, ref)
    assert _worst(errs) <= 1e-10, errs


@pytest.mark.parametrize("test", [1, 2])
def test_moloch_analytic_cases_full_size(cuda_lib, test):
    """The reference's moloch_static_test1/2 (Main/mod_bdycod.F90:5157-5199) on 300x260x30: the state
    at rest stays at rest, the uniform flow stays uniform."""
    case = _analytic_case(test, name="C5", jx=300, iy=260)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    t0 = core.get_state(capi.F_T)
    core.step(8)
    assert core.diag().nan_flag == 0
    assert np.abs(core.get_state(capi.F_U) - (0.0 if test == 1 else 10.0)).max() < 1e-6
    assert np.abs(core.get_state(capi.F_V)).max() < 1e-6 and np.abs(core.get_state(capi.F_W)).max() < 1e-5
    assert np.abs(core.get_state(capi.F_T) - t0).max() < 1e-5


def test_moloch_error_codes(cuda_lib):
    case = synth.make_case("tinym", ibdyfrq=1)
    core = capi.DynCore(cuda_lib, case.cfg)
    with pytest.raises(capi.RcmError) as ei:
        core.step(1)
    assert ei.value.code == -7
    synth.load_case(core, case, stage_next=False)
    core.step(39)
    with pytest.raises(capi.RcmError) as ei:
        core.step(1)
    assert ei.value.code == -6
