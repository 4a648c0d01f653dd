"""rcm_dyn_step_async / rcm_dyn_sync / rcm_dyn_get_state_async on the GPU: the asynchronous path must give exactly
the state of rcm_dyn_step, the export taken in the middle of a batch is the state at that point (device snapshot in
stream order) although later steps run while it is copied, and errors surface at rcm_dyn_sync."""
import numpy as np
import pytest

from regcm_b200 import capi, synth
from tests.helpers import prognostics

pytestmark = pytest.mark.gpu


def _pinned(shape):
    import torch
    t = torch.empty(tuple(shape)[::-1], dtype=torch.float64).pin_memory()
    return t, t.numpy().T


@pytest.mark.parametrize("name,kw", [("C3", {}), ("small", {"idynamic": 1}), ("smallm", {})])
def test_async_equals_sync(cuda_lib, name, kw):
    case = synth.make_case(name, **kw)
    a = capi.DynCore(cuda_lib, case.cfg); synth.load_case(a, case)
    b = capi.DynCore(cuda_lib, case.cfg); synth.load_case(b, case)
    c = capi.DynCore(cuda_lib, case.cfg); synth.load_case(c, case)
    nsteps, mid = 12, 5
    a.step(nsteps)
    c.step(mid)
    fields = prognostics(b)
    keep, outs = [], {}
    seen = []
    for n in range(nsteps):
        b.step_async(1)
        seen.append(b.diag().lcount)
        if n + 1 == mid:
            for f in fields:
                shp = b.field_shape(f)
                t, v = _pinned(shp)
                keep.append(t); outs[f] = v
                b.get_state_async(f, capi.LEVEL_A, v, (1, 1, 1), shp)
    b.sync()
    b.get_state_wait()
    assert all(0 <= s <= n + 1 for n, s in enumerate(seen)) and seen == sorted(seen)
    for f in fields:
        np.testing.assert_array_equal(outs[f], c.get_state(f, capi.LEVEL_A), err_msg="snapshot " + capi.FIELD_NAMES[f])
        for lv in ((capi.LEVEL_A,) if case.cfg.idynamic == 3 else (capi.LEVEL_A, capi.LEVEL_B)):
            np.testing.assert_array_equal(a.get_state(f, lv), b.get_state(f, lv), err_msg=capi.FIELD_NAMES[f])
    da, db = a.diag(), b.diag()
    assert da.lcount == db.lcount == nsteps and da.max_cfl == db.max_cfl and da.neg_q_points == db.neg_q_points


def test_async_error_surfaces_at_sync(cuda_lib):
    case = synth.make_case("tiny", dt=2000.0)      # far beyond the acoustic CFL limit
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    for _ in range(20):
        core.step_async(1)                          # enqueueing does not fail
    with pytest.raises(capi.RcmError) as ei:
        core.sync()
    assert ei.value.code in (-3, -4)


@pytest.mark.parametrize("name,kw", [("tiny", {}), ("tiny", {"idynamic": 1}), ("tinym", {})])
def test_async_boundary_staging_through_a_record_change(cuda_lib, name, kw):
    """rcm_dyn_stage_bdy_async (one contiguous H2D copy into a landing buffer + device unpack) while steps are in
    flight, consumed by the record change inside the batch: same state as the synchronous path."""
    case = synth.make_case(name, ibdyfrq=1, **kw)
    a = capi.DynCore(cuda_lib, case.cfg); synth.load_case(a, case, stage_next=True)     # set_bdy(BDY_NEXT): synchronous
    b = capi.DynCore(cuda_lib, case.cfg); synth.load_case(b, case, stage_next=False)
    nsteps = int(round(3600.0 / case.cfg.dt)) + 4
    a.step(nsteps)
    b.step_async(3)
    keep = []
    for fid, arr in case.bdy[capi.BDY_NEXT].items():
        shp = arr.shape if arr.ndim == 3 else arr.shape + (1,)
        t, v = _pinned(shp)
        v[...] = arr.reshape(shp, order="F")
        keep.append(t)
        b.stage_bdy_async(fid, v if arr.ndim == 3 else v[:, :, 0])
    b.step_async(nsteps - 3)
    b.sync(); b.stage_wait()
    for f in prognostics(a):
        np.testing.assert_array_equal(a.get_state(f), b.get_state(f), err_msg=capi.FIELD_NAMES[f])
