"""rcm_dyn_step_async / rcm_dyn_sync / rcm_dyn_get_state_async: host logic through the test-only host harness.
Asynchronous stepping must give exactly the state of synchronous stepping, accessors synchronise first, and the
non-blocking diag reports a step that has completed."""
import numpy as np
import pytest

from regcm_b200 import capi, synth
from tests.helpers import prognostics


@pytest.mark.parametrize("name,kw", [("tiny", {}), ("tiny", {"idynamic": 1}), ("tinym", {})])
def test_emu_async_equals_sync(emu_lib, name, kw):
    case = synth.make_case(name, ibdyfrq=1, **kw)
    a = capi.DynCore(emu_lib, case.cfg); synth.load_case(a, case)
    b = capi.DynCore(emu_lib, case.cfg); synth.load_case(b, case)
    nsteps = int(round(3600.0 / case.cfg.dt)) + 3          # crosses a boundary-record change
    a.step(nsteps)
    snap = None
    for n in range(nsteps):
        b.step_async(1)
        d = b.diag()                                        # does not need a sync
        assert 0 <= d.lcount <= n + 1
        if n == 4:                                          # export in the middle of the batch
            f0 = prognostics(b)[0]
            shp = b.field_shape(f0)
            snap = np.zeros(shp, order="F")
            b.get_state_async(f0, capi.LEVEL_A, snap, (1, 1, 1), shp)
    b.sync()
    b.get_state_wait()
    assert b.diag().lcount == a.diag().lcount == nsteps
    for f in prognostics(a):
        for lv in ((capi.LEVEL_A,) if case.cfg.idynamic == 3 else (capi.LEVEL_A, capi.LEVEL_B)):
            np.testing.assert_array_equal(a.get_state(f, lv), b.get_state(f, lv), err_msg=capi.FIELD_NAMES[f])
    # the export taken after 5 steps equals a synchronous run stopped there
    c = capi.DynCore(emu_lib, case.cfg); synth.load_case(c, case)
    c.step(5)
    np.testing.assert_array_equal(snap, c.get_state(prognostics(c)[0], capi.LEVEL_A))
    da, db = a.diag(), b.diag()
    assert da.max_cfl == db.max_cfl and da.neg_q_points == db.neg_q_points


def test_emu_accessors_synchronise(emu_lib):
    case = synth.make_case("tiny")
    a = capi.DynCore(emu_lib, case.cfg); synth.load_case(a, case)
    b = capi.DynCore(emu_lib, case.cfg); synth.load_case(b, case)
    a.step(3)
    b.step_async(2); b.step_async(1)
    np.testing.assert_array_equal(a.get_state(capi.F_T), b.get_state(capi.F_T))   # get_state finishes the batch first
    assert b.diag().lcount == 3
