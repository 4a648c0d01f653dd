"""Negative-moisture repair on the device (k_negq.cu; Main/mod_tendency.F90:381-393) against
the oracle, on inputs whose forecast DOES go negative: moisture with sharp fronts (tests/helpers.sharp_moisture).
Both sides must report the same, non-zero, number of repaired points; states agree to 1e-11."""
import numpy as np
import pytest

from regcm_b200 import capi
from tests.helpers import make_pair, rel_errors, sharp_moisture

pytestmark = pytest.mark.gpu


def _load_sharp(case, cores):
    qv, qc = sharp_moisture(case)
    for c in cores:
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            c.set_state(capi.F_QV, lv, qv)
            c.set_state(capi.F_QC, lv, qc)


@pytest.mark.parametrize("name,idyn", [("small", 2), ("C2", 2), ("small", 1), ("C1", 1), ("C3", 2)])
def test_negative_moisture_repair_vs_oracle(cuda_lib, name, idyn):
    case, core, ref = make_pair(cuda_lib, name, idynamic=idyn)
    _load_sharp(case, (core, ref))
    total = 0
    for step in range(3):
        core.step(1); ref.step(1)
        dc, dr = core.diag(), ref.diag()
        assert dc.neg_q_points == dr.neg_q_points, (step, dc.neg_q_points, dr.neg_q_points)
        total += dr.neg_q_points
        errs = rel_errors(core, ref)
        assert max(e[0] for e in errs.values()) <= 1e-11, (step, errs)
    assert total > 50
    assert core.get_state(capi.F_QV).min() >= 0.0 and core.get_state(capi.F_QC).min() >= 0.0


def test_negative_moisture_multi_step_call(cuda_lib):
    """rcm_dyn_step(n > 1): the second pass of the slow moisture kernel runs on every step of the call."""
    case, core, ref = make_pair(cuda_lib, "small")
    _load_sharp(case, (core, ref))
    core.step(5); ref.step(5)
    assert core.diag().neg_q_points == ref.diag().neg_q_points > 0
    assert max(e[0] for e in rel_errors(core, ref).values()) <= 1e-11


def test_smooth_moisture_takes_no_repair(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "small")
    core.step(4); ref.step(4)
    assert core.diag().neg_q_points == ref.diag().neg_q_points == 0
