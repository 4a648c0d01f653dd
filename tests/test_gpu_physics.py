"""Physics coupling boundary on the GPU (see tests/test_physics_coupling.py for the semantics):
slices of mkslice and a step with parametrization tendencies against the oracle, through the C ABI."""
import numpy as np
import pytest

from regcm_b200 import capi
from tests.helpers import make_pair, prognostics, rel_errors
from tests.test_physics_coupling import _phys_tendencies

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,idyn", [("small", 2), ("C2", 2), ("small", 1), ("C1", 1)])
def test_slices_vs_oracle(cuda_lib, name, idyn):
    case, core, ref = make_pair(cuda_lib, name, idynamic=idyn)
    core.step(2); ref.step(2)
    core.step_pre_physics(); ref.step_pre_physics()
    for sid in range(19):
        got, want = core.get_slice(sid), ref.get_slice(sid)
        sc = max(float(np.abs(want).max()), 1e-300)
        err = float(np.abs(got - want).max()) / sc
        tol = 1e-10 if (idyn == 1 and sid in (capi.SL_WPX3D, capi.SL_WB3D)) else 1e-11
        assert err <= tol, (capi.SLICE_NAMES[sid], err)
    core.step_post_physics(); ref.step_post_physics()
    assert max(e[0] for e in rel_errors(core, ref).values()) <= 1e-11


@pytest.mark.parametrize("name,idyn", [("small", 2), ("C3", 2), ("small", 1)])
def test_step_with_physics_tendencies_vs_oracle(cuda_lib, name, idyn):
    case, core, ref = make_pair(cuda_lib, name, idynamic=idyn)
    _, plain, _ = make_pair(cuda_lib, name, idynamic=idyn)
    ten = _phys_tendencies(case, idyn == 1)
    for n in range(3):
        for c in (core, ref):
            c.step_pre_physics()
            for f, a in ten.items():
                c.set_phys_tend(f, a * (1.0 + 0.25 * n))
            c.step_post_physics()
        plain.step_pre_physics(); plain.step_post_physics()
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) <= 1e-11, errs
    moved = rel_errors(core, plain)
    assert min(moved[k][0] for k in ("ta", "ua", "va", "qva")) > 1e-8, moved
    # the split step without tendencies is the plain step, bit for bit
    _, again, _ = make_pair(cuda_lib, name, idynamic=idyn)
    again.step(3)
    for f in prognostics(plain):
        np.testing.assert_array_equal(plain.get_state(f), again.get_state(f), err_msg=capi.FIELD_NAMES[f])


# ---------------------------------------------------------------------------- MOLOCH core (idynamic = 3)
def test_moloch_slices_vs_oracle(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "smallm")
    core.step(2); ref.step(2)
    core.step_pre_physics(); ref.step_pre_physics()
    for sid in range(19):
        got, want = core.get_slice(sid), ref.get_slice(sid)
        sc = max(float(np.abs(want).max()), 1e-300)
        assert float(np.abs(got - want).max()) / sc <= 1e-11, capi.SLICE_NAMES[sid]
    core.step_post_physics(); ref.step_post_physics()
    assert max(e[0] for e in rel_errors(core, ref).values()) <= 1e-12


def test_moloch_step_with_physics_tendencies_vs_oracle(cuda_lib):
    from tests.test_physics_coupling import _mo_tendencies
    case, core, ref = make_pair(cuda_lib, "smallm")
    _, plain, _ = make_pair(cuda_lib, "smallm")
    ten = _mo_tendencies(case)
    for n in range(3):
        for c in (core, ref):
            c.step_pre_physics()
            for f, a in ten.items():
                c.set_phys_tend(f, a * (1.0 + 0.25 * n))
            c.step_post_physics()
        plain.step_pre_physics(); plain.step_post_physics()
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) <= 1e-12, errs
    moved = rel_errors(core, plain)
    assert min(moved[k][0] for k in ("ta", "ua", "va", "qva")) > 1e-8, moved
    _, again, _ = make_pair(cuda_lib, "smallm")
    again.step(3)
    for f in prognostics(plain):
        np.testing.assert_array_equal(plain.get_state(f), again.get_state(f), err_msg=capi.FIELD_NAMES[f])
