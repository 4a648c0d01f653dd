"""Tracer transport on the GPU through the C ABI against the oracle: same cases as the host-harness
tests of tests/test_tracers.py.  Tolerance 1e-11 relative L-inf per tracer and time level (the GPU's libm pow/exp
differ from the host's in the last bits, as for the other prognostics)."""
import numpy as np
import pytest

from regcm_b200 import capi
from tests.helpers import make_pair
from tests.test_tracers import NTR, _spiky, _tracers

pytestmark = pytest.mark.gpu
TOL = 1e-11


@pytest.mark.parametrize("kw", [{}, {"iboudy": 1}, {"iboudy": 4}, {"upstream_mode": 0}, {"idynamic": 1}],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_gpu_tracers_vs_oracle(cuda_lib, kw):
    case, core, ref = make_pair(cuda_lib, "small", ntr=NTR, **kw)
    core.step(8); ref.step(8)
    for n, ((a, b), (ra, rb)) in enumerate(zip(_tracers(core), _tracers(ref))):
        assert ra.max() > 1e-6
        assert np.abs(a - ra).max() <= TOL * ra.max(), f"chi{n} level a"
        assert np.abs(b - rb).max() <= TOL * rb.max(), f"chi{n} level b"
    core.close()


def test_gpu_tracers_record_change(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "tiny", ntr=2, ibdyfrq=1)
    core.step(34); ref.step(34)
    for (a, b), (ra, rb) in zip(_tracers(core, 2), _tracers(ref, 2)):
        assert np.abs(a - ra).max() <= 1e-10 * ra.max() and np.abs(b - rb).max() <= 1e-10 * rb.max()
    core.close()


@pytest.mark.parametrize("idynamic", [2, 1])
def test_gpu_tracer_negative_fix(cuda_lib, idynamic):
    """The device-side resolution of the sequential 3x3 sweep (k_tr_repair), adjacent negatives included."""
    case, core, ref = make_pair(cuda_lib, "tiny", ntr=2, idynamic=idynamic)
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
        for (a, b), (ra, rb) in zip(_tracers(core, 2), _tracers(ref, 2)):
            assert np.abs(a - ra).max() <= TOL * ra.max() and np.abs(b - rb).max() <= TOL * rb.max()
    assert total > 50
    core.close()
