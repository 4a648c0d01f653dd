"""Oracle comparisons at the sizes the benchmark runs ("the headline configuration is never
compared with the oracle"): size-dependent paths — the tile-edge clamping of the TMA-staged kernels, index
products near 42 M elements per field, the register-capped builds at full occupancy — are exercised here
against the CPU restatement, through the C ABI.

  C4q  500 x 500 x 41 non-hydrostatic, 2 steps      (oracle ~10 s/step, 12 GB)
  C5   500 x 500 x 30 MOLOCH, 1 step                (BASELINE configs[4] at full size)
  C4   1000 x 1000 x 41 non-hydrostatic, 1 step     (BASELINE configs[3], the bench default; needs ~60 GB of host
                                                     memory for the oracle, skipped on smaller hosts)
Tolerance: 1e-11 relative L-inf per prognostic (normalised by the field maximum), both time levels."""
import numpy as np
import pytest

from regcm_b200 import capi
from tests.helpers import make_pair, rel_errors

pytestmark = pytest.mark.gpu
TOL = 1e-11


def _host_gb():
    try:
        import psutil
        return psutil.virtual_memory().available / 2 ** 30
    except Exception:
        return 0.0


def _worst(errs):
    return max(e[0] for e in errs.values())


def test_c4_quarter_vs_oracle(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "C4q")
    core.step(2); ref.step(2)
    errs = rel_errors(core, ref)
    assert _worst(errs) <= TOL, errs
    dc, dr = core.diag(), ref.diag()
    assert dc.istep == dr.istep and dc.neg_q_points == dr.neg_q_points and abs(dc.max_cfl - dr.max_cfl) <= 1e-9 * dr.max_cfl


def test_c5_full_size_vs_oracle(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "C5")
    core.step(1); ref.step(1)
    errs = rel_errors(core, ref)
    # w of the MOLOCH core is a small residual of large terms (|w| ~ cm/s on this near-balanced state): 1e-10 for it
    assert all(e[0] <= (1e-10 if k.startswith("w") else TOL) for k, e in errs.items()), errs
    dc, dr = core.diag(), ref.diag()
    assert abs(dc.ps_max - dr.ps_max) <= 1e-9 * dr.ps_max and abs(dc.ps_min - dr.ps_min) <= 1e-9 * dr.ps_min


@pytest.mark.skipif(_host_gb() < 100.0, reason="the oracle needs ~60 GB of host memory at 1000 x 1000 x 41")
def test_c4_headline_config_vs_oracle(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "C4")
    core.step(1); ref.step(1)
    errs = rel_errors(core, ref)
    assert _worst(errs) <= TOL, errs
    assert core.diag().istep == ref.diag().istep
