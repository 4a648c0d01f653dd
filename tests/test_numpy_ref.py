"""Independent pins of the oracle: NumPy restatements written from the
Fortran text (tests/numpy_ref.py) against the C++ oracle on the same inputs, <= 1e-12 relative."""
import numpy as np
import pytest

from oracle import pyoracle
from regcm_b200 import capi, synth
from tests import numpy_ref as nr

TOL = 1e-12


def _rel(a, b):
    return float(np.abs(a - b).max()) / max(float(np.abs(b).max()), 1e-300)


@pytest.mark.parametrize("kw", [{}, {"upstream_mode": 0}], ids=["upstream", "centred"])
def test_hadvuv_vadvuv_numpy_vs_oracle(kw):
    """Main/mod_advection.F90:124-303 on the state after 3 steps (so that w, p' and the divergence are non-trivial)."""
    case = synth.make_case("small", **kw)
    o = pyoracle.make_oracle(case.cfg)
    synth.load_case(o, case)
    o.step(3)
    nr.call(o, "advect_uv")     # start_advect; hadvuv; vadvuv on zeroed udyn, vdyn, current decoupled state
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    idx = (2, jx - 1, 2, iy - 1)   # jdi1, jdi2, idi1, idi2 (Main/mod_atm_interface.F90:203-353, single rank)
    t = {n: nr.tap(o, n) for n in ("ax_umc", "ax_vmc", "ax_ud", "ax_vd", "ax_uc", "ax_vc", "cr", "dmsf", "qdot", "twt", "xds",
                                   "udyn", "vdyn")}
    ut, vt = nr.hadvuv_vadvuv(t["ax_umc"], t["ax_vmc"], t["ax_ud"], t["ax_vd"], t["ax_uc"], t["ax_vc"], t["cr"], t["dmsf"],
                              t["qdot"], t["twt"], t["xds"], nr.scalar(o, "ul"), idx, kz, upstream=case.cfg.upstream_mode != 0)
    want_u = t["udyn"].s((idx[0], idx[1]), (idx[2], idx[3]), (1, kz))
    want_v = t["vdyn"].s((idx[0], idx[1]), (idx[2], idx[3]), (1, kz))
    assert np.abs(want_u).max() > 0 and np.abs(want_v).max() > 0
    assert _rel(ut, want_u) <= TOL and _rel(vt, want_v) <= TOL
