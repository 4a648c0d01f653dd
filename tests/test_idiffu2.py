"""idiffu = 2: the nine-point "LeVeque" horizontal diffusion operator, applied on every interior
point with no second-order boundary ring (Main/mod_diffusion.F90:63-65,386-411,592-601,726-735,881-894; halo width
idif = 1, Main/mod_params.F90:1968-1971).

CPU suite: the CUDA sources through the host harness (tests/emu) against the oracle, bit for bit for idynamic = 2 and
to the eigen-solver tolerance for idynamic = 1; pins of the oracle's operator against an independent NumPy statement
of the nine-point stencil (tests/test_numpy_ref.py::test_tracer_step_numpy_vs_oracle[idiffu2]) and of the stencil
against the analytic Laplacian it approximates."""
import numpy as np
import pytest

from oracle import pyoracle
from regcm_b200 import capi, synth
from tests.helpers import make_pair, rel_errors

NTR = 2


@pytest.mark.parametrize("kw", [{}, {"iboudy": 1}, {"iboudy": 0}, {"iboudy": 4}, {"ntr": NTR}, {"idynamic": 1},
                                {"idynamic": 1, "iboudy": 1, "ntr": NTR}],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_emu_idiffu2(emu_lib, kw):
    case, core, ref = make_pair(emu_lib, "tiny", oracle_checked=True, idiffu=2, **kw)
    for c in (core, ref):
        c.step(6)
    tol = 1e-11 if kw.get("idynamic") == 1 else 0.0
    for name, (linf, _) in rel_errors(core, ref).items():
        assert linf <= tol, (name, linf)
    for n in range(kw.get("ntr", 0)):
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            a, r = core.get_state(capi.F_CHI + n, lv), ref.get_state(capi.F_CHI + n, lv)
            assert np.abs(a - r).max() <= tol * np.abs(r).max()


def test_idiffu2_differs_from_idiffu1():
    """The switch is live in the oracle: both operators damp, but not by the same amount."""
    out = []
    for idiffu in (1, 2):
        case = synth.make_case("tiny", idiffu=idiffu)
        o = pyoracle.make_oracle(case.cfg)
        synth.load_case(o, case)
        o.step(4)
        out.append([o.get_state(f) for f in range(7)])
    for a, b in zip(*out):
        assert np.isfinite(a).all() and np.isfinite(b).all()
    assert any(np.abs(a - b).max() > 1e-9 * np.abs(a).max() for a, b in zip(*out))


def _nine_point(f):
    """o4_c1*(sides) + o4_c2*(corners) + o4_c3*centre on the interior of a (j, i) array."""
    c1, c2, c3 = 4.0 / 6.0, 1.0 / 6.0, -20.0 / 6.0
    s = f[2:, 1:-1] + f[:-2, 1:-1] + f[1:-1, 2:] + f[1:-1, :-2]
    c = f[2:, 2:] + f[:-2, :-2] + f[:-2, 2:] + f[2:, :-2]
    return c1 * s + c2 * c + c3 * f[1:-1, 1:-1]


def test_nine_point_operator_is_the_laplacian_to_fourth_order():
    """The stencil of Main/mod_diffusion.F90:63-65 is h^2 * (Laplacian + h^2/12 * biharmonic + O(h^4)): on
    sin(ax) sin(by) it returns -(a^2+b^2) h^2 f times (1 - (a^2+b^2) h^2 / 12) up to O(h^6)."""
    errs = []
    for n in (32, 64, 128):
        h = 1.0 / n
        x = np.arange(n + 1) * h
        X, Y = np.meshgrid(x, x, indexing="ij")
        a, b = 2 * np.pi, 4 * np.pi
        f = np.sin(a * X) * np.cos(b * Y)
        lam = a * a + b * b
        want = -lam * h * h * (1.0 - lam * h * h / 12.0) * f[1:-1, 1:-1]
        errs.append(np.abs(_nine_point(f) - want).max() / (lam * h * h))
    assert errs[1] < errs[0] / 12 and errs[2] < errs[1] / 12      # O(h^4) relative to the leading term
