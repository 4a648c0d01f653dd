"""idiffu = 3: the flux-limited 6th-order diffusion of Main/mod_diffusion.F90:412-517,602-652,736-790,
895-960 AS THIS SNAPSHOT OF THE REFERENCE HAS IT: every loop over j runs `do j = jdi2 , jdi2` / `do j = jci2 , jci2`
(:421,611,745,902), so the operator acts on one column of the domain, with the stencil indices clamped to the grid and
constant coefficients xkc = diff_6th_coef * p* (:154,174-183).  Restated and built as written (one rank: the stencil is three
cells wide).

CPU suite: host harness against the oracle, bit for bit for idynamic = 2; pins of the oracle: the operator touches column
jci2 only, and there it equals a NumPy statement of the limited 6th-order flux difference."""
import numpy as np
import pytest

from oracle import pyoracle
from regcm_b200 import capi, synth
from tests.helpers import make_pair, rel_errors


@pytest.mark.parametrize("kw", [{}, {"iboudy": 1}, {"iboudy": 4}, {"ntr": 2}, {"ibltyp": 2}, {"idynamic": 1}, {"idynamic": 1, "ntr": 1}],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_emu_idiffu3(emu_lib, kw):
    case, core, ref = make_pair(emu_lib, "tiny", oracle_checked=True, idiffu=3, **kw)
    if kw.get("ibltyp") == 2:
        from tests.test_tke import load_uw, uw_inputs
        for c in (core, ref):
            load_uw(c, case, *uw_inputs(case))
    for c in (core, ref):
        c.step(6)
    tol = 1e-11 if kw.get("idynamic") == 1 else 0.0
    for name, (linf, _) in rel_errors(core, ref).items():
        assert linf <= tol, (name, linf)
    extra = [capi.F_CHI + n for n in range(kw.get("ntr", 0))] + ([capi.F_TKE] if kw.get("ibltyp") == 2 else [])
    for fld in extra:
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            a, r = core.get_state(fld, lv), ref.get_state(fld, lv)
            assert np.abs(a - r).max() <= tol * np.abs(r).max(), (fld, lv)


def test_create_rejects_idiffu3_under_mpi(emu_lib):
    cfg = synth.make_case("tiny", records=()).cfg
    cfg.idiffu, cfg.nproc, cfg.cpus_j, cfg.cpus_i = 3, 4, 2, 2
    with pytest.raises(capi.RcmError):
        capi.DynCore(emu_lib, cfg)
    cfg = synth.make_case("tiny", records=()).cfg
    cfg.idiffu = 4
    with pytest.raises(capi.RcmError):
        capi.DynCore(emu_lib, cfg)


def _limited6(F, G, j, i, jmax, imax):
    """((x1 - x0) + (y1 - y0)) of :618-650 at the 1-based point (j, i) of the 2-D arrays F (diffused) and G (sign test)."""
    c = lambda v, hi: min(max(v, 1), hi) - 1      # noqa: E731  clamped, 0-based
    J = [c(j + d, jmax) for d in range(-3, 4)]    # j-3 .. j+3
    I = [c(i + d, imax) for d in range(-3, 4)]
    j0, i0 = j - 1, i - 1

    def flux(a, b, c2, d, e, f2, ga, gb):
        x = 10.0 * (a - b) + (-5.0) * (c2 - d) + 1.0 * (e - f2)
        return 0.0 if x * (ga - gb) <= 0.0 else x
    x0 = flux(F[j0, i0], F[J[2], i0], F[J[4], i0], F[J[1], i0], F[J[5], i0], F[J[0], i0], G[j0, i0], G[J[2], i0])
    x1 = flux(F[J[4], i0], F[j0, i0], F[J[5], i0], F[J[2], i0], F[J[6], i0], F[J[1], i0], G[J[4], i0], G[j0, i0])
    y0 = flux(F[j0, i0], F[j0, I[2]], F[j0, I[4]], F[j0, I[1]], F[j0, I[5]], F[j0, I[0]], G[j0, i0], G[j0, I[2]])
    y1 = flux(F[j0, I[4]], F[j0, i0], F[j0, I[5]], F[j0, I[2]], F[j0, I[6]], F[j0, I[1]], G[j0, I[4]], G[j0, i0])
    return (x1 - x0) + (y1 - y0)


def test_oracle_idiffu3_acts_on_one_column_and_equals_numpy():
    """One step with idiffu = 3 against one step without any horizontal diffusion (idiffu = 1 with ckh = adyndif = 0), on
    water vapour — the one prognostic nothing but its own tendency and time filter touches within a step (T is also moved by
    the acoustic steps / the split-explicit correction, which carry a change of the winds at j = jdi2 to the neighbouring
    columns).  The filtered level a differs on the cross column j = jci2 = jx - 2 only, and there by
    dt * (1 + (beta - 1) * gnu1) * xkc * limited6(qvb3d)   (RAW filter: Main/mod_timefilter.F90:231-278)."""
    from tests import numpy_ref as nr
    outs = {}
    for tag, kw in (("d3", {"idiffu": 3}), ("none", {"idiffu": 1, "ckh": 0.0, "adyndif": 0.0})):
        case = synth.make_case("tiny", iboudy=0, **kw)
        o = pyoracle.make_oracle(case.cfg); synth.load_case(o, case)
        qb = o.get_state(capi.F_QV, capi.LEVEL_B).copy()
        o.step(1)
        outs[tag] = (o.get_state(capi.F_QV, capi.LEVEL_A), qb, nr.tap(o, "psb").a.copy(), nr.tap(o, "msfd").a.copy(), case.cfg)
    qa, qb, psb, msfd, cfg = outs["d3"]
    q0 = outs["none"][0]
    jx, iy, kz, dt = cfg.jx, cfg.iy, cfg.kz, cfg.dt
    d = (qa - q0) / (dt * (1.0 + (0.53 - 1.0) * cfg.gnu1))
    jc = jx - 2                                   # jci2, 1-based
    mask = np.ones(d.shape[:2], bool); mask[jc - 1, 1:-1] = False
    assert np.abs(d[mask]).max() == 0.0           # nothing anywhere else
    assert np.abs(d[jc - 1, 1:-1]).max() > 0.0
    coef = 0.12 * 0.015625 / (2.0 * dt)
    worst = 0.0
    for k in range(kz):
        F = np.maximum(qb[:, :, k] / psb[:jx - 1, :iy - 1], 1.0e-8)         # qxb3d = max(atm2%qx * rpsb, minqq)
        G = F / msfd[:jx - 1, :iy - 1]
        for i in range(2, iy - 1):                                          # ici1 .. ici2
            want = coef * psb[jc - 1, i - 1] * _limited6(F, G, jc, i, jx - 1, iy - 1)
            worst = max(worst, abs(d[jc - 1, i - 1, k] - want))
    assert worst <= 1e-9 * np.abs(d[jc - 1]).max(), worst
