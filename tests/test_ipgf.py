"""ipgf = 1 (idynamic = 1): the pressure gradient force computed from the deviation of the virtual
temperature from the reference profile t00pg (p/p00pg)^pgfaa1 (Main/mod_tendency.F90:1893-1964,2039-2068;
Share/mod_constants.F90:359-362).  The reference's first loop runs `do k = - 1 , kz` (:1894): levels -1 and 0 lie outside the
arrays it writes; the levels 1..kz are what is built and restated.

CPU suite: the CUDA sources through the host harness against the oracle (1e-11: the two sides' split-explicit tables come
from independent eigen-solvers, tests/test_hydro.py), and an analytic pin of the oracle: an atmosphere at rest that follows
the reference profile over a 1500 m mountain stays at rest with ipgf = 1 up to the curvature of the profile across one
grid cell, while ipgf = 0 leaves the full truncation error of the two large cancelling terms."""
import numpy as np
import pytest

from oracle import pyoracle
from regcm_b200 import capi, synth
from tests.helpers import make_pair, rel_errors, work_errors

T00PG, P00PG, ALAM, EGRAV = 287.0, 101.325, 6.5e-3, 9.80665
RGAS = 6.02214129e23 * 1.3806504e-23 / 28.96454 * 1000.0     # Share/mod_constants.F90:92-134
PGFAA1 = ALAM * RGAS / EGRAV


@pytest.mark.parametrize("kw", [{}, {"iboudy": 1}, {"iboudy": 4}, {"nsplit": 3}, {"idiffu": 2}],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_emu_ipgf1(emu_lib, kw):
    case, core, ref = make_pair(emu_lib, "tiny", oracle_checked=True, idynamic=1, ipgf=1, **kw)
    core.step(6); ref.step(6)
    assert max(e[0] for e in rel_errors(core, ref).values()) <= 1e-11
    assert max(work_errors(core, ref).values()) <= 1e-10          # phi, omega, pten, qdot, xkc


def test_ipgf_switch_is_live_and_checked(emu_lib):
    out = []
    for ipgf in (0, 1):
        case = synth.make_case("tiny", idynamic=1, ipgf=ipgf)
        o = pyoracle.make_oracle(case.cfg); synth.load_case(o, case)
        o.step(4)
        out.append(o.get_state(capi.F_U))
    assert np.abs(out[0] - out[1]).max() > 1e-3 * np.abs(out[0]).max()
    cfg = synth.make_case("tiny", idynamic=1, records=()).cfg
    cfg.ipgf = 2
    with pytest.raises(capi.RcmError):
        capi.DynCore(emu_lib, cfg)


def _reference_profile_at_rest(ipgf):
    """One step of the oracle from rest on the reference profile (handed in as the boundary records, so that init and
    spinit see it); returns max(|u|, |v|) [m/s] in the interior afterwards and the height of the mountain."""
    case = synth.make_case("small", idynamic=1, ipgf=ipgf, iboudy=0, noise=False, unit_map=True)
    cfg = case.cfg
    kz, ptop = cfg.kz, cfg.ptop
    sigma = np.asarray(case.statics[capi.S_SIGMA]).ravel()
    hs = 0.5 * (sigma[1:] + sigma[:-1])
    ht = np.asarray(case.statics[capi.S_HT])[:cfg.jx - 1, :cfg.iy - 1] * EGRAV          # geopotential of the terrain
    # surface pressure in hydrostatic balance with the profile: ht + (R T00 / a) (ps/p00)^a = R T00 / a
    c = RGAS * T00PG / PGFAA1
    ps = P00PG * ((c - ht) / c) ** (1.0 / PGFAA1)                                    # [cb]
    pstar = ps - ptop
    p = hs[None, None, :] * pstar[:, :, None] + ptop
    t = T00PG * (p / P00PG) ** PGFAA1
    for slot in case.bdy:
        rec = case.bdy[slot]
        rec[capi.F_U] = np.zeros_like(rec[capi.F_U]); rec[capi.F_V] = np.zeros_like(rec[capi.F_V])
        rec[capi.F_T] = np.asfortranarray(t)
        rec[capi.F_QV] = np.zeros_like(rec[capi.F_QV])
        rec[capi.F_PS] = np.asfortranarray(ps * 10.0).reshape(rec[capi.F_PS].shape, order="F")      # ICBC ps in hPa
    o = pyoracle.make_oracle(cfg); synth.load_case(o, case)
    assert np.abs(o.get_state(capi.F_PS)[:, :, 0] - pstar).max() < 1e-10
    o.step(1)
    u, v = o.get_state(capi.F_U), o.get_state(capi.F_V)
    inner = (slice(3, -3), slice(3, -3))
    return max(np.abs(u[inner]).max(), np.abs(v[inner]).max()) / pstar.mean(), float(ht.max() / EGRAV)


def test_reference_profile_at_rest_stays_at_rest_with_ipgf1():
    u1, hmax = _reference_profile_at_rest(1)
    u0, _ = _reference_profile_at_rest(0)
    assert hmax > 1000.0                       # there is a mountain
    assert u0 > 1.0e-3                         # ipgf = 0: the truncation error of the two cancelling terms drives a wind
    assert u1 < 0.02 * u0                      # ipgf = 1: what is left is the curvature of the profile across a cell


@pytest.mark.gpu
@pytest.mark.parametrize("kw", [{}, {"iboudy": 1}, {"nsplit": 3}], ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_gpu_ipgf1_vs_oracle(cuda_lib, kw):
    case, core, ref = make_pair(cuda_lib, "small", idynamic=1, ipgf=1, **kw)
    core.step(8); ref.step(8)
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) <= 1e-11, errs
    core.close()
