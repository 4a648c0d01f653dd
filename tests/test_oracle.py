"""CPU tests of the oracle restatement (test infrastructure).  The reference
holds no golden vectors for the dyn-core ("parity unpinned"); these tests pin
the restatement through closed forms, physical invariants and committed
regression fixtures."""
import ctypes as C
import glob
import os

import numpy as np
import pytest

from oracle import pyoracle
from regcm_b200 import capi, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _aux(o, name, shape):
    fn = o.lib.lib.orc_get_aux
    fn.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double), C.c_int]
    n = int(np.prod(shape))
    buf = np.zeros(n, dtype=np.float64)
    got = fn(o.h, name.encode(), buf.ctypes.data_as(C.POINTER(C.c_double)), n)
    assert got == n, (name, got, n)
    return buf.reshape(shape, order="F")


def test_reference_atmosphere_closed_form(built):
    """nhbase (Share/mod_nhinterp.F90:88-104) against an independent numpy
    evaluation of the same closed form."""
    case = synth.make_case("small")
    o = pyoracle.make_oracle(case.cfg, checked=True)
    synth.load_case(o, case)
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    sig = case.statics[capi.S_SIGMA]
    hs = 0.5 * (sig[1:] + sig[:-1])
    ht = case.statics[capi.S_HT][: jx - 1, : iy - 1]
    rgas = 6.02214129e23 * 1.3806504e-23 / 28.96454 * 1000.0
    g = 9.80665
    p0, tlp, ts0, ptop = 101325.0, 47.70, 288.15, case.cfg.ptop * 1000.0
    b = ts0 / tlp
    ps0 = p0 * np.exp(-b + np.sqrt(b * b - 4.0 * (0.5 * g / rgas * ht / tlp))) - ptop
    pr0 = ps0[:, :, None] * hs + ptop
    t0 = np.maximum(ts0 + tlp * np.log(pr0 / p0), 216.65)
    np.testing.assert_allclose(_aux(o, "a0_ps", (jx - 1, iy - 1)), ps0, rtol=1e-13)
    np.testing.assert_allclose(_aux(o, "a0_pr", (jx - 1, iy - 1, kz)), pr0, rtol=1e-13)
    np.testing.assert_allclose(_aux(o, "a0_t", (jx - 1, iy - 1, kz)), t0, rtol=1e-13)
    rho = _aux(o, "a0_rho", (jx - 1, iy - 1, kz))
    np.testing.assert_allclose(rho, pr0 / rgas / t0, rtol=1e-13)
    # heights decrease with k and vanish at the surface full level
    zf = _aux(o, "a0_zf", (jx - 1, iy - 1, kz + 1))
    assert np.all(np.diff(zf, axis=2) < 0) and np.allclose(zf[:, :, -1], 0.0, atol=1e-9)


def test_relaxation_zone_map(built):
    """setup_boundaries (Main/mod_atm_interface.F90:383-542): the four zones
    tile the frame of width nspg-2 exactly once, with the diagonal corner split."""
    case = synth.make_case("small")
    o = pyoracle.make_oracle(case.cfg, checked=True)
    jx, iy, nsp = case.cfg.jx, case.cfg.iy, case.cfg.nspgx
    fn = o.lib.lib.orc_get_zone
    fn.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32)]
    for dot in (0, 1):
        z = np.zeros(jx * iy, dtype=np.int32)
        assert fn(o.h, dot, z.ctypes.data_as(C.POINTER(C.c_int32))) == 0
        z = z.reshape((jx, iy), order="F")
        flags, ib = z & 0xF, z >> 8
        nj, ni = (jx, iy) if dot else (jx - 1, iy - 1)
        # exactly one zone flag wherever any is set
        assert np.all(np.isin(flags, (0, 1, 2, 4, 8)))
        J, I = np.meshgrid(np.arange(1, jx + 1), np.arange(1, iy + 1), indexing="ij")
        dist = np.minimum(np.minimum(J, nj + 1 - J), np.minimum(I, ni + 1 - I))  # 1 = boundary row
        frame = (dist >= 2) & (dist <= nsp - 1) & (J <= nj) & (I <= ni)
        assert np.array_equal(flags > 0, frame)
        assert np.array_equal(ib[frame], dist[frame])


def test_rest_state_stays_at_rest(built):
    """Behavioural pin in the spirit of moloch_static_test1
    (Main/mod_bdycod.F90:5157-5199): a horizontally uniform, hydrostatically
    balanced, dry atmosphere at rest over flat terrain stays at rest."""
    case = synth.make_case("tiny", flat=True, unit_map=True, noise=False)
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    o0 = pyoracle.make_oracle(case.cfg)
    synth.load_case(o0, case)
    t0 = _aux(o0, "a0_t", (jx - 1, iy - 1, kz))
    for slot in case.bdy:
        b = case.bdy[slot]
        b[capi.F_U] = np.zeros_like(b[capi.F_U]); b[capi.F_V] = np.zeros_like(b[capi.F_V])
        b[capi.F_T] = np.asfortranarray(t0.copy())
        b[capi.F_QV] = np.full_like(b[capi.F_QV], 1.0e-8)
    o = pyoracle.make_oracle(case.cfg, checked=True)
    synth.load_case(o, case)
    o.step(20)
    ps = 1.0e2  # p* ~ 96 cb
    # residual: virtual-temperature buoyancy of the 1e-8 kg/kg moisture floor
    assert np.abs(o.get_state(capi.F_U)).max() / ps < 1e-4   # m/s
    assert np.abs(o.get_state(capi.F_V)).max() / ps < 1e-4
    assert np.abs(o.get_state(capi.F_W)).max() / ps < 1e-5
    assert np.abs(o.get_state(capi.F_PP)).max() / ps < 1e-2  # Pa


def test_checked_build_full_run(built):
    """Bounds-checked oracle runs through a boundary-record change."""
    case = synth.make_case("tiny", ibdyfrq=1)
    o = pyoracle.make_oracle(case.cfg, checked=True)
    synth.load_case(o, case)
    o.step(31)  # 1 h = 30 steps of 120 s
    d = o.diag()
    assert d.lcount == 31 and d.xbctime == 240.0 and 0 < d.max_cfl < 1 and d.neg_q_points == 0
    with pytest.raises(capi.RcmError) as ei:
        o.step(30)  # next record not staged
    assert ei.value.code == -6


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "nh_*.npz")) +
                                        glob.glob(os.path.join(ROOT, "tests", "golden", "hy_*.npz")) +
                                        glob.glob(os.path.join(ROOT, "tests", "golden", "mo_*.npz"))))
def test_oracle_matches_golden(built, path):
    """Regression pin: the oracle reproduces the committed fixtures exactly."""
    from tests.golden_util import load_golden
    case, nsteps, fields = load_golden(path)
    o = pyoracle.make_oracle(case.cfg)
    synth.load_case(o, case)
    o.step(nsteps)
    for key, ref in fields.items():
        if isinstance(key, tuple):
            np.testing.assert_array_equal(o.get_state(*key), ref)
