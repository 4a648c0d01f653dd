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


SOUND_IN = ("psb", "psdotb", "ac_u", "ac_v", "ac_pp", "ac_w", "ac_t", "a2_u", "a2_v", "a2_pp", "a2_w", "a2_t", "a2_qx",
            "a1_u", "a1_v", "a1_pp", "a1_w", "a1_t", "a1_pr", "a1_rho", "uten", "vten", "ppten", "wten",
            "a0_pr", "a0_rho", "a0_t", "a0_ps", "a0_rhof", "a0_dprddx", "a0_dprddy", "msfd", "msfx", "ht",
            "dsigma", "sigma", "twt", "dpsdxm", "dpsdym", "s_estore", "s_astore", "tmask")
SOUND_OUT = ("a1_u", "a2_u", "a1_v", "a2_v", "a1_pp", "a2_pp", "a1_w", "a2_w", "a1_t", "a2_t", "s_estore", "s_astore")


@pytest.mark.parametrize("kw", [{}, {"ifupr": 0}], ids=["ifupr1", "ifupr0"])
def test_sound_numpy_vs_oracle(kw):
    """The whole of `sound` (Main/mod_sound.F90:163-718: istep sub-steps of explicit u,v, the column-implicit w/p'
    solve with the 13x13 radiative upper BC, CFL, temperature correction, epilogue filters) on the inputs the
    oracle's 4th step hands it."""
    case = synth.make_case("small", **kw)
    o = pyoracle.make_oracle(case.cfg)
    synth.load_case(o, case)
    o.step(3)
    o.lib.lib.orc_break_before_sound(o.h, 1)
    o.step(1)                                   # tend() up to `call sound`
    o.lib.lib.orc_break_before_sound(o.h, 0)
    assert nr.scalar(o, "alarm_day") == 0.0     # the daily tmask update is not part of this restatement
    t = {n: nr.F(nr.tap(o, n).a.copy(), nr.tap(o, n).lb) for n in SOUND_IN}
    rgas = (6.02214129e23 * 1.3806504e-23 / 28.96454) * 1000.0       # Share/mod_constants.F90:92-134
    p = dict(jx=case.cfg.jx, iy=case.cfg.iy, kz=case.cfg.kz, dt=nr.scalar(o, "dt"), dx=nr.scalar(o, "dx"),
             xkd=nr.scalar(o, "xkd"), bet=nr.scalar(o, "bet"), xgamma=nr.scalar(o, "xgamma"), dtsmax=nr.scalar(o, "dtsmax"),
             lcount=nr.scalar(o, "lcount"), gnu1=case.cfg.gnu1, gnu2=case.cfg.gnu2, cpd=3.5 * rgas, ifupr=case.cfg.ifupr)
    istep, cfl = nr.sound(t, p)
    nr.call(o, "sound")
    d = o.diag()
    assert istep == d.istep and istep >= 4
    assert abs(cfl - d.max_cfl) <= 1e-12 * d.max_cfl      # diag reports the maximum over the sub-steps
    for n in SOUND_OUT:
        want = nr.tap(o, n).a
        assert np.abs(want).max() > 0 or n.startswith("s_"), n
        assert _rel(t[n].a, want) <= TOL, (n, _rel(t[n].a, want))


def test_wafone_numpy_vs_oracle():
    """One `wafone` call of the MOLOCH core (Main/mod_moloch.F90:1098-1361, generic projection branch) on theta_v with
    the velocities, metric factors and generalized vertical velocity `s` the oracle holds after 2 steps."""
    case = synth.make_case("smallm")
    o = pyoracle.make_oracle(case.cfg)
    synth.load_case(o, case)
    o.step(2)
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    t = {n: nr.tap(o, n) for n in ("m_tetav", "m_s", "m_fmz", "m_fmzf", "m_u", "m_v", "rmu", "rmv", "mx2")}
    assert np.abs(t["m_s"].a).max() > 0 and np.abs(t["m_u"].a).max() > 1.0
    pp = nr.F(t["m_tetav"].a.copy(), t["m_tetav"].lb)
    before = pp.a.copy()
    dta = case.cfg.dt / case.cfg.mo_nadv
    nr.wafone(pp, t["m_s"], t["m_fmz"], t["m_fmzf"], t["m_u"], t["m_v"], t["rmu"], t["rmv"], t["mx2"], dta,
              case.cfg.ds * 1000.0, nr.scalar(o, "mo_dz"), jx, iy, kz)
    nr.call(o, "wafone_tetav", dta)
    want = nr.tap(o, "m_tetav").a
    change = np.abs(want - before).max()
    assert change > 1e-3                         # the pass does something
    assert np.abs(pp.a - want).max() <= 1e-12 * np.abs(want).max()
    assert np.abs(pp.a - want).max() <= 1e-9 * change   # and agrees on the increment, not just on the 300 K background


@pytest.mark.parametrize("kw", [{}, {"upstream_mode": 0}, {"idiffu": 2}, {"ibltyp": 2, "iuwvadv": 1}, {"isladvec": 1}, {"isladvec": 1, "iqmsl": 0},
                                {"isladvec": 1, "idynamic": 1}], ids=["default", "centred", "idiffu2", "ind3", "sl", "sl-noclip", "sl-hydro"])
def test_tracer_step_numpy_vs_oracle(kw):
    """hadvtr + vadv4d(ind = 2) + diffu_x4d + forecast + RAW filter of one tracer, away from the lateral
    zones, on the flow of the oracle's 4th step.  Variants: centred advection, idiffu = 2, vadv4d ind = 3 (iuwvadv = 1),
    semi-Lagrangian horizontal advection (isladvec = 1, f4) with and without its limiter and in the hydrostatic core."""
    case = synth.make_case("small", ntr=1, **kw)
    o = pyoracle.make_oracle(case.cfg)
    synth.load_case(o, case)
    jx, iy, kz, nsp = case.cfg.jx, case.cfg.iy, case.cfg.kz, case.cfg.nspgx
    kpb = sigma = None
    if kw.get("iuwvadv"):   # vadv4d ind = 3: a PBL top that takes every value 0..kz (>= 4 reshapes the flux)
        k2 = np.random.default_rng(3).integers(0, kz + 1, size=(jx - 3, iy - 3)).astype(np.int32)
        o.set_kpbl(k2, lb=(2, 2))
        kpb, sigma = nr.F(k2, (2, 2)), nr.tap(o, "sigma")
    o.step(3)
    keep = lambda n: nr.F(nr.tap(o, n).a.copy(), nr.tap(o, n).lb)      # noqa: E731
    chia, chib = keep("chia0"), keep("chib0")
    psa0, psb0 = keep("psa"), keep("psb")     # idynamic = 1 filters p* at the end of the step; the tendencies saw these
    dt = nr.scalar(o, "dt")
    o.step(1)
    # arrays `tend` leaves behind: decoupled fluxes, sigma-dot and the scaled Smagorinsky coefficient of THIS step
    t = {n: nr.tap(o, n) for n in ("ax_umc", "ax_vmc", "ax_umd", "ax_vmd", "msfx", "msfd", "xmsf", "qdot", "twt", "xds", "xkc", "chic0", "chia0", "chib0")}
    t["psa"], t["rpsa"], t["rpsb"] = psa0, nr.F(1.0 / psa0.a, psa0.lb), nr.F(1.0 / psb0.a, psb0.lb)
    box = (nsp + 2, jx - 1 - nsp - 2, nsp + 2, iy - 1 - nsp - 2)
    sl = None
    if kw.get("isladvec"):
        sl = dict(umd=t["ax_umd"], vmd=t["ax_vmd"], msfx=t["msfx"], msfd=t["msfd"], dx=nr.scalar(o, "dx"), iqmsl=case.cfg.iqmsl)
    args = (chia, chib, t["ax_umc"], t["ax_vmc"], t["psa"], t["rpsa"], t["rpsb"], t["xmsf"], t["qdot"], t["twt"], t["xds"], t["xkc"],
            nr.scalar(o, "ul"), dt, case.cfg.gnu2, box, kz)
    fc, na, nb = nr.tracer_step(*args, upstream=case.cfg.upstream_mode != 0, idiffu=case.cfg.idiffu, kpb=kpb, sigma=sigma, sl=sl)
    if kpb is not None or sl is not None:     # and the option is live: the default scheme gives a different forecast
        fc2, _, _ = nr.tracer_step(*args)
        assert np.abs(fc2 - fc).max() > 1e-6 * np.abs(fc).max()
    box3 = ((box[0], box[1]), (box[2], box[3]), (1, kz))
    want_c, want_a, want_b = t["chic0"].s(*box3), t["chia0"].s(*box3), t["chib0"].s(*box3)
    assert fc.min() > 0.0 and np.abs(want_c - chib.s(*box3)).max() > 1e-3 * np.abs(want_c).max()   # the step did move the tracer
    sc = np.abs(want_c).max()
    assert np.abs(fc - want_c).max() <= 1e-12 * sc
    assert np.abs(na - want_a).max() <= 1e-12 * sc and np.abs(nb - want_b).max() <= 1e-12 * sc


@pytest.mark.parametrize("kw", [{}, {"upstream_mode": 0}, {"idiffu": 2}, {"idynamic": 1}], ids=["default", "centred", "idiffu2", "hydro"])
def test_tke_step_numpy_vs_oracle(kw):
    """hadv3d(ind = 1) + vadv3d(kz + 1) + diffu_x3df(nuk) + floored forecast + RA filter of the UW-PBL TKE,
    away from the lateral boundary, on the flow of the oracle's 4th step, with a physics source."""
    from tests.test_tke import load_uw, uw_inputs
    case = synth.make_case("small", ibltyp=2, **kw)
    o = pyoracle.make_oracle(case.cfg)
    synth.load_case(o, case)
    tke, src, kpbl, qc = uw_inputs(case)
    load_uw(o, case, tke, src, kpbl, qc)
    o.step(3)
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    ta, tb = nr.tap(o, "tke_a"), nr.tap(o, "tke_b")
    tkea, tkeb = nr.F(ta.a.copy(), ta.lb), nr.F(tb.a.copy(), tb.lb)
    dt = nr.scalar(o, "dt")
    psa0 = nr.F(nr.tap(o, "psa").a.copy(), nr.tap(o, "psa").lb)    # idynamic = 1: p* of level a is filtered at the end of the step
    o.step(1)
    t = {n: nr.tap(o, n) for n in ("ax_umc", "ax_vmc", "xmsf", "qdot", "twt", "dds", "xkcf", "tke_a", "tke_b", "tke_c")}
    t["psa"], t["rpsa"] = psa0, nr.F(1.0 / psa0.a, psa0.lb)        # rpsa = 1/psa (Main/mod_tendency.F90:1001-1003)
    box = (4, jx - 4, 4, iy - 4)
    fc, na, nb = nr.tke_step(tkea, tkeb, t["ax_umc"], t["ax_vmc"], t["psa"], t["rpsa"], t["xmsf"], t["qdot"], t["twt"], t["dds"], t["xkcf"],
                             nr.F(src, (1, 1, 1)), nr.scalar(o, "ul"), dt, case.cfg.gnu2, case.cfg.uw_nuk, case.cfg.uw_tkemin, box, kz,
                             upstream=case.cfg.upstream_mode != 0, idiffu=case.cfg.idiffu)
    sl = ((box[0], box[1]), (box[2], box[3]), (1, kz + 1))
    want_c, want_a, want_b = t["tke_c"].s(*sl), t["tke_a"].s(*sl), t["tke_b"].s(*sl)
    sc = np.abs(want_c).max()
    assert np.abs(want_c - tkeb.s(*sl)).max() > 1e-3 * sc            # the step did change the TKE
    assert np.abs(fc - want_c).max() <= 1e-12 * sc
    assert np.abs(na - want_a).max() <= 1e-12 * sc and np.abs(nb - want_b).max() <= 1e-12 * sc


@pytest.mark.parametrize("kw", [{}, {"upstream_mode": 0}, {"stability_enhance": 0}, {"t_extrema": 1.0e-3}], ids=["default", "centred", "nolimiter", "limiter-active"])
def test_advect_theta_numpy_vs_oracle(kw):
    """hadvt with the local-extrema limiter (Main/mod_advection.F90:311-394) + vadv3d (:743-753) on potential temperature,
    whole interior, on the flow of the oracle's 4th step.  `limiter-active` lowers t_extrema so that the limiter fires."""
    case = synth.make_case("small", **kw)
    o = pyoracle.make_oracle(case.cfg)
    synth.load_case(o, case)
    o.step(4)
    nr.call(o, "advect_theta")
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    idx = (2, jx - 2, 2, iy - 2)    # jci1, jci2, ici1, ici2
    t = {n: nr.tap(o, n) for n in ("th", "tha", "ax_umc", "ax_vmc", "psa", "xmsf", "qdot", "twt", "xds", "thten")}
    got = nr.advect_theta(t["th"], t["tha"], t["ax_umc"], t["ax_vmc"], t["psa"], t["xmsf"], t["qdot"], t["twt"], t["xds"], nr.scalar(o, "ul"),
                          case.cfg.t_extrema, idx, kz, upstream=case.cfg.upstream_mode != 0, stability_enhance=case.cfg.stability_enhance != 0)
    want = t["thten"].s((idx[0], idx[1]), (idx[2], idx[3]), (1, kz))
    sc = np.abs(want).max()
    assert sc > 0 and np.abs(got - want).max() <= 1e-12 * sc
    if kw.get("t_extrema"):
        # the limiter did act: without it the result differs
        free = nr.advect_theta(t["th"], t["tha"], t["ax_umc"], t["ax_vmc"], t["psa"], t["xmsf"], t["qdot"], t["twt"], t["xds"], nr.scalar(o, "ul"),
                               case.cfg.t_extrema, idx, kz, upstream=True, stability_enhance=False)
        assert np.abs(free - want).max() > 1e-6 * sc
