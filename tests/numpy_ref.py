"""TEST INFRASTRUCTURE — second, independent restatement of the riskiest dyn-core routines in NumPy.

SURVEY.md §8(c) / VERDICT r1: the parity oracle (oracle/, C++) was written by the same hand that wrote the CUDA
kernels, so a misreading of the Fortran would be common-mode.  The functions below were written from the
Fortran text of the reference alone (file:line cited per function), vectorised over whole (j,i) planes — a
different code shape from the loop-for-loop C++ — and are compared with the oracle in tests/test_numpy_ref.py.

Array convention: `F` wraps a NumPy array with Fortran lower bounds, so that the slices below can be written
with the reference's own index ranges: a[jdi1:jdi2, i-1, k] of the Fortran is `a.s((jdi1, jdi2), (idi1-1, idi2-1), k)`.
"""
import ctypes as C

import numpy as np


class F:
    """NumPy array with Fortran lower bounds; s(...) takes inclusive (lo, hi) ranges or single indices."""

    def __init__(self, a, lb):
        self.a, self.lb = a, tuple(lb)

    def s(self, *idx):
        sl = []
        for d, ix in enumerate(idx):
            if isinstance(ix, tuple):
                sl.append(slice(ix[0] - self.lb[d], ix[1] - self.lb[d] + 1))
            else:
                sl.append(ix - self.lb[d])
        return self.a[tuple(sl)]


def tap(o, name):
    """View (no copy) of an internal array of the oracle as an F with its Fortran bounds."""
    p = C.POINTER(C.c_double)()
    lb, n = (C.c_int32 * 4)(), (C.c_int32 * 4)()
    rc = o.lib.lib.orc_array(o.h, name.encode(), C.byref(p), lb, n)
    assert rc == 0, name
    shape = tuple(int(x) for x in n)
    nd = 4
    while nd > 1 and shape[nd - 1] == 1 and lb[nd - 1] == 1:
        nd -= 1
    a = np.ctypeslib.as_array(p, shape=(int(np.prod(shape)),)).reshape(shape[:nd] + (-1,), order="F")
    a = a.reshape(shape[:nd], order="F")
    return F(a, [int(x) for x in lb][:nd])


def scalar(o, name):
    v = C.c_double()
    assert o.lib.lib.orc_scalar(o.h, name.encode(), C.byref(v)) == 0, name
    return v.value


def call(o, routine, arg=0.0):
    o.lib.lib.orc_call.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
    assert o.lib.lib.orc_call(o.h, routine.encode(), C.c_double(arg)) == 0, routine


# ---------------------------------------------------------------------------------------------------------
def hadvuv_vadvuv(ua, va, u, v, uc, vc, divx, dmapf, svv, twt, xds, ul, idx, kz, upstream=True):
    """Main/mod_advection.F90:124-269 (hadvuv, idynamic = 2 branches) followed by :271-303 (vadvuv), applied to
    zero tendencies.  ua, va = atmx%umc, atmx%vmc (:81-82); u, v = atmx%ud, atmx%vd (hadv) and uc, vc =
    atmx%uc, atmx%vc (vadv) (Main/mod_tendency.F90:1293,1304); divx = mdv%cr; dmapf = mddom%dmsf; svv = qdot.
    Returns uten, vten on (jdi1:jdi2, idi1:idi2, 1:kz)."""
    jdi1, jdi2, idi1, idi2 = idx
    J, I = (jdi1, jdi2), (idi1, idi2)
    K = (1, kz)

    def sh(a, dj, di):
        return a.s((jdi1 + dj, jdi2 + dj), (idi1 + di, idi2 + di), K)

    dm = dmapf.s(J, I)[:, :, None]
    divd = 0.25 * (sh(divx, 0, 0) + sh(divx, 0, -1) + sh(divx, -1, 0) + sh(divx, -1, -1))
    ucmona = sh(ua, 0, 1) + 2.0 * sh(ua, 0, 0) + sh(ua, 0, -1)
    ucmonb = sh(ua, 1, 1) + 2.0 * sh(ua, 1, 0) + sh(ua, 1, -1)
    ucmonc = sh(ua, -1, 1) + 2.0 * sh(ua, -1, 0) + sh(ua, -1, -1)
    vcmona = sh(va, 1, 0) + 2.0 * sh(va, 0, 0) + sh(va, -1, 0)
    vcmonb = sh(va, 1, 1) + 2.0 * sh(va, 0, 1) + sh(va, -1, 1)
    vcmonc = sh(va, 1, -1) + 2.0 * sh(va, 0, -1) + sh(va, -1, -1)
    diag = divd - dm * ((ucmonb - ucmonc) + (vcmonb - vcmonc))
    if upstream:   # :225-262
        ff1 = ul * (sh(u, 1, 0) + sh(u, 0, 0))
        ff2 = ul * (sh(u, -1, 0) + sh(u, 0, 0))
        ff3 = ul * (sh(v, 0, 1) + sh(v, 0, 0))
        ff4 = ul * (sh(v, 0, -1) + sh(v, 0, 0))
        ucmonb = (1.0 + ff1) * ucmona + (1.0 - ff1) * ucmonb
        ucmonc = (1.0 + ff2) * ucmonc + (1.0 - ff2) * ucmona
        vcmonb = (1.0 + ff3) * vcmona + (1.0 - ff3) * vcmonb
        vcmonc = (1.0 + ff4) * vcmonc + (1.0 - ff4) * vcmona
    else:          # :166-190
        ucmonb = ucmona + ucmonb
        ucmonc = ucmonc + ucmona
        vcmonb = vcmona + vcmonb
        vcmonc = vcmonc + vcmona
    uten = 0.0 + sh(u, 0, 0) * diag - dm * (sh(u, 1, 0) * ucmonb - sh(u, -1, 0) * ucmonc + sh(u, 0, 1) * vcmonb - sh(u, 0, -1) * vcmonc)
    vten = 0.0 + sh(v, 0, 0) * diag - dm * (sh(v, 1, 0) * ucmonb - sh(v, -1, 0) * ucmonc + sh(v, 0, 1) * vcmonb - sh(v, 0, -1) * vcmonc)
    # vadvuv :283-296 — the k loop carries a dependence through uten(k-1), uten(k): keep its order
    for k in range(2, kz + 1):
        qq = 0.25 * (svv.s(J, I, k) + svv.s(J, (idi1 - 1, idi2 - 1), k) + svv.s((jdi1 - 1, jdi2 - 1), I, k) +
                     svv.s((jdi1 - 1, jdi2 - 1), (idi1 - 1, idi2 - 1), k))
        uu = qq * (twt.s(k, 1) * uc.s(J, I, k) + twt.s(k, 2) * uc.s(J, I, k - 1))
        vv = qq * (twt.s(k, 1) * vc.s(J, I, k) + twt.s(k, 2) * vc.s(J, I, k - 1))
        uten[:, :, k - 2] = uten[:, :, k - 2] - uu * xds.s(k - 1)
        uten[:, :, k - 1] = uten[:, :, k - 1] + uu * xds.s(k)
        vten[:, :, k - 2] = vten[:, :, k - 2] - vv * xds.s(k - 1)
        vten[:, :, k - 1] = vten[:, :, k - 1] + vv * xds.s(k)
    return uten, vten


# ---------------------------------------------------------------------------------------------------------
EGRAV = 9.80665                       # Share/mod_constants.F90:85
DLOWVAL = 1.0e-20                     # :68


def sound(t, p):
    """Main/mod_sound.F90:163-718 (`sound`, single rank, alarm_day not active), all `istep` sub-steps and the
    epilogue.  t: dict of F arrays (modified in place: ac_u, ac_v, ac_pp, ac_w, ac_t, a1_t, a2_t, a1_u ... a2_w,
    uten, vten, ppten, wten, s_estore, s_astore); p: scalars.  Returns (istep, max CFL number over the sub-steps)."""
    jx, iy, kz = p["jx"], p["iy"], p["kz"]
    kzp1 = kz + 1
    Jd, Id, Jdi, Idi = (1, jx), (1, iy), (2, jx - 1), (2, iy - 1)
    Jc, Ic, Jci, Ici = (1, jx - 1), (1, iy - 1), (2, jx - 2), (2, iy - 2)
    K, Kp = (1, kz), (1, kzp1)
    dt, dx, xkd, xgamma, gnu1, gnu2, cpd = p["dt"], p["dx"], p["xkd"], p["xgamma"], p["gnu1"], p["gnu2"], p["cpd"]
    bp, bm = (1.0 + p["bet"]) * 0.5, (1.0 - p["bet"]) * 0.5     # :157-160
    bpxbp, bpxbm = bp * bp, bp * bm
    regrav = 1.0 / EGRAV
    istep = max(int(dt / p["dtsmax"]), 2)                        # :201-205
    if p["lcount"] > 0:
        istep = max(4, istep)
    dts = dt / float(istep)

    def rng(r, d):
        return (r[0] + d, r[1] + d)

    def ci(a, dj=0, di=0, k=K, dk=0):      # a(jci1+dj:jci2+dj, ici1+di:ici2+di, k+dk)
        kk = rng(k, dk) if isinstance(k, tuple) else k + dk
        return a.s(rng(Jci, dj), rng(Ici, di), kk)

    def ci2(a, dj=0, di=0):
        return a.s(rng(Jci, dj), rng(Ici, di))

    psb, psdotb = t["psb"], t["psdotb"]
    rpsb_ce = 1.0 / psb.s(Jc, Ic)                                # :221-223
    rpsb = 1.0 / ci2(psb)
    u, v, pp, w, tt = t["ac_u"], t["ac_v"], t["ac_pp"], t["ac_w"], t["ac_t"]
    u.s(Jd, Id, K)[...] = t["a2_u"].s(Jd, Id, K) / psdotb.s(Jd, Id)[:, :, None]
    v.s(Jd, Id, K)[...] = t["a2_v"].s(Jd, Id, K) / psdotb.s(Jd, Id)[:, :, None]
    t["uten"].s(Jdi, Idi, K)[...] *= dts
    t["vten"].s(Jdi, Idi, K)[...] *= dts
    qv = ci(F(t["a2_qx"].a[..., 0], t["a2_qx"].lb[:3])) * rpsb[:, :, None]     # atmc%qx(:,:,:,iqv), :232-234
    pp.s(Jc, Ic, K)[...] = t["a2_pp"].s(Jc, Ic, K) * rpsb_ce[:, :, None]
    t["ppten"].s(Jci, Ici, K)[...] *= dts
    w.s(Jc, Ic, Kp)[...] = t["a2_w"].s(Jc, Ic, Kp) * rpsb_ce[:, :, None]
    t["wten"].s(Jci, Ici, Kp)[...] *= dts
    uten, vten, ppten, wten = t["uten"], t["vten"], t["ppten"], t["wten"]
    pr0, rho0, t0, ps0, rhof0 = t["a0_pr"], t["a0_rho"], t["a0_t"], t["a0_ps"], t["a0_rhof"]
    pr1, rho1 = t["a1_pr"], t["a1_rho"]
    msfd, msfx, ht = t["msfd"], t["msfx"], t["ht"]
    dsig, sigma, twt = t["dsigma"], t["sigma"], t["twt"]
    nj, ni = Jci[1] - Jci[0] + 1, Ici[1] - Ici[0] + 1
    pi = np.zeros((nj, ni, kz))
    cfl = 0.0
    m = {q: ci2(msfd, *d)[:, :, None] for q, d in (("00", (0, 0)), ("01", (0, 1)), ("10", (1, 0)), ("11", (1, 1)))}
    mx = ci2(msfx)[:, :, None]
    ps0c = ci2(ps0)[:, :, None]
    for it in range(1, istep + 1):
        if it > 1:
            ci(pp)[...] += xkd * pi                                                    # :252-256
        for k in range(1, kz + 1):                                                     # :257-264
            kp1, km1 = min(kz, k + 1), max(1, k - 1)
            tt.s(Jc, Ic, k)[...] = (pp.s(Jc, Ic, km1) - pp.s(Jc, Ic, kp1)) / (pr0.s(Jc, Ic, km1) - pr0.s(Jc, Ic, kp1))
        # ---- advance u and v (:270-299) ----
        def di4(a):   # a(j,i) + a(j-1,i) + a(j,i-1) + a(j-1,i-1) on the interior dots
            return a.s(Jdi, Idi, K), a.s(rng(Jdi, -1), Idi, K), a.s(Jdi, rng(Idi, -1), K), a.s(rng(Jdi, -1), rng(Idi, -1), K)
        r = di4(rho1); rho = 0.25 * (r[0] + r[1] + r[2] + r[3])
        d = di4(tt); dppdp0 = 0.25 * (d[0] + d[1] + d[2] + d[3])
        chh = 0.5 * dts / (rho * dx) / msfd.s(Jdi, Idi)[:, :, None]
        q = di4(pp)
        u.s(Jdi, Idi, K)[...] = u.s(Jdi, Idi, K) - chh * (q[0] - q[1] + q[2] - q[3] - t["a0_dprddx"].s(Jdi, Idi, K) * dppdp0)
        v.s(Jdi, Idi, K)[...] = v.s(Jdi, Idi, K) - chh * (q[0] - q[2] + q[1] - q[3] - t["a0_dprddy"].s(Jdi, Idi, K) * dppdp0)
        u.s(Jdi, Idi, K)[...] = u.s(Jdi, Idi, K) + uten.s(Jdi, Idi, K)
        v.s(Jdi, Idi, K)[...] = v.s(Jdi, Idi, K) + vten.s(Jdi, Idi, K)
        if it > 1:
            ci(pp)[...] -= xkd * pi                                                    # :302-306
        wo = ci(w, k=Kp).copy()                                                        # :317-319
        # ---- bottom boundary condition (:323-334) ----
        def four(a, k):   # a(j,i+1)+a(j,i)+a(j+1,i+1)+a(j+1,i) at level k
            return ci(a, 0, 1, k) + ci(a, 0, 0, k) + ci(a, 1, 1, k) + ci(a, 1, 0, k)
        wbot = 0.5 * 0.25 * regrav * (four(v, kz) * (ci2(ht, 0, 1) - ci2(ht, 0, -1)) +
                                      four(u, kz) * (ci2(ht, 1, 0) - ci2(ht, -1, 0))) / (dx * ci2(msfx))
        ci(w, k=kzp1)[...] = wbot
        e = np.zeros((nj, ni, kz)); f = np.zeros((nj, ni, kz))
        f[:, :, kz - 1] = wbot
        t["_work"] = dict(wo=wo, e=e, f=f, pi=pi, it=it, dts=dts, rpsb=rpsb, qv=qv)
        _sound_column(t, p, t["_work"], ci, ci2, m, mx, ps0c, bp, bm, bpxbp, bpxbm)
        cfl = max(cfl, t["_work"]["cfl"])
    # ---- epilogue: recouple and Robert-Asselin filters (:693-712, Main/mod_timefilter.F90:92-110,132-156) ----
    u.s(Jdi, Idi, K)[...] *= psdotb.s(Jdi, Idi)[:, :, None]
    v.s(Jdi, Idi, K)[...] *= psdotb.s(Jdi, Idi)[:, :, None]

    def ra(an, anm1, anp1, J, I, KK, alpha):
        dd = alpha * (anp1.s(J, I, KK) + anm1.s(J, I, KK) - 2.0 * an.s(J, I, KK))
        anm1.s(J, I, KK)[...] = an.s(J, I, KK) + dd
        an.s(J, I, KK)[...] = anp1.s(J, I, KK)
    ra(t["a1_u"], t["a2_u"], u, Jdi, Idi, K, gnu1)
    ra(t["a1_v"], t["a2_v"], v, Jdi, Idi, K, gnu1)
    ci(pp)[...] *= ci2(psb)[:, :, None]
    ra(t["a1_pp"], t["a2_pp"], pp, Jci, Ici, K, gnu1)
    w.a[np.abs(w.a) < DLOWVAL] = 0.0
    ci(w, k=Kp)[...] *= ci2(psb)[:, :, None]
    ra(t["a1_w"], t["a2_w"], w, Jci, Ici, Kp, gnu2)
    for nm in ("a2_w", "a1_w"):
        t[nm].a[np.abs(t[nm].a) < DLOWVAL] = 0.0
    return istep, cfl


def _sound_column(t, p, wk, ci, ci2, m, mx, ps0c, bp, bm, bpxbp, bpxbm):
    """Main/mod_sound.F90:335-444: the k-local coefficients cc, cdd, cj, tk, ca, g1, g2, the tridiagonal aa, b, c,
    the terrain terms pxup, pyvp and ptend — one sub-step, whole (j,i) planes at a time."""
    kz = p["kz"]
    dts, dx, xgamma = wk["dts"], p["dx"], p["xgamma"]
    u, v = t["ac_u"], t["ac_v"]
    pr0, rho0, t0 = t["a0_pr"], t["a0_rho"], t["a0_t"]
    pr1, rho1, a2t, dsig = t["a1_pr"], t["a1_rho"], t["a2_t"], t["dsigma"]
    rpsb = wk["rpsb"]
    ds = np.array([dsig.s(k) for k in range(1, kz + 1)])          # dsigma(1:kz)

    def four(a, k):     # a(j,i) + a(j+1,i) + a(j,i+1) + a(j+1,i+1) at level k
        return ci(a, 0, 0, k) + ci(a, 1, 0, k) + ci(a, 0, 1, k) + ci(a, 1, 1, k)

    def four_diff(a, ka, kb):   # the eight-term bracket of pxup/pyvp: + four at ka, - four at kb, in the reference's order
        return (ci(a, 0, 0, ka) + ci(a, 1, 0, ka) + ci(a, 0, 1, ka) + ci(a, 1, 1, ka) -
                ci(a, 0, 0, kb) - ci(a, 1, 0, kb) - ci(a, 0, 1, kb) - ci(a, 1, 1, kb))

    cc = xgamma * ci(pr1) * dts / (dx * mx)                                             # :336, :386
    cdd = xgamma * ci(pr1) * ci(rho0) * EGRAV * dts / (ps0c * ds[None, None, :])        # :337-338, :387-388
    cj = 0.5 * ci(rho0) * EGRAV * dts                                                   # :339, :389
    tk = (0.5 * ps0c * ci(t0)) / (xgamma * ci(pr0) * ci(a2t) * rpsb[:, :, None])        # :370-371, :377-378
    nj, ni = cc.shape[:2]
    z = lambda: np.zeros((nj, ni, kz))   # noqa: E731
    pxup, pyvp, ca, g1, g2, aa, b, c = z(), z(), z(), z(), z(), z(), z(), z()
    den1 = ci(pr0, k=1) - ci(pr0, k=2)                                                  # :340-352
    pxup[:, :, 0] = 0.0625 * (ci(pr0, 1, 0, 1) - ci(pr0, -1, 0, 1)) * four_diff(u, 1, 2) / den1
    pyvp[:, :, 0] = 0.0625 * (ci(pr0, 0, 1, 1) - ci(pr0, 0, -1, 1)) * four_diff(v, 1, 2) / den1
    for k in range(2, kz + 1):                                                          # :373-416
        kp1, km1 = min(k + 1, kz), k - 1
        q, qm = k - 1, km1 - 1                                                          # 0-based positions of k, km1
        rofac = (ds[qm] * ci(rho0, k=k) + ds[q] * ci(rho0, k=km1)) / (ds[qm] * ci(rho1, k=k) + ds[q] * ci(rho1, k=km1))
        ca[:, :, q] = EGRAV * dts / (ci(pr0, k=k) - ci(pr0, k=km1)) * rofac
        g1[:, :, q] = 1.0 - ds[qm] * tk[:, :, q]
        g2[:, :, q] = 1.0 + ds[q] * tk[:, :, qm]
        c[:, :, q] = -ca[:, :, q] * (cdd[:, :, qm] - cj[:, :, qm]) * g2[:, :, q] * bpxbp
        b[:, :, q] = 1.0 + ca[:, :, q] * (g1[:, :, q] * (cdd[:, :, q] - cj[:, :, q]) + g2[:, :, q] * (cdd[:, :, qm] + cj[:, :, qm])) * bpxbp
        aa[:, :, q] = -ca[:, :, q] * (cdd[:, :, q] + cj[:, :, q]) * g1[:, :, q] * bpxbp
        den = ci(pr0, k=km1) - ci(pr0, k=kp1)
        pyvp[:, :, q] = 0.125 * (ci(pr0, 0, 1, k) - ci(pr0, 0, -1, k)) * four_diff(v, km1, kp1) / den
        pxup[:, :, q] = 0.125 * (ci(pr0, 1, 0, k) - ci(pr0, -1, 0, k)) * four_diff(u, km1, kp1) / den
    pyvp[:, :, kz - 1] = pyvp[:, :, kz - 1] * 0.5                                       # :424-427
    pxup[:, :, kz - 1] = pxup[:, :, kz - 1] * 0.5
    # ptend (:359-369 for k = 1, :433-444 for k >= 2: the same expression)
    div = (ci(v, 0, 1) * m["01"] - ci(v, 0, 0) * m["00"] + ci(v, 1, 1) * m["11"] - ci(v, 1, 0) * m["10"] +
           ci(u, 1, 0) * m["10"] - ci(u, 0, 0) * m["00"] + ci(u, 1, 1) * m["11"] - ci(u, 0, 1) * m["01"])
    ptend = ci(t["ppten"]) - 0.5 * cc * (div / mx - 2.0 * (pyvp + pxup))
    wk.update(cc=cc, cdd=cdd, cj=cj, tk=tk, ca=ca, g1=g1, g2=g2, aa=aa, b=b, c=c, ptend=ptend)
    _sound_solve(t, p, wk, ci, ci2, ps0c, bp, bm, bpxbm)


def _sound_solve(t, p, wk, ci, ci2, ps0c, bp, bm, bpxbm):
    """Main/mod_sound.F90:445-689: rhs, explicit p' update, upward elimination, radiative upper BC (13x13
    convolution of estore with tmask, clamped indices), downward sweep, CFL, new p', temperature correction."""
    jx, iy, kz = p["jx"], p["iy"], p["kz"]
    kzp1 = kz + 1
    dts, dt, gnu1, cpd, xgamma = wk["dts"], p["dt"], p["gnu1"], p["cpd"], p["xgamma"]
    u, v, pp, w = t["ac_u"], t["ac_v"], t["ac_pp"], t["ac_w"]
    rho0, rhof0, pr1, rho1 = t["a0_rho"], t["a0_rhof"], t["a1_pr"], t["a1_rho"]
    dsig, sigma, twt, wten, ppten = t["dsigma"], t["sigma"], t["twt"], t["wten"], t["ppten"]
    wo, e, f, pi, qv = wk["wo"], wk["e"], wk["f"], wk["pi"], wk["qv"]
    cdd, cj, ca, g1, g2, aa, b, c, ptend = (wk[n] for n in ("cdd", "cj", "ca", "g1", "g2", "aa", "b", "c", "ptend"))
    nj, ni = cdd.shape[:2]
    rhs = np.zeros((nj, ni, kz))
    ppc = ci(pp)                      # view of atmc%pp on the interior
    wc = ci(w, k=(1, kzp1))           # view of atmc%w on the interior (w(kzp1) already holds the bottom BC)
    for k in range(2, kz + 1):        # :445-456 (atmc%w(k) still equals wo(k) for k <= kz)
        q, qm = k - 1, k - 2
        rhs[:, :, q] = wc[:, :, q] + ci(wten, k=k) + ca[:, :, q] * (
            bpxbm * ((cdd[:, :, qm] - cj[:, :, qm]) * g2[:, :, q] * wo[:, :, qm] -
                     ((cdd[:, :, qm] + cj[:, :, qm]) * g2[:, :, q] + (cdd[:, :, q] - cj[:, :, q]) * g1[:, :, q]) * wo[:, :, q] +
                     (cdd[:, :, q] + cj[:, :, q]) * g1[:, :, q] * wo[:, :, q + 1]) +
            (ppc[:, :, q] * g1[:, :, q] - ppc[:, :, qm] * g2[:, :, q]) +
            (g1[:, :, q] * ptend[:, :, q] - g2[:, :, q] * ptend[:, :, qm]) * bp)
    pi[...] = ppc                                                                      # :458-459
    ppc[...] = ppc + ptend + (cj * (wo[:, :, 1:] + wo[:, :, :-1]) + cdd * (wo[:, :, 1:] - wo[:, :, :-1])) * bm   # :465-467
    for k in range(kz, 1, -1):                                                         # :472-480
        q = k - 1
        denom = aa[:, :, q] * e[:, :, q] + b[:, :, q]
        e[:, :, q - 1] = -c[:, :, q] / denom
        f[:, :, q - 1] = (rhs[:, :, q] - f[:, :, q] * aa[:, :, q]) / denom
    wpval = np.zeros((nj, ni))
    if p["ifupr"] == 1:                                                                # :489-561
        denom = (cdd[:, :, 0] + cj[:, :, 0]) * bp
        est, ast = t["s_estore"], t["s_astore"]
        ci2(est)[...] = ppc[:, :, 0] + f[:, :, 0] * denom
        ci2(ast)[...] = denom * e[:, :, 0] + (cj[:, :, 0] - cdd[:, :, 0]) * bp
        tmask = t["tmask"]
        icross1, icross2, jcross1, jcross2 = 1, iy - 1, 1, jx - 1
        jj = np.arange(2, jx - 2 + 1); ii = np.arange(2, iy - 2 + 1)
        for nsi in range(-6, 7):          # summation order of the reference: nsi outer, nsj inner
            inn = np.clip(ii + nsi, icross1 + 1, icross2 - 1)
            for nsj in range(-6, 7):
                jnn = np.clip(jj + nsj, jcross1 + 1, jcross2 - 1)
                wpval = wpval + est.a[np.ix_(jnn - est.lb[0], inn - est.lb[1])] * tmask.s(nsj, nsi)
    wc[:, :, 0] = wpval                                                                # :566-568
    for k in range(1, kz + 1):                                                         # :572-574
        wc[:, :, k] = e[:, :, k - 1] * wc[:, :, k - 1] + f[:, :, k - 1]
    # ---- CFL (:611-633) ----
    ucrs = ci(u, 0, 0) + ci(u, 0, 1) + ci(u, 1, 0) + ci(u, 1, 1)
    vcrs = ci(v, 0, 0) + ci(v, 0, 1) + ci(v, 1, 0) + ci(v, 1, 1)
    cfl = 0.0
    dpx, dpy = ci2(t["dpsdxm"]), ci2(t["dpsdym"])
    for k in range(kz, 1, -1):
        q = k - 1
        sigdot = -ci(rhof0, k=k) * EGRAV * wc[:, :, q] / ps0c[:, :, 0] - sigma.s(k) * (
            dpx * (twt.s(k, 1) * ucrs[:, :, q] + twt.s(k, 2) * ucrs[:, :, q - 1]) +
            dpy * (twt.s(k, 1) * vcrs[:, :, q] + twt.s(k, 2) * vcrs[:, :, q - 1]))
        check = np.abs(sigdot) * dt / (dsig.s(k) + dsig.s(k - 1))
        cfl = max(cfl, float(check.max()))
    wk["cfl"] = cfl
    # ---- new pressure and temperature correction (:669-689) ----
    ds = np.array([dsig.s(k) for k in range(1, kz + 1)])
    ppold = pi.copy()
    cddtmp = xgamma * ci(pr1) * ci(rho0) * EGRAV * dts / (ps0c * ds[None, None, :])
    cjtmp = ci(rho0) * EGRAV * dts * 0.5
    ppc[...] = ppc + (cjtmp * (wc[:, :, 1:] + wc[:, :, :-1]) + cddtmp * (wc[:, :, 1:] - wc[:, :, :-1])) * bp
    pi[...] = ppc - ppold - ci(ppten)
    cpm = cpd * (1.0 + 0.80 * qv)                                                      # Share/cpmf.inc
    dpterm = ci2(t["psb"])[:, :, None] * (ppc - ppold) / (cpm * ci(rho1))
    ci(t["a2_t"])[...] = ci(t["a2_t"]) + gnu1 * dpterm
    ci(t["a1_t"])[...] = ci(t["a1_t"]) + dpterm


# ---------------------------------------------------------------------------------------------------------
MINDEN = 1.0e-15                      # Main/mod_moloch.F90:96


def _rdeno(t1, t2, t3, t4):
    """Main/mod_moloch.F90:1090-1096."""
    zzden = t3 - t4
    return (t1 - t2) / np.copysign(np.maximum(np.abs(zzden), MINDEN), zzden)


def _waf_phi(r, zamu, isgn):
    """limiter and weight, :1135-1137."""
    b = np.maximum(0.0, np.minimum(2.0, np.maximum(r, np.minimum(2.0 * r, 1.0))))
    return isgn + zamu * b - isgn * b


def wafone(pp, s, fmz, fmzf, u, v, rmu, rmv, mx2, dta, dx, dz, jx, iy, kz):
    """Main/mod_moloch.F90:1098-1361, generic projection branch (iproj /= 'ROTLLR'), single rank: WAF advection of
    one cross-point field, directional split z -> y -> x.  pp: F array, modified in place on the interior."""
    Jci, Ici = (2, jx - 2), (2, iy - 2)
    jcross1, jcross2, icross1, icross2 = 1, jx - 1, 1, iy - 1
    zdtrdx, zdtrdy, zdtrdz = dta / dx, dta / dx, dta / dz
    A = pp.a                                    # (jx-1, iy-1, kz), lower bounds 1
    assert pp.lb == (1, 1, 1) or tuple(pp.lb) == (1, 1, 1)
    jj = slice(Jci[0] - 1, Jci[1])              # 0-based interior slices
    ii = slice(Ici[0] - 1, Ici[1])
    P = A[jj, ii, :]
    S = s.s(Jci, Ici, (1, kz + 1))
    FZ, FZF = fmz.s(Jci, Ici, (1, kz)), fmzf.s(Jci, Ici, (1, kz + 1))
    nj, ni = P.shape[:2]
    # ---- vertical advection (:1118-1160) ----
    wfw = np.zeros((nj, ni, kz + 1))            # wfw(j,1) = wfw(j,kzp1) = 0
    for k in range(1, kz):                      # face k+1 between levels k and k+1
        zamu = S[:, :, k] * zdtrdz              # s(j,i,k+1)
        up = zamu >= 0.0
        isgn = np.where(up, 1.0, -1.0)
        k1_u, k1m1_u = k + 1, min(k + 2, kz)    # zamu >= 0
        k1m1_d, k1_d = k, max(k - 1, 1)         # zamu < 0
        p_k1 = np.where(up, P[:, :, k1_u - 1], P[:, :, k1_d - 1])
        p_k1m1 = np.where(up, P[:, :, k1m1_u - 1], P[:, :, k1m1_d - 1])
        r = _rdeno(p_k1, p_k1m1, P[:, :, k - 1], P[:, :, k])
        zphi = _waf_phi(r, zamu, isgn)
        wfw[:, :, k] = 0.5 * S[:, :, k] * ((1.0 + zphi) * P[:, :, k] + (1.0 - zphi) * P[:, :, k - 1])
    wz = np.zeros((jx - 1, iy - 1, kz))         # full cross extent; only interior j columns are used
    zrfmm = zdtrdz * FZ[:, :, 0] / FZF[:, :, 1]
    zdv = -S[:, :, 1] * zrfmm * P[:, :, 0]
    wz[jj, ii, 0] = P[:, :, 0] + wfw[:, :, 1] * zrfmm + zdv
    for k in range(2, kz + 1):
        zrfmm = zdtrdz * FZ[:, :, k - 1] / FZF[:, :, k]
        zrfmp = zdtrdz * FZ[:, :, k - 1] / FZF[:, :, k - 1]
        zdv = (S[:, :, k - 1] * zrfmp - S[:, :, k] * zrfmm) * P[:, :, k - 1]
        wz[jj, ii, k - 1] = P[:, :, k - 1] - wfw[:, :, k - 1] * zrfmp + wfw[:, :, k] * zrfmm + zdv
    wz[jj, 0, :] = wz[jj, Ici[0] - 1, :]        # has_bdybottom: wz(j,ice1,k) = wz(j,ici1,k)  (:1170-1184)
    wz[jj, iy - 2, :] = wz[jj, Ici[1] - 1, :]   # has_bdytop:    wz(j,ice2,k) = wz(j,ici2,k)
    return _wafone_horizontal(pp, wz, fmz, u, v, rmu, rmv, mx2, zdtrdx, zdtrdy, jx, iy, kz)


def _wafone_horizontal(pp, wz, fmz, u, v, rmu, rmv, mx2, zdtrdx, zdtrdy, jx, iy, kz):
    """Main/mod_moloch.F90:1271-1357: meridional then zonal pass of the generic branch."""
    Jci, Ici, K = (2, jx - 2), (2, iy - 2), (1, kz)
    jcross1, jcross2, icross1, icross2 = 1, jx - 1, 1, iy - 1
    A = pp.a
    jj = slice(Jci[0] - 1, Jci[1])
    ii = slice(Ici[0] - 1, Ici[1])
    P = A[jj, ii, :].copy()                     # pp of the divergence terms: the values before this call's update
    # ---- meridional advection: faces i = ici1 .. ice2 (= iy-1) ----
    If = np.arange(Ici[0], iy - 1 + 1)          # 1-based face index i
    vf = v.s(Jci, (If[0], If[-1]), K) * rmv.s(Jci, (If[0], If[-1]))[:, :, None]     # v(j,i,k)*rmv(j,i)
    zamu = vf * zdtrdy
    pos = zamu > 0.0
    isgn = np.where(pos, 1.0, -1.0)
    ih = np.where(pos, If[None, :, None] - 1, np.minimum(If[None, :, None] + 1, icross2))
    ihm1 = np.maximum(ih - 1, icross1)
    W = wz[jj, :, :]                            # (nj, iy-1, kz), second axis = i-1
    w_ih = np.take_along_axis(W, ih - 1, axis=1)
    w_ihm1 = np.take_along_axis(W, ihm1 - 1, axis=1)
    w_i = W[:, If - 1, :]
    w_im1 = W[:, If - 2, :]
    zphi = _waf_phi(_rdeno(w_ih, w_ihm1, w_i, w_im1), zamu, isgn)
    zpby = 0.5 * vf * ((1.0 + zphi) * w_im1 + (1.0 - zphi) * w_i)                  # zpby(j, i), i = If
    FZ = fmz.s(Jci, Ici, K)
    zrfmn = zdtrdy * 2.0 * FZ / (FZ + fmz.s(Jci, (Ici[0] + 1, Ici[1] + 1), K))
    zrfms = zdtrdy * 2.0 * FZ / (FZ + fmz.s(Jci, (Ici[0] - 1, Ici[1] - 1), K))
    ni = Ici[1] - Ici[0] + 1
    zdv = (vf[:, 1:ni + 1, :] * zrfmn - vf[:, 0:ni, :] * zrfms) * P
    p0 = np.zeros((jx - 1, iy - 1, kz))
    p0[jj, ii, :] = wz[jj, ii, :] + mx2.s(Jci, Ici)[:, :, None] * (zpby[:, 0:ni, :] * zrfms - zpby[:, 1:ni + 1, :] * zrfmn + zdv)
    p0[0, ii, :] = p0[Jci[0] - 1, ii, :]        # has_bdyleft : p0(jce1,i,k) = p0(jci1,i,k)
    p0[jx - 2, ii, :] = p0[Jci[1] - 1, ii, :]   # has_bdyright: p0(jce2,i,k) = p0(jci2,i,k)
    # ---- zonal advection: faces j = jci1 .. jce2 (= jx-1) ----
    Jf = np.arange(Jci[0], jx - 1 + 1)
    uf = u.s((Jf[0], Jf[-1]), Ici, K) * rmu.s((Jf[0], Jf[-1]), Ici)[:, :, None]     # u(j,i,k)*rmu(j,i)
    zamu = uf * zdtrdx
    pos = zamu > 0.0
    isgn = np.where(pos, 1.0, -1.0)
    jh = np.where(pos, Jf[:, None, None] - 1, np.minimum(Jf[:, None, None] + 1, jcross2))
    jhm1 = np.maximum(jh - 1, jcross1)
    Q = p0[:, ii, :]                            # (jx-1, ni, kz), first axis = j-1
    q_jh = np.take_along_axis(Q, jh - 1, axis=0)
    q_jhm1 = np.take_along_axis(Q, jhm1 - 1, axis=0)
    q_j = Q[Jf - 1, :, :]
    q_jm1 = Q[Jf - 2, :, :]
    zphi = _waf_phi(_rdeno(q_jh, q_jhm1, q_j, q_jm1), zamu, isgn)
    zpbw = 0.5 * uf * ((1.0 + zphi) * q_jm1 + (1.0 - zphi) * q_j)                  # zpbw(j, i), j = Jf
    zrfme = zdtrdx * 2.0 * FZ / (FZ + fmz.s((Jci[0] + 1, Jci[1] + 1), Ici, K))
    zrfmw = zdtrdx * 2.0 * FZ / (FZ + fmz.s((Jci[0] - 1, Jci[1] - 1), Ici, K))
    nj = Jci[1] - Jci[0] + 1
    zdv = (uf[1:nj + 1, :, :] * zrfme - uf[0:nj, :, :] * zrfmw) * P
    A[jj, ii, :] = p0[jj, ii, :] + mx2.s(Jci, Ici)[:, :, None] * (zpbw[0:nj, :, :] * zrfmw - zpbw[1:nj + 1, :, :] * zrfme + zdv)
    return pp


# ---------------------------------------------------------------------------------------------------------
def sl_tendency(varb, vara, umd, vmd, msfx, msfd, dt, dx, box, kz, iqmsl=1):
    """Semi-Lagrangian horizontal advection as `tend` applies it to a p*-coupled scalar (isladvec = 1), written from
    Main/mod_sladvection.F90: adv_velocity(.false.) (:92-114, with the repeated va(j+1,i,k) in vadym1_x), trajcalc_x
    (:139-221: third-order departure distance, integer cell and fractional part, the four stencil indices per direction,
    cubic weights), slhadv_x (:347-393: bilinear on the two outer rows, cubic on the two inner ones, cubic across,
    quasi-monotonic clip, tendency (value - var)/dt when it exceeds dlowval) on time level b, minus hdvg_x (:568-588:
    var * horizontal divergence of the map-scaled wind, for var > epsilon) on time level a.  umd, vmd = atmx%umd/vmd;
    msfx, msfd as stored.  Away from the lateral boundary (no index clamping).  Returns the tendency on the box."""
    j1, j2, i1, i2 = box

    def c(a, dj=0, di=0):
        return a.s((j1 + dj, j2 + dj), (i1 + di, i2 + di), (1, kz)) if len(a.lb) == 3 else a.s((j1 + dj, j2 + dj), (i1 + di, i2 + di))[:, :, None]

    uadvx = 0.25 * (c(umd) + c(umd, 0, 1) + c(umd, 1, 1) + c(umd, 1, 0)) / c(msfx)
    uadxp1 = 0.25 * (c(umd, 1, 0) + c(umd, 1, 1) + c(umd, 2, 1) + c(umd, 2, 0)) / c(msfx, 1, 0)
    uadxm1 = 0.25 * (c(umd) + c(umd, 0, 1) + c(umd, -1, 1) + c(umd, -1, 0)) / c(msfx, -1, 0)
    vadvy = 0.25 * (c(vmd) + c(vmd, 0, 1) + c(vmd, 1, 1) + c(vmd, 1, 0)) / c(msfx)
    vadyp1 = 0.25 * (c(vmd, 0, 1) + c(vmd, 1, 1) + c(vmd, 1, 2) + c(vmd, 0, 2)) / c(msfx, 0, 1)
    vadym1 = 0.25 * (c(vmd) + c(vmd, 0, -1) + c(vmd, 1, 0) + c(vmd, 1, 0)) / c(msfx, 0, -1)
    dtsq, dtcb = dt * dt, dt * dt * dt

    def depart(w, wp, wm):
        d1 = 0.5 * (wp - wm) / dx
        d2 = (wp - 2.0 * w + wm) / (dx * dx)
        dis = -w * dt + 0.5 * (dtsq * w * d1) - (dtcb * w) * (d1 * d1 + w * d2) / 6.0
        n = dis / dx
        npart = np.trunc(n).astype(int)
        assert np.abs(npart).max() <= 1
        frac = np.abs((npart * dx - dis) / dx)
        sgn = np.where(np.signbit(n), -1, 1)
        return npart, sgn, frac

    xnp, xsn, al = depart(uadvx, uadxp1, uadxm1)
    ynp, ysn, be = depart(vadvy, vadyp1, vadym1)
    J = np.arange(j1, j2 + 1)[:, None, None]
    I = np.arange(i1, i2 + 1)[None, :, None]
    K = np.arange(kz)[None, None, :]
    x0 = J + xnp; xm1 = x0 + xsn; xm2 = xm1 + xsn; xp1 = x0 - xsn
    y0 = I + ynp; ym1 = y0 + ysn; ym2 = ym1 + ysn; yp1 = y0 - ysn
    V = lambda x, y: varb.a[x - varb.lb[0], y - varb.lb[1], K]          # noqa: E731
    am2, am1, a0, ap1 = -(al * (1 - al * al)) / 6.0, (al * (1 + al) * (2 - al)) / 2.0, ((1 - al * al) * (2 - al)) / 2.0, -(al * (1 - al) * (2 - al)) / 6.0
    bm2, bm1, b0, bp1 = -(be * (1 - be * be)) / 6.0, (be * (1 + be) * (2 - be)) / 2.0, ((1 - be * be) * (2 - be)) / 2.0, -(be * (1 - be) * (2 - be)) / 6.0
    bl1 = al * V(xm1, yp1) + (1.0 - al) * V(x0, yp1)
    bl2 = al * V(xm1, ym2) + (1.0 - al) * V(x0, ym2)
    cb1 = am2 * V(xm2, y0) + am1 * V(xm1, y0) + a0 * V(x0, y0) + ap1 * V(xp1, y0)
    cb2 = am2 * V(xm2, ym1) + am1 * V(xm1, ym1) + a0 * V(x0, ym1) + ap1 * V(xp1, ym1)
    val = bm2 * bl2 + bm1 * cb2 + b0 * cb1 + bp1 * bl1
    if iqmsl == 1:
        four = np.stack([V(x0, y0), V(x0, ym1), V(xm1, y0), V(xm1, ym1)])
        val = np.clip(val, four.min(axis=0), four.max(axis=0))
    vb = c(varb)
    ten = np.where(np.abs(val - vb) > 1.0e-20, (val - vb) / dt, 0.0)
    ucapf = (c(umd, 1, 1) * c(msfd, 1, 1) + c(umd, 1, 0) * c(msfd, 1, 0)) * 0.5
    ucapi = (c(umd, 0, 1) * c(msfd, 0, 1) + c(umd) * c(msfd)) * 0.5
    vcapf = (c(vmd, 1, 1) * c(msfd, 1, 1) + c(vmd, 0, 1) * c(msfd, 0, 1)) * 0.5
    vcapi = (c(vmd, 1, 0) * c(msfd, 1, 0) + c(vmd) * c(msfd)) * 0.5
    hdvg = ((ucapf - ucapi) / dx + (vcapf - vcapi) / dx) / (c(msfx) * c(msfx))
    va = c(vara)
    return ten - np.where(va > np.finfo(float).eps, va * hdvg, 0.0)


def tracer_step(chia, chib, umc, vmc, psa, rpsa, rpsb, xmsf, svv, twt, xds, xkc, ul, dt, gnu2, box, kz, upstream=True, idiffu=1,
                kpb=None, sigma=None, sl=None):
    """One leapfrog step of ONE tracer away from the lateral zones, written from the Fortran text:
    decouple (Main/mod_tendency.F90:1029-1034), start_advect (Main/mod_advection.F90:109-122: uavg/vavg from the
    p*.u.m fluxes), hadvtr (:666-720), vadv4d ind = 2 (:895-914,946-950), diffu_x4d interior stencil
    (Main/mod_diffusion.F90:658-725 with chib3d = max(atm2%chi*rpsb, 0), Main/mod_slice.F90:198; idiffu = 2: the
    nine-point operator of :881-894 with the weights 4/6, 1/6, -20/6 of :63-65), the sums and the
    forecast (Main/mod_tendency.F90:548-569) and the RAW filter (Main/mod_timefilter.F90:231-278, alpha = gnu2,
    beta = 0.53, floor 0).  With `kpb` (F array of the PBL top on cross points) and `sigma` (full levels, 1-based F) the
    vertical flux is vadv4d ind = 3 (:915-957) instead of ind = 2.  With `sl` (dict: umd, vmd, msfx, msfd, dx, iqmsl) the
    horizontal advection is the semi-Lagrangian one (sl_tendency) instead of hadvtr.  `box` = (j1, j2, i1, i2): interior points at least three rows away from the zones.
    Returns (forecast atmc%chi, filtered atm1%chi, filtered atm2%chi) on the box."""
    j1, j2, i1, i2 = box
    K = (1, kz)

    def c(a, dj=0, di=0, k=K):
        return a.s((j1 + dj, j2 + dj), (i1 + di, i2 + di), k) if len(a.lb) == 3 else a.s((j1 + dj, j2 + dj), (i1 + di, i2 + di))[:, :, None]

    # start_advect: averages of the coupled, map-scaled winds on the cell faces
    uavg1 = c(umc, 0, 1) + c(umc, 0, 0)
    uavg2 = c(umc, 1, 1) + c(umc, 1, 0)
    vavg1 = c(vmc, 1, 0) + c(vmc, 0, 0)
    vavg2 = c(vmc, 1, 1) + c(vmc, 0, 1)
    x = lambda dj=0, di=0: c(chia, dj, di) * c(rpsa, dj, di)          # atmx%chi
    if upstream:
        f1 = 0.5 * ul * (uavg2 + uavg1) / c(psa)
        f2 = 0.5 * ul * (vavg2 + vavg1) / c(psa)
        fx1 = (1.0 + f1) * x(-1, 0) + (1.0 - f1) * x()
        fx2 = (1.0 + f1) * x() + (1.0 - f1) * x(1, 0)
        fy1 = (1.0 + f2) * x(0, -1) + (1.0 - f2) * x()
        fy2 = (1.0 + f2) * x() + (1.0 - f2) * x(0, 1)
    else:
        fx1, fx2, fy1, fy2 = x(-1, 0) + x(), x() + x(1, 0), x(0, -1) + x(), x() + x(0, 1)
    ten = -c(xmsf) * (uavg2 * fx2 - uavg1 * fx1 + vavg2 * fy2 - vavg1 * fy1)
    if sl is not None:
        ten = sl_tendency(chib, chia, sl["umd"], sl["vmd"], sl["msfx"], sl["msfd"], dt, sl["dx"], box, kz, sl.get("iqmsl", 1))
    # vadv4d, ind = 2: flux through interface k = 2..kz, zero unless the upwind coupled value exceeds mintr*ps
    fk = c(chia, k=(2, kz))
    fkm1 = c(chia, k=(1, kz - 1))
    sv = c(svv, k=(2, kz))
    thr = 1.0e-30 * c(psa)
    ok = np.where(sv > 0.0, fkm1 > thr, fk > thr)
    tw1 = twt.s((2, kz), 1)[None, None, :]
    tw2 = twt.s((2, kz), 2)[None, None, :]
    fg = np.where(ok, sv * (tw1 * fk + tw2 * fkm1), 0.0)
    if kpb is not None:
        fg = vadv_ind3_faces(c(chia), kpb.s((j1, j2), (i1, i2)), sigma, twt, kz) * sv
    xd = xds.s((1, kz))[None, None, :]
    ten[:, :, :-1] -= fg * xd[:, :, :-1]
    ten[:, :, 1:] += fg * xd[:, :, 1:]
    # diffu_x4d on chib3d
    b = lambda dj=0, di=0: np.maximum(c(chib, dj, di) * c(rpsb, dj, di), 0.0)
    if idiffu == 2:
        ten += c(xkc) * (4.0 / 6.0 * (b(1, 0) + b(-1, 0) + b(0, 1) + b(0, -1)) + 1.0 / 6.0 * (b(1, 1) + b(-1, -1) + b(-1, 1) + b(1, -1))
                         - 20.0 / 6.0 * b())
    else:
        ten -= c(xkc) * ((b(2, 0) + b(-2, 0) + b(0, 2) + b(0, -2)) - 4.0 * (b(1, 0) + b(-1, 0) + b(0, 1) + b(0, -1)) + 12.0 * b())
    fc = c(chib) + dt * ten
    fcp = np.where(fc < 0.0, np.nan, fc)        # the negative fix is not part of this restatement
    d = gnu2 * (fcp + c(chib) - 2.0 * c(chia))
    nb = np.maximum(c(chia) + 0.53 * d, 0.0)
    na = np.maximum(fcp + (0.53 - 1.0) * d, 0.0)
    return fc, na, nb


# ---------------------------------------------------------------------------------------------------------
def vadv_ind3_faces(f, kpb, sigma, twt, kz):
    """vadv4d ind = 3 (Main/mod_advection.F90:915-957): the value carried through interfaces 2..kz (before the product
    with sigma-dot) of the p*-coupled scalar f (nj, ni, kz): the twt interpolation, replaced at interfaces kpb-1 and kpb
    (kpb >= 4) by an extrapolation from level kpb-2 along the limited slope of the layer above.  sigma: F array, full
    levels 1..kz+1."""
    sg = lambda k: float(sigma.s(k))                               # noqa: E731
    hs = lambda k: (sg(k + 1) + sg(k)) * 0.5                       # noqa: E731  (Main/mod_params.F90:2008)
    fg = np.zeros(f.shape[:2] + (kz + 1,))                         # index = interface number, 2..kz used
    for k in range(2, kz + 1):
        fg[:, :, k] = float(twt.s(k, 1)) * f[:, :, k - 1] + float(twt.s(k, 2)) * f[:, :, k - 2]
    for jj in range(f.shape[0]):
        for ii in range(f.shape[1]):
            kp = int(kpb[jj, ii])
            if kp < 4:
                continue
            q = lambda k: f[jj, ii, k - 1]                          # noqa: E731
            k = kp - 2
            up, dn = q(k + 1) - q(k), q(k) - q(k - 1)
            if up > 0.0 and dn > 0.0:
                slope = min(up / (hs(k + 1) - hs(k)), dn / (hs(k) - hs(k - 1)))
            elif up < 0.0 and dn < 0.0:
                slope = max(up / (hs(k + 1) - hs(k)), dn / (hs(k) - hs(k - 1)))
            else:
                slope = 0.0
            k = kp
            fg[jj, ii, k - 1] = q(k - 2) + slope * (sg(k - 1) - hs(k - 2))
            if abs(q(k - 2) + slope * (hs(k - 1) - hs(k - 2)) - q(k)) > abs(q(k - 1) - q(k)):
                fg[jj, ii, k] = q(k)
            else:
                fg[jj, ii, k] = q(k - 2) + slope * (sg(k) - hs(k - 2))
    return fg[:, :, 2:]


def tke_step(tkea, tkeb, umc, vmc, psa, rpsa, xmsf, svv, twt, dds, xkcf, tkephy, ul, dt, gnu2, nuk, tkemin, box, kz, upstream=True,
             idiffu=1):
    """One leapfrog step of the UW-PBL TKE away from the lateral boundary, written from the Fortran text: hadv3d ind = 1
    (Main/mod_advection.F90:395-515: winds interpolated to the full levels 2..kz with twt, upstream weights from the
    half-level winds of level k), vadv3d nk = kzp1 (:755-766) on tke*psa (Main/mod_tendency.F90:1414-1425), diffu_x3df
    with fac = nuk on atm2%tke (Main/mod_diffusion.F90:523-601, interior stencil; xkcf = the scaled coefficient), the sum
    tkedyn*rpsa + tkephy, the floored forecast (Main/mod_tendency.F90:518-541) and the RA filter
    (Main/mod_timefilter.F90:92-110, alpha = gnu2).  tkea / tkeb: atm1 / atm2 TKE (uncoupled) before the step.
    Returns (atmc%tke, filtered atm1%tke, filtered atm2%tke) on the box, levels 1..kz+1."""
    j1, j2, i1, i2 = box
    K, K1 = (1, kz), (1, kz + 1)

    def c(a, dj=0, di=0, k=K):
        return a.s((j1 + dj, j2 + dj), (i1 + di, i2 + di), k) if len(a.lb) == 3 else a.s((j1 + dj, j2 + dj), (i1 + di, i2 + di))[:, :, None]

    uavg1 = c(umc, 0, 1) + c(umc, 0, 0)
    uavg2 = c(umc, 1, 1) + c(umc, 1, 0)
    vavg1 = c(vmc, 1, 0) + c(vmc, 0, 0)
    vavg2 = c(vmc, 1, 1) + c(vmc, 0, 1)
    a = lambda dj=0, di=0: c(tkea, dj, di, K1)                     # noqa: E731
    b = lambda dj=0, di=0: c(tkeb, dj, di, K1)                     # noqa: E731
    ps = c(psa)
    ten = np.zeros(a().shape)
    t1 = twt.s((2, kz), 1)[None, None, :]
    t2 = twt.s((2, kz), 2)[None, None, :]
    full = lambda u: t1 * u[:, :, 1:] + t2 * u[:, :, :-1]          # noqa: E731  winds on full levels 2..kz
    lv = slice(1, kz)                                              # full levels 2..kz
    f, fjm, fjp, fim, fip = a()[:, :, lv], a(-1, 0)[:, :, lv], a(1, 0)[:, :, lv], a(0, -1)[:, :, lv], a(0, 1)[:, :, lv]
    if upstream:
        f1 = (0.5 * ul * (uavg2 + uavg1) / ps)[:, :, 1:]
        f2 = (0.5 * ul * (vavg2 + vavg1) / ps)[:, :, 1:]
        fx1, fx2 = (1.0 + f1) * fjm + (1.0 - f1) * f, (1.0 + f1) * f + (1.0 - f1) * fjp
        fy1, fy2 = (1.0 + f2) * fim + (1.0 - f2) * f, (1.0 + f2) * f + (1.0 - f2) * fip
    else:
        fx1, fx2, fy1, fy2 = fjm + f, f + fjp, fim + f, f + fip
    ten[:, :, lv] -= c(xmsf) * (full(uavg2) * fx2 - full(uavg1) * fx1 + full(vavg2) * fy2 - full(vavg1) * fy1)
    sv = c(svv, k=K1)
    tkeps = a() * ps
    fx = 0.5 * (sv[:, :, :-1] + sv[:, :, 1:]) * (tkeps[:, :, :-1] + tkeps[:, :, 1:])
    dd = dds.s((1, kz + 1))[None, None, :]
    ten[:, :, 1:] += fx * dd[:, :, 1:]
    ten[:, :, :-1] -= fx * dd[:, :, :-1]
    xk = c(xkcf, k=K1)
    if idiffu == 2:
        ten += nuk * xk * (4.0 / 6.0 * (b(1, 0) + b(-1, 0) + b(0, 1) + b(0, -1)) + 1.0 / 6.0 * (b(1, 1) + b(-1, -1) + b(-1, 1) + b(1, -1))
                           - 20.0 / 6.0 * b())
    else:
        ten -= nuk * xk * ((b(2, 0) + b(-2, 0) + b(0, 2) + b(0, -2)) - 4.0 * (b(1, 0) + b(-1, 0) + b(0, 1) + b(0, -1)) + 12.0 * b())
    fc = np.maximum(tkemin, b() + dt * (ten * c(rpsa) + c(tkephy, k=K1)))
    d = gnu2 * (fc + b() - 2.0 * a())
    return fc, fc.copy(), a() + d


# ---------------------------------------------------------------------------------------------------------
def advect_theta(th, tha, umc, vmc, psa, xmsf, svv, twt, xds, ul, t_extrema, idx, kz, upstream=True, stability_enhance=True):
    """hadvt (Main/mod_advection.F90:311-394) of the uncoupled potential temperature `th` with its local-extrema
    limiter — the i-direction test first, then the j-direction test, both on the flux left by the previous test —
    followed by vadv3d(ind = 0, nk = kz) (:743-753) of the coupled `tha`, as `advection` calls them with ithadv = 1
    (Main/mod_tendency.F90:1348-1357).  idx = (jci1, jci2, ici1, ici2).  Returns thten on that box."""
    j1, j2, i1, i2 = idx
    K = (1, kz)

    def c(a, dj=0, di=0, k=K):
        return a.s((j1 + dj, j2 + dj), (i1 + di, i2 + di), k) if len(a.lb) == 3 else a.s((j1 + dj, j2 + dj), (i1 + di, i2 + di))[:, :, None]

    uavg1 = c(umc, 0, 1) + c(umc, 0, 0)      # start_advect (:109-122)
    uavg2 = c(umc, 1, 1) + c(umc, 1, 0)
    vavg1 = c(vmc, 1, 0) + c(vmc, 0, 0)
    vavg2 = c(vmc, 1, 1) + c(vmc, 0, 1)
    f, fjm, fjp, fim, fip = c(th), c(th, -1, 0), c(th, 1, 0), c(th, 0, -1), c(th, 0, 1)
    ps = c(psa)
    if upstream:
        f1 = 0.5 * ul * (uavg2 + uavg1) / ps
        f2 = 0.5 * ul * (vavg2 + vavg1) / ps
        fx1, fx2 = (1.0 + f1) * fjm + (1.0 - f1) * f, (1.0 + f1) * f + (1.0 - f1) * fjp
        fy1, fy2 = (1.0 + f2) * fim + (1.0 - f2) * f, (1.0 + f2) * f + (1.0 - f2) * fip
    else:
        fx1, fx2, fy1, fy2 = fjm + f, f + fjp, fim + f, f + fip
    fg = -c(xmsf) * (uavg2 * fx2 - uavg1 * fx1 + vavg2 * fy2 - vavg1 * fy1)
    if stability_enhance:
        for lo, hi in ((fim, fip), (fjm, fjp)):        # i-direction first
            big = np.abs(hi + lo - 2.0 * f) / ps > t_extrema
            fg = np.where(big & (f > hi) & (f > lo), np.minimum(fg, 0.0), fg)
            fg = np.where(big & (f < hi) & (f < lo), np.maximum(fg, 0.0), fg)
    ten = fg.copy()
    fx = c(svv, k=(2, kz)) * (twt.s((2, kz), 1)[None, None, :] * c(tha, k=(2, kz)) + twt.s((2, kz), 2)[None, None, :] * c(tha, k=(1, kz - 1)))
    xd = xds.s((1, kz))[None, None, :]
    ten[:, :, :-1] -= fx * xd[:, :, :-1]
    ten[:, :, 1:] += fx * xd[:, :, 1:]
    return ten
