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
    uten, vten, ppten, wten, s_estore, s_astore); p: scalars.  Returns (istep, max cfl of the last sub-step)."""
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
        cfl = t["_work"]["cfl"]
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
