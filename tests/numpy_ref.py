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
