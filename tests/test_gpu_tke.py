"""UW-PBL coupling of the transport (ibltyp = 2: TKE, vadv4d ind = 3) on the GPU through the C ABI against
the oracle, same cases as tests/test_tke.py.  TKE alone: 1e-11 relative L-inf after 8 steps, as for the other prognostics
(tests/test_gpu_parity.py).  With vadv4d ind = 3 the reference's flux selection is a rounding-level tie in about half of
the columns (tests/test_tke.py, module docstring), so a free-running comparison across libm implementations measures
that tie, not the kernel: the GPU is compared one step at a time along the oracle's trajectory, from identical input."""
import numpy as np
import pytest

from regcm_b200 import capi
from tests.helpers import make_pair, rel_errors
from tests.test_tke import load_uw, uw_inputs

pytestmark = pytest.mark.gpu
TOL = 1e-11


def _compare(core, ref, tol, ntr):
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) <= tol, errs
    for fld in (capi.F_TKE,) + tuple(capi.F_CHI + n for n in range(ntr)):
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            a, r = core.get_state(fld, lv), ref.get_state(fld, lv)
            assert np.abs(a - r).max() <= tol * np.abs(r).max(), (fld, lv)


def _sync_state(core, ref):
    from tests.test_restart import fields_of
    flds, levels = fields_of(ref.cfg)
    for f in flds:
        for lv in levels:
            core.set_state(f, lv, ref.get_state(f, lv))


@pytest.mark.parametrize("idynamic", [2, 1])
def test_gpu_ind3_along_the_oracle_trajectory(cuda_lib, idynamic):
    case, core, ref = make_pair(cuda_lib, "small", ibltyp=2, iuwvadv=1, ntr=2, idynamic=idynamic)
    inp = uw_inputs(case)
    for c in (core, ref):
        load_uw(c, case, *inp)
    for n in range(6):
        if n:
            _sync_state(core, ref)
        core.step(1); ref.step(1)
        _compare(core, ref, 1e-12, 2)
    # and the reshaping was active: the same six steps without it end somewhere else
    case0, core0, ref0 = make_pair(cuda_lib, "small", ibltyp=2, iuwvadv=0, ntr=2, idynamic=idynamic)
    load_uw(ref0, case0, *inp)
    ref0.step(6)
    qa, qb = ref.get_state(capi.F_QC), ref0.get_state(capi.F_QC)
    assert np.abs(qa - qb).max() > 1e-4 * np.abs(qa).max()
    core.close(); core0.close()


@pytest.mark.parametrize("kw", [{}, {"iboudy": 1}, {"iboudy": 4}, {"idiffu": 2}, {"idynamic": 1}],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_gpu_tke_vs_oracle(cuda_lib, kw):
    case, core, ref = make_pair(cuda_lib, "small", ibltyp=2, **kw)
    inp = uw_inputs(case)
    for c in (core, ref):
        load_uw(c, case, *inp)
        c.step(8)
    _compare(core, ref, TOL, kw.get("ntr", 0))
    core.close()


def test_gpu_tke_record_change(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "tiny", ibltyp=2, ntr=1, ibdyfrq=1)
    inp = uw_inputs(case)
    for c in (core, ref):
        load_uw(c, case, *inp)
        c.step(34)
    _compare(core, ref, 1e-10, 1)
    core.close()


def test_gpu_set_kpbl_is_checked(cuda_lib):
    case, core, ref = make_pair(cuda_lib, "tiny")
    with pytest.raises(capi.RcmError):
        core.set_kpbl(np.zeros((4, 4), np.int32))          # ibltyp /= 2: no kpbl
    core.close()
    case, core, ref = make_pair(cuda_lib, "tiny", ibltyp=2, iuwvadv=1)
    with pytest.raises(capi.RcmError):
        core.set_kpbl(np.full((4, 4), case.cfg.kz + 1, np.int32))   # the reference calls fatal
    core.close()
