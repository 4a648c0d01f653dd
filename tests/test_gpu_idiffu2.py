"""idiffu = 2 (Main/mod_diffusion.F90:386-411,592-601,726-735,881-894) on the GPU through the C ABI
against the oracle, same cases as tests/test_idiffu2.py.  Tolerance 1e-11 relative L-inf after 6-8 steps, as for the
other option variants (tests/test_gpu_parity.py: CUDA libm pow/exp/log vs glibc in the last bits)."""
import numpy as np
import pytest

from regcm_b200 import capi, synth
from tests.helpers import make_pair, rel_errors

pytestmark = pytest.mark.gpu
TOL = 1e-11


@pytest.mark.parametrize("kw", [{}, {"iboudy": 1}, {"iboudy": 4}, {"ntr": 2}, {"idynamic": 1}, {"idynamic": 1, "ntr": 2}],
                         ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_gpu_idiffu2_vs_oracle(cuda_lib, kw):
    case, core, ref = make_pair(cuda_lib, "small", idiffu=2, **kw)
    core.step(8); ref.step(8)
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) <= TOL, errs
    for n in range(kw.get("ntr", 0)):
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            a, r = core.get_state(capi.F_CHI + n, lv), ref.get_state(capi.F_CHI + n, lv)
            assert np.abs(a - r).max() <= TOL * np.abs(r).max()
    core.close()


def test_gpu_idiffu2_staged_equal_plain(cuda_lib, monkeypatch):
    """The corner operands come out of the same shared-memory tiles in the staged kernels: bit-identical to the plain ones."""
    case = synth.make_case("small", idiffu=2)
    cores = []
    for on in ("1", "0"):
        for v in ("RCM_UV_TMA", "RCM_TQ_TMA", "RCM_PPW_TMA", "RCM_WFWD_TMA"):
            monkeypatch.setenv(v, on)
        c = capi.DynCore(cuda_lib, case.cfg); synth.load_case(c, case)
        cores.append(c)
    a, b = cores
    a.step(5); b.step(5)
    for f in range(7):
        for lv in (0, 1):
            assert np.array_equal(a.get_state(f, lv), b.get_state(f, lv)), (f, lv)
    a.close(); b.close()


def test_gpu_idiffu2_differs_from_idiffu1(cuda_lib):
    out = []
    for idiffu in (1, 2):
        case = synth.make_case("tiny", idiffu=idiffu)
        c = capi.DynCore(cuda_lib, case.cfg); synth.load_case(c, case)
        c.step(4)
        out.append([c.get_state(f) for f in range(7)])
        c.close()
    assert any(np.abs(a - b).max() > 1e-9 * np.abs(a).max() for a, b in zip(*out))


@pytest.mark.parametrize("kw", [{}, {"ntr": 2, "ibltyp": 2}, {"idynamic": 1}], ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_gpu_idiffu3_vs_oracle(cuda_lib, kw):
    """idiffu = 3 as this snapshot of the reference has it (one column, clamped stencil; tests/test_idiffu3.py)."""
    case, core, ref = make_pair(cuda_lib, "small", idiffu=3, **kw)
    if kw.get("ibltyp") == 2:
        from tests.test_tke import load_uw, uw_inputs
        for c in (core, ref):
            load_uw(c, case, *uw_inputs(case))
    core.step(8); ref.step(8)
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) <= TOL, errs
    extra = [capi.F_CHI + n for n in range(kw.get("ntr", 0))] + ([capi.F_TKE] if kw.get("ibltyp") == 2 else [])
    for fld in extra:
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            a, r = core.get_state(fld, lv), ref.get_state(fld, lv)
            assert np.abs(a - r).max() <= TOL * np.abs(r).max()
    core.close()
