"""Kernel LOGIC check without a GPU: the same .cu sources, compiled with g++
against tests/emu/cuda_emu.h and executed thread by thread, must reproduce the
oracle bit for bit (no libm difference on the host).  This is test
infrastructure, not a product path: the shipped library is the nvcc build."""
import numpy as np
import pytest

from regcm_b200 import capi
from tests.helpers import make_pair, rel_errors, smooth_qc, work_errors

VARIANTS = [
    {}, {"iboudy": 1}, {"iboudy": 0}, {"iboudy": 2}, {"iboudy": 3}, {"iboudy": 4}, {"ifupr": 0}, {"ifrayd": 0}, {"ipptls": 0}, {"diffu_hgtf": 1},
    {"upstream_mode": 0}, {"stability_enhance": 0}, {"flat": True, "unit_map": True}, {"noise": False},
]


@pytest.mark.parametrize("kw", VARIANTS, ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_emu_bit_exact(emu_lib, kw):
    case, core, ref = make_pair(emu_lib, "tiny", oracle_checked=True, **kw)
    core.step(1); ref.step(1)
    assert max(work_errors(core, ref).values()) == 0.0
    core.step(7); ref.step(7)
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) == 0.0, errs
    assert core.diag().max_cfl == ref.diag().max_cfl and core.diag().istep == ref.diag().istep


def test_emu_boundary_record_change(emu_lib):
    case, core, ref = make_pair(emu_lib, "tiny", ibdyfrq=1)
    core.step(35); ref.step(35)
    assert core.diag().xbctime == ref.diag().xbctime == 6 * 120.0
    assert max(e[0] for e in rel_errors(core, ref).values()) == 0.0


def test_emu_c2_shape(emu_lib):
    case, core, ref = make_pair(emu_lib, "C2")
    core.step(3); ref.step(3)
    assert max(e[0] for e in rel_errors(core, ref).values()) == 0.0


def test_emu_cloud_water_path(emu_lib):
    case, core, ref = make_pair(emu_lib, "tiny")
    qc = smooth_qc(case)
    for c in (core, ref):
        c.set_state(capi.F_QC, capi.LEVEL_A, qc)
        c.set_state(capi.F_QC, capi.LEVEL_B, qc)
    core.step(8); ref.step(8)
    assert max(e[0] for e in rel_errors(core, ref).values()) == 0.0
    assert np.abs(ref.get_state(capi.F_QC)).max() > 0 and ref.diag().neg_q_points == 0
