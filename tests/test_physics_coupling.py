"""Physics coupling boundary: rcm_dyn_step_pre_physics / get_slice / set_phys_tend /
step_post_physics.  The reference calls physical_parametrizations in the middle of `tend`
(Main/mod_tendency.F90:271): they read the slices of mkslice (Main/mod_slice.F90:167-308) and return
aten(:,:,:,pc_physic), summed at :285-314,331-335,404-411.  Kernel logic through the host harness vs the oracle."""
import numpy as np
import pytest

from regcm_b200 import capi
from tests.helpers import make_pair, prognostics, rel_errors

SLICES_ANY = [s for s in range(19)]


def _phys_tendencies(case, hydro):
    """Synthetic parametrization output, coupled with p* (~96 cb): a heating bubble, moistening near the surface, a
    condensate source aloft, a drag on the wind and (NH) small p', w sources."""
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    ps = 96.0
    J, I, K = np.meshgrid(np.arange(jx - 1), np.arange(iy - 1), np.arange(kz), indexing="ij")
    bub = np.exp(-(((J - jx / 2) / 4.0) ** 2 + ((I - iy / 2) / 3.0) ** 2)) * np.sin(np.pi * (K + 0.5) / kz)
    out = {capi.F_T: ps * 2.0e-3 * bub, capi.F_QV: ps * 1.0e-7 * bub * (K > kz // 2), capi.F_QC: ps * 2.0e-8 * bub * (K < kz // 2)}
    Jd, Id, Kd = np.meshgrid(np.arange(jx), np.arange(iy), np.arange(kz), indexing="ij")
    out[capi.F_U] = -ps * 1.0e-4 * np.sin(2 * np.pi * Jd / jx) * (Kd == kz - 1)
    out[capi.F_V] = ps * 5.0e-5 * np.cos(2 * np.pi * Id / iy) * (Kd >= kz - 2)
    if not hydro:
        out[capi.F_PP] = ps * 0.05 * bub
        Kw = np.arange(kz + 1)[None, None, :]
        out[capi.F_W] = ps * 1.0e-4 * np.exp(-(((J[:, :, :1] - jx / 2) / 4.0) ** 2 + ((I[:, :, :1] - iy / 2) / 3.0) ** 2)) * np.sin(np.pi * Kw / kz)
    return {f: np.asfortranarray(a) for f, a in out.items()}


@pytest.mark.parametrize("idyn", [2, 1])
def test_emu_split_step_without_physics_equals_plain_step(emu_lib, idyn):
    case, a, _ = make_pair(emu_lib, "tiny", idynamic=idyn)
    _, b, _ = make_pair(emu_lib, "tiny", idynamic=idyn)
    for n in range(5):
        a.step(1)
        b.step_pre_physics()
        b.get_slice(capi.SL_TB3D)
        b.step_post_physics()
    for f in prognostics(a):
        for lv in (capi.LEVEL_A, capi.LEVEL_B):
            np.testing.assert_array_equal(a.get_state(f, lv), b.get_state(f, lv), err_msg=capi.FIELD_NAMES[f])
    with pytest.raises(capi.RcmError):
        b.step_post_physics()        # without pre_physics


@pytest.mark.parametrize("idyn", [2, 1])
def test_emu_slices_match_oracle(emu_lib, idyn):
    case, core, ref = make_pair(emu_lib, "small", oracle_checked=True, idynamic=idyn)
    core.step(3); ref.step(3)
    core.step_pre_physics(); ref.step_pre_physics()
    for sid in SLICES_ANY:
        got, want = core.get_slice(sid), ref.get_slice(sid)
        sc = float(np.abs(want).max())
        assert sc > 0 or sid in (capi.SL_QCB3D,), capi.SLICE_NAMES[sid]
        err = float(np.abs(got - want).max()) / max(sc, 1e-300)
        # idynamic = 1: the two sides have run 3 steps on vertical-mode tables from two independent eigen-solvers
        # (1e-13 in the state); omega is a small residual of large terms and carries that at ~1e-11
        tol = 1e-12 if idyn == 2 else (1e-10 if sid in (capi.SL_WPX3D, capi.SL_WB3D) else 1e-11)
        assert err <= tol, (capi.SLICE_NAMES[sid], err)
    # physical sanity of the hand-over: pressures increase downwards, heights decrease, rh in its clamp
    pb, pf, zq = ref.get_slice(capi.SL_PB3D), ref.get_slice(capi.SL_PF3D), ref.get_slice(capi.SL_ZQ)
    assert np.all(np.diff(pb, axis=2) > 0) and np.all(np.diff(pf, axis=2) > 0) and np.all(np.diff(zq, axis=2) <= 0)
    rh = ref.get_slice(capi.SL_RHB3D)
    assert rh.min() >= 0.01 and rh.max() <= 1.01


@pytest.mark.parametrize("idyn", [2, 1])
def test_emu_step_with_physics_tendencies_matches_oracle(emu_lib, idyn):
    case, core, ref = make_pair(emu_lib, "small", idynamic=idyn)
    _, plain, _ = make_pair(emu_lib, "small", idynamic=idyn)
    ten = _phys_tendencies(case, idyn == 1)
    for n in range(4):
        for c in (core, ref):
            c.step_pre_physics()
            for f, a in ten.items():
                c.set_phys_tend(f, a * (1.0 + 0.25 * n))     # the parametrizations answer differently every step
            c.step_post_physics()
        plain.step(1)
    errs = rel_errors(core, ref)
    tol = 0.0 if idyn == 2 else 1e-11
    assert max(e[0] for e in errs.values()) <= tol, errs
    # and the tendencies did something: every prognostic moved away from the dynamics-only run
    moved = rel_errors(core, plain)
    assert min(moved[k][0] for k in ("ta", "ua", "va", "qva")) > 1e-7, moved
    # cleared tendencies give the dynamics-only step again
    core.clear_phys_tend(); ref.clear_phys_tend()
    core.step(2); ref.step(2)
    assert max(e[0] for e in rel_errors(core, ref).values()) <= tol


# ---------------------------------------------------------------------------- MOLOCH core (idynamic = 3)
# `moloch` calls mkslice and physical_parametrizations after the sound/advection loops and before `boundary`
# (Main/mod_moloch.F90:393-426); the parametrizations return uncoupled tendencies that advance t, u, v, qx by dtsec
# (:1548-1557), hydrometeors floored at 0 and qv at minqq.
def _mo_tendencies(case):
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    J, I, K = np.meshgrid(np.arange(jx - 1), np.arange(iy - 1), np.arange(kz), indexing="ij")
    bub = np.exp(-(((J - jx / 2) / 4.0) ** 2 + ((I - iy / 2) / 3.0) ** 2)) * np.sin(np.pi * (K + 0.5) / kz)
    out = {capi.F_T: 2.0e-3 * bub, capi.F_QV: 1.0e-7 * bub * (K > kz // 2) - 3.0e-6 * (K < 2), capi.F_QC: 2.0e-8 * bub - 1.0e-9}
    Ju, Iu, Ku = np.meshgrid(np.arange(jx), np.arange(iy - 1), np.arange(kz), indexing="ij")
    out[capi.F_U] = -1.0e-4 * np.sin(2 * np.pi * Ju / jx) * (Ku == kz - 1)
    Jv, Iv, Kv = np.meshgrid(np.arange(jx - 1), np.arange(iy), np.arange(kz), indexing="ij")
    out[capi.F_V] = 5.0e-5 * np.cos(2 * np.pi * Iv / iy) * (Kv >= kz - 2)
    return {f: np.asfortranarray(a) for f, a in out.items()}


def test_emu_moloch_split_step_without_physics_equals_plain_step(emu_lib):
    case, a, _ = make_pair(emu_lib, "tinym")
    _, b, _ = make_pair(emu_lib, "tinym")
    for n in range(4):
        a.step(1)
        b.step_pre_physics()
        b.get_slice(capi.SL_TB3D)
        b.step_post_physics()
    for f in prognostics(a):
        np.testing.assert_array_equal(a.get_state(f), b.get_state(f), err_msg=capi.FIELD_NAMES[f])
    with pytest.raises(capi.RcmError):
        b.step_post_physics()


def test_emu_moloch_slices_match_oracle(emu_lib):
    case, core, ref = make_pair(emu_lib, "smallm", oracle_checked=True)
    core.step(2); ref.step(2)
    core.step_pre_physics(); ref.step_pre_physics()
    for sid in SLICES_ANY:
        got, want = core.get_slice(sid), ref.get_slice(sid)
        sc = float(np.abs(want).max())
        assert sc > 0 or sid in (capi.SL_QCB3D,), capi.SLICE_NAMES[sid]
        assert float(np.abs(got - want).max()) <= 1e-12 * max(sc, 1e-300), capi.SLICE_NAMES[sid]
    pb, pf, zq, dzq = (ref.get_slice(s) for s in (capi.SL_PB3D, capi.SL_PF3D, capi.SL_ZQ, capi.SL_DZQ))
    assert np.all(np.diff(pb, axis=2) > 0) and np.all(np.diff(pf, axis=2) > 0) and np.all(np.diff(zq, axis=2) < 0) and np.all(dzq > 0)
    assert np.all(pf[:, :, 0] == 100.0) and np.array_equal(pf[:, :, -1], ref.get_slice(capi.SL_PS2D)[:, :, 0])
    za = ref.get_slice(capi.SL_ZA)
    assert np.all(za < zq[:, :, :-1]) and np.all(za > zq[:, :, 1:])        # half levels between the full levels


def test_emu_moloch_step_with_physics_tendencies_matches_oracle(emu_lib):
    case, core, ref = make_pair(emu_lib, "smallm")
    _, plain, _ = make_pair(emu_lib, "smallm")
    ten = _mo_tendencies(case)
    for n in range(4):
        for c in (core, ref):
            c.step_pre_physics()
            for f, a in ten.items():
                c.set_phys_tend(f, a * (1.0 + 0.25 * n))
            c.step_post_physics()
        plain.step(1)
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) <= 0.0, errs
    moved = rel_errors(core, plain)
    assert min(moved[k][0] for k in ("ta", "ua", "va", "qva")) > 1e-7, moved
    qv, qc = ref.get_state(capi.F_QV), ref.get_state(capi.F_QC)
    assert qc[1:-1, 1:-1].min() >= 0.0 and qv[1:-1, 1:-1].min() >= 1.0e-8 and (qc[1:-1, 1:-1] == 0.0).any()   # the floors were reached
    core.clear_phys_tend(); ref.clear_phys_tend()
    core.step(2); ref.step(2)
    assert max(e[0] for e in rel_errors(core, ref).values()) <= 0.0
