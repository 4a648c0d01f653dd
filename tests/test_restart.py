"""Restart (the SAV path of Main/mod_savefile.F90:564-1264): a run that is stopped, whose state is
carried over through the C ABI exactly as `write_savefile` / `read_savefile` would carry it (prognostics of both time
levels, p*, dstor/hstor, tmask, the tracers, ktau and xbctime), and continued in a NEW handle must equal the
uninterrupted run bit for bit.  Host harness of the CUDA sources (tests/emu); the GPU version is tests/test_gpu_restart.py."""
import numpy as np
import pytest

from regcm_b200 import capi, synth


def fields_of(cfg):
    if cfg.idynamic == 3:
        return (capi.F_U, capi.F_V, capi.F_W, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PAI), (capi.LEVEL_A,)
    if cfg.idynamic == 1:
        f = (capi.F_U, capi.F_V, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PS)
    else:
        f = (capi.F_U, capi.F_V, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PP, capi.F_W)
    f = f + ((capi.F_TKE,) if cfg.ibltyp == 2 else ())
    return f + tuple(capi.F_CHI + n for n in range(cfg.ntr)), (capi.LEVEL_A, capi.LEVEL_B)


def save(core):
    """What the SAV file would hold."""
    flds, levels = fields_of(core.cfg)
    d = core.diag()
    sav = {"ktau": d.lcount, "xbctime": d.xbctime, "state": {(f, lv): core.get_state(f, lv) for f in flds for lv in levels}}
    if core.cfg.idynamic == 1:
        sav["dstor"], sav["hstor"] = core.get_restart(capi.R_DSTOR), core.get_restart(capi.R_HSTOR)
    if core.cfg.idynamic == 2:
        sav["tmask"] = core.get_restart(capi.R_TMASK)
    return sav


def restore(lib, case, sav, prep=None):
    core = capi.DynCore(lib, case.cfg)
    synth.load_case(core, case)                       # statics, the boundary records that bracket the restart date, init
    if prep:
        prep(core, case)                              # what the caller owns: kpbl (kpbl_io of the SAV file), physics tendencies
    for (f, lv), a in sav["state"].items():
        core.set_state(f, lv, a)
    if "dstor" in sav:
        core.set_restart(capi.R_DSTOR, sav["dstor"]); core.set_restart(capi.R_HSTOR, sav["hstor"])
    if "tmask" in sav:
        core.set_restart(capi.R_TMASK, sav["tmask"])
    core.set_clock(sav["ktau"], sav["xbctime"])
    return core


def check_restart(lib, name, n1, n2, prep=None, **kw):
    case = synth.make_case(name, **kw)
    ref = capi.DynCore(lib, case.cfg)
    synth.load_case(ref, case)
    if prep:
        prep(ref, case)
    ref.step(n1)
    sav = save(ref)
    ref.step(n2)
    new = restore(lib, case, sav, prep)
    assert new.diag().lcount == n1
    new.step(n2)
    flds, levels = fields_of(case.cfg)
    for f in flds:
        for lv in levels:
            np.testing.assert_array_equal(new.get_state(f, lv), ref.get_state(f, lv), err_msg=f"field {f} level {lv}")
    dn, dr = new.diag(), ref.diag()
    assert (dn.lcount, dn.xbctime, dn.dt) == (dr.lcount, dr.xbctime, dr.dt)
    ref.close(); new.close()


@pytest.mark.parametrize("name,kw", [("tiny", {}), ("tiny", {"ntr": 2}), ("tiny", {"idynamic": 1}), ("tiny", {"idynamic": 1, "nsplit": 3, "ntr": 1}),
                                     ("tinym", {})],
                         ids=["nh", "nh-tracers", "hydro", "hydro-nsplit3-tracers", "moloch"])
def test_restart_continues_bit_for_bit(emu_lib, name, kw):
    check_restart(emu_lib, name, 5, 6, **kw)


@pytest.mark.parametrize("idynamic", [2, 1])
def test_restart_with_tke(emu_lib, idynamic):
    """ibltyp = 2: atm1%tke / atm2%tke travel through get_state / set_state like the other prognostics
    (atm1_tke_io, atm2_tke_io of Main/mod_savefile.F90); kpbl and the physics tendency are the caller's."""
    from tests.test_tke import load_uw, uw_inputs
    check_restart(emu_lib, "tiny", 5, 6, prep=lambda c, case: load_uw(c, case, *uw_inputs(case)), idynamic=idynamic, ibltyp=2,
                  iuwvadv=1, ntr=1)


def test_restart_in_the_start_up_steps(emu_lib):
    """ktau = 1: the leapfrog still runs with dt = dtsec for one more step (Main/mod_tendency.F90:611-616)."""
    check_restart(emu_lib, "tiny", 1, 4)


def test_restart_arrays_are_checked(emu_lib):
    case = synth.make_case("tiny")
    core = capi.DynCore(emu_lib, case.cfg)
    synth.load_case(core, case)
    with pytest.raises(capi.RcmError):
        core.get_restart(capi.R_DSTOR)              # no split-explicit storage in the non-hydrostatic core
    with pytest.raises(capi.RcmError):
        core.set_clock(-1, 0.0)
    core.close()
