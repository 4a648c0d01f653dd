"""On-disk formats (regcm_b200/ncio.py): DOMAIN, ICBC and SAV files with the reference's names, dimensions,
units and checks (Main/mod_ncio.F90:81-381,583-1128, Share/mod_domain.F90:263-405, Main/mod_savefile.F90:564-1264).

No file written by the reference exists in this image, so the pins are: the names / dimension order / unit conventions asserted
here against what the Fortran reads, a run fed from files equal to the run fed from arrays bit for bit, and a restart through
a SAV file equal to the uninterrupted run bit for bit.  Host harness of the CUDA sources."""
import numpy as np
import pytest
from scipy.io import netcdf_file

from regcm_b200 import capi, ncio, synth

START = (1990, 1, 31, 22)       # so that the third record falls into the next month's file


def _hours(date):
    import datetime as dt
    return (dt.datetime(*date) - dt.datetime(1949, 12, 1)).total_seconds() / 3600.0


def _write_case(tmp_path, case, dtype="f8"):
    """DOMAIN + two ICBC files (records 1, 2 in the file of the start date, record 3 in the file of the next month)."""
    cfg = case.cfg
    ncio.write_domain(str(tmp_path / "case_DOMAIN000.nc"), cfg, case.statics)
    h0 = _hours(START)
    recs = [case.bdy[s] for s in (capi.BDY_B0, capi.BDY_B1, capi.BDY_NEXT)]
    ncio.write_icbc(ncio.icbc_name(str(tmp_path), "case", START), cfg, [h0, h0 + cfg.ibdyfrq], recs[:2], dtype=dtype)
    nxt = ncio.timeval2date(h0 + 2 * cfg.ibdyfrq, "hours since 1949-12-01 00:00:00 UTC")
    assert nxt[:2] == (1990, 2)                      # the third record lies in the next month's file
    ncio.write_icbc(ncio.icbc_name(str(tmp_path), "case", ncio.monfirst(nxt)), cfg, [h0 + 2 * cfg.ibdyfrq, h0 + 3 * cfg.ibdyfrq], [recs[2], recs[2]], dtype=dtype)
    return nxt


def _core_from_files(lib, tmp_path, cfg, nxt):
    core = capi.DynCore(lib, cfg)
    ncio.load_domain(core, ncio.read_domain_info(str(tmp_path / "case_DOMAIN000.nc"), cfg))
    rd = ncio.IcbcReader(str(tmp_path), "case", cfg, START)
    h0 = _hours(START)
    d1 = ncio.timeval2date(h0 + cfg.ibdyfrq, "hours since 1949-12-01 00:00:00 UTC")
    rd.feed(core, capi.BDY_B0, START)
    rd.feed(core, capi.BDY_B1, d1)
    rd.feed(core, capi.BDY_NEXT, nxt)                 # opens the February file (bdyin, Main/mod_bdycod.F90:669-690)
    core.init()
    rd.close()
    return core


@pytest.mark.parametrize("name,kw", [("tiny", {}), ("tiny", {"idynamic": 1}), ("tinym", {})], ids=["nh", "hydro", "moloch"])
def test_run_from_files_equals_run_from_arrays(emu_lib, tmp_path, name, kw):
    case = synth.make_case(name, ibdyfrq=1, **kw)
    nxt = _write_case(tmp_path, case)
    a = _core_from_files(emu_lib, tmp_path, case.cfg, nxt)
    b = capi.DynCore(emu_lib, case.cfg); synth.load_case(b, case)
    steps = int(round(3600.0 / case.cfg.dt)) + 3       # across the change of boundary record
    a.step(steps); b.step(steps)
    flds = (capi.F_U, capi.F_V, capi.F_T, capi.F_QV) + ((capi.F_PP, capi.F_W) if case.cfg.idynamic == 2 else ())
    for f in flds:
        np.testing.assert_array_equal(a.get_state(f), b.get_state(f), err_msg=capi.FIELD_NAMES[f])
    a.close(); b.close()


def test_file_layout_is_the_reference_layout(tmp_path):
    """What the Fortran reader relies on: dimension names and order, `kz` = kz + 1 in DOMAIN and kz in ICBC, ps in hPa, `w`
    on levels 2..kz+1 with `wtop` apart, the time coordinate in hours with units and calendar."""
    case = synth.make_case("tiny", ibdyfrq=6)
    cfg = case.cfg
    _write_case(tmp_path, case, dtype="f4")
    d = netcdf_file(str(tmp_path / "case_DOMAIN000.nc"), "r", mmap=False)
    assert d.dimensions["jx"] == cfg.jx and d.dimensions["iy"] == cfg.iy and d.dimensions["kz"] == cfg.kz + 1
    assert d.variables["xmap"].dimensions == ("iy", "jx") and d.variables["kz"].shape == (cfg.kz + 1,)
    assert abs(float(d.variables["ptop"][0]) * 0.1 - cfg.ptop) < 1e-12 and abs(d.grid_size_in_meters - cfg.ds * 1000.0) < 1e-9
    for v in ("xlat", "xlon", "dlat", "xmap", "dmap", "topo", "coriol"):
        assert v in d.variables
    d.close()
    f = netcdf_file(ncio.icbc_name(str(tmp_path), "case", START), "r", mmap=False)
    assert f.dimensions["kz"] == cfg.kz and f.dimensions["time"] is None
    assert f.variables["u"].dimensions == ("time", "kz", "iy", "jx") and f.variables["ps"].dimensions == ("time", "iy", "jx")
    assert f.variables["w"].shape[1] == cfg.kz and f.variables["wtop"].dimensions == ("time", "iy", "jx")
    assert f.variables["time"].units.decode().startswith("hours since") and f.variables["time"].calendar == b"gregorian"
    assert int(round(f.variables["time"][1] - f.variables["time"][0])) == cfg.ibdyfrq
    f.close()
    # single-precision files (what PreProc writes): the reader hands the library the rounded values
    rec = ncio.IcbcFile(ncio.icbc_name(str(tmp_path), "case", START), cfg).read(1)
    want = case.bdy[capi.BDY_B0]
    np.testing.assert_array_equal(rec[capi.F_T], want[capi.F_T].astype(np.float32).astype(np.float64))
    np.testing.assert_array_equal(rec[capi.F_W][:, :, 0], want[capi.F_W][:, :, 0].astype(np.float32).astype(np.float64))
    assert rec[capi.F_U].shape == (cfg.jx, cfg.iy, cfg.kz) and rec[capi.F_W].shape == (cfg.jx - 1, cfg.iy - 1, cfg.kz + 1)


def test_checks_of_the_reference(tmp_path):
    case = synth.make_case("tiny", ibdyfrq=6)
    _write_case(tmp_path, case)
    dom = str(tmp_path / "case_DOMAIN000.nc")
    for member, value in (("jx", case.cfg.jx + 1), ("kz", case.cfg.kz + 1), ("ds", case.cfg.ds * 2), ("ptop", case.cfg.ptop + 1.0)):
        cfg = synth.make_case("tiny", ibdyfrq=6).cfg
        setattr(cfg, member, value)
        with pytest.raises(ncio.NcioError):
            ncio.read_domain_info(dom, cfg)
    cfg = synth.make_case("tiny", ibdyfrq=3).cfg                    # spacing of the records /= ibdyfrq (mod_ncio.F90:632-640)
    with pytest.raises(ncio.NcioError, match="ibdyfrq"):
        ncio.IcbcFile(ncio.icbc_name(str(tmp_path), "case", START), cfg)
    f = ncio.IcbcFile(ncio.icbc_name(str(tmp_path), "case", START), case.cfg)
    assert f.search(START) == 1 and f.search((1990, 2, 1, 4)) == 2 and f.search((1990, 2, 1, 10)) == -1 and f.search((1990, 1, 1, 0)) == -1
    with pytest.raises(ncio.NcioError):
        f.search((1990, 2, 1, 1))                                    # inside the file's range but not a record
    f.close()
    with pytest.raises(ncio.NcioError):
        ncio.IcbcReader(str(tmp_path), "nosuch", case.cfg, START).record(START)


def test_calendars():
    u = "hours since 1949-12-01 00:00:00 UTC"
    assert ncio.timeval2date(0, u) == (1949, 12, 1, 0)
    assert ncio.timeval2date(24 * 31 + 5, u) == (1950, 1, 1, 5)
    assert ncio.timeval2date(24 * (31 + 31 + 28), u, "gregorian") == (1950, 3, 1, 0)
    assert ncio.timeval2date(24 * (31 + 365 + 31 + 28 + 31 + 30 + 31 + 30 + 31 + 31 + 30 + 31 + 30 + 31 + 31 + 29), u, "gregorian") == (1952, 3, 1, 0)
    assert ncio.timeval2date(24 * (31 + 365 + 365 + 31 + 28), u, "noleap") == (1952, 3, 1, 0)       # no 29 February
    assert ncio.timeval2date(24 * 30 * 3, u, "360_day") == (1950, 3, 1, 0)
    assert ncio.timeval2date(24 * 29 + 23, "hours since 1949-12-01 00:00:00 UTC", "360_day") == (1949, 12, 30, 23)
    assert ncio.toint10((1990, 1, 31, 22)) == 1990013122 and ncio.monfirst((1990, 1, 31, 22)) == (1990, 1, 1, 0)


@pytest.mark.parametrize("name,kw", [("tiny", {}), ("tiny", {"idynamic": 1}), ("tiny", {"ntr": 2, "ibltyp": 2}), ("tinym", {})],
                         ids=["nh", "hydro", "nh-tracers-tke", "moloch"])
def test_restart_through_a_sav_file(emu_lib, tmp_path, name, kw):
    from tests.test_restart import fields_of
    case = synth.make_case(name, **kw)
    ref = capi.DynCore(emu_lib, case.cfg); synth.load_case(ref, case)
    ref.step(5)
    path = ncio.sav_name(str(tmp_path), "case", (1990, 2, 1, 0))
    ncio.write_savefile(path, ref, (1990, 2, 1, 0))
    ref.step(4)
    new = capi.DynCore(emu_lib, case.cfg); synth.load_case(new, case)
    with pytest.raises(ncio.NcioError):
        ncio.read_savefile(path, new, date=(1990, 2, 1, 6))           # idatex /= the date asked for
    ncio.read_savefile(path, new, date=(1990, 2, 1, 0))
    assert new.diag().lcount == 5
    new.step(4)
    flds, levels = fields_of(case.cfg)
    for f in flds:
        for lv in levels:
            np.testing.assert_array_equal(new.get_state(f, lv), ref.get_state(f, lv), err_msg=f"{f} level {lv}")
    f = netcdf_file(path, "r", mmap=False)
    first = "atm_u" if case.cfg.idynamic == 3 else "atm1_u"
    assert f.variables[first].dimensions == (("khalf", "icross", "jdot") if case.cfg.idynamic == 3 else ("khalf", "idot", "jdot"))
    qx = "atm_qx" if case.cfg.idynamic == 3 else "atm2_qx"
    assert f.variables[qx].dimensions == ("nqx", "khalf", "icross", "jcross") and int(f.idatex) == 1990020100
    if case.cfg.idynamic == 1:
        assert f.variables["dstor"].dimensions == ("nsplit", "idot", "jdot") and "psa" in f.variables and "psb" in f.variables
    if case.cfg.idynamic == 2:
        assert f.variables["tmask"].shape == (13, 13)
    if case.cfg.ntr:
        assert f.variables["chia"].dimensions == ("ntr", "khalf", "icross", "jcross") and "atm1_tke" in f.variables
    f.close(); ref.close(); new.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw", [("small", {}), ("small", {"idynamic": 1}), ("smallm", {})], ids=["nh", "hydro", "moloch"])
def test_gpu_files_in_sav_out(cuda_lib, tmp_path, name, kw):
    """The same two paths on the GPU library: a run fed from DOMAIN / ICBC files equals the run fed from arrays, and a restart
    through a SAV file equals the uninterrupted run, bit for bit."""
    from tests.test_restart import fields_of
    case = synth.make_case(name, ibdyfrq=1, **kw)
    nxt = _write_case(tmp_path, case)
    a = _core_from_files(cuda_lib, tmp_path, case.cfg, nxt)
    b = capi.DynCore(cuda_lib, case.cfg); synth.load_case(b, case)
    a.step(28); b.step(28)                            # the record change (step 30 at dt = 120 s) falls after the restart
    path = ncio.sav_name(str(tmp_path), "case", START)
    ncio.write_savefile(path, a, START)
    a.step(5); b.step(5)
    c = capi.DynCore(cuda_lib, case.cfg); synth.load_case(c, case)
    ncio.read_savefile(path, c, date=START)
    c.step(5)
    flds, levels = fields_of(case.cfg)
    for f in flds:
        for lv in levels:
            np.testing.assert_array_equal(a.get_state(f, lv), b.get_state(f, lv), err_msg=f"files vs arrays: {f} level {lv}")
            np.testing.assert_array_equal(c.get_state(f, lv), b.get_state(f, lv), err_msg=f"restart: {f} level {lv}")
    a.close(); b.close(); c.close()


# ---------------------------------------------------------------------------- the time loop with files on both sides
def _write_long_case(tmp_path, case):
    """Five hourly records: 22, 23 h of 31 January in the file of the start date, 0, 1, 2 h of 1 February in the next one."""
    cfg = case.cfg
    ncio.write_domain(str(tmp_path / "case_DOMAIN000.nc"), cfg, case.statics)
    r = [case.bdy[s] for s in (capi.BDY_B0, capi.BDY_B1, capi.BDY_NEXT)]
    mix = lambda a, b, w: {f: np.asfortranarray((1 - w) * a[f] + w * b[f]) for f in a}
    recs = [r[0], r[1], r[2], mix(r[1], r[2], 0.4), mix(r[0], r[2], 0.7)]
    h0 = _hours(START)
    ncio.write_icbc(ncio.icbc_name(str(tmp_path), "case", START), cfg, [h0, h0 + 1], recs[:2])
    ncio.write_icbc(ncio.icbc_name(str(tmp_path), "case", (1990, 2, 1, 0)), cfg, [h0 + 2, h0 + 3, h0 + 4], recs[2:])
    return recs


@pytest.mark.parametrize("name,kw", [("tiny", {}), ("tiny", {"idynamic": 1}), ("tinym", {})], ids=["nh", "hydro", "moloch"])
def test_time_loop_with_files_and_restart(emu_lib, tmp_path, name, kw):
    """driver.Run = RCM_initialize / RCM_run for a dynamics-only run: three boundary intervals across the month change of the
    ICBC files equal the same run fed from arrays record by record; stopped at a SAV date and continued from the SAV file in
    a new process image (ifrest), it equals the uninterrupted run bit for bit."""
    _time_loop(emu_lib, tmp_path, name, kw)


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw", [("small", {}), ("small", {"idynamic": 1}), ("smallm", {})], ids=["nh", "hydro", "moloch"])
def test_gpu_time_loop_with_files_and_restart(cuda_lib, tmp_path, name, kw):
    """The same on the GPU library: here the read-ahead record really travels on the staging stream while the steps run."""
    _time_loop(cuda_lib, tmp_path, name, kw)


def _time_loop(emu_lib, tmp_path, name, kw):
    from regcm_b200 import driver
    from tests.test_restart import fields_of
    case = synth.make_case(name, ibdyfrq=1, **kw)
    cfg = case.cfg
    recs = _write_long_case(tmp_path, case)
    spr = int(round(3600.0 / cfg.dt))
    total = 3 * spr + 2
    run = driver.Run(emu_lib, cfg, str(tmp_path), str(tmp_path), "case", START, dirout=str(tmp_path))
    run.run(total, sav_dates=[(1990, 2, 1, 0)])
    # arrays, record by record
    ref = capi.DynCore(emu_lib, cfg)
    for sid, a in case.statics.items():
        ref.set_static(sid, a)
    for slot, rec in ((capi.BDY_B0, recs[0]), (capi.BDY_B1, recs[1]), (capi.BDY_NEXT, recs[2])):
        for fid in ncio.bdy_fields(cfg):
            ref.set_bdy(fid, slot, rec[fid])
    ref.init()
    ref.step(spr)
    for k in (3, 4):
        for fid in ncio.bdy_fields(cfg):
            ref.set_bdy(fid, capi.BDY_NEXT, recs[k][fid])
        ref.step(spr if k == 3 else spr + 2)
    flds, levels = fields_of(cfg)
    for f in flds:
        for lv in levels:
            np.testing.assert_array_equal(run.core.get_state(f, lv), ref.get_state(f, lv), err_msg=f"loop: {f} level {lv}")
    # restart from the SAV file written two intervals in
    again = driver.Run(emu_lib, cfg, str(tmp_path), str(tmp_path), "case", START, dirout=str(tmp_path), restart_from=(1990, 2, 1, 0))
    assert again.lcount == 2 * spr
    again.run(total - 2 * spr)
    for f in flds:
        for lv in levels:
            np.testing.assert_array_equal(again.core.get_state(f, lv), ref.get_state(f, lv), err_msg=f"restart: {f} level {lv}")
    with pytest.raises(ncio.NcioError):
        again.run(2 * spr)                              # the boundary data end before the run does
    run.close(); again.close(); ref.close()


@pytest.mark.parametrize("name,idynamic,world,layout", [("tiny", 2, 4, (2, 2)), ("tiny", 1, 2, (2, 1)), ("tinym", 3, 2, (1, 2))], ids=["nh-2x2", "hydro-2x1", "moloch-1x2"])
def test_ranks_reading_the_files_themselves(emu_lib, tmp_path, name, idynamic, world, layout):
    """Decomposed run, every rank reading DOMAIN and ICBC itself and handing the library the global arrays (it keeps its
    window): the assembled result equals the single rank fed from arrays."""
    import os
    from tests.test_multirank_emu import run_ranks
    case = synth.make_case(name, ibdyfrq=1, idynamic=idynamic)
    _write_case(tmp_path, case)
    steps = 5
    run_ranks(world, tmp_path, ["--case", name, "--idynamic", str(idynamic), "--ibdyfrq", "1", "--steps", str(steps), "--cpus-j", str(layout[0]),
                                "--cpus-i", str(layout[1]), "--from-files", str(tmp_path)])
    ref = capi.DynCore(emu_lib, case.cfg); synth.load_case(ref, case); ref.step(steps)
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    for fld in (capi.F_T, capi.F_QV):
        glob = np.full((jx - 1, iy - 1, kz), np.nan)
        for r in range(world):
            d = np.load(os.path.join(tmp_path, f"rank{r}.npz"))
            lb, a = d[f"lb{fld}"], d[f"f{fld}a"]
            glob[lb[0] - 1:lb[0] - 1 + a.shape[0], lb[1] - 1:lb[1] - 1 + a.shape[1], :] = a
        want = ref.get_state(fld, 0)
        assert not np.isnan(glob).any()
        assert np.abs(glob - want).max() <= (0.0 if idynamic == 2 else 1e-11) * np.abs(want).max(), fld
    ref.close()
