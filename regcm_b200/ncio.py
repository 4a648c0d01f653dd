"""On-disk formats either side of the dyn-core: the DOMAIN file, the ICBC files and the SAV restart file
of the reference, read and written on the host and handed to / taken from the library through the C ABI.

What it mirrors (names, dimensions, units and checks are the reference's; the arithmetic that follows — ps*0.1 - ptop,
coupling with p*, the time interpolation — is the library's `init_bdy` / `bdyin`):
  read_domain_info     Main/mod_ncio.F90:81-381 (variables) + check_domain, Share/mod_domain.F90:263-405 (dims jx, iy, kz;
                       `ptop` within 0.001 cb; global attribute `grid_size_in_meters` within 1 m of ds)
  IcbcFile             open_icbc / icbc_search / read_icbc / close_icbc, Main/mod_ncio.F90:583-687,548-571,835-1128: file
                       `<domname>_ICBC.<YYYYMMDDHH>.nc`, dimension `time`, its `units` / `calendar`, the spacing of the
                       records checked against ibdyfrq, variables ps, ts, u, v, t, qv (+ pp, w, wtop for idynamic = 2)
                       stored (time, kz, iy, jx); `w` fills levels 2..kz+1 and `wtop` level 1 (:953-964)
  write_savefile /     Main/mod_savefile.F90:764-1264 / :564-762: dimensions jcross, icross, jdot, idot, khalf, kfull,
  read_savefile        nsplit, nqx, ntr, ikern; variables atm1_u … atm2_pp, psa, psb, atm1_tke, atm2_tke, chia, chib, dstor,
                       hstor, tmask (MOLOCH: atm_u, atm_v, atm_w, atm_t, atm_pai, atm_qx, ps); global attributes idatex,
                       calendar.  The surface / physics state the reference also stores belongs to its physics packages and
                       is not the library's to write.
The writers of the DOMAIN and ICBC files (what PreProc/Terrain and PreProc/ICBC produce) exist so that a case can be
exchanged with the reference and for the tests; their layout is also the one the reference's own Python writer of idealised
cases produces (PreProc/CRM/initialize_icbc.py:75-252: dimensions jx, iy, kz, time (unlimited); ps [hPa] and ts on
(time, iy, jx); t, qv, u, v, w, pp on (time, kz, iy, jx); wtop on (time, iy, jx); time in "hours since 1949-12-01 00:00:00
UTC", calendar gregorian; `ptop` [hPa]; global attribute grid_size_in_meters).

Container: NetCDF classic / 64-bit offset through scipy.io.netcdf_file — what the reference writes unless it is built
with NETCDF4_HDF5 (Share/mod_dynparam.F90:377-389).  There is no HDF5 library in this image: NetCDF-4 files need
`nccopy -k classic` first.  Host side in Python because the image has no NetCDF C library to bind a compiled reader to."""
from __future__ import annotations

import datetime as _dt
import os
import re

import numpy as np
from scipy.io import netcdf_file

from . import capi

# DOMAIN variable -> static id (Main/mod_ncio.F90:138-191; idynamic = 3 reads the u/v-point variants, :146-159)
_DOMAIN_VARS = {capi.S_XLAT: "xlat", capi.S_XLON: "xlon", capi.S_DLAT: "dlat", capi.S_MSFX: "xmap", capi.S_MSFD: "dmap",
                capi.S_HT: "topo", capi.S_CORIOL: "coriol"}
_DOMAIN_VARS_MOLOCH = {capi.S_XLAT: "xlat", capi.S_XLON: "xlon", capi.S_ULAT: "ulat", capi.S_VLAT: "vlat", capi.S_MSFU: "umap",
                       capi.S_MSFV: "vmap", capi.S_MSFX: "xmap", capi.S_HT: "topo"}
_CROSS_STATICS = (capi.S_XLAT, capi.S_XLON, capi.S_MSFX, capi.S_HT)


class NcioError(RuntimeError):
    """What the reference reports through `fatal` (a missing variable, a mismatch between file and namelist)."""


def _static_shape(cfg, sid):
    jx, iy = cfg.jx, cfg.iy
    if sid in _CROSS_STATICS:
        return (jx - 1, iy - 1)
    if sid in (capi.S_MSFU, capi.S_ULAT):
        return (jx, iy - 1)
    if sid in (capi.S_MSFV, capi.S_VLAT):
        return (jx - 1, iy)
    return (jx, iy)


def _get2d(nc, name, what):
    if name not in nc.variables:
        raise NcioError(f"variable {name} miss ({what})")
    return np.asfortranarray(np.array(nc.variables[name][:], dtype=np.float64).T)      # file (iy, jx) -> (jx, iy)


# ---------------------------------------------------------------------------------------------------------- DOMAIN
def check_domain(nc, cfg, lmod=False, what="DOMAIN FILE"):
    """Share/mod_domain.F90:263-405: the file must describe the grid of the namelist.  Dimension `kz` counts the full sigma
    levels (kz + 1) in a DOMAIN file and the half levels (kz) in a model input file such as ICBC (`lmod`, :323-324)."""
    for dim, want in (("jx", cfg.jx), ("iy", cfg.iy), ("kz", cfg.kz if lmod else cfg.kz + 1)):
        if dim not in nc.dimensions:
            raise NcioError(f"Dimension {dim} miss ({what})")
        have = nc.dimensions[dim]
        have = have if have is not None else nc.variables[dim].shape[0]
        if have != want:
            raise NcioError(f"Mismatch: {dim.upper()} in {what} {have} /= namelist {want}")
    if "ptop" in nc.variables:
        ptsp = float(np.array(nc.variables["ptop"][...]).reshape(-1)[0])
        if abs(ptsp * 0.1 - cfg.ptop) > 0.001:
            raise NcioError(f"Mismatch: PTOP in {what} {ptsp} hPa /= namelist {cfg.ptop} cb")
    dsx = getattr(nc, "grid_size_in_meters", None)
    if dsx is None:
        raise NcioError(f"attribute grid_size_in_meters miss ({what})")
    if abs(float(dsx) * 1.0e-3 - cfg.ds) > 0.001:
        raise NcioError(f"Mismatch: DS in {what} {float(dsx) / 1000.0} km /= namelist {cfg.ds} km")


def read_domain_info(path, cfg):
    """-> {static id: array} ready for DynCore.set_static (sigma from the coordinate variable `kz`, else from `sigma`, as
    the reference tries them, :137-141)."""
    nc = netcdf_file(path, "r", mmap=False)
    try:
        check_domain(nc, cfg)
        out = {}
        sig = nc.variables["kz"] if "kz" in nc.variables else nc.variables.get("sigma")
        if sig is None:
            raise NcioError("variable sigma miss (DOMAIN FILE)")
        sigma = np.array(sig[:], dtype=np.float64)
        if sigma.shape != (cfg.kz + 1,):
            raise NcioError(f"sigma has {sigma.shape[0]} values, kz + 1 = {cfg.kz + 1} expected")
        out[capi.S_SIGMA] = sigma
        for sid, name in (_DOMAIN_VARS_MOLOCH if cfg.idynamic == 3 else _DOMAIN_VARS).items():
            a = _get2d(nc, name, "DOMAIN FILE")
            n1, n2 = _static_shape(cfg, sid)
            out[sid] = np.asfortranarray(a[:n1, :n2])
        return out
    finally:
        nc.close()


def _put_ptop(nc, cfg):
    """`ptop` [hPa] as a one-element variable: scipy's writer places a 0-d variable behind the record section, on top of the
    second record; nf90_get_var into a scalar reads the first element of either form (Share/mod_domain.F90:334-338)."""
    nc.createDimension("one", 1)
    v = nc.createVariable("ptop", "f8", ("one",)); v[:] = cfg.ptop * 10.0; v.units = "hPa"


def load_domain(core, dom):
    for sid, a in dom.items():
        core.set_static(sid, a)


def write_domain(path, cfg, statics):
    """The part of a DOMAIN file the dyn-core reads, every 2-D variable on the file's (iy, jx) grid."""
    nc = netcdf_file(path, "w", version=2)
    try:
        nc.createDimension("jx", cfg.jx); nc.createDimension("iy", cfg.iy); nc.createDimension("kz", cfg.kz + 1)
        nc.grid_size_in_meters = np.float64(cfg.ds * 1000.0)
        _put_ptop(nc, cfg)
        v = nc.createVariable("kz", "f8", ("kz",)); v[:] = statics[capi.S_SIGMA]; v.long_name = "Sigma at full model layers"
        for sid, name in (_DOMAIN_VARS_MOLOCH if cfg.idynamic == 3 else _DOMAIN_VARS).items():
            full = np.zeros((cfg.jx, cfg.iy))
            a = statics[sid]
            full[:a.shape[0], :a.shape[1]] = a
            v = nc.createVariable(name, "f8", ("iy", "jx")); v[:] = full.T
    finally:
        nc.close()


# ---------------------------------------------------------------------------------------------------------- calendar
def _parse_units(units):
    m = re.match(r"\s*hours\s+since\s+(\d+)-(\d+)-(\d+)(?:[ T](\d+):(\d+):(\d+(?:\.\d*)?))?", units)
    if not m:
        raise NcioError(f"time units not understood: {units!r}")
    y, mo, d = int(m.group(1)), int(m.group(2)), int(m.group(3))
    h = int(m.group(4) or 0)
    return y, mo, d, h


_CUM365 = np.cumsum([0, 31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31])


def timeval2date(hours, units, calendar="gregorian"):
    """hours since a reference date -> (year, month, day, hour) in the file's calendar (timeval2date of Share/mod_date.F90 for
    the three calendars the model knows: gregorian, noleap, 360_day)."""
    y, mo, d, h = _parse_units(units)
    hours = int(round(float(hours)))
    cal = calendar.strip().lower()
    if cal in ("gregorian", "standard", "proleptic_gregorian"):
        t = _dt.datetime(y, mo, d, h) + _dt.timedelta(hours=hours)
        return t.year, t.month, t.day, t.hour
    if cal in ("noleap", "365_day"):
        tot = ((y * 365 + int(_CUM365[mo - 1]) + d - 1) * 24 + h) + hours
        days, hh = divmod(tot, 24)
        yy, doy = divmod(days, 365)
        mm = int(np.searchsorted(_CUM365, doy, side="right"))
        return yy, mm, doy - int(_CUM365[mm - 1]) + 1, hh
    if cal == "360_day":
        tot = (((y * 12 + mo - 1) * 30 + d - 1) * 24 + h) + hours
        days, hh = divmod(tot, 24)
        months, dd = divmod(days, 30)
        yy, mm = divmod(months, 12)
        return yy, mm + 1, dd + 1, hh
    raise NcioError(f"calendar not understood: {calendar!r}")


def toint10(date):
    """(y, m, d, h) -> YYYYMMDDHH, the integer form of a model date (tochar10 / toint10 of Share/mod_date.F90)."""
    y, m, d, h = date
    return ((y * 100 + m) * 100 + d) * 100 + h


# ---------------------------------------------------------------------------------------------------------- ICBC
def bdy_fields(cfg):
    """Boundary fields of the dyn-core per core (what init_bdy / bdyin hand to the time interpolation,
    Main/mod_bdycod.F90:325-369,452-496) and the ICBC variable each one comes from."""
    f = {capi.F_U: "u", capi.F_V: "v", capi.F_T: "t", capi.F_QV: "qv"}
    if cfg.idynamic == 2:
        f.update({capi.F_PP: "pp", capi.F_W: "w"})
    else:
        f[capi.F_PS] = "ps"
    return f


class IcbcFile:
    """One `<domname>_ICBC.<YYYYMMDDHH>.nc` (open_icbc, Main/mod_ncio.F90:583-687)."""

    def __init__(self, path, cfg):
        if not os.path.exists(path):
            raise NcioError(f"Cannot open ICBC file {path}")
        self.cfg, self.path = cfg, path
        self.nc = netcdf_file(path, "r", mmap=False)
        check_domain(self.nc, cfg, lmod=True, what="ICBC FILE")
        if "time" not in self.nc.dimensions or "time" not in self.nc.variables:
            raise NcioError("Dimension time miss (ICBC FILE)")
        tv = self.nc.variables["time"]
        for att in ("units", "calendar"):
            if not hasattr(tv, att):
                raise NcioError(f"variable time {att} miss (ICBC FILE)")
        units, self.calendar = tv.units.decode(), tv.calendar.decode()
        t = np.array(tv[:], dtype=np.float64)
        if t.size < 1:
            raise NcioError("Time var in ICBC has zero dim.")
        if t.size > 1 and int(round(t[1] - t[0])) != cfg.ibdyfrq:                      # :632-640
            raise NcioError(f"Time var in ICBC inconsistency: expecting ibdyfrq = {cfg.ibdyfrq}, found {int(round(t[1] - t[0]))}")
        self.dates = [timeval2date(x, units, self.calendar) for x in t]
        self.hours = t
        names = ["ps", "ts", "u", "v", "t", "qv"] + (["pp", "w", "wtop"] if cfg.idynamic == 2 else [])      # :642-666
        for n in names:
            if n not in self.nc.variables:
                raise NcioError(f"variable {n} miss (ICBC FILE)")
        self.has_qc, self.has_qi = "qc" in self.nc.variables, "qi" in self.nc.variables

    def search(self, date):
        """icbc_search (:548-571): record number (1-based) of `date`, -1 when it lies outside this file."""
        d10 = toint10(date)
        if d10 > toint10(self.dates[-1]) or d10 < toint10(self.dates[0]):
            return -1
        for n, d in enumerate(self.dates):
            if toint10(d) == d10:
                return n + 1
        raise NcioError(f"Record is not found in ICBC file for {d10}")

    def read(self, rec):
        """read_icbc (:835-1128) of record `rec` (1-based): {field id: array in the shape and unit rcm_dyn_set_bdy takes}, plus
        "ts" (surface temperature, not a dyn-core field)."""
        cfg, v, r = self.cfg, self.nc.variables, rec - 1
        jx, iy, kz = cfg.jx, cfg.iy, cfg.kz

        def a3(name):
            return np.array(v[name][r], dtype=np.float64).transpose(2, 1, 0)          # file (kz, iy, jx) -> (jx, iy, kz)

        def a2(name):
            return np.array(v[name][r], dtype=np.float64).T

        moloch = cfg.idynamic == 3
        out = {"ts": np.asfortranarray(a2("ts")[:jx - 1, :iy - 1])}
        out[capi.F_U] = np.asfortranarray(a3("u")[:, :iy - 1] if moloch else a3("u"))
        out[capi.F_V] = np.asfortranarray(a3("v")[:jx - 1] if moloch else a3("v"))
        out[capi.F_T] = np.asfortranarray(a3("t")[:jx - 1, :iy - 1])
        out[capi.F_QV] = np.asfortranarray(a3("qv")[:jx - 1, :iy - 1])
        if cfg.idynamic == 2:
            out[capi.F_PP] = np.asfortranarray(a3("pp")[:jx - 1, :iy - 1])
            w = np.empty((jx - 1, iy - 1, kz + 1), order="F")
            w[:, :, 1:] = a3("w")[:jx - 1, :iy - 1]                                     # ww(:,:,2:kzp1) = w, :953
            w[:, :, 0] = a2("wtop")[:jx - 1, :iy - 1]                                   # ww(:,:,1) = wtop, :964
            out[capi.F_W] = w
        else:
            out[capi.F_PS] = np.asfortranarray(a2("ps")[:jx - 1, :iy - 1])
        return out

    def close(self):
        self.nc.close()


def icbc_name(dirglob, domname, date):
    return os.path.join(dirglob, f"{domname}_ICBC.{toint10(date):010d}.nc")


def monfirst(date):
    return (date[0], date[1], 1, 0)


class IcbcReader:
    """The file logic of init_bdy / bdyin (Main/mod_bdycod.F90:283-330,647-690): the file named after the start date of the
    run when the date asked for is that date, after the first of the month otherwise; when a date lies beyond the open file
    the next month's file is opened."""

    def __init__(self, dirglob, domname, cfg, globidate1):
        self.dirglob, self.domname, self.cfg, self.glob1 = dirglob, domname, cfg, tuple(globidate1)
        self.f = None

    def record(self, date):
        date = tuple(date)
        if self.f is None:
            self.f = IcbcFile(icbc_name(self.dirglob, self.domname, date if date == self.glob1 else monfirst(date)), self.cfg)
        rec = self.f.search(date)
        if rec < 0:
            self.f.close()
            self.f = IcbcFile(icbc_name(self.dirglob, self.domname, monfirst(date)), self.cfg)
            rec = self.f.search(date)
            if rec < 0:
                raise NcioError(f"ICBC for {toint10(date)} not found")
        return self.f.read(rec)

    def feed(self, core, slot, date, stage=False):
        """Hand the record of `date` to the library: rcm_dyn_set_bdy(slot) or, for the record read ahead, the asynchronous
        staging entry.  Returns the record (the caller keeps it alive while a staged copy is in flight)."""
        rec = self.record(date)
        for fid in bdy_fields(self.cfg):
            if stage:
                core.stage_bdy_async(fid, rec[fid])
            else:
                core.set_bdy(fid, slot, rec[fid])
        return rec

    def close(self):
        if self.f is not None:
            self.f.close()
            self.f = None


def write_icbc(path, cfg, dates_hours, records, units="hours since 1949-12-01 00:00:00 UTC", calendar="gregorian", dtype="f8"):
    """An ICBC file as PreProc/ICBC writes it: every variable on the file's (time, [kz,] iy, jx) grid.  `records` = a list
    of {field id: array as rcm_dyn_set_bdy takes it} (+ optional "ts"); `dates_hours` = the time coordinate."""
    jx, iy, kz = cfg.jx, cfg.iy, cfg.kz
    nc = netcdf_file(path, "w", version=2)
    try:
        nc.createDimension("time", None); nc.createDimension("jx", jx); nc.createDimension("iy", iy); nc.createDimension("kz", kz)
        nc.grid_size_in_meters = np.float64(cfg.ds * 1000.0)
        _put_ptop(nc, cfg)
        tv = nc.createVariable("time", "f8", ("time",)); tv.units = units; tv.calendar = calendar
        names3 = {"u": capi.F_U, "v": capi.F_V, "t": capi.F_T, "qv": capi.F_QV}
        names2 = {"ps": capi.F_PS, "ts": "ts"}
        if cfg.idynamic == 2:
            names3["pp"] = capi.F_PP
        var = {n: nc.createVariable(n, dtype, ("time", "kz", "iy", "jx")) for n in list(names3) + (["w"] if cfg.idynamic == 2 else [])}
        var.update({n: nc.createVariable(n, dtype, ("time", "iy", "jx")) for n in list(names2) + (["wtop"] if cfg.idynamic == 2 else [])})

        def full(a, nk=None):
            b = np.zeros((jx, iy) if nk is None else (jx, iy, nk))
            b[:a.shape[0], :a.shape[1]] = a
            return b.T if nk is None else b.transpose(2, 1, 0)

        # whole variables at once (scipy's record variables grow by slice assignment)
        tv[:] = np.asarray(dates_hours, dtype=np.float64)
        for n, fid in names3.items():
            var[n][:] = np.stack([full(rec[fid], kz) for rec in records])
        for n, fid in names2.items():
            var[n][:] = np.stack([full(rec[fid]) if fid in rec else np.zeros((iy, jx)) for rec in records])
        if cfg.idynamic == 2:
            var["w"][:] = np.stack([full(rec[capi.F_W][:, :, 1:], kz) for rec in records])
            var["wtop"][:] = np.stack([full(rec[capi.F_W][:, :, 0]) for rec in records])
    finally:
        nc.close()


# ---------------------------------------------------------------------------------------------------------- SAV
def sav_name(dirout, domname, date, prestr=""):
    """Main/mod_savefile.F90:790-792."""
    return os.path.join(dirout, f"{prestr}{domname}_SAV.{toint10(date):010d}.nc")


def _sav_layout(cfg):
    """-> [(variable, field id, level, dims in Fortran order)] of the dyn-core part of a SAV file (:836-1070)."""
    if cfg.idynamic == 3:
        v = [("atm_u", capi.F_U, 0, ("jdot", "icross", "khalf")), ("atm_v", capi.F_V, 0, ("jcross", "idot", "khalf")),
             ("atm_w", capi.F_W, 0, ("jcross", "icross", "kfull")), ("atm_t", capi.F_T, 0, ("jcross", "icross", "khalf")),
             ("atm_pai", capi.F_PAI, 0, ("jcross", "icross", "khalf")), ("ps", capi.F_PS, 0, ("jcross", "icross"))]
        return v
    v = []
    for lv, tag in ((capi.LEVEL_A, "atm1"), (capi.LEVEL_B, "atm2")):
        v += [(f"{tag}_u", capi.F_U, lv, ("jdot", "idot", "khalf")), (f"{tag}_v", capi.F_V, lv, ("jdot", "idot", "khalf")),
              (f"{tag}_t", capi.F_T, lv, ("jcross", "icross", "khalf"))]
        if cfg.ibltyp == 2:
            v.append((f"{tag}_tke", capi.F_TKE, lv, ("jcross", "icross", "kfull")))
        if cfg.idynamic == 2:
            v += [(f"{tag}_w", capi.F_W, lv, ("jcross", "icross", "kfull")), (f"{tag}_pp", capi.F_PP, lv, ("jcross", "icross", "khalf"))]
    if cfg.idynamic == 1:
        v += [("psa", capi.F_PS, capi.LEVEL_A, ("jcross", "icross")), ("psb", capi.F_PS, capi.LEVEL_B, ("jcross", "icross"))]
    return v


def _qx_names(cfg):
    return ("atm_qx",) if cfg.idynamic == 3 else ("atm1_qx", "atm2_qx")


def write_savefile(path, core, date, calendar="gregorian", ps0=None):
    """write_savefile (Main/mod_savefile.F90:764-1264), the dyn-core's share: both time levels of every prognostic, p*, the
    split-explicit stores, the mask of the upper radiative condition, the tracers; `idatex` and `calendar` as global
    attributes (:1880-1910).  The step count and the time since the last boundary record, which the reference rebuilds from
    idatex and its timer on restart, are stored as two more global attributes (`rcm_lcount`, `rcm_xbctime`) the reference
    ignores.  For idynamic = 2 psa / psb are the constant reference p* that the library rebuilds from the DOMAIN file; they
    are written when the caller passes them (`ps0`)."""
    cfg = core.cfg
    jx, iy, kz = cfg.jx, cfg.iy, cfg.kz
    nqx = 2                                   # qv, qc: the species the dyn-core carries (ipptls = 1 of the reference)
    d = core.diag()
    nc = netcdf_file(path, "w", version=2)
    try:
        for n, ln in (("jcross", jx - 1), ("icross", iy - 1), ("jdot", jx), ("idot", iy), ("khalf", kz), ("kfull", kz + 1), ("nqx", nqx)):
            nc.createDimension(n, ln)
        if cfg.idynamic == 1:
            nc.createDimension("nsplit", cfg.nsplit)
        if cfg.ntr > 0:
            nc.createDimension("ntr", cfg.ntr)
        if cfg.idynamic == 2 and cfg.ifupr == 1:
            nc.createDimension("ikern", 13)
        nc.idatex = np.int32(toint10(date)); nc.calendar = calendar
        nc.rcm_lcount = np.float64(d.lcount); nc.rcm_xbctime = np.float64(d.xbctime)

        def put(name, a, dims):
            v = nc.createVariable(name, "f8", tuple(reversed(dims)))                 # Fortran order -> file order
            v[:] = np.transpose(a)

        for name, fid, lv, dims in _sav_layout(cfg):
            put(name, core.get_state(fid, lv), dims)
        for name, lv in zip(_qx_names(cfg), (capi.LEVEL_A, capi.LEVEL_B)):
            put(name, np.stack([core.get_state(capi.F_QV, lv), core.get_state(capi.F_QC, lv)], axis=-1), ("jcross", "icross", "khalf", "nqx"))
        if cfg.idynamic == 2 and ps0 is not None:
            put("psa", ps0, ("jcross", "icross")); put("psb", ps0, ("jcross", "icross"))
        if cfg.ntr > 0 and cfg.idynamic != 3:                                          # chia / chib, :714-715
            for name, lv in (("chia", capi.LEVEL_A), ("chib", capi.LEVEL_B)):
                put(name, np.stack([core.get_state(capi.F_CHI + n, lv) for n in range(cfg.ntr)], axis=-1), ("jcross", "icross", "khalf", "ntr"))
        if cfg.idynamic == 1:
            for name, rid in (("dstor", capi.R_DSTOR), ("hstor", capi.R_HSTOR)):       # allocated on the dot extents, :326-327
                a = core.get_restart(rid)
                full = np.zeros((jx, iy, cfg.nsplit)); full[:a.shape[0], :a.shape[1]] = a
                put(name, full, ("jdot", "idot", "nsplit"))
        if cfg.idynamic == 2 and cfg.ifupr == 1:
            put("tmask", core.get_restart(capi.R_TMASK)[:, :, 0], ("ikern", "ikern"))
    finally:
        nc.close()


def read_savefile(path, core, date=None):
    """read_savefile (Main/mod_savefile.F90:564-762) into a core that has been created, given its DOMAIN fields and the
    boundary records that bracket the restart date, and initialised (the order of `init` on a restart,
    Main/mod_init.F90:418-796): restores the state, the stores, the mask and the clock.  A file whose idatex differs from
    `date` is refused as the reference refuses it (:1823-1829)."""
    cfg = core.cfg
    nc = netcdf_file(path, "r", mmap=False)
    try:
        if date is not None and int(nc.idatex) != toint10(date):
            raise NcioError(f"Mismatch in dates: idate1 in namelist is {toint10(date)}, idatex in SAV file is {int(nc.idatex)}")

        def get(name):
            if name not in nc.variables:
                raise NcioError(f"variable {name} miss (SAV FILE)")
            return np.asfortranarray(np.transpose(np.array(nc.variables[name][:], dtype=np.float64)))

        for name, fid, lv, _ in _sav_layout(cfg):
            if cfg.idynamic == 3 and fid == capi.F_PS:
                continue            # diagnosed from Pi at the surface by the library every step (include/regcm_dyn.h, RCM_F_PS)
            core.set_state(fid, lv, get(name))
        for name, lv in zip(_qx_names(cfg), (capi.LEVEL_A, capi.LEVEL_B)):
            qx = get(name)
            core.set_state(capi.F_QV, lv, np.asfortranarray(qx[..., 0])); core.set_state(capi.F_QC, lv, np.asfortranarray(qx[..., 1]))
        if cfg.ntr > 0 and cfg.idynamic != 3:
            for name, lv in (("chia", capi.LEVEL_A), ("chib", capi.LEVEL_B)):
                chi = get(name)
                for n in range(cfg.ntr):
                    core.set_state(capi.F_CHI + n, lv, np.asfortranarray(chi[..., n]))
        if cfg.idynamic == 1:
            for name, rid in (("dstor", capi.R_DSTOR), ("hstor", capi.R_HSTOR)):
                n1, n2, _ = core._restart_shape(rid)
                core.set_restart(rid, np.asfortranarray(get(name)[:n1, :n2]))
        if cfg.idynamic == 2 and cfg.ifupr == 1:
            core.set_restart(capi.R_TMASK, np.asfortranarray(get("tmask")[:, :, None]))
        core.set_clock(int(round(float(nc.rcm_lcount))), float(nc.rcm_xbctime))
    finally:
        nc.close()
