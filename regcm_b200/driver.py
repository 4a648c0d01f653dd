"""The time loop around the dyn-core with the reference's files on both sides: what RCM_initialize /
RCM_run do for a dynamics-only run (Main/mod_regcm_interface.F90:59-160,171-260) — read the DOMAIN file, open the ICBC file of
the start date and read the two records that bracket it (`init_bdy`, Main/mod_bdycod.F90:283-369), `init`, then per step
`tend` / `moloch`, on the boundary alarm `bdyin`, `bdyval` (all inside rcm_dyn_step), and on the SAV alarm `write_savefile`.
The one thing the host loop owns is I/O: the ICBC record after next is read while the steps of the current interval run
(rcm_dyn_stage_bdy_async; the reference reads it synchronously inside `bdyin`, :669-690) and the SAV file is written at the
dates asked for.  A restart (`ifrest`) reads the records that bracket the restart date again and then the SAV file, as
`init_bdy` + `init` do (Main/mod_init.F90:418-796).

Dates are (year, month, day, hour) tuples; the gregorian calendar of the ICBC files."""
from __future__ import annotations

import datetime as _dt

from . import capi, ncio


def _add_hours(date, hours):
    t = _dt.datetime(*date) + _dt.timedelta(hours=hours)
    return (t.year, t.month, t.day, t.hour)


class Run:
    def __init__(self, lib, cfg, dirter, dirglob, domname, globidate1, dirout=None, restart_from=None):
        """RCM_initialize.  `restart_from` = the date of a SAV file in `dirout` to continue from (ifrest = .true.)."""
        self.cfg, self.domname, self.dirout = cfg, domname, dirout
        self.steps_per_record = cfg.ibdyfrq * 3600.0 / cfg.dt
        if abs(self.steps_per_record - round(self.steps_per_record)) > 1e-9:
            raise ncio.NcioError("ibdyfrq hours must be a whole number of steps (Main/mod_params.F90 checks the same)")
        self.steps_per_record = int(round(self.steps_per_record))
        self.core = capi.DynCore(lib, cfg)
        ncio.load_domain(self.core, ncio.read_domain_info(f"{dirter}/{domname}_DOMAIN000.nc", cfg))
        self.icbc = ncio.IcbcReader(dirglob, domname, cfg, globidate1)
        start = tuple(restart_from) if restart_from else tuple(globidate1)
        hrs = (_dt.datetime(*start) - _dt.datetime(*globidate1)).total_seconds() / 3600.0
        if hrs < 0 or abs(hrs * 3600.0 / cfg.dt - round(hrs * 3600.0 / cfg.dt)) > 1e-9:
            raise ncio.NcioError("the restart date is not a step of the run")
        b0 = _add_hours(tuple(globidate1), int(hrs // cfg.ibdyfrq) * cfg.ibdyfrq)       # the record at or before the start date
        self.bdydate2 = _add_hours(b0, cfg.ibdyfrq)
        self.icbc.feed(self.core, capi.BDY_B0, b0)
        self.icbc.feed(self.core, capi.BDY_B1, self.bdydate2)
        self.core.init()
        self.date0 = tuple(globidate1)
        self.lcount = 0
        if restart_from:
            ncio.read_savefile(ncio.sav_name(dirout, domname, start), self.core, date=start)
            self.lcount = self.core.diag().lcount
        self._staged = None
        self._stage_next()

    def _stage_next(self):
        """Read the record after b1 ahead of time; it is consumed by the record change at the end of the current interval."""
        nxt = _add_hours(self.bdydate2, self.cfg.ibdyfrq)
        try:
            self._staged = self.icbc.feed(self.core, capi.BDY_NEXT, nxt, stage=True)
        except ncio.NcioError:
            self._staged = None                      # end of the boundary data: the run may still finish the current interval
        self.bdydate2 = nxt

    def date(self):
        secs = self.lcount * self.cfg.dt
        t = _dt.datetime(*self.date0) + _dt.timedelta(seconds=secs)
        return (t.year, t.month, t.day, t.hour), t.minute * 60 + t.second

    def run(self, nsteps, sav_dates=()):
        """RCM_run for `nsteps` steps; a SAV file is written when the model date reaches one of `sav_dates`."""
        sav = {}
        for d in sav_dates:                          # SAV alarm: the step count at which the model date is d
            k = (_dt.datetime(*d) - _dt.datetime(*self.date0)).total_seconds() / self.cfg.dt
            if abs(k - round(k)) < 1e-9:
                sav[int(round(k))] = tuple(d)
        end = self.lcount + nsteps
        while self.lcount < end:
            to_alarm = self.steps_per_record - self.lcount % self.steps_per_record
            n = min([to_alarm, end - self.lcount] + [k - self.lcount for k in sav if k > self.lcount])
            if n == to_alarm and self._staged is None:
                raise ncio.NcioError("ICBC data ends before the run does")
            self.core.step(n)
            self.lcount += n
            if n == to_alarm:                        # the boundary alarm fired inside the last step: b1 <- the staged record
                self.core.stage_wait()
                self._stage_next()
            if self.lcount in sav:
                d = sav[self.lcount]
                ncio.write_savefile(ncio.sav_name(self.dirout, self.domname, d), self.core, d)

    def close(self):
        self.icbc.close()
        self.core.close()
