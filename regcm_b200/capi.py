"""ctypes binding of the C ABI declared in include/regcm_dyn.h.

`DynCore` is the host-side mirror of the reference's interface for the dyn-core
path: the reference drives it with argument-less module procedures
(`call tend`, `call bdyin`, `call bdyval`, Main/mod_regcm_interface.F90:172-228)
acting on the public arrays of mod_atm_interface (Main/mod_atm_interface.F90:39-60);
here the same life cycle is create -> set_static -> set_bdy -> init -> step ->
get_state.  All compute happens in the CUDA library; this module contains no
numerical fallback and raises if the library cannot be loaded.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

import numpy as np

# enum rcm_field / rcm_static / rcm_work (include/regcm_dyn.h)
F_U, F_V, F_T, F_QV, F_QC, F_PP, F_W, F_PS, F_PAI = range(9)
F_TKE = 9        # atm1%tke / atm2%tke (ibltyp = 2), uncoupled, kz + 1 levels
F_CHI = 16       # tracer n is field F_CHI + n
R_DSTOR, R_HSTOR, R_TMASK = range(3)
MAXTR = 8
FIELD_NAMES = ("u", "v", "t", "qv", "qc", "pp", "w", "ps", "pai")
S_HT, S_MSFX, S_MSFD, S_CORIOL, S_XLAT, S_XLON, S_DLAT, S_SIGMA, S_MSFU, S_MSFV, S_ULAT, S_VLAT = range(12)
LEVEL_A, LEVEL_B = 0, 1
BDY_B0, BDY_B1, BDY_NEXT = 0, 1, 2
(W_CR, W_QDOT, W_XKC, W_TH, W_RHO1, W_PR1, W_UTEN, W_VTEN, W_PPTEN, W_WTEN, W_OMEGA, W_PHI, W_PTEN,
 W_MO_P, W_MO_TVIRT, W_MO_FMZ, W_MO_FMZF, W_MO_ZETA, W_MO_S, W_CQV, W_CQC) = range(21)
WORK_NAMES = ("cr", "qdot", "xkc", "th", "rho1", "pr1", "uten", "vten", "ppten", "wten", "omega", "phi", "pten",
              "mo_p", "mo_tvirt", "mo_fmz", "mo_fmzf", "mo_zeta", "mo_s", "cqv", "cqc")
# enum rcm_slice: what mkslice hands to the physical parametrizations (Main/mod_slice.F90:167-308)
(SL_TB3D, SL_QVB3D, SL_QCB3D, SL_UBX3D, SL_VBX3D, SL_PB3D, SL_PF3D, SL_PS2D, SL_RHOB3D, SL_TH3D, SL_TV3D, SL_TP3D,
 SL_WPX3D, SL_WB3D, SL_ZQ, SL_ZA, SL_DZQ, SL_QSB3D, SL_RHB3D) = range(19)
SLICE_NAMES = ("tb3d", "qvb3d", "qcb3d", "ubx3d", "vbx3d", "pb3d", "pf3d", "ps2d", "rhob3d", "th3d", "tv3d", "tp3d",
               "wpx3d", "wb3d", "zq", "za", "dzq", "qsb3d", "rhb3d")
# enum rcm_table: vertical-mode tables of the split-explicit scheme (idynamic = 1)
T_HBAR, T_ZMATX, T_ZMATXR, T_TAU, T_VARPA1, T_AM, T_AN, T_TAU0 = range(8)

ERRORS = {0: "RCM_OK", -1: "RCM_ERR_ARG", -2: "RCM_ERR_CUDA", -3: "RCM_ERR_CFL",
          -4: "RCM_ERR_NAN", -5: "RCM_ERR_NCCL", -6: "RCM_ERR_BDY", -7: "RCM_ERR_STATE"}


class RcmCfg(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32),
        ("jx", C.c_int32), ("iy", C.c_int32), ("kz", C.c_int32), ("nqx", C.c_int32),
        ("idynamic", C.c_int32),
        ("iboudy", C.c_int32), ("idiffu", C.c_int32), ("ipptls", C.c_int32),
        ("ifupr", C.c_int32), ("ifrayd", C.c_int32), ("rayndamp", C.c_int32),
        ("nspgx", C.c_int32), ("nspgd", C.c_int32),
        ("ibdyfrq", C.c_int32),
        ("upstream_mode", C.c_int32), ("stability_enhance", C.c_int32), ("diffu_hgtf", C.c_int32),
        ("nproc", C.c_int32), ("myid", C.c_int32),
        ("cpus_j", C.c_int32), ("cpus_i", C.c_int32),
        ("device", C.c_int32),
        ("nsplit", C.c_int32), ("lstand", C.c_int32),
        ("mo_nadv", C.c_int32), ("mo_nsound", C.c_int32), ("mo_nzfilt", C.c_int32), ("mo_rotllr", C.c_int32),
        ("ntr", C.c_int32),
        ("ibltyp", C.c_int32), ("iuwvadv", C.c_int32),
        ("ipgf", C.c_int32),
        ("isladvec", C.c_int32), ("iqmsl", C.c_int32),
        ("ds", C.c_double), ("ptop", C.c_double),
        ("dt", C.c_double),
        ("gnu1", C.c_double), ("gnu2", C.c_double), ("ckh", C.c_double), ("adyndif", C.c_double),
        ("uoffc", C.c_double), ("t_extrema", C.c_double), ("q_rel_extrema", C.c_double),
        ("nhbet", C.c_double), ("nhxkd", C.c_double), ("rayalpha0", C.c_double), ("rayhd", C.c_double),
        ("high_nudge", C.c_double), ("medium_nudge", C.c_double), ("low_nudge", C.c_double),
        ("bdy_nm", C.c_double), ("bdy_dm", C.c_double),
        ("base_state_pressure", C.c_double), ("logp_lrate", C.c_double), ("base_state_ts0", C.c_double),
        ("mo_anu2", C.c_double), ("mo_a0", C.c_double), ("mo_b0", C.c_double),
        ("uw_nuk", C.c_double), ("uw_tkemin", C.c_double),
        ("reserved_d", C.c_double * 3),
    ]


class RcmDiag(C.Structure):
    _fields_ = [
        ("lcount", C.c_int64), ("xbctime", C.c_double), ("dt", C.c_double),
        ("max_cfl", C.c_double), ("istep", C.c_int32), ("nan_flag", C.c_int32),
        ("neg_q_points", C.c_int64),
        ("ptntot", C.c_double), ("pt2tot", C.c_double),
        ("ps_max", C.c_double), ("ps_min", C.c_double),
        ("neg_chi_points", C.c_int64),
    ]


def default_cfg(idynamic: int = 2, **kw) -> RcmCfg:
    """Reference defaults (Main/mod_params.F90:190-300,645-665;
    Share/mod_dynparam.F90:147-154,502-508).  Pure data, no library call."""
    c = RcmCfg()
    c.struct_size = C.sizeof(RcmCfg)
    c.nqx = 2
    c.idynamic = idynamic
    c.iboudy, c.idiffu, c.ipptls = 5, 1, 1
    c.ifupr, c.ifrayd, c.rayndamp = 1, 1, 5
    c.nspgx = c.nspgd = 12
    c.ibdyfrq = 6
    c.upstream_mode = c.stability_enhance = 1
    c.nproc, c.myid, c.cpus_j, c.cpus_i, c.device = 1, 0, 0, 0, -1
    c.ptop = 5.0
    c.nsplit, c.lstand = 2, 1   # Main/mod_params.F90:283-284
    c.mo_nadv, c.mo_nsound, c.mo_nzfilt = 3, 6, 3      # Main/mod_params.F90:296-300
    c.mo_anu2, c.mo_a0, c.mo_b0 = 0.6, 0.0, 0.5        # Share/mod_dynparam.F90:189-190
    c.iqmsl = 1                                        # Main/mod_params.F90:244
    c.uw_nuk, c.uw_tkemin = 5.0, 1.0e-3                # Main/mod_params.F90:480; Main/pbllib/mod_pbl_uwtcm.F90:86
    if idynamic == 2:
        c.gnu1 = c.gnu2 = 0.1
        c.diffu_hgtf = 0
    else:
        c.gnu1 = c.gnu2 = 0.0625
        c.diffu_hgtf = 1
    c.ckh = c.adyndif = 1.0
    c.uoffc, c.t_extrema, c.q_rel_extrema = 0.25, 5.0, 0.20
    c.nhbet, c.nhxkd, c.rayalpha0, c.rayhd = 0.4, 0.1, 0.0003, 10000.0
    c.high_nudge, c.medium_nudge, c.low_nudge = 3.0, 2.0, 1.0
    c.bdy_nm = c.bdy_dm = -1.0
    c.base_state_pressure, c.logp_lrate, c.base_state_ts0 = 101325.0, 47.70, 288.15
    for k, v in kw.items():
        if not hasattr(c, k):
            raise AttributeError(f"rcm_cfg has no field {k}")
        setattr(c, k, v)
    return c


class RcmError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"{ERRORS.get(code, code)}: {msg}")
        self.code = code


_I3 = C.c_int32 * 3
_DP = C.POINTER(C.c_double)


def _bounds(lb: Sequence[int], ub: Sequence[int]):
    return _I3(*lb), _I3(*ub)


class CoreLib:
    """One shared library exporting the regcm_dyn C ABI under `prefix`.

    prefix = "rcm_dyn_" for the CUDA product library.  (The CPU restatement
    under oracle/ exports the same call shapes with prefix "orc_"; only the
    tests and the bench's cpu_baseline leg construct that one.)"""

    def __init__(self, path: str, prefix: str):
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} not found: build it first (python -c 'import __graft_entry__ as g; g.build()')")
        self.path, self.prefix = path, prefix
        self.lib = C.CDLL(path, mode=C.RTLD_LOCAL)
        p = prefix
        vp = C.c_void_p
        sig = {
            "create": ([C.POINTER(RcmCfg), C.POINTER(vp)], C.c_int),
            "set_static": ([vp, C.c_int, _DP, _I3, _I3], C.c_int),
            "set_bdy": ([vp, C.c_int, C.c_int, _DP, _I3, _I3], C.c_int),
            "init": ([vp], C.c_int),
            "step": ([vp, C.c_int], C.c_int),
            "get_state": ([vp, C.c_int, C.c_int, _DP, _I3, _I3], C.c_int),
            "get_work": ([vp, C.c_int, _DP, _I3, _I3], C.c_int),
            "diag": ([vp, C.POINTER(RcmDiag)], C.c_int),
            "last_error": ([vp, C.c_char_p, C.c_int], C.c_int),
            "destroy": ([vp], C.c_int),
        }
        for name, (args, res) in sig.items():
            fn = getattr(self.lib, p + name)
            fn.argtypes, fn.restype = args, res
            setattr(self, name, fn)
        for opt, (args, res) in {
            "set_state": ([vp, C.c_int, C.c_int, _DP, _I3, _I3], C.c_int),
            "launch_count": ([vp], C.c_int64),
            "last_step_ms": ([vp], C.c_double),
            "get_tile": ([vp, C.POINTER(C.c_int32)], C.c_int),
            "get_table": ([vp, C.c_int, _DP, C.c_int], C.c_int),
            "profile": ([vp, C.c_int], C.c_int),
            "profile_get": ([vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_char_p, C.c_int], C.c_int),
            "profile_kernels": ([], C.c_int),
            "comm_unique_id": ([C.c_void_p, C.c_int], C.c_int),
            "comm_init": ([vp, C.c_void_p, C.c_int], C.c_int),
            "stage_bdy_async": ([vp, C.c_int, _DP, _I3, _I3], C.c_int),
            "stage_wait": ([vp], C.c_int),
            "step_async": ([vp, C.c_int], C.c_int),
            "sync": ([vp], C.c_int),
            "get_state_async": ([vp, C.c_int, C.c_int, _DP, _I3, _I3], C.c_int),
            "get_state_wait": ([vp], C.c_int),
            # physics coupling boundary
            "step_pre_physics": ([vp], C.c_int),
            "step_post_physics": ([vp], C.c_int),
            "set_phys_tend": ([vp, C.c_int, _DP, _I3, _I3], C.c_int),
            "clear_phys_tend": ([vp], C.c_int),
            "get_slice": ([vp, C.c_int, _DP, _I3, _I3], C.c_int),
            "set_kpbl": ([vp, C.POINTER(C.c_int32), C.c_int32 * 2, C.c_int32 * 2], C.c_int),
            # restart
            "get_restart": ([vp, C.c_int, _DP, _I3, _I3], C.c_int),
            "set_restart": ([vp, C.c_int, _DP, _I3, _I3], C.c_int),
            "set_clock": ([vp, C.c_int64, C.c_double], C.c_int),
            # handle-less, allocation-free operator (device pointers): include/regcm_dyn.h
            "negq_repair_static": ([C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p], C.c_int),
        }.items():
            if hasattr(self.lib, p + opt):
                fn = getattr(self.lib, p + opt)
                fn.argtypes, fn.restype = args, res
                setattr(self, opt, fn)


class DynCore:
    """Handle on one dyn-core instance (one rank / one GPU)."""

    def __init__(self, lib: CoreLib, cfg: RcmCfg):
        self.lib, self.cfg = lib, cfg
        self.h = C.c_void_p()
        rc = lib.create(C.byref(cfg), C.byref(self.h))
        if rc != 0 or not self.h:
            raise RcmError(rc, f"{lib.prefix}create failed")
        self.jx, self.iy, self.kz = cfg.jx, cfg.iy, cfg.kz

    # -- helpers ---------------------------------------------------------
    def _check(self, rc: int):
        if rc != 0:
            buf = C.create_string_buffer(512)
            self.lib.last_error(self.h, buf, 512)
            raise RcmError(rc, buf.value.decode(errors="replace"))

    @staticmethod
    def _f(a: np.ndarray) -> np.ndarray:
        a = np.asarray(a, dtype=np.float64)
        if not a.flags.f_contiguous:
            a = np.asfortranarray(a)
        return a

    def field_shape(self, field: int):
        jx, iy, kz = self.jx, self.iy, self.kz
        if self.cfg.idynamic == 3 and field in (F_U, F_V):   # Arakawa C grid
            return (jx, iy - 1, kz) if field == F_U else (jx - 1, iy, kz)
        if field in (F_U, F_V):
            return (jx, iy, kz)
        if field in (F_W, F_TKE):
            return (jx - 1, iy - 1, kz + 1)
        if field == F_PS:
            return (jx - 1, iy - 1, 1)
        return (jx - 1, iy - 1, kz)

    # -- API ---------------------------------------------------------------
    def set_static(self, sid: int, a: np.ndarray, lb=None):
        a = self._f(a)
        if sid == S_SIGMA:
            lb3, ub3 = (1, 1, 1), (1, 1, a.size)
        else:
            l = lb or (1, 1)
            lb3, ub3 = (l[0], l[1], 1), (l[0] + a.shape[0] - 1, l[1] + a.shape[1] - 1, 1)
        self._check(self.lib.set_static(self.h, sid, a.ctypes.data_as(_DP), *_bounds(lb3, ub3)))

    def set_bdy(self, field: int, slot: int, a: np.ndarray, lb=(1, 1, 1)):
        a = self._f(a)
        if a.ndim == 2:
            a = a.reshape(a.shape + (1,), order="F")
        ub = tuple(lb[d] + a.shape[d] - 1 for d in range(3))
        self._check(self.lib.set_bdy(self.h, field, slot, a.ctypes.data_as(_DP), *_bounds(lb, ub)))

    def stage_bdy_async(self, field: int, a: np.ndarray, lb=(1, 1, 1)):
        """Queue the H2D copy of one field of the NEXT record; `a` must be Fortran-ordered
        float64 and stay alive until stage_wait() or the record change."""
        assert a.dtype == np.float64 and a.flags.f_contiguous
        shp = a.shape if a.ndim == 3 else a.shape + (1,)
        ub = tuple(lb[d] + shp[d] - 1 for d in range(3))
        self._check(self.lib.stage_bdy_async(self.h, field, a.ctypes.data_as(_DP), *_bounds(lb, ub)))

    def stage_wait(self):
        self._check(self.lib.stage_wait(self.h))

    def init(self):
        self._check(self.lib.init(self.h))

    def step(self, n: int = 1):
        self._check(self.lib.step(self.h, n))

    # -- physics coupling boundary: the step splits where `tend` calls physical_parametrizations -----------------
    def step_pre_physics(self):
        self._check(self.lib.step_pre_physics(self.h))

    def step_post_physics(self):
        self._check(self.lib.step_post_physics(self.h))

    def set_phys_tend(self, field: int, a: np.ndarray, lb=(1, 1, 1)):
        """aten(:,:,:,pc_physic) of one prognostic (coupled with p*), in effect until replaced or cleared."""
        a = self._f(a)
        ub = tuple(lb[d] + a.shape[d] - 1 for d in range(3))
        self._check(self.lib.set_phys_tend(self.h, field, a.ctypes.data_as(_DP), *_bounds(lb, ub)))

    def clear_phys_tend(self):
        self._check(self.lib.clear_phys_tend(self.h))

    def set_kpbl(self, kpbl: np.ndarray, lb=(2, 2)):
        """Top of the PBL (level index) on cross points, lb = global index of kpbl[0, 0]; read by vadv4d ind = 3."""
        k = np.asfortranarray(kpbl, dtype=np.int32)
        ub = (lb[0] + k.shape[0] - 1, lb[1] + k.shape[1] - 1)
        I2 = C.c_int32 * 2
        self._check(self.lib.set_kpbl(self.h, k.ctypes.data_as(C.POINTER(C.c_int32)), I2(*lb), I2(*ub)))

    def get_slice(self, sid: int) -> np.ndarray:
        """One slice of `atms` on the whole cross grid (jx-1, iy-1, nk)."""
        nk = self.kz + 1 if sid in (SL_PF3D, SL_WB3D, SL_ZQ) else 1 if sid == SL_PS2D else self.kz
        ub = (self.jx - 1, self.iy - 1, nk)
        out = np.zeros(ub, dtype=np.float64, order="F")
        self._check(self.lib.get_slice(self.h, sid, out.ctypes.data_as(_DP), *_bounds((1, 1, 1), ub)))
        return out

    def step_async(self, n: int = 1):
        """Enqueue n steps without waiting (rcm_dyn_step_async); sync() returns the verdict of the batch."""
        self._check(self.lib.step_async(self.h, n))

    def sync(self):
        self._check(self.lib.sync(self.h))

    def get_state_async(self, field: int, level: int, out: np.ndarray, lb, ub):
        """Queue the export of one field into `out` (Fortran-ordered float64, ideally page-locked); it is complete
        after get_state_wait()."""
        assert out.dtype == np.float64 and out.flags.f_contiguous and out.shape == tuple(ub[d] - lb[d] + 1 for d in range(3))
        self._check(self.lib.get_state_async(self.h, field, level, out.ctypes.data_as(_DP), *_bounds(lb, ub)))

    def get_state_wait(self):
        self._check(self.lib.get_state_wait(self.h))

    def get_state(self, field: int, level: int = LEVEL_A, lb=None, ub=None, out=None) -> np.ndarray:
        shp = self.field_shape(field)
        lb = lb or (1, 1, 1)
        ub = ub or shp
        if out is None:
            out = np.zeros(tuple(ub[d] - lb[d] + 1 for d in range(3)), dtype=np.float64, order="F")
        assert out.flags.f_contiguous and out.shape == tuple(ub[d] - lb[d] + 1 for d in range(3))
        self._check(self.lib.get_state(self.h, field, level, out.ctypes.data_as(_DP), *_bounds(lb, ub)))
        return out

    def set_state(self, field: int, level: int, a: np.ndarray, lb=(1, 1, 1)):
        a = self._f(a)
        if a.ndim == 2:
            a = a.reshape(a.shape + (1,), order="F")
        ub = tuple(lb[d] + a.shape[d] - 1 for d in range(3))
        self._check(self.lib.set_state(self.h, field, level, a.ctypes.data_as(_DP), *_bounds(lb, ub)))

    def get_work(self, wid: int) -> np.ndarray:
        jx, iy, kz = self.jx, self.iy, self.kz
        if wid in (W_UTEN, W_VTEN):
            lb, ub = (2, 2, 1), (jx - 1, iy - 1, kz)
        elif wid in (W_PPTEN,):
            lb, ub = (2, 2, 1), (jx - 2, iy - 2, kz)
        elif wid == W_WTEN:
            lb, ub = (2, 2, 1), (jx - 2, iy - 2, kz + 1)
        elif wid in (W_QDOT, W_MO_FMZF, W_MO_S):
            lb, ub = (1, 1, 1), (jx - 1, iy - 1, kz + 1)
        elif wid == W_OMEGA:
            lb, ub = (2, 2, 1), (jx - 2, iy - 2, kz)
        elif wid == W_PTEN:
            lb, ub = (1, 1, 1), (jx - 1, iy - 1, 1)
        else:
            lb, ub = (1, 1, 1), (jx - 1, iy - 1, kz)
        out = np.zeros(tuple(ub[d] - lb[d] + 1 for d in range(3)), dtype=np.float64, order="F")
        self._check(self.lib.get_work(self.h, wid, out.ctypes.data_as(_DP), *_bounds(lb, ub)))
        return out

    # -- restart: what the SAV file holds beside the prognostics (include/regcm_dyn.h) ------------------------------
    def _restart_shape(self, rid: int):
        return (13, 13, 1) if rid == R_TMASK else (self.jx - 1, self.iy - 1, self.cfg.nsplit)

    def get_restart(self, rid: int) -> np.ndarray:
        shp = self._restart_shape(rid)
        out = np.zeros(shp, dtype=np.float64, order="F")
        self._check(self.lib.get_restart(self.h, rid, out.ctypes.data_as(_DP), *_bounds((1, 1, 1), shp)))
        return out

    def set_restart(self, rid: int, a: np.ndarray, lb=(1, 1, 1)):
        a = self._f(a)
        ub = tuple(lb[d] + a.shape[d] - 1 for d in range(3))
        self._check(self.lib.set_restart(self.h, rid, a.ctypes.data_as(_DP), *_bounds(lb, ub)))

    def set_clock(self, lcount: int, xbctime: float):
        self._check(self.lib.set_clock(self.h, int(lcount), float(xbctime)))

    def get_table(self, tid: int) -> np.ndarray:
        """Vertical-mode table `tid` (enum rcm_table) as a Fortran-ordered array."""
        kz, ns = self.kz, self.cfg.nsplit
        shape = {T_HBAR: (kz,), T_ZMATX: (kz, kz), T_ZMATXR: (kz, kz), T_TAU: (kz, kz), T_VARPA1: (kz, kz + 1),
                 T_AM: (kz, ns), T_AN: (ns,), T_TAU0: (kz, kz)}[tid]
        n = int(np.prod(shape))
        out = np.zeros(n, dtype=np.float64)
        rc = self.lib.get_table(self.h, tid, out.ctypes.data_as(_DP), n)
        if rc < 0:
            self._check(rc)
        assert rc == n, (rc, n)
        return out.reshape(shape, order="F")

    def diag(self) -> RcmDiag:
        d = RcmDiag()
        self._check(self.lib.diag(self.h, C.byref(d)))
        return d

    def tile(self):
        """(jde1, jde2, ide1, ide2): this rank's dot-point extent in global indices."""
        out = (C.c_int32 * 4)()
        self._check(self.lib.get_tile(self.h, out))
        return tuple(int(x) for x in out)

    def tile_cells(self) -> int:
        j1, j2, i1, i2 = self.tile()
        return (j2 - j1 + 1) * (i2 - i1 + 1) * self.kz

    def profile(self, enable: bool):
        self._check(self.lib.profile(self.h, 1 if enable else 0))

    def profile_report(self):
        """{kernel name: (total ms, launches)} accumulated since profile(True)."""
        out = {}
        for kid in range(self.lib.profile_kernels()):
            ms, n, nm = C.c_double(), C.c_int64(), C.create_string_buffer(64)
            self._check(self.lib.profile_get(self.h, kid, C.byref(ms), C.byref(n), nm, 64))
            out[nm.value.decode()] = (ms.value, n.value)
        return out

    def comm_init(self, unique_id: bytes):
        buf = C.create_string_buffer(unique_id, 128)
        self._check(self.lib.comm_init(self.h, buf, 128))

    def launch_count(self) -> int:
        return int(self.lib.launch_count(self.h))

    def last_step_ms(self) -> float:
        return float(self.lib.last_step_ms(self.h))

    def close(self):
        if self.h:
            self.lib.destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nccl_unique_id(lib: "CoreLib") -> bytes:
    buf = C.create_string_buffer(128)
    rc = lib.comm_unique_id(buf, 128)
    if rc != 0:
        raise RcmError(rc, "rcm_dyn_comm_unique_id failed (libnccl.so.2 not loadable?)")
    return buf.raw


_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
# RCM_DYN_LIB selects another build of the same library (kernel tuning runs only)
CUDA_LIB_PATH = os.environ.get("RCM_DYN_LIB") or os.path.join(_PKG_DIR, "csrc", "libregcm_dyn.so")
_cuda_lib = None


def cuda_lib() -> CoreLib:
    """The product library.  Fails loudly when it has not been built: there is
    no CPU fallback behind this API."""
    global _cuda_lib
    if _cuda_lib is None:
        _cuda_lib = CoreLib(CUDA_LIB_PATH, "rcm_dyn_")
    return _cuda_lib
