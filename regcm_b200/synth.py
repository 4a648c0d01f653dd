"""Synthetic analytic DOMAIN + ICBC generator.

Replaces the reference's pre-processing chain (PreProc/Terrain, PreProc/ICBC,
read_domain_info / read_icbc in Main/mod_ncio.F90) with in-memory analytic
fields: Gaussian terrain, Lambert-conformal map factors, a balanced reference
temperature profile plus a smooth jet / thermal / moisture perturbation and
seeded round-off-scale noise, and boundary records 6 h apart so that the
boundary tendency `bt` is non-zero.  The reference's own precedent for
"analytic ICBC, physics off" is moloch_static_test1/2
(Main/mod_bdycod.F90:5157-5199).

Everything is generated on the GLOBAL grid so that every rank count sees
identical data.  Arrays are Fortran-ordered (j,i[,k]) float64.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict

import numpy as np

from . import capi

SEED = 20261104

# hard-coded full-sigma tables of the reference (data, Share/mod_sigma.F90:92-205)
_SIGMA = {
    14: [0.0, 0.04, 0.10, 0.17, 0.25, 0.35, 0.46, 0.56, 0.67, 0.77, 0.86, 0.93, 0.97, 0.99, 1.0],
    18: [0.0, 0.05, 0.10, 0.16, 0.23, 0.31, 0.39, 0.47, 0.55, 0.63, 0.71, 0.78, 0.84, 0.89, 0.93,
         0.96, 0.98, 0.99, 1.0],
    23: [0.0, 0.05, 0.1, 0.15, 0.2, 0.25, 0.3, 0.35, 0.4, 0.45, 0.5, 0.55, 0.6, 0.65, 0.7, 0.75,
         0.8, 0.85, 0.89, 0.93, 0.96, 0.98, 0.99, 1.0],
    41: [0.0000, 0.0500, 0.0978, 0.1436, 0.1875, 0.2295, 0.2697, 0.3082, 0.3451, 0.3804, 0.4143,
         0.4468, 0.4779, 0.5078, 0.5364, 0.5639, 0.5903, 0.6156, 0.6399, 0.6632, 0.6856, 0.7071,
         0.7277, 0.7476, 0.7667, 0.7850, 0.8027, 0.8196, 0.8359, 0.8516, 0.8667, 0.8812, 0.8952,
         0.9087, 0.9216, 0.9341, 0.9461, 0.9577, 0.9689, 0.9796, 0.9900, 1.0000],
}


def sigma_levels(kz: int) -> np.ndarray:
    if kz in _SIGMA:
        return np.array(_SIGMA[kz], dtype=np.float64)
    s = np.linspace(0.0, 1.0, kz + 1)
    return 1.0 - (1.0 - s) ** 1.5  # finer spacing near the surface


# named workloads (C2, C3, C4) + small cases for the test-suite
CASES = {
    # name: jx, iy, kz, ds[km], dt[s], nspg
    "tiny": dict(jx=20, iy=16, kz=8, ds=60.0, dt=120.0, nspg=4),
    "small": dict(jx=40, iy=30, kz=14, ds=60.0, dt=120.0, nspg=6),
    "C1": dict(jx=64, iy=34, kz=18, ds=60.0, dt=150.0, nspg=12, idynamic=1),  # Testing/test_001.in shape, hydrostatic
    "C2": dict(jx=64, iy=34, kz=18, ds=60.0, dt=120.0, nspg=8),
    "C3": dict(jx=250, iy=200, kz=23, ds=50.0, dt=120.0, nspg=12),
    "C4": dict(jx=1000, iy=1000, kz=41, ds=3.0, dt=6.0, nspg=12),
    "C4q": dict(jx=500, iy=500, kz=41, ds=3.0, dt=6.0, nspg=12),
    "C4e": dict(jx=500, iy=250, kz=41, ds=3.0, dt=6.0, nspg=12),   # one tile of C4 on 8 ranks (4 x 2): tuning the N = 8 tile size on one GPU
    # hydrostatic core at sizes where the kernels are HBM-bound (kernel roofline; not BASELINE configs)
    "C3h": dict(jx=250, iy=200, kz=23, ds=50.0, dt=120.0, nspg=12, idynamic=1),
    "C4h": dict(jx=1000, iy=1000, kz=41, ds=3.0, dt=6.0, nspg=12, idynamic=1),
    "C4hq": dict(jx=500, iy=500, kz=41, ds=3.0, dt=6.0, nspg=12, idynamic=1),
    # MOLOCH core (idynamic = 3): C5 is BASELINE configs[4] (12 km, ~500x500x30, mo_nadv=3, mo_nsound=6)
    "tinym": dict(jx=20, iy=16, kz=10, ds=12.0, dt=90.0, nspg=4, idynamic=3),
    "smallm": dict(jx=40, iy=30, kz=14, ds=12.0, dt=90.0, nspg=6, idynamic=3),
    "C5": dict(jx=500, iy=500, kz=30, ds=12.0, dt=90.0, nspg=12, idynamic=3),
    "C5q": dict(jx=250, iy=250, kz=30, ds=12.0, dt=90.0, nspg=12, idynamic=3),
}


@dataclass
class Case:
    name: str
    cfg: capi.RcmCfg
    statics: Dict[int, np.ndarray] = field(default_factory=dict)
    bdy: Dict[int, Dict[int, np.ndarray]] = field(default_factory=dict)  # slot -> field -> array
    lb: tuple = (1, 1)

    @property
    def cells(self) -> int:
        return self.cfg.jx * self.cfg.iy * self.cfg.kz


def _lambert_msf(lat_deg: np.ndarray, tl1=30.0, tl2=60.0) -> np.ndarray:
    d2r = np.pi / 180.0
    p1, p2 = tl1 * d2r, tl2 * d2r
    n = np.log(np.cos(p1) / np.cos(p2)) / np.log(np.tan(np.pi / 4 + p2 / 2) / np.tan(np.pi / 4 + p1 / 2))
    phi = lat_deg * d2r
    return (np.cos(p1) / np.cos(phi)) * (np.tan(np.pi / 4 + p1 / 2) / np.tan(np.pi / 4 + phi / 2)) ** n


def _hash_noise(fid: int, j, i, k) -> np.ndarray:
    """Counter-based uniform noise in [-1,1): a pure function of the GLOBAL grid
    index (splitmix64 finaliser), so that any window of any decomposition sees
    identical data."""
    with np.errstate(over="ignore"):
        x = (np.uint64(SEED) + np.uint64(fid) * np.uint64(0x9E3779B97F4A7C15)
             + j.astype(np.uint64) * np.uint64(0xBF58476D1CE4E5B9)
             + i.astype(np.uint64) * np.uint64(0x94D049BB133111EB)
             + k.astype(np.uint64) * np.uint64(0xD6E8FEB86659FD93))
        x ^= x >> np.uint64(30); x *= np.uint64(0xBF58476D1CE4E5B9)
        x ^= x >> np.uint64(27); x *= np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
    return (x >> np.uint64(11)).astype(np.float64) * (2.0 / 9007199254740992.0) - 1.0


def make_case(name: str = "C2", *, flat: bool = False, unit_map: bool = False, noise: bool = True,
              ibdyfrq: int = 6, window=None, records=(capi.BDY_B0, capi.BDY_B1, capi.BDY_NEXT),
              **cfg_over) -> Case:
    """`window` = (j1, j2, i1, i2): inclusive GLOBAL dot-index range to generate
    (default: the whole domain).  Cross-point arrays cover j1..min(j2, jx-1)."""
    p = dict(CASES[name])
    for k in ("jx", "iy", "kz", "ds", "dt"):
        if k in cfg_over:
            p[k] = cfg_over.pop(k)
    jx, iy, kz = p["jx"], p["iy"], p["kz"]
    idynamic = cfg_over.pop("idynamic", p.get("idynamic", 2))
    hydro = idynamic == 1
    cfg = capi.default_cfg(idynamic, jx=jx, iy=iy, kz=kz, ds=p["ds"], dt=p["dt"],
                           nspgx=p["nspg"], nspgd=p["nspg"], ibdyfrq=ibdyfrq)
    for k, v in cfg_over.items():
        setattr(cfg, k, v)
    if idynamic == 3:
        case = _make_moloch_case(name, cfg, flat=flat, unit_map=unit_map, noise=noise, ibdyfrq=ibdyfrq, records=records)
        return _window_moloch(case, window) if window is not None else case
    j1, j2, i1, i2 = window or (1, jx, 1, iy)
    j1, i1, j2, i2 = max(1, j1), max(1, i1), min(jx, j2), min(iy, i2)
    jc2, ic2 = min(j2, jx - 1), min(i2, iy - 1)
    dx = cfg.ds * 1000.0
    rearth = 6.371229e6
    clat, clon = 45.0, 13.0
    # dot points j=1..jx, i=1..iy ; cross point (j,i) = centre of dots (j..j+1, i..i+1)
    jd, idd = np.arange(j1, j2 + 1), np.arange(i1, i2 + 1)
    jc, ic = np.arange(j1, jc2 + 1), np.arange(i1, ic2 + 1)
    xd = (jd - 0.5 * (jx + 1)) * dx
    yd = (idd - 0.5 * (iy + 1)) * dx
    xc, yc = (jc - 0.5 * (jx + 1)) * dx + 0.5 * dx, (ic - 0.5 * (iy + 1)) * dx + 0.5 * dx
    Xd, Yd = np.meshgrid(xd, yd, indexing="ij")
    Xc, Yc = np.meshgrid(xc, yc, indexing="ij")
    # latitude span limited to +-20 deg about 45N so that the analytic Lambert map
    # factor stays in a realistic range (~0.97-1.07) on continental-size domains
    half_span = np.degrees(0.5 * iy * dx / rearth)
    fac = min(1.0, 20.0 / half_span)
    dlat = clat + np.degrees(Yd / rearth) * fac
    xlat = clat + np.degrees(Yc / rearth) * fac
    xlon = clon + np.degrees(Xc / (rearth * np.cos(np.radians(xlat))))
    if unit_map:
        msfd = np.ones_like(Xd); msfx = np.ones_like(Xc); coriol = np.zeros_like(Xd)
    else:
        msfd, msfx = _lambert_msf(dlat), _lambert_msf(xlat)
        coriol = 2.0 * 7.2921159e-5 * np.sin(np.radians(dlat))
    if flat:
        ht = np.zeros_like(Xc)
    else:
        L = 0.15 * min(jx, iy) * dx
        ht = 1500.0 * np.exp(-(Xc ** 2 + Yc ** 2) / L ** 2)
    sigma = sigma_levels(kz)
    case = Case(name=name, cfg=cfg)
    case.lb = (j1, i1)
    F = np.asfortranarray
    case.statics = {capi.S_HT: F(ht), capi.S_MSFX: F(msfx), capi.S_MSFD: F(msfd), capi.S_CORIOL: F(coriol),
                    capi.S_XLAT: F(xlat), capi.S_XLON: F(xlon), capi.S_DLAT: F(dlat), capi.S_SIGMA: sigma}

    # ---- reference-like thermodynamic profile (same closed form as nhbase) ----
    rgas = 6.02214129e23 * 1.3806504e-23 / 28.96454 * 1000.0
    g = 9.80665
    p0, tlp, ts0, ptop_pa = cfg.base_state_pressure, cfg.logp_lrate, cfg.base_state_ts0, cfg.ptop * 1000.0
    hs = 0.5 * (sigma[1:] + sigma[:-1])
    b = ts0 / tlp
    alnp = -b + np.sqrt(b * b - 4.0 * (0.5 * g / rgas * ht / tlp))
    ps0 = p0 * np.exp(alnp) - ptop_pa
    pr0 = ps0[:, :, None] * hs[None, None, :] + ptop_pa
    t0 = np.maximum(ts0 + tlp * np.log(pr0 / p0), 216.65)
    a = np.log(pr0 / p0)
    z = np.maximum(-(0.5 * rgas / g * tlp * a * a + rgas / g * ts0 * a), 0.0)
    del a, pr0

    Lx, Ly = jx * dx, iy * dx
    sh = np.sin(np.pi * hs)[None, None, :]
    kk = np.arange(1, kz + 1)
    if noise:
        Jc, Ic, Kc = np.meshgrid(jc, ic, kk, indexing="ij", sparse=True)
        Jd, Id, Kd = np.meshgrid(jd, idd, kk, indexing="ij", sparse=True)
        nz_t = _hash_noise(1, Jc + 0 * Ic + 0 * Kc, Ic + 0 * Jc + 0 * Kc, Kc + 0 * Jc + 0 * Ic)
        nz_u = _hash_noise(2, Jd + 0 * Id + 0 * Kd, Id + 0 * Jd + 0 * Kd, Kd + 0 * Jd + 0 * Id)
        nz_v = _hash_noise(3, Jd + 0 * Id + 0 * Kd, Id + 0 * Jd + 0 * Kd, Kd + 0 * Jd + 0 * Id)
    else:
        nz_t = nz_u = nz_v = 0.0

    def record(phase: float):
        """fields of one boundary record; `phase` shifts the smooth modes (U*6h)."""
        xs_c, xs_d = Xc[:, :, None] - phase, Xd[:, :, None] - phase
        wave = 2.0 * np.sin(2 * np.pi * xs_c / Lx) * np.cos(2 * np.pi * Yc[:, :, None] / Ly) * sh
        r2 = ((xs_c - 0.2 * Lx) ** 2 + (Yc[:, :, None] + 0.15 * Ly) ** 2) / (0.08 * min(Lx, Ly)) ** 2
        bubble = 1.0 * np.exp(-r2) * sh
        t = t0 + wave + bubble
        t = t * (1.0 + 1.0e-3 * 1.0e-2 * nz_t)   # +-1e-5 relative (~3 mK) noise
        u = 20.0 * sh ** 2 * np.ones(Xd.shape + (1,)) + \
            2.0 * np.sin(2 * np.pi * xs_d / Lx + 0.7) * np.sin(2 * np.pi * Yd[:, :, None] / Ly + 0.3)
        v = 5.0 * np.sin(2 * np.pi * xs_d / Lx) * np.ones((1, 1, kz))
        u = u + 2.0e-2 * nz_u
        v = v + 2.0e-2 * nz_v
        qv = np.maximum(0.012 * np.exp(-z / 2500.0) * (1.0 + 0.2 * np.sin(2 * np.pi * xs_c / Lx)), 1.0e-8)
        tr = {}
        for n in range(cfg.ntr):
            # tracer n: a plume travelling with the record phase over a weak background that decays with height
            # (mixing ratio, kg/kg); plumes of different tracers sit at different places and heights
            cx, cy = (0.25 - 0.15 * n) * Lx, (-0.2 + 0.13 * n) * Ly
            rr = ((xs_c - cx) ** 2 + (Yc[:, :, None] - cy) ** 2) / (0.10 * min(Lx, Ly)) ** 2
            zc = 1500.0 + 1200.0 * n
            tr[capi.F_CHI + n] = F(1.0e-9 * np.exp(-z / 4000.0) + 2.0e-7 * (n + 1) * np.exp(-rr) * np.exp(-((z - zc) / 1500.0) ** 2))
        if hydro:
            # surface pressure [hPa] on the terrain (same closed form as the profile above) with a
            # smooth travelling perturbation, so that the p* boundary tendency is non-zero
            ps = (ps0 + ptop_pa) * 0.01 * (1.0 + 1.0e-3 * np.sin(2 * np.pi * (Xc - phase) / Lx) * np.cos(2 * np.pi * Yc / Ly))
            return {capi.F_U: F(u), capi.F_V: F(v), capi.F_T: F(t), capi.F_QV: F(qv), capi.F_PS: F(ps), **tr}
        pp = np.zeros(Xc.shape + (kz,))
        ww = np.zeros(Xc.shape + (kz + 1,))
        return {capi.F_U: F(u), capi.F_V: F(v), capi.F_T: F(t), capi.F_QV: F(qv),
                capi.F_PP: F(pp), capi.F_W: F(ww), **tr}

    shift = 20.0 * 3600.0 * ibdyfrq * 0.25  # a quarter of the jet advection distance per record
    phases = {capi.BDY_B0: 0.0, capi.BDY_B1: shift, capi.BDY_NEXT: 2 * shift}
    case.bdy = {slot: record(phases[slot]) for slot in records}
    return case


def moloch_levels(cfg):
    """zita coordinate of the MOLOCH core (Share/mod_zita.F90:30-92, Main/mod_params.F90:1169,2802-2813):
    (hzita, dz, zita[kz+1], zitah[kz], ak[kz], bk[kz]), k = 1 at the top."""
    rgas = 6.02214129e23 * 1.3806504e-23 / 28.96454 * 1000.0
    kz = cfg.kz
    hzita = rgas * 280.0 / 9.80665
    dz = hzita / kz
    zita = (kz - np.arange(0, kz + 1)) * dz
    zitah = zita[:-1] - 0.5 * dz
    r = zitah / hzita
    gz = 1.0 - cfg.mo_a0 * r - (3.0 - 2.0 * cfg.mo_a0) * r ** 2 + (2.0 - cfg.mo_a0) * r ** 3
    bz = cfg.mo_b0 + (1.0 - cfg.mo_b0) * r
    ak = -hzita * bz * np.log(1.0 - r)
    return hzita, dz, zita, zitah, ak, gz


def _make_moloch_case(name, cfg, *, flat, unit_map, noise, ibdyfrq, records) -> Case:
    """Synthetic DOMAIN + ICBC for idynamic = 3 on the Arakawa C grid: u on (jx, iy-1), v on (jx-1, iy),
    T/qv on (jx-1, iy-1), ps [hPa] 2-D.  Temperature follows the standard lapse rate in the height of the
    zita levels (as the reference's own analytic cases do, Main/mod_bdycod.F90:5157-5199) plus the same smooth
    wave / bubble / jet perturbations and hashed noise as the sigma-coordinate cases."""
    jx, iy, kz = cfg.jx, cfg.iy, cfg.kz
    dx = cfg.ds * 1000.0
    rearth = 6.371229e6
    clat, clon = 45.0, 13.0
    jd, idd = np.arange(1, jx + 1), np.arange(1, iy + 1)
    jc, ic = np.arange(1, jx), np.arange(1, iy)
    xd, yd = (jd - 0.5 * (jx + 1)) * dx, (idd - 0.5 * (iy + 1)) * dx
    xc, yc = (jc - 0.5 * (jx + 1)) * dx + 0.5 * dx, (ic - 0.5 * (iy + 1)) * dx + 0.5 * dx
    Xc, Yc = np.meshgrid(xc, yc, indexing="ij")
    Xu, Yu = np.meshgrid(xd, yc, indexing="ij")      # u points: dot in x, cross in y
    Xv, Yv = np.meshgrid(xc, yd, indexing="ij")      # v points
    half_span = np.degrees(0.5 * iy * dx / rearth)
    fac = min(1.0, 20.0 / half_span)
    lat = lambda Y: clat + np.degrees(Y / rearth) * fac
    xlat, ulat, vlat = lat(Yc), lat(Yu), lat(Yv)
    xlon = clon + np.degrees(Xc / (rearth * np.cos(np.radians(xlat))))
    if unit_map:
        msfx, msfu, msfv = np.ones_like(Xc), np.ones_like(Xu), np.ones_like(Xv)
        ulat, vlat = np.zeros_like(Xu), np.zeros_like(Xv)      # zero Coriolis (test2 of the reference)
    else:
        msfx, msfu, msfv = _lambert_msf(xlat), _lambert_msf(ulat), _lambert_msf(vlat)
    if flat:
        ht = np.zeros_like(Xc)
    else:
        L = 0.15 * min(jx, iy) * dx
        ht = 1500.0 * np.exp(-(Xc ** 2 + Yc ** 2) / L ** 2)
    hzita, dz, zita, zitah, ak, bk = moloch_levels(cfg)
    sigma = 1.0 - zita / hzita
    case = Case(name=name, cfg=cfg)
    F = np.asfortranarray
    case.statics = {capi.S_HT: F(ht), capi.S_MSFX: F(msfx), capi.S_MSFU: F(msfu), capi.S_MSFV: F(msfv),
                    capi.S_XLAT: F(xlat), capi.S_XLON: F(xlon), capi.S_ULAT: F(ulat), capi.S_VLAT: F(vlat),
                    capi.S_SIGMA: sigma}
    rgas = 6.02214129e23 * 1.3806504e-23 / 28.96454 * 1000.0
    g = 9.80665
    zeta = ak[None, None, :] + (bk[None, None, :] - 1.0) * ht[:, :, None]      # height above the terrain
    zmsl = zeta + ht[:, :, None]
    t0 = np.maximum(288.15 - 0.00649 * zmsl, 210.0)
    ps0 = 1013.25 * (1.0 - 0.00649 * ht / 288.15) ** (g / (rgas * 0.00649))    # hPa on the terrain
    Lx, Ly = jx * dx, iy * dx
    hs = np.clip(zitah / hzita, 0.0, 1.0)
    sh = np.sin(np.pi * hs)[None, None, :]
    kk = np.arange(1, kz + 1)
    if noise:
        def nz(fid, ja, ia):
            J, I, K = np.meshgrid(ja, ia, kk, indexing="ij", sparse=True)
            return _hash_noise(fid, J + 0 * I + 0 * K, I + 0 * J + 0 * K, K + 0 * J + 0 * I)
        nz_t, nz_u, nz_v = nz(1, jc, ic), nz(2, jd, ic), nz(3, jc, idd)
    else:
        nz_t = nz_u = nz_v = 0.0

    def record(phase: float):
        xs_c = Xc[:, :, None] - phase
        wave = 2.0 * np.sin(2 * np.pi * xs_c / Lx) * np.cos(2 * np.pi * Yc[:, :, None] / Ly) * sh
        r2 = ((xs_c - 0.2 * Lx) ** 2 + (Yc[:, :, None] + 0.15 * Ly) ** 2) / (0.08 * min(Lx, Ly)) ** 2
        t = (t0 + wave + 1.0 * np.exp(-r2) * sh) * (1.0 + 1.0e-5 * nz_t)
        u = 15.0 * sh ** 2 * np.ones(Xu.shape + (1,)) + \
            2.0 * np.sin(2 * np.pi * (Xu[:, :, None] - phase) / Lx + 0.7) * np.sin(2 * np.pi * Yu[:, :, None] / Ly + 0.3)
        v = 4.0 * np.sin(2 * np.pi * (Xv[:, :, None] - phase) / Lx) * np.ones((1, 1, kz))
        u = u + 2.0e-2 * nz_u
        v = v + 2.0e-2 * nz_v
        qv = np.maximum(0.010 * np.exp(-zmsl / 2500.0) * (1.0 + 0.2 * np.sin(2 * np.pi * xs_c / Lx)), 1.0e-8)
        ps = ps0 * (1.0 + 1.0e-3 * np.sin(2 * np.pi * (Xc - phase) / Lx) * np.cos(2 * np.pi * Yc / Ly))
        return {capi.F_U: F(u), capi.F_V: F(v), capi.F_T: F(t), capi.F_QV: F(qv), capi.F_PS: F(ps)}

    shift = 15.0 * 3600.0 * ibdyfrq * 0.25
    phases = {capi.BDY_B0: 0.0, capi.BDY_B1: shift, capi.BDY_NEXT: 2 * shift}
    case.bdy = {slot: record(phases[slot]) for slot in records}
    return case


def _window_moloch(case: Case, window) -> Case:
    """Cut the GLOBAL C-grid arrays of a MOLOCH case down to the inclusive dot-index window (j1, j2, i1, i2);
    every array keeps its own grid's upper limit (u: iy-1 rows, v: jx-1 columns, cross: both)."""
    jx, iy = case.cfg.jx, case.cfg.iy
    j1, j2, i1, i2 = window
    j1, i1, j2, i2 = max(1, j1), max(1, i1), min(jx, j2), min(iy, i2)

    def cut(a):
        nj, ni = a.shape[0], a.shape[1]          # the array's own global extent
        return np.asfortranarray(a[j1 - 1:min(j2, nj), i1 - 1:min(i2, ni), ...])
    out = Case(name=case.name, cfg=case.cfg)
    out.lb = (j1, i1)
    out.statics = {sid: (a if sid == capi.S_SIGMA else cut(a)) for sid, a in case.statics.items()}
    out.bdy = {slot: {fid: cut(a) for fid, a in rec.items()} for slot, rec in case.bdy.items()}
    return out


def load_case(core: capi.DynCore, case: Case, stage_next: bool = True) -> None:
    """Feed statics and boundary records into a core and initialise it
    (= read_domain_info + init_bdy + init + first bdyval of the reference)."""
    lb = getattr(case, "lb", (1, 1))
    for sid, a in case.statics.items():
        core.set_static(sid, a, lb=lb)
    for slot in (capi.BDY_B0, capi.BDY_B1) + ((capi.BDY_NEXT,) if stage_next else ()):
        if slot not in case.bdy:
            continue
        for fid, a in case.bdy[slot].items():
            core.set_bdy(fid, slot, a, lb=(lb[0], lb[1], 1))
    core.init()
