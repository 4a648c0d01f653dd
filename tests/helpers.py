import numpy as np

from regcm_b200 import capi, synth


HYDRO_FIELDS = (capi.F_U, capi.F_V, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PS)
HYDRO_WORK = (capi.W_QDOT, capi.W_XKC, capi.W_OMEGA, capi.W_PHI, capi.W_PTEN)


MOLOCH_FIELDS = (capi.F_U, capi.F_V, capi.F_W, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PAI, capi.F_PS)
MOLOCH_WORK = (capi.W_MO_P, capi.W_MO_TVIRT, capi.W_MO_FMZ, capi.W_MO_FMZF, capi.W_MO_ZETA, capi.W_MO_S)


def prognostics(core):
    """Prognostic fields of the core in use (idynamic = 1 carries p*, no p', w; idynamic = 3: the
    single-level C-grid state of mo_atm)."""
    if core.cfg.idynamic == 3:
        return MOLOCH_FIELDS
    return HYDRO_FIELDS if core.cfg.idynamic == 1 else tuple(range(7))


def rel_errors(core, ref, levels=(capi.LEVEL_A, capi.LEVEL_B)):
    """Per prognostic field: L-inf and L2 error normalised by the reference."""
    out = {}
    if core.cfg.idynamic == 3:
        levels = (capi.LEVEL_A,)
    for f in prognostics(core):
        for lv in levels:
            a, b = core.get_state(f, lv), ref.get_state(f, lv)
            sc = max(float(np.abs(b).max()), 1e-300)
            l2 = float(np.sqrt(np.mean((a - b) ** 2))) / sc
            out[capi.FIELD_NAMES[f] + "ab"[lv]] = (float(np.abs(a - b).max()) / sc, l2)
    return out


def work_errors(core, ref):
    out = {}
    for w in (MOLOCH_WORK if core.cfg.idynamic == 3 else HYDRO_WORK if core.cfg.idynamic == 1 else range(10)):
        a, b = core.get_work(w), ref.get_work(w)
        sc = max(float(np.abs(b).max()), 1e-300)
        out[capi.WORK_NAMES[w]] = float(np.abs(a - b).max()) / sc
    return out


def make_pair(lib, name, oracle_checked=False, **kw):
    from oracle import pyoracle
    case = synth.make_case(name, **kw)
    core = capi.DynCore(lib, case.cfg)
    synth.load_case(core, case)
    ref = pyoracle.make_oracle(case.cfg, checked=oracle_checked)
    synth.load_case(ref, case)
    return case, core, ref


def smooth_qc(case):
    """Smooth positive coupled cloud-water field (kept away from the negative-q guard)."""
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    J, I, K = np.meshgrid(np.arange(jx - 1), np.arange(iy - 1), np.arange(kz), indexing="ij")
    return np.asfortranarray(96.0 * 1e-4 * (1.0 + 0.5 * np.sin(2 * np.pi * J / (jx - 1)) * np.cos(2 * np.pi * I / (iy - 1)))
                             * np.sin(np.pi * (K + 0.5) / kz))


def sharp_moisture(case, seed=7):
    """Coupled qv and qc with sharp fronts: top-hat patches, single-cell spikes and checkerboard patches next to
    dry air.  Advection and the 4th-order diffusion undershoot around them, so the leapfrog forecast goes negative
    at isolated points AND in adjacent groups — the sequential case of Main/mod_tendency.F90:381-393."""
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    rng = np.random.default_rng(seed)
    ps = 96.0
    qv = np.full((jx - 1, iy - 1, kz), ps * 2.0e-6)
    qc = np.zeros((jx - 1, iy - 1, kz))
    for _ in range(max(3, (jx * iy) // 120)):
        j0, i0 = int(rng.integers(2, jx - 6)), int(rng.integers(2, iy - 6))
        w, h = int(rng.integers(1, 4)), int(rng.integers(1, 4))
        k0 = int(rng.integers(0, kz - 2))
        qv[j0:j0 + w, i0:i0 + h, k0:k0 + 3] = ps * 8.0e-3 * rng.uniform(0.5, 1.0)
        qc[j0:j0 + w, i0:i0 + h, k0:k0 + 2] = ps * 1.0e-3 * rng.uniform(0.5, 1.0)
    # checkerboard patch: neighbouring cells alternate between moist and dry
    J, I = np.meshgrid(np.arange(jx - 1), np.arange(iy - 1), indexing="ij")
    cb = ((J + I) % 2 == 0) & (J > jx // 2) & (I > iy // 3) & (J < jx - 4) & (I < iy - 4)
    qc[cb, kz // 2] = ps * 2.0e-3
    qv[cb, kz // 2 + 1] = ps * 6.0e-3
    return np.asfortranarray(qv), np.asfortranarray(qc)
