"""Hydrostatic core (idynamic = 1, SURVEY.md §8 a3, a20-a22 and the hydrostatic
branches of a4-a18): pins of the CPU restatement, and the kernel LOGIC check of
the CUDA sources through the test-only host harness (tests/emu).  The oracle and the
product solve the init-time vertical-mode eigenproblem with DIFFERENT solvers
(oracle/linalg_ref.hpp vs regcm_b200/csrc/linalg_small.hpp), so the tables agree
to round-off (<= 1e-11 relative) and the hydrostatic state is compared at 1e-11 instead
of bit for bit; everything that does not depend on the tables still is bit-exact.

The reference ships no golden vectors for this path; the in-tree behavioural
pins it does hold are the `vmodes` self-tests (Main/mod_vmodes.F90:314-334:
linearised hydrostatic identity to 1000 eps; :571-593: real positive equivalent
depths).  They are reproduced here, and the eigen-solver / matrix inverse that
stand in for the vendored EISPACK `rg` / LINPACK `sgefa,sgedi` are checked
against LAPACK (numpy.linalg), which is independent of both implementations."""
import numpy as np
import pytest

from oracle import pyoracle
from regcm_b200 import capi, synth
from tests.helpers import make_pair, rel_errors, smooth_qc, work_errors


def _oracle(name="tiny", **kw):
    case = synth.make_case(name, idynamic=1, **kw)
    o = pyoracle.make_oracle(case.cfg, checked=True)
    synth.load_case(o, case)
    return case, o


def _normalise(z, sigma):
    """vorder/vnorml conventions (Main/mod_vmodes.F90:514-567): largest |component| positive,
    sum_k dsigma_k z_k^2 = 1."""
    ds = np.diff(sigma)
    out = z.copy()
    for l in range(z.shape[1]):
        kmax = np.argmax(np.abs(z[:, l]))
        a = np.sign(z[kmax, l]) / np.sqrt(np.sum(ds * z[:, l] ** 2))
        out[:, l] = a * z[:, l]
    return out


@pytest.mark.parametrize("name,kz", [("tiny", 8), ("small", 14), ("C1", 18)])
def test_vmodes_against_lapack(name, kz):
    case, o = _oracle(name)
    sigma = case.statics[capi.S_SIGMA]
    tau0 = o.get_table(capi.T_TAU0)
    w, v = np.linalg.eig(tau0)
    assert np.abs(w.imag).max() == 0.0 and w.real.min() > 0.0     # vcheke: real positive equivalent depths
    order = np.argsort(-w.real)
    hbar = o.get_table(capi.T_HBAR)
    assert np.abs(hbar - w.real[order]).max() / hbar.max() < 1e-12
    assert np.all(np.diff(hbar) < 0)
    # eigenvectors: first nsplit columns (scaled by fac = 2 dt / (2 aam + 1) in spinit)
    zl = _normalise(v.real[:, order], sigma)
    ns = case.cfg.nsplit
    dt = case.cfg.dt
    zm = o.get_table(capi.T_ZMATX)
    for l in range(ns):
        dtau = dt * 0.5 / (ns - l)
        fac = 2.0 * dt / (2.0 * round(dt / dtau) + 1.0)
        assert np.abs(zm[:, l] / fac - zl[:, l]).max() < 1e-10
    # the remaining columns are unscaled: zmatxr is the inverse of the unscaled matrix
    zfull = zm.copy()
    for l in range(ns):
        dtau = dt * 0.5 / (ns - l)
        zfull[:, l] /= 2.0 * dt / (2.0 * round(dt / dtau) + 1.0)
    zr = o.get_table(capi.T_ZMATXR)
    assert np.abs(zr @ zfull - np.eye(kz)).max() < 1e-9
    assert np.abs(zr[:ns] - np.linalg.inv(zfull)[:ns]).max() / np.abs(zr[:ns]).max() < 1e-9
    # the modes diagonalise tau: zr tau z = diag(hbar) for the leading modes
    d = zr @ tau0 @ zfull
    assert np.abs(d[:ns, :ns] - np.diag(hbar[:ns])).max() / hbar[0] < 1e-10


def test_oracle_hydro_rest_state_stays_at_rest():
    """Horizontally uniform p*, T(sigma), no wind, flat terrain, f = 0: every tendency vanishes
    (cf. the expected outcome of moloch_static_test1, Main/mod_bdycod.F90:5157-5199)."""
    case = synth.make_case("tiny", idynamic=1, flat=True, unit_map=True, noise=False)
    jx, iy, kz = case.cfg.jx, case.cfg.iy, case.cfg.kz
    tcol = case.bdy[capi.BDY_B0][capi.F_T][0:1, 0:1, :].copy()
    qcol = case.bdy[capi.BDY_B0][capi.F_QV][0:1, 0:1, :].copy()
    for slot in case.bdy:   # the same columns in every record: no boundary tendency
        b = case.bdy[slot]
        b[capi.F_U][...] = 0.0
        b[capi.F_V][...] = 0.0
        b[capi.F_T][...] = tcol
        b[capi.F_QV][...] = qcol
        b[capi.F_PS][...] = 1000.0
    o = pyoracle.make_oracle(case.cfg, checked=True)
    synth.load_case(o, case)
    t0, ps0 = o.get_state(capi.F_T), o.get_state(capi.F_PS)
    o.step(10)
    assert np.abs(o.get_state(capi.F_U)).max() < 1e-9 and np.abs(o.get_state(capi.F_V)).max() < 1e-9
    assert np.abs(o.get_state(capi.F_PS) - ps0).max() < 1e-10
    assert np.abs(o.get_state(capi.F_T) - t0).max() / np.abs(t0).max() < 1e-12


def test_oracle_hydro_noise_parameters_decay():
    """Bleck (1977) noise parameters (Main/mod_tendency.F90:1449-1459) are finite, positive and
    decay while the synthetic state adjusts; p* stays within physical bounds."""
    case, o = _oracle("small")
    vals = []
    for _ in range(6):
        o.step(10)
        d = o.diag()
        vals.append((d.ptntot, d.pt2tot))
        ps = o.get_state(capi.F_PS)
        assert 60.0 < ps.min() and ps.max() < 110.0 and d.neg_q_points == 0
    assert all(np.isfinite(v).all() and v[0] > 0 for v in vals)
    assert vals[-1][1] < vals[0][1]


TOL_TABLE = 1e-11   # two independent eigen-solvers / inverses
TOL_STATE = 1e-11   # relative L-inf per prognostic after 8 steps

HYDRO_VARIANTS = [
    {}, {"iboudy": 1}, {"iboudy": 0}, {"iboudy": 2}, {"iboudy": 3}, {"iboudy": 4}, {"ipptls": 0}, {"diffu_hgtf": 0}, {"upstream_mode": 0},
    {"stability_enhance": 0}, {"flat": True, "unit_map": True}, {"noise": False}, {"lstand": 0}, {"nsplit": 1}, {"nsplit": 3},
]


@pytest.mark.parametrize("kw", HYDRO_VARIANTS, ids=lambda k: "-".join(f"{a}{b}" for a, b in k.items()) or "default")
def test_emu_hydro_parity(emu_lib, kw):
    case, core, ref = make_pair(emu_lib, "tiny", oracle_checked=True, idynamic=1, **kw)
    assert np.array_equal(core.get_table(capi.T_TAU0), ref.get_table(capi.T_TAU0))   # the matrix both solvers start from
    for t in (capi.T_HBAR, capi.T_ZMATX, capi.T_ZMATXR, capi.T_TAU, capi.T_VARPA1, capi.T_AM, capi.T_AN):
        a, b = core.get_table(t), ref.get_table(t)
        assert np.abs(a - b).max() <= TOL_TABLE * np.abs(b).max(), t
    core.step(1); ref.step(1)
    assert max(work_errors(core, ref).values()) == 0.0   # the slow tendencies do not read the tables: bit for bit
    core.step(7); ref.step(7)
    errs = rel_errors(core, ref)
    assert max(e[0] for e in errs.values()) <= TOL_STATE, errs
    dc, dr = core.diag(), ref.diag()
    assert abs(dc.ptntot - dr.ptntot) <= 1e-9 * abs(dr.ptntot) and abs(dc.pt2tot - dr.pt2tot) <= 1e-9 * abs(dr.pt2tot)
    assert dc.lcount == dr.lcount == 8


def test_emu_hydro_boundary_record_change(emu_lib):
    case, core, ref = make_pair(emu_lib, "tiny", idynamic=1, ibdyfrq=1)
    core.step(35); ref.step(35)   # record change at step 30 (dt = 120 s)
    assert core.diag().xbctime == ref.diag().xbctime == 6 * 120.0
    assert max(e[0] for e in rel_errors(core, ref).values()) <= 1e-10


def test_emu_hydro_c1_shape(emu_lib):
    case, core, ref = make_pair(emu_lib, "C1")
    assert case.cfg.idynamic == 1 and (case.cfg.jx, case.cfg.iy, case.cfg.kz) == (64, 34, 18)
    core.step(3); ref.step(3)
    assert max(e[0] for e in rel_errors(core, ref).values()) <= TOL_STATE


def test_emu_hydro_cloud_water_path(emu_lib):
    case, core, ref = make_pair(emu_lib, "tiny", idynamic=1)
    qc = smooth_qc(case)
    for c in (core, ref):
        c.set_state(capi.F_QC, capi.LEVEL_A, qc)
        c.set_state(capi.F_QC, capi.LEVEL_B, qc)
    core.step(8); ref.step(8)
    assert max(e[0] for e in rel_errors(core, ref).values()) <= TOL_STATE
    assert np.abs(ref.get_state(capi.F_QC)).max() > 0 and ref.diag().neg_q_points == 0


def test_hydro_rejects_unsupported(emu_lib):
    case = synth.make_case("tiny", idynamic=1)
    case.cfg.nproc, case.cfg.cpus_j, case.cfg.cpus_i = 2, 3, 1   # layout does not match nproc
    with pytest.raises(capi.RcmError):
        capi.DynCore(emu_lib, case.cfg)
    case = synth.make_case("tiny", idynamic=1)
    case.cfg.nsplit = 9             # beyond RCM_MAXSPLIT
    with pytest.raises(capi.RcmError):
        capi.DynCore(emu_lib, case.cfg)
    case = synth.make_case("tiny", idynamic=1)
    case.cfg.idynamic = 4          # no such core
    with pytest.raises(capi.RcmError):
        capi.DynCore(emu_lib, case.cfg)
