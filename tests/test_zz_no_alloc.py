"""rcm_dyn_negq_repair_static: the allocation-free, handle-less variant of the negative-moisture repair
(include/regcm_dyn.h; csrc/k_negq_static.cu; no counterpart in the reference)
against the EXISTING entry points, through the C ABI, on two calls in a row.

How the existing path is made to run the same sweep on a known input: with ipptls = 0 the cloud-water tendency is
forced to zero (Main/mod_tendency.F90:331-335), so the leapfrog forecast atmc%qx(iqc) of rcm_dyn_step equals the
level-b field the test stored with rcm_dyn_set_state — including its negative values — and the repaired forecast is
read back with rcm_dyn_get_work(RCM_W_CQC).  The new entry point gets the same field as a device array.
Tolerance: the two paths do the same arithmetic in the same order: bit-equal is expected; the test allows the
1e-11 relative L-inf the other GPU parity tests use.

This file is collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from regcm_b200 import capi, synth

pytestmark = pytest.mark.gpu
TOL = 1e-11


def _field(jx, iy, kz, seed):
    """Cloud water with isolated and ADJACENT negative points (patches, stripes, a checkerboard), some of them on
    the outermost ring and right next to it."""
    rng = np.random.default_rng(seed)
    q = 96.0 * 1e-4 * rng.uniform(0.0, 1.0, size=(jx - 1, iy - 1, kz))
    neg = rng.uniform(size=q.shape) < 0.08
    q[neg] = -q[neg]
    q[3:9, 4:7, :: 2] = -1.0e-3                       # a solid negative patch: the longest dependency chains
    J, I = np.meshgrid(np.arange(jx - 1), np.arange(iy - 1), indexing="ij")
    q[((J + I) % 2 == 0) & (J > jx // 2), kz // 2] *= -1.0
    q[0, :, 0] = -2.0e-3                              # ring values are inputs of the sweep, never repaired
    q[:, -1, 1] = -2.0e-3
    return np.asfortranarray(q)


def _existing_path(lib, q):
    """Repaired forecast and repair count through rcm_dyn_create / set_state / step / get_work / diag."""
    case = synth.make_case("small", ipptls=0)
    core = capi.DynCore(lib, case.cfg)
    synth.load_case(core, case)
    core.set_state(capi.F_QC, capi.LEVEL_A, q)
    core.set_state(capi.F_QC, capi.LEVEL_B, q)
    core.step(1)
    out = core.get_work(capi.W_CQC)
    n = core.diag().neg_q_points
    qv_neg = int((core.get_work(capi.W_CQV) < 0).sum())
    core.close()
    return out, n, qv_neg


def _new_path(lib, torch, q, out_t, cnt_t):
    jm, im, kz = q.shape
    qin = torch.from_numpy(np.ascontiguousarray(q.transpose(2, 1, 0))).cuda()   # (k, i, j) C order = (j, i, k) j fastest
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.negq_repair_static(C.c_void_p(qin.data_ptr()), C.c_void_p(out_t.data_ptr()), jm, im, kz,
                                C.c_void_p(cnt_t.data_ptr()), C.c_void_p(st))
    assert rc == 0
    torch.cuda.synchronize()
    return np.asfortranarray(out_t.cpu().numpy().transpose(2, 1, 0)), int(cnt_t.cpu().numpy().view(np.uint64)[0])


def test_static_scratch_variant_matches_existing_entry_points(cuda_lib):
    import torch
    lib = cuda_lib
    cfg = synth.make_case("small", records=()).cfg
    jx, iy, kz = cfg.jx, cfg.iy, cfg.kz
    out_t = torch.empty((kz, iy - 1, jx - 1), dtype=torch.float64, device="cuda")   # the caller's output, reused
    cnt_t = torch.zeros(1, dtype=torch.int64, device="cuda")
    for call, seed in enumerate((11, 12)):      # two calls in a row: the second must not see the first one's scratch
        q = _field(jx, iy, kz, seed)
        want, nwant, qv_neg = _existing_path(lib, q)
        assert qv_neg == 0                       # so that diag's count is the cloud-water count
        got, ngot = _new_path(lib, torch, q, out_t, cnt_t)
        interior = (slice(1, jx - 2), slice(1, iy - 2))
        assert nwant > 100 and ngot == nwant, (call, ngot, nwant)
        assert want[interior].min() >= 0.0 and got[interior].min() >= 0.0
        # the regular path stores the forecast on the interior only (its sweep reads the ring from level b)
        sc = float(np.abs(want[interior]).max())
        err = float(np.abs(got[interior] - want[interior]).max())
        assert err <= TOL * sc, (call, err / sc)
        # the ring is copied unchanged; untouched points are bit-identical to the input
        assert np.array_equal(got[0], q[0]) and np.array_equal(got[:, -1], q[:, -1])
        keep = q >= 0
        assert np.array_equal(got[keep], q[keep])


def test_static_scratch_variant_rejects_sizes_beyond_the_scratch(cuda_lib):
    import torch
    lib = cuda_lib
    a = torch.zeros(64, dtype=torch.float64, device="cuda")
    b = torch.zeros(64, dtype=torch.float64, device="cuda")
    pa, pb = C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr())
    assert lib.negq_repair_static(pa, pb, 1024, 1024, 2, None, None) == -1      # 2 Mi cells > RCM_NQS_MAX_CELLS
    assert lib.negq_repair_static(pa, pb, 4, 4, 1025, None, None) == -1          # nk > RCM_NQS_MAX_PLANES
    assert lib.negq_repair_static(pa, pb, 2, 4, 1, None, None) == -1             # no interior
    assert lib.negq_repair_static(pa, pa, 4, 4, 1, None, None) == -1             # aliased
    assert lib.negq_repair_static(None, pb, 4, 4, 1, None, None) == -1
    assert lib.negq_repair_static(pa, pb, 4, 4, 4, None, None) == 0              # 64 cells: fits
    torch.cuda.synchronize()
