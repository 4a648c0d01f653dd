"""The C-ABI library loads and exports every symbol include/regcm_dyn.h declares
(no compute calls: runs without a GPU)."""
import ctypes as C
import os
import re

from regcm_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "regcm_dyn.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(rcm_dyn_\w+)\s*\(", hdr)))


def test_header_symbols_exported(built):
    lib = C.CDLL(capi.CUDA_LIB_PATH)
    syms = _declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/regcm_dyn.h but not exported"


def test_cfg_layout_matches_c(built):
    lib = C.CDLL(capi.CUDA_LIB_PATH)
    c = capi.RcmCfg()
    lib.rcm_dyn_default_cfg.argtypes = [C.POINTER(capi.RcmCfg), C.c_int]
    assert lib.rcm_dyn_default_cfg(C.byref(c), 2) == 0
    assert c.struct_size == C.sizeof(capi.RcmCfg)
    p = capi.default_cfg(2)
    for name, _ in capi.RcmCfg._fields_:
        if name.startswith("reserved") or name in ("jx", "iy", "kz", "ds", "dt"):
            continue
        assert getattr(c, name) == getattr(p, name), name


def test_create_rejects_bad_config(built):
    lib = capi.cuda_lib()
    bad = capi.default_cfg(2, jx=64, iy=34, kz=18, ds=60.0, dt=120.0)
    bad.struct_size = 4
    h = C.c_void_p()
    assert lib.create(C.byref(bad), C.byref(h)) == -1 and not h
    bad = capi.default_cfg(2, jx=64, iy=34, kz=18, ds=60.0, dt=120.0, iboudy=7)   # no such boundary option
    assert lib.create(C.byref(bad), C.byref(h)) == -1 and not h
    bad = capi.default_cfg(3, jx=64, iy=34, kz=18, ds=12.0, dt=90.0, iboudy=3)    # inflow/outflow without sponge: sigma cores only
    assert lib.create(C.byref(bad), C.byref(h)) == -1 and not h
    bad = capi.default_cfg(4, jx=64, iy=34, kz=18, ds=60.0, dt=120.0)   # no such core
    assert lib.create(C.byref(bad), C.byref(h)) == -1 and not h
    mol = capi.default_cfg(3, jx=64, iy=34, kz=18, ds=12.0, dt=90.0, mo_nsound=0)
    assert lib.create(C.byref(mol), C.byref(h)) == -1 and not h
    hyd = capi.default_cfg(1, jx=64, iy=34, kz=18, ds=60.0, dt=150.0, nsplit=9)
    assert lib.create(C.byref(hyd), C.byref(h)) == -1 and not h
    # a valid hydrostatic configuration passes the argument checks: without a device the
    # library refuses with RCM_ERR_CUDA (no CPU fallback), with one it creates a handle
    hyd = capi.default_cfg(1, jx=64, iy=34, kz=18, ds=60.0, dt=150.0)
    rc = lib.create(C.byref(hyd), C.byref(h))
    assert rc in (0, -2)
    if rc == 0:
        lib.destroy(h)


def test_product_has_no_cpu_fallback():
    """The package never imports the oracle or the host harness."""
    pkg = os.path.join(ROOT, "regcm_b200")
    for dp, _, fs in os.walk(pkg):
        for fn in fs:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dp, fn)).read()
                assert "oracle" not in src.replace("the CPU oracle", "").replace("CPU oracle", "") or fn in ("capi.py",), fn
                assert "libregcm_dyn_emu" not in src, fn
