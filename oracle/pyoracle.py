"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes loader of the CPU restatement (oracle/liboracle*.so).  May be imported
only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; the product package never imports it."""
import os
import subprocess

from regcm_b200 import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs = {}


def build(force: bool = False) -> None:
    if force or not all(os.path.exists(os.path.join(_HERE, n)) for n in ("liboracle.so", "liboracle_chk.so")):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-j4"])


def oracle_lib(checked: bool = False) -> capi.CoreLib:
    name = "liboracle_chk.so" if checked else "liboracle.so"
    if name not in _libs:
        build()
        _libs[name] = capi.CoreLib(os.path.join(_HERE, name), "orc_")
    return _libs[name]


def make_oracle(cfg, checked: bool = False) -> capi.DynCore:
    return capi.DynCore(oracle_lib(checked), cfg)
