"""Generates tests/golden/*.npz from the CPU oracle (oracle/).

The reference ships no golden vectors for the dyn-core ("parity unpinned",
SURVEY.md §8c) and cannot be built here (no Fortran compiler), so these
fixtures pin the oracle restatement against itself over time (regression pin)
and let the GPU tests compare against committed numbers without importing the
oracle.  Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from regcm_b200 import capi, synth  # noqa: E402
from oracle import pyoracle  # noqa: E402

CASES = [("tiny", 12, {}), ("tiny", 6, {"iboudy": 1, "ifupr": 0}), ("small", 4, {})]
# hydrostatic core (idynamic = 1): files hy_*.npz carry u, v, t, qv, qc, ps and the noise parameters
HYDRO_CASES = [("tiny", 12, {}), ("tiny", 6, {"iboudy": 1}), ("small", 4, {})]
HYDRO_FIELDS = (capi.F_U, capi.F_V, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PS)
# MOLOCH core (idynamic = 3): files mo_*.npz carry the single-level C-grid state u, v, w, t, qv, qc, pai, ps
MOLOCH_CASES = [("tinym", 10, {}), ("tinym", 6, {"iboudy": 1}), ("smallm", 3, {})]
MOLOCH_FIELDS = (capi.F_U, capi.F_V, capi.F_W, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PAI, capi.F_PS)


def main():
    for name, nsteps, kw in CASES:
        case = synth.make_case(name, **kw)
        o = pyoracle.make_oracle(case.cfg)
        synth.load_case(o, case)
        o.step(nsteps)
        out = {"nsteps": nsteps}
        for f in range(7):
            for lv in (0, 1):
                # float64, stored exactly
                out[capi.FIELD_NAMES[f] + "_" + "ab"[lv]] = o.get_state(f, lv)
        d = o.diag()
        out["max_cfl"] = d.max_cfl
        out["istep"] = d.istep
        tag = name + "".join(f"_{k}{v}" for k, v in sorted(kw.items()))
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", f"nh_{tag}_{nsteps}.npz"), **out)
        print("wrote", tag, nsteps)
    for name, nsteps, kw in HYDRO_CASES:
        case = synth.make_case(name, idynamic=1, **kw)
        o = pyoracle.make_oracle(case.cfg)
        synth.load_case(o, case)
        o.step(nsteps)
        out = {"nsteps": nsteps}
        for f in HYDRO_FIELDS:
            for lv in (0, 1):
                out[capi.FIELD_NAMES[f] + "_" + "ab"[lv]] = o.get_state(f, lv)
        d = o.diag()
        out["ptntot"] = d.ptntot
        out["pt2tot"] = d.pt2tot
        out["hbar"] = o.get_table(capi.T_HBAR)
        tag = name + "".join(f"_{k}{v}" for k, v in sorted(kw.items()))
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", f"hy_{tag}_{nsteps}.npz"), **out)
        print("wrote hydro", tag, nsteps)
    for name, nsteps, kw in MOLOCH_CASES:
        case = synth.make_case(name, **kw)
        o = pyoracle.make_oracle(case.cfg)
        synth.load_case(o, case)
        o.step(nsteps)
        out = {"nsteps": nsteps}
        for f in MOLOCH_FIELDS:
            out[capi.FIELD_NAMES[f] + "_a"] = o.get_state(f)
        d = o.diag()
        out["ps_max"], out["ps_min"] = d.ps_max, d.ps_min
        tag = name + "".join(f"_{k}{v}" for k, v in sorted(kw.items()))
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", f"mo_{tag}_{nsteps}.npz"), **out)
        print("wrote moloch", tag, nsteps)


if __name__ == "__main__":
    main()
