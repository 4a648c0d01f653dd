import os
import re

import numpy as np

from regcm_b200 import capi, synth


def load_golden(path):
    """-> (case, nsteps, {(field, level): array})  for a tests/golden fixture."""
    hydro = os.path.basename(path).startswith("hy_")
    if os.path.basename(path).startswith("mo_"):
        return _load_moloch(path)
    base = os.path.basename(path)[len("nh_"):-len(".npz")]
    parts = base.split("_")
    name, nsteps = parts[0], int(parts[-1])
    kw = {}
    for p in parts[1:-1]:
        m = re.match(r"([a-z]+)(\d+)$", p)
        kw[m.group(1)] = int(m.group(2))
    z = np.load(path)
    fields = {}
    for f in ((capi.F_U, capi.F_V, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PS) if hydro else range(7)):
        for lv in (0, 1):
            fields[(f, lv)] = z[capi.FIELD_NAMES[f] + "_" + "ab"[lv]]
    if hydro:
        kw["idynamic"] = 1
        fields["ptntot"], fields["pt2tot"], fields["hbar"] = float(z["ptntot"]), float(z["pt2tot"]), z["hbar"]
    return synth.make_case(name, **kw), nsteps, fields


def _load_moloch(path):
    parts = os.path.basename(path)[len("mo_"):-len(".npz")].split("_")
    name, nsteps = parts[0], int(parts[-1])
    kw = {}
    for p in parts[1:-1]:
        m = re.match(r"([a-z]+)(\d+)$", p)
        kw[m.group(1)] = int(m.group(2))
    z = np.load(path)
    fields = {(f, 0): z[capi.FIELD_NAMES[f] + "_a"]
              for f in (capi.F_U, capi.F_V, capi.F_W, capi.F_T, capi.F_QV, capi.F_QC, capi.F_PAI, capi.F_PS)}
    fields["ps_max"], fields["ps_min"] = float(z["ps_max"]), float(z["ps_min"])
    return synth.make_case(name, **kw), nsteps, fields
