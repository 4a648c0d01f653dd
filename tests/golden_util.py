import os
import re

import numpy as np

from regcm_b200 import capi, synth


def load_golden(path):
    """-> (case, nsteps, {(field, level): array})  for a tests/golden fixture."""
    base = os.path.basename(path)[len("nh_"):-len(".npz")]
    parts = base.split("_")
    name, nsteps = parts[0], int(parts[-1])
    kw = {}
    for p in parts[1:-1]:
        m = re.match(r"([a-z]+)(\d+)$", p)
        kw[m.group(1)] = int(m.group(2))
    z = np.load(path)
    fields = {}
    for f in range(7):
        for lv in (0, 1):
            fields[(f, lv)] = z[capi.FIELD_NAMES[f] + "_" + "ab"[lv]]
    return synth.make_case(name, **kw), nsteps, fields
