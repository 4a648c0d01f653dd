"""Multi-GPU parity (needs >= 2 GPUs; skipped otherwise): the NCCL halo exchange of the
product library against the same library on one GPU, and against the CPU oracle."""
import numpy as np
import pytest

from regcm_b200 import capi, synth
from test_multirank_emu import assemble, compare, run_ranks

pytestmark = pytest.mark.gpu


def ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def single_gpu(cuda_lib, case_name, steps):
    case = synth.make_case(case_name)
    core = capi.DynCore(cuda_lib, case.cfg)
    synth.load_case(core, case)
    core.step(steps)
    ref = {(f, "ab"[lv]): core.get_state(f, lv) for f in range(7) for lv in (0, 1)}
    core.close()
    return case, ref


@pytest.mark.parametrize("world,layout,case_name", [(2, (0, 0), "small"), (2, (1, 2), "small"), (4, (2, 2), "C2"), (8, (0, 0), "C2")])
def test_nccl_halo_matches_single_gpu(cuda_lib, tmp_path, world, layout, case_name):
    if ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    steps = 6
    run_ranks(world, tmp_path, ["--case", case_name, "--steps", str(steps), "--cpus-j", str(layout[0]), "--cpus-i", str(layout[1])],
              lib="cuda")
    case, ref = single_gpu(cuda_lib, case_name, steps)
    glob = assemble(tmp_path, world, case.cfg)
    compare(glob, ref, 1e-12)
    # and against the CPU oracle (same tolerance as the single-GPU parity tests)
    from oracle import pyoracle
    orc = pyoracle.make_oracle(case.cfg)
    synth.load_case(orc, case)
    orc.step(steps)
    oref = {(f, "ab"[lv]): orc.get_state(f, lv) for f in range(7) for lv in (0, 1)}
    compare(glob, oref, 1e-10)
