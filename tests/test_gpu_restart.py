"""Restart through the C ABI on the GPU (tests/test_restart.py on the host harness): the continued
run equals the uninterrupted one bit for bit — same library, same arithmetic, only the state made a round trip
through host memory."""
import pytest

from tests.test_restart import check_restart

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,kw", [("small", {}), ("small", {"ntr": 2}), ("small", {"idynamic": 1}), ("smallm", {})],
                         ids=["nh", "nh-tracers", "hydro", "moloch"])
def test_gpu_restart_continues_bit_for_bit(cuda_lib, name, kw):
    check_restart(cuda_lib, name, 5, 6, **kw)
