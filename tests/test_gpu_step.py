"""-m gpu parity tests: the CUDA step path (liblt_b200.so through the C ABI) against the golden fixtures recorded
from the unmodified reference, and against the oracle port on larger seeded inputs.

Tolerances (north_star): termination masks, reset ids and heightfield cell indices bit-exact; fp32 outputs within
1e-5 relative (+2e-6 absolute for quantities that pass through zero; torques: +2e-5, i.e. 1e-5 of the actuator
net's contraction magnitude)."""
import numpy as np
import pytest
import torch

from tests import helpers as T

pytestmark = pytest.mark.gpu


def _engine(fx, **kw):
    from legged_tracking_b200.engine import StepEngine
    eng = StepEngine(fx["flat"], T.sub(fx, "terrain/"), T.load_actuator(), device="cuda:0", export_height_indices=True, **kw)
    eng.load_state(T.sub(fx, "init/"))
    return eng


def _xp(a):
    return torch.as_tensor(a).cuda()


@pytest.mark.parametrize("epw", [0, 4, 8, 16, 32])
@pytest.mark.parametrize("name", T.fixture_names())
def test_cuda_step_reproduces_reference(name, epw):
    fx = T.load_fixture(name)
    eng = _engine(fx, envs_per_warp=epw)
    T.check_engine_state(eng, T.sub(fx, "init/"), "init")
    for k in range(T.num_steps(fx)):
        tq = T.engine_feed_and_step(eng, fx, k, _xp)
        torch.cuda.synchronize()
        T.check_engine_step(eng, fx, k, tq)


def test_library_was_loaded_from_tree():
    from legged_tracking_b200 import _abi
    lib = _abi.load()
    assert _abi.LIB_PATH.endswith("legged_tracking_b200/liblt_b200.so")
    assert lib.lt_abi_version() == _abi.DEFINES["LT_ABI_VERSION"]
