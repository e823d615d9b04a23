"""The per-env arithmetic of csrc/lt_core.cuh (shared by the CUDA kernels) compiled for the host and replayed
against the golden fixtures recorded from the unmodified reference.  CPU-only scaffolding: it pins the ordering /
rounding semantics of the step logic in the GPU-less container; the kernels themselves are covered by -m gpu."""
import numpy as np
import pytest

from tests import helpers as T
from tests.hostsim_runner import HostSim


@pytest.mark.parametrize("name", T.fixture_names())
def test_hostsim_reproduces_reference(name):
    fx = T.load_fixture(name)
    eng = HostSim(fx["flat"], T.sub(fx, "terrain/"), T.load_actuator())
    eng.load_state(T.sub(fx, "init/"))
    T.check_engine_state(eng, T.sub(fx, "init/"), "init")
    for k in range(T.num_steps(fx)):
        tq = T.engine_feed_and_step(eng, fx, k, lambda a: a)
        T.check_engine_step(eng, fx, k, tq)
