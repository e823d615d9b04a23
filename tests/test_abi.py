"""CPU checks of the C-ABI boundary: the library builds/loads, exports every function the header declares, and the
generated ctypes mirror agrees with the compiled struct sizes.  No compute calls (no GPU here)."""
import ctypes as C
import os

import pytest

from legged_tracking_b200 import _abi, build, cconfig
from tests import helpers as T


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _abi.load()


def test_exports_every_declared_symbol(lib):
    declared = _abi.declared_functions()
    assert set(declared) == set(_abi.EXPORTS), (sorted(set(declared) ^ set(_abi.EXPORTS)))
    for name in declared:
        assert hasattr(lib, name), f"liblt_b200.so does not export {name}"


def test_struct_sizes_agree(lib):
    assert lib.lt_config_bytes() == C.sizeof(_abi.lt_config)
    assert lib.lt_abi_version() == _abi.DEFINES["LT_ABI_VERSION"]
    # pointer structs: 8 bytes per member
    for s in (_abi.lt_state_ptrs, _abi.lt_sim_ptrs, _abi.lt_terrain_ptrs, _abi.lt_step_args, _abi.lt_tables):
        assert C.sizeof(s) == 8 * len(s._fields_)


@pytest.mark.parametrize("name", T.fixture_names())
def test_config_from_fixture(lib, name):
    fx = T.load_fixture(name)
    cfg = cconfig.make_lt_config(fx["flat"])
    assert lib.lt_num_uniform_slots(C.byref(cfg)) == cconfig.rng_n_slots(fx["flat"])
    spec = cconfig.state_spec(fx["flat"])
    assert [s[0] for s in spec] == [m for m, _ in _abi.lt_state_ptrs._fields_]


def test_create_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    fx = T.load_fixture(T.fixture_names()[0])
    cfg = cconfig.make_lt_config(fx["flat"])
    tables = cconfig.Tables(fx["flat"], T.load_actuator())
    h = C.c_void_p()
    rc = lib.lt_create(C.byref(cfg), C.byref(tables.c), C.byref(h))
    assert rc == _abi.ENUMS["LT_ERR_NO_DEVICE"] and not h.value
    assert b"no CPU fallback" in lib.lt_last_error(None)
    from legged_tracking_b200.engine import StepEngine
    with pytest.raises(_abi.LtError):
        StepEngine(fx["flat"], T.sub(fx, "terrain/"), T.load_actuator(), device="cpu")


def test_rejects_abi_drift(lib):
    fx = T.load_fixture(T.fixture_names()[0])
    cfg = cconfig.make_lt_config(fx["flat"])
    cfg.struct_bytes -= 4
    tables = cconfig.Tables(fx["flat"], T.load_actuator())
    h = C.c_void_p()
    assert lib.lt_create(C.byref(cfg), C.byref(tables.c), C.byref(h)) == _abi.ENUMS["LT_ERR_INVALID"]
