"""Product-side config classes (legged_tracking_b200.cfg): the flat step configuration derived from the product's own
`Cfg` presets must equal, key by key, the one recorded off the LIVE reference env for the same scripts/train.py arguments
(stored in each golden fixture).  When /root/reference is present the same derivation is also run on the reference's own
`Cfg` object (drop-in check: the reference's config tree is accepted unchanged)."""
import numpy as np
import pytest

from legged_tracking_b200 import cfg as ltcfg
from legged_tracking_b200 import flatcfg
from tests.helpers import fixture_names, load_fixture


def diff(a, b):
    a, b = flatcfg.normalise(a), flatcfg.normalise(b)
    return sorted(k for k in set(a) | set(b) if a.get(k) != b.get(k))


@pytest.mark.parametrize("name", fixture_names())
def test_preset_flat_equals_reference_flat(name):
    ref = load_fixture(name)["flat"]
    c = ltcfg.preset_cfg(name, ref["num_envs"])
    assert diff(ltcfg.flat_from_cfg(c), ref) == []
    flatcfg.check_supported(ltcfg.flat_from_cfg(c))


def test_cfg_tree_has_reference_surface():
    c = ltcfg.config_go1(ltcfg.make_cfg())
    for section in ("env", "terrain", "commands", "init_state", "control", "asset", "domain_rand", "rewards", "reward_scales",
                    "normalization", "obs_scales", "noise", "noise_scales", "sim", "curriculum_thresholds"):
        assert isinstance(getattr(c, section), type), section
    assert c.control.decimation == 4 and c.env.num_actions == 12
    scales = vars(c.reward_scales)                   # dict COPY of the public attributes (LR:1863)
    scales["torques"] = 123.0
    assert c.reward_scales.torques != 123.0
    fresh = ltcfg.make_cfg()                         # trees are independent: mutating one leaves the others pristine
    c.env.num_envs = 7
    assert fresh.env.num_envs != 7


def test_unknown_container_and_unfused_terms_are_loud():
    c = ltcfg.preset_cfg("cfg2_single_path_64", 64)
    c.rewards.reward_container_name = "CoRLRewards"
    with pytest.raises(NotImplementedError):
        ltcfg.flat_from_cfg(c)
    c = ltcfg.preset_cfg("cfg2_single_path_64", 64)
    c.reward_scales.my_term = 2.0                    # not a method of the container -> skipped like LR:1392-1394
    assert "my_term" not in ltcfg.flat_from_cfg(c)["reward_names"]
    assert "my_term" in ltcfg.flat_from_cfg(c, extra_reward_names=("my_term",))["reward_names"]


def test_reference_cfg_object_is_accepted_unchanged():
    from oracle import ref_harness
    if not ref_harness.reference_available():
        pytest.skip("reference tree not present on this machine")
    from oracle.gen_golden import FIXTURES
    name = "cfg3_vel_64"
    fx = FIXTURES[name]
    ref = load_fixture(name)["flat"]
    Cfg, _ = ref_harness.build_cfg(fx["argv"], ref["num_envs"], fx.get("num_rows"), fx.get("num_cols"), fx.get("overrides"))
    assert diff(ltcfg.flat_from_cfg(Cfg), ref) == []
