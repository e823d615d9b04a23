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


@pytest.mark.parametrize("name", fixture_names(planning=True))
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


def test_algorithmic_bytes_are_derived_from_the_buffer_list_and_match_the_survey():
    """1,516 B read + 2,605 B written = 4,121 B per env-step (post-physics), 816 B per torque sub-step,
    7,385 B per full step at cfg 2 / 5.  The figure is computed from cconfig.state_spec's shapes, so it moves when a buffer does."""
    from legged_tracking_b200 import cconfig, cfg, workload
    flat = cfg.flat_from_cfg(cfg.preset_cfg("cfg2_single_path_64", num_envs=4096), num_envs=4096)
    a = workload.algorithmic_bytes(flat)
    assert sum(a["post_physics_reads"].values()) == 1516
    assert sum(a["post_physics_writes"].values()) + a["post_physics_in_torque_kernel"] == 2605
    assert (a["post_physics"], a["torque_substep"], a["full_step"]) == (4121, 816, 7385)
    names = {n for n, _, _ in cconfig.state_spec(flat)}
    assert set(cconfig.POST_PHYSICS_TRAFFIC) <= names
    # drift: a wider observation row or one more reward term must move the figure
    wider = dict(flat, num_obs=flat["num_obs"] + 1)
    assert workload.algorithmic_bytes(wider)["post_physics"] == 4121 + 8
    # H > 1: the history shift is read + the whole history row written (cfg 4 remark)
    h3 = dict(flat, obs_history_length=3)
    assert workload.algorithmic_bytes(h3)["post_physics"] == 4121 + 2 * 4 * flat["num_obs"] * 2
