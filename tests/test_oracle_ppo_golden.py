"""oracle/ppo_port.py pinned against the vectors recorded from the unmodified reference PPO (tests/golden/ppo_*.npz,
made by oracle/gen_golden_ppo.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import ppo_port

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["ppo_adaptive_64x8", "ppo_fixed_32x6", "ppo_cnn_48x6"]


def load_case(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


def port_from_golden(g):
    init = {k[5:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("init/")}
    args = {k[9:]: (v.item() if v.ndim == 0 else v.tolist()) for k, v in g.items() if k.startswith("ppo_args/")}
    args.pop("selective_adaptation_module_loss", None)
    return ppo_port.PpoPort(init, g["meta/sizes"], activation=str(g["ac_args/activation"]), **args)


def run_rollout(port, g):
    T = int(g["meta/sizes"][1])
    tn = lambda a: torch.from_numpy(np.ascontiguousarray(a))
    for t in range(T):
        port.store(tn(g["in/obs"][t]), tn(g["in/priv"][t]), tn(g["in/hist"][t]), tn(g["in/actions"][t]), tn(g["in/rewards"][t]),
                   tn(g["in/dones"][t]), tn(g["in/time_outs"][t]) if bool(g["meta/time_outs"]) else None)
    port.compute_returns(tn(g["in/hist"][T]), tn(g["in/priv"][T]))


@pytest.mark.parametrize("name", CASES)
def test_rollout_and_gae_match_reference(name):
    g = load_case(name)
    port = port_from_golden(g)
    run_rollout(port, g)
    for k in ("observations", "privileged_observations", "observation_histories", "actions", "dones"):
        assert np.array_equal(port.s[k].numpy(), g["storage/" + k]), k
    for k in ("rewards", "values", "actions_log_prob", "mu", "sigma", "returns", "advantages"):
        np.testing.assert_allclose(port.s[k].numpy(), g["storage/" + k], rtol=1e-5, atol=2e-6, err_msg=k)


@pytest.mark.parametrize("name", CASES)
def test_update_matches_reference(name):
    g = load_case(name)
    port = port_from_golden(g)
    run_rollout(port, g)
    res = port.update(torch.from_numpy(g["update/perm"]))
    np.testing.assert_allclose(res["lrs"], g["update/lrs"], rtol=1e-12)            # every KL-schedule branch taken alike
    np.testing.assert_allclose(res["grad_norms"], g["update/grad_norms"], rtol=2e-4)
    np.testing.assert_allclose(res["losses"], g["update/losses"], rtol=1e-4, atol=1e-7)
    assert port.learning_rate == pytest.approx(float(g["update/final_lr"]), rel=1e-12)
    for k, v in port.params.items():
        np.testing.assert_allclose(v.detach().numpy(), g["final/" + k], rtol=1e-4, atol=2e-5, err_msg=k)
