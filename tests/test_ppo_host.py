"""Host-side logic of the PPO rows against the reference's recorded vectors (CPU, no kernels): ActorCritic layout and
forward, PPO.update's flat-bucket Adam / clip / KL schedule / adaptation step."""
import numpy as np
import pytest
import torch

from legged_tracking_b200 import ppo
from tests.helpers_ppo import PPO_CASES, HostStorage, actor_critic_from_golden, golden_args, load_case


@pytest.mark.parametrize("name", PPO_CASES)
def test_actor_critic_state_dict_and_forward(name):
    g = load_case(name)
    with golden_args(g):
        ac = actor_critic_from_golden(g)          # strict load: same keys and shapes as the reference's checkpoint
        T = int(g["meta/sizes"][1])
        with torch.no_grad():
            for t in (0, T - 1):
                hist, priv = torch.from_numpy(g["in/hist"][t]), torch.from_numpy(g["in/priv"][t])
                ac.act(hist)
                np.testing.assert_allclose(ac.action_mean.numpy(), g["storage/mu"][t], rtol=1e-5, atol=2e-6)
                np.testing.assert_allclose(ac.action_std.numpy(), g["storage/sigma"][t], rtol=1e-6)
                lp = ac.get_actions_log_prob(torch.from_numpy(g["in/actions"][t]))
                np.testing.assert_allclose(lp.numpy(), g["storage/actions_log_prob"][t, :, 0], rtol=1e-5, atol=2e-6)
                v = ac.evaluate(hist, priv)
                np.testing.assert_allclose(v.numpy(), g["storage/values"][t], rtol=1e-5, atol=2e-6)


@pytest.mark.parametrize("name", PPO_CASES)
def test_update_matches_reference(name):
    g = load_case(name)
    with golden_args(g):
        ac = actor_critic_from_golden(g)
        alg = ppo.PPO(ac, device="cpu")
        alg.storage = HostStorage.from_golden(g)
        losses = alg.update(indices=torch.from_numpy(g["update/perm"]))
        tr = alg.last_update_trace
        np.testing.assert_allclose(tr["lrs"], g["update/lrs"], rtol=1e-12)
        np.testing.assert_allclose(tr["grad_norms"], g["update/grad_norms"], rtol=2e-4)
        np.testing.assert_allclose(losses, g["update/losses"], rtol=1e-4, atol=1e-7)
        assert alg.learning_rate == pytest.approx(float(g["update/final_lr"]), rel=1e-12)
        assert alg.storage.cleared
        for k, v in ac.state_dict().items():
            np.testing.assert_allclose(v.numpy(), g["final/" + k], rtol=1e-4, atol=2e-5, err_msg=k)


def test_flat_params_alias_and_checkpoint_roundtrip():
    g = load_case(PPO_CASES[0])
    with golden_args(g):
        ac = actor_critic_from_golden(g)
        before = {k: v.clone() for k, v in ac.state_dict().items()}
        fp = ppo.FlatParams(ac)
        for k, v in ac.state_dict().items():
            assert torch.equal(v, before[k])
        lo, hi = fp.slice_of(ac.adaptation_module)
        assert hi - lo == sum(p.numel() for p in ac.adaptation_module.parameters())
        fp.flat.data.zero_()                          # the flat buffer IS the parameters
        assert all(float(p.detach().abs().sum()) == 0.0 for p in ac.parameters())
        ac.load_state_dict(before)                    # loading a checkpoint writes through the views
        assert float(fp.flat.data.abs().sum()) > 0.0


def test_storage_has_no_cpu_path():
    from legged_tracking_b200._abi import LtError
    with pytest.raises(LtError):
        ppo.RolloutStorage(4, 2, [3], [1], [6], [2], device="cpu")
