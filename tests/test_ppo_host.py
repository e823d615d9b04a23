"""Host side of the PPO rows without a GPU: the flat parameter layout under the reference's state_dict keys (checked against
the checkpoint recorded from the unmodified reference), initialisation order, and that nothing computes on the CPU."""
import numpy as np
import pytest
import torch

from legged_tracking_b200 import ppo
from legged_tracking_b200._abi import LtError
from tests.helpers_ppo import PPO_CASES, actor_critic_from_golden, golden_args, load_case


@pytest.mark.parametrize("name", PPO_CASES)
def test_state_dict_keys_shapes_and_flat_layout(name):
    g = load_case(name)
    with golden_args(g):
        ac = actor_critic_from_golden(g)          # strict load: same keys and shapes as the reference's checkpoint
        ref_keys = [k[5:] for k in g if k.startswith("init/")]
        sd = ac.state_dict()
        assert sorted(sd) == sorted(ref_keys)
        o = 0
        for k, v in sd.items():                   # views of ONE vector, in lt_policy.h's order, no gaps
            assert np.array_equal(v.numpy(), g["init/" + k]), k
            assert v.data_ptr() == ac.flat.data_ptr() + 4 * o, k
            o += v.numel()
        assert o == ac.num_params == ac.flat.numel()
        assert list(sd)[0] == "std"
        assert [k.split(".")[0] for k in list(sd)[1:]] == sorted((k.split(".")[0] for k in list(sd)[1:]),
                                                                 key=["height_map_encoder", "adaptation_module", "actor_body", "critic_body"].index)
        lo, hi = ac.body_slices["adaptation_module"]
        first = ac.body_slices["height_map_encoder"][0] if ac.encoder else lo       # the encoder sits in front of the adaptation module
        assert first == ac.num_actions and hi == ac.body_slices["actor_body"][0]


def test_default_initialisation_follows_the_reference_draw_order():
    """nn.Linear default init, layers drawn adaptation -> actor -> critic (actor_critic.py:43-82): same seed, same weights."""
    torch.manual_seed(7)
    ac = ppo.ActorCritic(10, 3, 20, 4)
    torch.manual_seed(7)
    dims = {"adaptation_module": [20] + ppo.AC_Args.adaptation_module_branch_hidden_dims + [3],
            "actor_body": [23] + ppo.AC_Args.actor_hidden_dims + [4], "critic_body": [23] + ppo.AC_Args.critic_hidden_dims + [1]}
    sd = ac.state_dict()
    for name in ("adaptation_module", "actor_body", "critic_body"):
        for l, (a, b) in enumerate(zip(dims[name][:-1], dims[name][1:])):
            lin = torch.nn.Linear(a, b)
            assert torch.equal(sd[f"{name}.{2 * l}.weight"], lin.weight.detach())
            assert torch.equal(sd[f"{name}.{2 * l}.bias"], lin.bias.detach())
    assert torch.equal(sd["std"], torch.full((4,), float(ppo.AC_Args.init_noise_std)))


def test_load_state_dict_is_strict_and_writes_through():
    ac = ppo.ActorCritic(10, 3, 20, 4)
    sd = {k: torch.zeros_like(v) for k, v in ac.state_dict().items()}
    ac.load_state_dict(sd)
    assert float(ac.flat.abs().sum()) == 0.0
    bad = dict(sd)
    bad.pop("std")
    with pytest.raises(RuntimeError):
        ac.load_state_dict(bad)
    bad = dict(sd)
    bad["actor_body.0.weight"] = torch.zeros(3, 3)
    with pytest.raises(RuntimeError):
        ac.load_state_dict(bad)


def test_exported_module_matches_reference_layout():
    """Runner.save's TorchScript export: Linear / ELU pairs with the reference's module indices."""
    ac = ppo.ActorCritic(10, 3, 20, 4)
    seq = ac.export_module("actor_body")
    assert [type(m).__name__ for m in seq] == ["Linear", "ELU", "Linear", "ELU", "Linear", "ELU", "Linear"]
    for k, v in seq.state_dict().items():
        assert torch.equal(v, ac.state_dict()["actor_body." + k])
    torch.jit.script(seq)


def test_cnn_variant_geometry_and_unbuilt_forms_are_loud():
    """ppo_cse_cnn (row f4): policy_input = [s, s, emb] (ppo_cse_cnn/actor_critic.py:88-91, 194-197), draw order encoder first;
    the Conv2d and GRU forms raise instead of falling back to torch."""
    from legged_tracking_b200 import ppo_cnn
    A = ppo_cnn.AC_Args
    saved = {k: getattr(A, k) for k in vars(A)}
    try:
        A.height_map_shape, A.cnn_num_embedding = (2, 3, 4), 16
        torch.manual_seed(3)
        ac = ppo_cnn.ActorCritic(30, 2, 60, 5)
        assert ac.policy_input_dim == 2 * (30 - 24) + 16 and ac.lstm_input_dim == 30 - 24 + 16
        sd = ac.state_dict()
        assert sd["height_map_encoder.model.0.weight"].shape == (256, 24) and sd["height_map_encoder.model.2.weight"].shape == (16, 256)
        assert sd["adaptation_module.0.weight"].shape[1] == 28 and sd["actor_body.0.weight"].shape[1] == 30
        torch.manual_seed(3)
        lin = torch.nn.Linear(24, 256)
        assert torch.equal(sd["height_map_encoder.model.0.weight"], lin.weight.detach())
        assert [type(m).__name__ for m in ac.export_module("height_map_encoder.model")] == ["Linear", "ELU", "Linear", "ELU"]
        for flag in ("use_cnn", "use_gru"):
            setattr(A, flag, True)
            with pytest.raises(NotImplementedError):
                ppo_cnn.ActorCritic(30, 2, 60, 5)
            setattr(A, flag, False)
        A.height_map_shape = (2, 21, 11)                # 462 cells out of a 261-wide row: the reference's slice / reshape breaks too
        with pytest.raises(ValueError):
            ppo_cnn.ActorCritic(261, 2, 261, 12)
    finally:
        for k, v in saved.items():
            setattr(A, k, v)


def test_no_cpu_path():
    ac = ppo.ActorCritic(10, 3, 20, 4)
    with pytest.raises(LtError):
        ac.act(torch.zeros(2, 20))
    with pytest.raises(LtError):
        ac.evaluate(torch.zeros(2, 20), torch.zeros(2, 3))
    with pytest.raises(LtError):
        ppo.PPO(ac, device="cpu")
    with pytest.raises(LtError):
        ppo.RolloutStorage(4, 2, [3], [1], [6], [2], device="cpu")
