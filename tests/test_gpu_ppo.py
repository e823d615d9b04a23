"""GPU parity of the PPO rows through the C ABI: lt_rollout_store, lt_gae,
lt_normalize_advantages and the PPO update, against the vectors recorded from the unmodified reference
(tests/golden/ppo_*.npz) and against oracle/ppo_port.py on other seeds and sizes.
Tolerances: copies and masks bit-exact; fp32 arithmetic 1e-5 relative (+2e-6 absolute); quantities that went through
20 Adam steps of tcgen05 split-fp16 vs MKL fp32 GEMMs 1e-3."""
import numpy as np
import pytest
import torch

from tests.helpers_ppo import PPO_CASES, actor_critic_from_golden, golden_args, load_case

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-5, 2e-6


def dev():
    return torch.device("cuda", 0)


def replay_rollout(alg, g):
    d = dev()
    N, T, no, npv, H, na = (int(x) for x in g["meta/sizes"])
    alg.init_storage(N, T, [no], [npv], [no * H], [na])
    tn = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(d)
    for t in range(T):
        a = alg.act(tn(g["in/obs"][t]), tn(g["in/priv"][t]), tn(g["in/hist"][t]), actions=tn(g["in/actions"][t]))
        assert a.data_ptr() == alg.storage.actions[t].data_ptr()          # written in place, row t of the rollout buffer
        infos = {"time_outs": tn(g["in/time_outs"][t])} if bool(g["meta/time_outs"]) else {}
        alg.process_env_step(tn(g["in/rewards"][t]), tn(g["in/dones"][t]), infos)
    alg.compute_returns(tn(g["in/hist"][T]), tn(g["in/priv"][T]))


@pytest.mark.parametrize("name", PPO_CASES)
def test_rollout_store_and_gae_match_reference(name):
    from legged_tracking_b200 import ppo
    g = load_case(name)
    with golden_args(g):
        alg = ppo.PPO(actor_critic_from_golden(g, dev()), device=dev())
        replay_rollout(alg, g)
        st = alg.storage
        for k in ("observations", "privileged_observations", "observation_histories", "actions", "dones"):
            assert np.array_equal(getattr(st, k).cpu().numpy(), g["storage/" + k]), k
        for k in ("rewards", "values", "actions_log_prob", "mu", "sigma", "returns"):
            np.testing.assert_allclose(getattr(st, k).cpu().numpy(), g["storage/" + k], rtol=RTOL, atol=ATOL, err_msg=k)
        np.testing.assert_allclose(st.advantages.cpu().numpy(), g["storage/advantages"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("name", PPO_CASES)
def test_ppo_update_matches_reference(name):
    from legged_tracking_b200 import ppo
    g = load_case(name)
    with golden_args(g):
        ac = actor_critic_from_golden(g, dev())
        alg = ppo.PPO(ac, device=dev())
        replay_rollout(alg, g)
        losses = alg.update(indices=torch.from_numpy(g["update/perm"]).to(dev()))
        tr = alg.last_update_trace
        np.testing.assert_allclose(tr["lrs"], g["update/lrs"], rtol=1e-6)           # same KL-schedule branch every minibatch (the
        #                                                                             learning rate lives on the device as fp32)
        np.testing.assert_allclose(tr["grad_norms"], g["update/grad_norms"], rtol=1e-3)
        np.testing.assert_allclose(losses, g["update/losses"], rtol=1e-3, atol=1e-6)
        for k, v in ac.state_dict().items():
            np.testing.assert_allclose(v.cpu().numpy(), g["final/" + k], rtol=1e-3, atol=1e-4, err_msg=k)


@pytest.mark.parametrize("T,N,seed", [(24, 4096, 0), (5, 37, 1), (1, 1, 2), (24, 1000, 3)])
def test_gae_kernel_matches_oracle(T, N, seed):
    import ctypes as C
    from legged_tracking_b200 import _abi
    from oracle import ppo_port
    lib = _abi.load()
    gen = torch.Generator().manual_seed(seed)
    rewards = torch.randn(T, N, 1, generator=gen) * 0.3
    values = torch.randn(T, N, 1, generator=gen)
    dones = (torch.rand(T, N, 1, generator=gen) < 0.1).to(torch.uint8)
    last = torch.randn(N, 1, generator=gen)
    d = dev()
    R, V, D, L = rewards.to(d), values.to(d), dones.to(d), last.to(d)
    ret, adv = torch.empty_like(R), torch.empty_like(R)
    stats = torch.zeros(3, dtype=torch.float64, device=d)
    p = lambda t: C.c_void_p(t.data_ptr())
    _abi.check(lib.lt_gae(p(R), p(V), p(D), p(L), p(ret), p(adv), p(stats), T, N, C.c_float(0.99), C.c_float(0.95), None))
    raw = adv.clone()
    if T * N > 1:
        _abi.check(lib.lt_normalize_advantages(p(adv), p(stats), T * N, None))
    torch.cuda.synchronize()
    # returns: same op order as the reference -> bit-exact
    T_ret, T_adv = ppo_port.gae(rewards, values, dones, last, 0.99, 0.95)
    assert torch.equal(ret.cpu(), T_ret)
    assert torch.equal(raw.cpu(), T_ret - values)
    assert float(stats[2]) == T * N
    np.testing.assert_allclose(float(stats[0]), float((T_ret - values).double().sum()), rtol=1e-12, atol=1e-12)
    if T * N > 1:
        np.testing.assert_allclose(adv.cpu().numpy(), T_adv.numpy(), rtol=1e-5, atol=1e-5)


def test_gae_full_size_properties():
    """BASELINE size (24 x 65536): normalised advantages have zero mean / unit std; done steps cut the recursion."""
    import ctypes as C
    from legged_tracking_b200 import _abi
    lib = _abi.load()
    T, N = 24, 65536
    d = dev()
    gen = torch.Generator(device=d).manual_seed(5)
    R = torch.randn(T, N, 1, generator=gen, device=d) * 0.3
    V = torch.randn(T, N, 1, generator=gen, device=d)
    D = (torch.rand(T, N, 1, generator=gen, device=d) < 0.05).to(torch.uint8)
    L = torch.randn(N, 1, generator=gen, device=d)
    ret, adv = torch.empty_like(R), torch.empty_like(R)
    stats = torch.zeros(3, dtype=torch.float64, device=d)
    p = lambda t: C.c_void_p(t.data_ptr())
    _abi.check(lib.lt_gae(p(R), p(V), p(D), p(L), p(ret), p(adv), p(stats), T, N, C.c_float(0.99), C.c_float(0.95), None))
    _abi.check(lib.lt_normalize_advantages(p(adv), p(stats), T * N, None))
    torch.cuda.synchronize()
    assert abs(float(adv.double().mean())) < 1e-6
    assert abs(float(adv.double().std()) - 1.0) < 1e-5
    done = D.bool()
    # terminal step: return == (reward - value) + value (no bootstrap, no carry from later steps)
    assert torch.equal(ret[done], (R[done] - V[done]) + V[done])


@pytest.mark.parametrize("N,no,npv,H,na", [(64, 42, 6, 5, 12), (33, 7, 3, 3, 5), (1, 1, 1, 1, 1)])
def test_rollout_store_ragged_sizes(N, no, npv, H, na):
    """Segments whose byte size is not a multiple of 16 take the scalar path; rows before/after stay untouched."""
    from legged_tracking_b200 import ppo
    d = dev()
    st = ppo.RolloutStorage(N, 3, [no], [npv], [no * H], [na], device=d)
    gen = torch.Generator().manual_seed(N)
    tr = ppo.RolloutStorage.Transition()
    tr.observations = torch.randn(N, no, generator=gen).to(d)
    tr.privileged_observations = torch.randn(N, npv, generator=gen).to(d)
    tr.observation_histories = torch.randn(N, no * H, generator=gen).to(d)
    tr.actions = torch.randn(N, na, generator=gen).to(d)
    tr.rewards = torch.randn(N, generator=gen).to(d)
    tr.values = torch.randn(N, 1, generator=gen).to(d)
    tr.actions_log_prob = torch.randn(N, generator=gen).to(d)
    tr.action_mean = torch.randn(N, na, generator=gen).to(d)
    tr.action_sigma = torch.rand(N, na, generator=gen).to(d)
    tr.dones = (torch.rand(N, generator=gen) < 0.3).to(d)
    tr.time_outs = tr.dones & (torch.rand(N, generator=gen) < 0.5).to(d)
    st.step = 1
    st.add_transitions(tr)
    torch.cuda.synchronize()
    assert st.step == 2
    for name, src in (("observations", tr.observations), ("privileged_observations", tr.privileged_observations),
                      ("observation_histories", tr.observation_histories), ("actions", tr.actions), ("values", tr.values),
                      ("mu", tr.action_mean), ("sigma", tr.action_sigma)):
        buf = getattr(st, name)
        assert torch.equal(buf[1], src.view_as(buf[1])), name
        assert float(buf[0].abs().sum()) == 0.0 and float(buf[2].abs().sum()) == 0.0, name
    want = tr.rewards + ppo.PPO_Args.gamma * torch.squeeze(tr.values * tr.time_outs.unsqueeze(1), 1)
    assert torch.equal(st.rewards[1, :, 0], want)
    assert torch.equal(st.dones[1, :, 0], tr.dones.to(torch.uint8))
    assert torch.equal(st.actions_log_prob[1, :, 0], tr.actions_log_prob)
    with pytest.raises(AssertionError):
        st.step = 3
        st.add_transitions(tr)
