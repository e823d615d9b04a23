"""-m gpu parity of rows a14 / a15 through the C ABI of include/lt_policy.h: the tcgen05 split-fp16
contraction kernel against a float64 product, the actor-critic forward against the oracle port on the reference-recorded
weights / inputs (tests/golden/ppo_*.npz), and the minibatch gradient against torch autograd of the oracle's loss.

Tolerances: north_star's 1e-5 relative for fp32 outputs, measured against the magnitude of the contraction (sum |a||b|), which
is how fp32 GEMM error scales; gradients 1e-4 relative to the largest entry of each tensor."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


@pytest.mark.parametrize("M,N,K,splits", [(128, 128, 64, 1), (128, 16, 64, 1), (300, 200, 263, 1), (1000, 512, 263, 1), (257, 16, 128, 1),
                                          (64, 12, 512, 1), (4096, 256, 512, 1), (512, 144, 4096, 3), (12, 129, 10000, 7), (1, 1, 1, 1),
                                          (65536, 512, 265, 1)])
def test_split_fp16_gemm_matches_float64(M, N, K, splits):
    """lt_linear_f32: out = a @ b.T + bias with every product as three tcgen05 fp16 MMAs (hi*hi + hi*lo + lo*hi), fp32 accumulate.
    Ragged M / N / K (TMA zero fill + epilogue masks), bn = 16 .. 256, split-K."""
    from legged_tracking_b200 import _abi
    lib = _abi.load()
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda", generator=g)
    b = torch.randn(N, K, device="cuda", generator=g) * 0.1
    bias = torch.randn(N, device="cuda", generator=g)
    out = torch.full((M, N), float("nan"), device="cuda")
    _abi.check_policy(lib.lt_linear_f32(_p(a), _p(b), _p(bias), _p(out), M, N, K, splits, None), None, "lt_linear_f32")
    torch.cuda.synchronize()
    want = a.double() @ b.double().T + bias.double()
    mag = a.double().abs() @ b.double().abs().T + bias.double().abs()
    rel = ((out.double() - want).abs() / mag).max().item()
    assert rel < 2e-6, f"max error / sum|a||b| = {rel:.3e}"      # measured 2e-7 .. 1e-6 (torch fp32: 2e-7 .. 4e-7)


@pytest.mark.parametrize("M,N,K,splits", [(128, 64, 64, 1), (128, 256, 128, 1), (512, 266, 1000, 1), (12, 129, 777, 3), (256, 513, 4096, 5),
                                          (1, 1, 1, 1), (130, 17, 65, 1), (512, 266, 393216, 64)])
def test_split_fp16_gemm_mn_major_matches_float64(M, N, K, splits):
    """lt_linear_tn_f32: out = a.T @ b with both operands read MN-major (contraction over rows) -- the form of every weight /
    bias gradient product.  Ragged M / N / K, several 64-wide operand chunks, split-K."""
    from legged_tracking_b200 import _abi
    lib = _abi.load()
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(K, M, device="cuda", generator=g)
    b = torch.randn(K, N, device="cuda", generator=g) * 0.1
    out = torch.full((M, N), float("nan"), device="cuda")
    _abi.check_policy(lib.lt_linear_tn_f32(_p(a), _p(b), _p(out), M, N, K, splits, None), None, "lt_linear_tn_f32")
    torch.cuda.synchronize()
    want = a.double().T @ b.double()
    mag = a.double().abs().T @ b.double().abs()
    rel = ((out.double() - want).abs() / mag).max().item()
    assert rel < 2e-6, f"max error / sum|a||b| = {rel:.3e}"


# ------------------------------------------------------------------------------------------------------------------------
# a14 / a15 through legged_tracking_b200.ppo (which only marshals pointers into include/lt_policy.h) against the oracle port
# evaluated in float64 on the same weights and inputs.
from tests.helpers_ppo import PPO_CASES, actor_critic_from_golden, golden_args, load_case      # noqa: E402


def _dev():
    return torch.device("cuda", 0)


def _params64(ac):
    return {k: v.detach().cpu().double().clone().requires_grad_(True) for k, v in ac.state_dict().items()}


def _close(got, want, rtol, atol, what):
    got, want = got.detach().cpu().double().numpy(), want.detach().cpu().double().numpy()
    err = np.abs(got - want)
    lim = atol + rtol * np.abs(want)
    assert np.all(err <= lim), f"{what}: max err {err.max():.3e} (|want| up to {np.abs(want).max():.3e}), worst excess {(err - lim).max():.3e}"


@pytest.mark.parametrize("name", PPO_CASES)          # ELU and tanh bodies
def test_policy_forward_matches_oracle(name):
    """latent / action mean / value / teacher action of lt_policy_forward: 1e-5 relative (+1e-5 of the largest output, the
    scale fp32 GEMM error lives on) against the float64 oracle."""
    from oracle import ppo_port
    g = load_case(name)
    with golden_args(g):
        ac = actor_critic_from_golden(g, _dev())
        p64 = _params64(ac)
        act = ppo_port.ACTIVATIONS[ac.activation]
        T = int(g["meta/sizes"][1])
        for t in (0, T - 1):
            hist, priv = torch.from_numpy(g["in/hist"][t]), torch.from_numpy(g["in/priv"][t])
            with torch.no_grad():
                mean, _, value = ppo_port.policy_forward(p64, act, hist.double(), priv.double(), ac.num_obs)
                pin = ppo_port.policy_input(p64, act, hist.double(), ac.num_obs)       # the history itself unless ppo_cse_cnn
                latent = ppo_port.mlp(p64, "adaptation_module", pin, act)
                teacher = ppo_port.mlp(p64, "actor_body", torch.cat((pin, priv.double()), dim=-1), act)
            info = {}
            got_mean = ac.act_student(hist.to(_dev()), policy_info=info)
            for got, want, what in ((got_mean, mean, "mean"), (torch.from_numpy(info["latents"]), latent, "latent"),
                                    (ac.get_student_latent(hist.to(_dev())), latent, "latent2"),
                                    (ac.evaluate(hist.to(_dev()), priv.to(_dev())), value, "value"),
                                    (ac.act_teacher(hist.to(_dev()), priv.to(_dev())), teacher, "teacher")):
                _close(got, want, 1e-5, 1e-5 * float(want.abs().max()), f"{name} t={t} {what}")
            # and the fp32 reference recording itself
            np.testing.assert_allclose(got_mean.cpu().numpy(), g["storage/mu"][t], rtol=1e-5, atol=2e-6)


def test_act_sampling_log_prob_and_injected_noise():
    """lt_policy_act: actions = mean + sigma * noise exactly for injected noise; log-prob is Normal.log_prob of the returned
    actions; the in-kernel Philox + Box-Muller draws are standard normal and differ per (step, row, action)."""
    from legged_tracking_b200 import ppo
    torch.manual_seed(3)
    ac = ppo.ActorCritic(20, 3, 40, 12).to(_dev())
    ac.load_state_dict({k: (v * 0 + 0.7 if k == "std" else v) for k, v in ac.state_dict().items()})
    rows = 4096
    gen = torch.Generator(device="cuda").manual_seed(1)
    hist = torch.randn(rows, 40, device="cuda", generator=gen)
    priv = torch.randn(rows, 3, device="cuda", generator=gen)
    noise = torch.randn(rows, 12, device="cuda", generator=gen)
    new = lambda *s: torch.full(s, float("nan"), device="cuda")
    out = dict(actions=new(rows, 12), actions_log_prob=new(rows), action_mean=new(rows, 12), action_sigma=new(rows, 12), value=new(rows, 1))
    ac.act_into(hist, priv, out, normal_noise=noise)
    torch.cuda.synchronize()
    assert torch.equal(out["action_sigma"], ac.std.expand(rows, 12))
    assert torch.equal(out["actions"], out["action_mean"] + out["action_sigma"] * noise) or \
        torch.allclose(out["actions"], out["action_mean"] + out["action_sigma"] * noise, rtol=0, atol=1e-6)      # fma contraction
    want_lp = torch.distributions.Normal(out["action_mean"].double(), out["action_sigma"].double()).log_prob(out["actions"].double()).sum(-1)
    _close(out["actions_log_prob"], want_lp, 1e-5, 1e-5, "log_prob")
    _close(out["value"], ac.evaluate(hist, priv), 0, 0, "value")
    _close(out["action_mean"], ac.act_student(hist), 0, 0, "mean")
    # Philox path
    a1, a2 = dict(actions=new(rows, 12), action_mean=new(rows, 12)), dict(actions=new(rows, 12))
    ac.act_into(hist, priv, a1, seed=5, step=1)
    ac.act_into(hist, priv, a2, seed=5, step=2)
    z1 = (a1["actions"] - a1["action_mean"]) / 0.7
    z2 = (a2["actions"] - a1["action_mean"]) / 0.7
    assert abs(float(z1.mean())) < 0.02 and abs(float(z1.std()) - 1.0) < 0.02
    assert abs(float((z1 * z2).mean())) < 0.02                      # steps are independent
    assert abs(float((z1[:, 0] * z1[:, 1]).mean())) < 0.05         # action columns are independent
    a3 = dict(actions=new(rows, 12))
    ac.act_into(hist, priv, a3, seed=5, step=1)
    assert torch.equal(a3["actions"], a1["actions"])              # keyed, not stateful
    # ActorCritic.act / get_actions_log_prob / entropy (the reference's distribution API)
    a = ac.act(hist)
    lp = ac.get_actions_log_prob(a)
    want = torch.distributions.Normal(ac.action_mean, ac.action_std)
    assert torch.allclose(lp, want.log_prob(a).sum(-1), rtol=1e-5, atol=1e-5)
    assert torch.allclose(ac.entropy, want.entropy().sum(-1), rtol=1e-6)


def _random_batch(ac, B, seed, device):
    gen = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=gen)
    A, P, D = ac.num_actions, ac.num_privileged_obs, ac.num_obs_history
    b = dict(obs_history=r(B, D), privileged_obs=r(B, P), actions=r(B, A), values=r(B, 1), returns=r(B, 1), advantages=r(B, 1),
             old_actions_log_prob=r(B, 1) * 0.3 - 1.1 * A, old_mu=r(B, A) * 0.5, old_sigma=torch.rand(B, A, generator=gen) + 0.5)
    return {k: v.to(device) for k, v in b.items()}


def _batch_ptrs(b):
    from legged_tracking_b200 import _abi
    bp = _abi.lt_ppo_batch_ptrs()
    for k, v in b.items():
        setattr(bp, k, v.data_ptr())
    return bp


def _oracle_grads(ac, b, idx, args):
    from oracle import ppo_port
    p64 = _params64(ac)
    act = ppo_port.ACTIVATIONS[ac.activation]
    s = {k: v.cpu().double()[idx.cpu()] for k, v in b.items()}
    loss, surr, vl, kl = ppo_port.minibatch_loss(p64, act, s["obs_history"], s["privileged_obs"], s["actions"], s["values"], s["advantages"],
                                                 s["returns"], s["old_actions_log_prob"], s["old_mu"], s["old_sigma"], args, ac.num_obs)
    loss.backward()
    flat = torch.cat([p64[k].grad.reshape(-1) for k in ac.state_dict()])
    return flat, float(surr), float(vl), float(kl)


def _grad_close(ac, got, want, tol, what):
    """per tensor: error relative to the largest entry of that tensor's gradient"""
    got, want = got.detach().cpu().double(), want.detach().cpu().double()
    for key, _, lo, hi in ac.layout:
        scale = float(want[lo:hi].abs().max())
        err = float((got[lo:hi] - want[lo:hi]).abs().max())
        assert err <= tol * scale + 1e-12, f"{what} {key}: max err {err:.3e} vs largest entry {scale:.3e}"


@pytest.mark.parametrize("activation,B,count", [("elu", 700, 517), ("tanh", 300, 300), ("elu", 130, 1), ("relu", 400, 256)])
def test_minibatch_grad_matches_autograd(activation, B, count):
    """lt_ppo_minibatch_grad: gather by index, forward, surrogate / clipped value / entropy loss, KL, full backward (tcgen05
    GEMMs for dX and dW, bias gradient through the constant-one row) against float64 autograd of the oracle's loss."""
    from legged_tracking_b200 import _abi, ppo
    from oracle import ppo_port
    lib = _abi.load()
    saved = ppo.AC_Args.activation
    ppo.AC_Args.activation = activation
    try:
        torch.manual_seed(11)
        ac = ppo.ActorCritic(30, 5, 90, 12).to(_dev())
    finally:
        ppo.AC_Args.activation = saved
    b = _random_batch(ac, B, 5, _dev())
    idx = torch.randperm(B, generator=torch.Generator().manual_seed(2))[:count].to(_dev())
    ac.grads = torch.full((ac.num_params,), float("nan"), device=_dev())
    ac._ready(count, training=True)
    stats = torch.zeros(16, device=_dev())
    hp = ppo.PPO.hparams()
    bp = _batch_ptrs(b)
    _abi.check_policy(lib.lt_ppo_minibatch_grad(ac._h, C.byref(bp), _p(idx), count, C.byref(hp), 1, _p(stats), None), ac._h)
    torch.cuda.synchronize()
    args = dict(ppo_port.PPO_DEFAULTS)
    want, surr, vl, kl = _oracle_grads(ac, b, idx, args)
    assert torch.isfinite(ac.grads).all()
    _grad_close(ac, ac.grads, want, 1e-4, f"{activation} B={count}")
    st = stats.cpu().double()
    assert st[3] == count
    np.testing.assert_allclose([st[0] / count, st[1] / count, st[2] / count], [surr, vl, kl], rtol=2e-5, atol=1e-6)
    # adaptation regression step on the same rows (PPO:150-176)
    num_train = count // 5 * 4
    ac.grads.zero_()
    _abi.check_policy(lib.lt_adaptation_grad(ac._h, C.byref(bp), _p(idx), count, num_train, 1, _p(stats), None), ac._h)
    torch.cuda.synchronize()
    p64 = _params64(ac)
    hist, priv = b["obs_history"].cpu().double()[idx.cpu()], b["privileged_obs"].cpu().double()[idx.cpu()]
    pred = ppo_port.mlp(p64, "adaptation_module", hist, ppo_port.ACTIVATIONS[activation])
    lo, hi = ac.body_slices["adaptation_module"]
    if num_train > 0:
        loss = torch.nn.functional.mse_loss(pred[:num_train], priv[:num_train])
        loss.backward()
        want_ad = torch.cat([p64[k].grad.reshape(-1) for k in ac.state_dict() if k.startswith("adaptation_module.")])
        got = ac.grads[lo:hi].cpu().double()
        for key, _, a, e in ac.layout:
            if key.startswith("adaptation_module."):
                scale = float(want_ad[a - lo:e - lo].abs().max())
                assert float((got[a - lo:e - lo] - want_ad[a - lo:e - lo]).abs().max()) <= 1e-4 * scale + 1e-12, key
        st = stats.cpu().double()
        np.testing.assert_allclose(st[4] / st[5], float(loss), rtol=2e-5)
        if count > num_train:
            np.testing.assert_allclose(st[6] / st[7], float(torch.nn.functional.mse_loss(pred[num_train:], priv[num_train:])), rtol=2e-5)
    assert float(ac.grads[hi:].abs().sum()) == 0.0 and float(ac.grads[:lo].abs().sum()) == 0.0      # rest of the bucket untouched


def test_policy_with_the_config_4_history_matches_the_oracle():
    """BASELINE config 4 feeds the policy a 30-row history of 70-wide observations (scripts/train_velocity_tracking.py:80): 2,100 + 4
    input columns (33 k-blocks of the first-layer contraction, a ragged last one), 4,096 rows forward and one minibatch gradient."""
    from legged_tracking_b200 import _abi, ppo
    from oracle import ppo_port
    lib = _abi.load()
    torch.manual_seed(31)
    ac = ppo.ActorCritic(70, 2, 2100, 12).to(_dev())
    rows = 4096
    b = _random_batch(ac, rows, 8, _dev())
    b["obs_history"].mul_(0.5)
    p64 = _params64(ac)
    act = ppo_port.ACTIVATIONS["elu"]
    with torch.no_grad():
        mean, _, value = ppo_port.policy_forward(p64, act, b["obs_history"].cpu().double(), b["privileged_obs"].cpu().double())
    _close(ac.act_student(b["obs_history"]), mean, 1e-5, 1e-5 * float(mean.abs().max()), "cfg4 mean")
    _close(ac.evaluate(b["obs_history"], b["privileged_obs"]), value, 1e-5, 1e-5 * float(value.abs().max()), "cfg4 value")
    idx = torch.randperm(rows, generator=torch.Generator().manual_seed(1)).to(_dev())
    ac.grads = torch.zeros(ac.num_params, device=_dev())
    ac._ready(rows, training=True)
    stats = torch.zeros(16, device=_dev())
    hp = ppo.PPO.hparams()
    bp = _batch_ptrs(b)
    _abi.check_policy(lib.lt_ppo_minibatch_grad(ac._h, C.byref(bp), _p(idx), rows, C.byref(hp), 1, _p(stats), None), ac._h)
    torch.cuda.synchronize()
    want, surr, vl, kl = _oracle_grads(ac, b, idx, dict(ppo_port.PPO_DEFAULTS))
    _grad_close(ac, ac.grads, want, 1e-4, "cfg4 history")


def _cnn_actor_critic(num_obs, P, H, A, shape, emb, seed, **ac_over):
    from legged_tracking_b200 import ppo_cnn
    args = ppo_cnn.AC_Args
    saved = {k: getattr(args, k) for k in vars(args)}
    try:
        args.height_map_shape, args.cnn_num_embedding = shape, emb
        for k, v in ac_over.items():
            setattr(args, k, v)
        torch.manual_seed(seed)
        return ppo_cnn.ActorCritic(num_obs, P, num_obs * H, A).to(_dev())
    finally:
        for k, v in saved.items():
            setattr(args, k, v)


@pytest.mark.parametrize("activation,H,count", [("elu", 2, 517), ("tanh", 1, 300), ("elu", 3, 1)])
def test_cnn_variant_minibatch_grad_matches_autograd(activation, H, count):
    """Row f4 (go1_gym_learn/ppo_cse_cnn, MLP height-map encoder): the PPO minibatch gradient and the adaptation regression
    gradient -- both reach the encoder through the three bodies' first layers -- against float64 autograd of the oracle."""
    from legged_tracking_b200 import _abi, ppo
    from oracle import ppo_port
    lib = _abi.load()
    ac = _cnn_actor_critic(44, 3, H, 12, (2, 4, 3), 40, 13, activation=activation, actor_hidden_dims=[96, 64], critic_hidden_dims=[80, 48],
                           adaptation_module_branch_hidden_dims=[72, 24])
    B = max(count, 64) + 100
    b = _random_batch(ac, B, 6, _dev())
    idx = torch.randperm(B, generator=torch.Generator().manual_seed(4))[:count].to(_dev())
    ac.grads = torch.full((ac.num_params,), float("nan"), device=_dev())
    ac._ready(count, training=True)
    stats = torch.zeros(16, device=_dev())
    hp = ppo.PPO.hparams()
    bp = _batch_ptrs(b)
    _abi.check_policy(lib.lt_ppo_minibatch_grad(ac._h, C.byref(bp), _p(idx), count, C.byref(hp), 1, _p(stats), None), ac._h)
    torch.cuda.synchronize()
    want, surr, vl, kl = _oracle_grads(ac, b, idx, dict(ppo_port.PPO_DEFAULTS))
    assert torch.isfinite(ac.grads).all()
    _grad_close(ac, ac.grads, want, 1e-4, f"cnn {activation} H={H} B={count}")
    st = stats.cpu().double()
    np.testing.assert_allclose([st[0] / count, st[1] / count, st[2] / count], [surr, vl, kl], rtol=2e-5, atol=1e-6)
    # adaptation step: encoder + adaptation module hold the gradient, nothing else is touched (ppo_cse_cnn/ppo.py:170-186)
    num_train = count // 5 * 4
    ac.grads.zero_()
    _abi.check_policy(lib.lt_adaptation_grad(ac._h, C.byref(bp), _p(idx), count, num_train, 1, _p(stats), None), ac._h)
    torch.cuda.synchronize()
    lo, hi = ac.body_slices["height_map_encoder"][0], ac.body_slices["adaptation_module"][1]
    assert float(ac.grads[hi:].abs().sum()) == 0.0 and float(ac.grads[:lo].abs().sum()) == 0.0
    if num_train > 0:
        p64 = _params64(ac)
        act = ppo_port.ACTIVATIONS[activation]
        hist, priv = b["obs_history"].cpu().double()[idx.cpu()], b["privileged_obs"].cpu().double()[idx.cpu()]
        pred = ppo_port.mlp(p64, "adaptation_module", ppo_port.policy_input(p64, act, hist, ac.num_obs), act)
        loss = torch.nn.functional.mse_loss(pred[:num_train], priv[:num_train])
        loss.backward()
        got = ac.grads.cpu().double()
        for key, _, a, e in ac.layout:
            if lo <= a < hi:
                w = p64[key].grad.reshape(-1)
                scale = float(w.abs().max())
                assert float((got[a:e] - w).abs().max()) <= 1e-4 * scale + 1e-12, key
        st = stats.cpu().double()
        np.testing.assert_allclose(st[4] / st[5], float(loss), rtol=2e-5)
    s_lo, s_hi = C.c_int64(), C.c_int64()
    _abi.check_policy(lib.lt_policy_adaptation_slice(ac._h, C.byref(s_lo), C.byref(s_hi)), ac._h)
    assert (s_lo.value, s_hi.value) == (lo, hi)


def test_cnn_variant_at_train_py_size_matches_fp32_torch():
    """scripts/train.py without --old_ppo and without --measure_front_half: 503-wide observations (41 scalars + 2x21x11 heights),
    encoder 462 -> 256 -> 256, default bodies on the 338-wide policy input.  16,384 rows forward and one gradient, against torch
    fp32 autograd (TF32 off) of the oracle's loss on the same GPU."""
    from legged_tracking_b200 import _abi, ppo
    from oracle import ppo_port
    lib = _abi.load()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    ac = _cnn_actor_critic(503, 2, 1, 12, (2, 21, 11), 256, 5)
    assert ac.policy_input_dim == 338
    rows = 16384
    b = _random_batch(ac, rows, 3, _dev())
    b["obs_history"].clamp_(-3, 3)
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in ac.state_dict().items()}
    act = ppo_port.ACTIVATIONS["elu"]
    with torch.no_grad():
        mean, _, value = ppo_port.policy_forward(sd, act, b["obs_history"], b["privileged_obs"], 503)
    _close(ac.act_student(b["obs_history"]), mean, 1e-5, 1e-5 * float(mean.abs().max()), "cnn mean")
    _close(ac.evaluate(b["obs_history"], b["privileged_obs"]), value, 1e-5, 1e-5 * float(value.abs().max()), "cnn value")
    idx = torch.randperm(rows, device=_dev())
    ac.grads = torch.zeros(ac.num_params, device=_dev())
    ac._ready(rows, training=True)
    stats = torch.zeros(16, device=_dev())
    hp = ppo.PPO.hparams()
    bp = _batch_ptrs(b)
    _abi.check_policy(lib.lt_ppo_minibatch_grad(ac._h, C.byref(bp), _p(idx), rows, C.byref(hp), 1, _p(stats), None), ac._h)
    torch.cuda.synchronize()
    s = {k: v[idx] for k, v in b.items()}
    loss, *_ = ppo_port.minibatch_loss(sd, act, s["obs_history"], s["privileged_obs"], s["actions"], s["values"], s["advantages"], s["returns"],
                                       s["old_actions_log_prob"], s["old_mu"], s["old_sigma"], dict(ppo_port.PPO_DEFAULTS), 503)
    loss.backward()
    want = torch.cat([sd[k].grad.reshape(-1) for k in ac.state_dict()])
    _grad_close(ac, ac.grads, want, 2e-4, "cnn @16384")          # fp32 torch on the other side: its own rounding is in the budget


def test_sharded_gradients_sum_to_the_full_minibatch_gradient():
    """Env sharding: two ranks each run half of a minibatch with grad_world = 2; the SUM of their [stats | grads]
    buffers -- what the all-reduce of dist.GradExchange produces -- equals the single-process result."""
    from legged_tracking_b200 import _abi, ppo
    lib = _abi.load()
    torch.manual_seed(4)
    ac = ppo.ActorCritic(30, 5, 90, 12).to(_dev())
    B = 512
    b = _random_batch(ac, B, 9, _dev())
    idx = torch.randperm(B, generator=torch.Generator().manual_seed(3)).to(_dev())
    hp = ppo.PPO.hparams()
    bp = _batch_ptrs(b)

    def run(ix, world):
        ac.grads = torch.zeros(ac.num_params, device=_dev())
        ac._bind()
        ac._ready(B, training=True)
        stats = torch.zeros(16, device=_dev())
        _abi.check_policy(lib.lt_ppo_minibatch_grad(ac._h, C.byref(bp), _p(ix), ix.numel(), C.byref(hp), world, _p(stats), None), ac._h)
        torch.cuda.synchronize()
        return ac.grads.clone().double(), stats.clone().double()
    full_g, full_s = run(idx, 1)
    g0, s0 = run(idx[0::2].contiguous(), 2)
    g1, s1 = run(idx[1::2].contiguous(), 2)
    _grad_close(ac, g0 + g1, full_g, 2e-5, "sharded")
    np.testing.assert_allclose((s0 + s1)[:4].cpu().numpy(), full_s[:4].cpu().numpy(), rtol=1e-5)


def test_minibatch_apply_is_clip_plus_adam_with_the_kl_schedule():
    """lt_ppo_minibatch_apply against nn.utils.clip_grad_norm_ + torch.optim.Adam on the CPU, three steps, and the three
    branches of the adaptive-KL learning-rate rule (PPO:111-125)."""
    from legged_tracking_b200 import _abi, ppo
    lib = _abi.load()
    torch.manual_seed(8)
    ac = ppo.ActorCritic(12, 2, 24, 3).to(_dev())
    n = ac.num_params
    ac.grads = torch.zeros(n, device=_dev())
    ac._ready(8, training=True)
    ref = torch.nn.Parameter(ac.flat.detach().cpu().clone())
    opt = torch.optim.Adam([ref], lr=1e-3)
    m, v = torch.zeros(n, device=_dev()), torch.zeros(n, device=_dev())
    lr = torch.tensor([1e-3], device=_dev())
    norm = torch.zeros(1, device=_dev())
    hp = ppo.PPO.hparams()
    gen = torch.Generator().manual_seed(0)
    want_lr = 1e-3
    for step, (scale, kl) in enumerate([(1e-3, 0.001), (5.0, 0.05), (0.2, 0.01)], start=1):
        g = torch.randn(n, generator=gen) * scale
        ac.grads.copy_(g.to(_dev()))
        stats = torch.tensor([0, 0, kl * 10, 10] + [0] * 12, dtype=torch.float32, device=_dev())
        _abi.check_policy(lib.lt_ppo_minibatch_apply(ac._h, C.byref(hp), _p(stats), _p(lr), _p(m), _p(v), step, _p(norm), None), ac._h)
        torch.cuda.synchronize()
        want_lr = max(1e-5, want_lr / 1.5) if kl > 0.02 else (min(1e-2, want_lr * 1.5) if kl < 0.005 else want_lr)
        for grp in opt.param_groups:
            grp["lr"] = want_lr
        ref.grad = g.clone()
        tn = torch.nn.utils.clip_grad_norm_([ref], 1.0)
        opt.step()
        np.testing.assert_allclose(float(lr), want_lr, rtol=1e-6)
        np.testing.assert_allclose(float(norm), float(tn), rtol=1e-5)
        np.testing.assert_allclose(ac.flat.cpu().numpy(), ref.detach().numpy(), rtol=1e-5, atol=1e-7)


def test_policy_at_baseline_size_matches_fp32_torch():
    """BASELINE.json cfg 5 shapes: 65,536 rows through act (obs_history 261, priv 2, the default 512-256-128 / 256-128 bodies)
    and one 393,216-row minibatch gradient, against torch fp32 (TF32 off) on the same GPU."""
    from legged_tracking_b200 import _abi, ppo
    from oracle import ppo_port
    lib = _abi.load()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(21)
    ac = ppo.ActorCritic(261, 2, 261, 12).to(_dev())
    rows, B = 65536, 65536 * 24 // 4
    gen = torch.Generator(device="cuda").manual_seed(2)
    hist = torch.randn(rows, 261, device="cuda", generator=gen).clamp_(-3, 3)
    priv = torch.rand(rows, 2, device="cuda", generator=gen) * 2 - 1
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in ac.state_dict().items()}
    act = ppo_port.ACTIVATIONS["elu"]
    with torch.no_grad():
        mean, _, value = ppo_port.policy_forward(sd, act, hist, priv)
    got_mean, got_value = ac.act_student(hist), ac.evaluate(hist, priv)
    _close(got_mean, mean, 1e-5, 1e-5 * float(mean.abs().max()), "mean @65536")
    _close(got_value, value, 1e-5, 1e-5 * float(value.abs().max()), "value @65536")
    # one full-size minibatch: rows drawn (with the permutation) out of a 6 x 65,536 slice of a rollout
    b = _random_batch(ac, B, 1, "cpu")
    b = {k: v.to(_dev()) for k, v in b.items()}
    idx = torch.randperm(B, device="cuda", generator=gen)
    ac.grads = torch.zeros(ac.num_params, device=_dev())
    ac._ready(B, training=True)
    stats = torch.zeros(16, device=_dev())
    hp = ppo.PPO.hparams()
    bp = _batch_ptrs(b)
    _abi.check_policy(lib.lt_ppo_minibatch_grad(ac._h, C.byref(bp), _p(idx), B, C.byref(hp), 1, _p(stats), None), ac._h)
    torch.cuda.synchronize()
    s = {k: v[idx] for k, v in b.items()}
    loss, surr, vl, kl = ppo_port.minibatch_loss(sd, act, s["obs_history"], s["privileged_obs"], s["actions"], s["values"], s["advantages"],
                                                 s["returns"], s["old_actions_log_prob"], s["old_mu"], s["old_sigma"], dict(ppo_port.PPO_DEFAULTS))
    loss.backward()
    want = torch.cat([sd[k].grad.reshape(-1) for k in ac.state_dict()])
    _grad_close(ac, ac.grads, want, 2e-4, "393,216-row minibatch")
    st = stats.cpu().double()
    np.testing.assert_allclose([st[0] / B, st[1] / B, st[2] / B], [float(surr), float(vl), float(kl)], rtol=1e-4)
