"""The reference's env API (TrajectoryTrackingEnv + HistoryWrapper, Cfg classes, `_reward_*` plugins) over the CUDA path:
API-level parity against the golden fixtures recorded through the reference's own wrapped env, the plugin protocol,
reset(), and a short PPO Runner loop."""
import numpy as np
import pytest
import torch

from tests.helpers import ATOL, RTOL, fixture_names, load_fixture, sub


def terrain_of(fx):
    return sub(fx, "terrain/")

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def frames_of(fx):
    out, k = [], 0
    while f"step{k}/in/actions" in fx:
        d = sub(fx, f"step{k}/in/")
        out.append(dict(dof_state=d["dof_state"], root=d["root"], contact=d["contact"], rigid=d["rigid"]))
        k += 1
    return out


def make_env(name, fx, **kw):
    from legged_tracking_b200 import cfg as ltcfg
    from legged_tracking_b200 import env as ltenv
    flat = fx["flat"]
    c = ltcfg.preset_cfg(name, flat["num_envs"])
    e = ltenv.TrajectoryTrackingEnv(sim_device=DEV, headless=True, cfg=c, terrain=terrain_of(fx), **kw)
    return e, ltenv.HistoryWrapper(e)


@pytest.mark.parametrize("name", fixture_names())
def test_wrapped_env_step_matches_reference(name):
    """Same call sequence as the reference's Runner: wrapped.step(actions) -> (obs_dict, rew, done, extras)."""
    from legged_tracking_b200 import env as ltenv
    fx = load_fixture(name)
    frames = frames_of(fx)
    e, w = make_env(name, fx, sim_backend=ltenv.ReplaySim(frames=frames))
    e.engine.load_state(sub(fx, "init/"))
    N = fx["flat"]["num_envs"]
    for k in range(len(frames)):
        d = sub(fx, f"step{k}/in/")
        U = np.full((N, e.engine.n_slots), 0.5, np.float32)
        U[:, :d["uniforms"].shape[1]] = d["uniforms"]
        inject = dict(uniforms=torch.from_numpy(U).to(DEV))
        if "noise_uniforms" in d:
            inject["noise_uniforms"] = torch.from_numpy(np.ascontiguousarray(d["noise_uniforms"], np.float32)).to(DEV)
        if "gravity_uniform" in d:
            inject["gravity_uniform"] = d["gravity_uniform"]
        obs, rew, done, extras = w.step(torch.from_numpy(d["actions"]).to(DEV), inject)
        out = sub(fx, f"step{k}/out/")
        assert np.array_equal(done.cpu().numpy().astype(bool), out["done"])
        np.testing.assert_allclose(rew.cpu().numpy(), out["rew"], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(obs["obs"].cpu().numpy(), out["obs"], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(obs["privileged_obs"].cpu().numpy(), out["privileged_obs"], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(obs["obs_history"].cpu().numpy(), out["obs_history"], rtol=RTOL, atol=ATOL)
        assert "time_outs" in extras and extras["time_outs"].shape == (N,)
        n_reset = int(out["done"].sum())
        if n_reset:
            assert len(extras["train/episode"]["episode_length"]) >= n_reset
        assert extras["joint_pos"].shape == (N, 12) and extras["torques"].shape == (N, 12)     # lazy host copies


def test_outputs_are_fresh_tensors_unless_aliased():
    name = "cfg2_nonoise_64"
    fx = load_fixture(name)
    e, w = make_env(name, fx)
    a = torch.zeros(e.num_envs, 12, device=DEV)
    o1, _, _, _ = w.step(a)
    keep = o1["obs"].clone()
    w.step(a)
    assert torch.equal(o1["obs"], keep)             # a reference held across step() keeps its values (reference semantics)
    e2, w2 = make_env(name, fx, alias_outputs=True)
    o, _, _, _ = w2.step(a)
    assert o["obs"].data_ptr() == e2.obs_buf.data_ptr()


def test_reward_plugin_term_is_added_like_a_fused_term():
    from legged_tracking_b200 import cfg as ltcfg
    from legged_tracking_b200 import env as ltenv
    name = "cfg2_nonoise_64"
    fx = load_fixture(name)
    N = fx["flat"]["num_envs"]

    class MyRewards(ltenv.RewardsCrawling):
        def _reward_hip_motion(self):                      # reads env attributes exactly like the shipped containers do
            return torch.sum(torch.square(self.env.dof_vel[:, [0, 3, 6, 9]]), dim=1)

        def _reward_alive(self):
            return torch.ones(self.env.num_envs, device=self.env.device)

    def run(plugin):
        c = ltcfg.preset_cfg(name, N)
        if plugin:
            c.reward_scales.hip_motion = -0.002
            c.reward_scales.alive = 0.5
        e = ltenv.TrajectoryTrackingEnv(sim_device=DEV, cfg=c, terrain=terrain_of(fx), reward_container=MyRewards if plugin else None,
                                        sim_backend=ltenv.ReplaySim(frames=frames_of(fx)))
        e.engine.load_state(sub(fx, "init/"))
        _, rew, _, _ = e.step(torch.from_numpy(fx["step0/in/actions"]).to(DEV))
        return e, rew.clone()

    e0, r0 = run(False)
    e1, r1 = run(True)
    assert e1.reward_names[-2:] == ["hip_motion", "alive"] and e1.engine.n_external == 2
    hip = torch.sum(torch.square(torch.from_numpy(fx["step0/in/dof_state"][-1]).view(N, 12, 2)[:, [0, 3, 6, 9], 1]), dim=1).to(DEV)
    dt = e1.dt
    want = r0 + hip * (-0.002 * dt) + 0.5 * dt
    torch.testing.assert_close(r1, want, rtol=1e-5, atol=2e-6)
    sums = e1.episode_sums
    keep = ~e1.reset_buf
    torch.testing.assert_close(sums["alive"][keep], torch.full((int(keep.sum()),), 0.5 * dt, device=DEV), rtol=1e-6, atol=0)


class NullSim:
    """Leaves the simulator tensors as the kernels wrote them."""
    def attach(self, env): pass
    def simulate(self, substep): pass
    def refresh(self): pass
    def push_resets(self, env_ids): pass


def test_reset_respawns_every_env():
    name = "cfg2_single_path_64"
    fx = load_fixture(name)
    e, w = make_env(name, fx, sim_backend=NullSim())
    od = w.reset()
    f = e.flat
    assert od["obs"].shape == (e.num_envs, e.num_obs) and od["obs_history"].shape == (e.num_envs, w.num_obs_history)
    assert torch.isfinite(od["obs"]).all()
    root = e.root_states.view(e.num_envs, -1, 13)[:, 0]
    want_xy = e.env_origins[:, :2] + torch.tensor(f["base_init_state"][:2], device=DEV)
    torch.testing.assert_close(root[:, :2], want_xy, rtol=0, atol=f["x_init_range"] + f["y_init_range"] + 1e-6)
    torch.testing.assert_close(root[:, 2], torch.full((e.num_envs,), f["base_init_state"][2], device=DEV))
    ratio = e.dof_pos / e.default_dof_pos
    assert float(ratio.min()) >= 0.5 - 1e-6 and float(ratio.max()) <= 1.5 + 1e-6
    assert (e.dof_vel == 0).all()
    ep = e.episode_length_buf
    assert int(ep.min()) >= 0 and int(ep.max()) <= int(e.max_episode_length) + 1
    assert set(e.extras.keys()) >= {"train/episode", "eval/episode", "timeouts"}


def test_runner_learns_two_iterations():
    from legged_tracking_b200 import ppo
    name = "cfg2_single_path_64"
    fx = load_fixture(name)
    e, w = make_env(name, fx, alias_outputs=True)
    saved = (ppo.RunnerArgs.num_steps_per_env, ppo.AC_Args.actor_hidden_dims, ppo.AC_Args.critic_hidden_dims)
    ppo.RunnerArgs.num_steps_per_env = 6
    ppo.AC_Args.actor_hidden_dims = [64, 32]
    ppo.AC_Args.critic_hidden_dims = [64, 32]
    try:
        logs = []
        r = ppo.Runner(w, device=DEV, log_fn=lambda d, step: logs.append((step, d)), save_dir=None)
        before = {k: v.clone() for k, v in r.alg.actor_critic.state_dict().items()}
        r.learn(2)
        assert all(np.isfinite(v) for k, v in r.last_metrics.items())
        after = r.alg.actor_critic.state_dict()
        assert any(not torch.equal(before[k], after[k]) for k in before)
        assert any("metrics" in d for _, d in logs)
        policy = r.get_inference_policy(device=DEV)
        a = policy(w.get_observations())
        assert a.shape == (e.num_envs, 12)
    finally:
        ppo.RunnerArgs.num_steps_per_env, ppo.AC_Args.actor_hidden_dims, ppo.AC_Args.critic_hidden_dims = saved
