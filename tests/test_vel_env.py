"""BASELINE config 4: the velocity-tracking env's step as the liblt_b200 launch chain
(legged_tracking_b200/vel_env.py) against tests/golden/vel_cfg4_64.npz -- 10 steps of the UNMODIFIED reference env
(VelocityTrackingEasyEnv + HistoryWrapper, scripts/train_velocity_tracking.py's configuration on plane terrain; oracle/gen_golden_vel.py)
with 62 resets, command resampling on the interval and at reset, one gravity randomisation and the interval DR.

The engine runs FREE from the recorded initial state: per step it only gets the actions, the simulator tensors, the reference's
uniform draws (in call order) and what its curricula returned.  Bars: integer / boolean state bit-exact; fp32 1e-5 relative
(+2e-6; exp / sin / erf differ from torch's CPU library in the last bits); torques as in the trajectory tests."""
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "vel_cfg4_64.npz")
RTOL, ATOL = 1e-5, 2e-6


def load():
    z = np.load(GOLDEN, allow_pickle=False)
    return {k: z[k] for k in z.files}


def sub(g, prefix):
    return {k[len(prefix):]: v for k, v in g.items() if k.startswith(prefix)}


def cfg_of(g):
    cfg = {}
    for k, v in sub(g, "cfg/").items():
        if v.dtype.kind in "US":
            cfg[k] = [str(x) for x in v.tolist()] if v.ndim else str(v)
        else:
            cfg[k] = v.item() if v.ndim == 0 else v.tolist()
    return cfg


def test_product_config_table_is_the_recorded_one():
    from legged_tracking_b200 import vel_env
    g = load()
    want = cfg_of(g)
    want.pop("env_origins")
    want.pop("num_envs")
    assert vel_env.cfg4_config() == want
    assert want["num_obs"] == 70 and want["obs_history_length"] == 30 and len(want["reward_names"]) == 19


def test_recording_exercises_the_rare_paths():
    g = load()
    steps = int(g["meta/steps"])
    assert sum(int(g[f"step{k}/out/done"].sum()) for k in range(steps)) >= 40
    assert any(f"step{k}/curriculum/0/commands" in g for k in range(steps))
    grav = [g[f"step{k}/state/gravity_vec"][0] for k in range(steps)]
    assert any(not np.array_equal(grav[0], x) for x in grav)                       # _randomize_gravity fired
    ep = g["init/episode_length_buf"]
    assert ((ep % 500) > 480).any() and (ep > 990).any()                           # resampling interval and time-outs inside the run


class ReplayRng:
    def __init__(self, draws, device):
        self.q = [torch.from_numpy(d) for d in draws]
        self.device = device

    def __call__(self, *shape):
        while self.q and self.q[0].numel() == 0:                                   # torch.rand(0) of an empty interval set
            self.q.pop(0)
        t = self.q.pop(0)
        assert tuple(t.shape) == tuple(shape), (tuple(t.shape), shape)
        return t.to(self.device)


class ReplaySource:
    def __init__(self):
        self.q = []

    def update(self, *a):
        pass

    def sample(self, category, batch_size):
        cmds, bins = self.q.pop(0)
        assert cmds.shape[0] == batch_size
        return cmds, bins


def close(got, want, what, rtol=RTOL, atol=ATOL):
    got, want = np.asarray(got), np.asarray(want)
    if want.dtype.kind in "biu":
        assert np.array_equal(got.astype(want.dtype), want), what
    else:
        np.testing.assert_allclose(got, want, rtol=rtol, atol=atol, err_msg=what)


@pytest.mark.gpu
def test_cfg4_step_chain_free_runs_the_reference_recording():
    from legged_tracking_b200 import vel_env, workload
    g = load()
    cfg = cfg_of(g)
    N, dev = int(cfg["num_envs"]), torch.device("cuda", 0)
    src = ReplaySource()
    eng = vel_env.VelocityStepEngine(cfg, N, np.asarray(cfg["env_origins"]), workload.load_actuator(), src, device=dev)
    eng.load_reference_state(sub(g, "init/"))
    eng._hist[eng._hist_cur].copy_(torch.from_numpy(g["init/obs_history"]).to(dev))
    names = eng.reward_names
    for k in range(int(g["meta/steps"])):
        t = f"step{k}/"
        draws = [g[key] for key in sorted(x for x in g if x.startswith(t + "rng/"))]
        kinds = [key.split("_", 1)[1] for key in sorted(x for x in g if x.startswith(t + "rng/"))]
        assert kinds[-1] == "rand_like" and "rand_like" not in kinds[:-1]          # the observation noise is the last draw of a step
        eng.rng = ReplayRng(draws[:-1], dev)
        j = 0
        src.q = []
        while t + f"curriculum/{j}/commands" in g:
            src.q.append((g[t + f"curriculum/{j}/commands"], g[t + f"curriculum/{j}/bins"]))
            j += 1
        tn = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)            # noqa: E731
        sim = {"dof_state": [tn(g[t + f"in/dof_state/{s}"]) for s in range(int(cfg["decimation"]))], "root": tn(g[t + "in/root"]),
               "contact": tn(g[t + "in/contact"]), "rigid": tn(g[t + "in/rigid"])}
        obs, rew, done, extras = eng.step(tn(g[t + "in/actions"]), sim, noise_uniforms=tn(draws[-1]))
        torch.cuda.synchronize()
        assert not [d for d in eng.rng.q if d.numel()] and not src.q, f"step {k}: unconsumed draws / samples"
        st = sub(g, t + "state/")
        cpu = lambda x: x.detach().cpu().numpy()                                     # noqa: E731
        close(cpu(done), g[t + "out/done"], f"{k} done")
        close(cpu(eng.episode_length_buf), st["episode_length_buf"], f"{k} episode_length_buf")
        close(cpu(eng.time_out_buf).astype(bool), st["time_out_buf"], f"{k} time_out_buf")
        close(cpu(eng.tq.state["torques"]), g[t + "out/all_torques"][-1], f"{k} torques", atol=2e-5)
        for name in ("base_lin_vel", "base_ang_vel", "projected_gravity", "foot_positions", "foot_velocities", "prev_foot_velocities", "commands",
                     "rew_buf_pos", "rew_buf_neg", "last_actions", "last_last_actions", "last_dof_vel", "last_root_vel", "last_joint_pos_target",
                     "last_last_joint_pos_target", "root_states", "gravity_vec"):
            close(cpu(getattr(eng, name)), st[name].reshape(cpu(getattr(eng, name)).shape), f"{k} {name}", atol=2e-5 if "joint_pos_target" in name else ATOL)
        close(cpu(eng.dof_state).reshape(-1, 2), st["dof_state"], f"{k} dof_state")
        for name in ("gait_indices", "foot_indices", "clock_inputs", "doubletime_clock_inputs", "halftime_clock_inputs", "desired_contact_states"):
            close(cpu(getattr(eng.gait, name)), st[name], f"{k} {name}")
        close(cpu(eng.last_contacts), st["last_contacts"], f"{k} last_contacts")
        close(cpu(eng.tq.state["motor_strengths"]), st["motor_strengths"], f"{k} motor_strengths")
        close(cpu(eng.tq.state["motor_offsets"]), st["motor_offsets"], f"{k} motor_offsets")
        close(cpu(rew), g[t + "out/rew"], f"{k} rew", rtol=2e-5, atol=5e-6)
        for i, n in enumerate(names + ["total"]):
            close(cpu(eng.episode_sums[i]), st["episode_sums/" + n], f"{k} episode_sums/{n}", rtol=2e-5, atol=5e-6)
        for i, n in enumerate(names + vel_env.RAW_SUMS):
            close(cpu(eng.command_sums[i]), st["command_sums/" + n], f"{k} command_sums/{n}", rtol=2e-5, atol=5e-6)
        close(cpu(obs["obs"]), g[t + "out/obs"], f"{k} obs")
        close(cpu(obs["privileged_obs"]), g[t + "out/privileged_obs"], f"{k} privileged_obs")
        close(cpu(obs["obs_history"]), g[t + "out/obs_history"], f"{k} obs_history")
        assert np.array_equal(eng.env_command_categories, st["env_command_categories"]), f"{k} categories"
        assert eng.common_step_counter == int(st["common_step_counter"])


@pytest.mark.gpu
def test_measure_heights_kernel_matches_the_reference_method():
    """lt_vel_measure_heights against LeggedRobot._get_heights (LRV:1790-1824) recorded on a synthetic int16 heightfield: 17 x 11
    points, yaw rotation, border offset, truncation, index clipping (robots past the border), minimum of three cells.  Cells are
    integers x 0.005: bit-exact; the body-height mean (LRV:164) to 1e-6."""
    import ctypes as C
    from legged_tracking_b200 import _abi
    g = load()
    lib = _abi.load()
    dev = torch.device("cuda", 0)
    pts, hs, root = (torch.from_numpy(g["heights/" + k]).to(dev).contiguous() for k in ("height_points", "height_samples", "root_states"))
    border, hscale, vscale = (float(x) for x in g["heights/scales"])
    N, P = root.shape[0], pts.shape[0]
    assert P == 17 * 11
    out, mean = torch.empty(N, P, device=dev), torch.empty(N, device=dev)
    p = lambda t: C.c_void_p(t.data_ptr())                                           # noqa: E731
    _abi.check(lib.lt_vel_measure_heights(p(root), p(pts), p(hs), N, P, hs.shape[0], hs.shape[1], C.c_float(border), C.c_float(hscale),
                                          C.c_float(vscale), p(out), p(mean), None), None, "lt_vel_measure_heights")
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), g["heights/measured_heights"])
    np.testing.assert_allclose(mean.cpu().numpy(), g["heights/body_height_mean"], rtol=1e-6, atol=1e-6)
