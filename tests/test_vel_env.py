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


def test_oracle_restatement_reproduces_the_recording_step_by_step():
    """oracle/vel_port.py, teacher-forced: each function is given the recorded state the reference method saw and must return what
    the reference left (steps without a reset for the post-reset quantities)."""
    from oracle import gait_port, vel_port
    g = load()
    c = cfg_of(g)
    names = c["reward_names"]
    scales = dict(zip(names, c["reward_scales"]))
    feet, term = torch.tensor(c["feet_indices"]), torch.tensor(c["termination_contact_indices"])
    tn = lambda a: torch.from_numpy(np.ascontiguousarray(a))                       # noqa: E731
    prev = sub(g, "init/")
    checked_obs = 0
    for k in range(int(g["meta/steps"])):
        t = f"step{k}/"
        st = sub(g, t + "state/")
        rigid = tn(g[t + "in/rigid"])
        pre = vel_port.prepare(tn(g[t + "in/root"]), rigid, tn(g[t + "in/contact"]), tn(prev["gravity_vec"]), tn(prev["episode_length_buf"]), feet, term,
                               c["max_episode_length"], c["use_terminal_body_height"], c["terminal_body_height"])
        done = g[t + "out/done"]
        assert np.array_equal(pre["reset_buf"].numpy(), done), k
        assert np.array_equal(pre["time_out_buf"].numpy(), st["time_out_buf"]), k
        keep = ~done                                                                # reset envs are overwritten afterwards
        for name in ("base_lin_vel", "base_ang_vel", "projected_gravity"):
            np.testing.assert_allclose(pre[name].numpy(), st[name], rtol=1e-6, atol=1e-7, err_msg=f"{k} {name}")
        assert np.array_equal(pre["foot_positions"].numpy(), st["foot_positions"]) and np.array_equal(pre["foot_velocities"].numpy(), st["foot_velocities"])
        assert np.array_equal(pre["episode_length_buf"].numpy()[keep], st["episode_length_buf"][keep])
        # compute_reward on the CoRL terms evaluated by the (already pinned) gait_port restatement, pre-reset state
        dof = tn(g[t + f"in/dof_state/{c['decimation'] - 1}"]).view(-1, 12, 2)
        # commands as the reward saw them are the recorded post-step commands for every env that neither reset nor resampled this step
        due = (pre["episode_length_buf"].numpy() % int(c["resampling_time"] / c["dt"])) == 0
        cmp = keep & ~due
        assert cmp.sum() >= 32
        if True:
            e = dict(commands=tn(st["commands"]), base_lin_vel=pre["base_lin_vel"], base_ang_vel=pre["base_ang_vel"],
                     projected_gravity=pre["projected_gravity"], gravity_vec=tn(st["gravity_vec"]), base_pos=pre["base_pos"], base_quat=pre["base_quat"],
                     torques=tn(g[t + "out/all_torques"][-1]), dof_pos=dof[..., 0], dof_vel=dof[..., 1], last_dof_vel=tn(prev["last_dof_vel"]),
                     actions=tn(st["actions"]), last_actions=tn(prev["last_actions"]), last_last_actions=tn(prev["last_last_actions"]),
                     joint_pos_target=tn(st["joint_pos_target"]), last_joint_pos_target=tn(prev["last_joint_pos_target"]),
                     last_last_joint_pos_target=tn(prev["last_last_joint_pos_target"]), contact_forces=tn(g[t + "in/contact"]),
                     foot_positions=pre["foot_positions"], foot_velocities=pre["foot_velocities"], prev_foot_velocities=tn(prev["foot_velocities"]),
                     foot_indices=tn(st["foot_indices"]), desired_contact_states=tn(st["desired_contact_states"]),
                     default_dof_pos=tn(np.asarray(c["default_dof_pos"], np.float32)), dof_pos_limits=tn(np.asarray(c["dof_pos_limits"], np.float32)),
                     last_contacts=tn(prev["last_contacts"]), feet_indices=feet, penalised_contact_indices=torch.tensor(c["penalised_contact_indices"]),
                     dt=c["dt"])
            vals, _ = gait_port.corl_terms(e, {kk[len("rewards/"):]: v for kk, v in c.items() if kk.startswith("rewards/")}, names)
            es = {n: tn(prev["episode_sums/" + n]).clone() for n in names + ["total"]}
            cs = {n: tn(prev["command_sums/" + n]).clone() for n in names + ["lin_vel_raw", "ang_vel_raw", "lin_vel_residual", "ang_vel_residual", "ep_timesteps"]}
            rew, pos, neg = vel_port.compute_reward(vals, names, scales, es, cs, pre["base_lin_vel"], pre["base_ang_vel"], e["commands"],
                                                    c["only_positive_rewards"], c["only_positive_rewards_ji22_style"], c["sigma_rew_neg"])
            np.testing.assert_allclose(rew.numpy()[cmp], g[t + "out/rew"][cmp], rtol=1e-5, atol=1e-6, err_msg=f"{k} rew")
            np.testing.assert_allclose(pos.numpy()[cmp], st["rew_buf_pos"][cmp], rtol=1e-5, atol=1e-6, err_msg=f"{k} rew_buf_pos")
            np.testing.assert_allclose(neg.numpy()[cmp], st["rew_buf_neg"][cmp], rtol=1e-5, atol=1e-6, err_msg=f"{k} rew_buf_neg")
            for n in es:
                np.testing.assert_allclose(es[n].numpy()[cmp], st["episode_sums/" + n][cmp], rtol=1e-5, atol=1e-6, err_msg=f"{k} episode_sums/{n}")
            for n in cs:
                np.testing.assert_allclose(cs[n].numpy()[cmp], st["command_sums/" + n][cmp], rtol=1e-5, atol=1e-6, err_msg=f"{k} command_sums/{n}")
        # compute_observations on the post-reset state the reference left
        noise = [g[key] for key in sorted(x for x in g if x.startswith(t + "rng/"))][-1]
        post_dof = tn(st["dof_state"]).view(-1, 12, 2)
        obs, priv = vel_port.compute_observations(c, tn(st["projected_gravity"]), tn(st["commands"]), tn(np.asarray(c["commands_scale"], np.float32)),
                                                  post_dof[..., 0], post_dof[..., 1], tn(np.asarray(c["default_dof_pos"], np.float32)), tn(st["actions"]),
                                                  tn(prev["last_actions"]) * tn((~done).astype(np.float32)).unsqueeze(1), tn(st["gait_indices"]),
                                                  tn(st["clock_inputs"]), tn(noise), tn(np.asarray(c["noise_scale_vec"], np.float32)),
                                                  tn(st["friction_coeffs"])[:, 0], tn(st["restitutions"])[:, 0])
        np.testing.assert_allclose(obs.numpy(), g[t + "out/obs"], rtol=1e-6, atol=1e-7, err_msg=f"{k} obs")
        np.testing.assert_allclose(priv.numpy(), g[t + "out/privileged_obs"], rtol=1e-6, atol=1e-7)
        checked_obs += 1
        prev = st
    assert checked_obs == int(g["meta/steps"])
    h = vel_port.get_heights(tn(g["heights/root_states"]), tn(g["heights/height_points"]), tn(g["heights/height_samples"]), *[float(x) for x in g["heights/scales"]])
    assert np.array_equal(h.numpy(), g["heights/measured_heights"])


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
        close(cpu(eng.tq.state["torques"]), g[t + "out/all_torques"][-1], f"{k} torques", atol=1e-5)
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


@pytest.mark.gpu
def test_cfg4_kernels_match_the_oracle_at_config_4_size():
    """4,096 envs (BASELINE config 4), random state: lt_vel_pre, lt_vel_reward and lt_vel_observe through the C ABI against
    oracle/vel_port.py (itself pinned to the reference recording), including the 30-row history shift and a ragged tail."""
    import ctypes as C
    from legged_tracking_b200 import _abi, vel_env, workload
    from oracle import vel_port
    g = load()
    c = cfg_of(g)
    N, dev = 4096 + 37, torch.device("cuda", 0)
    gen = torch.Generator().manual_seed(17)
    eng = vel_env.VelocityStepEngine(c, N, np.zeros((N, 3), np.float32), workload.load_actuator(), ReplaySource(), device=dev)
    lib = _abi.load()
    r = lambda *s: torch.randn(*s, generator=gen)                                  # noqa: E731
    B = int(c["num_bodies"])
    root = r(N, 13)
    root[:, 3:7] /= root[:, 3:7].norm(dim=1, keepdim=True)
    root[:, 2] = 0.02 + 0.4 * torch.rand(N, generator=gen)
    rigid, contact = r(N, B, 13), r(N, B, 3) * 0.8
    ep = torch.randint(0, 1003, (N,), generator=gen)
    grav = torch.tensor([0.1, -0.05, -0.99]).repeat(N, 1)
    for dst, src in ((eng.root_states, root), (eng.rigid_body_state, rigid), (eng.contact_forces, contact), (eng.episode_length_buf, ep), (eng.gravity_vec, grav)):
        dst.copy_(src.to(dev))
    _abi.check(lib.lt_vel_pre(C.byref(eng._vc), C.byref(eng._ptrs()), None), None, "lt_vel_pre")
    torch.cuda.synchronize()
    want = vel_port.prepare(root, rigid, contact, grav, ep, torch.tensor(c["feet_indices"]), torch.tensor(c["termination_contact_indices"]),
                            c["max_episode_length"], c["use_terminal_body_height"], c["terminal_body_height"])
    for k in ("base_lin_vel", "base_ang_vel", "projected_gravity", "foot_positions", "foot_velocities", "episode_length_buf"):
        close(getattr(eng, k).cpu().numpy(), want[k].numpy(), k)
    for k in ("reset_buf", "time_out_buf", "body_height_buf"):
        assert np.array_equal(getattr(eng, k).cpu().numpy().astype(bool), want[k].numpy()), k
    assert 0 < int(want["reset_buf"].sum()) < N
    iv = int(c["resampling_time"] / c["dt"])
    assert np.array_equal(eng.resample_due.cpu().numpy().astype(bool), (want["episode_length_buf"].numpy() % iv) == 0)
    # ---- compute_reward on random term values, mixed-sign sums
    names = c["reward_names"]
    scales = dict(zip(names, c["reward_scales"]))
    vals = torch.rand(len(names), N, generator=gen) * 3
    vals[3] *= 0                                                                   # an identically-zero term lands in the positive bucket
    cmd = r(N, int(c["num_commands"]))
    eng.commands.copy_(cmd.to(dev))
    es0, cs0 = r(len(names) + 1, N), r(len(names) + 5, N)
    eng.episode_sums.copy_(es0.to(dev))
    eng.command_sums.copy_(cs0.to(dev))
    vd = vals.to(dev)
    _abi.check(lib.lt_vel_reward(C.byref(eng._vc), C.byref(eng._ptrs()), C.c_void_p(vd.data_ptr()), None), None, "lt_vel_reward")
    torch.cuda.synchronize()
    es = {n: es0[i].clone() for i, n in enumerate(names + ["total"])}
    cs = {n: cs0[i].clone() for i, n in enumerate(names + vel_env.RAW_SUMS)}
    rew, pos, neg = vel_port.compute_reward({n: vals[i] for i, n in enumerate(names)}, names, scales, es, cs, want["base_lin_vel"], want["base_ang_vel"],
                                            cmd, c["only_positive_rewards"], c["only_positive_rewards_ji22_style"], c["sigma_rew_neg"])
    close(eng.rew_buf_pos.cpu().numpy(), pos.numpy(), "rew_buf_pos")
    close(eng.rew_buf_neg.cpu().numpy(), neg.numpy(), "rew_buf_neg")
    close(eng.rew_buf.cpu().numpy(), rew.numpy(), "rew_buf", rtol=2e-5, atol=1e-30)
    for i, n in enumerate(names + ["total"]):
        close(eng.episode_sums[i].cpu().numpy(), es[n].numpy(), "episode_sums/" + n)
    for i, n in enumerate(names + vel_env.RAW_SUMS):
        close(eng.command_sums[i].cpu().numpy(), cs[n].numpy(), "command_sums/" + n)
    # ---- compute_observations + history shift
    dof = r(N, 12, 2)
    acts, last, gait, clock = r(N, 12), r(N, 12), torch.rand(N, generator=gen), r(N, 4)
    fric, rest = torch.rand(N, generator=gen) * 3, torch.rand(N, generator=gen) * 0.4
    hist0 = r(N, 30 * 70)
    noise = torch.rand(N, 70, generator=gen)
    for dst, src in ((eng.dof_state, dof), (eng.tq.state["actions"], acts), (eng.last_actions, last), (eng.gait.gait_indices, gait),
                     (eng.gait.clock_inputs, clock), (eng.friction_coeffs, fric), (eng.restitutions, rest), (eng._hist[eng._hist_cur], hist0)):
        dst.copy_(src.to(dev))
    nz = noise.to(dev)
    _abi.check(lib.lt_vel_observe(C.byref(eng._vc), C.byref(eng._ptrs()), C.c_void_p(nz.data_ptr()), C.c_uint32(1), None), None, "lt_vel_observe")
    torch.cuda.synchronize()
    f32 = lambda k: torch.tensor(np.asarray(c[k], np.float32))                      # noqa: E731
    obs, priv = vel_port.compute_observations(c, want["projected_gravity"], cmd, f32("commands_scale"), dof[..., 0], dof[..., 1], f32("default_dof_pos"),
                                              acts, last, gait, clock, noise, f32("noise_scale_vec"), fric, rest)
    close(eng.obs_buf.cpu().numpy(), obs.numpy(), "obs")
    close(eng.privileged_obs_buf.cpu().numpy(), priv.numpy(), "priv")
    new_hist = eng._hist[1 - eng._hist_cur].cpu()
    assert torch.equal(new_hist[:, :29 * 70], hist0[:, 70:]) and torch.equal(new_hist[:, 29 * 70:], eng.obs_buf.cpu())
    assert torch.equal(eng.last_actions.cpu(), acts) and torch.equal(eng.last_last_actions.cpu(), last)
    assert torch.equal(eng.last_dof_vel.cpu(), dof[..., 1]) and torch.equal(eng.last_root_vel.cpu(), root[:, 7:13])
    # in-kernel Philox noise: uniform, column-independent, keyed by step
    eng.obs_buf.zero_()
    _abi.check(lib.lt_vel_observe(C.byref(eng._vc), C.byref(eng._ptrs()), None, C.c_uint32(2), None), None, "lt_vel_observe")
    torch.cuda.synchronize()
    got = eng.obs_buf.cpu()
    scale = f32("noise_scale_vec")
    noisy = scale > 0
    clean, _ = vel_port.compute_observations(dict(c, add_noise=False), want["projected_gravity"], cmd, f32("commands_scale"), dof[..., 0], dof[..., 1],
                                             f32("default_dof_pos"), acts, acts, gait, clock, noise, scale, fric, rest)
    u = ((got - clean)[:, noisy] / scale[noisy] + 1) / 2
    inside = (clean[:, noisy].abs() < c["clip_observations"] - 1)                   # un-clipped entries only
    uu = u[inside]
    assert 0.0 <= float(uu.min()) and float(uu.max()) <= 1.0 and abs(float(uu.mean()) - 0.5) < 0.01 and abs(float(uu.std()) - 0.2887) < 0.01


# ------------------------------------------------------------------------------------------------------------------------
# device mode: the curriculum and the three indexed paths as kernels (lt_vel_indexed_paths), no host read in the step
def _device_engine(g, N, seed=3):
    from legged_tracking_b200 import vel_env, workload
    c = cfg_of(g)
    origins = np.zeros((N, 3), np.float32)
    origins[:, 0] = np.arange(N) % 97
    origins[:, 1] = np.arange(N) // 97
    return vel_env.VelocityStepEngine(c, N, origins, workload.load_actuator(), "device", device=torch.device("cuda", 0), seed=seed), c


@pytest.mark.gpu
def test_device_curriculum_update_equals_the_reference_class():
    """RewardThresholdCurriculum.update (curriculum.py:133-152) recorded from the unmodified class on seeded inputs, two calls in a
    row: the device-side weights after lt_vel_indexed_paths equal the reference's (float64, 1e-12)."""
    from legged_tracking_b200 import _abi
    g = load()
    eng, c = _device_engine(g, 64)
    np.testing.assert_array_equal(eng.cur_weights[0].cpu().numpy(), g["curriculum/weights0"])
    for call in range(2):
        bins, rewards, thr = g[f"curriculum/update{call}/bins"], g[f"curriculum/update{call}/rewards"], g[f"curriculum/update{call}/thresholds"]
        nb = len(bins)
        eng._dc.ep_len = 1.0
        for k in range(4):
            eng._dc.key_thresholds[k] = float(thr[k])
            eng.command_sums[eng._dc.key_rows[k]].zero_()
            eng.command_sums[eng._dc.key_rows[k], :nb] = torch.from_numpy(rewards[k]).cuda()
        eng.env_bins_dev.zero_()
        eng.env_bins_dev[:nb] = torch.from_numpy(bins.astype(np.int32)).cuda()
        eng.env_categories_dev.zero_()
        eng.resample_due.zero_()
        eng.resample_due[:nb] = 1
        eng.dr_due.zero_()
        eng._device_paths(_abi.DEFINES["LT_VEL_PATH_CALLBACK"])
        torch.cuda.synchronize()
        np.testing.assert_allclose(eng.cur_weights[0].cpu().numpy(), g[f"curriculum/update{call}/weights"], rtol=0, atol=1e-12)
        assert float(eng.cur_weights[1:].sub(torch.from_numpy(g["curriculum/weights0"]).cuda()).abs().max()) == 0.0      # other categories untouched
        assert int(eng._hits.abs().sum()) == 0


@pytest.mark.gpu
def test_device_sampler_draws_bins_by_weight_and_points_inside_the_bin():
    """Curriculum.sample (curriculum.py:68-91) on the device: bins in proportion to the weights, a point uniform inside the cell,
    categories uniform (LRV:760-766), the gait rewrite of lt_resample_commands applied, command_sums zeroed."""
    g = load()
    N = 65536
    eng, c = _device_engine(g, N)
    grid, sizes = g["curriculum/grid"], g["curriculum/bin_sizes"]
    w = np.zeros(grid.shape[1])
    active = np.nonzero(g["curriculum/update1/weights"] > 0)[0]
    w[active] = g["curriculum/update1/weights"][active] * (1 + (np.arange(len(active)) % 3))        # unequal weights
    w = np.minimum(w, 1.0)
    eng.cur_weights.copy_(torch.from_numpy(np.tile(w, (4, 1))).cuda())
    eng.command_sums.fill_(-1e9)                                                                      # never a curriculum success
    eng.reset_all()
    torch.cuda.synchronize()
    bins, cats, cmd = eng.env_bins_dev.cpu().numpy(), eng.env_categories_dev.cpu().numpy(), eng.commands.cpu().numpy().astype(np.float64)
    assert set(np.unique(bins)) <= set(active.tolist())
    freq = np.bincount(bins, minlength=len(w)) / N
    np.testing.assert_allclose(freq[active], w[active] / w.sum(), atol=4 * np.sqrt(1.0 / (len(active) * N)) + 2e-3)
    np.testing.assert_allclose(np.bincount(cats, minlength=4) / N, 0.25, atol=0.01)
    ctr = grid[:, bins].T                                                                             # [N, 15]
    untouched = [0, 1, 2, 3, 4, 8, 9, 10, 11, 12, 13, 14]                                             # columns the gait rewrite leaves alone
    moving = (np.linalg.norm(cmd[:, :2], axis=1) > 0)                                                 # small planar commands were zeroed (LRV:836)
    for k in untouched:
        rows = moving if k < 2 else slice(None)
        assert np.all(np.abs(cmd[rows, k] - ctr[rows, k]) <= sizes[k] / 2 + 1e-6), k
    frac = (cmd[moving, 0] - (ctr[moving, 0] - sizes[0] / 2)) / sizes[0]                              # uniform inside the cell
    assert abs(frac.mean() - 0.5) < 0.01 and abs(frac.std() - 0.2887) < 0.01
    assert set(np.unique(cmd[:, 5:8]).tolist()) <= {0.0, 0.5}                                         # binary phases after the gait rewrite
    trot = cats == 1
    assert np.all(cmd[trot, 6] == 0) and np.all(cmd[trot, 7] == 0)
    assert float(eng.command_sums.abs().max()) == 0.0
    assert float(eng.episode_length_buf.abs().max()) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("custom_origins", [False, True])
def test_device_reset_path_equals_the_host_orchestrated_one(custom_origins):
    """The reset path as a kernel against the torch path that free-runs the reference recording (same uniforms, injected per env):
    DR tables, dof and root rows, buffers, lag ring, episode bookkeeping; and the callback's interval DR."""
    from legged_tracking_b200 import _abi, vel_env, workload
    g = load()
    N = 1000
    c = dict(cfg_of(g), custom_origins=custom_origins)          # True: heightfield / trimesh terrains, x / y init offsets drawn (LRV:991-1001)
    origins = np.zeros((N, 3), np.float32)
    origins[:, 0], origins[:, 1] = np.arange(N) % 97, np.arange(N) // 97
    dev_eng = vel_env.VelocityStepEngine(c, N, origins, workload.load_actuator(), "device", device=torch.device("cuda", 0), seed=3)
    host_eng = vel_env.VelocityStepEngine(c, N, origins, workload.load_actuator(), ReplaySource(), device=torch.device("cuda", 0))
    gen = torch.Generator().manual_seed(9)
    U = torch.rand(N, _abi.DEFINES["LT_VEL_SLOTS"], generator=gen)
    reset = torch.rand(N, generator=gen) < 0.3
    dr = torch.rand(N, generator=gen) < 0.2
    ids = reset.nonzero().flatten()
    for e in (dev_eng, host_eng):                                                   # the same dirty state on both
        g2 = torch.Generator().manual_seed(4)
        for t in (e.dof_state, e.root_states, e.last_actions, e.last_last_actions, e.last_dof_vel, e.feet_air_time, e.episode_sums,
                  e.tq.state["lag_buffer"], e.tq.state["motor_strengths"], e.tq.state["motor_offsets"], e.gait.gait_indices):
            t.copy_(torch.randn(t.shape, generator=g2).to(t.device))
        e.episode_length_buf.copy_(torch.randint(1, 900, (N,), generator=g2).cuda())
        e.reset_buf.copy_(reset.to(torch.uint8).cuda())
        e.dr_due.copy_(dr.to(torch.uint8).cuda())
        e.resample_due.zero_()
    want_means = {"rew_" + n: float(host_eng.episode_sums[i][ids.cuda()].mean()) for i, n in enumerate(host_eng.reward_names + ["total"])}
    # host path: reset_idx draws, in the reference's order -- category, motor strength, motor offsets, dof positions, yaw, root velocity
    draws = [U[ids, 0], U[ids, 18], U[ids, 19:31], U[ids, 31:43]] + ([U[ids, 50:51], U[ids, 51:52]] if custom_origins else []) + \
        [U[ids, 43:44], U[ids, 44:50]]
    host_eng.rng = ReplayRng([d.numpy() for d in draws], torch.device("cuda", 0))
    host_eng.source.q = [(np.zeros((int((host_eng.resampler.categories_from_uniform(U[ids, 0].cuda()) == k).sum()), 15)), np.zeros(0, np.int64))
                         for k in range(4) if int((host_eng.resampler.categories_from_uniform(U[ids, 0].cuda()) == k).sum())]
    host_eng.source.sample = lambda cat, n, q=host_eng.source.q: (q.pop(0)[0], np.zeros(n, np.int64))
    host_eng.reset_idx(ids.cuda())
    dev_eng._device_paths(_abi.DEFINES["LT_VEL_PATH_RESET"], U)
    torch.cuda.synchronize()
    for name in ("dof_state", "last_actions", "last_last_actions", "last_dof_vel", "feet_air_time", "episode_sums", "episode_length_buf"):
        assert torch.equal(getattr(dev_eng, name), getattr(host_eng, name)), name
    np.testing.assert_allclose(dev_eng.root_states.cpu().numpy(), host_eng.root_states.cpu().numpy(), rtol=1e-6, atol=1e-7)     # sin / cos of the yaw
    for name in ("lag_buffer", "motor_strengths", "motor_offsets"):
        assert torch.equal(dev_eng.tq.state[name], host_eng.tq.state[name]), name
    assert torch.equal(dev_eng.gait.gait_indices, host_eng.gait.gait_indices)
    assert np.array_equal(dev_eng.env_categories_dev.cpu().numpy()[ids], host_eng.env_command_categories[ids])
    got = dev_eng.episode_means()
    for k, v in want_means.items():
        assert got[k] == pytest.approx(v, rel=1e-4, abs=1e-5), k
    # the callback path: interval DR for the dr_due envs only
    before = {k: host_eng.tq.state[k].clone() for k in ("motor_strengths", "motor_offsets")}
    dr_ids = dr.nonzero().flatten()
    host_eng.rng = ReplayRng([U[dr_ids, 18].numpy(), U[dr_ids, 19:31].numpy()], torch.device("cuda", 0))
    host_eng._randomize_dof_props(dr_ids.cuda())
    dev_eng._device_paths(_abi.DEFINES["LT_VEL_PATH_CALLBACK"], U)
    torch.cuda.synchronize()
    for name in ("motor_strengths", "motor_offsets"):
        assert torch.equal(dev_eng.tq.state[name], host_eng.tq.state[name]), name
        assert not torch.equal(before[name], host_eng.tq.state[name])


@pytest.mark.gpu
def test_device_mode_runs_without_a_host_source_and_keeps_the_invariants():
    """Twenty full steps in device mode on synthetic simulator state: resets and resampling happen, everything stays finite,
    reset envs come out of the step with episode length 0 and zeroed buffers, commands stay inside the curriculum's limits, the
    curriculum's weights only grow and stay in [0, 1], and the observation history is the shifted one."""
    from legged_tracking_b200 import synthetic
    g = load()
    N = 2048
    eng, c = _device_engine(g, N)
    synth = synthetic.SyntheticState(N, eng.env_origins.cpu().numpy(), num_bodies=int(c["num_bodies"]), seed=5, feet_indices=list(c["feet_indices"]))
    eng.episode_length_buf.copy_(synth.initial_episode_length(c["max_episode_length"]).cuda())
    eng.reset_all()
    w0 = eng.cur_weights.clone()
    lim = np.asarray(c["curriculum_limits"])
    n_resets = 0
    for k in range(20):
        root = synth.root(k)
        sim = {"dof_state": [synth.dof_state(k, s).cuda() for s in range(int(c["decimation"]))], "root": root.cuda(), "contact": synth.contact(k).cuda(),
               "rigid": synth.rigid(k, root).cuda()}
        prev_hist = eng.obs_history.clone()
        obs, rew, done, extras = eng.step(synth.actions(k).contiguous().cuda(), sim)
        torch.cuda.synchronize()
        n_resets += int(done.sum())
        for t in (obs["obs"], obs["obs_history"], obs["privileged_obs"], rew, eng.commands, eng.episode_sums, eng.command_sums):
            assert torch.isfinite(t).all()
        assert torch.equal(obs["obs_history"][:, :-70], prev_hist[:, 70:]) and torch.equal(obs["obs_history"][:, -70:], obs["obs"])
        assert int(eng.episode_length_buf[done].abs().sum()) == 0 and float(eng.last_dof_vel[done].abs().sum()) == 0.0
        assert float(eng.episode_sums[:, done].abs().sum()) == 0.0
        cmd = eng.commands.cpu().numpy()
        for d in (0, 1, 2, 3, 4, 9, 10, 12, 13):
            assert np.all(cmd[:, d] >= lim[d, 0] - 1e-6) and np.all(cmd[:, d] <= lim[d, 1] + 1e-6), d
    assert n_resets > 0.02 * N * 20
    assert bool((eng.cur_weights >= w0).all()) and float(eng.cur_weights.max()) <= 1.0
    means = eng.episode_means()
    assert set(means) == {"rew_" + n for n in eng.reward_names + ["total"]} and all(np.isfinite(v) for v in means.values())


@pytest.mark.gpu
def test_device_mode_graph_replay_equals_eager_steps():
    """Two steps captured as one CUDA graph (device-side Philox step counter, ping-ponged history) replayed twice give, bit for bit,
    what four eager steps give."""
    from legged_tracking_b200 import synthetic
    g = load()
    N = 1536
    engines = []
    for _ in range(2):
        eng, c = _device_engine(g, N, seed=21)
        synth = synthetic.SyntheticState(N, eng.env_origins.cpu().numpy(), num_bodies=int(c["num_bodies"]), seed=6, feet_indices=list(c["feet_indices"]))
        eng.episode_length_buf.copy_(synth.initial_episode_length(c["max_episode_length"]).cuda())
        eng.reset_all()
        frames = []
        for k in range(2):
            root = synth.root(k)
            frames.append((synth.actions(k).contiguous().cuda(),
                           {"dof_state": [synth.dof_state(k, s).cuda() for s in range(int(c["decimation"]))], "root": root.cuda(),
                            "contact": synth.contact(k).cuda(), "rigid": synth.rigid(k, root).cuda()}))
        for a, sim in frames:                                                    # two eager warm-up steps on both
            eng.step(a, sim)
        engines.append((eng, frames))
    (eager, fe), (graphed, fg) = engines
    for _ in range(2):
        for a, sim in fe:
            eager.step(a, sim)
    assert graphed.graph_replay_ok(2)
    graph = graphed.capture(fg)
    for _ in range(2):
        graphed.replay(graph)
    torch.cuda.synchronize()
    assert eager.common_step_counter == graphed.common_step_counter and int(eager.step_counter_dev) == int(graphed.step_counter_dev) == 6
    for name in ("obs_buf", "obs_history", "rew_buf", "commands", "episode_sums", "command_sums", "episode_length_buf", "root_states", "dof_state",
                 "last_actions", "last_last_actions", "cur_weights", "env_bins_dev", "reset_buf"):
        assert torch.equal(getattr(eager, name), getattr(graphed, name)), name
    assert torch.equal(eager.tq.state["lag_buffer"], graphed.tq.state["lag_buffer"]) and torch.equal(eager.tq.counters, graphed.tq.counters)
    assert int(eager.reset_buf.sum()) > 0


# ------------------------------------------------------------------------------------------------------------------------
# facade and configuration mapping
def test_config_from_the_reference_cfg_tree_equals_the_recorded_table():
    """vel_env.config_from_cfg on the reference's own Cfg class, edited by the config block of scripts/train_velocity_tracking.py
    (executed from the reference file), gives the table recorded off the live reference env."""
    from oracle import ref_harness
    if not ref_harness.reference_available():
        pytest.skip("reference tree not present on this machine")
    from legged_tracking_b200 import vel_env
    from oracle import gen_golden_vel
    ref_harness.install()
    Cfg = gen_golden_vel.build_cfg(64)
    got, want = vel_env.config_from_cfg(Cfg), vel_env.cfg4_config()
    assert set(got) == set(want), sorted(set(got) ^ set(want))
    for k, w in want.items():
        g_ = got[k]
        if isinstance(w, (list, float)) and not (isinstance(w, list) and w and isinstance(w[0], str)):
            np.testing.assert_allclose(np.asarray(g_, np.float64), np.asarray(w, np.float64), rtol=1e-7, atol=1e-9, err_msg=k)
        else:
            assert g_ == w, k


@pytest.mark.gpu
def test_facade_steps_like_the_engine_and_keeps_the_vecenv_surface():
    """VelocityTrackingEasyEnv(sim_device, headless, cfg=None, sim=VelReplaySim()): reset() + steps through the SimBackend protocol
    give what the engine gives when it is handed the same frames directly; the VecEnv attributes the runner reads exist."""
    from legged_tracking_b200 import vel_env, workload
    N = 1024
    env = vel_env.VelocityTrackingEasyEnv("cuda:0", True, num_envs=N, sim=vel_env.VelReplaySim(seed=44), seed=5)
    obs = env.reset()
    assert obs["obs"].shape == (N, 70) and obs["obs_history"].shape == (N, 2100) and obs["privileged_obs"].shape == (N, 2)
    assert env.num_obs == 70 and env.num_obs_history == 2100 and env.num_privileged_obs == 2 and env.num_actions == 12
    assert float(obs["obs_history"].abs().sum()) == 0.0
    ref = vel_env.VelocityStepEngine(vel_env.cfg4_config(), N, env.engine.env_origins.cpu().numpy(), workload.load_actuator(), "device",
                                     device=torch.device("cuda", 0), seed=5)
    sim = vel_env.VelReplaySim(seed=44)
    sim.attach(type("E", (), {"engine": ref, "dof_state": ref.dof_state, "root_states": ref.root_states, "contact_forces": ref.contact_forces,
                              "rigid_body_state": ref.rigid_body_state})())
    ref.reset_buf.fill_(1)
    ref._device_paths(1)
    zero = torch.zeros(N, 12, device="cuda")
    f0 = sim.frames[0]
    ref.step(zero, {"dof_state": f0["dof"], "root": f0["root"], "contact": f0["contact"], "rigid": f0["rigid"]})
    ref.obs_history.zero_()
    gen = torch.Generator(device="cuda").manual_seed(2)
    for k in range(1, 5):
        a = torch.randn(N, 12, device="cuda", generator=gen) * 0.3
        o, r, d, info = env.step(a)
        f = sim.frames[k % 3]
        o2, r2, d2, _ = ref.step(a, {"dof_state": f["dof"], "root": f["root"], "contact": f["contact"], "rigid": f["rigid"]})
        torch.cuda.synchronize()
        assert torch.equal(o["obs"], o2["obs"]) and torch.equal(o["obs_history"], o2["obs_history"]) and torch.equal(r, r2) and torch.equal(d, d2)
        assert "time_outs" in info and info["privileged_obs"].shape == (N, 2)
    assert env.rew_buf.shape == (N,) and env.episode_length_buf.dtype == torch.int64 and env.commands.shape == (N, 15)


@pytest.mark.gpu
def test_config_4_trains_end_to_end_with_the_ppo_kernels():
    """workload.make_cfg4 + ppo.Runner: two learning iterations (24-step rollouts of the cfg 4 env in device mode, GAE, the PPO
    update on the 2,100-wide history) run, stay finite and move the parameters -- the reference's train_velocity_tracking.py loop
    (Runner(HistoryWrapper(env)).learn) on the hand-written path end to end."""
    from legged_tracking_b200 import ppo, workload
    torch.manual_seed(0)
    env = workload.make_cfg4(num_envs=1024)
    runner = ppo.Runner(env, device="cuda:0")
    before = runner.alg.actor_critic.flat.clone()
    runner.learn(num_learning_iterations=2, init_at_random_ep_len=True)
    torch.cuda.synchronize()
    after = runner.alg.actor_critic.flat
    assert torch.isfinite(after).all() and float((after - before).abs().max()) > 0
    assert runner.alg.actor_critic.num_obs_history == 2100
    assert torch.isfinite(env.obs_buf).all() and torch.isfinite(env.rew_buf).all()


@pytest.mark.gpu
def test_bound_simulator_tensors_give_the_same_steps_as_copied_ones():
    """bind_inputs: stepping on the simulator's own tensors (re-pointed, rewritten in place on reset) equals stepping on copies."""
    from legged_tracking_b200 import synthetic
    g = load()
    N = 1024
    engs = []
    for bind in (False, True):
        eng, c = _device_engine(g, N, seed=8)
        eng.bind_inputs = bind
        synth = synthetic.SyntheticState(N, eng.env_origins.cpu().numpy(), num_bodies=int(c["num_bodies"]), seed=3, feet_indices=list(c["feet_indices"]))
        eng.episode_length_buf.copy_(synth.initial_episode_length(c["max_episode_length"]).cuda())
        eng.reset_all()
        for k in range(6):
            root = synth.root(k)
            sim = {"dof_state": [synth.dof_state(k, s).cuda() for s in range(int(c["decimation"]))], "root": root.cuda(), "contact": synth.contact(k).cuda(),
                   "rigid": synth.rigid(k, root).cuda()}
            out = eng.step(synth.actions(k).contiguous().cuda(), sim)
        torch.cuda.synchronize()
        engs.append((eng, out))
    (a, oa), (b, ob) = engs
    for name in ("obs_buf", "obs_history", "rew_buf", "commands", "episode_sums", "command_sums", "episode_length_buf", "root_states", "dof_state",
                 "last_actions", "cur_weights", "reset_buf"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    assert torch.equal(a.tq.counters, b.tq.counters) and int(a.tq.counters[0]) == 6 and torch.equal(a.tq.state["lag_buffer"], b.tq.state["lag_buffer"])
