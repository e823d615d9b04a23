"""Records tests/golden/vel_cfg4_64.npz: the UNMODIFIED reference velocity-tracking env (go1_gym/envs/go1/velocity_tracking
VelocityTrackingEasyEnv + HistoryWrapper; BASELINE config 4) stepped on CPU under the isaacgym / params_proto / gym shims with
replayed synthetic simulator state.  Build container only (needs /root/reference).

Configuration: the config block of scripts/train_velocity_tracking.py, EXECUTED from the reference file (not restated), with one
change: terrain.mesh_type = "plane" (the trimesh generators live in isaacgym.terrain_utils, which is not in this image; cfg 4
observes no heights, so the terrain only places the env origins).

Recorded per step: the actions, the simulator tensors the step reads, every torch.rand / rand_like draw in call order, what the
command curricula returned, the 4-tuple and obs_history the wrapper returns, and a full state dump.

    python -m oracle.gen_golden_vel
"""
import contextlib
import io
import os
import textwrap

import numpy as np
import torch

from oracle import ref_harness as H

OUT = os.path.join(H.REPO, "tests", "golden", "vel_cfg4_64.npz")
STATE = ["episode_length_buf", "base_lin_vel", "base_ang_vel", "projected_gravity", "foot_positions", "foot_velocities", "prev_foot_velocities",
         "commands", "gait_indices", "foot_indices", "clock_inputs", "doubletime_clock_inputs", "halftime_clock_inputs", "desired_contact_states",
         "rew_buf", "rew_buf_pos", "rew_buf_neg", "reset_buf", "time_out_buf", "actions", "last_actions", "last_last_actions", "last_dof_vel",
         "last_root_vel", "joint_pos_target", "last_joint_pos_target", "last_last_joint_pos_target", "root_states", "dof_state", "motor_strengths",
         "motor_offsets", "torques", "joint_pos_err_last", "joint_pos_err_last_last", "joint_vel_last", "joint_vel_last_last", "gravity_vec",
         "gravities", "last_contacts", "obs_buf", "privileged_obs_buf", "friction_coeffs", "restitutions"]


def build_cfg(num_envs):
    from go1_gym.envs.base.legged_robot_velocity_tracking_config import Cfg
    from go1_gym.envs.go1.go1_config import config_go1
    src = open(os.path.join(H.REF_ROOT, "scripts", "train_velocity_tracking.py")).read()
    blk = src[src.index("    config_go1(Cfg)"):src.index("    env = VelocityTrackingEasyEnv(")]
    exec(textwrap.dedent(blk), {"Cfg": Cfg, "config_go1": config_go1})
    Cfg.terrain.mesh_type = "plane"
    Cfg.env.num_envs = num_envs
    return Cfg


def dump_state(env, fx, tag):
    for k in STATE:
        if not hasattr(env, k):                         # created by the first _step_contact_targets / compute_observations
            continue
        v = getattr(env, k)
        fx[tag + k] = (v.detach().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)).copy()
    for k, v in env.episode_sums.items():
        fx[tag + "episode_sums/" + k] = v.detach().numpy().copy()
    for k, v in env.command_sums.items():
        fx[tag + "command_sums/" + k] = v.detach().numpy().copy()
    for s, buf in enumerate(env.lag_buffer):
        fx[tag + f"lag_buffer/{s}"] = buf.detach().numpy().copy()
    fx[tag + "common_step_counter"] = np.int64(env.common_step_counter)
    fx[tag + "env_command_categories"] = np.asarray(env.env_command_categories).copy()


def record_heights(fx, Cfg):
    """LeggedRobot._init_height_points / _get_heights (LRV:1773-1824), unbound, on a namespace with a synthetic int16 heightfield
    (cfg 4's real terrain is a trimesh whose generators are not in this image; the lookup itself only needs the field)."""
    import types
    import go1_gym.envs.base.legged_robot_velocity_tracking as LRV
    g = torch.Generator().manual_seed(12)
    N, rows, cols = 48, 160, 120
    ns = types.SimpleNamespace(device="cpu", num_envs=N)
    saved = Cfg.terrain.mesh_type
    Cfg.terrain.mesh_type = "trimesh"
    try:
        ns.height_points = LRV.LeggedRobot._init_height_points(ns, torch.arange(N), Cfg)
        ns.height_samples = torch.randint(-40, 200, (rows, cols), generator=g, dtype=torch.int16)
        ns.terrain = types.SimpleNamespace(cfg=types.SimpleNamespace(border_size=2.5, horizontal_scale=0.1, vertical_scale=0.005))
        ns.root_states = torch.zeros(N, 13)
        ns.root_states[:, 0] = torch.rand(N, generator=g) * 16 - 3.0          # some robots past the border: indices get clipped
        ns.root_states[:, 1] = torch.rand(N, generator=g) * 12 - 3.0
        ns.root_states[:, 2] = 0.3 + 0.1 * torch.rand(N, generator=g)
        q = torch.randn(N, 4, generator=g)
        ns.root_states[:, 3:7] = q / q.norm(dim=1, keepdim=True)
        ns.base_quat = ns.root_states[:, 3:7]
        h = LRV.LeggedRobot._get_heights(ns, torch.arange(N), Cfg)
    finally:
        Cfg.terrain.mesh_type = saved
    fx["heights/height_points"] = ns.height_points[0].numpy().copy()
    fx["heights/height_samples"] = ns.height_samples.numpy().copy()
    fx["heights/root_states"] = ns.root_states.numpy().copy()
    fx["heights/scales"] = np.asarray([2.5, 0.1, 0.005])
    fx["heights/measured_heights"] = h.numpy().copy()
    fx["heights/body_height_mean"] = torch.mean(ns.root_states[:, 2].unsqueeze(1) - h, dim=1).numpy().copy()      # LRV:164


def main(num_envs=64, steps=10):
    H.install()
    if not hasattr(np, "int"):
        np.int = int                                    # legged_robot_velocity_tracking.py:1380
    from isaacgym import gymapi
    from legged_tracking_b200.synthetic import SyntheticState
    Cfg = build_cfg(num_envs)
    import random
    random.seed(11); np.random.seed(11); torch.manual_seed(11)
    gym = gymapi.acquire_gym()
    gym.set_replay(None)
    from go1_gym.envs.go1.velocity_tracking import VelocityTrackingEasyEnv
    from go1_gym.envs.wrappers.history_wrapper import HistoryWrapper
    with contextlib.redirect_stdout(io.StringIO()):
        env = VelocityTrackingEasyEnv(sim_device="cpu", headless=True, cfg=Cfg)
    wrapped = HistoryWrapper(env)
    N = env.num_envs
    synth = SyntheticState(N, env.env_origins.numpy(), num_bodies=gym.num_bodies_per_env, seed=1000, feet_indices=[int(i) for i in env.feet_indices])
    replay = H.Replay(synth, decimation=Cfg.control.decimation, num_actor=1)
    gym.set_replay(replay)
    fx = {}
    # ---- static configuration the product needs (what _parse_cfg / _init_buffers / _prepare_reward_function derived)
    c = {"num_envs": N, "num_bodies": int(gym.num_bodies_per_env), "num_commands": int(Cfg.commands.num_commands), "num_obs": int(env.num_obs),
         "num_privileged_obs": int(env.num_privileged_obs), "obs_history_length": int(Cfg.env.num_observation_history),
         "dt": float(env.dt), "max_episode_length": float(env.max_episode_length), "decimation": int(Cfg.control.decimation),
         "terminal_body_height": float(Cfg.rewards.terminal_body_height), "use_terminal_body_height": bool(Cfg.rewards.use_terminal_body_height),
         "resampling_time": float(Cfg.commands.resampling_time), "rand_interval": float(Cfg.domain_rand.rand_interval),
         "gravity_rand_interval": float(Cfg.domain_rand.gravity_rand_interval), "gravity_rand_duration": float(Cfg.domain_rand.gravity_rand_duration),
         "gravity_range": [float(x) for x in Cfg.domain_rand.gravity_range], "randomize_gravity": bool(Cfg.domain_rand.randomize_gravity),
         "only_positive_rewards": bool(Cfg.rewards.only_positive_rewards), "only_positive_rewards_ji22_style": bool(Cfg.rewards.only_positive_rewards_ji22_style),
         "sigma_rew_neg": float(Cfg.rewards.sigma_rew_neg), "clip_observations": float(Cfg.normalization.clip_observations),
         "clip_actions": float(Cfg.normalization.clip_actions), "obs_scale_dof_pos": float(env.obs_scales.dof_pos), "obs_scale_dof_vel": float(env.obs_scales.dof_vel),
         "add_noise": bool(env.add_noise), "observe_command": bool(Cfg.env.observe_command), "observe_two_prev_actions": bool(Cfg.env.observe_two_prev_actions),
         "observe_timing_parameter": bool(Cfg.env.observe_timing_parameter), "observe_clock_inputs": bool(Cfg.env.observe_clock_inputs),
         "priv_observe_friction": bool(Cfg.env.priv_observe_friction), "priv_observe_restitution": bool(Cfg.env.priv_observe_restitution),
         "friction_range": [float(x) for x in Cfg.normalization.friction_range], "restitution_range": [float(x) for x in Cfg.normalization.restitution_range],
         "kappa_gait_probs": float(Cfg.rewards.kappa_gait_probs), "pacing_offset": bool(Cfg.commands.pacing_offset),
         "gaitwise_curricula": bool(Cfg.commands.gaitwise_curricula), "exclusive_phase_offset": bool(Cfg.commands.exclusive_phase_offset),
         "balance_gait_distribution": bool(Cfg.commands.balance_gait_distribution), "binary_phases": bool(Cfg.commands.binary_phases),
         "randomize_motor_strength": bool(Cfg.domain_rand.randomize_motor_strength), "motor_strength_range": [float(x) for x in Cfg.domain_rand.motor_strength_range],
         "randomize_motor_offset": bool(Cfg.domain_rand.randomize_motor_offset), "motor_offset_range": [float(x) for x in Cfg.domain_rand.motor_offset_range],
         "x_init_range": float(Cfg.terrain.x_init_range), "y_init_range": float(Cfg.terrain.y_init_range), "yaw_init_range": float(Cfg.terrain.yaw_init_range),
         "x_init_offset": float(Cfg.terrain.x_init_offset), "y_init_offset": float(Cfg.terrain.y_init_offset), "custom_origins": bool(env.custom_origins),
         "action_scale": float(Cfg.control.action_scale), "hip_scale_reduction": float(Cfg.control.hip_scale_reduction),
         "control_type": str(Cfg.control.control_type),
         "lag_timesteps": int(Cfg.domain_rand.lag_timesteps), "randomize_lag_timesteps": bool(Cfg.domain_rand.randomize_lag_timesteps)}
    for k, v in c.items():
        fx["cfg/" + k] = np.asarray(v)
    for k in ("tracking_sigma", "tracking_sigma_yaw", "base_height_target", "gait_force_sigma", "gait_vel_sigma", "max_contact_force"):
        fx["cfg/rewards/" + k] = np.float64(getattr(Cfg.rewards, k))
    cur0 = env.curricula[0]
    fx["cfg/curriculum_limits"] = np.array([[cur0.lows[i], cur0.highs[i], cur0.ls[k]] for i, k in enumerate(cur0.keys)], np.float64)    # LRV:1328-1374
    cc = Cfg.commands
    fx["cfg/curriculum_init"] = np.array([cc.lin_vel_x, cc.lin_vel_y, cc.ang_vel_yaw, cc.body_height_cmd, cc.gait_frequency_cmd_range,
                                          cc.gait_phase_cmd_range, cc.gait_offset_cmd_range, cc.gait_bound_cmd_range, cc.gait_duration_cmd_range,
                                          cc.footswing_height_range, cc.body_pitch_range, cc.body_roll_range, cc.stance_width_range,
                                          cc.stance_length_range, cc.aux_reward_coef_range], np.float64)                              # LRV:1382-1400
    fx["cfg/curriculum_thresholds"] = np.array([Cfg.curriculum_thresholds.tracking_lin_vel, Cfg.curriculum_thresholds.tracking_ang_vel,
                                                Cfg.curriculum_thresholds.tracking_contacts_shaped_force,
                                                Cfg.curriculum_thresholds.tracking_contacts_shaped_vel], np.float64)
    fx["cfg/reward_names"] = np.array(env.reward_names)
    fx["cfg/reward_scales"] = np.array([float(env.reward_scales[n]) for n in env.reward_names])
    fx["cfg/category_names"] = np.array(env.category_names)
    fx["cfg/feet_indices"] = env.feet_indices.detach().numpy().copy()
    fx["cfg/penalised_contact_indices"] = env.penalised_contact_indices.detach().numpy().copy()
    fx["cfg/termination_contact_indices"] = env.termination_contact_indices.detach().numpy().copy()
    fx["cfg/commands_scale"] = env.commands_scale.detach().numpy().copy()
    fx["cfg/default_dof_pos"] = env.default_dof_pos.detach().numpy().copy()
    fx["cfg/dof_pos_limits"] = env.dof_pos_limits.detach().numpy().copy()
    fx["cfg/noise_scale_vec"] = env.noise_scale_vec.detach().numpy().copy()
    fx["cfg/torque_limits"] = env.torque_limits.detach().numpy().copy()
    fx["cfg/base_init_state"] = env.base_init_state.detach().numpy().copy()
    fx["cfg/env_origins"] = env.env_origins.detach().numpy().copy()
    # ---- the command curriculum as the reference built it (curriculum.py:27-60, LRV:1317-1400), and two of its update() calls on
    # seeded inputs (curriculum.py:133-152) for the device-side restatement
    import copy
    cur = copy.deepcopy(env.curricula[0])
    fx["curriculum/grid"] = cur.grid.copy()                                  # [15, n_bins] bin centres
    fx["curriculum/bin_sizes"] = np.array([*cur.bin_sizes.values()])
    fx["curriculum/weights0"] = cur.weights.copy()
    fx["curriculum/local_range"] = np.array([0.55, 0.55, 0.55, 0.55, 0.35, 0.25, 0.25, 0.25, 0.25, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0])    # LRV:755-757
    gc = torch.Generator().manual_seed(21)
    for call in range(2):
        nb = 40
        active = np.nonzero(cur.weights > 0)[0]
        bins = active[torch.randint(0, len(active), (nb,), generator=gc).numpy()]
        rewards = [torch.rand(nb, generator=gc) for _ in range(4)]
        thresholds = [0.35, 0.3, 0.4, 0.25]
        cur.update(bins, rewards, thresholds, local_range=fx["curriculum/local_range"])
        fx[f"curriculum/update{call}/bins"] = bins.copy()
        fx[f"curriculum/update{call}/rewards"] = torch.stack(rewards).numpy().copy()
        fx[f"curriculum/update{call}/thresholds"] = np.array(thresholds)
        fx[f"curriculum/update{call}/weights"] = cur.weights.copy()
    # ---- episode phases: timeouts (max_episode_length) and command resampling (every int(resampling_time / dt) steps) inside the recording
    g = torch.Generator().manual_seed(4)
    sample_interval = int(Cfg.commands.resampling_time / env.dt)
    ep = torch.randint(20, 400, (N,), generator=g)
    ep[::5] = sample_interval - 1 - torch.randint(0, steps, (len(ep[::5]),), generator=g)
    ep[1::7] = int(env.max_episode_length) - torch.randint(0, steps, (len(ep[1::7]),), generator=g)
    env.episode_length_buf[:] = ep
    env.common_step_counter = int(Cfg.domain_rand.gravity_rand_interval) - 3          # _randomize_gravity fires at the third step
    dump_state(env, fx, "init/")
    fx["init/obs_history"] = wrapped.obs_history.detach().numpy().copy()
    samples = []
    for cur in env.curricula:
        def wrapped_sample(batch_size, _orig=cur.sample, **kw):
            out = _orig(batch_size=batch_size, **kw)
            samples.append((np.asarray(out[0]).copy(), np.asarray(out[1]).copy()))
            return out
        cur.sample = wrapped_sample
    torques = []
    _ct = env._compute_torques

    def compute_torques(actions):
        t = _ct(actions)
        torques.append(t.detach().numpy().copy())
        return t
    env._compute_torques = compute_torques
    ga = torch.Generator().manual_seed(8)
    with H.RngRecorder() as rec:
        # the synthetic simulator draws with torch.rand too: keep its draws out of the log of the REFERENCE's draws
        for name in ("on_simulate", "on_refresh_root"):
            def quiet(gym_, _orig=getattr(replay, name)):
                was, rec.active = rec.active, False
                try:
                    return _orig(gym_)
                finally:
                    rec.active = was
            setattr(replay, name, quiet)
        for k in range(steps):
            tag = f"step{k}/"
            actions = torch.randn(N, 12, generator=ga) * 0.4
            samples.clear(); torques.clear()
            fx[tag + "in/actions"] = actions.detach().numpy().copy()
            for s in range(Cfg.control.decimation):
                fx[tag + f"in/dof_state/{s}"] = synth.dof_state(k, s).detach().numpy().copy()
            root = synth.root(k)
            fx[tag + "in/root"] = root.detach().numpy().copy()
            fx[tag + "in/contact"] = synth.contact(k).detach().numpy().copy()
            fx[tag + "in/rigid"] = synth.rigid(k, root).detach().numpy().copy()
            rec.active = True
            with contextlib.redirect_stdout(io.StringIO()):
                obs_dict, rew, done, info = wrapped.step(actions)
            rec.active = False
            for j, (kind, t) in enumerate(rec.take()):
                fx[tag + f"rng/{j:02d}_{kind}"] = t.detach().numpy().copy()
            for j, (cmd, bins) in enumerate(samples):
                fx[tag + f"curriculum/{j}/commands"] = cmd
                fx[tag + f"curriculum/{j}/bins"] = bins
            fx[tag + "out/all_torques"] = np.stack(torques)
            fx[tag + "out/obs"] = obs_dict["obs"].detach().numpy().copy()
            fx[tag + "out/privileged_obs"] = obs_dict["privileged_obs"].detach().numpy().copy()
            fx[tag + "out/obs_history"] = obs_dict["obs_history"].detach().numpy().copy()
            fx[tag + "out/rew"] = rew.detach().numpy().copy()
            fx[tag + "out/done"] = done.detach().numpy().copy()
            dump_state(env, fx, tag + "state/")
    fx["meta/steps"] = np.int64(steps)
    record_heights(fx, Cfg)
    np.savez_compressed(OUT, **fx)
    n_reset = sum(int(fx[f"step{k}/out/done"].sum()) for k in range(steps))
    n_samp = sum(1 for k in fx if k.endswith("/commands") and "/curriculum/" in k)
    print("wrote", OUT, len(fx), "arrays;", n_reset, "resets,", n_samp, "curriculum samples;", os.path.getsize(OUT) // 1024, "KiB")


if __name__ == "__main__":
    main()
