"""Generate the golden fixtures in tests/golden/ by running the UNMODIFIED reference on CPU.

Run in the build container only (needs /root/reference):   python oracle/gen_golden.py
Each fixture `tests/golden/<name>.npz` holds
  flat            JSON of the flat config read off the live reference env (ref_harness.flat_from_reference_env)
  terrain/*       tile table + per-env origins
  init/*          full env state after reset() (what OraclePort.load_state / the CUDA env consume)
  step<k>/in/*    what PhysX would have produced (synthetic), the clipped-or-not raw actions, and the reference's
                  own random draws scattered into the per-env slot layout of legged_tracking_b200/rng_slots.py
  step<k>/out/*   observations, rewards, resets, height indices, torques of all substeps, full state
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_harness as H                     # noqa: E402
from legged_tracking_b200 import rng_slots as RS        # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(HERE), "tests", "golden")

F32_ATTRS = ["actions", "last_actions", "last_last_actions", "joint_pos_target", "last_joint_pos_target",
             "last_last_joint_pos_target", "last_dof_vel", "last_root_vel", "torques", "joint_pos_err_last",
             "joint_pos_err_last_last", "joint_vel_last", "joint_vel_last_last", "motor_strengths", "motor_offsets",
             "Kp_factors", "Kd_factors", "com_displacements", "gravities", "base_lin_vel", "base_ang_vel",
             "projected_gravity", "base_rotation", "relative_linear", "relative_rotation", "local_relative_linear",
             "local_relative_rotation", "target_poses", "feet_air_time", "rew_buf", "rew_buf_pos", "rew_buf_neg",
             "trajectories", "root_states", "dof_state"]
I64_ATTRS = ["episode_length_buf", "curr_pose_index", "collision_count", "collision_env_buf"]
BOOL_ATTRS = ["reset_buf", "time_out_buf", "body_height_buf", "switched_buf", "reached_buf", "plan_buf",
              "reached_env_buf", "last_contacts"]


def snapshot(env, wrapped, prefix, out):
    """Copy the reference env's state into out[prefix + name] using the port's state names."""
    n = env.num_envs
    for a in F32_ATTRS:
        if not hasattr(env, a) and a.startswith("joint_"):      # actuator-net histories only exist for control_type actuator_net (LR:1307-1327)
            out[prefix + a] = np.zeros((n, env.num_dof), np.float32)
            continue
        out[prefix + a] = getattr(env, a).detach().float().numpy().copy()
    out[prefix + "friction_coeffs"] = env.friction_coeffs[:, :1].numpy().copy()
    out[prefix + "restitutions"] = env.restitutions[:, :1].numpy().copy()
    out[prefix + "payloads"] = env.payloads.view(n, 1).numpy().copy()
    for a in I64_ATTRS:
        out[prefix + a] = getattr(env, a).numpy().astype(np.int64).copy()
    for a in BOOL_ATTRS:
        v = getattr(env, a, None)
        out[prefix + a] = (np.zeros(n, bool) if v is None else v.numpy().astype(bool).copy())
    to = env.extras.get("time_outs", None)
    out[prefix + "time_outs_latched"] = np.zeros(n, bool) if to is None else to.numpy().astype(bool).copy()
    cpa = getattr(env, "camera_pitch_angle", None)
    out[prefix + "camera_pitch_angle"] = np.zeros(n, np.float32) if cpa is None else cpa.numpy().copy()
    out[prefix + "lag_buffer"] = torch.stack(env.lag_buffer, 0).numpy().copy()
    out[prefix + "gravity_vec"] = env.gravity_vec[0].numpy().copy()
    out[prefix + "obs_history"] = wrapped.obs_history.numpy().copy()
    out[prefix + "common_step_counter"] = np.int64(env.common_step_counter)
    if getattr(env.cfg.commands, "sampling_based_planning", False):
        out[prefix + "plan_length_buf"] = env.plan_length_buf.numpy().astype(np.int64).copy()
        out[prefix + "local_target_poses"] = env.local_target_poses.detach().float().numpy().copy()
    for k, v in env.episode_sums.items():
        out[prefix + "episode_sums/" + k] = v.numpy().copy()
    for k, v in env.command_sums.items():
        out[prefix + "command_sums/" + k] = v.numpy().copy()


class TaggedRng:
    """Records the reference's uniform draws per call site by wrapping *bound methods on the instance*
    (the reference source is untouched) and torch.rand / rand_like."""

    def __init__(self, env):
        self.env = env
        self.ctx = []
        self.in_reset = False
        self.log = []                       # (tag, env_ids or None, [tensors])
        self._rand, self._rand_like = torch.rand, torch.rand_like
        for name in ("_randomize_dof_props", "_reset_dofs", "_reset_root_states"):
            self._wrap(name, takes_cfg=True)
        self._wrap("trajectory_function", takes_cfg=False)
        self._wrap_gravity()
        orig_reset = env.reset_idx

        def reset_idx(env_ids):
            self.in_reset = True
            try:
                return orig_reset(env_ids)
            finally:
                self.in_reset = False
        env.reset_idx = reset_idx

    def _wrap(self, name, takes_cfg):
        orig = getattr(self.env, name)

        def wrapped(env_ids, *a):
            entry = [("reset/" if self.in_reset else "interval/") + name, env_ids.clone(), []]
            self.log.append(entry)
            self.ctx.append(entry)
            try:
                return orig(env_ids, *a)
            finally:
                self.ctx.pop()
        setattr(self.env, name, wrapped)

    def _wrap_gravity(self):
        orig = self.env._randomize_gravity

        def wrapped(external_force=None):
            entry = ["gravity", None, []]
            self.log.append(entry)
            self.ctx.append(entry)
            try:
                return orig(external_force)
            finally:
                self.ctx.pop()
        self.env._randomize_gravity = wrapped

    def __enter__(self):
        def rand(*a, **k):
            out = self._rand(*a, **k)
            if k.get("generator") is not None:
                return out                      # the harness's own seeded draws (SyntheticState), not the reference's
            (self.ctx[-1][2] if self.ctx else self._orphan("rand")).append(out.detach().clone())
            return out

        def rand_like(*a, **k):
            out = self._rand_like(*a, **k)
            (self.ctx[-1][2] if self.ctx else self._orphan("rand_like")).append(out.detach().clone())
            return out
        torch.rand, torch.rand_like = rand, rand_like
        return self

    def _orphan(self, kind):
        entry = ["orphan/" + kind, None, []]
        self.log.append(entry)
        return entry[2]

    def __exit__(self, *exc):
        torch.rand, torch.rand_like = self._rand, self._rand_like

    def take(self):
        log, self.log = self.log, []
        return log


def scatter_draws(log, flat, n):
    """Reference call log of one step -> (uniforms[n, n_slots], noise[n, num_obs] or None, gravity[3] or None)."""
    S = RS.n_slots(flat)
    U = np.full((n, S), 0.5, np.float32)
    noise, gravity = None, None
    for tag, env_ids, draws in log:
        ids = None if env_ids is None else env_ids.numpy()
        d = [t.numpy() for t in draws]
        if tag.endswith("_randomize_dof_props"):
            base = RS.RESET_STRENGTH if tag.startswith("reset/") else RS.INTERVAL_STRENGTH
            it = iter(d)
            if flat["randomize_motor_strength"]:
                U[ids, base] = next(it)
            if flat["randomize_motor_offset"]:
                U[ids, base + 1: base + 13] = next(it)
            if flat["randomize_Kp_factor"]:
                U[ids, base + 13] = next(it)
            if flat["randomize_Kd_factor"]:
                U[ids, base + 14] = next(it)
        elif tag.endswith("_reset_dofs"):
            U[ids, RS.RESET_DOF:RS.RESET_DOF + 12] = d[0]
        elif tag.endswith("_reset_root_states"):
            it = iter(d)
            if flat["custom_origins"]:
                U[ids, RS.RESET_X] = next(it)[:, 0]
                U[ids, RS.RESET_Y] = next(it)[:, 0]
            U[ids, RS.RESET_YAW] = next(it)[:, 0]
            U[ids, RS.RESET_VEL:RS.RESET_VEL + 6] = next(it)
        elif tag.endswith("trajectory_function"):
            if flat["traj_function"] == "random_goal":
                for j, t in enumerate(d):
                    U[ids, RS.TRAJ + j] = t[:, 0]
            elif flat["traj_function"] == "random_target":
                nt = flat["traj_length"] // flat["num_interpolation"] + 1
                for j, t in enumerate(d):
                    U[ids, RS.TRAJ + j * nt: RS.TRAJ + (j + 1) * nt] = t
            else:
                assert not d
        elif tag == "gravity":
            if d:
                gravity = d[0].astype(np.float32)
        elif tag == "orphan/rand_like":
            assert noise is None and len(d) == 1
            noise = d[0].astype(np.float32)
        else:
            raise RuntimeError(f"unexpected random draw at {tag}: {[t.shape for t in d]}")
    return U, noise, gravity


def terrain_from_reference_env(env):
    """Tile table + per-env origins exactly as the reference derives them (LR:1808-1847)."""
    out = {"env_origins": env.env_origins.numpy().astype(np.float32)}
    if env.cfg.terrain.mesh_type != "plane":
        t = env.terrain
        rows, cols = env.cfg.terrain.num_rows, env.cfg.terrain.num_cols
        tiles = torch.from_numpy(t.height_samples_by_row_col * t.vertical_scale).to(torch.float)   # LR:1838-1840
        out["tile_heights"] = tiles.reshape(rows * cols, 2, tiles.shape[3], tiles.shape[4]).numpy()
        out["env_tile"] = (env.grid_r * cols + env.grid_c).numpy().astype(np.int32)
        out["env_terrain_origin"] = env.env_terrain_origin.numpy().astype(np.float32)
        # cross-check: the per-env copy equals table[env_tile]
        assert np.array_equal(out["tile_heights"][out["env_tile"]], env.env_height_samples.numpy())
    return out


def actuator_from_reference():
    """Weights of resources/actuator_nets/unitree_go1.pt (TorchScript Sequential, softsign)."""
    H.install()
    m = torch.jit.load(os.path.join(H.REF_ROOT, "resources", "actuator_nets", "unitree_go1.pt"), map_location="cpu")
    sd = m.state_dict()
    return {"w0": sd["0.weight"].numpy(), "b0": sd["0.bias"].numpy(), "w1": sd["2.weight"].numpy(),
            "b1": sd["2.bias"].numpy(), "w2": sd["4.weight"].numpy(), "b2": sd["4.bias"].numpy()}


def record_run(argv, num_envs, num_rows=None, num_cols=None, steps=6, overrides=None, synth_seed=1000,
               full_state_every=2, near_timeout_frac=0.3, start_counter=0, synth_kwargs=None, init_hook=None):
    """Run the reference for `steps` policy steps; return the fixture dict (numpy arrays + 'flat' JSON string)."""
    import isaacgym.gymapi as gymapi
    wrapped, env, replay, Cfg, args = H.make_env(argv, num_envs, num_rows, num_cols, synth_seed=synth_seed,
                                                 overrides=overrides, synth_kwargs=synth_kwargs)
    import go1_gym.envs.base.legged_robot_trajectory_tracking as LRmod
    rng = TaggedRng(env)
    hidx = {}
    orig_stack = LRmod.stack_heights

    def stack_heights(samples, px, py):
        hidx["px"], hidx["py"] = px.clone(), py.clone()
        return orig_stack(samples, px, py)
    LRmod.stack_heights = stack_heights
    all_torques = []
    orig_ct = env._compute_torques

    def compute_torques(actions):
        t = orig_ct(actions)
        all_torques.append(t.detach().clone())
        return t
    env._compute_torques = compute_torques
    fx = {}
    try:
        with rng, torch.no_grad():
            wrapped.reset()
            # replayed state setup: episode phases, some close to the time-out so that resets happen in-fixture
            mel = int(env.max_episode_length)
            g = torch.Generator().manual_seed(synth_seed + 17)
            ep = torch.randint(mel, (env.num_envs,), generator=g)
            near = torch.rand(env.num_envs, generator=g) < near_timeout_frac
            ep = torch.where(near, mel - torch.randint(steps + 2, (env.num_envs,), generator=g), ep)
            ri = int(Cfg.domain_rand.rand_interval)
            ep[1::16] = ri - 1 - (torch.arange(len(ep[1::16])) % 3)      # interval-DR (LR:822-824) fires in-fixture
            env.episode_length_buf[:] = ep
            env.common_step_counter = int(start_counter)
            if init_hook is not None:
                init_hook(env)
            rng.take()
            # start the interval-DR clock where it will fire for a few envs too
            flat = H.flat_from_reference_env(env)
            fx["flat"] = json.dumps(flat, sort_keys=True)
            fx["argv"] = json.dumps(list(argv))
            for k, v in terrain_from_reference_env(env).items():
                fx["terrain/" + k] = v
            snapshot(env, wrapped, "init/", fx)
            n = env.num_envs
            for k in range(steps):
                del all_torques[:]
                hidx.clear()
                s = replay.step
                raw_actions = replay.synth.actions(s)
                pre = f"step{k}/in/"
                fx[pre + "actions"] = raw_actions.numpy().copy()
                fx[pre + "dof_state"] = np.stack([replay.synth.dof_state(s, j).numpy() for j in range(flat["decimation"])])
                root = replay.synth.root(s)
                fx[pre + "root"] = root.numpy().copy()
                fx[pre + "contact"] = replay.synth.contact(s).numpy().copy()
                fx[pre + "rigid"] = replay.synth.rigid(s, root).numpy().copy()
                obs, rew, done, info = wrapped.step(raw_actions.clone())
                U, noise, gravity = scatter_draws(rng.take(), flat, n)
                fx[pre + "uniforms"] = U
                if noise is not None:
                    fx[pre + "noise_uniforms"] = noise
                if gravity is not None:
                    fx[pre + "gravity_uniform"] = gravity
                post = f"step{k}/out/"
                fx[post + "obs"] = obs["obs"].numpy().copy()
                fx[post + "privileged_obs"] = obs["privileged_obs"].numpy().copy()
                fx[post + "obs_history"] = obs["obs_history"].numpy().copy()
                fx[post + "rew"] = rew.numpy().copy()
                fx[post + "done"] = done.numpy().astype(bool).copy()
                fx[post + "reset_env_ids"] = done.nonzero().flatten().numpy().astype(np.int64)
                fx[post + "all_torques"] = torch.stack(all_torques, 0).numpy().copy()
                if "px" in hidx:
                    fx[post + "px"] = hidx["px"].numpy().astype(np.int64)
                    fx[post + "py"] = hidx["py"].numpy().astype(np.int64)
                if (k % full_state_every) == 0 or k == steps - 1:
                    snapshot(env, wrapped, post + "state/", fx)
    finally:
        LRmod.stack_heights = orig_stack
        gymapi.acquire_gym().set_replay(None)
    return fx


def _planning_init(env):
    """Start the planning recording where the replan rule fires (LR:861-865): every env believes it reached its local target, and
    the interval counters sit at 7..9, so that the first three steps run the candidate search for a third of the envs each."""
    env.plan_buf = torch.ones(env.num_envs, dtype=torch.long)
    env.plan_length_buf = 7 + (torch.arange(env.num_envs) % 3)


COMMON = ["--old_ppo", "--measure_front_half"]
FIXTURES = {
    # BASELINE config 1 (plane; camera_zero must stay off there); noise + DR as train.py defaults
    "cfg1_plane_256": dict(argv=COMMON + ["--strategy", "e2e", "--terrain", "plane"], num_envs=256, steps=4, full_state_every=3,
                           overrides={"env.camera_zero": False}),
    # BASELINE config 2 shape (single_path, camera_zero, e2e) on a 4x4 tile grid
    "cfg2_single_path_64": dict(argv=COMMON + ["--camera_zero", "--strategy", "e2e", "--terrain", "single_path",
                                               "--penalty_scaler", "1.0"], num_envs=64, num_rows=4, num_cols=4, steps=8,
                                # gravity DR: counter 397 -> impulse ends (LR:828-830), 401 -> new gravity (LR:826-827)
                                start_counter=396),
    # same, deterministic suite (noise and all randomize_* off)
    "cfg2_nonoise_64": dict(argv=COMMON + ["--camera_zero", "--strategy", "e2e", "--terrain", "single_path",
                                           "--no_domain_rand"], num_envs=64, num_rows=4, num_cols=4, steps=6,
                            overrides={"noise.add_noise": False}),
    # BASELINE config 3 shape (vel strategy, terminal body height 0.25 -> many resets)
    "cfg3_vel_64": dict(argv=COMMON + ["--camera_zero", "--strategy", "vel", "--terrain", "single_path",
                                       "--terminal_body_height", "0.25"], num_envs=64, num_rows=4, num_cols=4, steps=8),
    # update_curriculum (LR:171-183): exploration scales start decaying after 3 global steps
    "cfg2_curriculum_32": dict(argv=COMMON + ["--camera_zero", "--strategy", "e2e", "--terrain", "single_path",
                                              "--exploration_steps", "3"], num_envs=32, num_rows=4, num_cols=4, steps=7),
    # option coverage: random_target trajectories, timestep in obs, positive-only reward, 3-step history,
    # prod-form exploration reward, terminate at end of trajectory, full (non front-half) grid
    "opts_random_target_32": dict(argv=["--old_ppo", "--camera_zero", "--strategy", "e2e", "--terrain", "single_path",
                                        "--random_target", "--timestep_in_obs", "--only_positive", "--num_history", "3",
                                        "--lin_vel_form", "prod", "--terminate_after_reach", "--r_orientation", "1.0",
                                        "--r_large_vel", "1.0"],
                                  num_envs=32, num_rows=4, num_cols=4, steps=6,
                                  overrides={"commands.switch_dist": 0.6, "terrain.x_init_range": 0.2,
                                             "terrain.y_init_range": 0.1, "terrain.yaw_init_range": 0.5,
                                             "env.observe_two_prev_actions": True,
                                             "env.num_observations": 504 + 12, "env.num_scalar_observations": 504 + 12}),
    # the OTHER shipped reward container (trajectory_tracking_reward.py:17-171; selected by
    # runs/trajectory_tracking/run-20230904_112307-rhi1my71/config.yaml): every term of it that the fused kernel claims,
    # with the shipped run's scales where it has them; thresholds chosen so that both branches of each gate fire on the
    # synthetic state (T_reach below / above the episode phases, small_vel / large_dist in the middle of the ranges)
    "ttr_container_32": dict(argv=COMMON + ["--camera_zero", "--strategy", "e2e", "--terrain", "single_path"],
                             num_envs=32, num_rows=4, num_cols=4, steps=6,
                             overrides={"rewards.reward_container_name": "TrajectoryTrackingRewards",
                                        "rewards.T_reach": 200, "rewards.small_vel_threshold": 0.4,
                                        "rewards.large_dist_threshold": 1.5, "rewards.target_ang_vel": np.pi / 2.0,
                                        "rewards.tracking_sigma_ang": 0.5, "commands.switch_dist": 0.6,
                                        "reward_scales.e2e": 0.0, "reward_scales.exploration_lin": 0.0,
                                        "reward_scales.exploration_yaw": 0.0, "reward_scales.base_height": 0.0,
                                        "reward_scales.large_vel": 0.0,
                                        "reward_scales.torques": -2e-5, "reward_scales.dof_acc": -5e-7,
                                        "reward_scales.collision": -1.0, "reward_scales.action_rate": -0.01,
                                        "reward_scales.dof_pos_limits": -20.0, "reward_scales.orientation": -0.5,
                                        "reward_scales.dof_vel": -1e-4, "reward_scales.dof_pos": -0.01,
                                        "reward_scales.task_old": 1.0, "reward_scales.reach_goal": 200.0,
                                        "reward_scales.reach_goal_t": 0.1, "reward_scales.reach_goal_T": 5.0,
                                        "reward_scales.task": 1.0, "reward_scales.exploration": 1.0,
                                        "reward_scales.stalling": 1.0, "reward_scales.linear_vel": -1.0,
                                        "reward_scales.lin_vel_z": -2.0, "reward_scales.ang_vel_xy": -0.05,
                                        "reward_scales.feet_air_time": 1.0, "reward_scales.survive": 0.1,
                                        "reward_scales.reaching_linear_vel": 0.6, "reward_scales.reaching_z": -0.5,
                                        "reward_scales.reaching_roll": -0.5,
                                        "reward_scales.reaching_pitch": -0.5, "reward_scales.reaching_yaw_abs": -0.3,
                                        "reward_scales.reaching_yaw": 0.3}),
    # same container with the l1 form of reaching_linear_vel (trajectory_tracking_reward.py:146-147) and the
    # ji22 positive/negative combination (LR:343-344)
    "ttr_l1_ji22_32": dict(argv=COMMON + ["--camera_zero", "--strategy", "e2e", "--terrain", "single_path",
                                          "--lin_vel_form", "l1"],
                           num_envs=32, num_rows=4, num_cols=4, steps=4,
                           overrides={"rewards.reward_container_name": "TrajectoryTrackingRewards",
                                      "rewards.only_positive_rewards": False,
                                      "rewards.only_positive_rewards_ji22_style": True, "rewards.sigma_rew_neg": 5.0,
                                      "reward_scales.e2e": 0.0, "reward_scales.exploration_lin": 0.0,
                                      "reward_scales.exploration_yaw": 0.0, "reward_scales.base_height": 0.0,
                                      "reward_scales.large_vel": 0.0, "reward_scales.reaching_linear_vel": -0.6,
                                      "reward_scales.exploration": 1.0, "reward_scales.survive": 0.5,
                                      "reward_scales.feet_air_time": 1.0, "reward_scales.stalling": 1.0,
                                      "rewards.small_vel_threshold": 0.4, "rewards.large_dist_threshold": 1.5}),
    # option coverage 2: random_goal trajectories (trajectory_function.py:28-40), the PD controller instead of the
    # actuator net (LR:989-991) with Kp/Kd factor randomisation, termination on body rotation (LR:215-216), the ji22
    # reward combination on the crawling container
    # sampling-based local planner inside the env step (LR:856-916): every env plans at its first step and on the
    # 10-step interval; the tunnel's hanging / standing wedges invalidate part of the 1,575 candidate poses
    "planning_32": dict(argv=COMMON + ["--camera_zero", "--strategy", "e2e", "--terrain", "single_path", "--penalty_scaler", "1.0"],
                        num_envs=32, num_rows=4, num_cols=4, steps=12, full_state_every=1, near_timeout_frac=0.7,
                        overrides={"commands.sampling_based_planning": True}, init_hook=_planning_init),
    # the same with the TrajectoryTrackingRewards container and its planning-only term reaching_local_goal = plan_buf & replan
    # (trajectory_tracking_reward.py:95-97; scale as in the archived hierarchical training scripts)
    "planning_ttr_32": dict(argv=COMMON + ["--camera_zero", "--strategy", "e2e", "--terrain", "single_path", "--penalty_scaler", "1.0"],
                            num_envs=32, num_rows=4, num_cols=4, steps=12, full_state_every=1, near_timeout_frac=0.7,
                            overrides={"commands.sampling_based_planning": True,
                                       "rewards.reward_container_name": "TrajectoryTrackingRewards",
                                       "reward_scales.e2e": 0.0, "reward_scales.exploration_lin": 0.0,
                                       "reward_scales.exploration_yaw": 0.0, "reward_scales.base_height": 0.0,
                                       "reward_scales.large_vel": 0.0, "reward_scales.reaching_local_goal": 100.0,
                                       "reward_scales.task": 1.0, "reward_scales.stalling": 1.0,
                                       "rewards.small_vel_threshold": 0.4, "rewards.large_dist_threshold": 1.5},
                            init_hook=_planning_init),
    "opts_random_goal_P_32": dict(argv=COMMON + ["--camera_zero", "--strategy", "e2e", "--terrain", "single_path"],
                                  num_envs=32, num_rows=4, num_cols=4, steps=6, synth_kwargs={"rp_range": 1.9},
                                  overrides={"commands.traj_function": "random_goal", "commands.x_mean": 1.5,
                                             "commands.x_range": 1.0, "commands.y_mean": 0.0, "commands.y_range": 0.6,
                                             "commands.yaw_range": 0.5, "control.control_type": "P",
                                             "domain_rand.randomize_Kp_factor": True,
                                             "domain_rand.randomize_Kd_factor": True,
                                             "env.use_terminal_body_rotation": True,
                                             "rewards.only_positive_rewards": False,
                                             "rewards.only_positive_rewards_ji22_style": True}),
}


def main(names=None):
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    act = actuator_from_reference()
    np.savez_compressed(os.path.join(GOLDEN_DIR, "actuator_net_go1.npz"), **act)
    for name, spec in FIXTURES.items():
        if names and name not in names:
            continue
        fx = record_run(**spec)
        path = os.path.join(GOLDEN_DIR, name + ".npz")
        np.savez_compressed(path, **fx)
        print(f"\n{name}: {len(fx)} arrays, {os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main(sys.argv[1:])
