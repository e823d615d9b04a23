"""Run the UNMODIFIED reference (/root/reference) on CPU under the shims -- test infrastructure only.

This file can only work in the build container, where /root/reference exists.  It is used by
oracle/gen_golden.py to produce the committed fixtures under tests/golden/ and by
tests/test_oracle_vs_reference.py (skipped when the reference is absent).  The product never
imports it.

What it does
  * puts oracle/shims (fake isaacgym / params_proto / gym / matplotlib) and the reference on sys.path;
  * builds the reference's `Cfg` by exec'ing the reference's own argparse block and config-mutation
    block of scripts/train.py (read at run time, lines located by marker text: scripts/train.py:46-241
    and :287-340), so the configuration is the reference's, not a transcription;
  * replaces PhysX by `Replay`, which writes legged_tracking_b200.synthetic.SyntheticState tensors into
    the fake gym's sim tensors (dof_state after every gym.simulate(); root/contact/rigid-body state at
    gym.refresh_actor_root_state_tensor());
  * records every torch.rand / torch.rand_like draw the reference makes, so the same uniforms can be
    injected into the CUDA path (the reference draws on a *compacted* env list; see scatter helpers).
"""
import importlib
import os
import sys
import textwrap
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF_ROOT = os.environ.get("LT_REFERENCE_ROOT", "/root/reference")


def reference_available():
    return os.path.isdir(os.path.join(REF_ROOT, "go1_gym"))


_installed = False


def install():
    """Idempotent: path setup + the monkeypatches SURVEY.md section 8c item 6 lists."""
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError(f"reference not found at {REF_ROOT}")
    for p in (REF_ROOT, os.path.join(HERE, "shims"), REPO):
        if p not in sys.path:
            sys.path.insert(0, p)
    if "wandb" not in sys.modules:
        try:
            import wandb  # noqa: F401
        except Exception:
            sys.modules["wandb"] = types.ModuleType("wandb")
    # the actuator net was saved with CUDA tensors; the reference passes no map_location
    _orig_load = torch.jit.load
    if not getattr(torch.jit.load, "_lt_patched", False):
        def _load(f, map_location=None, *a, **k):
            return _orig_load(f, map_location="cpu" if not torch.cuda.is_available() else map_location, *a, **k)
        _load._lt_patched = True
        torch.jit.load = _load
    if not hasattr(np, "int"):
        np.int = int
    if not hasattr(np, "float"):
        np.float = float
    import go1_gym.utils.tunnel as tunnel
    tunnel.visual_elevation_map = lambda *a, **k: None
    _installed = True


def _train_py_blocks():
    src = open(os.path.join(REF_ROOT, "scripts", "train.py")).read().splitlines()

    def find(marker, start=0):
        for i in range(start, len(src)):
            if marker in src[i]:
                return i
        raise RuntimeError(f"marker {marker!r} not found in reference scripts/train.py")
    c0 = find("config_go1(Cfg)")
    c1 = find("Cfg.normalization.clip_actions")
    cfg_block = textwrap.dedent("\n".join(src[c0:c1 + 1]))
    a0 = find("parser = argparse.ArgumentParser()")
    a1 = find("args = parser.parse_args()", a0)
    arg_block = textwrap.dedent("\n".join(src[a0:a1]))
    return arg_block, cfg_block


def parse_reference_args(argv):
    """argv -> args namespace using the reference's own argparse definitions (scripts/train.py:287-340)."""
    install()
    import argparse
    from go1_gym import MINI_GYM_ROOT_DIR
    arg_block, _ = _train_py_blocks()
    ns = dict(argparse=argparse, os=os, MINI_GYM_ROOT_DIR=MINI_GYM_ROOT_DIR)
    exec(arg_block, ns)
    return ns["parser"].parse_args(list(argv))


def build_cfg(argv, num_envs, num_rows=None, num_cols=None, overrides=None):
    """Fresh reference Cfg configured exactly as scripts/train.py:46-241 does for `argv`."""
    install()
    import go1_gym.envs.base.legged_robot_trajectory_tracking_config as cfgmod
    importlib.reload(cfgmod)                      # Cfg is mutated in place: start from pristine class attrs
    import go1_gym.envs.go1.go1_crawling as crawl
    import go1_gym_learn.ppo_cse.actor_critic as acmod
    args = parse_reference_args(argv)
    _, cfg_block = _train_py_blocks()
    ns = dict(args=args, Cfg=cfgmod.Cfg, config_go1=crawl.config_go1, np=np, AC_Args=acmod.AC_Args)
    exec(cfg_block, ns)
    Cfg = cfgmod.Cfg
    Cfg.env.num_envs = int(num_envs)
    if num_rows is not None:
        Cfg.terrain.num_rows = int(num_rows)
    if num_cols is not None:
        Cfg.terrain.num_cols = int(num_cols)
    Cfg.env.record_video = False
    for path, val in (overrides or {}).items():
        obj = Cfg
        parts = path.split(".")
        for p in parts[:-1]:
            obj = getattr(obj, p)
        setattr(obj, parts[-1], val)
    return Cfg, args


class RngRecorder:
    """Wraps torch.rand / torch.rand_like: the values still come from torch's generator, but every
    call is logged as (tag, tensor) so the draws can be replayed into the CUDA path."""

    def __init__(self):
        self.calls = []
        self.active = False
        self._rand = torch.rand
        self._rand_like = torch.rand_like

    def __enter__(self):
        rec = self

        def rand(*a, **k):
            out = rec._rand(*a, **k)
            if rec.active:
                rec.calls.append(("rand", out.detach().clone()))
            return out

        def rand_like(*a, **k):
            out = rec._rand_like(*a, **k)
            if rec.active:
                rec.calls.append(("rand_like", out.detach().clone()))
            return out
        torch.rand = rand
        torch.rand_like = rand_like
        return self

    def __exit__(self, *exc):
        torch.rand = self._rand
        torch.rand_like = self._rand_like

    def take(self):
        c, self.calls = self.calls, []
        return c


class Replay:
    """PhysX stand-in: feeds SyntheticState into the fake gym's tensors at the points where the reference
    expects fresh simulator output (legged_robot_trajectory_tracking.py:85-88 and :119-121)."""

    def __init__(self, synth, decimation=4, num_actor=2):
        self.synth = synth
        self.decimation = decimation
        self.num_actor = num_actor
        self.step = 0          # index of the policy step being simulated
        self.substep = 0
        self.enabled = True

    def on_simulate(self, gym):
        if not self.enabled:
            return
        gym.tensors["dof"].copy_(self.synth.dof_state(self.step, self.substep))
        self.substep += 1

    def on_refresh_root(self, gym):
        if not self.enabled or self.substep == 0:
            return                                  # the init-time refresh in _init_buffers: nothing simulated yet
        n = self.synth.n
        root = self.synth.root(self.step)
        gym.tensors["root"].view(n, self.num_actor, 13)[:, 0, :] = root
        gym.tensors["contact"].view(n, -1, 3).copy_(self.synth.contact(self.step))
        gym.tensors["rigid"].view(n, -1, 13).copy_(self.synth.rigid(self.step, root))
        self.step += 1
        self.substep = 0


def make_env(argv, num_envs, num_rows=None, num_cols=None, seed=11, synth_seed=1000, overrides=None,
             history_wrapper=True, synth_kwargs=None):
    """Build the reference TrajectoryTrackingEnv on CPU, seeded like scripts/train.py:31-40."""
    install()
    import random
    from isaacgym import gymapi
    Cfg, args = build_cfg(argv, num_envs, num_rows, num_cols, overrides)
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    from go1_gym.envs.go1.trajectory_tracking import TrajectoryTrackingEnv
    from go1_gym.envs.wrappers.history_wrapper import HistoryWrapper
    from legged_tracking_b200.synthetic import SyntheticState
    gym = gymapi.acquire_gym()
    gym.set_replay(None)
    env = TrajectoryTrackingEnv(sim_device="cpu", headless=True, cfg=Cfg)
    synth = SyntheticState(env.num_envs, env.env_origins.numpy(), num_bodies=gym.num_bodies_per_env, seed=synth_seed,
                           **(synth_kwargs or {}))
    replay = Replay(synth, decimation=Cfg.control.decimation, num_actor=env.num_actor)
    gym.set_replay(replay)
    wrapped = HistoryWrapper(env) if history_wrapper else env
    return wrapped, env, replay, Cfg, args


def flat_from_reference_env(env):
    """Flat config (legged_tracking_b200.flatcfg keys) read off a LIVE reference env, i.e. after the
    reference's own _parse_cfg / _create_envs / _init_buffers / _prepare_reward_function ran."""
    cfg = env.cfg
    c = cfg
    g = lambda obj, name, default=None: getattr(obj, name, default)   # noqa: E731
    fl = lambda t: [float(v) for v in torch.as_tensor(t).detach().reshape(-1).double()]   # noqa: E731
    il = lambda t: [int(v) for v in torch.as_tensor(t).reshape(-1)]   # noqa: E731
    plane = c.terrain.mesh_type == "plane"
    flat = dict(
        num_envs=int(env.num_envs), num_actors=int(env.num_actor), num_bodies=int(env.contact_forces.shape[1]),
        num_dof=int(env.num_dof), num_actions=int(env.num_actions),
        num_obs=int(env.num_obs), num_privileged_obs=int(env.num_privileged_obs),
        obs_history_length=int(c.env.num_observation_history),
        dt=float(env.dt), sim_dt=float(env.sim_params.dt), decimation=int(c.control.decimation),
        clip_actions=float(c.normalization.clip_actions), clip_observations=float(c.normalization.clip_observations),
        control_type=str(c.control.control_type), action_scale=float(c.control.action_scale),
        hip_scale_reduction=float(c.control.hip_scale_reduction),
        randomize_lag_timesteps=bool(c.domain_rand.randomize_lag_timesteps), lag_timesteps=int(c.domain_rand.lag_timesteps),
        default_dof_pos=fl(env.default_dof_pos), torque_limits=fl(env.torque_limits),
        p_gains=fl(env.p_gains), d_gains=fl(env.d_gains),
        dof_pos_limits_lower=fl(env.dof_pos_limits[:, 0]), dof_pos_limits_upper=fl(env.dof_pos_limits[:, 1]),
        feet_indices=il(env.feet_indices), penalised_contact_indices=il(env.penalised_contact_indices),
        termination_contact_indices=il(env.termination_contact_indices),
        mesh_type="plane" if plane else "trimesh", observe_heights=bool(c.env.observe_heights),
        measured_points_x=fl(torch.tensor(np.asarray(c.terrain.measured_points_x), dtype=torch.float)),
        measured_points_y=fl(torch.tensor(np.asarray(c.terrain.measured_points_y), dtype=torch.float)),
        measure_front_half=bool(c.terrain.measure_front_half), camera_zero=bool(c.env.camera_zero),
        rotate_camera=bool(c.env.rotate_camera), camera_offset=fl(env.camera_offset),
        horizontal_scale=float(c.terrain.horizontal_scale),
        tile_rows=0 if plane else int(env.env_height_samples.shape[2]),
        tile_cols=0 if plane else int(env.env_height_samples.shape[3]),
        terrain_num_rows=int(c.terrain.num_rows), terrain_num_cols=int(c.terrain.num_cols),
        ceiling_height=float(c.terrain.ceiling_height),
        command_type=str(c.env.command_type), command_xy_only=bool(c.env.command_xy_only),
        traj_function=str(c.commands.traj_function), traj_length=int(c.commands.traj_length),
        num_interpolation=int(c.commands.num_interpolation),
        base_x=float(c.commands.base_x), base_y=float(c.commands.base_y), base_z=float(c.commands.base_z),
        base_roll=float(c.commands.base_roll), base_pitch=float(c.commands.base_pitch), base_yaw=float(c.commands.base_yaw),
        x_mean=float(c.commands.x_mean), x_range=float(c.commands.x_range), y_mean=float(c.commands.y_mean),
        y_range=float(c.commands.y_range), z_range=float(c.commands.z_range), roll_range=float(c.commands.roll_range),
        pitch_range=float(c.commands.pitch_range), yaw_range=float(c.commands.yaw_range),
        switch_upon_reach=bool(c.commands.switch_upon_reach), switch_dist=float(c.commands.switch_dist),
        switch_yaw=float(c.commands.switch_yaw), switch_interval=float(c.commands.switch_interval),
        sampling_based_planning=bool(c.commands.sampling_based_planning), global_reference=bool(c.commands.global_reference),
        plan_interval=int(c.commands.plan_interval),
        candidate_target_poses=[[float(v) for v in row] for row in np.asarray(c.commands.candidate_target_poses).reshape(-1, 6)]
        if c.commands.sampling_based_planning else [],
        max_episode_length=float(env.max_episode_length),
        use_terminal_body_height=bool(c.rewards.use_terminal_body_height),
        terminal_body_height=float(c.rewards.terminal_body_height),
        terminate_end_of_trajectory=bool(c.env.terminate_end_of_trajectory), T_reach=float(g(c.rewards, "T_reach", 0)),
        use_terminal_body_rotation=bool(c.env.use_terminal_body_rotation),
        reward_container_name=str(c.rewards.reward_container_name),
        reward_names=list(env.reward_names),
        reward_scales={k: float(v) for k, v in env.reward_scales.items()},
        reward_scales_cfg={k: float(v) for k, v in vars(c.reward_scales).items()},
        only_positive_rewards=bool(c.rewards.only_positive_rewards),
        only_positive_rewards_ji22_style=bool(c.rewards.only_positive_rewards_ji22_style),
        sigma_rew_neg=float(c.rewards.sigma_rew_neg),
        base_height_target=float(g(c.rewards, "base_height_target", 0.0)),
        tracking_sigma_lin=float(c.rewards.tracking_sigma_lin), tracking_sigma_ang=float(c.rewards.tracking_sigma_ang),
        target_lin_vel=float(c.rewards.target_lin_vel), target_ang_vel=float(c.rewards.target_ang_vel),
        lin_reaching_criterion=float(c.rewards.lin_reaching_criterion),
        ang_reaching_criterion=float(c.rewards.ang_reaching_criterion),
        lin_vel_form=str(g(c.rewards, "lin_vel_form", "exp")),
        small_vel_threshold=float(g(c.rewards, "small_vel_threshold", 0.0)),
        large_dist_threshold=float(g(c.rewards, "large_dist_threshold", 0.0)),
        exploration_steps=float(g(c.rewards, "exploration_steps", float("inf"))),
        soft_dof_pos_limit=float(g(c.rewards, "soft_dof_pos_limit", 1.0)),
        observe_command=bool(c.env.observe_command), timestep_in_obs=bool(g(c.env, "timestep_in_obs", False)),
        observe_two_prev_actions=bool(c.env.observe_two_prev_actions),
        observe_timing_parameter=bool(c.env.observe_timing_parameter), observe_clock_inputs=bool(c.env.observe_clock_inputs),
        observe_vel=bool(c.env.observe_vel), observe_only_ang_vel=bool(c.env.observe_only_ang_vel),
        observe_only_lin_vel=bool(c.env.observe_only_lin_vel), observe_yaw=bool(c.env.observe_yaw),
        observe_contact_states=bool(c.env.observe_contact_states),
        obs_scale_lin_vel=float(c.obs_scales.lin_vel), obs_scale_ang_vel=float(c.obs_scales.ang_vel),
        obs_scale_dof_pos=float(c.obs_scales.dof_pos), obs_scale_dof_vel=float(c.obs_scales.dof_vel),
        obs_scale_height=float(c.obs_scales.height_measurements),
        add_noise=bool(env.add_noise), noise_scale_vec=fl(env.noise_scale_vec),
        priv_observe_friction=bool(c.env.priv_observe_friction), priv_observe_restitution=bool(c.env.priv_observe_restitution),
        priv_observe_base_mass=bool(c.env.priv_observe_base_mass),
        priv_observe_com_displacement=bool(c.env.priv_observe_com_displacement),
        priv_observe_motor_strength=bool(c.env.priv_observe_motor_strength),
        priv_observe_motor_offset=bool(c.env.priv_observe_motor_offset),
        priv_observe_body_height=bool(c.env.priv_observe_body_height),
        priv_observe_body_velocity=bool(c.env.priv_observe_body_velocity),
        priv_observe_gravity=bool(c.env.priv_observe_gravity),
        priv_observe_ground_friction=bool(c.env.priv_observe_ground_friction),
        priv_observe_clock_inputs=bool(c.env.priv_observe_clock_inputs),
        priv_observe_desired_contact_states=bool(c.env.priv_observe_desired_contact_states),
        norm_friction_range=fl(c.normalization.friction_range), norm_restitution_range=fl(c.normalization.restitution_range),
        norm_added_mass_range=fl(c.normalization.added_mass_range),
        norm_com_displacement_range=fl(c.normalization.com_displacement_range),
        norm_motor_strength_range=fl(c.normalization.motor_strength_range),
        norm_motor_offset_range=fl(c.normalization.motor_offset_range),
        norm_body_height_range=fl(c.normalization.body_height_range),
        norm_body_velocity_range=fl(c.normalization.body_velocity_range),
        norm_gravity_range=fl(c.normalization.gravity_range),
        base_init_state=fl(env.base_init_state), custom_origins=bool(env.custom_origins),
        x_init_range=float(c.terrain.x_init_range), y_init_range=float(c.terrain.y_init_range),
        x_init_offset=float(c.terrain.x_init_offset), y_init_offset=float(c.terrain.y_init_offset),
        yaw_init_range=float(c.terrain.yaw_init_range),
        randomize_motor_strength=bool(c.domain_rand.randomize_motor_strength),
        motor_strength_range=fl(c.domain_rand.motor_strength_range),
        randomize_motor_offset=bool(c.domain_rand.randomize_motor_offset),
        motor_offset_range=fl(c.domain_rand.motor_offset_range),
        randomize_Kp_factor=bool(c.domain_rand.randomize_Kp_factor), Kp_factor_range=fl(c.domain_rand.Kp_factor_range),
        randomize_Kd_factor=bool(c.domain_rand.randomize_Kd_factor), Kd_factor_range=fl(c.domain_rand.Kd_factor_range),
        rand_interval=int(c.domain_rand.rand_interval),
        randomize_gravity=bool(c.domain_rand.randomize_gravity), gravity_range=fl(c.domain_rand.gravity_range),
        gravity_rand_interval=int(c.domain_rand.gravity_rand_interval),
        gravity_rand_duration=int(c.domain_rand.gravity_rand_duration),
        push_robots=bool(c.domain_rand.push_robots),
        randomize_rigids_after_start=bool(c.domain_rand.randomize_rigids_after_start),
        send_timeouts=bool(c.env.send_timeouts),
    )
    return flat
