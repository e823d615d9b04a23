"""Config classes of the path, with the reference's names and defaults.

`Cfg` is the same class tree as the reference's go1_gym/envs/base/legged_robot_trajectory_tracking_config.py:6-373 (attribute
names and default values are loaded from data/cfg_defaults.json, which oracle/gen_cfg_defaults.py dumped from the reference
itself, so nothing is transcribed by hand), `config_go1(Cfg)` applies what go1_gym/envs/go1/go1_crawling.py:8-106 sets, and
`apply_preset(Cfg, name)` what scripts/train.py:46-241 sets for the argument sets of the golden fixtures.  Any class tree
with these attribute names works -- the reference's own `Cfg` object can be passed to `flat_from_cfg` unchanged.

`flat_from_cfg` derives the flat step configuration (flatcfg.py) the way the reference's constructor does:
_parse_cfg LR:1860-1878, _process_dof_props LR:680-708, _init_buffers LR:1221-1305, _create_envs LR:1540-1664,
_get_noise_scale_vec LR:1086-1166, _prepare_reward_function LR:1368-1418 (LR = legged_robot_trajectory_tracking.py).
"""
import json
import math
import os

import numpy as np

from . import flatcfg
from .params import PrefixProto

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "cfg_defaults.json")
_defaults = None


def defaults():
    global _defaults
    if _defaults is None:
        with open(_DATA) as f:
            _defaults = json.load(f)
    return _defaults


def _decode(v):
    if isinstance(v, dict):
        if "__ndarray__" in v:
            return np.asarray(v["__ndarray__"], dtype=v["dtype"])
        if "__float__" in v:
            return float(v["__float__"])
        if "__dict__" in v:
            return {k: _decode(x) for k, x in v["__dict__"].items()}
    if isinstance(v, list):
        return [_decode(x) for x in v]
    return v


def _build(name, node):
    ns = {}
    for k, v in node.items():
        ns[k] = _build(k, v["__class__"]) if isinstance(v, dict) and "__class__" in v else _decode(v)
    return type(PrefixProto)(name, (PrefixProto,), ns)


def _apply(cls, node):
    for k, v in node.items():
        if isinstance(v, dict) and "__class__" in v:
            if not isinstance(getattr(cls, k, None), type):
                setattr(cls, k, _build(k, {}))
            _apply(getattr(cls, k), v["__class__"])
        else:
            setattr(cls, k, _decode(v))


def make_cfg():
    """A fresh, pristine `Cfg` tree (the module-level `Cfg` is one such tree; the reference mutates its Cfg in place)."""
    return _build("Cfg", defaults()["base"])


Cfg = make_cfg()


def config_go1(Cnfg):
    """go1_gym/envs/go1/go1_crawling.py:8-106."""
    _apply(Cnfg, defaults()["go1"])
    return Cnfg


def preset_names():
    return list(defaults()["presets"])


def apply_preset(Cnfg, name):
    """What scripts/train.py:46-241 changes for one of the recorded argument sets (defaults()['preset_argv'][name])."""
    _apply(Cnfg, defaults()["presets"][name])
    return Cnfg


def preset_cfg(name, num_envs=None):
    c = apply_preset(config_go1(make_cfg()), name)
    if num_envs is not None:
        c.env.num_envs = int(num_envs)
    return c


class Go1Model:
    """What the reference reads from the robot asset through Isaac Gym (resources/robots/go1/urdf/go1.urdf):
    body/dof names in asset order and the URDF joint limits.  Index sets are derived from the names exactly as
    LR:1545-1551 / :1647-1664 do, so another asset ordering only needs another instance of this class."""
    legs = ("FL", "FR", "RL", "RR")
    body_names = ["base"] + [f"{l}_{p}" for l in legs for p in ("hip", "thigh", "calf", "foot")]
    dof_names = [f"{l}_{p}_joint" for l in legs for p in ("hip", "thigh", "calf")]
    dof_lower = [-0.802851455917, -1.0471975512, -2.69653369433] * 4
    dof_upper = [0.802851455917, 4.18879020479, -0.916297857297] * 4
    dof_velocity = [50.0, 28.0, 28.0] * 4
    dof_effort = [33.5] * 12
    num_extra_actor_bodies = 1          # the goal arrow actor (LR:934-947)


def _f32(x):
    return float(np.float32(x))


def _f32l(xs):
    return [float(v) for v in np.asarray(xs, dtype=np.float32).reshape(-1)]


def noise_scale_vec(cfg, num_actions=12, num_actuated_dof=12, n_feet=4):
    """LR:1086-1166, evaluated in float32 like the reference's torch.ones(n) * a * b * c chains."""
    ns, lvl, osc, e = cfg.noise_scales, cfg.noise.noise_level, cfg.obs_scales, cfg.env
    f = np.float32

    def ones(n, *factors):
        v = np.ones(n, np.float32)
        for x in factors:
            v = v * f(x)
        return v
    zeros = lambda n: np.zeros(int(n), np.float32)   # noqa: E731
    body = [ones(3, ns.gravity, lvl)]
    if e.observe_command:
        body.append(zeros(2 if e.command_xy_only else 6))
    body += [ones(num_actuated_dof, ns.dof_pos, lvl, osc.dof_pos), ones(num_actuated_dof, ns.dof_vel, lvl, osc.dof_vel),
             zeros(num_actions)]
    if getattr(e, "timestep_in_obs", False):
        body.append(zeros(1))
    if e.observe_heights:
        nx, ny = len(cfg.terrain.measured_points_x), len(cfg.terrain.measured_points_y)
        body.append(zeros((nx // 2) * ny * 2 if cfg.terrain.measure_front_half else 2 * nx * ny))
    if e.observe_two_prev_actions:
        body.append(zeros(num_actions))
    if e.observe_timing_parameter:
        body.append(zeros(1))
    if e.observe_clock_inputs:
        body.append(zeros(4))
    if e.observe_vel:
        body = [ones(3, ns.lin_vel, lvl, osc.lin_vel), ones(3, ns.ang_vel, lvl, osc.ang_vel)] + body
    if e.observe_only_lin_vel:
        body = [ones(3, ns.lin_vel, lvl, osc.lin_vel)] + body
    if e.observe_yaw:
        body.append(zeros(1))
    if e.observe_contact_states:
        body.append(ones(n_feet, ns.contact_states, lvl))
    return np.concatenate(body)


def fused_reward_names(container_name):
    """`_reward_<name>` methods of the two shipped containers that have a fused implementation
    (reward_crawling.py:18-123, trajectory_tracking_reward.py:17-171)."""
    crawling = ["dof_acc", "torques", "dof_pos_limits", "collision", "action_rate", "base_height", "ang_vel_xy", "orientation",
                "large_vel", "e2e", "exploration_lin", "exploration_yaw", "reaching_z", "reaching_roll", "reaching_pitch"]
    tracking = ["torques", "dof_vel", "dof_acc", "dof_pos", "collision", "action_rate", "dof_pos_limits", "orientation",
                "task_old", "reach_goal", "reach_goal_t", "reach_goal_T", "task", "exploration", "stalling", "linear_vel",
                "lin_vel_z", "ang_vel_xy", "feet_air_time", "survive", "reaching_linear_vel", "reaching_z", "reaching_roll",
                "reaching_pitch", "reaching_yaw_abs", "reaching_yaw", "reaching_local_goal"]
    table = {"RewardsCrawling": crawling, "TrajectoryTrackingRewards": tracking}
    if container_name not in table:
        raise NotImplementedError(f"reward container {container_name!r} has no fused implementation (CoRLRewards: cfg 4, DESIGN.md section 8)")
    return table[container_name]


def flat_from_cfg(cfg, model=Go1Model, sim_dt=None, num_envs=None, extra_reward_names=()):
    """Flat step configuration from a reference-style `cfg` tree.  `sim_dt` is SimParams.dt as PhysX holds it (a float32:
    0.005 -> 0.004999999888...).  `extra_reward_names`: `_reward_*` methods of a user container that are not
    fused (they are evaluated on the device in torch and enter through LT_TERM_EXTERNAL)."""
    c = cfg
    g = lambda obj, name, default=None: getattr(obj, name, default)   # noqa: E731
    sim_dt = _f32(0.005) if sim_dt is None else float(sim_dt)
    dt = c.control.decimation * sim_dt
    nd = len(model.dof_names)
    plane = c.terrain.mesh_type == "plane"
    # _process_dof_props (LR:692-706), float32 arithmetic
    f = np.float32
    lo, hi = [], []
    for a, b in zip(model.dof_lower, model.dof_upper):
        a, b = f(a), f(b)
        m = (a + b) / f(2)
        r = b - a
        half = f(f(0.5) * r) * f(c.rewards.soft_dof_pos_limit)
        lo.append(float(m - half))
        hi.append(float(m + half))
    # _init_buffers (LR:1288-1305)
    default_dof_pos, p_gains, d_gains = [], [], []
    for name in model.dof_names:
        default_dof_pos.append(_f32(c.init_state.default_joint_angles[name]))
        kp = kd = 0.0
        for key in c.control.stiffness.keys():
            if key in name:
                kp, kd = c.control.stiffness[key], c.control.damping[key]
        p_gains.append(_f32(kp))
        d_gains.append(_f32(kd))
    # _create_envs (LR:1545-1551)
    bodies = model.body_names
    feet = [bodies.index(s) for s in bodies if c.asset.foot_name in s]
    pen = [bodies.index(s) for n in c.asset.penalize_contacts_on for s in bodies if n in s]
    term = [bodies.index(s) for n in c.asset.terminate_after_contacts_on for s in bodies if n in s]
    # _parse_cfg (LR:1860-1878)
    ceil = lambda x: int(np.ceil(x))   # noqa: E731
    max_episode_length = float(np.ceil(c.env.episode_length_s / dt))
    rand_interval = ceil(c.domain_rand.rand_interval_s / dt)
    gravity_rand_interval = ceil(c.domain_rand.gravity_rand_interval_s / dt)
    gravity_rand_duration = ceil(gravity_rand_interval * c.domain_rand.gravity_impulse_duration)
    # _prepare_reward_function (LR:1379-1397)
    scales_cfg = {k: float(v) for k, v in vars(c.reward_scales).items()}
    fused = set(fused_reward_names(c.rewards.reward_container_name))
    scales, names = {}, []
    for k, v in scales_cfg.items():
        if v == 0:
            continue
        scales[k] = v * dt
        if k != "termination" and (k in fused or k in extra_reward_names):
            names.append(k)
    nsv = noise_scale_vec(c, c.env.num_actions, nd, len(feet))
    init = c.init_state
    flat = dict(
        num_envs=int(c.env.num_envs if num_envs is None else num_envs), num_actors=1 + model.num_extra_actor_bodies,
        num_bodies=len(bodies) + model.num_extra_actor_bodies, num_dof=nd, num_actions=int(c.env.num_actions),
        num_obs=int(c.env.num_observations), num_privileged_obs=int(c.env.num_privileged_obs),
        obs_history_length=int(c.env.num_observation_history),
        dt=float(dt), sim_dt=float(sim_dt), decimation=int(c.control.decimation),
        clip_actions=float(c.normalization.clip_actions), clip_observations=float(c.normalization.clip_observations),
        control_type=str(c.control.control_type), action_scale=float(c.control.action_scale),
        hip_scale_reduction=float(c.control.hip_scale_reduction),
        randomize_lag_timesteps=bool(c.domain_rand.randomize_lag_timesteps), lag_timesteps=int(c.domain_rand.lag_timesteps),
        default_dof_pos=default_dof_pos, torque_limits=_f32l(model.dof_effort), p_gains=p_gains, d_gains=d_gains,
        dof_pos_limits_lower=lo, dof_pos_limits_upper=hi,
        feet_indices=feet, penalised_contact_indices=pen, termination_contact_indices=term,
        mesh_type="plane" if plane else "trimesh", observe_heights=bool(c.env.observe_heights),
        measured_points_x=_f32l(c.terrain.measured_points_x), measured_points_y=_f32l(c.terrain.measured_points_y),
        measure_front_half=bool(c.terrain.measure_front_half), camera_zero=bool(c.env.camera_zero),
        rotate_camera=bool(c.env.rotate_camera), camera_offset=_f32l([0.12, 0.0, 0.0]),          # LR:1224
        horizontal_scale=float(c.terrain.horizontal_scale),
        tile_rows=0 if plane else int(c.terrain.terrain_length / c.terrain.horizontal_scale),      # tunnel.py:69-70
        tile_cols=0 if plane else int(c.terrain.terrain_width / c.terrain.horizontal_scale),
        terrain_num_rows=int(c.terrain.num_rows), terrain_num_cols=int(c.terrain.num_cols),
        ceiling_height=float(c.terrain.ceiling_height),
        command_type=str(c.env.command_type), command_xy_only=bool(c.env.command_xy_only),
        traj_function=str(c.commands.traj_function), traj_length=int(c.commands.traj_length),
        num_interpolation=int(c.commands.num_interpolation),
        **{k: float(getattr(c.commands, k)) for k in ("base_x", "base_y", "base_z", "base_roll", "base_pitch", "base_yaw", "x_mean",
                                                     "x_range", "y_mean", "y_range", "z_range", "roll_range", "pitch_range",
                                                     "yaw_range", "switch_dist", "switch_yaw", "switch_interval")},
        switch_upon_reach=bool(c.commands.switch_upon_reach),
        sampling_based_planning=bool(c.commands.sampling_based_planning), global_reference=bool(c.commands.global_reference),
        plan_interval=int(getattr(c.commands, "plan_interval", 10)),
        candidate_target_poses=[[float(v) for v in row] for row in np.asarray(getattr(c.commands, "candidate_target_poses", np.zeros((0, 6)))).reshape(-1, 6)]
        if c.commands.sampling_based_planning else [],
        max_episode_length=max_episode_length,
        use_terminal_body_height=bool(c.rewards.use_terminal_body_height), terminal_body_height=float(c.rewards.terminal_body_height),
        terminate_end_of_trajectory=bool(c.env.terminate_end_of_trajectory), T_reach=float(g(c.rewards, "T_reach", 0)),
        use_terminal_body_rotation=bool(c.env.use_terminal_body_rotation),
        reward_container_name=str(c.rewards.reward_container_name), reward_names=names, reward_scales=scales,
        reward_scales_cfg=scales_cfg,
        only_positive_rewards=bool(c.rewards.only_positive_rewards),
        only_positive_rewards_ji22_style=bool(c.rewards.only_positive_rewards_ji22_style),
        sigma_rew_neg=float(c.rewards.sigma_rew_neg), base_height_target=float(g(c.rewards, "base_height_target", 0.0)),
        tracking_sigma_lin=float(c.rewards.tracking_sigma_lin), tracking_sigma_ang=float(c.rewards.tracking_sigma_ang),
        target_lin_vel=float(c.rewards.target_lin_vel), target_ang_vel=float(c.rewards.target_ang_vel),
        lin_reaching_criterion=float(c.rewards.lin_reaching_criterion), ang_reaching_criterion=float(c.rewards.ang_reaching_criterion),
        lin_vel_form=str(g(c.rewards, "lin_vel_form", "exp")), small_vel_threshold=float(g(c.rewards, "small_vel_threshold", 0.0)),
        large_dist_threshold=float(g(c.rewards, "large_dist_threshold", 0.0)),
        exploration_steps=float(g(c.rewards, "exploration_steps", float("inf"))),
        soft_dof_pos_limit=float(g(c.rewards, "soft_dof_pos_limit", 1.0)),
        observe_command=bool(c.env.observe_command), timestep_in_obs=bool(g(c.env, "timestep_in_obs", False)),
        **{k: bool(getattr(c.env, k)) for k in ("observe_two_prev_actions", "observe_timing_parameter", "observe_clock_inputs",
                                                "observe_vel", "observe_only_ang_vel", "observe_only_lin_vel", "observe_yaw",
                                                "observe_contact_states")},
        obs_scale_lin_vel=float(c.obs_scales.lin_vel), obs_scale_ang_vel=float(c.obs_scales.ang_vel),
        obs_scale_dof_pos=float(c.obs_scales.dof_pos), obs_scale_dof_vel=float(c.obs_scales.dof_vel),
        obs_scale_height=float(c.obs_scales.height_measurements),
        add_noise=bool(c.noise.add_noise), noise_scale_vec=[float(v) for v in nsv],
        **{k: bool(getattr(c.env, k)) for k in ("priv_observe_friction", "priv_observe_restitution", "priv_observe_base_mass",
                                                "priv_observe_com_displacement", "priv_observe_motor_strength",
                                                "priv_observe_motor_offset", "priv_observe_body_height", "priv_observe_body_velocity",
                                                "priv_observe_gravity", "priv_observe_ground_friction", "priv_observe_clock_inputs",
                                                "priv_observe_desired_contact_states")},
        **{"norm_" + k: _f32l(getattr(c.normalization, k)) for k in
           ("friction_range", "restitution_range", "added_mass_range", "com_displacement_range", "motor_strength_range",
            "motor_offset_range", "body_height_range", "body_velocity_range", "gravity_range")},
        base_init_state=_f32l(list(init.pos) + list(init.rot) + list(init.lin_vel) + list(init.ang_vel)),     # LR:1553
        custom_origins=not plane and c.terrain.mesh_type in ("heightfield", "trimesh"),
        **{k: float(getattr(c.terrain, k)) for k in ("x_init_range", "y_init_range", "x_init_offset", "y_init_offset", "yaw_init_range")},
        randomize_motor_strength=bool(c.domain_rand.randomize_motor_strength), motor_strength_range=_f32l(c.domain_rand.motor_strength_range),
        randomize_motor_offset=bool(c.domain_rand.randomize_motor_offset), motor_offset_range=_f32l(c.domain_rand.motor_offset_range),
        randomize_Kp_factor=bool(c.domain_rand.randomize_Kp_factor), Kp_factor_range=_f32l(c.domain_rand.Kp_factor_range),
        randomize_Kd_factor=bool(c.domain_rand.randomize_Kd_factor), Kd_factor_range=_f32l(c.domain_rand.Kd_factor_range),
        rand_interval=rand_interval,
        randomize_gravity=bool(c.domain_rand.randomize_gravity), gravity_range=_f32l(c.domain_rand.gravity_range),
        gravity_rand_interval=gravity_rand_interval, gravity_rand_duration=gravity_rand_duration,
        push_robots=bool(c.domain_rand.push_robots), randomize_rigids_after_start=bool(c.domain_rand.randomize_rigids_after_start),
        send_timeouts=bool(c.env.send_timeouts),
    )
    if extra_reward_names:
        flat["external_reward_names"] = [n for n in names if n in extra_reward_names]
    return flat
