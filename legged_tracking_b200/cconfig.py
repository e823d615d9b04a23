"""flat config dict (flatcfg.py) -> lt_config / lt_tables of include/lt_b200.h, and the state buffer spec.

Python floats are passed through float32 exactly where the reference lets torch do it: torch multiplies a
float32 tensor by a Python float by first casting the scalar to float32, so `float32(x)` here reproduces the
reference's arithmetic (e.g. reward_scales * dt of LR:1385 is a double product that is cast once).
LR = reference go1_gym/envs/base/legged_robot_trajectory_tracking.py.
"""
import ctypes as C

import numpy as np

from . import _abi, flatcfg, rng_slots


def _bucket(name, scale, flat):
    """Which of rew_buf_pos / rew_buf_neg a term lands in (LR:332-335 tests the sign of the all-env sum)."""
    indefinite = name in flatcfg.SIGN_INDEFINITE_TERMS or name in flat.get("external_reward_names", ()) or (
        name == "exploration_lin" and flat["lin_vel_form"] == "prod")       # a normalised dot product (reward_crawling.py:95-101)
    if indefinite:
        return _abi.ENUMS["LT_BUCKET_BY_SUM"]
    sign = -1.0 if name in flatcfg.NONPOSITIVE_TERMS else 1.0
    return _abi.ENUMS["LT_BUCKET_POS"] if sign * scale >= 0 else _abi.ENUMS["LT_BUCKET_NEG"]


PRIV_ORDER = ["friction", "restitution", "added_mass", "com_displacement", "motor_strength", "motor_offset",
              "body_height", "body_velocity", "gravity"]


def make_lt_config(flat, seed=0, envs_per_warp=0):
    flatcfg.check_supported(flat)
    f = flat
    c = _abi.lt_config()
    c.struct_bytes = C.sizeof(_abi.lt_config)
    c.abi_version = _abi.DEFINES["LT_ABI_VERSION"]
    for k in ("num_envs", "num_actors", "num_bodies", "num_dof", "num_actions", "num_obs", "obs_history_length",
              "decimation", "lag_timesteps", "measure_front_half", "camera_zero", "tile_rows", "tile_cols", "traj_length",
              "num_interpolation", "switch_upon_reach", "use_terminal_body_height", "terminate_end_of_trajectory",
              "use_terminal_body_rotation", "only_positive_rewards", "only_positive_rewards_ji22_style", "observe_command",
              "timestep_in_obs", "observe_two_prev_actions", "observe_vel", "observe_only_ang_vel", "observe_only_lin_vel",
              "observe_yaw", "observe_contact_states", "global_reference", "add_noise", "observe_heights",
              "priv_observe_friction", "priv_observe_restitution", "priv_observe_base_mass",
              "priv_observe_com_displacement", "priv_observe_motor_strength", "priv_observe_motor_offset",
              "priv_observe_body_height", "priv_observe_body_velocity", "priv_observe_gravity", "custom_origins",
              "randomize_motor_strength", "randomize_motor_offset", "randomize_Kp_factor", "randomize_Kd_factor",
              "rand_interval", "send_timeouts", "randomize_lag_timesteps"):
        setattr(c, k, int(f[k]))
    c.num_privileged_obs = int(f["num_privileged_obs"])
    c.control_type = flatcfg.CONTROL_TYPES[f["control_type"]]
    c.traj_function = flatcfg.TRAJ_FUNCTIONS[f["traj_function"]]
    c.lin_vel_form = flatcfg.LIN_VEL_FORMS[f["lin_vel_form"]]
    c.plane = int(f["mesh_type"] == "plane")
    for k in ("dt", "clip_actions", "clip_observations", "action_scale", "hip_scale_reduction", "horizontal_scale",
              "ceiling_height", "base_x", "base_y", "base_z", "base_roll", "base_pitch", "base_yaw", "x_mean", "x_range",
              "y_mean", "y_range", "z_range", "roll_range", "pitch_range", "yaw_range", "switch_dist", "switch_yaw",
              "switch_interval", "max_episode_length", "terminal_body_height", "T_reach", "sigma_rew_neg",
              "base_height_target", "tracking_sigma_lin", "tracking_sigma_ang", "target_lin_vel", "target_ang_vel",
              "lin_reaching_criterion", "ang_reaching_criterion", "small_vel_threshold", "large_dist_threshold",
              "obs_scale_lin_vel", "obs_scale_ang_vel", "obs_scale_dof_pos", "obs_scale_dof_vel", "obs_scale_height",
              "x_init_range", "y_init_range", "x_init_offset", "y_init_offset", "yaw_init_range"):
        setattr(c, k, float(f[k]))
    for dst, src in (("default_dof_pos", "default_dof_pos"), ("torque_limits", "torque_limits"), ("p_gains", "p_gains"),
                     ("d_gains", "d_gains"), ("dof_pos_lower", "dof_pos_limits_lower"), ("dof_pos_upper", "dof_pos_limits_upper"),
                     ("camera_offset", "camera_offset"), ("base_init_state", "base_init_state"),
                     ("motor_strength_range", "motor_strength_range"), ("motor_offset_range", "motor_offset_range"),
                     ("Kp_factor_range", "Kp_factor_range"), ("Kd_factor_range", "Kd_factor_range")):
        arr = getattr(c, dst)
        vals = list(f[src])
        assert len(vals) <= len(arr), (dst, len(vals))
        for i, v in enumerate(vals):
            arr[i] = float(v)
    feet = list(f["feet_indices"])
    c.n_feet = len(feet)
    for i, b in enumerate(feet):
        c.feet_indices[i] = int(b)
    c.penalised_mask = sum(1 << int(b) for b in set(f["penalised_contact_indices"]))
    c.termination_mask = sum(1 << int(b) for b in set(f["termination_contact_indices"]))
    px, py = list(f["measured_points_x"]), list(f["measured_points_y"])
    c.n_points_x, c.n_points_y = len(px), len(py)
    for i, v in enumerate(px):
        c.points_x[i] = float(v)
    for i, v in enumerate(py):
        c.points_y[i] = float(v)
    names = list(f["reward_names"])
    c.n_terms = len(names)
    for i, n in enumerate(names):
        scale = float(f["reward_scales"][n])
        c.term_ids[i] = flatcfg.REWARD_TERM_IDS["external" if n in f.get("external_reward_names", ()) else n]
        c.term_scales[i] = scale
        c.term_bucket[i] = _bucket(n, scale, f)
    for i, n in enumerate(PRIV_ORDER):
        scale, shift = flatcfg.scale_shift(f["norm_" + n + "_range"])
        c.priv_scale[i] = scale
        c.priv_shift[i] = shift
    c.envs_per_warp = int(envs_per_warp)
    c.seed_lo = int(seed) & 0xFFFFFFFF
    c.seed_hi = (int(seed) >> 32) & 0xFFFFFFFF
    return c


class Tables:
    """Owns the host arrays lt_tables points to (they only need to live until lt_create returns)."""

    def __init__(self, flat, actuator):
        self.arrays = {
            "noise_scale_vec": np.ascontiguousarray(np.asarray(flat["noise_scale_vec"], dtype=np.float32)),
            "act_w0": np.ascontiguousarray(actuator["w0"], dtype=np.float32),
            "act_b0": np.ascontiguousarray(actuator["b0"], dtype=np.float32),
            "act_w1": np.ascontiguousarray(actuator["w1"], dtype=np.float32),
            "act_b1": np.ascontiguousarray(actuator["b1"], dtype=np.float32),
            "act_w2": np.ascontiguousarray(actuator["w2"], dtype=np.float32),
            "act_b2": np.ascontiguousarray(actuator["b2"], dtype=np.float32),
        }
        assert self.arrays["noise_scale_vec"].size == flat["num_obs"]
        assert self.arrays["act_w0"].shape == (32, 6) and self.arrays["act_w1"].shape == (32, 32)
        self.c = _abi.lt_tables()
        for k, a in self.arrays.items():
            setattr(self.c, k, a.ctypes.data)


def state_spec(flat):
    """[(member of lt_state_ptrs, dtype name, shape)] -- the attributes LeggedRobot._init_buffers allocates
    (LR:1203-1305) in the struct-of-arrays layout the kernels use."""
    f = flat
    N, D, A, F = f["num_envs"], f["num_dof"], f["num_actions"], len(f["feet_indices"])
    T = len(f["reward_names"])
    H = f["obs_history_length"]
    nx, ny = len(f["measured_points_x"]), len(f["measured_points_y"])
    spec = []
    for n in ("actions", "last_actions", "last_last_actions"):
        spec.append((n, "f32", (N, A)))
    for n in ("joint_pos_target", "last_joint_pos_target", "last_last_joint_pos_target", "last_dof_vel"):
        spec.append((n, "f32", (N, D)))
    spec.append(("last_root_vel", "f32", (N, 6)))
    for n in ("torques", "joint_pos_err_last", "joint_pos_err_last_last", "joint_vel_last", "joint_vel_last_last",
              "motor_strengths", "motor_offsets", "Kp_factors", "Kd_factors"):
        spec.append((n, "f32", (N, D)))
    spec += [("friction_coeffs", "f32", (N, 1)), ("restitutions", "f32", (N, 1)), ("payloads", "f32", (N, 1)),
             ("com_displacements", "f32", (N, 3)), ("gravities", "f32", (N, 3)), ("gravity_vec", "f32", (3,))]
    for n in ("base_lin_vel", "base_ang_vel", "projected_gravity", "base_rotation", "relative_linear", "relative_rotation",
              "local_relative_linear", "local_relative_rotation"):
        spec.append((n, "f32", (N, 3)))
    spec += [("target_poses", "f32", (N, 6)), ("feet_air_time", "f32", (N, F)), ("rew_buf", "f32", (N,)),
             ("rew_buf_pos", "f32", (N,)), ("rew_buf_neg", "f32", (N,)), ("camera_pitch_angle", "f32", (N,)),
             ("trajectories", "f32", (N, f["traj_length"], 6)), ("lag_buffer", "f32", (f["lag_timesteps"] + 1, N, D)),
             ("episode_sums", "f32", (T + 3, N)), ("command_sums", "f32", (T + 1, N)),
             ("obs_buf", "f32", (N, f["num_obs"])), ("privileged_obs_buf", "f32", (N, f["num_privileged_obs"])),
             ("obs_history", "f32", (N, H * f["num_obs"])), ("finished_sums", "f32", (T + 3, N)),
             ("finished_episode_length", "f32", (N,)), ("finished_goal_distance", "f32", (N,))]
    for n in ("episode_length_buf", "curr_pose_index", "collision_count", "collision_env_buf", "reset_env_ids"):
        spec.append((n, "i64", (N,)))
    for n in ("reset_buf", "time_out_buf", "body_height_buf", "switched_buf", "reached_buf", "plan_buf", "reached_env_buf",
              "time_outs_latched"):
        spec.append((n, "bool", (N,)))
    spec += [("last_contacts", "bool", (N, F)), ("reset_count", "i32", (1,)),
             ("height_px", "i32", (N, nx)), ("height_py", "i32", (N, ny))]
    members = [m for m, _ in _abi.lt_state_ptrs._fields_]
    assert [s[0] for s in spec] == members, "state_spec out of sync with lt_state_ptrs"
    return spec


# Which state buffers the post-physics phase of ONE env-step reads / writes (SURVEY.md section 8d's accounting): name -> role.
# workload.algorithmic_bytes derives the per-env byte counts from state_spec's shapes with this table, so that a buffer that
# is renamed, resized or dropped changes the figure (tests/test_cfg_flat.py pins it to SURVEY's 4,121 B at cfg 2 / 5).
# Buffers that are touched only on reset / interval DR (lag_buffer, finished_*, Kp_factors ...) are not part of the figure.
POST_PHYSICS_TRAFFIC = {
    "actions": "r", "last_actions": "rw", "last_last_actions": "w", "last_dof_vel": "rw", "torques": "r",
    "last_root_vel": "w", "base_lin_vel": "w", "base_ang_vel": "w", "projected_gravity": "w", "base_rotation": "rw_row1",
    "relative_linear": "w", "relative_rotation": "w", "episode_sums": "rw", "command_sums": "rw",
    "obs_buf": "w", "privileged_obs_buf": "w", "obs_history": "w", "rew_buf": "w",
    "episode_length_buf": "rw", "curr_pose_index": "rw", "collision_count": "rw",
    "reset_buf": "w", "time_out_buf": "w", "reached_buf": "w", "switched_buf": "w", "plan_buf": "w",
    "friction_coeffs": "r", "restitutions": "r",
    # moved into the torque kernel's last sub-step (LR:151-152): counted with the torque phase
    "last_joint_pos_target": "torque_w", "last_last_joint_pos_target": "torque_w",
}


# initial values that are not zero (LR:1203-1305)
STATE_INIT = {"motor_strengths": 1.0, "Kp_factors": 1.0, "Kd_factors": 1.0}
OPTIONAL_STATE = ("command_sums", "height_px", "height_py")


def rng_n_slots(flat):
    """Row width of the per-step uniform array (mirrors lt_num_uniform_slots / rng_slots.n_slots)."""
    return rng_slots.n_slots(flat)
