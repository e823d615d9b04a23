"""Flat (JSON-able) configuration of the post-physics step path.

The reference configures the env through a mutable class tree (`Cfg`, reference
go1_gym/envs/base/legged_robot_trajectory_tracking_config.py:6-373) plus values derived at
construction time (`_parse_cfg` LR:1860-1878, `_process_dof_props` LR:680-708, `_create_envs`
LR:1540-1664, `_prepare_reward_function` LR:1368-1418; LR = legged_robot_trajectory_tracking.py).
The CUDA library takes a plain C struct (include/lt_b200.h `lt_config`); this module defines the
intermediate flat dict both are built from, so that fixtures generated from the *live reference env*
(oracle/ref_harness.flat_from_reference_env) and the product's own Cfg restatement
(legged_tracking_b200.cfg.flat_from_cfg) can be compared key by key.
"""
import json
import math

import numpy as np

# Reward terms the fused kernel knows (id = position).  Order is irrelevant to results; evaluation
# order is `reward_names` of the flat config (the reference's dict order, LR:1389-1397).
REWARD_TERM_IDS = {
    # RewardsCrawling (go1_gym/envs/rewards/reward_crawling.py:18-123)
    "dof_acc": 0, "torques": 1, "dof_pos_limits": 2, "collision": 3, "action_rate": 4, "base_height": 5,
    "ang_vel_xy": 6, "orientation": 7, "large_vel": 8, "e2e": 9, "exploration_lin": 10, "exploration_yaw": 11,
    "reaching_z": 12, "reaching_roll": 13, "reaching_pitch": 14,
    # TrajectoryTrackingRewards extras (go1_gym/envs/rewards/trajectory_tracking_reward.py:17-171)
    "dof_vel": 15, "dof_pos": 16, "task_old": 17, "reach_goal": 18, "reach_goal_t": 19, "reach_goal_T": 20,
    "task": 21, "exploration": 22, "stalling": 23, "linear_vel": 24, "lin_vel_z": 25, "feet_air_time": 26,
    "survive": 27, "reaching_linear_vel": 28, "reaching_yaw_abs": 29, "reaching_yaw": 30,
    # extra addend evaluated by a user plugin in torch on the device (see env.py)
    "external": 31,
    # TrajectoryTrackingRewards term that exists only under sampling-based planning (trajectory_tracking_reward.py:95-97)
    "reaching_local_goal": 32,
}
LIN_VEL_FORMS = {"exp": 0, "l1": 1, "l2": 2, "prod": 3}
TRAJ_FUNCTIONS = {"fixed_target": 0, "random_goal": 1, "random_target": 2, "valid_goal": 3}
CONTROL_TYPES = {"actuator_net": 0, "P": 1}
COMMAND_TYPES = {"xy": 0, "xy_norm": 1, "6dof": 2}

# Terms whose sign is not fixed by construction (value can be positive or negative for one sign of the
# scale): the pos/neg bucket of LR:332-335 then depends on the all-env sum.
SIGN_INDEFINITE_TERMS = {"exploration", "feet_air_time", "external"}
# Terms that are <= 0 by construction
NONPOSITIVE_TERMS = {"stalling"}


def _f(x):
    return float(x)


def _lst(x):
    return [float(v) for v in np.asarray(x, dtype=np.float64).reshape(-1)]


# keys added after the first fixtures were recorded, with the value every recording made before them implies
LATE_KEYS = {"plan_interval": 10, "candidate_target_poses": []}


def normalise(flat):
    """Round-trip through JSON so that two flat configs compare equal iff their JSON is equal."""
    out = json.loads(json.dumps(flat, sort_keys=True))
    for k, v in LATE_KEYS.items():
        out.setdefault(k, v)
    return out


def dumps(flat):
    return json.dumps(flat, sort_keys=True)


def loads(s):
    return json.loads(s)


def scale_shift(rng):
    """reference go1_gym/utils/math_utils.py:35-38"""
    scale = 2.0 / (rng[1] - rng[0])
    shift = (rng[1] + rng[0]) / 2.0
    return scale, shift


def obs_layout(flat):
    """Column layout of obs_buf implied by the observe_* flags (LR:357-475).  Returns list of (name, width)."""
    nd = flat["num_actions"]
    body = [("gravity", 3)]
    if flat["observe_command"]:
        body += [("commands", 2)]      # command_type "xy": local_relative_linear[:, :2] (LR:801-802)
    body += [("dof_pos", nd), ("dof_vel", nd), ("actions", nd)]
    if flat["observe_command"] and flat["timestep_in_obs"]:
        body += [("timestep", 1)]
    if flat["observe_heights"]:
        nx = len(flat["measured_points_x"])
        ny = len(flat["measured_points_y"])
        x_start = nx // 2 + 1 if flat["measure_front_half"] else 0
        body += [("heights", 2 * (nx - x_start) * ny)]
    if flat["observe_two_prev_actions"]:
        body += [("last_actions", nd)]
    # observe_only_* are prepended after observe_vel in the reference, i.e. they end up in front
    pre = []
    if flat["observe_only_lin_vel"]:
        pre += [("lin_vel_only", 3)]
    if flat["observe_only_ang_vel"]:
        pre += [("ang_vel_only", 3)]
    if flat["observe_vel"]:
        pre += [("lin_vel", 3), ("ang_vel", 3)]
    post = []
    if flat["observe_yaw"]:
        post += [("yaw", 1)]
    if flat["observe_contact_states"]:
        post += [("contact_states", len(flat["feet_indices"]))]
    return pre + body + post


def obs_width(flat):
    return sum(w for _, w in obs_layout(flat))


def priv_layout(flat):
    """Column layout of privileged_obs_buf (LR:477-590)."""
    cols = []
    if flat["priv_observe_friction"]:
        cols.append(("friction", 1))
    if flat["priv_observe_restitution"]:
        cols.append(("restitution", 1))
    if flat["priv_observe_base_mass"]:
        cols.append(("base_mass", 1))
    if flat["priv_observe_com_displacement"]:
        cols.append(("com_displacement", 3))
    if flat["priv_observe_motor_strength"]:
        cols.append(("motor_strength", flat["num_dof"]))
    if flat["priv_observe_motor_offset"]:
        cols.append(("motor_offset", flat["num_dof"]))
    if flat["priv_observe_body_height"]:
        cols.append(("body_height", 1))
    if flat["priv_observe_body_velocity"]:
        cols.append(("body_velocity", 3))
    if flat["priv_observe_gravity"]:
        cols.append(("gravity", 3))
    return cols


def priv_width(flat):
    return sum(w for _, w in priv_layout(flat))


def check_supported(flat):
    """Fail loudly for configurations the fused path does not implement (no silent fallback)."""
    problems = []
    if flat["rotate_camera"]:
        problems.append("env.rotate_camera=True (reference itself says do not use, LR:1934)")
    if flat["sampling_based_planning"] and not flat.get("candidate_target_poses"):
        problems.append("commands.sampling_based_planning=True without commands.candidate_target_poses")
    if flat["command_type"] != "xy":
        problems.append(f"env.command_type={flat['command_type']!r} (only 'xy' is on the hot path)")
    for k in ("observe_timing_parameter", "observe_clock_inputs", "priv_observe_ground_friction",
              "priv_observe_clock_inputs", "priv_observe_desired_contact_states", "push_robots",
              "randomize_rigids_after_start"):
        if flat.get(k):
            problems.append(f"{k}=True")
    if flat["traj_function"] not in ("fixed_target", "random_goal", "random_target"):
        problems.append(f"commands.traj_function={flat['traj_function']!r}")
    if "reaching_local_goal" in flat["reward_names"] and not (flat["sampling_based_planning"] and flat.get("plan_interval", 0) > 0):
        problems.append("reward term 'reaching_local_goal' without sampling-based planning on a plan interval (the reference has no "
                        "`replan` attribute then, LR:856-861)")
    for name in flat["reward_names"]:
        if name not in REWARD_TERM_IDS and name not in flat.get("external_reward_names", ()):
            problems.append(f"reward term {name!r} has no fused implementation")
    if obs_width(flat) != flat["num_obs"]:
        problems.append(f"observation layout is {obs_width(flat)} wide but num_obs={flat['num_obs']} (LR:475)")
    if priv_width(flat) != flat["num_privileged_obs"]:
        problems.append(f"privileged layout is {priv_width(flat)} wide but num_privileged_obs="
                        f"{flat['num_privileged_obs']} (LR:589)")
    if problems:
        raise NotImplementedError("unsupported configuration for the fused step: " + "; ".join(problems))


def max_episode_length_from(episode_length_s, dt):
    return float(math.ceil(episode_length_s / dt))
