"""CPU restatement of the velocity-tracking specifics of SURVEY.md section 8 rows a9.3 and (f) 3.  TEST INFRASTRUCTURE ONLY.

  step_contact_targets   LeggedRobot._step_contact_targets, go1_gym/envs/base/legged_robot_velocity_tracking.py:844-923 (LRV)
  corl_terms             CoRLRewards._reward_*, go1_gym/envs/rewards/corl_rewards.py:15-201 (CR)

Pinned against tests/golden/gait_corl_64.npz, recorded from the unmodified reference by oracle/gen_golden_gait.py.  The quaternion
helpers are the isaacgym.torch_utils restatements of oracle/shims (parity unpinned at that boundary).
"""
import math

import numpy as np
import torch


def step_contact_targets(gait_indices, commands, dt, kappa, pacing_offset=False):
    """-> dict(gait_indices, foot_indices, clock_inputs, doubletime_clock_inputs, halftime_clock_inputs, desired_contact_states,
    desired_footswing_height)"""
    frequencies, phases, offsets, bounds, durations = (commands[:, k] for k in (4, 5, 6, 7, 8))
    gait = torch.remainder(gait_indices + dt * frequencies, 1.0)                                   # LRV:851
    if pacing_offset:                                                                             # LRV:853-862
        idx = [gait + phases + offsets + bounds, gait + bounds, gait + offsets, gait + phases]
    else:
        idx = [gait + phases + offsets + bounds, gait + offsets, gait + bounds, gait + phases]
    foot_indices = torch.remainder(torch.stack(idx, dim=1), 1.0)                                   # LRV:864
    warped = []
    for x in idx:                                                                                 # LRV:866-872
        rem = torch.remainder(x, 1)
        stance, swing = rem < durations, rem > durations
        y = x.clone()
        y[stance] = rem[stance] * (0.5 / durations[stance])
        y[swing] = 0.5 + (rem[swing] - durations[swing]) * (0.5 / (1 - durations[swing]))
        warped.append(y)
    w = torch.stack(warped, dim=1)
    cdf = torch.distributions.normal.Normal(0, kappa).cdf                                           # LRV:893-895
    r = torch.remainder(w, 1.0)
    smoothing = cdf(r) * (1 - cdf(r - 0.5)) + cdf(r - 1) * (1 - cdf(r - 0.5 - 1))                  # LRV:897-916
    return dict(gait_indices=gait, foot_indices=foot_indices, clock_inputs=torch.sin(2 * np.pi * w),
                doubletime_clock_inputs=torch.sin(4 * np.pi * w), halftime_clock_inputs=torch.sin(np.pi * w),
                desired_contact_states=smoothing, desired_footswing_height=commands[:, 9])


def _quat_from_angle_axis(angle, axis):
    theta = (angle / 2).unsqueeze(-1)
    xyz = (axis / axis.norm(p=2, dim=-1, keepdim=True).clamp(min=1e-9)) * theta.sin()
    return torch.cat([xyz, theta.cos()], dim=-1)
    # isaacgym.torch_utils.quat_from_angle_axis also normalises the result: done by the caller below


def _normalize(x):
    return x / x.norm(p=2, dim=-1).clamp(min=1e-9).unsqueeze(-1)


def _quat_mul(a, b):
    x1, y1, z1, w1 = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    x2, y2, z2, w2 = b[:, 0], b[:, 1], b[:, 2], b[:, 3]
    ww = (z1 + x1) * (x2 + y2)
    yy = (w1 - y1) * (w2 + z2)
    zz = (w1 + y1) * (w2 - z2)
    xx = ww + yy + zz
    qq = 0.5 * (xx + (z1 - x1) * (x2 - y2))
    w = qq - ww + (z1 - y1) * (y2 - z2)
    x = qq - xx + (x1 + w1) * (x2 + w2)
    y = qq - yy + (w1 - x1) * (y2 + z2)
    z = qq - zz + (z1 + y1) * (w2 - x2)
    return torch.stack([x, y, z, w], dim=-1)


def corl_terms(e, cfg, names):
    """e: dict of the env tensors (names as in the reference), cfg: dict of cfg.rewards scalars -> ({name: [N]}, new last_contacts)"""
    from oracle.planner_port import quat_rotate_inverse, quat_apply, yaw_quat
    N = e["commands"].shape[0]
    feet, pen = e["feet_indices"].long(), e["penalised_contact_indices"].long()
    cmd = e["commands"]
    foot_forces = torch.norm(e["contact_forces"][:, feet, :], dim=-1)
    out = {}
    last_contacts = e["last_contacts"]
    for n in names:
        if n == "tracking_lin_vel":                                                                # CR:15-18
            err = torch.sum(torch.square(cmd[:, :2] - e["base_lin_vel"][:, :2]), dim=1)
            v = torch.exp(-err / cfg["tracking_sigma"])
        elif n == "tracking_ang_vel":                                                              # CR:20-23
            v = torch.exp(-torch.square(cmd[:, 2] - e["base_ang_vel"][:, 2]) / cfg["tracking_sigma_yaw"])
        elif n == "lin_vel_z":
            v = torch.square(e["base_lin_vel"][:, 2])
        elif n == "ang_vel_xy":
            v = torch.sum(torch.square(e["base_ang_vel"][:, :2]), dim=1)
        elif n == "orientation":
            v = torch.sum(torch.square(e["projected_gravity"][:, :2]), dim=1)
        elif n == "torques":
            v = torch.sum(torch.square(e["torques"]), dim=1)
        elif n == "dof_acc":
            v = torch.sum(torch.square((e["last_dof_vel"] - e["dof_vel"]) / e["dt"]), dim=1)
        elif n == "action_rate":
            v = torch.sum(torch.square(e["last_actions"] - e["actions"]), dim=1)
        elif n == "collision":                                                                     # CR:49-52
            v = torch.sum(1. * (torch.norm(e["contact_forces"][:, pen, :], dim=-1) > 0.1), dim=1)
        elif n == "dof_pos_limits":                                                                # CR:54-58
            o = -(e["dof_pos"] - e["dof_pos_limits"][:, 0]).clip(max=0.)
            o = o + (e["dof_pos"] - e["dof_pos_limits"][:, 1]).clip(min=0.)
            v = torch.sum(o, dim=1)
        elif n == "jump":                                                                          # CR:60-65
            v = -torch.square(e["base_pos"][:, 2] - (cmd[:, 3] + cfg["base_height_target"]))
        elif n == "tracking_contacts_shaped_force":                                                # CR:67-75
            v = 0
            for i in range(4):
                v = v + -(1 - e["desired_contact_states"][:, i]) * (1 - torch.exp(-1 * foot_forces[:, i] ** 2 / cfg["gait_force_sigma"]))
            v = v / 4
        elif n == "tracking_contacts_shaped_vel":                                                  # CR:77-84
            fv = torch.norm(e["foot_velocities"], dim=2).view(N, -1)
            v = 0
            for i in range(4):
                v = v + -(e["desired_contact_states"][:, i] * (1 - torch.exp(-1 * fv[:, i] ** 2 / cfg["gait_vel_sigma"])))
            v = v / 4
        elif n == "dof_pos":
            v = torch.sum(torch.square(e["dof_pos"] - e["default_dof_pos"]), dim=1)
        elif n == "dof_vel":
            v = torch.sum(torch.square(e["dof_vel"]), dim=1)
        elif n == "action_smoothness_1":                                                           # CR:94-98
            d = torch.square(e["joint_pos_target"] - e["last_joint_pos_target"]) * (e["last_actions"] != 0)
            v = torch.sum(d, dim=1)
        elif n == "action_smoothness_2":                                                           # CR:100-105
            d = torch.square(e["joint_pos_target"] - 2 * e["last_joint_pos_target"] + e["last_last_joint_pos_target"])
            d = d * (e["last_actions"] != 0) * (e["last_last_actions"] != 0)
            v = torch.sum(d, dim=1)
        elif n == "feet_slip":                                                                     # CR:107-113 (stateful)
            contact = e["contact_forces"][:, feet, 2] > 1.
            filt = torch.logical_or(contact, last_contacts)
            last_contacts = contact
            fv = torch.square(torch.norm(e["foot_velocities"][:, :, 0:2], dim=2).view(N, -1))
            v = torch.sum(filt * fv, dim=1)
        elif n == "feet_contact_vel":                                                              # CR:115-120
            near = e["foot_positions"][:, :, 2] < 0.03
            fv = torch.square(torch.norm(e["foot_velocities"][:, :, 0:3], dim=2).view(N, -1))
            v = torch.sum(near * fv, dim=1)
        elif n == "feet_contact_forces":                                                           # CR:122-125
            v = torch.sum((foot_forces - cfg["max_contact_force"]).clip(min=0.), dim=1)
        elif n == "feet_clearance_cmd_linear":                                                     # CR:127-132
            ph = 1 - torch.abs(1.0 - torch.clip((e["foot_indices"] * 2.0) - 1.0, 0.0, 1.0) * 2.0)
            target = cmd[:, 9].unsqueeze(1) * ph + 0.02
            v = torch.sum(torch.square(target - e["foot_positions"][:, :, 2]) * (1 - e["desired_contact_states"]), dim=1)
        elif n == "feet_impact_vel":                                                               # CR:134-141
            v = torch.sum((foot_forces > 1.0) * torch.square(torch.clip(e["prev_foot_velocities"][:, :, 2], -100, 0)), dim=1)
        elif n == "orientation_control":                                                           # CR:149-161
            rp = cmd[:, 10:12]
            q_roll = _normalize(_quat_from_angle_axis(-rp[:, 1], torch.tensor([1., 0., 0.]).repeat(N, 1)))
            q_pitch = _normalize(_quat_from_angle_axis(-rp[:, 0], torch.tensor([0., 1., 0.]).repeat(N, 1)))
            want = quat_rotate_inverse(_quat_mul(q_roll, q_pitch), e["gravity_vec"])
            v = torch.sum(torch.square(e["projected_gravity"][:, :2] - want[:, :2]), dim=1)
        elif n == "raibert_heuristic":                                                             # CR:163-201
            rel = e["foot_positions"] - e["base_pos"].unsqueeze(1)
            conj = torch.cat([-e["base_quat"][:, :3], e["base_quat"][:, 3:]], dim=-1)
            body = torch.stack([quat_apply(yaw_quat(conj), rel[:, i, :]) for i in range(4)], dim=1)
            width, length = cmd[:, 12:13], cmd[:, 13:14]
            ys = torch.cat([width / 2, -width / 2, width / 2, -width / 2], dim=1)
            xs = torch.cat([length / 2, length / 2, -length / 2, -length / 2], dim=1)
            ph = torch.abs(1.0 - (e["foot_indices"] * 2.0)) * 1.0 - 0.5
            freq = cmd[:, 4]
            y_vel = cmd[:, 2:3] * length / 2
            yo = ph * y_vel * (0.5 / freq.unsqueeze(1))
            yo[:, 2:4] *= -1
            xo = ph * cmd[:, 0:1] * (0.5 / freq.unsqueeze(1))
            want = torch.cat(((xs + xo).unsqueeze(2), (ys + yo).unsqueeze(2)), dim=2)
            v = torch.sum(torch.square(torch.abs(want - body[:, :, 0:2])), dim=(1, 2))
        else:
            raise KeyError(n)
        out[n] = v
    return out, last_contacts
