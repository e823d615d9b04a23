"""CPU restatement of the reference's sampling-based local planner.  TEST INFRASTRUCTURE ONLY.

Follows LeggedRobot._plan_target_pose, go1_gym/envs/base/legged_robot_trajectory_tracking.py:850-920 (LR), and
_compute_relative_target_pose :922-932, with the per-env Python loop of :875-912 vectorised over candidates; math helpers are the
shim's isaacgym.torch_utils restatements and go1_gym/utils/math_utils.py:12-61 restated below.  Pinned against
tests/golden/planner_48.npz, recorded from the unmodified reference by oracle/gen_golden_planner.py.
"""
import numpy as np
import torch

ROBOT_SIZE = (0.3762, 0.0935, 0.114)          # LR:1212


def normalize(x, eps=1e-9):
    return x / x.norm(p=2, dim=-1).clamp(min=eps).unsqueeze(-1)


def quat_rotate_inverse(q, v):
    q_w, q_vec = q[:, -1], q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(-1, 1, 3), v.view(-1, 3, 1)).squeeze(-1) * 2.0
    return a - b + c


def quat_apply(a, b):                          # planner.py:17-23
    xyz = a[:, :3]
    t = xyz.cross(b, dim=-1) * 2
    return b + a[:, 3:] * t + xyz.cross(t, dim=-1)


def yaw_quat(q):
    qy = q.clone().view(-1, 4)
    qy[:, :2] = 0.0
    return normalize(qy)


def quat_from_euler_xyz(roll, pitch, yaw):     # planner.py:25-38
    cy, sy = torch.cos(yaw * 0.5), torch.sin(yaw * 0.5)
    cr, sr = torch.cos(roll * 0.5), torch.sin(roll * 0.5)
    cp, sp = torch.cos(pitch * 0.5), torch.sin(pitch * 0.5)
    return torch.stack([cy * sr * cp - sy * cr * sp, cy * cr * sp + sy * sr * cp, sy * cr * cp - cy * sr * sp,
                        cy * cr * cp + sy * sr * sp], dim=-1)


def get_euler_xyz(q):                          # planner.py:48-70
    qx, qy, qz, qw = 0, 1, 2, 3
    sinr_cosp = 2.0 * (q[:, qw] * q[:, qx] + q[:, qy] * q[:, qz])
    cosr_cosp = q[:, qw] * q[:, qw] - q[:, qx] * q[:, qx] - q[:, qy] * q[:, qy] + q[:, qz] * q[:, qz]
    roll = torch.atan2(sinr_cosp, cosr_cosp)
    sinp = 2.0 * (q[:, qw] * q[:, qy] - q[:, qz] * q[:, qx])
    pitch = torch.where(torch.abs(sinp) >= 1, torch.abs(torch.tensor(np.pi / 2.0)) * torch.sign(sinp), torch.asin(sinp))
    siny_cosp = 2.0 * (q[:, qw] * q[:, qz] + q[:, qx] * q[:, qy])
    cosy_cosp = q[:, qw] * q[:, qw] + q[:, qx] * q[:, qx] - q[:, qy] * q[:, qy] - q[:, qz] * q[:, qz]
    yaw = torch.atan2(siny_cosp, cosy_cosp)
    return roll % (2 * np.pi), pitch % (2 * np.pi), yaw % (2 * np.pi)


def wrap_to_pi(angles):                        # math_utils.py:20-23
    angles = angles % (2 * np.pi)
    return angles - 2 * np.pi * (angles > np.pi)


def rpy(q):                                    # math_utils.py:40-46
    return wrap_to_pi(torch.stack(get_euler_xyz(q), dim=-1))


def relative_pose(target, base_pos, base_quat):            # LR:922-932
    lin = quat_rotate_inverse(yaw_quat(base_quat), target[:, :3] - base_pos)
    rot = wrap_to_pi(target[:, 3:] - rpy(base_quat))
    return lin, rot


def plan(target_poses, base_pos, base_quat, points_xy, measured_heights, candidates, episode_length_buf, plan_buf,
         plan_length_buf, local_target_poses, plan_interval=10, stable=False):
    """-> dict(local_target_poses, plan_length_buf, local_relative_linear, local_relative_rotation, chosen, key_of_chosen).

    stable=False is the reference: torch.argsort without `stable` (LR:887), whose order among EQUAL keys is whatever this torch
    build's sort produces -- and equal keys are systematic here, the key ignores the candidate's z (five z values per tie group).
    stable=True resolves ties to the lowest candidate index, the rule lt_plan_target_pose implements."""
    N = target_poses.shape[0]
    rel_lin, _ = relative_pose(target_poses, base_pos, base_quat)
    pl = plan_length_buf + 1                                                # LR:858
    close = torch.norm(rel_lin[:, :2], dim=1) < 1.0                          # LR:859
    replan = (pl % plan_interval) == 0
    do_plan = (episode_length_buf == 1) | (replan & (plan_buf != 0))         # LR:861-865
    local = local_target_poses.clone()
    chosen = torch.full((N,), -3, dtype=torch.int32)
    key_of = torch.full((N,), float("nan"))
    cand = candidates.float()
    n_pts = points_xy.shape[0]
    base_rot = rpy(base_quat)
    for e in do_plan.nonzero().flatten().tolist():
        pl[e] = 0                                                           # LR:874
        if close[e]:
            local[e] = target_poses[e]
            chosen[e] = -2
            continue
        goal = target_poses[e].clone()
        goal[:2] -= base_pos[e, :2]                                          # LR:881
        key = torch.linalg.norm(cand[:, :2] - goal[:2], dim=1) + torch.linalg.norm(cand[:, 3:], dim=1) * 0.1      # LR:887-890
        quat = quat_from_euler_xyz(cand[:, 3], cand[:, 4], cand[:, 5])                                           # LR:892
        pts = torch.zeros(2 * n_pts, 3)
        pts[:, :2] = points_xy.repeat(2, 1)
        pts[:, 2] = measured_heights[e].reshape(-1)                                                               # LR:893-894
        d = pts[:, None, :] - cand[None, :, :3]                                                                   # LR:895-896
        r = quat_rotate_inverse(yaw_quat(quat).repeat(d.shape[0], 1), d.reshape(-1, 3)).view(d.shape)             # LR:897
        outside = torch.norm(r / torch.tensor(ROBOT_SIZE)[None, None, :], dim=-1) > 1.0                           # LR:898
        valid = torch.all(outside, dim=0)                                                                         # LR:899
        if not bool(valid.any()):
            local[e] = target_poses[e]                                                                            # LR:910
            chosen[e] = -1
            continue
        order = torch.argsort(key, dim=-1, stable=True) if stable else torch.argsort(key, dim=-1)
        c = int(order[valid[order]][0])                                                                           # LR:901
        key_of[e] = key[c]
        t = cand[c].clone()
        t[:2] = quat_apply(yaw_quat(base_quat[[e]]), t[None, :3])[0, :2] + base_pos[e, :2]                        # LR:905
        t[3:] = wrap_to_pi(t[3:] + base_rot[e])                                                                   # LR:906
        local[e] = t
        chosen[e] = c
    lin, rot = relative_pose(local, base_pos, base_quat)                                                          # LR:914-916
    return dict(local_target_poses=local, plan_length_buf=pl, local_relative_linear=lin, local_relative_rotation=rot, chosen=chosen,
                key_of_chosen=key_of, replan=replan)
