"""CPU restatement of the cfg 4 kernels of csrc/lt_vel.cu (go1_gym/envs/base/legged_robot_velocity_tracking.py, LRV below).
TEST INFRASTRUCTURE ONLY: only tests/ may import it.  Stateless torch functions over dicts of tensors named like the reference's
attributes; pinned step by step against tests/golden/vel_cfg4_64.npz (a recording of the unmodified env, oracle/gen_golden_vel.py)
in tests/test_vel_env.py, and used there as the comparator for the CUDA kernels at sizes the recording does not have."""
import torch

from oracle.port import quat_rotate_inverse


def quat_apply_yaw(quat, vec):
    """go1_gym/utils/math_utils.py:12-16 with isaacgym's normalize / quat_apply (planner.py:17-23)."""
    q = quat.clone().view(-1, 4)
    q[:, :2] = 0.0
    q = q / q.norm(p=2, dim=-1).clamp(min=1e-9).unsqueeze(-1)
    shape = vec.shape
    b = vec.reshape(-1, 3)
    xyz = q[:, :3]
    t = xyz.cross(b, dim=-1) * 2
    return (b + q[:, 3:] * t + xyz.cross(t, dim=-1)).view(shape)


def prepare(root_states, rigid_body_state, contact_forces, gravity_vec, episode_length_buf, feet_indices, termination_indices,
            max_episode_length, use_terminal_body_height, terminal_body_height, measured_heights=0.0):
    """LRV:117-130 + check_termination (LRV:156-167).  -> dict"""
    N = root_states.shape[0]
    ep = episode_length_buf + 1
    q = root_states[:, 3:7]
    out = dict(episode_length_buf=ep, base_pos=root_states[:, 0:3].clone(), base_quat=q.clone(),
               base_lin_vel=quat_rotate_inverse(q, root_states[:, 7:10]), base_ang_vel=quat_rotate_inverse(q, root_states[:, 10:13]),
               projected_gravity=quat_rotate_inverse(q, gravity_vec))
    rb = rigid_body_state.view(N, -1, 13)
    out["foot_velocities"], out["foot_positions"] = rb[:, feet_indices, 7:10].clone(), rb[:, feet_indices, 0:3].clone()
    reset = torch.any(torch.norm(contact_forces[:, termination_indices, :], dim=-1) > 1.0, dim=1)
    time_out = ep > max_episode_length
    reset = reset | time_out
    low = torch.zeros(N, dtype=torch.bool)
    if use_terminal_body_height:
        mh = measured_heights if isinstance(measured_heights, torch.Tensor) else torch.zeros(N, 1)
        low = torch.mean(root_states[:, 2].unsqueeze(1) - mh, dim=1) < terminal_body_height
        reset = torch.logical_or(low, reset)
    out.update(reset_buf=reset, time_out_buf=time_out, body_height_buf=low)
    return out


def get_heights(root_states, height_points, height_samples, border_size, horizontal_scale, vertical_scale):
    """LRV:1806-1824.  height_points [P,3] (the same grid for every env), height_samples int16 [rows, cols]."""
    N, P = root_states.shape[0], height_points.shape[0]
    pts = quat_apply_yaw(root_states[:, 3:7].repeat(1, P), height_points.unsqueeze(0).repeat(N, 1, 1)) + root_states[:, :3].unsqueeze(1)
    pts = pts + border_size
    pts = (pts / horizontal_scale).long()
    px = torch.clip(pts[:, :, 0].view(-1), 0, height_samples.shape[0] - 2)
    py = torch.clip(pts[:, :, 1].view(-1), 0, height_samples.shape[1] - 2)
    h = torch.min(torch.min(height_samples[px, py], height_samples[px + 1, py]), height_samples[px, py + 1])
    return h.view(N, -1) * vertical_scale


def compute_reward(term_values, names, scales, episode_sums, command_sums, base_lin_vel, base_ang_vel, commands, only_positive_rewards,
                   ji22, sigma_rew_neg):
    """LRV:281-318.  term_values {name: [N]} unscaled; scales {name: float} (x dt); episode_sums / command_sums dicts updated in place.
    -> (rew_buf, rew_buf_pos, rew_buf_neg)"""
    N = base_lin_vel.shape[0]
    rew_buf, pos, neg = torch.zeros(N), torch.zeros(N), torch.zeros(N)
    for n in names:
        rew = term_values[n] * scales[n]
        rew_buf += rew
        if torch.sum(rew) >= 0:
            pos += rew
        elif torch.sum(rew) <= 0:
            neg += rew
        episode_sums[n] += rew
        if n in ("tracking_contacts_shaped_force", "tracking_contacts_shaped_vel"):
            command_sums[n] += scales[n] + rew
        else:
            command_sums[n] += rew
    if only_positive_rewards:
        rew_buf[:] = torch.clip(rew_buf[:], min=0.0)
    elif ji22:
        rew_buf[:] = pos[:] * torch.exp(neg[:] / sigma_rew_neg)
    episode_sums["total"] += rew_buf
    command_sums["lin_vel_raw"] += base_lin_vel[:, 0]
    command_sums["ang_vel_raw"] += base_ang_vel[:, 2]
    command_sums["lin_vel_residual"] += (base_lin_vel[:, 0] - commands[:, 0]) ** 2
    command_sums["ang_vel_residual"] += (base_ang_vel[:, 2] - commands[:, 2]) ** 2
    command_sums["ep_timesteps"] += 1
    return rew_buf, pos, neg


def compute_observations(c, projected_gravity, commands, commands_scale, dof_pos, dof_vel, default_dof_pos, actions, last_actions,
                         gait_indices, clock_inputs, noise_uniforms, noise_scale_vec, friction_coeffs, restitutions):
    """LRV:320-426 for the switches cfg 4 has, + the clip of step() (LRV:102-105).  c: the recorded cfg dict.  -> (obs, privileged_obs)"""
    parts = [projected_gravity]
    if c["observe_command"]:
        parts.append(commands * commands_scale)
    parts += [(dof_pos - default_dof_pos) * c["obs_scale_dof_pos"], dof_vel * c["obs_scale_dof_vel"], actions]
    if c["observe_two_prev_actions"]:
        parts.append(last_actions)
    if c["observe_timing_parameter"]:
        parts.append(gait_indices.unsqueeze(1))
    if c["observe_clock_inputs"]:
        parts.append(clock_inputs)
    obs = torch.cat(parts, dim=-1)
    if c["add_noise"]:
        obs = obs + (2 * noise_uniforms - 1) * noise_scale_vec
    priv = []
    for flag, val, rng in (("priv_observe_friction", friction_coeffs, c["friction_range"]), ("priv_observe_restitution", restitutions, c["restitution_range"])):
        if c[flag]:
            scale, shift = 2.0 / (rng[1] - rng[0]), (rng[1] + rng[0]) / 2.0
            priv.append((val.unsqueeze(1) - shift) * scale)
    priv = torch.cat(priv, dim=1) if priv else torch.empty(obs.shape[0], 0)
    clip = c["clip_observations"]
    return torch.clip(obs, -clip, clip), torch.clip(priv, -clip, clip)
