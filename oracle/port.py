"""ORACLE (test infrastructure, not product): CPU restatement of the reference's per-env step path.

Plain torch-CPU fp32, written operation for operation in the order the reference executes, so that
float results agree with the reference to rounding and masks / indices agree bit for bit.  Every method
cites the reference lines it restates; LR = go1_gym/envs/base/legged_robot_trajectory_tracking.py.

Pinning: tests/test_oracle_golden.py checks this port against the fixtures in tests/golden/, which were
produced by running the UNMODIFIED reference under oracle/ref_harness.py (script: oracle/gen_golden.py).
The reference ships no tests or golden vectors of its own and its quaternion helpers
live in the un-vendored `isaacgym` package (restated in oracle/shims/isaacgym/torch_utils.py), so at that
third-party boundary parity is UNPINNED; everything inside /root/reference is pinned by execution.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module.  The product (legged_tracking_b200/) never does: it fails loudly without its CUDA library.

Differences from the reference that are deliberate and do not change results:
  * random numbers come from an explicit per-env slot array (legged_tracking_b200/rng_slots.py) instead
    of torch.rand on compacted lists;
  * the per-env terrain copy `env_height_samples[N,2,R,C]` (LR:1827-1840) is replaced by a tile table
    `[tiles,2,R,C]` + `env_tile[N]` (same values);
  * `faithful_gather=True` restates stack_heights (LR:1967-1970) as the reference's per-env loop (used for
    the CPU baseline timing); False uses one vectorised gather (same values).
"""
import math

import numpy as np
import torch

from legged_tracking_b200 import rng_slots as RS

TWO_PI = 2 * np.pi
EPSILON = 1e-6          # reward_crawling.py:7


# ----------------------------------------------------------------------------- quaternion helpers
def quat_rotate_inverse(q, v):
    """isaacgym.torch_utils.quat_rotate_inverse (call sites LR:132-134)."""
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(-1, 1, 3), v.view(-1, 3, 1)).squeeze(-1) * 2.0
    return a - b + c


def normalize(x, eps=1e-9):
    return x / x.norm(p=2, dim=-1).clamp(min=eps, max=None).unsqueeze(-1)


def quat_apply(a, b):
    """go1_gym/utils/planner.py:17-23"""
    xyz = a[:, :3]
    t = xyz.cross(b, dim=-1) * 2
    return b + a[:, 3:] * t + xyz.cross(t, dim=-1)


def wrap_to_pi(angles):
    """go1_gym/utils/math_utils.py:20-23 (in place)"""
    angles %= TWO_PI
    angles -= TWO_PI * (angles > np.pi)
    return angles


def get_euler_xyz(q):
    """go1_gym/utils/planner.py:47-66"""
    qx, qy, qz, qw = 0, 1, 2, 3
    sinr_cosp = 2.0 * (q[:, qw] * q[:, qx] + q[:, qy] * q[:, qz])
    cosr_cosp = q[:, qw] * q[:, qw] - q[:, qx] * q[:, qx] - q[:, qy] * q[:, qy] + q[:, qz] * q[:, qz]
    roll = torch.atan2(sinr_cosp, cosr_cosp)
    sinp = 2.0 * (q[:, qw] * q[:, qy] - q[:, qz] * q[:, qx])
    half_pi = torch.full_like(sinp, np.pi / 2.0)
    pitch = torch.where(torch.abs(sinp) >= 1, torch.abs(half_pi) * torch.sign(sinp), torch.asin(sinp))
    siny_cosp = 2.0 * (q[:, qw] * q[:, qz] + q[:, qx] * q[:, qy])
    cosy_cosp = q[:, qw] * q[:, qw] + q[:, qx] * q[:, qx] - q[:, qy] * q[:, qy] - q[:, qz] * q[:, qz]
    yaw = torch.atan2(siny_cosp, cosy_cosp)
    return roll % TWO_PI, pitch % TWO_PI, yaw % TWO_PI


def quaternion_to_roll_pitch_yaw(quat):
    """go1_gym/utils/math_utils.py:40-46"""
    roll, pitch, yaw = get_euler_xyz(quat)
    return wrap_to_pi(torch.stack([roll, pitch, yaw], dim=-1))


def quat_apply_yaw_inverse(quat, vec):
    """go1_gym/utils/math_utils.py:57-61"""
    quat_yaw = quat.clone().view(-1, 4)
    quat_yaw[:, :2] = 0.
    quat_yaw = normalize(quat_yaw)
    return quat_rotate_inverse(quat_yaw, vec)


def quat_from_euler_xyz(roll, pitch, yaw):
    """go1_gym/utils/planner.py:25-38"""
    cy, sy = torch.cos(yaw * 0.5), torch.sin(yaw * 0.5)
    cr, sr = torch.cos(roll * 0.5), torch.sin(roll * 0.5)
    cp, sp = torch.cos(pitch * 0.5), torch.sin(pitch * 0.5)
    qw = cy * cr * cp + sy * sr * sp
    qx = cy * sr * cp - sy * cr * sp
    qy = cy * cr * sp + sy * sr * cp
    qz = sy * cr * cp - cy * sr * sp
    return torch.stack([qx, qy, qz, qw], dim=-1)


def quat_from_yaw(yaw):
    """quat_from_angle_axis(yaw[K,1], z)[:, 0, :] as called at LR:1043: normalise([0,0,1]*sin(y/2), cos(y/2))."""
    theta = (yaw / 2).unsqueeze(-1)                                   # [K,1,1]
    axis = normalize(torch.tensor([0., 0., 1.]))
    xyz = axis * theta.sin()                                           # [K,1,3]
    w = theta.cos()
    return normalize(torch.cat([xyz, w], dim=-1))[:, 0, :]


def scale_shift(rng):
    """go1_gym/utils/math_utils.py:35-38"""
    return 2. / (rng[1] - rng[0]), (rng[1] + rng[0]) / 2.


STATE_F32 = [
    # name, trailing shape (per env)
    ("actions", "A"), ("last_actions", "A"), ("last_last_actions", "A"),
    ("joint_pos_target", "D"), ("last_joint_pos_target", "D"), ("last_last_joint_pos_target", "D"),
    ("last_dof_vel", "D"), ("last_root_vel", 6), ("torques", "D"),
    ("joint_pos_err_last", "D"), ("joint_pos_err_last_last", "D"), ("joint_vel_last", "D"), ("joint_vel_last_last", "D"),
    ("motor_strengths", "D"), ("motor_offsets", "D"), ("Kp_factors", "D"), ("Kd_factors", "D"),
    ("friction_coeffs", 1), ("restitutions", 1), ("payloads", 1), ("com_displacements", 3), ("gravities", 3),
    ("base_lin_vel", 3), ("base_ang_vel", 3), ("projected_gravity", 3), ("base_rotation", 3),
    ("relative_linear", 3), ("relative_rotation", 3), ("local_relative_linear", 3), ("local_relative_rotation", 3),
    ("target_poses", 6), ("feet_air_time", "F"), ("rew_buf", 0), ("rew_buf_pos", 0), ("rew_buf_neg", 0),
]
STATE_I64 = ["episode_length_buf", "curr_pose_index", "collision_count", "collision_env_buf"]
STATE_BOOL = ["reset_buf", "time_out_buf", "body_height_buf", "switched_buf", "reached_buf", "plan_buf",
              "reached_env_buf", "time_outs_latched"]


class OraclePort:
    """Stateful restatement of LeggedRobot.step for the trajectory-tracking env (LR:64-169)."""

    def __init__(self, flat, terrain, actuator, faithful_gather=False, num_threads=None):
        self.flat = flat
        f = flat
        self.N = N = int(f["num_envs"])
        self.A = int(f["num_actors"])
        self.B = int(f["num_bodies"])
        self.D = int(f["num_dof"])
        self.faithful_gather = faithful_gather
        t = lambda x, dt=torch.float32: torch.as_tensor(np.asarray(x), dtype=dt)    # noqa: E731
        # ---- constants (LR:1289-1305, 680-708, 1647-1664)
        self.default_dof_pos = t(f["default_dof_pos"]).view(1, -1)
        self.torque_limits = t(f["torque_limits"])
        self.p_gains = t(f["p_gains"])
        self.d_gains = t(f["d_gains"])
        self.dof_pos_limits = torch.stack([t(f["dof_pos_limits_lower"]), t(f["dof_pos_limits_upper"])], dim=1)
        self.feet_indices = t(f["feet_indices"], torch.long)
        self.penalised_contact_indices = t(f["penalised_contact_indices"], torch.long)
        self.termination_contact_indices = t(f["termination_contact_indices"], torch.long)
        self.base_init_state = t(f["base_init_state"])
        self.camera_offset = t(f["camera_offset"])
        self.noise_scale_vec = t(f["noise_scale_vec"])
        self.gravity_vec = torch.tensor([0., 0., -1.]).repeat(N, 1)        # LR:1221
        self.reward_scales = dict(f["reward_scales"])                        # already * dt (LR:1385)
        self.reward_names = list(f["reward_names"])
        self.common_step_counter = 0
        # ---- terrain (LR:1808-1847)
        self.plane = f["mesh_type"] == "plane"
        self.env_origins = t(terrain["env_origins"]).view(N, 3)
        if not self.plane:
            self.tile_heights = t(terrain["tile_heights"])                   # [T,2,R,C], metres
            self.env_tile = t(terrain["env_tile"], torch.long)               # [N]
            self.env_terrain_origin = t(terrain["env_terrain_origin"]).view(N, 3)
        px = t(f["measured_points_x"])
        py = t(f["measured_points_y"])
        self.nx, self.ny = len(px), len(py)
        gx, gy = torch.meshgrid(px, py, indexing="ij")                       # LR:1908-1916
        self.height_points = torch.zeros(N, self.nx, self.ny, 3)
        self.height_points[..., 0] = gx
        self.height_points[..., 1] = gy
        # ---- actuator net (LR:1307-1322; resources/actuator_nets/unitree_go1.pt)
        self.act_w = [t(actuator[k]) for k in ("w0", "b0", "w1", "b1", "w2", "b2")]
        # ---- sim tensors in Isaac Gym layout (LR:1173-1201)
        self.root_states = torch.zeros(N * self.A, 13)
        self.root_states[:, 6] = 1.0
        self.dof_state = torch.zeros(N * self.D, 2)
        self.contact_forces = torch.zeros(N, self.B, 3)
        self.rigid_body_state = torch.zeros(N, self.B, 13)
        self._alloc_state()
        self.last_px = None
        self.last_py = None

    # views, as LR:1189-1192
    @property
    def robot_root(self):
        return self.root_states.view(self.N, self.A, 13)[:, 0, :]

    @property
    def dof_pos(self):
        return self.dof_state.view(self.N, self.D, 2)[..., 0]

    @property
    def dof_vel(self):
        return self.dof_state.view(self.N, self.D, 2)[..., 1]

    def _alloc_state(self):
        N, f = self.N, self.flat
        dims = dict(A=f["num_actions"], D=self.D, F=len(f["feet_indices"]))
        for name, d in STATE_F32:
            d = dims.get(d, d)
            setattr(self, name, torch.zeros(N, d) if d else torch.zeros(N))
        for name in STATE_I64:
            setattr(self, name, torch.zeros(N, dtype=torch.long))
        for name in STATE_BOOL:
            setattr(self, name, torch.zeros(N, dtype=torch.bool))
        self.motor_strengths.fill_(1.0)
        self.Kp_factors.fill_(1.0)
        self.Kd_factors.fill_(1.0)
        L = f["lag_timesteps"] + 1
        self.lag_buffer = [torch.zeros(N, self.D) for _ in range(L)]          # LR:1199
        self.trajectories = torch.zeros(N, f["traj_length"], 6)               # LR:1255
        self.last_contacts = torch.zeros(N, dims["F"], dtype=torch.bool)
        self.camera_pitch_angle = torch.zeros(N)
        sum_names = list(self.reward_scales.keys()) + ["total", "total_pos", "total_neg"]      # LR:1400-1408
        self.episode_sums = {k: torch.zeros(N) for k in sum_names}
        self.command_sums = {k: torch.zeros(N) for k in list(self.reward_scales.keys()) +
                             ["lin_vel_raw", "ang_vel_raw", "lin_vel_residual", "ang_vel_residual", "ep_timesteps"]}
        self.obs_buf = torch.zeros(N, f["num_obs"])
        self.privileged_obs_buf = torch.zeros(N, f["num_privileged_obs"])
        self.obs_history = torch.zeros(N, f["num_obs"] * f["obs_history_length"])
        self.measured_heights = None
        self.finished_episodes = []     # per step: dict(env_ids, sums...) -- what the reference pushes into extras deques

    # ------------------------------------------------------------------ state import / export
    STATE_KEYS = ([n for n, _ in STATE_F32] + STATE_I64 + STATE_BOOL +
                  ["trajectories", "last_contacts", "root_states", "dof_state", "obs_history", "camera_pitch_angle"])

    def export_state(self):
        out = {k: getattr(self, k).clone() for k in self.STATE_KEYS}
        out["lag_buffer"] = torch.stack(self.lag_buffer, dim=0)
        for k, v in self.episode_sums.items():
            out["episode_sums/" + k] = v.clone()
        for k, v in self.command_sums.items():
            out["command_sums/" + k] = v.clone()
        out["common_step_counter"] = torch.tensor(self.common_step_counter)
        return out

    def load_state(self, st):
        for k in self.STATE_KEYS:
            if k in st:
                getattr(self, k).copy_(torch.as_tensor(st[k]).view_as(getattr(self, k)))
        if "lag_buffer" in st:
            lb = torch.as_tensor(st["lag_buffer"])
            self.lag_buffer = [lb[i].clone().float() for i in range(lb.shape[0])]
        for k in self.episode_sums:
            if "episode_sums/" + k in st:
                self.episode_sums[k].copy_(torch.as_tensor(st["episode_sums/" + k]))
        for k in self.command_sums:
            if "command_sums/" + k in st:
                self.command_sums[k].copy_(torch.as_tensor(st["command_sums/" + k]))
        if "common_step_counter" in st:
            self.common_step_counter = int(st["common_step_counter"])

    # ------------------------------------------------------------------ a3: torques
    def actuator_network(self, xs):
        """Sequential(Linear(6,32), softsign, Linear(32,32), softsign, Linear(32,1)) -- LR:1311-1320."""
        w0, b0, w1, b1, w2, b2 = self.act_w
        h = torch.nn.functional.softsign(torch.nn.functional.linear(xs, w0, b0))
        h = torch.nn.functional.softsign(torch.nn.functional.linear(h, w1, b1))
        return torch.nn.functional.linear(h, w2, b2)

    def compute_torques(self, actions):
        """LR:957-996"""
        f = self.flat
        actions_scaled = actions[:, :12] * f["action_scale"]
        actions_scaled[:, [0, 3, 6, 9]] *= f["hip_scale_reduction"]
        if f["randomize_lag_timesteps"]:
            self.lag_buffer = self.lag_buffer[1:] + [actions_scaled.clone()]
            self.joint_pos_target = self.lag_buffer[0] + self.default_dof_pos
        else:
            self.joint_pos_target = actions_scaled + self.default_dof_pos
        if f["control_type"] == "actuator_net":
            joint_pos_err = self.dof_pos - self.joint_pos_target + self.motor_offsets
            joint_vel = self.dof_vel
            xs = torch.cat((joint_pos_err.unsqueeze(-1), self.joint_pos_err_last.unsqueeze(-1),
                            self.joint_pos_err_last_last.unsqueeze(-1), joint_vel.unsqueeze(-1),
                            self.joint_vel_last.unsqueeze(-1), self.joint_vel_last_last.unsqueeze(-1)), dim=-1)
            torques = self.actuator_network(xs.view(self.N * 12, 6)).view(self.N, 12)
            self.joint_pos_err_last_last = torch.clone(self.joint_pos_err_last)
            self.joint_pos_err_last = torch.clone(joint_pos_err)
            self.joint_vel_last_last = torch.clone(self.joint_vel_last)
            self.joint_vel_last = torch.clone(joint_vel)
        elif f["control_type"] == "P":
            torques = self.p_gains * self.Kp_factors * (
                self.joint_pos_target - self.dof_pos + self.motor_offsets) - self.d_gains * self.Kd_factors * self.dof_vel
        else:
            raise NameError(f["control_type"])
        torques = torques * self.motor_strengths
        return torch.clip(torques, -self.torque_limits, self.torque_limits)

    # ------------------------------------------------------------------ a2: step
    def step(self, actions, sim, uniforms=None, noise_uniforms=None, gravity_uniform=None):
        """LR:64-112.  `sim` supplies what PhysX would: sim["dof_state"][s] after substep s, and
        sim["root"] (robot rows), sim["contact"], sim["rigid"] for the finished step."""
        f = self.flat
        with torch.no_grad():
            self.actions = torch.clip(torch.as_tensor(actions, dtype=torch.float32), -f["clip_actions"], f["clip_actions"])
            self.all_torques = []
            for s in range(f["decimation"]):
                self.torques = self.compute_torques(self.actions).view(self.torques.shape)
                self.all_torques.append(self.torques.clone())
                self.dof_state.copy_(torch.as_tensor(sim["dof_state"][s]).view_as(self.dof_state))   # PhysX substep
            self.robot_root.copy_(torch.as_tensor(sim["root"]))
            self.contact_forces.copy_(torch.as_tensor(sim["contact"]).view_as(self.contact_forces))
            self.rigid_body_state.copy_(torch.as_tensor(sim["rigid"]).view_as(self.rigid_body_state))
            self.post_physics_step(uniforms, noise_uniforms, gravity_uniform)
            clip_obs = f["clip_observations"]
            self.obs_buf = torch.clip(self.obs_buf, -clip_obs, clip_obs)
            self.privileged_obs_buf = torch.clip(self.privileged_obs_buf, -clip_obs, clip_obs)
            # HistoryWrapper.step (go1_gym/envs/wrappers/history_wrapper.py:18-24)
            self.obs_history = torch.cat((self.obs_history[:, f["num_obs"]:], self.obs_buf), dim=-1)
        return self.obs_buf, self.privileged_obs_buf, self.rew_buf, self.reset_buf

    # ------------------------------------------------------------------ a4: post_physics_step
    def post_physics_step(self, uniforms, noise_uniforms, gravity_uniform):
        """LR:114-169"""
        self.episode_length_buf += 1
        self.common_step_counter += 1
        root = self.robot_root
        base_quat = root[:, 3:7]
        self.base_lin_vel[:] = quat_rotate_inverse(base_quat, root[:, 7:10])
        self.base_ang_vel[:] = quat_rotate_inverse(base_quat, root[:, 10:13])
        self.projected_gravity[:] = quat_rotate_inverse(base_quat, self.gravity_vec)
        self.foot_positions = self.rigid_body_state[:, self.feet_indices, 0:3] - root[:, None, 0:3]
        self.post_physics_step_callback(uniforms, gravity_uniform)
        self.check_termination()
        self.compute_reward()
        env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()
        self.reset_env_ids = env_ids
        self.reset_idx(env_ids, uniforms)
        self.compute_observations(noise_uniforms)
        self.last_last_actions[:] = self.last_actions[:]
        self.last_actions[:] = self.actions[:]
        self.last_last_joint_pos_target[:] = self.last_joint_pos_target[:]
        self.last_joint_pos_target[:] = self.joint_pos_target[:]
        self.last_dof_vel[:] = self.dof_vel[:]
        self.last_root_vel[:] = self.robot_root[:, 7:13]
