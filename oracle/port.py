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
        self.plan_length_buf = torch.zeros(N, dtype=torch.long)               # LR:1214
        self.planned_target_poses = torch.zeros(N, 6)                         # local_target_poses, LR:1263
        self.planner_stable = False    # False = the reference's torch.argsort (unstable ties); True = the kernel's lowest-index rule
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
        out["plan_length_buf"] = self.plan_length_buf.clone()
        out["local_target_poses"] = self.planned_target_poses.clone()
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
        if "plan_length_buf" in st:
            self.plan_length_buf = torch.as_tensor(st["plan_length_buf"]).long().clone()
        if "local_target_poses" in st:
            self.planned_target_poses = torch.as_tensor(st["local_target_poses"]).float().clone()

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
        self.update_curriculum()
        self.last_last_actions[:] = self.last_actions[:]
        self.last_actions[:] = self.actions[:]
        self.last_last_joint_pos_target[:] = self.last_joint_pos_target[:]
        self.last_joint_pos_target[:] = self.joint_pos_target[:]
        self.last_dof_vel[:] = self.dof_vel[:]
        self.last_root_vel[:] = self.robot_root[:, 7:13]

    # ------------------------------------------------------------------ a5: heights
    def get_heights(self):
        """LR:1918-1965 (+ stack_heights LR:1967-1970).  Returns [N,2,nx,ny]."""
        f = self.flat
        N = self.N
        if self.plane:
            top = torch.ones(N, self.nx, self.ny)
            bottom = torch.zeros(N, self.nx, self.ny)
            return torch.stack([top, bottom], dim=1)
        # F3: the yaw-rotated points are discarded; F5: pitch is the previous step's (base_rotation not yet refreshed)
        self.camera_pitch_angle = self.base_rotation[:, 1].clone()
        points = self.height_points.clone()
        points[..., :2] += self.robot_root[:, None, None, :2]
        if f["camera_zero"]:
            camera_offset_xy = self.camera_offset[None, :2] * torch.cos(self.camera_pitch_angle[:, None])
            points[..., :2] += camera_offset_xy[:, None, None]
        points -= self.env_terrain_origin[:, None, None, :]
        points = (points / f["horizontal_scale"]).long()
        px = torch.clip(points[..., 0], 0, f["tile_rows"] - 2)
        py = torch.clip(points[..., 1], 0, f["tile_cols"] - 2)
        self.last_px, self.last_py = px, py
        if self.faithful_gather:
            samples = self.tile_heights
            tile = self.env_tile
            heights = torch.stack([samples[int(tile[e]), :, px[e], py[e]] for e in range(N)], dim=0)
        else:
            heights = self.tile_heights[self.env_tile[:, None, None], :, px, py].permute(0, 3, 1, 2)
        return heights

    # ------------------------------------------------------------------ a6: goal-relative pose
    def compute_relative_target_pose(self, target_poses):
        """LR:922-932"""
        root = self.robot_root
        relative_linear = target_poses[:, :3] - root[:, :3]
        relative_linear = quat_apply_yaw_inverse(root[:, 3:7], relative_linear)
        self.base_rotation = quaternion_to_roll_pitch_yaw(root[:, 3:7])
        relative_rotation = target_poses[:, 3:] - self.base_rotation
        relative_rotation = wrap_to_pi(relative_rotation)
        return relative_linear, relative_rotation

    # ------------------------------------------------------------------ a7: callback
    def randomize_dof_props(self, env_ids, uniforms, base):
        """LR:744-764.  `base` = first slot (rng_slots.INTERVAL_STRENGTH or RESET_STRENGTH)."""
        f = self.flat
        u = uniforms[env_ids]
        if f["randomize_motor_strength"]:
            lo, hi = f["motor_strength_range"]
            self.motor_strengths[env_ids, :] = u[:, base + 0].unsqueeze(1) * (hi - lo) + lo
        if f["randomize_motor_offset"]:
            lo, hi = f["motor_offset_range"]
            self.motor_offsets[env_ids, :] = u[:, base + 1: base + 13] * (hi - lo) + lo
        if f["randomize_Kp_factor"]:
            lo, hi = f["Kp_factor_range"]
            self.Kp_factors[env_ids, :] = u[:, base + 13].unsqueeze(1) * (hi - lo) + lo
        if f["randomize_Kd_factor"]:
            lo, hi = f["Kd_factor_range"]
            self.Kd_factors[env_ids, :] = u[:, base + 14].unsqueeze(1) * (hi - lo) + lo

    def randomize_gravity(self, external_force=None, gravity_uniform=None):
        """LR:645-660"""
        f = self.flat
        if external_force is not None:
            self.gravities[:, :] = external_force.unsqueeze(0)
        elif f["randomize_gravity"]:
            lo, hi = f["gravity_range"]
            ext = torch.as_tensor(gravity_uniform, dtype=torch.float32) * (hi - lo) + lo
            self.gravities[:, :] = ext.unsqueeze(0)
        gravity = self.gravities[0, :] + torch.Tensor([0, 0, -9.8])
        self.gravity_vec[:, :] = gravity.unsqueeze(0) / torch.norm(gravity)

    def update_curriculum(self):
        """LR:171-183 (the cl_fix_target branch, LR:186-196, only rewrites logging deques and commands.x_mean)."""
        f = self.flat
        for key in ("exploration_lin", "exploration_yaw"):
            if key in self.reward_scales:
                if self.common_step_counter > f["exploration_steps"]:
                    self.reward_scales[key] -= f["reward_scales_cfg"][key] * f["dt"] / f["exploration_steps"]
                    self.reward_scales[key] = max(self.reward_scales[key], 0)

    def post_physics_step_callback(self, uniforms, gravity_uniform):
        """LR:774-848"""
        f = self.flat
        if f["observe_heights"]:
            self.measured_heights = self.get_heights()
        idx = self.curr_pose_index
        self.target_poses = self.trajectories[torch.arange(self.N), idx, :]                 # LR:854
        self.relative_linear, self.relative_rotation = self.compute_relative_target_pose(self.target_poses)
        if f.get("sampling_based_planning"):                                                 # LR:856-916
            from oracle import planner_port
            root = self.robot_root
            out = planner_port.plan(self.target_poses, root[:, :3].clone(), root[:, 3:7].clone(), self.height_points[0].reshape(-1, 3)[:, :2],
                                    self.measured_heights.reshape(self.N, 2, -1), torch.tensor(f["candidate_target_poses"], dtype=torch.float64),
                                    self.episode_length_buf, self.plan_buf.long(), self.plan_length_buf, self.planned_target_poses,
                                    plan_interval=int(f["plan_interval"]), stable=self.planner_stable)
            self.plan_length_buf = out["plan_length_buf"]
            self.planned_target_poses = out["local_target_poses"]
            self.local_target_poses = self.planned_target_poses
            self.local_relative_linear, self.local_relative_rotation = out["local_relative_linear"], out["local_relative_rotation"]
            self.planner_chosen = out["chosen"]
            self.replan = out["replan"]                                                       # LR:861
        else:
            self.local_target_poses = self.target_poses                                      # LR:918-920
            self.local_relative_linear, self.local_relative_rotation = \
                self.compute_relative_target_pose(self.local_target_poses)
        # _set_the_target_pose_visual LR:934-947
        arrow = self.root_states.view(self.N, self.A, 13)[:, self.A - 1, :]
        arrow[:, :3] = self.local_target_poses[:, :3]
        arrow[:, 3:7] = quat_from_euler_xyz(self.local_target_poses[:, 3], self.local_target_poses[:, 4],
                                            self.local_target_poses[:, 5])
        # interval DR LR:822-824
        env_ids = (self.episode_length_buf % int(f["rand_interval"]) == 0).nonzero(as_tuple=False).flatten()
        if len(env_ids) > 0:
            self.randomize_dof_props(env_ids, uniforms, RS.INTERVAL_STRENGTH)
        # gravity LR:826-830
        if self.common_step_counter % int(f["gravity_rand_interval"]) == 0:
            self.randomize_gravity(gravity_uniform=gravity_uniform)
        if int(self.common_step_counter - f["gravity_rand_duration"]) % int(f["gravity_rand_interval"]) == 0:
            self.randomize_gravity(external_force=torch.tensor([0, 0, 0]))
        # waypoint switching LR:836-848
        if f["switch_upon_reach"]:
            self.switched_buf = torch.linalg.norm(self.relative_linear[:, :2], dim=1) < f["switch_dist"]
        else:
            self.switched_buf = (self.episode_length_buf % f["switch_interval"] == 0)
        env_ids = self.switched_buf.nonzero(as_tuple=False).flatten()
        self.curr_pose_index[env_ids] += 1
        self.curr_pose_index[env_ids] = torch.clip(self.curr_pose_index[env_ids], max=f["traj_length"] - 1)
        self.reached_buf = torch.logical_and(self.switched_buf, self.curr_pose_index == f["traj_length"] - 1)
        plan_buf_linear = torch.linalg.norm(self.local_relative_linear[:, :2], dim=1) < f["switch_dist"]
        plan_buf_yaw = torch.abs(self.local_relative_rotation[:, 2]) < f["switch_yaw"]
        self.plan_buf = torch.logical_and(plan_buf_linear, plan_buf_yaw)
        self.collision_count += torch.sum(
            (torch.norm(self.contact_forces[:, self.penalised_contact_indices, :], dim=-1) > 0.1).long(), dim=-1)

    # ------------------------------------------------------------------ a8: termination
    def check_termination(self):
        """LR:198-216"""
        f = self.flat
        self.reset_buf = torch.any(
            torch.norm(self.contact_forces[:, self.termination_contact_indices, :], dim=-1) > 1., dim=1)
        self.time_out_buf = self.episode_length_buf > f["max_episode_length"]
        self.reset_buf |= self.time_out_buf
        if f["use_terminal_body_height"]:
            self.body_height_buf = torch.mean(self.robot_root[:, 2].unsqueeze(1), dim=1) < f["terminal_body_height"]
            self.reset_buf = torch.logical_or(self.body_height_buf, self.reset_buf)
        if f["terminate_end_of_trajectory"]:
            reached = torch.logical_and(self.reached_buf, self.episode_length_buf > f["T_reach"])
            self.reset_buf = torch.logical_or(reached, self.reset_buf)
        if f["use_terminal_body_rotation"]:
            self.reset_buf = torch.logical_or(self.reset_buf, self.projected_gravity[:, 2] > 0.0)

    # ------------------------------------------------------------------ a9.1 / a9.2: reward terms
    # RewardsCrawling: go1_gym/envs/rewards/reward_crawling.py:18-123
    # TrajectoryTrackingRewards: go1_gym/envs/rewards/trajectory_tracking_reward.py:17-171
    def _target_lin_vel(self):
        f = self.flat
        target_linear_vel = self.relative_linear[:, :2]
        magnitude = torch.linalg.norm(target_linear_vel, dim=1, keepdim=True)
        target_linear_vel = target_linear_vel / (magnitude + EPSILON) * f["target_lin_vel"]
        target_linear_vel = target_linear_vel * (magnitude > f["lin_reaching_criterion"])
        return target_linear_vel, magnitude

    def _reward_dof_acc(self):
        return torch.sum(torch.square((self.last_dof_vel - self.dof_vel) / self.flat["dt"]), dim=1)

    def _reward_torques(self):
        return torch.sum(torch.square(self.torques), dim=1)

    def _reward_dof_pos_limits(self):
        out_of_limits = -(self.dof_pos - self.dof_pos_limits[:, 0]).clip(max=0.)
        out_of_limits += (self.dof_pos - self.dof_pos_limits[:, 1]).clip(min=0.)
        return torch.sum(out_of_limits, dim=1)

    def _reward_collision(self):
        return torch.sum(1. * (torch.norm(self.contact_forces[:, self.penalised_contact_indices, :], dim=-1) > 0.1), dim=1)

    def _reward_action_rate(self):
        return torch.sum(torch.square(self.last_actions - self.actions), dim=1)

    def _reward_base_height(self):
        return torch.square(self.robot_root[:, 2] - self.flat["base_height_target"])

    def _reward_ang_vel_xy(self):
        return torch.sum(torch.square(self.base_ang_vel[:, :2]), dim=1)

    def _reward_orientation(self):
        return torch.sum(torch.square(self.projected_gravity[:, :2]), dim=1)

    def _reward_large_vel(self):
        magnitude = torch.norm(self.base_lin_vel[:, :2], dim=1) > 0.5
        return torch.sum(torch.square(self.base_lin_vel[:, :2]), dim=1) * magnitude.float()

    def _reward_e2e(self):
        f = self.flat
        magnitude = torch.norm(self.relative_linear[:, :2], dim=1)
        if f["terminate_end_of_trajectory"]:
            return (magnitude < f["switch_dist"]) * f["max_episode_length"]
        reached = (magnitude < f["switch_dist"])
        after_t_reach = self.episode_length_buf > f["T_reach"]
        linear_vel_error = torch.sum(torch.square(self.base_lin_vel[:, :2]), dim=-1)
        return torch.exp(-linear_vel_error / f["tracking_sigma_lin"]) * reached.float() * after_t_reach.float()

    def _lin_vel_form(self, allow_prod):
        f = self.flat
        target_linear_vel, magnitude = self._target_lin_vel()
        form = f["lin_vel_form"]
        if form == "exp":
            err = torch.sum(torch.square(target_linear_vel - self.base_lin_vel[:, :2]), dim=-1)
            return torch.exp(-err / f["tracking_sigma_lin"])
        if form == "l1":
            return torch.sum(torch.abs(target_linear_vel - self.base_lin_vel[:, :2]), dim=-1)
        if form == "l2":
            return torch.sum(torch.square(target_linear_vel - self.base_lin_vel[:, :2]), dim=-1)
        if form == "prod" and allow_prod:
            base_vel_magnitude = torch.linalg.norm(self.base_lin_vel[:, :2], dim=1, keepdim=True)
            rew = target_linear_vel / f["target_lin_vel"] * self.base_lin_vel[:, :2] / (base_vel_magnitude + EPSILON)
            rew = torch.sum(rew, dim=-1)
            rew *= (base_vel_magnitude.flatten() > f["small_vel_threshold"])
            rew += torch.exp(-base_vel_magnitude.flatten() ** 2 / f["tracking_sigma_lin"]) * \
                (magnitude.flatten() < f["lin_reaching_criterion"])
            return rew
        raise NotImplementedError(form)

    def _reward_exploration_lin(self):
        return self._lin_vel_form(allow_prod=True)

    def _reward_reaching_linear_vel(self):
        return self._lin_vel_form(allow_prod=False)

    def _reward_exploration_yaw(self):
        f = self.flat
        target_angular_vel = self.relative_rotation[:, 2]
        magnitude = target_angular_vel.abs()
        target_angular_vel = target_angular_vel / (magnitude + EPSILON) * f["target_ang_vel"]
        target_angular_vel = target_angular_vel * (magnitude > f["ang_reaching_criterion"])
        ang_vel_error = torch.square(target_angular_vel - self.base_ang_vel[:, 2])
        return torch.exp(-ang_vel_error / f["tracking_sigma_ang"])

    _reward_reaching_yaw = _reward_exploration_yaw

    def _reward_reaching_z(self):
        return torch.square(self.relative_linear[:, 2])

    def _reward_reaching_roll(self):
        return torch.square(self.relative_rotation[:, 0])

    def _reward_reaching_pitch(self):
        return torch.square(self.relative_rotation[:, 1])

    def _reward_reaching_yaw_abs(self):
        return torch.square(self.relative_rotation[:, 2])

    def _reward_dof_vel(self):
        return torch.sum(torch.square(self.dof_vel), dim=1)

    def _reward_dof_pos(self):
        return torch.sum(torch.square(self.dof_pos - self.default_dof_pos), dim=1)

    def _reward_task_old(self):
        f = self.flat
        target_dist = 0.5
        r = target_dist / (target_dist + torch.norm(self.relative_linear[:, :2], dim=1)) / f["T_reach"]
        r = r * (self.episode_length_buf > f["T_reach"])
        return r

    def _reward_reach_goal(self):
        return self.reached_buf

    def _reward_reach_goal_t(self):
        return self.reached_buf * self.episode_length_buf.float()

    def _reward_reach_goal_T(self):
        return self.reached_buf * (self.episode_length_buf > self.flat["T_reach"]).float()

    def _reward_task(self):
        f = self.flat
        target_linear_vel, _ = self._target_lin_vel()
        err = torch.sum(torch.square(target_linear_vel - self.base_lin_vel[:, :2]), dim=-1)
        in_dist = torch.norm(self.relative_linear[:, :2], dim=1) < f["large_dist_threshold"]
        return torch.exp(-err / f["tracking_sigma_lin"]) * in_dist.float()

    def _reward_exploration(self):
        f = self.flat
        r = torch.sum(self.base_lin_vel[:, :2] * self.local_relative_linear[:, :2], dim=1)
        r /= (torch.norm(self.local_relative_linear[:, :2], dim=1) + EPSILON)
        r /= (torch.norm(self.base_lin_vel[:, :2], dim=1) + EPSILON)
        r *= (torch.norm(self.base_lin_vel[:, :2], dim=1) > f["small_vel_threshold"]).float()
        return r

    def _reward_reaching_local_goal(self):                                                    # trajectory_tracking_reward.py:95-97
        return torch.logical_and(self.plan_buf, self.replan)

    def _reward_stalling(self):
        f = self.flat
        small_vel = torch.norm(self.base_lin_vel[:, :2], dim=1) < f["small_vel_threshold"]
        large_dist = torch.norm(self.relative_linear[:, :2], dim=1) > f["large_dist_threshold"]
        return -(small_vel * large_dist).float()

    def _reward_linear_vel(self):
        return torch.norm(self.base_lin_vel[:, :3], dim=1) > 0.7

    def _reward_lin_vel_z(self):
        return torch.square(self.base_lin_vel[:, 2])

    def _reward_feet_air_time(self):
        contact = self.contact_forces[:, self.feet_indices, 2] > 1.
        contact_filt = torch.logical_or(contact, self.last_contacts)
        self.last_contacts = contact
        first_contact = (self.feet_air_time > 0.) * contact_filt
        self.feet_air_time += self.flat["dt"]
        rew = torch.sum((self.feet_air_time - 0.5) * first_contact, dim=1)
        self.feet_air_time *= ~contact_filt
        return rew

    def _reward_survive(self):
        return torch.ones(self.N)

    # ------------------------------------------------------------------ a9: reward sum
    def compute_reward(self, external=None):
        """LR:320-355"""
        f = self.flat
        self.rew_buf[:] = 0.
        self.rew_buf_pos[:] = 0.
        self.rew_buf_neg[:] = 0.
        self.term_values = {}
        for name in self.reward_names:
            if name == "external":
                val = external
            else:
                val = getattr(self, "_reward_" + name)()
            rew = val * self.reward_scales[name]
            self.term_values[name] = rew.clone() if torch.is_tensor(rew) else rew
            self.rew_buf += rew
            if torch.sum(rew) >= 0:
                self.rew_buf_pos += rew
            elif torch.sum(rew) <= 0:
                self.rew_buf_neg += rew
            self.episode_sums[name] += rew
            self.command_sums[name] += rew
        if f["only_positive_rewards"]:
            self.rew_buf[:] = torch.clip(self.rew_buf[:], min=0.)
        elif f["only_positive_rewards_ji22_style"]:
            self.rew_buf[:] = self.rew_buf_pos[:] * torch.exp(self.rew_buf_neg[:] / f["sigma_rew_neg"])
        self.episode_sums["total"] += self.rew_buf
        self.episode_sums["total_pos"] += self.rew_buf_pos
        self.episode_sums["total_neg"] += self.rew_buf_neg
        if "termination" in self.reward_scales:
            raise NotImplementedError("termination reward: no container in the reference defines _reward_termination")
        self.command_sums["ep_timesteps"] += 1

    # ------------------------------------------------------------------ a10: reset
    def trajectory_function(self, env_ids, uniforms):
        """go1_gym/envs/trajectories/trajectory_function.py:14-93"""
        f = self.flat
        fn = f["traj_function"]
        root = self.robot_root
        K = len(env_ids)
        L = f["traj_length"]
        if fn == "fixed_target":                                           # :14-26
            x = torch.arange(1, L + 1).repeat(K, 1) * f["base_x"]
            x += root[env_ids, 0:1]
            y = torch.arange(1, L + 1).repeat(K, 1) * f["base_y"]
            y += root[env_ids, 1:2]
            z = torch.zeros_like(x) + f["base_z"]
            yaw = torch.zeros_like(x) * f["base_yaw"]
            pitch = torch.zeros_like(x) + f["base_pitch"]
            roll = torch.zeros_like(x) + f["base_roll"]
            return torch.stack([x, y, z, roll, pitch, yaw], dim=2)
        u = uniforms[env_ids][:, RS.TRAJ:]
        if fn == "random_goal":                                            # :28-41
            x = (u[:, 0:1] - 0.5) * f["x_range"] + f["x_mean"]
            x += root[env_ids, 0:1]
            y = (u[:, 1:2] - 0.5) * f["y_range"] + f["y_mean"]
            y += root[env_ids, 1:2]
            z = torch.zeros_like(x) + f["base_z"]
            yaw = u[:, 2:3] * 2 * f["yaw_range"] - f["yaw_range"]
            pitch = torch.zeros_like(x)
            roll = torch.zeros_like(x)
            return torch.stack([x, y, z, roll, pitch, yaw], dim=2)
        if fn == "random_target":                                          # :70-93
            ni = f["num_interpolation"]
            nt = L // ni + 1
            blk = lambda i: u[:, i * nt:(i + 1) * nt]                      # noqa: E731
            x = blk(0) * 2 * f["x_range"] - f["x_range"]
            y = blk(1) * 2 * f["y_range"] - f["y_range"]
            z = blk(2) * 2 * f["z_range"] - f["z_range"]
            yaw = blk(3) * 2 * f["yaw_range"] - f["yaw_range"]
            pitch = blk(4) * 2 * f["pitch_range"] - f["pitch_range"]
            roll = blk(5) * 2 * f["roll_range"] - f["roll_range"]
            target_pose = torch.stack([x, y, z, roll, pitch, yaw], dim=2)
            target_pose[:, 0, :] = 0
            delta = (target_pose[:, 1:, :] - target_pose[:, :-1, :]) / ni
            interp = torch.stack([target_pose[:, :-1] + (i + 1) * delta for i in range(ni)], dim=2)
            interp = interp.reshape(K, -1, 6)
            interp[:, :, :3] += root[env_ids, :3][:, None, :]
            return interp
        raise NotImplementedError(fn)

    def reset_idx(self, env_ids, uniforms):
        """LR:218-296 with _reset_dofs LR:998-1013, _reset_root_states LR:1015-1052, _resample_trajectory LR:949-953"""
        f = self.flat
        if len(env_ids) == 0:
            return
        u = uniforms[env_ids]
        self.randomize_dof_props(env_ids, uniforms, RS.RESET_STRENGTH)
        # _reset_dofs
        self.dof_pos[env_ids] = self.default_dof_pos * ((1.5 - 0.5) * u[:, RS.RESET_DOF:RS.RESET_DOF + 12] + 0.5)
        self.dof_vel[env_ids] = 0.
        # _reset_root_states
        root = self.robot_root
        root[env_ids] = self.base_init_state
        root[env_ids, :3] += self.env_origins[env_ids]
        if f["custom_origins"]:
            xr, yr = f["x_init_range"], f["y_init_range"]
            root[env_ids, 0:1] += (xr - (-xr)) * u[:, RS.RESET_X:RS.RESET_X + 1] + (-xr)
            root[env_ids, 1:2] += (yr - (-yr)) * u[:, RS.RESET_Y:RS.RESET_Y + 1] + (-yr)
            root[env_ids, 0] += f["x_init_offset"]
            root[env_ids, 1] += f["y_init_offset"]
        yr = f["yaw_init_range"]
        init_yaws = (yr - (-yr)) * u[:, RS.RESET_YAW:RS.RESET_YAW + 1] + (-yr)
        root[env_ids, 3:7] = quat_from_yaw(init_yaws)
        root[env_ids, 7:13] = (0.5 - (-0.5)) * u[:, RS.RESET_VEL:RS.RESET_VEL + 6] + (-0.5)
        # _resample_trajectory
        self.curr_pose_index[env_ids] = 0
        self.trajectories[env_ids] = self.trajectory_function(env_ids, uniforms)
        episode_length_buf = self.episode_length_buf.float().clone()
        self.last_actions[env_ids] = 0.
        self.last_last_actions[env_ids] = 0.
        self.last_dof_vel[env_ids] = 0.
        self.feet_air_time[env_ids] = 0.
        self.episode_length_buf[env_ids] = 0
        self.reset_buf[env_ids] = 1
        self.local_relative_linear[env_ids] = 0.
        self.local_relative_rotation[env_ids] = 0.
        self.plan_buf[env_ids] = 1
        # extras (LR:256-273): what the reference pushes to host deques
        rec = {"env_ids": env_ids.clone(), "episode_length": episode_length_buf[env_ids].clone(),
               "reached": self.reached_buf[env_ids].clone(),
               "goal_distance": torch.norm(self.relative_linear[env_ids], dim=1)}
        for key in self.episode_sums.keys():
            rec["rew_" + key] = self.episode_sums[key][env_ids].clone()
            self.episode_sums[key][env_ids] = 0.
        self.finished_episodes.append(rec)
        self.reached_env_buf[env_ids] = self.reached_buf[env_ids]
        self.collision_env_buf[env_ids] = self.collision_count[env_ids] * self.reached_buf[env_ids]
        self.collision_count[env_ids] = 0
        if f["send_timeouts"]:
            self.time_outs_latched = self.time_out_buf.clone()                # extras["time_outs"], LR:289-291
        for i in range(len(self.lag_buffer)):
            self.lag_buffer[i][env_ids, :] = 0

    # ------------------------------------------------------------------ a11: observations
    def compute_observations(self, noise_uniforms):
        """LR:357-590"""
        f = self.flat
        root = self.robot_root
        commands = self.local_relative_linear[:, :2]                         # a VIEW in the reference (LR:801-802)
        parts = [self.projected_gravity]
        if f["observe_command"]:
            parts.append(commands * 1.0)                                     # commands_scale = ones (LR:1275)
        parts += [(self.dof_pos - self.default_dof_pos) * f["obs_scale_dof_pos"],
                  self.dof_vel * f["obs_scale_dof_vel"], self.actions]
        obs = torch.cat(parts, dim=-1)
        if f["observe_command"] and f["timestep_in_obs"]:
            obs = torch.cat((obs, self.episode_length_buf.unsqueeze(1) / f["max_episode_length"]), dim=-1)
        if f["observe_heights"]:
            x_start = int(self.measured_heights.shape[2] // 2 + 1) if f["measure_front_half"] else 0
            mh = self.measured_heights[:, :, x_start:, :]
            if f["camera_zero"]:
                mh -= root[..., 2:3, None, None]                             # POST-reset root z (in place on the view)
                camera_offset_z = torch.sin(self.camera_pitch_angle) * torch.norm(self.camera_offset[None, :], dim=-1)
                mh -= camera_offset_z[:, None, None, None]
                mh = torch.clip(mh, min=-0.3, max=0.3)
            else:
                mh = torch.clip(mh, min=0, max=f["ceiling_height"])
                mh /= f["ceiling_height"]
                mh -= 0.5
            mh = mh.reshape(self.N, -1) * f["obs_scale_height"]
            obs = torch.cat((obs, mh), dim=-1)
        if f["observe_two_prev_actions"]:
            obs = torch.cat((obs, self.last_actions), dim=-1)
        if f["observe_vel"]:
            if f["global_reference"]:
                obs = torch.cat((root[:, 7:10] * f["obs_scale_lin_vel"], self.base_ang_vel * f["obs_scale_ang_vel"], obs), dim=-1)
            else:
                obs = torch.cat((self.base_lin_vel * f["obs_scale_lin_vel"], self.base_ang_vel * f["obs_scale_ang_vel"], obs), dim=-1)
        if f["observe_only_ang_vel"]:
            obs = torch.cat((self.base_ang_vel * f["obs_scale_ang_vel"], obs), dim=-1)
        if f["observe_only_lin_vel"]:
            obs = torch.cat((self.base_lin_vel * f["obs_scale_lin_vel"], obs), dim=-1)
        if f["observe_yaw"]:
            forward = quat_apply(root[:, 3:7], torch.tensor([1., 0., 0.]).repeat(self.N, 1))
            heading = torch.atan2(forward[:, 1], forward[:, 0]).unsqueeze(1)
            obs = torch.cat((obs, heading), dim=-1)
        if f["observe_contact_states"]:
            obs = torch.cat((obs, (self.contact_forces[:, self.feet_indices, 2] > 1.).view(self.N, -1) * 1.0), dim=1)
        if f["add_noise"]:
            obs += (2 * torch.as_tensor(noise_uniforms, dtype=torch.float32) - 1) * self.noise_scale_vec
        assert obs.shape[1] == f["num_obs"], (obs.shape, f["num_obs"])
        self.obs_buf = obs
        # privileged observations LR:477-590
        priv = torch.empty(self.N, 0)
        def add(x, rng, divide=False):
            nonlocal priv
            scale, shift = scale_shift(rng)
            v = (x - shift) / scale if divide else (x - shift) * scale
            priv = torch.cat((priv, v), dim=1)
        if f["priv_observe_friction"]:
            add(self.friction_coeffs[:, 0].unsqueeze(1), f["norm_friction_range"])
        if f["priv_observe_restitution"]:
            add(self.restitutions[:, 0].unsqueeze(1), f["norm_restitution_range"])
        if f["priv_observe_base_mass"]:
            add(self.payloads[:, 0].unsqueeze(1), f["norm_added_mass_range"])
        if f["priv_observe_com_displacement"]:
            add(self.com_displacements, f["norm_com_displacement_range"])
        if f["priv_observe_motor_strength"]:
            add(self.motor_strengths, f["norm_motor_strength_range"])
        if f["priv_observe_motor_offset"]:
            add(self.motor_offsets, f["norm_motor_offset_range"])
        if f["priv_observe_body_height"]:
            add(root[:, 2].view(self.N, -1), f["norm_body_height_range"])
        if f["priv_observe_body_velocity"]:
            add(self.base_lin_vel.view(self.N, -1), f["norm_body_velocity_range"])
        if f["priv_observe_gravity"]:
            add(self.gravities, f["norm_gravity_range"], divide=True)        # LR:574-579 divides by the scale
        assert priv.shape[1] == f["num_privileged_obs"], (priv.shape, f["num_privileged_obs"])
        self.privileged_obs_buf = priv
