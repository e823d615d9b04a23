"""Seeded synthetic Isaac-Gym-layout state ("replayed PhysX output").

PhysX is out of scope (BASELINE.json north_star): the step consumes state in the layouts
of Isaac Gym's acquire_*_tensor calls (reference
go1_gym/envs/base/legged_robot_trajectory_tracking.py:1173-1201).  This module draws that
state from a seed so that the oracle (CPU) and the CUDA path see identical inputs, and so
that bench.py has a workload of the shape SURVEY.md section 8(d) describes.

Layouts produced (fp32):
  robot root rows  [N, 13]   pos3, quat xyzw, linvel3, angvel3
  dof_state        [N*12, 2] (pos, vel) interleaved
  contact_forces   [N, B, 3]
  rigid_body_state [N, B, 13]
  actions          [N, 12]
"""
import math

import numpy as np
import torch

NUM_DOF = 12
# resources/robots/go1/urdf/go1.urdf:96,138,166 (hip, thigh, calf) x (FL, FR, RL, RR)
DOF_LOWER = np.array([-0.802851455917, -1.0471975512, -2.69653369433] * 4, dtype=np.float32)
DOF_UPPER = np.array([0.802851455917, 4.18879020479, -0.916297857297] * 4, dtype=np.float32)
# go1_gym/envs/go1/go1_crawling.py:12-27 in Isaac Gym DOF order FL, FR, RL, RR x (hip, thigh, calf)
DEFAULT_DOF_POS = np.array([0.1, 0.8, -1.5, -0.1, 0.8, -1.5, 0.1, 1.0, -1.5, -0.1, 1.0, -1.5], dtype=np.float32)


def _quat_from_rpy(roll, pitch, yaw):
    cy, sy = torch.cos(yaw * 0.5), torch.sin(yaw * 0.5)
    cr, sr = torch.cos(roll * 0.5), torch.sin(roll * 0.5)
    cp, sp = torch.cos(pitch * 0.5), torch.sin(pitch * 0.5)
    qw = cy * cr * cp + sy * sr * sp
    qx = cy * sr * cp - sy * cr * sp
    qy = cy * cr * sp + sy * sr * cp
    qz = sy * cr * cp - cy * sr * sp
    return torch.stack([qx, qy, qz, qw], dim=-1)


class SyntheticState:
    """Deterministic function (seed, step[, substep]) -> state tensors, generated on CPU."""

    def __init__(self, num_envs, env_origins, num_bodies=18, seed=1000, feet_indices=(4, 8, 12, 16),
                 z_range=(0.15, 0.40), x_span=2.6, y_span=0.4, action_std=1.5, rp_range=0.3):
        self.n = int(num_envs)
        self.nb = int(num_bodies)
        self.seed = int(seed)
        self.env_origins = torch.as_tensor(np.asarray(env_origins), dtype=torch.float32).reshape(self.n, 3).cpu()
        self.feet = list(feet_indices)
        self.z_range = z_range
        self.x_span = x_span
        self.y_span = y_span
        self.action_std = action_std
        self.rp_range = float(rp_range)     # roll / pitch ~ U(+-rp_range); > pi/2 turns some robots over (LR:215-216)
        self.default_dof = torch.from_numpy(DEFAULT_DOF_POS)
        self.lo = torch.from_numpy(DOF_LOWER)
        self.hi = torch.from_numpy(DOF_UPPER)

    def _gen(self, step, substep, stream):
        g = torch.Generator(device="cpu")
        g.manual_seed((self.seed * 1000003 + step * 4099 + substep * 131 + stream) & 0x7FFFFFFFFFFF)
        return g

    def initial_episode_length(self, max_episode_length):
        g = self._gen(0, 0, 7)
        return torch.randint(int(max_episode_length), (self.n,), generator=g, dtype=torch.long)

    def actions(self, step):
        g = self._gen(step, 0, 1)
        a = torch.randn(self.n, NUM_DOF, generator=g) * self.action_std
        # a few large actions so that clip_actions (+-10) is exercised
        big = torch.rand(self.n, NUM_DOF, generator=g) < 0.01
        return torch.where(big, a * 10.0, a)

    def dof_state(self, step, substep):
        g = self._gen(step, substep, 2)
        pos = self.default_dof + torch.randn(self.n, NUM_DOF, generator=g) * 0.3
        mid = 0.5 * (self.lo + self.hi)
        half = 0.5 * (self.hi - self.lo) * 1.05
        pos = torch.max(torch.min(pos, mid + half), mid - half)
        vel = torch.randn(self.n, NUM_DOF, generator=g) * 5.0
        return torch.stack([pos, vel], dim=-1).reshape(self.n * NUM_DOF, 2).contiguous()

    def root(self, step):
        g = self._gen(step, 0, 3)
        u = torch.rand(self.n, 6, generator=g)
        pos = self.env_origins.clone()
        pos[:, 0] += u[:, 0] * self.x_span
        pos[:, 1] += (2 * u[:, 1] - 1) * self.y_span
        pos[:, 2] = self.z_range[0] + u[:, 2] * (self.z_range[1] - self.z_range[0])
        rpy = (2 * u[:, 3:6] - 1) * torch.tensor([self.rp_range, self.rp_range, math.pi])
        quat = _quat_from_rpy(rpy[:, 0], rpy[:, 1], rpy[:, 2])
        lin = torch.randn(self.n, 3, generator=g) * 0.5
        ang = torch.randn(self.n, 3, generator=g) * 1.0
        return torch.cat([pos, quat, lin, ang], dim=-1).contiguous()

    def contact(self, step):
        g = self._gen(step, 0, 4)
        hit = torch.rand(self.n, self.nb, 1, generator=g) < 0.1
        f = torch.randn(self.n, self.nb, 3, generator=g) * 20.0
        # a sprinkling of tiny forces that straddle the 0.1 N / 1 N thresholds
        small = torch.rand(self.n, self.nb, 1, generator=g) < 0.05
        f = torch.where(small, f * 0.01, f)
        return (f * (hit | small)).contiguous()

    def rigid(self, step, root):
        g = self._gen(step, 0, 5)
        rb = torch.zeros(self.n, self.nb, 13)
        rb[:, :, 0:3] = root[:, None, 0:3] + torch.randn(self.n, self.nb, 3, generator=g) * 0.2
        rb[:, :, 3:7] = root[:, None, 3:7]
        rb[:, :, 7:13] = root[:, None, 7:13] + torch.randn(self.n, self.nb, 6, generator=g) * 0.1
        return rb.contiguous()
