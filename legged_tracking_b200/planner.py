"""Sampling-based local planner: LeggedRobot._plan_target_pose with
cfg.commands.sampling_based_planning (go1_gym/envs/base/legged_robot_trajectory_tracking.py:850-920) as ONE kernel launch
(lt_plan_target_pose, csrc/lt_planner.cu) instead of a Python loop over the planning envs.  CUDA only.

engine.StepEngine creates one when the flat config has sampling_based_planning (cfg.commands.sampling_based_planning with its
candidate table): per step lt_measure_heights (the full 2 x 21 x 11 grid of _get_heights, LR:786-787) -> LocalPlanner.plan ->
lt_post_physics_step, which takes this step's local target and its relative pose through lt_step_args (commands, rewards, plan_buf
and the goal arrow then follow the LOCAL target as in the reference, LR:800-802, 845-847).  Ties between equal sort keys (the key
ignores the candidate's z) go to the lowest candidate index; the reference's torch.argsort leaves that order to its sort
implementation, so a tied pick can differ from the reference in the local target's z (tests/test_planner.py).
"""
import ctypes as C

import numpy as np
import torch

from . import _abi

ROBOT_SIZE = (0.3762, 0.0935, 0.114)          # LR:1212


class LocalPlanner:
    def __init__(self, num_envs, points_xy, candidate_target_poses, plan_interval=10, num_actors=2, device="cuda:0",
                 robot_size=ROBOT_SIZE):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _abi.LtError("LocalPlanner runs as an sm_100a kernel: there is no CPU path")
        self.lib = _abi.load()
        self.N = int(num_envs)
        f32 = lambda a: torch.as_tensor(np.ascontiguousarray(np.asarray(a, np.float32))).to(self.device)     # noqa: E731
        self.points_xy = f32(points_xy).reshape(-1, 2).contiguous()
        self.candidates = f32(candidate_target_poses).reshape(-1, 6).contiguous()        # `.float()` of the float64 config table (LR:880)
        self.plan_interval, self.root_stride = int(plan_interval), 13 * int(num_actors)
        self.robot_size = (C.c_float * 3)(*[float(x) for x in robot_size])
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=self.device)                         # noqa: E731
        self.plan_length_buf = z(self.N, dt=torch.int64)                                  # LR:1214
        self.local_target_poses = z(self.N, 6)                                            # LR:1263
        self.local_relative_linear, self.local_relative_rotation = z(self.N, 3), z(self.N, 3)
        self.chosen = z(self.N, dt=torch.int32)
        self.replan = z(self.N, dt=torch.uint8)                                           # LR:861

    def plan(self, target_poses, root_states, measured_heights, episode_length_buf, plan_buf, stream=None):
        """One call of _plan_target_pose for all envs.  -> local_target_poses (updated in place; also local_relative_* / chosen)."""
        N, n_pts = self.N, self.points_xy.shape[0]
        chk = lambda t, dt, numel, name: (t.device == self.device and t.dtype == dt and t.is_contiguous() and t.numel() == numel) or \
            (_ for _ in ()).throw(ValueError(f"{name}: expected a contiguous {dt} tensor of {numel} elements on {self.device}"))   # noqa: E731
        chk(target_poses, torch.float32, N * 6, "target_poses")
        chk(root_states, torch.float32, N * self.root_stride, "root_states")
        chk(measured_heights, torch.float32, N * 2 * n_pts, "measured_heights")
        chk(episode_length_buf, torch.int64, N, "episode_length_buf")
        pb = plan_buf if plan_buf.dtype == torch.uint8 else (plan_buf != 0).to(torch.uint8)
        a = _abi.lt_plan_args()
        for name, t in (("target_poses", target_poses), ("root_states", root_states), ("points_xy", self.points_xy),
                        ("measured_heights", measured_heights), ("candidates", self.candidates),
                        ("episode_length_buf", episode_length_buf), ("plan_buf", pb), ("plan_length_buf", self.plan_length_buf),
                        ("local_target_poses", self.local_target_poses), ("local_relative_linear", self.local_relative_linear),
                        ("local_relative_rotation", self.local_relative_rotation), ("chosen", self.chosen),
                        ("replan", self.replan)):
            setattr(a, name, t.data_ptr())
        st = stream if stream is not None else torch.cuda.current_stream(self.device)
        self._keep = pb
        _abi.check(self.lib.lt_plan_target_pose(C.byref(a), N, n_pts, self.candidates.shape[0], self.plan_interval, self.root_stride,
                                                self.robot_size, C.c_void_p(st.cuda_stream)), None, "lt_plan_target_pose")
        return self.local_target_poses
