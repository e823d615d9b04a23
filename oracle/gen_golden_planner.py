"""Records tests/golden/planner_48.npz from the UNMODIFIED reference: LeggedRobot._plan_target_pose with
cfg.commands.sampling_based_planning = True (legged_robot_trajectory_tracking.py:850-920) called on a reference env whose state
tensors were filled with seeded values (no PhysX: the function reads root_states, trajectories, measured_heights, the plan
buffers and the candidate table only).   python -m oracle.gen_golden_planner      (build container only: needs /root/reference)
"""
import contextlib
import io
import os

import numpy as np
import torch

from oracle import ref_harness as H

OUT = os.path.join(H.REPO, "tests", "golden", "planner_48.npz")


def main():
    N = 48
    argv = ["--old_ppo", "--measure_front_half", "--camera_zero", "--strategy", "e2e", "--terrain", "single_path"]
    with contextlib.redirect_stdout(io.StringIO()):
        wrapped, env, replay, Cfg, args = H.make_env(argv, N, 4, 4, overrides={"commands.sampling_based_planning": True})
    g = torch.Generator().manual_seed(4242)
    r = lambda *s: torch.rand(*s, generator=g)      # noqa: E731
    A = env.num_actor
    root = env.root_states.view(N, A, 13)
    root[:, 0, 0:2] = env.env_origins[:, :2] + torch.stack([r(N) * 2.0, r(N) * 0.6 - 0.3], dim=1)
    root[:, 0, 2] = 0.22 + 0.12 * r(N)
    from isaacgym.torch_utils import quat_from_euler_xyz
    root[:, 0, 3:7] = quat_from_euler_xyz((r(N) - 0.5) * 0.4, (r(N) - 0.5) * 0.4, (r(N) - 0.5) * 2 * np.pi)
    goal = torch.zeros(N, 6)
    dist = torch.cat([0.2 + 0.7 * r(N // 4), 1.0 + 3.0 * r(N - N // 4)])          # a quarter of the envs are close to their goal
    ang = (r(N) - 0.5) * 1.2
    goal[:, 0] = root[:, 0, 0] + dist * torch.cos(ang)
    goal[:, 1] = root[:, 0, 1] + dist * torch.sin(ang)
    goal[:, 2] = 0.34
    goal[:, 5] = (r(N) - 0.5) * 1.0
    env.trajectories[:] = goal[:, None, :]
    env.curr_pose_index[:] = 0
    mh = torch.zeros(N, 2, 21, 11)
    mh[:, 0] = 0.25 + 0.35 * r(N, 21, 11)                                          # ceiling: some points hang into the robot's height band
    mh[:, 1] = 0.20 * r(N, 21, 11)
    mh[::5, 0] = 0.8                                                               # open tiles: every candidate is valid
    mh[::5, 1] = 0.0
    mh[3::7, 0] = 0.28                                                             # blocked tiles: no candidate is valid
    env.measured_heights = mh.clone()
    env.plan_buf = (r(N) < 0.7).long()
    env.plan_length_buf = torch.randint(0, 25, (N,), generator=g)
    env.plan_length_buf[::3] = 9                                                   # +1 -> 10: the replan interval fires
    env.episode_length_buf[:] = torch.randint(2, 300, (N,), generator=g)
    env.episode_length_buf[1::6] = 1                                               # first step of an episode: always plan
    env.local_target_poses = torch.randn(N, 6, generator=g)
    fx = {"in/root_states": env.root_states.clone().numpy(), "in/target_poses": goal.numpy(),
          "in/points_xy": env.height_points[0].reshape(-1, 3)[:, :2].clone().numpy(), "in/measured_heights": mh.numpy(),
          "in/candidates": np.asarray(Cfg.commands.candidate_target_poses, np.float64),
          "in/episode_length_buf": env.episode_length_buf.clone().numpy(), "in/plan_buf": env.plan_buf.clone().numpy(),
          "in/plan_length_buf": env.plan_length_buf.clone().numpy(), "in/local_target_poses": env.local_target_poses.clone().numpy(),
          "meta/plan_interval": np.int64(Cfg.commands.plan_interval), "meta/num_actor": np.int64(A),
          "meta/robot_size": env.robot_size.numpy()}
    with contextlib.redirect_stdout(io.StringIO()) as log:
        env._plan_target_pose(torch.arange(N))
    fx["out/local_target_poses"] = env.local_target_poses.clone().numpy()
    fx["out/plan_length_buf"] = env.plan_length_buf.clone().numpy()
    fx["out/local_relative_linear"] = env.local_relative_linear.clone().numpy()
    fx["out/local_relative_rotation"] = env.local_relative_rotation.clone().numpy()
    fx["meta/no_valid_messages"] = np.int64(log.getvalue().count("no valid local goal"))
    np.savez_compressed(OUT, **fx)
    print("wrote", OUT, {k: v.shape for k, v in fx.items() if hasattr(v, "shape") and v.ndim})
    print("no-valid envs:", int(fx["meta/no_valid_messages"]))


if __name__ == "__main__":
    main()
