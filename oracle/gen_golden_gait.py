"""Records tests/golden/gait_corl_64.npz from the UNMODIFIED reference (build container only: needs /root/reference):

  * LeggedRobot._step_contact_targets (go1_gym/envs/base/legged_robot_velocity_tracking.py:844-923), called unbound on a
    namespace that carries exactly the attributes the method reads / writes, for 3 consecutive steps;
  * every CoRLRewards._reward_* (go1_gym/envs/rewards/corl_rewards.py:15-201) on a namespace env of seeded random state tensors
    (the container only reads attributes of `env`), twice in a row (feet_slip carries last_contacts from call to call).

Neither needs PhysX, a terrain or the env constructor.   python -m oracle.gen_golden_gait
"""
import os
import types

import numpy as np
import torch

from oracle import ref_harness as H

OUT = os.path.join(H.REPO, "tests", "golden", "gait_corl_64.npz")
CORL_TERMS = ["tracking_lin_vel", "tracking_ang_vel", "lin_vel_z", "ang_vel_xy", "orientation", "torques", "dof_acc", "action_rate",
              "collision", "dof_pos_limits", "jump", "tracking_contacts_shaped_force", "tracking_contacts_shaped_vel", "dof_pos", "dof_vel",
              "action_smoothness_1", "action_smoothness_2", "feet_slip", "feet_contact_vel", "feet_contact_forces",
              "feet_clearance_cmd_linear", "feet_impact_vel", "orientation_control", "raibert_heuristic"]


def main():
    H.install()
    import go1_gym.envs.base.legged_robot_velocity_tracking as LRV
    from go1_gym.envs.rewards.corl_rewards import CoRLRewards
    from go1_gym.envs.base.legged_robot_velocity_tracking_config import Cfg
    N = 64
    g = torch.Generator().manual_seed(77)
    r = lambda *s: torch.rand(*s, generator=g)          # noqa: E731
    rn = lambda *s: torch.randn(*s, generator=g)        # noqa: E731
    fx = {}
    # ---------------------------------------------------------------- gait: commands in the ranges of scripts/train_velocity_tracking.py
    cmd = torch.zeros(N, 15)
    cmd[:, 0] = (r(N) - 0.5) * 4
    cmd[:, 1] = (r(N) - 0.5) * 1.2
    cmd[:, 2] = (r(N) - 0.5) * 2
    cmd[:, 3] = (r(N) - 0.5) * 0.3
    cmd[:, 4] = 2.0 + 2.0 * r(N)                       # gait frequency
    cmd[:, 5:8] = r(N, 3)                              # phase, offset, bound
    cmd[:, 8] = 0.3 + 0.4 * r(N)                       # duration
    cmd[::9, 8] = 0.5
    cmd[:, 9] = 0.03 + 0.22 * r(N)                     # footswing height
    cmd[:, 10:12] = (r(N, 2) - 0.5) * 0.8              # pitch, roll
    cmd[:, 12] = 0.25 + 0.2 * r(N)                     # stance width
    cmd[:, 13] = 0.35 + 0.1 * r(N)                     # stance length
    cmd[:, 14] = 0.25
    for pacing in (False, True):
        ns = types.SimpleNamespace()
        ns.cfg = types.SimpleNamespace(env=types.SimpleNamespace(observe_gait_commands=True),
                                       commands=types.SimpleNamespace(pacing_offset=pacing, num_commands=15),
                                       rewards=types.SimpleNamespace(kappa_gait_probs=float(Cfg.rewards.kappa_gait_probs)))
        ns.dt = float(Cfg.control.decimation * np.float32(0.005))
        ns.commands = cmd.clone()
        ns.gait_indices = r(N).clone()
        if not pacing:
            ns.gait_indices[:4] = torch.tensor([0.0, 0.25, 0.5, 0.999])
        z4 = lambda: torch.zeros(N, 4)                  # noqa: E731
        ns.clock_inputs, ns.doubletime_clock_inputs, ns.halftime_clock_inputs, ns.desired_contact_states = z4(), z4(), z4(), z4()
        tag = f"gait/pacing{int(pacing)}/"
        fx[tag + "commands"] = cmd.numpy().copy()
        fx[tag + "gait_indices0"] = ns.gait_indices.numpy().copy()
        fx[tag + "dt"] = np.float64(ns.dt)
        fx[tag + "kappa"] = np.float64(ns.cfg.rewards.kappa_gait_probs)
        for k in range(3):
            LRV.LeggedRobot._step_contact_targets(ns)
            for name in ("gait_indices", "foot_indices", "clock_inputs", "doubletime_clock_inputs", "halftime_clock_inputs",
                         "desired_contact_states", "desired_footswing_height"):
                fx[f"{tag}step{k}/{name}"] = getattr(ns, name).numpy().copy()
    # ---------------------------------------------------------------- CoRL reward terms
    B, feet, pen = 17, [4, 8, 12, 16], [2, 6, 10, 14, 3, 7, 11, 15, 0]
    env = types.SimpleNamespace()
    env.device = "cpu"
    env.num_envs, env.num_dof, env.num_actuated_dof = N, 12, 12
    env.dt = ns.dt
    env.cfg = types.SimpleNamespace(
        rewards=types.SimpleNamespace(**{k: float(getattr(Cfg.rewards, k)) for k in
                                         ("tracking_sigma", "tracking_sigma_yaw", "base_height_target", "gait_force_sigma", "gait_vel_sigma",
                                          "max_contact_force")}),
        commands=types.SimpleNamespace(num_commands=15))
    env.commands = cmd.clone()
    env.base_lin_vel, env.base_ang_vel = rn(N, 3) * 0.8, rn(N, 3)
    q = rn(N, 4)
    env.base_quat = q / q.norm(dim=1, keepdim=True)
    env.base_pos = torch.cat([rn(N, 2) * 3, 0.25 + 0.1 * r(N, 1)], dim=1)
    from isaacgym.torch_utils import quat_rotate_inverse
    env.gravity_vec = torch.tensor([0.0, 0.0, -1.0]).repeat(N, 1)
    env.projected_gravity = quat_rotate_inverse(env.base_quat, env.gravity_vec)
    env.torques = rn(N, 12) * 8
    env.dof_vel, env.last_dof_vel = rn(N, 12) * 5, rn(N, 12) * 5
    env.actions, env.last_actions, env.last_last_actions = rn(N, 12), rn(N, 12), rn(N, 12)
    env.last_actions[::7] = 0.0                         # "first step" rows of action_smoothness_1 / 2
    env.last_last_actions[1::5] = 0.0
    env.default_dof_pos = torch.tensor([[0.1, 0.8, -1.5, -0.1, 0.8, -1.5, 0.1, 1.0, -1.5, -0.1, 1.0, -1.5]])
    env.dof_pos = env.default_dof_pos + rn(N, 12) * 0.4
    lim = torch.tensor([[-0.863, 0.863], [-0.686, 4.501], [-2.818, -0.888]]).repeat(4, 1)
    mid, rng = lim.mean(dim=1), lim[:, 1] - lim[:, 0]
    env.dof_pos_limits = torch.stack([mid - 0.45 * rng, mid + 0.45 * rng], dim=1)       # soft limits as _process_dof_props leaves them
    env.joint_pos_target, env.last_joint_pos_target, env.last_last_joint_pos_target = rn(N, 12) * 0.3, rn(N, 12) * 0.3, rn(N, 12) * 0.3
    cf = torch.zeros(N, B, 3)
    hit = r(N, B) < 0.35
    cf[hit] = rn(int(hit.sum()), 3) * 30
    cf[:, feet, 2] = torch.where(r(N, 4) < 0.6, 20 + 200 * r(N, 4), torch.zeros(N, 4))
    env.contact_forces = cf
    env.feet_indices, env.penalised_contact_indices = torch.tensor(feet), torch.tensor(pen)
    env.foot_velocities, env.prev_foot_velocities = rn(N, 4, 3), rn(N, 4, 3)
    env.foot_positions = env.base_pos.unsqueeze(1) + torch.cat([rn(N, 4, 2) * 0.25, -0.25 + 0.1 * r(N, 4, 1)], dim=2)
    env.foot_positions[:, :, 2] = 0.01 + 0.1 * r(N, 4)
    env.foot_indices = r(N, 4)
    env.desired_contact_states = r(N, 4)
    env.last_contacts = r(N, 4) < 0.5
    for k, v in vars(env).items():
        if isinstance(v, torch.Tensor):
            fx["corl/in/" + k] = v.numpy().copy()
    for k, v in vars(env.cfg.rewards).items():
        fx["corl/cfg/" + k] = np.float64(v)
    fx["corl/in/dt"] = np.float64(env.dt)
    container = CoRLRewards(env)
    names = [n for n in CORL_TERMS if hasattr(container, "_reward_" + n)]
    fx["corl/names"] = np.array(names)
    for call in range(2):
        for n in names:
            fx[f"corl/out{call}/{n}"] = getattr(container, "_reward_" + n)().detach().float().numpy().copy()
        fx[f"corl/out{call}/last_contacts"] = env.last_contacts.numpy().copy()
    np.savez_compressed(OUT, **fx)
    print("wrote", OUT, len(fx), "arrays; terms:", names)


if __name__ == "__main__":
    main()
