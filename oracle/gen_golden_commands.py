"""Records tests/golden/commands_64.npz from the UNMODIFIED reference (build container only: needs /root/reference):
LeggedRobot._init_command_distribution and LeggedRobot._resample_commands (go1_gym/envs/base/legged_robot_velocity_tracking.py:
1317-1400, 728-842) called unbound on a namespace that carries the attributes they read / write, with the reference's own
RewardThresholdCurriculum objects (go1_gym/envs/base/curriculum.py) and the command settings of scripts/train_velocity_tracking.py.
`curriculum.sample` and `torch.rand` are wrapped so that what they returned is recorded; their values stay the reference's.

    python -m oracle.gen_golden_commands
"""
import os
import types

import numpy as np
import torch

from oracle import ref_harness as H

OUT = os.path.join(H.REPO, "tests", "golden", "commands_64.npz")
# scripts/train_velocity_tracking.py:29, 158-214 (the command block) -- config values, not code
TRAIN_CMD = dict(distributional_commands=True, num_commands=15, resampling_time=10,
                 lin_vel_x=[-1.0, 1.0], lin_vel_y=[-0.6, 0.6], ang_vel_yaw=[-1.0, 1.0], body_height_cmd=[-0.25, 0.15],
                 gait_frequency_cmd_range=[2.0, 4.0], gait_phase_cmd_range=[0.0, 1.0], gait_offset_cmd_range=[0.0, 1.0],
                 gait_bound_cmd_range=[0.0, 1.0], gait_duration_cmd_range=[0.5, 0.5], footswing_height_range=[0.03, 0.35],
                 body_pitch_range=[-0.4, 0.4], body_roll_range=[-0.0, 0.0], stance_width_range=[0.10, 0.45], stance_length_range=[0.35, 0.45],
                 limit_vel_x=[-5.0, 5.0], limit_vel_y=[-0.6, 0.6], limit_vel_yaw=[-5.0, 5.0], limit_body_height=[-0.25, 0.15],
                 limit_gait_frequency=[2.0, 4.0], limit_gait_phase=[0.0, 1.0], limit_gait_offset=[0.0, 1.0], limit_gait_bound=[0.0, 1.0],
                 limit_gait_duration=[0.5, 0.5], limit_footswing_height=[0.03, 0.35], limit_body_pitch=[-0.4, 0.4],
                 limit_body_roll=[-0.0, 0.0], limit_stance_width=[0.10, 0.45], limit_stance_length=[0.35, 0.45],
                 num_bins_vel_x=21, num_bins_vel_y=1, num_bins_vel_yaw=21, num_bins_body_height=1, num_bins_gait_frequency=1,
                 num_bins_gait_phase=1, num_bins_gait_offset=1, num_bins_gait_bound=1, num_bins_gait_duration=1,
                 num_bins_footswing_height=1, num_bins_body_roll=1, num_bins_body_pitch=1, num_bins_stance_width=1)
# name: the four gait switches of cfg.commands (train_velocity_tracking.py:216-219 is the first line)
CASES = {"gaitwise_binary": dict(gaitwise_curricula=True, exclusive_phase_offset=False, balance_gait_distribution=False, binary_phases=True),
         "gaitwise": dict(gaitwise_curricula=True, exclusive_phase_offset=False, balance_gait_distribution=False, binary_phases=False),
         "exclusive": dict(gaitwise_curricula=False, exclusive_phase_offset=True, balance_gait_distribution=False, binary_phases=False),
         "balance_binary": dict(gaitwise_curricula=False, exclusive_phase_offset=False, balance_gait_distribution=True, binary_phases=True),
         "balance": dict(gaitwise_curricula=False, exclusive_phase_offset=False, balance_gait_distribution=True, binary_phases=False),
         "plain": dict(gaitwise_curricula=False, exclusive_phase_offset=False, balance_gait_distribution=False, binary_phases=False)}
SUM_KEYS = ["tracking_lin_vel", "tracking_ang_vel", "tracking_contacts_shaped_force", "tracking_contacts_shaped_vel", "lin_vel_raw",
            "ang_vel_raw", "lin_vel_residual", "ang_vel_residual", "ep_timesteps"]


def main():
    H.install()
    if not hasattr(np, "int"):
        np.int = int                                    # legged_robot_velocity_tracking.py:1380 uses the removed numpy alias
    import go1_gym.envs.base.legged_robot_velocity_tracking as LRV
    from go1_gym.envs.base.legged_robot_velocity_tracking_config import Cfg
    N = 64
    fx = {}
    for case, switches in CASES.items():
        saved = {k: getattr(Cfg.commands, k) for k in vars(Cfg.commands)}
        try:
            for k, v in {**TRAIN_CMD, **switches}.items():
                setattr(Cfg.commands, k, v)
            torch.manual_seed(31)
            g = torch.Generator().manual_seed(5)
            ns = types.SimpleNamespace()
            ns.cfg = Cfg
            ns.device = "cpu"
            ns.dt = float(Cfg.control.decimation * np.float32(0.005))
            Cfg.env.max_episode_length = np.ceil(Cfg.env.episode_length_s / ns.dt)       # what _parse_cfg derives (LRV:1495-1496)
            ns.num_envs = N
            LRV.LeggedRobot._init_command_distribution(ns, torch.arange(N))
            ns.commands = torch.zeros(N, Cfg.commands.num_commands)
            ns.command_sums = {k: torch.rand(N, generator=g) * 50 for k in SUM_KEYS}
            ns.curriculum_thresholds = dict(tracking_lin_vel=0.8, tracking_ang_vel=0.7, tracking_contacts_shaped_force=0.9,
                                            tracking_contacts_shaped_vel=0.9)
            ns.reward_scales = dict(tracking_lin_vel=1.0 * ns.dt, tracking_ang_vel=0.5 * ns.dt, tracking_contacts_shaped_force=4.0 * ns.dt,
                                    tracking_contacts_shaped_vel=4.0 * ns.dt)
            samples, rands = [], []
            for cur in ns.curricula:
                def wrapped(batch_size, _orig=cur.sample, **kw):
                    out = _orig(batch_size=batch_size, **kw)
                    samples.append((np.asarray(out[0]).copy(), np.asarray(out[1]).copy()))
                    return out
                cur.sample = wrapped
            _rand = torch.rand

            def rand(*a, **k):
                r = _rand(*a, **k)
                rands.append(r.numpy().copy())
                return r
            tag = f"{case}/"
            for k, v in switches.items():
                fx[tag + "cfg/" + k] = np.asarray(bool(v))
            fx[tag + "category_names"] = np.array(ns.category_names)
            # call 0: every env (the env constructor's first resample); call 1: a ragged subset, as after a reset
            for call, env_ids in enumerate([torch.arange(N), torch.tensor([3, 4, 9, 17, 18, 30, 31, 40, 41, 42, 50, 63])]):
                samples.clear()
                rands.clear()
                before = ns.commands.clone()
                torch.rand = rand
                try:
                    LRV.LeggedRobot._resample_commands(ns, env_ids)
                finally:
                    torch.rand = _rand
                t = f"{tag}call{call}/"
                fx[t + "env_ids"] = env_ids.numpy().copy()
                fx[t + "commands_before"] = before.numpy()
                fx[t + "commands_after"] = ns.commands.numpy().copy()
                fx[t + "category_uniform"] = rands[0]
                if len(rands) > 1:
                    fx[t + "gait_uniform"] = rands[1]
                # what the curricula returned, re-assembled per resampled env in env_ids order (category i takes u in [i/n, (i+1)/n))
                n_cat = len(ns.category_names)
                u = torch.from_numpy(rands[0])
                new = np.zeros((len(env_ids), Cfg.commands.num_commands), np.float64)
                cat = np.full(len(env_ids), -1, np.int64)
                it = iter(samples)
                for i in range(n_cat):
                    m = torch.logical_and(1.0 / n_cat * i <= u, u < 1.0 / n_cat * (i + 1)).numpy()
                    if m.sum() == 0:
                        continue
                    s_cmd, s_bins = next(it)
                    new[m] = s_cmd[:, :Cfg.commands.num_commands]
                    cat[m] = i
                fx[t + "new_commands"] = new
                fx[t + "category"] = cat
                fx[t + "env_command_categories"] = np.asarray(ns.env_command_categories).copy()
                for k2, v2 in ns.command_sums.items():
                    fx[t + "command_sums/" + k2] = v2.numpy().copy()
                for k2 in ns.command_sums:                                  # refill so that the next call's zeroing is visible
                    ns.command_sums[k2] = torch.rand(N, generator=g) * 50
                    fx[t + "command_sums_next_in/" + k2] = ns.command_sums[k2].numpy().copy()
        finally:
            for k, v in saved.items():
                setattr(Cfg.commands, k, v)
            if "max_episode_length" in vars(Cfg.env):
                delattr(Cfg.env, "max_episode_length")
    np.savez_compressed(OUT, **fx)
    print("wrote", OUT, len(fx), "arrays")


if __name__ == "__main__":
    main()
