"""CPU restatement of the device half of LeggedRobot._resample_commands (go1_gym/envs/base/legged_robot_velocity_tracking.py:
776-840; LRV below).  TEST INFRASTRUCTURE ONLY: only tests/ may import it.  Pinned against tests/golden/commands_64.npz, which
oracle/gen_golden_commands.py recorded from the unmodified method with the reference's own RewardThresholdCurriculum objects.

The curricula themselves (weight grids, numpy RandomState sampling, curriculum.py:113-159) stay host code in the product as in the
reference; this port starts where their output -- one command row and one category per resampled env -- is written to the env."""
import torch

GAITS = {"pronk": 0, "trot": 1, "pace": 2, "bound": 3}


def categories_from_uniform(u, n_categories):
    """LRV:760-766: category i takes the envs with i / n <= u < (i + 1) / n."""
    p = 1.0 / n_categories
    cat = torch.full(u.shape, -1, dtype=torch.int64)
    for i in range(n_categories):
        cat[torch.logical_and(p * i <= u, u < p * (i + 1))] = i
    return cat


def apply(commands, command_sums, env_ids, new_commands, category, category_names, gait_uniform=None, gaitwise_curricula=False,
          exclusive_phase_offset=False, balance_gait_distribution=False, binary_phases=False):
    """commands [N, nc] and command_sums {name: [N]} are modified in place like the reference's tensors."""
    nc = commands.shape[1]
    commands[env_ids, :] = new_commands[:, :nc].float()                              # LRV:776-777
    if nc > 5:
        if gaitwise_curricula:                                                       # LRV:780-797
            for i, name in enumerate(category_names):
                ids = env_ids[category == i]
                if name == "pronk":
                    for j in (5, 6, 7):
                        commands[ids, j] = (commands[ids, j] / 2 - 0.25) % 1
                elif name == "trot":
                    commands[ids, 5] = commands[ids, 5] / 2 + 0.25
                    commands[ids, 6] = 0
                    commands[ids, 7] = 0
                elif name == "pace":
                    commands[ids, 5] = 0
                    commands[ids, 6] = commands[ids, 6] / 2 + 0.25
                    commands[ids, 7] = 0
                elif name == "bound":
                    commands[ids, 5] = 0
                    commands[ids, 6] = 0
                    commands[ids, 7] = commands[ids, 7] / 2 + 0.25
        elif exclusive_phase_offset:                                                 # LRV:799-809
            u = gait_uniform
            trot, pace, bound = env_ids[u < 0.34], env_ids[torch.logical_and(0.34 <= u, u < 0.67)], env_ids[0.67 <= u]
            commands[pace, 5] = 0
            commands[bound, 5] = 0
            commands[trot, 6] = 0
            commands[bound, 6] = 0
            commands[trot, 7] = 0
            commands[pace, 7] = 0
        elif balance_gait_distribution:                                              # LRV:811-828
            u = gait_uniform
            pronk, trot = env_ids[u <= 0.25], env_ids[torch.logical_and(0.25 <= u, u < 0.50)]
            pace, bound = env_ids[torch.logical_and(0.50 <= u, u < 0.75)], env_ids[0.75 <= u]
            for j in (5, 6, 7):
                commands[pronk, j] = (commands[pronk, j] / 2 - 0.25) % 1
            commands[trot, 6] = 0
            commands[trot, 7] = 0
            commands[pace, 5] = 0
            commands[pace, 7] = 0
            commands[bound, 5] = 0
            commands[bound, 6] = 0
            commands[trot, 5] = commands[trot, 5] / 2 + 0.25
            commands[pace, 6] = commands[pace, 6] / 2 + 0.25
            commands[bound, 7] = commands[bound, 7] / 2 + 0.25
        if binary_phases:                                                            # LRV:830-833
            for j in (5, 6, 7):
                commands[env_ids, j] = (torch.round(2 * commands[env_ids, j])) / 2.0 % 1
    commands[env_ids, :2] *= (torch.norm(commands[env_ids, :2], dim=1) > 0.2).unsqueeze(1)      # LRV:836
    for key in command_sums:                                                         # LRV:839-840
        command_sums[key][env_ids] = 0.
    return commands
