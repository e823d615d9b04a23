"""SURVEY.md section 8 row (f) 3, second half: the device half of LeggedRobot._resample_commands (legged_robot_velocity_tracking.py:
776-840).  tests/golden/commands_64.npz is a recording of the UNMODIFIED method driven by the reference's own
RewardThresholdCurriculum objects (oracle/gen_golden_commands.py): six settings of the gait switches, a full resample and a ragged
one each.  CPU: the oracle restatement against the recording; GPU: lt_resample_commands through the C ABI against the same
recording.  Bar: bit-exact (copies, halvings, +-0.25, round-half-even, % 1, one sqrt)."""
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "commands_64.npz")
CASES = ["gaitwise_binary", "gaitwise", "exclusive", "balance_binary", "balance", "plain"]
SWITCHES = ("gaitwise_curricula", "exclusive_phase_offset", "balance_gait_distribution", "binary_phases")


def load():
    z = np.load(GOLDEN, allow_pickle=False)
    return {k: z[k] for k in z.files}


def _case(g, case):
    sw = {k: bool(g[f"{case}/cfg/{k}"]) for k in SWITCHES}
    names = [str(n) for n in g[f"{case}/category_names"]]
    sum_keys = sorted(k.split("/")[-1] for k in g if k.startswith(f"{case}/call0/command_sums/"))
    return sw, names, sum_keys


@pytest.mark.parametrize("case", CASES)
def test_oracle_reproduces_the_reference(case):
    from oracle import commands_port
    g = load()
    sw, names, sum_keys = _case(g, case)
    commands = torch.from_numpy(g[f"{case}/call0/commands_before"]).clone()
    for call in range(2):
        t = f"{case}/call{call}/"
        u = torch.from_numpy(g[t + "category_uniform"])
        cat = commands_port.categories_from_uniform(u, len(names))
        assert np.array_equal(cat.numpy(), g[t + "category"])
        sums = {k: torch.full((64,), 7.0) for k in sum_keys}
        gu = torch.from_numpy(g[t + "gait_uniform"]) if t + "gait_uniform" in g else None
        commands_port.apply(commands, sums, torch.from_numpy(g[t + "env_ids"]), torch.from_numpy(g[t + "new_commands"]), cat, names,
                            gait_uniform=gu, **sw)
        assert np.array_equal(commands.numpy(), g[t + "commands_after"]), (case, call)
        ids = g[t + "env_ids"]
        for k in sum_keys:
            want = np.full(64, 7.0, np.float32)
            want[ids] = 0.0
            assert np.array_equal(sums[k].numpy(), want)
            assert np.all(g[t + "command_sums/" + k][ids] == 0.0)


def test_recording_covers_every_gait_branch():
    g = load()
    assert set(g["gaitwise/call0/category"].tolist()) == {0, 1, 2, 3}
    u = g["balance/call0/gait_uniform"]
    assert (u <= 0.25).any() and ((0.25 <= u) & (u < 0.5)).any() and ((0.5 <= u) & (u < 0.75)).any() and (u >= 0.75).any()
    after = g["gaitwise_binary/call0/commands_after"]
    assert set(np.unique(after[:, 5:8]).tolist()) <= {0.0, 0.5}                 # binary phases
    assert (np.linalg.norm(g["plain/call0/commands_after"][:, :2], axis=1) == 0).any()      # small commands were zeroed


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_resample_commands_kernel_matches_the_reference(case):
    from legged_tracking_b200.gait import CommandResampler
    g = load()
    sw, names, sum_keys = _case(g, case)
    dev = torch.device("cuda", 0)
    commands = torch.from_numpy(g[f"{case}/call0/commands_before"]).to(dev).contiguous()
    sums = torch.full((len(sum_keys), 64), 7.0, device=dev)
    rs = CommandResampler(commands, sums, names, **sw)
    for call in range(2):
        t = f"{case}/call{call}/"
        sums.fill_(7.0)
        u = torch.from_numpy(g[t + "category_uniform"]).to(dev)
        cat = rs.categories_from_uniform(u)
        assert np.array_equal(cat.cpu().numpy(), g[t + "category"])
        gu = torch.from_numpy(g[t + "gait_uniform"]) if t + "gait_uniform" in g else None
        rs.apply(torch.from_numpy(g[t + "env_ids"]), g[t + "new_commands"], category=cat, gait_uniform=gu)
        torch.cuda.synchronize()
        assert np.array_equal(commands.cpu().numpy(), g[t + "commands_after"]), (case, call)
        want = np.full((len(sum_keys), 64), 7.0, np.float32)
        want[:, g[t + "env_ids"]] = 0.0
        assert np.array_equal(sums.cpu().numpy(), want)
    rs.apply(torch.zeros(0, dtype=torch.int64), np.zeros((0, 15)))              # LRV:730: nothing to do


@pytest.mark.gpu
def test_resample_commands_at_full_size_matches_the_oracle():
    """65,536 envs resampled at once (the first step of a cfg 4 run), random categories and draws, against the oracle."""
    from legged_tracking_b200.gait import CommandResampler
    from oracle import commands_port
    N, nc = 65536, 15
    gen = torch.Generator().manual_seed(3)
    names = ["pronk", "trot", "pace", "bound"]
    for sw in (dict(gaitwise_curricula=True, binary_phases=True), dict(balance_gait_distribution=True), dict(exclusive_phase_offset=True)):
        new = torch.rand(N, nc, generator=gen, dtype=torch.float64) * 2 - 0.5
        new[:, :2] *= 0.3
        cat = torch.randint(0, 4, (N,), generator=gen)
        gu = torch.rand(N, generator=gen)
        env_ids = torch.randperm(N, generator=gen)
        ref_cmd = torch.zeros(N, nc)
        commands_port.apply(ref_cmd, {}, env_ids, new, cat, names, gait_uniform=gu, **sw)
        cmd = torch.zeros(N, nc, device="cuda")
        CommandResampler(cmd, None, names, **sw).apply(env_ids, new, category=cat, gait_uniform=gu)
        torch.cuda.synchronize()
        assert torch.equal(cmd.cpu(), ref_cmd), sw
