"""SURVEY.md section 8 rows a9.3 (CoRLRewards) and (f) 3 (_step_contact_targets): the oracle restatements against the recording of
the unmodified reference (CPU), and the CUDA kernels lt_step_contact_targets / lt_corl_rewards through the C ABI against the same
recording (GPU).  Bars: boolean state (last_contacts) bit-exact; fp32 values 1e-5 relative (+2e-6 absolute; the Normal-CDF /
sin / exp values differ from torch's CPU math library in the last bits)."""
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gait_corl_64.npz")
RTOL, ATOL = 1e-5, 2e-6
GAIT_OUT = ("gait_indices", "foot_indices", "clock_inputs", "doubletime_clock_inputs", "halftime_clock_inputs", "desired_contact_states")


def load():
    z = np.load(GOLDEN, allow_pickle=False)
    return {k: z[k] for k in z.files}


def corl_inputs(g):
    e = {k[len("corl/in/"):]: torch.from_numpy(np.asarray(v)) for k, v in g.items() if k.startswith("corl/in/") and np.asarray(v).ndim}
    e["dt"] = float(g["corl/in/dt"])
    cfg = {k[len("corl/cfg/"):]: float(v) for k, v in g.items() if k.startswith("corl/cfg/")}
    return e, cfg, [str(n) for n in g["corl/names"]]


@pytest.mark.parametrize("pacing", [0, 1])
def test_oracle_step_contact_targets_reproduces_the_reference(pacing):
    from oracle import gait_port
    g = load()
    tag = f"gait/pacing{pacing}/"
    gait = torch.from_numpy(g[tag + "gait_indices0"])
    cmd = torch.from_numpy(g[tag + "commands"])
    for k in range(3):
        out = gait_port.step_contact_targets(gait, cmd, float(g[tag + "dt"]), float(g[tag + "kappa"]), bool(pacing))
        for name in GAIT_OUT + ("desired_footswing_height",):
            assert np.array_equal(out[name].numpy(), g[f"{tag}step{k}/{name}"]), (k, name)
        gait = out["gait_indices"]


def test_oracle_corl_terms_reproduce_the_reference():
    from oracle import gait_port
    g = load()
    e, cfg, names = corl_inputs(g)
    assert len(names) == 24
    for call in range(2):
        out, lc = gait_port.corl_terms(e, cfg, names)
        for n in names:
            np.testing.assert_allclose(out[n].numpy(), g[f"corl/out{call}/{n}"], rtol=1e-6, atol=1e-7, err_msg=n)
        assert np.array_equal(lc.numpy(), g[f"corl/out{call}/last_contacts"])
        e["last_contacts"] = lc


@pytest.mark.gpu
@pytest.mark.parametrize("pacing", [0, 1])
def test_step_contact_targets_kernel_matches_the_reference(pacing):
    from legged_tracking_b200.gait import GaitClock
    g = load()
    tag = f"gait/pacing{pacing}/"
    d = torch.device("cuda", 0)
    cmd = torch.from_numpy(g[tag + "commands"]).to(d)
    clk = GaitClock(cmd.shape[0], dt=float(g[tag + "dt"]), kappa=float(g[tag + "kappa"]), pacing_offset=bool(pacing), device=d)
    clk.gait_indices.copy_(torch.from_numpy(g[tag + "gait_indices0"]).to(d))
    for k in range(3):
        clk.step(cmd)
        torch.cuda.synchronize()
        for name in GAIT_OUT:
            np.testing.assert_allclose(getattr(clk, name).cpu().numpy(), g[f"{tag}step{k}/{name}"], rtol=RTOL, atol=ATOL, err_msg=f"step{k}/{name}")
        # the phase is a running sum modulo 1: re-seat it on the reference's value so that a last-bit difference cannot walk
        clk.gait_indices.copy_(torch.from_numpy(g[f"{tag}step{k}/gait_indices"]).to(d))


@pytest.mark.gpu
def test_corl_reward_kernel_matches_the_reference():
    from legged_tracking_b200.gait import CoRLRewardKernel
    g = load()
    e, cfg, names = corl_inputs(g)
    d = torch.device("cuda", 0)
    ker = CoRLRewardKernel(names, cfg, dt=e["dt"], feet_indices=e["feet_indices"].tolist(),
                           penalised_contact_indices=e["penalised_contact_indices"].tolist(), device=d)
    dev = {k: (v.to(d) if isinstance(v, torch.Tensor) else v) for k, v in e.items()}
    for call in range(2):
        out = ker(dev)                                   # [n_terms, N]; updates dev["last_contacts"] in place like the reference
        torch.cuda.synchronize()
        for i, n in enumerate(names):
            np.testing.assert_allclose(out[i].cpu().numpy(), g[f"corl/out{call}/{n}"], rtol=RTOL, atol=ATOL, err_msg=f"call{call}/{n}")
        assert np.array_equal(dev["last_contacts"].cpu().numpy().astype(bool), g[f"corl/out{call}/last_contacts"])
