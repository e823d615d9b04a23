"""Pin the oracle: OraclePort must reproduce the fixtures recorded from the UNMODIFIED reference.

Runs without a GPU and without /root/reference.  Free-running: the port is only given the initial state and
the per-step inputs, never the reference's intermediate state.
"""
import numpy as np
import pytest
import torch

from tests import helpers as T

STATE_SKIP = {"gravity_vec", "common_step_counter"}


def compare_state(port, want, tag):
    got = port.export_state()
    for k, w in want.items():
        if k in STATE_SKIP:
            continue
        if k == "obs_history":
            g = port.obs_history
        elif k not in got:
            raise AssertionError(f"{tag}: oracle has no state {k}")
        else:
            g = got[k]
        T.assert_close(f"{tag}/{k}", g.detach().numpy().reshape(np.asarray(w).shape), w)
    T.assert_close(f"{tag}/gravity_vec", port.gravity_vec[0].numpy(), want["gravity_vec"])
    assert port.common_step_counter == int(want["common_step_counter"])


@pytest.mark.parametrize("name", T.fixture_names(planning=True))
def test_port_reproduces_reference(name):
    fx = T.load_fixture(name)
    port = T.make_oracle(fx)
    compare_state(port, T.sub(fx, "init/"), "init")           # load/export round trip
    for k in range(T.num_steps(fx)):
        inp = T.step_inputs(fx, k)
        obs, priv, rew, reset = port.step(inp["actions"], inp["sim"], inp["uniforms"], inp["noise_uniforms"],
                                          inp["gravity_uniform"])
        out = T.sub(fx, f"step{k}/out/")
        T.assert_close(f"step{k}/done", reset.numpy(), out["done"])
        T.assert_close(f"step{k}/reset_env_ids", port.reset_env_ids.numpy(), out["reset_env_ids"])
        if "px" in out:
            T.assert_close(f"step{k}/px", port.last_px.numpy(), out["px"])
            T.assert_close(f"step{k}/py", port.last_py.numpy(), out["py"])
        T.assert_close(f"step{k}/all_torques", torch.stack(port.all_torques).numpy(), out["all_torques"])
        T.assert_close(f"step{k}/rew", rew.numpy(), out["rew"])
        T.assert_close(f"step{k}/obs", obs.numpy(), out["obs"])
        T.assert_close(f"step{k}/privileged_obs", priv.numpy(), out["privileged_obs"])
        T.assert_close(f"step{k}/obs_history", port.obs_history.numpy(), out["obs_history"])
        st = T.sub(fx, f"step{k}/out/state/")
        if st:
            compare_state(port, st, f"step{k}/state")
