"""Shared test plumbing: fixture loading, oracle construction, comparison with stated tolerances."""
import glob
import json
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# north_star tolerance: fp32 outputs within 1e-5 relative; masks / indices bit-exact.
# The absolute floor covers quantities that pass through zero (wrapped angles, differences).
RTOL = 1e-5
ATOL = 2e-6


def fixture_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                  if not os.path.basename(p).startswith("actuator"))


def load_fixture(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    fx = {k: z[k] for k in z.files}
    fx["flat"] = json.loads(str(fx["flat"]))
    return fx


def load_actuator():
    z = np.load(os.path.join(GOLDEN, "actuator_net_go1.npz"))
    return {k: z[k] for k in z.files}


def sub(fx, prefix):
    return {k[len(prefix):]: v for k, v in fx.items() if isinstance(k, str) and k.startswith(prefix)}


def num_steps(fx):
    k = 0
    while f"step{k}/in/actions" in fx:
        k += 1
    return k


def make_oracle(fx, **kw):
    from oracle.port import OraclePort
    port = OraclePort(fx["flat"], sub(fx, "terrain/"), load_actuator(), **kw)
    port.load_state({k: torch.as_tensor(v) for k, v in sub(fx, "init/").items()})
    return port


def step_inputs(fx, k):
    d = sub(fx, f"step{k}/in/")
    sim = dict(dof_state=torch.as_tensor(d["dof_state"]), root=torch.as_tensor(d["root"]),
               contact=torch.as_tensor(d["contact"]), rigid=torch.as_tensor(d["rigid"]))
    return dict(actions=torch.as_tensor(d["actions"]), sim=sim, uniforms=torch.as_tensor(d["uniforms"]),
                noise_uniforms=None if "noise_uniforms" not in d else torch.as_tensor(d["noise_uniforms"]),
                gravity_uniform=None if "gravity_uniform" not in d else torch.as_tensor(d["gravity_uniform"]))


def assert_close(name, got, want, rtol=RTOL, atol=ATOL):
    got = np.asarray(got)
    want = np.asarray(want)
    assert got.shape == want.shape, f"{name}: shape {got.shape} vs {want.shape}"
    if want.dtype == bool or np.issubdtype(want.dtype, np.integer):
        bad = np.argwhere(got.astype(want.dtype) != want)
        assert bad.size == 0, f"{name}: {len(bad)} mismatching entries (bit-exact required), first at {bad[0]}: " \
                              f"{got[tuple(bad[0])]} vs {want[tuple(bad[0])]}"
        return
    err = np.abs(got.astype(np.float64) - want.astype(np.float64))
    tol = atol + rtol * np.abs(want.astype(np.float64))
    bad = np.argwhere(err > tol)
    assert bad.size == 0, f"{name}: {len(bad)}/{want.size} entries out of tolerance (rtol={rtol}, atol={atol}); worst " \
                          f"abs err {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}: " \
                          f"{got[np.unravel_index(err.argmax(), err.shape)]} vs {want[np.unravel_index(err.argmax(), err.shape)]}"
