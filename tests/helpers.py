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


def fixture_names(planning=False):
    """Env-step fixtures.  planning=True adds the sampling-based-planning recording, which only the oracle / config tests and
    tests/test_planner.py take: the per-env host mirror has no planner, and the CUDA planner resolves the reference's tied sort
    keys by its own documented rule (tests/test_planner.py)."""
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                  if not os.path.basename(p).startswith(("actuator", "ppo_", "terrain_", "planner_", "gait_", "commands_", "vel_") + (() if planning else ("planning_",))))


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


# ------------------------------------------------------------------------------------------------------------
# Shared driver for the engines that expose the StepEngine call surface (product: legged_tracking_b200.engine on
# CUDA; test scaffolding: tests/hostsim_runner.HostSim).
STATE_SKIP = {"gravity_vec", "common_step_counter", "camera_pitch_angle"}
TORQUE_ATOL = 1e-5      # north_star: 1e-5 relative for torques; absolute floor 1e-5 N m on a +-33.5 N m range (measured worst 9e-6)


def to_np(x):
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.asarray(x)


def engine_feed_and_step(eng, fx, k, xp):
    """Replays step k of fixture fx through `eng`.  xp(x) converts a numpy array to the engine's array type."""
    d = sub(fx, f"step{k}/in/")
    flat = fx["flat"]
    N = flat["num_envs"]
    actions = xp(np.ascontiguousarray(d["actions"], np.float32))
    torques = []
    for s in range(flat["decimation"]):
        t = eng.compute_torques(actions, s)
        torques.append(to_np(t).copy())
        eng.dof_state[...] = xp(np.ascontiguousarray(d["dof_state"][s], np.float32).reshape(N * flat["num_dof"], 2))
    eng.root_states.reshape(N, -1, 13)[:, 0, :] = xp(np.ascontiguousarray(d["root"], np.float32))
    eng.contact_forces[...] = xp(np.ascontiguousarray(d["contact"], np.float32).reshape(N, -1, 3))
    eng.rigid_body_state[...] = xp(np.ascontiguousarray(d["rigid"], np.float32).reshape(-1, 13))
    U = np.full((N, eng.n_slots), 0.5, np.float32)
    U[:, :d["uniforms"].shape[1]] = d["uniforms"]
    eng.post_physics_step(uniforms=xp(U),
                          noise_uniforms=xp(np.ascontiguousarray(d["noise_uniforms"], np.float32)) if "noise_uniforms" in d else None,
                          gravity_uniform=d.get("gravity_uniform"))
    return np.stack(torques)


def check_engine_step(eng, fx, k, all_torques):
    out = sub(fx, f"step{k}/out/")
    st = eng.state
    assert_close(f"step{k}/done", to_np(st["reset_buf"]).astype(bool), out["done"])
    assert_close(f"step{k}/reset_env_ids", to_np(eng.reset_env_ids()), out["reset_env_ids"])
    if "px" in out and "height_px" in st:
        L = fx["flat"]
        nx = len(L["measured_points_x"])
        x0 = nx // 2 + 1 if L["measure_front_half"] else 0
        assert_close(f"step{k}/px", to_np(st["height_px"])[:, x0:].astype(np.int64), out["px"][:, x0:, 0])
        assert_close(f"step{k}/py", to_np(st["height_py"]).astype(np.int64), out["py"][:, 0, :])
    assert_close(f"step{k}/all_torques", all_torques, out["all_torques"], atol=TORQUE_ATOL)
    assert_close(f"step{k}/rew", to_np(st["rew_buf"]), out["rew"])
    assert_close(f"step{k}/obs", to_np(st["obs_buf"]), out["obs"])
    assert_close(f"step{k}/privileged_obs", to_np(st["privileged_obs_buf"]), out["privileged_obs"])
    assert_close(f"step{k}/obs_history", to_np(st["obs_history"]), out["obs_history"])
    want = sub(fx, f"step{k}/out/state/")
    if want:
        check_engine_state(eng, want, f"step{k}/state")


def check_engine_state(eng, want, tag):
    got = eng.export_state()
    for key, w in want.items():
        if key in STATE_SKIP:
            continue
        if key.startswith("command_sums/") and key not in got:
            assert not np.any(w), f"{tag}/{key}: reference row is non-zero but the engine does not track it"
            continue
        assert key in got, f"{tag}: engine has no state {key}"
        g = to_np(got[key])
        atol = TORQUE_ATOL if key in ("torques",) else ATOL
        assert_close(f"{tag}/{key}", g.reshape(np.asarray(w).shape), w, atol=atol)
    assert_close(f"{tag}/gravity_vec", to_np(got["gravity_vec"]).reshape(-1), np.asarray(want["gravity_vec"]).reshape(-1))
    assert int(got["common_step_counter"]) == int(want["common_step_counter"])
