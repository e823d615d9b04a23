"""-m gpu parity tests: the CUDA step path (liblt_b200.so through the C ABI) against the golden fixtures recorded
from the unmodified reference, and against the oracle port on larger seeded inputs.

Tolerances (north_star): termination masks, reset ids and heightfield cell indices bit-exact; fp32 outputs within
1e-5 relative (+2e-6 absolute for quantities that pass through zero; torques: +2e-5, i.e. 1e-5 of the actuator
net's contraction magnitude)."""
import numpy as np
import pytest
import torch

from tests import helpers as T

pytestmark = pytest.mark.gpu


def _engine(fx, **kw):
    from legged_tracking_b200.engine import StepEngine
    eng = StepEngine(fx["flat"], T.sub(fx, "terrain/"), T.load_actuator(), device="cuda:0", export_height_indices=True, **kw)
    eng.load_state(T.sub(fx, "init/"))
    return eng


def _xp(a):
    return torch.as_tensor(a).cuda()


@pytest.mark.parametrize("epw", [0, 4, 8, 16, 32])
@pytest.mark.parametrize("name", T.fixture_names())
def test_cuda_step_reproduces_reference(name, epw):
    fx = T.load_fixture(name)
    eng = _engine(fx, envs_per_warp=epw)
    T.check_engine_state(eng, T.sub(fx, "init/"), "init")
    for k in range(T.num_steps(fx)):
        tq = T.engine_feed_and_step(eng, fx, k, _xp)
        torch.cuda.synchronize()
        T.check_engine_step(eng, fx, k, tq)


def test_library_was_loaded_from_tree():
    from legged_tracking_b200 import _abi
    lib = _abi.load()
    assert _abi.LIB_PATH.endswith("legged_tracking_b200/liblt_b200.so")
    assert lib.lt_abi_version() == _abi.DEFINES["LT_ABI_VERSION"]


@pytest.mark.parametrize("name", ["cfg2_single_path_64", "opts_random_target_32"])
def test_in_kernel_philox_matches_host_mirror(name):
    """No injected uniforms: the kernel's counter-based generator (reset / DR slots and observation noise) must agree
    draw for draw with the host-compiled mirror of the same arithmetic (tests/hostsim)."""
    from tests.hostsim_runner import HostSim
    fx = T.load_fixture(name)
    eng = _engine(fx, seed=1234)
    host = HostSim(fx["flat"], T.sub(fx, "terrain/"), T.load_actuator(), seed=1234)
    host.load_state(T.sub(fx, "init/"))
    flat = fx["flat"]
    N = flat["num_envs"]
    for k in range(T.num_steps(fx)):
        d = T.sub(fx, f"step{k}/in/")
        for e, xp in ((eng, _xp), (host, lambda a: a)):
            a = xp(np.ascontiguousarray(d["actions"], np.float32))
            for s in range(flat["decimation"]):
                e.compute_torques(a, s)
                e.dof_state[...] = xp(np.ascontiguousarray(d["dof_state"][s], np.float32).reshape(N * 12, 2))
            e.root_states.reshape(N, -1, 13)[:, 0, :] = xp(np.ascontiguousarray(d["root"], np.float32))
            e.contact_forces[...] = xp(np.ascontiguousarray(d["contact"], np.float32).reshape(N, -1, 3))
            e.post_physics_step(gravity_uniform=np.array([0.25, 0.5, 0.75], np.float32))
        torch.cuda.synchronize()
        for key in ("obs_buf", "rew_buf", "motor_strengths", "motor_offsets", "trajectories", "obs_history"):
            T.assert_close(f"step{k}/{key}", T.to_np(eng.state[key]), host.state[key])
        T.assert_close(f"step{k}/root_states", T.to_np(eng.root_states), host.root_states)
        T.assert_close(f"step{k}/reset_buf", T.to_np(eng.state["reset_buf"]).astype(bool), host.state["reset_buf"].astype(bool))
    if flat["add_noise"]:      # the noise actually is noise: obs differs from the noise-free value, within the scale
        assert np.abs(T.to_np(eng.state["obs_buf"]) - T.sub(fx, f"step{T.num_steps(fx) - 1}/out/")["obs"]).max() > 1e-4


def test_philox_noise_statistics():
    """Uniformity of the in-kernel observation noise: mean 0 and variance scale^2/3 per column over 16k envs."""
    from legged_tracking_b200 import workload
    from legged_tracking_b200.engine import StepEngine
    flat, terrain, act, synth = workload.make("cfg2", num_envs=16384)
    flat_off = dict(flat, add_noise=False)
    outs = []
    for f in (flat, flat_off):
        eng = StepEngine(f, terrain, act, device="cuda:0", seed=7)
        a = synth.actions(0).cuda()
        for s in range(4):
            eng.dof_state.copy_(synth.dof_state(0, s).cuda())
            eng.compute_torques(a, s)
        eng.root_states.view(eng.N, -1, 13)[:, 0, :] = synth.root(0).cuda()
        eng.contact_forces.copy_(synth.contact(0).cuda())
        eng.post_physics_step()
        outs.append(eng.state["obs_buf"].double().cpu())
    noise = outs[0] - outs[1]
    scale = torch.tensor(flat["noise_scale_vec"], dtype=torch.float64)
    cols = scale > 0
    z = noise[:, cols] / scale[cols]                   # should be U(-1, 1)
    assert z.abs().max() <= 1.0 + 1e-3
    assert z.mean(0).abs().max() < 0.03                # sigma of the mean = 0.577 / sqrt(16384) = 0.0045
    assert (z.var(0) - 1.0 / 3.0).abs().max() < 0.02
    # neighbouring columns / envs are uncorrelated
    zc = z - z.mean(0)
    corr = (zc[:, :-1] * zc[:, 1:]).mean(0) * 3.0
    assert corr.abs().max() < 0.05


@pytest.mark.parametrize("n_envs", [3, 67, 1024])
def test_torque_kernel_ragged_tiles_and_float64_reference(n_envs):
    """lt_compute_torques (tensor-pipe actuator net, 32-sample warp tiles) on env counts whose 12*N samples do not fill
    the last tile, over enough sub-steps for the error/velocity histories and the 7-slot lag ring to wrap: torques,
    joint targets and the shifted histories against the oracle port; the network itself against a float64 evaluation
    of the same weights on the same fp32 inputs (what the 3xTF32 split has to reproduce to 1e-5 of the contraction)."""
    from legged_tracking_b200 import workload
    from legged_tracking_b200.engine import StepEngine
    from oracle.port import OraclePort
    flat, terrain, act, synth = workload.make("cfg2", num_envs=n_envs, tile_grid=(1, 1))
    eng = StepEngine(flat, terrain, act, device="cuda:0")
    port = OraclePort(flat, terrain, act)
    W = {k: torch.as_tensor(v, dtype=torch.float64) for k, v in act.items()}
    worst = 0.0
    for k in range(3):
        a = synth.actions(k)
        a_dev = a.cuda().contiguous()
        port.actions = torch.clip(a, -flat["clip_actions"], flat["clip_actions"])
        for s in range(flat["decimation"]):
            ds = synth.dof_state(k, s)
            eng.dof_state.copy_(ds.cuda())
            port.dof_state.copy_(ds)
            x64 = None
            if flat["control_type"] == "actuator_net":
                hist = [T.to_np(eng.state[q]).copy() for q in ("joint_pos_err_last", "joint_pos_err_last_last", "joint_vel_last",
                                                               "joint_vel_last_last")]
            want = port.compute_torques(port.actions)
            got = eng.compute_torques(a_dev, flat["decimation"] * k + s)    # no post-physics step in between: the ring head advances with the sub-step index
            torch.cuda.synchronize()
            T.assert_close(f"step{k}.{s}/torques", T.to_np(got), want.numpy(), atol=T.TORQUE_ATOL)
            T.assert_close(f"step{k}.{s}/joint_pos_target", T.to_np(eng.state["joint_pos_target"]), port.joint_pos_target.numpy())
            for q in ("joint_pos_err_last", "joint_pos_err_last_last", "joint_vel_last", "joint_vel_last_last"):
                T.assert_close(f"step{k}.{s}/{q}", T.to_np(eng.state[q]), getattr(port, q).numpy())
            # float64 network on the exact fp32 inputs the kernel saw
            x = np.stack([T.to_np(eng.state["joint_pos_err_last"]), hist[0], hist[1], T.to_np(eng.state["joint_vel_last"]), hist[2],
                          hist[3]], -1).reshape(-1, 6)
            x64 = torch.as_tensor(x, dtype=torch.float64)
            ss = torch.nn.functional.softsign
            y = ss(ss(x64 @ W["w0"].T + W["b0"]) @ W["w1"].T + W["b1"]) @ W["w2"].reshape(1, -1).T + W["b2"]
            y = y.view(n_envs, 12) * eng.state["motor_strengths"].double().cpu()
            lim = torch.as_tensor(flat["torque_limits"], dtype=torch.float64)
            y = torch.minimum(torch.maximum(y, -lim), lim)
            worst = max(worst, float((y - got.double().cpu()).abs().max()))
    print(f"torque kernel vs float64 network, N={n_envs}: max abs err {worst:.3e}")
    assert worst < T.TORQUE_ATOL          # measured: 9e-6 on torques of up to 33.5 N m
