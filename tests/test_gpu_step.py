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


@pytest.fixture(params=["mma", "tc5"])
def torque_impl(request, monkeypatch):
    """lt_create reads LT_TORQUE_TC5 (csrc/lt_kernels.cu): "0" forces the mma.sync actuator-net kernel, "1" the tcgen05 one
    (the kernel bench.py runs at >= 40,960 envs); every parity test that takes this fixture runs under both."""
    monkeypatch.setenv("LT_TORQUE_TC5", "1" if request.param == "tc5" else "0")
    return request.param


@pytest.mark.parametrize("epw", [0, 4, 8, 16, 32])
@pytest.mark.parametrize("name", T.fixture_names())
def test_cuda_step_reproduces_reference(name, epw, torque_impl):
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
    draw for draw with the host-compiled mirror of the same arithmetic (tests/hostsim).  NOT parity evidence: hostsim is
    the same csrc/lt_core.cuh compiled by g++, i.e. the code against itself -- this only pins the RNG draw order / counter
    layout between the device build and the host build."""
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


@pytest.mark.parametrize("n_envs", [3, 11, 43, 67, 1024])
def test_torque_kernel_ragged_tiles_and_float64_reference(n_envs, torque_impl):
    """lt_compute_torques (tensor-pipe actuator net; 32-sample warp tiles in the mma.sync kernel, 128-sample CTA tiles in the
    tcgen05 kernel) on env counts whose 12*N samples do not fill the last tile (12*11 = 128+4, 12*43 = 4*128+4), over enough sub-steps for the error/velocity histories and the 7-slot lag ring to wrap: torques,
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


@pytest.mark.parametrize("config,n_envs", [("cfg2", 4096), ("cfg3", 16384), ("cfg5", 65536)])
def test_cuda_step_matches_port_at_baseline_sizes(config, n_envs):
    """BASELINE.json sizes with the DEFAULT kernel selection (envs_per_warp heuristic, env_tile = i mod 1024 on the 32x32 tile
    grid, lt_obs_kernel's full grid, and at 65,536 envs the tcgen05 torque kernel): three steps of seeded synthetic state with
    injected uniforms, CUDA vs the oracle port.  Masks / reset ids / height cell indices bit-exact, fp32 outputs at the stated
    tolerance.  Episode phases are placed so that time-outs (and for cfg3 the 0.25 m body-height rule) reset envs in every step."""
    from legged_tracking_b200 import cconfig, workload
    from legged_tracking_b200.engine import StepEngine
    from oracle.port import OraclePort
    flat, terrain, act, synth = workload.make(config, num_envs=n_envs)
    assert flat["terrain_num_rows"] * flat["terrain_num_cols"] == 1024
    eng = StepEngine(flat, terrain, act, device="cuda:0", export_height_indices=True)
    port = OraclePort(flat, terrain, act)
    N = n_envs
    ep = synth.initial_episode_length(flat["max_episode_length"])
    ep[::7] = int(flat["max_episode_length"]) - 1
    ep[3::11] = int(flat["max_episode_length"])
    port.episode_length_buf.copy_(ep)
    eng.state["episode_length_buf"].copy_(ep.cuda())
    g = torch.Generator().manual_seed(5)
    n_resets = 0
    for k in range(3):
        U = torch.rand(N, cconfig.rng_n_slots(flat), generator=g)
        noise = torch.rand(N, flat["num_obs"], generator=g)
        gu = torch.rand(3, generator=g)
        actions, root = synth.actions(k), synth.root(k)
        sim = dict(dof_state=[synth.dof_state(k, s) for s in range(flat["decimation"])], root=root, contact=synth.contact(k),
                   rigid=synth.rigid(k, root))
        obs, priv, rew, reset = port.step(actions, sim, U, noise, gu)
        a_dev = actions.cuda().contiguous()
        tq = []
        for s in range(flat["decimation"]):
            tq.append(eng.compute_torques(a_dev, s).clone())
            eng.dof_state.copy_(sim["dof_state"][s].cuda())
        eng.root_states.view(N, -1, 13)[:, 0, :] = root.cuda()
        eng.contact_forces.copy_(sim["contact"].cuda())
        eng.post_physics_step(uniforms=U.cuda(), noise_uniforms=noise.cuda(), gravity_uniform=gu)
        torch.cuda.synchronize()
        tag = f"{config}@{n_envs}/step{k}"
        T.assert_close(f"{tag}/reset_buf", T.to_np(eng.state["reset_buf"]).astype(bool), reset.numpy().astype(bool))
        T.assert_close(f"{tag}/reset_env_ids", T.to_np(eng.reset_env_ids()), port.reset_env_ids.numpy())
        for m in ("time_out_buf", "body_height_buf", "switched_buf", "reached_buf"):
            T.assert_close(f"{tag}/{m}", T.to_np(eng.state[m]).astype(bool), getattr(port, m).numpy().astype(bool))
        for m in ("episode_length_buf", "curr_pose_index", "collision_count"):
            T.assert_close(f"{tag}/{m}", T.to_np(eng.state[m]), getattr(port, m).numpy())
        nx = len(flat["measured_points_x"])
        x0 = nx // 2 + 1 if flat["measure_front_half"] else 0
        T.assert_close(f"{tag}/px", T.to_np(eng.state["height_px"])[:, x0:].astype(np.int64),
                       port.last_px.numpy()[:, x0:, 0].astype(np.int64))
        T.assert_close(f"{tag}/py", T.to_np(eng.state["height_py"]).astype(np.int64), port.last_py.numpy()[:, 0, :].astype(np.int64))
        T.assert_close(f"{tag}/torques", T.to_np(tq[-1]), port.torques.numpy(), atol=T.TORQUE_ATOL)
        T.assert_close(f"{tag}/rew", T.to_np(eng.state["rew_buf"]), rew.numpy())
        T.assert_close(f"{tag}/obs", T.to_np(eng.state["obs_buf"]), obs.numpy())
        T.assert_close(f"{tag}/privileged_obs", T.to_np(eng.state["privileged_obs_buf"]), priv.numpy())
        for m in ("base_lin_vel", "projected_gravity", "relative_linear", "last_actions", "last_dof_vel", "trajectories",
                  "motor_strengths", "motor_offsets"):
            T.assert_close(f"{tag}/{m}", T.to_np(eng.state[m]).reshape(getattr(port, m).shape), getattr(port, m).numpy())
        n_resets += int(reset.sum())
    assert n_resets > N // 20


def test_sharded_engines_bucket_by_the_all_env_sum():
    """LR:332-335 puts a reward term into rew_buf_pos or rew_buf_neg by the sign of its sum over ALL envs.  Two engines that
    each hold a shard must exchange the shard-local sums (StepEngine.term_sum_exchange; across ranks: an all-reduce) or the
    buckets differ from the single-process reference.  The shards are chosen so that the `exploration` term's shard sums have
    OPPOSITE signs; with the exchange both reproduce the reference's rows, without it at least one shard does not."""
    from legged_tracking_b200.engine import StepEngine
    fx = T.load_fixture("ttr_container_32")
    flat = fx["flat"]
    N = flat["num_envs"]
    init, out0, d = T.sub(fx, "init/"), T.sub(fx, "step0/out/state/"), T.sub(fx, "step0/in/")
    done = fx["step0/out/done"]
    term = np.where(done, 0.0, out0["episode_sums/exploration"] - init["episode_sums/exploration"])   # resets flushed their sums
    pos = np.flatnonzero(term > 0)
    shards = [pos, np.setdiff1d(np.arange(N), pos)]
    assert len(pos) >= 2 and term[shards[1]].sum() < 0 < term[shards[0]].sum()

    def sliced(dct, ids):
        o = {}
        for key, v in dct.items():
            v = np.asarray(v)
            if key == "tile_heights":
                o[key] = v
            elif key in ("root_states", "dof_state"):            # Isaac Gym layouts: [N * actors, 13], [N * dofs, 2]
                o[key] = v.reshape(N, -1, v.shape[-1])[ids].reshape(-1, v.shape[-1])
            elif v.ndim >= 1 and v.shape[0] == N:
                o[key] = v[ids]
            elif v.ndim >= 2 and v.shape[1] == N:
                o[key] = v[:, ids]
            else:
                o[key] = v
        return o

    def run(exchange):
        engs = []
        for ids in shards:
            n = len(ids)
            e = StepEngine(dict(flat, num_envs=n), sliced(T.sub(fx, "terrain/"), ids), T.load_actuator(), device="cuda:0")
            e.load_state(sliced(init, ids))
            a = _xp(np.ascontiguousarray(d["actions"][ids], np.float32))
            for s in range(flat["decimation"]):
                e.compute_torques(a, s)
                e.dof_state[...] = _xp(np.ascontiguousarray(d["dof_state"][s].reshape(N, 12, 2)[ids], np.float32).reshape(n * 12, 2))
            e.root_states.reshape(n, -1, 13)[:, 0, :] = _xp(np.ascontiguousarray(d["root"][ids], np.float32))
            e.contact_forces[...] = _xp(np.ascontiguousarray(d["contact"].reshape(N, -1, 3)[ids], np.float32))
            engs.append(e)
        if exchange:                      # what the all-reduce does, with both "ranks" in one process
            total = torch.zeros_like(engs[0].term_sums)
            for e in engs:
                e.sign_prepass()
                total += e.term_sums
            for e in engs:
                e.term_sum_exchange = lambda buf: buf.copy_(total)
        outs = []
        for e, ids in zip(engs, shards):
            U = np.full((len(ids), e.n_slots), 0.5, np.float32)
            U[:, :d["uniforms"].shape[1]] = d["uniforms"][ids]
            e.post_physics_step(uniforms=_xp(U), noise_uniforms=_xp(np.ascontiguousarray(d["noise_uniforms"][ids], np.float32)),
                                gravity_uniform=d.get("gravity_uniform"))
            torch.cuda.synchronize()
            outs.append({m: T.to_np(e.state[m]) for m in ("rew_buf", "rew_buf_pos", "rew_buf_neg")})
        return outs

    want = {m: out0[m] for m in ("rew_buf", "rew_buf_pos", "rew_buf_neg")}
    for o, ids in zip(run(exchange=True), shards):
        for m in want:
            T.assert_close(f"exchange/{m}", o[m], want[m][ids])
    differs = any(np.abs(o[m] - want[m][ids]).max() > 1e-4 for o, ids in zip(run(exchange=False), shards)
                  for m in ("rew_buf_pos", "rew_buf_neg"))
    assert differs, "the shards were meant to disagree with the all-env bucket when they do not exchange their sums"
