#!/usr/bin/env python
"""bench.py -- env-steps/s of the fused step path (BASELINE.json metric) on N B200s of one node.

A "step" is one pass of LeggedRobot.step's data-parallel work over all envs of the rank:
4 x lt_compute_torques (actuator net, one per physics sub-step) + lt_post_physics_step (fused step kernel + ordered
reset compaction), on seeded synthetic simulator state in Isaac Gym's tensor layout (PhysX itself is out of scope).
Default workload: BASELINE config 5 shape (single_path tunnel, 2x10x11 front-half heights, camera_zero, e2e reward
set, noise + domain randomisation on), 65,536 envs PER GPU (weak scaling: env shards are independent).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--envs-per-gpu E]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 ... bench.py --gpus N ...
"""
import argparse
import json
import os
import sys
import threading
import time

# rank 0 prints exactly ONE line on stdout.  NCCL_DEBUG is left as the caller set it: _quiet_stdout() below points fd 1 at stderr
# for the whole run, so NCCL's banner / INFO log cannot land on stdout.

import numpy as np
import torch

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

from legged_tracking_b200 import workload  # noqa: E402

METRIC = "env_steps_per_s"
UNIT = "env-steps/s"
N_REPLAY_SETS = 3


_STDOUT_FD = None


def _quiet_stdout():
    """Libraries (NCCL's version banner, torch's C++ logging) write to fd 1; the contract is ONE JSON line on stdout.  Point fd 1
    at stderr for the duration of the run and keep the real stdout for _emit."""
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)


def _emit(text):
    sys.stdout.flush()
    if _STDOUT_FD is None:
        print(text, flush=True)
    else:
        os.write(_STDOUT_FD, (text + "\n").encode())


def ncu_traffic():
    """DRAM bytes per launch of each kernel from the committed ncu --set full capture (profiles/ncu_traffic.json, written by
    profiles/ncu_summary.py --traffic); {} if the file is absent."""
    p = os.path.join(REPO, "profiles", "ncu_traffic.json")
    return json.load(open(p)) if os.path.exists(p) else {}


def measured_peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index, period=0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self.stop_flag = False
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                 getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                 getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                 getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


class ReplaySets:
    """A few pre-generated 'PhysX outputs' resident in HBM; the step is re-pointed at one of them per step."""

    def __init__(self, eng, synth, flat, n_sets, pinned_host=False):
        N, A, D = eng.N, flat["num_actors"], flat["num_dof"]
        dev = eng.device
        self.sets = []
        self.host = []
        for k in range(n_sets):
            root_all = torch.zeros(N, A, 13)
            root_all[:, :, 6] = 1.0
            root_all[:, 0, :] = synth.root(k)
            h = {"actions": synth.actions(k).contiguous(),
                 "dof": [synth.dof_state(k, s) for s in range(flat["decimation"])],
                 "root": root_all.view(N * A, 13).contiguous(), "contact": synth.contact(k)}
            if pinned_host:
                h = {"actions": h["actions"].pin_memory(), "dof": [d.pin_memory() for d in h["dof"]],
                     "root": h["root"].pin_memory(), "contact": h["contact"].pin_memory()}
                self.host.append(h)
            self.sets.append({"actions": h["actions"].to(dev), "dof": [d.to(dev) for d in h["dof"]],
                              "root": h["root"].to(dev), "contact": h["contact"].to(dev)})


def launch_step(eng, s, eager_gravity=True):
    """One full step on replay set s (decimation torque launches + step, obs and compaction kernels)."""
    for sub, dof in enumerate(s["dof"]):
        eng.bind_sim(dof_state=dof)
        eng.compute_torques(s["actions"], sub)
    eng.bind_sim(root_states=s["root"], contact_forces=s["contact"])
    eng.post_physics_step(advance_gravity=eager_gravity)


def init_engine_state(eng, synth, flat):
    """Spread episode phases like TrajectoryTrackingEnv.reset (trajectory_tracking/__init__.py:49) and give every env
    a goal so that the relative-pose math runs on meaningful numbers."""
    N = eng.N
    eng.state["episode_length_buf"].copy_(synth.initial_episode_length(flat["max_episode_length"]).to(eng.device))
    g = torch.Generator().manual_seed(5)
    eng.state["friction_coeffs"].copy_((0.05 + 4.45 * torch.rand(N, 1, generator=g)).to(eng.device))
    eng.state["restitutions"].copy_((0.4 * torch.rand(N, 1, generator=g)).to(eng.device))
    traj = torch.zeros(N, flat["traj_length"], 6)
    traj[:, :, 0] = eng.env_origins[:, 0:1].cpu() + 2.5
    traj[:, :, 1] = eng.env_origins[:, 1:2].cpu()
    traj[:, :, 2] = 0.34
    eng.state["trajectories"].copy_(traj.to(eng.device))


def time_region(fn, steps, stream_sync=True):
    """CUDA-event time (ms) of `steps` calls of fn(k) on the current stream, synchronised on both sides."""
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(steps):
        fn(k)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def cpu_oracle_rate(steps, n_envs, config, warmup=1):
    """The reference's torch step path on the host cores (oracle port, faithful per-env height gather): env-steps/s."""
    from oracle.port import OraclePort
    from legged_tracking_b200 import cconfig
    torch.set_num_threads(os.cpu_count() or 1)
    flat, terrain, act, synth = workload.make(config, num_envs=n_envs)
    port = OraclePort(flat, terrain, act, faithful_gather=True)
    n = flat["num_envs"]
    U = torch.full((n, cconfig.rng_n_slots(flat)), 0.5)
    inputs = []
    for k in range(2):
        root = synth.root(k)
        inputs.append((synth.actions(k), dict(dof_state=[synth.dof_state(k, s) for s in range(flat["decimation"])], root=root,
                                              contact=synth.contact(k), rigid=synth.rigid(k, root))))
    port.episode_length_buf.copy_(synth.initial_episode_length(flat["max_episode_length"]))

    def one(k):
        a, sim = inputs[k % 2]
        port.step(a, sim, U, torch.rand(n, flat["num_obs"]) if flat["add_noise"] else None, torch.rand(3))
    for k in range(max(warmup, 1)):
        one(k)
    t0 = time.perf_counter()
    for k in range(steps):
        one(k + 1)
    dt = time.perf_counter() - t0
    return n * steps / dt, dt / steps, torch.get_num_threads()


def workload_config(args, world, n_env, touched_mb=None):
    """The `config` object of the JSON line: names the workload; identical for the B200 arm and the reference arm."""
    return {"workload": f"{args.config}: single_path tunnel, 2x10x11 front-half heights, camera_zero, e2e reward set, "
                        f"noise+DR on; full step = 4 x actuator-net torques + fused post-physics step",
            "envs_per_gpu": n_env, "total_envs": n_env * world, "parallelism": f"env-sharded x{world}, no data-path collective",
            "cuda_graph": not args.no_graph, "envs_per_warp": args.epw,
            "l2": f"inputs larger than L2: {touched_mb if touched_mb is not None else 484:.0f} MB touched per step, "
                  f"{N_REPLAY_SETS} replay sets cycled"}


def run_reference(args, rank):
    """--impl reference: the reference's own CPU implementation of the path on the box's host cores, all threads.  The UNMODIFIED
    reference (staged by oracle/make_ref.sh under oracle/_ref/, which travels to the GPU box) when present, else the oracle port.
    Same configuration and metric as the B200 arm; each step is a BOUNDED sample of the workload (--cpu-envs envs instead of
    65,536: the reference gathers heights with one kernel launch per env, ~50 us per env-step on these cores, so a full-size
    step would take > 3 s and the run many minutes).  --steps / --warmup are honoured as given."""
    if rank != 0:
        return
    from oracle import ref_arm
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    # size the per-step sample so that the whole run stays within a few minutes (~50 us per env-step on these cores)
    # in multiples of the 32 x 32 tile grid (the reference asserts num_envs % 1024 == 0, LR:1819)
    n_envs = int(max(1024, min(args.cpu_envs, (200.0 / ((steps + warmup) * 50e-6))) // 1024 * 1024))
    if ref_arm.available():
        rate, s_per_step, cores = ref_arm.step_rate(steps, warmup, n_envs)
        kind, how = "reference", "unmodified reference (oracle/_ref: go1_gym HistoryWrapper(TrajectoryTrackingEnv).step under the isaacgym shims)"
    else:
        rate, s_per_step, cores = cpu_oracle_rate(steps, n_envs, "cfg2", warmup=warmup)
        kind, how = "port", "oracle/port.py with the reference's per-env height gather (oracle/_ref not staged)"
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": s_per_step * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, max(1, args.gpus), args.envs_per_gpu),
            "sample": f"bounded sample: {n_envs} of the {args.envs_per_gpu} envs per step (CPU time is linear in envs)",
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"{n_envs} envs x {steps} steps after {warmup} warm-up steps, {how}"},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    _emit(json.dumps(line))


def ppo_iteration(eng, rs, graphs, flat, dev, dist, iters):
    """BASELINE metric part 2: ms of one PPO iteration = T=24-step rollout (policy forward, transition store, full env step)
    + GAE + 5 epochs x 4 minibatches of the actor-critic update (+ gradient / KL / advantage-moment all-reduces for N>1)."""
    from legged_tracking_b200 import ppo
    N, T = eng.N, ppo.RunnerArgs.num_steps_per_env
    no, npv, H = flat["num_obs"], flat["num_privileged_obs"], flat["obs_history_length"]
    torch.manual_seed(0)
    ac = ppo.ActorCritic(no, npv, no * H, flat["num_actions"]).to(dev)
    alg = ppo.PPO(ac, device=dev)
    alg.init_storage(N, T, [no], [npv], [no * H], [flat["num_actions"]])
    st = eng.state
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

    def iteration(k0):
        ev[0].record()
        with torch.inference_mode():
            for t in range(T):
                actions = alg.act(st["obs_buf"], st["privileged_obs_buf"], st["obs_history"])
                s = (k0 + t) % N_REPLAY_SETS
                rs.sets[s]["actions"].copy_(actions)
                if graphs and eng.graph_replay_ok():
                    graphs[s].replay()
                    eng.gravity.common_step_counter += 1
                else:
                    launch_step(eng, rs.sets[s])
                alg.process_env_step(st["rew_buf"], st["reset_buf"], {"time_outs": st["time_outs_latched"]})
            ev[1].record()
            alg.compute_returns(st["obs_history"], st["privileged_obs_buf"])
        ev[2].record()
        losses = alg.update()
        ev[3].record()
        torch.cuda.synchronize()
        return [ev[i].elapsed_time(ev[i + 1]) for i in range(3)] + [alg.exchange.seconds() * 1e3], losses
    alg.exchange.timed = True
    iteration(0)                                             # warm-up (workspaces, allocator, NCCL channels)
    if dist:
        dist.barrier()
    alg.exchange.calls = alg.exchange.bytes = 0
    parts, losses = [], None
    for it in range(iters):
        p_, losses = iteration(1 + it * T)
        parts.append(p_)
    parts = np.asarray(parts).mean(axis=0)
    tot = torch.tensor([float(parts[:3].sum())] + [float(x) for x in parts], device=dev)
    if dist:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    tot = tot.tolist()
    calls, nbytes = alg.exchange.calls // max(iters, 1), alg.exchange.bytes // max(iters, 1)
    n_params = ac.num_params
    flop_fwd = 2.0 * sum(int(np.prod(shape)) for key, shape, _, _ in ac.layout if key.endswith(".weight"))
    mb_rows = N * T // ppo.PPO_Args.num_mini_batches
    n_mb = ppo.PPO_Args.num_learning_epochs * ppo.PPO_Args.num_mini_batches
    ad_w = 2.0 * sum(int(np.prod(shape)) for key, shape, _, _ in ac.layout if key.startswith("adaptation_module") and key.endswith(".weight"))
    update_flop = n_mb * mb_rows * (3.0 * flop_fwd + 3.0 * ad_w)         # fwd + dX + dW of the three bodies, + the adaptation step
    del alg, ac
    torch.cuda.empty_cache()
    return {"iter_ms": tot[0], "rollout_ms": tot[1], "gae_ms": tot[2], "update_ms": tot[3], "iters_timed": iters,
            "allreduce_ms": tot[4], "allreduce_calls": calls, "allreduce_bytes": nbytes,
            "num_steps_per_env": T, "epochs": ppo.PPO_Args.num_learning_epochs, "mini_batches": ppo.PPO_Args.num_mini_batches,
            "samples_per_iter_per_gpu": N * T, "params": n_params,
            "gemm": "hand-written tcgen05 kernel (csrc/lt_gemm.cuh): fp16-split x3 products, fp32 accumulation in TMEM, TMA-fed",
            "update_fp32_equiv_tflops": update_flop / (tot[3] * 1e-3) / 1e12 if tot[3] > 0 else None,
            "update_tensor_tflops": 3.0 * update_flop / (tot[3] * 1e-3) / 1e12 if tot[3] > 0 else None,
            "finite": bool(all(np.isfinite(x) for x in losses))}


def api_step_timing(args, dev, n_env, steps):
    """env-steps/s through the facade the north-star says stays the same: HistoryWrapper(TrajectoryTrackingEnv).step(actions) with a
    device-resident simulator stand-in (env.ReplaySim cycling frames in HBM).  Eager launches from Python, per-step extras, the
    4-tuple / obs-dict API: what a Runner sees.  Device-timed and wall-timed (the wall figure includes the Python overhead the
    device hides behind the asynchronous queue)."""
    from legged_tracking_b200 import cfg as ltcfg, env as ltenv
    preset = workload.CONFIGS[args.config][0]
    c = ltcfg.preset_cfg(preset, num_envs=n_env)
    _, terrain, _, _ = workload.make(args.config, num_envs=n_env)
    if c.terrain.mesh_type != "plane":
        c.terrain.num_rows, c.terrain.num_cols = 32, 32
    actions = torch.randn(n_env, 12, device=dev).clamp_(-3, 3)
    out = {"unit": UNIT, "steps": steps,
           "path": "HistoryWrapper(TrajectoryTrackingEnv).step, eager launches, device-resident ReplaySim (3 frames cycled)"}
    # (a) the reference's behaviour: finished-episode statistics appended to the extras deques on the host every step (LR:256-291);
    #     the synthetic state resets ~14 % of the envs per step (a trained policy: < 1 %), so this figure is host-logging bound
    # (b) log_episodes=False, alias_outputs=True: the facade's own cost (Python launch sequence, 4-tuple / obs-dict, lazy extras)
    for key, kw in (("with_episode_logging", dict()), ("facade_only", dict(log_episodes=False, alias_outputs=True))):
        e = ltenv.TrajectoryTrackingEnv(sim_device=dev, headless=True, cfg=c, terrain=terrain,
                                        sim_backend=ltenv.ReplaySim(seed=77, device_frames=3), **kw)
        w = ltenv.HistoryWrapper(e)
        w.reset()
        for _ in range(3):
            w.step(actions)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ms = time_region(lambda k: w.step(actions), steps)
        wall = time.perf_counter() - t0
        out[key] = {"value": n_env * steps / (ms * 1e-3), "ms_per_step": ms / steps, "wall_ms_per_step": wall * 1e3 / steps}
        del w, e
        torch.cuda.empty_cache()
    out["value"] = out["facade_only"]["value"]
    return out


def rollout_kernel_timing(flat, dev, n_env, peak):
    """lt_rollout_store (observation third of a transition + the env third), lt_gae + lt_normalize_advantages, timed alone with
    CUDA events (20 launches each, inputs larger than L2 across the loop: 24 different rows of the [T, N, *] buffers)."""
    import ctypes as C
    from legged_tracking_b200 import _abi, ppo
    no, npv, H, na = flat["num_obs"], flat["num_privileged_obs"], flat["obs_history_length"], flat["num_actions"]
    T = ppo.RunnerArgs.num_steps_per_env
    st = ppo.RolloutStorage(n_env, T, [no], [npv], [no * H], [na], device=dev)
    g = torch.Generator(device=dev).manual_seed(3)
    tr = ppo.RolloutStorage.Transition()
    tr.observations = torch.randn(n_env, no, device=dev, generator=g)
    tr.privileged_observations = torch.randn(n_env, npv, device=dev, generator=g)
    tr.observation_histories = torch.randn(n_env, no * H, device=dev, generator=g)
    tr.rewards = torch.randn(n_env, device=dev, generator=g)
    tr.dones = torch.rand(n_env, device=dev, generator=g) < 0.05
    tr.time_outs = tr.dones.clone()
    st.values.normal_(generator=g)
    st.rewards.normal_(generator=g)

    def store(k):
        st.step = k % T
        tr.values = st.values[st.step]
        st.store_observations(tr)
        st.store_env_step(tr)
    store(0)
    ms_store = time_region(store, T) / T
    last = torch.randn(n_env, 1, device=dev, generator=g)

    def gae(k):
        st.compute_returns(last, 0.99, 0.95)
    gae(0)
    ms_gae = time_region(gae, 10) / 10
    b_store = 4 * (no + npv + no * H) * 2 + (4 + 4 + 1 + 1 + 4 + 1)            # read + write of the copied rows; rewards / values / masks
    b_gae = T * (4 + 4 + 1 + 4 + 4) + T * 8                                      # lt_gae: r, V, done read; R, A written; normalise: A read + written
    out = {}
    for name, ms, b, what in (("rollout_store", ms_store, b_store, "lt_rollout_store_kernel x2 (observation rows at act(), rewards / dones after the step)"),
                              ("gae", ms_gae, b_gae, "lt_gae_kernel + lt_normalize_kernel over [24, N]")):
        ach = b * n_env / (ms * 1e-3) / 1e9
        out[name] = {"kernel": what, "ms_per_launch": ms, "achieved": ach, "bound": "hbm", "peak": peak, "unit": "GB/s", "frac": ach / peak,
                     "algorithmic_bytes_per_env": b}
    del st
    torch.cuda.empty_cache()
    return out


def terrain_build_timing(dev):
    """the 32 x 32 single_path tile table built on the device (the reference: a 3.8 s Python loop)."""
    from legged_tracking_b200 import cfg as ltcfg
    from legged_tracking_b200.terrain import TunnelTerrain
    cfg = ltcfg.preset_cfg("cfg2_single_path_64").terrain
    cfg.num_rows, cfg.num_cols = 32, 32
    t0 = time.perf_counter()
    t = TunnelTerrain(cfg, device=dev, rng=np.random.RandomState(11))
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    # kernel time: 20 launches captured in one CUDA graph (an eager Python loop would time ctypes, not the 10-us kernel)
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for _ in range(20):
            t.rebuild()
    g.replay()
    ms = time_region(lambda k: g.replay(), 10) / 200
    nbytes = t.tile_heights.numel() * 4
    return {"tiles": 1024, "kernel_us": ms * 1e3, "written_gb_per_s": nbytes / (ms * 1e-3) / 1e9,
            "host_draws_plus_kernel_ms": wall * 1e3, "bytes": nbytes}


def cfg4_step_timing(dev, n_envs, steps=60):
    """BASELINE.json configs[3]: the velocity-tracking env's full step (4 x torques + the post-physics launch chain of
    legged_tracking_b200/vel_env.py, 17 x 11 height grid measured on a synthetic heightfield, 70 x 30 observation history) at
    `n_envs`, driven from Python like the reference drives its own (three nonzero() host reads per step, resets and command
    resampling through indexed torch writes).  Device time per step, inputs resident in HBM."""
    from legged_tracking_b200 import synthetic, vel_env
    cfg = vel_env.cfg4_config()
    side = int(np.ceil(np.sqrt(n_envs)))
    xx, yy = np.meshgrid(np.arange(side), np.arange(side), indexing="ij")
    origins = np.zeros((n_envs, 3), np.float32)
    origins[:, 0], origins[:, 1] = 3.0 * xx.flatten()[:n_envs], 3.0 * yy.flatten()[:n_envs]
    rs = np.random.RandomState(5)
    field = dict(height_samples=rs.randint(-10, 10, size=(int(3.0 * side / 0.1) + 64, int(3.0 * side / 0.1) + 64)).astype(np.int16),
                 border_size=2.0, horizontal_scale=0.1, vertical_scale=0.005, measured_points_x=vel_env.CFG4_MEASURED_POINTS_X,
                 measured_points_y=vel_env.CFG4_MEASURED_POINTS_Y)
    eng = vel_env.VelocityStepEngine(cfg, n_envs, origins, workload.load_actuator(),
                                     vel_env.UniformCommandSource(vel_env.CFG4_COMMAND_LOW, vel_env.CFG4_COMMAND_HIGH, seed=1), device=dev,
                                     seed=9, heightfield=field)
    synth = synthetic.SyntheticState(n_envs, origins, num_bodies=int(cfg["num_bodies"]), seed=77, feet_indices=list(cfg["feet_indices"]))
    eng.episode_length_buf.copy_(synth.initial_episode_length(cfg["max_episode_length"]).to(dev))
    frames = []
    for k in range(3):
        root = synth.root(k)
        frames.append({"actions": synth.actions(k).contiguous().to(dev), "root": root.to(dev), "contact": synth.contact(k).to(dev),
                       "rigid": synth.rigid(k, root).to(dev), "dof_state": [synth.dof_state(k, s).to(dev) for s in range(int(cfg["decimation"]))]})
    eng.reset_idx(torch.arange(n_envs, device=dev))                    # first commands for every env, as env.reset() does

    def one(k):
        f = frames[k % 3]
        eng.step(f["actions"], f)
    for k in range(5):
        one(k)
    resets0 = int(eng.reset_buf.sum())
    t0 = time.perf_counter()
    ms = time_region(one, steps)
    wall = time.perf_counter() - t0
    ms_chain = time_region(lambda k: eng.step(frames[k % 3]["actions"], frames[k % 3], host_paths=False), steps)
    # device mode: curriculum, resampling, DR and reset as kernels over masks -- no host read in the step
    dev_eng = vel_env.VelocityStepEngine(cfg, n_envs, origins, workload.load_actuator(), "device", device=dev, seed=9, heightfield=field)
    dev_eng.episode_length_buf.copy_(synth.initial_episode_length(cfg["max_episode_length"]).to(dev))
    dev_eng.bind_inputs = True                 # the replayed frames are read (and, on reset, rewritten) in place, as Isaac Gym's tensors are
    dev_eng.reset_all()
    for k in range(5):
        dev_eng.step(frames[k % 3]["actions"], frames[k % 3])
    resets_dev = int(dev_eng.reset_buf.sum())
    t0 = time.perf_counter()
    ms_dev = time_region(lambda k: dev_eng.step(frames[k % 3]["actions"], frames[k % 3]), steps)
    wall_dev = time.perf_counter() - t0
    # ... and the same step replayed as CUDA graphs of two steps each (eager on the steps around a gravity event)
    pair = [(frames[0]["actions"], frames[0]), (frames[1]["actions"], frames[1])]
    while not dev_eng.graph_replay_ok(2):
        dev_eng.step(*pair[0])
    graph2 = dev_eng.capture(pair)

    def two_steps(k):
        if dev_eng.graph_replay_ok(2):
            dev_eng.replay(graph2)
        else:
            for a_, f_ in pair:
                dev_eng.step(a_, f_)
    two_steps(0)
    ms_graph = time_region(two_steps, steps // 2) / 2
    no, H = int(cfg["num_obs"]), int(cfg["obs_history_length"])
    hist_bytes = 2 * 4 * no * H
    return {"unit": UNIT, "envs": n_envs, "steps": steps, "value": n_envs * steps / (ms * 1e-3), "ms_per_step": ms / steps,
            "device_mode": {"ms_per_step": ms_dev / steps, "value": n_envs * steps / (ms_dev * 1e-3), "wall_ms_per_step": wall_dev * 1e3 / steps,
                            "resets_last_warmup_step": resets_dev,
                            "graph_replay": {"ms_per_step": ms_graph / (steps // 2), "value": n_envs * (steps // 2) / (ms_graph * 1e-3),
                                             "note": "two steps per CUDA graph (the history buffers ping-pong), device-side Philox step counter"},
                            "note": "command curriculum (update + sampling), interval resampling / DR and reset_idx as mask-driven kernels "
                                    "(lt_vel_indexed_paths): the full step with no host read, launches issued eagerly from Python"},
            "launch_chain_only": {"ms_per_step": ms_chain / steps, "value": n_envs * steps / (ms_chain * 1e-3),
                                  "note": "the same step without the host-orchestrated indexed paths (command resampling, interval DR, "
                                          "reset_idx) and their three host reads: launches only, still issued eagerly from Python"},
            "wall_ms_per_step": wall * 1e3 / steps, "resets_last_warmup_step": resets0,
            "history_bytes_per_env_step": hist_bytes, "history_gb_per_s": hist_bytes * n_envs * steps / (ms * 1e-3) / 1e9,
            "path": "vel_env.VelocityStepEngine.step: lt_compute_torques x4, lt_vel_measure_heights, lt_vel_pre, lt_resample_commands, "
                    "lt_step_contact_targets, lt_corl_rewards, lt_vel_reward, reset_idx (indexed torch writes), lt_vel_observe; eager, "
                    "three nonzero() host reads per step as in the reference"}


def cfg4_ppo_iteration(dev, n_envs=4096, iters=2):
    """BASELINE.json configs[3] end to end: Runner.learn on the config-4 env in device mode (24-step rollout of the env + the policy on
    the 30 x 70 history, GAE, 5 x 4 minibatch update), wall time per iteration after one warm-up iteration."""
    from legged_tracking_b200 import ppo
    env = workload.make_cfg4(num_envs=n_envs, device=dev)
    runner = ppo.Runner(env, device=dev, save_dir=None)               # no checkpoint writes inside the timed iterations
    runner.learn(num_learning_iterations=1, init_at_random_ep_len=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    runner.learn(num_learning_iterations=iters)
    torch.cuda.synchronize()
    dt_ms = (time.perf_counter() - t0) * 1e3 / iters
    T = runner.num_steps_per_env
    return {"envs": n_envs, "iter_ms": dt_ms, "env_steps_per_s": n_envs * T / (dt_ms * 1e-3), "num_steps_per_env": T,
            "obs_history_width": int(env.num_obs_history), "finite": bool(torch.isfinite(runner.alg.actor_critic.flat).all())}


def strong_block(args, rank, world, dev, dist, peak):
    """BASELINE.json configs[4] as the north-star words it: 65,536 envs IN TOTAL, sharded over the `world` GPUs (strong scaling),
    the full step, and the PPO iteration (T = 24 rollout + GAE + 5 x 4 minibatch update) with its NCCL exchanges timed apart.
    All times are device times, max over ranks."""
    from legged_tracking_b200.engine import StepEngine
    total = 65536
    n_env = total // world
    flat, terrain, act, synth = workload.make(args.config, num_envs=n_env, seed=3000 + rank)
    eng = StepEngine(flat, terrain, act, device=dev, seed=211 + rank, envs_per_warp=args.epw)
    init_engine_state(eng, synth, flat)
    rs = ReplaySets(eng, synth, flat, N_REPLAY_SETS)
    for k in range(3):
        launch_step(eng, rs.sets[k % N_REPLAY_SETS])
    torch.cuda.synchronize()
    graphs = []
    if not args.no_graph:
        side = torch.cuda.Stream()
        for st in rs.sets:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                launch_step(eng, st, eager_gravity=False)
            graphs.append(g)

    def step_fn(k):
        i = k % N_REPLAY_SETS
        if graphs and eng.graph_replay_ok():
            graphs[i].replay()
            eng.gravity.common_step_counter += 1
        else:
            launch_step(eng, rs.sets[i])

    def torque_only(k):
        st = rs.sets[k % N_REPLAY_SETS]
        for sub, dof in enumerate(st["dof"]):
            eng.bind_sim(dof_state=dof)
            eng.compute_torques(st["actions"], sub)

    def post_only(k):
        st = rs.sets[k % N_REPLAY_SETS]
        eng.bind_sim(root_states=st["root"], contact_forces=st["contact"], dof_state=st["dof"][-1])
        eng.post_physics_step(advance_gravity=False)
    for k in range(3):
        step_fn(k)
    steps = max(20, args.steps)
    dist.barrier()
    ms = time_region(step_fn, steps) / steps
    ksub = max(20, min(args.steps, 300))
    ms_torque = time_region(torque_only, ksub) / (ksub * flat["decimation"])
    ms_post = time_region(post_only, ksub) / ksub
    t = torch.tensor([ms, ms_torque, ms_post], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_torque, ms_post = [float(x) for x in t.tolist()]
    alg = workload.algorithmic_bytes(flat)
    out = {"total_envs": total, "envs_per_gpu": n_env, "scaling": "strong", "steps": steps, "ms_per_step": ms,
           "value": total / (ms * 1e-3), "unit": UNIT,
           "kernels_us": {"torque_substep": ms_torque * 1e3, "post_physics_chain": ms_post * 1e3},
           "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s",
                        "post_physics": {"achieved": alg["post_physics"] * n_env / (ms_post * 1e-3) / 1e9,
                                         "frac": alg["post_physics"] * n_env / (ms_post * 1e-3) / 1e9 / peak},
                        "full_step": {"achieved": alg["full_step"] * n_env / (ms * 1e-3) / 1e9,
                                      "frac": alg["full_step"] * n_env / (ms * 1e-3) / 1e9 / peak}}}
    if not args.no_ppo:
        out["ppo_iteration"] = ppo_iteration(eng, rs, graphs, flat, dev, dist, args.ppo_iters)
    return out


def run_b200(args, rank, world, local_rank):
    from legged_tracking_b200.engine import StepEngine
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    host_binding = None
    if world > 1:                              # before any pinned allocation: the staging blocks then land on the GPU's NUMA node
        from legged_tracking_b200 import dist as lt_dist
        host_binding = lt_dist.bind_host_to_gpu(local_rank)
    n_env = args.envs_per_gpu
    flat, terrain, act, synth = workload.make(args.config, num_envs=n_env, seed=1000 + rank)
    eng = StepEngine(flat, terrain, act, device=dev, seed=11 + rank, envs_per_warp=args.epw)
    init_engine_state(eng, synth, flat)
    rs = ReplaySets(eng, synth, flat, N_REPLAY_SETS, pinned_host=True)
    alg = workload.algorithmic_bytes(flat)
    launches_per_step = flat["decimation"] + 3

    # ---- warm-up (eager), then one CUDA graph per replay set
    for k in range(max(args.warmup, 3)):
        launch_step(eng, rs.sets[k % N_REPLAY_SETS])
    torch.cuda.synchronize()
    graphs = []
    if not args.no_graph:
        side = torch.cuda.Stream()
        for s in rs.sets:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                launch_step(eng, s, eager_gravity=False)
            graphs.append(g)

    def step_fn(k):
        s = k % N_REPLAY_SETS
        if graphs and eng.graph_replay_ok():
            graphs[s].replay()
            eng.gravity.common_step_counter += 1
        else:
            launch_step(eng, rs.sets[s])
    for k in range(3):
        step_fn(k)

    # ---- timed region: exactly K steps, barrier + synchronize on both sides, device timed, max over ranks
    sampler = ClockSampler(local_rank)
    sampler.start()
    if dist:
        dist.barrier()
    ms = time_region(step_fn, args.steps)
    if dist:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    # per-kernel shares (same launches, each kernel class alone), for the roofline of the dominant kernel
    s0 = rs.sets[0]

    def torque_only(k):
        s = rs.sets[k % N_REPLAY_SETS]
        for sub, dof in enumerate(s["dof"]):
            eng.bind_sim(dof_state=dof)
            eng.compute_torques(s["actions"], sub)

    def post_only(k):
        s = rs.sets[k % N_REPLAY_SETS]
        eng.bind_sim(root_states=s["root"], contact_forces=s["contact"], dof_state=s["dof"][-1])
        eng.post_physics_step(advance_gravity=False)
    ksub = max(20, min(args.steps, 300))
    # best of three: these sub-timings only apportion the step between its kernel classes, and a single short region now and then
    # catches a clock / scheduling blip (SCALE_r01 N=4 reported one torque figure at twice every other N)
    ms_torque = min(time_region(torque_only, ksub) for _ in range(3)) / (ksub * flat["decimation"])
    ms_post = min(time_region(post_only, ksub) for _ in range(3)) / ksub
    sampler.stop_flag = True
    sampler.join()

    ppo_res = None
    if not args.no_ppo:
        ppo_res = ppo_iteration(eng, rs, graphs, flat, dev, dist, args.ppo_iters)

    # ---- e2e: the same step through the public call with HOST buffers (pinned): H2D inputs, kernels, D2H results
    N, A = eng.N, flat["num_actors"]
    out_h = {"obs": torch.empty(N, flat["num_obs"]).pin_memory(), "priv": torch.empty(N, flat["num_privileged_obs"]).pin_memory(),
             "rew": torch.empty(N).pin_memory(), "done": torch.empty(N, dtype=torch.bool).pin_memory()}
    own = {"actions": torch.empty(N, 12, device=dev), "dof": torch.empty(N * 12, 2, device=dev),
           "root": torch.empty(N * A, 13, device=dev), "contact": torch.empty(N, flat["num_bodies"], 3, device=dev)}
    eng.bind_sim(root_states=own["root"], dof_state=own["dof"], contact_forces=own["contact"])

    def e2e_step(k):
        h = rs.host[k % N_REPLAY_SETS]
        own["actions"].copy_(h["actions"], non_blocking=True)
        for sub, dof in enumerate(h["dof"]):
            own["dof"].copy_(dof, non_blocking=True)                  # what PhysX would hand back after sub-step `sub`
            eng.compute_torques(own["actions"], sub)
        own["root"].copy_(h["root"], non_blocking=True)
        own["contact"].copy_(h["contact"], non_blocking=True)
        eng.post_physics_step()
        out_h["obs"].copy_(eng.state["obs_buf"], non_blocking=True)
        out_h["priv"].copy_(eng.state["privileged_obs_buf"], non_blocking=True)
        out_h["rew"].copy_(eng.state["rew_buf"], non_blocking=True)
        out_h["done"].copy_(eng.state["reset_buf"], non_blocking=True)
        torch.cuda.current_stream().synchronize()                     # the caller needs obs/rew/done before acting again
    for k in range(3):
        e2e_step(k)
    ke = max(10, min(args.steps, 200))
    if dist:
        dist.barrier()
    ms_e2e_single = time_region(e2e_step, ke)

    # the product's host-buffer step: env shards pipelined over CUDA streams so that the up- and the down-link work at once
    # (engine.HostStepPipeline); same bytes per step, same kernels
    n_chunks = max(1, args.e2e_chunks)
    while n_env % n_chunks or (n_env // n_chunks) % 1024:
        n_chunks -= 1
    shard_sizes = [n_env // n_chunks] * n_chunks
    if args.e2e_shards:                       # explicit (uneven) shard sizes: a small first shard shortens the unoverlapped lead-in
        shard_sizes = [int(x) for x in args.e2e_shards.split(",")]
        if sum(shard_sizes) != n_env or any(m % 1024 for m in shard_sizes):
            raise SystemExit("--e2e-shards must be multiples of 1024 that add up to --envs-per-gpu")
        n_chunks = len(shard_sizes)
    ms_e2e = ms_e2e_single
    pipe = engines = None
    if n_chunks > 1:
        from legged_tracking_b200.engine import HostStepPipeline
        engines, h_in, h_out = [], [], []
        for c, M in enumerate(shard_sizes):
            flat_c, terrain_c, act_c, synth_c = workload.make(args.config, num_envs=M, seed=2000 + 16 * rank + c)
            terrain_c = dict(terrain_c, tile_heights=eng.tile_heights) if hasattr(eng, "tile_heights") else terrain_c
            e_c = StepEngine(flat_c, terrain_c, act_c, device=dev, seed=101 + 16 * rank + c, envs_per_warp=args.epw)
            init_engine_state(e_c, synth_c, flat_c)
            engines.append(e_c)
            rs_c = ReplaySets(e_c, synth_c, flat_c, 2, pinned_host=True)
            h_in.append(rs_c.host)
            del rs_c
        pipe = HostStepPipeline(engines)
        h_out = pipe.alloc_host_outputs()                             # obs / priv / rew / done + the reset write-back rows
        packed = [pipe.alloc_host_inputs() for _ in range(2)]          # the replayed simulator state, kept in the pipeline's pinned layout
        for r in range(2):
            for c in range(n_chunks):
                dst, src = packed[r][c], h_in[c][r]
                dst["actions"].copy_(src["actions"]); dst["root"].copy_(src["root"]); dst["contact"].copy_(src["contact"])
                for a, b in zip(dst["dof"], src["dof"]):
                    a.copy_(b)
        h_in = None

        for k in range(3):
            pipe.step(packed[k % 2], h_out)
        pipe_graphs = [] if args.no_graph else [pipe.capture(packed[r], h_out) for r in range(2)]

        def e2e_pipe(k):
            if pipe_graphs and pipe.graph_replay_ok():
                pipe.replay(pipe_graphs[k % 2])
            else:
                pipe.step(packed[k % 2], h_out)
            for ho in h_out:                                          # what a caller does next: more resets than write-back rows?
                if int(ho["reset_count"][0]) > ho["reset_ids"].shape[0]:
                    raise RuntimeError("reset write-back capacity exceeded: the bench would have to read the full state tensors")
        for k in range(3):
            e2e_pipe(k)
        if dist:
            dist.barrier()
        ms_e2e = time_region(e2e_pipe, ke)
    if dist:
        t = torch.tensor([ms_e2e, ms_e2e_single], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e, ms_e2e_single = [float(x) for x in t.tolist()]
    h2d = sum(int(x.numel()) * 4 for x in (own["actions"], own["root"], own["contact"])) + flat["decimation"] * own["dof"].numel() * 4
    d2h = sum(int(v.numel()) * v.element_size() for v in out_h.values())
    d2h_writeback = 0
    if n_chunks > 1:                                                  # the pipelined path also brings the reset rows down
        d2h_writeback = sum(int(v.numel()) * v.element_size() for ho in h_out for k, v in ho.items() if k.startswith("reset_"))

    peak, peak_src = measured_peaks()
    strong = None
    if world > 1 and 65536 % world == 0 and (65536 // world) % 1024 == 0:
        del pipe, engines
        torch.cuda.empty_cache()
        strong = strong_block(args, rank, world, dev, dist, peak)
    if rank != 0:
        return
    ms_step = ms / args.steps
    value = world * n_env * args.steps / (ms * 1e-3)
    torque_total = ms_torque * flat["decimation"]
    dominant = "lt_torque_kernel" if torque_total > ms_post else "lt_step_kernel"          # class of the dominant kernel
    tr = ncu_traffic() if n_env == 65536 else {}          # the capture was taken at 65,536 envs
    tr_post = [tr.get(k) for k in ("lt_step_kernel", "lt_obs_kernel", "lt_compact_kernel")]
    k_post = {"kernel": "post-physics chain: lt_step_kernel + lt_obs_kernel + lt_compact_kernel", "ms_per_launch": ms_post,
              "achieved": alg["post_physics"] * n_env / (ms_post * 1e-3) / 1e9,
              "traffic": None if None in tr_post else sum(tr_post)}
    # lt_compute_torques picks the tcgen05 kernel from LT_TC5_MIN_ENVS (csrc/lt_kernels.cu) envs per GPU, else the mma.sync kernel
    forced = os.environ.get("LT_TORQUE_TC5")
    tq_name = "lt_torque_tc5_kernel" if (forced == "1" or (forced is None and n_env >= 40960)) else "lt_torque_mma_kernel"
    k_tq = {"kernel": tq_name, "ms_per_launch": ms_torque,
            "achieved": alg["torque_substep"] * n_env / (ms_torque * 1e-3) / 1e9,
            "fp32_equiv_tflops": alg["torque_flops_substep"] * n_env / (ms_torque * 1e-3) / 1e12,
            "traffic": tr.get(tq_name),
            "note": "actuator net on the tensor cores (fp16-split x3, fp32 accumulate; tcgen05.mma with A and D in TMEM for large "
                    "shards, mma.sync m16n8k16 for small ones); bounded by XU (64 MUFU.RCP per sample) + issue, not by HBM: see "
                    "profiles/ and DESIGN.md section 4"}
    for d in (k_post, k_tq):
        d.update({"bound": "hbm", "peak": peak, "unit": "GB/s", "frac": d["achieved"] / peak})
    dom = k_tq if dominant == "lt_torque_kernel" else k_post
    roofline = {"bound": "hbm", "achieved": dom["achieved"], "peak": peak, "unit": "GB/s", "frac": dom["frac"],
                "traffic": dom.get("traffic"), "traffic_source": "profiles/ncu_traffic.json (ncu --set full, 65,536 envs, per launch)",
                "kernel": dom["kernel"], "peak_source": peak_src, "ms_per_launch": dom["ms_per_launch"],
                "algorithmic_bytes_per_env": alg["torque_substep"] if dom is k_tq else alg["post_physics"],
                "kernels": {"post_physics": k_post, "torque_substep": k_tq},
                "share_of_step": {"torque_x%d" % flat["decimation"]: torque_total / (torque_total + ms_post),
                                  "post_physics": ms_post / (torque_total + ms_post)}}
    if world == 1:
        try:
            roofline["kernels"].update(rollout_kernel_timing(flat, dev, n_env, peak))
        except Exception as e:
            roofline["kernels"]["rollout_store"] = {"error": str(e)[:300]}
    cpu = None
    if world == 1 and not args.no_cpu:
        from oracle import ref_arm
        if ref_arm.available():
            rate, s_per_step, cores = ref_arm.step_rate(args.cpu_steps, 2, args.cpu_envs)
            kind, how = "reference", "the unmodified reference (oracle/_ref) stepping HistoryWrapper(TrajectoryTrackingEnv) under the isaacgym shims"
        else:
            rate, s_per_step, cores = cpu_oracle_rate(args.cpu_steps, args.cpu_envs, "cfg2")
            kind, how = "port", "oracle/port.py (torch CPU, reference's per-env height gather)"
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"{args.cpu_envs} envs x {args.cpu_steps} steps of the same workload shape, {how}, {s_per_step * 1e3:.1f} ms/step"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": workload_config(args, world, n_env, alg['full_step'] * n_env / 1e6),
            "clocks": sampler.summary(), "gpu_launches": launches_per_step * args.steps,
            "e2e": {"value": world * n_env * ke / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h + d2h_writeback,
                    "ms_per_step": ms_e2e / ke, "steps": ke,
                    "reset_writeback_bytes_per_step": d2h_writeback,      # fixed-capacity rows of the envs the step reset (root + dof state)
                    "link_gb_per_s_per_rank": (h2d + d2h + d2h_writeback) * ke / (ms_e2e * 1e-3) / 1e9,
                    "path": f"engine.HostStepPipeline: {n_chunks} env shards on {n_chunks} streams (H2D, kernels and D2H of different shards overlap)"
                            + ("" if args.no_graph else ", the whole pipelined step replayed as one CUDA graph")
                            if n_chunks > 1 else "single engine, one stream",
                    "single_stream": {"value": world * n_env * ke / (ms_e2e_single * 1e-3), "ms_per_step": ms_e2e_single / ke},
                    "host_binding_rank0": host_binding},
            "post_physics_only": {"value": world * n_env / (ms_post * 1e-3), "unit": UNIT, "ms_per_step": ms_post},
            "roofline": roofline}
    if world == 1:
        try:
            line["api_step"] = api_step_timing(args, dev, n_env, max(20, min(args.steps, 200)))
        except Exception as e:                    # auxiliary number
            line["api_step"] = {"error": str(e)[:300]}
    if world == 1:
        try:
            line["cfg4_step"] = {"n4096": cfg4_step_timing(dev, 4096), "n65536": cfg4_step_timing(dev, 65536, steps=30),
                                 "ppo_iteration_n4096": cfg4_ppo_iteration(dev, 4096)}
        except Exception as e:                    # auxiliary number
            line["cfg4_step"] = {"error": str(e)[:300]}
    try:
        line["terrain_build"] = terrain_build_timing(dev)
    except Exception as e:                        # auxiliary number: never costs the headline line
        line["terrain_build"] = {"error": str(e)[:200]}
    if ppo_res:
        line["ppo_iteration"] = ppo_res
    if strong:
        line["strong_65536_total"] = strong
    if cpu:
        line["cpu_baseline"] = cpu
    _emit(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="cfg5")
    ap.add_argument("--envs-per-gpu", type=int, default=65536)
    ap.add_argument("--epw", type=int, default=0, help="envs per warp of the step kernel (0 = auto)")
    ap.add_argument("--e2e-chunks", type=int, default=4, help="env shards of the pipelined host-buffer step (1 = single stream)")
    ap.add_argument("--e2e-shards", default="", help="explicit env-shard sizes of the pipelined host-buffer step, e.g. 4096,12288,16384,16384,16384")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ppo", action="store_true", help="skip the PPO-iteration timing (rollout + GAE + update)")
    ap.add_argument("--ppo-iters", type=int, default=2)
    ap.add_argument("--cpu-envs", type=int, default=4096)
    ap.add_argument("--cpu-steps", type=int, default=40)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    _quiet_stdout()
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU baseline")
    run_b200(args, rank, world, local_rank)
    if world > 1 and torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
