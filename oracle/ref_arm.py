"""The reference arm of bench.py: the UNMODIFIED reference's TrajectoryTrackingEnv.step on the host cores, from the staged
copy under oracle/_ref/ (oracle/make_ref.sh), driven through oracle/ref_harness.py (shims for isaacgym / params_proto / gym,
PhysX replaced by replayed synthetic state).  TEST INFRASTRUCTURE: only bench.py --impl reference / cpu_baseline may call it.
"""
import os
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")

# BASELINE.json configs[1] / [4]: scripts/train.py flags of the single_path e2e run
CFG5_ARGV = ["--old_ppo", "--measure_front_half", "--camera_zero", "--strategy", "e2e", "--terrain", "single_path",
             "--penalty_scaler", "1.0"]


def available():
    return os.path.isdir(os.path.join(REF_DIR, "go1_gym"))


def step_rate(steps, warmup, num_envs, threads=None):
    """-> (env-steps/s, seconds per step, threads).  HistoryWrapper(TrajectoryTrackingEnv).step(actions), `warmup` untimed
    steps then `steps` timed ones, on `num_envs` envs of the cfg 2 / 5 configuration."""
    os.environ["LT_REFERENCE_ROOT"] = REF_DIR
    import torch
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    from oracle import ref_harness as H
    wrapped, env, replay, Cfg, args = H.make_env(CFG5_ARGV, num_envs)
    synth = replay.synth
    with torch.no_grad():
        wrapped.reset()
        env.episode_length_buf[:] = synth.initial_episode_length(int(env.max_episode_length))
        k = 0
        for _ in range(max(warmup, 0)):
            wrapped.step(synth.actions(k))
            k += 1
        t0 = time.perf_counter()
        for _ in range(steps):
            wrapped.step(synth.actions(k))
            k += 1
        dt = time.perf_counter() - t0
    return env.num_envs * steps / dt, dt / steps, torch.get_num_threads()
