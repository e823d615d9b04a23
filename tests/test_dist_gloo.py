"""world_size-2 `gloo` tests (CPU) of the env-sharded N>1 path: shard bounds, advantage-moment merge, the [stats | gradient]
exchange buffer of the PPO update, the parameter broadcast and the reward-sign exchange.  The arithmetic on either side of
these collectives is CUDA-only; tests/test_gpu_policy.py checks on one GPU that two half-minibatch gradients produced with
grad_world = 2 sum to the full-minibatch gradient."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from legged_tracking_b200 import dist as ltd


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(fn, world, *args):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_entry, args=(fn, r, world, port, q) + args) for r in range(world)]
    for p in procs:
        p.start()
    out = {}
    try:
        for _ in range(world):                # drain before join: a child blocks in put() until its payload is read
            r, v = q.get(timeout=240)
            out[r] = v
    finally:
        for p in procs:
            p.join(30)
            if p.is_alive():
                p.kill()
    for p in procs:
        assert p.exitcode == 0, f"rank exited with {p.exitcode}"
    for r, v in out.items():
        if isinstance(v, str) and v.startswith("ERROR"):
            raise AssertionError(f"rank {r}: {v}")
    return out


def _entry(fn, r, world, port, q, *args):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r))
    torch.set_num_threads(2)
    try:
        ltd.init_from_env("gloo")
        q.put((r, fn(r, world, *args)))
    except Exception:
        import traceback
        q.put((r, "ERROR " + traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def test_shard_bounds_cover_everything():
    for total in (1, 7, 64, 65536, 100001):
        for world in (1, 2, 3, 8):
            spans = [ltd.shard_bounds(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans[:-1], spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _moments(r, world):
    g = torch.Generator().manual_seed(3)
    adv = torch.randn(8, 64, generator=g, dtype=torch.float64)
    lo, hi = ltd.shard_bounds(64, world, r)
    mine = adv[:, lo:hi]
    stats = torch.tensor([mine.sum(), (mine * mine).sum(), float(mine.numel())], dtype=torch.float64)
    ltd.merge_moments(stats)
    n = stats[2]
    mean = stats[0] / n
    std = torch.sqrt((stats[1] - n * mean * mean) / (n - 1))
    return float(mean - adv.mean()), float(std - adv.std())


def test_advantage_moments_merge_to_global_statistics():
    out = _spawn(_moments, 2)
    for r in (0, 1):
        assert abs(out[r][0]) < 1e-14 and abs(out[r][1]) < 1e-13


def _exchange(r, world):
    """dist.GradExchange: the [stats | gradient] buffer of the PPO update.  Rank r holds (r + 1) everywhere."""
    n, hi = 40, 17
    ex = ltd.GradExchange(n, "cpu", stats=16)
    assert ex.stats.data_ptr() == ex.buf.data_ptr() and ex.grads.data_ptr() == ex.buf.data_ptr() + 64
    ex.buf.fill_(float(r + 1))
    ex.exchange_main()
    total = float(sum(range(1, world + 1)))
    ok_main = bool((ex.buf == total).all())
    # adaptation step: stats[4..16) and gradient[0..hi) are summed again; stats[0..4) (already global) and the rest are not
    ex.buf.fill_(float(r + 1))
    ex.stats[:4] = 100.0
    ex.exchange_adaptation(hi)
    ok_ad = (bool((ex.stats[:4] == 100.0).all()) and bool((ex.stats[4:] == total).all()) and bool((ex.grads[:hi] == total).all())
             and bool((ex.grads[hi:] == float(r + 1)).all()))
    # replicas start from rank 0's parameters whatever each rank drew
    p = torch.full((5,), float(r + 7))
    ltd.broadcast_(p, src=0)
    return ok_main, ok_ad, p.tolist(), ex.calls, ex.bytes


def test_grad_exchange_sums_the_right_slices_and_broadcast_aligns_replicas():
    out = _spawn(_exchange, 2)
    for r in (0, 1):
        ok_main, ok_ad, p, calls, nbytes = out[r]
        assert ok_main and ok_ad
        assert p == [7.0] * 5
        assert calls == 2 and nbytes == 4 * (56 + 12 + 17)
    single = _spawn(_exchange, 1)[0]
    assert single[0] and single[3] == 0            # world 1: no collective is issued


def _sign_exchange(r, world):
    """StepEngine's all-env reward-sign exchange (legged_robot_trajectory_tracking.py:332-335) on the host side: shard sums
    of opposite sign must produce the whole job's sign on every rank."""
    sums = torch.tensor([3.0, -1.0, 0.5] if r == 0 else [-5.0, 4.0, -0.5])
    ltd.allreduce_sum_(sums)
    return (sums >= 0).tolist()


def test_reward_sign_uses_the_global_sum():
    out = _spawn(_sign_exchange, 2)
    assert out[0] == out[1] == [False, True, True]
