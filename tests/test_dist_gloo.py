"""world_size-2 `gloo` tests (CPU) of the env-sharded N>1 path: shard bounds, advantage-moment merge, the flat gradient
bucket exchange, and a full PPO.update on two ranks that must reproduce the single-process update."""
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


def _bucket(r, world):
    torch.manual_seed(0)
    lin = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Tanh(), torch.nn.Linear(7, 2))
    x = torch.randn(16, 5)
    ref = torch.autograd.grad(lin(x).pow(2).mean(), list(lin.parameters()))
    b = ltd.GradBucket(lin.parameters())
    b.attach()
    lin(x[r::world]).pow(2).mean().backward()
    b.allreduce_mean()
    return max(float((p.grad - g).abs().max()) for p, g in zip(lin.parameters(), ref))


def test_grad_bucket_mean_allreduce_equals_full_batch_gradient():
    out = _spawn(_bucket, 2)
    assert max(out.values()) < 1e-7


def _ppo_update(r, world, name):
    from legged_tracking_b200 import ppo
    from tests.helpers_ppo import HostStorage, actor_critic_from_golden, golden_args, load_case
    g = load_case(name)
    with golden_args(g):
        # minibatch 160 -> adaptation train split 128; per rank 80 -> 64: every rank's split is its half of the global one
        ppo.PPO_Args.num_mini_batches = 2
        fields = {k: torch.from_numpy(g["storage/" + k][:5]) for k in
                  ("observations", "privileged_observations", "observation_histories", "actions", "values", "returns",
                   "actions_log_prob", "advantages", "mu", "sigma")}
        perm = torch.from_numpy(np.random.default_rng(1).permutation(320))
        ac = actor_critic_from_golden(g)
        alg = ppo.PPO(ac, device="cpu")
        alg.storage = HostStorage(fields, part=(r, world))
        losses = alg.update(indices=perm)
        return dict(losses=losses, lrs=alg.last_update_trace["lrs"], sd={k: v.numpy().copy() for k, v in ac.state_dict().items()})


def test_two_rank_ppo_update_equals_single_process():
    name = "ppo_adaptive_64x8"
    single = _spawn(_ppo_update, 1, name)[0]
    two = _spawn(_ppo_update, 2, name)
    for r in (0, 1):
        assert two[r]["lrs"] == single["lrs"]
        for k, v in single["sd"].items():
            np.testing.assert_allclose(two[r]["sd"][k], v, rtol=2e-4, atol=2e-5, err_msg=k)
    for k, v in two[0]["sd"].items():
        assert np.array_equal(v, two[1]["sd"][k]), k            # replicas stay bit-identical
    # loss means: surrogate/value are means of half-batch means
    np.testing.assert_allclose(0.5 * (np.asarray(two[0]["losses"]) + np.asarray(two[1]["losses"]))[:3],
                               np.asarray(single["losses"])[:3], rtol=2e-4, atol=1e-6)
