"""Multi-GPU plumbing for the env-sharded path: one process per GPU, environments partitioned across ranks with no
data-path collective in the simulation step.  Collectives appear only where the reference's single-process math
couples all environments:

  * the sign of reward terms summed over all envs (legged_robot_trajectory_tracking.py:341-347) -> all-reduce of the
    LT_MAX_TERMS float sums between lt_reward_sign_prepass and lt_post_physics_step;
  * advantage normalisation over the whole [T,N] rollout (rollout_storage.py:89-90) -> all-reduce of the
    {sum, sum of squares, count} doubles lt_gae leaves on the device;
  * the PPO gradient step (ppo.py:144-148) -> mean all-reduce of one flat gradient bucket;
  * the adaptive-KL learning-rate decision (ppo.py:111-125) -> mean all-reduce of the scalar KL so that every rank
    takes the same branch.

Everything here works on whatever device the tensors live on (NCCL for CUDA, gloo for the CPU tests)."""
import os

import torch
import torch.distributed as dist


def initialized():
    return dist.is_available() and dist.is_initialized()


def world_size():
    return dist.get_world_size() if initialized() else 1


def rank():
    return dist.get_rank() if initialized() else 0


def init_from_env(backend=None):
    """torchrun-style rendezvous (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT / LOCAL_RANK).  -> (rank, world, local_rank)"""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank(), world_size(), local


def shard_bounds(total, world=None, r=None):
    """Contiguous env shard [lo, hi) of rank r: the first (total % world) ranks get one extra env."""
    world = world_size() if world is None else world
    r = rank() if r is None else r
    base, extra = divmod(int(total), world)
    lo = r * base + min(r, extra)
    return lo, lo + base + (1 if r < extra else 0)


def allreduce_sum_(t):
    if world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def allreduce_mean_scalar(t):
    """Mean over ranks of a 0-d/1-element tensor (equal shard sizes assumed, as in the env-sharded bench)."""
    if world_size() > 1:
        t = t.clone()
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        t /= world_size()
    return t


class GradBucket:
    """One flat fp32 buffer aliased by every parameter's .grad, so the gradient exchange is ONE all-reduce per
    minibatch instead of one per tensor (26 tensors in the default ActorCritic)."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        p0 = self.params[0]
        self.flat = torch.zeros(n, dtype=p0.dtype, device=p0.device)
        self.views = []
        o = 0
        for p in self.params:
            self.views.append(self.flat[o:o + p.numel()].view_as(p))
            o += p.numel()

    def attach(self):
        """Point every .grad at its slice (zeroed).  Call instead of optimizer.zero_grad()."""
        self.flat.zero_()
        for p, v in zip(self.params, self.views):
            p.grad = v

    def allreduce_mean(self):
        if world_size() > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
            self.flat /= world_size()


def merge_moments(stats):
    """stats: double[3] = {sum A, sum A^2, count} of the local shard (written by lt_gae) -> global, in place."""
    return allreduce_sum_(stats)
