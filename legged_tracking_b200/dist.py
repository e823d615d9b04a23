"""Multi-GPU plumbing for the env-sharded path: one process per GPU, environments partitioned across ranks with no
data-path collective in the simulation step.  Collectives appear only where the reference's single-process math
couples all environments:

  * the sign of reward terms summed over all envs (legged_robot_trajectory_tracking.py:341-347) -> all-reduce of the
    LT_MAX_TERMS float sums between lt_reward_sign_prepass and lt_post_physics_step;
  * advantage normalisation over the whole [T,N] rollout (rollout_storage.py:89-90) -> all-reduce of the
    {sum, sum of squares, count} doubles lt_gae leaves on the device;
  * the PPO gradient step (ppo.py:144-148) and the adaptive-KL learning-rate decision (ppo.py:111-125) -> ONE sum all-reduce
    of the [loss / KL sums | flat gradient] buffer per minibatch (GradExchange), so every rank takes the same branch;
  * the initial parameters -> broadcast from rank 0 (replicas must not depend on per-rank seeds).

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


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_host_to_gpu(device_index, sysfs="/sys"):
    """Pins the calling process to the CPUs of the NUMA node the GPU's PCIe function hangs off, so that the pinned staging blocks a
    rank allocates afterwards (first touch) and the thread that issues its copies sit on the GPU's side of the host: with one
    process per GPU the ranks then stop sharing one memory controller for their H2D / D2H traffic.  Nothing is changed when the
    host exposes no NUMA placement for the device (numa_node < 0: single-node hosts, most VMs) or when the node's CPU list does
    not intersect the process's current affinity mask.  -> {"pci", "numa_node", "cpus", "bound"}"""
    info = {"pci": None, "numa_node": None, "cpus": None, "bound": False}
    try:
        p = torch.cuda.get_device_properties(device_index)
        pci = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        info["pci"] = pci
        with open(os.path.join(sysfs, "bus/pci/devices", pci, "numa_node")) as f:
            node = int(f.read().strip())
        info["numa_node"] = node
        if node < 0:
            return info
        with open(os.path.join(sysfs, "devices/system/node", f"node{node}", "cpulist")) as f:
            cpus = _parse_cpulist(f.read()) & set(os.sched_getaffinity(0))
        info["cpus"] = len(cpus)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["bound"] = True
    except (OSError, ValueError, AttributeError) as e:
        info["error"] = f"{type(e).__name__}: {e}"
    return info


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


def broadcast_(t, src=0):
    if world_size() > 1:
        dist.broadcast(t, src=src)
    return t


class GradExchange:
    """The one buffer the PPO update exchanges between env-sharded ranks: [stats (LT_PPO_STATS floats) | flat gradient].
    lt_ppo_minibatch_grad / lt_adaptation_grad (include/lt_policy.h) write per-rank SUMS scaled by 1 / (count * world) into it,
    so a single SUM all-reduce per call yields the whole job's mean gradient, loss sums, KL sum and row count -- every rank then
    takes the same learning-rate branch and applies the same Adam step (replicas stay bit-identical).

    main step        all-reduce of the whole buffer (stats[0..4) + every gradient)
    adaptation step  all-reduce of stats[4..16) + gradient[0 .. adaptation_hi): the slice is contiguous with the tail of the
                     stats because std and the adaptation module come first in the flat layout; stats[0..4) of the main step
                     are NOT touched again (they are already global sums)
    `timed = True` brackets every exchange with CUDA events (bench.py reads seconds() after a synchronize)."""

    MAIN_STATS = 4

    def __init__(self, n_params, device, stats=16):
        self.n_stats = int(stats)
        self.buf = torch.zeros(self.n_stats + int(n_params), dtype=torch.float32, device=device)
        self.stats = self.buf[:self.n_stats]
        self.grads = self.buf[self.n_stats:]
        self.timed = False
        self.calls, self.bytes = 0, 0
        self._events = []

    def _allreduce(self, view):
        if world_size() <= 1:
            return
        self.calls += 1
        self.bytes += view.numel() * view.element_size()
        if self.timed and view.is_cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            dist.all_reduce(view, op=dist.ReduceOp.SUM)
            e1.record()
            self._events.append((e0, e1))
        else:
            dist.all_reduce(view, op=dist.ReduceOp.SUM)

    def exchange_main(self):
        self._allreduce(self.buf)

    def exchange_adaptation(self, adaptation_hi):
        self._allreduce(self.buf[self.MAIN_STATS:self.n_stats + int(adaptation_hi)])

    def seconds(self):
        """Device time spent inside the all-reduces since the last call (needs timed = True and a prior synchronize)."""
        t = sum(a.elapsed_time(b) for a, b in self._events) * 1e-3
        self._events = []
        return t


def merge_moments(stats):
    """stats: double[3] = {sum A, sum A^2, count} of the local shard (written by lt_gae) -> global, in place."""
    return allreduce_sum_(stats)
