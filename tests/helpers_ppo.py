"""Test scaffolding for the PPO rows: loads tests/golden/ppo_*.npz and provides a host-memory stand-in for the
rollout buffers so that PPO.update's host logic (flat parameter bucket, two Adams, KL schedule, cross-rank exchange)
can be exercised without a GPU.  Not product code: legged_tracking_b200.ppo.RolloutStorage itself is CUDA-only."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PPO_CASES = ["ppo_adaptive_64x8", "ppo_fixed_32x6", "ppo_cnn_48x6"]      # the last one: go1_gym_learn/ppo_cse_cnn (row f4)


def variant_module(g):
    """legged_tracking_b200.ppo or .ppo_cnn, by the reference package the fixture was recorded from"""
    from legged_tracking_b200 import ppo, ppo_cnn
    return ppo_cnn if "meta/variant" in g and str(g["meta/variant"]) == "ppo_cse_cnn" else ppo


def load_case(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


def _val(v):
    return v.item() if v.ndim == 0 else v.tolist()


class golden_args:
    """Context manager: set legged_tracking_b200.ppo.AC_Args / PPO_Args to what the reference held for the fixture."""

    def __init__(self, g):
        self.g = g

    def __enter__(self):
        mod = variant_module(self.g)
        self.saved = [(cls, k, getattr(cls, k)) for cls in (mod.AC_Args, mod.PPO_Args) for k in vars(cls)]
        for k, v in self.g.items():
            if k.startswith("ac_args/"):
                setattr(mod.AC_Args, k[8:], _val(v))
            elif k.startswith("ppo_args/"):
                setattr(mod.PPO_Args, k[9:], _val(v))
        return self

    def __exit__(self, *a):
        for cls, k, v in self.saved:
            setattr(cls, k, v)


def actor_critic_from_golden(g, device="cpu"):
    N, T, no, npv, H, na = (int(x) for x in g["meta/sizes"])
    ac = variant_module(g).ActorCritic(no, npv, no * H, na)
    sd = {k[5:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("init/")}
    ac.load_state_dict(sd, strict=True)
    return ac.to(device)


class HostStorage:
    """Same mini_batch_generator contract as RolloutStorage, over given [T,N,*] tensors on any device; `part=(r, w)`
    makes this rank see every w-th sample of each minibatch (for the world_size-2 equivalence test)."""

    def __init__(self, fields, part=(0, 1)):
        self.f = {k: v.flatten(0, 1) for k, v in fields.items()}
        self.part = part
        self.cleared = False

    @classmethod
    def from_golden(cls, g, device="cpu", part=(0, 1)):
        names = ("observations", "privileged_observations", "observation_histories", "actions", "values", "returns",
                 "actions_log_prob", "advantages", "mu", "sigma")
        return cls({k: torch.from_numpy(g["storage/" + k]).to(device) for k in names}, part)

    def mini_batch_generator(self, num_mini_batches, num_epochs=8, indices=None):
        f = self.f
        B = f["values"].shape[0]
        mb = B // num_mini_batches
        if indices is None:
            indices = torch.randperm(num_mini_batches * mb)
        r, w = self.part
        for _ in range(num_epochs):
            for i in range(num_mini_batches):
                b = indices[i * mb:(i + 1) * mb][r::w]
                yield (f["observations"][b], f["observations"][b], f["privileged_observations"][b], f["observation_histories"][b],
                       f["actions"][b], f["values"][b], f["advantages"][b], f["returns"][b], f["actions_log_prob"][b],
                       f["mu"][b], f["sigma"][b], None, None)

    def clear(self):
        self.cleared = True
