"""Scratch driver for profiling rows a14 / a15 alone: one PPO.act at N envs and `--minibatches` minibatch steps at the cfg-5
minibatch size on random rollout contents.  Used under ncu (launch list / --set full); prints CUDA-event times otherwise."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from legged_tracking_b200 import ppo  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=65536)
ap.add_argument("--T", type=int, default=24)
ap.add_argument("--epochs", type=int, default=1)
ap.add_argument("--acts", type=int, default=2)
ap.add_argument("--reps", type=int, default=1)
a = ap.parse_args()
dev = torch.device("cuda", 0)
torch.manual_seed(0)
ppo.PPO_Args.num_learning_epochs = a.epochs
ac = ppo.ActorCritic(261, 2, 261, 12).to(dev)
alg = ppo.PPO(ac, device=dev)
alg.init_storage(a.envs, a.T, [261], [2], [261], [12])
st = alg.storage
g = torch.Generator(device=dev).manual_seed(1)
for name in ("observation_histories", "privileged_observations", "actions", "values", "returns", "advantages", "actions_log_prob", "mu"):
    getattr(st, name).normal_(generator=g)
st.sigma.fill_(1.0)
st.actions_log_prob.mul_(0.3).sub_(12 * 1.4)
obs = torch.randn(a.envs, 261, device=dev, generator=g)
priv = torch.randn(a.envs, 2, device=dev, generator=g)


def timed(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def acts():
    for _ in range(a.acts):
        st.step = 0
        alg.act(obs, priv, obs)


def update():
    st.step = a.T
    alg.compute_returns(obs, priv)          # lt_gae_kernel + lt_normalize_kernel in the launch list too
    alg.update()


acts()
update()
for _ in range(a.reps):
    t_act = timed(acts) / a.acts
    t_upd = timed(update)
    print(f"act {t_act:.3f} ms per call ({a.envs} rows); update {t_upd:.2f} ms for {a.epochs * ppo.PPO_Args.num_mini_batches} minibatches "
          f"of {a.envs * a.T // ppo.PPO_Args.num_mini_batches} rows = {t_upd / (a.epochs * ppo.PPO_Args.num_mini_batches):.2f} ms each")
