"""PPO rows of the path behind the reference's own class names and call signatures.

  AC_Args / ActorCritic      go1_gym_learn/ppo_cse/actor_critic.py:9-160   (same state_dict keys: a checkpoint of either
                                                                           loads into the other)
  RolloutStorage             go1_gym_learn/ppo_cse/rollout_storage.py       ([T,N,*] buffers in HBM; add_transitions,
                                                                           compute_returns run as liblt_b200 kernels)
  PPO_Args / PPO             go1_gym_learn/ppo_cse/ppo.py                   (act / process_env_step / compute_returns / update)
  RunnerArgs / Runner        go1_gym_learn/ppo_cse/__init__.py:46-357       (learn loop; logging back-ends left to the caller)

Storage and GAE have no CPU fallback: RolloutStorage raises on a non-CUDA device.  The MLP forward/backward GEMMs go
through torch autograd (cuBLAS) -- plumbing the reference shares; what is new here is the storage/GAE kernels, the
single flat parameter/gradient bucket (one Adam launch, one clip, one all-reduce per minibatch) and env sharding.
"""
import copy
import ctypes as C
import os
import time
from collections import deque

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.distributions import Normal

from . import _abi, dist as ltd
from .params import PrefixProto


class AC_Args(PrefixProto, cli=False):          # actor_critic.py:9-21
    init_noise_std = 1.0
    actor_hidden_dims = [512, 256, 128]
    critic_hidden_dims = [512, 256, 128]
    activation = 'elu'
    adaptation_module_branch_hidden_dims = [256, 128]
    use_decoder = False
    normalize_obs = False


class PPO_Args(PrefixProto, cli=False):         # ppo.py:13-30
    value_loss_coef = 1.0
    use_clipped_value_loss = True
    clip_param = 0.2
    entropy_coef = 0.01
    num_learning_epochs = 5
    num_mini_batches = 4
    learning_rate = 1.e-3
    adaptation_module_learning_rate = 1.e-3
    num_adaptation_module_substeps = 1
    schedule = 'adaptive'
    gamma = 0.99
    lam = 0.95
    desired_kl = 0.01
    max_grad_norm = 1.
    selective_adaptation_module_loss = False


class RunnerArgs(PrefixProto, cli=False):       # ppo_cse/__init__.py:46-62
    algorithm_class_name = 'RMA'
    num_steps_per_env = 24
    max_iterations = 1500
    save_interval = 400
    save_video_interval = 100
    log_freq = 10
    resume = False
    load_run = -1
    checkpoint = -1
    resume_path = None
    resume_curriculum = True


def get_activation(act_name):                   # actor_critic.py:162-178
    table = {"elu": nn.ELU, "selu": nn.SELU, "relu": nn.ReLU, "crelu": nn.ReLU, "lrelu": nn.LeakyReLU,
             "tanh": nn.Tanh, "sigmoid": nn.Sigmoid}
    if act_name not in table:
        raise ValueError(f"invalid activation function {act_name!r}")
    return table[act_name]()


def _mlp(n_in, hidden, n_out, activation):
    """Linear/activation pairs with a bare last Linear; module indices 0,2,4,... as in actor_critic.py:43-82."""
    dims = [n_in] + list(hidden)
    layers = []
    for a, b in zip(dims[:-1], dims[1:]):
        layers += [nn.Linear(a, b), activation]
    layers.append(nn.Linear(dims[-1], n_out))
    return nn.Sequential(*layers)


class ActorCritic(nn.Module):
    is_recurrent = False

    def __init__(self, num_obs, num_privileged_obs, num_obs_history, num_actions, **kwargs):
        if kwargs:
            print("ActorCritic.__init__ got unexpected arguments, which will be ignored: " + str(list(kwargs)))
        if AC_Args.use_decoder or AC_Args.normalize_obs:
            raise NotImplementedError("use_decoder / normalize_obs are off in every reference config (actor_critic.py:18-20)")
        self.decoder = AC_Args.use_decoder
        super().__init__()
        self.num_obs_history = num_obs_history
        self.num_privileged_obs = num_privileged_obs
        act = get_activation(AC_Args.activation)
        self.adaptation_module = _mlp(num_obs_history, AC_Args.adaptation_module_branch_hidden_dims, num_privileged_obs, act)
        self.actor_body = _mlp(num_privileged_obs + num_obs_history, AC_Args.actor_hidden_dims, num_actions, act)
        self.critic_body = _mlp(num_privileged_obs + num_obs_history, AC_Args.critic_hidden_dims, 1, act)
        self.std = nn.Parameter(AC_Args.init_noise_std * torch.ones(num_actions))
        self.distribution = None
        Normal.set_default_validate_args = False
        self.normalize_obs = False

    def reset(self, dones=None):
        pass

    def forward(self):
        raise NotImplementedError

    @property
    def action_mean(self):
        return self.distribution.mean

    @property
    def action_std(self):
        return self.distribution.stddev

    @property
    def entropy(self):
        return self.distribution.entropy().sum(dim=-1)

    def update_distribution(self, observation_history):
        latent = self.adaptation_module(observation_history)
        mean = self.actor_body(torch.cat((observation_history, latent), dim=-1))
        self.distribution = Normal(mean, mean * 0. + self.std)

    def act(self, observation_history, **kwargs):
        self.update_distribution(observation_history)
        return self.distribution.sample()

    def get_actions_log_prob(self, actions):
        return self.distribution.log_prob(actions).sum(dim=-1)

    def act_expert(self, ob, policy_info={}):
        return self.act_teacher(ob["obs_history"], ob["privileged_obs"])

    def act_inference(self, ob, policy_info={}):
        return self.act_student(ob["obs_history"], policy_info=policy_info)

    def act_student(self, observation_history, policy_info={}):
        latent = self.adaptation_module(observation_history)
        actions_mean = self.actor_body(torch.cat((observation_history, latent), dim=-1))
        policy_info["latents"] = latent.detach().cpu().numpy()
        return actions_mean

    def act_teacher(self, observation_history, privileged_info, policy_info={}):
        actions_mean = self.actor_body(torch.cat((observation_history, privileged_info), dim=-1))
        policy_info["latents"] = privileged_info
        return actions_mean

    def evaluate(self, observation_history, privileged_observations, **kwargs):
        return self.critic_body(torch.cat((observation_history, privileged_observations), dim=-1))

    def get_student_latent(self, observation_history):
        return self.adaptation_module(observation_history)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class RolloutStorage:
    """[T, N, *] rollout buffers resident in HBM (rollout_storage.py:26-52).  add_transitions is ONE launch of
    lt_rollout_store (the reference issues 11 copy_ kernels + 3 for the time-out bootstrap); compute_returns is
    lt_gae (thread per env, T steps backwards) + an optional cross-rank moment all-reduce + lt_normalize_advantages."""

    class Transition:
        def __init__(self):
            self.observations = None
            self.privileged_observations = None
            self.observation_histories = None
            self.critic_observations = None
            self.actions = None
            self.rewards = None
            self.dones = None
            self.values = None
            self.actions_log_prob = None
            self.action_mean = None
            self.action_sigma = None
            self.env_bins = None
            self.time_outs = None          # extras["time_outs"]; bootstrap fused into the store kernel (ppo.py:85-87)

        def clear(self):
            self.__init__()

    def __init__(self, num_envs, num_transitions_per_env, obs_shape, privileged_obs_shape, obs_history_shape, actions_shape,
                 device='cuda:0'):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _abi.LtError("RolloutStorage lives in GPU memory and is written by liblt_b200 kernels: there is no CPU path")
        self.lib = _abi.load()
        self.obs_shape, self.privileged_obs_shape = obs_shape, privileged_obs_shape
        self.obs_history_shape, self.actions_shape = obs_history_shape, actions_shape
        T, N = int(num_transitions_per_env), int(num_envs)
        z = lambda *s, **k: torch.zeros(T, N, *s, device=self.device, **k)
        self.observations = z(*obs_shape)
        self.privileged_observations = z(*privileged_obs_shape)
        self.observation_histories = z(*obs_history_shape)
        self.rewards = z(1)
        self.actions = z(*actions_shape)
        self.dones = z(1, dtype=torch.uint8)
        self.actions_log_prob = z(1)
        self.values = z(1)
        self.returns = z(1)
        self.advantages = z(1)
        self.mu = z(*actions_shape)
        self.sigma = z(*actions_shape)
        self.env_bins = z(1)
        self.stats = torch.zeros(3, dtype=torch.float64, device=self.device)       # {sum A, sum A^2, count}
        self.num_transitions_per_env, self.num_envs = T, N
        self.step = 0
        self._ptrs = _abi.lt_rollout_ptrs()
        for name, _ in _abi.lt_rollout_ptrs._fields_:
            setattr(self._ptrs, name, getattr(self, name).data_ptr())
        self._keep = None

    @staticmethod
    def _f32(t, n, w):
        if t is None:
            return None
        t = t.detach()
        if t.dtype != torch.float32 or not t.is_contiguous():
            t = t.float().contiguous()
        if t.numel() != n * w:
            raise ValueError(f"transition field has {t.numel()} elements, expected {n}x{w}")
        return t

    @staticmethod
    def _u8(t, n):
        if t is None:
            return None
        t = t.detach()
        if t.dtype == torch.bool:
            t = t.contiguous().view(torch.uint8)
        elif t.dtype != torch.uint8:
            t = (t != 0).view(torch.uint8)
        if t.numel() != n:
            raise ValueError(f"transition mask has {t.numel()} elements, expected {n}")
        return t.contiguous()

    def _launch_store(self, transition, fields):
        n = self.num_envs
        no, npv, nh, na = (int(s[0]) for s in (self.obs_shape, self.privileged_obs_shape, self.obs_history_shape, self.actions_shape))
        widths = dict(observations=no, privileged_observations=npv, observation_histories=nh, actions=na, rewards=1, values=1,
                      actions_log_prob=1, action_mean=na, action_sigma=na)
        tr = _abi.lt_transition_ptrs()
        keep = {}
        for k in fields:
            src = getattr(transition, k, None)
            v = self._u8(src, n) if k in ("dones", "time_outs") else self._f32(src, n, widths[k])
            if v is not None and v.device != self.device:
                raise ValueError(f"transition.{k} is on {v.device}, storage on {self.device}")
            keep[k] = v
            setattr(tr, k, v.data_ptr() if v is not None else None)
        self._keep = keep            # sources stay referenced until the next call (the launch is asynchronous)
        _abi.check(self.lib.lt_rollout_store(C.byref(self._ptrs), C.byref(tr), self.step, n, no, npv, nh, na,
                                             C.c_float(PPO_Args.gamma), _stream(self.device)), None, "lt_rollout_store")

    _POLICY_FIELDS = ("observations", "privileged_observations", "observation_histories", "actions", "values", "actions_log_prob",
                      "action_mean", "action_sigma")
    _ENV_FIELDS = ("rewards", "values", "dones", "time_outs")

    def add_transitions(self, transition):
        """rollout_storage.py:57-71 (+ the time-out bootstrap of ppo.py:85-87 when transition.time_outs is set)."""
        if self.step >= self.num_transitions_per_env:
            raise AssertionError("Rollout buffer overflow")
        self._launch_store(transition, self._POLICY_FIELDS + ("rewards", "dones", "time_outs"))
        self.step += 1

    def store_policy_step(self, transition):
        """First half of a transition, at act() time: everything known before env.step().  The env's observation buffers
        are rewritten in place by the step kernels, so they must be captured now, not at process_env_step() time."""
        if self.step >= self.num_transitions_per_env:
            raise AssertionError("Rollout buffer overflow")
        self._launch_store(transition, self._POLICY_FIELDS)

    def store_env_step(self, transition):
        """Second half, after env.step(): rewards (+ time-out bootstrap from the stored values) and dones."""
        self._launch_store(transition, self._ENV_FIELDS)
        self.step += 1

    def clear(self):
        self.step = 0

    def compute_returns(self, last_values, gamma, lam):
        T, n = self.num_transitions_per_env, self.num_envs
        lv = self._f32(last_values, n, 1)
        s = _stream(self.device)
        _abi.check(self.lib.lt_gae(_ptr(self.rewards), _ptr(self.values), _ptr(self.dones), _ptr(lv), _ptr(self.returns),
                                   _ptr(self.advantages), _ptr(self.stats), T, n, C.c_float(gamma), C.c_float(lam), s), None, "lt_gae")
        ltd.merge_moments(self.stats)       # env-sharded ranks normalise with the whole job's mean / std
        _abi.check(self.lib.lt_normalize_advantages(_ptr(self.advantages), _ptr(self.stats), T * n, s), None,
                   "lt_normalize_advantages")
        self._keep_lv = lv

    def get_statistics(self):               # rollout_storage.py:92-98
        done = self.dones
        done[-1] = 1
        flat_dones = done.permute(1, 0, 2).reshape(-1, 1)
        done_indices = torch.cat((flat_dones.new_tensor([-1], dtype=torch.int64), flat_dones.nonzero(as_tuple=False)[:, 0]))
        trajectory_lengths = (done_indices[1:] - done_indices[:-1])
        return trajectory_lengths.float().mean(), self.rewards.mean()

    def flat_views(self):
        return dict(observations=self.observations.flatten(0, 1), privileged_obs=self.privileged_observations.flatten(0, 1),
                    obs_history=self.observation_histories.flatten(0, 1), actions=self.actions.flatten(0, 1),
                    values=self.values.flatten(0, 1), returns=self.returns.flatten(0, 1),
                    old_actions_log_prob=self.actions_log_prob.flatten(0, 1), advantages=self.advantages.flatten(0, 1),
                    old_mu=self.mu.flatten(0, 1), old_sigma=self.sigma.flatten(0, 1), env_bins=self.env_bins.flatten(0, 1))

    def mini_batch_generator(self, num_mini_batches, num_epochs=8, indices=None):
        """rollout_storage.py:99-137: ONE permutation per update, reused by every epoch.  `indices` injects it (tests)."""
        batch_size = self.num_envs * self.num_transitions_per_env
        mini_batch_size = batch_size // num_mini_batches
        if indices is None:
            indices = torch.randperm(num_mini_batches * mini_batch_size, requires_grad=False, device=self.device)
        f = self.flat_views()
        for epoch in range(num_epochs):
            for i in range(num_mini_batches):
                b = indices[i * mini_batch_size:(i + 1) * mini_batch_size]
                yield (f["observations"][b], f["observations"][b], f["privileged_obs"][b], f["obs_history"][b], f["actions"][b],
                       f["values"][b], f["advantages"][b], f["returns"][b], f["old_actions_log_prob"][b], f["old_mu"][b],
                       f["old_sigma"][b], None, f["env_bins"][b])


def ppo_losses(actor_critic, batch, args=PPO_Args):
    """Surrogate / value / entropy terms of one minibatch (ppo.py:104-141).  Pure torch, device agnostic.
    -> (loss, surrogate_loss, value_loss, kl_mean or None)"""
    (_, _, privileged_obs_batch, obs_history_batch, actions_batch, target_values_batch, advantages_batch, returns_batch,
     old_actions_log_prob_batch, old_mu_batch, old_sigma_batch, masks_batch, _) = batch
    actor_critic.act(obs_history_batch, masks=masks_batch)
    actions_log_prob_batch = actor_critic.get_actions_log_prob(actions_batch)
    value_batch = actor_critic.evaluate(obs_history_batch, privileged_obs_batch, masks=masks_batch)
    mu_batch, sigma_batch, entropy_batch = actor_critic.action_mean, actor_critic.action_std, actor_critic.entropy
    kl_mean = None
    if args.desired_kl is not None and args.schedule == 'adaptive':
        with torch.no_grad():
            kl = torch.sum(torch.log(sigma_batch / old_sigma_batch + 1.e-5)
                           + (torch.square(old_sigma_batch) + torch.square(old_mu_batch - mu_batch))
                           / (2.0 * torch.square(sigma_batch)) - 0.5, axis=-1)
            kl_mean = torch.mean(kl)
    ratio = torch.exp(actions_log_prob_batch - torch.squeeze(old_actions_log_prob_batch))
    adv = torch.squeeze(advantages_batch)
    surrogate = -adv * ratio
    surrogate_clipped = -adv * torch.clamp(ratio, 1.0 - args.clip_param, 1.0 + args.clip_param)
    surrogate_loss = torch.max(surrogate, surrogate_clipped).mean()
    if args.use_clipped_value_loss:
        value_clipped = target_values_batch + (value_batch - target_values_batch).clamp(-args.clip_param, args.clip_param)
        value_loss = torch.max((value_batch - returns_batch).pow(2), (value_clipped - returns_batch).pow(2)).mean()
    else:
        value_loss = (returns_batch - value_batch).pow(2).mean()
    loss = surrogate_loss + args.value_loss_coef * value_loss - args.entropy_coef * entropy_batch.mean()
    return loss, surrogate_loss, value_loss, kl_mean


def next_learning_rate(lr, kl_mean, args=PPO_Args):
    """Adaptive-KL schedule (ppo.py:118-121)."""
    if kl_mean > args.desired_kl * 2.0:
        return max(1e-5, lr / 1.5)
    if kl_mean < args.desired_kl / 2.0 and kl_mean > 0.0:
        return min(1e-2, lr * 1.5)
    return lr


class FlatParams:
    """All parameters of a module re-homed as views of ONE flat buffer (and their .grad as views of one flat
    gradient bucket): Adam, the gradient-norm clip and the cross-rank gradient exchange each become a single
    launch over ~0.6M floats instead of 26 small ones.  state_dict()/load_state_dict() are unaffected."""

    def __init__(self, module):
        self.module = module
        self.plist = list(module.parameters())
        n = sum(p.numel() for p in self.plist)
        p0 = self.plist[0]
        flat = torch.empty(n, dtype=p0.dtype, device=p0.device)
        self.offsets = {}
        o = 0
        for p in self.plist:
            flat[o:o + p.numel()].copy_(p.data.reshape(-1))
            p.data = flat[o:o + p.numel()].view_as(p)
            self.offsets[id(p)] = (o, o + p.numel())
            o += p.numel()
        self.flat = nn.Parameter(flat)
        self.bucket = ltd.GradBucket(self.plist)
        self.flat.grad = self.bucket.flat

    def slice_of(self, submodule):
        """(lo, hi) of a submodule's parameters inside the flat buffer (must be contiguous)."""
        spans = sorted(self.offsets[id(p)] for p in submodule.parameters())
        for (a, b), (c, d) in zip(spans[:-1], spans[1:]):
            assert b == c, "submodule parameters are not contiguous in the flat buffer"
        return spans[0][0], spans[-1][1]


class PPO:
    """ppo.py:33-206."""

    def __init__(self, actor_critic, device='cuda:0'):
        self.device = torch.device(device)
        self.actor_critic = actor_critic
        self.actor_critic.to(self.device)
        self.storage = None
        self.fp = FlatParams(self.actor_critic)
        self.optimizer = torch.optim.Adam([self.fp.flat], lr=PPO_Args.learning_rate)
        # ppo.py:45-46 builds the second Adam over ALL parameters, but only the adaptation module ever holds a gradient
        # when it steps (zero_grad() leaves the others None, which Adam skips): an Adam over that slice is the same update.
        lo, hi = self.fp.slice_of(self.actor_critic.adaptation_module)
        self._ad_slice = (lo, hi)
        self.ad_flat = nn.Parameter(self.fp.flat.data[lo:hi])
        self.ad_flat.grad = self.fp.bucket.flat[lo:hi]
        self.adaptation_module_optimizer = torch.optim.Adam([self.ad_flat], lr=PPO_Args.adaptation_module_learning_rate)
        self.transition = RolloutStorage.Transition()
        self.learning_rate = PPO_Args.learning_rate
        self.last_update_trace = None

    def init_storage(self, num_envs, num_transitions_per_env, actor_obs_shape, privileged_obs_shape, obs_history_shape, action_shape):
        self.storage = RolloutStorage(num_envs, num_transitions_per_env, actor_obs_shape, privileged_obs_shape,
                                      obs_history_shape, action_shape, self.device)

    def test_mode(self):
        self.actor_critic.eval()

    def train_mode(self):
        self.actor_critic.train()

    def act(self, obs, privileged_obs, obs_history, actions=None):
        """ppo.py:65-78.  `actions` (tests only) replaces the sample so that a recorded rollout can be replayed."""
        sampled = self.actor_critic.act(obs_history).detach()
        self.transition.actions = sampled if actions is None else actions
        self.transition.values = self.actor_critic.evaluate(obs_history, privileged_obs).detach()
        self.transition.actions_log_prob = self.actor_critic.get_actions_log_prob(self.transition.actions).detach()
        self.transition.action_mean = self.actor_critic.action_mean.detach()
        self.transition.action_sigma = self.actor_critic.action_std.detach()
        self.transition.observations = obs
        self.transition.critic_observations = obs
        self.transition.privileged_observations = privileged_obs
        self.transition.observation_histories = obs_history
        self.storage.store_policy_step(self.transition)       # captured now: env.step() rewrites its buffers in place
        return self.transition.actions

    def process_env_step(self, rewards, dones, infos):
        """ppo.py:80-93.  The time-out bootstrap rewards += gamma * values * time_outs happens inside lt_rollout_store."""
        self.transition.rewards = rewards
        self.transition.dones = dones
        self.transition.time_outs = infos['time_outs'].to(self.device) if 'time_outs' in infos else None
        self.storage.store_env_step(self.transition)
        self.transition.clear()
        self.actor_critic.reset(dones)

    def compute_returns(self, last_critic_obs, last_critic_privileged_obs):
        last_values = self.actor_critic.evaluate(last_critic_obs, last_critic_privileged_obs).detach()
        self.storage.compute_returns(last_values, PPO_Args.gamma, PPO_Args.lam)

    def update(self, indices=None):
        """ppo.py:95-206.  `indices` (tests only) injects the minibatch permutation."""
        A = PPO_Args
        bucket = self.fp.bucket
        lo, hi = self._ad_slice
        # per-minibatch scalars stay on the device; one host read at the end (the reference syncs 4x per minibatch)
        acc = torch.zeros(4, device=self.device)
        trace = dict(grad_norms=[], lrs=[])
        n_mb = 0
        for batch in self.storage.mini_batch_generator(A.num_mini_batches, A.num_learning_epochs, indices=indices):
            loss, surrogate_loss, value_loss, kl_mean = ppo_losses(self.actor_critic, batch, A)
            if kl_mean is not None:
                kl_mean = ltd.allreduce_mean_scalar(kl_mean)          # every rank takes the same branch
                self.learning_rate = next_learning_rate(self.learning_rate, kl_mean.item(), A)
                for g in self.optimizer.param_groups:
                    g['lr'] = self.learning_rate
            bucket.attach()
            loss.backward()
            bucket.allreduce_mean()
            total_norm = torch.linalg.vector_norm(bucket.flat)
            bucket.flat.mul_(torch.clamp(A.max_grad_norm / (total_norm + 1e-6), max=1.0))   # nn.utils.clip_grad_norm_
            self.optimizer.step()
            trace["grad_norms"].append(total_norm)
            trace["lrs"].append(self.learning_rate)
            acc[0] += value_loss.detach()
            acc[1] += surrogate_loss.detach()
            privileged_obs_batch, obs_history_batch = batch[2], batch[3]
            num_train = int(privileged_obs_batch.shape[0] // 5 * 4)
            for _ in range(A.num_adaptation_module_substeps):
                pred = self.actor_critic.adaptation_module(obs_history_batch)
                adaptation_loss = F.mse_loss(pred[:num_train], privileged_obs_batch[:num_train])
                with torch.no_grad():
                    adaptation_test_loss = F.mse_loss(pred[num_train:], privileged_obs_batch[num_train:])
                bucket.flat[lo:hi].zero_()
                adaptation_loss.backward()
                if ltd.world_size() > 1:
                    ltd.allreduce_sum_(bucket.flat[lo:hi]).div_(ltd.world_size())
                self.adaptation_module_optimizer.step()
                acc[2] += adaptation_loss.detach()
                acc[3] += adaptation_test_loss
            n_mb += 1
        num_updates = A.num_learning_epochs * A.num_mini_batches
        acc = acc.tolist()
        n_sub = num_updates * A.num_adaptation_module_substeps
        trace["grad_norms"] = [float(x) for x in torch.stack(trace["grad_norms"]).tolist()] if trace["grad_norms"] else []
        self.last_update_trace = trace
        self.storage.clear()
        return acc[0] / num_updates, acc[1] / num_updates, acc[2] / n_sub, 0.0, 0.0, acc[3] / n_sub, 0.0, 0.0


class Runner:
    """ppo_cse/__init__.py:65-357.  wandb / video logging of the reference is replaced by an optional `log_fn(dict, step)`
    callback; checkpoints (ac_weights.pt + the two TorchScript modules) go to `save_dir`."""

    def __init__(self, env, device='cuda:0', runner_args=RunnerArgs, ac_args=AC_Args, log_fn=None, save_dir="last_run/checkpoints"):
        self.device = device
        self.env = env
        self.runner_args = runner_args
        self.log_fn = log_fn
        self.save_dir = save_dir
        actor_critic = ActorCritic(self.env.num_obs, self.env.num_privileged_obs, self.env.num_obs_history,
                                   self.env.num_actions).to(self.device)
        if runner_args.resume:
            actor_critic.load_state_dict(torch.load(runner_args.resume, map_location=self.device))
        self.alg = PPO(actor_critic, device=self.device)
        self.num_steps_per_env = runner_args.num_steps_per_env
        self.alg.init_storage(self.env.num_train_envs, self.num_steps_per_env, [self.env.num_obs],
                              [self.env.num_privileged_obs], [self.env.num_obs_history], [self.env.num_actions])
        self.tot_timesteps = 0
        self.tot_time = 0
        self.current_learning_iteration = 0
        self.last_metrics = None
        self.env.reset()

    def learn(self, num_learning_iterations, init_at_random_ep_len=False, eval_freq=100, curriculum_dump_freq=500,
              eval_expert=False, update_model=True):
        num_train_envs = self.env.num_train_envs
        obs_dict = self.env.get_observations()
        obs, privileged_obs, obs_history = (obs_dict[k].to(self.device) for k in ("obs", "privileged_obs", "obs_history"))
        self.alg.actor_critic.train()
        self.rewbuffer, self.lenbuffer = deque(maxlen=100), deque(maxlen=100)
        cur_reward_sum = torch.zeros(self.env.num_envs, dtype=torch.float, device=self.device)
        cur_episode_length = torch.zeros(self.env.num_envs, dtype=torch.float, device=self.device)
        if init_at_random_ep_len:
            self.env.episode_length_buf[:] = torch.randint_like(self.env.episode_length_buf, high=int(self.env.max_episode_length))
        very_start = time.time()
        for it in range(num_learning_iterations):
            start = time.time()
            with torch.inference_mode():
                for i in range(self.num_steps_per_env):
                    actions_train = self.alg.act(obs[:num_train_envs], privileged_obs[:num_train_envs], obs_history[:num_train_envs])
                    if num_train_envs < self.env.num_envs:
                        if eval_expert:
                            actions_eval = self.alg.actor_critic.act_teacher(obs_history[num_train_envs:], privileged_obs[num_train_envs:])
                        else:
                            actions_eval = self.alg.actor_critic.act_student(obs_history[num_train_envs:])
                        actions_all = torch.cat((actions_train, actions_eval), dim=0)
                    else:
                        actions_all = actions_train
                    obs_dict, rewards, dones, infos = self.env.step(actions_all)
                    obs, privileged_obs, obs_history = (obs_dict[k].to(self.device) for k in ("obs", "privileged_obs", "obs_history"))
                    rewards, dones = rewards.to(self.device), dones.to(self.device)
                    if update_model:
                        self.alg.process_env_step(rewards[:num_train_envs], dones[:num_train_envs], infos)
                    if self.log_fn is not None and 'train/episode' in infos:
                        self.log_fn({"train": dict(infos['train/episode'])}, it * self.num_steps_per_env + i)
                    if 'curriculum' in infos:
                        cur_reward_sum += rewards
                        cur_episode_length += 1
                        new_ids = (dones > 0).nonzero(as_tuple=False)
                        new_ids_train = new_ids[new_ids < num_train_envs]
                        if new_ids_train.numel():
                            self.rewbuffer.extend(cur_reward_sum[new_ids_train].cpu().numpy().tolist())
                            self.lenbuffer.extend(cur_episode_length[new_ids_train].cpu().numpy().tolist())
                        cur_reward_sum[new_ids] = 0
                        cur_episode_length[new_ids] = 0
                collection_time = time.time() - start
                start = time.time()
                if update_model:
                    self.alg.compute_returns(obs_history[:num_train_envs], privileged_obs[:num_train_envs])
            losses = self.alg.update() if update_model else (0,) * 8
            learn_time = time.time() - start
            names = ("mean_value_loss", "mean_surrogate_loss", "adaptation_loss", "mean_decoder_loss", "mean_decoder_loss_student",
                     "mean_adaptation_module_test_loss", "mean_decoder_test_loss", "mean_decoder_test_loss_student")
            self.last_metrics = dict(zip(names, losses), collection_time=collection_time, learn_time=learn_time,
                                     learning_rate=self.alg.learning_rate,
                                     fps=(it + 1) * self.env.num_envs * self.num_steps_per_env * ltd.world_size() / (time.time() - very_start))
            if self.log_fn is not None:
                self.log_fn({"metrics": self.last_metrics}, it * self.num_steps_per_env + self.num_steps_per_env - 1)
            self.tot_timesteps += self.num_steps_per_env * self.env.num_envs
            self.current_learning_iteration += 1
            if it % self.runner_args.save_interval == 0 and ltd.rank() == 0 and self.save_dir:
                self.save(self.save_dir)
        if ltd.rank() == 0 and self.save_dir:
            self.save(self.save_dir)

    def save(self, save_path):
        """ac_weights.pt + adaptation_module_latest.jit + body_latest.jit (ppo_cse/__init__.py:282-303)."""
        os.makedirs(save_path, exist_ok=True)
        torch.save(self.alg.actor_critic.state_dict(), os.path.join(save_path, "ac_weights.pt"))
        for name, mod in (("adaptation_module_latest.jit", self.alg.actor_critic.adaptation_module),
                          ("body_latest.jit", self.alg.actor_critic.actor_body)):
            m = copy.deepcopy(mod).to('cpu')
            torch.jit.script(m).save(os.path.join(save_path, name))

    def get_inference_policy(self, device=None):
        self.alg.actor_critic.eval()
        if device is not None and torch.device(device) != self.alg.device:
            return copy.deepcopy(self.alg.actor_critic).to(device).act_inference      # the live copy stays flat-bucketed
        return self.alg.actor_critic.act_inference

    def get_expert_policy(self, device=None):
        self.alg.actor_critic.eval()
        if device is not None and torch.device(device) != self.alg.device:
            return copy.deepcopy(self.alg.actor_critic).to(device).act_expert
        return self.alg.actor_critic.act_expert
