"""CPU restatement of the reference's PPO rows.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module; the
product (legged_tracking_b200/) never does.  Pinned against tests/golden/ppo_*.npz, which oracle/gen_golden_ppo.py
recorded from the unmodified reference (tests/test_oracle_ppo_golden.py).

Follows, function by function:
  go1_gym_learn/ppo_cse/rollout_storage.py:57-71   add_transitions       -> PpoPort.store
  go1_gym_learn/ppo_cse/ppo.py:80-93               process_env_step      -> PpoPort.store (time-out bootstrap)
  go1_gym_learn/ppo_cse/rollout_storage.py:76-90   compute_returns       -> gae
  go1_gym_learn/ppo_cse/rollout_storage.py:99-137  mini_batch_generator  -> PpoPort.update (one permutation, reused)
  go1_gym_learn/ppo_cse/actor_critic.py:118-160    update_distribution / evaluate / log-prob / entropy -> policy_forward
  go1_gym_learn/ppo_cse/ppo.py:99-206              update                -> PpoPort.update
  go1_gym_learn/ppo_cse_cnn/actor_critic.py:27-62, 181-198   HeightMapEncoder (MLP form) / process_obs_history -> policy_input
      (row f4: the state_dict then holds 'height_map_encoder.model.*'; everything else of ppo_cse_cnn is ppo_cse's code)
"""
import math

import torch
import torch.nn.functional as F

ACTIVATIONS = {"elu": F.elu, "selu": F.selu, "relu": F.relu, "crelu": F.relu, "lrelu": F.leaky_relu,
               "tanh": torch.tanh, "sigmoid": torch.sigmoid}          # actor_critic.py:162-178

PPO_DEFAULTS = dict(value_loss_coef=1.0, use_clipped_value_loss=True, clip_param=0.2, entropy_coef=0.01,
                    num_learning_epochs=5, num_mini_batches=4, learning_rate=1e-3, adaptation_module_learning_rate=1e-3,
                    num_adaptation_module_substeps=1, schedule="adaptive", gamma=0.99, lam=0.95, desired_kl=0.01,
                    max_grad_norm=1.0)                                  # ppo.py:13-30


def mlp(params, prefix, x, act):
    """nn.Sequential of Linear / activation pairs, last Linear bare (actor_critic.py:43-82).  Keys are the
    reference's state_dict names: '<prefix>.<2k>.weight' / '.bias'."""
    idx = sorted({int(k.split(".")[1]) for k in params if k.startswith(prefix + ".")})
    for j, i in enumerate(idx):
        x = F.linear(x, params[f"{prefix}.{i}.weight"], params[f"{prefix}.{i}.bias"])
        if j != len(idx) - 1:
            x = act(x)
    return x


ENCODER = "height_map_encoder.model"


def policy_input(params, act, obs_history, num_obs=None):
    """What the three bodies read.  ppo_cse: the observation history itself.  ppo_cse_cnn (params hold the encoder,
    use_cnn = use_gru = False): process_obs_history, ppo_cse_cnn/actor_critic.py:181-198 -- every history row's height columns go
    through the encoder (Linear, act, Linear, act; :46-52), recurrent_latent_embedding is the Identity, and the result is
    cat(last row's scalars, [last row's scalars, last row's embedding])."""
    if ENCODER + ".0.weight" not in params:
        return obs_history
    hm = params[ENCODER + ".0.weight"].shape[1]
    emb_w = params[ENCODER + ".2.weight"].shape[0]
    batch = obs_history.shape[0]
    hist = obs_history.view(batch, -1, num_obs)
    height_map = hist[:, :, -hm:].reshape(-1, hm)
    emb = act(F.linear(act(F.linear(height_map, params[ENCODER + ".0.weight"], params[ENCODER + ".0.bias"])),
                       params[ENCODER + ".2.weight"], params[ENCODER + ".2.bias"])).reshape(batch, -1, emb_w)
    lstm_input = torch.cat([hist[:, :, :-hm], emb], dim=-1)
    recurrent_latent = lstm_input                                                  # nn.Identity (:103)
    return torch.cat([hist[:, -1, :-hm], recurrent_latent[:, -1, :]], dim=-1)


def policy_forward(params, act, obs_history, privileged_obs=None, num_obs=None):
    """-> (mean, sigma, value).  update_distribution (actor_critic.py:118-121): the actor reads cat(obs_history,
    adaptation_module(obs_history)); evaluate (:155-157): the critic reads cat(obs_history, privileged_obs)."""
    obs_history = policy_input(params, act, obs_history, num_obs)
    latent = mlp(params, "adaptation_module", obs_history, act)
    mean = mlp(params, "actor_body", torch.cat((obs_history, latent), dim=-1), act)
    sigma = mean * 0.0 + params["std"]
    value = None
    if privileged_obs is not None:
        value = mlp(params, "critic_body", torch.cat((obs_history, privileged_obs), dim=-1), act)
    return mean, sigma, value


def normal_log_prob(actions, mean, sigma):
    """torch.distributions.Normal.log_prob summed over the action dim (actor_critic.py:131-132)."""
    var = sigma ** 2
    return (-((actions - mean) ** 2) / (2 * var) - sigma.log() - math.log(math.sqrt(2 * math.pi))).sum(dim=-1)


def normal_entropy(sigma):
    return (0.5 + 0.5 * math.log(2 * math.pi) + torch.log(sigma)).sum(dim=-1)        # actor_critic.py:114-116


def gae(rewards, values, dones, last_values, gamma, lam):
    """rollout_storage.py:76-90.  rewards/values [T,N,1] float, dones [T,N,1] uint8, last_values [N,1].
    -> (returns, normalised advantages)."""
    T = rewards.shape[0]
    returns = torch.zeros_like(rewards)
    advantage = 0
    for step in reversed(range(T)):
        next_values = last_values if step == T - 1 else values[step + 1]
        not_terminal = 1.0 - dones[step].float()
        delta = rewards[step] + not_terminal * gamma * next_values - values[step]
        advantage = delta + not_terminal * gamma * lam * advantage
        returns[step] = advantage + values[step]
    adv = returns - values
    return returns, (adv - adv.mean()) / (adv.std() + 1e-8)


def minibatch_loss(params, act, hist, priv, actions, target_values, advantages, returns, old_logp, old_mu, old_sigma, a, num_obs=None):
    """ppo.py:104-141 for one minibatch.  -> (loss, surrogate_loss, value_loss, kl_mean or None)"""
    mu, sigma, value = policy_forward(params, act, hist, priv, num_obs)
    logp = normal_log_prob(actions, mu, sigma)
    entropy = normal_entropy(sigma)
    kl_mean = None
    if a["desired_kl"] is not None and a["schedule"] == "adaptive":
        with torch.no_grad():
            kl = torch.sum(torch.log(sigma / old_sigma + 1.e-5)
                           + (torch.square(old_sigma) + torch.square(old_mu - mu)) / (2.0 * torch.square(sigma)) - 0.5, axis=-1)
            kl_mean = torch.mean(kl)
    ratio = torch.exp(logp - torch.squeeze(old_logp))
    adv = torch.squeeze(advantages)
    surrogate_loss = torch.max(-adv * ratio, -adv * torch.clamp(ratio, 1.0 - a["clip_param"], 1.0 + a["clip_param"])).mean()
    if a["use_clipped_value_loss"]:
        value_clipped = target_values + (value - target_values).clamp(-a["clip_param"], a["clip_param"])
        value_loss = torch.max((value - returns).pow(2), (value_clipped - returns).pow(2)).mean()
    else:
        value_loss = (returns - value).pow(2).mean()
    loss = surrogate_loss + a["value_loss_coef"] * value_loss - a["entropy_coef"] * entropy.mean()
    return loss, surrogate_loss, value_loss, kl_mean


class PpoPort:
    """State: params (dict of leaf tensors named as the reference's state_dict), two Adam optimisers over ALL
    parameters (ppo.py:44-46 -- the second one only ever sees adaptation-module gradients), the [T,N,*] storage."""

    def __init__(self, state_dict, sizes, activation="elu", **ppo_args):
        self.N, self.T, self.num_obs, self.num_priv, self.H, self.num_actions = (int(s) for s in sizes)
        self.args = dict(PPO_DEFAULTS)
        self.args.update(ppo_args)
        self.act = ACTIVATIONS[activation]
        self.params = {k: torch.as_tensor(v).clone().requires_grad_(True) for k, v in state_dict.items()}
        plist = list(self.params.values())
        self.optimizer = torch.optim.Adam(plist, lr=self.args["learning_rate"])
        self.adaptation_optimizer = torch.optim.Adam(plist, lr=self.args["adaptation_module_learning_rate"])
        self.learning_rate = self.args["learning_rate"]
        T, N = self.T, self.N
        z = lambda w: torch.zeros(T, N, w)
        self.s = dict(observations=z(self.num_obs), privileged_observations=z(self.num_priv),
                      observation_histories=z(self.num_obs * self.H), actions=z(self.num_actions), rewards=z(1),
                      values=z(1), actions_log_prob=z(1), mu=z(self.num_actions), sigma=z(self.num_actions),
                      dones=torch.zeros(T, N, 1, dtype=torch.uint8))
        self.step = 0

    @torch.no_grad()
    def store(self, obs, priv, hist, actions, rewards, dones, time_outs=None):
        """PPO.act's bookkeeping (ppo.py:65-78) for GIVEN actions + process_env_step (:80-93) + add_transitions."""
        mean, sigma, value = policy_forward(self.params, self.act, hist, priv, self.num_obs)
        t = self.step
        if t >= self.T:
            raise AssertionError("Rollout buffer overflow")
        r = rewards.clone()
        if time_outs is not None:
            r += self.args["gamma"] * torch.squeeze(value * time_outs.unsqueeze(1), 1)
        s = self.s
        s["observations"][t] = obs
        s["privileged_observations"][t] = priv
        s["observation_histories"][t] = hist
        s["actions"][t] = actions
        s["rewards"][t] = r.view(-1, 1)
        s["dones"][t] = dones.view(-1, 1)
        s["values"][t] = value
        s["actions_log_prob"][t] = normal_log_prob(actions, mean, sigma).view(-1, 1)
        s["mu"][t] = mean
        s["sigma"][t] = sigma
        self.step += 1

    @torch.no_grad()
    def compute_returns(self, last_hist, last_priv):
        _, _, last_values = policy_forward(self.params, self.act, last_hist, last_priv, self.num_obs)
        self.s["last_values"] = last_values
        self.s["returns"], self.s["advantages"] = gae(self.s["rewards"], self.s["values"], self.s["dones"], last_values,
                                                      self.args["gamma"], self.args["lam"])

    def update(self, perm):
        """ppo.py:99-206 with the permutation given (the reference draws torch.randperm once per update and reuses
        it for every epoch, rollout_storage.py:102).  -> dict(losses, grad_norms, lrs)."""
        a = self.args
        flat = {k: v.flatten(0, 1) for k, v in self.s.items() if k != "last_values"}
        B = self.N * self.T
        nmb = a["num_mini_batches"]
        mb = B // nmb
        plist = list(self.params.values())
        sums = dict(value=0.0, surrogate=0.0, adaptation=0.0, adaptation_test=0.0)
        norms, lrs = [], []
        for _ in range(a["num_learning_epochs"]):
            for i in range(nmb):
                idx = perm[i * mb:(i + 1) * mb]
                hist, priv, actions = flat["observation_histories"][idx], flat["privileged_observations"][idx], flat["actions"][idx]
                target_values, advantages, returns = flat["values"][idx], flat["advantages"][idx], flat["returns"][idx]
                old_logp, old_mu, old_sigma = flat["actions_log_prob"][idx], flat["mu"][idx], flat["sigma"][idx]
                loss, surrogate_loss, value_loss, kl_mean = minibatch_loss(
                    self.params, self.act, hist, priv, actions, target_values, advantages, returns, old_logp, old_mu, old_sigma, a, self.num_obs)
                if kl_mean is not None:
                    if kl_mean > a["desired_kl"] * 2.0:
                        self.learning_rate = max(1e-5, self.learning_rate / 1.5)
                    elif kl_mean < a["desired_kl"] / 2.0 and kl_mean > 0.0:
                        self.learning_rate = min(1e-2, self.learning_rate * 1.5)
                    for g in self.optimizer.param_groups:
                        g["lr"] = self.learning_rate
                self.optimizer.zero_grad()
                loss.backward()
                norms.append(float(torch.nn.utils.clip_grad_norm_(plist, a["max_grad_norm"])))
                lrs.append(self.optimizer.param_groups[0]["lr"])
                self.optimizer.step()
                sums["value"] += value_loss.item()
                sums["surrogate"] += surrogate_loss.item()
                num_train = int(priv.shape[0] // 5 * 4)                     # ppo.py:160-161
                for _sub in range(a["num_adaptation_module_substeps"]):
                    pred = mlp(self.params, "adaptation_module", policy_input(self.params, self.act, hist, self.num_obs), self.act)
                    ad_loss = F.mse_loss(pred[:num_train], priv[:num_train])
                    ad_test = F.mse_loss(pred[num_train:], priv[num_train:])
                    self.adaptation_optimizer.zero_grad()
                    ad_loss.backward()
                    self.adaptation_optimizer.step()
                    sums["adaptation"] += ad_loss.item()
                    sums["adaptation_test"] += ad_test.item()
        nup = a["num_learning_epochs"] * nmb
        nsub = nup * a["num_adaptation_module_substeps"]
        self.step = 0
        losses = (sums["value"] / nup, sums["surrogate"] / nup, sums["adaptation"] / nsub, 0.0, 0.0,
                  sums["adaptation_test"] / nsub, 0.0, 0.0)
        return dict(losses=losses, grad_norms=norms, lrs=lrs)
