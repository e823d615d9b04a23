"""PPO rows of the path behind the reference's own class names and call signatures.

  AC_Args / ActorCritic      go1_gym_learn/ppo_cse/actor_critic.py:9-160   (same constructor, method names and state_dict keys: a
                                                                           checkpoint of either loads into the other)
  RolloutStorage             go1_gym_learn/ppo_cse/rollout_storage.py       ([T,N,*] buffers in HBM; add_transitions,
                                                                           compute_returns run as liblt_b200 kernels)
  PPO_Args / PPO             go1_gym_learn/ppo_cse/ppo.py                   (act / process_env_step / compute_returns / update)
  RunnerArgs / Runner        go1_gym_learn/ppo_cse/__init__.py:46-357       (learn loop; logging back-ends left to the caller)

Nothing here computes: every forward, loss, backward, clip and Adam step is a call into liblt_b200.so (include/lt_policy.h --
tcgen05 split-fp16 contractions with fp32 accumulation, fused heads), and the rollout buffers are written by
lt_rollout_store / lt_policy_act.  There is no CPU path and no torch autograd path: on a non-CUDA device the classes raise.
This module owns the memory (one flat parameter vector, one [stats | gradient] exchange buffer, Adam moments, the rollout
buffers), maps the reference's state_dict keys onto the flat vector, and runs the cross-rank exchange (dist.GradExchange).
"""
import ctypes as C
import math
import os
import time
from collections import OrderedDict, deque

import torch

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


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _f32c(t, device, what):
    """The kernels read float32, contiguous, on the policy's device; anything else is the caller's bug, not a silent copy
    chain -- except dtype / layout, which the reference also accepts."""
    if t is None:
        return None
    if t.device != device:
        raise _abi.LtError(f"{what} is on {t.device}, the policy on {device}")
    t = t.detach()
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.float().contiguous()
    return t


_BODIES = (("adaptation_module", "adaptation_module_branch_hidden_dims"), ("actor_body", "actor_hidden_dims"),
           ("critic_body", "critic_hidden_dims"))
_LOG_SQRT_2PI = math.log(math.sqrt(2.0 * math.pi))
# get_activation (actor_critic.py:162-178) -> LT_ACT_* of include/lt_policy.h
_ACTIVATIONS = dict(elu="LT_ACT_ELU", selu="LT_ACT_SELU", relu="LT_ACT_RELU", crelu="LT_ACT_RELU", lrelu="LT_ACT_LRELU",
                    tanh="LT_ACT_TANH", sigmoid="LT_ACT_SIGMOID")


class ActorCritic:
    """actor_critic.py:23-159 as a handle on an lt_policy (include/lt_policy.h).

    Parameters live in ONE flat float32 vector laid out as lt_policy.h states (std, then weight/bias of every Linear of the
    adaptation module, the actor body and the critic body): state_dict() returns views of it under the reference's keys
    ('std', 'adaptation_module.0.weight', ... 'critic_body.6.bias'), load_state_dict() writes through them.  The MLPs
    themselves exist only as tensor-core operand planes inside the library."""
    is_recurrent = False
    ARGS = AC_Args                        # the params_proto class this variant reads its geometry from

    def __init__(self, num_obs, num_privileged_obs, num_obs_history, num_actions, **kwargs):
        args = kwargs.pop("ac_args", None) or self.ARGS
        if kwargs:
            print("ActorCritic.__init__ got unexpected arguments, which will be ignored: " + str(list(kwargs)))
        if args.use_decoder or getattr(args, "normalize_obs", False):
            raise NotImplementedError("use_decoder / normalize_obs are off in every reference config (actor_critic.py:18-20)")
        if args.activation not in _ACTIVATIONS:
            raise ValueError(f"invalid activation function {args.activation!r}")         # actor_critic.py:177-178
        self.ac_args = args
        self.activation = args.activation
        self.decoder = False
        self.normalize_obs = False
        self.num_obs, self.num_obs_history = int(num_obs), int(num_obs_history)
        self.num_privileged_obs, self.num_actions = int(num_privileged_obs), int(num_actions)
        self.hidden = {name: [int(h) for h in getattr(args, attr)] for name, attr in _BODIES}
        self.encoder = self._encoder_spec(args)          # None, or dict(height_map_size, hidden, embedding): ppo_cnn.ActorCritic
        # what the three bodies read in place of the observation history (ppo_cse: the history itself)
        self.policy_input_dim = self.num_obs_history if self.encoder is None else \
            2 * (self.num_obs - self.encoder["height_map_size"]) + self.encoder["embedding"]
        n_in = dict(adaptation_module=self.policy_input_dim, actor_body=self.num_privileged_obs + self.policy_input_dim,
                    critic_body=self.num_privileged_obs + self.policy_input_dim)
        n_out = dict(adaptation_module=self.num_privileged_obs, actor_body=self.num_actions, critic_body=1)
        # flat layout (lt_policy.h): [(key, shape, lo, hi)]
        self.layout = [("std", (self.num_actions,), 0, self.num_actions)]
        o = self.num_actions

        def add_linears(name, dims):
            nonlocal o
            for l, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
                self.layout.append((f"{name}.{2 * l}.weight", (b, a), o, o + a * b))
                o += a * b
                self.layout.append((f"{name}.{2 * l}.bias", (b,), o, o + b))
                o += b
        self.body_slices = {}
        if self.encoder is not None:
            lo_body = o
            add_linears("height_map_encoder.model", [self.encoder["height_map_size"], self.encoder["hidden"], self.encoder["embedding"]])
            self.body_slices["height_map_encoder"] = (lo_body, o)
        for name, _ in _BODIES:
            lo_body = o
            add_linears(name, [n_in[name]] + self.hidden[name] + [n_out[name]])
            self.body_slices[name] = (lo_body, o)
        self.num_params = o
        self.device = torch.device("cpu")
        self.flat = torch.empty(o, dtype=torch.float32)
        self._init_parameters()
        self.grads = None                 # bound by PPO (dist.GradExchange.grads)
        self._h, self._cap, self._training, self._dirty = None, 0, False, True
        self._dist = None                 # (mean, sigma) of the last act(): the reference's self.distribution

    def _encoder_spec(self, args):
        return None

    def _init_parameters(self):
        """torch.nn.Linear's default initialisation, drawn in the order the reference constructs its layers ((encoder,)
        adaptation module, actor, critic; weight then bias; actor_critic.py:43-82), so a seeded run starts from the same weights."""
        views = self._views()
        for key, shape, _, _ in self.layout:
            if key.endswith(".weight"):
                lin = torch.nn.Linear(shape[1], shape[0])          # used as the initialiser only
                views[key].copy_(lin.weight.detach())
                views[key[:-6] + "bias"].copy_(lin.bias.detach())
        views["std"].fill_(float(self.ac_args.init_noise_std))

    def _views(self):
        return OrderedDict((key, self.flat[lo:hi].view(shape)) for key, shape, lo, hi in self.layout)

    # ------------------------------------------------------------------------------------------ nn.Module look-alikes
    def state_dict(self):
        return self._views()

    def load_state_dict(self, state_dict, strict=True):
        views = self._views()
        missing, unexpected = [k for k in views if k not in state_dict], [k for k in state_dict if k not in views]
        if strict and (missing or unexpected):
            raise RuntimeError(f"ActorCritic.load_state_dict: missing keys {missing}, unexpected keys {unexpected}")
        for k, v in views.items():
            if k in state_dict:
                src = torch.as_tensor(state_dict[k])
                if tuple(src.shape) != tuple(v.shape):
                    raise RuntimeError(f"ActorCritic.load_state_dict: {k} has shape {tuple(src.shape)}, expected {tuple(v.shape)}")
                v.copy_(src)
        self._dirty = True

    def parameters(self):
        return iter(self._views().values())

    @property
    def std(self):
        return self.flat[:self.num_actions]

    def to(self, device):
        device = torch.device(device)
        if device.type == "cuda" and device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        if device != self.device:
            self._release()
            self.flat = self.flat.to(device)
            self.grads = None
            self.device = device
        return self

    def train(self, mode=True):
        return self

    def eval(self):
        return self

    def reset(self, dones=None):
        pass

    def forward(self):
        raise NotImplementedError

    def parameters_changed(self):
        """Call after writing self.flat from outside (broadcast, checkpoint surgery): the operand planes are re-derived
        before the next launch."""
        self._dirty = True

    # ------------------------------------------------------------------------------------------ library handle
    def _release(self):
        if self._h is not None:
            _abi.load().lt_policy_destroy(self._h)
        self._h, self._cap, self._training = None, 0, False

    def __del__(self):
        try:
            self._release()
        except Exception:
            pass

    def reserve(self, rows, training=False):
        """Size the library's workspaces for `rows` samples per call (and for gradients)."""
        if self.device.type != "cuda":
            raise _abi.LtError("the actor-critic runs as sm_100a kernels only: move it to a CUDA device (there is no CPU path)")
        rows = int(rows)
        if self._h is not None and rows <= self._cap and (self._training or not training):
            return
        training = training or self._training
        rows = max(rows, self._cap)
        lib = _abi.load()
        if self._h is not None:
            torch.cuda.synchronize(self.device)
            self._release()
        cfg = _abi.lt_policy_config()
        cfg.struct_bytes = C.sizeof(_abi.lt_policy_config)
        cfg.num_obs_history, cfg.num_privileged_obs, cfg.num_actions = self.num_obs_history, self.num_privileged_obs, self.num_actions
        for cname, name in (("adaptation", "adaptation_module"), ("actor", "actor_body"), ("critic", "critic_body")):
            h = self.hidden[name]
            if not 1 <= len(h) <= _abi.DEFINES["LT_POLICY_MAX_HIDDEN"]:
                raise _abi.LtError(f"{name}: {len(h)} hidden layers, the kernels take 1..{_abi.DEFINES['LT_POLICY_MAX_HIDDEN']}")
            setattr(cfg, f"n_{cname}_hidden", len(h))
            arr = getattr(cfg, f"{cname}_hidden")
            for i, w in enumerate(h):
                arr[i] = w
        cfg.max_rows, cfg.training = rows, 1 if training else 0
        cfg.activation = _abi.DEFINES[_ACTIVATIONS[self.activation]]
        if self.encoder is not None:
            cfg.height_map_size, cfg.num_obs = self.encoder["height_map_size"], self.num_obs
            cfg.encoder_hidden, cfg.encoder_embedding = self.encoder["hidden"], self.encoder["embedding"]
        handle = C.c_void_p()
        with torch.cuda.device(self.device):
            _abi.check_policy(lib.lt_policy_create(C.byref(cfg), C.byref(handle)), None, "lt_policy_create")
        self._h, self._cap, self._training = handle, rows, training
        if lib.lt_policy_num_params(handle) != self.num_params:
            raise _abi.LtError("flat parameter layout differs between ppo.py and the library")
        if training and self.grads is None:
            self.grads = torch.zeros(self.num_params, dtype=torch.float32, device=self.device)
        self._bind()

    def _bind(self):
        if self._h is not None:
            _abi.check_policy(_abi.load().lt_policy_bind_params(self._h, _ptr(self.flat), _ptr(self.grads)), self._h, "lt_policy_bind_params")
            self._dirty = True

    def bind_grads(self, grads):
        self.grads = grads
        self._bind()

    def _ready(self, rows, training=False):
        self.reserve(rows, training)
        if self._dirty:
            _abi.check_policy(_abi.load().lt_policy_refresh(self._h, _stream(self.device)), self._h, "lt_policy_refresh")
            self._dirty = False
        return _abi.load()

    # ------------------------------------------------------------------------------------------ forward entry points
    def _forward(self, obs_history, privileged_obs, teacher, want_latent, want_mean, want_value):
        oh = _f32c(obs_history, self.device, "obs_history")
        pv = _f32c(privileged_obs, self.device, "privileged_obs")
        rows = oh.shape[0]
        lib = self._ready(rows)
        new = lambda w: torch.empty(rows, w, dtype=torch.float32, device=self.device)
        latent = new(self.num_privileged_obs) if want_latent else None
        mean = new(self.num_actions) if want_mean else None
        value = new(1) if want_value else None
        _abi.check_policy(lib.lt_policy_forward(self._h, _ptr(oh), _ptr(pv), rows, 1 if teacher else 0, _ptr(latent), _ptr(mean),
                                                _ptr(value), _stream(self.device)), self._h, "lt_policy_forward")
        return latent, mean, value

    def act_into(self, obs_history, privileged_obs, out, actions_in=None, normal_noise=None, seed=0, step=0):
        """PPO.act fused (ppo.py:65-77): forward, sample, log-prob, value, written straight into the tensors of `out`
        (keys actions, actions_log_prob, action_mean, action_sigma, value; any may be missing)."""
        oh = _f32c(obs_history, self.device, "obs_history")
        pv = _f32c(privileged_obs, self.device, "privileged_obs")
        rows = oh.shape[0]
        lib = self._ready(rows)
        given = _f32c(actions_in, self.device, "actions")
        noise = _f32c(normal_noise, self.device, "normal_noise")
        o = lambda k: _ptr(out.get(k))
        _abi.check_policy(lib.lt_policy_act(self._h, _ptr(oh), _ptr(pv), rows, _ptr(noise), _ptr(given), C.c_uint64(seed), C.c_uint64(step),
                                            o("actions"), o("actions_log_prob"), o("action_mean"), o("action_sigma"), o("value"),
                                            _stream(self.device)), self._h, "lt_policy_act")
        self._keep = (oh, pv, given, noise)          # inputs stay referenced until the next call (asynchronous launch)

    def update_distribution(self, observation_history):
        _, mean, _ = self._forward(observation_history, None, False, False, True, False)
        self._dist = (mean, self.std.expand_as(mean))

    def act(self, observation_history, **kwargs):
        rows = observation_history.shape[0]
        new = lambda: torch.empty(rows, self.num_actions, dtype=torch.float32, device=self.device)
        out = dict(actions=new(), action_mean=new(), action_sigma=new())
        self._act_step = getattr(self, "_act_step", 0) + 1
        self.act_into(observation_history, None, out, seed=torch.initial_seed() & (2 ** 63 - 1), step=self._act_step)
        self._dist = (out["action_mean"], out["action_sigma"])
        return out["actions"]

    @property
    def action_mean(self):
        return self._dist[0]

    @property
    def action_std(self):
        return self._dist[1]

    @property
    def entropy(self):
        return (0.5 + _LOG_SQRT_2PI + torch.log(self._dist[1])).sum(dim=-1)

    def get_actions_log_prob(self, actions):
        mean, sigma = self._dist
        return (-((actions - mean) ** 2) / (2 * sigma ** 2) - torch.log(sigma) - _LOG_SQRT_2PI).sum(dim=-1)

    def act_expert(self, ob, policy_info={}):
        return self.act_teacher(ob["obs_history"], ob["privileged_obs"])

    def act_inference(self, ob, policy_info={}):
        return self.act_student(ob["obs_history"], policy_info=policy_info)

    def act_student(self, observation_history, policy_info={}):
        latent, mean, _ = self._forward(observation_history, None, False, True, True, False)
        policy_info["latents"] = latent.cpu().numpy()
        return mean

    def act_teacher(self, observation_history, privileged_info, policy_info={}):
        _, mean, _ = self._forward(observation_history, privileged_info, True, False, True, False)
        policy_info["latents"] = privileged_info
        return mean

    def evaluate(self, observation_history, privileged_observations, **kwargs):
        return self._forward(observation_history, privileged_observations, False, False, False, True)[2]

    def get_student_latent(self, observation_history):
        return self._forward(observation_history, None, False, True, False, False)[0]

    # ------------------------------------------------------------------------------------------ deployment export
    def export_module(self, name):
        """One body as a torch.nn.Sequential on the CPU (Linear / ELU pairs, module indices as in the reference) for the
        TorchScript files the deployment code loads (ppo_cse/__init__.py:282-303).  Export only: nothing here calls it."""
        import torch.nn as nn
        sd = {k: v.detach().cpu().clone() for k, v in self.state_dict().items() if k.startswith(name + ".")}
        n_lin = len(sd) // 2
        layers = []
        for l in range(n_lin):
            w = sd[f"{name}.{2 * l}.weight"]
            layers.append(nn.Linear(w.shape[1], w.shape[0]))
            if l != n_lin - 1:
                layers.append(dict(elu=nn.ELU, selu=nn.SELU, relu=nn.ReLU, crelu=nn.ReLU, lrelu=nn.LeakyReLU, tanh=nn.Tanh,
                                   sigmoid=nn.Sigmoid)[self.activation]())
        seq = nn.Sequential(*layers)
        seq.load_state_dict({k[len(name) + 1:]: v for k, v in sd.items()})
        return seq


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

    _OBS_FIELDS = ("observations", "privileged_observations", "observation_histories")

    def store_observations(self, transition):
        """The observation third of a transition, at act() time (PPO.act has lt_policy_act write actions, log-prob, mean,
        sigma and value into row t itself)."""
        if self.step >= self.num_transitions_per_env:
            raise AssertionError("Rollout buffer overflow")
        self._launch_store(transition, self._OBS_FIELDS)

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

    def batch_ptrs(self):
        """lt_ppo_batch_ptrs over the flat [T*N, w] views (rollout_storage.py:104-116): a minibatch is a list of row indices."""
        b = _abi.lt_ppo_batch_ptrs()
        for field, buf in (("obs_history", self.observation_histories), ("privileged_obs", self.privileged_observations),
                           ("actions", self.actions), ("values", self.values), ("returns", self.returns),
                           ("advantages", self.advantages), ("old_actions_log_prob", self.actions_log_prob),
                           ("old_mu", self.mu), ("old_sigma", self.sigma)):
            setattr(b, field, buf.data_ptr())
        return b

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


class PPO:
    """ppo.py:33-206.  act() is ONE fused library call that writes the sample, its log-prob, mean, sigma and the value straight
    into row t of the rollout buffers; update() is, per minibatch, lt_ppo_minibatch_grad -> [cross-rank exchange] ->
    lt_ppo_minibatch_apply -> lt_adaptation_grad -> [exchange] -> lt_adaptation_apply, with the learning rate, the KL decision,
    the clip coefficient and every loss sum staying on the device (one host read at the end of the update; the reference
    synchronises four times per minibatch)."""

    def __init__(self, actor_critic, device='cuda:0'):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _abi.LtError("PPO runs as sm_100a kernels only (liblt_b200.so): there is no CPU path")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.lib = _abi.load()
        self.actor_critic = actor_critic.to(self.device)
        n = self.actor_critic.num_params
        if ltd.world_size() > 1:           # replicas start from rank 0's weights whatever each rank's seed was
            ltd.broadcast_(self.actor_critic.flat, src=0)
            self.actor_critic.parameters_changed()
        self.exchange = ltd.GradExchange(n, self.device, stats=_abi.DEFINES["LT_PPO_STATS"])
        self.actor_critic.bind_grads(self.exchange.grads)
        z = lambda: torch.zeros(n, dtype=torch.float32, device=self.device)
        self.adam_m, self.adam_v = z(), z()                  # torch.optim.Adam state of self.optimizer (ppo.py:44)
        # ppo.py:45-46 builds the second Adam over ALL parameters, but only the adaptation module ever holds a gradient when it
        # steps (zero_grad() leaves the others None, which Adam skips): moments over that slice are the same update.
        self.ad_m, self.ad_v = z(), z()
        self.adam_step, self.ad_step = 0, 0
        self.learning_rate = PPO_Args.learning_rate
        self.lr_dev = torch.tensor([self.learning_rate], dtype=torch.float32, device=self.device)
        self.storage = None
        self.transition = RolloutStorage.Transition()
        self.last_update_trace = None
        self.sample_seed = (torch.initial_seed() + 0x9E3779B97F4A7C15 * (ltd.rank() + 1)) & (2 ** 63 - 1)
        self.act_calls = 0

    def init_storage(self, num_envs, num_transitions_per_env, actor_obs_shape, privileged_obs_shape, obs_history_shape, action_shape):
        self.storage = RolloutStorage(num_envs, num_transitions_per_env, actor_obs_shape, privileged_obs_shape,
                                      obs_history_shape, action_shape, self.device)
        mb = num_envs * num_transitions_per_env // PPO_Args.num_mini_batches
        self.actor_critic.reserve(max(int(num_envs), int(mb)), training=True)

    def test_mode(self):
        self.actor_critic.eval()

    def train_mode(self):
        self.actor_critic.train()

    def act(self, obs, privileged_obs, obs_history, actions=None):
        """ppo.py:65-78.  `actions` (tests only) replaces the sample so that a recorded rollout can be replayed."""
        st, tr = self.storage, self.transition
        t = st.step
        if t >= st.num_transitions_per_env:
            raise AssertionError("Rollout buffer overflow")
        if obs_history.shape[0] != st.num_envs:
            raise ValueError(f"act() got {obs_history.shape[0]} rows, the storage holds {st.num_envs} envs")
        tr.actions, tr.values, tr.actions_log_prob = st.actions[t], st.values[t], st.actions_log_prob[t]
        tr.action_mean, tr.action_sigma = st.mu[t], st.sigma[t]
        self.act_calls += 1
        self.actor_critic.act_into(obs_history, privileged_obs,
                                   dict(actions=tr.actions, actions_log_prob=tr.actions_log_prob, action_mean=tr.action_mean,
                                        action_sigma=tr.action_sigma, value=tr.values),
                                   actions_in=actions, seed=self.sample_seed, step=self.act_calls)
        tr.observations = obs
        tr.critic_observations = obs
        tr.privileged_observations = privileged_obs
        tr.observation_histories = obs_history
        st.store_observations(tr)                 # captured now: env.step() rewrites its buffers in place
        return tr.actions

    def process_env_step(self, rewards, dones, infos):
        """ppo.py:80-93.  The time-out bootstrap rewards += gamma * values * time_outs happens inside lt_rollout_store."""
        self.transition.rewards = rewards
        self.transition.dones = dones
        self.transition.time_outs = infos['time_outs'].to(self.device) if 'time_outs' in infos else None
        self.storage.store_env_step(self.transition)
        self.transition.clear()
        self.actor_critic.reset(dones)

    def compute_returns(self, last_critic_obs, last_critic_privileged_obs):
        last_values = self.actor_critic.evaluate(last_critic_obs, last_critic_privileged_obs)
        self.storage.compute_returns(last_values, PPO_Args.gamma, PPO_Args.lam)

    @staticmethod
    def hparams(A=PPO_Args):
        hp = _abi.lt_ppo_hparams()
        hp.clip_param, hp.value_loss_coef, hp.entropy_coef = A.clip_param, A.value_loss_coef, A.entropy_coef
        hp.max_grad_norm = A.max_grad_norm
        hp.adaptive_schedule = 1 if (A.desired_kl is not None and A.schedule == 'adaptive') else 0
        hp.desired_kl = A.desired_kl if A.desired_kl is not None else 0.0
        hp.adam_beta1, hp.adam_beta2, hp.adam_eps = 0.9, 0.999, 1e-8          # torch.optim.Adam defaults (ppo.py:44-46)
        hp.lr_min, hp.lr_max, hp.lr_factor = 1e-5, 1e-2, 1.5                   # ppo.py:118-121
        hp.use_clipped_value_loss = 1 if A.use_clipped_value_loss else 0
        return hp

    def update(self, indices=None):
        """ppo.py:95-206.  `indices` (tests only) injects the minibatch permutation (rollout_storage.py:102 draws ONE
        torch.randperm per update and reuses it for every epoch)."""
        A = PPO_Args
        st, ac, ex, lib = self.storage, self.actor_critic, self.exchange, self.lib
        batch_size = st.num_envs * st.num_transitions_per_env
        mb = batch_size // A.num_mini_batches
        if indices is None:
            indices = torch.randperm(A.num_mini_batches * mb, device=self.device)
        indices = indices.to(device=self.device, dtype=torch.int64).contiguous()
        batch = st.batch_ptrs()
        hp = self.hparams(A)
        world = ltd.world_size()
        ac._ready(mb, training=True)
        h, s = ac._h, _stream(self.device)
        n_sub = int(A.num_adaptation_module_substeps)
        num_updates = A.num_learning_epochs * A.num_mini_batches
        ad_lo, ad_hi = ac.body_slices["adaptation_module"]
        # per-minibatch records, all on the device: [k] = stats[0..4) | lr | grad norm, and the adaptation sums per substep
        rec = torch.zeros(num_updates, 6, dtype=torch.float32, device=self.device)
        rec_ad = torch.zeros(num_updates, max(n_sub, 1), 4, dtype=torch.float32, device=self.device)
        num_train = int(mb // 5 * 4)                                            # ppo.py:160-161
        self.lr_dev.fill_(self.learning_rate)
        k = 0
        for _epoch in range(A.num_learning_epochs):
            for i in range(A.num_mini_batches):
                idx = C.c_void_p(indices.data_ptr() + 8 * i * mb)
                _abi.check_policy(lib.lt_ppo_minibatch_grad(h, C.byref(batch), idx, mb, C.byref(hp), world, _ptr(ex.stats), s), h,
                                  "lt_ppo_minibatch_grad")
                ex.exchange_main()
                self.adam_step += 1
                _abi.check_policy(lib.lt_ppo_minibatch_apply(h, C.byref(hp), _ptr(ex.stats), _ptr(self.lr_dev), _ptr(self.adam_m),
                                                             _ptr(self.adam_v), self.adam_step, C.c_void_p(rec[k, 5:].data_ptr()), s), h,
                                  "lt_ppo_minibatch_apply")
                rec[k, :4].copy_(ex.stats[:4])
                rec[k, 4:5].copy_(self.lr_dev)
                for sub in range(n_sub):
                    _abi.check_policy(lib.lt_adaptation_grad(h, C.byref(batch), idx, mb, num_train, world, _ptr(ex.stats), s), h,
                                      "lt_adaptation_grad")
                    ex.exchange_adaptation(ad_hi)
                    self.ad_step += 1
                    _abi.check_policy(lib.lt_adaptation_apply(h, C.byref(hp), C.c_float(A.adaptation_module_learning_rate), _ptr(self.ad_m),
                                                              _ptr(self.ad_v), self.ad_step, s), h, "lt_adaptation_apply")
                    rec_ad[k, sub].copy_(ex.stats[4:8])
                k += 1
        ac._dirty = False                          # the apply calls refreshed the operand planes themselves
        rec_h, ad_h = rec.double().cpu(), rec_ad.double().cpu()                 # the one host read of the update
        surr, vl, rows = rec_h[:, 0], rec_h[:, 1], rec_h[:, 3]
        self.learning_rate = float(rec_h[-1, 4]) if num_updates else self.learning_rate
        self.last_update_trace = dict(grad_norms=rec_h[:, 5].tolist(), lrs=[float(x) for x in rec_h[:, 4].float().tolist()],
                                      kl_means=(rec_h[:, 2] / rec_h[:, 3]).tolist())
        mean_value_loss = float((vl / rows).sum()) / num_updates
        mean_surrogate_loss = float((surr / rows).sum()) / num_updates
        n_ad = max(num_updates * n_sub, 1)
        ad_train = float((ad_h[:, :n_sub, 0] / ad_h[:, :n_sub, 1].clamp(min=1)).sum()) / n_ad
        ad_test = float((ad_h[:, :n_sub, 2] / ad_h[:, :n_sub, 3].clamp(min=1)).sum()) / n_ad
        st.clear()
        return mean_value_loss, mean_surrogate_loss, ad_train, 0.0, 0.0, ad_test, 0.0, 0.0


class Runner:
    """ppo_cse/__init__.py:65-357.  wandb / video logging of the reference is replaced by an optional `log_fn(dict, step)`
    callback; checkpoints (ac_weights.pt + the two TorchScript modules) go to `save_dir`."""

    ACTOR_CRITIC = ActorCritic             # ppo_cnn.Runner swaps in the height-map-encoder variant

    def __init__(self, env, device='cuda:0', runner_args=RunnerArgs, ac_args=AC_Args, log_fn=None, save_dir="last_run/checkpoints"):
        self.device = device
        self.env = env
        self.runner_args = runner_args
        self.log_fn = log_fn
        self.save_dir = save_dir
        actor_critic = self.ACTOR_CRITIC(self.env.num_obs, self.env.num_privileged_obs, self.env.num_obs_history,
                                         self.env.num_actions).to(self.device)
        if runner_args.resume:
            actor_critic.load_state_dict(torch.load(runner_args.resume, map_location=self.device))
        self.alg = PPO(actor_critic, device=self.device)         # broadcasts rank 0's weights when env-sharded
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
        for name, body in (("adaptation_module_latest.jit", "adaptation_module"), ("body_latest.jit", "actor_body")):
            torch.jit.script(self.alg.actor_critic.export_module(body)).save(os.path.join(save_path, name))

    def get_inference_policy(self, device=None):
        """-> callable(obs_dict) on the training GPU.  For another device (deployment) load the TorchScript files save() writes."""
        self._same_device(device)
        return self.alg.actor_critic.act_inference

    def get_expert_policy(self, device=None):
        self._same_device(device)
        return self.alg.actor_critic.act_expert

    def _same_device(self, device):
        if device is not None and torch.device(device).type != "cuda":
            raise _abi.LtError("the policy runs as sm_100a kernels: for CPU / robot deployment load adaptation_module_latest.jit "
                               "and body_latest.jit written by Runner.save()")
