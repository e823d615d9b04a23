"""The reference's environment API over the CUDA step path.

  LeggedRobot             go1_gym/envs/base/legged_robot_trajectory_tracking.py (LR) :19-169 + base_task.py:14-103
  TrajectoryTrackingEnv   go1_gym/envs/go1/trajectory_tracking/__init__.py:11-56
  HistoryWrapper          go1_gym/envs/wrappers/history_wrapper.py:6-41
  reward containers       go1_gym/envs/rewards/reward_crawling.py, trajectory_tracking_reward.py (`_reward_<name>` protocol,
                          LR:1368-1397)

Same constructor arguments, attributes and return values, so go1_gym_learn's Runner (or ppo.Runner here) drives it
unchanged.  What differs is underneath: the per-env work of step() is `decimation` lt_compute_torques launches and one
lt_post_physics_step (StepEngine), and PhysX is a `SimBackend` object -- the three places LeggedRobot talks to the simulator
(LR:84-88 actuate+simulate+refresh dof state, LR:119-121 refresh root/contact/rigid-body state, LR:1010-1013/1049-1052 push
reset states).  `ReplaySim` replays seeded synthetic or recorded Isaac Gym tensors; a live Isaac Gym binds the same tensors
(INTEGRATION.md).  No CPU path: the engine raises without a CUDA device.
"""
import inspect
from collections import defaultdict, deque

import numpy as np
import torch

from . import cfg as ltcfg
from . import flatcfg, workload
from .engine import StepEngine
from .synthetic import SyntheticState


# ------------------------------------------------------------------------------------------------ simulator back-ends
class SimBackend:
    """What the step path needs from the simulator.  All tensors are the env's (Isaac Gym layouts, on the device)."""

    def attach(self, env):
        self.env = env

    def simulate(self, substep):
        """gym.set_dof_actuation_force_tensor(env.torques); gym.simulate(); gym.refresh_dof_state_tensor() (LR:84-88):
        must leave the new dof_state in env.dof_state."""
        raise NotImplementedError

    def refresh(self):
        """refresh_actor_root_state / net_contact_force / rigid_body_state tensors (LR:119-121)."""
        raise NotImplementedError

    def push_resets(self, env_ids):
        """set_actor_root_state_tensor_indexed / set_dof_state_tensor_indexed for the envs the step kernel re-spawned
        (it already wrote their rows in env.root_states / env.dof_state; LR:1010-1013, 1049-1052)."""


class ReplaySim(SimBackend):
    """Seeded synthetic 'PhysX output' (synthetic.SyntheticState), or recorded per-step arrays:
    frames[k] = dict(dof_state=[decimation, N*12, 2], root=[N,13], contact=[N,B,3], rigid=[N*B,13])."""

    def __init__(self, seed=1000, frames=None, device_frames=0):
        """device_frames > 0: that many synthetic frames are generated once, kept in HBM and cycled with device-to-device
        copies (what a GPU simulator hands over), instead of a fresh host-generated frame per step."""
        self.seed = seed
        self.frames = frames
        self.device_frames = int(device_frames)
        self._dev = None
        self.k = 0

    def attach(self, env):
        super().attach(env)
        if self.frames is None:
            self.gen = SyntheticState(env.num_envs, env.env_origins.cpu().numpy(), num_bodies=env.num_bodies, seed=self.seed,
                                      feet_indices=env.feet_indices.tolist())
        self._root = None
        if self.frames is None and self.device_frames > 0:
            d, n = env.device, env.num_envs
            self._dev = []
            for k in range(1, self.device_frames + 1):
                root = self.gen.root(k)
                self._dev.append(dict(dof=[self.gen.dof_state(k, s).reshape(env.dof_state.shape).to(d) for s in range(env.flat["decimation"])],
                                      root=root.reshape(n, 13).to(d), contact=self.gen.contact(k).reshape(env.contact_forces.shape).to(d),
                                      rigid=self.gen.rigid(k, root).reshape(env.rigid_body_state.shape).to(d)))

    def simulate(self, substep):
        e = self.env
        if self._dev is not None:
            e.dof_state.copy_(self._dev[self.k % len(self._dev)]["dof"][substep])
            return
        if self.frames is not None:
            src = torch.as_tensor(np.ascontiguousarray(self.frames[self.k]["dof_state"][substep], np.float32))
        else:
            src = self.gen.dof_state(self.k + 1, substep)
        e.dof_state.copy_(src.reshape(e.dof_state.shape), non_blocking=True)

    def refresh(self):
        e = self.env
        if self._dev is not None:
            fr = self._dev[self.k % len(self._dev)]
            e.root_states.view(e.num_envs, -1, 13)[:, 0, :].copy_(fr["root"])
            e.contact_forces.copy_(fr["contact"])
            e.rigid_body_state.copy_(fr["rigid"])
            self.k += 1
            return
        if self.frames is not None:
            fr = self.frames[self.k]
            root, contact, rigid = (torch.as_tensor(np.ascontiguousarray(fr[k], np.float32)) for k in ("root", "contact", "rigid"))
        else:
            root = self.gen.root(self.k + 1)
            contact = self.gen.contact(self.k + 1)
            rigid = self.gen.rigid(self.k + 1, root)
        e.root_states.view(e.num_envs, -1, 13)[:, 0, :].copy_(root.reshape(e.num_envs, 13))
        e.contact_forces.copy_(contact.reshape(e.contact_forces.shape))
        e.rigid_body_state.copy_(rigid.reshape(e.rigid_body_state.shape))
        self.k += 1


# ------------------------------------------------------------------------------------------------ reward containers
class _FusedContainer:
    """Marker base of the shipped containers: their `_reward_<name>` terms run inside lt_step_kernel.  A subclass that adds
    or overrides a `_reward_<name>(self) -> Tensor[N]` method gets that term evaluated in torch on the device and summed
    through LT_TERM_EXTERNAL (same scale / bucket / episode-sum treatment as a fused term)."""
    container_name = None

    def __init__(self, env):
        self.env = env

    @classmethod
    def fused_names(cls):
        return set(ltcfg.fused_reward_names(cls.container_name))


class RewardsCrawling(_FusedContainer):
    container_name = "RewardsCrawling"


class TrajectoryTrackingRewards(_FusedContainer):
    container_name = "TrajectoryTrackingRewards"


REWARD_CONTAINERS = {"RewardsCrawling": RewardsCrawling, "TrajectoryTrackingRewards": TrajectoryTrackingRewards}


def external_terms(container_cls, cfg):
    """Names (nonzero scale) whose `_reward_<name>` is a Python method of `container_cls` rather than a fused term."""
    out = []
    fused = container_cls.fused_names() if issubclass(container_cls, _FusedContainer) else set()
    for name, scale in vars(cfg.reward_scales).items():
        if scale == 0 or name == "termination":
            continue
        fn = getattr(container_cls, "_reward_" + name, None)
        if fn is not None and callable(fn):
            out.append(name)                     # user code wins, also over a fused term of the same name
        elif name not in fused:
            pass                                 # LR:1392-1394: nonzero scale without a method -> warning, skipped
    return out


class LazyExtras(dict):
    """extras dict whose heavy entries (host copies of state, trajectory_tracking/__init__.py:28-42) are produced on
    first access instead of on every step (the reference does 9 device->host copies per step for them)."""

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self._lazy = {}

    def set_lazy(self, key, fn):
        self._lazy[key] = fn
        super().pop(key, None)

    def __missing__(self, key):
        if key in self._lazy:
            v = self._lazy[key]()
            self[key] = v
            return v
        raise KeyError(key)

    def __contains__(self, key):
        return super().__contains__(key) or key in self._lazy

    def keys(self):
        return list(super().keys()) + [k for k in self._lazy if not super().__contains__(k)]

    def get(self, key, default=None):
        try:
            return self[key]
        except KeyError:
            return default


# ------------------------------------------------------------------------------------------------ the env
class LeggedRobot:
    """LR:19-169.  `eval_cfg` must be None (every shipped train script runs without eval envs)."""

    def __init__(self, cfg, sim_params=None, physics_engine="SIM_PHYSX", sim_device="cuda:0", headless=True, eval_cfg=None,
                 initial_dynamics_dict=None, sim_backend=None, terrain=None, reward_container=None, actuator=None, seed=0,
                 robot_model=ltcfg.Go1Model, log_episodes=True, alias_outputs=False):
        if eval_cfg is not None:
            raise NotImplementedError("eval_cfg: only the train half of the train/eval split is implemented")
        self.cfg, self.eval_cfg = cfg, None
        self.sim_params, self.physics_engine, self.sim_device, self.headless = sim_params, physics_engine, sim_device, headless
        self.device = torch.device(sim_device)
        self.initial_dynamics_dict = initial_dynamics_dict
        self.log_episodes = log_episodes
        # The reference builds NEW obs / privileged_obs / obs_history tensors every step (torch.cat), and its PPO keeps
        # references to them across env.step().  The kernels rewrite persistent buffers in place, so by default step()
        # hands out clones; alias_outputs=True returns the buffers themselves (ppo.PPO here captures them at act() time).
        self.alias_outputs = alias_outputs
        # ---- reward container / plugin discovery (LR:1373-1397)
        if reward_container is None:
            reward_container = REWARD_CONTAINERS.get(cfg.rewards.reward_container_name)
            if reward_container is None:
                raise NotImplementedError(f"reward container {cfg.rewards.reward_container_name!r}")
        self._external_names = external_terms(reward_container, cfg)
        sim_dt = getattr(sim_params, "dt", None)
        self.flat = ltcfg.flat_from_cfg(cfg, robot_model, sim_dt=sim_dt, extra_reward_names=tuple(self._external_names))
        flatcfg.check_supported(self.flat)
        f = self.flat
        # ---- BaseTask attributes (base_task.py:37-66)
        self.num_envs = self.num_train_envs = f["num_envs"]
        self.num_eval_envs = 0
        self.num_obs, self.num_privileged_obs, self.num_actions = f["num_obs"], f["num_privileged_obs"], f["num_actions"]
        self.num_dof = self.num_dofs = self.num_actuated_dof = f["num_dof"]
        self.num_actor, self.num_bodies = f["num_actors"], f["num_bodies"]
        self.dt, self.max_episode_length = f["dt"], f["max_episode_length"]
        cfg.env.max_episode_length = self.max_episode_length                      # LR:1869
        self.obs_scales, self.reward_scales = cfg.obs_scales, dict(f["reward_scales"])
        self.reward_names = list(f["reward_names"])
        # ---- terrain: plane grid (LR:1848-1858) or a tile table {tile_heights, env_tile, env_terrain_origin, env_origins}
        if terrain is None:
            if f["mesh_type"] != "plane":
                raise ValueError("a heightfield terrain needs terrain=dict(tile_heights, env_tile, env_terrain_origin, env_origins); "
                                 "terrain.TunnelTerrain(cfg.terrain).table(num_envs) builds one of the single_path shape")
            terrain = {"env_origins": plane_env_origins(self.num_envs, cfg.env.env_spacing)}
        self.engine = StepEngine(f, terrain, actuator if actuator is not None else workload.load_actuator(), device=self.device,
                                 seed=seed, alias_history=False)      # HistoryWrapper zeroes obs_history rows on reset without
        #                                                                      touching obs_buf (history_wrapper.py:33-41): separate buffers
        e = self.engine
        self.root_states, self.dof_state = e.root_states, e.dof_state
        self.contact_forces, self.rigid_body_state = e.contact_forces, e.rigid_body_state
        self.env_origins = e.env_origins
        ids = lambda k: torch.tensor(f[k], dtype=torch.long, device=self.device)   # noqa: E731
        self.feet_indices, self.penalised_contact_indices = ids("feet_indices"), ids("penalised_contact_indices")
        self.termination_contact_indices = ids("termination_contact_indices")
        t = lambda k: torch.tensor(f[k], dtype=torch.float, device=self.device)   # noqa: E731
        self.default_dof_pos = t("default_dof_pos").unsqueeze(0)
        self.torque_limits, self.p_gains, self.d_gains = t("torque_limits"), t("p_gains"), t("d_gains")
        self.dof_pos_limits = torch.stack((t("dof_pos_limits_lower"), t("dof_pos_limits_upper")), dim=1)
        self.noise_scale_vec, self.add_noise = t("noise_scale_vec"), f["add_noise"]
        self.reward_container = reward_container(self)
        self.extras = self._fresh_extras()
        self.common_step_counter = 0
        self.sim = sim_backend if sim_backend is not None else ReplaySim(seed=1000 + seed)
        self.sim.attach(self)
        self.reset_buf.fill_(True)                                               # base_task.py:60

    # ---- attributes with the reference's names: engine state tensors and views of the simulator tensors
    _ALIASES = {"commands": "local_relative_linear"}                              # LR:801-802 (a view in the reference too)

    def __getattr__(self, name):
        eng = self.__dict__.get("engine")
        if eng is not None:
            st = eng.state
            if name in st:
                return st[name]
            if name in LeggedRobot._ALIASES:
                return st[LeggedRobot._ALIASES[name]]
            pl = getattr(eng, "planner", None)              # sampling-based planning state (LR:1212-1214, 1263)
            if pl is not None and name in ("local_target_poses", "plan_length_buf"):
                return getattr(pl, name)
            if pl is not None and name == "measured_heights":
                return eng.measured_heights.view(eng.N, 2, len(eng.flat["measured_points_x"]), len(eng.flat["measured_points_y"]))
        raise AttributeError(name)

    def __setattr__(self, name, value):
        eng = self.__dict__.get("engine")
        if eng is not None and name in eng.state and isinstance(value, torch.Tensor):
            eng.state[name].copy_(value)              # the kernels hold the pointers: assignment writes through
            return
        object.__setattr__(self, name, value)

    @property
    def base_pos(self):
        return self.root_states.view(self.num_envs, -1, 13)[:, 0, 0:3]

    @property
    def base_quat(self):
        return self.root_states.view(self.num_envs, -1, 13)[:, 0, 3:7]

    @property
    def dof_pos(self):
        return self.dof_state.view(self.num_envs, self.num_dof, 2)[..., 0]

    @property
    def dof_vel(self):
        return self.dof_state.view(self.num_envs, self.num_dof, 2)[..., 1]

    @property
    def episode_sums(self):
        return {n: self.engine.state["episode_sums"][r] for r, n in enumerate(self.engine.sum_names)}

    @property
    def lag_buffer(self):
        return list(self.engine.lag_buffer_logical())

    def _fresh_extras(self):
        ex = LazyExtras({"train/episode": defaultdict(lambda: deque([], 4000)), "eval/episode": defaultdict(lambda: deque([], 4000)),
                         "timeouts": deque([], 4000)})
        return ex

    def _out(self, t):
        return t if self.alias_outputs else t.clone()

    def get_observations(self):
        return self._out(self.obs_buf)

    def get_privileged_observations(self):
        return self._out(self.privileged_obs_buf)

    # ---- stepping
    def step(self, actions, inject=None):
        """LR:64-112 -> (obs_buf, privileged_obs_buf, rew_buf, reset_buf, extras).  `inject` (parity tests only): dict with
        any of uniforms / noise_uniforms / gravity_uniform replacing the in-kernel Philox draws of this step."""
        if actions.device != self.device or actions.dtype != torch.float32 or not actions.is_contiguous():
            actions = actions.to(self.device, torch.float32).contiguous()
        eng = self.engine
        for s in range(self.flat["decimation"]):
            eng.compute_torques(actions, s)               # clip + lag + actuator net; writes self.torques (LR:82-83)
            self.sim.simulate(s)                          # LR:84-88
        self.post_physics_step(**(inject or {}))
        return self.obs_buf, self.privileged_obs_buf, self.rew_buf, self.reset_buf, self.extras

    def post_physics_step(self, uniforms=None, noise_uniforms=None, gravity_uniform=None):
        """LR:114-169."""
        self.sim.refresh()
        self.common_step_counter += 1
        extra = self._evaluate_plugins() if self._external_names else None
        self.engine.post_physics_step(uniforms=uniforms, noise_uniforms=noise_uniforms, extra_reward=extra,
                                      gravity_uniform=gravity_uniform)
        self._fill_extras()
        self.sim.push_resets(self.engine.state["reset_env_ids"])

    def _evaluate_plugins(self):
        """User `_reward_<name>` methods (LR:1393-1397), evaluated in torch on the device BEFORE the fused step so that
        they see this step's simulator state.  Quantities the reference refreshes at LR:124-134 are refreshed here the
        same way for the plugin's benefit; goal-relative quantities (relative_linear, ...) still hold the previous
        step's values at this point (documented deviation)."""
        q = self.base_quat
        st = self.engine.state
        root = self.root_states.view(self.num_envs, -1, 13)[:, 0]
        st["base_lin_vel"].copy_(quat_rotate_inverse(q, root[:, 7:10]))
        st["base_ang_vel"].copy_(quat_rotate_inverse(q, root[:, 10:13]))
        st["projected_gravity"].copy_(quat_rotate_inverse(q, st["gravity_vec"].expand(self.num_envs, 3)))
        st["episode_length_buf"] += 1                    # LR:124 (the kernel adds it again: undo below)
        rows = [getattr(self.reward_container, "_reward_" + n)().to(torch.float32).reshape(self.num_envs)
                for n in self._external_names]
        st["episode_length_buf"] -= 1
        return torch.stack(rows).contiguous()

    def _fill_extras(self):
        """extras of reset_idx (LR:256-291) from the finished-episode record the step kernel leaves on the device."""
        ex, st = self.extras, self.engine.state
        if self.flat["send_timeouts"]:
            ex["time_outs"] = st["time_outs_latched"][:self.num_train_envs]
        if not self.log_episodes:
            return
        n = int(st["reset_count"].item())                 # one host read per step (the reference's nonzero() syncs too)
        if n == 0:
            return
        ids = st["reset_env_ids"][:n]
        sums = st["finished_sums"][:, ids].cpu().numpy()
        ep = ex["train/episode"]
        for r, key in enumerate(self.engine.sum_names):
            ep["rew_" + key].extend(sums[r])
        ep["episode_length"].extend(st["finished_episode_length"][ids].cpu().numpy())
        ep["reached"].extend(st["reached_env_buf"][ids].cpu().numpy())
        ep["goal_distance"].extend(st["finished_goal_distance"][ids].cpu().numpy())
        for key in ("exploration_lin", "exploration_yaw"):                     # update_curriculum logging, LR:176,182
            if key in self.engine.curriculum.scales:
                ep[key] = self.engine.curriculum.scales[key]
        if self.flat["send_timeouts"]:
            ex["timeouts"].extend(st["time_out_buf"][:self.num_train_envs])

    def reset_idx(self, env_ids):
        """LR:218-296 for an explicit id list (outside a step)."""
        env_ids = torch.as_tensor(env_ids, device=self.device, dtype=torch.int64)
        if env_ids.numel() == 0:
            return
        self.engine.reset_idx(env_ids)
        self.sim.push_resets(env_ids)

    def reset(self):
        """base_task.py:97-103."""
        self.reset_idx(torch.arange(self.num_envs, device=self.device))
        obs, privileged_obs, _, _, _ = self.step(torch.zeros(self.num_envs, self.num_actions, device=self.device))
        return obs, privileged_obs


def plane_env_origins(num_envs, spacing):
    """LR:1848-1858."""
    num_cols = np.floor(np.sqrt(num_envs))
    num_rows = np.ceil(num_envs / num_cols)
    xx, yy = torch.meshgrid(torch.arange(num_rows), torch.arange(num_cols), indexing="ij")
    o = torch.zeros(num_envs, 3)
    o[:, 0] = spacing * xx.flatten()[:num_envs]
    o[:, 1] = spacing * yy.flatten()[:num_envs]
    return o.numpy()


def quat_rotate_inverse(q, v):
    """isaacgym.torch_utils.quat_rotate_inverse (host-side torch, only for the plugin slow path)."""
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * (q_vec * v).sum(-1, keepdim=True) * 2.0
    return a - b + c


class TrajectoryTrackingEnv(LeggedRobot):
    """go1_gym/envs/go1/trajectory_tracking/__init__.py:11-56: 4-tuple step, extras with the per-step state dumps."""

    def __init__(self, sim_device="cuda:0", headless=True, num_envs=None, prone=False, deploy=False, cfg=None, eval_cfg=None,
                 initial_dynamics_dict=None, physics_engine="SIM_PHYSX", **kw):
        if cfg is None:
            raise ValueError("cfg is required (a Cfg tree: legged_tracking_b200.cfg.preset_cfg(...) or the reference's own)")
        if num_envs is not None:
            cfg.env.num_envs = num_envs
        sim_params = kw.pop("sim_params", None)
        super().__init__(cfg, sim_params, physics_engine, sim_device, headless, eval_cfg, initial_dynamics_dict, **kw)

    def step(self, actions, inject=None):
        super().step(actions, inject)
        ex = self.extras
        ex["privileged_obs"] = self._out(self.privileged_obs_buf)
        npy = lambda t: t.detach().cpu().numpy()   # noqa: E731
        feet = self.feet_indices
        ex.set_lazy("joint_pos", lambda: npy(self.dof_pos))
        ex.set_lazy("joint_vel", lambda: npy(self.dof_vel))
        ex.set_lazy("joint_pos_target", lambda: npy(self.joint_pos_target))
        ex.set_lazy("joint_vel_target", lambda: torch.zeros(12))
        ex.set_lazy("body_linear_vel", lambda: npy(self.base_lin_vel))
        ex.set_lazy("body_angular_vel", lambda: npy(self.base_ang_vel))
        ex.set_lazy("body_linear_vel_cmd", lambda: npy(self.commands)[:, 0:2])
        ex.set_lazy("body_angular_vel_cmd", lambda: npy(self.commands)[:, 2:])
        ex.set_lazy("contact_states", lambda: npy(self.contact_forces[:, feet, 2] > 1.).copy())
        ex.set_lazy("foot_positions", lambda: npy(self.rigid_body_state.view(self.num_envs, -1, 13)[:, feet, 0:3]).copy())
        ex.set_lazy("body_pos", lambda: npy(self.root_states[:, 0:3]))
        ex.set_lazy("torques", lambda: npy(self.torques))
        return self._out(self.obs_buf), self.rew_buf, self.reset_buf, ex

    def reset(self):
        self.reset_idx(torch.arange(self.num_envs, device=self.device))
        self.episode_length_buf = torch.randint(int(self.max_episode_length), (self.num_envs,), device=self.device, dtype=torch.long)
        self.extras = self._fresh_extras()
        obs, _, _, _ = self.step(torch.zeros(self.num_envs, self.num_actions, device=self.device, requires_grad=False))
        return obs


class HistoryWrapper:
    """go1_gym/envs/wrappers/history_wrapper.py:6-41.  The observation history is shifted by lt_obs_kernel in the same
    pass that writes obs_buf; this class only presents it (and keeps get_observations()'s shifting side effect)."""

    def __init__(self, env):
        self.env = env
        self.obs_history_length = env.cfg.env.num_observation_history
        self.num_obs_history = self.obs_history_length * env.num_obs
        self.obs_history = env.engine.state["obs_history"]
        self.num_privileged_obs = env.num_privileged_obs

    def __getattr__(self, name):                     # gym.Wrapper attribute forwarding
        return getattr(self.__dict__["env"], name)

    def step(self, action, inject=None):
        obs, rew, done, info = self.env.step(action, inject) if inject is not None else self.env.step(action)
        return {'obs': obs, 'privileged_obs': info["privileged_obs"], 'obs_history': self.env._out(self.obs_history)}, rew, done, info

    def get_observations(self):
        obs = self.env.get_observations()
        privileged_obs = self.env.get_privileged_observations()
        n = self.env.num_obs
        if self.obs_history_length > 1:
            self.obs_history[:, :-n] = self.obs_history[:, n:].clone()
        self.obs_history[:, -n:] = self.env.obs_buf
        return {'obs': obs, 'privileged_obs': privileged_obs, 'obs_history': self.env._out(self.obs_history)}

    def reset_idx(self, env_ids):
        ret = self.env.reset_idx(env_ids)
        self.obs_history[env_ids, :] = 0
        return ret

    def reset(self):
        ret = self.env.reset()
        privileged_obs = self.env.get_privileged_observations()
        self.obs_history[:, :] = 0
        return {"obs": ret, "privileged_obs": privileged_obs, "obs_history": self.env._out(self.obs_history)}
