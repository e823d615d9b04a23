"""BASELINE config 4: the post-physics step of the velocity-tracking env (walk-these-ways; go1_gym/envs/base/
legged_robot_velocity_tracking.py, LRV below; scripts/train_velocity_tracking.py) as a chain of liblt_b200 launches.

  step(actions, sim)   LRV:60-105 (LeggedRobot.step) + :108-154 (post_physics_step) + history_wrapper.py:19-24
      lt_compute_torques x decimation   the same kernel and state as the trajectory env (_compute_torques is identical in both files)
      lt_vel_pre                        episode counter, body-frame velocities, foot states, check_termination, interval masks
      lt_resample_commands              for the envs whose resampling interval elapsed (commands drawn by the caller's source)
      lt_step_contact_targets           gait clock
      lt_corl_rewards + lt_vel_reward   the CoRLRewards terms and compute_reward
      reset_idx                         indexed torch writes on the device for reset_buf.nonzero() -- the rare path, host-orchestrated
                                        exactly like the reference's (it needs the host curricula anyway)
      lt_vel_observe                    70-column observation + noise, last_* shifts, obs clip, the 30-row history shift

What stays on the host, as in the reference: the command curricula (`command_source`: anything with update() / sample(), e.g.
CurriculumAdapter around the reference's own RewardThresholdCurriculum objects), the gravity-randomisation schedule, the choice of
env ids for the three indexed paths (one nonzero() each, a host sync the reference has too).

Scope: the configuration scripts/train_velocity_tracking.py sets (no height observations, push_robots off, no rigid-body
re-randomisation, flat origins); other switches raise.  There is no CPU path.
"""
import ctypes as C
import json
import math
import os

import numpy as np
import torch

from . import _abi
from .gait import CommandResampler, CoRLRewardKernel, GaitClock

DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
RAW_SUMS = ["lin_vel_raw", "ang_vel_raw", "lin_vel_residual", "ang_vel_residual", "ep_timesteps"]       # LRV:312-316
CURRICULUM_KEYS = ["tracking_lin_vel", "tracking_ang_vel", "tracking_contacts_shaped_force", "tracking_contacts_shaped_vel"]   # LRV:746-747
LOCAL_RANGE = [0.55, 0.55, 0.55, 0.55, 0.35, 0.25, 0.25, 0.25, 0.25, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0]       # LRV:755-757


# command ranges of scripts/train_velocity_tracking.py:158-171 (x, y, yaw velocity, body height, gait frequency, phase, offset, bound,
# duration, footswing height, pitch, roll, stance width, stance length, aux) -- configuration values for UniformCommandSource
CFG4_COMMAND_LOW = [-1.0, -0.6, -1.0, -0.25, 2.0, 0.0, 0.0, 0.0, 0.5, 0.03, -0.4, -0.0, 0.10, 0.35, 0.0]
CFG4_COMMAND_HIGH = [1.0, 0.6, 1.0, 0.15, 4.0, 1.0, 1.0, 1.0, 0.5, 0.35, 0.4, 0.0, 0.45, 0.45, 0.01]
CFG4_MEASURED_POINTS_X = [round(-0.8 + 0.1 * i, 1) for i in range(17)]      # legged_robot_velocity_tracking_config.py: 17 x 11 grid
CFG4_MEASURED_POINTS_Y = [round(-0.5 + 0.1 * i, 1) for i in range(11)]


def cfg4_config():
    """What _parse_cfg / _init_buffers / _prepare_reward_function derive from scripts/train_velocity_tracking.py's Cfg (recorded off
    the live reference env by oracle/gen_golden_vel.py; a config table, not code)."""
    return json.load(open(os.path.join(DATA, "vel_cfg4.json")))


def scale_shift(rng):
    return 2.0 / (rng[1] - rng[0]), (rng[1] + rng[0]) / 2.0            # go1_gym/utils/math_utils.py:35-38


def curriculum_tables(limits, init_region):
    """The command grid of go1_gym/envs/base/curriculum.py Curriculum.__init__ (:27-60) and the initial weights of set_to (:16-25,
    LRV:1382-1400): limits [(low, high, n_bins)] per command dimension, init_region [(low, high)].
    -> (grid [n_dims, n_bins_total] bin centres, bin_sizes [n_dims], weights0 [n_bins_total])"""
    axes, sizes = [], []
    for lo, hi, n in limits:
        n = int(n)
        size = (hi - lo) / n
        axes.append(np.linspace(lo + size / 2, hi - size / 2, n))
        sizes.append(size)
    grid = np.stack(np.meshgrid(*axes, indexing="ij")).reshape(len(axes), -1)
    low, high = np.asarray(init_region, np.float64)[:, 0], np.asarray(init_region, np.float64)[:, 1]
    inside = np.logical_and(grid >= low[:, None], grid <= high[:, None]).all(axis=0)
    weights = np.zeros(grid.shape[1])
    weights[inside] = 1.0
    return grid, np.asarray(sizes, np.float64), weights


class UniformCommandSource:
    """A command source without a curriculum: every command uniform inside [low, high] (what curriculum.sample degenerates to when
    every bin has weight 1).  For benchmarks and smoke runs."""

    def __init__(self, low, high, seed=0):
        self.low, self.high = np.asarray(low, np.float64), np.asarray(high, np.float64)
        self.rng = np.random.RandomState(seed)

    def update(self, category, old_bins, task_rewards, success_thresholds):
        pass

    def sample(self, category, batch_size):
        return self.rng.uniform(self.low, self.high, size=(batch_size, len(self.low))), np.zeros(batch_size, np.int64)


class CurriculumAdapter:
    """The reference's curricula (go1_gym/envs/base/curriculum.py objects, one per category) behind the source protocol."""

    def __init__(self, curricula):
        self.curricula = list(curricula)

    def update(self, category, old_bins, task_rewards, success_thresholds):
        self.curricula[category].update(old_bins, task_rewards, success_thresholds, local_range=np.array(LOCAL_RANGE))

    def sample(self, category, batch_size):
        return self.curricula[category].sample(batch_size=batch_size)


class VelocityStepEngine:
    def __init__(self, cfg, num_envs, env_origins, actuator, command_source, device="cuda:0", seed=0, rng=None, heightfield=None):
        """command_source: a host command source (update / sample, see CurriculumAdapter) -- the reference's arrangement, three host
        reads per step -- or the string "device": the RewardThresholdCurriculum itself (weights, success update, bin sampling) and the
        three indexed paths (interval resampling and DR, reset_idx) run as kernels over masks, with no host read in the step.
        heightfield: None (plane: measured heights are zero, LRV:1803-1804) or a dict(height_samples int16 [rows, cols], border_size,
        horizontal_scale, vertical_scale, measured_points_x, measured_points_y) -- the terrain's sample grid as
        cfg.terrain / Terrain.height_field_raw hold it (LRV:1806-1824)."""
        from . import cfg as ltcfg
        from .engine import StepEngine
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _abi.LtError("VelocityStepEngine runs as sm_100a kernels: there is no CPU path")
        self.lib = _abi.load()
        self.cfg, self.N = dict(cfg), int(num_envs)
        c, N, dev = self.cfg, self.N, self.device
        if c["control_type"] != "actuator_net" or not c["randomize_lag_timesteps"]:
            raise NotImplementedError("cfg 4 as scripts/train_velocity_tracking.py sets it: actuator_net with lag randomisation")
        self.source = command_source
        self.device_paths = isinstance(command_source, str) and command_source == "device"
        self.rng = rng or (lambda *shape: torch.rand(*shape, device=dev))
        self.dt = float(c["dt"])
        # ---- torques: the trajectory env's kernel and state (LRV:925-964 == LR:957-996), through a plane-terrain StepEngine
        t = ltcfg.preset_cfg("cfg1_plane_256", num_envs=N)
        flat = ltcfg.flat_from_cfg(t, num_envs=N)
        flat.update(clip_actions=float(c["clip_actions"]), action_scale=float(c["action_scale"]), hip_scale_reduction=float(c["hip_scale_reduction"]),
                    lag_timesteps=int(c["lag_timesteps"]), randomize_lag_timesteps=True, control_type="actuator_net", decimation=int(c["decimation"]),
                    default_dof_pos=[float(x) for x in np.asarray(c["default_dof_pos"]).reshape(-1)],
                    torque_limits=[float(x) for x in c["torque_limits"]])
        self.tq = StepEngine(flat, {"env_origins": np.asarray(env_origins, np.float32)}, actuator, device=dev, seed=seed)
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=dev)                 # noqa: E731
        f32 = lambda a: torch.as_tensor(np.asarray(a, np.float32)).to(dev).contiguous()        # noqa: E731
        B, nc, no, H = int(c["num_bodies"]), int(c["num_commands"]), int(c["num_obs"]), int(c["obs_history_length"])
        self.env_origins = f32(env_origins)
        self.base_init_state = f32(c["base_init_state"])
        self.root_states, self.rigid_body_state, self.contact_forces = z(N, 13), z(N, B, 13), z(N, B, 3)
        self.dof_state = z(N, 12, 2)
        self.gravity_vec = f32([0.0, 0.0, -1.0]).repeat(N, 1).contiguous()
        self.gravities = z(N, 3)
        self.episode_length_buf = z(N, dt=torch.int64)
        self.base_pos, self.base_quat = z(N, 3), z(N, 4)
        self.base_lin_vel, self.base_ang_vel, self.projected_gravity = z(N, 3), z(N, 3), z(N, 3)
        self.foot_positions, self.foot_velocities, self.prev_foot_velocities = z(N, 4, 3), z(N, 4, 3), z(N, 4, 3)
        u8 = lambda: z(N, dt=torch.uint8)                                                       # noqa: E731
        self.reset_buf, self.time_out_buf, self.body_height_buf, self.resample_due, self.dr_due = u8(), u8(), u8(), u8(), u8()
        self.commands = z(N, nc)
        self.commands_scale, self.default_dof_pos = f32(c["commands_scale"]), f32(c["default_dof_pos"]).reshape(-1)
        self.dof_pos_limits, self.noise_scale_vec = f32(c["dof_pos_limits"]), f32(c["noise_scale_vec"])
        self.reward_names = list(c["reward_names"])
        T = len(self.reward_names)
        self.rew_buf, self.rew_buf_pos, self.rew_buf_neg = z(N), z(N), z(N)
        self.episode_sums = z(T + 1, N)                                   # rows: reward_names, "total"
        self.command_sums = z(T + len(RAW_SUMS), N)                       # rows: reward_names, RAW_SUMS
        self.term_sums = z(max(T, 1), dt=torch.float64)
        self.last_actions, self.last_last_actions = z(N, 12), z(N, 12)
        self.last_joint_pos_target, self.last_last_joint_pos_target = z(N, 12), z(N, 12)
        self.last_dof_vel, self.last_root_vel = z(N, 12), z(N, 6)
        self.feet_air_time = z(N, 4)
        self.last_contacts = z(N, 4, dt=torch.bool)
        self.friction_coeffs, self.restitutions = z(N), z(N)
        self.obs_buf, self.privileged_obs_buf = z(N, no), z(N, int(c["num_privileged_obs"]))
        self._hist = [z(N, H * no), z(N, H * no)]
        self._hist_cur = 0
        self.common_step_counter = 0
        self.env_command_bins = np.zeros(N, dtype=np.int64)
        self.env_command_categories = np.zeros(N, dtype=np.int64)
        self.category_names = list(c["category_names"])
        self.extras = {}
        self.gait = GaitClock(N, self.dt, float(c["kappa_gait_probs"]), pacing_offset=bool(c["pacing_offset"]), device=dev)
        self.corl = CoRLRewardKernel(self.reward_names, {k: c["rewards/" + k] for k in ("tracking_sigma", "tracking_sigma_yaw", "base_height_target",
                                                                                          "gait_force_sigma", "gait_vel_sigma", "max_contact_force")},
                                     self.dt, c["feet_indices"], c["penalised_contact_indices"], device=dev)
        self.resampler = CommandResampler(self.commands, self.command_sums, self.category_names,
                                          gaitwise_curricula=c["gaitwise_curricula"], exclusive_phase_offset=c["exclusive_phase_offset"],
                                          balance_gait_distribution=c["balance_gait_distribution"], binary_phases=c["binary_phases"])
        self.reward_scales = {n: float(s) for n, s in zip(self.reward_names, c["reward_scales"])}      # already x dt (LRV:1243)
        self.curriculum_thresholds = dict(tracking_lin_vel=0.8, tracking_ang_vel=0.7, tracking_contacts_shaped_force=0.9,
                                          tracking_contacts_shaped_vel=0.9)     # train_velocity_tracking.py:24-27
        self.sample_interval = int(float(c["resampling_time"]) / self.dt)
        self.step_count = 0
        self.bind_inputs = False                  # step(): use the simulator's tensors in place instead of copying them (see step)
        self.heightfield = None
        if heightfield is not None:
            hf = dict(heightfield)
            xs, ys = np.asarray(hf["measured_points_x"], np.float32), np.asarray(hf["measured_points_y"], np.float32)
            gx, gy = np.meshgrid(xs, ys, indexing="ij")                              # _init_height_points (LRV:1773-1788)
            pts = np.stack([gx.reshape(-1), gy.reshape(-1), np.zeros(gx.size, np.float32)], axis=1)
            hs = torch.as_tensor(np.asarray(hf["height_samples"])).to(device=dev, dtype=torch.int16).contiguous()
            self.heightfield = dict(points=f32(pts), samples=hs, border=float(hf["border_size"]), hscale=float(hf["horizontal_scale"]),
                                    vscale=float(hf["vertical_scale"]))
            self.measured_heights, self.measured_height_mean = z(N, pts.shape[0]), z(N)
        self._vc = self._make_cfg(seed)
        self._keep = None
        if self.device_paths:
            self._init_device_paths()

    # ------------------------------------------------------------------------------------------------ plumbing
    @property
    def obs_history(self):
        return self._hist[self._hist_cur]

    def _make_cfg(self, seed):
        c, v = self.cfg, _abi.lt_vel_cfg()
        v.num_envs, v.num_bodies, v.num_commands = self.N, int(c["num_bodies"]), int(c["num_commands"])
        v.num_obs, v.num_privileged_obs, v.obs_history_length = int(c["num_obs"]), int(c["num_privileged_obs"]), int(c["obs_history_length"])
        for i, b in enumerate(c["feet_indices"]):
            v.feet_indices[i] = int(b)
        term = list(c["termination_contact_indices"])
        v.n_termination = len(term)
        for i, b in enumerate(term):
            v.termination_indices[i] = int(b)
        v.use_terminal_body_height, v.terminal_body_height = int(bool(c["use_terminal_body_height"])), float(c["terminal_body_height"])
        v.max_episode_length = float(c["max_episode_length"])
        v.sample_interval, v.rand_interval = self.sample_interval, int(float(c["rand_interval"]))
        v.decimation, v.lag_slots = int(c["decimation"]), int(self.tq.state["lag_buffer"].shape[0])
        v.n_terms = len(self.reward_names)
        for i, n in enumerate(self.reward_names):
            v.term_scales[i] = self.reward_scales[n]
            v.command_sum_adds_scale[i] = 1 if n in ("tracking_contacts_shaped_force", "tracking_contacts_shaped_vel") else 0
        v.only_positive_rewards = int(bool(c["only_positive_rewards"]))
        v.only_positive_rewards_ji22_style = int(bool(c["only_positive_rewards_ji22_style"]))
        v.sigma_rew_neg = float(c["sigma_rew_neg"])
        for k in ("observe_command", "observe_two_prev_actions", "observe_timing_parameter", "observe_clock_inputs", "add_noise",
                  "priv_observe_friction", "priv_observe_restitution"):
            setattr(v, k, int(bool(c[k])))
        v.clip_observations, v.obs_scale_dof_pos, v.obs_scale_dof_vel = float(c["clip_observations"]), float(c["obs_scale_dof_pos"]), float(c["obs_scale_dof_vel"])
        v.friction_scale, v.friction_shift = scale_shift(c["friction_range"])
        v.restitution_scale, v.restitution_shift = scale_shift(c["restitution_range"])
        v.seed_lo, v.seed_hi = seed & 0xffffffff, (seed >> 32) & 0xffffffff
        return v

    def _ptrs(self):
        p, tq = _abi.lt_vel_ptrs(), self.tq.state
        d = dict(root_states=self.root_states, rigid_body_state=self.rigid_body_state, contact_forces=self.contact_forces, dof_state=self.dof_state,
                 gravity_vec=self.gravity_vec, episode_length_buf=self.episode_length_buf, base_pos=self.base_pos, base_quat=self.base_quat,
                 base_lin_vel=self.base_lin_vel, base_ang_vel=self.base_ang_vel, projected_gravity=self.projected_gravity,
                 foot_positions=self.foot_positions, foot_velocities=self.foot_velocities, prev_foot_velocities=self.prev_foot_velocities,
                 reset_buf=self.reset_buf, time_out_buf=self.time_out_buf, body_height_buf=self.body_height_buf, resample_due=self.resample_due,
                 dr_due=self.dr_due, commands=self.commands, rew_buf=self.rew_buf, rew_buf_pos=self.rew_buf_pos, rew_buf_neg=self.rew_buf_neg,
                 episode_sums=self.episode_sums, command_sums=self.command_sums, term_sums=self.term_sums, commands_scale=self.commands_scale,
                 default_dof_pos=self.default_dof_pos, noise_scale_vec=self.noise_scale_vec, actions=tq["actions"],
                 joint_pos_target=tq["joint_pos_target"], gait_indices=self.gait.gait_indices, clock_inputs=self.gait.clock_inputs,
                 friction_coeffs=self.friction_coeffs, restitutions=self.restitutions, last_actions=self.last_actions,
                 last_last_actions=self.last_last_actions, last_joint_pos_target=self.last_joint_pos_target,
                 last_last_joint_pos_target=self.last_last_joint_pos_target, last_dof_vel=self.last_dof_vel, last_root_vel=self.last_root_vel,
                 obs_buf=self.obs_buf, privileged_obs_buf=self.privileged_obs_buf, obs_history_in=self._hist[self._hist_cur],
                 obs_history_out=self._hist[1 - self._hist_cur], torque_counters=self.tq.counters)
        if self.heightfield is not None:
            d["measured_height_mean"] = self.measured_height_mean
        if self.device_paths:
            d["step_counter"] = self.step_counter_dev
        for k, t in d.items():
            setattr(p, k, t.data_ptr())
        return p

    def _init_device_paths(self):
        c, dev, N = self.cfg, self.device, self.N
        grid, sizes, w0 = curriculum_tables(c["curriculum_limits"], c["curriculum_init"])
        n_cat, nb = len(self.category_names), grid.shape[1]
        f64 = lambda a: torch.as_tensor(np.ascontiguousarray(a, np.float64)).to(dev)           # noqa: E731
        self.cur_grid, self.cur_bin_sizes = f64(grid), f64(sizes)
        self.cur_weights = f64(np.tile(w0, (n_cat, 1)))
        self._hits = torch.zeros(2, n_cat, nb, dtype=torch.int32, device=dev)
        self.env_bins_dev = torch.zeros(N, dtype=torch.int32, device=dev)
        self.env_categories_dev = torch.zeros(N, dtype=torch.int32, device=dev)
        self.episode_acc = torch.zeros(len(self.reward_names) + 2, device=dev)
        self._work = torch.zeros(2, N, dtype=torch.int32, device=dev)
        self._work_count = torch.zeros(3, dtype=torch.int32, device=dev)        # {collecting, published, ticket}: lt_vel_scan_kernel
        self.step_counter_dev = torch.zeros(1, dtype=torch.int32, device=dev)       # Philox step key of the device mode (uint32 bits)
        d = _abi.lt_vel_dev_cfg()
        d.n_bins, d.n_dims = nb, grid.shape[0]
        keys = [n for n in CURRICULUM_KEYS if n in self.reward_names]
        d.n_keys = len(keys)
        for i, n in enumerate(keys):
            d.key_rows[i] = self.reward_names.index(n)
            d.key_thresholds[i] = float(c["curriculum_thresholds"][CURRICULUM_KEYS.index(n)]) * self.reward_scales[n]
        d.ep_len = min(float(c["max_episode_length"]), self.sample_interval)
        for i, v in enumerate(LOCAL_RANGE[:grid.shape[0]]):
            d.local_range[i] = v
        d.randomize_motor_strength, d.randomize_motor_offset = int(bool(c["randomize_motor_strength"])), int(bool(c["randomize_motor_offset"]))
        d.motor_strength_lo, d.motor_strength_span = c["motor_strength_range"][0], c["motor_strength_range"][1] - c["motor_strength_range"][0]
        d.motor_offset_lo, d.motor_offset_span = c["motor_offset_range"][0], c["motor_offset_range"][1] - c["motor_offset_range"][0]
        d.yaw_init_range = float(c["yaw_init_range"])
        d.custom_origins = int(bool(c["custom_origins"]))
        d.x_init_range, d.y_init_range = float(c["x_init_range"]), float(c["y_init_range"])
        d.x_init_offset, d.y_init_offset = float(c["x_init_offset"]), float(c["y_init_offset"])
        for i, v in enumerate(np.asarray(c["base_init_state"], np.float32).reshape(-1)):
            d.base_init_state[i] = float(v)
        d.lag_slots = int(self.tq.state["lag_buffer"].shape[0])
        d.cmd = self.resampler.cfg
        self._dc = d
        q, st = _abi.lt_vel_dev_ptrs(), self.tq.state
        for k, t in dict(grid=self.cur_grid, bin_sizes=self.cur_bin_sizes, weights=self.cur_weights, hit_self=self._hits[0], hit_adj=self._hits[1],
                         env_bins=self.env_bins_dev, env_categories=self.env_categories_dev, commands=self.commands, dof_state_rw=self.dof_state,
                         root_states_rw=self.root_states, env_origins=self.env_origins, motor_strengths=st["motor_strengths"],
                         motor_offsets=st["motor_offsets"], lag_buffer=st["lag_buffer"], gait_indices=self.gait.gait_indices,
                         feet_air_time=self.feet_air_time, episode_acc=self.episode_acc, work_env=self._work[0], work_code=self._work[1],
                         work_count=self._work_count).items():
            setattr(q, k, t.data_ptr())
        self._dq = q

    def _device_paths(self, which, uniform_rows=None):
        u = None
        if uniform_rows is not None:
            u = uniform_rows.to(device=self.device, dtype=torch.float32).contiguous()
            self._keep_u = u
        _abi.check(self.lib.lt_vel_indexed_paths(C.byref(self._vc), C.byref(self._ptrs()), C.byref(self._dc), C.byref(self._dq), which,
                                                 C.c_void_p(u.data_ptr()) if u is not None else None, C.c_uint32(self.step_count & 0x7fffffff),
                                                 self._stream()), None, "lt_vel_indexed_paths")

    # ---- CUDA-graph replay of the device-mode step (an even number of steps per graph: the history buffers ping-pong)
    def graph_replay_ok(self, n_steps=2):
        """No gravity event (LRV:719-723: a host decision baked into a captured graph) within the next n_steps steps."""
        c = self.cfg
        interval, duration = int(float(c["gravity_rand_interval"])), float(c["gravity_rand_duration"])
        return not any((k % interval == 0) or (int(k - duration) % interval == 0)
                       for k in range(self.common_step_counter + 1, self.common_step_counter + n_steps + 1))

    def capture(self, frames):
        """frames: an even number of (actions, sim) pairs whose tensors stay where they are (the simulator writes into them before every
        replay).  -> a CUDA graph of those steps.  Capture after a few eager steps, on steps for which graph_replay_ok() holds."""
        if not self.device_paths or len(frames) % 2 or not self.graph_replay_ok(len(frames)):
            raise _abi.LtError("capture: device mode, an even number of steps, and no gravity event inside them")
        saved = (self.common_step_counter, self.step_count)
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(device=self.device)
        with torch.cuda.graph(g, stream=side):
            for actions, sim in frames:
                self.step(actions, sim)
        self.common_step_counter, self.step_count = saved            # capturing ran the host bookkeeping, not the kernels
        self._graph_steps = len(frames)
        return g

    def replay(self, graph):
        graph.replay()
        self.common_step_counter += self._graph_steps
        self.step_count += self._graph_steps

    def reset_all(self):
        """env.reset() of the device mode: every env through the reset path (first commands, DR, initial state)."""
        self.reset_buf.fill_(1)
        self._device_paths(_abi.DEFINES["LT_VEL_PATH_RESET"])

    def episode_means(self):
        """extras['train/episode'] of the device-side reset path (LRV:200-205): mean of every episode-sum row over the envs reset since
        the last call (one host read, on demand)."""
        acc = self.episode_acc.cpu()
        self.episode_acc.zero_()
        n = max(float(acc[-1]), 1.0)
        return {"rew_" + name: float(acc[i]) / n for i, name in enumerate(self.reward_names + ["total"])}

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ------------------------------------------------------------------------------------------------ the step
    def step(self, actions, sim, noise_uniforms=None, host_paths=True, uniform_rows=None, bind=None):
        """actions [N,12]; sim: what PhysX hands back -- dof_state: one [N*12,2] (or [N,12,2]) tensor per sub-step, root [N,13],
        contact [N,B,3], rigid [N,B,13], all on the device.  -> ({'obs','privileged_obs','obs_history'}, rew, done, extras)
        host_paths=False (measurement only) leaves out the three host-orchestrated indexed paths -- command resampling, interval
        DR, reset_idx -- and with them the three nonzero() host reads: what remains is the launch chain alone.
        uniform_rows (device mode, tests): two [N, LT_VEL_SLOTS] tensors of uniforms for the callback and the reset path.
        bind=True: the simulator's tensors are used in place (a reset then rewrites rows of sim["root"] / the last sim["dof_state"], as
        the reference rewrites Isaac Gym's tensors) instead of being copied into the engine's own."""
        bind = self.bind_inputs if bind is None else bind
        for sub, dof in enumerate(sim["dof_state"]):                                  # LRV:76-82: torques, THEN simulate + refresh
            self.apply_torques(actions, sub)
            if bind:
                self.dof_state = dof.view(self.N, 12, 2)                              # the simulator's tensor itself, as with Isaac Gym
            else:
                self.dof_state.copy_(dof.view(self.N, 12, 2))
        if bind:
            self.root_states, self.contact_forces = sim["root"], sim["contact"]
            self.rigid_body_state = sim["rigid"].view(self.N, -1, 13)
            if self.device_paths:
                self._dq.dof_state_rw, self._dq.root_states_rw = self.dof_state.data_ptr(), self.root_states.data_ptr()
        else:
            self.root_states.copy_(sim["root"])
            self.contact_forces.copy_(sim["contact"])
            self.rigid_body_state.copy_(sim["rigid"].view(self.rigid_body_state.shape))
        return self.post_physics_step(noise_uniforms=noise_uniforms, host_paths=host_paths, uniform_rows=uniform_rows)

    def begin_step(self, actions):
        """Nothing to prepare: the action clip happens inside the first torque launch (LRV:68-69)."""

    def apply_torques(self, actions, sub):
        """_compute_torques (LRV:925-964) on the dof_state the last refresh left (post reset at sub-step 0)."""
        self.tq.bind_sim(dof_state=self.dof_state.view(self.N * 12, 2))
        self.tq.compute_torques(actions, sub)

    def post_physics_step(self, noise_uniforms=None, host_paths=True, uniform_rows=None):
        """LRV:108-154 on the simulator tensors as they stand (root_states, dof_state, contact_forces, rigid_body_state)."""
        tq, lib, s = self.tq, self.lib, self._stream()        # (lt_vel_pre advances the torque handle's sub-step / lag-ring counters)
        self.common_step_counter += 1                                                 # LRV:118
        if self.heightfield is not None:                                              # LRV:708-709 (reads this step's root state only)
            hf = self.heightfield
            _abi.check(lib.lt_vel_measure_heights(C.c_void_p(self.root_states.data_ptr()), C.c_void_p(hf["points"].data_ptr()),
                                                  C.c_void_p(hf["samples"].data_ptr()), self.N, int(hf["points"].shape[0]),
                                                  int(hf["samples"].shape[0]), int(hf["samples"].shape[1]), C.c_float(hf["border"]),
                                                  C.c_float(hf["hscale"]), C.c_float(hf["vscale"]),
                                                  C.c_void_p(self.measured_heights.data_ptr()), C.c_void_p(self.measured_height_mean.data_ptr()), s),
                       None, "lt_vel_measure_heights")
        _abi.check(lib.lt_vel_pre(C.byref(self._vc), C.byref(self._ptrs()), s), None, "lt_vel_pre")
        # ---- _post_physics_step_callback (LRV:693-726)
        if self.device_paths:
            self._device_paths(_abi.DEFINES["LT_VEL_PATH_CALLBACK"], uniform_rows[0] if uniform_rows is not None else None)
        elif host_paths:
            self._resample_commands(self.resample_due.nonzero(as_tuple=False).flatten())
        self.gait.step(self.commands)
        if host_paths and not self.device_paths:
            self._randomize_dof_props(self.dr_due.nonzero(as_tuple=False).flatten())
        self._gravity_schedule()
        # ---- compute_reward (LRV:281-318)
        env = dict(dof_state=self.dof_state, commands=self.commands, base_lin_vel=self.base_lin_vel, base_ang_vel=self.base_ang_vel, projected_gravity=self.projected_gravity,
                   gravity_vec=self.gravity_vec, base_pos=self.base_pos, base_quat=self.base_quat, torques=tq.state["torques"],
                   last_dof_vel=self.last_dof_vel, actions=tq.state["actions"], last_actions=self.last_actions,
                   last_last_actions=self.last_last_actions, joint_pos_target=tq.state["joint_pos_target"],
                   last_joint_pos_target=self.last_joint_pos_target, last_last_joint_pos_target=self.last_last_joint_pos_target,
                   contact_forces=self.contact_forces, foot_positions=self.foot_positions, foot_velocities=self.foot_velocities,
                   prev_foot_velocities=self.prev_foot_velocities, foot_indices=self.gait.foot_indices,
                   desired_contact_states=self.gait.desired_contact_states, default_dof_pos=self.default_dof_pos,
                   dof_pos_limits=self.dof_pos_limits, last_contacts=self.last_contacts)
        terms = self.corl(env)
        _abi.check(lib.lt_vel_reward(C.byref(self._vc), C.byref(self._ptrs()), C.c_void_p(terms.data_ptr()), s), None, "lt_vel_reward")
        # ---- reset_idx (LRV:169-265)
        if self.device_paths:
            self._device_paths(_abi.DEFINES["LT_VEL_PATH_RESET"], uniform_rows[1] if uniform_rows is not None else None)
            self.extras["time_outs"] = self.time_out_buf.view(torch.bool)
        elif host_paths:
            self.reset_idx(self.reset_buf.nonzero(as_tuple=False).flatten())
        # ---- compute_observations + last_* shifts + obs clip + history (LRV:320-470, 143-148, 102-105; history_wrapper.py:23)
        self.step_count += 1
        nz = None
        if noise_uniforms is not None:
            nz = noise_uniforms.to(device=self.device, dtype=torch.float32).contiguous()
        _abi.check(lib.lt_vel_observe(C.byref(self._vc), C.byref(self._ptrs()), C.c_void_p(nz.data_ptr()) if nz is not None else None,
                                      C.c_uint32(self.step_count & 0xffffffff), s), None, "lt_vel_observe")
        self._hist_cur = 1 - self._hist_cur
        self._keep = (nz, terms)
        self.extras["privileged_obs"] = self.privileged_obs_buf
        return ({"obs": self.obs_buf, "privileged_obs": self.privileged_obs_buf, "obs_history": self.obs_history}, self.rew_buf,
                self.reset_buf.bool(), self.extras)

    # ------------------------------------------------------------------------------------------------ host-orchestrated rare paths
    def _resample_commands(self, env_ids):
        """LRV:728-842: curriculum update and sampling through the command source (host), everything else in lt_resample_commands."""
        K = int(env_ids.numel())
        if K == 0:
            return
        c = self.cfg
        ep_len = min(float(c["max_episode_length"]), self.sample_interval)                  # LRV:732-733
        ids_cpu = env_ids.cpu()
        cats_old = self.env_command_categories[ids_cpu.numpy()]
        rows = {n: self.command_sums[self.reward_names.index(n)] for n in CURRICULUM_KEYS if n in self.reward_names}
        for i in range(len(self.category_names)):                                           # LRV:736-757
            sel = ids_cpu[torch.from_numpy(np.atleast_1d(cats_old == i))]
            if sel.numel() == 0:
                continue
            sel_dev = sel.to(self.device)
            task_rewards = [rows[n][sel_dev] / ep_len for n in rows]
            thresholds = [self.curriculum_thresholds[n] * self.reward_scales[n] for n in rows]
            if thresholds:
                self.source.update(i, self.env_command_bins[sel.numpy()], task_rewards, thresholds)
        u = self.rng(K)                                                                     # LRV:760
        cat = self.resampler.categories_from_uniform(u)
        cat_cpu = cat.cpu().numpy()
        nc = int(c["num_commands"])
        new = np.zeros((K, nc), np.float64)
        for i in range(len(self.category_names)):                                           # LRV:768-777
            m = cat_cpu == i
            if not m.any():
                continue
            cmds, bins = self.source.sample(i, int(m.sum()))
            new[m] = np.asarray(cmds)[:, :nc]
            sel = ids_cpu.numpy()[m]
            self.env_command_bins[sel] = bins
            self.env_command_categories[sel] = i
        gait_u = None
        if nc > 5 and not c["gaitwise_curricula"] and (c["exclusive_phase_offset"] or c["balance_gait_distribution"]):
            gait_u = self.rng(K)                                                            # LRV:800, 812
        self.resampler.apply(env_ids, new, category=cat, gait_uniform=gait_u)

    def _randomize_dof_props(self, env_ids):
        """LRV:663-683 (the two switches cfg 4 sets); writes the torque kernel's motor tables."""
        K, c, st = int(env_ids.numel()), self.cfg, self.tq.state
        if K == 0:
            return
        if c["randomize_motor_strength"]:
            lo, hi = c["motor_strength_range"]
            st["motor_strengths"][env_ids, :] = self.rng(K).unsqueeze(1) * (hi - lo) + lo
        if c["randomize_motor_offset"]:
            lo, hi = c["motor_offset_range"]
            st["motor_offsets"][env_ids, :] = self.rng(K, 12) * (hi - lo) + lo

    def _gravity_schedule(self):
        """LRV:719-723 + _randomize_gravity (LRV:565-580)."""
        c = self.cfg
        interval, duration = int(float(c["gravity_rand_interval"])), float(c["gravity_rand_duration"])
        if self.common_step_counter % interval == 0:
            self._randomize_gravity(None)
        if int(self.common_step_counter - duration) % interval == 0:
            self._randomize_gravity(torch.zeros(3, device=self.device))

    def _randomize_gravity(self, external_force):
        c = self.cfg
        if external_force is not None:
            self.gravities[:, :] = external_force.unsqueeze(0)
        elif c["randomize_gravity"]:
            lo, hi = c["gravity_range"]
            self.gravities[:, :] = (self.rng(3) * (hi - lo) + lo).unsqueeze(0)
        g = self.gravities[0, :] + torch.tensor([0.0, 0.0, -9.8], device=self.device)
        self.gravity_vec[:, :] = g.unsqueeze(0) / torch.norm(g)

    def reset_idx(self, env_ids):
        """LRV:169-265."""
        K = int(env_ids.numel())
        if K == 0:
            return
        c, st = self.cfg, self.tq.state
        self._resample_commands(env_ids)
        self._randomize_dof_props(env_ids)
        # _reset_dofs (LRV:966-981)
        self.dof_state[env_ids, :, 0] = self.default_dof_pos.unsqueeze(0) * (self.rng(K, 12) * (1.5 - 0.5) + 0.5)
        self.dof_state[env_ids, :, 1] = 0.0
        # _reset_root_states (LRV:983-1019)
        self.root_states[env_ids] = self.base_init_state
        self.root_states[env_ids, :3] += self.env_origins[env_ids]
        if c["custom_origins"]:                                                              # heightfield / trimesh terrains (LRV:991-1001)
            xr, yr = float(c["x_init_range"]), float(c["y_init_range"])
            self.root_states[env_ids, 0:1] += self.rng(K, 1) * (xr - (-xr)) + (-xr)
            self.root_states[env_ids, 1:2] += self.rng(K, 1) * (yr - (-yr)) + (-yr)
            self.root_states[env_ids, 0] += float(c["x_init_offset"])
            self.root_states[env_ids, 1] += float(c["y_init_offset"])
        yaw_rng = float(c["yaw_init_range"])
        yaw = self.rng(K, 1) * (yaw_rng - (-yaw_rng)) + (-yaw_rng)
        half = yaw / 2
        quat = torch.cat([torch.zeros(K, 2, device=self.device), half.sin(), half.cos()], dim=1)     # quat_from_angle_axis(yaw, z)
        self.root_states[env_ids, 3:7] = quat
        self.root_states[env_ids, 7:13] = self.rng(K, 6) * (0.5 - (-0.5)) + (-0.5)
        # buffers (LRV:192-198)
        self.last_actions[env_ids] = 0.0
        self.last_last_actions[env_ids] = 0.0
        self.last_dof_vel[env_ids] = 0.0
        self.feet_air_time[env_ids] = 0.0
        self.episode_length_buf[env_ids] = 0
        self.reset_buf[env_ids] = 1
        ep = {}
        for i, n in enumerate(self.reward_names + ["total"]):                                # LRV:200-205
            ep["rew_" + n] = torch.mean(self.episode_sums[i][env_ids])
            self.episode_sums[i][env_ids] = 0.0
        self.extras["train/episode"] = ep
        self.extras["time_outs"] = self.time_out_buf.bool()                                  # LRV:255-256
        self.gait.gait_indices[env_ids] = 0.0                                                # LRV:258
        st["lag_buffer"][:, env_ids, :] = 0.0                                                # LRV:260-261

    # ------------------------------------------------------------------------------------------------ state import (tests, resume)
    def load_reference_state(self, st):
        """`init/*` arrays of tests/golden/vel_cfg4_64.npz (the reference env's attribute names)."""
        dev = self.device
        put = lambda dst, src: dst.copy_(torch.as_tensor(np.asarray(src)).to(dst.dtype).reshape(dst.shape).to(dev))      # noqa: E731
        for k in ("episode_length_buf", "commands", "last_actions", "last_last_actions", "last_dof_vel", "last_root_vel", "last_joint_pos_target",
                  "last_last_joint_pos_target", "root_states", "dof_state", "gravity_vec", "gravities", "last_contacts", "foot_velocities",
                  "prev_foot_velocities", "foot_positions", "base_lin_vel", "base_ang_vel", "projected_gravity"):
            if k in st:
                put(getattr(self, k), st[k])
        put(self.gait.gait_indices, st["gait_indices"])
        put(self.friction_coeffs, np.asarray(st["friction_coeffs"])[:, 0])
        put(self.restitutions, np.asarray(st["restitutions"])[:, 0])
        ts = self.tq.state
        for k in ("motor_strengths", "motor_offsets", "joint_pos_err_last", "joint_pos_err_last_last", "joint_vel_last", "joint_vel_last_last",
                  "joint_pos_target", "torques", "actions"):
            put(ts[k], st[k])
        L = ts["lag_buffer"].shape[0]
        for s_ in range(L):
            put(ts["lag_buffer"][s_], st[f"lag_buffer/{s_}"])
        for i, n in enumerate(self.reward_names + ["total"]):
            put(self.episode_sums[i], st["episode_sums/" + n])
        for i, n in enumerate(self.reward_names + RAW_SUMS):
            put(self.command_sums[i], st["command_sums/" + n])
        self.common_step_counter = int(st["common_step_counter"])
        self.env_command_categories = np.asarray(st["env_command_categories"]).astype(np.int64).copy()


# ------------------------------------------------------------------------------------------------ Cfg tree -> the engine's table
CORL_TERMS_SET = None


def config_from_cfg(Cfg, model=None, sim_dt=None):
    """What LeggedRobot._parse_cfg (LRV:1488-1530), _init_buffers (:1124-1300), _get_noise_scale_vec (:1071-1135), _process_dof_props,
    _create_envs and _prepare_reward_function (:1402-1450) derive from a velocity-tracking `Cfg` tree (the reference's own class after
    scripts/train_velocity_tracking.py edited it, or any object with the same attributes) -- the table VelocityStepEngine takes.
    cfg4_config() is this function's output for that script, recorded off the live reference env."""
    from . import cfg as ltcfg
    from .gait import CORL_TERMS
    model = model or ltcfg.Go1Model
    c = Cfg
    f32 = lambda x: float(np.float32(x))                                                   # noqa: E731
    sim_dt = f32(0.005) if sim_dt is None else float(sim_dt)
    dt = c.control.decimation * sim_dt
    bodies = model.body_names
    feet = [bodies.index(s) for s in bodies if c.asset.foot_name in s]
    pen = [bodies.index(s) for n in c.asset.penalize_contacts_on for s in bodies if n in s]
    term = [bodies.index(s) for n in c.asset.terminate_after_contacts_on for s in bodies if n in s]
    f = np.float32
    lim = []
    for a, b in zip(model.dof_lower, model.dof_upper):                                     # soft limits, float32 (LRV _process_dof_props)
        a, b = f(a), f(b)
        m, r = (a + b) / f(2), b - a
        half = f(f(0.5) * r) * f(c.rewards.soft_dof_pos_limit)
        lim.append([float(m - half), float(m + half)])
    default_dof_pos = [f32(c.init_state.default_joint_angles[n]) for n in model.dof_names]
    osc, ns, lvl = c.obs_scales, c.noise_scales, c.noise.noise_level
    nc, na = int(c.commands.num_commands), int(c.env.num_actions)

    def ones(n, *factors):
        v = np.ones(n, np.float32)
        for x in factors:
            v = v * np.float32(x)
        return v
    body = [ones(3, ns.gravity, lvl)]
    if c.env.observe_command:
        body.append(np.zeros(nc, np.float32))
    body += [ones(12, ns.dof_pos, lvl, osc.dof_pos), ones(12, ns.dof_vel, lvl, osc.dof_vel), np.zeros(na, np.float32)]
    if c.env.observe_two_prev_actions:
        body.append(np.zeros(na, np.float32))
    if c.env.observe_timing_parameter:
        body.append(np.zeros(1, np.float32))
    if c.env.observe_clock_inputs:
        body.append(np.zeros(4, np.float32))
    for flag in ("observe_vel", "observe_only_lin_vel", "observe_only_ang_vel", "observe_yaw", "observe_contact_states"):
        if getattr(c.env, flag, False):
            raise NotImplementedError(f"env.{flag}: not among the observation columns of the cfg 4 kernels")
    noise_vec = np.concatenate(body)
    commands_scale = [osc.lin_vel, osc.lin_vel, osc.ang_vel, osc.body_height_cmd, osc.gait_freq_cmd, osc.gait_phase_cmd, osc.gait_phase_cmd,
                      osc.gait_phase_cmd, osc.gait_phase_cmd, osc.footswing_height_cmd, osc.body_pitch_cmd, osc.body_roll_cmd,
                      osc.stance_width_cmd, osc.stance_length_cmd, osc.aux_reward_cmd][:nc]
    names, scales = [], []
    for k, v in vars(c.reward_scales).items():                                             # LRV:1412-1430: non-zero scales x dt, container order
        if k.startswith("_") or v == 0 or k == "termination":
            continue
        if k not in CORL_TERMS:
            continue                                                                       # not a method of the container: skipped like LRV:1425-1427
        names.append(k)
        scales.append(float(v) * dt)
    cc = c.commands
    gri = int(np.ceil(c.domain_rand.gravity_rand_interval_s / dt))
    init = c.init_state
    limits = [(cc.limit_vel_x, cc.num_bins_vel_x), (cc.limit_vel_y, cc.num_bins_vel_y), (cc.limit_vel_yaw, cc.num_bins_vel_yaw),
              (cc.limit_body_height, cc.num_bins_body_height), (cc.limit_gait_frequency, cc.num_bins_gait_frequency),
              (cc.limit_gait_phase, cc.num_bins_gait_phase), (cc.limit_gait_offset, cc.num_bins_gait_offset),
              (cc.limit_gait_bound, cc.num_bins_gait_bound), (cc.limit_gait_duration, cc.num_bins_gait_duration),
              (cc.limit_footswing_height, cc.num_bins_footswing_height), (cc.limit_body_pitch, cc.num_bins_body_pitch),
              (cc.limit_body_roll, cc.num_bins_body_roll), (cc.limit_stance_width, cc.num_bins_stance_width),
              (cc.limit_stance_length, cc.num_bins_stance_length), (cc.limit_aux_reward_coef, cc.num_bins_aux_reward_coef)]
    init_region = [cc.lin_vel_x, cc.lin_vel_y, cc.ang_vel_yaw, cc.body_height_cmd, cc.gait_frequency_cmd_range, cc.gait_phase_cmd_range,
                   cc.gait_offset_cmd_range, cc.gait_bound_cmd_range, cc.gait_duration_cmd_range, cc.footswing_height_range, cc.body_pitch_range,
                   cc.body_roll_range, cc.stance_width_range, cc.stance_length_range, cc.aux_reward_coef_range]
    th = c.curriculum_thresholds
    out = {
        "num_bodies": len(bodies), "num_commands": nc, "num_obs": int(c.env.num_observations), "num_privileged_obs": int(c.env.num_privileged_obs),
        "obs_history_length": int(c.env.num_observation_history), "dt": float(dt), "decimation": int(c.control.decimation),
        "max_episode_length": float(np.ceil(c.env.episode_length_s / dt)), "terminal_body_height": float(c.rewards.terminal_body_height),
        "use_terminal_body_height": bool(c.rewards.use_terminal_body_height), "resampling_time": float(cc.resampling_time),
        "rand_interval": float(np.ceil(c.domain_rand.rand_interval_s / dt)), "gravity_rand_interval": float(gri),
        "gravity_rand_duration": float(np.ceil(gri * c.domain_rand.gravity_impulse_duration)),
        "gravity_range": [float(x) for x in c.domain_rand.gravity_range], "randomize_gravity": bool(c.domain_rand.randomize_gravity),
        "only_positive_rewards": bool(c.rewards.only_positive_rewards), "only_positive_rewards_ji22_style": bool(c.rewards.only_positive_rewards_ji22_style),
        "sigma_rew_neg": float(c.rewards.sigma_rew_neg), "clip_observations": float(c.normalization.clip_observations),
        "clip_actions": float(c.normalization.clip_actions), "obs_scale_dof_pos": float(osc.dof_pos), "obs_scale_dof_vel": float(osc.dof_vel),
        "add_noise": bool(c.noise.add_noise), "observe_command": bool(c.env.observe_command), "observe_two_prev_actions": bool(c.env.observe_two_prev_actions),
        "observe_timing_parameter": bool(c.env.observe_timing_parameter), "observe_clock_inputs": bool(c.env.observe_clock_inputs),
        "priv_observe_friction": bool(c.env.priv_observe_friction), "priv_observe_restitution": bool(c.env.priv_observe_restitution),
        "friction_range": [float(x) for x in c.normalization.friction_range], "restitution_range": [float(x) for x in c.normalization.restitution_range],
        "kappa_gait_probs": float(c.rewards.kappa_gait_probs), "pacing_offset": bool(cc.pacing_offset), "gaitwise_curricula": bool(cc.gaitwise_curricula),
        "exclusive_phase_offset": bool(cc.exclusive_phase_offset), "balance_gait_distribution": bool(cc.balance_gait_distribution),
        "binary_phases": bool(cc.binary_phases), "randomize_motor_strength": bool(c.domain_rand.randomize_motor_strength),
        "motor_strength_range": [float(x) for x in c.domain_rand.motor_strength_range], "randomize_motor_offset": bool(c.domain_rand.randomize_motor_offset),
        "motor_offset_range": [float(x) for x in c.domain_rand.motor_offset_range], "x_init_range": float(c.terrain.x_init_range),
        "y_init_range": float(c.terrain.y_init_range), "yaw_init_range": float(c.terrain.yaw_init_range), "x_init_offset": float(c.terrain.x_init_offset),
        "y_init_offset": float(c.terrain.y_init_offset), "custom_origins": c.terrain.mesh_type in ("heightfield", "trimesh"),
        "action_scale": float(c.control.action_scale), "hip_scale_reduction": float(c.control.hip_scale_reduction), "control_type": str(c.control.control_type),
        "lag_timesteps": int(c.domain_rand.lag_timesteps), "randomize_lag_timesteps": bool(c.domain_rand.randomize_lag_timesteps),
        **{"rewards/" + k: float(getattr(c.rewards, k)) for k in ("tracking_sigma", "tracking_sigma_yaw", "base_height_target", "gait_force_sigma",
                                                                   "gait_vel_sigma", "max_contact_force")},
        "curriculum_limits": [[float(r[0]), float(r[1]), float(n)] for r, n in limits], "curriculum_init": [[float(r[0]), float(r[1])] for r in init_region],
        "curriculum_thresholds": [float(th.tracking_lin_vel), float(th.tracking_ang_vel), float(th.tracking_contacts_shaped_force),
                                  float(th.tracking_contacts_shaped_vel)],
        "reward_names": names, "reward_scales": scales, "category_names": ["pronk", "trot", "pace", "bound"] if cc.gaitwise_curricula else ["nominal"],
        "feet_indices": feet, "penalised_contact_indices": pen, "termination_contact_indices": term,
        "commands_scale": [f32(x) for x in commands_scale], "default_dof_pos": [default_dof_pos], "dof_pos_limits": lim,
        "noise_scale_vec": [float(x) for x in noise_vec], "torque_limits": [f32(x) for x in model.dof_effort],
        "base_init_state": [f32(x) for x in list(init.pos) + list(init.rot) + list(init.lin_vel) + list(init.ang_vel)],
    }
    return out


class VelocityTrackingEasyEnv:
    """go1_gym/envs/go1/velocity_tracking/__init__.py:12-60 (VelocityTrackingEasyEnv) with HistoryWrapper's obs dict folded in, around
    VelocityStepEngine.  PhysX is replaced by a SimBackend (env.SimBackend protocol: simulate(substep) leaves the new dof_state,
    refresh() the root / contact / rigid-body tensors); the rest of the constructor signature is the reference's.

      env = VelocityTrackingEasyEnv(sim_device="cuda:0", headless=True, cfg=Cfg, sim=my_backend)
      obs_dict = env.reset();  obs_dict, rew, done, info = env.step(actions)

    cfg: the reference's Cfg tree (config_from_cfg derives the engine's table), an already derived table (dict), or None for
    scripts/train_velocity_tracking.py's configuration.  command_source: "device" (default) or a host source (CurriculumAdapter)."""

    def __init__(self, sim_device, headless, num_envs=None, prone=False, deploy=False, cfg=None, eval_cfg=None, initial_dynamics_dict=None,
                 physics_engine="SIM_PHYSX", sim=None, command_source="device", env_origins=None, heightfield=None, seed=0):
        from . import workload
        if eval_cfg is not None or initial_dynamics_dict is not None or prone or deploy:
            raise NotImplementedError("eval_cfg / initial_dynamics_dict / prone / deploy: not on the cfg 4 step path")
        if sim is None:
            raise ValueError("a SimBackend is required (there is no PhysX here): see legged_tracking_b200.env.SimBackend / VelReplaySim")
        table = cfg4_config() if cfg is None else (dict(cfg) if isinstance(cfg, dict) else config_from_cfg(cfg))
        if num_envs is None:
            num_envs = int(cfg.env.num_envs) if (cfg is not None and not isinstance(cfg, dict)) else 4096
        self.cfg, self.num_envs, self.num_train_envs, self.num_eval_envs = cfg, int(num_envs), int(num_envs), 0
        self.device = torch.device(sim_device)
        self.headless = headless
        if env_origins is None:                                        # _get_env_origins on a plane (LRV:1700-1716): a grid, env_spacing apart
            side = int(np.floor(np.sqrt(self.num_envs)))
            rows = int(np.ceil(self.num_envs / side))
            xx, yy = np.meshgrid(np.arange(rows), np.arange(side), indexing="ij")
            spacing = float(getattr(getattr(cfg, "env", None), "env_spacing", 3.0)) if cfg is not None and not isinstance(cfg, dict) else 3.0
            env_origins = np.zeros((self.num_envs, 3), np.float32)
            env_origins[:, 0], env_origins[:, 1] = spacing * xx.flatten()[:self.num_envs], spacing * yy.flatten()[:self.num_envs]
        self.engine = VelocityStepEngine(table, self.num_envs, env_origins, workload.load_actuator(), command_source, device=self.device, seed=seed,
                                         heightfield=heightfield)
        e = self.engine
        self.num_obs, self.num_privileged_obs, self.num_actions = int(table["num_obs"]), int(table["num_privileged_obs"]), 12
        self.num_obs_history = self.num_obs * int(table["obs_history_length"])
        self.max_episode_length, self.dt = float(table["max_episode_length"]), float(table["dt"])
        self.sim = sim
        self.sim.attach(self)
        # the tensors a backend writes into / reads (Isaac Gym layouts)
        self.root_states, self.dof_state, self.contact_forces, self.rigid_body_state = e.root_states, e.dof_state, e.contact_forces, e.rigid_body_state
        self.torques = e.tq.state["torques"]
        self.extras = e.extras

    # VecEnv attributes the runner reads (go1_gym_learn/env/vec_env.py:10-39)
    @property
    def obs_buf(self):
        return self.engine.obs_buf

    @property
    def privileged_obs_buf(self):
        return self.engine.privileged_obs_buf

    @property
    def rew_buf(self):
        return self.engine.rew_buf

    @property
    def reset_buf(self):
        return self.engine.reset_buf

    @property
    def episode_length_buf(self):
        return self.engine.episode_length_buf

    @property
    def commands(self):
        return self.engine.commands

    def _obs_dict(self):
        e = self.engine
        return {"obs": e.obs_buf, "privileged_obs": e.privileged_obs_buf, "obs_history": e.obs_history}

    def get_observations(self):
        return self._obs_dict()

    def get_privileged_observations(self):
        return self.engine.privileged_obs_buf

    def step(self, actions):
        """LRV:60-105 through the backend: torques -> simulate -> refreshed dof_state, decimation times; then the post-physics chain."""
        e = self.engine
        actions = actions.to(device=self.device, dtype=torch.float32).contiguous()
        e.begin_step(actions)
        for sub in range(int(e.cfg["decimation"])):
            e.apply_torques(actions, sub)
            self.sim.simulate(sub)
        self.sim.refresh()
        obs, rew, done, extras = e.post_physics_step()
        if not e.device_paths:
            ids = done.nonzero(as_tuple=False).flatten()
            if ids.numel():
                self.sim.push_resets(ids)
        else:
            self.sim.push_resets(None)                                 # the rows are in root_states / dof_state; which ones: reset_buf
        return obs, rew, done, extras

    def reset_idx(self, env_ids):
        e = self.engine
        if e.device_paths:
            e.reset_buf.zero_()
            e.reset_buf[env_ids] = 1
            e._device_paths(_abi.DEFINES["LT_VEL_PATH_RESET"])
        else:
            e.reset_idx(env_ids)

    def reset(self):
        """velocity_tracking/__init__.py:56-59: reset every env, then one step with zero actions; history cleared (history_wrapper.py:36-40)."""
        self.reset_idx(torch.arange(self.num_envs, device=self.device))
        self.step(torch.zeros(self.num_envs, self.num_actions, device=self.device))
        self.engine.obs_history.zero_()
        return self._obs_dict()


class VelReplaySim:
    """SimBackend of synthetic 'PhysX output' for the velocity env: a few seeded frames kept in HBM and cycled (env.ReplaySim's
    counterpart for the single-actor tensors of cfg 4)."""

    def __init__(self, seed=1000, n_frames=3):
        self.seed, self.n_frames, self.k = seed, n_frames, 0

    def attach(self, env):
        from . import synthetic
        self.env = env
        e = env.engine
        synth = synthetic.SyntheticState(e.N, e.env_origins.cpu().numpy(), num_bodies=int(e.cfg["num_bodies"]), seed=self.seed,
                                         feet_indices=list(e.cfg["feet_indices"]))
        dev = e.device
        self.frames = []
        for k in range(self.n_frames):
            root = synth.root(k)
            self.frames.append({"dof": [synth.dof_state(k, s).to(dev) for s in range(int(e.cfg["decimation"]))], "root": root.to(dev),
                                "contact": synth.contact(k).to(dev), "rigid": synth.rigid(k, root).to(dev)})
        e.episode_length_buf.copy_(synth.initial_episode_length(e.cfg["max_episode_length"]).to(dev))

    def simulate(self, substep):
        self.env.dof_state.copy_(self.frames[self.k % self.n_frames]["dof"][substep].view(self.env.dof_state.shape))

    def refresh(self):
        f = self.frames[self.k % self.n_frames]
        self.env.root_states.copy_(f["root"])
        self.env.contact_forces.copy_(f["contact"])
        self.env.rigid_body_state.copy_(f["rigid"].view(self.env.rigid_body_state.shape))
        self.k += 1

    def push_resets(self, env_ids):
        pass
