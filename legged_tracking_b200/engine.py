"""StepEngine: torch-owned device buffers + the liblt_b200 handle that steps them.

This is the host side of the drop-in boundary: it owns what LeggedRobot._init_buffers owns in the reference
(go1_gym/envs/base/legged_robot_trajectory_tracking.py:1169-1327, LR below) -- the four simulator tensors in
Isaac Gym's acquire_*_tensor layout and one tensor per env attribute -- and forwards step()'s two phases to the
C ABI (include/lt_b200.h).  PyTorch is used for device memory and streams only; all per-env arithmetic happens
in the CUDA kernels.  No CPU fallback: constructing an engine on a CPU device raises.
"""
import ctypes as C

import numpy as np
import torch

from . import _abi, cconfig
from . import dist as ltd

_TORCH_DTYPES = {"f32": torch.float32, "i64": torch.int64, "bool": torch.bool, "i32": torch.int32}


class GravitySchedule:
    """Host-side control flow of _randomize_gravity (LR:645-660, 826-830): WHEN gravity changes is decided by the
    global step counter, identical for every env, so it stays on the host; the per-env rows are written by the
    step kernel.  Arithmetic on the 3-vector uses torch CPU float32 ops in the reference's order."""

    def __init__(self, flat):
        self.enabled = bool(flat["randomize_gravity"])
        self.lo, self.hi = [float(v) for v in flat["gravity_range"]]
        self.interval = int(flat["gravity_rand_interval"])
        self.duration = int(flat["gravity_rand_duration"])
        self.gravities = torch.zeros(3)
        self.common_step_counter = 0

    @staticmethod
    def gravity_vec_of(gravities):
        gravity = gravities + torch.tensor([0.0, 0.0, -9.8])
        return gravity / torch.norm(gravity)

    def fires_next(self):
        """True if advance() will return a new gravity on the next step (lets a caller keep such steps out of a CUDA graph)."""
        n = self.common_step_counter + 1
        return n % self.interval == 0 or int(n - self.duration) % self.interval == 0

    def advance(self, draw_uniform3):
        """Called once per policy step BEFORE the kernel launch.  Returns the new `gravities` row (float32[3]
        tensor) if _randomize_gravity fires during this step's callback, else None."""
        self.common_step_counter += 1
        new = None
        if self.common_step_counter % self.interval == 0:
            if self.enabled:
                u = torch.as_tensor(draw_uniform3(), dtype=torch.float32)
                new = u * (self.hi - self.lo) + self.lo
            else:
                new = self.gravities.clone()              # gravity_vec is recomputed from the unchanged rows
        if int(self.common_step_counter - self.duration) % self.interval == 0:
            new = torch.zeros(3)
        if new is not None:
            self.gravities = new.clone()
        return new


class ExplorationCurriculum:
    """Host side of LeggedRobot.update_curriculum (LR:171-183): once the global step counter exceeds
    rewards.exploration_steps the exploration_lin / exploration_yaw scales decay linearly to zero, in Python double
    arithmetic exactly as the reference mutates its reward_scales dict; the kernels pick the new scales up from the
    next step (the update runs after compute_observations, LR:146)."""

    KEYS = ("exploration_lin", "exploration_yaw")

    def __init__(self, flat):
        self.scales = {k: float(v) for k, v in flat["reward_scales"].items()}
        self.cfg_scales = flat["reward_scales_cfg"]
        self.dt = float(flat["dt"])
        self.steps = float(flat["exploration_steps"])
        self.names = list(flat["reward_names"])

    def advance(self, common_step_counter):
        """-> True if a scale changed."""
        changed = False
        for key in self.KEYS:
            if key in self.scales and common_step_counter > self.steps:
                self.scales[key] -= self.cfg_scales[key] * self.dt / self.steps
                self.scales[key] = max(self.scales[key], 0)
                changed = True
        return changed

    def term_scales(self):
        return [self.scales[n] for n in self.names]

    def changes_at(self, common_step_counter):
        """Would advance(common_step_counter) change a scale?  (A captured CUDA graph holds the scales it was captured with.)"""
        return any(key in self.scales and common_step_counter > self.steps and self.scales[key] > 0 for key in self.KEYS)


class StepEngine:
    def __init__(self, flat, terrain, actuator, device="cuda:0", seed=0, envs_per_warp=0, track_command_sums=True,
                 export_height_indices=False, alias_history=True):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _abi.LtError("StepEngine needs a CUDA device (sm_100a); there is no CPU implementation in this package")
        self.lib = _abi.load()
        self.flat = flat
        self.N = int(flat["num_envs"])
        self.cfg = cconfig.make_lt_config(flat, seed=seed, envs_per_warp=envs_per_warp)
        self.n_slots = self.lib.lt_num_uniform_slots(C.byref(self.cfg))
        tables = cconfig.Tables(flat, actuator)
        self.handle = C.c_void_p()
        with torch.cuda.device(self.device):
            _abi.check(self.lib.lt_create(C.byref(self.cfg), C.byref(tables.c), C.byref(self.handle)), None, "lt_create")
        dev = self.device
        N, A, B, D = self.N, int(flat["num_actors"]), int(flat["num_bodies"]), int(flat["num_dof"])
        # ---- simulator tensors, Isaac Gym layout (LR:1173-1201)
        self.root_states = torch.zeros(N * A, 13, device=dev)
        self.root_states[:, 6] = 1.0
        self.dof_state = torch.zeros(N * D, 2, device=dev)
        self.contact_forces = torch.zeros(N, B, 3, device=dev)
        self.rigid_body_state = torch.zeros(N * B, 13, device=dev)
        sim = _abi.lt_sim_ptrs(self.root_states.data_ptr(), self.dof_state.data_ptr(), self.contact_forces.data_ptr(),
                               self.rigid_body_state.data_ptr())
        _abi.check(self.lib.lt_bind_sim(self.handle, C.byref(sim)), self.handle, "lt_bind_sim")
        # ---- terrain (LR:1808-1847)
        def t(x, dt):          # numpy / list / torch tensor on any device (terrain.TunnelTerrain hands over a device-resident table)
            if isinstance(x, torch.Tensor):
                return x.to(device=dev, dtype=dt).contiguous()
            return torch.as_tensor(np.ascontiguousarray(np.asarray(x)), dtype=dt).to(dev).contiguous()
        self.env_origins = t(terrain["env_origins"], torch.float32).view(N, 3)
        self.plane = flat["mesh_type"] == "plane"
        ter = _abi.lt_terrain_ptrs()
        ter.env_origins = self.env_origins.data_ptr()
        n_tiles = 0
        if not self.plane:
            self.tile_heights = t(terrain["tile_heights"], torch.float32)
            self.env_tile = t(terrain["env_tile"], torch.int32)
            self.env_terrain_origin = t(terrain["env_terrain_origin"], torch.float32).view(N, 3)
            n_tiles = int(self.tile_heights.shape[0])
            assert tuple(self.tile_heights.shape[1:]) == (2, flat["tile_rows"], flat["tile_cols"])
            assert int(self.env_tile.max()) < n_tiles
            ter.tile_heights = self.tile_heights.data_ptr()
            ter.env_tile = self.env_tile.data_ptr()
            ter.env_terrain_origin = self.env_terrain_origin.data_ptr()
        _abi.check(self.lib.lt_bind_terrain(self.handle, C.byref(ter), n_tiles), self.handle, "lt_bind_terrain")
        # ---- env state, one tensor per LeggedRobot attribute (LR:1203-1305)
        self.state = {}
        ptrs = _abi.lt_state_ptrs()
        skip = set()
        if not track_command_sums:
            skip.add("command_sums")
        if not export_height_indices:
            skip |= {"height_px", "height_py"}
        for name, dt, shape in cconfig.state_spec(flat):
            if name in skip:
                setattr(ptrs, name, None)
                continue
            if name == "obs_history" and alias_history and flat["obs_history_length"] == 1 and "obs_buf" in self.state:
                # H = 1 (every trajectory-tracking config): obs_history = cat(obs_history[:, num_obs:], obs) IS obs
                # (history_wrapper.py:22), so the two names share one buffer and the observation row is written once
                self.state[name] = self.state["obs_buf"]
                setattr(ptrs, name, self.state["obs_buf"].data_ptr())
                continue
            buf = torch.full(shape, cconfig.STATE_INIT.get(name, 0), dtype=_TORCH_DTYPES[dt], device=dev)
            self.state[name] = buf
            setattr(ptrs, name, buf.data_ptr())
        self.state["gravity_vec"].copy_(torch.tensor([0.0, 0.0, -1.0]))
        _abi.check(self.lib.lt_bind_state(self.handle, C.byref(ptrs)), self.handle, "lt_bind_state")
        self._ptrs = ptrs
        self.counters = self._wrap_i32(self.lib.lt_counters(self.handle), 4)
        self.planner = None
        if flat.get("sampling_based_planning"):
            from .planner import LocalPlanner
            px, py = np.asarray(flat["measured_points_x"], np.float32), np.asarray(flat["measured_points_y"], np.float32)
            gx, gy = np.meshgrid(px, py, indexing="ij")                                        # _init_height_points, LR:1902-1916
            self.planner = LocalPlanner(N, np.stack([gx.reshape(-1), gy.reshape(-1)], axis=1), flat["candidate_target_poses"],
                                        plan_interval=flat.get("plan_interval", 10), num_actors=flat["num_actors"], device=dev)
            self.measured_heights = torch.zeros(N, 2, px.size * py.size, device=dev)
            self._arange = torch.arange(N, device=dev)
        self.gravity = GravitySchedule(flat)
        self.curriculum = ExplorationCurriculum(flat)
        self.reward_names = list(flat["reward_names"])
        self.n_external = sum(1 for n in self.reward_names if n == "external" or n in flat.get("external_reward_names", ()))
        self.sum_names = self.reward_names + ["total", "total_pos", "total_neg"]
        self._keepalive = []
        # LR:332-335 buckets a term by the sign of its ALL-ENV sum.  When this engine holds only a shard of the envs (one rank
        # of an env-sharded job, or one chunk of a HostStepPipeline) the shard-local sums of the sign-indefinite terms must be
        # combined before the step kernel reads them: lt_reward_sign_prepass -> exchange -> lt_post_physics_step.
        self.any_bucket_by_sum = any(int(self.cfg.term_bucket[i]) == _abi.ENUMS["LT_BUCKET_BY_SUM"] for i in range(self.cfg.n_terms))
        self.term_sums = self._wrap_f32(self.lib.lt_term_sum_buffer(self.handle), _abi.DEFINES["LT_MAX_TERMS"])
        self.term_sum_exchange = None         # callable(float32[LT_MAX_TERMS] device tensor) -> None (in place); None = dist all-reduce

    def _wrap_i32(self, ptr, n):
        """torch view of a library-owned device int32 array (read-only use)."""
        class _Holder:
            pass
        h = _Holder()
        h.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (int(ptr), False), "version": 2}
        return torch.as_tensor(h, device=self.device)

    def _wrap_f32(self, ptr, n):
        class _Holder:
            pass
        h = _Holder()
        h.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}
        return torch.as_tensor(h, device=self.device)

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.lib.lt_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    # ---------------------------------------------------------------- views with the reference's attribute names
    def __getattr__(self, name):
        st = self.__dict__.get("state")
        if st is not None and name in st:
            return st[name]
        raise AttributeError(name)

    @property
    def robot_root(self):
        return self.root_states.view(self.N, -1, 13)[:, 0, :]

    @property
    def dof_pos(self):
        return self.dof_state.view(self.N, -1, 2)[..., 0]

    @property
    def dof_vel(self):
        return self.dof_state.view(self.N, -1, 2)[..., 1]

    def lag_buffer_logical(self):
        """lag_buffer in the reference's list order (slot 0 = oldest, LR:972-974)."""
        head = int(self.counters[2].item())
        L = self.state["lag_buffer"].shape[0]
        idx = [(head + j) % L for j in range(L)]
        return self.state["lag_buffer"][idx]

    def bind_sim(self, root_states=None, dof_state=None, contact_forces=None, rigid_body_state=None):
        """Re-point the step at other simulator tensors (host-only; e.g. a replay set).  LR:1173-1201."""
        for name, tns in (("root_states", root_states), ("dof_state", dof_state), ("contact_forces", contact_forces),
                          ("rigid_body_state", rigid_body_state)):
            if tns is not None:
                assert tns.is_cuda and tns.dtype == torch.float32 and tns.is_contiguous() and tns.numel() == getattr(self, name).numel(), name
                setattr(self, name, tns)
        sim = _abi.lt_sim_ptrs(self.root_states.data_ptr(), self.dof_state.data_ptr(), self.contact_forces.data_ptr(),
                               self.rigid_body_state.data_ptr())
        _abi.check(self.lib.lt_bind_sim(self.handle, C.byref(sim)), self.handle, "lt_bind_sim")

    # ---------------------------------------------------------------- stepping
    @staticmethod
    def _stream_ptr(stream):
        s = stream if stream is not None else torch.cuda.current_stream()
        return C.c_void_p(s.cuda_stream)

    def compute_torques(self, actions, substep, stream=None):
        """LeggedRobot._compute_torques (LR:957-996).  `actions`: raw policy output [N,12] on the device."""
        assert actions.is_cuda and actions.dtype == torch.float32 and actions.is_contiguous()
        _abi.check(self.lib.lt_compute_torques(self.handle, C.c_void_p(actions.data_ptr()), int(substep),
                                               self._stream_ptr(stream)), self.handle, "lt_compute_torques")
        return self.state["torques"]

    def post_physics_step(self, uniforms=None, noise_uniforms=None, extra_reward=None, gravity_uniform=None, stream=None,
                          advance_gravity=True):
        """LeggedRobot.post_physics_step (LR:114-169) on the bound simulator tensors."""
        args = _abi.lt_step_args()
        for name, tns, shape in (("uniforms", uniforms, (self.N, self.n_slots)),
                                 ("noise_uniforms", noise_uniforms, (self.N, self.flat["num_obs"])),
                                 ("extra_reward", extra_reward, (self.n_external, self.N))):
            if tns is not None:
                if name == "extra_reward" and tns.dim() == 1:
                    tns = tns.view(1, -1)
                assert tns.is_cuda and tns.dtype == torch.float32 and tns.is_contiguous() and tuple(tns.shape) == shape, name
                setattr(args, name, tns.data_ptr())
        if self.planner is not None:
            self._plan(args, stream)
        new_g = None
        if advance_gravity:
            draw = (lambda: gravity_uniform) if gravity_uniform is not None else (lambda: torch.rand(3))
            new_g = self.gravity.advance(draw)
        if new_g is not None:
            g_host = (C.c_float * 3)(*[float(v) for v in new_g])
            args.new_gravity = C.cast(g_host, C.c_void_p)
        if self.any_bucket_by_sum and (self.term_sum_exchange is not None or ltd.world_size() > 1):
            self.sign_prepass(args, stream)
            with torch.cuda.stream(stream if stream is not None else torch.cuda.current_stream()):
                if self.term_sum_exchange is not None:
                    self.term_sum_exchange(self.term_sums)
                else:
                    ltd.allreduce_sum_(self.term_sums)
        _abi.check(self.lib.lt_post_physics_step(self.handle, C.byref(args), self._stream_ptr(stream)), self.handle,
                   "lt_post_physics_step")
        if new_g is not None:
            # gravity_vec changes AFTER this step's projected_gravity was computed (LR:134 precedes LR:826-830)
            self.state["gravity_vec"].copy_(GravitySchedule.gravity_vec_of(new_g), non_blocking=False)
        if advance_gravity and self.curriculum.advance(self.gravity.common_step_counter):
            self.set_reward_scales(self.curriculum.term_scales())
            self._scales_changed = True           # graphs captured before this point hold stale scales (graph_replay_ok)

    def _plan(self, args, stream):
        """Sampling-based planning (LR:786-799, 856-916) ahead of the fused step: full-grid heights (lt_measure_heights), the planner
        kernel (lt_plan_target_pose), and this step's local target handed to lt_post_physics_step through lt_step_args."""
        st, pl = self.state, self.planner
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        with torch.cuda.stream(s):
            _abi.check(self.lib.lt_measure_heights(self.handle, C.c_void_p(self.measured_heights.data_ptr()), C.c_void_p(s.cuda_stream)),
                       self.handle, "lt_measure_heights")
            target = st["trajectories"][self._arange, st["curr_pose_index"]].contiguous()        # LR:854
            ep = st["episode_length_buf"] + 1                                                  # LR:126 precedes the callback
            pl.plan(target, self.root_states, self.measured_heights, ep, st["plan_buf"].view(torch.uint8), stream=s)
        self._plan_keep = (target, ep)
        args.local_target_poses = pl.local_target_poses.data_ptr()
        args.local_relative_linear = pl.local_relative_linear.data_ptr()
        args.local_relative_rotation = pl.local_relative_rotation.data_ptr()
        args.replan = pl.replan.data_ptr()

    def graph_replay_ok(self):
        """May the NEXT step be a replay of a CUDA graph captured from post_physics_step(advance_gravity=False)?  A captured
        graph freezes what the host passes by value at capture time: the gravity event flags and the per-term reward scales
        (lt_config travels as a kernel parameter).  So a step on which _randomize_gravity fires (LR:826-830) or on which
        update_curriculum changes a scale (LR:171-183) must be launched eagerly (and the graph re-captured once the scales
        have changed, since later replays would still carry the old ones).  After a replay the caller advances
        gravity.common_step_counter itself."""
        return not self.gravity.fires_next() and not self.scales_stale_in_graph()

    def scales_stale_in_graph(self):
        n = self.gravity.common_step_counter + 1
        return self.curriculum.changes_at(n) or self.curriculum.changes_at(n - 1) or getattr(self, "_scales_changed", False)

    def sign_prepass(self, args=None, stream=None):
        """lt_reward_sign_prepass: shard-local sums of the sign-indefinite terms -> self.term_sums (device float32[LT_MAX_TERMS]).
        The next lt_post_physics_step uses whatever that buffer then holds instead of recomputing it."""
        if args is None:
            args = _abi.lt_step_args()
        _abi.check(self.lib.lt_reward_sign_prepass(self.handle, C.byref(args), self._stream_ptr(stream)), self.handle,
                   "lt_reward_sign_prepass")

    def set_reward_scales(self, scales):
        """New per-term scales (already multiplied by dt, order of reward_names) for the following steps (LR:171-183)."""
        arr = (C.c_float * len(scales))(*[float(v) for v in scales])
        _abi.check(self.lib.lt_set_term_scales(self.handle, arr, len(scales)), self.handle, "lt_set_term_scales")

    def reset_idx(self, env_ids, uniforms=None, stream=None):
        """LeggedRobot.reset_idx (LR:218-296) for an explicit id list, outside a step."""
        env_ids = torch.as_tensor(env_ids, dtype=torch.int64, device=self.device).contiguous()
        if env_ids.numel() == 0:
            return
        u = None
        if uniforms is not None:
            assert uniforms.is_cuda and uniforms.dtype == torch.float32 and tuple(uniforms.shape) == (self.N, self.n_slots)
            u = C.c_void_p(uniforms.data_ptr())
        _abi.check(self.lib.lt_reset_idx(self.handle, C.c_void_p(env_ids.data_ptr()), int(env_ids.numel()), None, u,
                                         self._stream_ptr(stream)), self.handle, "lt_reset_idx")
        self._keepalive = [env_ids, uniforms]

    def reset_env_ids(self):
        """env_ids = reset_buf.nonzero().flatten() (LR:143); synchronises to read the count."""
        n = int(self.state["reset_count"].item())
        return self.state["reset_env_ids"][:n]

    # ---------------------------------------------------------------- state import / export (oracle naming)
    def load_state(self, st):
        """Accepts the keys of oracle.port.OraclePort.export_state / the `init/*` arrays of the golden fixtures."""
        def put(dst, src):
            src = torch.as_tensor(np.asarray(src))
            dst.copy_(src.to(dst.dtype).reshape(dst.shape).to(dst.device))
        for name, buf in self.state.items():
            if name in st and name not in ("lag_buffer", "episode_sums", "command_sums"):
                put(buf, st[name])
        if "root_states" in st:
            put(self.root_states, st["root_states"])
        if "dof_state" in st:
            put(self.dof_state, st["dof_state"])
        if "lag_buffer" in st:
            put(self.state["lag_buffer"], st["lag_buffer"])          # logical order == physical while head == 0
            assert int(self.counters[2].item()) == 0, "load_state: lag ring already rotated"
        for r, name in enumerate(self.sum_names):
            if "episode_sums/" + name in st:
                put(self.state["episode_sums"][r], st["episode_sums/" + name])
        if "command_sums" in self.state:
            for r, name in enumerate(self.reward_names + ["ep_timesteps"]):
                if "command_sums/" + name in st:
                    put(self.state["command_sums"][r], st["command_sums/" + name])
        if "gravity_vec" in st:
            put(self.state["gravity_vec"], np.asarray(st["gravity_vec"]).reshape(-1)[:3])
        if "gravities" in st:
            self.gravity.gravities = torch.as_tensor(np.asarray(st["gravities"]), dtype=torch.float32).reshape(-1, 3)[0].clone()
        if "common_step_counter" in st:
            self.gravity.common_step_counter = int(st["common_step_counter"])
        if self.planner is not None:
            if "plan_length_buf" in st:
                put(self.planner.plan_length_buf, st["plan_length_buf"])
            if "local_target_poses" in st:
                put(self.planner.local_target_poses, st["local_target_poses"])

    def export_state(self):
        out = {k: v.detach().cpu() for k, v in self.state.items()
               if k not in ("lag_buffer", "episode_sums", "command_sums", "finished_sums")}
        out["root_states"] = self.root_states.cpu()
        out["dof_state"] = self.dof_state.cpu()
        out["lag_buffer"] = self.lag_buffer_logical().cpu()
        for r, name in enumerate(self.sum_names):
            out["episode_sums/" + name] = self.state["episode_sums"][r].cpu()
        if "command_sums" in self.state:
            for r, name in enumerate(self.reward_names + ["ep_timesteps"]):
                out["command_sums/" + name] = self.state["command_sums"][r].cpu()
        out["common_step_counter"] = torch.tensor(self.gravity.common_step_counter)
        if self.planner is not None:
            out["plan_length_buf"] = self.planner.plan_length_buf.cpu()
            out["local_target_poses"] = self.planner.local_target_poses.cpu()
        return out


class HostStepPipeline:
    """The step for callers whose simulator state lives in HOST memory (the reference-facing path bench.py reports as `e2e`).

    One policy step moves ~0.75 KB per env up (actions, 4 x dof_state, root_states, contact forces) and ~1.05 KB per env down
    (obs, privileged obs, reward, done), so with a single engine the step is three serial phases -- H2D, kernels, D2H -- and the
    PCIe link idles in one direction at any time.  Envs are independent, so the batch is cut into
    `n_chunks` contiguous env shards, each with its own StepEngine on its own CUDA stream: while shard k's results go down,
    shard k+1's inputs come up and shard k+2's kernels run.  Results are identical to n_chunks independent engines (the same
    sharding rule as across GPUs: shard sizes must keep env -> tile = i mod tiles, i.e. be multiples of rows * cols).

    host_in (pinned, per shard k):  actions [M,12], dof [decimation][M*12,2], root [M*A,13], contact [M,B,3]
    host_out (pinned, per shard k): obs [M,num_obs], priv [M,num_priv], rew [M], done [M] bool
      and, optionally, the write-back of what a reset changed inside the simulator state (the reference pushes it with
      set_actor_root_state_tensor_indexed / set_dof_state_tensor_indexed, LR:1010-1013, 1049-1052) -- alloc_host_outputs(cap):
      reset_count int32[1], reset_ids int64[cap] (ascending; the first reset_count are valid), reset_root [cap,13], reset_dof
      [cap,24]: the robot root state and (pos, vel) rows of the envs that were reset this step.  A step with more than `cap`
      resets in a shard is reported by reset_count > cap; the caller then reads the full tensors (StepEngine.root_states / dof_state).
    """

    def __init__(self, engines):
        self.engines = list(engines)
        dev = self.engines[0].device
        self.streams = [torch.cuda.Stream(device=dev) for _ in self.engines]
        self.dev_in, self._packed_dev, self._layout = [], [], []
        for e in self.engines:
            f = e.flat
            # one contiguous device block per shard (actions | dof_state x decimation | root_states | contact forces), so that a
            # caller who keeps its simulator tensors in the matching pinned block (alloc_host_inputs) uploads a shard with ONE copy
            sizes = [("actions", (e.N, 12))] + [(f"dof{k}", (e.N * 12, 2)) for k in range(f["decimation"])] + \
                    [("root", (e.N * f["num_actors"], 13)), ("contact", (e.N, f["num_bodies"], 3))]
            total = sum(int(np.prod(shp)) for _, shp in sizes)
            block = torch.empty(total, device=dev)
            views, off = {}, 0
            for name, shp in sizes:
                k = int(np.prod(shp))
                views[name] = block[off:off + k].view(*shp)
                off += k
            own = {"actions": views["actions"], "dof": [views[f"dof{k}"] for k in range(f["decimation"])], "root": views["root"],
                   "contact": views["contact"]}
            e.bind_sim(root_states=own["root"], dof_state=own["dof"][-1], contact_forces=own["contact"])
            self.dev_in.append(own)
            self._packed_dev.append(block)
            self._layout.append(sizes)

    def alloc_host_inputs(self):
        """Pinned host blocks in the device layout, one per shard: {"packed": flat tensor, "actions", "dof": [...], "root", "contact":
        views into it}.  A simulator that writes its state into these views gets single-copy uploads in step()."""
        out = []
        for sizes in self._layout:
            total = sum(int(np.prod(shp)) for _, shp in sizes)
            block = torch.empty(total).pin_memory()
            views, off = {}, 0
            for name, shp in sizes:
                k = int(np.prod(shp))
                views[name] = block[off:off + k].view(*shp)
                off += k
            n_sub = sum(1 for name, _ in sizes if name.startswith("dof"))
            out.append({"packed": block, "actions": views["actions"], "dof": [views[f"dof{k}"] for k in range(n_sub)],
                        "root": views["root"], "contact": views["contact"]})
        return out

    @staticmethod
    def _tail_views(block, layout):
        views = {}
        for name, off, nbytes, dt, shape in layout:
            views[name] = block[off:off + nbytes].view(dt).view(*shape)
        return views

    def alloc_host_outputs(self, reset_capacity=None):
        """Pinned result blocks, one dict per shard: "obs", and "tail" = ONE byte block holding priv | rew | done | reset_count |
        reset_ids | reset_root | reset_dof (the other keys are views into it), so that a shard's results come down in two
        transfers -- every extra small copy costs the down-link a hand-over.  reset_capacity = write-back rows per shard
        (default N/8)."""
        out = []
        self._tail_dev = []
        for e in self.engines:
            f = e.flat
            cap = int(reset_capacity) if reset_capacity is not None else max(e.N // 8, 1)
            cap = min(cap, e.N)
            segs = [("priv", torch.float32, (e.N, f["num_privileged_obs"])), ("rew", torch.float32, (e.N,)), ("done", torch.bool, (e.N,)),
                    ("reset_count", torch.int32, (1,)), ("reset_ids", torch.int64, (cap,)), ("reset_root", torch.float32, (cap, 13)),
                    ("reset_dof", torch.float32, (cap, 2 * f["num_dof"]))]
            layout, off = [], 0
            for name, dt, shape in segs:
                nbytes = int(np.prod(shape)) * torch.empty((), dtype=dt).element_size()
                layout.append((name, off, nbytes, dt, shape))
                off = (off + nbytes + 15) // 16 * 16
            host = torch.zeros(off, dtype=torch.uint8).pin_memory()
            dev = torch.zeros(off, dtype=torch.uint8, device=e.device)
            ho = {"obs": torch.zeros(e.N, f["num_obs"]).pin_memory(), "tail": host}
            ho.update(self._tail_views(host, layout))
            out.append(ho)
            self._tail_dev.append((host.data_ptr(), dev, self._tail_views(dev, layout)))
        return out

    @property
    def num_envs(self):
        return sum(e.N for e in self.engines)

    def _issue(self, host_in, host_out, advance_gravity=True):
        """Fork from the current stream to the shard streams, issue every shard's copies and kernels, join back."""
        cur = torch.cuda.current_stream()
        for s in self.streams:
            s.wait_stream(cur)
        for c, (e, s, own, blk, hi, ho) in enumerate(zip(self.engines, self.streams, self.dev_in, self._packed_dev, host_in, host_out)):
            with torch.cuda.stream(s):
                if "packed" in hi:
                    blk.copy_(hi["packed"], non_blocking=True)         # the whole shard in one transfer
                else:
                    # the shard's replayed inputs go up as ONE uninterrupted burst (a copy that waits for the previous sub-step's
                    # kernel leaves the up-link idle for a copy-engine / compute hand-over every sub-step: 34 GB/s instead of ~50)
                    own["actions"].copy_(hi["actions"], non_blocking=True)
                    for sub, dof in enumerate(hi["dof"]):
                        own["dof"][sub].copy_(dof, non_blocking=True)     # what PhysX hands back after sub-step `sub`
                    own["root"].copy_(hi["root"], non_blocking=True)
                    own["contact"].copy_(hi["contact"], non_blocking=True)
                for sub in range(len(own["dof"])):
                    e.bind_sim(dof_state=own["dof"][sub])
                    e.compute_torques(own["actions"], sub)
                e.post_physics_step(advance_gravity=advance_gravity)
                ho["obs"].copy_(e.state["obs_buf"], non_blocking=True)
                if "tail" in ho:                                        # alloc_host_outputs: everything else in ONE transfer
                    key, dev_block, tv = self._tail_dev[c]
                    if key != ho["tail"].data_ptr():
                        raise ValueError("host_out was not made by this pipeline's last alloc_host_outputs()")
                    cap = tv["reset_ids"].shape[0]
                    ids = e.state["reset_env_ids"][:cap]
                    safe = ids.clamp(0, e.N - 1)                        # entries past reset_count are stale
                    tv["priv"].copy_(e.state["privileged_obs_buf"])
                    tv["rew"].copy_(e.state["rew_buf"])
                    tv["done"].copy_(e.state["reset_buf"])
                    tv["reset_count"].copy_(e.state["reset_count"])
                    tv["reset_ids"].copy_(ids)
                    # rows the step rewrote on reset, for a host-resident simulator (LR:1010-1013, 1049-1052)
                    torch.index_select(own["root"].view(e.N, -1, 13)[:, 0, :], 0, safe, out=tv["reset_root"])
                    torch.index_select(own["dof"][-1].view(e.N, -1), 0, safe, out=tv["reset_dof"])
                    ho["tail"].copy_(dev_block, non_blocking=True)
                else:
                    ho["priv"].copy_(e.state["privileged_obs_buf"], non_blocking=True)
                    ho["rew"].copy_(e.state["rew_buf"], non_blocking=True)
                    ho["done"].copy_(e.state["reset_buf"], non_blocking=True)
        for s in self.streams:
            cur.wait_stream(s)

    def step(self, host_in, host_out):
        """Issue every shard's copies and kernels on its stream, then wait for all of them (the caller needs obs / rew / done
        of every env before it can act again)."""
        self._issue(host_in, host_out)
        torch.cuda.current_stream().synchronize()

    def capture(self, host_in, host_out):
        """The whole pipelined step -- every shard's upload, kernels and downloads, forked over the shard streams -- as ONE CUDA
        graph bound to these pinned buffers (a simulator that keeps writing its state into the same alloc_host_inputs() blocks
        replays it every step).  One graph launch replaces ~15 Python-issued calls per shard, which is what lets more, smaller
        shards pay off: the first shard's upload + kernels and the last shard's download are the only unoverlapped parts."""
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(device=self.engines[0].device)
        with torch.cuda.graph(g, stream=side):
            self._issue(host_in, host_out, advance_gravity=False)
        return g

    def graph_replay_ok(self):
        return all(e.graph_replay_ok() for e in self.engines)

    def replay(self, graph):
        """One step by graph replay (only when graph_replay_ok(): no gravity event, no reward-scale change on this step)."""
        graph.replay()
        for e in self.engines:
            e.gravity.common_step_counter += 1
        torch.cuda.current_stream().synchronize()
