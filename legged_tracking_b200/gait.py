"""Velocity-tracking specifics as CUDA kernels behind the reference's names:

  GaitClock          LeggedRobot._step_contact_targets, go1_gym/envs/base/legged_robot_velocity_tracking.py:844-923
  CoRLRewardKernel   the CoRLRewards container, go1_gym/envs/rewards/corl_rewards.py:15-201 -- all requested _reward_<name> terms
                     in ONE launch, written as the [n_terms, N] block of unscaled values that lt_step_args.extra_reward takes

  CommandResampler   the tensor half of LeggedRobot._resample_commands, :776-840

All are pointer marshalling around liblt_b200.so (csrc/lt_gait.cu); CUDA only.  vel_env.VelocityStepEngine chains them into the
cfg 4 (VelocityTrackingEasyEnv) step.
"""
import ctypes as C

import torch

from . import _abi

CORL_TERMS = ["tracking_lin_vel", "tracking_ang_vel", "lin_vel_z", "ang_vel_xy", "orientation", "torques", "dof_acc", "action_rate",
              "collision", "dof_pos_limits", "jump", "tracking_contacts_shaped_force", "tracking_contacts_shaped_vel", "dof_pos", "dof_vel",
              "action_smoothness_1", "action_smoothness_2", "feet_slip", "feet_contact_vel", "feet_contact_forces",
              "feet_clearance_cmd_linear", "feet_impact_vel", "orientation_control", "raibert_heuristic"]


def _need_cuda(device, what):
    device = torch.device(device)
    if device.type != "cuda":
        raise _abi.LtError(f"{what} runs as an sm_100a kernel: there is no CPU path")
    return device


class GaitClock:
    """State and outputs of _step_contact_targets: gait_indices [N] (advanced in place by step()), foot_indices, clock_inputs,
    doubletime_clock_inputs, halftime_clock_inputs, desired_contact_states [N,4]."""

    def __init__(self, num_envs, dt, kappa, pacing_offset=False, device="cuda:0"):
        self.device = _need_cuda(device, "GaitClock")
        self.lib = _abi.load()
        self.N, self.dt, self.kappa, self.pacing_offset = int(num_envs), float(dt), float(kappa), bool(pacing_offset)
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=self.device)          # noqa: E731
        self.gait_indices = z(self.N)
        self.foot_indices, self.clock_inputs, self.doubletime_clock_inputs = z(self.N, 4), z(self.N, 4), z(self.N, 4)
        self.halftime_clock_inputs, self.desired_contact_states = z(self.N, 4), z(self.N, 4)

    def step(self, commands, stream=None):
        if commands.device != self.device or commands.dtype != torch.float32 or not commands.is_contiguous() or commands.shape[0] != self.N:
            raise ValueError(f"commands: expected a contiguous float32 [N, num_commands] tensor on {self.device}")
        p = _abi.lt_gait_ptrs()
        p.commands = commands.data_ptr()
        for k in ("gait_indices", "foot_indices", "clock_inputs", "doubletime_clock_inputs", "halftime_clock_inputs", "desired_contact_states"):
            setattr(p, k, getattr(self, k).data_ptr())
        st = stream if stream is not None else torch.cuda.current_stream(self.device)
        _abi.check(self.lib.lt_step_contact_targets(C.byref(p), self.N, int(commands.shape[1]), C.c_float(self.dt), C.c_float(self.kappa),
                                                    1 if self.pacing_offset else 0, C.c_void_p(st.cuda_stream)), None, "lt_step_contact_targets")
        self.desired_footswing_height = commands[:, 9] if commands.shape[1] > 9 else None          # LRV:921-922


class CoRLRewardKernel:
    """names: the reward terms to evaluate (order = rows of the result); cfg: dict with the cfg.rewards scalars the terms read
    (tracking_sigma, tracking_sigma_yaw, base_height_target, gait_force_sigma, gait_vel_sigma, max_contact_force)."""

    _PTRS = [n for n, _ in _abi.lt_corl_ptrs._fields_ if n not in ("out",)]

    def __init__(self, names, cfg, dt, feet_indices, penalised_contact_indices, device="cuda:0"):
        self.device = _need_cuda(device, "CoRLRewardKernel")
        self.lib = _abi.load()
        unknown = [n for n in names if n not in CORL_TERMS]
        if unknown:
            raise KeyError(f"not a CoRLRewards term: {unknown}")
        self.names = list(names)
        c = _abi.lt_corl_cfg()
        c.n_terms = len(self.names)
        for i, n in enumerate(self.names):
            c.term_ids[i] = _abi.DEFINES["LT_CORL_" + n.upper()]
        if len(feet_indices) != 4:
            raise ValueError("CoRLRewards assumes four feet (corl_rewards.py:71, 81)")
        for i, b in enumerate(feet_indices):
            c.feet_indices[i] = int(b)
        c.n_penalised = len(penalised_contact_indices)
        for i, b in enumerate(penalised_contact_indices):
            c.penalised[i] = int(b)
        c.dt = float(dt)
        for k in ("tracking_sigma", "tracking_sigma_yaw", "base_height_target", "gait_force_sigma", "gait_vel_sigma", "max_contact_force"):
            setattr(c, k, float(cfg[k]))
        self.cfg = c
        self.out = None

    def __call__(self, env, stream=None):
        """env: mapping (or object) with the tensors CoRLRewards reads, under the reference's attribute names.  -> [n_terms, N]"""
        get = env.__getitem__ if isinstance(env, dict) else (lambda k: getattr(env, k))
        cmd = get("commands")
        N = int(cmd.shape[0])
        c, p = self.cfg, _abi.lt_corl_ptrs()
        c.num_envs, c.num_commands = N, int(cmd.shape[1])
        c.num_bodies = int(get("contact_forces").shape[1])
        keep = []
        c.dof_stride = 1
        dof_state = env.get("dof_state") if isinstance(env, dict) else getattr(env, "dof_state", None)
        for name in self._PTRS:
            if name in ("dof_pos", "dof_vel") and dof_state is not None:
                # Isaac Gym's interleaved [N,12,2] tensor read in place: no contiguous copies of its two columns
                if dof_state.dtype != torch.float32 or dof_state.device != self.device or not dof_state.is_contiguous():
                    raise ValueError(f"dof_state: expected a contiguous float32 tensor on {self.device}")
                c.dof_stride = 2
                keep.append(dof_state)
                setattr(p, name, dof_state.data_ptr() + (4 if name == "dof_vel" else 0))
                continue
            t = get(name)
            if name == "last_contacts":
                if t.dtype == torch.bool:
                    t = t.view(torch.uint8)
                ok = t.dtype == torch.uint8
            else:
                ok = t.dtype == torch.float32
            if not ok or t.device != self.device or not t.is_contiguous():
                raise ValueError(f"{name}: expected a contiguous {'bool' if name == 'last_contacts' else 'float32'} tensor on {self.device}")
            keep.append(t)
            setattr(p, name, t.data_ptr())
        if self.out is None or self.out.shape[1] != N:
            self.out = torch.empty(len(self.names), N, dtype=torch.float32, device=self.device)
        p.out = self.out.data_ptr()
        st = stream if stream is not None else torch.cuda.current_stream(self.device)
        self._keep = keep
        _abi.check(self.lib.lt_corl_rewards(C.byref(c), C.byref(p), C.c_void_p(st.cuda_stream)), None, "lt_corl_rewards")
        return self.out


GAIT_IDS = {"pronk": "LT_GAIT_PRONK", "trot": "LT_GAIT_TROT", "pace": "LT_GAIT_PACE", "bound": "LT_GAIT_BOUND"}


class CommandResampler:
    """The device half of LeggedRobot._resample_commands (legged_robot_velocity_tracking.py:776-840) as ONE launch of
    lt_resample_commands: new command rows into `commands`, the gait-category rewrite of the phase / offset / bound columns,
    binary-phase rounding, small-velocity zeroing, command_sums reset.  The curricula that produce the rows (RewardThresholdCurriculum,
    numpy) stay with the caller, as in the reference.

    category_names: the reference's list (['pronk', 'trot', 'pace', 'bound'] under gaitwise_curricula, else ['nominal']);
    commands [N, num_commands] and command_sums [n_sums, N] (or None) are the env's own tensors, updated in place."""

    def __init__(self, commands, command_sums, category_names, gaitwise_curricula=False, exclusive_phase_offset=False,
                 balance_gait_distribution=False, binary_phases=False):
        self.device = _need_cuda(commands.device, "CommandResampler")
        self.lib = _abi.load()
        ok = lambda t: t.dtype == torch.float32 and t.is_contiguous() and t.device == self.device      # noqa: E731
        if not ok(commands) or commands.dim() != 2:
            raise ValueError("commands: expected a contiguous float32 [N, num_commands] CUDA tensor")
        if command_sums is not None and (not ok(command_sums) or command_sums.dim() != 2 or command_sums.shape[1] != commands.shape[0]):
            raise ValueError("command_sums: expected a contiguous float32 [n_sums, N] tensor on the same device")
        if len(category_names) > _abi.DEFINES["LT_CMD_MAX_CATEGORIES"]:
            raise ValueError("too many command categories")
        self.commands, self.command_sums, self.category_names = commands, command_sums, list(category_names)
        c = _abi.lt_cmd_cfg()
        c.num_envs, c.num_commands = int(commands.shape[0]), int(commands.shape[1])
        c.n_sums = int(command_sums.shape[0]) if command_sums is not None else 0
        c.gaitwise_curricula, c.exclusive_phase_offset = int(bool(gaitwise_curricula)), int(bool(exclusive_phase_offset))
        c.balance_gait_distribution, c.binary_phases = int(bool(balance_gait_distribution)), int(bool(binary_phases))
        c.n_categories = len(self.category_names)
        for i, name in enumerate(self.category_names):
            c.category_gait[i] = _abi.DEFINES[GAIT_IDS[name]] if name in GAIT_IDS else _abi.DEFINES["LT_GAIT_NONE"]
        self.cfg = c

    def categories_from_uniform(self, u):
        """LRV:760-766 on the device: category i takes the envs whose draw lies in [i / n, (i + 1) / n)."""
        n = len(self.category_names)
        p = 1.0 / n
        cat = torch.full(u.shape, -1, dtype=torch.int32, device=u.device)
        for i in range(n):
            cat[torch.logical_and(p * i <= u, u < p * (i + 1))] = i
        return cat

    def apply(self, env_ids, new_commands, category=None, gait_uniform=None, stream=None):
        """env_ids int64 [K]; new_commands [K, >= num_commands] (what curriculum.sample returned, any float dtype, host or device);
        category int32 [K] (needed under gaitwise_curricula); gait_uniform float32 [K] (the second torch.rand draw of the
        exclusive_phase_offset / balance_gait_distribution branches)."""
        K = int(env_ids.numel())
        if K == 0:
            return
        dev = self.device
        ids = env_ids.to(device=dev, dtype=torch.int64).contiguous()
        new = torch.as_tensor(new_commands)[:, :self.cfg.num_commands].to(device=dev, dtype=torch.float32).contiguous()
        if new.shape[0] != K:
            raise ValueError("new_commands: one row per resampled env")
        cat = None if category is None else torch.as_tensor(category).to(device=dev, dtype=torch.int32).contiguous()
        gu = None if gait_uniform is None else torch.as_tensor(gait_uniform).to(device=dev, dtype=torch.float32).contiguous()
        p = _abi.lt_cmd_ptrs()
        p.env_ids, p.new_commands = ids.data_ptr(), new.data_ptr()
        p.category = cat.data_ptr() if cat is not None else None
        p.gait_uniform = gu.data_ptr() if gu is not None else None
        p.commands = self.commands.data_ptr()
        p.command_sums = self.command_sums.data_ptr() if self.command_sums is not None else None
        st = stream if stream is not None else torch.cuda.current_stream(dev)
        self._keep = (ids, new, cat, gu)
        _abi.check(self.lib.lt_resample_commands(C.byref(self.cfg), C.byref(p), K, C.c_void_p(st.cuda_stream)), None, "lt_resample_commands")
