"""TEST SCAFFOLDING: drives tests/hostsim/libhostsim.so (lt_core.cuh compiled for the host) on numpy buffers.

Mirrors legged_tracking_b200.engine.StepEngine call for call so that the per-env arithmetic shared with the
CUDA kernels can be checked against the golden fixtures without a GPU.  Never imported by the product.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import torch

from legged_tracking_b200 import _abi, cconfig
from legged_tracking_b200.engine import GravitySchedule, ExplorationCurriculum

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "hostsim", "hostsim.cpp")
LIB = os.path.join(HERE, "hostsim", "libhostsim.so")
_NP = {"f32": np.float32, "i64": np.int64, "bool": np.uint8, "i32": np.int32}


def build():
    deps = [SRC, os.path.join(_abi.PKG_DIR, "csrc", "lt_core.cuh"), _abi.HEADER]
    if os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in deps):
        return LIB
    cmd = ["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", LIB, SRC]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
    return LIB


class HostSim:
    def __init__(self, flat, terrain, actuator, seed=0):
        self.lib = C.CDLL(build())
        self.flat = flat
        self.N = N = int(flat["num_envs"])
        self.cfg = cconfig.make_lt_config(flat, seed=seed)
        assert self.lib.lths_config_bytes() == C.sizeof(_abi.lt_config)
        self.tables = cconfig.Tables(flat, actuator)
        A, B, D = int(flat["num_actors"]), int(flat["num_bodies"]), int(flat["num_dof"])
        self.root_states = np.zeros((N * A, 13), np.float32)
        self.root_states[:, 6] = 1.0
        self.dof_state = np.zeros((N * D, 2), np.float32)
        self.contact_forces = np.zeros((N, B, 3), np.float32)
        self.rigid_body_state = np.zeros((N * B, 13), np.float32)
        self.sim = _abi.lt_sim_ptrs(self.root_states.ctypes.data, self.dof_state.ctypes.data, self.contact_forces.ctypes.data,
                                    self.rigid_body_state.ctypes.data)
        self.ter_arrays = {"env_origins": np.ascontiguousarray(terrain["env_origins"], np.float32).reshape(N, 3)}
        self.ter = _abi.lt_terrain_ptrs()
        self.ter.env_origins = self.ter_arrays["env_origins"].ctypes.data
        if flat["mesh_type"] != "plane":
            self.ter_arrays["tile_heights"] = np.ascontiguousarray(terrain["tile_heights"], np.float32)
            self.ter_arrays["env_tile"] = np.ascontiguousarray(terrain["env_tile"], np.int32)
            self.ter_arrays["env_terrain_origin"] = np.ascontiguousarray(terrain["env_terrain_origin"], np.float32).reshape(N, 3)
            for k in ("tile_heights", "env_tile", "env_terrain_origin"):
                setattr(self.ter, k, self.ter_arrays[k].ctypes.data)
        self.state = {}
        self.ptrs = _abi.lt_state_ptrs()
        for name, dt, shape in cconfig.state_spec(flat):
            buf = np.full(shape, cconfig.STATE_INIT.get(name, 0), dtype=_NP[dt])
            self.state[name] = buf
            setattr(self.ptrs, name, buf.ctypes.data)
        self.state["gravity_vec"][:] = [0.0, 0.0, -1.0]
        self.counters = np.zeros(4, np.int32)
        self.gravity = GravitySchedule(flat)
        self.curriculum = ExplorationCurriculum(flat)
        self.reward_names = list(flat["reward_names"])
        self.sum_names = self.reward_names + ["total", "total_pos", "total_neg"]
        self.n_slots = cconfig.rng_n_slots(flat)
        self.pending = 0

    def compute_torques(self, actions, substep):
        a = np.ascontiguousarray(actions, np.float32)
        rc = self.lib.lths_torques(C.byref(self.cfg), C.byref(self.tables.c), C.byref(self.ptrs),
                                   C.c_void_p(self.dof_state.ctypes.data), C.c_void_p(a.ctypes.data), int(substep),
                                   C.c_void_p(self.counters.ctypes.data))
        assert rc == 0
        self.pending += 1
        return self.state["torques"]

    def post_physics_step(self, uniforms=None, noise_uniforms=None, extra_reward=None, gravity_uniform=None):
        args = _abi.lt_step_args()
        keep = []
        for name, arr in (("uniforms", uniforms), ("noise_uniforms", noise_uniforms), ("extra_reward", extra_reward)):
            if arr is not None:
                a = np.ascontiguousarray(arr, np.float32)
                keep.append(a)
                setattr(args, name, a.ctypes.data)
        draw = (lambda: gravity_uniform) if gravity_uniform is not None else (lambda: torch.rand(3))
        new_g = self.gravity.advance(draw)
        if new_g is not None:
            g = np.ascontiguousarray(new_g.numpy(), np.float32)
            keep.append(g)
            args.new_gravity = g.ctypes.data
        rc = self.lib.lths_post_physics(C.byref(self.cfg), C.byref(self.tables.c), C.byref(self.ptrs), C.byref(self.sim),
                                        C.byref(self.ter), C.byref(args), C.c_void_p(self.counters.ctypes.data), self.pending)
        assert rc == 0
        self.pending = 0
        if new_g is not None:
            self.state["gravity_vec"][:] = GravitySchedule.gravity_vec_of(new_g).numpy()
        if self.curriculum.advance(self.gravity.common_step_counter):
            for t, v in enumerate(self.curriculum.term_scales()):
                self.cfg.term_scales[t] = float(v)

    def reset_env_ids(self):
        return self.state["reset_env_ids"][: int(self.state["reset_count"][0])]

    def lag_buffer_logical(self):
        head = int(self.counters[2])
        L = self.state["lag_buffer"].shape[0]
        return self.state["lag_buffer"][[(head + j) % L for j in range(L)]]

    def load_state(self, st):
        def put(dst, src):
            dst[...] = np.asarray(src).astype(dst.dtype).reshape(dst.shape)
        for name, buf in self.state.items():
            if name in st and name not in ("lag_buffer", "episode_sums", "command_sums"):
                put(buf, st[name])
        if "root_states" in st:
            put(self.root_states, st["root_states"])
        if "dof_state" in st:
            put(self.dof_state, st["dof_state"])
        if "lag_buffer" in st:
            put(self.state["lag_buffer"], st["lag_buffer"])
        for r, name in enumerate(self.sum_names):
            if "episode_sums/" + name in st:
                put(self.state["episode_sums"][r], st["episode_sums/" + name])
        for r, name in enumerate(self.reward_names + ["ep_timesteps"]):
            if "command_sums/" + name in st:
                put(self.state["command_sums"][r], st["command_sums/" + name])
        if "gravity_vec" in st:
            put(self.state["gravity_vec"], np.asarray(st["gravity_vec"]).reshape(-1)[:3])
        if "gravities" in st:
            self.gravity.gravities = torch.as_tensor(np.asarray(st["gravities"]), dtype=torch.float32).reshape(-1, 3)[0].clone()
        if "common_step_counter" in st:
            self.gravity.common_step_counter = int(st["common_step_counter"])

    def export_state(self):
        out = {k: torch.as_tensor(v.copy()) for k, v in self.state.items()
               if k not in ("lag_buffer", "episode_sums", "command_sums", "finished_sums")}
        out["root_states"] = torch.as_tensor(self.root_states.copy())
        out["dof_state"] = torch.as_tensor(self.dof_state.copy())
        out["lag_buffer"] = torch.as_tensor(self.lag_buffer_logical().copy())
        for r, name in enumerate(self.sum_names):
            out["episode_sums/" + name] = torch.as_tensor(self.state["episode_sums"][r].copy())
        for r, name in enumerate(self.reward_names + ["ep_timesteps"]):
            out["command_sums/" + name] = torch.as_tensor(self.state["command_sums"][r].copy())
        out["common_step_counter"] = torch.tensor(self.gravity.common_step_counter)
        return out
