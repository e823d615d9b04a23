"""Synthetic workloads of BASELINE.json's configs at full size, for bench.py / scale tests / smoke().

PhysX is outside the hot path, so a workload is: the flat config of one of the reference's argument sets
(cfg.flat_from_cfg(cfg.preset_cfg(...)): the reference's own defaults + what scripts/train.py changes for those flags), the
`single_path` tile table ([tiles, 2, 80, 40] float32 metres: layer 0 ceiling, layer 1 floor) built on the device by
terrain.TunnelTerrain from the reference's random stream (np.random seed 11 like scripts/train.py:33-40; every tile
differs), and seeded simulator state (synthetic.SyntheticState).  Nothing here reads tests/.
"""
import os

import numpy as np

from . import cfg as ltcfg, synthetic

# BASELINE.json configs -> (argument-set preset of cfg.py, num_envs, tile grid)
CONFIGS = {
    "cfg1": ("cfg1_plane_256", 256, None),
    "cfg2": ("cfg2_single_path_64", 4096, (32, 32)),
    "cfg3": ("cfg3_vel_64", 16384, (32, 32)),
    "cfg5": ("cfg2_single_path_64", 65536, (32, 32)),
}


def make_cfg4(num_envs=4096, device="cuda:0", seed=1000, command_source="device", heightfield=True):
    """BASELINE.json configs[3]: the velocity-tracking env (vel_env.VelocityTrackingEasyEnv) on a synthetic simulator backend and,
    for the 17 x 11 height grid the config names, a synthetic int16 heightfield under the env grid.  -> the env (call reset())."""
    from . import vel_env
    n = int(num_envs)
    side = int(np.floor(np.sqrt(n)))
    rows = int(np.ceil(n / side))
    xx, yy = np.meshgrid(np.arange(rows), np.arange(side), indexing="ij")
    origins = np.zeros((n, 3), np.float32)
    origins[:, 0], origins[:, 1] = 3.0 * xx.flatten()[:n], 3.0 * yy.flatten()[:n]
    field = None
    if heightfield:
        rs = np.random.RandomState(5)
        field = dict(height_samples=rs.randint(-10, 10, size=(int(3.0 * rows / 0.1) + 64, int(3.0 * side / 0.1) + 64)).astype(np.int16),
                     border_size=2.0, horizontal_scale=0.1, vertical_scale=0.005, measured_points_x=vel_env.CFG4_MEASURED_POINTS_X,
                     measured_points_y=vel_env.CFG4_MEASURED_POINTS_Y)
    return vel_env.VelocityTrackingEasyEnv(device, True, num_envs=n, sim=vel_env.VelReplaySim(seed=seed), command_source=command_source,
                                           env_origins=origins, heightfield=field, seed=seed)


def load_actuator():
    """Weights of the reference's actuator network (resources/actuator_nets/unitree_go1.pt, LR:1307-1309) as plain arrays:
    w0 [32,6] b0 [32] w1 [32,32] b1 [32] w2 [1,32] b2 [1] (extracted by oracle/gen_golden.actuator_from_reference)."""
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "actuator_net_go1.npz"))
    return {k: z[k] for k in z.files}


_TABLES = {}


def tunnel_table(terrain_cfg, rows, cols, device="cuda:0", seed=11):
    """Host copy of the device-built tile table + tile origins for a rows x cols grid (cached: the bench builds several engines)."""
    import torch
    from .terrain import TunnelTerrain
    key = (rows, cols, seed, str(device))
    if key not in _TABLES:
        terrain_cfg.num_rows, terrain_cfg.num_cols = rows, cols
        t = TunnelTerrain(terrain_cfg, device=device, rng=np.random.RandomState(seed))
        torch.cuda.synchronize()
        _TABLES[key] = (t.tile_heights.cpu().numpy(), t.env_origins.reshape(rows * cols, 3).astype(np.float32),
                        t.all_terrain_origins.reshape(rows * cols, 3).astype(np.float32))
    return _TABLES[key]


def make(config="cfg5", num_envs=None, seed=1000, tile_grid=None, device="cuda:0"):
    """-> (flat, terrain, actuator, SyntheticState).  Needs the CUDA device for heightfield configs (the tile table is built by
    lt_build_tunnel_tiles; there is no CPU generator in the product)."""
    preset, n_default, grid = CONFIGS[config]
    n = int(num_envs or n_default)
    c = ltcfg.preset_cfg(preset, num_envs=n)
    grid = tile_grid or grid
    if c.terrain.mesh_type == "plane":
        flat = ltcfg.flat_from_cfg(c, num_envs=n)
        side = int(np.floor(np.sqrt(n)))
        rows_ = int(np.ceil(n / side))
        xx, yy = np.meshgrid(np.arange(rows_), np.arange(side), indexing="ij")
        origins = np.zeros((n, 3), np.float32)
        origins[:, 0] = 3.0 * xx.flatten()[:n]                            # env_spacing 3.0 (LR:1850-1858)
        origins[:, 1] = 3.0 * yy.flatten()[:n]
        terrain = {"env_origins": origins}
    else:
        rows, cols = grid
        while n % (rows * cols):                                          # LR:1819 requires divisibility
            rows, cols = max(rows // 2, 1), max(cols // 2, 1)
        c.terrain.num_rows, c.terrain.num_cols = rows, cols
        flat = ltcfg.flat_from_cfg(c, num_envs=n)
        tiles, env_o, tile_o = tunnel_table(c.terrain, rows, cols, device=device)
        env_tile = (np.arange(n) % (rows * cols)).astype(np.int32)       # LR:1816-1823
        terrain = {"tile_heights": tiles, "env_tile": env_tile, "env_terrain_origin": np.ascontiguousarray(tile_o[env_tile]),
                   "env_origins": np.ascontiguousarray(env_o[env_tile])}
    synth = synthetic.SyntheticState(n, terrain["env_origins"], num_bodies=flat["num_bodies"], seed=seed,
                                     feet_indices=flat["feet_indices"])
    return flat, terrain, load_actuator(), synth


# ---------------------------------------------------------------------------------------------------------------
# Algorithmic traffic per env (SURVEY.md section 8d accounting), regenerated from the configuration.
def algorithmic_bytes(flat):
    """Per-env bytes of one step, derived from the real buffer list (cconfig.state_spec) + cconfig.POST_PHYSICS_TRAFFIC, plus
    the simulator tensors of the input ABI.  At cfg 2 / 3 / 5 this reproduces SURVEY.md section 8d: 4,121 B post-physics (of
    which 96 B, the joint-target history shift, are executed by the torque kernel's last sub-step), 816 B per torque sub-step."""
    from . import cconfig
    itemsize = {"f32": 4, "i64": 8, "i32": 4, "bool": 1}
    n = flat["num_envs"]
    per_env = {}
    for name, dt, shape in cconfig.state_spec(flat):
        total = itemsize[dt]
        for d in shape:
            total *= d
        per_env[name] = total // n if n in shape else None
    nd = flat["num_dof"]
    nx, ny = len(flat["measured_points_x"]), len(flat["measured_points_y"])
    x0 = nx // 2 + 1 if flat["measure_front_half"] else 0
    n_heights = 2 * (nx - x0) * ny if flat["observe_heights"] else 0
    H = flat["obs_history_length"]
    # simulator tensors (a1) and tables: not in lt_state_ptrs
    reads = {"root row": 52, "dof_state": 8 * nd, "contact (penalised bodies)": 12 * len(flat["penalised_contact_indices"]),
             "trajectory row": 24, "terrain origin + tile id": 16, "heights": 4 * n_heights}
    writes = {"arrow root row": 28}
    torque_shift = 0
    for name, role in cconfig.POST_PHYSICS_TRAFFIC.items():
        b = per_env[name]                       # KeyError / None here = the table went out of sync with state_spec
        if b is None:
            raise KeyError(name)
        if name == "command_sums":
            b += 16                             # SURVEY counts 15 rows: the reference keeps 4 more per-command sums than the fused path
        if role == "rw_row1":                   # base_rotation: the previous pitch is read (4 B), the row is written
            reads[name] = 4
            writes[name] = b
        elif role == "torque_w":
            torque_shift += b
        else:
            if "r" in role:
                reads[name] = b
            if "w" in role:
                writes[name] = b
    if H > 1:
        reads["obs_history shift"] = 4 * flat["num_obs"] * (H - 1)
    post = sum(reads.values()) + sum(writes.values()) + torque_shift
    torque = (8 * nd + 4 * nd + 4 * nd + 8 * nd + 16 * nd) + (4 * nd + 4 * nd + 16 * nd + 4 * nd)
    return {"post_physics": post, "post_physics_reads": reads, "post_physics_writes": writes,
            "post_physics_in_torque_kernel": torque_shift, "torque_substep": torque,
            "full_step": post + flat["decimation"] * torque,
            "torque_flops_substep": nd * 2 * (6 * 32 + 32 * 32 + 32)}
