"""Synthetic workloads of BASELINE.json's configs at full size, for bench.py / scale tests / smoke().

PhysX and the terrain generator are outside the hot path, so a workload is: a flat config (the reference's own,
recorded off the live reference env into tests/golden/*.npz by oracle/gen_golden.py), a tile table of the
`single_path` shape ([tiles, 2, 80, 40] float32 metres: layer 0 ceiling, layer 1 floor) and seeded simulator state
(synthetic.SyntheticState).  The tile table here is SYNTHETIC: the 16 tiles the reference's generator produced for
the fixture are cycled over the 32x32 grid with a seeded rescaling of every obstacle so that all 1024 tiles differ;
origins follow the reference's grid (tile (r, c) at (4 r, 2 c) m; robot spawn at +(0.72, 1.0), LR:1808-1847).
"""
import json
import os

import numpy as np

from . import synthetic

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(REPO, "tests", "golden")

# BASELINE.json configs -> (fixture whose flat config is used, num_envs, tile grid)
CONFIGS = {
    "cfg1": ("cfg1_plane_256", 256, None),
    "cfg2": ("cfg2_single_path_64", 4096, (32, 32)),
    "cfg3": ("cfg3_vel_64", 16384, (32, 32)),
    "cfg5": ("cfg2_single_path_64", 65536, (32, 32)),
}


def load_fixture_flat(fixture):
    z = np.load(os.path.join(GOLDEN, fixture + ".npz"), allow_pickle=False)
    flat = json.loads(str(z["flat"]))
    terrain = {k[len("terrain/"):]: z[k] for k in z.files if k.startswith("terrain/")}
    return flat, terrain


def load_actuator():
    """Weights of the reference's actuator network (resources/actuator_nets/unitree_go1.pt, LR:1307-1309) as plain arrays:
    w0 [32,6] b0 [32] w1 [32,32] b1 [32] w2 [1,32] b2 [1] (extracted by oracle/gen_golden.actuator_from_reference)."""
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "actuator_net_go1.npz"))
    return {k: z[k] for k in z.files}


def synth_terrain(base_terrain, num_envs, rows, cols, seed=11):
    """Tile table + per-env origins for a rows x cols grid; env i lives on tile i % (rows*cols) (LR:1816-1823)."""
    rng = np.random.RandomState(seed)
    base = np.asarray(base_terrain["tile_heights"], np.float32)           # [16,2,R,C]
    n_tiles = rows * cols
    R, C = base.shape[2], base.shape[3]
    tiles = base[np.arange(n_tiles) % base.shape[0]].copy()
    ceil_ref = tiles[:, 0].max(axis=(1, 2), keepdims=True)                # free ceiling height of each tile
    k = rng.uniform(0.5, 1.0, size=(n_tiles, 1, 1)).astype(np.float32)
    tiles[:, 0] = ceil_ref - (ceil_ref - tiles[:, 0]) * k                 # shrink every hanging obstacle
    tiles[:, 1] = tiles[:, 1] * rng.uniform(0.5, 1.0, size=(n_tiles, 1, 1)).astype(np.float32)
    length = R * 0.05
    width = C * 0.05
    tr, tc = np.divmod(np.arange(n_tiles), cols)
    t_origin = np.stack([tr * length, tc * width, np.zeros(n_tiles)], axis=1).astype(np.float32)
    spawn = np.asarray(base_terrain["env_origins"][0] - base_terrain["env_terrain_origin"][0], np.float32)
    env_tile = (np.arange(num_envs) % n_tiles).astype(np.int32)
    return {"tile_heights": np.ascontiguousarray(tiles), "env_tile": env_tile,
            "env_terrain_origin": np.ascontiguousarray(t_origin[env_tile]),
            "env_origins": np.ascontiguousarray(t_origin[env_tile] + spawn[None, :])}


def make(config="cfg5", num_envs=None, seed=1000, tile_grid=None):
    """-> (flat, terrain, actuator, SyntheticState)"""
    fixture, n_default, grid = CONFIGS[config]
    flat, base_terrain = load_fixture_flat(fixture)
    n = int(num_envs or n_default)
    flat = dict(flat)
    flat["num_envs"] = n
    grid = tile_grid or grid
    if flat["mesh_type"] == "plane":
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
        flat["terrain_num_rows"], flat["terrain_num_cols"] = rows, cols
        terrain = synth_terrain(base_terrain, n, rows, cols)
    synth = synthetic.SyntheticState(n, terrain["env_origins"], num_bodies=flat["num_bodies"], seed=seed,
                                     feet_indices=flat["feet_indices"])
    return flat, terrain, load_actuator(), synth


# ---------------------------------------------------------------------------------------------------------------
# Algorithmic traffic per env (SURVEY.md section 8d accounting), regenerated from the configuration.
def algorithmic_bytes(flat):
    nd = flat["num_dof"]
    n_terms = len(flat["reward_names"])
    nx, ny = len(flat["measured_points_x"]), len(flat["measured_points_y"])
    x0 = nx // 2 + 1 if flat["measure_front_half"] else 0
    n_heights = 2 * (nx - x0) * ny if flat["observe_heights"] else 0
    H = flat["obs_history_length"]
    reads = {"root row": 52, "dof_state": 8 * nd, "contact (penalised bodies)": 12 * len(flat["penalised_contact_indices"]),
             "actions": 4 * nd, "last_actions": 4 * nd, "last_dof_vel": 4 * nd, "torques": 4 * nd, "trajectory row": 24,
             "curr_pose_index": 8, "episode_length_buf": 8, "terrain origin + tile id": 16, "previous pitch": 4,
             "privileged sources": 4 * flat["num_privileged_obs"], "episode_sums": 4 * (n_terms + 3),
             "command_sums": 4 * (n_terms + 5), "collision_count": 8, "heights": 4 * n_heights}
    writes = {"obs_buf": 4 * flat["num_obs"], "privileged_obs_buf": 4 * flat["num_privileged_obs"],
              "obs_history": 4 * flat["num_obs"] * H, "rew_buf": 4, "masks": 5, "episode_length_buf": 8,
              "history rows (5 x 12)": 5 * 4 * nd, "last_root_vel": 24, "base_lin_vel/ang_vel/projected_gravity": 36,
              "base_rotation": 12, "relative_linear/rotation": 24, "episode_sums": 4 * (n_terms + 3),
              "command_sums": 4 * (n_terms + 5), "curr_pose_index": 8, "collision_count": 8, "arrow root row": 28}
    if H > 1:
        reads["obs_history shift"] = 4 * flat["num_obs"] * (H - 1)
    post = sum(reads.values()) + sum(writes.values())
    torque = (8 * nd + 4 * nd + 4 * nd + 8 * nd + 16 * nd) + (4 * nd + 4 * nd + 16 * nd + 4 * nd)
    return {"post_physics": post, "post_physics_reads": reads, "post_physics_writes": writes, "torque_substep": torque,
            "full_step": post + flat["decimation"] * torque,
            "torque_flops_substep": nd * 2 * (6 * 32 + 32 * 32 + 32)}
