"""TEST INFRASTRUCTURE (build container only): record tunnel tile tables from the UNMODIFIED reference generator.

  python -m oracle.gen_golden_terrain          -> tests/golden/terrain_single_path_*.npz

Runs go1_gym.utils.tunnel.Terrain (tunnel.py:51-217) with terrain_type single_path (tunnel_fn.py:99-163) exactly as
scripts/train.py configures it (np.random.seed first, scripts/train.py:33-40), and stores what
LeggedRobot._get_env_origins derives from it (LR:1827-1840): the per-tile height table in metres as float32, the env
origins and the terrain origins.  The product's device-side builder (legged_tracking_b200/terrain.py + lt_build_tunnel_tiles)
and the oracle restatement (oracle/terrain_port.py) are both checked against these files.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, REPO)

from oracle import ref_harness  # noqa: E402

ARGV = ["--camera_zero", "--strategy", "e2e", "--terrain", "single_path", "--measure_front_half", "--penalty_scaler", "1.0"]
CASES = {"terrain_single_path_4x4_seed11": dict(rows=4, cols=4, seed=11, overrides={}),
         "terrain_single_path_6x5_seed5": dict(rows=6, cols=5, seed=5, overrides={"terrain.p_double": 0.6, "terrain.p_flat": 0.5})}


def terrain_cfg_dict(t):
    keys = ["terrain_length", "terrain_width", "horizontal_scale", "vertical_scale", "num_rows", "num_cols", "terrain_ratio_x",
            "terrain_ratio_y", "start_loc", "ceiling_height", "p_flat", "p_double", "terrain_type"]
    return {k: getattr(t, k) for k in keys}


def record(name, rows, cols, seed, overrides):
    Cfg, _ = ref_harness.build_cfg(ARGV, num_envs=rows * cols, num_rows=rows, num_cols=cols, overrides=overrides)
    from go1_gym.utils.tunnel import Terrain
    np.random.seed(seed)
    t = Terrain(Cfg.terrain, rows * cols)
    # LR:1827-1840
    tiles = torch.from_numpy(t.height_samples_by_row_col * t.vertical_scale).to(torch.float).numpy()
    L, W = t.length_per_env_pixels, t.width_per_env_pixels
    out = {"tile_heights": tiles.reshape(rows * cols, 2, L, W),
           "env_origins": torch.from_numpy(t.env_origins).to(torch.float).numpy().reshape(rows * cols, 3),
           "env_terrain_origin": torch.from_numpy(t.all_terrain_origins).to(torch.float).numpy().reshape(rows * cols, 3),
           "seed": np.int64(seed)}
    import json
    out["cfg"] = np.array(json.dumps(terrain_cfg_dict(Cfg.terrain)))
    path = os.path.join(REPO, "tests", "golden", name + ".npz")
    np.savez_compressed(path, **out)
    print(name, out["tile_heights"].shape, os.path.getsize(path), "bytes; dtype of the reference map:", t.height_field_raw.dtype)


if __name__ == "__main__":
    for name, kw in CASES.items():
        record(name, **kw)
