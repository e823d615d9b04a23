"""TEST INFRASTRUCTURE: CPU (numpy, float64) restatement of the reference's tunnel terrain generator for
`terrain_type = single_path` -- the checker of the device-side tile-table builder.

Follows go1_gym/utils/tunnel.py:51-217 (Terrain.__init__, make_terrain, add_terrain_to_map),
go1_gym/utils/tunnel_fn.py:3-21 (vec_plane_from_points) and :99-163 (TerrainFunctions.single_path), and the conversion
LeggedRobot._get_env_origins applies (legged_robot_trajectory_tracking.py:1827-1840).  Pinned against
tests/golden/terrain_single_path_*.npz, which oracle/gen_golden_terrain.py recorded from the unmodified reference
(tests/test_oracle_terrain.py).  Only tests/ import this file; the product never does.

The random draws come from numpy's GLOBAL generator in the reference (scripts/train.py:33-40 seeds it); `rng` here is any
object with the same `uniform` method (np.random itself, or a RandomState).
"""
import numpy as np


def planes_heights(p1, p2, p3, xy):
    """tunnel_fn.py:3-21.  p*: [n_pyr, 4, 3]; xy: [px, py, 2] -> [px, py] = max over pyramids of min over the 4 faces."""
    px, py = xy.shape[:2]
    flat = xy.reshape(-1, 2)
    v1, v2 = p3 - p1, p3 - p2
    cp = np.cross(v1, v2)
    a, b, c = cp[..., 0], cp[..., 1], cp[..., 2]
    d = np.sum(cp * p3, axis=-1)
    h = np.clip((d / c)[..., None] - (a / c)[..., None] * flat[..., 0] - (b / c)[..., None] * flat[..., 1], 0, a_max=np.inf)
    return np.max(np.min(h, axis=1), axis=0).reshape(px, py)


def single_path_raw(shape, horizontal_scale, vertical_scale, p_flat, p_double, top, rng):
    """tunnel_fn.py:99-163 -> integer height field (units of vertical_scale), shape = (pixel_x, pixel_y)."""
    pixel_x, pixel_y = shape
    l, w = pixel_x * horizontal_scale, pixel_y * horizontal_scale
    p1 = rng.uniform()
    p2 = rng.uniform()
    num_y = 2 if p2 < p_double else 1
    if top:
        offset_y = rng.uniform(-0.6, 0.6, size=(num_y, 1))
        offset_x = rng.uniform(-0.3, 0.3, size=(num_y, 1))
        height_max, height_min = (0.4, 0.7) if p1 < p_flat else (0.0, 0.0)
        lw_low, lw_high = 0.2, 0.4
    else:
        offset_y = rng.uniform(-0.4, 0.4, size=(num_y, 1))
        offset_x = rng.uniform(-0.2, 0.2, size=(num_y, 1))
        height_max, height_min = (0.15, 0.3) if p1 < p_flat else (0.0, 0.0)
        lw_low, lw_high = 0.1, 0.3
    mean_x = np.linspace(-w / 2, w / 2, 3)[1:-1]
    mean_y = np.zeros(num_y)
    mean_x, mean_y = np.meshgrid(mean_x, mean_y)
    mean_y += offset_y
    mean_x += offset_x
    mean_z = rng.uniform(mean_x) * (height_max - height_min) + height_min          # uniform(low=mean_x, high=1.0)
    means = np.stack([mean_x.flatten(), mean_y.flatten(), mean_z.flatten()], axis=1)
    pw, pl = rng.uniform(low=lw_low, high=lw_high, size=(2, means.shape[0]))
    z = np.zeros_like(pw)
    corners = np.stack([np.stack([pw + means[:, 0], pl + means[:, 1], z], axis=1),
                        np.stack([-pw + means[:, 0], pl + means[:, 1], z], axis=1),
                        np.stack([-pw + means[:, 0], -pl + means[:, 1], z], axis=1),
                        np.stack([pw + means[:, 0], -pl + means[:, 1], z], axis=1),
                        means], axis=1)                                             # [n_pyr, 5, 3]
    faces = corners[:, [[0, 1, -1], [1, 2, -1], [2, 3, -1], [3, 0, -1]], :]           # [n_pyr, 4 faces, 3 points, 3]
    xy = np.stack(np.meshgrid(np.linspace(-w / 2, w / 2, pixel_y), np.linspace(-l / 2, l / 2, pixel_x)), axis=-1)
    h = planes_heights(faces[:, :, 0, :], faces[:, :, 1, :], faces[:, :, 2, :], xy)
    if not top:
        h[0, :] = 0.5
        h[-1, :] = 0.5
        h[:, 0] = 0.5
        h[:, -1] = 0.5
    return (h / vertical_scale).astype(int)


def build(cfg, rng=np.random):
    """tunnel.py:51-217 for single_path + LR:1827-1840.  cfg: dict with the reference's Cfg.terrain fields.
    -> dict(tile_heights [tiles,2,L,W] f32 metres, env_origins [tiles,3] f32, env_terrain_origin [tiles,3] f32)."""
    rows, cols = int(cfg["num_rows"]), int(cfg["num_cols"])
    hs, vs = cfg["horizontal_scale"], cfg["vertical_scale"]
    env_length, env_width = cfg["terrain_length"], cfg["terrain_width"]
    rx, ry = cfg["terrain_ratio_x"], cfg["terrain_ratio_y"]
    Wp, Lp = int(env_width / hs), int(env_length / hs)
    ceiling = cfg["ceiling_height"]
    tot_rows, tot_cols = int(rows * Lp), int(cols * Wp)
    field = np.ones((2, tot_rows, tot_cols), dtype=np.int16) * int(1.0 / vs) * ceiling      # float64 (tunnel.py:81)
    field[1, :, :] = 0.5 * int(1.0 / vs)
    samples = np.zeros((rows, cols, 2, Lp, Wp))
    env_origins = np.zeros((rows, cols, 3))
    all_origins = np.zeros((rows, cols, 3))
    sub_shape = (int(Wp * ry), int(Lp * rx))                                       # SubTerrain(width, length): tunnel.py:147-151
    for k in range(rows * cols):
        i, j = np.unravel_index(k, (rows, cols))
        rng.uniform(0.0, 1.0)                                                       # difficulty (unused by single_path)
        top = single_path_raw(sub_shape, hs, vs, cfg["p_flat"], cfg["p_double"], True, rng)
        bottom = single_path_raw(sub_shape, hs, vs, cfg["p_flat"], cfg["p_double"], False, rng)
        top = np.clip(ceiling / vs - top, a_max=None, a_min=0.05 / vs)             # tunnel.py:95-97
        sx, ex = int(round((i + 0.5 - rx / 2.0) * Lp, 4)), int(round((i + 0.5 + rx / 2.0) * Lp, 4))
        sy, ey = int((j + 0.5 - ry / 2.0) * Wp), int((j + 0.5 + ry / 2.0) * Wp)
        field[0, sx:ex, sy:ey] = top.T
        field[1, sx:ex, sy:ey] = bottom.T
        samples[i, j, :] = field[:, i * Lp:(i + 1) * Lp, j * Wp:(j + 1) * Wp]
        env_origins[i, j] = [(i + 0.5 - cfg["start_loc"]) * env_length, (j + 0.5) * env_width, 0.0]
        all_origins[i, j] = [i * env_length, j * env_width, 0.0]
    return {"tile_heights": (samples * vs).astype(np.float32).reshape(rows * cols, 2, Lp, Wp),
            "env_origins": env_origins.astype(np.float32).reshape(-1, 3),
            "env_terrain_origin": all_origins.astype(np.float32).reshape(-1, 3)}
