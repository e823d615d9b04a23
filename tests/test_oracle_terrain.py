"""CPU: the oracle restatement of the tunnel terrain generator (oracle/terrain_port.py) against the tables recorded from the
unmodified reference (tests/golden/terrain_single_path_*.npz, written by oracle/gen_golden_terrain.py), and the HOST half of the
product's builder (legged_tracking_b200.terrain.draw_single_path: the random draws in the reference's order).  Bit-exact."""
import glob
import json
import os

import numpy as np
import pytest

from oracle import terrain_port

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAMES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "terrain_single_path_*.npz")))


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return z, json.loads(str(z["cfg"])), int(z["seed"])


@pytest.mark.parametrize("name", NAMES)
def test_port_reproduces_reference_tables(name):
    z, cfg, seed = load(name)
    out = terrain_port.build(cfg, np.random.RandomState(seed))
    for k in ("tile_heights", "env_origins", "env_terrain_origin"):
        assert out[k].dtype == z[k].dtype and np.array_equal(out[k], z[k]), k


@pytest.mark.parametrize("name", NAMES)
def test_product_draws_follow_the_reference_order(name):
    """draw_single_path consumes the generator exactly like the reference (same stream position afterwards), and its wedges,
    pushed through the oracle's plane evaluation, give the reference's patches."""
    from legged_tracking_b200 import terrain
    z, cfg, seed = load(name)
    r1, r2 = np.random.RandomState(seed), np.random.RandomState(seed)
    n_pyr, wedges = terrain.draw_single_path(cfg, r1)
    terrain_port.build(cfg, r2)
    assert r1.uniform() == r2.uniform()
    hs, vs = cfg["horizontal_scale"], cfg["vertical_scale"]
    Lp, Wp = int(cfg["terrain_length"] / hs), int(cfg["terrain_width"] / hs)
    sub_rows, sub_cols = int(Lp * cfg["terrain_ratio_x"]), int(Wp * cfg["terrain_ratio_y"])
    w, l = sub_rows * hs, sub_cols * hs
    xy = np.stack(np.meshgrid(np.linspace(-w / 2, w / 2, sub_rows), np.linspace(-l / 2, l / 2, sub_cols)), axis=-1)
    r0, c0 = int(round((0.5 - cfg["terrain_ratio_x"] / 2) * Lp, 4)), int((0.5 - cfg["terrain_ratio_y"] / 2) * Wp)
    for k in range(n_pyr.shape[0]):
        n = n_pyr[k, 1]                                        # floor layer
        m, pw, pl = wedges[k, 1, :n, :3], wedges[k, 1, :n, 3], wedges[k, 1, :n, 4]
        zc = np.zeros_like(pw)
        corners = np.stack([np.stack([s0 * pw + m[:, 0], s1 * pl + m[:, 1], zc], 1) for s0, s1 in ((1, 1), (-1, 1), (-1, -1), (1, -1))]
                           + [m], axis=1)
        faces = corners[:, [[0, 1, -1], [1, 2, -1], [2, 3, -1], [3, 0, -1]], :]
        h = terrain_port.planes_heights(faces[:, :, 0], faces[:, :, 1], faces[:, :, 2], xy)
        h[0, :] = h[-1, :] = 0.5
        h[:, 0] = h[:, -1] = 0.5
        want = z["tile_heights"][k, 1, r0:r0 + sub_rows, c0:c0 + sub_cols]
        got = ((h / vs).astype(int).T.astype(np.float64) * vs).astype(np.float32)
        assert np.array_equal(got, want), k
