"""Tunnel terrain for `terrain_type = single_path`, built on the device.

Host mirror of the reference's `Terrain` (go1_gym/utils/tunnel.py:51-217) for the fields the step path consumes:
`env_origins`, `all_terrain_origins`, `length_per_env_pixels`, `width_per_env_pixels`, `vertical_scale`,
`horizontal_scale`, and -- instead of the `[rows, cols, 2, L, W]` float64 `height_samples_by_row_col` plus the per-env
copy LeggedRobot makes of it (LR:1827-1840, 25.6 KB per env) -- ONE float32 table `tile_heights [tiles, 2, L, W]` in HBM.

Split of the work:
  * here (host, numpy): the random draws, in the reference's order on the reference's generator (numpy's global one unless
    `rng` is given; scripts/train.py:33-40 seeds it), a few float64 scalars per wedge, the patch placement integers
    (tunnel.py:190-194) and the two linspace tables (tunnel_fn.py:152);
  * `lt_build_tunnel_tiles` (csrc/lt_terrain.cu): the face planes of every wedge and every cell of every tile.
The reference spends 3.8 s in a Python loop for 1,024 tiles; the result here is bit-identical (tests/test_gpu_terrain.py
against tests/golden/terrain_single_path_*.npz, recorded from the unmodified reference).

Not built: the trimesh/vertex export for PhysX (out of scope: PhysX), `valid_tunnel_only` (needs OMPL, absent in the
reference's own default run) and the other `terrain_type`s.
"""
import ctypes as C

import numpy as np
import torch

from . import _abi


def _get(cfg, name):
    return cfg[name] if isinstance(cfg, dict) else getattr(cfg, name)


def draw_single_path(cfg, rng=np.random):
    """The draws of Terrain.__init__ (tunnel.py:88-92) and TerrainFunctions.single_path (tunnel_fn.py:105-135) for every tile,
    ceiling first, then floor.  -> n_pyramids [tiles,2] int32, wedges [tiles,2,2,5] float64 (mean_x, mean_y, mean_z, pw, pl)."""
    rows, cols = int(_get(cfg, "num_rows")), int(_get(cfg, "num_cols"))
    hs = _get(cfg, "horizontal_scale")
    p_flat, p_double = _get(cfg, "p_flat"), _get(cfg, "p_double")
    Wp, Lp = int(_get(cfg, "terrain_width") / hs), int(_get(cfg, "terrain_length") / hs)
    pixel_y = int(Lp * _get(cfg, "terrain_ratio_x"))            # SubTerrain(width, length): raw field is (pixel_x, pixel_y)
    w = pixel_y * hs
    tiles = rows * cols
    n_pyr = np.zeros((tiles, 2), np.int32)
    wedges = np.zeros((tiles, 2, 2, 5), np.float64)
    for k in range(tiles):
        rng.uniform(0.0, 1.0)                                   # difficulty: drawn, unused by single_path
        for layer, top in ((0, True), (1, False)):
            p1 = rng.uniform()
            p2 = rng.uniform()
            n = 2 if p2 < p_double else 1
            oy_r, ox_r, lw = (0.6, 0.3, (0.2, 0.4)) if top else (0.4, 0.2, (0.1, 0.3))
            offset_y = rng.uniform(-oy_r, oy_r, size=(n, 1))
            offset_x = rng.uniform(-ox_r, ox_r, size=(n, 1))
            flat_vals = (0.4, 0.7) if top else (0.15, 0.3)
            height_max, height_min = flat_vals if p1 < p_flat else (0.0, 0.0)
            mean_x = np.linspace(-w / 2, w / 2, 3)[1:-1] + np.zeros((n, 1)) + offset_x        # one column of wedges (tunnel_fn.py:127-130)
            mean_y = np.zeros((n, 1)) + offset_y
            mean_z = rng.uniform(mean_x) * (height_max - height_min) + height_min            # uniform(low=mean_x, high=1)
            pw, pl = rng.uniform(low=lw[0], high=lw[1], size=(2, n))
            n_pyr[k, layer] = n
            wedges[k, layer, :n, 0] = mean_x[:, 0]
            wedges[k, layer, :n, 1] = mean_y[:, 0]
            wedges[k, layer, :n, 2] = mean_z[:, 0]
            wedges[k, layer, :n, 3] = pw
            wedges[k, layer, :n, 4] = pl
    return n_pyr, wedges


class TunnelTerrain:
    """Drop-in for the fields of the reference's `Terrain` that the env reads, with the tile table resident on `device`."""

    def __init__(self, cfg, num_robots=None, device="cuda:0", rng=np.random, stream=None):
        if _get(cfg, "terrain_type") != "single_path":
            raise NotImplementedError(f"terrain_type {_get(cfg, 'terrain_type')!r}: only single_path is built on the device")
        self.cfg, self.num_robots = cfg, num_robots
        rows, cols = int(_get(cfg, "num_rows")), int(_get(cfg, "num_cols"))
        self.num_rows, self.num_cols = rows, cols
        self.horizontal_scale = hs = _get(cfg, "horizontal_scale")
        self.vertical_scale = vs = _get(cfg, "vertical_scale")
        self.env_length, self.env_width = _get(cfg, "terrain_length"), _get(cfg, "terrain_width")
        rx, ry = _get(cfg, "terrain_ratio_x"), _get(cfg, "terrain_ratio_y")
        self.terrain_ratio_x, self.terrain_ratio_y = rx, ry
        self.width_per_env_pixels = Wp = int(self.env_width / hs)
        self.length_per_env_pixels = Lp = int(self.env_length / hs)
        ceiling = _get(cfg, "ceiling_height")
        tiles = rows * cols
        n_pyr, wedges = draw_single_path(cfg, rng)
        sub_rows, sub_cols = int(Lp * rx), int(Wp * ry)
        # patch placement (tunnel.py:190-194), tile-local; origins (tunnel.py:209-216)
        window = np.zeros((tiles, 2), np.int32)
        self.env_origins = np.zeros((rows, cols, 3))
        self.all_terrain_origins = np.zeros((rows, cols, 3))
        for k in range(tiles):
            i, j = np.unravel_index(k, (rows, cols))
            sx, ex = int(round((i + 0.5 - rx / 2.0) * Lp, 4)), int(round((i + 0.5 + rx / 2.0) * Lp, 4))
            sy, ey = int((j + 0.5 - ry / 2.0) * Wp), int((j + 0.5 + ry / 2.0) * Wp)
            if (ex - sx, ey - sy) != (sub_rows, sub_cols):
                raise ValueError("terrain patch does not fit its tile (the reference's slice assignment would fail too)")
            window[k] = (sx - i * Lp, sy - j * Wp)
            self.env_origins[i, j] = [(i + 0.5 - _get(cfg, "start_loc")) * self.env_length, (j + 0.5) * self.env_width, 0.0]
            self.all_terrain_origins[i, j] = [i * self.env_length, j * self.env_width, 0.0]
        w, l = sub_rows * hs, sub_cols * hs                     # tunnel_fn.py:103 with (pixel_x, pixel_y) = (sub_cols, sub_rows)
        xs = np.linspace(-w / 2, w / 2, sub_rows)
        ys = np.linspace(-l / 2, l / 2, sub_cols)
        border_top = float((np.ones((), np.int16) * int(1.0 / vs) * ceiling))            # tunnel.py:81
        border_bottom = float(0.5 * int(1.0 / vs))                                       # tunnel.py:82
        consts = np.array([vs, ceiling / vs, 0.05 / vs, border_top, border_bottom], np.float64)
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _abi.LtError("TunnelTerrain builds the tile table with a CUDA kernel: there is no CPU fallback")
        lib = _abi.load()
        self._inputs = [torch.as_tensor(a, device=dev) for a in (n_pyr, wedges, window, xs, ys, consts)]
        self.tile_heights = torch.empty(tiles, 2, Lp, Wp, dtype=torch.float32, device=dev)
        st = stream if stream is not None else torch.cuda.current_stream(dev)
        with torch.cuda.device(dev):
            _abi.check(lib.lt_build_tunnel_tiles(*[C.c_void_p(t.data_ptr()) for t in self._inputs],
                                                 C.c_void_p(self.tile_heights.data_ptr()), tiles, Lp, Wp, sub_rows, sub_cols,
                                                 C.c_void_p(st.cuda_stream)), None, "lt_build_tunnel_tiles")
        self.n_pyramids, self.wedges = n_pyr, wedges
        self._shape = (tiles, Lp, Wp, sub_rows, sub_cols)

    def rebuild(self, stream=None):
        """Re-run the device kernel on the stored draws (timing / re-materialisation after the table was freed)."""
        tiles, Lp, Wp, sub_rows, sub_cols = self._shape
        st = stream if stream is not None else torch.cuda.current_stream(self.tile_heights.device)
        _abi.check(_abi.load().lt_build_tunnel_tiles(*[C.c_void_p(t.data_ptr()) for t in self._inputs],
                                                     C.c_void_p(self.tile_heights.data_ptr()), tiles, Lp, Wp, sub_rows, sub_cols,
                                                     C.c_void_p(st.cuda_stream)), None, "lt_build_tunnel_tiles")

    def table(self, num_envs):
        """The `terrain` dict StepEngine / the env facade take: tile table + per-env tile ids and origins, env i -> tile i mod
        (rows * cols) as the reference's grid_r / grid_c.repeat(m) does (LR:1816-1823)."""
        tiles = self.num_rows * self.num_cols
        if num_envs % tiles:
            raise ValueError(f"num_envs {num_envs} must be a multiple of num_rows * num_cols = {tiles} (LR:1819)")
        env_tile = np.arange(num_envs, dtype=np.int32) % tiles
        return {"tile_heights": self.tile_heights, "env_tile": env_tile,
                "env_origins": self.env_origins.reshape(tiles, 3).astype(np.float32)[env_tile],
                "env_terrain_origin": self.all_terrain_origins.reshape(tiles, 3).astype(np.float32)[env_tile]}
