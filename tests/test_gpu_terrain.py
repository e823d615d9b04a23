"""-m gpu: the device-side tile-table builder (lt_build_tunnel_tiles through legged_tracking_b200.terrain.TunnelTerrain) against
the tables recorded from the unmodified reference generator, and against the oracle port at the BASELINE size (32 x 32 tiles).
Bit-exact (float64 arithmetic with numpy's op order, float32 result)."""
import numpy as np
import pytest
import torch

from tests.test_oracle_terrain import NAMES, load

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", NAMES)
def test_device_tiles_equal_reference(name):
    from legged_tracking_b200.terrain import TunnelTerrain
    z, cfg, seed = load(name)
    t = TunnelTerrain(cfg, device="cuda:0", rng=np.random.RandomState(seed))
    torch.cuda.synchronize()
    got = t.tile_heights.cpu().numpy()
    assert got.dtype == np.float32 and got.shape == z["tile_heights"].shape
    bad = np.argwhere(got != z["tile_heights"])
    assert bad.size == 0, f"{len(bad)} cells differ, first {bad[0]}: {got[tuple(bad[0])]} vs {z['tile_heights'][tuple(bad[0])]}"
    tiles = cfg["num_rows"] * cfg["num_cols"]
    assert np.array_equal(t.env_origins.reshape(tiles, 3).astype(np.float32), z["env_origins"])
    assert np.array_equal(t.all_terrain_origins.reshape(tiles, 3).astype(np.float32), z["env_terrain_origin"])


def test_device_tiles_equal_oracle_at_full_size_and_feed_the_engine():
    """1,024 tiles (scripts/train.py:129-143 grid), global numpy generator seeded like scripts/train.py:33-40."""
    from legged_tracking_b200 import workload
    from legged_tracking_b200.engine import StepEngine
    from legged_tracking_b200.terrain import TunnelTerrain
    from oracle import terrain_port
    _, cfg, _ = load(NAMES[0])
    cfg = dict(cfg, num_rows=32, num_cols=32)
    np.random.seed(11)
    t = TunnelTerrain(cfg, device="cuda:0")
    want = terrain_port.build(cfg, np.random.RandomState(11))
    assert np.array_equal(t.tile_heights.cpu().numpy(), want["tile_heights"])
    # the table is what the step engine consumes
    flat, _, act, synth = workload.make("cfg2", num_envs=2048)
    table = t.table(2048)
    assert np.array_equal(table["env_origins"][:1024], want["env_origins"])
    eng = StepEngine(flat, table, act, device="cuda:0")
    a = synth.actions(0).cuda()
    for s in range(flat["decimation"]):
        eng.dof_state.copy_(synth.dof_state(0, s).cuda())
        eng.compute_torques(a, s)
    eng.root_states.view(eng.N, -1, 13)[:, 0, :] = synth.root(0).cuda()
    eng.contact_forces.copy_(synth.contact(0).cuda())
    eng.post_physics_step()
    torch.cuda.synchronize()
    assert torch.isfinite(eng.state["obs_buf"]).all()
