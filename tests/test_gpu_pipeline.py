"""-m gpu: engine.HostStepPipeline (env shards on separate CUDA streams, host buffers in / out) returns exactly what the same
shards give when stepped one after the other on the default stream."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _shard(seed, n):
    from legged_tracking_b200 import workload
    from legged_tracking_b200.engine import StepEngine
    flat, terrain, act, synth = workload.make("cfg2", num_envs=n, seed=seed, tile_grid=(4, 4))
    eng = StepEngine(flat, terrain, act, device="cuda:0", seed=seed)
    eng.state["episode_length_buf"].copy_(synth.initial_episode_length(flat["max_episode_length"]).cuda())
    return flat, eng, synth


def _host_inputs(flat, synth, k, n):
    A = flat["num_actors"]
    root = torch.zeros(n, A, 13)
    root[:, :, 6] = 1.0
    root[:, 0, :] = synth.root(k)
    return {"actions": synth.actions(k).contiguous().pin_memory(),
            "dof": [synth.dof_state(k, s).pin_memory() for s in range(flat["decimation"])],
            "root": root.view(n * A, 13).contiguous().pin_memory(), "contact": synth.contact(k).pin_memory()}


def test_pipeline_equals_sequential_shards():
    from legged_tracking_b200.engine import HostStepPipeline
    n, shards = 512, 3
    a = [_shard(40 + c, n) for c in range(shards)]
    b = [_shard(40 + c, n) for c in range(shards)]
    pipe = HostStepPipeline([e for _, e, _ in a])
    assert pipe.num_envs == n * shards
    for k in range(3):
        h_in = [_host_inputs(flat, synth, k, n) for flat, _, synth in a]
        h_out = [{"obs": torch.empty(n, flat["num_obs"]).pin_memory(), "priv": torch.empty(n, flat["num_privileged_obs"]).pin_memory(),
                  "rew": torch.empty(n).pin_memory(), "done": torch.empty(n, dtype=torch.bool).pin_memory()} for flat, _, _ in a]
        if k == 1:                                             # the single-copy path: inputs kept in the pipeline's pinned layout
            packed = pipe.alloc_host_inputs()
            for dst, src in zip(packed, h_in):
                dst["actions"].copy_(src["actions"]); dst["root"].copy_(src["root"]); dst["contact"].copy_(src["contact"])
                for a_, b_ in zip(dst["dof"], src["dof"]):
                    a_.copy_(b_)
            pipe.step(packed, h_out)
        else:
            pipe.step(h_in, h_out)
        for (flat, eng, _), hi, ho in zip(b, h_in, h_out):
            act = hi["actions"].cuda()
            for s, dof in enumerate(hi["dof"]):
                eng.dof_state.copy_(dof.cuda())
                eng.compute_torques(act, s)
            eng.root_states.copy_(hi["root"].cuda())
            eng.contact_forces.copy_(hi["contact"].cuda())
            eng.post_physics_step()
            torch.cuda.synchronize()
            assert torch.equal(ho["obs"], eng.state["obs_buf"].cpu()), k
            assert torch.equal(ho["rew"], eng.state["rew_buf"].cpu()), k
            assert torch.equal(ho["done"], eng.state["reset_buf"].cpu().bool()), k
            assert torch.equal(ho["priv"], eng.state["privileged_obs_buf"].cpu()), k


def test_captured_pipeline_step_equals_the_eager_one():
    """HostStepPipeline.capture / replay: the whole pipelined step (uploads, kernels and downloads of uneven shards forked over
    their streams) as one CUDA graph gives what the eagerly issued shards give, step after step, with new host data written
    into the same pinned blocks before every replay."""
    from legged_tracking_b200.engine import HostStepPipeline
    sizes = (256, 768, 512)
    a = [_shard(60 + c, n) for c, n in enumerate(sizes)]
    b = [_shard(60 + c, n) for c, n in enumerate(sizes)]
    pa, pb = HostStepPipeline([e for _, e, _ in a]), HostStepPipeline([e for _, e, _ in b])
    mk_out = lambda: [{"obs": torch.empty(n, flat["num_obs"]).pin_memory(), "priv": torch.empty(n, flat["num_privileged_obs"]).pin_memory(),   # noqa: E731
                       "rew": torch.empty(n).pin_memory(), "done": torch.empty(n, dtype=torch.bool).pin_memory()}
                      for (flat, _, _), n in zip(a, sizes)]
    out_a, out_b = mk_out(), mk_out()
    packed_a, packed_b = pa.alloc_host_inputs(), pb.alloc_host_inputs()

    def fill(packed, k):
        for dst, (flat, _, synth), n in zip(packed, a, sizes):
            src = _host_inputs(flat, synth, k, n)
            dst["actions"].copy_(src["actions"]); dst["root"].copy_(src["root"]); dst["contact"].copy_(src["contact"])
            for a_, b_ in zip(dst["dof"], src["dof"]):
                a_.copy_(b_)
    fill(packed_a, 0)
    fill(packed_b, 0)
    pa.step(packed_a, out_a)                                   # warm-up step on both, eagerly
    pb.step(packed_b, out_b)
    graph = pa.capture(packed_a, out_a)                        # capture does not execute: state is untouched
    replays = 0
    for k in range(1, 5):
        fill(packed_a, k)
        fill(packed_b, k)
        if pa.graph_replay_ok():
            pa.replay(graph)
            replays += 1
        else:
            pa.step(packed_a, out_a)
        pb.step(packed_b, out_b)
        for oa, ob in zip(out_a, out_b):
            for key in ("obs", "priv", "rew", "done"):
                assert torch.equal(oa[key], ob[key]), (k, key)
    assert replays >= 3
    for (_, ea, _), (_, eb, _) in zip(a, b):
        assert torch.equal(ea.state["episode_sums"], eb.state["episode_sums"])
        assert ea.gravity.common_step_counter == eb.gravity.common_step_counter


def test_reset_write_back_rows_are_what_the_step_wrote():
    """alloc_host_outputs: the rows a reset rewrote inside the simulator state (robot root state, dof pos / vel) come down with
    the step's results, for exactly the envs of reset_buf.nonzero(), in ascending order (LR:143, 1010-1013, 1049-1052)."""
    from legged_tracking_b200.engine import HostStepPipeline
    n = 1024
    flat, eng, synth = _shard(77, n)
    eng.state["episode_length_buf"].fill_(int(flat["max_episode_length"]) - 1)          # two steps from a time-out for everyone ...
    eng.state["episode_length_buf"][::7] = int(flat["max_episode_length"])              # ... and every 7th env times out now
    pipe = HostStepPipeline([eng])
    out = pipe.alloc_host_outputs(reset_capacity=256)
    pipe.step([_host_inputs(flat, synth, 0, n)], out)
    ho = out[0]
    cnt = int(ho["reset_count"][0])
    want_ids = ho["done"].nonzero().flatten()
    assert cnt == want_ids.numel() and 0 < cnt <= 256
    assert torch.equal(ho["reset_ids"][:cnt], want_ids)
    root = eng.root_states.view(n, -1, 13)[:, 0, :].cpu()
    dof = eng.dof_state.view(n, -1).cpu()
    assert torch.equal(ho["reset_root"][:cnt], root[want_ids])
    assert torch.equal(ho["reset_dof"][:cnt], dof[want_ids])
    assert float(ho["reset_dof"][:cnt, 1::2].abs().max()) == 0.0                          # dof_vel = 0 after _reset_dofs (LR:1008)
