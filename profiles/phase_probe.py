"""Scratch probe: S env shards on S streams, each free-running its own captured step graph, optionally phase-shifted, against
the single-engine step.  python profiles/phase_probe.py [total_envs] [steps]"""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch
import bench
from legged_tracking_b200 import workload
from legged_tracking_b200.engine import StepEngine

N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
K = int(sys.argv[2]) if len(sys.argv) > 2 else 300


def build(n, seed):
    flat, terrain, act, synth = workload.make("cfg5", num_envs=n, seed=1000 + seed)
    eng = StepEngine(flat, terrain, act, device="cuda:0", seed=11 + seed)
    bench.init_engine_state(eng, synth, flat)
    rs = bench.ReplaySets(eng, synth, flat, 3)
    for k in range(3):
        bench.launch_step(eng, rs.sets[k % 3])
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    graphs = []
    for s in rs.sets:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            bench.launch_step(eng, s, eager_gravity=False)
        graphs.append(g)
    return eng, rs, graphs


def run(S, shift):
    shards = [build(N // S, i) for i in range(S)]
    streams = [torch.cuda.Stream() for _ in range(S)]
    torch.cuda.synchronize()

    def region(steps):
        cur = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for s in streams:
            s.wait_stream(cur)
        if shift:
            # shard i starts after shard i-1's torque launches of the first step: a one-time phase offset
            for i, (eng, rs, graphs) in enumerate(shards):
                with torch.cuda.stream(streams[i]):
                    if i:
                        streams[i].wait_event(prev_ev)
                    s0 = rs.sets[0]
                    for sub, dof in enumerate(s0["dof"]):
                        eng.bind_sim(dof_state=dof)
                        eng.compute_torques(s0["actions"], sub)
                    prev_ev = torch.cuda.Event()
                    prev_ev.record(streams[i])
        for k in range(steps):
            for i, (eng, rs, graphs) in enumerate(shards):
                with torch.cuda.stream(streams[i]):
                    graphs[k % 3].replay()
        for s in streams:
            cur.wait_stream(s)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)
    region(10)
    ms = min(region(K) for _ in range(3)) / K
    print(f"shards {S} shift {shift}: {ms * 1e3:.1f} us per step of {N} envs -> {N / ms / 1e3:.1f} M env-steps/s", flush=True)
    del shards


for S, shift in ((1, 0), (2, 0), (2, 1), (4, 0), (4, 1), (8, 0)):
    run(S, shift)
