"""Where the time of one pipelined host-buffer step goes: per-shard CUDA-event timestamps (ms from the start of the step)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from legged_tracking_b200 import workload
from legged_tracking_b200.engine import StepEngine, HostStepPipeline

n_chunks = int(sys.argv[1]) if len(sys.argv) > 1 else 4
M = 65536 // n_chunks
engines, h_in, h_out = [], [], []
for c in range(n_chunks):
    flat, terrain, act, synth = workload.make("cfg5", num_envs=M, seed=2000 + c)
    e = StepEngine(flat, terrain, act, device="cuda:0", seed=101 + c)
    bench.init_engine_state(e, synth, flat)
    engines.append(e)
    rs = bench.ReplaySets(e, synth, flat, 2, pinned_host=True)
    h_in.append(rs.host)
    h_out.append({"obs": torch.empty(M, flat["num_obs"]).pin_memory(), "priv": torch.empty(M, flat["num_privileged_obs"]).pin_memory(),
                  "rew": torch.empty(M).pin_memory(), "done": torch.empty(M, dtype=torch.bool).pin_memory()})
pipe = HostStepPipeline(engines)
for k in range(5):
    pipe.step([h[k % 2] for h in h_in], h_out)
torch.cuda.synchronize()
# instrumented copy of HostStepPipeline.step
ev = lambda: torch.cuda.Event(enable_timing=True)
t0 = ev(); marks = []
cpu0 = time.perf_counter()
t0.record()
cur = torch.cuda.current_stream()
for s in pipe.streams:
    s.wait_stream(cur)
cpu_marks = []
for e, s, own, hi, ho in zip(pipe.engines, pipe.streams, pipe.dev_in, [h[0] for h in h_in], h_out):
    with torch.cuda.stream(s):
        a, b, c_, d = ev(), ev(), ev(), ev()
        a.record()
        own["actions"].copy_(hi["actions"], non_blocking=True)
        for sub, dof in enumerate(hi["dof"]):
            own["dof"][sub].copy_(dof, non_blocking=True)
        own["root"].copy_(hi["root"], non_blocking=True)
        own["contact"].copy_(hi["contact"], non_blocking=True)
        for sub in range(len(hi["dof"])):
            e.bind_sim(dof_state=own["dof"][sub])
            e.compute_torques(own["actions"], sub)
        b.record()
        e.post_physics_step()
        c_.record()
        ho["obs"].copy_(e.state["obs_buf"], non_blocking=True)
        ho["priv"].copy_(e.state["privileged_obs_buf"], non_blocking=True)
        ho["rew"].copy_(e.state["rew_buf"], non_blocking=True)
        ho["done"].copy_(e.state["reset_buf"], non_blocking=True)
        d.record()
        marks.append((a, b, c_, d))
    cpu_marks.append(time.perf_counter() - cpu0)
for s in pipe.streams:
    s.synchronize()
total = time.perf_counter() - cpu0
print("cpu enqueue done per shard (ms):", [round(x * 1e3, 3) for x in cpu_marks], "step wall ms", round(total * 1e3, 3))
for k, (a, b, c_, d) in enumerate(marks):
    print(f"shard {k}: start {t0.elapsed_time(a):.3f}  inputs+torques done {t0.elapsed_time(b):.3f}  chain done {t0.elapsed_time(c_):.3f}  D2H done {t0.elapsed_time(d):.3f}")
