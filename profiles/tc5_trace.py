"""Timing study of lt_torque_tc5_kernel: per-CTA %globaltimer stamps (build with -DTC5_TRACE: profiles/variants.py)."""
import ctypes as C
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import numpy as np
import torch
import bench
from legged_tracking_b200 import workload, _abi
from legged_tracking_b200.engine import StepEngine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
flat, terrain, act, synth = workload.make("cfg5", num_envs=n)
eng = StepEngine(flat, terrain, act, device="cuda:0", seed=3)
bench.init_engine_state(eng, synth, flat)
rs = bench.ReplaySets(eng, synth, flat, 3)
for k in range(40):
    s = rs.sets[k % 3]
    eng.bind_sim(dof_state=s["dof"][k % 4])
    eng.compute_torques(s["actions"], k % 4)
torch.cuda.synchronize()
lib = _abi.load()
buf = (C.c_ulonglong * (2048 * 8))()
lib.lt_tc5_trace_read.argtypes = [C.c_void_p]
assert lib.lt_tc5_trace_read(buf) == 0
t = np.frombuffer(buf, dtype=np.uint64).reshape(2048, 8).astype(np.int64)
grid = int((t[:, 0] > 0).sum())
t = t[:grid]
t0 = t[:, 0].min()
rel = (t - t0) / 1000.0
names = ["entry", "setup done", "loop start", "tile0 pre done", "tile1 start", "loop end", "exit sync"]
print("CTAs", grid, "kernel span us", (t[:, 6].max() - t0) / 1000.0)
for k, nm in enumerate(names):
    v = rel[:, k]
    print(f"{nm:16s} min {v.min():7.2f}  p50 {np.median(v):7.2f}  p90 {np.percentile(v, 90):7.2f}  max {v.max():7.2f}")
d = rel[:, 5] - rel[:, 4]
print("tiles 1..end per CTA us: p50", np.median(d), "max", d.max())
