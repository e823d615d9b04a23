"""Small driver for ncu captures: a few full steps of the cfg5 workload (65,536 envs unless told otherwise).

  python profiles/run_step.py [n_envs] [steps]
"""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main():
    import torch
    import bench
    from legged_tracking_b200 import workload
    from legged_tracking_b200.engine import StepEngine
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    flat, terrain, act, synth = workload.make("cfg5", num_envs=n)
    eng = StepEngine(flat, terrain, act, device="cuda:0", seed=3)
    bench.init_engine_state(eng, synth, flat)
    rs = bench.ReplaySets(eng, synth, flat, 3)
    for k in range(steps):
        bench.launch_step(eng, rs.sets[k % 3])
    torch.cuda.synchronize()
    print("ok", n, steps, float(eng.state["rew_buf"].sum()))


if __name__ == "__main__":
    main()
