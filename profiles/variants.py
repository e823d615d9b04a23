"""Tuning harness: build liblt_b200 variants (-D switches) here, time them on the GPU box.

  python profiles/variants.py build            # in the build container: writes build/variants/<name>.so
  python profiles/variants.py run [N ...]      # on the GPU box (gpurun): post-physics + torque time per variant
"""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
VDIR = os.path.join(REPO, "build", "variants")
MMA_ONLY = ("lt_torque_mma.cu", "lt_torque_tc5.cu")          # variants whose defines touch only the tensor-pipe torque kernel
VARIANTS = {
    "tc5_base": [],
    "tc5_exp1_noalu": ["TC5_EXP=1"],
    "tc5_exp2_nomma": ["TC5_EXP=2"],
    "tc5_exp3_neither": ["TC5_EXP=3"],
    "tc5_trace": ["TC5_TRACE=1"],          # per-CTA phase stamps: LT_TORQUE_TC5=1 LT_B200_LIB=build/variants/tc5_trace.so python profiles/tc5_trace.py
}
# round-2 experiments on this kernel (64-column TMEM layout at 4..8 CTAs per SM, paired reciprocals, early first fetch, equal tile
# counts, staggered start): profiles/r02_tc5.md, code in profiles/r02_tc5_experiments.patch
# earlier ablations: TC5_EXP=1 (no softsign / split arithmetic), 2 (no MMAs), 3 (neither): profiles/r01e_tc5.md
# FP32-pipe torque kernel variants tuned in an earlier pass (build with only=None): LT_TORQUE_SPT / LT_TORQUE_MINB / LT_TORQUE_THREADS


def build():
    from concurrent.futures import ThreadPoolExecutor
    from legged_tracking_b200 import build as b
    os.makedirs(VDIR, exist_ok=True)

    def one(item):
        name, defs = item
        b.build(force=True, defines=defs, out=os.path.join(VDIR, name + ".so"), only=MMA_ONLY if name.startswith(("mma_", "tc5_")) else None)
        print("built", name, flush=True)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        list(ex.map(one, VARIANTS.items()))


def run_one(n_envs):
    import torch
    from legged_tracking_b200 import workload
    from legged_tracking_b200.engine import StepEngine
    import bench
    flat, terrain, act, synth = workload.make("cfg5", num_envs=n_envs)
    eng = StepEngine(flat, terrain, act, device="cuda:0", seed=3)
    bench.init_engine_state(eng, synth, flat)
    rs = bench.ReplaySets(eng, synth, flat, 3)

    def post(k):
        s = rs.sets[k % 3]
        eng.bind_sim(root_states=s["root"], contact_forces=s["contact"], dof_state=s["dof"][-1])
        eng.post_physics_step(advance_gravity=False)

    def torque(k):
        s = rs.sets[k % 3]
        eng.bind_sim(dof_state=s["dof"][k % 4])
        eng.compute_torques(s["actions"], k % 4)
    for k in range(10):
        torque(k)
        post(k)
    return {"n": n_envs, "post_us": bench.time_region(post, 200) / 200 * 1e3, "torque_us": bench.time_region(torque, 200) / 200 * 1e3}


def run(ns):
    for name in VARIANTS:
        lib = os.path.join(VDIR, name + ".so")
        if not os.path.exists(lib):
            continue
        for n in ns:
            env = dict(os.environ, LT_B200_LIB=lib)
            r = subprocess.run([sys.executable, __file__, "one", str(n)], env=env, capture_output=True, text=True)
            print(name, r.stdout.strip().splitlines()[-1] if r.returncode == 0 else r.stderr[-400:], flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build()
    elif sys.argv[1] == "one":
        print(json.dumps(run_one(int(sys.argv[2]))))
    else:
        run([int(a) for a in sys.argv[2:]] or [65536, 8192])
