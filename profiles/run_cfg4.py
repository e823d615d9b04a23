"""Driver for ncu launch lists of the cfg 4 step (device mode):  python profiles/run_cfg4.py [n_envs] [steps]"""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main():
    import numpy as np
    import torch
    from legged_tracking_b200 import synthetic, vel_env, workload
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    cfg = vel_env.cfg4_config()
    side = int(np.ceil(np.sqrt(n)))
    xx, yy = np.meshgrid(np.arange(side), np.arange(side), indexing="ij")
    origins = np.zeros((n, 3), np.float32)
    origins[:, 0], origins[:, 1] = 3.0 * xx.flatten()[:n], 3.0 * yy.flatten()[:n]
    rs = np.random.RandomState(5)
    field = dict(height_samples=rs.randint(-10, 10, size=(int(3.0 * side / 0.1) + 64, int(3.0 * side / 0.1) + 64)).astype(np.int16), border_size=2.0,
                 horizontal_scale=0.1, vertical_scale=0.005, measured_points_x=vel_env.CFG4_MEASURED_POINTS_X, measured_points_y=vel_env.CFG4_MEASURED_POINTS_Y)
    eng = vel_env.VelocityStepEngine(cfg, n, origins, workload.load_actuator(), "device", device="cuda:0", seed=9, heightfield=field)
    synth = synthetic.SyntheticState(n, origins, num_bodies=int(cfg["num_bodies"]), seed=77, feet_indices=list(cfg["feet_indices"]))
    eng.episode_length_buf.copy_(synth.initial_episode_length(cfg["max_episode_length"]).cuda())
    eng.bind_inputs = True
    eng.reset_all()
    for k in range(steps):
        root = synth.root(k % 3)
        sim = {"dof_state": [synth.dof_state(k % 3, s).cuda() for s in range(int(cfg["decimation"]))], "root": root.cuda(),
               "contact": synth.contact(k % 3).cuda(), "rigid": synth.rigid(k % 3, root).cuda()}
        eng.step(synth.actions(k % 3).contiguous().cuda(), sim)
    torch.cuda.synchronize()
    print("ok", n, steps, float(eng.rew_buf.sum()), int(eng.reset_buf.sum()))


if __name__ == "__main__":
    main()
