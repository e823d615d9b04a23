"""Golden vectors for the PPO rows recorded from the UNMODIFIED reference.

Test infrastructure only; runs in the build container (needs /root/reference).  It drives the reference's own
`go1_gym_learn.ppo_cse.PPO` (ppo.py), `ActorCritic` (actor_critic.py) and `RolloutStorage` (rollout_storage.py)
on CPU through their public API -- act / process_env_step / compute_returns / update -- on seeded synthetic
transitions and records everything a re-implementation needs to be checked step by step:

  * the initial ActorCritic state_dict and the hyper-parameters (AC_Args / PPO_Args);
  * per rollout step: obs, privileged_obs, obs_history, the sampled actions, rewards, dones, time_outs and what
    the reference stored (values, log-probs, mu, sigma, bootstrapped rewards);
  * after compute_returns: returns and normalised advantages;
  * during update: the minibatch permutation (torch.randperm is wrapped, values still torch's), the total gradient
    norm that nn.utils.clip_grad_norm_ saw for every minibatch, the learning rate after every minibatch;
  * after update: the returned mean losses and the final state_dict.

    python -m oracle.gen_golden_ppo            # writes tests/golden/ppo_*.npz
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
if REPO not in sys.path:
    sys.path.insert(0, REPO)

from oracle import ref_harness  # noqa: E402

CASES = {
    # name: sizes, AC_Args overrides, PPO_Args overrides
    "ppo_adaptive_64x8": dict(N=64, T=8, num_obs=42, num_priv=6, H=5, num_actions=12, seed=5,
                              ac=dict(actor_hidden_dims=[64, 48, 32], critic_hidden_dims=[64, 48, 32],
                                      adaptation_module_branch_hidden_dims=[32, 16]),
                              ppo=dict(), time_outs=True),
    "ppo_fixed_32x6": dict(N=32, T=6, num_obs=30, num_priv=2, H=3, num_actions=12, seed=9,
                           ac=dict(actor_hidden_dims=[48, 32], critic_hidden_dims=[40, 24],
                                   adaptation_module_branch_hidden_dims=[24], activation="tanh", init_noise_std=0.7),
                           ppo=dict(schedule="fixed", use_clipped_value_loss=False, num_learning_epochs=2,
                                    num_mini_batches=2, learning_rate=5e-4, entropy_coef=0.0, gamma=0.97, lam=0.9,
                                    max_grad_norm=0.5),
                           time_outs=False),
    # go1_gym_learn/ppo_cse_cnn (row f4), MLP height-map encoder + Identity: 2 history rows of 20 scalars + 2x5x4 heights
    "ppo_cnn_48x6": dict(variant="ppo_cse_cnn", N=48, T=6, num_obs=60, num_priv=3, H=2, num_actions=12, seed=21,
                         ac=dict(actor_hidden_dims=[64, 48, 32], critic_hidden_dims=[64, 32], adaptation_module_branch_hidden_dims=[48, 24],
                                 height_map_shape=(2, 5, 4), cnn_num_embedding=32, use_cnn=False, use_gru=False),
                         ppo=dict(num_learning_epochs=3, num_mini_batches=3), time_outs=True),
}


def synth_transitions(case):
    """Seeded synthetic env outputs (numpy Generator, independent of torch's RNG stream)."""
    g = np.random.default_rng(case["seed"])
    N, T, no, npv, H = case["N"], case["T"], case["num_obs"], case["num_priv"], case["H"]
    out = dict(
        obs=g.standard_normal((T + 1, N, no)).astype(np.float32) * 0.6,
        priv=g.standard_normal((T + 1, N, npv)).astype(np.float32) * 0.4,
        hist=g.standard_normal((T + 1, N, no * H)).astype(np.float32) * 0.6,
        rewards=(g.standard_normal((T, N)) * 0.2 + 0.05).astype(np.float32),
        dones=(g.random((T, N)) < 0.12),
    )
    out["time_outs"] = out["dones"] & (g.random((T, N)) < 0.5)
    return out


def record(case):
    ref_harness.install()
    import importlib
    variant = case.get("variant", "ppo_cse")
    acmod = importlib.import_module(f"go1_gym_learn.{variant}.actor_critic")
    ppomod = importlib.import_module(f"go1_gym_learn.{variant}.ppo")
    import torch.nn as nn

    saved_ac = {k: getattr(acmod.AC_Args, k) for k in vars(acmod.AC_Args)}
    saved_ppo = {k: getattr(ppomod.PPO_Args, k) for k in vars(ppomod.PPO_Args)}
    for k, v in case["ac"].items():
        setattr(acmod.AC_Args, k, v)
    for k, v in case["ppo"].items():
        setattr(ppomod.PPO_Args, k, v)
    try:
        N, T, no, npv, H, na = (case[k] for k in ("N", "T", "num_obs", "num_priv", "H", "num_actions"))
        torch.manual_seed(case["seed"])
        ac = acmod.ActorCritic(no, npv, no * H, na)
        out = {"init/" + k: v.detach().numpy().copy() for k, v in ac.state_dict().items()}
        alg = ppomod.PPO(ac, device="cpu")
        alg.init_storage(N, T, [no], [npv], [no * H], [na])
        tr = synth_transitions(case)
        for k, v in tr.items():
            out["in/" + k] = v
        actions = []
        for t in range(T):
            a = alg.act(torch.from_numpy(tr["obs"][t]), torch.from_numpy(tr["priv"][t]), torch.from_numpy(tr["hist"][t]))
            actions.append(a.numpy().copy())
            infos = {"time_outs": torch.from_numpy(tr["time_outs"][t])} if case["time_outs"] else {}
            alg.process_env_step(torch.from_numpy(tr["rewards"][t]), torch.from_numpy(tr["dones"][t]), infos)
        out["in/actions"] = np.stack(actions)
        alg.compute_returns(torch.from_numpy(tr["hist"][T]), torch.from_numpy(tr["priv"][T]))
        st = alg.storage
        for k in ("observations", "privileged_observations", "observation_histories", "actions", "rewards", "dones",
                  "values", "actions_log_prob", "mu", "sigma", "returns", "advantages"):
            out["storage/" + k] = getattr(st, k).detach().numpy().copy()
        out["storage/last_values"] = ac.evaluate(torch.from_numpy(tr["hist"][T]), torch.from_numpy(tr["priv"][T])).detach().numpy()

        perms, norms, lrs = [], [], []
        _randperm, _clip = torch.randperm, nn.utils.clip_grad_norm_

        def randperm(*a, **k):
            p = _randperm(*a, **k)
            perms.append(p.numpy().copy())
            return p

        def clip(params, max_norm, *a, **k):
            tn = _clip(params, max_norm, *a, **k)
            norms.append(float(tn))
            lrs.append(alg.optimizer.param_groups[0]["lr"])
            return tn

        torch.randperm, nn.utils.clip_grad_norm_ = randperm, clip
        try:
            losses = alg.update()
        finally:
            torch.randperm, nn.utils.clip_grad_norm_ = _randperm, _clip
        assert len(perms) == 1
        out["update/perm"] = perms[0]
        out["update/grad_norms"] = np.asarray(norms, np.float64)
        out["update/lrs"] = np.asarray(lrs, np.float64)
        out["update/losses"] = np.asarray([float(x) for x in losses], np.float64)
        out["update/final_lr"] = np.float64(alg.learning_rate)
        for k, v in ac.state_dict().items():
            out["final/" + k] = v.detach().numpy().copy()
        # hyper-parameters as the reference held them during the run
        for k in vars(acmod.AC_Args):
            out["ac_args/" + k] = np.asarray(getattr(acmod.AC_Args, k))
        for k in vars(ppomod.PPO_Args):
            out["ppo_args/" + k] = np.asarray(getattr(ppomod.PPO_Args, k))
        out["meta/sizes"] = np.asarray([N, T, no, npv, H, na], np.int64)
        out["meta/time_outs"] = np.asarray(bool(case["time_outs"]))
        out["meta/variant"] = np.asarray(variant)
        return out
    finally:
        for k, v in saved_ac.items():
            setattr(acmod.AC_Args, k, v)
        for k, v in saved_ppo.items():
            setattr(ppomod.PPO_Args, k, v)


def main(names=None):
    os.makedirs(os.path.join(REPO, "tests", "golden"), exist_ok=True)
    for name, case in CASES.items():
        if names and name not in names:
            continue
        out = record(case)
        path = os.path.join(REPO, "tests", "golden", name + ".npz")
        np.savez_compressed(path, **out)
        print(f"{name}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB, losses {out['update/losses'][:3]}, "
              f"lr {out['update/lrs'][0]:.2e} -> {float(out['update/final_lr']):.2e}")


if __name__ == "__main__":
    main(sys.argv[1:])
