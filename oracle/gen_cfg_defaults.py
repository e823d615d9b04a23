"""Dump the reference's configuration DATA (attribute names and values of its `Cfg` tree) so that the product's
`legged_tracking_b200.cfg` can offer the same config classes with the same defaults without transcribing them.

Test/build infrastructure; needs /root/reference (build container only).  Output: legged_tracking_b200/data/cfg_defaults.json
  base     pristine `Cfg` (go1_gym/envs/base/legged_robot_trajectory_tracking_config.py)
  go1      what `config_go1(Cfg)` changes (go1_gym/envs/go1/go1_crawling.py), as a delta
  presets  what scripts/train.py:46-241 changes on top for the argument sets of the golden fixtures, as deltas

    python -m oracle.gen_cfg_defaults
"""
import importlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
if REPO not in sys.path:
    sys.path.insert(0, REPO)

from oracle import ref_harness  # noqa: E402

def presets():
    """One preset per golden fixture: same scripts/train.py argv (+ overrides) as oracle/gen_golden.FIXTURES, so the flat
    config derived from a preset can be compared key by key with the one recorded off the live reference env."""
    from oracle.gen_golden import FIXTURES
    return {name: dict(argv=list(fx["argv"]), overrides=dict(fx.get("overrides") or {}), num_rows=fx.get("num_rows"),
                       num_cols=fx.get("num_cols")) for name, fx in FIXTURES.items()}


def encode(v):
    if isinstance(v, np.ndarray):
        return {"__ndarray__": v.tolist(), "dtype": str(v.dtype)}
    if isinstance(v, (np.floating, np.integer)):
        return v.item()
    if isinstance(v, float) and v == float("inf"):
        return {"__float__": "inf"}
    if isinstance(v, dict):
        return {"__dict__": {k: encode(x) for k, x in v.items()}}
    if isinstance(v, (list, tuple)):
        return [encode(x) for x in v]
    return v


def dump(cls):
    out = {}
    for k, v in vars(cls).items():
        if isinstance(v, type):
            out[k] = {"__class__": dump(v)}
        else:
            out[k] = encode(v)
    return out


def delta(new, old):
    d = {}
    for k, v in new.items():
        if isinstance(v, dict) and "__class__" in v:
            sub = delta(v["__class__"], old.get(k, {}).get("__class__", {}))
            if sub:
                d[k] = {"__class__": sub}
        elif k not in old or json.dumps(old[k], sort_keys=True) != json.dumps(v, sort_keys=True):
            d[k] = v
    return d


def main():
    ref_harness.install()
    import go1_gym.envs.base.legged_robot_trajectory_tracking_config as cfgmod
    import go1_gym.envs.go1.go1_crawling as crawl
    importlib.reload(cfgmod)
    base = dump(cfgmod.Cfg)
    crawl.config_go1(cfgmod.Cfg)
    go1_full = dump(cfgmod.Cfg)
    P = presets()
    out = {"base": base, "go1": delta(go1_full, base), "presets": {}, "preset_argv": {k: v["argv"] for k, v in P.items()}}
    for name, p in P.items():
        Cfg, _ = ref_harness.build_cfg(p["argv"], num_envs=4096, num_rows=p["num_rows"], num_cols=p["num_cols"],
                                       overrides=p["overrides"])
        full = dump(Cfg)
        d = delta(full, go1_full)
        d.get("env", {}).get("__class__", {}).pop("num_envs", None)
        out["presets"][name] = d
    path = os.path.join(REPO, "legged_tracking_b200", "data", "cfg_defaults.json")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))      # insertion order kept: reward order = class-attribute order
    print(path, os.path.getsize(path), "bytes;", {k: len(json.dumps(v)) for k, v in out["presets"].items()})


if __name__ == "__main__":
    main()
