"""Fake `isaacgym.gymutil` (oracle shim, test infrastructure only)."""


def parse_device_str(s):
    """'cuda:0' -> ('cuda', 0); 'cpu' -> ('cpu', 0)  (base_task.py:25-32 picks the env device from this)."""
    s = str(s)
    if ":" in s:
        kind, idx = s.split(":")
        return kind, int(idx)
    return s, 0


def parse_sim_config(cfg_dict, sim_params):
    """Copy scalar entries of vars(cfg.sim) onto sim_params; dt stays a C float."""
    import numpy as np
    for k, v in cfg_dict.items():
        if k == "dt":
            sim_params.dt = float(np.float32(v))
        elif k in ("substeps", "use_gpu_pipeline", "up_axis"):
            setattr(sim_params, k, v)


class WireframeSphereGeometry:
    def __init__(self, *a, **k):
        pass


def draw_lines(*a, **k):
    pass
