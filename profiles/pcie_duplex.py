"""PCIe ceiling of the box: one-direction and simultaneous two-direction copy bandwidth from pinned memory (what bounds bench.py's e2e)."""
import torch

n = 64 << 20
h_up, h_dn = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
d_up, d_dn = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    s1.synchronize(); s2.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def up():
    with torch.cuda.stream(s1):
        d_up.copy_(h_up, non_blocking=True)


def dn():
    with torch.cuda.stream(s2):
        h_dn.copy_(d_dn, non_blocking=True)


def both():
    up(); dn()


import time
for name, fn, nbytes in (("H2D", up, n), ("D2H", dn, n), ("H2D + D2H together", both, 2 * n)):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 10
    print(f"{name}: {nbytes / dt / 1e9:.1f} GB/s")
