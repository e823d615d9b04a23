"""Scratch probe: pinned-memory link bandwidth up / down / both ways at once, at the shard sizes of the e2e pipeline."""
import sys
import torch

def t(fn, n=20):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n

for mb in (3, 6, 12, 17, 49, 69):
    n = mb * 1024 * 1024 // 4
    h_up, h_dn = torch.empty(n).pin_memory(), torch.empty(n).pin_memory()
    d_up, d_dn = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    up = t(lambda: d_up.copy_(h_up, non_blocking=True))
    dn = t(lambda: h_dn.copy_(d_dn, non_blocking=True))
    def both():
        cur = torch.cuda.current_stream()
        s1.wait_stream(cur); s2.wait_stream(cur)
        with torch.cuda.stream(s1):
            d_up.copy_(h_up, non_blocking=True)
        with torch.cuda.stream(s2):
            h_dn.copy_(d_dn, non_blocking=True)
        cur.wait_stream(s1); cur.wait_stream(s2)
    bo = t(both)
    g = lambda ms: mb * 1.048576 / ms
    print(f"{mb:3d} MiB  up {g(up):5.1f} GB/s  down {g(dn):5.1f} GB/s  both at once {g(bo):5.1f} GB/s each way ({bo:.3f} ms vs {max(up, dn):.3f})", flush=True)
