#!/usr/bin/env python3
"""Raw pinned-memory copy bandwidth of the box (the floor of the e2e leg): H2D, D2H, both at once."""
import torch

n = 464486400  # bytes per step of the 4K workload (H2D); D2H is 199065600
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(199065600, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(199065600, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=5):
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


def h2d():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


import time
for name, fn, nb in (("h2d", h2d, n), ("d2h", d2h, 199065600), ("both", lambda: (h2d(), d2h()), n + 199065600)):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / 5 * 1e3
    print(f"{name}: {ms:.3f} ms per step, {nb / ms / 1e6:.1f} GB/s")
