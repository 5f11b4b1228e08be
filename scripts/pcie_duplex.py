#!/usr/bin/env python
"""Is the host link full duplex on this box?  H2D alone, D2H alone, both at once (two streams)."""
import time
import torch

n = 1 << 30
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, reps=4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                d_a.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_b, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


run(True, True, 1)
a, b, c = run(True, False), run(False, True), run(True, True)
print("H2D alone %.1f GB/s, D2H alone %.1f GB/s, both at once: %.1f + %.1f GB/s (%.1f ms vs %.1f ms serial)"
      % (n / a / 1e9, n / b / 1e9, n / c / 1e9, n / c / 1e9, c * 1e3, (a + b) * 1e3))
