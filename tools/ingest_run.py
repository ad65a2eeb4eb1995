#!/usr/bin/env python3
"""Experiment driver: one warm-up and two timed ingests of C2's shape (50 M scan-order rows, 1 M groups, 5 features);
run under `ncu --metrics gpu__time_duration.sum` for the per-kernel times."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
G, n, p = 1_000_000, 50, 5
N = G * n
gen = torch.Generator(device="cuda")
gen.manual_seed(5)
gid = torch.randint(0, G, (N,), generator=gen, dtype=torch.int64, device="cuda")
y = torch.rand(N, generator=gen, dtype=torch.float64, device="cuda")
x = torch.rand((N, p), generator=gen, dtype=torch.float64, device="cuda")
ing = pkg.DeviceIngest(N, p, G)
for _ in range(3):
    ing.run(gid, y, x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ing.run(gid, y, x)
e1.record()
torch.cuda.synchronize()
print("ingest ms", e0.elapsed_time(e1))
