#!/usr/bin/env python3
"""Experiment driver: fit + HC1 on C2's shape, device-resident inputs."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from importlib import import_module  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
synth = import_module("anofox_statistics_b200.synth")
G, n, p = 1_000_000, 50, 5
d = synth.make_torch(G, n, p, seed=2)
fit = pkg.DeviceFit(G, G * n, p, inference=True)


def timed(fn, steps=5, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


base = timed(lambda: fit.fit(d["offsets"], d["y"], d["x"], None, method="ols", compute_inference=True))
for hc in ("hc0", "hc1", "hc3"):
    t = timed(lambda: fit.fit(d["offsets"], d["y"], d["x"], None, method="ols", compute_inference=True, hc_type=hc))
    print(hc, "step ms", round(t, 3), "hc pass ms", round(t - base, 3))
