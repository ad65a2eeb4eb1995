#!/usr/bin/env python3
"""Experiment driver: fit step of 1 M groups x 50 rows for p = 4 and p = 8 (small-group kernels), device-resident."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from importlib import import_module  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
synth = import_module("anofox_statistics_b200.synth")
shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]] or [(8, 50), (4, 50), (8, 200)]
for p, n in shapes:
    G = 1_000_000 if n == 50 else 250_000
    d = synth.make_torch(G, n, p, seed=2)
    fit = pkg.DeviceFit(G, G * n, p, inference=True)
    for _ in range(3):
        fit.fit(d["offsets"], d["y"], d["x"], None, method="ols", compute_inference=True)
    torch.cuda.synchronize()
    pkg.stage_timing(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        fit.fit(d["offsets"], d["y"], d["x"], None, method="ols", compute_inference=True)
    e1.record()
    torch.cuda.synchronize()
    calls, acc_ms, solve_ms = pkg.stage_timing_read()
    pkg.stage_timing(False)
    print("p", p, "n", n, "step ms", round(e0.elapsed_time(e1) / 5, 4), "accumulate ms", round(acc_ms / max(calls, 1), 4),
          "ok", int((fit.out["status"] == 0).sum().item()))
    del d, fit
    torch.cuda.empty_cache()
