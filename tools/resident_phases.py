#!/usr/bin/env python3
"""Experiment: per-phase clock totals of the resident fit-predict kernel on C5 (ANOFOX_B200_RESIDENT_DEBUG=1)."""
import os
import sys
os.environ["ANOFOX_B200_RESIDENT_DEBUG"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from importlib import import_module  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
synth = import_module("anofox_statistics_b200.synth")
G, n, p = 100_000, 1000, 8
d = synth.make_torch(G, n, p, seed=2, holdout=0.2)
pkg.force_path("resident")
fit = pkg.DeviceFit(G, G * n, p, inference=False, predict=True)
for _ in range(3):
    fit.fit_predict(d["offsets"], d["y"], d["x"], None, None, method="ols")
torch.cuda.synchronize()
raw = fit.debug_raw()
for tw in (0, 1):
    v = raw[12 + 8 * tw: 12 + 8 * tw + 6].astype(float)
    groups = max(v[5], 1)
    names = ["wait_full", "fit(moments+exchange+solve)", "prep", "barrier", "predict+tail_barrier"]
    print("warp", tw, "groups", int(v[5]), " cycles per group:", {k: round(v[i] / groups) for i, k in enumerate(names)},
          "sum", round(sum(v[:5]) / groups))
