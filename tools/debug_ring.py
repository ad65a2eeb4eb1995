import sys, os
import numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/oracle")
import __graft_entry__ as ge
lib = ge.load_package()
import oracle
from test_gpu_parity import make_case
p, nr, G, method = 8, (900, 1100), 40, "ols"
rng = np.random.default_rng(4000 + p)
c = make_case(rng, G, nr[0], nr[1], p, dirty=False, weights=False)
N = len(c["y"])
y, x = c["y"].copy(), c["x"].copy()
offs = c["offsets"].copy()
train = (rng.uniform(size=N) < 0.8).astype(np.uint8)
lib.force_path("record")
for trial in range(3):
    pd, fd = lib.fit_predict_batch_host(offs, y, x, None, train, method=method)
    po, fo = oracle.fit_predict_batch(offs, y, x, None, train, method=method)
    bad = np.flatnonzero(fd["n_observations"] != fo["n_observations"])
    print("trial", trial, "bad groups", bad, fd["n_observations"][bad], fo["n_observations"][bad])
    for g in bad:
        print("  group", g, "rows", offs[g], offs[g + 1], "n", offs[g + 1] - offs[g], "train sum", train[offs[g]:offs[g + 1]].sum())
print("N", N, "grid", min(148, (N + 4095) // 4096), "rows per tile", -(-N // min(148, (N + 4095) // 4096)))
