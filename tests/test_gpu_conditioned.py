"""GPU parity OUTSIDE the well-conditioned regime: controlled cond(X) from large column offsets, polynomial
features, rho = 1 - 10^-k pairs and a column collinear to within r of earlier ones (straddling R's 1e-7 rule).

Bit-exact: status, n_observations, NaN / NULL pattern (i.e. the aliasing decision of every column).
Values: max(1e-9, COND_C * eps * kappa_LS) per group (tests/parity.py: "tolerance scaled by condition number").
The Gram-Cholesky solve alone cannot meet this (cond(X'X) = cond(X)^2); the groups it flags are re-fitted by
the Householder QR fallback (csrc/qr.cuh), whose use the tests also assert through the device counters.
"""
import numpy as np
import pytest

from conditioned import KINDS, group_tolerances, make_conditioned
from parity import assert_fit_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(b200):
    b200.device_check()
    return b200


@pytest.fixture(params=["auto", "stash", "record", "fused", "stages", "qr", "resident"])
def path(request, lib):
    lib.force_path(request.param)
    yield request.param
    lib.force_path("auto")


CASES = [(k, lv) for k, lvs in KINDS.items() for lv in lvs]


@pytest.mark.parametrize("kind,level", CASES, ids=[f"{k}-{lv}" for k, lv in CASES])
@pytest.mark.parametrize("noise", [2.0, 1e-6], ids=["noisy", "nearexact"])
def test_conditioned_small_groups(lib, oracle, path, kind, level, noise):
    """p = 5 (the small-group kernels and, forced, the record / stage paths), OLS with inference."""
    rng = np.random.default_rng(int(9000 + 13 * CASES.index((kind, level)) + (noise < 1)))
    c = make_conditioned(rng, kind, level, G=96, n=40, p=5, noise=noise)
    opt = dict(method="ols", compute_inference=True)
    dev = lib.fit_batch_host(c["offsets"], c["y"], c["x"], None, **opt)
    ora = oracle.fit_batch(c["offsets"], c["y"], c["x"], None, **opt)
    assert (ora["status"] == 0).all()
    tol, kap = group_tolerances(c, ora)
    assert_fit_parity(dev, ora, inference=True, label=f"{kind}-{level}-{noise} kappa<={kap.max():.3g}", **tol)


WIDE = [("poly", 6, 12, "ols"), ("rho", 8, 16, "ridge"), ("rho", 6, 16, "wls"), ("nearcol", 4e-7, 16, "ols"),
        ("nearcol", 2.5e-8, 24, "ols"), ("offset", 6, 40, "ols"), ("nearcol", 1e-6, 64, "ols"), ("rho", 8, 33, "wls")]


@pytest.mark.parametrize("kind,level,p,method", WIDE, ids=[f"{w[0]}-{w[1]}-p{w[2]}-{w[3]}" for w in WIDE])
def test_conditioned_wide_groups(lib, oracle, kind, level, p, method):
    """9 <= p <= 64: the ring accumulation with the fixed / mid / wide solve kernels, all three methods."""
    rng = np.random.default_rng(9500 + p)
    c = make_conditioned(rng, kind, level, G=24, n=4 * p + 40, p=p, noise=1e-3, weights=(method == "wls"))
    opt = dict(method=method, compute_inference=True, alpha=1e-3 if method == "ridge" else 0.0)
    ora = oracle.fit_batch(c["offsets"], c["y"], c["x"], c["w"], **opt)
    assert (ora["status"] == 0).all()
    tol, kap = group_tolerances(c, ora)
    for path in ("auto", "qr"):
        lib.force_path(path)
        try:
            dev = lib.fit_batch_host(c["offsets"], c["y"], c["x"], c["w"], **opt)
        finally:
            lib.force_path("auto")
        assert_fit_parity(dev, ora, inference=True, label=f"{kind}-{level} p={p} {method} [{path}] kappa<={kap.max():.3g}", **tol)


def test_fallback_is_taken_where_needed_and_only_there(lib, oracle):
    """Device counters: well-conditioned synthetic data of the BASELINE shapes queue (almost) nothing for the QR
    fallback; ill-conditioned groups are queued; duplicated columns (exact rank deficiency, the C3 mix) are not."""
    import torch
    synth = __import__("anofox_statistics_b200.synth", fromlist=["synth"])

    def run(c, p, **opt):
        G, N = len(c["offsets"]) - 1, len(c["y"])
        f = lib.DeviceFit(G, N, p, inference=bool(opt.get("compute_inference")))
        t = {k: (None if c[k] is None else torch.tensor(c[k], device="cuda")) for k in ("offsets", "y", "x", "w")}
        f.fit(t["offsets"], t["y"], t["x"], t["w"], **opt)
        return f.counters()

    d = synth.make_numpy(20000, 50, 5, seed=2)                       # C2's distribution
    assert run(d, 5, method="ols", compute_inference=True)["qr"] == 0
    d = synth.make_numpy(4000, 200, 16, seed=3, weights=True, rank_deficient=True)  # C3's mix: constant + duplicated columns
    assert run(d, 16, method="wls")["qr"] <= 4
    assert run(d, 16, method="ridge", alpha=1.0)["qr"] == 0
    rng = np.random.default_rng(5)
    c = make_conditioned(rng, "rho", 8, G=64, n=40, p=5)
    assert run(c, 5, method="ols", compute_inference=True)["qr"] == 64
    c = make_conditioned(rng, "nearcol", 1.5e-7, G=64, n=40, p=5)    # inside the decision band of the aliasing rule
    assert run(c, 5, method="ols")["qr"] == 64
    c = make_conditioned(rng, "offset", 6, G=64, n=40, p=5)          # large offsets alone: the shift row removes them
    assert run(c, 5, method="ols")["qr"] == 0
