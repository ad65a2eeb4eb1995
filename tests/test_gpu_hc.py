"""HC0 .. HC3 robust standard errors (csrc/hc.cuh) against the oracle, which is itself
pinned to an independent numpy restatement of the estimators (tests/test_oracle_hc.py). Reference:
crates/anofox-stats-core/src/models/ols.rs:209-245, wls.rs:230-266."""
import numpy as np
import pytest
from parity import assert_fit_parity

pytestmark = pytest.mark.gpu

SHAPES = [(300, 40, 3), (257, 50, 5), (64, 200, 16), (12, 400, 40), (5, 300, 64)]


@pytest.fixture(scope="module")
def lib(b200):
    b200.device_check()
    return b200


@pytest.mark.parametrize("hc", ["hc0", "hc1", "hc2", "hc3"])
@pytest.mark.parametrize("method", ["ols", "wls"])
@pytest.mark.parametrize("G,n,p", SHAPES, ids=[f"p{s[2]}" for s in SHAPES])
def test_hc_parity_vs_oracle(lib, oracle, b200, hc, method, G, n, p):
    from importlib import import_module
    synth = import_module("anofox_statistics_b200.synth")
    d = synth.make_numpy(G, n, p, seed=100 + p, ragged=True, weights=(method == "wls"), rank_deficient=True)
    # heteroscedastic noise, so that the robust and the classical standard errors really differ
    rng = np.random.default_rng(p)
    d["y"] = d["y"] + rng.normal(size=d["y"].shape) * np.abs(d["x"][:, 0])
    d["y"][::37] = np.nan  # NULL responses: rows dropped by both passes
    opt = dict(method=method, compute_inference=True, hc_type=hc)
    dev = lib.fit_batch_host(d["offsets"], d["y"], d["x"], d["w"], **opt)
    ora = oracle.fit_batch(d["offsets"], d["y"], d["x"], d["w"], **opt)
    assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=f"{hc}/{method}/p{p}")
    base = lib.fit_batch_host(d["offsets"], d["y"], d["x"], d["w"], method=method, compute_inference=True)
    m = np.isfinite(base["std_errors"])
    assert np.abs(dev["std_errors"][m] / base["std_errors"][m] - 1).max() > 1e-3  # not the classical numbers
    assert np.array_equal(dev["f_statistic"], base["f_statistic"], equal_nan=True)  # ols.rs:222-223
    assert np.array_equal(dev["coefficients"], base["coefficients"], equal_nan=True)


@pytest.mark.parametrize("hc", ["hc1", "hc3"])
def test_hc_without_intercept_and_single_fit(lib, oracle, hc):
    rng = np.random.default_rng(3)
    n, p = 80, 3
    x = rng.normal(size=(n, p)) + 1.0
    y = x @ np.array([1.0, -2.0, 0.5]) + rng.normal(size=n) * (0.1 + np.abs(x[:, 1]))
    off = np.array([0, n], dtype=np.int64)
    for ic in (True, False):
        opt = dict(method="ols", compute_inference=True, hc_type=hc, fit_intercept=ic)
        dev = lib.fit_batch_host(off, y, x, None, **opt)
        ora = oracle.fit_batch(off, y, x, None, **opt)
        assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=f"{hc}/ic={ic}")
    # the drop-in symbol takes hc_type in its options struct (anofox_stats_ffi.h:130-141)
    r = lib.single_fit("ols", y, [x[:, j] for j in range(p)], None, compute_inference=True, hc_type=hc)
    o = oracle.single_fit("ols", y, [x[:, j] for j in range(p)], None, compute_inference=True, hc_type=hc)
    assert r.ok and o.ok
    assert np.allclose(r.std_errors, o.std_errors, rtol=1e-9) and np.allclose(r.p_values, o.p_values, rtol=1e-8)


def test_ridge_ignores_hc_and_bad_value_is_an_error(lib, b200):
    rng = np.random.default_rng(5)
    x = rng.normal(size=(60, 2))
    y = x[:, 0] + rng.normal(size=60)
    off = np.array([0, 60], dtype=np.int64)
    a = lib.fit_batch_host(off, y, x, None, method="ridge", alpha=1.0, compute_inference=True, hc_type="hc3")
    b = lib.fit_batch_host(off, y, x, None, method="ridge", alpha=1.0, compute_inference=True)
    assert np.array_equal(a["std_errors"], b["std_errors"])  # ridge.rs has no hc_type


@pytest.mark.parametrize("hc,method", [("hc1", "ols"), ("hc3", "wls")])
def test_hc_small_features_with_long_groups_mixed_in(lib, oracle, hc, method):
    """p <= 8: groups of up to 2 048 rows take the four-lanes-per-group kernel, longer ones are listed by it and
    handled by the CTA-per-group kernel (list mode) -- both kinds in one batch, and the boundary sizes themselves."""
    rng = np.random.default_rng(31)
    sizes = np.array([100, 3000, 50, 2500, 2048, 2049, 7, 2047])
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    N, p = int(off[-1]), 4
    x = rng.normal(size=(N, p)) * 2 + 1
    y = 0.5 + x @ np.array([1.0, -2.0, 0.25, 3.0]) + rng.normal(size=N) * (0.2 + np.abs(x[:, 1]))
    x[int(off[3]):int(off[4]), 2] = 1.75  # a constant column inside one of the long groups
    y[::53] = np.nan
    w = rng.uniform(0.5, 2.0, size=N) if method == "wls" else None
    opt = dict(method=method, compute_inference=True, hc_type=hc)
    dev = lib.fit_batch_host(off, y, x, w, **opt)
    ora = oracle.fit_batch(off, y, x, w, **opt)
    assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=f"{hc}/{method}/long groups")
