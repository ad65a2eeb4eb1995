"""Window form of the fit-predict aggregates on the device (csrc/window.cuh) against the
oracle, which refits every output row's frame from its raw rows as the reference does
(src/window_functions/ols_fit_predict.cpp:243-326)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(b200):
    b200.device_check()
    return b200


def make(G, n, p, seed, weights=False, holes=True, ragged=True):
    rng = np.random.default_rng(seed)
    sizes = rng.integers(max(n // 2, 1), n + n // 2 + 1, size=G) if ragged else np.full(G, n)
    off = np.zeros(G + 1, dtype=np.int64)
    np.cumsum(sizes, out=off[1:])
    N = int(off[-1])
    x = rng.uniform(-10, 10, size=(N, p))
    gid = np.repeat(np.arange(G), sizes)
    beta = rng.uniform(-5, 5, size=(G, p))
    y = rng.uniform(-10, 10, size=G)[gid] + np.einsum("ij,ij->i", x, beta[gid]) + rng.normal(0, 2.0, size=N)
    if p >= 3:
        cst = rng.uniform(size=G) < 0.1
        x[cst[gid], 1] = 3.25  # a constant column in some partitions
        dup = rng.uniform(size=G) < 0.1
        x[dup[gid], p - 1] = x[dup[gid], 0]  # a duplicated column in others
    if holes:
        y[rng.uniform(size=N) < 0.1] = np.nan  # rows to predict only (y IS NULL)
        x[rng.uniform(size=N) < 0.01, 0] = np.nan  # NULL feature: dropped inside the fit, NaN prediction
    w = rng.uniform(0.5, 1.5, size=N) if weights else None
    return off, y, x, w


def check(dev, ora, label, rtol=1e-9):
    for k in ("yhat", "yhat_lower", "yhat_upper"):
        a, b = dev[k], ora[k]
        assert np.array_equal(np.isnan(a), np.isnan(b)), (label, k, np.flatnonzero(np.isnan(a) != np.isnan(b))[:10])
        m = np.isfinite(b)
        if m.any():
            scale = np.maximum(np.abs(b[m]), 1e-3 * np.abs(ora["yhat"][m]).max())
            err = np.abs(a[m] - b[m]) / scale
            assert err.max() <= rtol, (label, k, err.max(), int(np.flatnonzero(m)[err.argmax()]))


FRAMES = [(-1, 1), (-1, 0), (4, 0), (7, 3), (-1, 3), (15, 1)]


@pytest.mark.parametrize("preceding,end_offset", FRAMES, ids=[f"prec{a}_end{b}" for a, b in FRAMES])
@pytest.mark.parametrize("method", ["ols", "ridge", "wls"])
@pytest.mark.parametrize("p", [1, 3, 5, 8])
def test_window_parity_vs_oracle(lib, oracle, p, method, preceding, end_offset):
    off, y, x, w = make(40, 60, p, seed=1000 + 17 * p + end_offset, weights=(method == "wls"))
    if method == "wls":
        w[::29] = np.nan  # NULL weight: the row does not train (wls_fit_predict.cpp:152)
    opt = dict(method=method, alpha=0.7 if method == "ridge" else 0.0)
    dev = lib.window_fit_predict_host(off, y, x, w, preceding=preceding, end_offset=end_offset, **opt)
    ora = oracle.window_fit_predict(off, y, x, w, preceding=preceding, end_offset=end_offset, **opt)
    check(dev, ora, f"{method} p={p} frame=({preceding},{end_offset})", rtol=2e-9)


def test_window_reference_kat(lib):
    # test/sql/fit_predict/test_ols_fit_predict_basic.test: y = 2 i + 1 for i <= 6, NULL after; OVER (ORDER BY id)
    i = np.arange(1, 11, dtype=np.float64)
    y = np.where(i <= 6, 2.0 * i + 1.0, np.nan)
    off = np.array([0, 10], dtype=np.int64)
    r = lib.window_fit_predict_host(off, y, i.reshape(-1, 1), None, preceding=-1, end_offset=0, method="ols")
    ok = ~np.isnan(r["yhat"])
    assert ok.sum() == 8 and ok[2:].all()
    assert np.allclose(r["yhat"][ok], 2.0 * i[ok] + 1.0, rtol=1e-12)
    assert (r["yhat_upper"][ok] >= r["yhat_lower"][ok]).all()
    # exact fit: the interval collapses to (nearly) nothing relative to yhat
    assert np.abs(r["yhat_upper"][ok] - r["yhat"][ok]).max() < 1e-9
    r0 = lib.window_fit_predict_host(off, y, i.reshape(-1, 1), None, preceding=-1, end_offset=0, method="ols",
                                     fit_intercept=False)
    assert (~np.isnan(r0["yhat"])).sum() == 9


def test_window_benchmark_shape(lib, oracle):
    """examples/performance_1m_groups/benchmark_ols.sql at reduced scale: many partitions x 100 rows x 3 features,
    UNBOUNDED PRECEDING .. 1 PRECEDING; all partitions the same length (the generator's i % G layout)."""
    off, y, x, w = make(3000, 100, 3, seed=7, holes=False, ragged=False)
    dev = lib.window_fit_predict_host(off, y, x, None, preceding=-1, end_offset=1, method="ols")
    ora = oracle.window_fit_predict(off, y, x, None, preceding=-1, end_offset=1, method="ols", precision=1)
    check(dev, ora, "benchmark shape", rtol=2e-9)
    assert (~np.isnan(dev["yhat"])).sum() == 3000 * (100 - 5)  # rows 0..4 of a partition have <= 4 training rows


def test_window_rejects_wide_designs(lib, b200):
    off = np.array([0, 30], dtype=np.int64)
    with pytest.raises(b200.B200Error):
        lib.window_fit_predict_host(off, np.zeros(30), np.zeros((30, 9)), None, method="ols")


def test_window_degenerate_inputs(lib, oracle):
    """Empty input, empty partitions between full ones, one-row partitions, a frame that never holds enough rows,
    and the argument checks of the entry point (frame end, feature limit)."""
    # no rows at all
    dev = lib.window_fit_predict_host(np.zeros(1, dtype=np.int64), np.zeros(0), np.zeros((0, 2)), None, preceding=-1, end_offset=1)
    assert all(dev[k].size == 0 for k in ("yhat", "yhat_lower", "yhat_upper"))
    # empty and one-row partitions next to ordinary ones
    rng = np.random.default_rng(77)
    sizes = np.array([0, 1, 0, 25, 1, 2, 0, 40, 3, 0])
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    N = int(off[-1])
    x = rng.uniform(-3, 3, size=(N, 2))
    y = 1.0 + x @ np.array([2.0, -1.0]) + rng.normal(0, 0.5, size=N)
    for prec, end in ((-1, 1), (1, 0), (2, 1)):  # the last two frames never reach p + 2 rows: every prediction NULL
        dev = lib.window_fit_predict_host(off, y, x, None, preceding=prec, end_offset=end)
        ora = oracle.window_fit_predict(off, y, x, None, preceding=prec, end_offset=end)
        check(dev, ora, f"degenerate prec={prec} end={end}")
        if prec > 0:
            assert np.isnan(dev["yhat"]).all()
    # argument checks
    with pytest.raises(lib.B200Error, match="frame end"):
        lib.window_fit_predict_host(off, y, x, None, preceding=-1, end_offset=-1)
    with pytest.raises(lib.B200Error, match="up to 8 features"):
        lib.window_fit_predict_host(np.array([0, 20], dtype=np.int64), np.zeros(20), np.zeros((20, 9)), None, preceding=-1, end_offset=1)
