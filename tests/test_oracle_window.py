"""The oracle's window form ({ols,ridge,wls}_fit_predict OVER (...)) against the reference's own known answers
(test/sql/fit_predict/test_ols_fit_predict_basic.test, test_ridge_fit_predict_rolling.test) and against a direct
numpy restatement of one frame."""
import numpy as np


def _ols_data():
    i = np.arange(1, 11, dtype=np.float64)
    y = np.where(i <= 6, 2.0 * i + 1.0, np.nan)
    return np.array([0, 10], dtype=np.int64), y, i.reshape(-1, 1)


def test_reference_kat_ols_fit_predict_basic(oracle):
    off, y, x = _ols_data()
    # OVER (ORDER BY id): default frame = UNBOUNDED PRECEDING .. CURRENT ROW
    r = oracle.window_fit_predict(off, y, x, None, preceding=-1, end_offset=0, method="ols")
    ok = ~np.isnan(r["yhat"])
    assert ok.sum() == 8 and ok[2:].all() and not ok[:2].any()      # tests 2 and 3
    assert (r["yhat_upper"][ok] >= r["yhat_lower"][ok]).all()        # test 4
    assert np.allclose(r["yhat"][ok], 2.0 * x[ok, 0] + 1.0, rtol=1e-12)  # exact line, extrapolated to the NULL-y rows
    r0 = oracle.window_fit_predict(off, y, x, None, preceding=-1, end_offset=0, method="ols", fit_intercept=False)
    assert (~np.isnan(r0["yhat"])).sum() == 9                        # test 5


def _ridge_series():
    # test_ridge_fit_predict_rolling.test: 15 rows, y = 2 x + 5 + noise-free pattern is not needed for the counts
    t = np.arange(1, 16, dtype=np.float64)
    y = 2.0 * t + 5.0 + np.sin(t)
    return np.array([0, 15], dtype=np.int64), y, t.reshape(-1, 1)


def test_frame_counts_follow_the_min_obs_rule(oracle):
    off, y, x = _ridge_series()
    opt = dict(method="ridge", alpha=0.1)
    # expanding, current row: NULL until more than p + intercept = 2 training rows -> 13 of 15
    r = oracle.window_fit_predict(off, y, x, None, preceding=-1, end_offset=0, **opt)
    assert (~np.isnan(r["yhat"])).sum() == 13
    # ROWS BETWEEN 4 PRECEDING AND CURRENT ROW -> 13 as well (frames of 3+ rows from row 3 on)
    r = oracle.window_fit_predict(off, y, x, None, preceding=4, end_offset=0, **opt)
    assert (~np.isnan(r["yhat"])).sum() == 13
    # ROWS BETWEEN 7 PRECEDING AND 3 PRECEDING: frames [i-7, i-3]; at least 3 rows from i = 6 on (0-based 5)
    r = oracle.window_fit_predict(off, y, x, None, preceding=7, end_offset=3, **opt)
    assert (~np.isnan(r["yhat"][7:])).sum() == 8 and np.isnan(r["yhat"][:5]).all()


def test_one_frame_against_numpy(oracle):
    rng = np.random.default_rng(4)
    n, p = 30, 3
    x = rng.normal(size=(n, p))
    y = 1.0 + x @ np.array([0.5, -1.0, 2.0]) + 0.1 * rng.normal(size=n)
    y[5] = np.nan
    off = np.array([0, n], dtype=np.int64)
    r = oracle.window_fit_predict(off, y, x, None, preceding=-1, end_offset=1, method="ols")
    i = 20  # frame = rows 0..19, predicted at x[19] (the x of the last frame row, ols_fit_predict.cpp:157-163)
    tr = [q for q in range(0, 20) if not np.isnan(y[q])]
    Z = np.column_stack([np.ones(len(tr)), x[tr]])
    b, *_ = np.linalg.lstsq(Z, y[tr], rcond=None)
    assert np.isclose(r["yhat"][i], b[0] + x[19] @ b[1:], rtol=1e-10)
    from scipy import stats
    e = y[tr] - Z @ b
    rse = np.sqrt(e @ e / (len(tr) - 4))
    m = stats.t.ppf(0.975, len(tr) - 4) * rse * np.sqrt(1 + 1 / len(tr))
    assert np.isclose(r["yhat_upper"][i] - r["yhat"][i], m, rtol=1e-9)
    assert np.isnan(r["yhat"][0])  # empty frame
