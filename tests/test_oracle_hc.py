"""HC0 .. HC3 robust standard errors of the oracle against an independent numpy restatement of the published
estimators (MacKinnon & White 1985; R sandwich::vcovHC for weighted fits). The reference delegates to an
un-vendored crate (ols.rs:209-245, wls.rs:230-266) and its tests only ask for non-NULL lists
(test/sql/regression/test_map_options.test:88-157), so this pins the oracle to the textbook definition."""
import numpy as np
import pytest
from scipy import stats


def numpy_hc(y, X, w, hc, intercept=True, level=0.95):
    n, p = X.shape
    keep = [j for j in range(p) if not np.all(np.abs(X[:, j] - X[0, j]) < 1e-10)]
    Z = np.column_stack(([np.ones(n)] if intercept else []) + [X[:, j] for j in keep])
    sw = np.sqrt(w)
    Zt, yt = Z * sw[:, None], y * sw
    beta, *_ = np.linalg.lstsq(Zt, yt, rcond=None)
    e = yt - Zt @ beta
    k = Z.shape[1]
    bread = np.linalg.inv(Zt.T @ Zt)
    h = np.einsum("ij,jk,ik->i", Zt, bread, Zt)
    om = {"hc0": e ** 2, "hc1": e ** 2 * n / (n - k), "hc2": e ** 2 / (1 - h), "hc3": e ** 2 / (1 - h) ** 2}[hc]
    V = bread @ (Zt.T * om) @ Zt @ bread
    se = np.sqrt(np.diag(V))
    t = beta / se
    df = n - k
    pv = 2 * stats.t.sf(np.abs(t), df)
    tc = stats.t.ppf((1 + level) / 2, df)
    ic = 1 if intercept else 0
    out = {q: np.full(p, np.nan) for q in ("coef", "se", "t", "p", "lo", "hi")}
    for c, j in enumerate(keep):
        out["coef"][j], out["se"][j], out["t"][j], out["p"][j] = beta[c + ic], se[c + ic], t[c + ic], pv[c + ic]
        out["lo"][j], out["hi"][j] = beta[c + ic] - tc * se[c + ic], beta[c + ic] + tc * se[c + ic]
    return out


@pytest.mark.parametrize("hc", ["hc0", "hc1", "hc2", "hc3"])
@pytest.mark.parametrize("method", ["ols", "wls"])
@pytest.mark.parametrize("intercept", [True, False])
def test_oracle_hc_matches_numpy(oracle, hc, method, intercept):
    rng = np.random.default_rng(11)
    n, p = 60, 4
    X = rng.normal(size=(n, p)) + 2.0
    X[:, 2] = 1.5  # a constant column: NaN everywhere
    y = 1.0 + X[:, 0] - 2 * X[:, 1] + rng.normal(size=n) * (0.2 + np.abs(X[:, 0]))  # heteroscedastic noise
    w = rng.uniform(0.5, 2.0, n) if method == "wls" else np.ones(n)
    off = np.array([0, n], dtype=np.int64)
    got = oracle.fit_batch(off, y, X, w if method == "wls" else None, method=method, compute_inference=True,
                           fit_intercept=intercept, hc_type=hc)
    want = numpy_hc(y, X, w, hc, intercept)
    assert got["status"][0] == 0
    for q, key in (("coef", "coefficients"), ("se", "std_errors"), ("t", "t_values"), ("p", "p_values"),
                   ("lo", "ci_lower"), ("hi", "ci_upper")):
        g = got[key].reshape(1, p)[0]
        assert np.array_equal(np.isnan(g), np.isnan(want[q])), (q, g, want[q])
        m = ~np.isnan(g)
        assert np.allclose(g[m], want[q][m], rtol=2e-9, atol=1e-300), (q, g, want[q])
    # classical F statistic is passed through unchanged (ols.rs:222-223)
    base = oracle.fit_batch(off, y, X, w if method == "wls" else None, method=method, compute_inference=True,
                            fit_intercept=intercept)
    assert got["f_statistic"][0] == base["f_statistic"][0]
    assert not np.allclose(got["std_errors"][~np.isnan(got["std_errors"])],
                           base["std_errors"][~np.isnan(base["std_errors"])], rtol=1e-3)
