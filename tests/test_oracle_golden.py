"""Pin the CPU oracle against every golden vector / known answer the reference holds for this path.

R fixtures: /root/reference/test/data/{ols,wls,ridge,inference}_tests (tolerances from their
metadata.json: 1e-10 coefficients / R^2, 1e-8 p-values and intervals). Known answers: transcribed
from test/sql/ols_validation.test and validation/r_seesion_output.txt (tests/golden/make_golden.py).
"""
import json
import math
import os

import numpy as np
import pytest
from scipy import stats

HERE = os.path.dirname(os.path.abspath(__file__))
R_FIX = json.load(open(os.path.join(HERE, "golden", "r_fixtures.json")))["cases"]
KAT = json.load(open(os.path.join(HERE, "golden", "kat.json")))["cases"]


def fit_kwargs(c):
    kw = dict(fit_intercept=c["fit_intercept"], compute_inference=c.get("compute_inference", False),
              confidence_level=c.get("confidence_level", 0.95))
    if c["method"] == "ridge":
        kw.update(alpha=c["alpha"], lambda_scaling=c.get("lambda_scaling", "raw"))
    return kw


def close(a, b, tol):
    return abs(a - b) <= tol * max(1.0, abs(b))


def check_against_r(r, c):
    """Shared by the oracle tests (CPU) and the device tests (GPU): r is a SingleFit-like object."""
    e, tol = c["expected"], c["tol"]
    assert r.ok
    if e.get("intercept") is None:
        assert math.isnan(r.intercept)  # ols.rs:172-176 -> NaN when no intercept
    else:
        assert close(r.intercept, e["intercept"], tol), (r.intercept, e["intercept"])
    assert len(r.coefficients) == len(e["coefficients"])
    for got, want in zip(r.coefficients, e["coefficients"]):
        if want is None:
            assert math.isnan(got)  # aliased / constant column -> NaN at the same position as R's NA
        else:
            assert close(got, want, tol), (got, want)
    for k in ("r_squared", "adj_r_squared", "residual_std_error"):
        if k in e:
            assert close(getattr(r, k), e[k], tol), (k, getattr(r, k), e[k])
    if "n_observations" in e:
        assert r.n_observations == e["n_observations"]
    if c.get("compute_inference"):
        ti = c["tol_inference"]
        for k in ("std_errors", "t_values", "ci_lower", "ci_upper"):
            for got, want in zip(getattr(r, k), e[k]):
                assert close(got, want, ti), (k, got, want)
        for got, want in zip(r.p_values, e["p_values"]):
            assert abs(got - want) <= ti * want, (got, want)  # relative: p-values go down to 1e-85
        assert close(r.f_statistic, e["f_statistic"], ti)
        _, d1, d2 = c["model_stats_raw"]["fstatistic"]
        want = stats.f.sf(e["f_statistic"], d1, d2)
        assert abs(r.f_pvalue - want) <= 1e-8 * want


@pytest.mark.parametrize("c", R_FIX, ids=[c["name"] for c in R_FIX])
def test_oracle_matches_r_golden(oracle, c):
    r = oracle.single_fit(c["method"], c["y"], c["x"], c["w"], **fit_kwargs(c))
    check_against_r(r, c)
    if "x_new" in c:  # prediction fixture: yhat is a pin; the interval follows lib.rs:2340-2347
        e = c["expected"]
        n = len(c["y"])
        for i, want in enumerate(e["yhat"]):
            ok, (yh, lo, hi) = oracle.predict_with_interval(r.coefficients, r.intercept, [c["x_new"][0][i]],
                                                            r.residual_std_error, n, 0.95)
            assert ok and close(yh, want, 1e-10)
            margin = stats.t.ppf(0.975, n - 2) * r.residual_std_error * math.sqrt(1 + 1 / n)
            assert close(lo, yh - margin, 1e-10) and close(hi, yh + margin, 1e-10)
            # R's exact interval adds the leverage term: the simplified one is narrower (by up to ~5 % at the extrapolated x = 20)
            assert e["r_prediction_lower"][i] <= lo <= yh <= hi <= e["r_prediction_upper"][i]
            assert (hi - lo) > 0.94 * (e["r_prediction_upper"][i] - e["r_prediction_lower"][i])


def check_kat(r, c):
    ex = c["expect"]
    assert r.ok
    for k, v in ex.items():
        if k == "intercept_is_nan":
            assert math.isnan(r.intercept)
        elif k.endswith("_is_nan"):
            assert math.isnan(r.coefficients[int(k[4])])
        elif k == "n_nan_coef":
            assert int(np.isnan(r.coefficients).sum()) == v
        elif k == "rse_below":
            assert r.residual_std_error < v
        elif k == "n_observations":
            assert r.n_observations == v
        else:
            absolute = k.endswith("_abs")
            name = k[:-4] if absolute else k
            got = r.coefficients[int(name[4])] if name.startswith("coef") else getattr(r, name)
            if absolute:
                assert abs(got - v[0]) < v[1], (k, got, v)
            else:  # the .test files compare round(value, digits)
                assert round(got, v[1]) + 0.0 == v[0], (k, got, v)


@pytest.mark.parametrize("c", KAT, ids=[c["name"] for c in KAT])
def test_oracle_known_answers(oracle, c):
    r = oracle.single_fit(c["method"], c["y"], c["x"], None, fit_intercept=c["fit_intercept"],
                          alpha=c.get("alpha", 0.0))
    check_kat(r, c)


def test_oracle_special_functions_vs_scipy(oracle):
    rng = np.random.default_rng(0)
    for df in [1, 2, 3, 5, 10, 44, 98, 146, 183, 1000, 100000]:
        for t in np.concatenate([rng.uniform(0, 6, 20), [1e-3, 0.5, 12.0, 40.0, 300.0]]):
            want = 2 * stats.t.sf(t, df)
            got = oracle.t_two_sided_p(t, df)
            assert abs(got - want) <= 1e-11 * want + 1e-300, (t, df, got, want)
        for lvl in [0.5, 0.8, 0.9, 0.95, 0.99, 0.999]:
            want = stats.t.ppf((1 + lvl) / 2, df)
            got = oracle.t_critical(lvl, df)
            assert abs(got - want) <= 1e-10 * want, (lvl, df, got, want)
    for d1, d2 in [(1, 5), (3, 146), (5, 44), (16, 183), (8, 991), (2, 2)]:
        for f in [1e-3, 0.5, 1.0, 3.0, 25.0, 1e3, 1e5]:
            want = stats.f.sf(f, d1, d2)
            got = oracle.f_upper_p(f, d1, d2)
            assert abs(got - want) <= 1e-10 * want + 1e-300, (f, d1, d2, got, want)
    # tiny |t| (scipy's t.sf loses digits there): closed form for df = 1, mpmath otherwise
    assert abs(oracle.t_two_sided_p(1e-8, 1) - (1 - 2 / math.pi * math.atan(1e-8))) < 1e-15
    import mpmath
    mpmath.mp.dps = 40
    for t, df in [(1e-8, 7), (1e-5, 44), (60.0, 44), (25.0, 3)]:
        x = mpmath.mpf(df) / (df + mpmath.mpf(t) ** 2)
        want = float(mpmath.betainc(mpmath.mpf(df) / 2, 0.5, 0, x, regularized=True))
        assert abs(oracle.t_two_sided_p(t, df) - want) <= 1e-13 * want
    assert math.isnan(oracle.t_critical(0.95, 0))  # lib.rs:2218-2220
    assert math.isnan(oracle.t_critical(1.0, 5)) and math.isnan(oracle.t_critical(0.0, 5))


def test_oracle_error_taxonomy(oracle):
    """ols.rs:38-139, ridge.rs:38-40, wls.rs:76-90, lib.rs:65-84."""
    A = oracle.abi
    y = [1.0, 2.0, 3.5, 4.0]
    x = [[1.0, 2.0, 3.0, 4.5]]
    assert oracle.single_fit("ridge", y, x, alpha=-1.0).error_code == A.INVALID_ALPHA
    assert oracle.single_fit("ols", y, []).error_code == A.INVALID_INPUT
    assert oracle.single_fit("ols", [], [[]]).error_code == A.INVALID_INPUT
    assert oracle.single_fit("ols", y, [[1.0, 2.0]]).error_code == A.DIMENSION_MISMATCH
    nan = float("nan")
    assert oracle.single_fit("ols", [nan] * 4, x).error_code == A.NO_VALID_DATA
    assert oracle.single_fit("wls", y, x, [0.0, -1.0, nan, float("inf")]).error_code == A.NO_VALID_DATA
    # n_valid < p_eff + intercept
    r = oracle.single_fit("ols", [1.0, 2.0], [[1.0, 2.0], [2.0, 5.0], [1.0, 7.0]])
    assert r.error_code == A.INSUFFICIENT_DATA
    # exact fit n == p + 1 allowed (ols.rs:132-139)
    r = oracle.single_fit("ols", [1.0, 2.0, 4.0], [[1.0, 2.0, 3.0], [1.0, 0.0, 2.0]])
    assert r.ok and r.n_observations == 3
    # all columns constant: intercept-only shortcut (ols.rs:101-130) / error without intercept
    r = oracle.single_fit("ols", y, [[5.0] * 4], compute_inference=True)
    assert r.ok and math.isnan(r.coefficients[0]) and r.r_squared == 0.0 and r.adj_r_squared == 0.0
    assert abs(r.intercept - np.mean(y)) < 1e-15 and abs(r.residual_std_error - np.std(y, ddof=1)) < 1e-15
    assert r.inference_len == 0 and math.isnan(r.f_statistic)  # lib.rs:251-253
    assert oracle.single_fit("ols", y, [[5.0] * 4], fit_intercept=False).error_code == A.INSUFFICIENT_DATA
    # zero / negative weights are dropped (wls.rs tests: n_observations == 8 of 10)
    yy = np.arange(10.0) * 2 + 1
    xx = [np.arange(10.0)]
    w = [1.0] * 10
    w[3], w[7] = 0.0, -2.0
    r = oracle.single_fit("wls", yy, xx, w)
    assert r.ok and r.n_observations == 8
    # NULL (validity bit 0) is NaN is dropped (types.rs:79-89)
    r = oracle.single_fit("ols", yy, xx, y_validity=[1] * 9 + [0])
    assert r.ok and r.n_observations == 9


def test_oracle_information_criteria(oracle):
    import ctypes as C
    lib, A = oracle.lib(), oracle.abi
    out, err = C.c_double(), A.AnofoxError()
    assert lib.oracle_compute_aic(10.0, 100, 3, C.byref(out), C.byref(err))
    assert abs(out.value - (100 * math.log(0.1) + 6)) < 1e-12
    assert lib.oracle_compute_bic(10.0, 100, 3, C.byref(out), C.byref(err))
    assert abs(out.value - (100 * math.log(0.1) + 3 * math.log(100))) < 1e-12
    assert lib.oracle_compute_aic(0.0, 10, 2, C.byref(out), C.byref(err)) and out.value == -math.inf
    assert not lib.oracle_compute_aic(1.0, 0, 2, C.byref(out), C.byref(err)) and err.code == A.INVALID_INPUT
    assert not lib.oracle_compute_bic(-1.0, 5, 2, C.byref(out), C.byref(err))
