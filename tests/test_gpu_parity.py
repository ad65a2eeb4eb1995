"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on identical inputs.

Bit-exact: per-group status (NULL-struct pattern), n_observations, NaN/NULL positions.
Values: relative 1e-9 (north star: "1e-9 for well-conditioned f64"); p-values relative 1e-8.
Run on a B200 with:  python -m pytest tests -m gpu
"""
import json
import math
import os

import numpy as np
import pytest

from parity import assert_fit_parity, max_rel_err
from test_oracle_golden import KAT, R_FIX, check_against_r, check_kat, fit_kwargs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(b200):
    b200.device_check()  # fails loudly without an sm_100 device: there is no fallback
    return b200


@pytest.fixture(params=["auto", "stash", "record", "fused", "stages", "qr", "resident"])
def path(request, lib):
    """The kernel paths: "auto" takes the small-group kernels where they apply (p <= 8, small groups), "record"
    forces the DMMA accumulation kernel + solve kernel everywhere, "fused" / "stages" are the A/B variants, and
    "qr" is "auto" followed by the QR fallback kernel (csrc/qr.cuh) re-fitting EVERY group from its rows."""
    lib.force_path(request.param)
    yield request.param
    lib.force_path("auto")


# ------------------------------------------------------------------ drop-in symbols vs goldens
@pytest.mark.parametrize("c", R_FIX, ids=[c["name"] for c in R_FIX])
def test_single_fit_matches_r_golden(lib, path, c):
    r = lib.single_fit(c["method"], c["y"], c["x"], c["w"], **fit_kwargs(c))
    check_against_r(r, c)


@pytest.mark.parametrize("c", KAT, ids=[c["name"] for c in KAT])
def test_single_fit_known_answers(lib, path, c):
    r = lib.single_fit(c["method"], c["y"], c["x"], None, fit_intercept=c["fit_intercept"], alpha=c.get("alpha", 0.0))
    check_kat(r, c)


def test_single_fit_error_taxonomy(lib, oracle, path):
    """Same calls as test_oracle_error_taxonomy: identical ok / error code from both implementations."""
    nan, inf = float("nan"), float("inf")
    y = [1.0, 2.0, 3.5, 4.0]
    x = [[1.0, 2.0, 3.0, 4.5]]
    yy = np.arange(10.0) * 2 + 1 + np.sin(np.arange(10.0))
    xx = [np.arange(10.0)]
    w = [1.0] * 10
    w[3], w[7] = 0.0, -2.0
    cases = [
        ("ridge", y, x, None, dict(alpha=-1.0)),
        ("ols", y, [], None, {}),
        ("ols", [], [[]], None, {}),
        ("ols", y, [[1.0, 2.0]], None, {}),
        ("ols", [nan] * 4, x, None, {}),
        ("wls", y, x, [0.0, -1.0, nan, inf], {}),
        ("ols", [1.0, 2.0], [[1.0, 2.0], [2.0, 5.0], [1.0, 7.0]], None, {}),
        ("ols", [1.0, 2.0, 4.0], [[1.0, 2.0, 3.0], [1.0, 0.0, 2.0]], None, {}),
        ("ols", y, [[5.0] * 4], None, dict(compute_inference=True)),
        ("ols", y, [[5.0] * 4], None, dict(fit_intercept=False)),
        ("wls", yy, xx, w, {}),
        ("ols", yy, xx, None, dict(y_validity=[1] * 9 + [0])),
        ("ols", [3.0], [[1.0]], None, {}),
        ("ols", [3.0], [[1.0]], None, dict(fit_intercept=False)),
        ("wls", [3.0], [[1.0]], [2.0], {}),
    ]
    for method, yv, xv, wv, kw in cases:
        a = lib.single_fit(method, yv, xv, wv, **kw)
        b = oracle.single_fit(method, yv, xv, wv, **kw)
        assert a.ok == b.ok and a.error_code == b.error_code, (method, kw, a.error_code, b.error_code, a.error_message)
        if a.ok:
            assert a.n_observations == b.n_observations
            assert np.array_equal(np.isnan(a.coefficients), np.isnan(b.coefficients))
            assert max_rel_err(a.coefficients, b.coefficients) < 1e-9
            for k in ("intercept", "r_squared", "adj_r_squared", "residual_std_error"):
                va, vb = getattr(a, k), getattr(b, k)
                assert (math.isnan(va) and math.isnan(vb)) or abs(va - vb) <= 1e-9 * max(1, abs(vb)), (k, va, vb)
            if kw.get("compute_inference"):
                assert a.inference_len == b.inference_len


def test_free_functions_null_the_pointers(lib):
    import ctypes as C
    A = lib.abi
    l = lib.load_library()
    y = np.array([1.0, 2.0, 2.5, 4.1, 5.0])
    x = np.array([1.0, 2.0, 3.0, 4.0, 5.0])
    xs = (A.DataArray * 1)(A.DataArray(x.ctypes.data_as(A.c_double_p), None, 5))
    core, inf, err = A.FitResultCore(), A.FitResultInference(), A.AnofoxError()
    ok = l.anofox_ols_fit(A.DataArray(y.ctypes.data_as(A.c_double_p), None, 5), xs, 1,
                          A.OlsOptions(True, True, 0.95, 1, 0), C.byref(core), C.byref(inf), C.byref(err))
    assert ok and bool(core.coefficients) and bool(inf.p_values) and inf.len == 1
    l.anofox_free_result_core(C.byref(core))
    l.anofox_free_result_inference(C.byref(inf))
    assert not core.coefficients and not inf.std_errors and not inf.ci_upper  # lib.rs:272-311
    l.anofox_free_result_core(None)  # NULL-safe
    l.anofox_free_result_inference(None)
    # an hc_type outside NONE .. HC3 is a call-level error ("Invalid hc_type", map_options_parser.cpp)
    ok = l.anofox_ols_fit(A.DataArray(y.ctypes.data_as(A.c_double_p), None, 5), xs, 1,
                          A.OlsOptions(True, True, 0.95, 1, 7), C.byref(core), C.byref(inf), C.byref(err))
    assert not ok and err.code == A.INVALID_INPUT


# ------------------------------------------------------------------ special functions
def test_device_special_functions(lib, oracle):
    rng = np.random.default_rng(1)
    ts, dfs = [], []
    for df in [1, 2, 3, 4, 5, 7, 10, 25, 44, 98, 146, 183, 991, 5000, 100000]:
        for t in np.concatenate([rng.uniform(0, 8, 25), [1e-6, 0.3, 15.0, 45.0, 400.0]]):
            ts.append(t)
            dfs.append(df)
    got = lib.eval_specfun(0, ts, dfs)
    want = np.array([oracle.t_two_sided_p(t, d) for t, d in zip(ts, dfs)])
    rel = np.abs(got - want) / np.maximum(want, 1e-300)
    assert rel.max() < 1e-10, (rel.max(), ts[int(rel.argmax())], dfs[int(rel.argmax())])
    lv, dd = [], []
    for df in [1, 2, 3, 4, 5, 8, 13, 44, 98, 183, 991, 4096, 4097, 100000]:
        for level in [0.5, 0.8, 0.9, 0.95, 0.99, 0.999, 0.999999]:
            lv.append(level)
            dd.append(df)
    got = lib.eval_specfun(1, lv, dd)
    want = np.array([oracle.t_critical(l, d) for l, d in zip(lv, dd)])
    rel = np.abs(got - want) / want
    assert rel.max() < 1e-11, (rel.max(), lv[int(rel.argmax())], dd[int(rel.argmax())])
    fs, d1, d2 = [], [], []
    for a, b in [(1, 5), (3, 146), (5, 44), (16, 183), (8, 991), (2, 2), (64, 100000), (1, 1)]:
        for f in [1e-3, 0.5, 1.0, 3.0, 25.0, 1e3, 1e5]:
            fs.append(f)
            d1.append(a)
            d2.append(b)
    got = lib.eval_specfun(2, fs, d1, d2)
    want = np.array([oracle.f_upper_p(f, a, b) for f, a, b in zip(fs, d1, d2)])
    rel = np.abs(got - want) / np.maximum(want, 1e-300)
    assert rel.max() < 1e-10, rel.max()
    # table-driven fast paths used by the solve kernel (integer df): body sum, tail series, CF fallback
    ts, dfs = [], []
    for df in list(range(1, 41)) + [44, 45, 98, 99, 146, 183, 200, 399, 400, 401, 991, 4096, 4097, 100000]:
        for t in np.concatenate([rng.uniform(0, 9, 30), [1e-9, 1e-4, 0.05, 12.0, 30.0, 80.0, 500.0, 1e5]]):
            ts.append(t)
            dfs.append(df)
    got = lib.eval_specfun(3, ts, dfs)
    want = np.array([oracle.t_two_sided_p(t, d) for t, d in zip(ts, dfs)])
    ok = want > 1e-290
    rel = np.abs(got[ok] - want[ok]) / want[ok]
    assert rel.max() < 2e-10, (rel.max(), np.array(ts)[ok][int(rel.argmax())], np.array(dfs)[ok][int(rel.argmax())])
    assert (got[~ok] < 1e-280).all()
    fs, d1, d2 = [], [], []
    for a in [1, 2, 3, 4, 5, 7, 8, 16, 17, 32, 63, 64]:
        for b in [1, 2, 3, 5, 10, 44, 45, 146, 183, 991, 4096, 5000]:
            for f in [1e-4, 0.2, 1.0, 2.5, 9.0, 40.0, 300.0, 1e4, 1e7]:
                fs.append(f)
                d1.append(a)
                d2.append(b)
    got = lib.eval_specfun(4, fs, d1, d2)
    want = np.array([oracle.f_upper_p(f, a, b) for f, a, b in zip(fs, d1, d2)])
    ok = want > 1e-290
    rel = np.abs(got[ok] - want[ok]) / want[ok]
    i = int(rel.argmax())
    assert rel.max() < 2e-10, (rel.max(), np.array(fs)[ok][i], np.array(d1)[ok][i], np.array(d2)[ok][i])
    # drop-in scalar symbols
    assert math.isnan(lib.t_critical(0.95, 0)) and math.isnan(lib.t_critical(1.0, 3))
    assert abs(lib.t_critical(0.95, 98) - oracle.t_critical(0.95, 98)) < 1e-12
    coef, x = [2.0, float("nan"), -1.0], [1.5, 100.0, 2.0]
    for icpt in (0.5, float("nan")):
        for rse, n in [(0.7, 40), (float("nan"), 40), (0.7, 4), (0.0, 40), (0.7, 5)]:
            a = lib.predict_with_interval(coef, icpt, x, rse, n, 0.9)
            b = oracle.predict_with_interval(coef, icpt, x, rse, n, 0.9)
            assert a[0] == b[0] and np.allclose(a[1], b[1], rtol=1e-12, atol=0), (icpt, rse, n, a, b)
    assert lib.predict_with_interval(coef, 0.0, [1.0], 1.0, 10, 0.9)[0] is False  # length mismatch


def test_device_special_functions_large_df(lib):
    """df above the table range (groups of more than 4096 rows, up to the 10^9-row single group of config 4):
    Hill's normalising expansion for the t tail and the finite beta recurrence for the F tail, against
    60-digit mpmath values."""
    import mpmath as mp
    mp.mp.dps = 60

    def t_tail(t, n):
        t, n = mp.mpf(t), mp.mpf(n)
        if t * t > 3000 and n > 4000:
            return mp.mpf(0)  # < 1e-600: below the double range (and mpmath's series gives up out there)
        return mp.betainc(n / 2, mp.mpf(1) / 2, 0, n / (n + t * t), regularized=True)

    def f_tail(f, d1, d2):
        # I_x(a, b), a = d2/2, b = d1/2, from I_x(a, 1) = x^a or I_x(a, 1/2) (the t tail) by the exact recurrence
        # I_x(a, b+1) = I_x(a, b) + x^a y^b / (b B(a, b)), in 60-digit arithmetic (mpmath's own betainc does not
        # converge for a ~ 10^8); the identity itself is checked against betainc for d2 <= 10^5 below
        f, a, b = mp.mpf(f), mp.mpf(d2) / 2, mp.mpf(d1) / 2
        x, y = d2 / (d2 + d1 * f), d1 * f / (d2 + d1 * f)
        if d1 % 2 == 0:
            acc, bp = x ** a, mp.mpf(1)
        else:
            acc, bp = t_tail(mp.sqrt(d1 * f), d2), mp.mpf(1) / 2
        while bp < b:
            acc += mp.exp(a * mp.log(x) + bp * mp.log(y) - mp.log(bp) - mp.log(mp.beta(a, bp)))
            bp += 1
        return acc

    for d1, d2, f in [(3, 4097, 2.0), (64, 10**5, 1.2), (63, 20001, 30.0), (2, 4097, 0.01)]:
        direct = mp.betainc(mp.mpf(d2) / 2, mp.mpf(d1) / 2, 0, mp.mpf(d2) / (d2 + d1 * mp.mpf(f)), regularized=True)
        assert abs(f_tail(f, d1, d2) / direct - 1) < 1e-30

    dfl = [4097, 20001, 10**5, 999_999, 124_999_935, 10**9]
    ts, dfs = [], []
    for df in dfl:
        for t in [1e-5, 0.02, 0.5, 1.0, 1.96, 3.0, 6.0, 10.0, 20.0, 37.0]:
            ts.append(t)
            dfs.append(df)
    want = np.array([float(t_tail(t, d)) for t, d in zip(ts, dfs)])
    for mode in (0, 3):
        got = lib.eval_specfun(mode, ts, dfs)
        rel = np.abs(got - want) / want
        assert rel.max() < 1e-10, (mode, rel.max(), ts[int(rel.argmax())], dfs[int(rel.argmax())])
    assert (lib.eval_specfun(3, [1e3, 1e6], [10**8, 5000]) == 0.0).all()
    lv, dd = [], []
    for df in dfl:
        for level in [0.5, 0.9, 0.95, 0.99, 0.999999]:
            lv.append(level)
            dd.append(df)
    got = lib.eval_specfun(1, lv, dd)
    back = np.array([float(t_tail(t, d)) for t, d in zip(got, dd)])
    alpha = 1.0 - np.array(lv)
    assert (np.abs(back - alpha) / alpha).max() < 1e-10
    fs, d1, d2 = [], [], []
    for b in dfl:
        for a in [1, 2, 3, 5, 16, 63, 64]:
            for f in [0.001, 0.01, 0.3, 1.0, 1.5, 3.0, 10.0, 25.0, 40.0, 1e5]:
                fs.append(f)
                d1.append(a)
                d2.append(b)
    want = np.array([float(f_tail(f, a, b)) for f, a, b in zip(fs, d1, d2)])
    for mode in (2, 4):
        got = lib.eval_specfun(mode, fs, d1, d2)
        ok = want > 1e-290
        rel = np.abs(got[ok] - want[ok]) / want[ok]
        i = int(rel.argmax())
        assert rel.max() < 1e-10, (mode, rel.max(), np.array(fs)[ok][i], np.array(d1)[ok][i], np.array(d2)[ok][i])


# ------------------------------------------------------------------ batch layer vs oracle
def make_case(rng, G, n_lo, n_hi, p, dirty=True, weights=False, collinear=True):
    sizes = rng.integers(n_lo, n_hi + 1, size=G)
    if dirty and G >= 8:
        sizes[1], sizes[3], sizes[5] = 0, 1, 2  # empty / single-row / two-row groups
    offsets = np.zeros(G + 1, dtype=np.int64)
    np.cumsum(sizes, out=offsets[1:])
    N = int(offsets[-1])
    x = rng.uniform(-10, 10, size=(N, p))
    gid = np.repeat(np.arange(G), sizes)
    beta = rng.uniform(-5, 5, size=(G, p))
    y = rng.uniform(-10, 10, size=G)[gid] + np.einsum("ij,ij->i", x, beta[gid]) + rng.normal(0, 2.0, size=N)
    w = rng.uniform(0.5, 1.5, size=N) if weights else None
    if dirty:
        u = rng.uniform(size=G)
        for g in np.flatnonzero(u < 0.15):  # a constant column
            x[offsets[g]:offsets[g + 1], rng.integers(0, p)] = 3.25
        if collinear and p >= 2:
            for g in np.flatnonzero((u >= 0.15) & (u < 0.30)):  # duplicated / affine-dependent column
                a, b = rng.choice(p, size=2, replace=False)
                x[offsets[g]:offsets[g + 1], max(a, b)] = 2.0 * x[offsets[g]:offsets[g + 1], min(a, b)] + 3.0
        for g in np.flatnonzero((u >= 0.30) & (u < 0.33)):  # every column constant
            x[offsets[g]:offsets[g + 1], :] = 1.5
        rows = rng.integers(0, max(N, 1), size=max(N // 40, 1))
        x[rows, rng.integers(0, p, size=rows.size)] = np.nan  # NULL elements -> NaN -> row dropped
        rows = rng.integers(0, max(N, 1), size=max(N // 60, 1))
        y[rows] = np.where(rng.uniform(size=rows.size) < 0.5, np.nan, np.inf)
        if G >= 8 and sizes[7] > 0:
            x[offsets[7], 0] = np.nan  # the first row of a group is invalid: shift row = a later one
            y[offsets[7]:offsets[8]][:1] = np.nan
        if G >= 10:
            y[offsets[9]:offsets[10]] = np.nan  # a group with no valid row at all
        if weights:
            rows = rng.integers(0, max(N, 1), size=max(N // 30, 1))
            w[rows] = rng.choice([0.0, -1.0, np.nan, np.inf], size=rows.size)
    return dict(offsets=offsets, y=y, x=x, w=w)


CASES = [
    # method, p, intercept, inference, (n_lo, n_hi), G
    ("ols", 1, True, True, (3, 40), 300),
    ("ols", 3, True, True, (5, 120), 400),
    ("ols", 5, True, True, (30, 70), 600),
    ("ols", 5, False, True, (30, 70), 300),
    ("ols", 6, True, False, (10, 90), 300),
    ("ols", 7, True, True, (10, 90), 300),   # SIDE mode (p % 8 == 7)
    ("ols", 8, True, True, (12, 300), 300),  # SIDE mode (p % 8 == 0)
    ("ols", 9, True, True, (12, 100), 200),
    ("ols", 13, False, True, (20, 100), 200),
    ("wls", 5, True, True, (30, 70), 400),
    ("wls", 16, True, True, (40, 260), 200),
    ("wls", 8, False, False, (12, 100), 200),
    ("ridge", 5, True, True, (30, 70), 400),
    ("ridge", 16, True, True, (40, 260), 200),
    ("ridge", 16, False, False, (40, 260), 100),
    ("ols", 17, True, True, (40, 200), 100),
    ("ols", 24, True, False, (60, 200), 60),
    ("ols", 32, True, True, (80, 200), 40),
    ("ols", 40, True, False, (100, 200), 20),
    ("ols", 64, True, True, (150, 260), 12),
]


@pytest.mark.parametrize("method,p,icpt,inf,nr,G", CASES,
                         ids=[f"{c[0]}-p{c[1]}-{'i' if c[2] else 'noi'}-{'inf' if c[3] else 'core'}" for c in CASES])
def test_batch_parity_vs_oracle(lib, oracle, path, method, p, icpt, inf, nr, G):
    if path not in ("auto", "qr") and p > 8:
        pytest.skip("p > 8 always takes the record path")
    rng = np.random.default_rng(1000 + p * 7 + len(method))
    c = make_case(rng, G, nr[0], nr[1], p, dirty=True, weights=(method == "wls"))
    opt = dict(method=method, fit_intercept=icpt, compute_inference=inf, confidence_level=0.9,
               alpha=1.0 if method == "ridge" else 0.0)
    dev = lib.fit_batch_host(c["offsets"], c["y"], c["x"], c["w"], **opt)
    ora = oracle.fit_batch(c["offsets"], c["y"], c["x"], c["w"], **opt)
    assert ora["rc"] == 0
    st = ora["status"]
    # the case generator must actually exercise the NULL patterns it claims to
    assert (st == 1000).sum() >= 2 and (st == 10).sum() >= 1 and (st == 0).sum() > G // 2
    assert_fit_parity(dev, ora, rtol=1e-9, inference=inf, label=f"{method} p={p}")


def test_batch_ridge_glmnet_scaling(lib, oracle, path):
    rng = np.random.default_rng(5)
    c = make_case(rng, 100, 30, 80, 4, dirty=False)
    opt = dict(method="ridge", alpha=0.3, lambda_scaling="glmnet", compute_inference=True)
    dev = lib.fit_batch_host(c["offsets"], c["y"], c["x"], None, **opt)
    ora = oracle.fit_batch(c["offsets"], c["y"], c["x"], None, **opt)
    assert_fit_parity(dev, ora, rtol=1e-9, inference=True)
    with pytest.raises(lib.B200Error) as e:
        lib.fit_batch_host(c["offsets"], c["y"], c["x"], None, method="ridge", alpha=-0.5)
    assert e.value.code == lib.abi.INVALID_ALPHA  # ridge.rs:38-40: before anything else


def test_batch_solver_option_is_result_neutral(lib):
    """test_map_options.test: svd / qr / cholesky agree to 1e-10 on r_squared."""
    rng = np.random.default_rng(6)
    c = make_case(rng, 50, 20, 60, 3, dirty=False)
    r = [lib.fit_batch_host(c["offsets"], c["y"], c["x"], None, solver=s)["r_squared"] for s in ("svd", "qr", "cholesky")]
    assert np.abs(r[0] - r[1]).max() < 1e-10 and np.abs(r[0] - r[2]).max() < 1e-10


def test_batch_empty_and_degenerate_inputs(lib, oracle, path):
    # no groups at all
    out = lib.fit_batch_host(np.zeros(1, dtype=np.int64), np.zeros(0), np.zeros((0, 3)), None)
    assert out["status"].shape == (0,)
    # only empty groups
    out = lib.fit_batch_host(np.zeros(6, dtype=np.int64), np.zeros(0), np.zeros((0, 2)), None)
    assert (out["status"] == 1000).all() and np.isnan(out["coefficients"]).all()
    # error_handling.test:102-131: a group with n < p + 1 is NULL, its neighbour is fine
    y = np.array([1.0, 2.0, 1.0, 2.2, 2.9, 4.1, 5.2])
    x = np.array([[1.0, 2.0], [2.0, 1.0], [1.0, 5.0], [2.0, 3.0], [3.0, 4.5], [4.0, 1.0], [5.0, 2.0]])
    off = np.array([0, 2, 7], dtype=np.int64)
    dev = lib.fit_batch_host(off, y, x, None)
    ora = oracle.fit_batch(off, y, x, None)
    assert dev["status"].tolist() == ora["status"].tolist() == [6, 0]
    assert_fit_parity(dev, ora)
    # many tiny groups (more groups than a stage header holds) and offsets with repeats
    rng = np.random.default_rng(7)
    c = make_case(rng, 5000, 0, 4, 2, dirty=False)
    dev = lib.fit_batch_host(c["offsets"], c["y"], c["x"], None, compute_inference=True)
    ora = oracle.fit_batch(c["offsets"], c["y"], c["x"], None, compute_inference=True)
    assert_fit_parity(dev, ora, rtol=1e-8, inference=True, ptol=1e-6, check_ic=False)


def test_batch_large_groups_take_the_streaming_path(lib, oracle, path):
    """Groups larger than a shared-memory stage are accumulated by the streaming kernel."""
    rng = np.random.default_rng(8)
    sizes = np.array([30, 40000, 25, 0, 9000, 60, 30000, 17])
    offsets = np.zeros(len(sizes) + 1, dtype=np.int64)
    np.cumsum(sizes, out=offsets[1:])
    N, p = int(offsets[-1]), 5
    x = rng.uniform(-10, 10, size=(N, p))
    y = 3.0 + x @ rng.uniform(-5, 5, size=p) + rng.normal(0, 2.0, size=N)
    w = rng.uniform(0.5, 1.5, size=N)
    x[offsets[1], 2] = np.nan  # first row of a big group invalid
    x[offsets[4]:offsets[5], 3] = -7.0  # constant column inside a big group
    y[rng.integers(0, N, size=50)] = np.nan
    for method in ("ols", "wls"):
        opt = dict(method=method, compute_inference=True)
        dev = lib.fit_batch_host(offsets, y, x, w, **opt)
        ora = oracle.fit_batch(offsets, y, x, w, **opt)
        assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=method)
    # one very wide big group
    N, p = 6000, 64
    x = rng.normal(size=(N, p))
    y = x @ rng.uniform(-1, 1, size=p) + rng.normal(0, 0.5, size=N)
    off = np.array([0, N], dtype=np.int64)
    dev = lib.fit_batch_host(off, y, x, None, compute_inference=True)
    ora = oracle.fit_batch(off, y, x, None, compute_inference=True)
    assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label="p64")


RING_CASES = [
    # p, (n_lo, n_hi), G, method   -- group sizes chosen to exercise team sizes 1 .. 16 of the ring kernel
    (3, (3000, 5000), 12, "ols"),
    (8, (900, 1100), 40, "ols"),
    (8, (100, 140), 200, "wls"),
    (12, (500, 700), 40, "ols"),
    (16, (180, 420), 120, "wls"),
    (16, (150, 260), 150, "ridge"),
    (23, (260, 300), 60, "ols"),
    (33, (300, 500), 30, "ols"),
    (64, (300, 600), 10, "ols"),
]


@pytest.mark.parametrize("p,nr,G,method", RING_CASES, ids=[f"{c[3]}-p{c[0]}-n{c[1][0]}" for c in RING_CASES])
def test_ring_kernel_teams(lib, oracle, p, nr, G, method):
    """The streaming ring kernel (ring.cuh): clean batches (no redo), masked rows (w <= 0, training flags),
    a few non-finite values (redo list), empty groups between large ones."""
    lib.force_path("record")
    try:
        rng = np.random.default_rng(4000 + p)
        c = make_case(rng, G, nr[0], nr[1], p, dirty=False, weights=(method == "wls"))
        opt = dict(method=method, compute_inference=True, alpha=1.0 if method == "ridge" else 0.0)
        dev = lib.fit_batch_host(c["offsets"], c["y"], c["x"], c["w"], **opt)
        ora = oracle.fit_batch(c["offsets"], c["y"], c["x"], c["w"], **opt)
        assert (ora["status"] == 0).all()
        assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=f"ring clean p={p}")
        # the same result wherever the batch starts (odd row offset -> misaligned bulk copies)
        sub = slice(int(c["offsets"][1]), None)
        off2 = c["offsets"][1:] - c["offsets"][1]
        dev2 = lib.fit_batch_host(off2, c["y"][sub].copy(), c["x"][sub].copy(), None if c["w"] is None else c["w"][sub].copy(), **opt)
        assert np.array_equal(dev2["coefficients"], dev["coefficients"][1:])
        assert np.array_equal(dev2["p_values"], dev["p_values"][1:], equal_nan=True)
        # masks, non-finite values, degenerate columns, empty groups
        N = len(c["y"])
        y, x = c["y"].copy(), c["x"].copy()
        w = c["w"].copy() if c["w"] is not None else None
        offs = c["offsets"].copy()
        offs = np.concatenate([offs[:3], [offs[2]], offs[3:]])  # an empty group in the middle
        g0, g1, g2 = int(offs[0]), int(offs[1]), int(offs[2])
        x[g0, 0] = np.nan                     # first row of group 0 invalid
        y[g1:g1 + 40] = np.nan                # the first 40 rows of group 1 are held out
        x[g2:int(offs[4]), p - 1] = 2.5       # constant column
        x[int(offs[5]):int(offs[6]), 0] = x[int(offs[5]), 0] + 1e-10 * np.sign(rng.normal(size=int(offs[6] - offs[5])))  # the 1e-10 band
        if p >= 2:
            x[int(offs[6]):int(offs[7]), 1] = 3.0 * x[int(offs[6]):int(offs[7]), 0] - 1.0
        y[rng.integers(0, N, size=5)] = np.inf
        if w is not None:
            w[rng.integers(0, N, size=N // 20)] = 0.0
            w[g2:g2 + 50] = -1.0
        dev = lib.fit_batch_host(offs, y, x, w, **opt)
        ora = oracle.fit_batch(offs, y, x, w, **opt)
        assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=f"ring dirty p={p}")
        # fit-predict engine: training flags (HAST) with whole blocks of held-out rows
        train = (rng.uniform(size=N) < 0.8).astype(np.uint8)
        train[g0:g0 + 70] = 0
        train[~np.isfinite(y)] = 0
        popt = dict(method=method, confidence_level=0.95, alpha=1.0 if method == "ridge" else 0.0)
        pd, fd = lib.fit_predict_batch_host(offs, y, x, w, train, **popt)
        po, fo = oracle.fit_predict_batch(offs, y, x, w, train, **popt)
        assert_fit_parity(fd, fo, rtol=1e-9, label=f"ring predict p={p}")
        m = np.isfinite(po["yhat_upper"])
        assert np.array_equal(np.isnan(pd["yhat_upper"]), np.isnan(po["yhat_upper"]))
        assert (np.abs(pd["yhat_upper"][m] - po["yhat_upper"][m]) / np.maximum(np.abs(po["yhat"][m]), 1)).max() < 1e-9
    finally:
        lib.force_path("auto")


def test_exact_fit_groups_use_the_residual_pass(lib, oracle, path):
    """Noise-free groups: the moment form of RSS cancels completely, the queued residual pass must restore
    r_squared == 1 to the last digits and a residual standard error at rounding level (anofox_stats.test:12-15,
    ols_validation.test:121-133)."""
    rng = np.random.default_rng(12)
    c = make_case(rng, 300, 10, 60, 4, dirty=False)
    gid = np.repeat(np.arange(300), np.diff(c["offsets"]))
    beta = rng.uniform(-5, 5, size=(300, 4))
    y = 2.0 + np.einsum("ij,ij->i", c["x"], beta[gid])
    y[c["offsets"][150]:] += rng.normal(0, 1e-5, size=len(y) - c["offsets"][150])  # half: tiny but real noise
    for method, w in (("ols", None), ("wls", rng.uniform(0.5, 1.5, size=len(y)))):
        opt = dict(method=method, compute_inference=True)
        dev = lib.fit_batch_host(c["offsets"], y, c["x"], w, **opt)
        ora = oracle.fit_batch(c["offsets"], y, c["x"], w, **opt)
        assert (dev["status"] == 0).all()
        assert np.abs(dev["coefficients"] - ora["coefficients"]).max() < 1e-9
        assert (dev["r_squared"][:150] >= 1 - 1e-14).all()
        assert (dev["residual_std_error"][:150] < 1e-11).all()  # |y| ~ 100: rounding-level residuals
        # tiny-noise half: the statistic is well defined and must agree to 1e-9 although RSS/TSS ~ 1e-14
        e = max_rel_err(dev["residual_std_error"][150:], ora["residual_std_error"][150:])
        assert e < 1e-6, e  # limited by the coefficients' own 1e-16 * cond rounding, not by cancellation
        assert np.abs(dev["r_squared"][150:] - ora["r_squared"][150:]).max() < 1e-15


def test_fit_predict_parity(lib, oracle, path):
    """Engine of ols/ridge/wls_fit_predict_by: every row predicted, held-out rows (y NULL) not trained on."""
    rng = np.random.default_rng(9)
    # p > 8 always takes the ring kernel: default training flags from y alone + the checked path for NaN features
    for method, p in [("ols", 3), ("ols", 8), ("ridge", 5), ("wls", 4), ("ols", 16), ("ridge", 12), ("wls", 9)]:
        c = make_case(rng, 120, 8, 90 if p <= 8 else 300, p, dirty=True, weights=(method == "wls"), collinear=True)
        N = len(c["y"])
        hold = rng.uniform(size=N) < 0.2
        c["y"][hold] = np.nan
        opt = dict(method=method, confidence_level=0.95, alpha=1.0 if method == "ridge" else 0.0)
        pd, fd = lib.fit_predict_batch_host(c["offsets"], c["y"], c["x"], c["w"], None, **opt)
        po, fo = oracle.fit_predict_batch(c["offsets"], c["y"], c["x"], c["w"], None, **opt)
        assert np.array_equal(pd["is_training"], po["is_training"])
        assert_fit_parity(fd, fo, rtol=1e-9, label=f"predict/{method}")
        for k in ("yhat", "yhat_lower", "yhat_upper"):
            assert np.array_equal(np.isnan(pd[k]), np.isnan(po[k])), k  # NULL triples at the same rows
            m = np.isfinite(po[k])
            scale = np.maximum(np.abs(po["yhat"][m]), 1.0)
            assert (np.abs(pd[k][m] - po[k][m]) / scale).max() < 1e-9, k
        ok = np.isfinite(po["yhat"])
        assert (pd["yhat_lower"][ok] <= pd["yhat"][ok]).all() and (pd["yhat"][ok] <= pd["yhat_upper"][ok]).all()
        # explicit training flags (split column semantics): only flagged rows train
        train = (rng.uniform(size=N) < 0.7).astype(np.uint8)
        train[~np.isfinite(c["y"])] = 0
        pd, fd = lib.fit_predict_batch_host(c["offsets"], c["y"], c["x"], c["w"], train, **opt)
        po, fo = oracle.fit_predict_batch(c["offsets"], c["y"], c["x"], c["w"], train, **opt)
        assert np.array_equal(pd["is_training"], train)
        assert_fit_parity(fd, fo, rtol=1e-9, label=f"predict-flags/{method}")
        m = np.isfinite(po["yhat_upper"])
        assert np.array_equal(np.isnan(pd["yhat_upper"]), np.isnan(po["yhat_upper"]))
        assert (np.abs(pd["yhat_upper"][m] - po["yhat_upper"][m]) / np.maximum(np.abs(po["yhat"][m]), 1)).max() < 1e-9


def test_row_sharded_single_group(lib, oracle):
    """Config-4 path on one GPU: two row shards -> moment blocks -> SUM / MAX combine -> solve."""
    import ctypes as C
    import torch
    A = lib.abi
    l = lib.load_library()
    rng = np.random.default_rng(10)
    for p, N in [(5, 20000), (64, 30000), (16, 9001)]:
        x = rng.normal(size=(N, p))
        for j in range(1, p):
            x[:, j] = 0.3 * x[:, j - 1] + math.sqrt(1 - 0.09) * x[:, j]  # AR(1) correlation across features
        y = 1.5 + x @ rng.uniform(-1, 1, size=p) + rng.normal(0, 0.7, size=N)
        y[5] = np.nan
        x[17, p - 1] = np.inf
        off = np.array([0, N], dtype=np.int64)
        ora = oracle.fit_batch(off, y, x, None, compute_inference=True)
        xd, yd = torch.tensor(x, device="cuda"), torch.tensor(y, device="cuda")
        shift = torch.cat([xd[0], yd[:1]]).contiguous()  # rank 0's first valid row, "broadcast"
        L = lib.moment_block_len(p)
        cut = N // 3
        blocks, cdevs = [], []
        wb = l.anofox_b200_workspace_bytes(1, N, p)
        ws = torch.empty(wb + 256, dtype=torch.uint8, device="cuda")
        wp = (ws.data_ptr() + 255) & ~255
        err = A.AnofoxError()
        for a, b in [(0, cut), (cut, N)]:
            m = torch.zeros(L, dtype=torch.float64, device="cuda")
            cd = torch.zeros(p, dtype=torch.float64, device="cuda")
            xs, ys = xd[a:b].contiguous(), yd[a:b].contiguous()
            rc = l.anofox_b200_moments_device(C.c_void_p(ys.data_ptr()), C.c_void_p(xs.data_ptr()), None, b - a, p,
                                              C.c_void_p(shift.data_ptr()), C.c_void_p(m.data_ptr()),
                                              C.c_void_p(cd.data_ptr()), C.c_void_p(wp), wb, None, C.byref(err))
            assert rc == 0, err.message
            torch.cuda.synchronize()
            blocks.append(m)
            cdevs.append(cd)
        msum = blocks[0] + blocks[1]            # what ncclAllReduce(SUM) produces
        cmax = torch.maximum(cdevs[0], cdevs[1])  # ncclAllReduce(MAX)
        fit = lib.DeviceFit(1, N, p, inference=True)
        o = A.make_fit_options(compute_inference=True)
        rc = l.anofox_b200_solve_moments_device(C.c_void_p(msum.data_ptr()), C.c_void_p(cmax.data_ptr()),
                                                C.c_void_p(shift.data_ptr()), p, C.byref(o), C.byref(fit.out_struct),
                                                None, C.byref(err))
        assert rc == 0, err.message
        torch.cuda.synchronize()
        dev = {k: v.cpu().numpy() for k, v in fit.out.items()}
        assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=f"sharded p={p}")


def _fit_from_row_shards(lib, x, y, w, cuts, **opt):
    """Row shards -> moment blocks (anofox_b200_moments_device) -> SUM / MAX -> anofox_b200_solve_moments_device."""
    import ctypes as C
    import torch
    A = lib.abi
    l = lib.load_library()
    N, p = x.shape
    xd, yd = torch.tensor(x, device="cuda"), torch.tensor(y, device="cuda")
    wd = None if w is None else torch.tensor(w, device="cuda")
    shift = torch.cat([xd[0], yd[:1]]).contiguous()
    L = lib.moment_block_len(p)
    wb = l.anofox_b200_workspace_bytes(1, N, p)
    ws = torch.empty(wb + 256, dtype=torch.uint8, device="cuda")
    wp = (ws.data_ptr() + 255) & ~255
    err = A.AnofoxError()
    msum = torch.zeros(L, dtype=torch.float64, device="cuda")
    cmax = torch.zeros(p, dtype=torch.float64, device="cuda")
    for a, b in zip(cuts[:-1], cuts[1:]):
        m = torch.zeros(L, dtype=torch.float64, device="cuda")
        cd = torch.zeros(p, dtype=torch.float64, device="cuda")
        # views into the big arrays: shard starts are not 16-byte aligned when `a` is odd
        xs, ys = xd[a:b], yd[a:b]
        wsd = None if wd is None else C.c_void_p(wd[a:b].data_ptr())
        rc = l.anofox_b200_moments_device(C.c_void_p(ys.data_ptr()), C.c_void_p(xs.data_ptr()), wsd, b - a, p,
                                          C.c_void_p(shift.data_ptr()), C.c_void_p(m.data_ptr()),
                                          C.c_void_p(cd.data_ptr()), C.c_void_p(wp), wb, None, C.byref(err))
        assert rc == 0, err.message
        torch.cuda.synchronize()
        msum += m
        cmax = torch.maximum(cmax, cd)
    fit = lib.DeviceFit(1, N, p, inference=True)
    o = A.make_fit_options(compute_inference=True, **opt)
    rc = l.anofox_b200_solve_moments_device(C.c_void_p(msum.data_ptr()), C.c_void_p(cmax.data_ptr()),
                                            C.c_void_p(shift.data_ptr()), p, C.byref(o), C.byref(fit.out_struct),
                                            None, C.byref(err))
    assert rc == 0, err.message
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in fit.out.items()}


SHARD_CASES = [
    # p, N, method, constant column: clean shards stay on the ring kernel's slice mode (no checked pass)
    (64, 300_000, "ols", False),
    (5, 1_000_003, "ols", False),
    (16, 250_001, "wls", False),
    (33, 120_000, "ols", True),
    (8, 77, "ols", False),
    (48, 150_000, "ridge", False),
    # more slice-kernel shapes: non-SIDE (y and the 1-column ride in the last tile), p % 16 == 8, ragged tails
    (34, 90_003, "ols", True),
    (40, 100_001, "ols", False),
    (64, 70_007, "ols", True),
]


@pytest.mark.parametrize("p,N,method,const", SHARD_CASES, ids=[f"{c[2]}-p{c[0]}-n{c[1]}" for c in SHARD_CASES])
def test_row_sharded_clean_shards(lib, oracle, p, N, method, const):
    rng = np.random.default_rng(77 + p)
    x = rng.normal(size=(N, p)) + 3.0
    for j in range(1, p):
        x[:, j] = 0.3 * x[:, j - 1] + math.sqrt(1 - 0.09) * x[:, j]
    if const:
        x[:, 2] = -4.25
        x[:, p - 1] = x[:, 0]  # and an exactly duplicated one: the later column is aliased
    y = 1.5 + x @ rng.uniform(-1, 1, size=p) + rng.normal(0, 0.7, size=N)
    w = rng.uniform(0.5, 1.5, size=N) if method == "wls" else None
    opt = dict(method=method, alpha=2.0 if method == "ridge" else 0.0)
    cuts = [0, N // 3 + 1, N]
    dev = _fit_from_row_shards(lib, x, y, w, cuts, **opt)
    ora = oracle.fit_batch(np.array([0, N], dtype=np.int64), y, x, w, compute_inference=True, **opt)
    assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=f"clean shards p={p} {method}")
    if const:
        assert np.isnan(dev["coefficients"][0, 2]) and np.isnan(dev["coefficients"][0, p - 1])
    # the same rows in different shard cuts: the sums differ in the last bits only
    dev2 = _fit_from_row_shards(lib, x, y, w, [0, N // 2, N], **opt)
    m = np.isfinite(dev["coefficients"])
    assert np.allclose(dev["coefficients"][m], dev2["coefficients"][m], rtol=1e-10, atol=0)


def test_row_sharded_exact_fits_are_refined(lib, oracle):
    """Row-sharded group whose fit is exact or nearly exact (R^2 = 1, 1 - 1e-9): RSS = TSS - ESS of the reduced
    moment block cancels; the refine protocol (flag -> per-shard residual sums -> SUM -> second solve) must restore
    r_squared, the residual standard error and every statistic derived from them."""
    import torch
    from importlib import import_module
    D = import_module(lib.__name__ + ".distributed")
    rng = np.random.default_rng(31)
    for p, N, noise, method in [(6, 50_000, 0.0, "ols"), (64, 30_000, 1e-6, "ols"), (12, 40_000, 3e-5, "wls"), (5, 9_000, 0.0, "ridge")]:
        x = rng.normal(size=(N, p)) * 3.0 + 1.0
        beta = rng.uniform(-2, 2, size=p)
        y = 0.75 + x @ beta + noise * rng.normal(size=N)
        w = rng.uniform(0.5, 1.5, size=N) if method == "wls" else None
        f = D.ShardedSingleGroupFit(N, p, inference=True)
        t = lambda v: None if v is None else torch.tensor(v, device="cuda")  # noqa: E731
        opt = dict(method=method, compute_inference=True, alpha=1e-9 if method == "ridge" else 0.0)
        out = f.fit(t(y), t(x), t(w), **opt)
        torch.cuda.synchronize()
        assert int(f.flag[0]) == 1, (p, noise)
        dev = {k: v.cpu().numpy() for k, v in out.items()}
        ora = oracle.fit_batch(np.array([0, N], dtype=np.int64), y, x, w, **opt)
        assert abs(dev["r_squared"][0] - ora["r_squared"][0]) < 1e-13
        if noise > 0:
            assert_fit_parity(dev, ora, rtol=1e-7, inference=True, check_ic=False, label=f"sharded exact p={p} noise={noise}")
        else:
            # pure rounding-level residuals: only their size is determined
            assert dev["residual_std_error"][0] < 1e-10 * np.abs(y).max()
            assert np.allclose(dev["coefficients"][0], ora["coefficients"][0], rtol=1e-9, atol=1e-12)


def test_sharded_single_group_fit_class(lib, oracle):
    """distributed.ShardedSingleGroupFit on one rank (no process group): first-valid-row search on the device,
    moment block, solve. The first rows of the shard are invalid, so the shift is a later row."""
    import torch
    from importlib import import_module
    D = import_module(lib.__name__ + ".distributed")
    rng = np.random.default_rng(21)
    for p, N, method in [(3, 70001, "ols"), (64, 20000, "ols"), (9, 5000, "wls"), (4, 300, "ridge")]:
        x = rng.normal(size=(N, p))
        y = 0.5 + x @ rng.uniform(-1, 1, size=p) + rng.normal(0, 0.4, size=N)
        w = rng.uniform(0.5, 1.5, size=N) if method == "wls" else None
        y[:3] = np.nan
        x[3, 0] = np.inf
        if w is not None:
            w[4] = 0.0
        x[:, p - 1] = 7.0 if p > 3 else x[:, p - 1]  # a constant column for the wider cases
        f = D.ShardedSingleGroupFit(N, p, inference=True)
        t = lambda v: None if v is None else torch.tensor(v, device="cuda")  # noqa: E731
        opt = dict(method=method, compute_inference=True, alpha=0.7 if method == "ridge" else 0.0)
        out = f.fit(t(y), t(x), t(w), **opt)
        torch.cuda.synchronize()
        dev = {k: v.cpu().numpy() for k, v in out.items()}
        ora = oracle.fit_batch(np.array([0, N], dtype=np.int64), y, x, w, **opt)
        assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=f"sharded-class p={p} {method}")
    # a shard without any valid row reports found = 0 and a zero shift
    yb = torch.full((100,), float("nan"), device="cuda", dtype=torch.float64)
    xb = torch.zeros((100, 2), device="cuda", dtype=torch.float64)
    be = D.CudaBackend(100, 2, torch.device("cuda"))
    cand = be.first_valid(yb, xb, None).cpu().numpy()
    assert cand.tolist() == [0.0, 0.0, 0.0, 0.0]
    f = D.ShardedSingleGroupFit(100, 2, inference=False)
    out = f.fit(yb, xb, None)
    torch.cuda.synchronize()
    assert int(out["status"][0]) == lib.abi.NO_VALID_DATA


def test_device_pointer_api_matches_host_api(lib):
    import torch
    rng = np.random.default_rng(11)
    c = make_case(rng, 700, 20, 80, 5, dirty=True, weights=True)
    host = lib.fit_batch_host(c["offsets"], c["y"], c["x"], c["w"], method="wls", compute_inference=True)
    G, N = len(c["offsets"]) - 1, len(c["y"])
    f = lib.DeviceFit(G, N, 5, inference=True)
    t = {k: torch.tensor(c[k], device="cuda") for k in ("offsets", "y", "x", "w")}
    before = lib.launch_count()
    out = f.fit(t["offsets"], t["y"], t["x"], t["w"], method="wls", compute_inference=True)
    torch.cuda.synchronize()
    assert lib.launch_count() - before >= 2  # window bounds + the fused kernel (record path: + solve, refine)
    for k, v in out.items():
        assert np.array_equal(v.cpu().numpy(), host[k], equal_nan=True), k  # same kernels, same bits


def test_host_api_chunked_pipeline(lib, oracle):
    """Host-pointer calls above 64 MB run as chunks of whole groups on three streams (copy in / kernels / copy
    out). The results must not depend on the chunking: same bits as the one-shot device-pointer call, and parity
    with the oracle on a sample of the groups. Covers fit and fit_predict."""
    import torch
    rng = np.random.default_rng(12)
    c = make_case(rng, 45_000, 30, 90, 5, dirty=True, weights=True)
    G, N = len(c["offsets"]) - 1, len(c["y"])
    assert N * 7 * 8 > 64 << 20
    opt = dict(method="wls", compute_inference=True)
    host = lib.fit_batch_host(c["offsets"], c["y"], c["x"], c["w"], **opt)
    f = lib.DeviceFit(G, N, 5, inference=True)
    t = {k: torch.tensor(c[k], device="cuda") for k in ("offsets", "y", "x", "w")}
    out = f.fit(t["offsets"], t["y"], t["x"], t["w"], **opt)
    torch.cuda.synchronize()
    for k, v in out.items():
        assert np.array_equal(v.cpu().numpy(), host[k], equal_nan=True), k
    # oracle parity on the groups around the chunk boundaries and at both ends
    sel = np.unique(np.concatenate([np.arange(0, 50), np.arange(G // 3 - 200, G // 3 + 200),
                                    np.arange(G // 2 - 200, G // 2 + 200), np.arange(2 * G // 3 - 200, 2 * G // 3 + 200),
                                    np.arange(G - 50, G)]))
    for a, b in [(sel[0], sel[49] + 1), (G // 3 - 200, G // 3 + 200), (G - 50, G)]:
        off = c["offsets"][a:b + 1]
        r0, r1 = off[0], off[-1]
        ora = oracle.fit_batch(off - r0, c["y"][r0:r1], c["x"][r0:r1], c["w"][r0:r1], **opt)
        dev = {k: v[a:b] for k, v in host.items()}
        assert_fit_parity(dev, ora, rtol=1e-9, inference=True, label=f"chunked host groups {a}:{b}")
    # fit_predict through the same pipeline
    y2 = c["y"].copy()
    y2[rng.uniform(size=N) < 0.2] = np.nan
    pd, fd = lib.fit_predict_batch_host(c["offsets"], y2, c["x"], None, None)
    a, b = G // 2 - 100, G // 2 + 100
    off = c["offsets"][a:b + 1]
    r0, r1 = off[0], off[-1]
    po, fo = oracle.fit_predict_batch(off - r0, y2[r0:r1], c["x"][r0:r1], None, None)
    assert_fit_parity({k: v[a:b] for k, v in fd.items()}, fo, rtol=1e-9, label="chunked predict fit")
    m = np.isfinite(po["yhat_upper"])
    assert np.array_equal(np.isnan(pd["yhat_upper"][r0:r1]), ~m)
    assert np.allclose(pd["yhat_upper"][r0:r1][m], po["yhat_upper"][m], rtol=1e-9, atol=0)
    assert np.array_equal(pd["is_training"][r0:r1], po["is_training"])


def test_full_size_properties_config2(lib):
    """BASELINE config 2 at full size (1 M groups x 50 rows x 5 features, inference on): size-independent
    properties instead of an oracle run -- scale equivariance, group-permutation invariance of the results,
    linearity in y and a checksum of checksums against sub-batch runs."""
    import torch
    synth = __import__("anofox_statistics_b200.synth", fromlist=["synth"])
    G, n, p = 1_000_000, 50, 5
    d = synth.make_torch(G, n, p, seed=2)
    f = lib.DeviceFit(G, G * n, p, inference=True)
    opt = dict(method="ols", compute_inference=True)
    base = {k: v.clone() for k, v in f.fit(d["offsets"], d["y"], d["x"], **opt).items()}
    torch.cuda.synchronize()
    assert int((base["status"] != 0).sum()) == 0 and int((base["n_observations"] != n).sum()) == 0
    assert bool(torch.isfinite(base["coefficients"]).all())
    assert bool(((base["r_squared"] > 0.0) & (base["r_squared"] <= 1.0)).all())
    assert bool(((base["p_values"] >= 0) & (base["p_values"] <= 1)).all())
    # (1) y -> 3 y + 2: slopes x3, intercept 3a+2, r2 / t / p unchanged
    y2 = d["y"] * 3.0 + 2.0
    o2 = f.fit(d["offsets"], y2, d["x"], **opt)
    torch.cuda.synchronize()
    assert float((o2["coefficients"] - 3 * base["coefficients"]).abs().max()) < 1e-9
    assert float((o2["intercept"] - (3 * base["intercept"] + 2)).abs().max()) < 1e-8
    assert float((o2["r_squared"] - base["r_squared"]).abs().max()) < 1e-11
    assert float(((o2["t_values"] - base["t_values"]).abs() / base["t_values"].abs().clamp_min(1e-3)).max()) < 1e-8
    # (2) the first 4096 groups fitted alone give bit-identical results (no cross-group leakage)
    Gs = 4096
    fs = lib.DeviceFit(Gs, Gs * n, p, inference=True)
    os_ = fs.fit(d["offsets"][:Gs + 1].contiguous(), d["y"][:Gs * n].contiguous(), d["x"][:Gs * n].contiguous(), **opt)
    torch.cuda.synchronize()
    for k in ("coefficients", "intercept", "r_squared", "p_values", "f_pvalue", "residual_std_error"):
        assert torch.equal(os_[k], base[k][:Gs]), k
    # (3) checksum of checksums: sum of per-group RSS equals the RSS of the stacked residuals
    xb = d["x"].view(G, n, p)
    fitted = (xb * base["coefficients"][:, None, :]).sum(-1) + base["intercept"][:, None]
    rss = ((d["y"].view(G, n) - fitted) ** 2).sum(-1)
    assert float(((rss - base["rss"]).abs() / rss).max()) < 1e-9
