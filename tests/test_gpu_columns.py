"""Batched scalar array fits -- {ols,ridge,wls}_fit(y LIST, x LIST(LIST)) for all SQL rows of a chunk in one call
(src/table_functions/ols_fit.cpp:120-230) -- against the oracle's single-fit symbol, which
is what the reference calls once per row."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("method,p", [("ols", 3), ("wls", 5), ("ridge", 12), ("ols", 40)])
def test_fit_columns_batch_matches_per_row_single_fits(b200, oracle, method, p):
    b200.device_check()
    rng = np.random.default_rng(50 + p)
    F = 37
    sizes = rng.integers(1, 3 * p + 30, size=F)  # includes fits with a single row and n < p (error rows)
    off = np.zeros(F + 1, dtype=np.int64)
    np.cumsum(sizes, out=off[1:])
    N = int(off[-1])
    ys, blocks, ws, cols = [], [], [], []
    for i in range(F):
        n = int(sizes[i])
        X = rng.normal(size=(n, p)) * 3 + 1
        if i % 5 == 0 and p > 2:
            X[:, 1] = 2.0
        y = 0.5 + X @ rng.normal(size=p) + rng.normal(size=n)
        if n > 4:
            y[1] = np.nan  # NULL inside the list
        ys.append(y)
        blocks.append(X.T.reshape(-1))  # p columns of n values, back to back
        cols.append([X[:, j].copy() for j in range(p)])
        ws.append(rng.uniform(0.5, 2.0, size=n))
    y_all, x_cm, w_all = np.concatenate(ys), np.concatenate(blocks), np.concatenate(ws)
    opt = dict(method=method, compute_inference=True, alpha=0.3 if method == "ridge" else 0.0)
    fit = b200.DeviceFit(F, N, p, inference=True)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    out = fit.fit_columns(dev(off), dev(y_all), dev(x_cm), dev(w_all) if method == "wls" else None, **opt)
    torch.cuda.synchronize()
    out = {k: v.cpu().numpy() for k, v in out.items()}
    for i in range(F):
        sopt = {k: v for k, v in opt.items() if k != "method"}
        r = oracle.single_fit(method, ys[i], cols[i], ws[i] if method == "wls" else None, **sopt)
        if not r.ok:
            assert out["status"][i] == r.error_code, (i, out["status"][i], r.error_code)
            continue
        assert out["status"][i] == 0, (i, sizes[i], out["status"][i])
        c = np.asarray(r.coefficients)
        assert np.array_equal(np.isnan(out["coefficients"][i]), np.isnan(c)), i
        m = ~np.isnan(c)
        scale = max(np.abs(c[m]).max() if m.any() else 1.0, abs(r.intercept))
        if m.any():
            assert np.abs(out["coefficients"][i][m] - c[m]).max() <= 1e-9 * scale, i
        assert abs(out["intercept"][i] - r.intercept) <= 1e-9 * scale
        assert out["n_observations"][i] == r.n_observations
        if r.inference_len:
            se = np.asarray(r.std_errors)
            mm = ~np.isnan(se)
            assert np.allclose(out["std_errors"][i][mm], se[mm], rtol=1e-8), i
            assert np.allclose(out["p_values"][i][mm], np.asarray(r.p_values)[mm], rtol=1e-7, atol=1e-300), i
