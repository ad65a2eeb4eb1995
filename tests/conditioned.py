"""Workloads with CONTROLLED conditioning (shared by the CPU oracle tests and the GPU parity tests).

The regime the well-conditioned generators (x ~ U(-10,10)) never reach: large column offsets, polynomial
features, highly correlated pairs, and a column that is collinear with earlier ones to within a chosen relative
distance r -- including values on either side of R's aliasing rule r < 1e-7 (LINPACK dqrdc2 as used by lm.fit,
the convention the reference's golden vectors pin: test/data/ols_tests/expected/perfect_collinearity.json).
"""
import numpy as np

from parity import conditioned_tolerance, ls_condition

KINDS = {
    # kind: levels
    "offset": [3, 5, 6],               # x_j = 10^level + U(0,1) (j+1)
    "poly": [4, 6],                    # x_j = t^(j+1), t ~ U(1,2), level = number of powers
    "rho": [4, 6, 8],                  # x_1 = rho x_0 + sqrt(1-rho^2) z, rho = 1 - 10^-level
    "nearcol": [1e-5, 1e-6, 4e-7, 1.5e-7, 0.7e-7, 2.5e-8, 1e-9],  # relative distance r of the last column from span(1, x_0, x_1)
}


def make_conditioned(rng, kind, level, G, n, p, noise=2.0, weights=False):
    """G groups of n rows (ragged: n .. 1.5 n). Returns the CSR dict."""
    sizes = rng.integers(n, n + n // 2 + 1, size=G)
    offsets = np.zeros(G + 1, dtype=np.int64)
    np.cumsum(sizes, out=offsets[1:])
    N = int(offsets[-1])
    x = rng.uniform(-10, 10, size=(N, p))
    gid = np.repeat(np.arange(G), sizes)
    if kind == "offset":
        x = 10.0 ** level + rng.uniform(0, 1, size=(N, p)) * (np.arange(p) + 1.0)
    elif kind == "poly":
        t = rng.uniform(1, 2, size=N)
        q = min(p, int(level))
        for j in range(q):
            x[:, j] = t ** (j + 1)
    elif kind == "rho":
        rho = 1.0 - 10.0 ** (-level)
        x[:, 1] = rho * x[:, 0] + np.sqrt(1 - rho * rho) * rng.uniform(-10, 10, size=N)
    elif kind == "nearcol":
        r = float(level)
        for g in range(G):
            a, b = offsets[g], offsets[g + 1]
            base = x[a:b, 0] + x[a:b, 1]
            # unit vector orthogonal to the intercept and to every earlier column: the remaining norm of the last
            # column after dqrdc2 has passed the others is exactly r * |base|
            q_, _ = np.linalg.qr(np.column_stack([np.ones(b - a), x[a:b, :p - 1]]))
            z = rng.normal(size=b - a)
            z -= q_ @ (q_.T @ z)
            z /= np.linalg.norm(z)
            x[a:b, p - 1] = base + r * np.linalg.norm(base) * z
    else:
        raise ValueError(kind)
    beta = rng.uniform(-5, 5, size=(G, p))
    y = rng.uniform(-10, 10, size=G)[gid] + np.einsum("ij,ij->i", x, beta[gid]) + rng.normal(0, noise, size=N)
    w = rng.uniform(0.5, 1.5, size=N) if weights else None
    return dict(offsets=offsets, y=y, x=x, w=w, kind=kind, level=level)


def group_tolerances(c, ora, fit_intercept=True, floor=1e-9):
    """Per-group tolerances from the data and the ORACLE's aliasing pattern -> dict for assert_fit_parity:
    rtol       max(floor, COND_C eps kappa_LS)                                      (parity.conditioned_tolerance)
    coef_scale largest |slope| of the group: slope errors are NOT measured against a large intercept
    icpt_scale max(|b0|, largest |slope|, |ybar| + sum |xbar_j beta_j|): the intercept is the difference
               ybar - sum xbar_j beta_j, and no f64 implementation resolves it below eps times those terms
    rse_rtol   16 eps M / rse, M = largest per-row sum of |terms| of the (shifted) fitted value: the rounding floor
               of the residuals themselves, which dominates for near-exact fits
    """
    from parity import EPS
    G = len(c["offsets"]) - 1
    tol = np.full(G, floor)
    kap = np.ones(G)
    cs, isc, rr = np.ones(G), np.ones(G), np.zeros(G)
    for g in range(G):
        if ora["status"][g] != 0:
            continue
        a, b = c["offsets"][g], c["offsets"][g + 1]
        beta = ora["coefficients"][g]
        live = np.isfinite(beta)
        x, y = c["x"][a:b], c["y"][a:b]
        w = None if c["w"] is None else c["w"][a:b]
        bl = np.where(live, beta, 0.0)
        cs[g] = max(np.abs(bl).max(), 1e-300)
        ww = np.ones(len(y)) if w is None else w
        xbar, ybar = (ww[:, None] * x).sum(0) / ww.sum(), (ww * y).sum() / ww.sum()
        b0 = ora["intercept"][g]
        isc[g] = max(abs(b0) if np.isfinite(b0) else 0.0, cs[g], abs(ybar) + np.abs(xbar * bl).sum())
        xs, ys = (x - x[0], y - y[0]) if fit_intercept else (x, y)
        M = (np.abs(ys) + np.abs(xs * bl).sum(1)).max()
        rse = ora["residual_std_error"][g]
        rr[g] = 16 * EPS * M / rse if np.isfinite(rse) and rse > 0 else 0.0
        if not live.any():
            continue
        k, tn = ls_condition(x, y, w, live, fit_intercept, ora["r_squared"][g])
        kap[g] = k
        tol[g] = conditioned_tolerance(k, tn, floor)
    return dict(rtol=tol, coef_scale=cs, icpt_scale=isc, rse_rtol=rr), kap
