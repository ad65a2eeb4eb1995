"""Comparison helpers shared by the GPU parity tests and __graft_entry__.smoke()."""
import numpy as np

CORE = ["coefficients", "intercept", "r_squared", "adj_r_squared", "residual_std_error"]
INF = ["std_errors", "t_values", "ci_lower", "ci_upper", "f_statistic"]


def nan_pattern_equal(a, b):
    return np.array_equal(np.isnan(a), np.isnan(b))


def max_rel_err(a, b, floor=1e-12):
    """max |a-b| / max(|b|, floor * scale) over entries finite in both."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    m = np.isfinite(a) & np.isfinite(b)
    if not m.any():
        return 0.0
    scale = max(np.abs(b[m]).max(), 1e-300)
    return float((np.abs(a[m] - b[m]) / np.maximum(np.abs(b[m]), floor * scale + 1e-300)).max())


EPS = 2.220446049250313e-16
# Tolerance outside the well-conditioned regime (north star: "1e-9 for well-conditioned f64, with the tolerance
# scaled by condition number otherwise"): max(1e-9, COND_C * eps * kappa_LS), kappa_LS = kappa + kappa^2 tan(theta)
# the condition number of the least-squares PROBLEM (Wedin / Bjorck): kappa = 2-norm condition number of the
# centred (when an intercept is fitted), column-equilibrated design restricted to the non-aliased columns,
# tan(theta) = |residual| / |fitted part| = sqrt((1 - R^2) / R^2). A backward-stable QR in f64 attains a small
# multiple of eps * kappa_LS; the device's QR fallback is held to COND_C = 32 of it.
COND_C = 32.0


def ls_condition(x, y, w, live, fit_intercept, r2):
    """(kappa, tan theta) of one group's least-squares problem. x [n, p] valid rows only, live = bool mask of the
    columns that carry a coefficient, r2 = the oracle's R^2."""
    xs = np.asarray(x, dtype=np.float64)[:, live]
    if xs.shape[1] == 0 or xs.shape[0] == 0:
        return 1.0, 0.0
    ww = np.ones(xs.shape[0]) if w is None else np.asarray(w, dtype=np.float64)
    if fit_intercept:
        xs = xs - (ww[:, None] * xs).sum(0) / ww.sum()
    xs = xs * np.sqrt(ww)[:, None]
    nrm = np.sqrt((xs * xs).sum(0))
    xs = xs / np.where(nrm > 0, nrm, 1.0)
    sv = np.linalg.svd(xs, compute_uv=False)
    kappa = float(sv[0] / max(sv[-1], 1e-300))
    r2 = min(max(float(r2), 1e-300), 1.0)
    return kappa, float(np.sqrt(max(1.0 - r2, 0.0) / r2))


def conditioned_tolerance(kappa, tan_theta, floor=1e-9):
    return max(floor, COND_C * EPS * (kappa + kappa * kappa * tan_theta))


def assert_fit_parity(dev, ora, rtol=1e-9, inference=False, ptol=1e-8, check_ic=True, label="", coef_scale=None,
                      icpt_scale=None, rse_rtol=None):
    """Device result dict vs oracle result dict (same SoA layout).

    Bit-exact: status (NULL-struct pattern), n_observations, NaN positions of every array.
    Relative tolerance rtol (north star: 1e-9 for well-conditioned f64) on values; p-values relative ptol.
    rtol may be an array with one tolerance per group (conditioned suites: conditioned_tolerance above).
    coef_scale / icpt_scale [G]: the magnitudes the slope and intercept errors are taken relative to (default: the
    group's largest |coefficient| incl. the intercept); rse_rtol [G]: relative tolerance of the residual standard
    error and of what is proportional to it (SE, t, F) where the rounding floor of the residuals themselves --
    eps * |terms of the fitted value| / |residual| -- exceeds rtol (near-exact fits).
    """
    G = len(ora["status"])
    rt = np.broadcast_to(np.asarray(rtol, dtype=np.float64), (G,))
    rt2 = rt[:, None]
    scalar_tol = np.ndim(rtol) == 0
    assert np.array_equal(dev["status"], ora["status"]), (label, "status", np.flatnonzero(dev["status"] != ora["status"])[:10])
    bad = np.flatnonzero(dev["n_observations"] != ora["n_observations"])
    assert bad.size == 0, (label, "n_observations", bad[:10], dev["n_observations"][bad[:10]], ora["n_observations"][bad[:10]])
    for k in CORE:
        assert nan_pattern_equal(dev[k], ora[k]), (label, "NaN pattern", k, np.argwhere(np.isnan(dev[k]) != np.isnan(ora[k]))[:5])

    def worst(err, tol, mask):
        """largest err / tol over the masked entries (0 when none)"""
        if not mask.any():
            return 0.0, -1
        q = np.where(mask, err / tol, 0.0)
        i = int(np.argmax(q))
        return float(q.reshape(-1)[i]), i

    # coefficients are compared against the per-group coefficient scale (a slope of 1e-3 next to one of 5 is
    # only determined to 1e-9 * 5 by either implementation)
    c_dev, c_ora = dev["coefficients"], ora["coefficients"]
    scale = np.nanmax(np.abs(np.concatenate([c_ora, ora["intercept"][:, None]], axis=1)), axis=1, initial=0.0)
    scale = np.where(scale > 0, scale, 1.0)[:, None]
    cs = scale if coef_scale is None else np.asarray(coef_scale, dtype=np.float64)[:, None]
    isc = scale[:, 0] if icpt_scale is None else np.asarray(icpt_scale, dtype=np.float64)
    rr = rt if rse_rtol is None else np.maximum(rt, np.asarray(rse_rtol, dtype=np.float64))
    rr2 = rr[:, None]
    with np.errstate(invalid="ignore"):
        q, i = worst(np.abs(c_dev - c_ora) / cs, rt2, np.isfinite(c_ora))
        assert q <= 1.0, (label, "coefficients", q, "x tol at", i)
        q, i = worst(np.abs(dev["intercept"] - ora["intercept"]) / isc, rt, np.isfinite(ora["intercept"]))
        assert q <= 1.0, (label, "intercept", q, "x tol at group", i)
        for k in ["r_squared", "adj_r_squared"]:
            q, i = worst(np.abs(dev[k] - ora[k]), rt, np.isfinite(ora[k]))
            assert q <= 1.0, (label, k, q, "x tol at group", i)
        b = ora["residual_std_error"]
        m = np.isfinite(b) & np.isfinite(dev["residual_std_error"])
        sc = max(np.abs(b[m]).max(), 1e-300) if m.any() else 1.0
        q, i = worst(np.abs(dev["residual_std_error"] - b) / np.maximum(np.abs(b), 1e-12 * sc + 1e-300), rr, m)
        assert q <= 1.0, (label, "residual_std_error", q, "x tol at group", i)
        if inference:
            assert np.array_equal(dev["has_inference"], ora["has_inference"]), (label, "has_inference")
            for k in INF + ["p_values", "f_pvalue"]:
                assert nan_pattern_equal(dev[k], ora[k]), (label, "NaN pattern", k)
            for k in ["std_errors", "t_values", "f_statistic"]:
                a, b = dev[k], ora[k]
                m = np.isfinite(a) & np.isfinite(b)
                sc = max(np.abs(b[m]).max(), 1e-300) if m.any() else 1.0
                tol = 10 * (rr2 if a.ndim == 2 else rr)
                q, i = worst(np.abs(a - b) / np.maximum(np.abs(b), 1e-12 * sc + 1e-300), tol, m)
                assert q <= 1.0, (label, k, q, "x tol at", i)
            for k in ["ci_lower", "ci_upper"]:
                q, i = worst(np.abs(dev[k] - ora[k]) / np.maximum(cs, np.abs(ora[k])), 10 * rr2, np.isfinite(ora[k]))
                assert q <= 1.0, (label, k, q, "x tol at", i)
            for k in ["p_values", "f_pvalue"]:
                a, b = dev[k], ora[k]
                m = np.isfinite(b) & (b > 1e-290)
                # a p-value moves by ~ (1 + t^2) times the relative error of its statistic
                tol = ptol if scalar_tol else (rr2 if a.ndim == 2 else rr) * 10 * (1.0 + np.square(
                    np.nan_to_num(ora["t_values"] if a.ndim == 2 else np.sqrt(np.abs(ora["f_statistic"])), nan=0.0, posinf=0.0)))
                q, i = worst(np.abs(a - b) / np.where(m, b, 1.0), np.broadcast_to(tol, a.shape), m)
                assert q <= 1.0, (label, k, q, "x tol at", i)
                z = np.isfinite(b) & (b <= 1e-290)
                assert (a[z] <= 1e-280).all(), (label, k, "underflow region")
        if check_ic and "aic" in dev and "aic" in ora:
            for k in ["aic", "bic"]:
                m = np.isfinite(ora[k]) & np.isfinite(dev[k]) & (ora["r_squared"] < 1 - 1e-6)
                tol = np.maximum(1e-7, 100 * rt)
                q, i = worst(np.abs(dev[k] - ora[k]) / np.maximum(1.0, np.abs(ora[k])), tol, m)
                assert q <= 1.0, (label, k, q, "x tol at group", i)
