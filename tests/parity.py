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


def assert_fit_parity(dev, ora, rtol=1e-9, inference=False, ptol=1e-8, check_ic=True, label=""):
    """Device result dict vs oracle result dict (same SoA layout).

    Bit-exact: status (NULL-struct pattern), n_observations, NaN positions of every array.
    Relative tolerance rtol (north star: 1e-9 for well-conditioned f64) on values; p-values relative ptol.
    """
    assert np.array_equal(dev["status"], ora["status"]), (label, "status", np.flatnonzero(dev["status"] != ora["status"])[:10])
    bad = np.flatnonzero(dev["n_observations"] != ora["n_observations"])
    assert bad.size == 0, (label, "n_observations", bad[:10], dev["n_observations"][bad[:10]], ora["n_observations"][bad[:10]])
    for k in CORE:
        assert nan_pattern_equal(dev[k], ora[k]), (label, "NaN pattern", k)
    # coefficients are compared against the per-group coefficient scale (a slope of 1e-3 next to one of 5 is
    # only determined to 1e-9 * 5 by either implementation)
    c_dev, c_ora = dev["coefficients"], ora["coefficients"]
    scale = np.nanmax(np.abs(np.concatenate([c_ora, ora["intercept"][:, None]], axis=1)), axis=1, initial=0.0)
    scale = np.where(scale > 0, scale, 1.0)[:, None]
    m = np.isfinite(c_ora)
    err = np.abs(c_dev - c_ora)[m] / np.broadcast_to(scale, c_ora.shape)[m]
    assert err.size == 0 or err.max() <= rtol, (label, "coefficients", err.max())
    m = np.isfinite(ora["intercept"])
    err = np.abs(dev["intercept"] - ora["intercept"])[m] / scale[:, 0][m]
    assert err.size == 0 or err.max() <= rtol, (label, "intercept", err.max())
    for k in ["r_squared", "adj_r_squared"]:
        m = np.isfinite(ora[k])
        err = np.abs(dev[k] - ora[k])[m]
        assert err.size == 0 or err.max() <= rtol, (label, k, err.max())
    e = max_rel_err(dev["residual_std_error"], ora["residual_std_error"])
    assert e <= rtol, (label, "residual_std_error", e)
    if inference:
        assert np.array_equal(dev["has_inference"], ora["has_inference"]), (label, "has_inference")
        for k in INF + ["p_values", "f_pvalue"]:
            assert nan_pattern_equal(dev[k], ora[k]), (label, "NaN pattern", k)
        for k in ["std_errors", "t_values", "f_statistic"]:
            e = max_rel_err(dev[k], ora[k])
            assert e <= 10 * rtol, (label, k, e)
        for k in ["ci_lower", "ci_upper"]:
            m = np.isfinite(ora[k])
            err = np.abs(dev[k] - ora[k])[m] / np.broadcast_to(scale, ora[k].shape)[m]
            assert err.size == 0 or err.max() <= 10 * rtol, (label, k, err.max())
        for k in ["p_values", "f_pvalue"]:
            a, b = dev[k], ora[k]
            m = np.isfinite(b) & (b > 1e-290)
            err = np.abs(a[m] - b[m]) / b[m]
            assert err.size == 0 or err.max() <= ptol, (label, k, err.max())
            z = np.isfinite(b) & (b <= 1e-290)
            assert (a[z] <= 1e-280).all(), (label, k, "underflow region")
    if check_ic and "aic" in dev and "aic" in ora:
        for k in ["aic", "bic"]:
            m = np.isfinite(ora[k]) & np.isfinite(dev[k]) & (ora["r_squared"] < 1 - 1e-6)
            err = np.abs(dev[k] - ora[k])[m] / np.maximum(1.0, np.abs(ora[k][m]))
            assert err.size == 0 or err.max() <= 1e-7, (label, k, err.max())
