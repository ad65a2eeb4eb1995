"""The oracle outside the well-conditioned regime (CPU): its long-double dqrdc2 path against a 60-digit mpmath
least-squares solve, and its aliasing decisions on either side of R's 1e-7 rule."""
import numpy as np
import pytest

from conditioned import KINDS, group_tolerances, make_conditioned

mp = pytest.importorskip("mpmath")


def mp_lstsq(x, y, live):
    """[intercept, slopes of live columns] by 60-digit QR."""
    mp.mp.dps = 60
    n = x.shape[0]
    cols = [j for j in range(x.shape[1]) if live[j]]
    A = mp.matrix(n, 1 + len(cols))
    for i in range(n):
        A[i, 0] = 1
        for c, j in enumerate(cols):
            A[i, c + 1] = mp.mpf(float(x[i, j]))
    b = mp.matrix([mp.mpf(float(v)) for v in y])
    sol, _ = mp.qr_solve(A, b)
    return np.array([float(v) for v in sol])


CASES = [(k, lv) for k, lvs in KINDS.items() for lv in lvs]


@pytest.mark.parametrize("kind,level", CASES, ids=[f"{k}-{lv}" for k, lv in CASES])
def test_oracle_matches_multiprecision(oracle, kind, level):
    rng = np.random.default_rng(77)
    c = make_conditioned(rng, kind, level, G=3, n=24, p=4, noise=0.5)
    ora = oracle.fit_batch(c["offsets"], c["y"], c["x"], None, compute_inference=True)
    assert (ora["status"] == 0).all()
    tol, kap = group_tolerances(c, ora)
    for g in range(3):
        a, b = c["offsets"][g], c["offsets"][g + 1]
        live = np.isfinite(ora["coefficients"][g])
        ref = mp_lstsq(c["x"][a:b], c["y"][a:b], live)
        got = np.concatenate([[ora["intercept"][g]], ora["coefficients"][g][live]])
        # the long-double oracle must sit far inside the tolerance the f64 device path is held to: slopes against
        # the largest slope, the intercept against the terms it is a difference of
        assert np.abs(got[1:] - ref[1:]).max() / tol["coef_scale"][g] <= 0.05 * tol["rtol"][g] + 1e-14, (kind, level, g, kap[g])
        assert abs(got[0] - ref[0]) / tol["icpt_scale"][g] <= 0.05 * tol["rtol"][g] + 1e-14, (kind, level, g, kap[g])
    if kind == "nearcol":
        # dqrdc2: the column is aliased iff its remaining norm is below 1e-7 of its own norm
        aliased = np.isnan(ora["coefficients"][:, -1])
        assert aliased.all() if level < 1e-7 else not aliased.any(), (level, aliased)
        assert np.isfinite(ora["coefficients"][:, :-1]).all()
