"""The one-pass fit-predict kernel (csrc/resident.cuh: TMA-staged whole groups -> DMMA moments -> register solve ->
predictions from shared memory) against the oracle, which fits each group's training rows and predicts every row as the
reference's *_fit_predict_agg does (src/aggregate_functions/ols_predict_aggregate.cpp:322-426). Group sizes are those
the kernel is chosen for by itself (tens of KB per group, up to one shared-memory stage), ragged, so that some groups
exceed a stage (streaming path + list-mode predictions), with held-out rows, NULL features, constant / duplicated /
nearly collinear columns (QR queue -> predictions redone) and noise-free groups (residuals taken from the rows)."""
import numpy as np
import pytest
from parity import assert_fit_parity
from test_gpu_parity import make_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(b200):
    b200.device_check()
    return b200


@pytest.fixture
def resident(lib):
    """The one-pass kernel is opt-in (force_path "resident"): on C5 it is issue-bound and a little slower than the
    three-kernel path it replaces."""
    lib.force_path("resident")
    yield lib
    lib.force_path("auto")


PK = ("yhat", "yhat_lower", "yhat_upper")


def _check_pred(pd, po, label):
    for k in ("yhat", "yhat_lower", "yhat_upper"):
        assert np.array_equal(np.isnan(pd[k]), np.isnan(po[k])), (label, k, np.flatnonzero(np.isnan(pd[k]) != np.isnan(po[k]))[:8])
        m = np.isfinite(po[k])
        scale = np.maximum(np.abs(po["yhat"][m]), 1.0)
        err = np.abs(pd[k][m] - po[k][m]) / scale
        assert err.max() < 1e-9, (label, k, err.max(), int(np.flatnonzero(m)[err.argmax()]))


CASES = [("ols", 8, (600, 1100), 60), ("ols", 3, (1500, 2600), 40), ("ridge", 5, (800, 1500), 50), ("wls", 8, (500, 900), 50),
         ("ols", 1, (2000, 3400), 24), ("wls", 4, (800, 1500), 40), ("ols", 7, (700, 1000), 40), ("ols", 6, (900, 1400), 40),
         ("ols", 2, (1800, 3200), 24)]


@pytest.mark.parametrize("method,p,nr,G", CASES, ids=[f"{c[0]}_p{c[1]}" for c in CASES])
def test_resident_fit_predict_parity(resident, oracle, method, p, nr, G):
    lib = resident
    rng = np.random.default_rng(400 + 13 * p + len(method))
    c = make_case(rng, G, nr[0], nr[1], p, dirty=True, weights=(method == "wls"), collinear=True)
    off, y, x, w = c["offsets"], c["y"], c["x"], c["w"]
    N = len(y)
    # nearly collinear columns in one group (small pivot -> QR queue) and a noise-free group (exact residual pass)
    if p >= 2:
        a, b = int(off[11]), int(off[12])
        x[a:b, 1] = x[a:b, 0] * (1.0 + 1e-7 * rng.normal(size=b - a))
    a, b = int(off[13]), int(off[14])
    xs = np.nan_to_num(x[a:b])
    y[a:b] = 1.5 + xs @ np.arange(1, p + 1)
    hold = rng.uniform(size=N) < 0.2
    y[hold] = np.nan  # held-out rows: predicted, not trained on
    assert lib.plan(G, N, p, method == "wls", "predict")[0] == "resident"
    opt = dict(method=method, confidence_level=0.9, alpha=0.5 if method == "ridge" else 0.0)
    pd, fd = lib.fit_predict_batch_host(off, y, x, w, None, **opt)
    po, fo = oracle.fit_predict_batch(off, y, x, w, None, **opt)
    assert np.array_equal(pd["is_training"], po["is_training"])
    skip = np.zeros(G, dtype=bool)
    skip[11] = True  # conditioned group: compared with its own tolerance below
    skip[13] = True  # noise-free group: its residual statistics are rounding noise, checked on their own below
    keep = ~skip
    assert np.array_equal(fd["status"], fo["status"])
    assert_fit_parity({k: v[keep] for k, v in fd.items()}, {k: v[keep] for k, v in fo.items()}, rtol=1e-9, label=f"resident/{method}")
    rows = np.ones(N, dtype=bool)
    rows[int(off[11]):int(off[12])] = False
    _check_pred({k: pd[k][rows] for k in PK}, {k: po[k][rows] for k in PK}, f"resident/{method}")
    if fo["status"][13] == 0 and method != "ridge":
        assert fd["r_squared"][13] >= 1 - 1e-13 and fd["residual_std_error"][13] < 1e-9
        live = np.isfinite(fo["coefficients"][13])
        assert np.array_equal(np.isfinite(fd["coefficients"][13]), live)
        assert np.abs(fd["coefficients"][13][live] - fo["coefficients"][13][live]).max() < 1e-9 * p
    # the nearly collinear group: predictions are well determined even where single coefficients are not
    r = slice(int(off[11]), int(off[12]))
    m = np.isfinite(po["yhat"][r])
    assert np.array_equal(np.isnan(pd["yhat"][r]), np.isnan(po["yhat"][r]))
    if m.any():
        assert (np.abs(pd["yhat"][r][m] - po["yhat"][r][m]) / np.maximum(np.abs(po["yhat"][r][m]), 1.0)).max() < 1e-6
    # explicit training flags
    train = (rng.uniform(size=N) < 0.7).astype(np.uint8)
    train[~np.isfinite(y)] = 0
    pd, fd = lib.fit_predict_batch_host(off, y, x, w, train, **opt)
    po, fo = oracle.fit_predict_batch(off, y, x, w, train, **opt)
    assert np.array_equal(pd["is_training"], train)
    assert np.array_equal(fd["status"], fo["status"])
    assert_fit_parity({k: v[keep] for k, v in fd.items()}, {k: v[keep] for k, v in fo.items()}, rtol=1e-9, label=f"resident-flags/{method}")
    _check_pred({k: pd[k][rows] for k in PK}, {k: po[k][rows] for k in PK}, f"resident-flags/{method}")


def test_resident_matches_the_three_kernel_path_bit_for_status(lib):
    """Same batch through the resident kernel and through accumulate -> solve -> predict: same NULL pattern, same
    statuses, values within 1e-9 (two summation orders of the same sums)."""
    rng = np.random.default_rng(5)
    c = make_case(rng, 80, 700, 1000, 8, dirty=True, weights=False, collinear=True)
    c["y"][rng.uniform(size=len(c["y"])) < 0.2] = np.nan
    try:
        lib.force_path("record")
        p0, f0 = lib.fit_predict_batch_host(c["offsets"], c["y"], c["x"], None, None, method="ols")
        lib.force_path("resident")
        p1, f1 = lib.fit_predict_batch_host(c["offsets"], c["y"], c["x"], None, None, method="ols")
    finally:
        lib.force_path("auto")
    assert np.array_equal(f0["status"], f1["status"])
    assert np.array_equal(p0["is_training"], p1["is_training"])
    _check_pred(p1, p0, "resident vs record")
