"""Ingest: unsorted (group_id, y, x[, w]) rows -> group-sorted CSR by the library's stable radix partition
(csrc/ingest.cuh), checked bit-exactly against numpy's stable argsort, and end to end
against the oracle fitted on the numpy-sorted rows (what *AggUpdate / *AggCombine + Finalize produce,
src/aggregate_functions/ols_aggregate.cpp:120-338)."""
import numpy as np
import pytest
import torch
from parity import assert_fit_parity

pytestmark = pytest.mark.gpu

CASES = [(1, 1, 1), (1000, 7, 3), (50_000, 300, 5), (200_000, 70_000, 2), (300_001, 1 << 17, 4), (100_000, 3, 16)]


@pytest.mark.parametrize("N,G,p", CASES, ids=[f"N{c[0]}_G{c[1]}_p{c[2]}" for c in CASES])
def test_ingest_is_a_stable_partition(b200, N, G, p):
    b200.device_check()
    rng = np.random.default_rng(N + G)
    gid = rng.integers(0, G, size=N).astype(np.int64)
    if G > 10:
        gid[gid == 3] = 4  # an empty group in the middle
    y = rng.normal(size=N)
    x = rng.normal(size=(N, p))
    w = rng.uniform(0.5, 1.5, size=N)
    ing = b200.DeviceIngest(N, p, G, weights=True)
    dev = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    ing.run(dev(gid), dev(y), dev(x), dev(w))
    torch.cuda.synchronize()
    order = np.argsort(gid, kind="stable")
    counts = np.bincount(gid, minlength=G)
    off = np.concatenate([[0], np.cumsum(counts)])
    assert np.array_equal(ing.offsets.cpu().numpy(), off)
    assert np.array_equal(ing.row_index.cpu().numpy(), order)  # stable: arrival order inside every group
    assert np.array_equal(ing.y.cpu().numpy(), y[order])
    assert np.array_equal(ing.x.cpu().numpy(), x[order])
    assert np.array_equal(ing.w.cpu().numpy(), w[order])
    assert ing.bad_ids() == 0


def test_ingest_flags_ids_out_of_range(b200):
    gid = np.array([0, 5, 2, -1, 1], dtype=np.int64)
    ing = b200.DeviceIngest(5, 1, 3)
    dev = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    ing.run(dev(gid), dev(np.zeros(5)), dev(np.zeros((5, 1))))
    torch.cuda.synchronize()
    assert ing.bad_ids() == 1
    assert ing.offsets.cpu().numpy()[-1] == 5  # still a well-formed partition


def test_unsorted_rows_to_fits_match_the_oracle(b200, oracle):
    """The whole aggregate path: scan-order rows -> ingest -> batch fit, against the oracle on numpy-sorted rows."""
    rng = np.random.default_rng(9)
    G, p = 4000, 5
    sizes = rng.integers(0, 90, size=G)
    gid = np.repeat(np.arange(G), sizes)
    rng.shuffle(gid)
    N = gid.size
    x = rng.uniform(-10, 10, size=(N, p))
    beta = rng.uniform(-5, 5, size=(G, p))
    y = rng.uniform(-10, 10, size=G)[gid] + np.einsum("ij,ij->i", x, beta[gid]) + rng.normal(0, 2.0, size=N)
    y[::41] = np.nan
    cst = rng.uniform(size=G) < 0.05
    x[cst[gid], 2] = 1.0 + 3e-11 * rng.integers(0, 2, size=int(cst[gid].sum()))  # the 1e-10 rule looks at the FIRST row
    ing = b200.DeviceIngest(N, p, G)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    ing.run(dev(gid.astype(np.int64)), dev(y), dev(x))
    fit = b200.DeviceFit(G, N, p, inference=True)
    out = fit.fit(ing.offsets, ing.y, ing.x, None, method="ols", compute_inference=True)
    torch.cuda.synchronize()
    got = {k: v.cpu().numpy() for k, v in out.items()}
    order = np.argsort(gid, kind="stable")
    off = np.concatenate([[0], np.cumsum(np.bincount(gid, minlength=G))]).astype(np.int64)
    ora = oracle.fit_batch(off, y[order], x[order], None, method="ols", compute_inference=True)
    assert_fit_parity(got, ora, rtol=1e-9, inference=True, label="ingest -> fit")


def test_ingest_and_column_fits_degenerate_inputs(b200, oracle):
    """Zero rows, every row in one group, and a column-fit batch that holds empty fits."""
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    # no rows: offsets all zero
    ing = b200.DeviceIngest(0, 2, 5)
    ing.run(dev(np.zeros(0, dtype=np.int64)), dev(np.zeros(0)), dev(np.zeros((0, 2))))
    torch.cuda.synchronize()
    assert np.array_equal(ing.offsets.cpu().numpy(), np.zeros(6, dtype=np.int64)) and ing.bad_ids() == 0
    # every row in the last of many groups
    N, G = 5000, 1 << 16
    rng = np.random.default_rng(3)
    y, x = rng.normal(size=N), rng.normal(size=(N, 2))
    ing = b200.DeviceIngest(N, 2, G)
    ing.run(dev(np.full(N, G - 1, dtype=np.int64)), dev(y), dev(x))
    torch.cuda.synchronize()
    off = ing.offsets.cpu().numpy()
    assert off[-1] == N and (off[:-1] == 0).all()
    assert np.array_equal(ing.y.cpu().numpy(), y) and np.array_equal(ing.x.cpu().numpy(), x)
    # column fits: empty fits between ordinary ones give the reference's error code for them, the others fit
    sizes = np.array([0, 12, 0, 0, 30, 1])
    offc = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    p = 2
    ys, blocks, cols = [], [], []
    for n in sizes:
        X = rng.normal(size=(int(n), p))
        ys.append(1.0 + X @ np.array([0.5, -2.0]) + rng.normal(0, 0.1, size=int(n)))
        blocks.append(X.T.reshape(-1))
        cols.append([X[:, j].copy() for j in range(p)])
    fit = b200.DeviceFit(len(sizes), int(offc[-1]), p, inference=False)
    out = fit.fit_columns(dev(offc), dev(np.concatenate(ys)), dev(np.concatenate(blocks)), None, method="ols")
    torch.cuda.synchronize()
    st = out["status"].cpu().numpy()
    for i, n in enumerate(sizes):
        if n == 0:
            assert st[i] != 0, i
            continue
        r = oracle.single_fit("ols", ys[i], cols[i], None)
        assert st[i] == (0 if r.ok else r.error_code), (i, st[i])
        if r.ok:
            assert np.allclose(out["coefficients"][i].cpu().numpy(), np.asarray(r.coefficients), rtol=1e-9, atol=1e-12,
                               equal_nan=True)
