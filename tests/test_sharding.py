"""Multi-GPU host logic on CPU (no GPU needed): group-range sharding and the row-sharded single-group
protocol of anofox-statistics_b200/distributed.py run over gloo with world_size 2. The device compute is
replaced by a numpy backend that writes the moment block in the layout include/anofox_b200.h documents,
and the result is checked against the CPU oracle on the unsharded data."""
import os
import socket

import numpy as np
import pytest

from __graft_entry__ import load_package


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


# ------------------------------------------------------------------ group-range sharding
def test_group_ranges_partition_and_balance():
    pkg = load_package()
    from importlib import import_module
    sh = import_module(pkg.__name__ + ".sharding")
    rng = np.random.default_rng(0)
    for G, world in [(1000, 8), (7, 8), (1, 4), (0, 2), (100, 1), (64, 3)]:
        sizes = rng.integers(0, 200, size=G)
        offsets = np.zeros(G + 1, dtype=np.int64)
        np.cumsum(sizes, out=offsets[1:])
        cuts = sh.group_ranges(offsets, world)
        assert cuts[0] == 0 and cuts[-1] == G and np.all(np.diff(cuts) >= 0)
        seen = []
        for r in range(world):
            loc, (r0, r1), (g0, g1) = sh.local_csr(offsets, cuts, r)
            assert loc[0] == 0 and loc[-1] == r1 - r0 and len(loc) == g1 - g0 + 1
            assert np.array_equal(np.diff(loc), sizes[g0:g1])
            seen.extend(range(g0, g1))
        assert seen == list(range(G))  # every group exactly once, in order
        if G >= 100 * world:
            rows = np.array([offsets[cuts[r + 1]] - offsets[cuts[r]] for r in range(world)])
            assert rows.max() - rows.min() <= 2 * sizes.max()  # balanced by rows up to one group
    # one giant group among small ones is never split
    offsets = np.array([0, 10, 20, 100020, 100030], dtype=np.int64)
    cuts = sh.group_ranges(offsets, 4)
    assert cuts[-1] == 4 and sum(cuts[r + 1] > cuts[r] for r in range(4)) >= 1
    with pytest.raises(ValueError):
        sh.group_ranges(np.array([0, 5, 3]), 2)
    assert sh.row_ranges(10, 4) == [(0, 4), (4, 8), (8, 10), (10, 10)]
    rr = sh.row_ranges(1_000_000_007, 8)
    assert rr[0][0] == 0 and rr[-1][1] == 1_000_000_007 and all(a[1] == b[0] for a, b in zip(rr, rr[1:]))
    assert all((a[1] - a[0]) % 4 == 0 for a in rr[:-1])


# ------------------------------------------------------------------ numpy stand-in for the device compute
class NumpyBackend:
    """Writes exactly what anofox_b200_first_valid_row_device / anofox_b200_moments_device document."""

    def __init__(self, p):
        self.p = p

    @staticmethod
    def _valid(y, x, w):
        ok = np.isfinite(y) & np.isfinite(x).all(axis=1)
        if w is not None:
            ok &= np.isfinite(w) & (w > 0)
        return ok

    def first_valid(self, y, x, w):
        import torch
        yn, xn = y.numpy(), x.numpy()
        ok = self._valid(yn, xn, None if w is None else w.numpy())
        out = np.zeros(self.p + 2)
        if ok.any():
            f = int(np.argmax(ok))
            out[0] = 1.0
            out[1:self.p + 1] = xn[f]
            out[self.p + 1] = yn[f]
        return torch.from_numpy(out)

    def moments(self, y, x, w, shift):
        import torch
        p, k = self.p, self.p + 1
        yn, xn = y.numpy(), x.numpy()
        wn = None if w is None else w.numpy()
        ok = self._valid(yn, xn, wn)
        s = shift.numpy()
        d = np.concatenate([xn[ok] - s[:p], (yn[ok] - s[p])[:, None]], axis=1)
        ww = np.ones(ok.sum()) if wn is None else wn[ok]
        S1 = (ww[:, None] * d).sum(axis=0)
        S2 = (ww[:, None] * d).T @ d
        m = np.zeros(4 + k + k * (k + 1) // 2)
        m[0], m[1], m[2] = ok.sum(), ww.sum(), len(yn)
        m[4:4 + k] = S1
        m[4 + k:] = S2[np.triu_indices(k)]
        cd = np.abs(d[:, :p]).max(axis=0) if ok.any() else np.zeros(p)
        return torch.from_numpy(m), torch.from_numpy(cd)


def solve_from_block(m, cd, shift, p):
    """Plain normal equations from the reduced block (OLS with intercept, no aliasing): the block layout check."""
    k = p + 1
    n, sw = m[0], m[1]
    S1 = m[4:4 + k]
    S2 = np.zeros((k, k))
    S2[np.triu_indices(k)] = m[4 + k:]
    S2 = S2 + np.triu(S2, 1).T
    C = S2 - np.outer(S1, S1) / sw  # centred
    beta = np.linalg.solve(C[:p, :p], C[:p, p])
    mean = shift + S1 / sw
    b0 = mean[p] - mean[:p] @ beta
    rss = C[p, p] - C[:p, p] @ beta
    return beta, b0, 1.0 - rss / C[p, p], int(n)


def _worker(rank, world, port, case, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pkg = load_package()
        from importlib import import_module
        D = import_module(pkg.__name__ + ".distributed")
        sh = import_module(pkg.__name__ + ".sharding")
        y, x, w = case["y"], case["x"], case["w"]
        if case["kind"] == "rows":
            a, b = case["ranges"][rank]
            t = lambda v: None if v is None else torch.from_numpy(np.ascontiguousarray(v[a:b]))  # noqa: E731
            m, cd, shift = D.sharded_moments(NumpyBackend(x.shape[1]), t(y), t(x), t(w), dist)
            q.put((rank, m.numpy().copy(), cd.numpy().copy(), shift.numpy().copy()))
        else:  # group-range sharding: each rank fits its own groups (oracle as the stand-in compute), gather
            import oracle
            cuts = sh.group_ranges(case["offsets"], world)
            loc, (r0, r1), (g0, g1) = sh.local_csr(case["offsets"], cuts, rank)
            out = oracle.fit_batch(loc, y[r0:r1], x[r0:r1], None if w is None else w[r0:r1], compute_inference=True)
            q.put((rank, g0, g1, out["coefficients"].copy(), out["status"].copy(), out["p_values"].copy()))
    finally:
        dist.destroy_process_group()


def _run(world, case):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    return sorted(res, key=lambda t: t[0])


@pytest.mark.parametrize("variant", ["plain", "rank0_starts_invalid", "rank0_all_invalid", "weighted"])
def test_row_sharded_protocol_over_gloo(oracle, variant):
    rng = np.random.default_rng(3)
    N, p = 4001, 6
    x = rng.normal(size=(N, p))
    x[:, 3] = 2.5  # a constant column: cdev must stay 0 after the MAX reduce
    y = 1.0 + x @ rng.uniform(-2, 2, size=p) + rng.normal(0, 0.3, size=N)
    w = rng.uniform(0.5, 1.5, size=N) if variant == "weighted" else None
    ranges = [(0, 2000), (2000, N)]
    if variant == "rank0_starts_invalid":
        y[:5] = np.nan
        x[5, 1] = np.inf
    if variant == "rank0_all_invalid":
        y[:2000] = np.nan
    if variant == "weighted":
        w[0] = 0.0
        w[10] = -1.0
    res = _run(2, dict(kind="rows", y=y, x=x, w=w, ranges=ranges))
    (_, m0, c0, s0), (_, m1, c1, s1) = res
    assert np.array_equal(m0, m1) and np.array_equal(c0, c1) and np.array_equal(s0, s1)  # identical on every rank
    ok = NumpyBackend._valid(y, x, w)
    f = int(np.argmax(ok))
    assert np.array_equal(s0[:p], x[f]) and s0[p] == y[f]  # the group's first valid row, whichever rank holds it
    import torch
    full_m, full_c = NumpyBackend(p).moments(torch.from_numpy(y), torch.from_numpy(x),
                                             None if w is None else torch.from_numpy(w), torch.from_numpy(s0))
    assert np.allclose(m0, full_m.numpy(), rtol=1e-12, atol=1e-9)
    assert np.array_equal(c0, full_c.numpy()) and c0[3] == 0.0
    assert m0[0] == ok.sum() and m0[2] == N
    # the reduced block reproduces the oracle's fit of the unsharded group (constant column dropped by hand)
    keep = [j for j in range(p) if c0[j] >= 1e-10]
    k = p + 1
    S2 = np.zeros((k, k))
    S2[np.triu_indices(k)] = m0[4 + k:]
    idx = keep + [p]
    sub = np.concatenate([m0[:4], m0[4:4 + k][idx], S2[np.ix_(idx, idx)][np.triu_indices(len(idx))]])
    beta, b0, r2, n = solve_from_block(sub, c0[keep], s0[idx], len(keep))
    ora = oracle.fit_batch(np.array([0, N], dtype=np.int64), y, x, w, method="wls" if w is not None else "ols")
    assert ora["status"][0] == 0 and n == ora["n_observations"][0]
    assert np.isnan(ora["coefficients"][0, 3])
    assert np.allclose(beta, ora["coefficients"][0, keep], rtol=1e-9)
    assert abs(b0 - ora["intercept"][0]) < 1e-9 and abs(r2 - ora["r_squared"][0]) < 1e-10


def test_group_sharded_results_over_gloo(oracle):
    rng = np.random.default_rng(4)
    G, p = 301, 3
    sizes = rng.integers(0, 60, size=G)
    offsets = np.zeros(G + 1, dtype=np.int64)
    np.cumsum(sizes, out=offsets[1:])
    N = int(offsets[-1])
    x = rng.uniform(-10, 10, size=(N, p))
    y = x @ np.array([1.0, -2.0, 0.5]) + rng.normal(0, 1, size=N)
    res = _run(2, dict(kind="groups", y=y, x=x, w=None, offsets=offsets))
    full = oracle.fit_batch(offsets, y, x, None, compute_inference=True)
    covered = 0
    for _, g0, g1, coef, status, pv in res:
        assert np.array_equal(status, full["status"][g0:g1])
        assert np.array_equal(coef, full["coefficients"][g0:g1], equal_nan=True)  # same code, same rows -> same bits
        assert np.array_equal(pv, full["p_values"][g0:g1], equal_nan=True)
        covered += g1 - g0
    assert covered == G


def test_mixed_plan_splits_only_the_giants():
    """SURVEY 8(e) "mixed case": groups above the threshold are row-sharded over all ranks, the others stay whole."""
    pkg = load_package()
    from importlib import import_module
    sh = import_module(pkg.__name__ + ".sharding")
    rng = np.random.default_rng(8)
    for world in (1, 2, 4, 8):
        G = 500
        sizes = rng.integers(0, 100, size=G)
        sizes[[17, 250, 499]] = [50_000, 200_001, 77_777]
        offsets = np.zeros(G + 1, dtype=np.int64)
        np.cumsum(sizes, out=offsets[1:])
        plan = sh.mixed_plan(offsets, world, giant_rows=10_000)
        assert plan["giants"] == [17, 250, 499]
        owned = []
        for r in range(world):
            for g0, g1 in plan["segments"][r]:
                assert g0 < g1 and not any(g0 <= g < g1 for g in plan["giants"])
                owned.extend(range(g0, g1))
        assert sorted(owned) == [g for g in range(G) if g not in plan["giants"]]  # every small group exactly once
        small_rows = [sum(int(sizes[g]) for g0, g1 in plan["segments"][r] for g in range(g0, g1)) for r in range(world)]
        assert max(small_rows) - min(small_rows) <= 2 * 100  # balanced without the giants
        for g in plan["giants"]:
            rr = plan["giant_rows"][g]
            assert len(rr) == world and rr[0][0] == offsets[g] and rr[-1][1] == offsets[g + 1]
            assert all(a[1] == b[0] for a, b in zip(rr, rr[1:]))
    assert sh.mixed_plan(np.array([0, 5, 9]), 2, giant_rows=100)["giants"] == []
