"""Seeded synthetic workloads of the shapes BASELINE.json names.

Distributions follow the reference's own generators: x ~ U(-10,10), intercept ~ U(-10,10),
slopes ~ U(-5,5), noise ~ N(0, 2^2) (examples/performance_10k_groups_R/generate_test_data.sql:37-84),
WLS weights ~ U(0.5,1.5) (examples/performance_1m_groups/benchmark_wls.sql:14). Data are group-sorted
CSR: offsets[G+1], y[N], X[N,p] row-major, w[N].
"""
import numpy as np

CONFIGS = {
    # name: (groups, rows per group, features)
    "C1": (10_000, 100, 3),
    "C2": (1_000_000, 50, 5),
    "C3": (1_000_000, 200, 16),
    "C4": (1, 1_000_000_000, 64),
    "C5": (100_000, 1_000, 8),
}


def make_numpy(G, n, p, seed=0, ragged=False, weights=False, rank_deficient=False, holdout=0.0):
    """Host (numpy) workload. ragged: group sizes uniform in [n/2, 3n/2]."""
    rng = np.random.default_rng(seed)
    sizes = rng.integers(max(n // 2, 1), n + n // 2 + 1, size=G) if ragged else np.full(G, n)
    offsets = np.zeros(G + 1, dtype=np.int64)
    np.cumsum(sizes, out=offsets[1:])
    N = int(offsets[-1])
    x = rng.uniform(-10, 10, size=(N, p))
    gid = np.repeat(np.arange(G), sizes)
    b0 = rng.uniform(-10, 10, size=G)
    beta = rng.uniform(-5, 5, size=(G, p))
    if rank_deficient:
        # 5 % one constant column, 5 % one duplicated column (col 7 := col 2 when p allows), 1 % both
        u = rng.uniform(size=G)
        cst = (u < 0.05) | (u > 0.99)
        dup = ((u >= 0.05) & (u < 0.10)) | (u > 0.99)
        cc, d_to, d_from = (1 % p), min(7, p - 1), min(2, p - 1)
        rows_c = cst[gid]
        x[rows_c, cc] = 3.25
        if d_to != d_from:
            rows_d = dup[gid]
            x[rows_d, d_to] = x[rows_d, d_from]
    y = b0[gid] + np.einsum("ij,ij->i", x, beta[gid]) + rng.normal(0, 2.0, size=N)
    w = rng.uniform(0.5, 1.5, size=N) if weights else None
    train = None
    if holdout > 0:
        train = (rng.uniform(size=N) >= holdout).astype(np.uint8)
        y = np.where(train == 1, y, np.nan)
    return dict(offsets=offsets, y=y, x=x, w=w, train=train, G=G, N=N, p=p)


def make_torch(G, n, p, seed=0, weights=False, rank_deficient=False, holdout=0.0, device="cuda", chunk=1 << 24):
    """Device-resident workload of equal-size groups, generated in chunks (C3 is 27 GB)."""
    import torch
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    N = G * n
    dev = torch.device(device)
    offsets = torch.arange(0, G + 1, dtype=torch.int64, device=dev) * n
    x = torch.empty((N, p), dtype=torch.float64, device=dev)
    y = torch.empty(N, dtype=torch.float64, device=dev)
    b0 = torch.rand(G, generator=gen, dtype=torch.float64, device=dev) * 20 - 10
    beta = torch.rand((G, p), generator=gen, dtype=torch.float64, device=dev) * 10 - 5
    u = torch.rand(G, generator=gen, dtype=torch.float64, device=dev)
    gchunk = max(1, chunk // n)
    for g0 in range(0, G, gchunk):
        g1 = min(G, g0 + gchunk)
        r0, r1 = g0 * n, g1 * n
        xs = torch.rand((r1 - r0, p), generator=gen, dtype=torch.float64, device=dev) * 20 - 10
        if rank_deficient:
            uu = u[g0:g1].repeat_interleave(n)
            cst = (uu < 0.05) | (uu > 0.99)
            dup = ((uu >= 0.05) & (uu < 0.10)) | (uu > 0.99)
            cc, d_to, d_from = (1 % p), min(7, p - 1), min(2, p - 1)
            xs[cst, cc] = 3.25
            if d_to != d_from:
                xs[dup, d_to] = xs[dup, d_from]
        bb = beta[g0:g1].repeat_interleave(n, dim=0)
        ys = b0[g0:g1].repeat_interleave(n) + (xs * bb).sum(dim=1)
        ys += torch.randn(r1 - r0, generator=gen, dtype=torch.float64, device=dev) * 2.0
        x[r0:r1] = xs
        y[r0:r1] = ys
        del xs, ys, bb
    w = None
    if weights:
        w = torch.rand(N, generator=gen, dtype=torch.float64, device=dev) + 0.5
    train = None
    if holdout > 0:
        train = (torch.rand(N, generator=gen, dtype=torch.float64, device=dev) >= holdout).to(torch.uint8)
        y = torch.where(train == 1, y, torch.full_like(y, float("nan")))
    return dict(offsets=offsets, y=y, x=x, w=w, train=train, G=G, N=N, p=p)
