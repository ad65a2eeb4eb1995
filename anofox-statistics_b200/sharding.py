"""Multi-GPU partitioning of the group-fit path. Host logic only: no numerics.

Group-parallel workloads (ols/ridge/wls_fit_agg GROUP BY, *_fit_predict_by): the group-sorted CSR is cut
into `world` contiguous group-id ranges balanced by ROW count; every rank fits its own range and owns
the matching slice of the struct-of-arrays outputs. No collective is needed on the data path.

Single huge group: rows are cut into `world` contiguous row ranges (see distributed.py).
"""
import numpy as np


def group_ranges(offsets, world):
    """Cut points g[0..world] (g[0] = 0, g[world] = G) such that rank r owns groups g[r]:g[r+1].

    Cuts sit at the group boundary nearest to r * N / world rows, so that row counts -- the quantity the
    HBM-bound accumulation kernel is proportional to -- are balanced; groups are never split.
    """
    offsets = np.asarray(offsets, dtype=np.int64)
    G = len(offsets) - 1
    if world < 1:
        raise ValueError("world must be >= 1")
    if G < 0 or offsets[0] != 0 or np.any(np.diff(offsets) < 0):
        raise ValueError("offsets must start at 0 and be non-decreasing")
    N = int(offsets[-1])
    cuts = np.zeros(world + 1, dtype=np.int64)
    cuts[world] = G
    for r in range(1, world):
        target = (N * r) // world
        g = int(np.searchsorted(offsets, target, side="left"))  # first boundary at or after the target
        if g > 0 and target - offsets[g - 1] < offsets[min(g, G)] - target:
            g -= 1  # the boundary before the target is closer
        cuts[r] = min(max(g, cuts[r - 1]), G)
    return cuts


def local_csr(offsets, cuts, rank):
    """(local offsets rebased to 0, row range [r0, r1), group range [g0, g1)) of one rank."""
    offsets = np.asarray(offsets, dtype=np.int64)
    g0, g1 = int(cuts[rank]), int(cuts[rank + 1])
    r0, r1 = int(offsets[g0]), int(offsets[g1])
    return offsets[g0:g1 + 1] - r0, (r0, r1), (g0, g1)


def row_ranges(n_rows, world):
    """Contiguous row ranges of a single group, multiples of 4 rows except the last (4-row MMA steps)."""
    per = -(-n_rows // world)
    per = (per + 3) & ~3
    return [(min(r * per, n_rows), min((r + 1) * per, n_rows)) for r in range(world)]


def mixed_plan(offsets, world, giant_rows):
    """GROUP BY with a few giant groups among many ("mixed case").

    Groups of more than `giant_rows` rows are ROW-sharded over all ranks (every rank accumulates a slice of the
    group's rows, the moment blocks meet in one all-reduce: distributed.py); the remaining groups are cut into
    contiguous group-id ranges balanced by their row count WITHOUT the giants, and fitted whole by one rank each.

    Returns dict(giants=[g...], giant_rows={g: [(r0, r1) per rank]} (absolute row ranges),
                 cuts=[world + 1] group cut points, segments={rank: [(g0, g1), ...]} = the rank's range with the
                 giants cut out (each segment is a contiguous run of whole small groups)).
    """
    offsets = np.asarray(offsets, dtype=np.int64)
    G = len(offsets) - 1
    sizes = np.diff(offsets)
    giants = [int(g) for g in np.flatnonzero(sizes > giant_rows)]
    small = sizes.copy()
    small[giants] = 0
    so = np.zeros(G + 1, dtype=np.int64)
    np.cumsum(small, out=so[1:])
    cuts = group_ranges(so, world)
    segments = {}
    for r in range(world):
        g0, g1 = int(cuts[r]), int(cuts[r + 1])
        segs, start = [], g0
        for g in giants:
            if g0 <= g < g1:
                if g > start:
                    segs.append((start, g))
                start = g + 1
        if g1 > start:
            segs.append((start, g1))
        segments[r] = segs
    giant_ranges = {}
    for g in giants:
        base = int(offsets[g])
        giant_ranges[g] = [(base + a, base + b) for a, b in row_ranges(int(sizes[g]), world)]
    return dict(giants=giants, giant_rows=giant_ranges, cuts=cuts, segments=segments)
