"""anofox_b200_ingest_device_static: the ingest whose scratch is a file-scope device array inside the library (no
allocation; include/anofox_b200.h). No counterpart in the reference.
Checked through the C ABI against the existing workspace entry point (anofox_b200_ingest_device), bit for bit (the
tolerance of the ingest tests: integer / byte work is exact), on two calls in a row that share the array, and with a
third, smaller call after them (stale contents of the shared scratch must not leak into a later result).
Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(N, G, p, seed, weights):
    rng = np.random.default_rng(seed)
    gid = rng.integers(0, G, size=N).astype(np.int64)
    if G > 10:
        gid[gid == 3] = 4  # an empty group
    return gid, rng.normal(size=N), rng.normal(size=(N, p)), (rng.uniform(0.5, 1.5, size=N) if weights else None)


def _static(b200, gid, y, x, w, G):
    N, p = x.shape
    dev = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    d_gid, d_y, d_x, d_w = dev(gid), dev(y), dev(x), dev(w)
    out = dict(offsets=torch.full((G + 1,), -7, dtype=torch.int64, device="cuda"),
               y=torch.full((N,), np.nan, dtype=torch.float64, device="cuda"),
               x=torch.full((N, p), np.nan, dtype=torch.float64, device="cuda"),
               w=torch.full((N,), np.nan, dtype=torch.float64, device="cuda") if w is not None else None,
               row_index=torch.full((N,), -7, dtype=torch.int64, device="cuda"),
               bad=torch.full((1,), -7, dtype=torch.int32, device="cuda"))
    torch.cuda.synchronize()
    ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
    b200.ingest_device_static(ptr(d_gid), ptr(d_y), ptr(d_x), ptr(d_w), N, p, G, ptr(out["offsets"]), ptr(out["y"]),
                              ptr(out["x"]), ptr(out["w"]), ptr(out["row_index"]), ptr(out["bad"]),
                              torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return {k: (None if v is None else v.cpu().numpy()) for k, v in out.items()}


def _workspace(b200, gid, y, x, w, G):
    N, p = x.shape
    dev = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    ing = b200.DeviceIngest(N, p, G, weights=w is not None)
    ing.run(dev(gid), dev(y), dev(x), dev(w))
    torch.cuda.synchronize()
    return dict(offsets=ing.offsets.cpu().numpy(), y=ing.y.cpu().numpy(), x=ing.x.cpu().numpy(),
                w=None if w is None else ing.w.cpu().numpy(), row_index=ing.row_index.cpu().numpy(),
                bad=np.array([ing.bad_ids()], dtype=np.int32))


def _same(a, b, label):
    for k in a:
        if a[k] is None:
            assert b[k] is None, (label, k)
            continue
        assert np.array_equal(a[k], b[k]), (label, k)


def test_static_ingest_matches_the_workspace_entry_point_twice_in_a_row(b200):
    b200.device_check()
    # two different calls back to back on the shared array (the second has more groups -> more radix passes and a
    # different carve of the scratch), then a small one whose result must not see what the big ones left behind
    calls = [(200_000, 300, 5, 1, True), (150_001, 70_000, 3, 2, False), (1_000, 7, 2, 3, False)]
    for i, (N, G, p, seed, weights) in enumerate(calls):
        gid, y, x, w = _case(N, G, p, seed, weights)
        if i == 1:
            gid[17] = G + 5  # an id out of range: flagged, filed under group 0, in both entry points
        got = _static(b200, gid, y, x, w, G)
        ref = _workspace(b200, gid, y, x, w, G)
        _same(got, ref, f"call {i}")
        assert got["bad"][0] == (1 if i == 1 else 0)
        if i != 1:  # and against numpy's stable sort, so that the two entry points cannot be wrong together
            order = np.argsort(gid, kind="stable")
            assert np.array_equal(got["row_index"], order)
            assert np.array_equal(got["x"], x[order])
    # the same call again: identical result from the array the previous calls dirtied
    gid, y, x, w = _case(*calls[0][:3], calls[0][3], calls[0][4])
    _same(_static(b200, gid, y, x, w, calls[0][1]), _workspace(b200, gid, y, x, w, calls[0][1]), "repeat")


def test_static_ingest_rejects_sizes_beyond_the_built_in_array(b200):
    b200.device_check()
    N, G = 2_000_000, 1000
    assert b200.load_library().anofox_b200_ingest_workspace_bytes(N, G) > b200.INGEST_STATIC_MAX_BYTES
    with pytest.raises(b200.B200Error, match="built-in array holds"):
        b200.ingest_device_static(None, None, None, None, N, 2, G, None, None, None, None, None, None, 0)
    # the limit itself is about the scratch, not the rows: a call just inside it is accepted (no rows are touched
    # before the argument checks, so bad pointers give the ordinary error, not a fault)
    with pytest.raises(b200.B200Error, match="bad arguments"):
        b200.ingest_device_static(None, None, None, None, 1000, 2, 10, None, None, None, None, None, None, 0)
