"""l3d_visual_neighbors / l3d_fundamentals on the device against the same source compiled for the host (which
tests/test_geometry_host.py pins against the oracle without a GPU): random cameras, many co-visible views, similarity
ties, a baseline limit that rejects views.  Every FP64 value of F must match bit for bit, every list exactly."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from line3d_b200 import capi  # noqa: E402

pytestmark = pytest.mark.gpu


def random_cameras(n, rng, spread=30.0):
    cams = np.zeros((n, capi.CAMERA_DOUBLES))
    for v in range(n):
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        if np.linalg.det(q) < 0:
            q[:, 0] = -q[:, 0]
        f = rng.uniform(800.0, 2500.0)
        K = np.array([[f, rng.normal(0, 0.01), rng.uniform(400, 1200)], [0.0, f * rng.uniform(0.98, 1.02), rng.uniform(300, 900)], [0.0, 0.0, 1.0]])
        C = rng.uniform(-spread, spread, 3)
        t = -q @ C
        cams[v, 0:9] = K.ravel(); cams[v, 9:18] = q.ravel(); cams[v, 18:21] = t; cams[v, 21:24] = -(q.T @ t)
    return cams


@pytest.mark.parametrize("n,seed", [(9, 1), (64, 2), (300, 3)])
def test_fundamentals_device_equals_host_build(n, seed):
    rng = np.random.default_rng(seed)
    cams = random_cameras(n, rng)
    pairs = np.array([(a, b) for a in range(n) for b in rng.choice(n, size=min(n - 1, 12), replace=False) if a != b], np.int32)
    ctx = capi.Context(0)
    try:
        got = ctx.fundamentals(cams, pairs)
    finally:
        ctx.close()
    want = capi.fundamentals_host(cams, pairs)
    assert got.shape == want.shape and np.array_equal(got.view(np.uint64), want.view(np.uint64))


@pytest.mark.parametrize("n,k,baseline,seed", [(9, 4, 0.25, 1), (64, 10, 8.0, 2), (300, -1, 6.0, 3), (300, 20, 12.0, 4), (170, 12, 0.0, 5)])
def test_neighbour_selection_device_equals_host_build(n, k, baseline, seed):
    rng = np.random.default_rng(seed)
    cams = random_cameras(n, rng)
    start, other, sim = [0], [], []
    for v in range(n):
        m = int(rng.integers(0, n))                       # some views with no co-visible view at all, some with nearly all
        o = np.sort(rng.choice([x for x in range(n) if x != v], size=min(m, n - 1), replace=False))
        s = rng.integers(1, 40, size=o.shape[0]).astype(np.float32) / np.float32(40.0)       # many ties
        other += o.tolist(); sim += s.tolist(); start.append(len(other))
    ctx = capi.Context(0)
    try:
        got = ctx.visual_neighbors(cams, start, other, sim, baseline, k)
    finally:
        ctx.close()
    want = capi.visual_neighbors_host(cams, start, other, sim, baseline, k)
    assert len(got) == len(want) == n
    for a, b in zip(got, want):
        assert np.array_equal(a, b)
    if k > 0:
        assert max(len(a) for a in got) <= k


def _resident_hypotheses(ctx, rng, n_views, n_seg):
    """Synthetic views and one hypothesis per segment, unprojected on the device (they stay resident); returns the
    end points as l3d_unproject reports them and the camera of each hypothesis."""
    views = []
    for v in range(n_views):
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        f = rng.uniform(900.0, 2000.0)
        Kinv = np.linalg.inv(np.array([[f, 0.0, 640.0], [0.0, f, 480.0], [0.0, 0.0, 1.0]]))
        segs = rng.uniform(0.0, 1200.0, size=(n_seg, 4)).astype(np.float32)
        views.append(dict(id=v, segs=segs, RtKinv=q.T @ Kinv, C=rng.uniform(-2.0, 2.0, 3), P=np.zeros((3, 4)), spatial_k=0.01, k_lower=0.001, k_upper=0.002))
    ctx.set_views(views)
    hin = np.zeros(n_views * n_seg, capi.HYP_IN_DTYPE)
    hin["view"] = np.repeat(np.arange(n_views), n_seg); hin["seg"] = np.tile(np.arange(n_seg), n_views)
    hin["depth1"] = rng.uniform(1.0, 4.0, hin.shape[0]).astype(np.float32); hin["depth2"] = rng.uniform(1.0, 4.0, hin.shape[0]).astype(np.float32)
    out = ctx.unproject(hin)
    pts = np.concatenate([out["P1"], out["P2"]], axis=1)
    return pts, hin["view"].astype(np.uint32)


@pytest.mark.parametrize("n_views,n_seg,sizes,seed", [(6, 40, [4, 5, 9, 17, 30], 1), (12, 300, [4, 64, 200, 777, 1500, 33, 4], 2), (40, 50, [40] * 45 + [400], 3)])
def test_line_fit_device_equals_host_build(n_views, n_seg, sizes, seed):
    rng = np.random.default_rng(seed)
    ctx = capi.Context(0)
    try:
        pts, cams = _resident_hypotheses(ctx, rng, n_views, n_seg)
        start, member = [0], []
        for m in sizes:                                   # clusters: random hypotheses, put in (camera, segment) order = index order
            member += np.sort(rng.choice(pts.shape[0], size=m, replace=False)).tolist()
            start.append(len(member))
        back = capi.similarity_back(rng.uniform(0.05, 0.3), np.linalg.qr(rng.normal(size=(3, 3)))[0], rng.normal(size=3))
        got = ctx.fit_lines(back, start, member, cams[member])
    finally:
        ctx.close()
    want = capi.fit_lines_host(back, start, pts[member], cams[member])
    assert len(got) == len(want) == len(sizes)
    assert sum(w.shape[0] for w in want) > 0
    for a, b in zip(got, want):
        assert a.shape == b.shape and np.array_equal(a.view(np.uint64), b.view(np.uint64))


def test_bad_arguments_are_reported():
    ctx = capi.Context(0)
    try:
        cams = np.zeros((2, capi.CAMERA_DOUBLES))
        with pytest.raises(capi.L3DError, match="out of range"):
            ctx.fundamentals(cams, [[0, 5]])
        with pytest.raises(capi.L3DError, match="ascending|out of range"):
            ctx.visual_neighbors(np.zeros((3, capi.CAMERA_DOUBLES)), [0, 2, 2, 2], [2, 1], [0.5, 0.5], 0.1, 3)
    finally:
        ctx.close()
