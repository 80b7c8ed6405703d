"""The neighbour graph and the fundamental matrices (l3d_visual_neighbors, l3d_fundamentals) are one
source compiled twice: for the device (the product) and for the host (l3d_debug_*_host).  Here, without a GPU, the host
build is pinned against the oracle's dumps -- which are pinned against the reference's own sources: the fundamental
matrices must come out bit for bit in FP64, the neighbour lists identical.  tests/test_gpu_end_to_end.py then checks the
device build against the same dumps through Line3D."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from line3d_b200 import capi, scenes  # noqa: E402
from oracle import oracle_py  # noqa: E402


def cameras_from_dump(views):
    """views.npy rows: id, median, k_upper, k_lower, spatial_k, C[3], RtKinv[9], P[12], K[9], R[9], t[3] (after the scene
    normalisation, as the fundamental matrices see them)."""
    cams = np.zeros((views.shape[0], capi.CAMERA_DOUBLES))
    cams[:, 0:9] = views[:, 29:38]; cams[:, 9:18] = views[:, 38:47]; cams[:, 18:21] = views[:, 47:50]; cams[:, 21:24] = views[:, 5:8]
    return cams


@pytest.mark.parametrize("maker,kw,params", [
    (scenes.tiny, dict(), dict(neighbors=5)),
    (scenes.urban, dict(n_views=14, n_seg=60, seed=31, blocks=(2, 2)), dict(neighbors=4)),
    (scenes.dense_dome, dict(n_views=12, n_seg=50, seed=5), dict(neighbors=-1)),
])
def test_fundamentals_host_build_equals_the_oracle_bit_for_bit(tmp_path, maker, kw, params):
    scn = str(tmp_path / "s.scn")
    scenes.write_scene(maker(**kw), scn)
    d = str(tmp_path / "dump"); os.makedirs(d)
    _, _, _, dm = oracle_py.run_scene(scn, d, **params)
    views, fund = dm["views"], dm["fundamentals"]
    assert fund.shape[0] > 0
    index = {int(v): k for k, v in enumerate(views[:, 0])}
    # the host computes F for (view, neighbour) when it first meets the pair, in view order, and files the transpose
    # for the reverse pair (line3D.cc:1949-1965): which direction was computed follows from the neighbour lists
    computed, have = [], set()
    for a, b in dm["visual_neighbors"]:
        a, b = int(a), int(b)
        if (a, b) not in have:
            computed.append((a, b)); have.add((a, b)); have.add((b, a))
    assert len(have) == fund.shape[0]
    F = capi.fundamentals_host(cameras_from_dump(views), np.array([[index[a], index[b]] for a, b in computed], np.int32))
    got = {}
    for (a, b), f in zip(computed, F):
        got[(a, b)] = f; got[(b, a)] = f.T
    for row in fund:
        assert np.array_equal(got[(int(row[0]), int(row[1]))].ravel(), row[2:11])          # FP64, every bit


def _centre(R, t):
    """C = R^T (-t): the product accumulates from zero with ascending inner index (python floats are IEEE doubles)."""
    out = []
    for i in range(3):
        s = 0.0
        for k in range(3):
            s = s + float(R[k][i]) * (-float(t[k]))
        out.append(s)
    return out


def _similarities(sc):
    """Line3D::processWorldpoints + the similarity of findVisualNeighbors (line3D.cc:236-281, 483-495), restated: the
    counts are built incrementally in the order the images are added, and a worldpoint starts to count when its second
    view arrives."""
    wp2views, common, num = {}, {}, {}
    def bump(a, b):
        common.setdefault(a, {}); common.setdefault(b, {})
        common[a][b] = common[a].get(b, 0) + 1; common[b][a] = common[b].get(a, 0) + 1
    for v in sc.views:
        num[v.id] = 0
        for wp in v.wps.tolist():
            seen = wp2views.setdefault(wp, {})
            if len(seen) == 2:
                v1, v2 = sorted(seen)
                bump(v1, v2); num[v1] += 1; num[v2] += 1
            if len(seen) >= 2:
                for w in sorted(seen):
                    bump(w, v.id)
                num[v.id] += 1
            seen[v.id] = True
    sims = {}
    for a in sorted(common):
        for b in sorted(common[a]):
            s = np.float32(2.0) * np.float32(common[a][b]) / np.float32(num[a] + num[b])
            if s > 1e-12:
                sims.setdefault(a, {})[b] = s
    return sims


@pytest.mark.parametrize("maker,kw,k,min_baseline", [
    (scenes.tiny, dict(), 5, 0.25),
    (scenes.urban, dict(n_views=14, n_seg=60, seed=31, blocks=(2, 2)), 4, 0.25),
    (scenes.dense_dome, dict(n_views=12, n_seg=50, seed=5), -1, 0.25),
    (scenes.urban, dict(n_views=16, n_seg=60, seed=9, blocks=(2, 2)), 6, 25.0),          # a baseline limit that actually rejects views
    (scenes.dense_dome, dict(n_views=14, n_seg=50, seed=6), 5, 50.0),
])
def test_neighbour_selection_host_build_equals_the_oracle(tmp_path, maker, kw, k, min_baseline):
    sc = maker(**kw)
    scn = str(tmp_path / "s.scn")
    scenes.write_scene(sc, scn)
    d = str(tmp_path / "dump"); os.makedirs(d)
    _, _, _, dm = oracle_py.run_scene(scn, d, neighbors=k, min_baseline=min_baseline)
    want = {}
    for a, b in dm["visual_neighbors"]:
        want.setdefault(int(a), []).append(int(b))
    ids = sorted(v.id for v in sc.views)
    index = {vid: i for i, vid in enumerate(ids)}
    by_id = {v.id: v for v in sc.views}
    cams = np.zeros((len(ids), capi.CAMERA_DOUBLES))
    for vid in ids:
        v = by_id[vid]
        cams[index[vid], 0:9] = np.asarray(v.K, np.float64).ravel(); cams[index[vid], 9:18] = np.asarray(v.R, np.float64).ravel()
        cams[index[vid], 18:21] = np.asarray(v.t, np.float64).ravel(); cams[index[vid], 21:24] = _centre(np.asarray(v.R, np.float64), np.asarray(v.t, np.float64))
    sims = _similarities(sc)
    start, other, sim = [0], [], []
    for vid in ids:
        for o in sorted(sims.get(vid, {})):
            other.append(index[o]); sim.append(sims[vid][o])
        start.append(len(other))
    got = capi.visual_neighbors_host(cams, start, other, sim, min_baseline, k)
    got_by_id = {ids[i]: [ids[j] for j in lst] for i, lst in enumerate(got) if len(lst)}
    assert got_by_id == want
    if min_baseline > 1.0:                                    # the limit did something
        unlimited = capi.visual_neighbors_host(cams, start, other, sim, 0.0, k)
        assert any(len(a) != len(b) or (a != b).any() for a, b in zip(got, unlimited))


@pytest.mark.parametrize("maker,kw,params", [
    (scenes.tiny, dict(), dict(neighbors=5)),
    (scenes.tiny, dict(n_views=10, n_seg=200, seed=3), dict(neighbors=6, diffusion=True)),
    (scenes.urban, dict(n_views=16, n_seg=150, seed=31, blocks=(2, 2)), dict(neighbors=6)),
    (scenes.wireframe_building, dict(n_views=12, n_seg=200, seed=2), dict(neighbors=8)),
    (scenes.wireframe_building, dict(n_views=24, n_seg=400, seed=4), dict(neighbors=12, threads=8)),      # clusters of up to 87 hypotheses, 11 segments on a line
])
def test_line_fit_host_build_equals_the_oracle_bit_for_bit(tmp_path, maker, kw, params):
    """Every 3-D line the oracle emits, recomputed from its cluster (lines_seg2d), the hypotheses' end points
    (best_match) and the scene normalisation (transform): same number of segments, every FP64 coordinate identical."""
    scn = str(tmp_path / "s.scn")
    scenes.write_scene(maker(**kw), scn)
    d = str(tmp_path / "dump"); os.makedirs(d)
    n_lines, _, _, dm = oracle_py.run_scene(scn, d, **params)
    assert n_lines > 0
    hyp = {(int(r[0]), int(r[1])): r[6:12] for r in dm["best_match"]}
    tr = dm["transform"].ravel()
    back = capi.similarity_back(tr[0], tr[1:10], tr[10:13])
    start, pts, cams = [0], [], []
    for li in range(n_lines):
        rows = dm["lines_seg2d"][dm["lines_seg2d"][:, 0] == li]
        for _, cam, seg in rows:
            pts.append(hyp[(int(cam), int(seg))]); cams.append(int(cam))
        start.append(len(cams))
    got = capi.fit_lines_host(back, start, np.array(pts), cams)
    total = 0
    for li in range(n_lines):
        want = dm["lines_seg3d"][dm["lines_seg3d"][:, 0] == li][:, 1:7]
        assert got[li].shape == want.shape and np.array_equal(got[li].view(np.uint64), np.ascontiguousarray(want).view(np.uint64)), li
        total += want.shape[0]
    assert total == dm["lines_seg3d"].shape[0] > 0


def test_line_fit_degenerate_clusters():
    """No cluster at all; a cluster that fewer than three cameras see; hypotheses that are all the same point (zero
    scatter: the direction falls back to the first axis); three cameras seeing one stretch give exactly one segment."""
    back = capi.similarity_back(1.0, np.eye(3), np.zeros(3))
    assert capi.fit_lines_host(back, [0], np.zeros((0, 6)), []) == []
    two_cams = np.array([[0, 0, 0, 1, 0, 0], [0, 0.1, 0, 1, 0.1, 0]], float)
    assert capi.fit_lines_host(back, [0, 2], two_cams, [0, 1])[0].shape == (0, 6)
    same = np.tile(np.array([[1.0, 2.0, 3.0, 1.0, 2.0, 3.0]]), (4, 1))
    got = capi.fit_lines_host(back, [0, 4], same, [0, 1, 2, 3])[0]
    assert got.shape[0] <= 4 and np.isfinite(got).all()
    three = np.array([[0, 0, 0, 4, 0, 0], [1, 0.01, 0, 3, 0.01, 0], [0.5, -0.01, 0, 3.5, -0.01, 0]], float)
    seg = capi.fit_lines_host(back, [0, 3], three, [0, 1, 2])[0]
    assert seg.shape == (1, 6)
    # the stretch where all three are open: from x = 1 (the last to open) to the first to close, whichever end the sweep starts from
    xs = sorted([seg[0, 0], seg[0, 3]])
    assert xs == [1.0, 3.0]
    # scaling and rotating the scene back changes the coordinates, not the structure
    R = np.array([[0.0, -1.0, 0.0], [1.0, 0.0, 0.0], [0.0, 0.0, 1.0]])
    seg2 = capi.fit_lines_host(capi.similarity_back(0.5, R, [1.0, 0.0, 0.0]), [0, 3], three, [0, 1, 2])[0]
    assert seg2.shape == (1, 6)
    assert np.allclose(np.linalg.norm(seg2[0, :3] - seg2[0, 3:]), 2.0 * np.linalg.norm(seg[0, :3] - seg[0, 3:]))


# ---- an independent restatement in plain Python floats (IEEE doubles, no contraction), on random inputs -------------
def _mul33(a, b):
    return [[sum_from_zero(a[i][k] * b[k][j] for k in range(3)) for j in range(3)] for i in range(3)]


def sum_from_zero(terms):
    s = 0.0
    for t in terms:
        s = s + t
    return s


def _inv33(m):
    c00 = m[1][1] * m[2][2] - m[1][2] * m[2][1]
    c01 = m[1][2] * m[2][0] - m[1][0] * m[2][2]
    c02 = m[1][0] * m[2][1] - m[1][1] * m[2][0]
    det = m[0][0] * c00 + m[0][1] * c01 + m[0][2] * c02
    i = 1.0 / det
    return [[c00 * i, (m[0][2] * m[2][1] - m[0][1] * m[2][2]) * i, (m[0][1] * m[1][2] - m[0][2] * m[1][1]) * i],
            [c01 * i, (m[0][0] * m[2][2] - m[0][2] * m[2][0]) * i, (m[0][2] * m[1][0] - m[0][0] * m[1][2]) * i],
            [c02 * i, (m[0][1] * m[2][0] - m[0][0] * m[2][1]) * i, (m[0][0] * m[1][1] - m[0][1] * m[1][0]) * i]]


def _fundamental_py(c1, c2):
    K1, R1, t1 = [c1[0:9].reshape(3, 3).tolist(), c1[9:18].reshape(3, 3).tolist(), c1[18:21].tolist()]
    K2, R2, t2 = [c2[0:9].reshape(3, 3).tolist(), c2[9:18].reshape(3, 3).tolist(), c2[18:21].tolist()]
    T_ = lambda m: [[m[j][i] for j in range(3)] for i in range(3)]
    R = _mul33(R2, T_(R1))
    t = [t2[i] - sum_from_zero(R[i][k] * t1[k] for k in range(3)) for i in range(3)]
    T = [[0.0, -t[2], t[1]], [t[2], 0.0, -t[0]], [-t[1], t[0], 0.0]]
    return _mul33(_mul33(_inv33(T_(K2)), _mul33(T, R)), _inv33(K1))


def _random_cameras(n, rng, spread):
    cams = np.zeros((n, capi.CAMERA_DOUBLES))
    for v in range(n):
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        f = rng.uniform(500.0, 3000.0)
        K = np.array([[f, rng.normal(0, 0.5), rng.uniform(300, 2000)], [0.0, f * rng.uniform(0.9, 1.1), rng.uniform(200, 1500)], [0.0, 0.0, 1.0]])
        C = rng.uniform(-spread, spread, 3)
        cams[v, 0:9] = K.ravel(); cams[v, 9:18] = q.ravel(); cams[v, 18:21] = -q @ C; cams[v, 21:24] = C
    return cams


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_fundamentals_host_build_equals_a_plain_python_restatement(seed):
    rng = np.random.default_rng(seed)
    cams = _random_cameras(12, rng, 25.0)
    pairs = [(a, b) for a in range(12) for b in range(12) if a != b]
    F = capi.fundamentals_host(cams, pairs)
    for (a, b), f in zip(pairs, F):
        want = np.array(_fundamental_py(cams[a], cams[b]))
        assert np.array_equal(f.view(np.uint64), want.view(np.uint64)), (a, b)


def _select_py(cams, v, other, sim, min_baseline, k):
    def base(a, b):
        s = 0.0
        for i in range(3):
            d = float(cams[a, 21 + i]) - float(cams[b, 21 + i])
            s = s + d * d
        return np.float32(np.sqrt(s))
    mb = np.float32(min_baseline)
    kept = []
    for o, s in zip(other, sim):
        if not base(v, o) > mb:
            continue
        if all(base(e, o) > mb for e, _ in kept):
            kept.append((o, np.float32(s)))
    kept.sort(key=lambda e: -e[1])               # python's sort is stable
    if k > 0:
        kept = kept[:k]
    return sorted(o for o, _ in kept)


@pytest.mark.parametrize("n,k,baseline,seed", [(10, 3, 5.0, 1), (40, 8, 9.0, 2), (40, -1, 0.0, 3), (25, 5, 30.0, 4)])
def test_neighbour_selection_host_build_equals_a_plain_python_restatement(n, k, baseline, seed):
    rng = np.random.default_rng(seed)
    cams = _random_cameras(n, rng, 20.0)
    start, other, sim = [0], [], []
    for v in range(n):
        o = np.sort(rng.choice([x for x in range(n) if x != v], size=int(rng.integers(0, n)), replace=False))
        other += o.tolist(); sim += (rng.integers(1, 12, size=o.shape[0]).astype(np.float32) / np.float32(12.0)).tolist(); start.append(len(other))
    got = capi.visual_neighbors_host(cams, start, other, sim, baseline, k)
    for v in range(n):
        assert got[v].tolist() == _select_py(cams, v, other[start[v]: start[v + 1]], sim[start[v]: start[v + 1]], baseline, k), v


def test_bad_arguments_are_refused():
    cams = np.zeros((2, capi.CAMERA_DOUBLES))
    with pytest.raises(capi.L3DError):
        capi.visual_neighbors_host(cams, [0, 1, 2], [0, 0], [0.5, 0.5], 0.1, 3)       # a view listed as its own neighbour
    with pytest.raises(capi.L3DError):
        capi.fundamentals_host(cams, [[0, 2]])
    with pytest.raises(capi.L3DError):                      # members not in camera order
        capi.fit_lines_host(capi.similarity_back(1.0, np.eye(3), np.zeros(3)), [0, 2], np.zeros((2, 6)), [3, 1])
