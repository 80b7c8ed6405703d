"""Analytic known-answer tests of the oracle (the reference has no tests of its own,
SURVEY.md section 4): geometry whose answers are known in closed form."""
import numpy as np
import pytest

from oracle import oracle_py

K = np.array([[1600.0, 0, 960.0], [0, 1600.0, 640.0], [0, 0, 1.0]])


def _cam(C, yaw):
    c, s = np.cos(yaw), np.sin(yaw)
    R = np.array([[c, 0, -s], [0, 1, 0], [s, 0, c]])   # world -> camera, y down, z forward
    t = -R @ np.asarray(C, float)
    return R, t


def _tables(R1, t1, R2, t2):
    Kinv = np.linalg.inv(K)
    R = R2 @ R1.T; t = t2 - R @ t1
    T = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    F = np.linalg.inv(K.T) @ (T @ R) @ Kinv                      # line3D.cc:1982-1992
    return F, R1.T @ Kinv, R2.T @ Kinv, -R1.T @ t1, -R2.T @ t2


def _project(R, t, X):
    p = K @ (R @ X + t)
    return p[:2] / p[2]


def test_depths_equal_true_camera_distances_and_mirror_on_swap():
    R1, t1 = _cam((-1.0, 0.0, 0.0), 0.15); R2, t2 = _cam((1.2, 0.1, 0.0), -0.2)
    X1, X2 = np.array([-0.3, -0.4, 6.0]), np.array([0.5, 0.35, 7.0])
    s = np.concatenate([_project(R1, t1, X1), _project(R1, t1, X2)]); q = np.concatenate([_project(R2, t2, X1), _project(R2, t2, X2)])
    F, A1, A2, C1, C2 = _tables(R1, t1, R2, t2)
    acc, d, ov = oracle_py.pair_test(s, q, F, A1, A2, C1, C2)
    assert acc
    true = [np.linalg.norm(X1 - C1), np.linalg.norm(X2 - C1), np.linalg.norm(X1 - C2), np.linalg.norm(X2 - C2)]
    np.testing.assert_allclose(d, true, rtol=2e-4)
    np.testing.assert_allclose(ov, [1.0, 1.0], atol=2e-3)           # the same 3-D segment seen in both: full overlap
    # swapped roles use F^T and give the mirrored tuple (line3D.cc:847-856 relies on this)
    acc2, d2, _ = oracle_py.pair_test(q, s, F.T, A2, A1, C2, C1)
    assert acc2
    assert np.array_equal(d2, d[[2, 3, 0, 1]])


def test_partial_and_no_overlap():
    R1, t1 = _cam((-1.0, 0.0, 0.0), 0.1); R2, t2 = _cam((1.0, 0.0, 0.0), -0.1)
    F, A1, A2, C1, C2 = _tables(R1, t1, R2, t2)
    Xa, Xb, Xc, Xd = [np.array([0.0, -1.0 + 0.5 * k, 6.0]) for k in range(4)]   # collinear points on a vertical 3-D line
    s = np.concatenate([_project(R1, t1, Xa), _project(R1, t1, Xc)])            # src sees a..c
    q_half = np.concatenate([_project(R2, t2, Xb), _project(R2, t2, Xd)])       # tgt sees b..d: overlap b..c of a..d = 1/3
    acc, d, ov = oracle_py.pair_test(s, q_half, F, A1, A2, C1, C2)
    np.testing.assert_allclose(ov, [1 / 3, 1 / 3], atol=5e-3)
    assert acc and np.all(d > 0)
    Xe, Xf = np.array([0.0, 1.2, 6.0]), np.array([0.0, 1.9, 6.0])               # disjoint from a..c
    q_none = np.concatenate([_project(R2, t2, Xe), _project(R2, t2, Xf)])
    acc, d, ov = oracle_py.pair_test(s, q_none, F, A1, A2, C1, C2)
    assert not acc and np.all(d == 0) and np.all(ov == 0)


def test_short_segments_are_rejected():
    R1, t1 = _cam((-1.0, 0.0, 0.0), 0.0); R2, t2 = _cam((1.0, 0.0, 0.0), 0.0)
    F, A1, A2, C1, C2 = _tables(R1, t1, R2, t2)
    s = np.array([900.0, 600.0, 900.3, 600.4]); q = np.array([700.0, 600.0, 700.3, 600.4])    # < 1 px long (cudawrapper.cu:212)
    acc, d, ov = oracle_py.pair_test(s, q, F, A1, A2, C1, C2)
    assert not acc and np.all(d == 0)


def test_score_counts_supporting_cameras():
    """A hypothesis supported by n other cameras at zero residual scores n."""
    cams = [_cam((x, 0.05 * i, 0.0), -0.04 * x) for i, x in enumerate((-1.5, -0.5, 0.6, 1.4, 2.1))]
    X1, X2 = np.array([-0.2, -0.5, 7.0]), np.array([0.4, 0.6, 7.5])
    Kinv = np.linalg.inv(K)
    (Rs, ts) = cams[0]
    src = np.concatenate([_project(Rs, ts, X1), _project(Rs, ts, X2)])[None, :]
    Cs = -Rs.T @ ts
    tgt_segs, P, data, depths, offs = [], [], [], [], []
    for c, (R, t) in enumerate(cams[1:]):
        tgt_segs.append(np.concatenate([_project(R, t, X1), _project(R, t, X2)]))
        P.append((K @ np.hstack([R, t[:, None]])).ravel()); offs.append(c)
        Ct = -R.T @ t
        data.append([0, c, 0, 0]); depths.append([np.linalg.norm(X1 - Cs), np.linalg.norm(X2 - Cs), np.linalg.norm(X1 - Ct), np.linalg.norm(X2 - Ct)])
    score = oracle_py.score_matches(np.array(data), np.array(depths), src, Rs.T @ Kinv, Cs, np.array(tgt_segs), np.array(offs), np.array(P), 3.5, 10.0, 0.004)
    # each of the 4 matches is supported by the 3 other cameras with confidence ~1
    np.testing.assert_allclose(score, 3.0, atol=2e-3)


def test_collinearity_known_cases():
    segs = np.array([[100, 100, 200, 100], [260, 100, 400, 100.5],   # collinear, disjoint -> kept
                     [150, 100, 300, 100],                           # collinear with 0 but overlapping -> rejected (cudawrapper.cu:522-528)
                     [100, 300, 200, 340]], np.float32)              # unrelated
    rel = oracle_py.collinearity(segs, 2.0)
    assert rel[0, 1] > 0.9 and rel[1, 0] == rel[0, 1]
    assert rel[0, 2] == 0 and rel[0, 3] == 0 and rel[1, 3] == 0
    assert np.all(np.diag(rel) == 0)


def test_diffusion_two_node_graph_by_hand():
    # W = [[0, w], [w, 0]]: P row-normalised = [[0,1],[1,0]]; every step keeps P' = max(eps, P_ij * (P_row_j . W_col_i)) = w,
    # then normalises back to 1 except after the last of the 10 iterations (cudawrapper.cu:1157-1184)
    w = np.float32(0.37)
    out = oracle_py.diffusion(np.array([[0, 1], [1, 0]]), np.array([w, w]), 2)
    assert out[:, :2].tolist() == [[0, 1], [1, 0]]
    np.testing.assert_allclose(out[:, 2], [w, w], rtol=1e-6)


def test_diffusion_lockstep_quirk_three_nodes():
    """cudawrapper.cu:786-800 multiplies the k-th stored entry of P's row with the k-th stored entry of W's
    column whatever their indices.  Path graph 0-1-2: row 1 of P = (1,0),(1,2), column 0 of W = (1,0) only."""
    ij = np.array([[0, 1], [1, 0], [1, 2], [2, 1]]); w = np.array([0.8, 0.8, 0.4, 0.4], np.float32)
    out = oracle_py.diffusion(ij, w, 3)
    assert out[:, :2].tolist() == [[0, 1], [1, 0], [1, 2], [2, 1]]
    # one iteration by hand, float32 -----------------------------------------------------------
    f = np.float32
    P = {(0, 1): f(1.0), (1, 0): f(0.8) / (f(0.8) + f(0.4)), (1, 2): f(0.4) / (f(0.8) + f(0.4)), (2, 1): f(1.0)}
    W = {(0, 1): f(0.8), (1, 0): f(0.8), (1, 2): f(0.4), (2, 1): f(0.4)}
    rows = {0: [(0, 1)], 1: [(1, 0), (1, 2)], 2: [(2, 1)]}; cols = {0: [(1, 0)], 1: [(0, 1), (2, 1)], 2: [(1, 2)]}
    for it in range(10):
        Pn = {}
        for (c, r), val in P.items():               # entry y = (row c, col r): target (r, c)
            mul = f(0)
            for a, b in zip(rows[r], cols[c]):      # lockstep, stops at the shorter list
                mul = f(mul + f(P[a] * W[b]))
            mul = f(mul * val)
            Pn[(r, c)] = max(mul, f(1e-12))
        P = Pn
        if it < 9:
            for r in rows:
                s = f(0)
                for e in rows[r]:
                    s = f(s + P[e])
                for e in rows[r]:
                    P[e] = f(P[e] / s)
    np.testing.assert_array_equal(out[:, 2], np.array([P[(0, 1)], P[(1, 0)], P[(1, 2)], P[(2, 1)]], np.float32))
