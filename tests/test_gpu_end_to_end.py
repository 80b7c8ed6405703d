"""GPU parity, whole pipeline: Line3D(...) / addImage / compute3Dmodel / getResult through
the C++ host scaffold and the CUDA library, against the CPU oracle on the same scenes.
North-star bar: identical match sets, hypothesis coordinates / scores / affinities within
1e-4 relative, identical final clustered 3-D lines."""
import tempfile

import numpy as np
import pytest

import l3d_testutil as util

pytestmark = pytest.mark.gpu

EXACT = ["visual_neighbors", "views_static", "fundamentals", "collinearity", "match_log", "match_median", "potential",
         "cluster_nodes", "lines_seg2d", "transform"]


def _run(scene, **params):
    from line3d_b200 import host
    d = tempfile.mkdtemp(prefix="l3d_gpu_")
    n, tm, st, dm = host.run_scene(scene, d, **params)
    return n, st, dm


def _check(dm_o, n_o, dm_g, n_g):
    # host-side tables and integer structures: identical
    for k in ("visual_neighbors", "fundamentals", "collinearity", "match_log", "match_median", "potential", "cluster_nodes", "transform"):
        assert dm_o[k].shape == dm_g[k].shape, k
        assert np.array_equal(dm_o[k], dm_g[k]), k
    assert np.array_equal(dm_o["views"], dm_g["views"])
    # match sets, depths and scores: bit-exact
    fo, fg = dm_o["filtered_matches"], dm_g["filtered_matches"]
    assert fo.shape == fg.shape
    key = lambda a: np.lexsort((a[:, 3], a[:, 2], a[:, 1], a[:, 0]))
    fo, fg = fo[key(fo)], fg[key(fg)]
    assert np.array_equal(fo[:, :8], fg[:, :8])
    assert np.array_equal(fo[:, 8], fg[:, 8])
    # hypotheses: same segments, coordinates within 1e-4 relative (they are in fact ~1e-15)
    bo, bg = dm_o["best_match"], dm_g["best_match"]
    assert bo.shape == bg.shape
    assert np.array_equal(bo[:, :4], bg[:, :4])
    assert np.array_equal(bo[:, 4], bg[:, 4])
    np.testing.assert_allclose(bg[:, 6:15], bo[:, 6:15], rtol=1e-9, atol=1e-12)
    assert np.array_equal(bo[:, 15:], bg[:, 15:])
    # affinities feeding the clustering: same edges in the same order, weights within 1e-4 relative
    for k in ("edges_pre_diffusion", "edges_cluster"):
        eo, eg = dm_o[k], dm_g[k]
        assert eo.shape == eg.shape, k
        assert np.array_equal(eo[:, :2], eg[:, :2]), k
        np.testing.assert_allclose(eg[:, 2], eo[:, 2], rtol=1e-4, err_msg=k)
    # final lines identical
    assert n_o == n_g
    assert np.array_equal(dm_o["lines_seg2d"], dm_g["lines_seg2d"])
    assert dm_o["lines_seg3d"].shape == dm_g["lines_seg3d"].shape
    np.testing.assert_allclose(dm_g["lines_seg3d"], dm_o["lines_seg3d"], rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("params", [dict(neighbors=5), dict(neighbors=5, diffusion=True), dict(neighbors=-1),
                                    dict(neighbors=4, collinearity=False)])
def test_tiny_pipeline(tiny_scene, params):
    dm_o = util.oracle_dump(tiny_scene, **params)
    n_g, st, dm_g = _run(tiny_scene, **params)
    _check(dm_o, dm_o["_n_lines"], dm_g, n_g)
    assert st["pair_tests"] == dm_o["_counters"]["pair_tests"]
    assert st["raw_matches"] == dm_o["_counters"]["raw"]
    assert st["kernel_launches"] > 0


def test_small_pipeline(small_scene):
    params = dict(neighbors=6)
    dm_o = util.oracle_dump(small_scene, threads=8, **params)
    n_g, st, dm_g = _run(small_scene, **params)
    _check(dm_o, dm_o["_n_lines"], dm_g, n_g)
