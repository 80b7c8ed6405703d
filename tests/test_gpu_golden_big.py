"""The GPU product at BENCHMARK size against digests of the reference's own output (tests/golden/big_*.npz, made by
tests/golden/make_golden_big.py from oracle/_ref: the reference's sources compiled for the CPU).  Every per-view record
set, median, best hypothesis, potential correspondence, affinity edge (in order) and final line must be identical;
the 3-D coordinates of the lines to 1e-9.  This is the parity evidence for the very workload bench.py times (C2), for
C3 (C2 with diffusion), for slices of the shapes of C4 and C5 and for a view with more than 150 cameras."""
import importlib.util
import os
import tempfile

import numpy as np
import pytest

import l3d_testutil as util

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_spec = importlib.util.spec_from_file_location("make_golden_big", os.path.join(GOLD, "make_golden_big.py"))
CASES = ["c2", "c3", "c4slice", "c5slice", "manycams"]


def _cases_module():
    import sys
    saved = list(sys.argv)
    m = importlib.util.module_from_spec(_spec)
    _spec.loader.exec_module(m)
    sys.argv = saved
    return m


@pytest.mark.parametrize("case", CASES)
def test_product_matches_reference_digest(case):
    from line3d_b200 import host
    path = os.path.join(GOLD, f"big_{case}.npz")
    if not os.path.exists(path):
        pytest.skip(f"{path} has not been generated (tests/golden/make_golden_big.py {case})")
    g = np.load(path)
    mk = _cases_module()
    scn = mk.scene_path(case)
    assert np.array_equal(mk.file_sha(scn), g["scene_sha"]), "the scene generator no longer produces the scene the digest was made from"
    params = mk.CASES[case][1]
    n, tm, st, dm = host.run_scene(scn, tempfile.mkdtemp(), **params)
    d = util.digest_of_dumps(dm, n)
    assert n == int(g["n_lines"][0])
    assert np.array_equal(d["match_log"], g["match_log"])
    assert np.array_equal(d["match_median"], g["match_median"])
    bad = np.nonzero((d["filtered_sha"] != g["filtered_sha"]).any(axis=1))[0]
    assert bad.size == 0, f"filtered records differ in {bad.size} views, first: view {int(g['match_log'][bad[0], 0])}"
    for name in ("visual_neighbors", "collinearity", "best_match", "potential", "edges_pre_diffusion", "edges_cluster", "lines_seg2d"):
        assert np.array_equal(d[name + "_shape"], g[name + "_shape"]), name
        assert np.array_equal(d[name + "_sha"], g[name + "_sha"]), name
    assert int(d["cluster_nodes"][0]) == int(g["cluster_nodes"][0])
    assert d["lines_seg3d"].shape == g["lines_seg3d"].shape
    np.testing.assert_allclose(d["lines_seg3d"], g["lines_seg3d"], rtol=1e-9, atol=1e-9)
