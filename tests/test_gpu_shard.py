"""GPU: the product's per-rank shard (Line3D::setShard) against the oracle's, both ranks of a
two-rank split run one after the other on the same device."""
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_each_shard_matches_the_oracle(tmp_path):
    from line3d_b200 import host, scenes
    from oracle import oracle_py
    out = scenes.Scene()
    for d in range(2):
        sc = scenes.tiny(seed=7 + d)
        T = np.array([80.0 * d, 0.0, 0.0])
        for v in sc.views:
            out.views.append(scenes.View(v.id + d * len(sc.views), v.width, v.height, v.K, v.R, v.t - v.R @ T, v.segs, v.wps + d * 100000))
    scn = str(tmp_path / "two.scn")
    scenes.write_scene(out, scn)
    total = 0
    for rank in (0, 1):
        n_o, _, ct, dm_o = oracle_py.run_scene(scn, tempfile.mkdtemp(), neighbors=5, shard=(rank, 2))
        s = host.Session(scn, neighbors=5)
        s.set_shard(rank, 2)
        n_g, tm, st = s.compute(keep_debug=True)
        dm_g = s.dump()
        s.close()
        assert n_g == n_o and n_g > 0
        for k in ("match_log", "match_median", "filtered_matches", "best_match", "edges_cluster", "lines_seg2d"):
            assert np.array_equal(dm_o[k], dm_g[k]), k
        np.testing.assert_allclose(dm_g["lines_seg3d"], dm_o["lines_seg3d"], rtol=1e-9, atol=1e-9)
        assert st["pair_tests"] == ct["pair_tests"]
        total += n_g
    n_all, _, _, _ = oracle_py.run_scene(scn, None, neighbors=5)
    assert total == n_all
