"""The conservative pre-tests in front of the exact pair test (k_pretest) must never change a result: with them switched
off every pair of a computed view pair goes to the exact test, as in the reference (cudawrapper.cu:538-611).  Checked
where their error allowance is hardest to meet -- coordinates around 8 kpx and segments of a few pixels, for which the
reference's own FP32 line through the endpoints is off by more than the 0.25 px allowance -- and on degenerate epipolar
geometry, where every pair survives them."""
import numpy as np
import pytest

import l3d_testutil as util
from line3d_b200 import scenes

pytestmark = pytest.mark.gpu


def _large_short_scene(path):
    """An urban scene blown up to 7680 x 5120 px; a third of the segments cut down to 1.5 .. 12 px."""
    sc = scenes.urban(n_views=10, n_seg=900, seed=31, blocks=(2, 1))
    rng = np.random.default_rng(4)
    for v in sc.views:
        v.K = np.diag([4.0, 4.0, 1.0]) @ v.K
        v.width *= 4; v.height *= 4
        s = v.segs.astype(np.float64) * 4.0
        p1, p2 = s[:, :2], s[:, 2:]
        d = p2 - p1
        ln = np.maximum(np.hypot(d[:, 0], d[:, 1]), 1e-9)
        u = d / ln[:, None]
        cut = rng.random(s.shape[0]) < 0.34
        mid = p1 + rng.uniform(0.1, 0.9, s.shape[0])[:, None] * d
        L = rng.uniform(1.5, 12.0, s.shape[0])
        s[cut, :2] = (mid - 0.5 * L[:, None] * u)[cut]
        s[cut, 2:] = (mid + 0.5 * L[:, None] * u)[cut]
        v.segs = s.astype(np.float32)
    scenes.write_scene(sc, path)
    return sc


def _rows(got):
    rows = np.concatenate([util.filtered_as_rows(v, got[v]) for v in sorted(got)])
    return rows[np.lexsort((rows[:, 3], rows[:, 2], rows[:, 1], rows[:, 0]))]


def test_short_segments_at_8kpx_with_and_without_pretest(tmp_path):
    from line3d_b200 import capi
    p = str(tmp_path / "big_short.scn")
    sc = _large_short_scene(p)
    sig = dict(sigma_p=14.0, sigma_a=10.0)                  # 4 x the default: the image is 4 x larger
    dm = util.oracle_dump(p, neighbors=6, collinearity=False, threads=8, unc_upper=20.0, unc_lower=4.0, **sig)
    ids, views, F, vn = util.tables_from_dump(dm, sc)
    ctx = capi.Context(0)
    normal = util.run_matching(ctx, ids, views, F, vn, plan=True, **sig)
    ctx.debug_set_pretest(1)
    off = util.run_matching(ctx, ids, views, F, vn, plan=True, **sig)
    ctx.debug_set_pretest(0)
    want = dm["filtered_matches"]; want = want[np.lexsort((want[:, 3], want[:, 2], want[:, 1], want[:, 0]))]
    assert want.shape[0] > 300 and int(dm["_counters"]["raw"]) > 20000
    assert np.array_equal(_rows(off), want)                 # the exact tier alone equals the oracle ...
    assert np.array_equal(_rows(normal), want)              # ... and the pre-tests drop nothing it accepts
    tp = sum(r["pair_tests"] for r in normal.values())
    assert sum(r["n_candidates"] for r in off.values()) == tp
    assert sum(r["n_candidates"] for r in normal.values()) < 0.25 * tp
    assert sum(r["n_raw"] for r in normal.values()) == sum(r["n_raw"] for r in off.values()) == int(dm["_counters"]["raw"])
    for v in normal:
        assert normal[v]["median_depth"] == off[v]["median_depth"] and normal[v]["n_scored"] == off[v]["n_scored"]
    ctx.close()


def _rand_view(i, n):
    rng = np.random.default_rng(100 + i)
    segs = rng.uniform(10, 1900, size=(n, 4)).astype(np.float32)
    return dict(id=i, segs=segs, RtKinv=np.eye(3), C=np.array([float(i), 0.0, 0.0]), P=np.hstack([np.eye(3), np.zeros((3, 1))]),
                spatial_k=0.01, k_lower=0.001, k_upper=0.005)


def test_degenerate_geometry_keeps_every_pair_and_overflow_is_reported():
    """A fundamental matrix of NaNs: no float test can decide anything, every pair must reach the exact test (which
    rejects them all).  More candidates than a batch can index are an error with a message, not a wrap-around."""
    from line3d_b200 import capi
    ctx = capi.Context(0)
    ctx.set_views([_rand_view(i, 1500) for i in range(4)])
    F = np.full((3, 3), np.nan)
    res = ctx.match_view(0, [dict(view=1, F=F, to_be_matched=True, deliver=False), dict(view=2, F=F, to_be_matched=True, deliver=False)])
    assert res["pair_tests"] == 2 * 1500 * 1500 and res["n_candidates"] == res["pair_tests"] and res["n_raw"] == 0 and res["n_filtered"] == 0
    ctx.reset_matching()
    ctx.debug_set_candidate_limit(1_000_000)
    with pytest.raises(capi.L3DError, match="candidates"):
        ctx.match_view(0, [dict(view=1, F=F, to_be_matched=True, deliver=False)])
    ctx.debug_set_candidate_limit(0)
    ctx.reset_matching()
    res = ctx.match_view(0, [dict(view=1, F=F, to_be_matched=True, deliver=False)])      # the context survives
    assert res["n_candidates"] == 1500 * 1500
    ctx.close()


def test_a_batch_with_too_many_candidates_is_cut_smaller(tiny_scene):
    from line3d_b200 import capi
    sc = scenes.read_scene(tiny_scene)
    dm = util.oracle_dump(tiny_scene, neighbors=5, collinearity=False)
    ids, views, F, vn = util.tables_from_dump(dm, sc)
    ctx = capi.Context(0)
    plain = util.run_matching(ctx, ids, views, F, vn, plan=True)
    per_view = max(r["n_candidates"] for r in plain.values())
    total = sum(r["n_candidates"] for r in plain.values())
    assert total > 2 * per_view
    ctx.debug_set_candidate_limit(per_view + 1)               # the whole schedule does not fit in one batch, a single view does
    cut = util.run_matching(ctx, ids, views, F, vn, plan=True)
    ctx.debug_set_candidate_limit(0)
    assert np.array_equal(_rows(cut), _rows(plain))
    ctx.close()
