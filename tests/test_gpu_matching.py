"""GPU parity: matching + scoring through the C ABI against the CPU oracle, on the same
seeded scenes.  Integer and FP32 results are compared bit for bit: match sets, depths,
scores (the device evaluates expf/acosf with glibc's algorithms) and median depths."""
import numpy as np
import pytest

from line3d_b200 import scenes
import l3d_testutil as util

pytestmark = pytest.mark.gpu


def _compare(dm, got):
    exp = dm["filtered_matches"]
    log = dm["match_log"]; med = dm["match_median"].ravel()
    assert sorted(got) == [int(r[0]) for r in log]
    for k, r in enumerate(log):
        vid = int(r[0]); res = got[vid]
        rows = util.filtered_as_rows(vid, res)
        want = exp[exp[:, 0] == vid]
        assert res["n_delivered"] == int(r[2]), (vid, res["n_delivered"], r)
        assert rows.shape == want.shape, (vid, rows.shape, want.shape)
        if not res["scored"]:
            # no neighbour left to pair-test: the reference returns the delivered records unscored, in
            # delivery order (cudawrapper.cu:877-878); the C ABI returns the same set sorted by (src, cam, tgt)
            want = want[np.lexsort((want[:, 3], want[:, 2], want[:, 1]))]
        assert np.array_equal(rows[:, :4], want[:, :4]), f"view {vid}: match set differs"
        assert np.array_equal(rows[:, 4:8], want[:, 4:8]), f"view {vid}: depths differ"
        assert np.array_equal(rows[:, 8], want[:, 8]), f"view {vid}: scores differ"
        assert res["median_depth"] == pytest.approx(float(med[k]), rel=0, abs=0), (vid, res["median_depth"], med[k])


@pytest.mark.parametrize("neighbors", [5, -1])
def test_tiny_scene_matches_oracle(tiny_scene, neighbors):
    from line3d_b200 import capi
    dm = util.oracle_dump(tiny_scene, neighbors=neighbors)
    sc = scenes.read_scene(tiny_scene)
    ids, views, F, vn = util.tables_from_dump(dm, sc)
    ctx = capi.Context(0)
    got = util.run_matching(ctx, ids, views, F, vn)
    _compare(dm, got)
    # the pre-test is a superset filter: candidates >= raw matches, and it removes most pairs
    tot_c = sum(r["n_candidates"] for r in got.values()); tot_r = sum(r["n_raw"] for r in got.values())
    tot_p = sum(r["pair_tests"] for r in got.values())
    assert tot_r <= tot_c <= tot_p
    assert tot_p == int(dm["_counters"]["pair_tests"])
    assert tot_r == int(dm["_counters"]["raw"])
    ctx.close()


def test_small_scene_matches_oracle(small_scene):
    from line3d_b200 import capi
    dm = util.oracle_dump(small_scene, neighbors=6, threads=8)
    sc = scenes.read_scene(small_scene)
    ids, views, F, vn = util.tables_from_dump(dm, sc)
    ctx = capi.Context(0)
    got = util.run_matching(ctx, ids, views, F, vn)
    _compare(dm, got)
    # run again on the same context: no state may leak between runs
    got2 = util.run_matching(ctx, ids, views, F, vn)
    _compare(dm, got2)
    ctx.close()
