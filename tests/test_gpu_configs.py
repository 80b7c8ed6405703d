"""GPU parity on the shapes of BASELINE.json's configs: C1 at full size against the oracle,
the dense-neighbour (k = -1) and urban shapes at reduced size, and size-independent
properties at a larger size where the oracle would take too long: run-to-run determinism and
equality of the batched (l3d_plan_matching) and the view-by-view execution."""
import tempfile

import numpy as np
import pytest

import l3d_testutil as util

pytestmark = pytest.mark.gpu


def _pipeline_equal(scene_path, threads=16, **params):
    from line3d_b200 import host
    dm_o = util.oracle_dump(scene_path, threads=threads, **params)
    n_g, tm, st, dm_g = host.run_scene(scene_path, tempfile.mkdtemp(), **params)
    assert n_g == dm_o["_n_lines"]
    for k in ("visual_neighbors", "match_log", "match_median", "best_match", "potential", "edges_cluster", "lines_seg2d"):
        assert dm_o[k].shape == dm_g[k].shape, k
        assert np.array_equal(dm_o[k], dm_g[k]), k
    # a view with nothing left to pair-test returns its delivered records unscored: in delivery order in the reference
    # (cudawrapper.cu:877-878), sorted by (src, cam, tgt) here -- compare as sets
    fo, fg = dm_o["filtered_matches"], dm_g["filtered_matches"]
    assert fo.shape == fg.shape
    key = lambda a: np.lexsort((a[:, 3], a[:, 2], a[:, 1], a[:, 0]))
    assert np.array_equal(fo[key(fo)], fg[key(fg)])
    np.testing.assert_allclose(dm_g["lines_seg3d"], dm_o["lines_seg3d"], rtol=1e-9, atol=1e-9)
    return n_g, st


def test_c1_full_size(tmp_path):
    """BASELINE.json configs[0]: 24 views, ~800 segments/view, k = 12 -- whole pipeline identical to the oracle."""
    from line3d_b200 import scenes
    p = str(tmp_path / "c1.scn")
    scenes.write_scene(scenes.make("C1"), p)
    n, st = _pipeline_equal(p, neighbors=12)
    assert n > 50 and st["pair_tests"] > 5e7


def test_dense_neighbours_reduced(tmp_path):
    """configs[4] shape (all co-visible views are neighbours, k = -1) at 14 views x ~250 segments."""
    from line3d_b200 import scenes
    p = str(tmp_path / "dome.scn")
    scenes.write_scene(scenes.dense_dome(n_views=14, n_seg=250, seed=1005), p)
    _pipeline_equal(p, neighbors=-1)


def test_urban_reduced_with_diffusion(tmp_path):
    """configs[1]/[2] shape at 16 views x ~500 segments, with the diffusion step."""
    from line3d_b200 import scenes
    p = str(tmp_path / "urban.scn")
    scenes.write_scene(scenes.urban(n_views=16, n_seg=500, seed=1002, blocks=(1, 1)), p)
    _pipeline_equal(p, neighbors=6, diffusion=True)


def test_ragged_and_empty_views(tmp_path):
    """A view without segments, one with a single segment, one with 37 among views of 700: the whole pipeline against
    the oracle (which equals the reference build on the same scene, tests/test_oracle_vs_ref.py)."""
    from test_oracle_vs_ref import ragged_scene
    p = str(tmp_path / "ragged.scn")
    ragged_scene(p)
    n, st = _pipeline_equal(p, neighbors=6)
    assert n > 0


def test_batched_equals_view_by_view_and_is_deterministic(tmp_path):
    """Larger urban slice (no oracle): the scaffold announces the schedule (batched pair tests), the plain C-ABI loop
    does not; both must give the same filtered matches, and a second run the same again."""
    from line3d_b200 import capi, host, scenes
    p = str(tmp_path / "u.scn")
    sc = scenes.urban(n_views=24, n_seg=1500, seed=1002, blocks=(2, 1))
    scenes.write_scene(sc, p)
    s = host.Session(p, neighbors=8, collinearity=False)
    n1, _, st1 = s.compute(keep_debug=True); d1 = s.dump()
    n2, _, st2 = s.compute(keep_debug=True); d2 = s.dump()
    s.close()
    assert n1 == n2 and np.array_equal(d1["filtered_matches"], d2["filtered_matches"]) and np.array_equal(d1["lines_seg3d"], d2["lines_seg3d"])
    ids, views, F, vn = util.tables_from_dump(d1, sc)
    ctx = capi.Context(0)
    got = util.run_matching(ctx, ids, views, F, vn)
    ctx.close()
    rows = np.concatenate([util.filtered_as_rows(v, got[v]) for v in sorted(got)])
    assert np.array_equal(rows, d1["filtered_matches"])
    assert [got[v]["median_depth"] for v in sorted(got)] == [float(x) for x in d1["match_median"].ravel()]
    assert sum(r["pair_tests"] for r in got.values()) == st1["pair_tests"]
    # a raw match (v,i)->(n,j) delivered to n must show up there with mirrored depths if it survives: every filtered
    # record of a view whose camera was processed earlier as a mutual neighbour stems from that view's filtered list
    f = d1["filtered_matches"]
    keys = {(int(r[0]), int(r[1]), int(r[2]), int(r[3])): r[4:8] for r in f}
    checked = 0
    for (v, i, n, j), dep in list(keys.items())[:20000]:
        if n < v and (n, j, v, i) in keys:
            assert np.array_equal(keys[(n, j, v, i)][[2, 3, 0, 1]], dep)
            checked += 1
    assert checked > 0


def test_plan_run_ahead_and_retained_results(tmp_path):
    """C ABI only: with an announced schedule the library scores one view ahead of the caller, and with
    l3d_retain_results every view's records stay readable after the loop; both must not change a single record.
    A call that leaves the schedule while a view is running ahead is refused."""
    from line3d_b200 import capi, scenes
    sc = scenes.urban(n_views=16, n_seg=900, seed=77, blocks=(2, 1))
    p = str(tmp_path / "u.scn")
    scenes.write_scene(sc, p)
    dm = util.oracle_dump(p, neighbors=6, collinearity=False)
    ids, views, F, vn = util.tables_from_dump(dm, sc)
    ctx = capi.Context(0)
    plain = util.run_matching(ctx, ids, views, F, vn)
    for plan, retain in ((True, False), (True, True), (False, True)):
        got = util.run_matching(ctx, ids, views, F, vn, plan=plan, retain=retain)
        assert sorted(got) == sorted(plain)
        for v in plain:
            assert np.array_equal(got[v]["filtered"], plain[v]["filtered"]), (plan, retain, v)
            for k in ("pair_tests", "n_raw", "n_delivered", "n_scored", "n_filtered", "score_iters", "score_iters_other_cam", "median_depth", "scored"):
                assert got[v][k] == plain[v][k], (plan, retain, v, k)
    # many small batches: the pair tests of each batch run on a helper thread and a second stream while the views of the
    # batch before it are scored
    for pairs in (1, 3_000_000, 20_000_000):
        ctx.debug_set_batch_pairs(pairs)
        got = util.run_matching(ctx, ids, views, F, vn, plan=True, retain=(pairs == 1))
        for v in plain:
            assert np.array_equal(got[v]["filtered"], plain[v]["filtered"]), (pairs, v)
            assert got[v]["median_depth"] == plain[v]["median_depth"] and got[v]["n_delivered"] == plain[v]["n_delivered"], (pairs, v)
    ctx.debug_set_batch_pairs(0)
    rows = np.concatenate([util.filtered_as_rows(v, plain[v]) for v in sorted(plain)])
    key = lambda a: np.lexsort((a[:, 3], a[:, 2], a[:, 1], a[:, 0]))
    assert np.array_equal(rows[key(rows)], dm["filtered_matches"][key(dm["filtered_matches"])])
    # leaving the schedule: the second announced view is already on the GPU when the first call returns
    index = {vid: i for i, vid in enumerate(ids)}
    sched = util.schedule_of(ids, F, vn)
    ctx.set_views(views)
    ctx.plan_matching([(index[vid], lst) for vid, _, lst in sched])
    ctx.match_view(index[sched[0][0]], sched[0][2])
    with pytest.raises(capi.L3DError, match="schedule"):
        ctx.match_view(index[sched[2][0]], sched[2][2])
    ctx.reset_matching()
    again = util.run_matching(ctx, ids, views, F, vn, plan=True)
    assert all(np.array_equal(again[v]["filtered"], plain[v]["filtered"]) for v in plain)
    ctx.close()


def test_c2_full_size_properties():
    """BASELINE.json's C2 at full size (200 views x ~2000 segments, 5.9e9 pair tests; the oracle would need hours):
    properties that do not depend on the size.  Run-to-run identity of every record (the library runs views out of
    order, in groups, beside the next batch's pair tests, with atomics in several kernels), work accounting, and the
    mirror property of handed-over records: a match (v,i)->(n,j) kept by both views has mirrored depths."""
    import bench
    from line3d_b200 import host
    s = host.Session(bench.scene_file("C2"), neighbors=12, collinearity=False)
    n1, _, st1 = s.compute(keep_debug=True); d1 = s.dump()
    n2, _, st2 = s.compute(keep_debug=True); d2 = s.dump()
    s.close()
    assert n1 == n2 and n1 > 0
    for k in ("filtered_matches", "match_median", "match_log", "best_match", "edges_cluster", "lines_seg3d"):
        assert np.array_equal(d1[k], d2[k]), k
    assert st1["pair_tests"] == st2["pair_tests"] == 5875904553          # between V k N^2 / 2 and V k N^2
    assert st1["raw_matches"] == st2["raw_matches"] and st1["score_iters"] == st2["score_iters"]
    assert st1["raw_matches"] <= st1["candidates"] <= st1["pair_tests"] and st1["filtered_matches"] <= st1["scored_matches"]
    f = d1["filtered_matches"]
    assert f.shape[0] == st1["filtered_matches"]
    assert np.all(f[:, 8] > 0.5)                                       # stored confidence = score / 2 with score > 1
    fi = f[:, :4].astype(np.int64)
    assert fi[:, 0].max() < 256 and fi[:, 2].max() < 256 and fi[:, 1].max() < 65536 and fi[:, 3].max() < 65536
    key = ((fi[:, 0] * 65536 + fi[:, 1]) * 256 + fi[:, 2]) * 65536 + fi[:, 3]
    srt = np.argsort(key, kind="stable"); skey = key[srt]
    assert np.all(np.diff(skey) > 0)                                   # no record twice
    back = np.nonzero(fi[:, 2] < fi[:, 0])[0]                          # records towards an earlier view ...
    mirror = ((fi[back, 2] * 65536 + fi[back, 3]) * 256 + fi[back, 0]) * 65536 + fi[back, 1]
    pos = np.searchsorted(skey, mirror)
    hit = (pos < skey.size) & (skey[np.minimum(pos, skey.size - 1)] == mirror)
    assert hit.sum() > 1000                                            # ... that this earlier view kept as well
    a_rows = f[back[hit]]; b_rows = f[srt[pos[hit]]]
    assert np.array_equal(a_rows[:, 4:8], b_rows[:, [6, 7, 4, 5]])


@pytest.mark.parametrize("hyp", [32, 64])
def test_small_scoring_blocks(small_scene, hyp):
    """The scoring kernels keep one maximum per (camera, hypothesis) in shared memory, so a view with hundreds of
    neighbours (k = -1) gets fewer hypotheses per block: 128 up to 192 cameras, 64 up to 384, 32 up to 1600 (the first
    version failed above 320).  Forced here on a small scene, with and without the 3-D gate's window kernel: records,
    scores and medians must equal the oracle's."""
    from line3d_b200 import capi, scenes
    sc = scenes.read_scene(small_scene)
    dm = util.oracle_dump(small_scene, neighbors=-1, collinearity=False, threads=8)
    ids, views, F, vn = util.tables_from_dump(dm, sc)
    ctx = capi.Context(0)
    ctx.debug_set_score_block(hyp)
    for gate in (True, False):
        vv = [dict(v) for v in views]
        if not gate:
            for v in vv:
                v["spatial_k"] = np.float32(0.0)             # no 3-D gate: the quadratic kernel (k_score)
        got = util.run_matching(ctx, ids, vv, F, vn, plan=True)
        if gate:
            rows = np.concatenate([util.filtered_as_rows(v, got[v]) for v in sorted(got)])
            key = lambda a: np.lexsort((a[:, 3], a[:, 2], a[:, 1], a[:, 0]))
            ref = dm["filtered_matches"]
            assert ref.shape[0] > 1000 and rows.shape == ref.shape and np.array_equal(rows[key(rows)], ref[key(ref)])
            for r, m in zip(dm["match_log"], dm["match_median"].ravel()):
                assert got[int(r[0])]["median_depth"] == float(m)
        else:
            ctx.debug_set_score_block(0)
            base = util.run_matching(ctx, ids, vv, F, vn, plan=True)      # the same without the gate at 128 per block
            ctx.debug_set_score_block(hyp)
            for v in base:
                assert np.array_equal(base[v]["filtered"], got[v]["filtered"]) and base[v]["median_depth"] == got[v]["median_depth"]
    ctx.debug_set_score_block(0)
    ctx.close()
