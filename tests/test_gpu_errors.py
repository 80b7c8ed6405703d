"""Error behaviour of the C ABI on the GPU box: every call returns an error code and a message, nothing is done on
bad input, and the context stays usable (include/l3d_b200.h; the reference prints and carries on)."""
import numpy as np
import pytest

import l3d_testutil as util

pytestmark = pytest.mark.gpu


def _view(i, n):
    rng = np.random.default_rng(i)
    segs = rng.uniform(10, 900, size=(n, 4)).astype(np.float32)
    return dict(id=i, segs=segs, RtKinv=np.eye(3), C=np.array([float(i), 0.0, 0.0]), P=np.hstack([np.eye(3), np.zeros((3, 1))]),
                spatial_k=0.01, k_lower=0.001, k_upper=0.005)


def test_bad_arguments_are_refused_and_the_context_survives(tiny_scene):
    from line3d_b200 import capi, scenes
    ctx = capi.Context(0)
    # more segments in one view than the 16-bit ids of the candidate lists hold
    with pytest.raises(capi.L3DError, match="65535"):
        ctx.set_views([_view(0, 70000), _view(1, 10), _view(2, 10), _view(3, 10)])
    views = [_view(i, 40) for i in range(4)]
    ctx.set_views(views)
    F = np.eye(3)
    nb = lambda v, tbm=True: dict(view=v, F=F, to_be_matched=tbm, deliver=False)
    with pytest.raises(capi.L3DError, match="out of range"):
        ctx.match_view(7, [nb(1)])
    with pytest.raises(capi.L3DError, match="neighbour"):
        ctx.match_view(0, [nb(0)])                      # a view cannot be its own neighbour
    with pytest.raises(capi.L3DError, match="ascending"):
        ctx.match_view(0, [nb(2), nb(1)])
    with pytest.raises(capi.L3DError):
        ctx.match_view(0, [])                           # no neighbours at all
    # ... and the context still works: the tiny scene through it equals the oracle
    sc = scenes.read_scene(tiny_scene)
    dm = util.oracle_dump(tiny_scene, neighbors=5, collinearity=False)
    ids, vws, Fs, vn = util.tables_from_dump(dm, sc)
    got = util.run_matching(ctx, ids, vws, Fs, vn, plan=True)
    rows = np.concatenate([util.filtered_as_rows(v, got[v]) for v in sorted(got)])
    key = lambda a: np.lexsort((a[:, 3], a[:, 2], a[:, 1], a[:, 0]))
    assert np.array_equal(rows[key(rows)], dm["filtered_matches"][key(dm["filtered_matches"])])
    ctx.close()


def test_views_without_matches(tiny_scene):
    """Random segments under identity geometry: (almost) nothing matches; empty results come back as such."""
    from line3d_b200 import capi
    ctx = capi.Context(0)
    ctx.set_views([_view(i, 50) for i in range(4)])
    F = np.array([[0.0, -1e-3, 0.2], [1e-3, 0.0, -0.3], [-0.2, 0.3, 0.0]])
    res = ctx.match_view(0, [dict(view=1, F=F, to_be_matched=True, deliver=True), dict(view=2, F=F, to_be_matched=True, deliver=False)])
    assert res["pair_tests"] == 2 * 50 * 50 and res["n_filtered"] == res["filtered"].shape[0] and res["n_scored"] == res["n_raw"]
    res1 = ctx.match_view(1, [dict(view=0, F=F.T, to_be_matched=False, deliver=False), dict(view=3, F=F, to_be_matched=True, deliver=False)])
    assert res1["pair_tests"] == 50 * 50 and res1["n_delivered"] <= res["n_filtered"]
    ctx.close()
