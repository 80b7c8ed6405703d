"""The oracle against the reference itself (oracle/_ref, built from /root/reference by
oracle/Makefile) on fresh seeded scenes.  Skipped where that build is absent."""
import tempfile

import numpy as np
import pytest

from line3d_b200 import scenes
from oracle import oracle_py, refrun

pytestmark = pytest.mark.skipif(not refrun.available(), reason="oracle/_ref not built (needs /root/reference)")


def _both(path, **params):
    n1, _, d1 = refrun.run_scene(path, tempfile.mkdtemp(), **params)
    n2, _, _, d2 = oracle_py.run_scene(path, tempfile.mkdtemp(), **params)
    assert n1 == n2
    assert sorted(d1) == sorted(d2)
    for k in d1:
        assert d1[k].shape == d2[k].shape, k
        assert np.array_equal(d1[k], d2[k]), k
    return n1, d1


@pytest.mark.parametrize("params", [dict(neighbors=5), dict(neighbors=5, diffusion=True), dict(neighbors=-1), dict(neighbors=3, collinearity=False),
                                    dict(neighbors=5, sigma_p=2.5, sigma_a=7.0, unc_upper=4.0, unc_lower=1.5)])
def test_tiny_scene_bit_identical(tiny_scene, params):
    n, dm = _both(tiny_scene, **params)
    assert n > 0 and dm["filtered_matches"].shape[0] > 0


def test_other_seed_bit_identical(tmp_path):
    p = str(tmp_path / "s.scn")
    scenes.write_scene(scenes.tiny(n_views=9, n_seg=150, seed=99, step_deg=9.0), p)
    _both(p, neighbors=6)


def ragged_scene(path):
    """Views of very different sizes: one without any segment, one with a single segment, one with a few."""
    sc = scenes.urban(n_views=12, n_seg=700, seed=5, blocks=(2, 1))
    for k, n in ((3, 0), (5, 1), (7, 37)):
        sc.views[k].segs = sc.views[k].segs[:n]
    scenes.write_scene(sc, path)
    return sc


def test_ragged_and_empty_views_bit_identical(tmp_path):
    p = str(tmp_path / "ragged.scn")
    ragged_scene(p)
    n, dm = _both(p, neighbors=6)
    assert n > 0 and dm["filtered_matches"].shape[0] > 0


def test_too_few_views_is_refused(tmp_path):
    # line3D.cc:347-351: fewer than 4 views -> error line, no result
    p = str(tmp_path / "s3.scn")
    sc = scenes.tiny()
    scenes.write_scene(scenes.Scene(views=sc.views[:3]), p)
    n1, _, _ = refrun.run_scene(p, None, neighbors=3)
    n2, _, _, _ = oracle_py.run_scene(p, None, neighbors=3)
    assert n1 == 0 and n2 == 0


def test_operator_level(tiny_scene):
    sc = scenes.read_scene(tiny_scene)
    assert np.array_equal(refrun.collinearity(sc.views[1].segs), oracle_py.collinearity(sc.views[1].segs))
    rng = np.random.default_rng(0)
    n = 14
    und = [(i, j) for i in range(n) for j in range(i + 1, n) if rng.random() < 0.3] + [(i, (i + 1) % n) for i in range(n - 1)]
    und = sorted(set((min(a, b), max(a, b)) for a, b in und))
    ij, w = [], []
    for (i, j) in und:
        ww = np.float32(rng.uniform(0.02, 1.0)); ij += [(i, j), (j, i)]; w += [ww, ww]
    assert np.array_equal(refrun.diffusion(np.array(ij), np.array(w), n), oracle_py.diffusion(np.array(ij), np.array(w), n))
