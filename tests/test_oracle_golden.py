"""The oracle (oracle/oracle.cpp) against golden vectors produced by the reference's own
code (tests/golden/make_golden.py ran oracle/_ref, the CPU build of /root/reference).
Everything is compared bit for bit."""
import os
import tempfile

import numpy as np
import pytest

from oracle import oracle_py

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {"k4": dict(neighbors=4), "k4_diffusion": dict(neighbors=4, diffusion=True), "kall_nocollin": dict(neighbors=-1, collinearity=False)}


@pytest.mark.parametrize("tag", sorted(CASES))
def test_pipeline_matches_reference_golden(tag):
    gold = np.load(os.path.join(GOLD, f"ref_g1_{tag}.npz"))
    d = tempfile.mkdtemp()
    n, tm, ct, dm = oracle_py.run_scene(os.path.join(GOLD, "scene_g1.scn"), d, **CASES[tag])
    assert n == int(gold["n_lines"][0])
    for k in gold.files:
        if k == "n_lines":
            continue
        assert dm[k].shape == gold[k].shape, k
        assert np.array_equal(dm[k], gold[k]), k
    assert ct["raw"] > 0 and ct["pair_tests"] > ct["accepted"] >= ct["raw"]


def test_ragged_scene_matches_reference_golden():
    """scene_g2: views with 0, 1 and 37 segments among views of 700."""
    gold = np.load(os.path.join(GOLD, "ref_g2_k6.npz"))
    n, tm, ct, dm = oracle_py.run_scene(os.path.join(GOLD, "scene_g2.scn"), tempfile.mkdtemp(), neighbors=6)
    assert n == int(gold["n_lines"][0]) and n > 0
    for k in gold.files:
        if k == "n_lines":
            continue
        assert dm[k].shape == gold[k].shape, k
        assert np.array_equal(dm[k], gold[k]), k


def test_collinearity_matches_reference_golden():
    g = np.load(os.path.join(GOLD, "ref_ops.npz"))
    rel = oracle_py.collinearity(g["collin_segs"], 2.0)
    assert np.array_equal(rel, g["collin_rel"])
    assert (rel > 0).sum() > 0 and np.array_equal(rel, rel.T) and np.all(np.diag(rel) == 0)


def test_diffusion_matches_reference_golden():
    g = np.load(os.path.join(GOLD, "ref_ops.npz"))
    out = oracle_py.diffusion(g["diff_ij"], g["diff_w"], int(g["diff_n"][0]))
    assert np.array_equal(out, g["diff_out"])
