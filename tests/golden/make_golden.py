"""Regenerates the golden fixtures in this directory from the REFERENCE ITSELF: the
reference's own sources compiled for the CPU (oracle/_ref/libline3d_ref.so, see
oracle/Makefile).  Needs /root/reference (the build container); the fixtures are committed
so that the oracle can be pinned where the reference is absent.

    python tests/golden/make_golden.py
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from line3d_b200 import scenes  # noqa: E402
from oracle import refrun  # noqa: E402

CASES = {
    "k4": dict(neighbors=4),
    "k4_diffusion": dict(neighbors=4, diffusion=True),
    "kall_nocollin": dict(neighbors=-1, collinearity=False),
}


def golden_scene():
    return scenes.tiny(n_views=7, n_seg=70, seed=21, step_deg=10.0)


def ragged_scene():
    """Views of very different sizes: one without any segment, one with a single segment, one with a few."""
    sc = scenes.urban(n_views=12, n_seg=700, seed=5, blocks=(2, 1))
    for k, n in ((3, 0), (5, 1), (7, 37)):
        sc.views[k].segs = sc.views[k].segs[:n]
    return sc


def main():
    assert refrun.available(), "build oracle/_ref first (make -C oracle ref)"
    scn2 = os.path.join(HERE, "scene_g2.scn")
    scenes.write_scene(ragged_scene(), scn2)
    n, tm, dm = refrun.run_scene(scn2, tempfile.mkdtemp(), neighbors=6)
    dm["n_lines"] = np.array([n])
    np.savez_compressed(os.path.join(HERE, "ref_g2_k6.npz"), **dm)
    print("g2 (ragged)", "lines", n, {k: v.shape for k, v in dm.items() if k in ("filtered_matches", "edges_cluster", "lines_seg3d")})
    scn = os.path.join(HERE, "scene_g1.scn")
    scenes.write_scene(golden_scene(), scn)
    for tag, params in CASES.items():
        d = tempfile.mkdtemp()
        n, tm, dm = refrun.run_scene(scn, d, **params)
        dm["n_lines"] = np.array([n])
        np.savez_compressed(os.path.join(HERE, f"ref_g1_{tag}.npz"), **dm)
        print(tag, "lines", n, {k: v.shape for k, v in dm.items() if k in ("filtered_matches", "edges_cluster", "lines_seg3d")})
    # operator-level vectors
    sc = scenes.read_scene(scn)
    rel = refrun.collinearity(sc.views[0].segs, 2.0)
    rng = np.random.default_rng(5)
    n_nodes = 9
    und = [(i, j) for i in range(n_nodes) for j in range(i + 1, n_nodes) if rng.random() < 0.45]
    for i in range(n_nodes):                      # every node needs at least one entry
        if not any(i in e for e in und):
            und.append((i, (i + 1) % n_nodes) if i < (i + 1) % n_nodes else ((i + 1) % n_nodes, i))
    ij, w = [], []
    for (i, j) in und:
        ww = np.float32(rng.uniform(0.05, 1.0))
        ij += [(i, j), (j, i)]; w += [ww, ww]
    ij = np.array(ij, np.int32); w = np.array(w, np.float32)
    out = refrun.diffusion(ij, w, n_nodes)
    np.savez_compressed(os.path.join(HERE, "ref_ops.npz"), collin_segs=sc.views[0].segs, collin_rel=rel, diff_ij=ij, diff_w=w, diff_n=np.array([n_nodes]), diff_out=out)
    print("ops: collinearity nnz", int((rel > 0).sum()), "diffusion nnz", out.shape[0])


if __name__ == "__main__":
    main()
