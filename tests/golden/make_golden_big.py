"""Golden digests at BENCHMARK size, from the REFERENCE ITSELF (oracle/_ref/libline3d_ref.so: the reference's own
sources compiled for the CPU, see oracle/Makefile).  One run per case with all host cores; minutes to an hour each.

    python tests/golden/make_golden_big.py [case ...]          # default: all cases

Cases (scenes are regenerated from their seeds; the digest stores the scene file's sha256 so that a test can tell a
different generator from a different result):

    c2        BASELINE.json configs[1]: urban, 200 views x ~2000 segments, k = 12, collinearity on (what bench.py runs)
    c3        the same scene with diffusion (configs[2])
    c4slice   configs[3]'s shape: 60 views x ~4000 segments, k = 20
    c5slice   configs[4]'s shape: 40 views x ~3000 segments, every co-visible view a neighbour
    manycams  170 views x ~260 segments, every co-visible view a neighbour: more than 150 cameras per view

Each case writes tests/golden/big_<case>.npz (tests/l3d_testutil.digest_of_dumps) with the reference's wall times.
"""
import hashlib
import os
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import l3d_testutil as util  # noqa: E402
from line3d_b200 import scenes  # noqa: E402
from oracle import refrun  # noqa: E402

CASES = {
    "c2": (lambda: scenes.make("C2"), dict(neighbors=12)),
    "c3": (lambda: scenes.make("C3"), dict(neighbors=12, diffusion=True)),
    "c4slice": (lambda: scenes.urban(n_views=60, n_seg=4000, seed=1004, blocks=(2, 1)), dict(neighbors=20)),
    "c5slice": (lambda: scenes.dense_dome(n_views=40, n_seg=3000, seed=1005), dict(neighbors=-1)),
    "manycams": (lambda: scenes.dense_dome(n_views=170, n_seg=260, seed=1015), dict(neighbors=-1)),
}


def scene_path(case: str) -> str:
    """The case's scene, regenerated from its seed into the temp dir."""
    p = os.path.join(tempfile.gettempdir(), f"l3d_big_{case}.scn")
    if not os.path.exists(p):
        scenes.write_scene(CASES[case][0](), p + ".tmp")
        os.replace(p + ".tmp", p)
    return p


def file_sha(path: str) -> np.ndarray:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 24), b""):
            h.update(chunk)
    return np.frombuffer(h.digest(), np.uint8).copy()


def main():
    assert refrun.available(), "build oracle/_ref first (make -C oracle ref)"
    threads = int(os.environ.get("L3D_GOLDEN_THREADS", os.cpu_count() or 1))
    for case in (sys.argv[1:] or list(CASES)):
        p = scene_path(case)
        params = CASES[case][1]
        d = tempfile.mkdtemp(prefix=f"l3d_big_{case}_")
        t0 = time.time()
        n, tm, dm = refrun.run_scene(p, d, threads=threads, **params)
        wall = time.time() - t0
        dg = util.digest_of_dumps(dm, n)
        dg["scene_sha"] = file_sha(p)
        dg["ref_seconds"] = np.array([tm["matching"], tm["compute3Dmodel"], wall]); dg["ref_threads"] = np.array([threads])
        np.savez_compressed(os.path.join(HERE, f"big_{case}.npz"), **dg)
        print(case, "lines", n, "filtered", dm["filtered_matches"].shape[0], "edges", dm["edges_cluster"].shape[0],
              "matching %.1f s, compute3Dmodel %.1f s, %d threads" % (tm["matching"], tm["compute3Dmodel"], threads), flush=True)
        for f in os.listdir(d):
            os.remove(os.path.join(d, f))
        os.rmdir(d)


if __name__ == "__main__":
    main()
