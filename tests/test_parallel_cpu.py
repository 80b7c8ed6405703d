"""N > 1 host logic on the CPU: two gloo ranks each run their shard (connected components of the
neighbour graph) and the results are gathered on rank 0.  The compute inside each rank is the
oracle here (the product has no CPU path); the sharding rule and the gather are what is tested:
the union of the ranks' results must equal the single-process result."""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, pickle, tempfile
sys.path.insert(0, sys.argv[1])
from line3d_b200 import parallel
from oracle import oracle_py
rank, world, local, dist = parallel.init("gloo")
n, tm, ct, dm = oracle_py.run_scene(sys.argv[2], tempfile.mkdtemp(), neighbors=5, shard=(rank, world))
lines = parallel.gather_lines(parallel.lines_from_dumps(dm), dist)
tot = parallel.reduce_counters(dict(pair_tests=ct["pair_tests"], n_raw=ct["raw"], t_match=tm["matching"]), dist)
if rank == 0:
    pickle.dump(dict(lines=lines, tot=tot, views=sorted(set(int(v) for v in dm["match_log"][:, 0]))), open(sys.argv[3], "wb"))
else:
    pickle.dump(dict(views=sorted(set(int(v) for v in dm["match_log"][:, 0]))), open(sys.argv[3] + ".r1", "wb"))
dist.destroy_process_group()
'''


def _two_district_scene(path):
    sys.path.insert(0, ROOT)
    from line3d_b200 import scenes
    out = scenes.Scene()
    for d in range(2):
        sc = scenes.tiny(seed=7 + d)
        T = np.array([80.0 * d, 0.0, 0.0])
        for v in sc.views:
            out.views.append(scenes.View(v.id + d * len(sc.views), v.width, v.height, v.K, v.R, v.t - v.R @ T, v.segs, v.wps + d * 100000))
    scenes.write_scene(out, path)


def test_two_gloo_ranks_reproduce_the_single_process_result(tmp_path):
    import pickle
    from line3d_b200 import parallel
    from oracle import oracle_py
    scn = str(tmp_path / "two.scn")
    _two_district_scene(scn)
    worker = str(tmp_path / "worker.py")
    open(worker, "w").write(WORKER)
    out = str(tmp_path / "out.pkl")
    env = dict(os.environ, OMP_NUM_THREADS="1")
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                           "--master-port", "29611", worker, ROOT, scn, out], env=env, timeout=600)
    got = pickle.load(open(out, "rb")); r1 = pickle.load(open(out + ".r1", "rb"))
    n, tm, ct, dm = oracle_py.run_scene(scn, tempfile.mkdtemp(), neighbors=5)
    want = parallel.gather_lines(parallel.lines_from_dumps(dm), None)
    # every view processed by exactly one rank, whole components only
    assert sorted(got["views"] + r1["views"]) == sorted(set(int(v) for v in dm["match_log"][:, 0]))
    assert not (set(got["views"]) & set(r1["views"]))
    assert got["tot"]["pair_tests"] == ct["pair_tests"] and got["tot"]["n_raw"] == ct["raw"]
    assert len(got["lines"]) == len(want) == n
    for a, b in zip(got["lines"], want):
        assert a[0] == b[0]
        np.testing.assert_array_equal(a[1], b[1])
