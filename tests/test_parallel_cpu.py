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


# ---- bench.py's N > 1 plumbing that needs no GPU ------------------------------------------------
SCENE_WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
import bench
from line3d_b200 import scenes
scenes.CONFIGS["T4"] = dict(fn=scenes.urban, kw=dict(n_views=11, n_seg=40, seed=int(sys.argv[2]), blocks=(2, 2), per_view_rng=True), k=4, diffusion=False)
dist.init_process_group("gloo")
path = bench.scene_file_by_ranks("T4", dist.get_rank(), dist.get_world_size(), dist.barrier)
print("SCENE", dist.get_rank(), path, flush=True)
dist.destroy_process_group()
'''


def test_ranks_generate_the_scene_together_and_get_the_single_process_file(tmp_path):
    """bench.py at N > 1: every rank generates a range of the views (per-view random streams), rank 0 joins the parts;
    the file must be byte-identical with the one a single process writes."""
    import hashlib
    import bench
    from line3d_b200 import scenes
    seed = 424200 + os.getpid() % 1000                 # a fresh cache name, so the ranks really generate
    worker = str(tmp_path / "scene_worker.py")
    open(worker, "w").write(SCENE_WORKER)
    env = dict(os.environ, OMP_NUM_THREADS="1", TMPDIR=str(tmp_path))
    out = subprocess.check_output([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                                   "--master-port", "29613", worker, ROOT, str(seed)], env=env, timeout=600).decode()
    paths = sorted(set(l.split()[2] for l in out.splitlines() if l.startswith("SCENE")))
    assert len(paths) == 1 and os.path.exists(paths[0])
    scenes.CONFIGS["T4"] = dict(fn=scenes.urban, kw=dict(n_views=11, n_seg=40, seed=seed, blocks=(2, 2), per_view_rng=True), k=4, diffusion=False)
    try:
        single = str(tmp_path / "single.scn")
        c = scenes.CONFIGS["T4"]
        scenes.write_scene(c["fn"](**c["kw"]), single)
    finally:
        del scenes.CONFIGS["T4"]
    sha = lambda p: hashlib.sha256(open(p, "rb").read()).hexdigest()
    assert sha(paths[0]) == sha(single)
    assert not [f for f in os.listdir(os.path.dirname(paths[0])) if ".part" in f]      # the parts are removed


def test_reference_arm_under_torchrun_prints_one_line_from_rank_0(tmp_path):
    """`bench.py --impl reference` launched like the GPU arm (torchrun, N > 1): rank 0 alone runs the CPU code and prints
    the JSON line, the other ranks exit 0 without work.  A tiny configuration stands in for C4."""
    import json
    runner = str(tmp_path / "ref_runner.py")
    open(runner, "w").write(r'''
import sys
sys.path.insert(0, sys.argv[1])
import bench
from line3d_b200 import scenes
scenes.CONFIGS["T5"] = dict(fn=scenes.urban, kw=dict(n_views=6, n_seg=60, seed=99, blocks=(2, 2)), k=3, diffusion=False)
bench.CPU_SAMPLE["T5"] = dict(views=4, neighbors=3, stride=1)
sys.argv = ["bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--config", "T5"]
bench.main()
''')
    env = dict(os.environ, OMP_NUM_THREADS="1", TMPDIR=str(tmp_path))
    out = subprocess.check_output([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                                   "--master-port", "29615", runner, ROOT], env=env, timeout=600).decode()
    lines = [l for l in out.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["value"] > 0 and d["unit"] == "pairs/s"
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["same_k"] is True
