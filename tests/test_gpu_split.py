"""Several GPUs on ONE connected scene (l3d_comm_init / Line3D::setSplit): every view's segments are dealt out to the
ranks, which exchange over NCCL what later views need.  Two ranks, one process per GPU under torchrun; the split result
must equal the oracle's single-process result record for record.  Needs two GPUs (skipped otherwise)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, pickle
root, scn, out = sys.argv[1:4]
sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "tests"))
import numpy as np
import torch, torch.distributed as dist
import l3d_testutil as util
from line3d_b200 import capi, host, scenes
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo")
box = [capi.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
params = dict(neighbors=6, collinearity=True)
sc = scenes.read_scene(scn)
dm = util.oracle_dump(scn, threads=4, **params)
res = dict(rank=rank)
# ---- C ABI: every rank drives the same schedule; all ranks receive all records (result_root = -1)
ids, views, F, vn = util.tables_from_dump(dm, sc)
ctx = capi.Context(local)
ctx.comm_init(rank, world, box[0], -1)
for plan in (True, False):
    got = util.run_matching(ctx, ids, views, F, vn, plan=plan)
    rows = np.concatenate([util.filtered_as_rows(v, got[v]) for v in sorted(got)])
    key = lambda a: np.lexsort((a[:, 3], a[:, 2], a[:, 1], a[:, 0]))
    want = dm["filtered_matches"]
    res["abi_rows_equal_%d" % plan] = bool(rows.shape == want.shape and np.array_equal(rows[key(rows)], want[key(want)]))
    res["abi_medians_equal_%d" % plan] = bool(all(got[int(r[0])]["median_depth"] == float(m) for r, m in zip(dm["match_log"], dm["match_median"].ravel())))
    res["abi_pair_tests_%d" % plan] = int(sum(r["pair_tests"] for r in got.values()))
    res["abi_iters_%d" % plan] = int(sum(r["score_iters_other_cam"] for r in got.values()))      # (the oracle counts the iterations on another camera)
res["oracle_pair_tests"] = int(dm["_counters"]["pair_tests"]); res["oracle_iters"] = int(dm["_counters"]["score_iters"])
ctx.close()
# ---- the Line3D API: rank 0 gathers the records and clusters them
box = [capi.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
p2 = dict(params); p2["collinearity"] = params["collinearity"] and rank == 0      # only rank 0 clusters
s = host.Session(scn, device=local, **p2)
s.set_split(rank, world, box[0])
for rep in range(2):
    n, tm, st = s.compute(keep_debug=True)
res["n_lines"] = n; res["oracle_lines"] = int(dm["_n_lines"]); res["pair_tests"] = int(st["pair_tests"])
if rank == 0:
    d = s.dump()
    key = lambda a: np.lexsort((a[:, 3], a[:, 2], a[:, 1], a[:, 0]))
    res["filtered_equal"] = bool(d["filtered_matches"].shape == dm["filtered_matches"].shape and np.array_equal(d["filtered_matches"][key(d["filtered_matches"])], dm["filtered_matches"][key(dm["filtered_matches"])]))
    for k in ("match_median", "best_match", "potential", "edges_cluster", "lines_seg2d"):
        res[k + "_equal"] = bool(d[k].shape == dm[k].shape and np.array_equal(d[k], dm[k]))
    res["lines3d_close"] = bool(d["lines_seg3d"].shape == dm["lines_seg3d"].shape and np.allclose(d["lines_seg3d"], dm["lines_seg3d"], rtol=1e-9, atol=1e-9))
s.close()
pickle.dump(res, open(out + ".r%d" % rank, "wb"))
dist.barrier()
dist.destroy_process_group()
'''


def test_two_ranks_split_one_connected_scene(tmp_path):
    import pickle
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    sys.path.insert(0, ROOT)
    from line3d_b200 import scenes
    scn = str(tmp_path / "one.scn")
    scenes.write_scene(scenes.urban(n_views=14, n_seg=700, seed=91, blocks=(2, 1)), scn)      # one street: one connected component
    worker = str(tmp_path / "worker.py")
    open(worker, "w").write(WORKER)
    out = str(tmp_path / "out.pkl")
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                           "--master-port", "29631", worker, ROOT, scn, out], timeout=900)
    r0 = pickle.load(open(out + ".r0", "rb")); r1 = pickle.load(open(out + ".r1", "rb"))
    for r in (r0, r1):
        for plan in (0, 1):
            assert r["abi_rows_equal_%d" % plan] and r["abi_medians_equal_%d" % plan], (r["rank"], plan)
    for plan in (0, 1):                                                # the ranks' shares add up to the whole
        assert r0["abi_iters_%d" % plan] + r1["abi_iters_%d" % plan] == r0["oracle_iters"]
        assert r0["abi_pair_tests_%d" % plan] + r1["abi_pair_tests_%d" % plan] == r0["oracle_pair_tests"]
        assert abs(r0["abi_pair_tests_%d" % plan] - r1["abi_pair_tests_%d" % plan]) < 0.02 * r0["oracle_pair_tests"]
    assert r0["pair_tests"] + r1["pair_tests"] == r0["oracle_pair_tests"]
    assert r0["n_lines"] == r0["oracle_lines"] > 0 and r1["n_lines"] == 0
    for k in ("filtered_equal", "match_median_equal", "best_match_equal", "potential_equal", "edges_cluster_equal", "lines_seg2d_equal", "lines3d_close"):
        assert r0[k], k
