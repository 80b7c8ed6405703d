"""Shared helpers for the parity tests: drive the C ABI view by view from tables dumped by
the oracle (or the reference build), the way the host scaffold does (line3D.cc:698-885)."""
from __future__ import annotations

import os
import tempfile

import numpy as np


def oracle_dump(scene_path, **params):
    from oracle import oracle_py
    d = tempfile.mkdtemp(prefix="l3d_oracle_")
    n, tm, ct, dm = oracle_py.run_scene(scene_path, d, **params)
    dm["_n_lines"] = n; dm["_timings"] = tm; dm["_counters"] = ct
    return dm


def tables_from_dump(dm, scene):
    """views.npy columns: id, median, k_upper, k_lower, spatial_k, C(3), RtKinv(9), P(12), K(9), R(9), t(3)."""
    vw = dm["views"]
    ids = [int(x) for x in vw[:, 0]]
    by_id = {v.id: v for v in scene.views}
    views = []
    for r in vw:
        views.append(dict(id=int(r[0]), segs=by_id[int(r[0])].segs, k_upper=np.float32(r[2]), k_lower=np.float32(r[3]),
                          spatial_k=np.float32(r[4]), C=r[5:8], RtKinv=r[8:17], P=r[17:29]))
    F = {(int(r[0]), int(r[1])): r[2:11].reshape(3, 3) for r in dm["fundamentals"]}
    vn = {}
    for a, b in dm["visual_neighbors"]:
        vn.setdefault(int(a), []).append(int(b))
    return ids, views, F, vn


def schedule_of(ids, F, vn):
    """The reference's matched_ bookkeeping (line3D.cc:732-736, 844-845, 875-881) as a static schedule:
    [(view id, neighbour ids, neighbour dicts)] in matching order."""
    index = {vid: i for i, vid in enumerate(ids)}
    matched = {vid: set() for vid in ids}
    sched = []
    for vid in sorted(vn):
        nbs = sorted(vn[vid])
        if not nbs:
            continue
        lst = []
        for n in nbs:
            tbm = n not in matched[vid]
            deliver = (vid in vn.get(n, [])) and (vid not in matched[n])
            lst.append(dict(view=index[n], F=F[(vid, n)], to_be_matched=tbm, deliver=deliver))
        sched.append((vid, nbs, lst))
        for n in nbs:
            matched[vid].add(n)
            if vid in vn.get(n, []):
                matched[n].add(vid)
    return sched


def run_matching(ctx, ids, views, F, vn, sigma_p=3.5, sigma_a=10.0, resident=False, plan=False, retain=False):
    """The reference's view loop (line3D.cc:620-648).  plan: announce the schedule first (l3d_plan_matching);
    retain: keep every view's records in the library's memory (l3d_retain_results) and read them after the loop."""
    index = {vid: i for i, vid in enumerate(ids)}
    ctx.set_views(views)
    sched = schedule_of(ids, F, vn)
    ctx.retain_results(retain)
    if plan:
        ctx.plan_matching([(index[vid], lst) for vid, _, lst in sched])
    out = {}
    for vid, nbs, lst in sched:
        res = ctx.match_view(index[vid], lst, sigma_p, sigma_a, resident=resident, copy=not retain)
        res["neighbors"] = nbs
        res["timings"] = ctx.last_timings()
        out[vid] = res
    if retain and not resident:
        for res in out.values():
            res["filtered"] = res["filtered"].copy()      # only now: the buffers must have survived the later calls
    return out


def filtered_as_rows(vid, res):
    """l3d_match records -> the (vID, seg1, cam2 global, seg2, depths x4, conf) rows of filtered_matches.npy."""
    f = res["filtered"]
    nbs = np.asarray(res["neighbors"], dtype=np.float64)
    rows = np.zeros((f.shape[0], 9), dtype=np.float64)
    rows[:, 0] = vid; rows[:, 1] = f["src_seg"]
    rows[:, 2] = nbs[f["cam"]] if res["scored"] else f["cam"]
    rows[:, 3] = f["tgt_seg"]; rows[:, 4:8] = f["depths"]; rows[:, 8] = f["confidence"]
    return rows


# ---- digests of a whole-pipeline dump (tests/golden/make_golden_big.py, tests/test_gpu_golden_big.py) --------------
def _sha(a) -> np.ndarray:
    import hashlib
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8).copy()


def pack_filtered(rows: np.ndarray) -> np.ndarray:
    """filtered_matches.npy rows (vID, seg1, cam2, seg2, 4 depths, conf; float64) of ONE view as canonical bytes: sorted by
    (seg1, cam2, seg2) -- the reference returns the records of an unscored view in delivery order (cudawrapper.cu:877-878),
    the library in sorted order -- ids as int32, depths and confidence as the float32 they are."""
    rec = np.zeros(rows.shape[0], dtype=[("ids", "<i4", (4,)), ("val", "<f4", (5,))])
    if rows.shape[0]:
        o = np.lexsort((rows[:, 3], rows[:, 2], rows[:, 1]))
        rec["ids"] = rows[o, :4].astype(np.int32); rec["val"] = rows[o, 4:9].astype(np.float32)
    return rec


def digest_of_dumps(dm: dict, n_lines: int) -> dict:
    """Compact, order-exact digest of everything a run dumps: a few hundred KB for a 200-view scene."""
    out = dict(n_lines=np.array([n_lines], np.int64), match_log=dm["match_log"].astype(np.int32), match_median=dm["match_median"].astype(np.float32).reshape(-1),
               cluster_nodes=dm["cluster_nodes"].astype(np.int32).reshape(-1), lines_seg3d=dm["lines_seg3d"].astype(np.float64))
    f = dm["filtered_matches"]
    log = out["match_log"]
    sha = np.zeros((log.shape[0], 32), np.uint8)
    # the rows of one view are contiguous, in the order of the match log
    start = 0
    for k in range(log.shape[0]):
        n = int(log[k, 3])
        rows = f[start:start + n]
        assert n == 0 or (rows[:, 0] == log[k, 0]).all(), "filtered_matches is not grouped by view in match_log order"
        sha[k] = _sha(pack_filtered(rows)); start += n
    assert start == f.shape[0]
    out["filtered_sha"] = sha
    for name, dt in (("visual_neighbors", np.int32), ("best_match", np.float64), ("potential", np.int32), ("edges_pre_diffusion", np.float64),
                     ("edges_cluster", np.float64), ("lines_seg2d", np.int32), ("collinearity", np.float64)):
        a = dm[name].astype(dt)
        out[name + "_sha"] = _sha(a); out[name + "_shape"] = np.array(a.shape, np.int64)
    return out
