#!/usr/bin/env python
"""bench.py -- matched+scored segment pairs per second on synthetic wireframe scenes.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C2]

Metric (BASELINE.json): pair tests (one (source segment, target segment) evaluation of a
computed view pair) divided by the time of raw matching + scoring + filtering over all views.
A "step" is one pass of that hot path over the whole scene.

  value   device-resident: the scene's segment and camera tables are already in HBM, the
          timed region is the view loop of l3d_match_view_resident calls (host clock around
          calls that end in a stream synchronise; the CUDA-event device time of the same
          calls is reported beside it as device_ms_per_step).
  e2e     the same metric through the unchanged Line3D API with HOST buffers: upload of all
          tables, the view loop with its per-view device->host copy of the filtered matches.
  roofline  the dominant kernel (scoring) against the measured FP32 FFMA peak of the device,
          algorithmic flops per SURVEY.md section 8(d).
  cpu_baseline / --impl reference: the reference's own sources built for the CPU
          (oracle/_ref) -- or the oracle port if that build is absent -- on a bounded sample.

Multi-GPU (N > 1, one process per GPU under torchrun): the neighbour graph is partitioned
into connected components, every rank matches and scores the views of its components
(scaling = weak: one C2-sized district per GPU), results are reduced over NCCL.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from line3d_b200 import scenes  # noqa: E402

METRIC = "matched+scored segment pairs/s"
UNIT = "pairs/s"
# dram__bytes_read.sum + dram__bytes_write.sum of the k_pretest launch of this workload (C2), from the ncu --set full
# capture summarised in profiles/ (0.20 GB read: hoisted segments; 0.69 GB written: one candidate bit per pair test)
PRETEST_TRAFFIC = 0.896e9

# bounded CPU samples of the named workloads: the first views of the scene, all neighbours of each other
CPU_SAMPLE = {"C1": dict(views=24, neighbors=12), "C2": dict(views=5, neighbors=4), "C3": dict(views=5, neighbors=4),
              "C4": dict(views=4, neighbors=3), "C5": dict(views=4, neighbors=-1)}


def scene_file(config: str, rank: int = 0) -> str:
    """Seeded scene for a config; cached under the temp dir so both arms see the same bytes."""
    c = scenes.CONFIGS[config]
    kw = dict(c["kw"])
    if rank:
        kw["seed"] = kw["seed"] + 100 * rank      # one district per rank, same shape
    path = os.path.join(tempfile.gettempdir(), f"l3d_scene_{config}_{kw['seed']}_{kw.get('n_views')}x{kw.get('n_seg')}.scn")
    if not os.path.exists(path):
        sc = c["fn"](**kw)
        tmp = path + f".{os.getpid()}.tmp"
        scenes.write_scene(sc, tmp)
        os.replace(tmp, path)
    return path


def multi_scene_file(config: str, n: int) -> str:
    """n districts of the config in one scene (one connected component each): the job that is sharded over n GPUs."""
    path = os.path.join(tempfile.gettempdir(), f"l3d_scene_{config}_x{n}.scn")
    if not os.path.exists(path):
        tmp = path + f".{os.getpid()}.tmp"
        scenes.write_scene(scenes.multi_district(config, n), tmp)
        os.replace(tmp, path)
    return path


def sample_scene_file(config: str) -> tuple[str, dict]:
    s = CPU_SAMPLE[config]
    full = scenes.read_scene(scene_file(config))
    sub = scenes.Scene(views=full.views[: s["views"]], name=f"{config}-sample")
    path = os.path.join(tempfile.gettempdir(), f"l3d_scene_{config}_sample{s['views']}.scn")
    if not os.path.exists(path):
        scenes.write_scene(sub, path)
    n = [v.segs.shape[0] for v in sub.views]
    desc = dict(views=len(sub.views), segments_per_view=int(np.mean(n)), neighbors=s["neighbors"])
    return path, desc


class ClockSampler:
    """nvidia-smi SM clock / throttle reasons sampled during the timed region."""

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for k, nm in enumerate(names):
                if len(r) > 4 + k and r[4 + k].lower().startswith("active"):
                    reasons.add(nm)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=float(np.max(mx)) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def run_cpu(config: str, threads: int, prefer_ref: bool = True) -> dict:
    """The reference's CPU code for the path on a bounded sample of the workload."""
    path, desc = sample_scene_file(config)
    params = dict(neighbors=desc["neighbors"], collinearity=False, diffusion=False)
    from oracle import oracle_py, refrun
    # pair-test count of the sample comes from the oracle's counters (identical in both builds)
    if prefer_ref and refrun.available():
        t0 = time.perf_counter()
        n, tm, _ = refrun.run_scene(path, None, threads=threads, **params)
        wall = time.perf_counter() - t0
        _, _, ct, _ = oracle_py.run_scene(path, None, threads=threads, **params)
        kind, secs = "reference", tm["matching"]
    else:
        t0 = time.perf_counter()
        n, tm, ct, _ = oracle_py.run_scene(path, None, threads=threads, **params)
        wall = time.perf_counter() - t0
        kind, secs = "port", tm["matching"]
    return dict(value=ct["pair_tests"] / max(secs, 1e-9), unit=UNIT, cores=threads, kind=kind,
                sample=f"first {desc['views']} views of {config} (~{desc['segments_per_view']} segments/view, k={desc['neighbors']}): "
                       f"{int(ct['pair_tests'])} pair tests, {int(ct['score_iters'])} scoring iterations, {secs:.2f} s in matching+scoring",
                pair_tests=int(ct["pair_tests"]), seconds=secs, wall_seconds=wall)


def reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    per_step = []
    cb = None
    for _ in range(args.warmup):
        run_cpu(args.config, threads)
    for _ in range(args.steps):
        cb = run_cpu(args.config, threads)
        per_step.append(cb["seconds"])
    ms = 1e3 * float(np.mean(per_step))
    value = cb["pair_tests"] / (ms * 1e-3)
    cb = dict(value=value, unit=UNIT, cores=cb["cores"], kind=cb["kind"], sample=cb["sample"])
    c = scenes.CONFIGS[args.config]
    out = dict(impl="reference", metric=METRIC, value=value, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
               ms_per_step=ms, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
               config=dict(workload=f"{args.config}: {c['kw']['n_views']} views x ~{c['kw']['n_seg']} segments, k={c['k']}", sample=cb["sample"]),
               cpu_baseline=cb, e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(out), flush=True)


def ours(args) -> None:
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl")
    from line3d_b200 import capi, host

    c = scenes.CONFIGS[args.config]
    if world > 1:
        if rank == 0:
            multi_scene_file(args.config, world)      # written once, read by every rank
        dist.barrier()
        path = multi_scene_file(args.config, world)
    else:
        path = scene_file(args.config)
    params = dict(neighbors=c["k"], collinearity=True, diffusion=c["diffusion"])      # the reference's defaults (commons.h:49, 61)
    sess = host.Session(path, device=local, **params)
    sess.set_shard(rank, world)                       # this rank's connected components of the neighbour graph

    def barrier():
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()

    # ---- device-resident steps ---------------------------------------------------------------
    for _ in range(args.warmup):
        sess.compute(measure_only=True, flush_l2=True)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    stats = []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _, _, st = sess.compute(measure_only=True, flush_l2=True)
        stats.append(st)
    barrier()
    loop_wall = time.perf_counter() - t0
    clocks = sampler.stop()
    t_step = float(np.mean([s["t_matching"] for s in stats]))
    t_dev = float(np.mean([s["t_matching_device"] for s in stats]))
    pair_tests = stats[-1]["pair_tests"]
    launches = int(sum(s["kernel_launches"] for s in stats))
    # untimed extra pass with CUDA events between the stages and the pair tests in one batch with nothing running beside
    # them: the per-stage breakdown and the kernels' own durations.  The timed steps run without those events and with
    # the pair tests of batch i + 1 beside the per-view work of batch i.
    sess.compute(measure_only=True, flush_l2=True, stage_timing=True)      # (grows the single-batch buffers once)
    _, _, st_prof = sess.compute(measure_only=True, flush_l2=True, stage_timing=True)

    # ---- end to end through the Line3D API, host buffers ------------------------------------------
    e2e_stats = []
    sess.compute(measure_only=False, flush_l2=True)      # untimed: grows the pinned result buffers once
    for _ in range(max(1, min(args.steps, 3))):
        _, tm, st = sess.compute(measure_only=False, flush_l2=True)
        e2e_stats.append((tm, st))
    t_e2e = float(np.mean([s["t_upload"] + s["t_matching"] for _, s in e2e_stats]))
    st_e = e2e_stats[-1][1]

    # ---- reduce over ranks ---------------------------------------------------------------------------
    tot_pairs, t_max, t_dev_max, t_e2e_max = pair_tests, t_step, t_dev, t_e2e
    if dist is not None:
        import torch
        v = torch.tensor([pair_tests], dtype=torch.float64, device="cuda"); dist.all_reduce(v); tot_pairs = float(v.item())
        m = torch.tensor([t_step, t_dev, t_e2e], dtype=torch.float64, device="cuda"); dist.all_reduce(m, op=dist.ReduceOp.MAX)
        t_max, t_dev_max, t_e2e_max = [float(x) for x in m.tolist()]
        L = torch.tensor([launches], dtype=torch.float64, device="cuda"); dist.all_reduce(L); launches = int(L.item())
    if rank != 0:
        sess.close()
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ------------------------------------------------------------
    ctx = capi.Context(local)
    fp32_peak = ctx.fp32_peak_tflops()
    ctx.close()
    s = stats[-1]
    stage = {k: float(st_prof[k]) for k in ("t_pretest", "t_exact", "t_assemble", "t_score", "t_reduce")}
    # kernel times alone, from the profiling pass (one batch, nothing beside it); in the timed steps the pair tests of a
    # batch run beside the per-view kernels of the batch before, where they take longer (reported as *_overlapped)
    t_kp = float(st_prof["t_k_pretest"]); t_ke = float(st_prof["t_k_exact"])
    t_kp_ov = float(np.mean([x["t_k_pretest"] for x in stats])); t_ke_ov = float(np.mean([x["t_k_exact"] for x in stats]))
    # The largest single kernel is k_pretest (one launch per step on this workload): it is arithmetic-bound (HBM traffic
    # under 1 % of peak in profiles/), so its roof is the FP32 rate.  Algorithmic work per pair test:
    # 8 FFMA for the four residuals + 8 FADD + 1 FMUL for the sign test = 17 FMA-pipe instructions = 25 flops;
    # the interval test of the ~2 % survivors is left out (lower bound).  The peak is an all-FFMA figure, so this
    # instruction mix cannot exceed 25 / (2 * 17) = 74 % of it: issue_frac is the same measurement in instructions.
    PRETEST_FLOPS, PRETEST_INSTR = 25.0, 17.0
    achieved = PRETEST_FLOPS * s["pair_tests"] / max(t_kp, 1e-12) / 1e12
    # reference arithmetic per unit: matching 104 + 420 a flops per pair test with a = raw / pair tests;
    # scoring 30 flops per inner iteration on another camera (lower bounds)
    a = s["raw_matches"] / max(s["pair_tests"], 1.0)
    flops = {"t_score": 30.0 * s["score_iters_other_cam"], "t_pretest": (104.0 + 420.0 * a) * s["pair_tests"]}
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    roofline = dict(bound="fp32", kernel="k_pretest", achieved=achieved, peak=fp32_peak,
                    unit="TFLOP/s", frac=achieved / max(fp32_peak, 1e-9), traffic=PRETEST_TRAFFIC,
                    peak_source="FFMA-chain microkernel run in this process (l3d_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 figure",
                    algorithmic_flops_per_launch=PRETEST_FLOPS * s["pair_tests"], launches_per_step=1, kernel_ms_per_launch=1e3 * t_kp,
                    issue_frac=(PRETEST_INSTR * s["pair_tests"] / max(t_kp, 1e-12)) / max(0.5e12 * fp32_peak, 1e-9),
                    exact_kernel_ms=1e3 * t_ke, kernel_ms_overlapped=dict(k_pretest=1e3 * t_kp_ov, k_exact_flat=1e3 * t_ke_ov), stages_ms={k[2:]: 1e3 * v for k, v in stage.items()},
                    stages_note="device ms per stage, and the kernels' own times, from one extra pass with CUDA events between the stages and "
                                f"no overlap of pair tests and per-view work; that pass took {1e3 * st_prof['t_matching_device']:.1f} ms on the device",
                    reference_arithmetic=dict(note="SURVEY.md 8(d) flops of the reference's own code for the same units, over this library's device time",
                                              pair_test_path_tflops=flops["t_pretest"] / max(stage["t_pretest"] + stage["t_exact"], 1e-12) / 1e12,
                                              scoring_tflops=flops["t_score"] / max(stage["t_score"], 1e-12) / 1e12),
                    hbm_peak_gbs=peaks.get("hbm_gbs"))
    cb = run_cpu(args.config, os.cpu_count() or 1) if world == 1 else None
    if cb:
        cb = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}

    value = tot_pairs / t_max
    out = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup, ms_per_step=1e3 * t_max,
               higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
               config=dict(workload=f"{args.config}: {c['kw']['n_views']} views x ~{c['kw']['n_seg']} segments, k={c['k']}, no diffusion"
                                    + (f"; {world} such districts in one scene, sharded by connected component over {world} GPUs" if world > 1 else ""),
                           pair_tests_per_step=int(tot_pairs), candidates_per_step=int(s["candidates"]), raw_matches_per_step=int(s["raw_matches"]),
                           scoring_iterations_per_step=int(s["score_iters"]), l2="flushed between steps (256 MiB memset); the per-step working set "
                           "(match lists of all views) is far larger than L2 anyway"),
               device_ms_per_step=1e3 * t_dev_max, loop_wall_s=loop_wall, clocks=clocks, gpu_launches=launches,
               e2e=dict(value=tot_pairs / t_e2e_max, unit=UNIT, h2d_bytes_per_step=int(st_e["h2d_bytes"]), d2h_bytes_per_step=int(st_e["d2h_bytes"]),
                        ms_per_step=1e3 * t_e2e_max, inside_calls_ms=1e3 * float(st_e["t_upload"] + st_e["t_matching_calls"]),
                        timed="wall clock of the upload plus the whole matching loop of Line3D::compute3Dmodel, the host's bookkeeping between the calls included "
                              "(the library works ahead of the host, so the time inside the calls alone, inside_calls_ms, would flatter it)",
                        api="Line3D::compute3Dmodel -> l3d_set_views + l3d_match_view per view (host buffers)",
                        compute3Dmodel_ms=1e3 * float(np.mean([tm["compute3Dmodel"] for tm, _ in e2e_stats])),
                        host_stages_ms={k[2:]: 1e3 * st_e[k] for k in ("t_neighbors", "t_upload", "t_matching", "t_hypotheses", "t_affinity", "t_diffusion", "t_clustering", "t_lines")}),
               roofline=roofline)
    if cb:
        out["cpu_baseline"] = cb
    print(json.dumps(out), flush=True)
    sess.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=sorted(scenes.CONFIGS))
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours(args)


if __name__ == "__main__":
    main()
