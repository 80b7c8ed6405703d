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
  roofline  the dominant kernel (k_pretest: the conservative pair pre-test every (source, target) pair of
          a computed view pair goes through) against the FP32 FFMA peak measured in the same process.
  cpu_baseline / --impl reference: the reference's own sources built for the CPU
          (oracle/_ref) -- or the oracle port if that build is absent -- on a bounded sample.

Multi-GPU (N > 1, one process per GPU under torchrun): ONE connected scene -- by default C4, BASELINE.json's
1000-view configuration -- is split over the N GPUs (Line3D::setSplit / l3d_comm_init: every view's segments are dealt
out to the ranks, which exchange over NCCL what later views need; scaling = strong).  Rank 0 then runs the same
workload alone on its GPU, so the line carries the speed-up against one GPU on the same box and workload.
`--weak` gives the older mode instead: one C2-sized district per GPU, sharded by connected component, no data-path
collective.
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
# dram__bytes_read.sum + dram__bytes_write.sum of k_pretest per pair test, from the ncu --set full capture summarised in
# profiles/r2_kernels.md (one batch of C2: 7.345e8 pair tests).  The kernel writes one candidate bit per pair test
# (0.125 B) and reads the hoisted segments once per tile; bench.py scales the captured figure to the launch it times.
PRETEST_TRAFFIC_PER_PAIR = (23.44e6 + 35.02e6) / 7.345e8

# bounded CPU samples of the named workloads: the first views of the scene, all neighbours of each other.  C2/C3 keep
# the configuration's k = 12 (13 views) and take every second segment of each view (they are sorted by length, so the
# length distribution stays): 13 whole views are 80 s per step on 8 cores, and with every view a neighbour of every
# other they have 5.5 scoring iterations per pair test against 2.55 in the whole scene -- the thinned sample has about
# the whole scene's ratio.  C4's k = 20 would need 21 views, so its sample has k = 7 and says so.
CPU_SAMPLE = {"C1": dict(views=24, neighbors=12, stride=1), "C2": dict(views=13, neighbors=12, stride=2), "C3": dict(views=13, neighbors=12, stride=2),
              "C4": dict(views=8, neighbors=7, stride=2), "C5": dict(views=4, neighbors=-1, stride=1)}
_SAMPLE_COUNTERS: dict = {}


def full_config_reference(config: str):
    """The reference's CPU build on the WHOLE configuration, timed once when the benchmark-size golden digests were made
    (tests/golden/make_golden_big.py, in the build container): read from the digest file, not measured here."""
    path = os.path.join(ROOT, "tests", "golden", f"big_{config.lower()}.npz")
    if not os.path.exists(path):
        return None
    try:
        d = np.load(path)
        secs, thr = d["ref_seconds"], int(d["ref_threads"][0])
        out = dict(seconds_matching=float(secs[0]), seconds_compute3Dmodel=float(secs[1]), threads=thr,
                   where="build container, once, with the golden digests (tests/golden/make_golden_big.py)")
        return out
    except Exception:      # an older digest without timings
        return None


def scene_file(config: str, rank: int = 0) -> str:
    """Seeded scene for a config; cached under the temp dir so both arms see the same bytes."""
    c = scenes.CONFIGS[config]
    kw = dict(c["kw"])
    if rank:
        kw["seed"] = kw["seed"] + 100 * rank      # one district per rank, same shape
    path = os.path.join(tempfile.gettempdir(), f"l3d_scene_{config}_{kw['seed']}_{kw.get('n_views')}x{kw.get('n_seg')}.scn")
    if not os.path.exists(path):
        sc = scenes.make_parallel(config) if (rank == 0 and kw.get("per_view_rng")) else c["fn"](**kw)
        tmp = path + f".{os.getpid()}.tmp"
        scenes.write_scene(sc, tmp)
        os.replace(tmp, path)
    return path


def scene_file_by_ranks(config: str, rank: int, world: int, barrier) -> str:
    """The same file as scene_file(config), generated by all ranks together (every rank a range of the views)."""
    c = scenes.CONFIGS[config]
    kw = dict(c["kw"])
    path = os.path.join(tempfile.gettempdir(), f"l3d_scene_{config}_{kw['seed']}_{kw.get('n_views')}x{kw.get('n_seg')}.scn")
    have = os.path.exists(path)
    barrier()                                        # (nobody starts writing before everybody has looked)
    if have:
        return path
    if not kw.get("per_view_rng"):
        if rank == 0:
            scene_file(config)
        barrier()
        return path
    n = kw["n_views"]
    first, end = n * rank // world, n * (rank + 1) // world
    part = path + f".part{rank}"
    scenes.write_scene(scenes.Scene(views=c["fn"](view_range=(first, end), **kw).views), part)
    barrier()
    if rank == 0:
        views = [v for r in range(world) for v in scenes.read_scene(path + f".part{r}").views]
        scenes.write_scene(scenes.Scene(views=views, name=config), path + ".tmp")
        os.replace(path + ".tmp", path)
        for r in range(world):
            os.remove(path + f".part{r}")
    barrier()
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
    sub = scenes.Scene(views=[scenes.View(v.id, v.width, v.height, v.K, v.R, v.t, np.ascontiguousarray(v.segs[:: s["stride"]]), v.wps) for v in full.views[: s["views"]]],
                       name=f"{config}-sample")
    path = os.path.join(tempfile.gettempdir(), f"l3d_scene_{config}_sample{s['views']}s{s['stride']}.scn")
    if not os.path.exists(path):
        scenes.write_scene(sub, path)
    n = [v.segs.shape[0] for v in sub.views]
    desc = dict(views=len(sub.views), segments_per_view=int(np.mean(n)), neighbors=s["neighbors"], stride=s["stride"])
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
        if path not in _SAMPLE_COUNTERS:           # (once per sample: the counters do not change between steps)
            _SAMPLE_COUNTERS[path] = oracle_py.run_scene(path, None, threads=threads, **params)[2]
        ct = _SAMPLE_COUNTERS[path]
        kind, secs = "reference", tm["matching"]
    else:
        t0 = time.perf_counter()
        n, tm, ct, _ = oracle_py.run_scene(path, None, threads=threads, **params)
        wall = time.perf_counter() - t0
        kind, secs = "port", tm["matching"]
    k_full = scenes.CONFIGS[config]["k"]
    return dict(value=ct["pair_tests"] / max(secs, 1e-9), unit=UNIT, cores=threads, kind=kind, same_k=desc["neighbors"] == k_full,
                sample=f"first {desc['views']} views of {config}" + (f", every {desc['stride']}. segment" if desc["stride"] > 1 else "") + f" (~{desc['segments_per_view']} segments/view, k={desc['neighbors']}): "
                       f"{int(ct['pair_tests'])} pair tests, {int(ct['score_iters'])} scoring iterations, {secs:.2f} s in matching+scoring",
                pair_tests=int(ct["pair_tests"]), seconds=secs, wall_seconds=wall,
                # the reference's host stages on the same sample (line3D.cc, sparsematrix.cc, clustering.cc: everything of
                # compute3Dmodel that is not compute_pairwise_matches), same cores
                compute3Dmodel_seconds=float(tm["compute3Dmodel"]), host_stages_seconds=max(float(tm["compute3Dmodel"]) - float(secs), 0.0))


def default_config(args) -> str:
    """BASELINE.json: configs[1] (C2) is what the metric is quoted on at one GPU; configs[3] (C4) is the multi-GPU one."""
    if args.config:
        return args.config
    return "C2" if (args.gpus <= 1 or args.weak) else "C4"


def reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    args.config = default_config(args)
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
    cb = dict(value=value, unit=UNIT, cores=cb["cores"], kind=cb["kind"], sample=cb["sample"], same_k=cb["same_k"],
              compute3Dmodel_seconds=cb["compute3Dmodel_seconds"], host_stages_seconds=cb["host_stages_seconds"])
    full = full_config_reference(args.config)
    if full:
        cb["full_config"] = full
    c = scenes.CONFIGS[args.config]
    out = dict(impl="reference", metric=METRIC, value=value, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
               ms_per_step=ms, higher_is_better=True, scaling="weak" if (args.gpus <= 1 or args.weak) else "strong", vs_baseline=None, dtype="f32", data="synthetic",
               config=dict(workload=f"{args.config}: {c['kw']['n_views']} views x ~{c['kw']['n_seg']} segments, k={c['k']}", sample=cb["sample"],
                           note="the CPU arm runs on the host cores of one box whatever --gpus says"),
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

    def barrier():
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()

    args.config = default_config(args)
    c = scenes.CONFIGS[args.config]
    strong = world > 1 and not args.weak
    if strong:
        path = scene_file_by_ranks(args.config, rank, world, barrier)
    elif world > 1:
        if rank == 0:
            multi_scene_file(args.config, world)      # written once, read by every rank
        dist.barrier()
        path = multi_scene_file(args.config, world)
    else:
        path = scene_file(args.config)
    params = dict(neighbors=c["k"], collinearity=True, diffusion=c["diffusion"])      # the reference's defaults (commons.h:49, 61)
    if strong:
        params["collinearity"] = rank == 0            # only rank 0 clusters, so only it needs the collinearity lists
    sess = host.Session(path, device=local, **params)
    if strong:
        box = [capi.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        sess.set_split(rank, world, box[0])           # one scene on all GPUs: every view's segments dealt out to the ranks
    else:
        sess.set_shard(rank, world)                   # this rank's connected components of the neighbour graph

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
    # One GPU: the whole compute3Dmodel (matching, then hypotheses, affinities, clustering, lines on the host's side).
    # One scene on several GPUs: the matching stage with every view's records copied to rank 0's host memory, where the
    # later stages would run (they are rank 0's alone and are timed at one GPU).
    e2e_mode = 2 if strong else False
    e2e_stats = []
    sess.compute(measure_only=e2e_mode, flush_l2=True)      # untimed: grows the pinned result buffers once
    barrier()
    for _ in range(max(1, args.steps)):
        _, tm, st = sess.compute(measure_only=e2e_mode, flush_l2=True)
        e2e_stats.append((tm, st))
    barrier()
    t_e2e = float(np.mean([s["t_upload"] + s["t_matching"] for _, s in e2e_stats]))
    st_e = e2e_stats[-1][1]

    # ---- the same workload on ONE of the GPUs of this box, for the speed-up (strong scaling only) ------------------
    single = None
    if strong:
        sess.set_split(0, 1, None)                       # every rank leaves the communicator
        barrier()
        if rank == 0:
            sess.compute(measure_only=True, flush_l2=True)
            one = [sess.compute(measure_only=True, flush_l2=True)[2] for _ in range(2)]
            t1 = float(np.mean([x["t_matching"] for x in one]))
            single = dict(value=one[-1]["pair_tests"] / t1, unit=UNIT, ms_per_step=1e3 * t1, steps=2, warmup=1,
                          note="the same scene, matched by rank 0's GPU alone after the timed region")
        barrier()

    # ---- reduce over ranks ---------------------------------------------------------------------------
    tot_pairs, t_max, t_dev_max, t_e2e_max = pair_tests, t_step, t_dev, t_e2e
    if dist is not None:
        import torch
        v = torch.tensor([pair_tests], dtype=torch.float64, device="cuda"); dist.all_reduce(v); tot_pairs = float(v.item())
        m = torch.tensor([t_step, t_dev, t_e2e], dtype=torch.float64, device="cuda"); dist.all_reduce(m, op=dist.ReduceOp.MAX)
        t_max, t_dev_max, t_e2e_max = [float(x) for x in m.tolist()]
        L = torch.tensor([launches, st_e["h2d_bytes"], st_e["d2h_bytes"]], dtype=torch.float64, device="cuda"); dist.all_reduce(L)
        launches = int(L[0].item()); st_e["h2d_bytes"] = float(L[1].item()); st_e["d2h_bytes"] = float(L[2].item())
    if rank != 0:
        sess.close()
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ------------------------------------------------------------
    ctx = capi.Context(local)
    fp32_peak = ctx.fp32_peak_tflops()
    fp32x2_peak = ctx.fp32x2_peak_tflops()
    ctx.close()
    fp32_nominal = 148 * 128 * 2 * (clocks.get("sm_mhz") or 1965.0) * 1e6 / 1e12      # SMs x FP32 lanes x 2 flops x clock under load
    s = stats[-1]
    stage = {k: float(st_prof[k]) for k in ("t_pretest", "t_exact", "t_assemble", "t_score", "t_reduce")}
    # kernel times alone, from the profiling pass (one batch, nothing beside it); in the timed steps the pair tests of a
    # batch run beside the per-view kernels of the batch before, where they take longer (reported as *_overlapped)
    t_kp = float(st_prof["t_k_pretest"]); t_ke = float(st_prof["t_k_exact"])
    t_kp_ov = float(np.mean([x["t_k_pretest"] for x in stats])); t_ke_ov = float(np.mean([x["t_k_exact"] for x in stats]))
    # The largest single kernel is k_pretest (one launch per step on this workload): it is arithmetic-bound (HBM traffic
    # under 1 % of peak in profiles/), so its roof is the FP32 rate.  Algorithmic work per pair test:
    # sums and differences of the residuals of the two epipolar lines (8 FMA = 16 flops), |S| - |D| - w for both lines
    # (4 adds), (S_1 D_1)(S_2 D_2) (3 multiplications): 23 flops in 4 FFMA2 + 2 FADD + 1 FADD2 + 1 FMUL2 + 1 FMUL; the
    # interval test of the ~4 % survivors runs in its own kernel (k_pretest2) and is not counted.
    PRETEST_FLOPS, PRETEST_INSTR = 23.0, 9.0
    achieved = PRETEST_FLOPS * st_prof["pair_tests"] / max(t_kp, 1e-12) / 1e12      # (rank 0's launch and rank 0's pair tests)
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
                    unit="TFLOP/s", frac=achieved / max(fp32_peak, 1e-9), traffic=PRETEST_TRAFFIC_PER_PAIR * st_prof["pair_tests"],
                    traffic_note="ncu dram bytes of one captured batch (profiles/r2_kernels.md) scaled by pair tests to this launch",
                    peak_source="FFMA-chain microkernel run in this process (l3d_measure_fp32_peak: 16 chains per thread, operands in registers); MEASURED_PEAKS.json has no FP32 figure",
                    peak_fp32x2=fp32x2_peak, peak_nominal=fp32_nominal, frac_of_nominal=achieved / fp32_nominal,
                    algorithmic_flops_per_launch=PRETEST_FLOPS * st_prof["pair_tests"], launches_per_step=1, launches_note="the timing pass runs the whole step's pair tests as ONE batch (one k_pretest launch); the timed steps cut them into about eight",
                    kernel_ms_per_launch=1e3 * t_kp,
                    fma_pipe_instr_per_pair=PRETEST_INSTR,
                    exact_kernel_ms=1e3 * t_ke, pretest2_kernel_ms=1e3 * float(st_prof["t_k_pretest2"]), kernel_ms_overlapped=dict(k_pretest=1e3 * t_kp_ov, k_exact_flat=1e3 * t_ke_ov), stages_ms={k[2:]: 1e3 * v for k, v in stage.items()},
                    stages_note="device ms per stage, and the kernels' own times, from one extra pass with CUDA events between the stages and "
                                f"no overlap of pair tests and per-view work; that pass took {1e3 * st_prof['t_matching_device']:.1f} ms on the device",
                    reference_arithmetic=dict(note="SURVEY.md 8(d) flops of the reference's own code for the same units, over this library's device time",
                                              pair_test_path_tflops=flops["t_pretest"] / max(stage["t_pretest"] + stage["t_exact"], 1e-12) / 1e12,
                                              scoring_tflops=flops["t_score"] / max(stage["t_score"], 1e-12) / 1e12),
                    hbm_peak_gbs=peaks.get("hbm_gbs"))
    cb = run_cpu(args.config, os.cpu_count() or 1) if world == 1 else None
    if cb:
        cb = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "same_k", "compute3Dmodel_seconds", "host_stages_seconds")}
        full = full_config_reference(args.config)
        if full:                                     # the whole configuration, timed once in the build container (not on this box)
            full["value"] = float(tot_pairs / max(world, 1)) / full["seconds_matching"] if not strong and world == 1 else None
            full["unit"] = UNIT
            cb["full_config"] = full

    value = tot_pairs / t_max
    out = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup, ms_per_step=1e3 * t_max,
               higher_is_better=True, scaling="strong" if strong else "weak", vs_baseline=None, dtype="f32", data="synthetic",
               config=dict(workload=f"{args.config}: {c['kw']['n_views']} views x ~{c['kw']['n_seg']} segments, k={c['k']}, no diffusion"
                                    + (f"; ONE connected scene split over {world} GPUs (every view's segments dealt out to the ranks, NCCL exchange per group of views)" if strong else
                                       (f"; {world} such districts in one scene, sharded by connected component over {world} GPUs" if world > 1 else "")),
                           pair_tests_per_step=int(tot_pairs), candidates_per_step=int(s["candidates"]), raw_matches_per_step=int(s["raw_matches"]),
                           scoring_iterations_per_step=int(s["score_iters"]), l2="flushed between steps (256 MiB memset); the per-step working set "
                           "(match lists of all views) is far larger than L2 anyway"),
               device_ms_per_step=1e3 * t_dev_max, loop_wall_s=loop_wall, clocks=clocks, gpu_launches=launches,
               e2e=dict(value=tot_pairs / t_e2e_max, unit=UNIT, h2d_bytes_per_step=int(st_e["h2d_bytes"]), d2h_bytes_per_step=int(st_e["d2h_bytes"]),
                        ms_per_step=1e3 * t_e2e_max, inside_calls_ms=1e3 * float(st_e["t_upload"] + st_e["t_matching_calls"]),
                        timed="wall clock of the upload plus the whole matching loop of Line3D::compute3Dmodel, the host's bookkeeping between the calls included "
                              "(the library works ahead of the host, so the time inside the calls alone, inside_calls_ms, would flatter it)",
                        api="Line3D::compute3Dmodel -> l3d_set_views + l3d_match_view per view (host buffers)" + ("; stopped after matching, all views' records on rank 0's host" if strong else ""),
                        compute3Dmodel_ms=None if strong else 1e3 * float(np.mean([tm["compute3Dmodel"] for tm, _ in e2e_stats])),
                        host_stages_ms={k[2:]: 1e3 * st_e[k] for k in ("t_neighbors", "t_upload", "t_matching", "t_hypotheses", "t_affinity", "t_diffusion", "t_clustering", "t_lines")}),
               roofline=roofline)
    if cb:
        out["cpu_baseline"] = cb
    if c["diffusion"] and st_e.get("rdd_entries", 0) > 0:
        # C3: the diffusion's step kernel against the measured HBM copy bandwidth (per entry and iteration
        # 16 bytes plus 32 bytes per lock-step pair)
        gbs = st_e["rdd_step_bytes"] / max(st_e["rdd_step_ms"] * 1e-3, 1e-12) / 1e9
        out["diffusion"] = dict(kernel="k_rdd_step", bound="hbm", achieved=gbs, peak=peaks.get("hbm_gbs"), unit="GB/s", frac=gbs / max(peaks.get("hbm_gbs") or 6548.5, 1e-9),
                                entries=int(st_e["rdd_entries"]), algorithmic_bytes_per_iteration=st_e["rdd_step_bytes"], kernel_ms=st_e["rdd_step_ms"], iterations=10,
                                stage_ms=1e3 * st_e["t_diffusion"], note="the matrix (a few tens of MB) stays in L2 between the ten iterations, so the rate can exceed the HBM figure")
    if single:
        out["single_gpu_same_workload"] = single
        out["speedup_vs_1gpu"] = value / single["value"]
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
    ap.add_argument("--config", default=None, choices=sorted(scenes.CONFIGS), help="default: C2 at one GPU (and with --weak), C4 on several")
    ap.add_argument("--weak", action="store_true", help="several GPUs: one C2-sized district per GPU, sharded by connected component")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours(args)


if __name__ == "__main__":
    main()
