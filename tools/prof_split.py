"""Per-kernel device times of the per-view chain with ONE scene split over the GPUs of a torchrun job (or on one GPU):
one warm-up pass, one pass with a CUDA event after every kernel.  L3D_PROFILE=1 prints the table (rank 0's is enough).

    L3D_PROFILE=1 python -m torch.distributed.run --nproc-per-node 8 ... tools/prof_split.py [C4]
"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from line3d_b200 import capi, host, scenes

config = sys.argv[1] if len(sys.argv) > 1 else "C4"
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
if rank != 0: os.environ.pop("L3D_PROFILE", None)      # rank 0's table is enough
dist = None
if world > 1:
    import torch, torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
barrier = (lambda: dist.barrier()) if dist else (lambda: None)
c = scenes.CONFIGS[config]
path = bench.scene_file_by_ranks(config, rank, world, barrier) if world > 1 else bench.scene_file(config)
s = host.Session(path, device=local, neighbors=c["k"], collinearity=False)
if world > 1:
    box = [capi.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    s.set_split(rank, world, box[0])
s.compute(measure_only=True)
for timing in (False, True, False):
    barrier()
    t = time.time()
    n, tm, st = s.compute(measure_only=True, stage_timing=timing)
    if rank == 0:
        print("stage_timing", timing, "matching %.1f ms" % (1e3 * st["t_matching"]), {k: round(1e3 * st[k], 1) for k in ("t_pretest", "t_exact", "t_assemble", "t_score", "t_reduce")}, flush=True)
# how the pair tests share the GPU with the per-view chain: the knobs are read once per schedule (l3d_plan_matching)
if os.environ.get("L3D_SWEEP"):
    knobs = ("L3D_BG_BLOCKS", "L3D_P1_ROWS", "L3D_NO_OVERLAP", "L3D_BATCH_DIV", "L3D_GROUP_VIEWS")
    sweep = [{}, {"L3D_NO_OVERLAP": 1}, {"L3D_GROUP_VIEWS": 16}, {"L3D_BG_BLOCKS": 4}, {"L3D_BG_BLOCKS": 3}, {"L3D_BG_BLOCKS": 2}, {"L3D_P1_ROWS": 128}, {"L3D_P1_ROWS": 128, "L3D_BG_BLOCKS": 3},
             {"L3D_BATCH_DIV": 16}, {"L3D_BATCH_DIV": 4}, {"L3D_BATCH_DIV": 2}, {"L3D_BATCH_DIV": 16, "L3D_BG_BLOCKS": 3}, {"L3D_GROUP_VIEWS": 4}, {}]
    for cfg in sweep:
        for k in knobs: os.environ.pop(k, None)
        for k, v in cfg.items(): os.environ[k] = str(v)
        ms = []
        for it in range(3):
            barrier()
            n, tm, st = s.compute(measure_only=True)
            ms.append(1e3 * st["t_matching"])
        if rank == 0:
            print("sweep", cfg, "matching ms", [round(x, 1) for x in ms], flush=True)
s.close()
if dist:
    dist.barrier(); dist.destroy_process_group()
