"""Multi-GPU plumbing: one process per GPU (torchrun), torch.distributed for the collectives
(backend nccl on GPUs over NVLink, gloo in the CPU tests).

The matching-and-scoring path shards by connected component of the neighbour graph
(Line3D::setShard): every rank holds the full tables, matches and scores
the views of its components, and only results travel -- the final 3-D lines and the work
counters are gathered on rank 0.  There is no data-path collective inside the hot loop.
"""
from __future__ import annotations

import os

import numpy as np


def env_rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def init(backend: str | None = None):
    """Initialise torch.distributed from the torchrun environment; returns (rank, world, local_rank, dist or None)."""
    rank, world, local = env_rank_world()
    if world <= 1:
        return rank, world, local, None
    import torch
    import torch.distributed as dist
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(local)
    if not dist.is_initialized():
        dist.init_process_group(backend)
    return rank, world, local, dist


def lines_from_dumps(dm: dict) -> list:
    """Final lines of one rank as a list of (tuple of (cam, seg), array of 3-D sub-segments)."""
    s2, s3 = dm["lines_seg2d"], dm["lines_seg3d"]
    out = []
    n = int(max(s2[:, 0].max() if s2.size else -1, s3[:, 0].max() if s3.size else -1)) + 1
    for k in range(n):
        a = s2[s2[:, 0] == k][:, 1:]
        b = s3[s3[:, 0] == k][:, 1:]
        out.append((tuple(map(tuple, a.tolist())), b))
    return out


def gather_lines(lines: list, dist, dst: int = 0):
    """Gather every rank's final lines on `dst`; components are disjoint, so the union is the result.
    Lines are returned sorted by their smallest (cam, seg) member, which does not depend on the sharding."""
    if dist is None:
        merged = list(lines)
    else:
        bucket = [None] * dist.get_world_size() if dist.get_rank() == dst else None
        dist.gather_object(lines, bucket, dst=dst)
        if dist.get_rank() != dst:
            return None
        merged = [l for part in bucket for l in part]
    merged.sort(key=lambda l: min(l[0]) if l[0] else (1 << 30, 0))
    return merged


def reduce_counters(values: dict, dist, device=None):
    """Sum (keys starting with 'n_' or 'pair') / max (keys starting with 't_') over ranks; every rank gets the result."""
    if dist is None:
        return dict(values)
    import torch
    keys = sorted(values)
    sums = torch.tensor([float(values[k]) if not k.startswith("t_") else 0.0 for k in keys], dtype=torch.float64, device=device)
    maxs = torch.tensor([float(values[k]) if k.startswith("t_") else 0.0 for k in keys], dtype=torch.float64, device=device)
    dist.all_reduce(sums)
    dist.all_reduce(maxs, op=dist.ReduceOp.MAX)
    return {k: (float(maxs[i]) if k.startswith("t_") else float(sums[i])) for i, k in enumerate(keys)}
