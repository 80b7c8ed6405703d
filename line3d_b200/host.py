"""ctypes driver of libline3d_host.so: the L3D::Line3D C++ API (line3d_b200/csrc/line3D.h)
run on a scene file.  This is the call a user of the reference makes -- Line3D(...),
addImage per view, compute3Dmodel, getResult -- so bench.py's end-to-end number and the
end-to-end parity tests go through here."""
from __future__ import annotations

import ctypes as C
import os
import shutil
import tempfile

import numpy as np

from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB = os.path.join(_HERE, "libline3d_host.so")

DEFAULTS = dict(neighbors=10, unc_upper=5.0, unc_lower=1.0, sigma_p=3.5, sigma_a=10.0,
                min_baseline=0.25, collinearity=True, diffusion=False)  # commons.h:42-61

STAT_NAMES = ["pair_tests", "candidates", "raw_matches", "scored_matches", "filtered_matches", "score_iters", "score_iters_other_cam",
              "hypotheses", "affinity_pairs", "edges", "kernel_launches", "t_neighbors", "t_upload", "t_matching", "t_matching_device",
              "t_hypotheses", "t_affinity", "t_diffusion", "t_clustering", "t_lines", "t_total", "t_pretest", "t_exact", "t_assemble",
              "t_score", "t_reduce", "h2d_bytes", "d2h_bytes", "t_matching_calls", "t_k_pretest", "t_k_exact", "t_k_pretest2", "rdd_step_ms", "rdd_step_bytes", "rdd_entries", "_r1", "_r2", "_r3", "_r4", "_r5"]

_lib = None


def lib():
    global _lib
    if _lib is None:
        capi.lib()  # the CUDA library first: fail loudly if it is missing
        if not os.path.exists(HOST_LIB):
            raise capi.L3DError(f"{HOST_LIB} is missing: build it with __graft_entry__.build()")
        L = C.CDLL(HOST_LIB)
        L.l3dhost_open.restype = C.c_void_p
        L.l3dhost_open.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int]
        L.l3dhost_compute.restype = C.c_int
        L.l3dhost_compute.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.l3dhost_reset_geometry.restype = C.c_int; L.l3dhost_reset_geometry.argtypes = [C.c_void_p]
        L.l3dhost_set_shard.restype = C.c_int; L.l3dhost_set_shard.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.l3dhost_set_split.restype = C.c_int; L.l3dhost_set_split.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]
        L.l3dhost_dump.restype = C.c_int; L.l3dhost_dump.argtypes = [C.c_void_p, C.c_char_p]
        L.l3dhost_close.restype = None; L.l3dhost_close.argtypes = [C.c_void_p]
        _lib = L
    return _lib


class Session:
    """Line3D(...) + addImage for every view of a scene file; compute() = compute3Dmodel + getResult."""

    def __init__(self, scene_path: str, verbose: bool = False, device: int | None = None, **params):
        p = dict(DEFAULTS); p.update(params)
        self.params = p
        self._data_dir = tempfile.mkdtemp(prefix="l3d_data_")
        if device is not None:
            os.environ["L3D_DEVICE"] = str(device)
        self._h = lib().l3dhost_open(scene_path.encode(), self._data_dir.encode(), int(p["neighbors"]), p["unc_upper"], p["unc_lower"],
                                     p["sigma_p"], p["sigma_a"], p["min_baseline"], int(bool(p["collinearity"])), int(verbose))
        if not self._h:
            shutil.rmtree(self._data_dir, ignore_errors=True)
            raise capi.L3DError("l3dhost_open failed (no CUDA device, or unreadable scene)")

    def compute(self, measure_only: bool | int = False, flush_l2: bool = False, diffusion: bool | None = None, keep_debug: bool = False, fresh: bool = True, stage_timing: bool = False):
        """fresh: put the cameras back to their original poses first, so that every call sees the same input (a repeated
        compute3Dmodel otherwise re-normalises the normalised scene, as the reference does)."""
        if fresh:
            lib().l3dhost_reset_geometry(self._h)
        wall = (C.c_double * 2)(); st = (C.c_double * 40)()
        dif = self.params["diffusion"] if diffusion is None else diffusion
        n = lib().l3dhost_compute(self._h, int(bool(dif)), int(measure_only), int(flush_l2), int(keep_debug), int(stage_timing), wall, st)
        if n < 0:
            raise capi.L3DError(f"l3dhost_compute failed with code {n}")
        stats = dict(zip(STAT_NAMES, [float(x) for x in st]))
        return n, dict(add_image=wall[0], compute3Dmodel=wall[1]), stats

    def set_shard(self, rank: int, world: int):
        lib().l3dhost_set_shard(self._h, int(rank), int(world))

    def set_split(self, rank: int, world: int, nccl_id: bytes | None):
        """Several GPUs on ONE scene (Line3D::setSplit): nccl_id from capi.comm_unique_id() on rank 0, the same on all ranks."""
        if lib().l3dhost_set_split(self._h, int(rank), int(world), nccl_id) != 0:
            raise capi.L3DError("l3dhost_set_split failed (NCCL)")

    def dump(self) -> dict:
        d = tempfile.mkdtemp(prefix="l3d_dump_")
        lib().l3dhost_dump(self._h, d.encode())
        out = {f[:-4]: np.load(os.path.join(d, f)) for f in sorted(os.listdir(d)) if f.endswith(".npy")}
        shutil.rmtree(d, ignore_errors=True)
        return out

    def close(self):
        if getattr(self, "_h", None):
            lib().l3dhost_close(self._h)
            self._h = None
            shutil.rmtree(self._data_dir, ignore_errors=True)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def run_scene(scene_path: str, dump_dir: str | None = None, verbose: bool = False, **params):
    """One-shot convenience: returns (n_lines, timings, stats, dumps)."""
    s = Session(scene_path, verbose=verbose, **params)
    try:
        n, tm, st = s.compute(keep_debug=dump_dir is not None)
        dumps = s.dump() if dump_dir is not None else None
    finally:
        s.close()
    return n, tm, st, dumps
