"""ctypes driver of libline3d_host.so: the L3D::Line3D C++ API (line3d_b200/csrc/line3D.h)
run on a scene file.  This is the call a user of the reference makes -- Line3D(...),
addImage per view, compute3Dmodel, getResult -- so bench.py's end-to-end number and the
end-to-end parity tests go through here."""
from __future__ import annotations

import ctypes as C
import os
import tempfile

import numpy as np

from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB = os.path.join(_HERE, "libline3d_host.so")

DEFAULTS = dict(neighbors=10, unc_upper=5.0, unc_lower=1.0, sigma_p=3.5, sigma_a=10.0,
                min_baseline=0.25, collinearity=True, diffusion=False)  # commons.h:42-61

STAT_NAMES = ["pair_tests", "candidates", "raw_matches", "scored_matches", "filtered_matches", "score_iters", "hypotheses",
              "affinity_pairs", "edges", "kernel_launches", "t_neighbors", "t_upload", "t_matching", "t_matching_device",
              "t_hypotheses", "t_affinity", "t_diffusion", "t_clustering", "t_lines", "t_total"]

_lib = None


def lib():
    global _lib
    if _lib is None:
        capi.lib()  # the CUDA library first: fail loudly if it is missing
        if not os.path.exists(HOST_LIB):
            raise capi.L3DError(f"{HOST_LIB} is missing: build it with __graft_entry__.build()")
        L = C.CDLL(HOST_LIB)
        L.l3dhost_run_scene.restype = C.c_int
        L.l3dhost_run_scene.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float,
                                        C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        _lib = L
    return _lib


def run_scene(scene_path: str, dump_dir: str | None = None, verbose: bool = False, measure_only: bool = False, repeat: int = 1, **params):
    """Returns (n_lines, timings, stats, dumps)."""
    p = dict(DEFAULTS); p.update(params)
    tm = (C.c_double * 2)(); st = (C.c_double * 20)()
    with tempfile.TemporaryDirectory(prefix="l3d_data_") as data_dir:
        n = lib().l3dhost_run_scene(scene_path.encode(), data_dir.encode(), int(p["neighbors"]), p["unc_upper"], p["unc_lower"],
                                    p["sigma_p"], p["sigma_a"], p["min_baseline"], int(bool(p["collinearity"])), int(bool(p["diffusion"])),
                                    int(verbose), int(measure_only), int(repeat), (dump_dir or "").encode(), tm, st)
    if n < 0:
        raise capi.L3DError(f"l3dhost_run_scene failed with code {n}")
    timings = dict(add_image=tm[0], compute3Dmodel=tm[1])
    stats = dict(zip(STAT_NAMES, [float(x) for x in st]))
    dumps = None
    if dump_dir:
        dumps = {f[:-4]: np.load(os.path.join(dump_dir, f)) for f in sorted(os.listdir(dump_dir)) if f.endswith(".npy")}
    return n, timings, stats, dumps
