"""oracle/refrun.py -- TEST INFRASTRUCTURE.  ctypes driver for oracle/_ref/libline3d_ref.so
(the reference's own sources compiled for the CPU by oracle/Makefile).  Used by tests, by
tests/golden/make_golden.py and by bench.py's reference arm; never by the product path."""
from __future__ import annotations

import ctypes
import os
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libline3d_ref.so")
# the same reference host code and harness, with integration/cudawrapper_b200.cc + libl3d_b200.so in place of the
# reference's CUDA file (oracle/Makefile, target refhost): needs a GPU
REFHOST_B200_SO = os.path.join(_HERE, "_ref", "libline3d_refhost_b200.so")

DEFAULTS = dict(neighbors=10, unc_upper=5.0, unc_lower=1.0, sigma_p=3.5, sigma_a=10.0,
                min_baseline=0.25, collinearity=True, diffusion=False)  # commons.h:42-61


def available() -> bool:
    return os.path.exists(REF_SO)


_libs = {}


def lib(path: str = REF_SO):
    if path not in _libs:
        L = ctypes.CDLL(path)
        L.ref_run_scene.restype = ctypes.c_int
        L.ref_run_scene.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_float, ctypes.c_float,
                                    ctypes.c_float, ctypes.c_float, ctypes.c_float, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_char_p, ctypes.POINTER(ctypes.c_double)]
        if hasattr(L, "ref_emu_set_threads"):
            L.ref_emu_set_threads.argtypes = [ctypes.c_int]
        _libs[path] = L
    return _libs[path]


def run_scene(scene_path: str, dump_dir: str | None = None, threads: int = 1, verbose: bool = False, so: str = REF_SO, **params):
    """Run the reference on a scene file. Returns (n_lines, timings dict, dumps dict or None)."""
    p = dict(DEFAULTS); p.update(params)
    L = lib(so)
    if hasattr(L, "ref_emu_set_threads"):
        L.ref_emu_set_threads(int(threads))
    tm = (ctypes.c_double * 5)()
    with tempfile.TemporaryDirectory(prefix="l3dref_") as data_dir:
        n = L.ref_run_scene(scene_path.encode(), data_dir.encode(), int(p["neighbors"]), p["unc_upper"], p["unc_lower"],
                            p["sigma_p"], p["sigma_a"], p["min_baseline"], int(bool(p["collinearity"])),
                            int(bool(p["diffusion"])), int(verbose), (dump_dir or "").encode(), tm)
    timings = dict(add_image=tm[0], compute3Dmodel=tm[1], matching=tm[2], diffusion=tm[3], clustering=tm[4])
    dumps = None
    if dump_dir:
        dumps = {f[:-4]: np.load(os.path.join(dump_dir, f)) for f in sorted(os.listdir(dump_dir)) if f.endswith(".npy")}
    return n, timings, dumps


def collinearity(segs: np.ndarray, s: float = 2.0, so: str = REF_SO) -> np.ndarray:
    segs = np.ascontiguousarray(segs, dtype=np.float32)
    n = segs.shape[0]
    out = np.zeros((n, n), dtype=np.float32)
    L = lib(so)
    L.ref_collinearity(segs.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(n), ctypes.c_float(s), out.ctypes.data_as(ctypes.c_void_p))
    return out


def diffusion(ij: np.ndarray, w: np.ndarray, n: int, so: str = REF_SO) -> np.ndarray:
    ij = np.ascontiguousarray(ij, dtype=np.int32); w = np.ascontiguousarray(w, dtype=np.float32)
    out = np.zeros((w.shape[0], 3), dtype=np.float32)
    L = lib(so)
    L.ref_diffusion(ij.ctypes.data_as(ctypes.c_void_p), w.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(w.shape[0]), ctypes.c_int(n),
                    out.ctypes.data_as(ctypes.c_void_p))
    return out
