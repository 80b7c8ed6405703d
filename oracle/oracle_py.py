"""oracle/oracle_py.py -- TEST INFRASTRUCTURE.  ctypes wrapper around oracle/liboracle.so
(the CPU restatement in oracle.cpp).  Imported only by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / reference arm."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "liboracle.so")

DEFAULTS = dict(neighbors=10, unc_upper=5.0, unc_lower=1.0, sigma_p=3.5, sigma_a=10.0,
                min_baseline=0.25, collinearity=True, diffusion=False)  # commons.h:42-61

_lib = None
_f = ctypes.c_float
_vp = ctypes.c_void_p


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            build()
        L = ctypes.CDLL(ORACLE_SO)
        L.oracle_run_scene.restype = ctypes.c_int
        L.oracle_run_scene.argtypes = [ctypes.c_char_p, ctypes.c_int, _f, _f, _f, _f, _f, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
        L.oracle_set_shard.restype = None; L.oracle_set_shard.argtypes = [ctypes.c_int, ctypes.c_int]
        L.oracle_pair_test.restype = ctypes.c_int
        L.oracle_pair_test.argtypes = [_vp] * 8
        L.oracle_match_pair.restype = ctypes.c_long
        L.oracle_match_pair.argtypes = [_vp, ctypes.c_int, _vp, ctypes.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, ctypes.c_long, _vp]
        L.oracle_score_matches.restype = None
        L.oracle_score_matches.argtypes = [_vp, _vp, ctypes.c_long, _vp, ctypes.c_int, _vp, _vp, _vp, _vp, _vp, ctypes.c_int, _f, _f, _f, _vp]
        L.oracle_collinearity.restype = None
        L.oracle_collinearity.argtypes = [_vp, ctypes.c_int, _f, _vp]
        L.oracle_diffusion.restype = None
        L.oracle_diffusion.argtypes = [_vp, _vp, ctypes.c_int, ctypes.c_int, _vp]
        L.oracle_similarity.restype = _f
        L.oracle_similarity.argtypes = [_vp, _vp, _vp, _vp, _f]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(_vp)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def run_scene(scene_path, dump_dir=None, threads=1, shard=(0, 1), **params):
    p = dict(DEFAULTS); p.update(params)
    lib().oracle_set_shard(int(shard[0]), int(shard[1]))
    tm = (ctypes.c_double * 5)(); ct = (ctypes.c_double * 5)()
    n = lib().oracle_run_scene(scene_path.encode(), int(p["neighbors"]), p["unc_upper"], p["unc_lower"], p["sigma_p"], p["sigma_a"],
                               p["min_baseline"], int(bool(p["collinearity"])), int(bool(p["diffusion"])),
                               (dump_dir or "").encode(), int(threads), tm, ct)
    timings = dict(add_image=tm[0], compute3Dmodel=tm[1], matching=tm[2], diffusion=tm[3], clustering=tm[4])
    counters = dict(pair_tests=ct[0], accepted=ct[1], raw=ct[2], score_iters=ct[3], gate_pass=ct[4])
    dumps = None
    if dump_dir:
        dumps = {f[:-4]: np.load(os.path.join(dump_dir, f)) for f in sorted(os.listdir(dump_dir)) if f.endswith(".npy")}
    return n, timings, counters, dumps


def pair_test(seg_src, seg_tgt, F, RtKinv_src, RtKinv_tgt, C_src, C_tgt):
    a = [_f32(x) for x in (seg_src, seg_tgt, F, RtKinv_src, RtKinv_tgt, C_src, C_tgt)]
    out = np.zeros(6, np.float32)
    acc = lib().oracle_pair_test(*[_p(x) for x in a], _p(out))
    return bool(acc), out[:4].copy(), out[4:].copy()


def match_pair(segs_src, segs_tgt, F, RtKinv_src, RtKinv_tgt, C_src, C_tgt):
    ss, st = _f32(segs_src), _f32(segs_tgt)
    a = [_f32(x) for x in (F, RtKinv_src, RtKinv_tgt, C_src, C_tgt)]
    cap = max(ss.shape[0] * st.shape[0] // 4, 1024)
    while True:
        ij = np.zeros((cap, 2), np.int32); dp = np.zeros((cap, 4), np.float32)
        band = ctypes.c_long(0)
        n = lib().oracle_match_pair(_p(ss), ss.shape[0], _p(st), st.shape[0], *[_p(x) for x in a], _p(ij), _p(dp), cap, ctypes.addressof(band))
        if n <= cap:
            return ij[:n].copy(), dp[:n].copy(), band.value
        cap = n


def score_matches(data, depths, segs_src, RtKinv_src, C_src, tgt_segs, cam_offsets, projections, sigma_p, sigma_a, spatial_k):
    data, depths, segs_src, tgt_segs = _f32(data), _f32(depths), _f32(segs_src), _f32(tgt_segs)
    R, C, P = _f32(RtKinv_src), _f32(C_src), _f32(projections)
    off = np.ascontiguousarray(cam_offsets, dtype=np.int32)
    out = np.zeros(data.shape[0], np.float32)
    lib().oracle_score_matches(_p(data), _p(depths), data.shape[0], _p(segs_src), segs_src.shape[0], _p(R), _p(C), _p(tgt_segs), _p(off), _p(P),
                               off.shape[0], sigma_p, sigma_a, spatial_k, _p(out))
    return out


def collinearity(segs, s=2.0):
    segs = _f32(segs); n = segs.shape[0]
    out = np.zeros((n, n), np.float32)
    lib().oracle_collinearity(_p(segs), n, s, _p(out))
    return out


def diffusion(ij, w, n):
    ij = np.ascontiguousarray(ij, dtype=np.int32); w = _f32(w)
    out = np.zeros((w.shape[0], 3), np.float32)
    lib().oracle_diffusion(_p(ij), _p(w), w.shape[0], n, _p(out))
    return out


def similarity(h1, v1, h2, v2, sigma_a=10.0):
    a = [np.ascontiguousarray(x, dtype=np.float64) for x in (h1, v1, h2, v2)]
    return float(lib().oracle_similarity(*[_p(x) for x in a], sigma_a))
