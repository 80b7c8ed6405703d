"""ctypes binding of libl3d_b200.so (include/l3d_b200.h) -- the same C ABI the C++ host
scaffold (line3d_b200/csrc/line3D.cc) and a maintainer's own binding would use.

There is no CPU fallback: if the CUDA library is missing, or no CUDA device can be
opened, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libl3d_b200.so")


class L3DError(RuntimeError):
    pass


class View(C.Structure):
    _fields_ = [("id", C.c_int32), ("n_segments", C.c_int32), ("segments", C.POINTER(C.c_float)),
                ("RtKinv", C.c_float * 9), ("C", C.c_float * 3), ("P", C.c_float * 12), ("spatial_k", C.c_float),
                ("k_lower", C.c_float), ("k_upper", C.c_float), ("RtKinv_d", C.c_double * 9), ("C_d", C.c_double * 3)]


class Neighbor(C.Structure):
    _fields_ = [("view", C.c_int32), ("F", C.c_float * 9), ("to_be_matched", C.c_int32), ("deliver", C.c_int32)]


class ViewPlan(C.Structure):
    _fields_ = [("view", C.c_int32), ("n_neighbors", C.c_int32), ("neighbors", C.POINTER(Neighbor))]


class Match(C.Structure):
    _fields_ = [("src_seg", C.c_int32), ("cam", C.c_int32), ("tgt_seg", C.c_int32), ("confidence", C.c_float), ("depths", C.c_float * 4)]


MATCH_DTYPE = np.dtype([("src_seg", "<i4"), ("cam", "<i4"), ("tgt_seg", "<i4"), ("confidence", "<f4"), ("depths", "<f4", (4,))])
COLLIN_DTYPE = np.dtype([("i", "<i4"), ("j", "<i4"), ("value", "<f4")])
EDGE_DTYPE = np.dtype([("i", "<i4"), ("j", "<i4"), ("w", "<f4")])
HYP_IN_DTYPE = np.dtype([("view", "<i4"), ("seg", "<i4"), ("depth1", "<f4"), ("depth2", "<f4")])
SEG3D_DTYPE = np.dtype([("P1", "<f8", (3,)), ("P2", "<f8", (3,)), ("dir", "<f8", (3,))])
PAIR_DTYPE = np.dtype([("a", "<i4"), ("b", "<i4")])


class MatchResult(C.Structure):
    _fields_ = [("pair_tests", C.c_uint64), ("n_candidates", C.c_uint64), ("n_raw", C.c_uint64), ("n_delivered", C.c_uint64),
                ("n_scored", C.c_uint64), ("n_filtered", C.c_uint64), ("score_iters", C.c_uint64), ("score_iters_other_cam", C.c_uint64),
                ("scored", C.c_int32),
                ("median_depth", C.c_float), ("filtered", C.POINTER(Match))]


_lib = None


def lib():
    """Load the CUDA library; fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise L3DError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(nvcc, sm_100a). There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.l3d_create.restype = C.c_int; L.l3d_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.l3d_destroy.restype = None; L.l3d_destroy.argtypes = [vp]
        L.l3d_last_error.restype = C.c_char_p; L.l3d_last_error.argtypes = [vp]
        L.l3d_collinearity.restype = C.c_int; L.l3d_collinearity.argtypes = [vp, vp, C.c_int, C.c_float, C.POINTER(vp), C.POINTER(C.c_size_t)]
        L.l3d_collinearity_prealloc.restype = C.c_int
        L.l3d_collinearity_prealloc.argtypes = [vp, C.c_int, C.c_float, vp, C.c_int, vp, vp, C.c_char_p, C.c_size_t]
        L.l3d_set_views.restype = C.c_int; L.l3d_set_views.argtypes = [vp, C.c_int, C.POINTER(View)]
        for f in (L.l3d_match_view, L.l3d_match_view_resident):
            f.restype = C.c_int; f.argtypes = [vp, C.c_int, C.c_int, C.POINTER(Neighbor), C.c_float, C.c_float, C.POINTER(MatchResult)]
        L.l3d_reset_matching.restype = C.c_int; L.l3d_reset_matching.argtypes = [vp]
        L.l3d_unproject.restype = C.c_int; L.l3d_unproject.argtypes = [vp, C.c_size_t, vp, vp]
        L.l3d_visual_neighbors.restype = C.c_int; L.l3d_visual_neighbors.argtypes = [vp, C.c_int, vp, vp, vp, vp, C.c_float, C.c_int, vp, vp]
        L.l3d_fundamentals.restype = C.c_int; L.l3d_fundamentals.argtypes = [vp, C.c_int, vp, C.c_size_t, vp, vp]
        L.l3d_debug_visual_neighbors_host.restype = C.c_int; L.l3d_debug_visual_neighbors_host.argtypes = [C.c_int, vp, vp, vp, vp, C.c_float, C.c_int, vp, vp]
        L.l3d_fit_lines.restype = C.c_int; L.l3d_fit_lines.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp, vp]
        L.l3d_debug_fit_lines_host.restype = C.c_int; L.l3d_debug_fit_lines_host.argtypes = [vp, C.c_int, vp, vp, vp, vp, vp]
        L.l3d_debug_fundamentals_host.restype = C.c_int; L.l3d_debug_fundamentals_host.argtypes = [C.c_int, vp, C.c_size_t, vp, vp]
        L.l3d_affinity.restype = C.c_int; L.l3d_affinity.argtypes = [vp, vp, C.c_size_t, vp, C.c_float, vp]
        L.l3d_diffuse.restype = C.c_int; L.l3d_diffuse.argtypes = [vp, C.c_int, C.c_size_t, vp, vp]
        L.l3d_last_timings.restype = C.c_int; L.l3d_last_timings.argtypes = [vp, C.POINTER(C.c_float)]
        L.l3d_measure_fp32_peak.restype = C.c_int; L.l3d_measure_fp32_peak.argtypes = [vp, C.POINTER(C.c_float)]
        L.l3d_measure_fp32x2_peak.restype = C.c_int; L.l3d_measure_fp32x2_peak.argtypes = [vp, C.POINTER(C.c_float)]
        L.l3d_comm_unique_id.restype = C.c_int; L.l3d_comm_unique_id.argtypes = [C.c_char_p]
        L.l3d_comm_init.restype = C.c_int; L.l3d_comm_init.argtypes = [vp, C.c_int, C.c_int, C.c_char_p, C.c_int]
        L.l3d_comm_destroy.restype = C.c_int; L.l3d_comm_destroy.argtypes = [vp]
        L.l3d_debug_set_pretest.restype = C.c_int; L.l3d_debug_set_pretest.argtypes = [vp, C.c_int]
        L.l3d_debug_set_candidate_limit.restype = C.c_int; L.l3d_debug_set_candidate_limit.argtypes = [vp, C.c_uint64]
        L.l3d_flush_l2.restype = C.c_int; L.l3d_flush_l2.argtypes = [vp]
        L.l3d_debug_math.restype = C.c_int; L.l3d_debug_math.argtypes = [vp, C.c_int, C.c_int, vp, vp]
        L.l3d_kernel_launches.restype = C.c_uint64; L.l3d_kernel_launches.argtypes = [vp]
        _lib = L
    return _lib


EXPORTS = ["l3d_create", "l3d_destroy", "l3d_last_error", "l3d_collinearity", "l3d_collinearity_prealloc", "l3d_set_views", "l3d_match_view",
           "l3d_match_view_resident", "l3d_retain_results", "l3d_plan_matching", "l3d_debug_set_batch_pairs", "l3d_comm_unique_id", "l3d_comm_init", "l3d_comm_destroy", "l3d_debug_host_times", "l3d_debug_chain_times", "l3d_debug_set_score_block", "l3d_diffuse_stats", "l3d_deliver", "l3d_candidate_buffer", "l3d_affinity_edges", "l3d_debug_set_pretest", "l3d_debug_set_candidate_limit", "l3d_measure_fp32x2_peak", "l3d_reset_matching", "l3d_visual_neighbors", "l3d_fundamentals", "l3d_debug_visual_neighbors_host", "l3d_debug_fundamentals_host", "l3d_fit_lines", "l3d_debug_fit_lines_host", "l3d_unproject", "l3d_affinity", "l3d_diffuse",
           "l3d_last_timings", "l3d_last_kernel_timings", "l3d_set_stage_timing", "l3d_mark", "l3d_marks_elapsed", "l3d_measure_fp32_peak", "l3d_kernel_launches", "l3d_debug_math", "l3d_flush_l2"]


COMM_ID_BYTES = 128


CAMERA_DOUBLES = 24        # l3d_camera: K[9], R[9], t[3], C[3]


def _geometry_args(cameras, start, other, similarity):
    cams = np.ascontiguousarray(cameras, dtype=np.float64).reshape(-1, CAMERA_DOUBLES)
    start = np.ascontiguousarray(start, dtype=np.int32); other = np.ascontiguousarray(other, dtype=np.int32)
    sim = np.ascontiguousarray(similarity, dtype=np.float32)
    assert start.shape[0] == cams.shape[0] + 1 and other.shape[0] == sim.shape[0] == int(start[-1])
    return cams, start, other, sim


def visual_neighbors_host(cameras, start, other, similarity, min_baseline: float, k: int):
    """l3d_debug_visual_neighbors_host: the neighbour selection compiled for the host (tests; no GPU needed).  Returns a
    list of int32 arrays, one per view."""
    cams, start, other, sim = _geometry_args(cameras, start, other, similarity)
    count = np.zeros(cams.shape[0], np.int32); out = np.zeros(max(other.shape[0], 1), np.int32)
    if lib().l3d_debug_visual_neighbors_host(cams.shape[0], cams.ctypes.data, start.ctypes.data, other.ctypes.data, sim.ctypes.data, min_baseline, k, count.ctypes.data, out.ctypes.data) != 0:
        raise L3DError("l3d_debug_visual_neighbors_host: bad arguments")
    return [out[start[v]: start[v] + count[v]].copy() for v in range(cams.shape[0])]


def fundamentals_host(cameras, pairs) -> np.ndarray:
    """l3d_debug_fundamentals_host: F (n, 3, 3) float64 for (source, neighbour) index pairs, on the host (tests)."""
    cams = np.ascontiguousarray(cameras, dtype=np.float64).reshape(-1, CAMERA_DOUBLES)
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    F = np.zeros((pairs.shape[0], 3, 3), np.float64)
    if lib().l3d_debug_fundamentals_host(cams.shape[0], cams.ctypes.data, pairs.shape[0], pairs.ctypes.data, F.ctypes.data) != 0:
        raise L3DError("l3d_debug_fundamentals_host: bad arguments")
    return F


def similarity_back(scale: float, R, t) -> np.ndarray:
    """l3d_similarity for the way back from the normalised scene (line3D.cc:1782-1786): 1 / scale, R^T, -t."""
    R = np.asarray(R, np.float64).reshape(3, 3); t = np.asarray(t, np.float64).reshape(3)
    return np.concatenate([[1.0 / float(scale)], R.T.ravel(), -t]).astype(np.float64)


def _split_segments(start, count, seg):
    return [seg[start[c]: start[c] + count[c]].copy() for c in range(len(count))]


def fit_lines_host(back, start, points, member_camera):
    """l3d_debug_fit_lines_host: per cluster an (n, 6) float64 array of 3-D segments (tests; no GPU needed)."""
    back = np.ascontiguousarray(back, np.float64); start = np.ascontiguousarray(start, np.int32)
    points = np.ascontiguousarray(points, np.float64).reshape(-1, 6); cam = np.ascontiguousarray(member_camera, np.uint32)
    assert points.shape[0] == cam.shape[0] == int(start[-1])
    count = np.zeros(start.shape[0] - 1, np.int32); seg = np.zeros((max(points.shape[0], 1), 6), np.float64)
    if lib().l3d_debug_fit_lines_host(back.ctypes.data, count.shape[0], start.ctypes.data, points.ctypes.data, cam.ctypes.data, count.ctypes.data, seg.ctypes.data) != 0:
        raise L3DError("l3d_debug_fit_lines_host: bad arguments")
    return _split_segments(start, count, seg)


def comm_unique_id() -> bytes:
    """An NCCL id for l3d_comm_init: made on rank 0, handed to the other ranks by the caller (e.g. a gloo broadcast)."""
    buf = C.create_string_buffer(COMM_ID_BYTES)
    if lib().l3d_comm_unique_id(buf) != 0:
        raise L3DError(lib().l3d_last_error(None).decode())
    return buf.raw


class Context:
    """One l3d_ctx: one CUDA device, one host thread."""

    def __init__(self, device: int = 0):
        self._L = lib()
        h = C.c_void_p()
        if self._L.l3d_create(device, C.byref(h)) != 0:
            raise L3DError(self._L.l3d_last_error(None).decode())
        self._h = h
        self._keep = None

    def close(self):
        if getattr(self, "_h", None):
            self._L.l3d_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise L3DError(self._L.l3d_last_error(self._h).decode())

    # ---- neighbour graph and fundamental matrices ------------------------------------------
    def visual_neighbors(self, cameras, start, other, similarity, min_baseline: float, k: int):
        cams, start, other, sim = _geometry_args(cameras, start, other, similarity)
        count = np.zeros(cams.shape[0], np.int32); out = np.zeros(max(other.shape[0], 1), np.int32)
        self._check(self._L.l3d_visual_neighbors(self._h, cams.shape[0], cams.ctypes.data, start.ctypes.data, other.ctypes.data, sim.ctypes.data, min_baseline, k, count.ctypes.data, out.ctypes.data))
        return [out[start[v]: start[v] + count[v]].copy() for v in range(cams.shape[0])]

    def fit_lines(self, back, start, member, member_camera):
        """l3d_fit_lines on the hypotheses of the last unproject(): per cluster an (n, 6) float64 array of 3-D segments."""
        back = np.ascontiguousarray(back, np.float64); start = np.ascontiguousarray(start, np.int32)
        member = np.ascontiguousarray(member, np.int32); cam = np.ascontiguousarray(member_camera, np.uint32)
        count = np.zeros(start.shape[0] - 1, np.int32); seg = np.zeros((max(member.shape[0], 1), 6), np.float64)
        self._check(self._L.l3d_fit_lines(self._h, back.ctypes.data, count.shape[0], start.ctypes.data, member.ctypes.data, cam.ctypes.data, count.ctypes.data, seg.ctypes.data))
        return _split_segments(start, count, seg)

    def fundamentals(self, cameras, pairs) -> np.ndarray:
        cams = np.ascontiguousarray(cameras, dtype=np.float64).reshape(-1, CAMERA_DOUBLES)
        pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
        F = np.zeros((pairs.shape[0], 3, 3), np.float64)
        self._check(self._L.l3d_fundamentals(self._h, cams.shape[0], cams.ctypes.data, pairs.shape[0], pairs.ctypes.data, F.ctypes.data))
        return F

    # ---- collinearity ----------------------------------------------------------------
    def collinearity(self, segs: np.ndarray, s: float = 2.0) -> np.ndarray:
        segs = np.ascontiguousarray(segs, dtype=np.float32).reshape(-1, 4)
        p = C.c_void_p(); n = C.c_size_t()
        self._check(self._L.l3d_collinearity(self._h, segs.ctypes.data, segs.shape[0], s, C.byref(p), C.byref(n)))
        if n.value == 0:
            return np.zeros(0, COLLIN_DTYPE)
        buf = (C.c_char * (n.value * COLLIN_DTYPE.itemsize)).from_address(p.value)
        return np.frombuffer(buf, dtype=COLLIN_DTYPE).copy()

    # ---- views -------------------------------------------------------------------------
    def set_views(self, views: list[dict]):
        """views: dicts with id, segs (n,4) f32, RtKinv (3,3), C (3,), P (3,4) as float64 arrays, spatial_k, k_lower, k_upper."""
        arr = (View * len(views))()
        keep = []
        for i, v in enumerate(views):
            segs = np.ascontiguousarray(v["segs"], dtype=np.float32).reshape(-1, 4)
            keep.append(segs)
            a = arr[i]
            a.id = int(v["id"]); a.n_segments = segs.shape[0]
            a.segments = segs.ctypes.data_as(C.POINTER(C.c_float))
            R = np.asarray(v["RtKinv"], dtype=np.float64).reshape(9); Cc = np.asarray(v["C"], dtype=np.float64).reshape(3)
            P = np.asarray(v["P"], dtype=np.float64).reshape(12)
            a.RtKinv[:] = [float(np.float32(x)) for x in R]; a.C[:] = [float(np.float32(x)) for x in Cc]; a.P[:] = [float(np.float32(x)) for x in P]
            a.RtKinv_d[:] = list(R); a.C_d[:] = list(Cc)
            a.spatial_k = float(v["spatial_k"]); a.k_lower = float(v["k_lower"]); a.k_upper = float(v["k_upper"])
        self._check(self._L.l3d_set_views(self._h, len(views), arr))
        self._n_views = len(views)

    def comm_init(self, rank: int, world: int, nccl_id: bytes | None, result_root: int = -1):
        self._check(self._L.l3d_comm_init(self._h, int(rank), int(world), nccl_id, int(result_root)))

    def set_stage_timing(self, on: bool):
        self._check(self._L.l3d_set_stage_timing(self._h, int(bool(on))))

    def reset_matching(self):
        self._check(self._L.l3d_reset_matching(self._h))

    @staticmethod
    def _neighbors(neighbors):
        nb = (Neighbor * len(neighbors))()
        for i, n in enumerate(neighbors):
            nb[i].view = int(n["view"]); nb[i].to_be_matched = int(bool(n["to_be_matched"])); nb[i].deliver = int(bool(n["deliver"]))
            nb[i].F[:] = [float(np.float32(x)) for x in np.asarray(n["F"], dtype=np.float64).reshape(9)]
        return nb

    def plan_matching(self, schedule: list[tuple[int, list[dict]]]):
        """schedule: (view, neighbors) in the order the views will be passed to match_view."""
        keep = [self._neighbors(nbs) for _, nbs in schedule]
        plans = (ViewPlan * max(len(schedule), 1))()
        for k, (view, nbs) in enumerate(schedule):
            plans[k].view = int(view); plans[k].n_neighbors = len(nbs); plans[k].neighbors = C.cast(keep[k], C.POINTER(Neighbor))
        self._check(self._L.l3d_plan_matching(self._h, len(schedule), plans))

    def debug_set_batch_pairs(self, pairs: int):
        self._L.l3d_debug_set_batch_pairs.argtypes = [C.c_void_p, C.c_uint64]
        self._check(self._L.l3d_debug_set_batch_pairs(self._h, int(pairs)))

    def debug_set_pretest(self, mode: int):
        self._check(self._L.l3d_debug_set_pretest(self._h, int(mode)))

    def debug_set_score_block(self, hypotheses: int):
        self._L.l3d_debug_set_score_block.argtypes = [C.c_void_p, C.c_int]
        self._check(self._L.l3d_debug_set_score_block(self._h, int(hypotheses)))

    def debug_set_candidate_limit(self, limit: int):
        self._check(self._L.l3d_debug_set_candidate_limit(self._h, int(limit)))

    def retain_results(self, on: bool):
        self._check(self._L.l3d_retain_results(self._h, int(bool(on))))

    def match_view(self, view: int, neighbors: list[dict], sigma_p: float = 3.5, sigma_a: float = 10.0, resident: bool = False, copy: bool = True):
        """neighbors: dicts with view, F (3,3) float64, to_be_matched, deliver; ascending view order.
        copy=False returns a numpy view of the library's own buffer (what `filtered` points to)."""
        nb = self._neighbors(neighbors)
        res = MatchResult()
        fn = self._L.l3d_match_view_resident if resident else self._L.l3d_match_view
        self._check(fn(self._h, view, len(neighbors), nb, sigma_p, sigma_a, C.byref(res)))
        out = dict(pair_tests=res.pair_tests, n_candidates=res.n_candidates, n_raw=res.n_raw, n_delivered=res.n_delivered,
                   n_scored=res.n_scored, n_filtered=res.n_filtered, score_iters=res.score_iters,
                   score_iters_other_cam=res.score_iters_other_cam, scored=bool(res.scored),
                   median_depth=float(res.median_depth), filtered=None)
        if not resident:
            if res.n_filtered:
                buf = (C.c_char * (res.n_filtered * MATCH_DTYPE.itemsize)).from_address(C.addressof(res.filtered.contents))
                out["filtered"] = np.frombuffer(buf, dtype=MATCH_DTYPE)
                if copy:
                    out["filtered"] = out["filtered"].copy()
            else:
                out["filtered"] = np.zeros(0, MATCH_DTYPE)
        return out

    # ---- hypotheses / affinity / diffusion ------------------------------------------------
    def unproject(self, hyp_in: np.ndarray) -> np.ndarray:
        hyp_in = np.ascontiguousarray(hyp_in, dtype=HYP_IN_DTYPE)
        out = np.zeros(hyp_in.shape[0], SEG3D_DTYPE)
        self._check(self._L.l3d_unproject(self._h, hyp_in.shape[0], hyp_in.ctypes.data, out.ctypes.data))
        return out

    def affinity(self, median_depth: np.ndarray, pairs: np.ndarray, sigma_a: float = 10.0) -> np.ndarray:
        md = np.ascontiguousarray(median_depth, dtype=np.float32)
        assert md.shape[0] == self._n_views
        pairs = np.ascontiguousarray(pairs, dtype=PAIR_DTYPE)
        out = np.zeros(pairs.shape[0], np.float32)
        self._check(self._L.l3d_affinity(self._h, md.ctypes.data, pairs.shape[0], pairs.ctypes.data, sigma_a, out.ctypes.data))
        return out

    def diffuse(self, n_nodes: int, edges: np.ndarray) -> np.ndarray:
        edges = np.ascontiguousarray(edges, dtype=EDGE_DTYPE)
        out = np.zeros(edges.shape[0], EDGE_DTYPE)
        self._check(self._L.l3d_diffuse(self._h, n_nodes, edges.shape[0], edges.ctypes.data, out.ctypes.data))
        return out

    # ---- measurement ------------------------------------------------------------------------
    def last_timings(self):
        ms = (C.c_float * 6)()
        self._check(self._L.l3d_last_timings(self._h, ms))
        return dict(zip(("pretest", "exact", "assemble", "score", "reduce", "total"), [float(x) for x in ms]))

    def fp32_peak_tflops(self) -> float:
        t = C.c_float()
        self._check(self._L.l3d_measure_fp32_peak(self._h, C.byref(t)))
        return float(t.value)

    def fp32x2_peak_tflops(self) -> float:
        t = C.c_float()
        self._check(self._L.l3d_measure_fp32x2_peak(self._h, C.byref(t)))
        return float(t.value)

    def flush_l2(self):
        self._check(self._L.l3d_flush_l2(self._h))

    def debug_math(self, which: int, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.zeros_like(x)
        self._check(self._L.l3d_debug_math(self._h, which, x.shape[0], x.ctypes.data, out.ctypes.data))
        return out

    def kernel_launches(self) -> int:
        return int(self._L.l3d_kernel_launches(self._h))
