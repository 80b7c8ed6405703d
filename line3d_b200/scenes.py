"""Seeded synthetic wireframe scenes for the Line3D matching-and-scoring path.

The reference ships no sample data, so every workload in
BASELINE.json is synthesised here, deterministically from a seed:

* geometry: axis-aligned "building" boxes on a ground plane with a regular window grid on
  every facade and a roof grid (3-D line segments = the wireframe);
* cameras: 1920x1280 px, f = 1600 px, principal point at the image centre -- the way
  main_vsfm.cpp:232-241 builds K;
* per-view 2-D segments: projection of the front-facing edges, clipped to the image,
  N(0, 0.3 px) endpoint noise, 20 % fragmentation, 50 % endpoint-order swap, sorted by
  length descending (the detector's order, line3D.cc:1853-1868), trimmed or padded with
  clutter to the requested count;
* worldpoint-id lists: random facade points seen by the views, so that the reference's
  own similarity 2*common/(n_i+n_j) (line3D.cc:494) ranks geometric neighbours first.

A scene is a flat little-endian file (`write_scene` / `read_scene`):

    "L3DSCN01" | u32 V | u32 reserved
    per view: u32 id, width, height, nseg, nwp | f64 K[9] R[9] t[3] (row-major)
              | f32 seg[nseg][4] (p1x p1y p2x p2y) | u32 wp[nwp]

This module is pure numpy host code; it feeds tests, bench.py and the C++ harnesses.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field

import numpy as np

IMG_W, IMG_H, FOCAL = 1920, 1280, 1600.0
_REAL_COUNTS: list = []   # number of projected (non-clutter) segments of the views generated so far (diagnostics)
MAGIC = b"L3DSCN01"


@dataclass
class View:
    id: int
    width: int
    height: int
    K: np.ndarray  # (3,3) f64
    R: np.ndarray  # (3,3) f64
    t: np.ndarray  # (3,)  f64
    segs: np.ndarray  # (n,4) f32
    wps: np.ndarray  # (m,)  u32


@dataclass
class Scene:
    views: list = field(default_factory=list)
    name: str = ""

    @property
    def num_views(self):
        return len(self.views)


# --------------------------------------------------------------------------------------
# file format
# --------------------------------------------------------------------------------------
def write_scene(scene: Scene, path: str) -> None:
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<II", len(scene.views), 0))
        for v in scene.views:
            segs = np.ascontiguousarray(v.segs, dtype="<f4").reshape(-1, 4)
            wps = np.ascontiguousarray(v.wps, dtype="<u4").reshape(-1)
            f.write(struct.pack("<IIIII", v.id, v.width, v.height, segs.shape[0], wps.shape[0]))
            f.write(np.ascontiguousarray(v.K, dtype="<f8").tobytes())
            f.write(np.ascontiguousarray(v.R, dtype="<f8").tobytes())
            f.write(np.ascontiguousarray(v.t, dtype="<f8").tobytes())
            f.write(segs.tobytes())
            f.write(wps.tobytes())


def read_scene(path: str) -> Scene:
    with open(path, "rb") as f:
        buf = f.read()
    assert buf[:8] == MAGIC, "not a Line3D scene file"
    nv, _ = struct.unpack_from("<II", buf, 8)
    off = 16
    sc = Scene(name=path)
    for _ in range(nv):
        vid, w, h, ns, nw = struct.unpack_from("<IIIII", buf, off)
        off += 20
        K = np.frombuffer(buf, "<f8", 9, off).reshape(3, 3).copy(); off += 72
        R = np.frombuffer(buf, "<f8", 9, off).reshape(3, 3).copy(); off += 72
        t = np.frombuffer(buf, "<f8", 3, off).copy(); off += 24
        segs = np.frombuffer(buf, "<f4", ns * 4, off).reshape(ns, 4).copy(); off += ns * 16
        wps = np.frombuffer(buf, "<u4", nw, off).copy(); off += nw * 4
        sc.views.append(View(vid, w, h, K, R, t, segs, wps))
    return sc


# --------------------------------------------------------------------------------------
# geometry
# --------------------------------------------------------------------------------------
def _facade_edges(origin, u, v, nu, nv, du, dv, normal, margin=0.6, fill=0.55):
    """Window rectangles on the planar facade origin + a*u + b*v; returns (E,2,3), (E,3) normals."""
    edges = []
    for i in range(nu):
        for j in range(nv):
            a0 = margin + i * du
            b0 = margin + j * dv
            a1 = a0 + fill * du
            b1 = b0 + fill * dv
            c = [origin + a0 * u + b0 * v, origin + a1 * u + b0 * v,
                 origin + a1 * u + b1 * v, origin + a0 * u + b1 * v]
            for k in range(4):
                edges.append((c[k], c[(k + 1) % 4]))
    e = np.array(edges, dtype=np.float64).reshape(-1, 2, 3)
    n = np.tile(np.asarray(normal, dtype=np.float64), (e.shape[0], 1))
    return e, n


def _building(x0, y0, sx, sy, sz, win=2.4):
    """Axis-aligned box [x0,x0+sx]x[y0,y0+sy]x[0,sz] with window grids; z is up."""
    ex = np.array([1.0, 0, 0]); ey = np.array([0, 1.0, 0]); ez = np.array([0, 0, 1.0])
    o = np.array([x0, y0, 0.0])
    E, N = [], []
    faces = [  # origin, u (horizontal), v (vertical), extent_u, extent_v, outward normal
        (o, ex, ez, sx, sz, -ey),
        (o + sy * ey, ex, ez, sx, sz, ey),
        (o, ey, ez, sy, sz, -ex),
        (o + sx * ex, ey, ez, sy, sz, ex),
        (o + sz * ez, ex, ey, sx, sy, ez),  # roof
    ]
    for (fo, u, v, eu, ev, n) in faces:
        # outline of the face
        c = [fo, fo + eu * u, fo + eu * u + ev * v, fo + ev * v]
        for k in range(4):
            E.append(np.array([[c[k], c[(k + 1) % 4]]])); N.append(np.array([n]))
        nu = max(int((eu - 1.2) // win), 0)
        nv = max(int((ev - 1.2) // win), 0)
        if nu and nv:
            e, nn = _facade_edges(fo, u, v, nu, nv, (eu - 1.2) / nu, (ev - 1.2) / nv, n)
            E.append(e); N.append(nn)
    return np.concatenate(E), np.concatenate(N)


def _look_at(C, target, up=(0.0, 0.0, 1.0)):
    """World->camera rotation with +z forward, +x right, +y down; t = -R C."""
    C = np.asarray(C, dtype=np.float64)
    fwd = np.asarray(target, dtype=np.float64) - C
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, np.asarray(up, dtype=np.float64))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    R = np.stack([right, down, fwd])
    return R, -R @ C


def _clip_2d(p, q, w, h):
    """Liang-Barsky clip of segment p->q to [0,w]x[0,h]; returns clipped endpoints or None."""
    d = q - p
    t0, t1 = 0.0, 1.0
    for pi, qi in ((-d[0], p[0]), (d[0], w - p[0]), (-d[1], p[1]), (d[1], h - p[1])):
        if pi == 0.0:
            if qi < 0.0:
                return None
        else:
            r = qi / pi
            if pi < 0.0:
                if r > t1:
                    return None
                t0 = max(t0, r)
            else:
                if r < t0:
                    return None
                t1 = min(t1, r)
    return p + t0 * d, p + t1 * d


def _project_view(edges, normals, K, R, t, rng, n_target, noise=0.3, frag=0.2, min_len=12.0):
    C = -R.T @ t
    mid = edges.mean(axis=1)
    facing = np.einsum("ij,ij->i", normals, C[None, :] - mid) > 0.0
    E = edges[facing]
    near = 0.5
    Pc = (E @ R.T) + t  # (E,2,3) camera coordinates
    A, B = Pc[:, 0, :].copy(), Pc[:, 1, :].copy()
    keep = ~((A[:, 2] < near) & (B[:, 2] < near))
    A, B = A[keep], B[keep]
    # clip against the near plane
    with np.errstate(divide="ignore", invalid="ignore"):
        fa = np.nan_to_num((near - A[:, 2]) / (B[:, 2] - A[:, 2]), nan=0.0, posinf=0.0, neginf=0.0)
        fb = np.nan_to_num((near - B[:, 2]) / (A[:, 2] - B[:, 2]), nan=0.0, posinf=0.0, neginf=0.0)
    ca, cb = A[:, 2] < near, B[:, 2] < near
    A2 = np.where(ca[:, None], A + fa[:, None] * (B - A), A)
    B2 = np.where(cb[:, None], B + fb[:, None] * (A - B), B)
    pa = A2 @ K.T; pa = pa[:, :2] / pa[:, 2:3]
    pb = B2 @ K.T; pb = pb[:, :2] / pb[:, 2:3]
    # Liang-Barsky against [0, W-1] x [0, H-1], all segments at once
    w, h = IMG_W - 1.0, IMG_H - 1.0
    d = pb - pa
    t0 = np.zeros(pa.shape[0]); t1 = np.ones(pa.shape[0]); ok = np.ones(pa.shape[0], dtype=bool)
    for pi, qi in ((-d[:, 0], pa[:, 0]), (d[:, 0], w - pa[:, 0]), (-d[:, 1], pa[:, 1]), (d[:, 1], h - pa[:, 1])):
        par = pi == 0.0
        ok &= ~(par & (qi < 0.0))
        with np.errstate(divide="ignore", invalid="ignore"):
            r = np.where(par, 0.0, qi / np.where(par, 1.0, pi))
        neg = (pi < 0.0)
        pos = (pi > 0.0)
        ok &= ~(neg & (r > t1)); t0 = np.where(neg & ok, np.maximum(t0, r), t0)
        ok &= ~(pos & (r < t0)); t1 = np.where(pos & ok, np.minimum(t1, r), t1)
    ca_ = pa + t0[:, None] * d
    cb_ = pa + t1[:, None] * d
    out = np.concatenate([ca_, cb_], 1)[ok]
    segs = np.array(out, dtype=np.float64).reshape(-1, 4)
    # fragmentation: split 20 % of the segments at a random interior point, small gap
    if segs.shape[0]:
        split = rng.random(segs.shape[0]) < frag
        keep = segs[~split]
        sp = segs[split]
        s = rng.uniform(0.3, 0.7, sp.shape[0])[:, None]
        g = 0.02
        m1 = sp[:, :2] + (s - g) * (sp[:, 2:] - sp[:, :2])
        m2 = sp[:, :2] + (s + g) * (sp[:, 2:] - sp[:, :2])
        segs = np.concatenate([keep, np.concatenate([sp[:, :2], m1], 1), np.concatenate([m2, sp[:, 2:]], 1)])
        segs = segs + rng.normal(0.0, noise, segs.shape)
        ln = np.hypot(segs[:, 0] - segs[:, 2], segs[:, 1] - segs[:, 3])
        segs = segs[ln > min_len]
    _REAL_COUNTS.append(segs.shape[0])
    # trim (keep the longest, like the detector's cap) or pad with clutter
    if segs.shape[0] < n_target:
        k = n_target - segs.shape[0]
        c0 = np.stack([rng.uniform(0, IMG_W - 1, k), rng.uniform(0, IMG_H - 1, k)], 1)
        ang = rng.uniform(0, np.pi, k)
        ll = rng.uniform(min_len, 60.0, k)
        c1 = c0 + np.stack([np.cos(ang), np.sin(ang)], 1) * ll[:, None]
        c1[:, 0] = np.clip(c1[:, 0], 0, IMG_W - 1); c1[:, 1] = np.clip(c1[:, 1], 0, IMG_H - 1)
        segs = np.concatenate([segs, np.concatenate([c0, c1], 1)])
    swap = rng.random(segs.shape[0]) < 0.5
    segs[swap] = segs[swap][:, [2, 3, 0, 1]]
    ln = np.hypot(segs[:, 0] - segs[:, 2], segs[:, 1] - segs[:, 3])
    order = np.argsort(-ln, kind="stable")
    segs = segs[order][:n_target]
    return segs.astype(np.float32)


def _worldpoints(buildings, n, rng):
    """Random points on building facades with their outward normals."""
    pts, nrm = [], []
    per = max(n // max(len(buildings), 1), 1)
    for (x0, y0, sx, sy, sz) in buildings:
        for _ in range(per):
            f = rng.integers(0, 5)
            a, b = rng.random(2)
            if f == 0: p, nn = (x0 + a * sx, y0, b * sz), (0, -1, 0)
            elif f == 1: p, nn = (x0 + a * sx, y0 + sy, b * sz), (0, 1, 0)
            elif f == 2: p, nn = (x0, y0 + a * sy, b * sz), (-1, 0, 0)
            elif f == 3: p, nn = (x0 + sx, y0 + a * sy, b * sz), (1, 0, 0)
            else: p, nn = (x0 + a * sx, y0 + b * sy, sz), (0, 0, 1)
            pts.append(p); nrm.append(nn)
    return np.array(pts, dtype=np.float64), np.array(nrm, dtype=np.float64)


def _visible_wps(pts, nrm, K, R, t):
    C = -R.T @ t
    facing = np.einsum("ij,ij->i", nrm, C[None, :] - pts) > 0.0
    pc = pts @ R.T + t
    z = pc[:, 2]
    ok = facing & (z > 0.5)
    uv = (pc @ K.T)
    with np.errstate(divide="ignore", invalid="ignore"):
        u = uv[:, 0] / uv[:, 2]; v = uv[:, 1] / uv[:, 2]
    ok &= (u >= 0) & (u < IMG_W) & (v >= 0) & (v < IMG_H)
    return np.nonzero(ok)[0].astype(np.uint32)


def _assemble(name, buildings, cams, n_seg, seed, n_wp=4000, win=2.4, per_view_rng=False, view_range=None):
    """per_view_rng: every view draws from its own generator seeded with (seed, view id) instead of one generator
    for the whole scene, so that any range of views (view_range = (first, end)) can be generated on its own -- by
    another process -- and the parts put together give the same scene."""
    assert view_range is None or per_view_rng
    rng = np.random.default_rng(seed)
    E, N = [], []
    for (x0, y0, sx, sy, sz) in buildings:
        e, n = _building(x0, y0, sx, sy, sz, win=win)
        E.append(e); N.append(n)
    E = np.concatenate(E); N = np.concatenate(N)
    wp, wn = _worldpoints(buildings, n_wp, rng)
    K = np.array([[FOCAL, 0, IMG_W / 2.0], [0, FOCAL, IMG_H / 2.0], [0, 0, 1.0]])
    sc = Scene(name=name)
    for vid, (C, target) in enumerate(cams):
        if view_range is not None and not (view_range[0] <= vid < view_range[1]):
            continue
        if per_view_rng:
            rng = np.random.default_rng([seed, vid])
        R, t = _look_at(C, target)
        lo, hi = int(round(n_seg * 0.95)), int(round(n_seg * 1.05))
        n_t = int(rng.integers(lo, hi + 1))
        segs = _project_view(E, N, K, R, t, rng, n_t)
        wps = _visible_wps(wp, wn, K, R, t)
        if wps.size == 0:
            wps = np.array([0], dtype=np.uint32)
        sc.views.append(View(vid, IMG_W, IMG_H, K.copy(), R, t, segs, wps))
    return sc


# --------------------------------------------------------------------------------------
# named configurations (BASELINE.json "configs")
# --------------------------------------------------------------------------------------
def wireframe_building(n_views=24, n_seg=800, seed=1001, radius=34.0, height=9.0):
    """C1-shaped: one building, a ring of cameras around it."""
    b = [(-11.0, -7.0, 22.0, 14.0, 16.0)]
    cams = []
    for i in range(n_views):
        a = 2.0 * np.pi * i / n_views
        C = (radius * np.cos(a), radius * np.sin(a), height + 2.0 * np.sin(3 * a))
        cams.append((C, (0.0, 0.0, 7.0)))
    return _assemble("wireframe_building", b, cams, n_seg, seed, n_wp=3000, win=1.9)


def urban(n_views=200, n_seg=2000, seed=1002, blocks=(4, 3), per_view_rng=False, view_range=None):
    """C2/C4-shaped: city blocks, cameras walking the streets between them."""
    rng = np.random.default_rng(seed + 7)
    bx, by = blocks
    pitch_x, pitch_y = 46.0, 40.0
    buildings = []
    for i in range(bx):
        for j in range(by):
            sx = rng.uniform(24.0, 30.0); sy = rng.uniform(20.0, 26.0); sz = rng.uniform(14.0, 30.0)
            buildings.append((i * pitch_x, j * pitch_y, sx, sy, sz))
    # street centre lines run between block rows: a serpentine trajectory
    path = []
    xs0, xs1 = 4.0, (bx - 1) * pitch_x + 12.0     # keep the look-ahead target inside the blocks
    for j in range(by + 1):
        y = j * pitch_y - 8.0
        seg = np.linspace(xs0, xs1, 200)
        if j % 2:
            seg = seg[::-1]
        for x in seg:
            path.append((x, y, j))
    path = np.array(path)
    idx = np.linspace(0, len(path) - 1, n_views).round().astype(int)
    cams = []
    for k, i in enumerate(idx):
        x, y, street = path[i]
        side = 1.0 if (k % 2 == 0) else -1.0
        if street == 0:
            side = 1.0          # outer streets: only one side has buildings
        elif street == by:
            side = -1.0
        j = min(i + 5, len(path) - 1)
        dx, dy = path[j][:2] - path[i][:2]
        if path[j][2] != street:
            dx, dy = path[i][:2] - path[max(i - 5, 0)][:2]
        nrm = np.hypot(dx, dy)
        if nrm < 1e-9:
            dx, dy = 1.0, 0.0; nrm = 1.0
        dx, dy = dx / nrm, dy / nrm
        # look obliquely at the facades on alternating sides of the street
        # street direction alternates (serpentine): "side" is meant in world +y / -y terms
        sgn = 1.0 if dx >= 0 else -1.0
        tgt = (x + 12.0 * dx - sgn * side * 14.0 * dy, y + 12.0 * dy + sgn * side * 14.0 * dx, 9.0)
        cams.append(((x, y, 1.7 + 0.2 * np.sin(k)), tgt))
    return _assemble("urban", buildings, cams, n_seg, seed, n_wp=20000, win=1.5, per_view_rng=per_view_rng, view_range=view_range)


def dense_dome(n_views=300, n_seg=3000, seed=1005, radius=60.0):
    """C5-shaped: cameras on a dome, every pair of views co-visible."""
    b = [(-16.0, -10.0, 32.0, 20.0, 22.0), (22.0, -8.0, 14.0, 16.0, 12.0), (-36.0, -6.0, 14.0, 12.0, 10.0)]
    cams = []
    ga = np.pi * (3.0 - np.sqrt(5.0))
    for i in range(n_views):
        zf = 0.12 + 0.8 * (i + 0.5) / n_views
        r = np.sqrt(1.0 - zf * zf)
        a = ga * i
        cams.append(((radius * r * np.cos(a), radius * r * np.sin(a), radius * zf), (0.0, 0.0, 8.0)))
    return _assemble("dense_dome", b, cams, n_seg, seed, n_wp=6000, win=1.3)


def tiny(n_views=8, n_seg=120, seed=7, step_deg=11.0):
    """Small arc scene for unit tests (seconds on the CPU oracle)."""
    b = [(-6.0, -4.0, 12.0, 8.0, 9.0)]
    cams = []
    for i in range(n_views):
        a = np.deg2rad(25.0 + step_deg * i)
        cams.append(((22.0 * np.cos(a), 22.0 * np.sin(a), 6.0 + np.sin(2 * a)), (0.0, 0.0, 4.0)))
    return _assemble("tiny", b, cams, n_seg, seed, n_wp=800, win=2.2)


CONFIGS = {
    "C1": dict(fn=wireframe_building, kw=dict(n_views=24, n_seg=800, seed=1001), k=12, diffusion=False),
    "C2": dict(fn=urban, kw=dict(n_views=200, n_seg=2000, seed=1002), k=12, diffusion=False),
    "C3": dict(fn=urban, kw=dict(n_views=200, n_seg=2000, seed=1002), k=12, diffusion=True),
    "C4": dict(fn=urban, kw=dict(n_views=1000, n_seg=4000, seed=1004, blocks=(8, 6), per_view_rng=True), k=20, diffusion=False),
    "C5": dict(fn=dense_dome, kw=dict(n_views=300, n_seg=3000, seed=1005), k=-1, diffusion=False),
}


def multi_district(config: str, n: int, spacing: float = 260.0) -> Scene:
    """n independent districts of the named configuration in one scene: district d is generated with its own
    seed, moved by a world-space offset (cameras only; image measurements are unchanged by a rigid move of the
    whole district), its view ids are offset by d * V and its worldpoint ids by d * 10^6, so the districts share
    no worldpoints and form separate connected components of the neighbour graph."""
    c = CONFIGS[config]
    out = Scene(name=f"{config}x{n}")
    for d in range(n):
        kw = dict(c["kw"]); kw["seed"] = kw["seed"] + 100 * d
        sc = c["fn"](**kw)
        T = np.array([spacing * (d % 4), spacing * (d // 4), 0.0])
        nv = len(sc.views)
        for v in sc.views:
            t = v.t - v.R @ T                      # C' = C + T  ->  t' = -R C'
            out.views.append(View(v.id + d * nv, v.width, v.height, v.K, v.R, t, v.segs, (v.wps.astype(np.uint64) + d * 1000000).astype(np.uint32)))
    return out


def make(config: str) -> Scene:
    c = CONFIGS[config]
    return c["fn"](**c["kw"])


def _make_part(arg):
    config, first, end = arg
    c = CONFIGS[config]
    return c["fn"](view_range=(first, end), **c["kw"]).views


def make_parallel(config: str, processes: int | None = None) -> Scene:
    """A configuration whose views have their own generators (per_view_rng), made by several processes at once."""
    import multiprocessing as mp
    import os
    c = CONFIGS[config]
    if not c["kw"].get("per_view_rng"):
        return make(config)
    n = c["kw"]["n_views"]
    processes = processes or min(os.cpu_count() or 1, 32)
    step = max((n + 4 * processes - 1) // (4 * processes), 1)
    parts = [(config, a, min(a + step, n)) for a in range(0, n, step)]
    with mp.get_context("fork").Pool(processes) as pool:
        views = [v for part in pool.map(_make_part, parts) for v in part]
    return Scene(views=views, name=config)
