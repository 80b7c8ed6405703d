"""How sequential is the per-view work really?  Longest chain of hand-over dependencies in a scene's schedule."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from line3d_b200 import host
cfg = sys.argv[1] if len(sys.argv) > 1 else 'C2'
s = host.Session(bench.scene_file(cfg), neighbors=12, collinearity=False)
s.compute(measure_only=True, keep_debug=True)
vn = s.dump()["visual_neighbors"].astype(int)
s.close()
nb = {}
for a, b in vn: nb.setdefault(a, set()).add(b)
ids = sorted(nb)
# v depends on u < v if u hands its matches over to v: u matched v (v in nb[u]) and v has u as neighbour
level = {}
for v in ids:
    deps = [u for u in ids if u < v and v in nb.get(u, ()) and u in nb.get(v, ())]
    level[v] = 1 + max([level[u] for u in deps], default=0)
depth = max(level.values())
hist = np.bincount(list(level.values()))
print(cfg, "views", len(ids), "longest dependency chain", depth, "widest level", hist.max(), "mean width", len(ids) / depth)
