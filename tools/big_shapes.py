"""Robustness on the shapes of the large configs (slices of C4 and C5): runs, memory, timings."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from line3d_b200 import host, scenes

def run(name, sc, **params):
    p = f"/tmp/{name}.scn"
    scenes.write_scene(sc, p)
    t = time.time()
    s = host.Session(p, **params)
    t_add = time.time() - t
    for mo in (True, False):
        n, tm, st = s.compute(measure_only=mo)
        print(name, "measure_only" if mo else "full", "lines", n, "add %.2fs" % t_add, "compute %.3fs" % tm["compute3Dmodel"],
              {k: (round(v, 4) if k.startswith("t_") else int(v)) for k, v in st.items() if not k.startswith("_")}, flush=True)
    s.close()

which = sys.argv[1:] or ["c4", "c5"]
if "c4" in which:   # C4 shape: ~4000 segments/view, k = 20; 60-view slice of the trajectory
    run("c4slice", scenes.urban(n_views=60, n_seg=4000, seed=1004, blocks=(2, 1)), neighbors=20)
if "c5" in which:   # C5 shape: ~3000 segments/view, all co-visible views are neighbours; 40 views on the dome
    run("c5slice", scenes.dense_dome(n_views=40, n_seg=3000, seed=1005), neighbors=-1)
