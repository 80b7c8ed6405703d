"""One matching pass over C2 (no clustering) -- the command the ncu captures under profiles/ are taken on.
With --timing: two extra passes with per-stage CUDA events, printing the pair-test kernels' own durations."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from line3d_b200 import host
s = host.Session(bench.scene_file('C2'), neighbors=12, collinearity=False)
n, tm, st = s.compute(measure_only=True)
print(n, tm, st['pair_tests'])
if "--timing" in sys.argv:
    s.compute(measure_only=True, flush_l2=True, stage_timing=True)
    for _ in range(3):
        _, _, st = s.compute(measure_only=True, flush_l2=True, stage_timing=True)
        print("variant", os.environ.get("L3D_P1_VARIANT", "0"), "k_pretest %.3f ms  k_pretest2 %.3f ms  k_exact %.3f ms  device %.2f ms" %
              (1e3 * st["t_k_pretest"], 1e3 * st["t_k_pretest2"], 1e3 * st["t_k_exact"], 1e3 * st["t_matching_device"]))
s.close()
