"""One matching pass over C2 (no clustering) -- the command the ncu captures under profiles/ are taken on."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from line3d_b200 import host
s = host.Session(bench.scene_file('C2'), neighbors=12, collinearity=False)
n, tm, st = s.compute(measure_only=True)
print(n, tm, st['pair_tests'])
s.close()
