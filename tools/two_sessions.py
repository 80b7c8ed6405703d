"""Feasibility probe: two independent sessions on one GPU from two host threads.  If their steps overlap well (one's
pair tests filling the SMs the other's per-view chain leaves idle), two steps take much less than twice one step."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from line3d_b200 import host

p = bench.scene_file('C2')
ss = [host.Session(p, neighbors=12, collinearity=False) for _ in range(2)]
for s in ss:
    for _ in range(2):
        s.compute(measure_only=True)
def run(s, n, out):
    t0 = time.perf_counter()
    for _ in range(n):
        _, _, st = s.compute(measure_only=True)
    out.append((time.perf_counter() - t0, st['t_matching']))
N = 4
o = []; run(ss[0], N, o); print('one session : %.1f ms per compute call, matching %.1f ms' % (1e3 * o[0][0] / N, 1e3 * o[0][1]))
o = []; th = [threading.Thread(target=run, args=(s, N, o)) for s in ss]
t0 = time.perf_counter(); [t.start() for t in th]; [t.join() for t in th]; w = time.perf_counter() - t0
print('two sessions: %.1f ms wall per pair of compute calls; matching %.1f / %.1f ms' % (1e3 * w / N, 1e3 * o[0][1], 1e3 * o[1][1]))
