import sys, os
sys.path.insert(0, '/root/repo')
os.environ['L3D_PROFILE'] = '1'
import bench
from line3d_b200 import host
p = bench.scene_file('C2')
s = host.Session(p, neighbors=12, collinearity=False)
for _ in range(2):
    n, tm, st = s.compute()
    print(n, tm, {k: round(v, 4) for k, v in st.items() if k.startswith('t_')})
s.close()
s = host.Session(p, neighbors=12, collinearity=True)
s.compute()
n, tm, st = s.compute()
print('collin', n, tm, {k: round(v, 4) for k, v in st.items() if k.startswith('t_')}, st['affinity_pairs'], st['edges'])
