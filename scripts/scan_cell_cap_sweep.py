"""GPU box: a raw 131 072-point scan as its own target: index build time and k = 20 / k = 1 self-search against the cap on
the dense cell table (b200_index_opts.max_cells; default max(2^22, 64 per point))."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
sc = synth.Scene()
scan, _ = sc.scan((5.0, 2.0, 0.3), n_points=131072, seed=1)
f4 = synth.feat4(scan)
def t(fn, n=5):
    fn()
    t0 = time.perf_counter()
    for _ in range(n):
        r = fn()
    return (time.perf_counter() - t0) / n * 1e3, r
for cap in (0, 1 << 24, 1 << 25, 1 << 26):
    idx = B.Index(f4, max_cells=cap)
    bms = []
    for _ in range(5):
        idx.update(f4, max_cells=cap); bms.append(idx.info()["build_ms"])
    info = idx.info()
    k20, _ = t(lambda: idx.find_closests(f4, knn=20))
    sn, _ = t(lambda: idx.self_normals(knn=20))
    print("max_cells %9d: cell %.3f occupied %d build device %.3f ms, knn20 wall %.2f ms, self normals k20 wall %.2f ms, visits/query %.0f"
          % (cap, info["cell_size"], info["occupied_cells"], np.median(bms), k20, sn, idx.visit_counter / (6 * f4.shape[0])), flush=True)
    idx.close()
