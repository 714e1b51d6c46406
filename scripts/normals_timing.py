"""Where does the per-scan normal estimation spend its time? raw LiDAR scan (non-uniform density) vs a uniform crop."""
import time
import numpy as np
from open3d_slam_private_b200 import capi as B, synth

sc = synth.Scene()
scan, _ = sc.scan((5.0, 2.0, 0.3), n_points=131072, seed=1)
mp, _ = sc.sample_map(2_000_000)
crop = mp[np.argsort(np.linalg.norm(mp[:, :2] - np.array([5.0, 2.0]), axis=1))[:131072]]


def t(fn, n=3):
    fn()
    t0 = time.perf_counter()
    for _ in range(n):
        r = fn()
    return (time.perf_counter() - t0) / n * 1e3, r


for name, pts in (("raw scan", scan), ("uniform crop", crop)):
    f4 = synth.feat4(pts)
    for tpc in (0.0, 1.0, 8.0):
        ms, idx = t(lambda: B.Index(f4, target_per_cell=tpc))
        info = idx.info()
        msq, (ids, d) = t(lambda: idx.find_closests(f4, knn=20))
        ms1, _ = t(lambda: idx.find_closests(f4, knn=1))
        print("%-13s tpc %.0f: build wall %.2f ms (device %.2f) cell %.3f occ %d  knn20 wall %.2f ms  knn1 %.2f ms  visits/query %.0f"
              % (name, tpc, ms, info["build_ms"], info["cell_size"], info["occupied_cells"], msq, ms1,
                 idx.visit_counter / (4 * 2 * f4.shape[0])))
        idx.close()
    msn, _ = t(lambda: B.surface_normals(f4, knn=20))
    msn10, _ = t(lambda: B.surface_normals(f4, knn=10))
    print("%-13s surface_normals k20 wall %.2f ms, k10 %.2f ms" % (name, msn, msn10))

    idx = B.Index(f4, target_per_cell=8.0)
    for k in (10, 20):
        def again():
            idx.update(f4, target_per_cell=8.0)
            return idx.self_normals(knn=k)
        ms, out = t(again, 5)
        ms2, _ = t(lambda: idx.self_normals(knn=k), 5)
        ref = B.surface_normals(f4, knn=k)["normals"]
        print("%-13s kept index: update + normals k%d wall %.2f ms; normals only %.2f ms; equal to stateless: %s"
              % (name, k, ms, ms2, bool((out["normals"] == ref).all())))
    idx.close()
