"""GPU box: phase breakdown of the persistent kernel on the small golden pair (car_cloud401 -> car_cloud400)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from open3d_slam_private_b200 import capi as B
from oracle import oracle_py as O
g = os.path.join(ROOT, "tests", "golden")
c400 = np.load(os.path.join(g, "car_cloud400.npz")); c401 = np.load(os.path.join(g, "car_cloud401.npz"))
ref4, refn, read4 = O.feat4(c400["xyz"]), c400["normals"], O.feat4(c401["xyz"])
print("points", ref4.shape[0], read4.shape[0])
icp = B.ICP(B.default_icp_params(maxDist=0.5, trimmed_ratio=0.9, max_iterations=30, min_diff_trans=0.008))
icp.init_reference(ref4, refn)
for rep in range(2):
    r = icp.compute(read4)
    print("iters %d device_ms %.3f grid %d" % (r["iterations"], r["device_ms"], r["grid_ctas"]))
pt = icp.phase_times()
print("   match  select  linear    wait   solve   total (us)")
for row in pt[:6]:
    print("  " + " ".join("%7.1f" % v for v in row))
print(icp.index_info() if hasattr(icp, "index_info") else "")
