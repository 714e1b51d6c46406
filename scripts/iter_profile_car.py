"""GPU box: per-iteration phase times of the persistent kernel on the reference's own car_cloud400/401 pair
(25 k points, strongly non-uniform density) - the smoke() workload."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from open3d_slam_private_b200 import capi as B

def feat4(x):
    return np.concatenate([x.astype(np.float32), np.ones((len(x), 1), np.float32)], 1)
g = os.path.join(ROOT, "tests", "golden")
c400 = np.load(os.path.join(g, "car_cloud400.npz")); c401 = np.load(os.path.join(g, "car_cloud401.npz"))
ref4, refn, read4 = feat4(c400["xyz"]), c400["normals"], feat4(c401["xyz"])
N = len(read4)
icp = B.ICP(B.default_icp_params(maxDist=0.5, trimmed_ratio=0.9, max_iterations=30, min_diff_trans=0.008))
icp.init_reference(ref4, refn)
icp.upload_reading(read4)
for rep in range(3):
    r = icp.run_resident(np.eye(4, dtype=np.float32))
pt = icp.phase_times(); st = icp.debug_stats()
dbg = os.environ.get("B200_ICP_DEBUG") == "1"
print("N=%d M=%d: %d iterations, %.3f ms, grid %s" % (N, len(ref4), r["iterations"], r["device_ms"], r.get("grid")))
print(" it path ncand | match select linear wait solve total | sub: pre solve upd chk" + (" | rows/q cands/q matchmax matchmean" if dbg else ""))
for it in range(len(pt)):
    extra = " | %5.1f %5.1f %4.0f %4.0f" % (st[it, 0] / N, st[it, 1] / N, st[it, 2], st[it, 3]) if dbg else ""
    print(" %2d %4d %5d | %s | %s%s" % (it, int(st[it, 8]), int(st[it, 9]), " ".join("%5.1f" % v for v in pt[it]),
                                      " ".join("%4.1f" % v for v in st[it, 4:8]), extra))
