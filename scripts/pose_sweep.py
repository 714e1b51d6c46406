"""GPU box: per-pose cost of the persistent registration kernel over the 32 C2 poses (which poses are slow, and why)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
import bench

N = 131072
d = synth.make_c2(M=5_000_000, N=N, n_poses=int(os.environ.get("POSES", 32)))
icp = B.ICP(B.default_icp_params(**bench.icp_kwargs()))
icp.init_reference(synth.feat4(d["map_xyz"]), d["map_normals"])
dbg = os.environ.get("B200_ICP_DEBUG") == "1"
print("pose iters dev_ms | it0: match select linear wait solve total | rest_us/iter | unmatched%% | rows/q cands/q matchmax matchmean (it0)")
for i, s in enumerate(d["scans"]):
    r4 = synth.feat4(s["xyz"])
    icp.upload_reading(r4)
    icp.flush_l2()
    r = icp.run_resident(s["T_init"])
    pt = icp.phase_times()
    ids, dd, w = icp.matches()
    extra = ""
    if dbg:
        st = icp.debug_stats()
        extra = "| %.1f %.1f %.0f %.0f" % (st[0, 0] / N, st[0, 1] / N, st[0, 2], st[0, 3])
    rest = (pt[1:, 5].sum() / max(1, len(pt) - 1))
    print("%2d %2d %.3f | %s | %.1f | %.1f %s" % (i, r["iterations"], r["device_ms"], " ".join("%.0f" % v for v in pt[0]), rest,
                                            100.0 * (ids < 0).mean(), extra))
