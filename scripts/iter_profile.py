"""GPU box: per-ITERATION phase times of the persistent registration kernel for a few C2 poses
(CTA 0's stamps; with B200_ICP_DEBUG=1 also the slowest / mean CTA match time)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
import bench

N = 131072
poses = [int(x) for x in os.environ.get("POSES", "0,3,7").split(",")]
d = synth.make_c2(M=5_000_000, N=N, n_poses=max(poses) + 1)
icp = B.ICP(B.default_icp_params(**bench.icp_kwargs()))
icp.init_reference(synth.feat4(d["map_xyz"]), d["map_normals"])
dbg = os.environ.get("B200_ICP_DEBUG") == "1"
for i in poses:
    s = d["scans"][i]
    icp.upload_reading(synth.feat4(s["xyz"]))
    for rep in range(2):
        icp.flush_l2()
        r = icp.run_resident(s["T_init"])
    pt = icp.phase_times()
    st = icp.debug_stats()
    print("pose %d: %d iterations, %.3f ms" % (i, r["iterations"], r["device_ms"]))
    print(" it path ncand | match select linear wait solve total | sub: pre solve upd chk" + (" | rows/q cands/q matchmax matchmean" if dbg else ""))
    for it in range(len(pt)):
        extra = ""
        if dbg:
            extra = " | %5.1f %5.1f %4.0f %4.0f" % (st[it, 0] / N, st[it, 1] / N, st[it, 2], st[it, 3])
        print(" %2d %4d %5d | %s | %s%s" % (it, int(st[it, 8]), int(st[it, 9]), " ".join("%5.1f" % v for v in pt[it]),
                                          " ".join("%4.1f" % v for v in st[it, 4:8]), extra))
