"""GPU box: registration time vs index cell size on the C2 workload."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
import bench

M = int(os.environ.get("MAP", 5_000_000)); N = int(os.environ.get("SCAN", 131072))
d = synth.make_c2(M=M, N=N, n_poses=4)
ref4 = synth.feat4(d["map_xyz"])
for cell in [float(c) for c in os.environ.get("CELLS", "0,0.12,0.16,0.2,0.25,0.3,0.4").split(",")]:
    kw = bench.icp_kwargs(); kw["cell_size"] = cell
    icp = B.ICP(B.default_icp_params(**kw))
    icp.init_reference(ref4, d["map_normals"])
    tot = []; it = []; tq = []; ph = np.zeros(6)
    for rep in range(2):
        for s in d["scans"]:
            icp.upload_reading(synth.feat4(s["xyz"]))
            icp.flush_l2()
            r = icp.run_resident(s["T_init"])
            if rep:
                tot.append(r["device_ms"]); it.append(r["iterations"]); tq.append(r["points_touched"] / (N * r["iterations"]))
                pt = icp.phase_times(); ph += pt.sum(0)
    print("cell %.3f: device_ms %s iters %s cand/q/it %.1f | phase sums us match %.0f select %.0f lin %.0f wait %.0f solve %.0f" % (
        cell, " ".join("%.3f" % t for t in tot), it, np.mean(tq), *(ph[:5] / len(tot))), flush=True)
    icp.close()
