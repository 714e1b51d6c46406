"""GPU box: per-iteration phase breakdown of the persistent registration kernel on the C2 workload."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
import bench

M = int(os.environ.get("MAP", 5_000_000)); N = int(os.environ.get("SCAN", 131072))
d = synth.make_c2(M=M, N=N, n_poses=int(os.environ.get("POSES", 2)))
kw = bench.icp_kwargs(); kw["cell_size"] = float(os.environ.get("CELL", 0))
icp = B.ICP(B.default_icp_params(**kw))
icp.init_reference(synth.feat4(d["map_xyz"]), d["map_normals"])
print(B.Index.info.__doc__ or "", end="")
for rep in range(2):
    for s in d["scans"]:
        icp.upload_reading(synth.feat4(s["xyz"]))
        icp.flush_l2()
        r = icp.run_resident(s["T_init"])
        print("iters %d device_ms %.3f total_ms %.3f touched/query/iter %.1f" % (
            r["iterations"], r["device_ms"], r["total_ms"], r["points_touched"] / (N * r["iterations"])))
        pt = icp.phase_times()
        print("   match  select  linear    wait   solve   total (us)")
        for row in pt:
            print("  " + " ".join("%7.1f" % v for v in row))
        print("  sum " + " ".join("%7.1f" % v for v in pt.sum(0)))
        ds = icp.debug_stats()
        print("  shape/cands: " + " ".join("%d/%d" % (r[8], r[9]) for r in ds))
        print("  limit: " + " ".join("%.5f" % r[10] for r in ds) + " | delta: " + " ".join("%.4f" % r[11] for r in ds))
        if os.environ.get("B200_ICP_DEBUG") == "1":
            w = icp.debug_warps().astype(np.float64)
            if len(w):
                o = np.argsort(w[:, 3])
                print("   per-warp it0: ns pct 50/90/99/max %s | rows 50/99/max %s | cands 50/99/max %s" % (
                    np.percentile(w[:, 3], [50, 90, 99, 100]).astype(int), np.percentile(w[:, 0], [50, 99, 100]).astype(int),
                    np.percentile(w[:, 1], [50, 99, 100]).astype(int)))
                print("   corr(ns, rows) %.2f corr(ns, cands) %.2f ; slowest 5 warps (rows, cands, passes, ns): %s" % (
                    np.corrcoef(w[:, 3], w[:, 0])[0, 1], np.corrcoef(w[:, 3], w[:, 1])[0, 1], w[o[-5:]].astype(int).tolist()))
                print("   fastest 3: %s" % w[o[:3]].astype(int).tolist())
            for row in icp.debug_stats():
                print("   rows/q %.1f cands/q %.1f match max %.1f us mean %.1f us | reduce %.1f solve %.1f update %.1f check %.1f us" % (
                    row[0] / N, row[1] / N, row[2], row[3], row[4], row[5], row[6], row[7]))
np.save(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "H_log.npy"), icp.log()["H"])
