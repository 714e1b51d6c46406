"""GPU box: the Open3D-mode loop (o3d_match_lin_kernel) with the three query -> lane maps (B200_O3D_QMAP = 0 coherent,
1 scattered, 2 scattered until the RMSE is below 0.4 cells) on C3 (GICP, 131 072 vs 5 M), a C4 pair (200 k vs 200 k,
point-to-plane) and a C5 slab (131 072 vs 25 M)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth

def run(name, make, scans, reps=3):
    for qmap in (0, 1, 2):
        os.environ["B200_O3D_QMAP"] = str(qmap)
        h = make()
        ms, it = [], []
        for rep in range(reps):
            for src4, srcn, init in scans:
                r = h.register(src4, srcn, init=init)
                if rep:
                    ms.append(r["total_ms"]); it.append(r["iterations"])
        print("%s qmap %d: %.3f ms per registration (median %.3f), updates %.1f, per pass %.3f ms" % (
            name, qmap, np.mean(ms), np.median(ms), np.mean(it), np.mean(ms) / (np.mean(it) + 1)), flush=True)
        h.close()

which = os.environ.get("WHICH", "c3,c4,c5").split(",")
if "c3" in which:
    d = synth.make_c2(M=5_000_000, N=131072, n_poses=4)
    ref4, refn = synth.feat4(d["map_xyz"]), d["map_normals"]
    est = B.O3dICP(kind=B.O3D_GENERALIZED)
    scans = []
    for s in d["scans"]:
        s4 = synth.feat4(s["xyz"])
        scans.append((s4, est.estimate_normals(s4, knn=20, radius=1.0).astype(np.float32), s["T_init"].astype(np.float64)))
    est.close()
    def mk():
        h = B.O3dICP(kind=B.O3D_GENERALIZED, max_correspondence_distance=0.5, max_iteration=30)
        h.set_target(ref4, refn)
        return h
    run("c3 gicp", mk, scans)
if "c4" in which:
    p = synth.make_c4(n_distinct=1, N=200_000)[0]
    tgt4, src4 = synth.feat4(p["tgt_xyz"]), synth.feat4(p["src_xyz"])
    def mk():
        h = B.O3dICP(kind=B.O3D_POINT_TO_PLANE, max_correspondence_distance=0.5, max_iteration=30)
        h.set_target_estimate_normals(tgt4, knn=10, max_dist=2.0)
        return h
    run("c4 pair", mk, [(src4, None, p["T_init"].astype(np.float64))], reps=6)
if "c5" in which:
    street = synth.TiledStreet(1)
    xyz, nrm = street.tile_points(0, int(os.environ.get("TILE", 25_000_000)))
    tgt4 = synth.feat4(xyz)
    scans = [(synth.feat4(s["xyz"]), None, s["T_init"].astype(np.float64)) for s in synth.make_c5_scans(street, 4, 131072)]
    def mk():
        h = B.O3dICP(kind=B.O3D_POINT_TO_PLANE, max_correspondence_distance=0.5, max_iteration=30)
        h.set_target(tgt4, nrm)
        return h
    run("c5 slab", mk, scans)
