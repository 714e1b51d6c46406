"""GPU box: Open3D-mode registration time against the index cell size (0 = automatic, 3 points per occupied cell) on a
C4 pair (200 k vs 200 k) and a C5 slab (131 072 vs 25 M)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
def run(name, tgt4, tn, scans, cells):
    for cell in cells:
        h = B.O3dICP(kind=0, max_correspondence_distance=0.5, max_iteration=30, cell_size=cell)
        h.set_target(tgt4, tn)
        ms, dev, it = [], [], []
        for rep in range(4):
            for s4, Ti in scans:
                r = h.register(s4, init=Ti)
                if rep:
                    ms.append(r["total_ms"]); dev.append(r["device_ms"]); it.append(r["iterations"])
        print("%s cell %.3f: %.3f ms total, %.3f ms device (median), updates %.1f" % (name, cell, np.median(ms), np.median(dev), np.mean(it)), flush=True)
        h.close()
p = synth.make_c4(n_distinct=1, N=200_000)[0]
tgt4, src4 = synth.feat4(p["tgt_xyz"]), synth.feat4(p["src_xyz"])
est = B.O3dICP(kind=0)
tn = est.estimate_normals(tgt4, knn=10, radius=2.0).astype(np.float32)
est.close()
run("c4 pair", tgt4, tn, [(src4, p["T_init"].astype(np.float64))], (0.0, 0.08, 0.11, 0.14, 0.18, 0.24))
if os.environ.get("C5", "1") == "1":
    street = synth.TiledStreet(1)
    xyz, nrm = street.tile_points(0, 25_000_000)
    scans = [(synth.feat4(s["xyz"]), s["T_init"].astype(np.float64)) for s in synth.make_c5_scans(street, 4, 131072)]
    run("c5 slab", synth.feat4(xyz), nrm, scans, (0.0, 0.07, 0.09, 0.12, 0.16))
