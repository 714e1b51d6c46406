"""GPU box: C5-sized slab (25 M points of one tile) on one device, registration time against the index cell size."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
street = synth.TiledStreet(1)
xyz, nrm = street.tile_points(0, int(os.environ.get("TILE", 25_000_000)))
tgt4 = synth.feat4(xyz)
scans = synth.make_c5_scans(street, 4, 131072)
for cell in (0.0, 0.08, 0.12, 0.18, 0.25):
    h = B.O3dICP(kind=0, max_correspondence_distance=0.5, max_iteration=30, cell_size=cell)
    t0 = time.time(); h.set_target(tgt4, nrm); tb = time.time() - t0
    ms, it = [], []
    for rep in range(3):
        for s in scans:
            r = h.register(synth.feat4(s["xyz"]), init=s["T_init"].astype(np.float64))
            if rep: ms.append(r["total_ms"]); it.append(r["iterations"])
    print("cell %.3f: set_target %.2f s, registration %.3f ms (updates %.1f), per pass %.3f ms" % (
        cell, tb, np.mean(ms), np.mean(it), np.mean(ms) / (np.mean(it) + 1)), flush=True)
    h.close()
