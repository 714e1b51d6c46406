"""GPU box: k = 20 normals on the C2 map and scans: device ms and how many points fell back to the exact warp path."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
d = synth.make_c2(M=5_000_000, N=131072, n_poses=3)
ref4 = synth.feat4(d["map_xyz"])
idx = B.Index(ref4)
for knn in (10, 20):
    for rep in range(2):
        idx.self_normals(knn=knn)
        print("map 5M knn %d: device %.3f ms, fallback %d" % (knn, idx.last_device_ms(), idx.last_normals_fallback()))
idx.close()
s = None
for sc in d["scans"]:
    s4 = synth.feat4(sc["xyz"])
    if s is None: s = B.Index(s4)
    else: s.update(s4)
    for knn in (10, 20):
        s.self_normals(knn=knn)
        print("scan knn %d: device %.3f ms, fallback %d, index %s" % (knn, s.last_device_ms(), s.last_normals_fallback(), s.info()))
t0 = time.time(); s.self_normals(knn=10, smooth=1); print("scan smooth knn 10: device %.3f ms wall %.3f s" % (s.last_device_ms(), time.time() - t0))
