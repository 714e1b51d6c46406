"""GPU box: device time of the map index build (bbox -> histogram -> scan -> scatter) against its HBM roofline
(~116 B/point algorithmic)."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth

peak = 6550.1
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
for M in (5_000_000, 25_000_000):
    d = synth.make_c2(M=M, N=1024, n_poses=1)
    ref4 = synth.feat4(d["map_xyz"])
    idx = B.Index(ref4, d["map_normals"])
    best = 1e9
    for rep in range(5):
        t0 = time.perf_counter()
        idx.update(ref4, d["map_normals"])
        wall = time.perf_counter() - t0
        info = idx.info()
        best = min(best, info["build_ms"])
    gbs = M * 116 / (best * 1e-3) / 1e9
    print(json.dumps({"map_points": M, "build_ms_device_best_of_5": best, "last_wall_ms_incl_h2d": wall * 1e3,
                      "algorithmic_GBps": gbs, "frac_of_hbm_peak": gbs / peak, "info": {k: (v if not hasattr(v, "tolist") else v.tolist()) for k, v in info.items()}}))
