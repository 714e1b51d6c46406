"""GPU box: device time of the map index build on the same 5 M points in three input orders: shuffled (what the synthetic
generator produces and bench.py times), sorted along x (what a map grown scan by scan along a street roughly looks like),
and already cell-sorted (the order of a rebuilt patch) — the scatter's 16-byte writes coalesce only when neighbours in the
input are neighbours in space."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
peak = 6550.1
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
M = 5_000_000
d = synth.make_c2(M=M, N=1024, n_poses=1)
xyz, nrm = d["map_xyz"], d["map_normals"]
orders = {"shuffled (as generated)": np.arange(M), "sorted along x": np.argsort(xyz[:, 0], kind="stable")}
idx = B.Index(synth.feat4(xyz), nrm)
info = idx.info()
cell = info["cell_size"]
key = (np.floor((xyz - xyz.min(0)) / cell)).astype(np.int64)
dims = key.max(0) + 1
orders["cell-sorted (x fastest)"] = np.argsort(key[:, 0] + dims[0] * (key[:, 1] + dims[1] * key[:, 2]), kind="stable")
for name, o in orders.items():
    r4, n3 = synth.feat4(xyz[o]), np.ascontiguousarray(nrm[o])
    best = 1e9
    for rep in range(5):
        idx.update(r4, n3)
        best = min(best, idx.info()["build_ms"])
    gbs = M * 116 / (best * 1e-3) / 1e9
    print("%-26s build %.3f ms device (best of 5)  %.0f GB/s algorithmic = %.1f %% of the measured HBM bandwidth" % (name, best, gbs, 100 * gbs / peak), flush=True)
