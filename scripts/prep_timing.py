"""GPU box: device time of the pre-processing kernels (b200_prep_*) against their HBM roofline, next to the CPU oracle.
Algorithmic bytes: crop = 16 B read per point + 20 B written per kept point (point + index); voxel down-sample = 16 B
read per point + 24 B written per voxel."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from open3d_slam_private_b200 import capi as B, synth
from oracle import oracle_py as O

peak = 6550.1
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
d = synth.make_c2(M=5_000_000, N=131072, n_poses=1)
scan4 = synth.feat4(d["scans"][0]["xyz"])
map4 = synth.feat4(d["map_xyz"])
prep = B.Prep()
centre = tuple(float(v) for v in d["scans"][0]["T_true"][:3, 3])


def best(fn, reps=7):
    ms, wall, out = 1e9, 1e9, None
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        wall = min(wall, (time.perf_counter() - t0) * 1e3)
        ms = min(ms, prep.last_device_ms())
    return ms, wall, out


def cpu(fn, reps=2):
    t = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t = min(t, (time.perf_counter() - t0) * 1e3)
    return t


rows = []
for name, cloud, kw in (("crop MinMaxRadius 1..30 m, 131072-pt scan (sensor frame)", scan4, dict(radius_min=1.0, radius_max=30.0)),
                        ("crop MaxRadius 30 m, 5 M-pt map (cropSubmap)", map4, dict(radius_max=30.0, center=centre))):
    kind = B.CROP_MINMAX_RADIUS if "radius_min" in kw else B.CROP_MAX_RADIUS
    ms, wall, out = best(lambda: prep.crop(kind, cloud, **kw))
    n, k = cloud.shape[0], len(out[2])
    byts = 16 * n + 20 * k
    c_ms = cpu(lambda: O.crop(kind, cloud, **kw))
    rows.append(dict(op=name, points=n, kept=k, device_ms=ms, wall_ms_incl_h2d_d2h=wall, algorithmic_GBps=byts / ms / 1e6,
                     frac_of_measured_hbm=byts / ms / 1e6 / peak, cpu_oracle_ms_1_thread=c_ms))
for name, cloud, vs in (("voxel 0.10 m, 131072-pt scan", scan4, 0.10), ("voxel 0.50 m, 131072-pt scan", scan4, 0.50),
                        ("voxel 0.25 m, 5 M-pt map", map4, 0.25)):
    ms, wall, out = best(lambda: prep.voxel_downsample(vs, cloud))
    n, k = cloud.shape[0], len(out[2])
    byts = 16 * n + 24 * k
    c_ms = cpu(lambda: O.voxel_downsample(vs, cloud), reps=1)
    rows.append(dict(op=name, points=n, voxels=k, ordered_voxels=prep.last_ordered_voxels(), device_ms=ms,
                     wall_ms_incl_h2d_d2h=wall, algorithmic_GBps=byts / ms / 1e6,
                     frac_of_measured_hbm=byts / ms / 1e6 / peak, cpu_oracle_ms_1_thread=c_ms))
for r in rows:
    print(json.dumps(r))
