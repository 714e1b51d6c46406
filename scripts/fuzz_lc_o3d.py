"""GPU box: randomised parity sweep (not part of the suite): FPFH, feature 1-NN, RANSAC with several model sizes and
thresholds, overlap indices, and the Open3D-mode registration (all kinds, all query maps) against the oracle."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth
from oracle import oracle_py as O

rng = np.random.default_rng(int(os.environ.get("SEED", 1)))
lc = B.LoopClosure()
fails = 0
def check(ok, what):
    global fails
    if not ok:
        fails += 1
        print("FAIL", what, flush=True)

def cloud(n):
    kind = rng.integers(0, 3)
    if kind == 0:      # planes
        a = np.c_[rng.uniform(-5, 5, n // 2), rng.uniform(-5, 5, n // 2), 0.03 * rng.standard_normal(n // 2)]
        b = np.c_[rng.uniform(-5, 5, n - n // 2), 5 + 0.03 * rng.standard_normal(n - n // 2), rng.uniform(0, 4, n - n // 2)]
        p = np.vstack([a, b])
    elif kind == 1:    # blob with dense core
        p = rng.standard_normal((n, 3)) * rng.choice([0.3, 2.0, 6.0], size=(n, 1))
    else:              # uniform box
        p = rng.uniform(-4, 4, (n, 3))
    return p.astype(np.float32)

for trial in range(int(os.environ.get("TRIALS", 12))):
    n = int(rng.integers(300, 5000))
    p = cloud(n)
    p4 = O.feat4(p)
    radius, knn = float(rng.choice([0.5, 1.0, 2.5])), int(rng.choice([5, 30, 100, 250]))
    nrm, _ = O.o3d_estimate_normals(p4, 10, 1.0, threads=8)
    f, cnt = lc.compute_fpfh(p4, nrm, radius, knn)
    g, gcnt = O.o3d_fpfh(p4, nrm, radius, knn, threads=8)
    err = np.abs(f - g).max(1)
    check((cnt == gcnt).all(), "fpfh counts %d" % trial)
    bad = np.nonzero(err > 1e-9)[0]
    explained = True
    if len(bad):
        # a differing FPFH must trace back to a differing SPFH (a pair feature on a bin edge, libm vs libdevice): the set
        # of bad points has to be covered by the neighbourhoods' inverse images of a FEW source points
        ids_k, _ = O.KdTree(p4).knn(p4, knn, radius, threads=8)
        members = [set(row[row >= 0].tolist()) for row in ids_k[bad]]
        common = set.intersection(*members) if members else set()
        remaining, sources = set(range(len(bad))), []
        while remaining and len(sources) < 4:
            cand = {}
            for bi in remaining:
                for c in members[bi]:
                    cand[c] = cand.get(c, 0) + 1
            c = max(cand, key=cand.get)
            sources.append(c)
            remaining = {bi for bi in remaining if c not in members[bi]}
        explained = not remaining
        print("   fpfh trial %d: %d points differ (max %.3g of 200), all of them hold one of the %d source point(s) %s in their neighbourhood: %s"
              % (trial, len(bad), err.max(), len(sources), sources, explained), flush=True)
    check(explained and np.median(err) < 1e-9,   # one moved bin is worth 100 / (n - 1): 25 at knn 5
          "fpfh values %d (%.3g bad, max %.3g)" % (trial, (err > 1e-9).mean(), err.max()))
    # a moved, subsampled, noisy copy
    yaw = rng.uniform(-1, 1)
    T = np.eye(4); T[:3, :3] = [[np.cos(yaw), -np.sin(yaw), 0], [np.sin(yaw), np.cos(yaw), 0], [0, 0, 1]]; T[:3, 3] = rng.uniform(-2, 2, 3)
    keep = rng.random(n) < 0.8
    q = ((p[keep].astype(np.float64) - T[:3, 3]) @ T[:3, :3] + 0.005 * rng.standard_normal((keep.sum(), 3))).astype(np.float32)
    q4 = O.feat4(q)
    qn, _ = O.o3d_estimate_normals(q4, 10, 1.0, threads=8)
    fq, _ = O.o3d_fpfh(q4, qn, radius, knn, threads=8)
    ids, d2 = lc.feature_nn(fq, g)
    oi, od = O.o3d_feature_nn(fq, g, threads=8)
    check((ids == oi).all() and (d2 == od).all(), "feature nn %d" % trial)
    kw = dict(ransac_n=int(rng.choice([3, 4, 6])), max_correspondence_distance=float(rng.choice([0.05, 0.3])),
              distance_threshold=float(rng.choice([0.0, 0.1, 0.5])), edge_length_similarity=float(rng.choice([0.0, 0.5, 0.9])),
              max_iteration=int(rng.choice([500, 5000, 70000])), confidence=float(rng.choice([0.9, 0.999, 1.0])),
              seed=int(rng.integers(1, 1 << 30)), mutual_filter=int(rng.integers(0, 2)))
    r = lc.ransac(q4, p4, fq, g, batch=int(rng.choice([1024, 4096, 65536])), **kw)
    o = O.o3d_ransac(q4, p4, fq, g, threads=8, **kw)
    ok = (r["iterations_run"] == o["iterations_run"] and r["validations"] == o["validations"] and
          np.abs(r["T"] - o["T"]).max() < 1e-9 and r["fitness"] == o["fitness"] and
          r["correspondences"].shape == o["correspondences"].shape and (r["correspondences"] == o["correspondences"]).all())
    check(ok, "ransac %d %s: it %d/%d val %d/%d dT %.3g" % (trial, kw, r["iterations_run"], o["iterations_run"], r["validations"],
                                                          o["validations"], np.abs(r["T"] - o["T"]).max()))
    vox, mn = float(rng.choice([0.2, 0.5, 1.5])), int(rng.choice([1, 2, 5]))
    si, ti = lc.overlap_indices(q4, p4, T, vox, mn)
    osi, oti = O.overlap_indices(q4, p4, T, vox, mn)
    check(len(si) == len(osi) and (si == osi).all() and len(ti) == len(oti) and (ti == oti).all(), "overlap %d" % trial)
    print("trial %d ok so far: n %d radius %.1f knn %d ransac it %d fitness %.3f" % (trial, n, radius, knn, r["iterations_run"], r["fitness"]), flush=True)

# Open3D-mode registration on a C4 pair, every kind and query map
pair = synth.make_c4(n_distinct=1, N=60_000)[0]
tgt4, src4 = synth.feat4(pair["tgt_xyz"]), synth.feat4(pair["src_xyz"])
est = B.O3dICP(kind=0)
tn = est.estimate_normals(tgt4, knn=10, radius=2.0).astype(np.float32)
sn = est.estimate_normals(src4, knn=10, radius=2.0).astype(np.float32)
est.close()
for kind in (0, 1, 2):
    cov_s = cov_t = None
    o = None
    for qmap in (0, 1, 2):
        os.environ["B200_O3D_QMAP"] = str(qmap)
        h = B.O3dICP(kind=kind, max_correspondence_distance=0.5, max_iteration=30)
        h.set_target(tgt4, tn)
        r = h.register(src4, sn if kind == 2 else None, init=pair["T_init"].astype(np.float64))
        h.close()
        if o is None:
            o = r
        check(r["iterations"] == o["iterations"] and np.abs(r["T"] - o["T"]).max() < 1e-9, "o3d kind %d qmap %d vs qmap 0: %d/%d %.3g" % (
            kind, qmap, r["iterations"], o["iterations"], np.abs(r["T"] - o["T"]).max()))
    print("o3d kind %d: iterations %d fitness %.4f" % (kind, o["iterations"], o["fitness"]), flush=True)
print("FUZZ FAILS:", fails)
