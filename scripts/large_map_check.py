"""GPU box: one device holding a 100 M-point map (6.4 GB of points and normals, twice that with the cell-sorted copies):
index build time, exact k-NN against a numpy brute force over the tiles a neighbour can lie in, and a registration of a
131 072-point scan against the whole map (pose vs truth). The map is BASELINE config C2's 5 M-point street laid end to
end 20 times along x (tiles do not overlap), so a query's neighbour lies in its own tile or the next one."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from open3d_slam_private_b200 import capi as B, synth

TILES = int(os.environ.get("TILES", 20))
d = synth.make_c2(M=5_000_000, N=131072, n_poses=2)
xyz, nrm = d["map_xyz"].astype(np.float32), d["map_normals"].astype(np.float32)
L = float(np.ceil(xyz[:, 0].max() - xyz[:, 0].min()) + 10.0)
M1 = len(xyz)
big = np.empty((TILES * M1, 4), np.float32)
bign = np.empty((TILES * M1, 3), np.float32)
for t in range(TILES):
    big[t * M1:(t + 1) * M1, :3] = xyz
    big[t * M1:(t + 1) * M1, 0] += np.float32(t * L)
    big[t * M1:(t + 1) * M1, 3] = 1.0
    bign[t * M1:(t + 1) * M1] = nrm
M = len(big)
t0 = time.perf_counter()
idx = B.Index(big, bign)
wall = time.perf_counter() - t0
info = idx.info()
print("map %d points: H2D + index build %.2f s wall, build %.2f ms device, cell %.3f m, dims %s, occupied cells %d" % (
    M, wall, info["build_ms"], info["cell_size"], info.get("dims"), info["occupied_cells"]), flush=True)

rng = np.random.default_rng(2)
Q = 96
src = rng.integers(0, M, Q)
q = big[src].copy()
q[:, :3] += rng.normal(0, 0.05, (Q, 3)).astype(np.float32)
q[:8, :3] += 3.0                                   # a few queries off the surfaces
for k in (1, 5):
    ids, dist = idx.find_closests(q, knn=k, max_dist=np.inf)
    bad = 0
    for i in range(Q):
        t = int(src[i] // M1)
        lo, hi = max(0, t - 1) * M1, min(TILES, t + 2) * M1
        dd = (big[lo:hi, :3] - q[i, :3]) ** 2
        d2 = (dd[:, 0] + dd[:, 1]) + dd[:, 2]                       # float32, x -> y -> z
        order = np.lexsort((np.arange(lo, hi), d2))[:k] + lo
        # nothing outside the three tiles can be closer than the k-th neighbour found inside them
        kth = float(d2[order[-1] - lo])
        margin = min(abs(q[i, 0] - (big[lo, 0] if lo else -1e9)), 1e9)
        assert np.sqrt(kth) < L / 2, "neighbour search window too small"
        if not ((ids[i] == order).all() and (dist[i].view(np.uint32) == d2[order - lo].view(np.uint32)).all()):
            bad += 1
    print("k = %d: %d / %d queries bit-exact against the brute force (ids and float distances)" % (k, Q - bad, Q), flush=True)
    assert bad == 0

# registration of a scan taken in tile 7 against the whole 100 M-point map
icp = B.ICP(B.default_icp_params(knn=1, maxDist=0.5, use_trimmed=1, trimmed_ratio=0.9, max_iterations=30, use_differential=1,
                                 min_diff_rot=1e-3, min_diff_trans=8e-3, smooth_length=3))
t0 = time.perf_counter()
icp.init_reference(big, bign)
print("initReference (H2D + centroid + index) %.2f s wall" % (time.perf_counter() - t0), flush=True)
shift = np.eye(4); shift[0, 3] = 7 * L
for s in d["scans"]:
    Ti, Tt = shift @ s["T_init"], shift @ s["T_true"]
    r = None
    for rep in range(3):
        r = icp.compute(synth.feat4(s["xyz"]), None, Ti.astype(np.float32))
    err = np.abs(r["T"] - Tt)
    print("registration vs the %d-point map: status %d, %d iterations, %.3f ms device, |t - truth| %.4f m, |R - truth| %.2e" % (
        M, r["status"], r["iterations"], r["device_ms"], err[:3, 3].max(), err[:3, :3].max()), flush=True)
    assert r["status"] == 0 and err[:3, 3].max() < 0.02
print("large map check ok")
