"""Loop-closure front half: ComputeFPFHFeature (Submap.cpp:272) and
RegistrationRANSACBasedOnFeatureMatching (PlaceRecognition.cpp:81-85).

CPU: the oracle against an independent numpy restatement of FPFH and against known answers (Open3D v0.15.1 is not in
the reference tree and the reference holds no test for these calls: parity is pinned to the oracle, unpinned to the
reference binary). GPU: the device path (b200_lc_*, through the C ABI) against the oracle."""
import ctypes

import numpy as np
import pytest

from open3d_slam_private_b200 import capi as B
from open3d_slam_private_b200 import synth
from oracle import oracle_py as O


def yaw_T(yaw, t):
    T = np.eye(4)
    T[:3, :3] = [[np.cos(yaw), -np.sin(yaw), 0], [np.sin(yaw), np.cos(yaw), 0], [0, 0, 1]]
    T[:3, 3] = t
    return T


_PAIR = {}


def submap_pair(M=250_000, R=22.0, vox=0.5):
    """two independent samplings of the same street scene around two nearby poses, voxelised like
    Submap::computeFeatures does (featureVoxelSize_ 0.5); the source is expressed in a frame moved by T^-1"""
    key = (M, R, vox)
    if key not in _PAIR:
        sc = synth.Scene(seed=5)
        a, _ = sc.sample_map(M, seed=6)
        b, _ = sc.sample_map(M, seed=7)
        a = a[(a[:, 0] ** 2 + a[:, 1] ** 2) < R * R]
        b = b[((b[:, 0] - 1.0) ** 2 + (b[:, 1] - 0.5) ** 2) < R * R]
        T = yaw_T(0.3, [3.0, -2.0, 0.2])
        Ti = np.linalg.inv(T)
        bs = (b.astype(np.float64) @ Ti[:3, :3].T + Ti[:3, 3]).astype(np.float32)
        t4 = O.feat4(O.voxel_downsample(vox, O.feat4(a))[0])
        s4 = O.feat4(O.voxel_downsample(vox, O.feat4(bs))[0])
        tn, _ = O.o3d_estimate_normals(t4, 30, 1.5, threads=8)
        sn, _ = O.o3d_estimate_normals(s4, 30, 1.5, threads=8)
        tf, _ = O.o3d_fpfh(t4, tn, 2.5, 100, threads=8)
        sf, _ = O.o3d_fpfh(s4, sn, 2.5, 100, threads=8)
        _PAIR[key] = dict(t4=t4, s4=s4, tn=tn, sn=sn, tf=tf, sf=sf, T=T)
    return _PAIR[key]


# ---------------------------------------------------------------------------------------------------- CPU: oracle
def np_pair_features(p1, n1, p2, n2):
    dp = p2 - p1
    d = np.linalg.norm(dp)
    if d == 0.0:
        return np.zeros(4)
    a1, a2 = n1.dot(dp) / d, n2.dot(dp) / d
    if np.arccos(abs(a1)) > np.arccos(abs(a2)):
        n1, n2, dp, f2 = n2, n1, -dp, -a2
    else:
        f2 = a1
    v = np.cross(dp, n1)
    vn = np.linalg.norm(v)
    if vn == 0.0:
        return np.zeros(4)
    v = v / vn
    w = np.cross(n1, v)
    return np.array([np.arctan2(w.dot(n2), n1.dot(n2)), v.dot(n2), f2, d])


def np_fpfh(pts, nrm, radius, max_nn):
    """ComputeFPFHFeature restated with numpy only (brute-force neighbourhoods); independent of oracle/fpfh.cpp"""
    n = len(pts)
    p32 = pts.astype(np.float32)
    nb = []
    for i in range(n):
        d = ((p32 - p32[i]) ** 2)
        d2 = (d[:, 0] + d[:, 1]) + d[:, 2]                  # float32, x -> y -> z like the k-NN contract
        idx = np.nonzero(d2 <= np.float32(radius) * np.float32(radius))[0]
        idx = idx[np.lexsort((idx, d2[idx]))][:max_nn]
        nb.append(idx)
    P = p32.astype(np.float64)
    spfh = np.zeros((n, 33))
    for i in range(n):
        idx = nb[i]
        if len(idx) <= 1:
            continue
        inc = 100.0 / (len(idx) - 1)
        for j in idx[1:]:
            f = np_pair_features(P[i], nrm[i], P[j], nrm[j])
            for g, x in enumerate((11 * (f[0] + np.pi) / (2 * np.pi), 11 * (f[1] + 1.0) * 0.5, 11 * (f[2] + 1.0) * 0.5)):
                spfh[i, 11 * g + min(10, max(0, int(np.floor(x))))] += inc
    out = np.zeros((n, 33))
    for i in range(n):
        idx = nb[i]
        if len(idx) <= 1:
            continue
        s = np.zeros(3)
        for j in idx[1:]:
            d2 = ((P[j] - P[i]) ** 2).sum()
            if d2 == 0.0:
                continue
            val = spfh[j] / d2
            out[i] += val
            s += val.reshape(3, 11).sum(1)
        s = np.where(s != 0.0, 100.0 / np.where(s != 0.0, s, 1.0), 0.0)
        out[i] = out[i] * np.repeat(s, 11) + spfh[i]
    return out, np.array([len(x) for x in nb])


def small_cloud(n=260, seed=3):
    rng = np.random.default_rng(seed)
    # two planes meeting in an edge + a few isolated points and an exact duplicate
    a = np.c_[rng.uniform(-2, 2, n // 2), rng.uniform(-2, 2, n // 2), 0.02 * rng.standard_normal(n // 2)]
    b = np.c_[rng.uniform(-2, 2, n // 2), 2 + 0.02 * rng.standard_normal(n // 2), rng.uniform(0, 2, n // 2)]
    far = np.array([[30.0, 0, 0], [30.2, 0, 0], [-40.0, 5, 1]])
    pts = np.vstack([a, b, far, a[:1]]).astype(np.float32)
    nrm = np.vstack([np.tile([0, 0, 1.0], (n // 2, 1)), np.tile([0, -1.0, 0], (n // 2, 1)), np.tile([0, 0, 1.0], (3, 1)),
                     [[0, 0, 1.0]]])
    nrm = nrm + 0.05 * rng.standard_normal(nrm.shape)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    return pts, nrm


def test_oracle_fpfh_equals_numpy_restatement():
    pts, nrm = small_cloud()
    for radius, max_nn in ((0.8, 100), (1.5, 12)):
        f, cnt = O.o3d_fpfh(O.feat4(pts), nrm, radius, max_nn)
        g, gcnt = np_fpfh(pts, nrm, radius, max_nn)
        assert (cnt == gcnt).all()
        assert np.abs(f - g).max() < 1e-9, np.abs(f - g).max()
        has = cnt > 1
        # every 11-bin group of a point with neighbours sums to 100 (SPFH) + 100 (weighted neighbours)
        assert np.allclose(f[has].reshape(-1, 3, 11).sum(2), 200.0, atol=1e-9)
        assert (f[~has] == 0).all()


def test_oracle_feature_nn_is_exact():
    rng = np.random.default_rng(0)
    a, b = rng.uniform(0, 100, (300, 33)), rng.uniform(0, 100, (500, 33))
    b[77] = b[400]                       # a tie: the smaller index wins
    a[5] = b[400]
    ids, d2 = O.o3d_feature_nn(a, b)
    ref = ((a[:, None, :] - b[None, :, :]) ** 2).sum(2)
    assert (ids == ref.argmin(1)).all() and ids[5] == 77
    assert np.allclose(d2, ref.min(1), rtol=1e-12)


def test_oracle_ransac_known_answer():
    # exact correspondences: target = T * source, features = one-hot-ish codes that make the feature match the identity
    rng = np.random.default_rng(1)
    n = 400
    src = rng.uniform(-10, 10, (n, 3)).astype(np.float32)
    T = yaw_T(0.7, [1.0, 2.0, -0.5])
    tgt = (src.astype(np.float64) @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
    feat = rng.uniform(0, 100, (n, 33))
    perm = rng.permutation(n)
    r = O.o3d_ransac(O.feat4(src), O.feat4(tgt[perm]), feat, feat[perm], max_correspondence_distance=0.05,
                     distance_threshold=0.05, edge_length_similarity=0.9, max_iteration=1000, confidence=0.999, seed=3)
    assert r["mutual_used"] and r["n_feature_correspondences"] == n
    assert r["fitness"] == 1.0 and len(r["correspondences"]) == n
    assert (perm[r["correspondences"][:, 1]] == r["correspondences"][:, 0]).all()
    assert np.abs(r["T"] - T).max() < 1e-4
    # inlier ratio 1 => the exit estimate is 0 after the first validated hypothesis
    assert r["validations"] == 1 and r["iterations_run"] < 10
    # Open3D's early returns: ransac_n < 3 or no distance -> empty RegistrationResult (identity, fitness 0)
    e = O.o3d_ransac(O.feat4(src), O.feat4(tgt), feat, feat, ransac_n=2)
    assert (e["T"] == np.eye(4)).all() and e["fitness"] == 0 and len(e["correspondences"]) == 0


def test_oracle_ransac_recovers_the_submap_transform():
    p = submap_pair()
    r = O.o3d_ransac(p["s4"], p["t4"], p["sf"], p["tf"], max_iteration=1000000, confidence=0.999, seed=7, threads=8)
    assert r["mutual_used"] and r["n_feature_correspondences"] > 1000
    assert len(r["correspondences"]) >= 25                       # ransacMinCorrespondenceSetSize_
    dT = np.linalg.inv(p["T"]) @ r["T"]
    assert np.abs(dT[:3, 3]).max() < 0.5 and np.abs(dT[:3, :3] - np.eye(3)).max() < 0.03   # voxel-size accuracy
    # reproducible for a seed, different stream for another
    r2 = O.o3d_ransac(p["s4"], p["t4"], p["sf"], p["tf"], max_iteration=1000000, confidence=0.999, seed=7, threads=2)
    assert (r2["T"] == r["T"]).all() and r2["iterations_run"] == r["iterations_run"]


def test_oracle_overlap_indices_known_answer():
    # voxel 1.0: source points (after a +10 shift in x) and target points share voxel (10, 0, 0) only
    src = O.feat4(np.array([[0.2, 0.3, 0.4], [0.7, 0.1, 0.9], [5.5, 0.5, 0.5], [-0.5, 0.5, 0.5]], np.float32))
    tgt = O.feat4(np.array([[10.9, 0.9, 0.1], [3.5, 0.5, 0.5], [10.1, 0.2, 0.3], [9.5, 0.5, 0.5]], np.float32))
    T = np.eye(4); T[0, 3] = 10.0
    si, ti = O.overlap_indices(src, tgt, T, 1.0, 1)
    assert si.tolist() == [0, 1, 3] and ti.tolist() == [0, 2, 3]      # (-0.5 + 10) and 9.5 share voxel 9
    si, ti = O.overlap_indices(src, tgt, T, 1.0, 2)
    assert si.tolist() == [0, 1] and ti.tolist() == [0, 2]
    # negative coordinates floor downwards (VoxelHashMap.hpp:48-51)
    si, ti = O.overlap_indices(O.feat4(np.array([[-0.1, 0, 0]], np.float32)), O.feat4(np.array([[0.1, 0, 0]], np.float32)),
                               np.eye(4), 1.0, 1)
    assert len(si) == 0 and len(ti) == 0


def test_ransac_struct_sizes_match_header():
    lib = B.load()
    assert ctypes.sizeof(B.RansacParams) == lib.b200_struct_size(6)
    assert ctypes.sizeof(B.RansacResult) == lib.b200_struct_size(7)


# ------------------------------------------------------------------------------------------------- GPU: device path
@pytest.mark.gpu
def test_fpfh_small_cloud_vs_oracle():
    pts, nrm = small_cloud()
    lc = B.LoopClosure()
    # max_nn 800: the neighbourhood tile no longer fits into shared memory, the per-thread fallback kernel runs
    for radius, max_nn in ((0.8, 100), (1.5, 12), (0.3, 5), (100.0, 300), (100.0, 800)):
        f, cnt = lc.compute_fpfh(O.feat4(pts), nrm, radius, max_nn)
        g, gcnt = O.o3d_fpfh(O.feat4(pts), nrm, radius, max_nn)
        assert (cnt == gcnt).all()
        assert np.abs(f - g).max() < 1e-9, (radius, max_nn, np.abs(f - g).max())
    f, cnt = lc.compute_fpfh(np.zeros((0, 4), np.float32), np.zeros((0, 3)), 1.0, 10)
    assert f.shape == (0, 33)


@pytest.mark.gpu
def test_fpfh_submap_vs_oracle():
    # the reference's own configuration: featureRadius_ 2.5, featureKnn_ 100 on a 0.5 m voxelised submap
    p = submap_pair()
    lc = B.LoopClosure()
    for pts, nrm, ref in ((p["t4"], p["tn"], p["tf"]), (p["s4"], p["sn"], p["sf"])):
        f, cnt = lc.compute_fpfh(pts, nrm, 2.5, 100)
        # libm and libdevice round acos / atan2 differently in the last place: among the 1.6 M point pairs of this cloud
        # one or two sit on a bin edge (or on the |angle1| = |angle2| swap) and fall on the other side. One such SPFH
        # bin moves 100 / (n - 1) and reaches the FPFH of the ~100 points that have that point as a neighbour, weighted
        # down by their sums — so: nearly all points to 1e-9, the rest off by a small fraction of a group's 200.
        err = np.abs(f - ref).max(1)
        assert (err > 1e-9).mean() < 0.03, (err > 1e-9).mean()
        assert err.max() < 5.0 and np.median(err) < 1e-9, err.max()
        assert cnt.max() == 100 and cnt.min() >= 1
    assert lc.last_device_ms() > 0


@pytest.mark.gpu
def test_feature_nn_bit_exact():
    p = submap_pair()
    lc = B.LoopClosure()
    ids, d2 = lc.feature_nn(p["sf"], p["tf"])
    oi, od = O.o3d_feature_nn(p["sf"], p["tf"], threads=8)
    assert (ids == oi).all() and (d2 == od).all()
    rng = np.random.default_rng(0)
    a, b = rng.uniform(0, 100, (70, 33)), rng.uniform(0, 100, (5000, 33))
    b[4100] = b[31]
    a[5] = b[31]
    ids, d2 = lc.feature_nn(a, b)
    oi, od = O.o3d_feature_nn(a, b)
    assert (ids == oi).all() and ids[5] == 31 and (d2 == od).all()


def assert_ransac_equal(r, o):
    assert r["mutual_used"] == o["mutual_used"] and r["n_feature_correspondences"] == o["n_feature_correspondences"]
    assert r["iterations_run"] == o["iterations_run"] and r["validations"] == o["validations"]
    assert np.abs(r["T"] - o["T"]).max() < 1e-9, np.abs(r["T"] - o["T"]).max()
    assert r["fitness"] == o["fitness"] and abs(r["inlier_rmse"] - o["inlier_rmse"]) < 1e-12
    assert (r["correspondences"] == o["correspondences"]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [7, 8, 9])
def test_ransac_submap_vs_oracle(seed):
    p = submap_pair()
    lc = B.LoopClosure()
    kw = dict(max_iteration=1000000, confidence=0.999, seed=seed)
    r = lc.ransac(p["s4"], p["t4"], p["sf"], p["tf"], **kw)
    o = O.o3d_ransac(p["s4"], p["t4"], p["sf"], p["tf"], threads=8, **kw)
    assert_ransac_equal(r, o)
    dT = np.linalg.inv(p["T"]) @ r["T"]
    assert np.abs(dT[:3, 3]).max() < 0.5
    # the batch size does not change the answer
    r2 = lc.ransac(p["s4"], p["t4"], p["sf"], p["tf"], batch=1024, **kw)
    assert_ransac_equal(r2, o)


@pytest.mark.gpu
def test_ransac_runs_all_iterations_without_an_exit_estimate():
    # confidence 1 never lowers the exit estimate: every iteration is consumed, across several batches
    p = submap_pair()
    lc = B.LoopClosure()
    kw = dict(max_iteration=20000, confidence=1.0, seed=11, max_correspondence_distance=0.4)
    o = O.o3d_ransac(p["s4"], p["t4"], p["sf"], p["tf"], threads=8, **kw)
    for batch in (4096, 65536):
        r = lc.ransac(p["s4"], p["t4"], p["sf"], p["tf"], batch=batch, **kw)
        assert r["iterations_run"] == 20000
        assert_ransac_equal(r, o)
    # without the checkers every hypothesis is validated
    kw = dict(max_iteration=3000, confidence=1.0, seed=2, edge_length_similarity=0.0, distance_threshold=0.0)
    r = lc.ransac(p["s4"], p["t4"], p["sf"], p["tf"], batch=1024, **kw)
    o = O.o3d_ransac(p["s4"], p["t4"], p["sf"], p["tf"], threads=8, **kw)
    assert r["validations"] == 3000
    assert_ransac_equal(r, o)


@pytest.mark.gpu
def test_ransac_mutual_fallback_and_early_returns():
    p = submap_pair()
    lc = B.LoopClosure()
    # no mutual filter: every source point with its feature neighbour
    kw = dict(mutual_filter=0, max_iteration=50000, confidence=0.999, seed=5)
    r = lc.ransac(p["s4"], p["t4"], p["sf"], p["tf"], **kw)
    o = O.o3d_ransac(p["s4"], p["t4"], p["sf"], p["tf"], threads=8, **kw)
    assert not r["mutual_used"] and r["n_feature_correspondences"] == len(p["s4"])
    assert_ransac_equal(r, o)
    # too few mutual pairs (< 3 * ransac_n) -> falls back to all pairs
    s4, sf = p["s4"][:40], p["sf"][:40]
    kw = dict(max_iteration=2000, confidence=0.99, seed=1)
    r = lc.ransac(s4, p["t4"], sf, p["tf"], **kw)
    o = O.o3d_ransac(s4, p["t4"], sf, p["tf"], threads=8, **kw)
    assert_ransac_equal(r, o)
    # Open3D's early returns: identity, fitness 0, no correspondences
    e = lc.ransac(p["s4"], p["t4"], p["sf"], p["tf"], ransac_n=2)
    assert (e["T"] == np.eye(4)).all() and e["fitness"] == 0 and len(e["correspondences"]) == 0
    e = lc.ransac(p["s4"], p["t4"], p["sf"], p["tf"], max_correspondence_distance=0.0)
    assert (e["T"] == np.eye(4)).all() and e["fitness"] == 0
    with pytest.raises(B.B200Error):
        lc.ransac(p["s4"], p["t4"], p["sf"], p["tf"], ransac_n=9)


@pytest.mark.gpu
def test_ransac_exact_correspondences_known_answer():
    rng = np.random.default_rng(1)
    n = 400
    src = rng.uniform(-10, 10, (n, 3)).astype(np.float32)
    T = yaw_T(0.7, [1.0, 2.0, -0.5])
    tgt = (src.astype(np.float64) @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
    feat = rng.uniform(0, 100, (n, 33))
    perm = rng.permutation(n)
    kw = dict(max_correspondence_distance=0.05, distance_threshold=0.05, edge_length_similarity=0.9, max_iteration=1000,
              confidence=0.999, seed=3)
    r = B.LoopClosure().ransac(O.feat4(src), O.feat4(tgt[perm]), feat, feat[perm], **kw)
    o = O.o3d_ransac(O.feat4(src), O.feat4(tgt[perm]), feat, feat[perm], **kw)
    assert_ransac_equal(r, o)
    assert r["fitness"] == 1.0 and np.abs(r["T"] - T).max() < 1e-4


@pytest.mark.gpu
def test_overlap_indices_bit_exact():
    # PlaceRecognition.cpp:98-108: the dense submaps, the RANSAC pose, voxel = 3 x the map voxel size
    p = synth.make_c4(n_distinct=1, N=60_000)[0]
    src4, tgt4 = synth.feat4(p["src_xyz"]), synth.feat4(p["tgt_xyz"])
    lc = B.LoopClosure()
    for vox, mn, T in ((0.3, 1, p["T_true"]), (0.75, 3, p["T_init"]), (2.0, 40, np.eye(4))):
        si, ti = lc.overlap_indices(src4, tgt4, T, vox, mn)
        osi, oti = O.overlap_indices(src4, tgt4, T, vox, mn)
        assert len(si) == len(osi) and (si == osi).all() and len(ti) == len(oti) and (ti == oti).all()
    assert 0 < len(si) < len(src4)
    si, ti = lc.overlap_indices(src4[:0], tgt4, np.eye(4), 0.5, 1)
    assert len(si) == 0 and len(ti) == 0
    with pytest.raises(B.B200Error):
        lc.overlap_indices(src4, tgt4, np.eye(4), 0.0, 1)


def _splitmix(x):
    m = (1 << 64) - 1
    x &= m
    x ^= x >> 30; x = (x * 0xbf58476d1ce4e5b9) & m
    x ^= x >> 27; x = (x * 0x94d049bb133111eb) & m
    x ^= x >> 31
    return x


def np_ransac(src, tgt, cs, ct, rn, max2, edge, dist_thr, max_iteration, confidence, seed):
    """RegistrationRANSACBasedOnCorrespondence restated with numpy + python integers (independent of oracle/fpfh.cpp):
    the documented sample stream, both checkers, Umeyama by numpy's SVD, the (fitness, rmse) order and the exit estimate"""
    nc = len(cs)
    S, D = src[cs].astype(np.float64), tgt[ct].astype(np.float64)
    best_good, best_rmse, best_T = 0, 0.0, np.eye(4)
    exit_itr, itr, validations = max_iteration, 0, 0
    while itr < max_iteration and itr < exit_itr:
        idx = [(((_splitmix(seed + (itr * rn + j) * 0x9e3779b97f4a7c15) >> 32) * nc) >> 32) for j in range(rn)]
        itr += 1
        s, d = S[idx], D[idx]
        ok = True
        if edge > 0:
            for a in range(rn):
                for b in range(a + 1, rn):
                    ds, dt = np.linalg.norm(s[a] - s[b]), np.linalg.norm(d[a] - d[b])
                    if ds < dt * edge or dt < ds * edge:
                        ok = False
        if not ok:
            continue
        ms, md = s.mean(0), d.mean(0)
        U, _, Vt = np.linalg.svd((d - md).T @ (s - ms) / rn)
        R = U @ np.diag([1.0, 1.0, np.sign(np.linalg.det(U) * np.linalg.det(Vt)) or 1.0]) @ Vt
        T = np.eye(4); T[:3, :3] = R; T[:3, 3] = md - R @ ms
        if dist_thr > 0 and (np.linalg.norm(s @ R.T + T[:3, 3] - d, axis=1) > dist_thr).any():
            continue
        validations += 1
        e = ((S @ R.T + T[:3, 3] - D) ** 2).sum(1)
        inl = e < max2
        good = int(inl.sum())
        rmse = float(np.sqrt(e[inl].sum() / good)) if good else 0.0
        if good > best_good or (good == best_good and good > 0 and rmse < best_rmse):
            best_good, best_rmse, best_T = good, rmse, T
            pw = (good / nc) ** rn
            if confidence < 1.0 and pw > 0:
                k = 0.0 if pw >= 1.0 else np.log(1.0 - confidence) / np.log(1.0 - pw)
                if 0 <= k < exit_itr:
                    exit_itr = int(np.ceil(k))
    return dict(T=best_T, good=best_good, rmse=best_rmse, iterations_run=itr, validations=validations)


@pytest.mark.parametrize("seed,rn", [(5, 3), (6, 4), (7, 3)])
def test_oracle_ransac_equals_numpy_restatement(seed, rn):
    rng = np.random.default_rng(seed)
    n = 500
    src = rng.uniform(-10, 10, (n, 3)).astype(np.float32)
    T = yaw_T(-0.4, [0.5, -1.0, 0.3])
    tgt = (src.astype(np.float64) @ T[:3, :3].T + T[:3, 3] + 0.01 * rng.standard_normal((n, 3))).astype(np.float32)
    feat = rng.uniform(0, 100, (n, 33))
    tfeat = feat.copy()
    wrong = rng.choice(n, 200, replace=False)                 # 40 % of the feature matches point at the wrong target
    tfeat[wrong] = rng.uniform(0, 100, (200, 33))
    kw = dict(mutual_filter=False, max_correspondence_distance=0.08, ransac_n=rn, edge_length_similarity=0.9,
              distance_threshold=0.1, max_iteration=400, confidence=0.9999, seed=seed)
    o = O.o3d_ransac(O.feat4(src), O.feat4(tgt), feat, tfeat, **kw)
    ij, _ = O.o3d_feature_nn(feat, tfeat)
    g = np_ransac(src, tgt, np.arange(n), ij, rn, 0.08 ** 2, 0.9, 0.1, 400, 0.9999, seed)
    assert o["iterations_run"] == g["iterations_run"] and o["validations"] == g["validations"]
    assert len(o["correspondences"]) == g["good"] and g["good"] > 100
    assert np.abs(o["T"] - g["T"]).max() < 1e-9 and abs(o["inlier_rmse"] - g["rmse"]) < 1e-12


@pytest.mark.gpu
def test_compute_features_equals_normals_then_fpfh():
    """Submap::computeFeatures' normals + FPFH in one call on one index = b200_o3d_estimate_normals followed by
    b200_lc_compute_fpfh, bit for bit (the normals never leave the device in between)"""
    p = submap_pair()
    lc = B.LoopClosure()
    est = B.O3dICP(kind=0)
    for pts in (p["t4"], p["s4"][:3000]):
        nrm = est.estimate_normals(pts, knn=10, radius=1.0)
        f, _ = lc.compute_fpfh(pts, nrm, 2.5, 100)
        n2, f2 = lc.compute_features(pts, normal_radius=1.0, normal_knn=10, fpfh_radius=2.5, fpfh_max_nn=100)
        assert np.array_equal(n2, nrm) and np.array_equal(f2, f)
    est.close()
    with pytest.raises(B.B200Error):
        lc.compute_features(p["t4"], normal_knn=64)
