"""CPU tests: pin the oracle (oracle/) against the reference's own goldens and known-answer tests.

Goldens carried over from the reference (see scripts/make_golden_fixtures.py):
  validT3d on car_cloud401 -> car_cloud400        libpointmatcher/utest/utest.cpp:351-360, utest.h:62-87
  icpSingular / icpIdentity known answers          utest.cpp:163-221
  icp_data/defaultIdentityDataPointsFilter         utest.cpp:81-161 (5 % median criterion)
"""
import json
import os

import numpy as np
import pytest

from oracle import oracle_py as O


@pytest.fixture(scope="module")
def gold(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "reference_goldens.json")))
    g["c400"] = np.load(os.path.join(golden_dir, "car_cloud400.npz"))
    g["c401"] = np.load(os.path.join(golden_dir, "car_cloud401.npz"))
    g["cloud0"] = np.load(os.path.join(golden_dir, "cloud.00000.npz"))["xyz"]
    g["cloud1"] = np.load(os.path.join(golden_dir, "cloud.00001.npz"))["xyz"]
    return g


def _angle(Ra, Rb):
    R = Ra.astype(np.float64).T @ Rb.astype(np.float64)
    return float(np.arccos(np.clip((np.trace(R) - 1) / 2, -1, 1)))


def identity_yaml_params(**kw):
    # examples/data/default-identity.yaml
    base = dict(trimmed_ratio=1.0, max_iterations=40, min_diff_rot=0.001, min_diff_trans=0.01, smooth_length=4)
    base.update(kw)
    return O.default_params(**base)


def test_validT3d(gold):
    """IcpHelper::validate3dTransformation with the default chain (utest.h:62-87)."""
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    read4 = O.feat4(gold["c401"]["xyz"])
    V = np.array(gold["validT3d"])
    for accum, mean_mode in ((1, 1), (0, 0)):
        r = O.icp(O.default_params(accum=accum, mean_mode=mean_mode), ref4, refn, read4)
        assert r["status"] == 0
        assert abs(np.linalg.norm(V[:3, 3]) - np.linalg.norm(r["T"][:3, 3])) < 0.1
        assert _angle(V[:3, :3], r["T"][:3, :3]) < 0.1
    # reference-literal arithmetic (float GEMM order, float mean) and the parity arithmetic (double sums) agree
    a = O.icp(O.default_params(accum=1, mean_mode=1), ref4, refn, read4)
    b = O.icp(O.default_params(accum=0, mean_mode=0), ref4, refn, read4)
    assert a["iterations"] == b["iterations"]
    assert np.abs(a["T"] - b["T"]).max() < 1e-5


def test_icp_singular(gold):
    """utest.cpp:163-199: planar grid shifted by 1 m in z, rank-deficient 6x6 -> exact unit z translation."""
    nX, d = 10, np.float32(0.1)
    pts = np.array([[d * x - (nX * d / 2), d * y - (nX * d / 2), 0] for x in range(nX) for y in range(nX)], np.float32)
    pts0 = O.feat4(pts)
    pts1 = pts0.copy()
    pts1[:, 2] = 1
    rf, rn = O.sampling_surface_normal(pts1, knn=10, ratio=1.0, sampling_method=0)
    assert rf.shape[0] == 100
    r = O.icp(identity_yaml_params(), rf, rn, pts0)
    assert r["status"] == 0
    E = np.array(gold["icpSingular_expected"], np.float32)
    # Eigen isApprox, default precision 1e-5 for float
    assert np.linalg.norm(E - r["T"]) <= 1e-5 * min(np.linalg.norm(E), np.linalg.norm(r["T"]))


def test_icp_identity(gold):
    """utest.cpp:201-221: identical clouds -> identity within 1e-4."""
    c0 = O.feat4(gold["cloud0"])
    rf, rn = O.sampling_surface_normal(c0, knn=10, ratio=1.0, sampling_method=0)
    r = O.icp(identity_yaml_params(), rf, rn, c0)
    I = np.eye(4, dtype=np.float32)
    assert np.linalg.norm(r["T"] - I) <= gold["icpIdentity_epsilon"] * np.linalg.norm(I)


def test_ref_trans_default_identity_filter(gold):
    """utest.cpp:81-161 for icp_data/defaultIdentityDataPointsFilter.yaml (deterministic config)."""
    rf, rn = O.sampling_surface_normal(O.feat4(gold["cloud0"]), knn=10, ratio=0.666666, sampling_method=1)
    p = O.default_params(trimmed_ratio=0.75, max_iterations=40, min_diff_rot=0.001, min_diff_trans=0.01, smooth_length=4)
    data = O.feat4(gold["cloud1"])
    r = O.icp(p, rf, rn, data)
    assert r["status"] == 0
    refT = np.array(gold["defaultIdentityDataPointsFilter.ref_trans"])
    D = data.T.astype(np.float64)
    cur = r["T"].astype(np.float64) @ D
    rel = np.median(np.abs(cur - refT @ D)) / np.median(np.abs(cur))
    assert rel < 0.05
    assert rel < 1e-3  # in practice the oracle lands on the stored transform


def _reading_filter(name, f4):
    """the reading-side DataPointsFilter of the config as a mask (semantics: the cited inPlaceFilter bodies)"""
    x = f4[:, :3]
    if name == "defaultRemoveNaNDataPointsFilter":       # RemoveNaN.cpp: drop columns holding a NaN
        return f4[~np.isnan(x).any(1)]
    if name == "defaultMaxDistDataPointsFilter":         # MaxDist.cpp:60-71, dim = -1: keep norm < |maxDist| (200)
        return f4[np.linalg.norm(x, axis=1) < 200.0]
    if name == "defaultPointToPlaneMinDistDataPointsFilter":   # MinDist.cpp, dim = -1: keep norm > |minDist| (1)
        return f4[np.linalg.norm(x, axis=1) > 1.0]
    if name == "defaultBoundingBoxDataPointsFilter":     # BoundingBox.cpp:82-107: xMin 0.2, rest default (-1 .. 1), removeInside
        inside = (x[:, 0] > 0.2) & (x[:, 0] < 1) & (x[:, 1] > -1) & (x[:, 1] < 1) & (x[:, 2] > -1) & (x[:, 2] < 1)
        return f4[~inside]
    if name == "defaultDistanceLimitDataPointsFilter":   # DistanceLimit.cpp:75-93, dim = -1, removeInside 0: keep norm < |dist| (200)
        return f4[np.linalg.norm(x, axis=1) < 200.0]
    if name == "defaultMaxQuantileOnAxisDataPointsFilter":   # MaxQuantileOnAxis.cpp:62-92, dim 0, ratio 0.72
        n = f4.shape[0]
        limit = np.sort(x[:, 0])[int(n * 0.72)]              # nth_element at size * ratio, values[nbPointsOut]
        return f4[x[:, 0] < limit]
    if name == "defaultFixStepSamplingDataPointsFilter":     # FixStepSampling.cpp:72-86, step 10: columns phase, phase + 10, ...
        return f4[3::10]                                     # phase = rand() % 10: glibc's first unseeded rand() is 1804289383
    raise KeyError(name)


@pytest.mark.parametrize("name", ["defaultRemoveNaNDataPointsFilter", "defaultMaxDistDataPointsFilter",
                                  "defaultBoundingBoxDataPointsFilter", "defaultPointToPlaneMinDistDataPointsFilter",
                                  "defaultDistanceLimitDataPointsFilter", "defaultMaxQuantileOnAxisDataPointsFilter",
                                  "defaultFixStepSamplingDataPointsFilter"])
def test_ref_trans_more_deterministic_configs(gold, name):
    """utest.cpp:81-161 for seven more icp_data/*.yaml configs whose reading filter is a plain mask; the rest of the
    chain is the one of defaultIdentityDataPointsFilter.yaml. Criterion as in the reference: relative median error of
    curT * data against refT * data below 5 % (data = the UNFILTERED reading, utest.cpp:147-159)."""
    rf, rn = O.sampling_surface_normal(O.feat4(gold["cloud0"]), knn=10, ratio=0.666666, sampling_method=1)
    p = O.default_params(trimmed_ratio=0.75, max_iterations=40, min_diff_rot=0.001, min_diff_trans=0.01, smooth_length=4)
    data = O.feat4(gold["cloud1"])
    reading = np.ascontiguousarray(_reading_filter(name, data))
    assert 0 < reading.shape[0] <= data.shape[0]
    r = O.icp(p, rf, rn, reading)
    assert r["status"] == 0
    refT = np.array(gold[name + ".ref_trans"])
    D = data.T.astype(np.float64)
    cur = r["T"].astype(np.float64) @ D
    rel = np.median(np.abs(cur - refT @ D)) / np.median(np.abs(cur))
    assert rel < 0.05, rel


@pytest.mark.parametrize("name", ["defaultObservationDirectionDataPointsFilter", "defaultOrientNormalsDataPointsFilter",
                                  "defaultSimpleSensorNoiseDataPointsFilter"])
def test_ref_trans_surface_normal_reference_configs(gold, name):
    """utest.cpp:81-161 for three icp_data/*.yaml configs whose REFERENCE filter is the plain
    SurfaceNormalDataPointsFilter (knn 10: every point kept) and whose reading filters only add descriptors
    (SurfaceNormal, ObservationDirection, OrientNormals, SimpleSensorNoise leave the reading's points untouched and
    point-to-plane uses the reference's normals only): pins the oracle's surface normals + ICP against the reference's
    golden transforms."""
    ref = O.feat4(gold["cloud0"])
    st, sn = O.surface_normals(ref, 10)
    assert st == 0
    p = O.default_params(trimmed_ratio=0.75, max_iterations=40, min_diff_rot=0.001, min_diff_trans=0.01, smooth_length=4)
    data = O.feat4(gold["cloud1"])
    r = O.icp(p, ref, sn["normals"], data)
    assert r["status"] == 0
    refT = np.array(gold[name + ".ref_trans"])
    D = data.T.astype(np.float64)
    cur = r["T"].astype(np.float64) @ D
    rel = np.median(np.abs(cur - refT @ D)) / np.median(np.abs(cur))
    assert rel < 0.05, rel


def test_kdtree_matches_brute_force(gold):
    ref4 = O.feat4(gold["c400"]["xyz"])
    q4 = O.feat4(gold["c401"]["xyz"][:1500])
    kd = O.KdTree(ref4)
    for k, md in ((1, np.inf), (5, np.inf), (3, 0.05), (10, 0.2)):
        i1, d1 = kd.knn(q4, k, md)
        i2, d2 = O.knn_brute(ref4, q4, k, md)
        assert (i1 == i2).all() and (d1.view(np.uint32) == d2.view(np.uint32)).all()
    # threads do not change the result
    i3, d3 = kd.knn(q4, 5, np.inf, threads=4)
    i2, d2 = O.knn_brute(ref4, q4, 5, np.inf)
    assert (i3 == i2).all()


def test_knn_ties_smallest_index():
    ref = np.array([[0, 0, 0], [1, 0, 0], [0, 0, 0], [-1, 0, 0], [1, 0, 0]], np.float32)
    ids, d = O.knn_brute(O.feat4(ref), O.feat4(np.zeros((1, 3), np.float32)), 5)
    assert ids[0].tolist() == [0, 2, 1, 3, 4]
    ids, d = O.knn_brute(O.feat4(ref), O.feat4(np.zeros((1, 3), np.float32)), 3, 0.5)
    assert ids[0].tolist() == [0, 2, -1] and np.isinf(d[0, 2])


def test_quantile_index_arithmetic():
    """Matches.cpp:60-87: invalid entries dropped, index = size * quantile in float, max when quantile == 1."""
    d = np.array([5, 1, np.inf, 3, 2, 4, np.inf], np.float32)
    assert O.dists_quantile(d, 0.5) == (0, 3.0)   # 5 valid, index 2 of [1,2,3,4,5]
    assert O.dists_quantile(d, 1.0) == (0, 5.0)
    assert O.dists_quantile(d, 0.0) == (0, 1.0)
    assert O.dists_quantile(np.full(4, np.inf, np.float32), 0.5)[0] == O.ORC_E_NO_MATCHES
    st, w = O.outlier_trimmed(d, 0.5)
    assert w.tolist() == [0, 1, 0, 1, 1, 0, 0]


def test_solve6_rank_deficient_and_full():
    rng = np.random.default_rng(3)
    J = rng.normal(size=(50, 6)).astype(np.float32)
    H = (J.T @ J).astype(np.float32)
    x_true = rng.normal(size=6).astype(np.float32)
    b = H @ x_true
    x, info = O.solve6(H, b)
    assert info == 0 and np.allclose(x, x_true, rtol=1e-3, atol=1e-4)
    # planar case of icpSingular: rows/cols 2,3,4 exactly zero
    H2 = H.copy()
    H2[[2, 3, 4], :] = 0
    H2[:, [2, 3, 4]] = 0
    b2 = b.copy()
    b2[[2, 3, 4]] = 0
    x2, info2 = O.solve6(H2, b2)
    assert info2 == 1 and np.all(x2[[2, 3, 4]] == 0)
    assert np.allclose(H2 @ x2, b2, rtol=1e-3, atol=1e-3)


def test_p2point_recovers_rigid_motion():
    rng = np.random.default_rng(5)
    P = rng.normal(size=(200, 3)).astype(np.float32)
    th = 0.3
    R = np.array([[np.cos(th), -np.sin(th), 0], [np.sin(th), np.cos(th), 0], [0, 0, 1]], np.float32)
    t = np.array([0.1, -0.2, 0.3], np.float32)
    Q = P @ R.T + t
    T = O.p2point_T(O.feat4(P), O.feat4(Q), np.ones(200, np.float32))
    assert np.allclose(T[:3, :3], R, atol=1e-5) and np.allclose(T[:3, 3], t, atol=1e-5)


def test_surface_normals_on_plane():
    rng = np.random.default_rng(7)
    xy = rng.uniform(-1, 1, size=(500, 2))
    pts = np.c_[xy, 0.3 * xy[:, 0] + 0.1].astype(np.float32)
    st, out = O.surface_normals(O.feat4(pts), knn=8)
    n_true = np.array([-0.3, 0, 1]) / np.linalg.norm([-0.3, 0, 1])
    assert st == 0
    assert (np.abs(out["normals"] @ n_true) > 1 - 1e-4).all()
    assert (out["ids"][:, 0] == np.arange(500)).all()  # self match first (ALLOW_SELF_MATCH)


def test_o3d_estimate_normals_oracle_against_numpy(golden_dir):
    """the oracle's Open3D EstimateNormals restatement (count-normalised double covariance of the hybrid neighbourhood,
    smallest eigenvector, < 3 neighbours -> (0,0,1), oriented to the origin) against numpy on the same neighbourhoods"""
    f = O.feat4(np.load(os.path.join(golden_dir, "car_cloud400.npz"))["xyz"])
    f = np.concatenate([f, O.feat4(np.array([[50, 50, 50]], np.float32))])
    n, cnt = O.o3d_estimate_normals(f, 10, 2.0)
    ids, d = O.KdTree(f).knn(f, 10, 2.0, threads=4)
    checked = 0
    for i in range(0, f.shape[0] - 1, 41):
        nb = f[ids[i][ids[i] >= 0]][:, :3].astype(np.float64)
        assert cnt[i] == nb.shape[0] and ids[i, 0] == i          # the point itself is its first neighbour
        if nb.shape[0] < 3:                                       # Identity covariance -> (0, 0, 1), then oriented
            assert np.array_equal(np.abs(n[i]), [0.0, 0.0, 1.0])
            continue
        w, v = np.linalg.eigh(np.cov(nb.T, bias=True))
        if (w[1] - w[0]) < 1e-6 * w[2]:
            continue
        e = v[:, 0] if v[:, 0] @ (-f[i, :3].astype(np.float64)) >= 0 else -v[:, 0]
        assert np.abs(e - n[i]).max() < 1e-7
        checked += 1
    assert checked > 400
    assert cnt[-1] == 1 and np.array_equal(n[-1], [0.0, 0.0, -1.0])


# ---- the Open3D registration oracle (oracle/o3d.cpp, recalled from upstream) against an independent numpy restatement
def _np_rot_e1_to_x(n):
    """GetRotationFromE1ToX: the rotation that takes e1 to the unit vector n (Rodrigues; identity when e1 . n < -0.99)"""
    n = n / np.linalg.norm(n)
    c = n[0]
    if c < -0.99:
        return np.eye(3)
    v = np.cross([1.0, 0.0, 0.0], n)
    sk = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])
    return np.eye(3) + sk + sk @ sk / (1.0 + c)


def _np_vec6_to_T(x):
    """TransformVector6dToMatrix4d: R = Rz(x2) Ry(x1) Rx(x0), t = x[3:6]"""
    cx, sx, cy, sy, cz, sz = np.cos(x[0]), np.sin(x[0]), np.cos(x[1]), np.sin(x[1]), np.cos(x[2]), np.sin(x[2])
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    T = np.eye(4)
    T[:3, :3] = Rz @ Ry @ Rx
    T[:3, 3] = x[3:6]
    return T


def _np_o3d_update(kind, src, tgt, tn, sn, T, md, eps=1e-3):
    """one RegistrationICP / RegistrationGeneralizedICP update from pose T, with numpy only: brute-force 1-NN of the
    transformed source (rounded to float like the cloud handed to the kd-tree) within md, then the estimation step of
    TransformationEstimationPointToPlane / PointToPoint / ForGeneralizedICP; returns (ids, U @ T)"""
    q = (src.astype(np.float64) @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
    t32 = tgt.astype(np.float32)
    ids = np.full(len(q), -1, np.int64)
    for i in range(len(q)):
        d = (t32 - q[i]) ** 2
        d2 = (d[:, 0] + d[:, 1]) + d[:, 2]
        j = int(np.argmin(d2))
        if d2[j] <= np.float32(md) * np.float32(md):
            ids[i] = j
    m = ids >= 0
    s, t = q[m].astype(np.float64), tgt[ids[m]].astype(np.float64)
    if kind == 1:                                # Eigen::umeyama without scaling
        ms, mt = s.mean(0), t.mean(0)
        U_, _, Vt = np.linalg.svd((t - mt).T @ (s - ms) / len(s))
        S = np.diag([1.0, 1.0, np.sign(np.linalg.det(U_) * np.linalg.det(Vt))])
        R = U_ @ S @ Vt
        U = np.eye(4)
        U[:3, :3] = R
        U[:3, 3] = mt - R @ ms
        return ids, U @ T
    JTJ, JTr = np.zeros((6, 6)), np.zeros(6)
    if kind == 0:
        n = tn[ids[m]].astype(np.float64)
        J = np.hstack([np.cross(s, n), n])
        r = ((s - t) * n).sum(1)
        JTJ, JTr = J.T @ J, J.T @ r
    else:
        R = T[:3, :3]
        for a in range(len(s)):
            Ct = _np_rot_e1_to_x(tn[ids[m][a]].astype(np.float64)) @ np.diag([eps, 1, 1]) @ _np_rot_e1_to_x(tn[ids[m][a]].astype(np.float64)).T
            Rs = _np_rot_e1_to_x(sn[np.nonzero(m)[0][a]].astype(np.float64))
            Cs = R @ (Rs @ np.diag([eps, 1, 1]) @ Rs.T) @ R.T
            w, V = np.linalg.eigh(Ct + Cs)
            W = V @ np.diag(1.0 / np.sqrt(w)) @ V.T          # M^(-1/2)
            sx = np.array([[0, -s[a][2], s[a][1]], [s[a][2], 0, -s[a][0]], [-s[a][1], s[a][0], 0]])
            J = W @ np.hstack([-sx, np.eye(3)])              # three rows
            r = W @ (s[a] - t[a])
            JTJ += J.T @ J
            JTr += J.T @ r
    x = np.linalg.solve(JTJ, -JTr)
    return ids, _np_vec6_to_T(x) @ T


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_o3d_registration_oracle_against_numpy(gold, kind):
    """The oracle of boundary B is recalled from Open3D upstream (no copy of Open3D in the reference tree): here one
    update of each estimation kind is re-derived with numpy from the published formulas — correspondences identical,
    updated pose within 1e-9 — and the loop's fixed point is checked on a known motion."""
    rng = np.random.default_rng(4)
    c400 = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "car_cloud400.npz"))
    sel = rng.choice(len(c400["xyz"]), 1500, replace=False)
    tgt, tn = c400["xyz"][sel].astype(np.float32), c400["normals"][sel].astype(np.float32)
    T_true = _np_vec6_to_T(np.array([0.01, -0.02, 0.015, 0.05, -0.03, 0.02]))
    Ti = np.linalg.inv(T_true)
    keep = rng.choice(len(tgt), 700, replace=False)
    src = (tgt[keep].astype(np.float64) @ Ti[:3, :3].T + Ti[:3, 3]).astype(np.float32)
    sn = (tn[keep].astype(np.float64) @ Ti[:3, :3].T).astype(np.float32)
    init = _np_vec6_to_T(np.array([0.002, 0.001, -0.003, 0.01, 0.0, -0.01]))
    o = O.o3d_icp(kind, O.feat4(src), O.feat4(tgt), tn, sn if kind == 2 else None, max_dist=0.5, init=init, max_iteration=1)
    ids, T1 = _np_o3d_update(kind, src, tgt, tn, sn, init, 0.5)
    # the oracle reports the correspondences of its LAST match (after the update); those of the first are checked
    # through the update they produce
    assert o["iterations"] == 1
    assert np.abs(o["T"] - T1).max() < 1e-9, np.abs(o["T"] - T1).max()
    ids2, _ = _np_o3d_update(kind, src, tgt, tn, sn, o["T"], 0.5)
    assert (o["ids"] == ids2).all()
    assert abs(o["fitness"] - (ids2 >= 0).mean()) < 1e-15
    # free running: the exact copy is brought back onto the target
    r = O.o3d_icp(kind, O.feat4(src), O.feat4(tgt), tn, sn if kind == 2 else None, max_dist=0.5, init=init, max_iteration=50)
    assert np.abs(r["T"] - T_true).max() < (2e-3 if kind == 1 else 1e-4), np.abs(r["T"] - T_true).max()
    assert r["fitness"] == 1.0


def test_o3d_information_matrix_oracle_against_numpy():
    rng = np.random.default_rng(6)
    tgt = rng.uniform(-3, 3, (400, 3)).astype(np.float32)
    src = (tgt[:300] + 0.01 * rng.standard_normal((300, 3))).astype(np.float32)
    src[:20] += 5.0                                           # no correspondence within 0.3
    T = _np_vec6_to_T(np.array([0.001, 0.002, -0.001, 0.005, -0.004, 0.003]))
    info, fit, rmse, ids = O.o3d_information_matrix(O.feat4(src), O.feat4(tgt), T, 0.3)
    q = (src.astype(np.float64) @ T[:3, :3].T + T[:3, 3]).astype(np.float32)
    G = np.zeros((6, 6))
    n, s2 = 0, 0.0
    for i in range(len(q)):
        d = (tgt - q[i]) ** 2
        d2 = (d[:, 0] + d[:, 1]) + d[:, 2]
        j = int(np.argmin(d2))
        if d2[j] > np.float32(0.3) * np.float32(0.3):
            assert ids[i] == -1
            continue
        assert ids[i] == j
        x, y, z = tgt[j].astype(np.float64)
        A = np.array([[0, z, -y, 1, 0, 0], [-z, 0, x, 0, 1, 0], [y, -x, 0, 0, 0, 1.0]])
        G += A.T @ A
        n += 1
        s2 += float(d2[j])
    assert n == 280 and abs(fit - n / 300) < 1e-15 and abs(rmse - np.sqrt(s2 / n)) < 1e-9
    assert np.abs(info - G).max() < 1e-9 * np.abs(G).max()
