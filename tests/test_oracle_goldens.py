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
