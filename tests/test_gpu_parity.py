"""GPU parity tests: the CUDA path (through the C ABI, ctypes) against the CPU oracle on the same inputs.

Bars (north_star): kNN ids / dists and outlier masks BIT-EXACT; H, b, poses within float tolerance
(rotation < 1e-5 rad, translation < 1e-5 m after convergence) with identical iteration counts.
"""
import json
import os

import numpy as np
import pytest

from oracle import oracle_py as O
from open3d_slam_private_b200 import capi as B

pytestmark = pytest.mark.gpu

ROT_TOL = 1e-5   # rad
TRANS_TOL = 1e-5  # m


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def rot_angle(Ra, Rb):
    """angle between two rotations; via the chord |Ra - Rb|_F = 2*sqrt(2)*sin(theta/2), which stays accurate at 1e-7
    rad where arccos((tr - 1) / 2) loses everything to the float32 rounding of the inputs"""
    c = np.linalg.norm(Ra.astype(np.float64) - Rb.astype(np.float64))
    return float(2.0 * np.arcsin(min(1.0, c / (2.0 * np.sqrt(2.0)))))


@pytest.fixture(scope="module")
def gold(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "reference_goldens.json")))
    g["c400"] = np.load(os.path.join(golden_dir, "car_cloud400.npz"))
    g["c401"] = np.load(os.path.join(golden_dir, "car_cloud401.npz"))
    g["cloud0"] = np.load(os.path.join(golden_dir, "cloud.00000.npz"))["xyz"]
    g["cloud1"] = np.load(os.path.join(golden_dir, "cloud.00001.npz"))["xyz"]
    return g


def to_b(p):
    """oracle IcpParams -> C-ABI IcpParams"""
    return B.default_icp_params(knn=p.knn, maxDist=p.maxDist, use_trimmed=p.use_trimmed, trimmed_ratio=p.trimmed_ratio,
                                use_maxdist_filter=p.use_maxdist_filter, maxdist_filter=p.maxdist_filter,
                                use_surface_normal_filter=p.use_surface_normal_filter,
                                surface_normal_max_angle=p.surface_normal_max_angle, minimizer=p.minimizer,
                                max_iterations=p.max_iterations, use_differential=p.use_differential,
                                min_diff_rot=p.min_diff_rot, min_diff_trans=p.min_diff_trans,
                                smooth_length=p.smooth_length)


# ----------------------------------------------------------------------------------------------------- kNN
@pytest.mark.parametrize("k,md", [(1, np.inf), (1, 0.5), (1, 0.05), (2, np.inf), (3, 0.5), (5, np.inf), (10, 0.3),
                                  (20, np.inf), (33, 1.0)])
def test_knn_bit_exact_car_clouds(gold, k, md):
    """ui/Matcher.cpp:40-64 sweeps knn x epsilon x maxDist on these clouds; here ids/dists are pinned bit-exact."""
    ref4 = O.feat4(gold["c400"]["xyz"])
    q4 = O.feat4(gold["c401"]["xyz"])
    idx = B.Index(ref4)
    ids, d = idx.find_closests(q4, knn=k, max_dist=md)
    oi, od = O.KdTree(ref4).knn(q4, k, md, threads=4)
    assert (ids == oi).all()
    assert (bits(d) == bits(od)).all()
    assert idx.visit_counter > 0


@pytest.mark.parametrize("cell", [0.0, 0.02, 0.3, 5.0])
def test_knn_independent_of_cell_size(gold, cell):
    ref4 = O.feat4(gold["c400"]["xyz"][::3])
    q4 = O.feat4(gold["c401"]["xyz"][::5])
    ids, d = B.Index(ref4, cell_size=cell).find_closests(q4, knn=4, max_dist=np.inf)
    oi, od = O.KdTree(ref4).knn(q4, 4, np.inf)
    assert (ids == oi).all() and (bits(d) == bits(od)).all()


def test_knn_ties_duplicates_and_outside_queries():
    rng = np.random.default_rng(11)
    base = rng.integers(-20, 20, size=(3000, 3)).astype(np.float32) * 0.25  # lattice => many exact ties
    ref = np.concatenate([base, base[:500], base[:100]])                    # duplicate points
    ref4 = O.feat4(ref)
    q = np.concatenate([rng.integers(-24, 24, size=(2000, 3)).astype(np.float32) * 0.25,
                        rng.uniform(-200, 200, size=(200, 3)).astype(np.float32),       # far outside the grid
                        np.array([[1e6, 0, 0], [0, -1e6, 1e6]], np.float32)])
    q4 = O.feat4(q)
    idx = B.Index(ref4)
    for k, md in ((1, np.inf), (4, np.inf), (7, 1.0), (1, 0.1)):
        ids, d = idx.find_closests(q4, knn=k, max_dist=md)
        oi, od = O.knn_brute(ref4, q4, k, md)
        assert (ids == oi).all(), (k, md)
        assert (bits(d) == bits(od)).all()


def test_knn_edge_cases():
    ref4 = O.feat4(np.array([[0, 0, 0], [1, 1, 1]], np.float32))
    idx = B.Index(ref4)
    ids, d = idx.find_closests(O.feat4(np.array([[0.1, 0, 0]], np.float32)), knn=5)   # k > M
    assert ids[0].tolist() == [0, 1, -1, -1, -1] and np.isinf(d[0, 2:]).all()
    ids, d = idx.find_closests(np.zeros((0, 4), np.float32), knn=1)                    # empty query
    assert ids.shape == (0, 1)
    ids, d = idx.find_closests(O.feat4(np.array([[np.nan, 0, 0]], np.float32)), knn=1)  # NaN query: no match
    assert ids[0, 0] == -1
    one = B.Index(O.feat4(np.array([[2, 3, 4]], np.float32)))                          # single point, degenerate bbox
    ids, d = one.find_closests(O.feat4(np.array([[2, 3, 5]], np.float32)), knn=1)
    assert ids[0, 0] == 0 and d[0, 0] == 1.0
    with pytest.raises(B.B200Error):
        idx.find_closests(ref4, knn=65)


def test_knn_var_dist(gold):
    """KDTreeVarDistMatcher (MatchersImpl.cpp:133-150): per-point radius."""
    ref4 = O.feat4(gold["c400"]["xyz"])
    q4 = O.feat4(gold["c401"]["xyz"][:4000])
    rng = np.random.default_rng(2)
    radii = rng.uniform(0.01, 0.3, q4.shape[0]).astype(np.float32)
    ids, d = B.Index(ref4).find_closests(q4, knn=2, max_dist_per_point=radii)
    kd = O.KdTree(ref4)
    for i in range(0, 4000, 97):
        oi, od = kd.knn(q4[i:i + 1], 2, float(radii[i]))
        assert (ids[i] == oi[0]).all() and (bits(d[i]) == bits(od[0])).all()


# ----------------------------------------------------------------------------------------------------- outlier filters
def test_outlier_filters_bit_exact(gold):
    ref4 = O.feat4(gold["c400"]["xyz"])
    q4 = O.feat4(gold["c401"]["xyz"])
    ids, d = O.KdTree(ref4).knn(q4, 3, 0.08, threads=4)   # contains invalid (inf) entries
    assert np.isinf(d).any()
    for q in (0.0, 0.1, 0.5, 0.85, 0.9, 0.999999, 1.0):
        st, v = O.dists_quantile(d, q)
        assert st == 0 and B.dists_quantile(d, q) == v
        st, w = O.outlier_trimmed(d, q)
        assert (B.outlier_dist(B.OUTLIER_TRIMMEDDIST, q, d) == w).all()
    assert (B.outlier_dist(B.OUTLIER_MAXDIST, 0.05, d) == O.outlier_maxdist(d, 0.05)[1]).all()
    assert (B.outlier_dist(B.OUTLIER_MINDIST, 0.02, d) == O.outlier_mindist(d, 0.02)[1]).all()
    assert (B.outlier_dist(B.OUTLIER_MEDIANDIST, 3.0, d) == O.outlier_median(d, 3.0)[1]).all()
    assert (B.outlier_dist(B.OUTLIER_NULL, 0.0, d) == O.outlier_null(d)[1]).all()
    # all invalid -> ConvergenceError (Matches.cpp:75-76)
    with pytest.raises(B.ConvergenceError):
        B.outlier_dist(B.OUTLIER_TRIMMEDDIST, 0.9, np.full((10, 1), np.inf, np.float32))
    # many equal values around the quantile
    dd = np.repeat(np.array([0.5, 0.25, 0.25, 1.0], np.float32), 1000)
    for q in (0.3, 0.5, 0.74, 0.76):
        assert B.dists_quantile(dd, q) == O.dists_quantile(dd, q)[1]


def test_outlier_surface_normal_bit_exact(gold):
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    q4 = O.feat4(gold["c401"]["xyz"])
    ids, d = O.KdTree(ref4).knn(q4, 2, 0.1, threads=4)
    rng = np.random.default_rng(4)
    nread = rng.normal(size=(q4.shape[0], 3)).astype(np.float32)
    for ang in (0.3, 1.0, 1.57):
        st, w = O.outlier_surface_normal(nread, refn, ids, ang)
        assert (B.outlier_surface_normal(nread, refn, ids, ang) == w).all()
    assert (B.outlier_surface_normal(None, None, ids, 1.0) == 1).all()


# ----------------------------------------------------------------------------------------------------- error minimizer
@pytest.fixture(scope="module")
def matched(gold):
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    q4 = O.feat4(gold["c401"]["xyz"])
    ids, d = O.KdTree(ref4).knn(q4, 2, 0.15, threads=4)
    st, w = O.outlier_trimmed(d, 0.8)
    return ref4, refn, q4, ids, d, w


def test_error_elements_bit_exact(matched):
    ref4, refn, q4, ids, d, w = matched
    st, o = O.error_elements(q4, ref4, refn, w, ids, d)
    g = B.error_elements(q4, ref4, refn, w, ids, d)
    assert g["reading"].shape == o["reading"].shape
    for key in ("reading", "reference", "ref_normals", "weights", "dists"):
        assert (bits(g[key]) == bits(o[key])).all(), key
    assert (g["ids"] == o["ids"]).all()
    assert g["point_used_ratio"] == o["point_used_ratio"]
    assert g["weighted_point_used_ratio"] == o["weighted_point_used_ratio"]
    with pytest.raises(B.ConvergenceError):
        B.error_elements(q4, ref4, refn, np.zeros_like(w), ids, d)


def test_hessian_solve_and_p2point(matched):
    ref4, refn, q4, ids, d, w = matched
    st, o = O.error_elements(q4, ref4, refn, w, ids, d)
    H, b = O.p2plane_hb(o["reading"], o["reference"], o["ref_normals"], o["weights"], accum=1)
    Hg, bg = B.p2plane_hessian(o["reading"], o["reference"], o["ref_normals"], o["weights"])
    assert np.abs(Hg - H).max() <= 2e-6 * np.abs(H).max()
    assert np.abs(bg - b).max() <= 2e-6 * max(np.abs(b).max(), 1e-3)
    x, info = O.solve6(H, b)
    xg, Tg, infog = B.p2plane_solve(H, b)
    assert info == infog == 0
    assert np.allclose(xg, x, rtol=1e-5, atol=1e-7)
    T = O.p2plane_T(x)
    assert np.abs(Tg - T).max() < 1e-6
    # rank-deficient (icpSingular-like) system takes the same branch
    H2 = H.copy()
    H2[[2, 3, 4], :] = 0
    H2[:, [2, 3, 4]] = 0
    b2 = b.copy()
    b2[[2, 3, 4]] = 0
    x2, i2 = O.solve6(H2, b2)
    xg2, _, ig2 = B.p2plane_solve(H2, b2)
    assert i2 == ig2 == 1 and np.allclose(xg2, x2, rtol=1e-4, atol=1e-6)
    # zero system -> identity (NaN guard, PointToPlane.cpp:374-380)
    xz, Tz, _ = B.p2plane_solve(np.zeros((6, 6), np.float32), np.zeros(6, np.float32))
    assert (Tz == np.eye(4)).all()
    # device sums in double: tight against the oracle's double-accumulated variant, and within the stated pose
    # tolerance of the reference-literal float-sequential sums (whose order is an Eigen artefact)
    Tpg = B.p2point(o["reading"], o["reference"], o["weights"])
    assert np.abs(Tpg - O.p2point_T(o["reading"], o["reference"], o["weights"], accum=1)).max() < 2e-6
    assert np.abs(Tpg - O.p2point_T(o["reading"], o["reference"], o["weights"], accum=0)).max() < 5e-5


def test_solve_rank_threshold_band():
    """The warp solve's rank certificate (lambda_min(A) > 2.8 x 6 eps_f x the largest column norm) and, where it does
    not decide, the exact replay of the reference's float FullPivHouseholderQR rank test must reach the oracle's branch
    (PointToPlane.cpp:199-256) for matrices whose smallest eigenvalue sits anywhere around that threshold."""
    rng = np.random.default_rng(17)
    eps = np.finfo(np.float32).eps
    seen = set()
    for trial in range(40):
        Q, _ = np.linalg.qr(rng.normal(size=(6, 6)))
        lam = np.array([1.0, 0.8, 0.6, 0.5, 0.3, 0.0]) * rng.uniform(0.5, 50.0)
        cmax = lam[0]
        for t in (0.05, 0.3, 0.6, 0.9, 1.0, 1.1, 1.5, 2.0, 2.8, 3.2, 4.0, 8.0, 100.0):
            lam[5] = t * 6 * eps * cmax
            H = ((Q * lam) @ Q.T).astype(np.float32)
            H = ((H + H.T) * np.float32(0.5)).astype(np.float32)
            b = (H.astype(np.float64) @ rng.normal(size=6)).astype(np.float32)
            x, info = O.solve6(H, b)
            xg, _, infog = B.p2plane_solve(H, b)
            assert info == infog, (trial, t, info, infog)
            seen.add(info)
            if info == 0 and t >= 8.0:      # well-conditioned enough for the solutions to be comparable entry by entry
                assert np.allclose(xg, x, rtol=2e-3, atol=1e-5), (trial, t)
            else:                            # near-singular: compare through the residual the reference's isApprox uses
                r0 = np.linalg.norm(H.astype(np.float64) @ x - b)
                r1 = np.linalg.norm(H.astype(np.float64) @ xg - b)
                assert r1 <= r0 * 1.5 + 1e-4 * max(np.linalg.norm(b), 1e-6), (trial, t, r0, r1)
    assert 0 in seen and len(seen) >= 2      # both sides of the rank decision were exercised


def test_rigid_apply_bit_exact(gold):
    rng = np.random.default_rng(9)
    f = O.feat4(gold["cloud0"])
    n = rng.normal(size=(f.shape[0], 3)).astype(np.float32)
    T = O.p2plane_T(np.array([0.02, -0.01, 0.03, 0.5, -0.2, 0.1], np.float32))
    for Tm in (T, np.diag([1.1, 1, 1, 1]).astype(np.float32) @ T):   # second one triggers correctParameters
        fo, no = O.rigid_apply(Tm, f, n)
        fg, ng = B.rigid_apply(Tm, f, n)
        assert (bits(fg) == bits(fo)).all() and (bits(ng) == bits(no)).all()


# ----------------------------------------------------------------------------------------------------- surface normals
@pytest.mark.parametrize("knn,md,sort", [(5, np.inf, 0), (10, np.inf, 1), (20, 1.0, 0)])
def test_surface_normals(gold, knn, md, sort):
    f = O.feat4(gold["cloud0"])
    st, o = O.surface_normals(f, knn, md, sort_eigen=sort)
    g = B.surface_normals(f, knn, md, sort_eigen=sort,
                          keep=("normals", "densities", "eigValues", "eigVectors", "matchedIds", "meanDists"))
    assert (g["matchedIds"] == o["matchedIds"]).all()          # self-kNN ids bit-exact
    assert (bits(g["densities"]) == bits(o["densities"])).all()
    assert (bits(g["meanDists"]) == bits(o["meanDists"])).all()
    assert np.allclose(g["eigValues"], o["eigValues"], rtol=1e-5, atol=1e-9)
    # normals modulo sign (eigenvector sign is an artefact of the eigen solver)
    dots = np.abs(np.sum(g["normals"] * o["normals"], axis=1))
    nz = np.linalg.norm(o["normals"], axis=1) > 0
    assert (dots[nz] > 1 - 1e-5).mean() > 0.999   # near-degenerate eigenvalue pairs may rotate within the eigenspace
    assert (np.linalg.norm(g["normals"][~nz], axis=1) == 0).all()


# ----------------------------------------------------------------------------------------------------- whole ICP
def run_both(p, ref4, refn, read4, readn=None, T_init=None):
    o = O.icp(p, ref4, refn, read4, readn, T_init)
    icp = B.ICP(to_b(p))
    icp.init_reference(ref4, refn)
    g = icp.compute(read4, readn, T_init, raise_on_error=False)
    icp.inputs = (p, ref4, refn, read4, readn, T_init)
    return o, g, icp


STEP_TOL = 2e-6  # |T_iter(device) - T_iter(oracle)| after ONE iteration from the same pose (float32 4x4 entries)


def assert_icp_parity(o, g, icp, pose_tol_scale=1.0):
    """Two layers. (1) Free-running: same status, iteration count, stop reason, final pose within the north_star
    tolerance. (2) One-step ("teacher-forced", orc_icp_forced): the oracle recomputes every iteration from the pose the
    device held at that iteration, so the inputs really are the same: correspondences (ids, dists) and outlier masks
    of the last iteration are then BIT-EXACT, P matches at every iteration, and H, b and the updated pose agree within
    float rounding."""
    assert g["status"] == o["status"]
    assert g["iterations"] == o["iterations"], (g["iterations"], o["iterations"])
    assert g["max_iterations_reached"] == o["max_iterations_reached"]
    assert np.abs(g["T"][:3, 3] - o["T"][:3, 3]).max() < TRANS_TOL * pose_tol_scale
    assert rot_angle(g["T"][:3, :3], o["T"][:3, :3]) < ROT_TOL * pose_tol_scale
    lg = icp.log()
    n = len(lg["P"])
    assert n == min(g["iterations"], 64)
    if n == 0 or n < g["iterations"]:
        return
    forced = getattr(icp, "oracle_forced", None)   # a kept oracle session (large maps: no kd-tree rebuild)
    of = forced(lg["T_iters"]) if forced else O.icp(*icp.inputs, forced_T_iters=lg["T_iters"])
    assert of["iterations"] == n
    ids, d, w = icp.matches()
    assert (ids == of["ids"]).all()
    assert (bits(d) == bits(of["dists"])).all()
    assert (w == of["weights"]).all()
    assert (lg["P"] == of["P"]).all()
    assert g["n_matches_last"] == of["n_matches_last"]
    for i in range(n):
        assert np.abs(lg["T_iters"][i] - of["T_iters"][i]).max() < STEP_TOL, i
        if icp.inputs[0].minimizer == 0:
            hs = np.abs(of["H"][i]).max()
            assert np.abs(lg["H"][i] - of["H"][i]).max() <= 2e-6 * hs, i
            assert np.abs(lg["b"][i] - of["b"][i]).max() <= 2e-6 * max(np.abs(of["b"][i]).max(), 1e-3 * hs), i


def test_icp_validT3d_and_parity(gold):
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    read4 = O.feat4(gold["c401"]["xyz"])
    o, g, icp = run_both(O.default_params(), ref4, refn, read4)
    assert_icp_parity(o, g, icp)
    V = np.array(gold["validT3d"])
    assert abs(np.linalg.norm(V[:3, 3]) - np.linalg.norm(g["T"][:3, 3])) < 0.1   # utest.h:83-84
    assert rot_angle(V[:3, :3], g["T"][:3, :3]) < 0.1


@pytest.mark.parametrize("kw", [
    dict(maxDist=0.5, trimmed_ratio=0.9, max_iterations=30, min_diff_trans=0.008),           # param/icp.yaml (eps 0)
    dict(use_trimmed=0, use_maxdist_filter=1, maxdist_filter=0.3),
    dict(use_trimmed=0),                                                                      # empty filter chain
    dict(knn=3, trimmed_ratio=0.7),
    dict(knn=20, trimmed_ratio=0.8, max_iterations=6),                                        # k-best list of 32
    dict(knn=40, trimmed_ratio=0.8, maxDist=2.0, max_iterations=4),                           # k-best list of 64
    dict(minimizer=1, trimmed_ratio=0.8),                                                     # point-to-point
    dict(use_differential=0, max_iterations=7),
    dict(trimmed_ratio=1.0, maxDist=0.2),
])
def test_icp_chain_variants(gold, kw):
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    read4 = O.feat4(gold["c401"]["xyz"])
    o, g, icp = run_both(O.default_params(**kw), ref4, refn, read4)
    assert_icp_parity(o, g, icp)


def test_device_pipeline_reaches_the_reference_goldens(gold):
    """The reference's own end-to-end goldens (icp_data/*.ref_trans, criterion of utest.cpp:147-159) with NOTHING from
    the oracle on the path: SurfaceNormalDataPointsFilter (knn 10) on the device for the reference cloud, then the
    registration on the device — the three configs whose reading filters only add descriptors."""
    ref4, data = O.feat4(gold["cloud0"]), O.feat4(gold["cloud1"])
    nrm = B.surface_normals(ref4, knn=10)["normals"]
    icp = B.ICP(B.default_icp_params(trimmed_ratio=0.75, max_iterations=40, min_diff_rot=0.001, min_diff_trans=0.01,
                                     smooth_length=4))
    icp.init_reference(ref4, nrm)
    r = icp.compute(data)
    assert r["status"] == 0
    D = data.T.astype(np.float64)
    cur = r["T"].astype(np.float64) @ D
    for name in ("defaultObservationDirectionDataPointsFilter", "defaultOrientNormalsDataPointsFilter",
                 "defaultSimpleSensorNoiseDataPointsFilter"):
        refT = np.array(gold[name + ".ref_trans"])
        rel = np.median(np.abs(cur - refT @ D)) / np.median(np.abs(cur))
        assert rel < 0.05, (name, rel)


def test_icp_surface_normal_filter(gold):
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    read4 = O.feat4(gold["c401"]["xyz"])
    st, sn = O.surface_normals(read4, 10)
    p = O.default_params(maxDist=0.5, trimmed_ratio=0.9, use_surface_normal_filter=1, surface_normal_max_angle=1.0)
    o, g, icp = run_both(p, ref4, refn, read4, sn["normals"])
    assert_icp_parity(o, g, icp)


def test_icp_singular_and_identity(gold):
    pid = O.default_params(trimmed_ratio=1.0, max_iterations=40, min_diff_rot=0.001, min_diff_trans=0.01, smooth_length=4)
    nX, d = 10, np.float32(0.1)
    pts = np.array([[d * x - (nX * d / 2), d * y - (nX * d / 2), 0] for x in range(nX) for y in range(nX)], np.float32)
    pts0 = O.feat4(pts)
    pts1 = pts0.copy()
    pts1[:, 2] = 1
    rf, rn = O.sampling_surface_normal(pts1, knn=10, ratio=1.0, sampling_method=0)
    o, g, icp = run_both(pid, rf, rn, pts0)
    assert_icp_parity(o, g, icp)
    E = np.array(gold["icpSingular_expected"], np.float32)
    assert np.linalg.norm(E - g["T"]) <= 1e-5 * np.linalg.norm(E)          # utest.cpp:195-198
    c0 = O.feat4(gold["cloud0"])
    rf, rn = O.sampling_surface_normal(c0, knn=10, ratio=1.0, sampling_method=0)
    o, g, icp = run_both(pid, rf, rn, c0)
    assert_icp_parity(o, g, icp)
    assert np.linalg.norm(g["T"] - np.eye(4)) <= 1e-4 * 2                   # utest.cpp:219


def test_icp_ref_trans_golden(gold):
    rf, rn = O.sampling_surface_normal(O.feat4(gold["cloud0"]), knn=10, ratio=0.666666, sampling_method=1)
    p = O.default_params(trimmed_ratio=0.75, max_iterations=40, min_diff_rot=0.001, min_diff_trans=0.01, smooth_length=4)
    data = O.feat4(gold["cloud1"])
    o, g, icp = run_both(p, rf, rn, data)
    assert_icp_parity(o, g, icp)
    refT = np.array(gold["defaultIdentityDataPointsFilter.ref_trans"])
    D = data.T.astype(np.float64)
    cur = g["T"].astype(np.float64) @ D
    assert np.median(np.abs(cur - refT @ D)) / np.median(np.abs(cur)) < 0.05   # utest.cpp:147-159


def test_icp_errors(gold):
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    read4 = O.feat4(gold["c401"]["xyz"])
    icp = B.ICP(B.default_icp_params())
    with pytest.raises(B.B200Error):     # no reference yet
        icp.compute(read4)
    with pytest.raises(B.B200Error):     # point-to-plane without normals (InvalidField, ICP.cpp:1519)
        icp.init_reference(ref4, None)
    icp.init_reference(ref4, refn)
    far = read4.copy()
    far[:, :3] += 1000
    icp2 = B.ICP(B.default_icp_params(maxDist=0.1))
    icp2.init_reference(ref4, refn)
    with pytest.raises(B.ConvergenceError):   # no matches at all
        icp2.compute(far)
    with pytest.raises(B.B200Error):
        B.ICP(B.default_icp_params(max_iterations=0, use_differential=0))


def test_icp_with_initial_guess_and_reuse(gold):
    """Mapper.cpp:343,373 pattern: one initReference, many compute() calls with a prior."""
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    read4 = O.feat4(gold["c401"]["xyz"])
    p = O.default_params(maxDist=0.5, trimmed_ratio=0.9, max_iterations=30, min_diff_trans=0.008)
    icp = B.ICP(to_b(p))
    icp.init_reference(ref4, refn)
    for seed in range(3):
        rng = np.random.default_rng(seed)
        Ti = O.p2plane_T(np.r_[rng.normal(0, 0.02, 3), rng.normal(0, 0.05, 3)].astype(np.float32))
        o = O.icp(p, ref4, refn, read4, None, Ti)
        g = icp.compute(read4, None, Ti, raise_on_error=False)
        icp.inputs = (p, ref4, refn, read4, None, Ti)
        assert_icp_parity(o, g, icp)
        icp.upload_reading(read4)
        g2 = icp.run_resident(Ti)
        assert (g2["T"] == g["T"]).all() and g2["iterations"] == g["iterations"]   # deterministic


def test_icp_bound_checker(gold):
    """BoundTransformationChecker (TransformationCheckersImpl.cpp:165-222): generous bounds change nothing, tight
    bounds raise ConvergenceError at the iteration the oracle raises it"""
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    read4 = O.feat4(gold["c401"]["xyz"])
    base = O.icp(O.default_params(), ref4, refn, read4)

    def bounded(p):
        q = to_b(p)
        q.use_bound, q.max_rotation_norm, q.max_translation_norm = p.use_bound, p.max_rotation_norm, p.max_translation_norm
        icp = B.ICP(q)
        icp.init_reference(ref4, refn)
        g = icp.compute(read4, raise_on_error=False)
        icp.close()
        return g

    p = O.default_params(use_bound=1, max_rotation_norm=1.0, max_translation_norm=1.0)
    o, g = O.icp(p, ref4, refn, read4), bounded(p)
    assert o["status"] == 0 == g["status"] and g["iterations"] == o["iterations"] == base["iterations"]
    assert np.abs(g["T"] - o["T"]).max() < 1e-5
    moved = float(np.linalg.norm(base["T_iter"][:3, 3]))
    for kw in (dict(max_translation_norm=0.3 * moved), dict(max_rotation_norm=1e-4), dict(max_translation_norm=0.0)):
        p = O.default_params(use_bound=1, **kw)
        o, g = O.icp(p, ref4, refn, read4), bounded(p)
        assert o["status"] == O.ORC_E_BOUND and g["status"] == B.B200_E_BOUND
        assert g["iterations"] == o["iterations"] and not g["max_iterations_reached"]
    q = to_b(O.default_params())
    q.use_bound, q.max_rotation_norm = 1, -1.0
    with pytest.raises(B.B200Error):
        B.ICP(q)
    with pytest.raises(B.ConvergenceError):
        q.max_rotation_norm, q.max_translation_norm = 1e-4, 1.0
        icp = B.ICP(q)
        icp.init_reference(ref4, refn)
        icp.compute(read4)


@pytest.mark.parametrize("kw", [dict(use_trimmed=0, use_differential=0, max_iterations=40),
                                dict(use_trimmed=0, use_maxdist_filter=1, maxdist_filter=0.4, use_differential=0,
                                     max_iterations=25),
                                dict(trimmed_ratio=0.9, maxDist=0.5, use_differential=0, max_iterations=40)])
def test_icp_single_barrier_shape_is_deterministic(gold, kw, monkeypatch):
    """Consecutive single-barrier iterations (every iteration when TrimmedDist is absent) share no buffer across
    iterations any more (double-buffered by parity): many repetitions with B200_ICP_FAST=1 give bit-identical logs,
    and they equal the general shape's (B200_ICP_FAST=0) bit for bit — both sum the same values in the same order."""
    ref4, refn = O.feat4(gold["c400"]["xyz"]), gold["c400"]["normals"]
    read4 = O.feat4(gold["c401"]["xyz"])
    p = to_b(O.default_params(**kw))

    def run(fast, reps):
        monkeypatch.setenv("B200_ICP_FAST", fast)
        icp = B.ICP(p)
        icp.init_reference(ref4, refn)
        icp.upload_reading(read4)
        out = []
        for _ in range(reps):
            g = icp.run_resident()
            lg = icp.log()
            out.append((g["iterations"], g["T"].copy(), lg["H"].copy(), lg["b"].copy(), icp.debug_stats()[:, 8].copy()))
        icp.close()
        return out

    fastr, slow = run("1", 12), run("0", 1)[0]
    if not kw.get("use_trimmed", 1):                              # without TrimmedDist every iteration is single-barrier
        assert (fastr[0][4] == 1).all()
    assert (slow[4] == 0).all()
    for r in fastr:
        assert r[0] == fastr[0][0] and (bits(r[1]) == bits(fastr[0][1])).all()
        assert (bits(r[2]) == bits(fastr[0][2])).all() and (bits(r[3]) == bits(fastr[0][3])).all()
    assert fastr[0][0] == slow[0]
    assert np.abs(fastr[0][1] - slow[1]).max() < 1e-6
    assert np.abs(fastr[0][2] - slow[2]).max() <= 1e-6 * np.abs(slow[2]).max()


@pytest.mark.parametrize("knn,md", [(5, np.inf), (10, 0.5), (20, np.inf)])
def test_surface_normals_smooth(gold, knn, md):
    """smoothNormals (SurfaceNormal.cpp:265-288): in place, in index order — point i sees the already-smoothed normals
    of its lower-index neighbours. The device reproduces the chain exactly (ordered warps + done flags): where the raw
    normals agree in sign with the oracle's the smoothed ones are equal to float rounding."""
    f = O.feat4(gold["cloud0"])
    st, o = O.surface_normals(f, knn, md, smooth=1)
    st, raw = O.surface_normals(f, knn, md, smooth=0)
    g = B.surface_normals(f, knn, md, smooth=1, keep=("normals", "matchedIds"))
    graw = B.surface_normals(f, knn, md, smooth=0, keep=("normals",))
    assert (g["matchedIds"] == o["matchedIds"]).all()
    # the sign of an eigenvector is an artefact of the eigen solver and the smoothing is not sign-invariant: compare on
    # the points whose whole dependency cone (neighbours of neighbours ...) has the oracle's raw signs
    same = (np.abs(graw["normals"] - raw["normals"]).max(1) < 1e-5)
    if same.all():
        assert np.abs(g["normals"] - o["normals"]).max() < 1e-5
    else:   # replay the reference's recurrence on the DEVICE's raw normals: must match the device's smoothing
        n = graw["normals"].copy()
        ids, dd = o["ids"], o["dists"]
        for i in range(f.shape[0]):
            cur = n[i].copy()
            mean = np.zeros(3, np.float32)
            cnt = 0
            for j in range(knn):
                if np.isfinite(dd[i, j]):
                    nb = n[ids[i, j]]
                    dot = np.float32(np.float32(cur[0] * nb[0]) + np.float32(cur[1] * nb[1])) + np.float32(cur[2] * nb[2])
                    mean = (mean + nb) if dot > 0 else (mean - nb)
                    cnt += 1
            n[i] = mean / np.float32(cnt)
        assert np.abs(g["normals"] - n).max() < 1e-5
    with pytest.raises(B.B200Error):
        B.surface_normals(f, 40, np.inf, smooth=1)


@pytest.mark.parametrize("tile", ["0", "2"])
def test_surface_normals_uniform_cloud_both_kernels(tile):
    """400 000 uniformly sampled points, k = 20, through the warp-cooperative kernel (default) and through the cell-block
    tile kernel (B200_SN_TILE=2, with its exact fallback): neighbour ids bit-exact against the oracle's kd-tree either
    way. Run in a subprocess (the switch is read once per process)."""
    import subprocess, sys
    code = ("import sys; sys.path[:0] = ['.', 'tests']; import test_gpu_parity as t; t._uniform_cloud_normals(%r)" % tile)
    env = dict(os.environ, B200_SN_TILE=tile)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], env=env, cwd=root, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr


def _uniform_cloud_normals(tile):
    rng = np.random.default_rng(5)
    xy = rng.uniform(-20, 20, size=(400000, 2)).astype(np.float32)
    pts = np.concatenate([xy, (0.05 * np.sin(xy[:, :1]) + rng.normal(0, 0.005, (xy.shape[0], 1))).astype(np.float32)], 1)
    f = O.feat4(pts)
    idx = B.Index(f)
    out = idx.self_normals(knn=20, keep=("normals", "matchedIds", "densities"))
    fb = idx.last_normals_fallback()
    idx.close()
    assert fb < 0.05 * f.shape[0], fb   # the tiles settle nearly every point themselves (0 with the tiles off)
    assert (fb > 0) == (tile == "2") or fb == 0
    pick = np.sort(rng.choice(f.shape[0], 5000, replace=False))
    oi, od = O.KdTree(f).knn(np.ascontiguousarray(f[pick]), 20, np.inf, threads=4)
    assert (out["matchedIds"][pick].astype(np.int64) == oi).all()
    assert (np.abs(out["normals"][pick, 2]) > 0.95).mean() > 0.99


def test_icp_open3d_double_clouds(gold):
    """b200_icp_init_reference_o3d / b200_icp_compute_o3d: Open3D's double clouds in, open3dToPointmatcher's
    double -> float cast (open3d_conversions.cpp:57-118) on the device == the cast on the host, bit for bit"""
    rng = np.random.default_rng(0)
    ref_d = gold["c400"]["xyz"].astype(np.float64) + rng.uniform(-1e-9, 1e-9, gold["c400"]["xyz"].shape)   # true doubles
    refn_d = gold["c400"]["normals"].astype(np.float64)
    read_d = gold["c401"]["xyz"].astype(np.float64) + rng.uniform(-1e-9, 1e-9, gold["c401"]["xyz"].shape)
    p = B.default_icp_params(maxDist=0.5, trimmed_ratio=0.9, max_iterations=30, min_diff_trans=0.008)
    a = B.ICP(p)
    a.init_reference_o3d(ref_d, refn_d)
    ra = a.compute_o3d(read_d)
    ia, da, wa = a.matches()
    b = B.ICP(p)
    b.init_reference(O.feat4(ref_d.astype(np.float32)), refn_d.astype(np.float32))
    rb = b.compute(O.feat4(read_d.astype(np.float32)))
    ib, db, wb = b.matches()
    assert ra["iterations"] == rb["iterations"] and np.array_equal(ra["T"], rb["T"])
    assert (ia == ib).all() and (bits(da) == bits(db)).all() and (wa == wb).all()
    with pytest.raises(B.B200Error):
        B.ICP(p).compute_o3d(read_d)          # no reference yet
    a.close(); b.close()
