"""GPU parity ON THE BENCHMARKED CONFIGS THEMSELVES (BASELINE configs C2-C5 at their full sizes), through the C ABI,
against the CPU oracle — the workloads are the very generators bench.py times (open3d_slam_private_b200/synth.py).

C2  131 072-point scan vs 5 M-point map, param/icp.yaml chain: status, iteration count, pose < 1e-5 rad / 1e-5 m
    vs the oracle, one-step ("teacher-forced") ids / dists / weights BIT-EXACT and H, b within 2e-6, with both
    iteration shapes (B200_ICP_FAST=0/1), with and without X-ICP, asserting from the debug statistics that the
    single-barrier shape really ran; raw k-NN at k = 1 and k = 20 against the oracle's kd-tree over the 5 M map.
C3  GICP with covariances from k = 20 normals on the same clouds vs oracle/o3d.cpp (kind 2).
C4  one 200 k vs 200 k submap pair, target normals estimated on the device.
C5  a 10 M-point two-tile street on one device: un-partitioned vs 2 and 3 slabs exchanging through the mailboxes,
    and vs the oracle.
"""
import os

import numpy as np
import pytest

from oracle import oracle_py as O
from open3d_slam_private_b200 import capi as B
from open3d_slam_private_b200 import partition as P
from open3d_slam_private_b200 import synth

from test_gpu_parity import assert_icp_parity, bits, rot_angle, to_b

pytestmark = pytest.mark.gpu

N_POSES = 4
CPU = max(1, os.cpu_count() or 1)


def c2_params(**kw):
    """bench.py icp_kwargs(): param/icp.yaml modules with epsilon 0"""
    return O.default_params(knn=1, maxDist=0.5, use_trimmed=1, trimmed_ratio=0.9, max_iterations=30, use_differential=1,
                            min_diff_rot=1e-3, min_diff_trans=8e-3, smooth_length=3, minimizer=0, knn_threads=CPU, **kw)


@pytest.fixture(scope="module")
def c2():
    d = synth.make_c2(M=5_000_000, N=131072, n_poses=N_POSES)
    ref4 = synth.feat4(d["map_xyz"])
    refn = d["map_normals"]
    ses = O.IcpSession(ref4, refn)
    scans = [(synth.feat4(s["xyz"]), s["T_init"], s["T_true"]) for s in d["scans"]]
    yield dict(ref4=ref4, refn=refn, ses=ses, scans=scans)
    ses.close()


def _xicp_b(q, p):
    q.degeneracy = p.degeneracy
    q.xicp_enough_information = p.xicp_enough
    q.xicp_insufficient_information = p.xicp_insufficient
    q.xicp_min_alignment_angle_deg = 80.0
    q.xicp_strong_alignment_angle_deg = 45.0
    return q


@pytest.mark.parametrize("fast", ["1", "0"])
@pytest.mark.parametrize("degeneracy", [0, 1])
def test_c2_registration_parity(c2, fast, degeneracy, monkeypatch):
    """the loop of ICP.cpp:1027-1311 on the configuration the headline number is quoted on"""
    monkeypatch.setenv("B200_ICP_FAST", fast)
    p = c2_params(degeneracy=degeneracy, xicp_enough=250.0, xicp_insufficient=180.0)
    icp = B.ICP(_xicp_b(to_b(p), p))
    icp.init_reference(c2["ref4"], c2["refn"])
    shapes = []
    for r4, Ti, Tt in c2["scans"]:
        o = c2["ses"].icp(p, r4, None, Ti)
        g = icp.compute(r4, None, Ti, raise_on_error=False)
        icp.inputs = (p, c2["ref4"], c2["refn"], r4, None, Ti)
        icp.oracle_forced = lambda Ts, r4=r4, Ti=Ti: c2["ses"].icp(p, r4, None, Ti, forced_T_iters=Ts)
        assert g["status"] == 0 and g["grid_ctas"] == 256
        assert_icp_parity(o, g, icp)
        assert np.abs(g["T"][:3, 3] - Tt[:3, 3]).max() < 5e-3        # and it is the right answer
        if degeneracy:
            assert g["loc_xyz"].tolist() == o["loc_xyz"].tolist() and g["loc_rpy"].tolist() == o["loc_rpy"].tolist()
            assert not g["returned_prior"]
        st = icp.debug_stats()
        shapes += [int(v) for v in st[:g["iterations"], 8]]
    if fast == "1":
        assert 1 in shapes, "the single-barrier iteration shape never ran on C2: %r" % shapes
    else:
        assert set(shapes) == {0}, shapes
    icp.close()


def test_c2_knn_bit_exact_5M(c2):
    """findClosests over the 5 M-point index (10^8-cell table, uint32 offsets): ids / dists bit-exact, k = 1 for all
    131 072 queries at the initial guess (the expensive first iteration) and k = 20 for a subset"""
    r4, Ti, _ = c2["scans"][0]
    q4 = np.ascontiguousarray((r4.astype(np.float32) @ Ti.T.astype(np.float32)).astype(np.float32))
    q4[:, 3] = 1.0
    idx = B.Index(c2["ref4"])
    kd = O.KdTree(c2["ref4"])
    ids, d = idx.find_closests(q4, knn=1, max_dist=0.5)
    oi, od = kd.knn(q4, 1, 0.5, threads=CPU)
    assert (ids == oi).all() and (bits(d) == bits(od)).all()
    assert (ids >= 0).mean() > 0.9
    sub = np.ascontiguousarray(q4[::16])
    ids, d = idx.find_closests(sub, knn=20, max_dist=1.0)
    oi, od = kd.knn(sub, 20, 1.0, threads=CPU)
    assert (ids == oi).all() and (bits(d) == bits(od)).all()
    ids, d = idx.find_closests(np.ascontiguousarray(q4[::64]), knn=1, max_dist=np.inf)
    oi, od = kd.knn(np.ascontiguousarray(q4[::64]), 1, np.inf, threads=CPU)
    assert (ids == oi).all() and (bits(d) == bits(od)).all()
    idx.close()


def _normals_from_ids(pts4, ids):
    """normal = eigenvector of the smallest eigenvalue of the neighbourhood covariance, numpy double (independent of
    both the device and the oracle's own eigen solver); returns normals and the eigenvalue gap ratio"""
    nb = pts4[ids][:, :, :3].astype(np.float64)
    c = nb - nb.mean(1, keepdims=True)
    C = np.einsum("nki,nkj->nij", c, c)
    w, v = np.linalg.eigh(C)
    return v[:, :, 0], (w[:, 1] - w[:, 0]) / np.maximum(w[:, 2], 1e-300)


@pytest.fixture(scope="module")
def c3_normals(c2):
    """per-point normals (the source of GICP's covariances) from k = 20 neighbours, estimated on the device for the
    5 M-point map"""
    idx = B.Index(c2["ref4"])
    out = idx.self_normals(knn=20, keep=("normals", "matchedIds"))
    idx.close()
    return out


def test_c3_map_normals_k20_5M(c2, c3_normals):
    """SurfaceNormalDataPointsFilter at C3's size: self-kNN ids bit-exact against the oracle's kd-tree on a sample of
    the 5 M points, normals equal (modulo sign) to a numpy eigen-decomposition of the same neighbourhoods"""
    n, mid = c3_normals["normals"], c3_normals["matchedIds"]
    rng = np.random.default_rng(3)
    pick = np.sort(rng.choice(c2["ref4"].shape[0], 20000, replace=False))
    oi, od = O.KdTree(c2["ref4"]).knn(np.ascontiguousarray(c2["ref4"][pick]), 20, np.inf, threads=CPU)
    assert (mid[pick].astype(np.int64) == oi).all()
    want, gap = _normals_from_ids(c2["ref4"], oi)
    dots = np.abs((n[pick].astype(np.float64) * want).sum(1))
    well = gap > 1e-3                                   # a clear smallest eigenvalue
    assert well.mean() > 0.95
    assert (dots[well] > 1 - 1e-4).mean() > 0.999
    assert np.abs(np.linalg.norm(n[pick], axis=1) - 1).max() < 1e-3


def _oriented(n, pts4):
    n = np.ascontiguousarray(n, np.float32).copy()
    flip = (n * (-pts4[:, :3])).sum(1) < 0
    n[flip] *= -1
    return n


@pytest.mark.parametrize("max_iteration", [0, 1, 30])
def test_c3_gicp_parity(c2, c3_normals, max_iteration):
    """RegistrationGeneralizedICP (CloudRegistration.cpp:17-23) on 131 072 vs 5 M points: the same normals on both
    sides; correspondences bit-exact for a given pose, one update within 1e-9, free-running result within 1e-6"""
    tgtn = _oriented(c3_normals["normals"], c2["ref4"])
    h = B.O3dICP(kind=B.O3D_GENERALIZED, max_correspondence_distance=0.5, max_iteration=max_iteration)
    h.set_target(c2["ref4"], tgtn)
    sidx = None
    for r4, Ti, Tt in c2["scans"][:2]:
        if sidx is None:
            sidx = B.Index(r4)
        else:
            sidx.update(r4)
        srcn = _oriented(sidx.self_normals(knn=20)["normals"], r4)
        init = Ti.astype(np.float64)
        g = h.register(r4, srcn, init=init)
        ids, d = h.correspondences()
        o = O.o3d_icp(2, r4, c2["ref4"], tgtn, srcn, max_dist=0.5, init=init, max_iteration=max_iteration, threads=CPU)
        assert g["iterations"] == o["iterations"]
        if max_iteration == 0:
            assert (ids == o["ids"]).all() and g["fitness"] == o["fitness"]
            assert abs(g["inlier_rmse"] - o["inlier_rmse"]) <= 1e-12 * max(1.0, o["inlier_rmse"])
        elif max_iteration == 1:
            # 1e-9 RELATIVE to the pose's magnitude (translations are ~50 m here; test_gpu_o3d's 1e-9 is on ~1 m)
            assert np.abs(g["T"] - o["T"]).max() < 1e-9 * max(1.0, np.abs(o["T"]).max()) and (ids == o["ids"]).all()
        else:
            assert rot_angle(g["T"][:3, :3], o["T"][:3, :3]) < 1e-6
            assert np.abs(g["T"][:3, 3] - o["T"][:3, 3]).max() < 1e-6
            assert (ids != o["ids"]).mean() < 1e-3
            assert np.abs(g["T"][:3, 3] - Tt[:3, 3]).max() < 2e-2
    sidx.close()
    h.close()


def test_c4_submap_pair_parity():
    """one loop-closure candidate of C4 (200 000 vs 200 000 points, max correspondence distance 0.5, target normals
    from k = 10 neighbours within 2 m): device vs oracle with the oracle's own normals"""
    pair = synth.make_c4(n_distinct=1, N=200_000)[0]
    tgt4, src4 = synth.feat4(pair["tgt_xyz"]), synth.feat4(pair["src_xyz"])
    init = pair["T_init"].astype(np.float64)
    h = B.O3dICP(kind=B.O3D_POINT_TO_PLANE, max_correspondence_distance=0.5, max_iteration=30)
    h.set_target_estimate_normals(tgt4, knn=10, max_dist=2.0)
    g = h.register(src4, init=init)
    ids, d = h.correspondences()
    h.close()
    n, _ = O.o3d_estimate_normals(tgt4, 10, 2.0, threads=CPU)    # Open3D EstimateNormals restated (oracle/o3d.cpp)
    o = O.o3d_icp(0, src4, tgt4, n.astype(np.float32), None, max_dist=0.5, init=init, max_iteration=30, threads=CPU)
    assert g["iterations"] == o["iterations"]
    assert rot_angle(g["T"][:3, :3], o["T"][:3, :3]) < 1e-5
    assert np.abs(g["T"][:3, 3] - o["T"][:3, 3]).max() < 1e-5
    assert (ids != o["ids"]).mean() < 1e-3
    assert np.abs(g["T"][:3, 3] - pair["T_true"][:3, 3]).max() < 5e-2


@pytest.fixture(scope="module")
def c5():
    street = synth.TiledStreet(2)
    parts = [street.tile_points(g, 5_000_000) for g in range(2)]
    xyz = np.ascontiguousarray(np.concatenate([p[0] for p in parts]))
    nrm = np.ascontiguousarray(np.concatenate([p[1] for p in parts]))
    scans = synth.make_c5_scans(street, 2, 131072)
    return dict(tgt4=synth.feat4(xyz), tgtn=nrm, scans=[(synth.feat4(s["xyz"]), s["T_init"].astype(np.float64)) for s in scans])


@pytest.mark.parametrize("n_slabs", [2, 3])
def test_c5_slabs_equal_unpartitioned_10M(c5, n_slabs):
    """a 10 M-point map cut into slabs on one device (one handle and stream per slab, sums exchanged through the
    peer mailboxes) against the un-partitioned registration and against the oracle"""
    tgt4, tgtn = c5["tgt4"], c5["tgtn"]
    kw = dict(kind=0, max_correspondence_distance=0.5, max_iteration=30)
    whole = B.O3dICP(**kw)
    whole.set_target(tgt4, tgtn)
    axis, edges = P.plan_slabs(tgt4, n_slabs)
    hs, ptrs = [], []
    for g in range(n_slabs):
        gid = P.slab_members(tgt4, axis, edges, g, 0.5)
        h = B.O3dICP(**kw)
        h.set_target(np.ascontiguousarray(tgt4[gid]), np.ascontiguousarray(tgtn[gid]))
        h.set_partition(axis, float(edges[g]), float(edges[g + 1]), gid)
        ptrs.append(h.peer_init(g, n_slabs)[0])
        hs.append(h)
    for h in hs:
        h.peer_connect(pointers=ptrs, devices=[0] * n_slabs)
    for k, (s4, init) in enumerate(c5["scans"]):
        ref = whole.register(s4, init=init)
        ref_ids, _ = whole.correspondences()
        for h in hs:
            h.enqueue_peers(s4, init=init)
        res = [h.finish() for h in hs]
        own = [h.correspondences()[0] for h in hs]
        ids = P.merge_correspondences(own)
        if n_slabs == 2:   # the scans are taken astride the border of the two tiles; with 3 slabs the middle one owns them
            assert sum(int((o_ != -2).any()) for o_ in own) == 2
        for r in res:
            assert r["iterations"] == ref["iterations"] and r["n_correspondences"] == ref["n_correspondences"]
            assert np.abs(r["T"] - ref["T"]).max() < 1e-10
            assert np.array_equal(r["T"], res[0]["T"])
        assert (ids == ref_ids).all()
        if k == 0 and n_slabs == 2:
            o = O.o3d_icp(0, s4, tgt4, tgtn, None, max_dist=0.5, init=init, max_iteration=30, threads=CPU)
            assert ref["iterations"] == o["iterations"]
            assert rot_angle(ref["T"][:3, :3], o["T"][:3, :3]) < 1e-6
            assert np.abs(ref["T"][:3, 3] - o["T"][:3, 3]).max() < 1e-6
            assert (ref_ids != o["ids"]).mean() < 1e-3
    for h in hs + [whole]:
        h.close()


def test_c2_far_pose_takes_the_svd_fallback_branch(c2):
    """A scan 55 m from the map's centroid: the rotation block of H dwarfs the translation block, the reference's float
    full-pivot QR calls H rank deficient in EVERY iteration, the minimal-norm solution fails isApprox and the double
    SVD fallback solves (PointToPlane.cpp:199-246, info = 2). The device answers that branch from the double LDL^T it
    already holds (solve.cuh): same iteration count, pose and one-step parity as every other pose."""
    d = synth.make_c2(M=1000, N=131072, n_poses=17, only={16})
    s = d["scans"][16]
    r4, Ti = synth.feat4(s["xyz"]), s["T_init"]
    p = c2_params()
    o = c2["ses"].icp(p, r4, None, Ti)
    infos = [O.solve6(o["H"][k], o["b"][k])[1] for k in range(o["iterations"])]
    assert set(infos) == {2}, infos
    icp = B.ICP(to_b(p))
    icp.init_reference(c2["ref4"], c2["refn"])
    g = icp.compute(r4, None, Ti, raise_on_error=False)
    icp.inputs = (p, c2["ref4"], c2["refn"], r4, None, Ti)
    icp.oracle_forced = lambda Ts: c2["ses"].icp(p, r4, None, Ti, forced_T_iters=Ts)
    assert_icp_parity(o, g, icp)
    st = icp.debug_stats()
    assert st[1:g["iterations"], 5].max() < 80.0, st[:, 5]     # the solve no longer costs 110 us (serial Jacobi SVD)
    icp.close()
