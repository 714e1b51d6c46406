"""X-ICP localizability (degeneracyAwareness: OptimizedEqualityConstraints, the setting of
open3d_slam_ros/param/icp.yaml): oracle known answers on analytic scenes (CPU) and device parity (GPU).

Reference: ICP.cpp:1112-1141 (call site), :1580-1591 (eigenAnalysis), :2128-2155 (detectLocalizability), :2187-2444
(detectLocalizabilityWithOptimizedMethod); PointToPlane.cpp:461-505, 570-626, 757-778 (constrained solve). The
reference holds no test for this path (parity "pinned to the oracle"); the known answers below are geometric: a corridor
cannot observe translation along its axis, a single plane observes only its normal translation and the two tilts.
"""
import numpy as np
import pytest

from oracle import oracle_py as O


def _plane(rng, n, origin, u, v, normal):
    a, b = rng.uniform(0, 1, n), rng.uniform(0, 1, n)
    pts = np.asarray(origin)[None, :] + a[:, None] * np.asarray(u)[None, :] + b[:, None] * np.asarray(v)[None, :]
    return pts, np.tile(np.asarray(normal, np.float64), (n, 1))


def corridor(rng, n, L=40.0):
    """floor, ceiling and two walls of a corridor along x (4 m wide, 3 m high)"""
    per = n // 4
    parts = [_plane(rng, per, [-L / 2, -2, 0], [L, 0, 0], [0, 4, 0], [0, 0, 1]),
             _plane(rng, per, [-L / 2, -2, 3], [L, 0, 0], [0, 4, 0], [0, 0, -1]),
             _plane(rng, per, [-L / 2, -2, 0], [L, 0, 0], [0, 0, 3], [0, 1, 0]),
             _plane(rng, per, [-L / 2, 2, 0], [L, 0, 0], [0, 0, 3], [0, -1, 0])]
    return np.vstack([p for p, _ in parts]), np.vstack([m for _, m in parts])


def room(rng, n, L=12.0):
    """closed box: every direction is observable"""
    per = n // 6
    parts = [_plane(rng, per, [-L / 2, -4, 0], [L, 0, 0], [0, 8, 0], [0, 0, 1]),
             _plane(rng, per, [-L / 2, -4, 3], [L, 0, 0], [0, 8, 0], [0, 0, -1]),
             _plane(rng, per, [-L / 2, -4, 0], [L, 0, 0], [0, 0, 3], [0, 1, 0]),
             _plane(rng, per, [-L / 2, 4, 0], [L, 0, 0], [0, 0, 3], [0, -1, 0]),
             _plane(rng, per, [-L / 2, -4, 0], [0, 8, 0], [0, 0, 3], [1, 0, 0]),
             _plane(rng, per, [L / 2, -4, 0], [0, 8, 0], [0, 0, 3], [-1, 0, 0])]
    return np.vstack([p for p, _ in parts]), np.vstack([m for _, m in parts])


def ground(rng, n, L=30.0):
    return _plane(rng, n, [-L / 2, -L / 2, 0], [L, 0, 0], [0, L, 0], [0, 0, 1])


def make_pair(kind, seed=3, M=40000, N=8000, noise=0.01, nnoise=0.02):
    rng = np.random.default_rng(seed)
    gen = dict(corridor=corridor, room=room, ground=ground)[kind]
    ref, refn = gen(rng, M)
    ref = ref + rng.normal(0, noise, ref.shape)
    refn = refn + rng.normal(0, nnoise, refn.shape)
    refn /= np.linalg.norm(refn, axis=1, keepdims=True)
    read, _ = gen(rng, N, **({"L": 20.0} if kind != "room" else {}))
    a = 0.02
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    read = read @ R.T + np.array([0.30, 0.05, -0.04]) + rng.normal(0, noise, read.shape)
    return O.feat4(ref), refn.astype(np.float32), O.feat4(read)


def params(**kw):
    base = dict(maxDist=1.0, trimmed_ratio=0.9, max_iterations=30, min_diff_trans=0.008, degeneracy=1, knn_threads=4)
    base.update(kw)
    return O.default_params(**base)


# ------------------------------------------------------------------------------------------------ oracle (CPU)
def test_oracle_room_is_fully_localizable_and_equals_plain_icp():
    ref4, refn, read4 = make_pair("room")
    x = O.icp(params(), ref4, refn, read4)
    n = O.icp(params(degeneracy=0), ref4, refn, read4)
    assert x["status"] == 0 and not x["returned_prior"]
    assert (x["loc_xyz"] == 1).all() and (x["loc_rpy"] == 1).all()
    # no constraint => the very same solve (PointToPlane.cpp:473-477)
    assert x["iterations"] == n["iterations"]
    assert (x["T"] == n["T"]).all()
    # the sums that decided it are far above the thresholds (250 / 180)
    assert x["xicp_sums"][0::2].min() > 250


def test_oracle_corridor_constrains_translation_along_the_axis():
    ref4, refn, read4 = make_pair("corridor")
    x = O.icp(params(), ref4, refn, read4)
    n = O.icp(params(degeneracy=0), ref4, refn, read4)
    assert x["status"] == 0 and not x["returned_prior"]
    assert x["loc_xyz"].tolist() == [1, 1, 0] and x["loc_rpy"].tolist() == [1, 1, 1]
    v = x["eig_t"][:, 2]                       # the direction of the smallest singular value: the corridor axis
    assert abs(abs(v[0]) - 1) < 1e-3
    # every update obeys v_i . t = 0 (constraint value 0, ICP.cpp:2321-2324) with the eigenvector of its own
    # iteration (re-analysed while a direction is degenerate, :2205); v is the last one, hence the 1e-4
    assert np.abs(x["x"][:, 3:] @ v).max() < 1e-4
    assert abs(x["T"][0, 3]) < 1e-3
    # the unconstrained solve slides along the corridor on noise alone
    assert abs(n["T"][0, 3]) > 10 * abs(x["T"][0, 3])
    # the observable part is still registered: rotation 0.02 rad about z and the y / z offsets come back
    assert abs(np.arctan2(x["T"][0, 1], x["T"][0, 0]) - 0.02) < 2e-3
    assert abs(x["T"][1, 3] + 0.05) < 0.02 and abs(x["T"][2, 3] - 0.04) < 0.02


def test_oracle_single_plane_has_three_degenerate_directions():
    ref4, refn, read4 = make_pair("ground")
    x = O.icp(params(), ref4, refn, read4)
    assert x["status"] == 0 and not x["returned_prior"]
    assert x["loc_xyz"].tolist() == [1, 0, 0]     # only the translation along the normal is observable
    assert x["loc_rpy"].tolist() == [1, 1, 0]     # the two tilts are, the spin about the normal is not
    assert abs(abs(x["eig_t"][2, 0]) - 1) < 1e-2
    assert abs(abs(x["eig_r"][2, 2]) - 1) < 5e-2
    # the in-plane translation and the yaw stay at the initial guess
    assert np.abs(x["T"][:2, 3]).max() < 2e-3
    assert abs(np.arctan2(x["T"][0, 1], x["T"][0, 0])) < 2e-3
    assert abs(x["T"][2, 3] - 0.04) < 0.01


def test_oracle_constrained_solve_against_numpy():
    """solvePossiblyUnderdeterminedLinearSystemWithEqualityConstraints builds the KKT system [[A, C], [C^T, 0]]
    (PointToPlane.cpp:480-503); its solution minimises the quadratic subject to C^T x = d."""
    rng = np.random.default_rng(5)
    for c in (1, 2, 3, 6):
        J = rng.normal(size=(200, 6))
        A = (J.T @ J).astype(np.float32)
        b = rng.normal(size=6).astype(np.float32)
        Cm = np.linalg.qr(rng.normal(size=(6, 6)))[0][:, :c].astype(np.float32)
        n = 6 + c
        K = np.zeros((n, n), np.float32)
        K[:6, :6] = A
        K[:6, 6:] = Cm
        K[6:, :6] = Cm.T
        rhs = np.concatenate([b, np.zeros(c, np.float32)])
        x, info = O.solve_n(K, rhs)
        want = np.linalg.solve(K.astype(np.float64), rhs.astype(np.float64))
        assert info == 0
        assert np.abs(x - want).max() < 1e-5 * max(1.0, np.abs(want).max())
        assert np.abs(Cm.T @ x[:6]).max() < 1e-5
    x6, _ = O.solve_n(A, b)
    assert np.abs(x6 - O.solve6(A, b)[0]).max() == 0


# ------------------------------------------------------------------------------------------------ device (GPU)
def _to_b(B, p):
    from test_gpu_parity import to_b
    q = to_b(p)
    q.degeneracy = p.degeneracy
    q.xicp_enough_information = p.xicp_enough
    q.xicp_insufficient_information = p.xicp_insufficient
    q.xicp_min_alignment_angle_deg = 80.0
    q.xicp_strong_alignment_angle_deg = 45.0
    return q


def _same_up_to_sign(a, b, tol):
    for c in range(3):
        s = np.sign(np.dot(a[:, c], b[:, c])) or 1.0
        if np.abs(a[:, c] - s * b[:, c]).max() > tol:
            return False
    return True


@pytest.mark.gpu
@pytest.mark.parametrize("kind,flags", [("room", ([1, 1, 1], [1, 1, 1])), ("corridor", ([1, 1, 0], [1, 1, 1])),
                                        ("ground", ([1, 0, 0], [1, 1, 0]))])
@pytest.mark.parametrize("fast", ["1", "0"])
def test_device_xicp_parity(kind, flags, fast, monkeypatch):
    from open3d_slam_private_b200 import capi as B
    from test_gpu_parity import assert_icp_parity
    monkeypatch.setenv("B200_ICP_FAST", fast)
    ref4, refn, read4 = make_pair(kind)
    p = params()
    o = O.icp(p, ref4, refn, read4)
    icp = B.ICP(_to_b(B, p))
    icp.init_reference(ref4, refn)
    g = icp.compute(read4, None, None, raise_on_error=False)
    icp.inputs = (p, ref4, refn, read4, None, None)
    assert g["loc_xyz"].tolist() == o["loc_xyz"].tolist() == flags[0]
    assert g["loc_rpy"].tolist() == o["loc_rpy"].tolist() == flags[1]
    assert not g["returned_prior"] and g["xicp_borderline"] == 0
    # the alignment sums over all pairs: same float terms, summed in double on both sides
    assert np.abs(g["xicp_sums"] - o["xicp_sums"]).max() <= 1e-4 * max(1.0, np.abs(o["xicp_sums"]).max())
    assert _same_up_to_sign(g["eig_t"], o["eig_t"], 1e-4) and _same_up_to_sign(g["eig_r"], o["eig_r"], 1e-4)
    # a constrained direction is only as well conditioned as the constraint: 10x the plain pose tolerance there
    assert_icp_parity(o, g, icp, pose_tol_scale=1.0 if kind == "room" else 10.0)


@pytest.mark.gpu
@pytest.mark.parametrize("which", [2, 4, 8, 10])        # combined sums of translation e = 1, 2 and rotation e = 1, 2
@pytest.mark.parametrize("rel", [-3e-4, -1e-6, 0.0, 1e-6, 3e-4])
def test_device_xicp_borderline_sums_are_replayed_exactly(which, rel):
    """A threshold placed within a few ulps / 0.03 % of an alignment sum: the parallel double sums cannot certify what
    the reference's sequential float recurrence (ICP.cpp:2128-2155) would decide, so the device replays it. The flags
    must be the oracle's (which runs that recurrence literally) on either side of the threshold."""
    from open3d_slam_private_b200 import capi as B
    ref4, refn, read4 = make_pair("room", seed=9, M=30000, N=6000)
    base = O.icp(params(), ref4, refn, read4)
    thr = float(np.float32(base["xicp_sums"][which] * (1.0 + rel)))
    p = params(xicp_enough=thr, xicp_insufficient=1e9)
    o = O.icp(p, ref4, refn, read4)
    icp = B.ICP(_to_b(B, p))
    icp.init_reference(ref4, refn)
    g = icp.compute(read4, None, None, raise_on_error=False)
    assert g["xicp_borderline"] >= 1
    assert g["loc_xyz"].tolist() == o["loc_xyz"].tolist() and g["loc_rpy"].tolist() == o["loc_rpy"].tolist()
    assert g["status"] == o["status"] and g["iterations"] == o["iterations"]
    assert np.abs(g["T"] - o["T"]).max() < 1e-4


@pytest.mark.gpu
def test_device_xicp_rejects_knn_above_one():
    from open3d_slam_private_b200 import capi as B
    q = B.default_icp_params(knn=3)
    q.degeneracy = 1
    with pytest.raises(Exception):
        B.ICP(q)
