"""Map maintenance on the device (SURVEY 8(f) rank 3, b200_map_*): Submap::insertScan (Submap.cpp:39-96 ->
voxelizeWithinCroppingVolume, helpers.cpp:117-200), ScanToMapIcp::cropSubmap (ScanToMapRegistration.cpp:90-96) and the
hand-over of the patch to ICP::initReference without leaving the device.

CPU: the oracle's restatement against a plain numpy grouping. GPU: the device map against the oracle after every
insertion — sizes, the pass-through set and the output order identical; pass-through points bit-identical; voxel means
within 1e-12 (double atomics vs sequential double sums); crop indices bit-exact; a registration against the patch
handed over on the device equals, bit for bit, the one against the same patch uploaded from the host.
"""
import numpy as np
import pytest

from oracle import oracle_py as O


def _scene(seed, n, centre):
    """a wavy ground patch with an upright wall, seen from `centre` (sensor frame)"""
    rng = np.random.default_rng(seed)
    xy = rng.uniform(-12, 12, size=(n, 2))
    z = 0.05 * np.sin(xy[:, 0]) + 0.02 * rng.normal(size=n)
    pts = np.concatenate([xy, z[:, None]], 1)
    wall = np.stack([rng.uniform(-12, 12, n // 4), np.full(n // 4, 6.0) + 0.01 * rng.normal(size=n // 4),
                     rng.uniform(0, 3, n // 4)], 1)
    pts = np.concatenate([pts, wall]) - np.asarray(centre)[None, :]
    nrm = np.concatenate([np.tile([0.0, 0.0, 1.0], (n, 1)), np.tile([0.0, -1.0, 0.0], (n // 4, 1))])
    nrm = nrm + 0.05 * rng.normal(size=nrm.shape)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    return O.feat4(pts.astype(np.float32)), nrm.astype(np.float32)


def _pose(x, y, yaw):
    c, s = np.cos(yaw), np.sin(yaw)
    T = np.eye(4)
    T[:3, :3] = [[c, -s, 0], [s, c, 0], [0, 0, 1]]
    T[:3, 3] = [x, y, 0.3]
    return T


POSES = [(0.0, 0.0, 0.0), (0.8, -0.3, 0.05), (1.7, -0.5, 0.11), (2.4, -0.2, 0.2)]
CROP = dict(kind=3, radius_max=9.0)     # MaxRadiusCroppingVolume around the sensor (the map builder's cropper)
VOXEL = 0.2


def test_oracle_insert_scan_against_numpy():
    scan4, scann = _scene(1, 4000, (0, 0, 0))
    T = _pose(0.5, 0.2, 0.1)
    p, n, n_out = O.map_insert_scan(np.zeros((0, 3)), None, scan4, scann, T, VOXEL, center=T[:3, 3], **CROP)
    w = scan4[:, :3].astype(np.float64) @ T[:3, :3].T + T[:3, 3]
    wn = scann.astype(np.float64) @ T[:3, :3].T
    inside = np.linalg.norm(w - T[:3, 3], axis=1) <= CROP["radius_max"]
    assert n_out == int((~inside).sum()) and n_out > 0
    assert np.allclose(p[:n_out], w[~inside], rtol=0, atol=1e-12) and np.allclose(n[:n_out], wn[~inside], atol=1e-12)
    key = np.floor(w[inside] * (1.0 / VOXEL)).astype(np.int64)
    uniq, first, inv = np.unique(key, axis=0, return_index=True, return_inverse=True)
    order = np.argsort(first)                      # voxels by the index of their first member
    assert p.shape[0] == n_out + uniq.shape[0]
    for rank, u in enumerate(order[:200]):
        m = inv.ravel() == u
        assert np.allclose(p[n_out + rank], w[inside][m].mean(0), atol=1e-12)
        mn = wn[inside][m].mean(0)
        assert np.allclose(n[n_out + rank], mn / np.linalg.norm(mn), atol=1e-12)
    # voxel_size <= 0: plain concatenation; an empty scan changes nothing
    q, _, k = O.map_insert_scan(p, n, scan4[:10], scann[:10], T, 0.0)
    assert q.shape[0] == p.shape[0] + 10 and k == q.shape[0] and np.array_equal(q[:p.shape[0]], p)


@pytest.mark.gpu
def test_device_map_follows_the_oracle_over_several_insertions():
    from open3d_slam_private_b200 import capi as B
    m = B.Map()
    op, on = np.zeros((0, 3)), None
    for k, (x, y, yaw) in enumerate(POSES):
        T = _pose(x, y, yaw)
        scan4, scann = _scene(10 + k, 30000, (x, y, 0.3))
        crop = B.crop_params(center=T[:3, 3], **CROP)
        n_out, size = m.insert_scan(scan4, scann, T, VOXEL, crop)
        op, on, o_out = O.map_insert_scan(op, on, scan4, scann, T, VOXEL, center=T[:3, 3], **CROP)
        assert (n_out, size) == (o_out, op.shape[0]) and m.size() == (size, True)
        gp, gn = m.download()
        assert np.array_equal(gp[:n_out], op[:n_out]) and np.array_equal(gn[:n_out], on[:n_out])   # pass-through: copies
        assert np.abs(gp - op).max() < 1e-12 and np.abs(gn - on).max() < 1e-12                    # voxel means, same order
        if k > 0:
            assert n_out > 0           # older parts of the map lie outside the cropper and pass through untouched
        assert m.last_device_ms() > 0
        # keep the oracle on the device's map so that ulp-level differences cannot move a mean across a voxel face
        op, on = gp, gn
    # ---- cropSubmap: indices bit-exact, floats are the rounded doubles
    c2 = dict(kind=1, radius_max=6.0, min_z=-1.0, max_z=2.0)
    centre = _pose(*POSES[-1])[:3, 3]
    p4, n3, idx = m.crop(B.crop_params(center=centre, **c2))
    want = O.crop_d(1, op, radius_max=6.0, min_z=-1.0, max_z=2.0, center=centre)
    assert np.array_equal(idx, want) and idx.size > 1000
    assert np.array_equal(p4[:, :3], op[want].astype(np.float32)) and (p4[:, 3] == 1).all()
    assert np.array_equal(n3, on[want].astype(np.float32))
    # ---- the patch handed to the registration on the device == the same patch uploaded from the host
    T = _pose(*POSES[-1])
    scan4, _ = _scene(99, 20000, tuple(T[:3, 3]))
    Ti = (T @ _pose(0.05, -0.04, 0.01)).astype(np.float32)
    Ti[2, 3] = T[2, 3]
    prm = B.default_icp_params(maxDist=0.5, trimmed_ratio=0.9, max_iterations=30, min_diff_trans=0.008)
    a = B.ICP(prm)
    assert m.crop_to_icp(B.crop_params(center=centre, **c2), a) == idx.size
    ra = a.compute(scan4, None, Ti)
    b = B.ICP(prm)
    b.init_reference(p4, n3)
    rb = b.compute(scan4, None, Ti)
    assert ra["iterations"] == rb["iterations"] and np.array_equal(ra["T"], rb["T"])
    assert abs(ra["T"][2, 3] - T[2, 3]) < 0.05      # (the wavy ground pins z; x / y slide along it)
    a.close(); b.close(); m.close()


@pytest.mark.gpu
def test_device_map_edge_cases():
    from open3d_slam_private_b200 import capi as B
    m = B.Map()
    scan4, scann = _scene(3, 2000, (0, 0, 0))
    T = np.eye(4)
    crop = B.crop_params(**CROP)
    assert m.insert_scan(scan4[:0], None, T, VOXEL, crop) == (0, 0)            # empty scan: nothing happens
    with pytest.raises(B.B200Error):
        m.crop_to_icp(crop, B.ICP())                                          # empty map: no reference
    assert m.insert_scan(scan4, scann, T, 0.0, crop) == (2500, 2500)           # voxel size 0: plain concatenation
    p, n = m.download()
    assert np.array_equal(p, scan4[:, :3].astype(np.float64)) and np.array_equal(n, scann.astype(np.float64))
    with pytest.raises(B.B200Error):
        m.insert_scan(scan4, None, T, VOXEL, crop)                             # map with normals, scan without
    with pytest.raises(B.B200Error):
        m.insert_scan(scan4, scann, T, 1e-7, crop)                             # voxel index beyond 2^20
    assert m.size() == (2500, True)                                            # the failed calls left the map alone
    m.set(p[:100], None)
    assert m.size() == (100, False)
    with pytest.raises(B.B200Error):
        m.crop(B.crop_params(kind=9))
    far = B.crop_params(kind=3, radius_max=1.0, center=(500.0, 0.0, 0.0))
    assert m.crop(far)[2].size == 0
    with pytest.raises(B.B200Error):
        m.crop_to_icp(far, B.ICP(B.default_icp_params(minimizer=1)))           # nothing inside the volume
    m.close()
