"""libpointmatcher chain over a spatially partitioned map (b200_icp_slab_*): every slab matches the
queries it owns, the zero-filled exchange arrays are summed, and the persistent kernel runs the rest of the iteration
(TrimmedDist's quantile over ALL pairs, H/b, solve, checkers) on the complete set. The bar: the partitioned run IS the
un-partitioned one — pose, iteration count, matched ids (global), distances and outlier weights bit for bit.
All slabs share one device here (the exchange is a plain sum); the multi-rank driver is the same loop with an
all-reduce in its place (partition.register_partitioned_pm, run by bench.py's c5 mode on several GPUs)."""
import numpy as np
import pytest

from oracle import oracle_py as O

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _whole_then_slabs(map4, mapn, read4, n_slabs, T_init=None, plan=None, **kw):
    from open3d_slam_private_b200 import capi as B, partition as P
    prm = B.default_icp_params(**kw)
    whole = B.ICP(prm)
    whole.init_reference(map4, mapn)
    r0 = whole.compute(read4, None, T_init, raise_on_error=False)
    m0 = whole.matches()
    whole.close()
    mean = r0["T_refIn_refMean"][:3, 3].astype(np.float64)      # the float centroid the whole-map run subtracted
    r1 = P.register_slabs_pm_one_device(map4, mapn, read4, n_slabs, params=prm, T_init=T_init, mean=mean, plan=plan)
    return r0, m0, r1


@pytest.mark.parametrize("n_slabs", [1, 2, 3])
@pytest.mark.parametrize("kw", [dict(maxDist=0.5, trimmed_ratio=0.9, max_iterations=30, min_diff_trans=0.008),
                                dict(maxDist=1.0, trimmed_ratio=0.7, use_maxdist_filter=1, maxdist_filter=0.8, max_iterations=12,
                                     use_differential=0)])
def test_partitioned_chain_equals_whole_map_run(n_slabs, kw):
    import os
    g = os.path.join(os.path.dirname(__file__), "golden")
    c400, c401 = np.load(os.path.join(g, "car_cloud400.npz")), np.load(os.path.join(g, "car_cloud401.npz"))
    map4, mapn, read4 = O.feat4(c400["xyz"]), c400["normals"], O.feat4(c401["xyz"])
    r0, m0, r1 = _whole_then_slabs(map4, mapn, read4, n_slabs, **kw)
    assert r0["status"] == 0 == r1["status"]
    assert r1["iterations"] == r0["iterations"] and r1["iterations"] > 3
    assert (bits(r1["T"]) == bits(r0["T"])).all(), np.abs(r1["T"] - r0["T"]).max()
    for T in r1["all_T"]:                                       # every slab ends on the same pose
        assert (bits(T) == bits(r0["T"])).all()
    ids, d, w = r1["matches"]
    assert np.array_equal(ids, m0[0]) and (bits(d) == bits(m0[1])).all() and (bits(w) == bits(m0[2])).all()
    assert 0 < w.sum() < w.size                                 # TrimmedDist did trim, over all pairs
    if n_slabs > 1:
        assert sum(r1["slab_points"]) >= map4.shape[0] and min(r1["owned_first_iteration"]) > 0


def test_partitioned_chain_on_the_benchmarked_config():
    """C2's sizes (131 072-point scan, 5 M-point map) cut into two slabs: same bar"""
    from open3d_slam_private_b200 import synth
    import bench
    d = synth.make_c2(M=5_000_000, N=131072, n_poses=1)
    map4, mapn = synth.feat4(d["map_xyz"]), d["map_normals"]
    s = d["scans"][0]
    # two cuts through the scan (an equal-count cut of this long map would leave the scan to one slab)
    w = s["xyz"].astype(np.float64) @ np.asarray(s["T_init"], np.float64)[:3, :3].T + np.asarray(s["T_init"], np.float64)[:3, 3]
    axis = int(np.argmax(np.ptp(map4[:, :2], axis=0)))
    cuts = np.quantile(w[:, axis], [0.35, 0.7]).astype(np.float32)
    plan = (axis, np.array([-np.inf, cuts[0], cuts[1], np.inf], np.float32))
    r0, m0, r1 = _whole_then_slabs(map4, mapn, synth.feat4(s["xyz"]), 3, T_init=s["T_init"], plan=plan, **bench.icp_kwargs())
    assert r0["status"] == 0 == r1["status"] and r1["iterations"] == r0["iterations"]
    assert (bits(r1["T"]) == bits(r0["T"])).all()
    ids, dd, w = r1["matches"]
    assert np.array_equal(ids, m0[0]) and (bits(dd) == bits(m0[1])).all() and (bits(w) == bits(m0[2])).all()
    assert min(r1["owned_first_iteration"]) > 1000              # the scan really straddles the cut


def test_slab_handle_error_paths():
    from open3d_slam_private_b200 import capi as B
    import os
    g = os.path.join(os.path.dirname(__file__), "golden")
    c400 = np.load(os.path.join(g, "car_cloud400.npz"))
    map4, mapn = O.feat4(c400["xyz"]), c400["normals"]
    mean = map4[:, :3].astype(np.float64).mean(0)
    icp = B.ICP(B.default_icp_params())
    with pytest.raises(B.B200Error):
        icp.slab_begin(map4[:100])                               # no slab reference yet
    icp.init_reference_slab(map4, mapn, None, mean, 0, -np.inf, np.inf)
    with pytest.raises(B.B200Error):
        icp.compute(map4[:100])                                  # a slab handle does not run whole registrations
    with pytest.raises(B.B200Error):
        icp.slab_match(0, 0)                                     # null exchange arrays
    with pytest.raises(B.B200Error):
        icp.slab_finish()                                        # nothing has run
    with pytest.raises(B.B200Error):
        B.ICP(B.default_icp_params(knn=3)).init_reference_slab(map4, mapn, None, mean, 0, -np.inf, np.inf)
    icp.close()


def test_stream_ordered_driver_on_one_slab():
    """partition.SlabPmICP.register — the multi-rank driver: match enqueued on the handle's own stream
    (b200_icp_slab_match_async), the exchange ordered on that stream (b200_icp_slab_stream as torch's current stream), one
    host synchronisation per iteration — with a single slab and an in-place "all-reduce" that adds nothing: the result is
    the whole-map run bit for bit, twice in a row on the same handle."""
    import os
    import torch
    from open3d_slam_private_b200 import capi as B, partition as P
    g = os.path.join(os.path.dirname(__file__), "golden")
    c400, c401 = np.load(os.path.join(g, "car_cloud400.npz")), np.load(os.path.join(g, "car_cloud401.npz"))
    map4, mapn, read4 = O.feat4(c400["xyz"]), c400["normals"], O.feat4(c401["xyz"])
    prm = B.default_icp_params(maxDist=0.5, trimmed_ratio=0.9, max_iterations=30, min_diff_trans=0.008)
    whole = B.ICP(prm)
    whole.init_reference(map4, mapn)
    r0 = whole.compute(read4, None, None)
    m0 = whole.matches()
    whole.close()
    mean = r0["T_refIn_refMean"][:3, 3].astype(np.float64)
    calls = []

    def fake_allreduce(t):            # runs on the current stream = the handle's: ordering, not arithmetic, is under test
        calls.append(int(torch.cuda.current_stream().cuda_stream))
        t.add_(0)

    pm = P.SlabPmICP(map4, mapn, np.arange(map4.shape[0], dtype=np.int32), mean, 0, -np.inf, np.inf, params=prm,
                     allreduce=fake_allreduce)
    for rep in range(2):
        r1 = pm.register(read4)
        assert r1["status"] == 0 and r1["iterations"] == r0["iterations"]
        assert (bits(r1["T"]) == bits(r0["T"])).all()
        ids, d, w = pm.matches()
        assert np.array_equal(ids, m0[0]) and (bits(d) == bits(m0[1])).all() and (bits(w) == bits(m0[2])).all()
    assert calls and all(c == pm.icp.slab_stream() for c in calls)
    pm.close()
