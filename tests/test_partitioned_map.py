"""Spatially partitioned map.

CPU part (gloo, world_size 2): the slab plan is a partition, the halo makes each slab self-sufficient, and the
per-rank correspondences / H, b sums — computed here by the ORACLE standing in for the device — combine over the
process group to the single-process result (ids bit-exact, sums to rounding).
GPU part: two slabs driven step by step on one device with the sums added in between (what the allreduce does)
reproduce the un-partitioned device registration; with >= 2 GPUs the same through NCCL.
"""
import os
import time
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from open3d_slam_private_b200 import partition as P

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MD = 0.5


def _clouds():
    from oracle import oracle_py as O
    a = np.load(os.path.join(GOLD, "car_cloud400.npz"))
    b = np.load(os.path.join(GOLD, "car_cloud401.npz"))
    return O.feat4(a["xyz"]), np.ascontiguousarray(a["normals"], np.float32), O.feat4(b["xyz"])


def test_plan_is_a_partition_with_equal_counts():
    rng = np.random.default_rng(5)
    xyz = rng.uniform([-100, -10, 0], [100, 10, 5], size=(20000, 3)).astype(np.float32)
    for g in (1, 2, 3, 8):
        axis, edges = P.plan_slabs(xyz, g)
        assert axis == 0 and edges.size == g + 1 and np.isneginf(edges[0]) and np.isposinf(edges[-1])
        c = xyz[:, axis]
        counts = [int(((c >= edges[i]) & (c < edges[i + 1])).sum()) for i in range(g)]
        assert sum(counts) == xyz.shape[0] and max(counts) - min(counts) <= 2
        members = [P.slab_members(xyz, axis, edges, i, MD) for i in range(g)]
        for i, m in enumerate(members):
            assert (np.diff(m) > 0).all()            # map order kept: ties resolve as in the whole map
            assert m.size >= counts[i]
    with pytest.raises(ValueError):
        P.plan_slabs(np.zeros((100, 3), np.float32), 4)


def test_merge_requires_exactly_one_owner():
    a = np.array([3, -2, -1, -2], np.int32)
    b = np.array([-2, 7, -2, 5], np.int32)
    assert (P.merge_correspondences([a, b]) == [3, 7, -1, 5]).all()
    with pytest.raises(AssertionError):
        P.merge_correspondences([a, a])


def _oracle_rank_share(rank, world, tgt4, tgtn, q4):
    """what one rank's device computes for a given (already transformed) float scan, restated with the oracle"""
    from oracle import oracle_py as O
    axis, edges = P.plan_slabs(tgt4, world)
    gid = P.slab_members(tgt4, axis, edges, rank, MD)
    c = q4[:, axis]
    mine = (c >= edges[rank]) & (c < edges[rank + 1])
    ids = np.full(q4.shape[0], -2, np.int32)
    sums = np.zeros(P.N_SUMS)
    if mine.any():
        lid, d = O.KdTree(np.ascontiguousarray(tgt4[gid])).knn(np.ascontiguousarray(q4[mine]), 1, MD, threads=2)
        lid, d = lid.ravel(), d.ravel()
        ok = lid >= 0
        g = np.where(ok, gid[np.maximum(lid, 0)], -1).astype(np.int32)
        ids[mine] = g
        s = q4[mine][ok][:, :3].astype(np.float64)
        t = tgt4[g[ok]][:, :3].astype(np.float64)
        n = tgtn[g[ok]].astype(np.float64)
        J = np.concatenate([np.cross(s, n), n], 1)
        r = ((s - t) * n).sum(1)
        H = J.T @ J
        sums[:21] = [H[a, b] for b in range(6) for a in range(b + 1)]
        sums[21:27] = J.T @ r
        sums[27] = ok.sum()
        sums[28] = d[ok].astype(np.float64).sum()
    return ids, sums


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tgt4, tgtn, q4 = _clouds()
    ids, sums = _oracle_rank_share(rank, world, tgt4, tgtn, q4)
    ts = torch.from_numpy(sums)
    dist.all_reduce(ts, op=dist.ReduceOp.SUM)
    ti = torch.from_numpy(ids.copy())
    dist.all_reduce(ti, op=dist.ReduceOp.MAX)
    owners = torch.from_numpy((ids != -2).astype(np.int32))
    dist.all_reduce(owners, op=dist.ReduceOp.SUM)
    ids1, sums1 = _oracle_rank_share(0, 1, tgt4, tgtn, q4)
    ok = bool((owners.numpy() == 1).all()) and bool((ti.numpy() == ids1).all())
    ok = ok and bool(np.allclose(ts.numpy(), sums1, rtol=1e-12, atol=1e-9)) and ts.numpy()[27] == sums1[27]
    q.put((rank, ok))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_partitioned_sums_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(out) == [(0, True), (1, True)]


# ------------------------------------------------------------------------------------------------------ device
@pytest.mark.gpu
@pytest.mark.parametrize("n_slabs", [2, 3])
def test_slabs_on_one_device_equal_unpartitioned(n_slabs):
    from open3d_slam_private_b200 import capi as B
    tgt4, tgtn, src4 = _clouds()
    kw = dict(kind=0, max_correspondence_distance=MD, max_iteration=30)
    whole = B.O3dICP(**kw)
    whole.set_target(tgt4, tgtn)
    ref = whole.register(src4)
    ref_ids, ref_d = whole.correspondences()
    whole.close()
    axis, edges = P.plan_slabs(tgt4, n_slabs)
    hs, sums = [], []
    for g in range(n_slabs):
        gid = P.slab_members(tgt4, axis, edges, g, MD)
        h = B.O3dICP(**kw)
        h.set_target(np.ascontiguousarray(tgt4[gid]), np.ascontiguousarray(tgtn[gid]))
        h.set_partition(axis, float(edges[g]), float(edges[g + 1]), gid)
        with pytest.raises(B.B200Error):
            h.register(src4)                       # a slab must be driven stepwise
        hs.append(h)
        sums.append(torch.zeros(P.N_SUMS, dtype=torch.float64, device="cuda:0"))
    stream = torch.cuda.current_stream().cuda_stream or 1   # 0x1 = cudaStreamLegacy
    for h in hs:
        h.begin(src4, stream=stream)
    for it in range(31):
        for h, s in zip(hs, sums):
            h.step_linearize(s.data_ptr())
        tot = torch.stack(sums).sum(0)             # what the allreduce leaves on every rank
        for h in hs:
            h.step_update(tot.data_ptr())
        if all(h.poll()[0] for h in hs):
            break
    res = [h.finish() for h in hs]
    ids = P.merge_correspondences([h.correspondences()[0] for h in hs])
    for h in hs:
        h.close()
    for r in res:
        assert r["iterations"] == ref["iterations"]
        assert np.abs(r["T"] - ref["T"]).max() < 1e-10
        assert r["n_correspondences"] == ref["n_correspondences"]
        assert abs(r["fitness"] - ref["fitness"]) == 0.0 and abs(r["inlier_rmse"] - ref["inlier_rmse"]) < 1e-12
        assert np.array_equal(r["T"], res[0]["T"])  # every rank holds the same pose bit for bit
    assert (ids == ref_ids).all()


@pytest.mark.gpu
@pytest.mark.parametrize("n_slabs", [2, 3])
def test_slabs_exchanging_over_mailboxes_equal_unpartitioned(n_slabs):
    """The peer-memory exchange (o3d_exchange_solve_kernel) with all slabs on one device, one handle and one stream per
    slab: the kernels of the slabs run concurrently and hand each other their sums through the mailboxes; the result is
    the un-partitioned registration's (ids identical, pose to 1e-10), every slab holds the same pose bit for bit, and a
    second registration on the same handles (sequence numbers carry on) gives the same again."""
    from open3d_slam_private_b200 import capi as B
    tgt4, tgtn, src4 = _clouds()
    kw = dict(kind=0, max_correspondence_distance=MD, max_iteration=30)
    whole = B.O3dICP(**kw)
    whole.set_target(tgt4, tgtn)
    ref = whole.register(src4)
    ref_ids, _ = whole.correspondences()
    whole.close()
    axis, edges = P.plan_slabs(tgt4, n_slabs)
    hs, ptrs = [], []
    for g in range(n_slabs):
        gid = P.slab_members(tgt4, axis, edges, g, MD)
        h = B.O3dICP(**kw)
        h.set_target(np.ascontiguousarray(tgt4[gid]), np.ascontiguousarray(tgtn[gid]))
        h.set_partition(axis, float(edges[g]), float(edges[g + 1]), gid)
        ptrs.append(h.peer_init(g, n_slabs)[0])
        hs.append(h)
    for h in hs:
        h.peer_connect(pointers=ptrs, devices=[0] * n_slabs)
    for rep in range(2):
        for h in hs:
            h.enqueue_peers(src4)          # asynchronous, each on its handle's own stream
        res = [h.finish() for h in hs]
        ids = P.merge_correspondences([h.correspondences()[0] for h in hs])
        for r in res:
            assert r["iterations"] == ref["iterations"]
            assert np.abs(r["T"] - ref["T"]).max() < 1e-10
            assert r["n_correspondences"] == ref["n_correspondences"]
            assert np.array_equal(r["T"], res[0]["T"])
        assert (ids == ref_ids).all()
    for h in hs:
        h.close()


@pytest.mark.gpu
def test_peer_exchange_times_out_instead_of_hanging(monkeypatch):
    """A peer that never delivers (crashed rank, rank out of step): the exchange kernel gives up after the time limit
    (5 s of SM clock by default, 0.2 s here), the remaining enqueued iterations return at once, b200_o3d_finish reports
    the failure, and the device stays usable."""
    from open3d_slam_private_b200 import capi as B
    monkeypatch.setenv("B200_O3D_PEER_TIMEOUT_MS", "200")
    tgt4, tgtn, src4 = _clouds()
    kw = dict(kind=0, max_correspondence_distance=MD, max_iteration=30)
    axis, edges = P.plan_slabs(tgt4, 2)
    hs, ptrs = [], []
    for g in range(2):
        gid = P.slab_members(tgt4, axis, edges, g, MD)
        h = B.O3dICP(**kw)
        h.set_target(np.ascontiguousarray(tgt4[gid]), np.ascontiguousarray(tgtn[gid]))
        h.set_partition(axis, float(edges[g]), float(edges[g + 1]), gid)
        ptrs.append(h.peer_init(g, 2)[0])
        hs.append(h)
    for h in hs:
        h.peer_connect(pointers=ptrs, devices=[0, 0])
    t0 = time.perf_counter()
    hs[0].enqueue_peers(src4)              # slab 1 never takes part
    with pytest.raises(B.B200Error) as e:
        hs[0].finish()
    assert "did not arrive" in str(e.value)
    assert time.perf_counter() - t0 < 3.0
    for h in hs:
        h.close()
    whole = B.O3dICP(**kw)                 # the device is fine
    whole.set_target(tgt4, tgtn)
    assert whole.register(src4)["iterations"] > 0
    whole.close()


def _nccl_worker(rank, world, port, q, exchange="nccl"):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda:%d" % rank))
    tgt4, tgtn, src4 = _clouds()
    res = P.register_partitioned(tgt4, tgtn, src4, dist=dist, device=rank, kind=0, max_correspondence_distance=MD,
                                 max_iteration=30, exchange=exchange)
    q.put((rank, res["T"], res["iterations"], res["ids"], res["collectives"]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("exchange", ["nccl", "peer"])
def test_partitioned_nccl_two_gpus(exchange):
    from open3d_slam_private_b200 import capi as B
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    tgt4, tgtn, src4 = _clouds()
    whole = B.O3dICP(kind=0, max_correspondence_distance=MD, max_iteration=30)
    whole.set_target(tgt4, tgtn)
    ref = whole.register(src4)
    ref_ids, _ = whole.correspondences()
    whole.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q, exchange)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=300) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, T, iters, ids, coll in out:
        assert iters == ref["iterations"] and (coll >= iters + 1 if exchange == "nccl" else coll == 0)
        assert np.abs(T - ref["T"]).max() < 1e-10
        assert (ids == ref_ids).all()
    assert np.array_equal(out[0][1], out[1][1])
