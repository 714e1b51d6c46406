"""CPU tests of the multi-GPU host logic (one process per GPU, gloo, world_size 2): independent registrations are
sharded over ranks with no data-path collective and the results come back in input order on every rank."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from open3d_slam_private_b200 import batch


def test_shards_are_a_partition():
    for n in (0, 1, 7, 256):
        for world in (1, 2, 4, 8):
            seen = sorted(i for r in range(world) for i in batch.shard_round_robin(n, r, world))
            assert seen == list(range(n))
    costs = [5, 1, 9, 3, 3, 8, 2]
    parts = batch.shard_lpt(costs, 3)
    assert sorted(i for p in parts for i in p) == list(range(len(costs)))
    loads = [sum(costs[i] for i in p) for p in parts]
    assert max(loads) - min(loads) <= max(costs)
    assert parts == batch.shard_lpt(costs, 3)   # deterministic


def _fake_register(pair, params, device):
    # stands in for the device: a pose that encodes which pair was registered and by which "device"
    T = np.eye(4, dtype=np.float32)
    T[0, 3] = float(pair["read4"].shape[0])
    T[1, 3] = float(device)
    return dict(T=T, status=0, iterations=3, device_ms=0.0, point_used_ratio=1.0)


def _worker(rank, world, port, balance, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pairs = [dict(ref4=np.ones((10 + i, 4), np.float32), read4=np.ones((3 + i, 4), np.float32)) for i in range(9)]
    res = batch.register_batch(pairs, None, device=rank, rank=rank, world=world, dist=dist, balance=balance,
                               register_fn=_fake_register)
    P = batch.poses_as_array(res)
    ok = all(P[i, 0, 3] == 3 + i for i in range(9)) and set(P[:, 1, 3].tolist()) == {0.0, 1.0}
    if balance == "round_robin":
        ok = ok and all(P[i, 1, 3] == i % world for i in range(9))
    q.put((rank, bool(ok)))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(balance):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, balance, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(out) == [(0, True), (1, True)]


def test_register_batch_world2_gloo_round_robin():
    _run("round_robin")


def test_register_batch_world2_gloo_lpt():
    _run("lpt")
