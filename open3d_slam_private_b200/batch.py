"""Batched independent registrations sharded over GPUs (BASELINE config C4: loop-closure candidates
of PlaceRecognition.cpp:70-111 and odometry constraints of constraint_builders.cpp:92-118 are independent
registrations — the reference loops over them serially with its `omp parallel for` commented out).

One process per GPU (torch.distributed for the plumbing only). There is NO data-path collective: every rank registers
its own shard on its own device; only the small results (4x4 pose, iterations, fitness) are gathered at the end.
"""
import numpy as np


def shard_round_robin(n_items, rank, world):
    """registration r -> rank r mod world"""
    return list(range(rank, n_items, world))


def shard_lpt(costs, world):
    """size-balanced assignment (longest processing time first) by cost N*M; returns world lists of indices.
    Deterministic: ties broken by index."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    load = [0.0] * world
    out = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += costs[i]
    return [sorted(s) for s in out]


def default_register(pair, params, device):
    """one registration through the C ABI; pair = dict(ref4, ref_normals, read4, read_normals=None, T_init=None)"""
    from . import capi as B
    icp = B.ICP(params, device=device)
    try:
        icp.init_reference(pair["ref4"], pair.get("ref_normals"))
        r = icp.compute(pair["read4"], pair.get("read_normals"), pair.get("T_init"), raise_on_error=False)
    finally:
        icp.close()
    return dict(T=r["T"], status=r["status"], iterations=r["iterations"], device_ms=r["device_ms"],
                point_used_ratio=r["point_used_ratio"])


def register_batch(pairs, params=None, device=0, rank=0, world=1, dist=None, balance="round_robin",
                   register_fn=default_register):
    """Registers this rank's shard of `pairs` and returns the results of ALL pairs in input order on every rank
    (gathered with dist.all_gather_object when world > 1). `pairs` must be the same list on every rank (only the local
    shard's clouds are touched)."""
    n = len(pairs)
    if balance == "lpt":
        costs = [float(p["ref4"].shape[0]) * float(p["read4"].shape[0]) for p in pairs]
        mine = shard_lpt(costs, world)[rank]
    else:
        mine = shard_round_robin(n, rank, world)
    local = [(i, register_fn(pairs[i], params, device)) for i in mine]
    if world > 1:
        if dist is None:
            raise ValueError("world > 1 needs a torch.distributed handle")
        gathered = [None] * world
        dist.all_gather_object(gathered, local)
    else:
        gathered = [local]
    out = [None] * n
    for part in gathered:
        for i, r in part:
            out[i] = r
    if any(r is None for r in out):
        raise RuntimeError("a registration of the batch was not assigned to any rank")
    return out


def poses_as_array(results):
    return np.stack([np.asarray(r["T"], np.float32) for r in results])
