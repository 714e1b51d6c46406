#!/usr/bin/env python
"""Secondary benchmarks for the BASELINE configs that are not bench.py's headline (C2):

  --mode c3   GICP, covariances from k = 20 normals, 128k scan vs 5M map, one B200            (registrations/s)
  --mode c4   batch of independent submap-to-submap registrations sharded over the ranks, no collective (pairs/s)
  --mode c5   city-scale map spatially partitioned over the ranks, one NCCL allreduce of 32 doubles per iteration
              (registrations/s; --tile-points per rank, 25 M x 8 = 200 M at full size)

Launch like bench.py (python bench_modes.py --mode c4 --gpus 1, or torchrun ... bench_modes.py --mode c5 --gpus N).
Timing: CUDA events on the stream the kernels run on, max over ranks; rank 0 prints ONE JSON line. Inputs larger
than L2 where stated. Everything runs through the C ABI (ctypes); no CPU fallback.
"""
import argparse
import json
import os
import time

import numpy as np

from bench import ClockSampler, dist_env, pinned_copy


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", choices=["c3", "c4", "c5"], required=True)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--map-points", type=int, default=5_000_000)
    ap.add_argument("--scan-points", type=int, default=131072)
    ap.add_argument("--poses", type=int, default=4)
    ap.add_argument("--batch", type=int, default=256, help="c4: registrations in the batch")
    ap.add_argument("--pair-points", type=int, default=200_000, help="c4: points per submap")
    ap.add_argument("--tile-points", type=int, default=25_000_000, help="c5: map points per rank")
    ap.add_argument("--max-dist", type=float, default=0.5)
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="c5: sums over NVLink peer memory inside the solve kernel, or an NCCL allreduce between kernels")
    return ap.parse_args()


def setup():
    import torch
    from open3d_slam_private_b200 import capi as B
    rank, world, local = dist_env()
    B.require_device()
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return torch, B, dist, rank, world, local


def max_over_ranks(torch, dist, vals):
    if dist is None:
        return list(vals)
    t = torch.tensor(list(vals), device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def emit(rank, dist, out):
    if rank == 0:
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


_NIDX = {}


def device_normals(B, xyz4, knn, device):
    """estimateNormalsOrCovariancesIfNeeded on the device: k-NN normals, normalised, flipped towards the origin. The
    index handle is kept between calls (its device buffers are reused), as the C++ shims do."""
    idx = _NIDX.get(device)
    if idx is None:
        idx = _NIDX[device] = B.Index(xyz4, device=device)
    else:
        idx.update(xyz4)
    n = idx.self_normals(knn=knn)["normals"]
    flip = (n * (-xyz4[:, :3])).sum(1) < 0
    n[flip] *= -1
    return np.ascontiguousarray(n, np.float32)


# ------------------------------------------------------------------------------------------------------------ C3
def run_c3(args):
    torch, B, dist, rank, world, local = setup()
    from open3d_slam_private_b200 import synth
    d = synth.make_c2(M=args.map_points, N=args.scan_points, n_poses=args.poses)
    tgt4 = synth.feat4(d["map_xyz"])
    t0 = time.perf_counter()
    tgtn = device_normals(B, tgt4, 20, local)          # GICP "covariances from k = 20 neighbours" (normals -> C)
    normals_map_s = time.perf_counter() - t0
    h = B.O3dICP(kind=B.O3D_GENERALIZED, max_correspondence_distance=args.max_dist, max_iteration=30, device=local)
    h.set_target(tgt4, tgtn)
    scans = []
    for s in d["scans"]:
        s4 = synth.feat4(s["xyz"])
        scans.append((pinned_copy(s4), s["T_init"].astype(np.float64), s["T_true"]))
    clk = ClockSampler(local) if rank == 0 else None
    nrm_ms = reg_ms = 0.0
    iters, perr = [], 0.0
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    t_mark0 = None
    for i in range(args.warmup + args.steps):
        if i == args.warmup:
            torch.cuda.synchronize()
            t_mark0 = time.time()
        (keep, s4), Ti, Tt = scans[(rank + i * world) % len(scans)]
        t0 = time.perf_counter()
        sn = device_normals(B, s4, 20, local)          # per-scan covariances: part of the registration's cost
        t1 = time.perf_counter()
        r = h.register(s4, sn, init=Ti)
        if i >= args.warmup:
            nrm_ms += (t1 - t0) * 1e3
            reg_ms += r["total_ms"]
            iters.append(r["iterations"])
            perr = max(perr, float(np.abs(r["T"][:3, 3] - Tt[:3, 3]).max()))
    clocks = clk.stop(t_mark0, time.time()) if clk else None
    reg_ms, nrm_ms = max_over_ranks(torch, dist, [reg_ms, nrm_ms])
    K = args.steps
    emit(rank, dist, {
        "metric": "GICP registrations/sec (128k scan vs 5M map, covariances from k=20 normals)",
        "value": world * K / (reg_ms * 1e-3), "unit": "registrations/s", "n_gpus": world, "steps": K,
        "warmup": args.warmup, "ms_per_step": reg_ms / K, "higher_is_better": True, "scaling": "weak",
        "dtype": "f32 match / f64 linearisation", "data": "synthetic",
        "config": {"workload": "C3", "map_points": args.map_points, "scan_points": args.scan_points,
                   "iterations_per_step": iters, "max_abs_translation_error_vs_truth_m": perr,
                   "scan_normals_k20_wall_ms_per_step_incl_h2d_d2h": nrm_ms / K,
                   "map_normals_k20_wall_s_incl_h2d_d2h": normals_map_s,
                   "timed": "H2D of scan + normals, 31 x (match+linearise, solve) launches; CUDA events"},
        "clocks": clocks})
    h.close()


# ------------------------------------------------------------------------------------------------------------ C4
def run_c4(args):
    torch, B, dist, rank, world, local = setup()
    from open3d_slam_private_b200 import batch, synth
    distinct = synth.make_c4(n_distinct=min(8, args.batch), N=args.pair_points)
    hosts = []
    for p in distinct:
        hosts.append(dict(tgt=pinned_copy(synth.feat4(p["tgt_xyz"])), src=pinned_copy(synth.feat4(p["src_xyz"])),
                          T_init=p["T_init"], T_true=p["T_true"]))
    mine = batch.shard_round_robin(args.batch, rank, world)
    h = B.O3dICP(kind=B.O3D_POINT_TO_PLANE, max_correspondence_distance=args.max_dist, max_iteration=30, device=local)

    def one(idx):
        p = hosts[idx % len(hosts)]
        tgt4, src4 = p["tgt"][1], p["src"][1]
        h.set_target_estimate_normals(tgt4, knn=10, max_dist=2.0)   # loop-closure submaps arrive without normals
        r = h.register(src4, init=p["T_init"])
        return r, p

    for i in range(args.warmup):
        one(i)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    clk = ClockSampler(local) if rank == 0 else None
    t_mark0 = time.time()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    t0 = time.perf_counter()
    reg_ms, iters, perr = 0.0, [], 0.0
    for idx in mine:
        r, p = one(idx)
        reg_ms += r["total_ms"]
        iters.append(r["iterations"])
        perr = max(perr, float(np.abs(r["T"][:3, 3] - p["T_true"][:3, 3]).max()))
    ev1.record()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = clk.stop(t_mark0, time.time()) if clk else None
    wall_ms, reg_ms = max_over_ranks(torch, dist, [wall_ms, reg_ms])
    emit(rank, dist, {
        "metric": "submap-to-submap registrations/sec (batch sharded over GPUs, no collective)",
        "value": args.batch / (wall_ms * 1e-3), "unit": "registrations/s", "n_gpus": world, "steps": args.batch,
        "warmup": args.warmup, "ms_per_step": wall_ms / max(1, len(mine)), "higher_is_better": True, "scaling": "strong",
        "dtype": "f32 match / f64 linearisation", "data": "synthetic",
        "config": {"workload": "C4", "batch": args.batch, "points_per_submap": args.pair_points,
                   "per_registration": "H2D both submaps, index build, k=10 normals of the target on its own index, point-to-plane ICP "
                                       "(Open3D semantics), D2H result — wall clock per rank, max over ranks",
                   "registration_only_ms_max_rank_total": reg_ms, "mean_iterations": float(np.mean(iters)),
                   "max_abs_translation_error_vs_truth_m": perr},
        "clocks": clocks})
    h.close()


# ------------------------------------------------------------------------------------------------------------ C5
def run_c5(args):
    torch, B, dist, rank, world, local = setup()
    from open3d_slam_private_b200 import partition as P
    from open3d_slam_private_b200 import synth
    street = synth.TiledStreet(world)
    halo = float(P.halo_width(args.max_dist, world * street.L))
    t0 = time.perf_counter()
    xyz, nrm, gid = street.slab(rank, args.tile_points, halo)
    gen_s = time.perf_counter() - t0
    edges = street.edges()
    ar = None
    if dist is not None:
        def ar(t):
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
    t0 = time.perf_counter()
    slab = P.SlabICP(synth.feat4(xyz), nrm, gid, 0, float(edges[rank]), float(edges[rank + 1]), kind=B.O3D_POINT_TO_PLANE,
                     max_correspondence_distance=args.max_dist, max_iteration=30, device=local, allreduce=ar)
    if dist is not None and args.exchange == "peer":
        slab.connect_peers(dist)
    build_s = time.perf_counter() - t0
    scans = [(pinned_copy(synth.feat4(s["xyz"])), s["T_init"], s["T_true"])
             for s in synth.make_c5_scans(street, args.poses, args.scan_points)]
    clk = ClockSampler(local) if rank == 0 else None
    reg_ms, iters, perr, owned, coll = 0.0, [], 0.0, [], 0
    t_mark0 = None
    for i in range(args.warmup + args.steps):
        if i == args.warmup:
            if dist is not None:
                dist.barrier()
            torch.cuda.synchronize()
            t_mark0 = time.time()
        (keep, s4), Ti, Tt = scans[i % len(scans)]
        r = slab.register(s4, init=Ti)
        if i >= args.warmup:
            reg_ms += r["total_ms"]
            iters.append(r["iterations"])
            coll += r["collectives"]
            perr = max(perr, float(np.abs(r["T"][:3, 3] - Tt[:3, 3]).max()))
            owned.append(int((slab.correspondences()[0] != -2).sum()))
    clocks = clk.stop(t_mark0, time.time()) if clk else None
    (reg_ms,) = max_over_ranks(torch, dist, [reg_ms])
    own_all = [owned]
    if dist is not None:
        own_all = [None] * world
        dist.all_gather_object(own_all, owned)
    K = args.steps
    emit(rank, dist, {
        "metric": "scan-to-map registrations/sec, map spatially partitioned over the GPUs (H/b summed over the ranks)",
        "value": K / (reg_ms * 1e-3), "unit": "registrations/s", "n_gpus": world, "steps": K, "warmup": args.warmup,
        "ms_per_step": reg_ms / K, "higher_is_better": True, "scaling": "weak (map grows with the ranks, one scan)",
        "dtype": "f32 match / f64 linearisation", "data": "synthetic",
        "config": {"workload": "C5", "map_points_total": world * args.tile_points, "map_points_per_rank": args.tile_points,
                   "slab_points_rank0_incl_halo": int(gid.size), "scan_points": args.scan_points, "halo_m": halo,
                   "iterations_per_step": iters, "collectives_per_step": coll / K,
                   "owned_scan_points_per_rank_first_step": [o[0] if o else 0 for o in own_all],
                   "max_abs_translation_error_vs_truth_m": perr, "slab_generation_s": gen_s,
                   "slab_h2d_and_index_build_s": build_s,
                   "exchange": args.exchange if dist is not None else "none",
                   "timed": ("H2D of the scan, (match+linearise, exchange over peer memory + solve) per iteration, all "
                             "iterations enqueued at once, no host polling; " if (dist is not None and args.exchange == "peer")
                             else "H2D of the scan, (match+linearise, reduce, NCCL allreduce, solve) per iteration, polled "
                                  "every 4 iterations; ") + "CUDA events on the launching stream, max over ranks"},
        "clocks": clocks})
    slab.close()


if __name__ == "__main__":
    a = parse()
    {"c3": run_c3, "c4": run_c4, "c5": run_c5}[a.mode](a)
