#!/usr/bin/env python
"""The BASELINE configs beside bench.py's headline (C2), as functions bench.py calls and folds into its ONE JSON line
under "modes" (and a small CLI of its own: python bench_modes.py --mode c4 --gpus 1, or torchrun ... --mode c5).

  index_build   KDTreeMatcher::init replacement on the 5 M-point map (bbox -> histogram -> scan -> scatter)
  normals       SurfaceNormalDataPointsFilter k = 20 on the map (once) and on a scan (per registration)
  c3            GICP, covariances from k = 20 normals, 131 072-pt scan vs 5 M-pt map, one B200
  c4            batch of independent submap-to-submap registrations sharded over the ranks, no collective
  c5            city-scale map spatially partitioned over the ranks, sums exchanged over NVLink peer memory
  lc            loop-closure front half: FPFH of both submaps of a C4 pair + RANSAC feature matching

Every mode returns {"ms", "value", "unit", "roofline": {...}, "e2e": {...}, "cpu_baseline": {...} (rank 0, one GPU
only), ...}. Timing: CUDA events on the stream the kernels run on (inside the library), max over ranks. Algorithmic
bytes per unit are SURVEY.md §8(d)'s: 116 B/point index build, 376 B/point normals at k = 20, 92 B per GICP
correspondence-iteration, 64 B per point-to-plane correspondence-iteration. Everything goes through the C ABI.
"""
import argparse
import json
import os
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
B_INDEX_BUILD, B_NORMALS_K20, B_GICP, B_P2PLANE = 116, 376, 92, 64


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md)"


def roofline(kernel, alg_bytes, ms, per_unit):
    peak, src = hbm_peak()
    a = alg_bytes / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
    return {"kernel": kernel, "bound": "hbm", "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak,
            "traffic": None, "algorithmic_bytes_per_launch": alg_bytes, "algorithmic_bytes_per_unit": per_unit,
            "avg_launch_ms": ms, "peak_source": src}


def stats(v):
    v = np.asarray(v, np.float64)
    return {"median": float(np.median(v)), "p95": float(np.percentile(v, 95)), "mean": float(v.mean()), "n": int(v.size)}


class Ctx:
    """what bench.py hands to a mode: plumbing (torch, capi, process group) and the C2 clouds when already generated"""

    def __init__(self, args, torch, B, dist, rank, world, local, c2=None):
        self.args, self.torch, self.B, self.dist = args, torch, B, dist
        self.rank, self.world, self.local, self._c2 = rank, world, local, c2

    def c2(self):
        if self._c2 is None:
            from open3d_slam_private_b200 import synth
            d = synth.make_c2(M=self.args.map_points, N=self.args.scan_points, n_poses=min(4, self.args.poses))
            self._c2 = (synth.feat4(d["map_xyz"]), d["map_normals"],
                        [(synth.feat4(s["xyz"]), s["T_init"], s["T_true"]) for s in d["scans"]])
        return self._c2

    def max_over_ranks(self, vals):
        if self.dist is None:
            return [float(v) for v in vals]
        t = self.torch.tensor([float(v) for v in vals], device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    @property
    def with_cpu(self):
        return self.rank == 0 and self.world == 1 and not getattr(self.args, "no_cpu_baseline", False)


def pinned(torch, a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t, t.numpy()


def oriented(n, pts4):
    """NormalizeNormals + OrientNormalsTowardsCameraLocation(origin) (CloudRegistration.cpp:75-77) on the host"""
    n = np.ascontiguousarray(n, np.float32)
    flip = (n * (-pts4[:, :3])).sum(1) < 0
    n[flip] *= -1
    return n


# --------------------------------------------------------------------------------------------------- index build
def mode_index_build(ctx):
    B = ctx.B
    ref4, _, _ = ctx.c2()
    M = ref4.shape[0]
    keep, host = pinned(ctx.torch, ref4)
    t0 = time.perf_counter()
    idx = B.Index(host, device=ctx.local)
    first_wall = (time.perf_counter() - t0) * 1e3
    dev, wall = [], []
    for _ in range(5):
        t0 = time.perf_counter()
        idx.update(host)
        wall.append((time.perf_counter() - t0) * 1e3)
        dev.append(idx.info()["build_ms"])
    info = idx.info()
    idx.close()
    ms = float(np.median(dev))
    out = {"what": "KDTreeMatcher::init replacement: uniform-grid index over the %d-point map" % M,
           "ms": ms, "value": M / (ms * 1e-3), "unit": "points/s", "device_ms": stats(dev),
           "roofline": roofline("bbox/hist/scan/scatter (grid_index.cu)", float(M) * B_INDEX_BUILD, ms, B_INDEX_BUILD),
           "e2e": {"ms": float(np.median(wall)), "first_call_ms_incl_allocation": first_wall,
                   "h2d_bytes_per_step": M * 16, "d2h_bytes_per_step": 64,
                   "what": "b200_index_update(host map): H2D + build + host decisions, host wall clock"},
           "cell_size": info.get("cell_size"), "dims": info.get("dims")}
    if ctx.with_cpu:
        from oracle import oracle_py as O
        t0 = time.perf_counter()
        kd = O.KdTree(ref4)
        s = time.perf_counter() - t0
        del kd
        out["cpu_baseline"] = {"value": M / s, "unit": "points/s", "cores": 1, "kind": "port", "ms": s * 1e3,
                               "sample": "one kd-tree build over the same %d-point map (oracle, single thread like libnabo's)" % M}
    return out


# --------------------------------------------------------------------------------------------------- normals
def mode_normals(ctx, knn=20):
    B = ctx.B
    ref4, _, scans = ctx.c2()
    M, N = ref4.shape[0], scans[0][0].shape[0]
    midx = B.Index(ref4, device=ctx.local)
    map_dev, map_wall = [], []
    for _ in range(2):
        t0 = time.perf_counter()
        midx.self_normals(knn=knn)
        map_wall.append((time.perf_counter() - t0) * 1e3)
        map_dev.append(midx.last_device_ms())
    midx.close()
    sidx = B.Index(scans[0][0], device=ctx.local)
    scan_dev, scan_wall, upd = [], [], []
    for i in range(8):
        s4 = scans[i % len(scans)][0]
        t0 = time.perf_counter()
        sidx.update(s4)
        t1 = time.perf_counter()
        sidx.self_normals(knn=knn)
        t2 = time.perf_counter()
        upd.append((t1 - t0) * 1e3)
        scan_wall.append((t2 - t0) * 1e3)
        scan_dev.append(sidx.last_device_ms())
    sidx.close()
    mm, sm = float(np.min(map_dev)), float(np.median(scan_dev))
    out = {"what": "SurfaceNormalDataPointsFilter knn = %d (self k-NN + covariance + 3x3 eigen) on the device" % knn,
           "ms": mm, "value": M / (mm * 1e-3), "unit": "points/s",
           "map": {"points": M, "device_ms": mm, "wall_ms_incl_d2h_of_normals": float(np.min(map_wall)),
                   "roofline": roofline("surface_normal_warp_kernel", float(M) * B_NORMALS_K20, mm, B_NORMALS_K20)},
           "scan": {"points": N, "device_ms": stats(scan_dev), "index_update_wall_ms": float(np.median(upd)),
                    "wall_ms_incl_h2d_index_d2h": float(np.median(scan_wall)),
                    "roofline": roofline("surface_normal_warp_kernel", float(N) * B_NORMALS_K20, sm, B_NORMALS_K20)},
           "roofline": roofline("surface_normal_warp_kernel", float(M) * B_NORMALS_K20, mm, B_NORMALS_K20),
           "e2e": {"ms": float(np.min(map_wall)), "h2d_bytes_per_step": 0, "d2h_bytes_per_step": M * 12,
                   "what": "b200_index_self_normals on the resident map index, normals copied back"}}
    if ctx.with_cpu:
        from oracle import oracle_py as O
        s4 = scans[0][0]
        t0 = time.perf_counter()
        O.surface_normals(s4, knn)
        s = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": N / s, "unit": "points/s", "cores": 1, "kind": "port", "ms": s * 1e3,
                               "sample": "the %d-point scan (kd-tree build + k = %d normals), oracle on one thread as the "
                                         "reference filter runs; the 5 M map scales per point" % (N, knn)}
    return out


# --------------------------------------------------------------------------------------------------- C3: GICP
def mode_c3(ctx, steps=16, warmup=3):
    B, torch = ctx.B, ctx.torch
    ref4, _, scans = ctx.c2()
    M, N = ref4.shape[0], scans[0][0].shape[0]
    md = ctx.args.max_dist
    h = B.O3dICP(kind=B.O3D_GENERALIZED, max_correspondence_distance=md, max_iteration=30, device=ctx.local)
    # "covariances from k = 20 neighbours": estimateNormalsOrCovariancesIfNeeded on both clouds (hybrid radius 1.0,
    # knn 20, parameter_structure_definitions.lua:67-73), GICP derives C = R diag(eps, 1, 1) R^T from the normal
    t0 = time.perf_counter()
    h.set_target_estimate_normals(ref4, knn=20, max_dist=1.0)
    map_setup_ms = (time.perf_counter() - t0) * 1e3
    host = [(pinned(torch, s4), Ti.astype(np.float64), Tt) for s4, Ti, Tt in scans]
    reg, wall, iters, perr = [], [], [], 0.0
    last = None
    for i in range(warmup + steps):
        (keep, s4), Ti, Tt = host[i % len(host)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = h.register_estimate_normals(s4, knn=20, radius=1.0, init=Ti)   # per-scan covariances: part of the cost
        t1 = time.perf_counter()
        if i >= warmup:
            reg.append(r["total_ms"]); wall.append((t1 - t0) * 1e3); iters.append(r["iterations"])
            perr = max(perr, float(np.abs(r["T"][:3, 3] - Tt[:3, 3]).max()))
            last = (s4, Ti, r)
    ms = float(np.mean(reg))
    # every update is preceded by a match+linearise pass, and one more evaluates the final pose
    passes = float(np.mean(iters)) + 1.0
    out = {"what": "C3: GICP (Open3D semantics), covariances from k = 20 neighbours within 1 m on both clouds, %d-pt scan "
                   "vs %d-pt map, max corr. distance %.2f; a step = H2D of the scan, index + k = 20 normals of the scan on "
                   "the device, all iterations" % (N, M, md),
           "ms": ms, "value": 1e3 / ms, "unit": "registrations/s", "registration_ms": stats(reg),
           "map_h2d_index_normals_k20_wall_ms_once": map_setup_ms,
           "updates_per_registration": stats(iters), "max_abs_translation_error_vs_truth_m": perr,
           "roofline": roofline("o3d_match_lin_kernel (GICP rows) + scan normals", passes * N * B_GICP + N * B_NORMALS_K20, ms, B_GICP),
           "e2e": {"ms": float(np.mean(wall)), "value": 1e3 / float(np.mean(wall)), "unit": "registrations/s",
                   "h2d_bytes_per_step": N * 16 + 128, "d2h_bytes_per_step": 200,
                   "what": "b200_o3d_register_estimate_normals(host scan): host wall clock around the call"}}
    if ctx.with_cpu and last is not None:
        from oracle import oracle_py as O
        s4, Ti, r = last
        cores = os.cpu_count() or 1
        tgtn = O.o3d_estimate_normals(ref4[:1], 1, 1.0)   # warm the library
        t0 = time.perf_counter()
        sn = O.o3d_estimate_normals(s4, 20, 1.0, threads=cores)[0].astype(np.float32)
        t_sn = time.perf_counter() - t0
        t0 = time.perf_counter()
        tn = O.o3d_estimate_normals(ref4, 20, 1.0, threads=cores)[0].astype(np.float32)
        t_tn = time.perf_counter() - t0
        t0 = time.perf_counter()
        o = O.o3d_icp(2, s4, ref4, tn, sn, max_dist=md, init=Ti, max_iteration=30, threads=cores)
        s = time.perf_counter() - t0 + t_sn
        out["cpu_baseline"] = {"value": 1.0 / s, "unit": "registrations/s", "cores": cores, "kind": "port", "ms": s * 1e3,
                               "map_normals_s_once": t_tn,
                               "sample": "one registration of the same scan: k = 20 normals of the scan + GICP incl. the "
                                         "kd-tree build over the map, oracle (Open3D v0.15.1 restated), %d threads for "
                                         "the neighbour searches" % cores}
        out["parity"] = {"iterations_equal": bool(o["iterations"] == r["iterations"]),
                         "max_abs_dT_vs_oracle": float(np.abs(o["T"] - r["T"]).max())}
    h.close()
    return out


# --------------------------------------------------------------------------------------------------- C4: submap pairs
def mode_c4(ctx, warmup=3):
    B, torch, args = ctx.B, ctx.torch, ctx.args
    from open3d_slam_private_b200 import batch, synth
    distinct = synth.make_c4(n_distinct=min(8, args.batch), N=args.pair_points)
    hosts = [dict(tgt=pinned(torch, synth.feat4(p["tgt_xyz"])), src=pinned(torch, synth.feat4(p["src_xyz"])),
                  T_init=p["T_init"], T_true=p["T_true"]) for p in distinct]
    mine = batch.shard_round_robin(args.batch, ctx.rank, ctx.world)
    h = B.O3dICP(kind=B.O3D_POINT_TO_PLANE, max_correspondence_distance=args.max_dist, max_iteration=30, device=ctx.local)

    def one(idx):
        p = hosts[idx % len(hosts)]
        h.set_target_estimate_normals(p["tgt"][1], knn=10, max_dist=2.0)   # loop-closure submaps arrive without normals
        return h.register(p["src"][1], init=p["T_init"]), p

    for i in range(warmup):
        one(i)
    ctx.barrier()
    t0 = time.perf_counter()
    reg_ms, iters, perr = 0.0, [], 0.0
    for idx in mine:
        r, p = one(idx)
        reg_ms += r["total_ms"]
        iters.append(r["iterations"])
        perr = max(perr, float(np.abs(r["T"][:3, 3] - p["T_true"][:3, 3]).max()))
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    wall_ms, reg_ms = ctx.max_over_ranks([wall_ms, reg_ms])
    it_sum = float(np.sum(iters) + len(iters))
    per = reg_ms / max(1, len(mine))
    Np = args.pair_points
    out = {"what": "C4: %d independent %d-vs-%d-point submap registrations (point-to-plane, Open3D semantics), round-robin "
                   "over %d GPU(s), no collective" % (args.batch, Np, Np, ctx.world),
           "ms": wall_ms / max(1, len(mine)), "value": args.batch / (wall_ms * 1e-3), "unit": "registrations/s",
           "n_gpus": ctx.world, "scaling": "strong", "collective": "none",
           "registration_only_ms": per, "mean_updates": float(np.mean(iters)),
           "max_abs_translation_error_vs_truth_m": perr,
           "roofline": roofline("o3d_match_lin_kernel (point-to-plane)", it_sum * Np * B_P2PLANE / max(1, len(mine)), per, B_P2PLANE),
           "e2e": {"ms": wall_ms / max(1, len(mine)), "value": args.batch / (wall_ms * 1e-3), "unit": "registrations/s",
                   "h2d_bytes_per_step": 2 * Np * 16, "d2h_bytes_per_step": 200,
                   "what": "per pair: H2D both submaps, index build, k = 10 target normals on the device, registration, "
                           "result back; host wall clock over the rank's share, max over ranks"}}
    if ctx.with_cpu:
        from oracle import oracle_py as O
        p = distinct[0]
        tgt4, src4 = synth.feat4(p["tgt_xyz"]), synth.feat4(p["src_xyz"])
        cores = os.cpu_count() or 1
        t0 = time.perf_counter()
        st, on = O.surface_normals(tgt4, 10, 2.0)
        n = on["normals"].astype(np.float64)
        n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-300)
        n = oriented(n.astype(np.float32), tgt4)
        o = O.o3d_icp(0, src4, tgt4, n, None, max_dist=args.max_dist, init=p["T_init"].astype(np.float64),
                      max_iteration=30, threads=cores)
        s = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": 1.0 / s, "unit": "registrations/s", "cores": cores, "kind": "port", "ms": s * 1e3,
                               "sample": "one pair: k = 10 normals of the target (one thread) + point-to-plane ICP (kd-tree "
                                         "match on %d threads), oracle" % cores}
    h.close()
    return out


# --------------------------------------------------------------------------------------------------- C5: partitioned map
def mode_c5(ctx, steps=16, warmup=3):
    B, torch, args, dist = ctx.B, ctx.torch, ctx.args, ctx.dist
    from open3d_slam_private_b200 import partition as P
    from open3d_slam_private_b200 import synth
    world, rank = ctx.world, ctx.rank
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
                     max_correspondence_distance=args.max_dist, max_iteration=30, device=ctx.local, allreduce=ar)
    peer = dist is not None and args.exchange == "peer"
    if peer:
        slab.connect_peers(dist)
    build_s = time.perf_counter() - t0
    N = args.scan_points
    scans = [(pinned(torch, synth.feat4(s["xyz"])), s["T_init"], s["T_true"]) for s in synth.make_c5_scans(street, 4, N)]
    reg, iters, perr, owned = [], [], 0.0, []
    for i in range(warmup + steps):
        if i == warmup:
            ctx.barrier()
        (keep, s4), Ti, Tt = scans[i % len(scans)]
        r = slab.register(s4, init=Ti)
        if i >= warmup:
            reg.append(r["total_ms"]); iters.append(r["iterations"])
            perr = max(perr, float(np.abs(r["T"][:3, 3] - Tt[:3, 3]).max()))
            if len(owned) < len(scans):
                owned.append(int((slab.correspondences()[0] != -2).sum()))
    (tot,) = ctx.max_over_ranks([float(np.sum(reg))])
    own_all = [owned]
    if dist is not None:
        own_all = [None] * world
        dist.all_gather_object(own_all, owned)
    ms = tot / steps
    passes = float(np.mean(iters)) + 1.0
    out = {"what": "C5: %d-pt scan vs a %d-point map cut into %d x-slabs of %d points (+ halo %.2f m), one slab per GPU"
                   % (N, world * args.tile_points, world, args.tile_points, halo),
           "ms": ms, "value": 1e3 / ms, "unit": "registrations/s", "n_gpus": world, "scaling": "weak (the map grows with the ranks)",
           "collective": ("32 doubles per iteration over NVLink peer memory inside o3d_exchange_solve_kernel (no NCCL)"
                          if peer else ("NCCL allreduce(SUM) of 32 doubles per iteration" if dist is not None else "none")),
           "updates_per_registration": stats(iters), "max_abs_translation_error_vs_truth_m": perr,
           "owned_scan_points_per_rank": [o[0] if o else 0 for o in own_all],
           "slab_generation_s": gen_s, "slab_h2d_and_index_build_s": build_s,
           "roofline": roofline("o3d_match_lin_kernel (point-to-plane, owned points)", passes * N * B_P2PLANE, ms, B_P2PLANE),
           "e2e": {"ms": ms, "value": 1e3 / ms, "unit": "registrations/s", "h2d_bytes_per_step": N * 16 + 128,
                   "d2h_bytes_per_step": 200,
                   "what": "H2D of the scan on every rank + all iterations + result; CUDA events on the launching stream, "
                           "max over ranks"}}
    slab.close()
    # ---- the libpointmatcher chain of caller A (TrimmedDist 0.9 over ALL pairs) on the same slabs: b200_icp_slab_*
    try:
        own = (xyz[:, 0] >= edges[rank]) & (xyz[:, 0] < edges[rank + 1])
        sums = torch.tensor(list(xyz[own].astype(np.float64).sum(0)) + [float(own.sum())], dtype=torch.float64,
                            device="cuda:%d" % ctx.local)
        if dist is not None:
            dist.all_reduce(sums, op=dist.ReduceOp.SUM)
        mean = (sums[:3] / sums[3]).cpu().numpy()
        prm = B.default_icp_params(knn=1, maxDist=args.max_dist, use_trimmed=1, trimmed_ratio=0.9, max_iterations=30,
                                   use_differential=1, min_diff_rot=1e-3, min_diff_trans=8e-3, smooth_length=3)
        pm = P.SlabPmICP(synth.feat4(xyz), nrm, gid, mean, 0, float(edges[rank]), float(edges[rank + 1]), params=prm,
                         device=ctx.local, allreduce=ar)
        wall, its, perr2, col = [], [], 0.0, 0
        for i in range(2 + 8):
            if i == 2:
                ctx.barrier()
            (keep, s4), Ti, Tt = scans[i % len(scans)]
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = pm.register(s4, None, Ti, raise_on_error=False)
            t1 = time.perf_counter()
            if i >= 2:
                wall.append((t1 - t0) * 1e3); its.append(r["iterations"]); col = r["collectives"]
                perr2 = max(perr2, float(np.abs(r["T"][:3, 3] - Tt[:3, 3]).max()))
        (tot2,) = ctx.max_over_ranks([float(np.sum(wall))])
        out["pm_chain"] = {"what": "same slabs, libpointmatcher chain (knn 1, maxDist %.2f, TrimmedDist 0.90 over all pairs, "
                                   "Differential + Counter): per iteration a slab match kernel, an all-reduce of 9 N floats + N "
                                   "int32, one launch of the persistent kernel on the summed matches; host-driven, wall clock, "
                                   "max over ranks" % args.max_dist,
                           "ms": tot2 / len(wall), "iterations": stats(its), "status": int(r["status"]),
                           "collectives_last_registration": col, "max_abs_translation_error_vs_truth_m": perr2,
                           "parity": "bit-identical to the whole-map registration by construction; asserted on one device in "
                                     "tests/test_gpu_slab_pm.py"}
        pm.close()
    except Exception as e:   # the o3d line above stands on its own
        out["pm_chain"] = {"error": "%s: %s" % (type(e).__name__, e)}
    return out


# --------------------------------------------------------------------------------------------------- map maintenance
def mode_map(ctx, steps=12, warmup=2):
    """a mapping loop on the resident map — insert the scan (Submap::insertScan), cut the patch
    around the robot into the registration handle (cropSubmap + initReference), register the next scan — against what
    the reference does at every reference update: patch on the host -> H2D -> index build."""
    B, torch = ctx.B, ctx.torch
    ref4, refn, scans = ctx.c2()
    M, N = ref4.shape[0], scans[0][0].shape[0]
    voxel, r_build, r_match = 0.1, 40.0, 30.0
    m = B.Map(device=ctx.local)
    t0 = time.perf_counter()
    m.set(ref4[:, :3].astype(np.float64), refn.astype(np.float64))
    set_ms = (time.perf_counter() - t0) * 1e3
    h = B.O3dICP(kind=0, device=ctx.local)
    icp = B.ICP(B.default_icp_params(knn=1, maxDist=0.5, use_trimmed=1, trimmed_ratio=0.9, max_iterations=30,
                                     use_differential=1, min_diff_rot=1e-3, min_diff_trans=8e-3, smooth_length=3), device=ctx.local)
    base = B.ICP(icp.params, device=ctx.local)
    prepared = []
    for s4, Ti, Tt in scans:
        sn = h.estimate_normals(s4, knn=10, radius=1.0).astype(np.float32)   # merged scans carry normals (isMergeScanValid)
        prepared.append((pinned(torch, s4), pinned(torch, sn), Ti, Tt))
    h.close()
    ins, crp, reg, wall, sizes, patch, base_wall, outside = [], [], [], [], [], [], [], []
    for i in range(warmup + steps):
        (k1, s4), (k2, sn), Ti, Tt = prepared[i % len(prepared)]
        c_build = B.crop_params(kind=3, radius_max=r_build, center=Tt[:3, 3])
        c_match = B.crop_params(kind=1, radius_max=r_match, min_z=-5.0, max_z=20.0, center=Tt[:3, 3])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n_out, size = m.insert_scan(s4, sn, Tt, voxel, c_build)
        a = m.last_device_ms()
        K = m.crop_to_icp(c_match, icp)
        b = m.last_device_ms()
        r = icp.compute(s4, None, Ti)
        t1 = time.perf_counter()
        if i >= warmup:
            ins.append(a); crp.append(b); reg.append(r["total_ms"]); wall.append((t1 - t0) * 1e3)
            sizes.append(size); patch.append(K); outside.append(n_out)
        if i >= warmup and len(base_wall) < 4:   # the reference's way: the patch crosses PCIe and is indexed from scratch
            p4, n3, _ = m.crop(c_match)
            kp, hp = pinned(torch, p4)
            kn, hn = pinned(torch, n3)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            base.init_reference(hp, hn)
            base_wall.append((time.perf_counter() - t0) * 1e3)
    ms = float(np.mean(ins))
    # insert: read the map (48 B/pt) + write it (48) + hash traffic; crop: read 24 B/pt + write 32 B per kept point
    alg = float(np.mean(sizes)) * 96.0 + N * 28.0
    out = {"what": "map maintenance on the resident %d-point map: insert a %d-pt scan (voxel %.2f m inside a %.0f m "
                   "radius), cut the %.0f m patch into the ICP handle, register" % (M, N, voxel, r_build, r_match),
           "ms": ms, "value": 1e3 / float(np.mean(wall)), "unit": "map updates + registrations/s",
           "insert_scan_device_ms": stats(ins), "crop_to_icp_device_ms": stats(crp), "registration_ms": stats(reg),
           "loop_wall_ms": stats(wall), "map_points": stats(sizes), "patch_points": stats(patch),
           "pass_through_points": stats(outside), "map_upload_once_ms": set_ms,
           "host_patch_init_reference_wall_ms": stats(base_wall) if base_wall else None,
           "roofline": roofline("classify_insert / write_map (map_maint.cu)", alg, ms, 96),
           "e2e": {"ms": float(np.mean(wall)), "h2d_bytes_per_step": N * 28 + 128 + N * 16, "d2h_bytes_per_step": 600,
                   "what": "host wall clock of insert_scan + crop_to_icp + b200_icp_compute; no map point crosses PCIe"}}
    if ctx.with_cpu:
        from oracle import oracle_py as O
        (k1, s4), (k2, sn), Ti, Tt = prepared[0]
        mp, mn = m.download()
        t0 = time.perf_counter()
        O.map_insert_scan(mp, mn, s4, sn, Tt, voxel, kind=3, radius_max=r_build, center=Tt[:3, 3])
        s = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": 1.0 / s, "unit": "insertions/s", "cores": 1, "kind": "port", "ms": s * 1e3,
                               "sample": "one Submap::insertScan on the same map (oracle, single thread like the reference's loop)"}
    icp.close(); base.close(); m.close()
    return out


# --------------------------------------------------------------------------------------------------- loop closure
def mode_lc(ctx, steps=5):
    """Submap::computeFeatures (voxelise 0.5 m -> normals -> FPFH, Submap.cpp:255-275) on both
    submaps of a C4 pair, then RegistrationRANSACBasedOnFeatureMatching (PlaceRecognition.cpp:81-85) with the
    reference's defaults (Parameters.hpp:122-133)."""
    B, args = ctx.B, ctx.args
    from open3d_slam_private_b200 import synth
    p = synth.make_c4(n_distinct=1, N=args.pair_points)[0]
    prep, est, lc = B.Prep(device=ctx.local), B.O3dICP(kind=B.O3D_POINT_TO_POINT, device=ctx.local), B.LoopClosure(device=ctx.local)

    def features(xyz):
        t0 = time.perf_counter()
        sp, _, _ = prep.voxel_downsample(0.5, synth.feat4(xyz))          # featureVoxelSize_
        vox_ms = prep.last_device_ms()
        sp4 = synth.feat4(sp.astype(np.float32))
        # normalKnn_ / normalEstimationRadius_ and featureRadius_ / featureKnn_ on one index, normals stay on the device
        nrm, f = lc.compute_features(sp4, normal_radius=1.0, normal_knn=10, fpfh_radius=2.5, fpfh_max_nn=100)
        return sp4, nrm, f, None, vox_ms, lc.last_device_ms(), (time.perf_counter() - t0) * 1e3

    fp_ms, wall_ms, rs_ms, rs_wall = [], [], [], []
    for it in range(steps + 1):
        s4, sn, sf, scnt, _, ms_s, w_s = features(p["src_xyz"])
        t4, tn, tf, tcnt, _, ms_t, w_t = features(p["tgt_xyz"])
        t0 = time.perf_counter()
        r = lc.ransac(s4, t4, sf, tf, seed=1 + it)
        w = (time.perf_counter() - t0) * 1e3
        if it:   # the first pass allocates
            fp_ms += [ms_s, ms_t]; wall_ms += [w_s, w_t]; rs_ms.append(r["device_ms"]); rs_wall.append(w)
    n_sp = (len(s4) + len(t4)) / 2
    dT = np.linalg.inv(p["T_true"]) @ r["T"]
    ms = float(np.median(fp_ms))
    per_pt = 16 + 24 + 4 * 100 + 33 * 8 * 2    # point, normal, neighbour ids, SPFH written + FPFH written
    out = {"what": "loop-closure front half on a C4 pair (%d-point submaps -> %d / %d sparse points): normals (radius 1, knn 10) + FPFH (radius 2.5, "
                   "knn 100) per submap in one call, then RANSAC feature matching (1 M iterations max, confidence 0.99, mutual filter, "
                   "distance + edge-length checkers)" % (args.pair_points, len(s4), len(t4)),
           "ms": ms, "value": n_sp / (ms * 1e-3), "unit": "points/s (normals + FPFH)", "fpfh_device_ms": stats(fp_ms),
           "ransac_device_ms": stats(rs_ms), "ransac_iterations_run": int(r["iterations_run"]),
           "ransac_validations": int(r["validations"]), "ransac_feature_correspondences": int(r["n_feature_correspondences"]),
           "ransac_inlier_correspondences": int(len(r["correspondences"])), "ransac_fitness": float(r["fitness"]),
           "ransac_kernel_launches": int(r["kernel_launches"]),
           "ransac_translation_error_vs_truth_m": float(np.abs(dT[:3, 3]).max()),
           "roofline": roofline("hybrid_knn + spfh + fpfh kernels (loop_closure.cu)", n_sp * per_pt, ms, per_pt),
           "e2e": {"ms": float(np.median(wall_ms)), "ransac_ms": float(np.median(rs_wall)),
                   "h2d_bytes_per_step": int(args.pair_points * 16), "d2h_bytes_per_step": int(n_sp * 33 * 8),
                   "what": "per submap: H2D of the 200 k points, voxel down-sample, normals and FPFH on the device, features "
                           "back (host wall clock); ransac_ms: b200_lc_ransac_feature_matching with host clouds and features"}}
    out["roofline"]["limiter"] = ("L2 gathers and double-precision transcendentals (acos, atan2 per pair), not HBM: the "
                                  "neighbour rows are re-read 100 times from L2")
    if ctx.with_cpu:
        from oracle import oracle_py as O
        cores = os.cpu_count() or 1
        t0 = time.perf_counter()
        on, _ = O.o3d_estimate_normals(s4, 10, 1.0, threads=cores)
        of, _ = O.o3d_fpfh(s4, on, 2.5, 100, threads=cores)
        s_f = time.perf_counter() - t0
        t0 = time.perf_counter()
        o = O.o3d_ransac(s4, t4, sf, tf, seed=steps + 1, threads=cores)
        s_r = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": len(s4) / s_f, "unit": "points/s (normals + FPFH)", "cores": cores, "kind": "port",
                               "ms": s_f * 1e3, "ransac_ms": s_r * 1e3,
                               "sample": "normals + FPFH of the source submap and one RANSAC feature matching, oracle on %d threads "
                                         "(Open3D's own loops are OpenMP-parallel)" % cores}
        out["parity"] = {"fpfh_max_abs_diff_median": float(np.median(np.abs(of - sf).max(1))),
                         "ransac_T_max_abs_diff": float(np.abs(o["T"] - r["T"]).max()),
                         "ransac_iterations_equal": bool(o["iterations_run"] == r["iterations_run"])}
    lc.close(); prep.close(); est.close()
    return out


MODES = {"index_build": mode_index_build, "normals": mode_normals, "c3": mode_c3, "c4": mode_c4, "c5": mode_c5,
         "map": mode_map, "lc": mode_lc}


def add_mode_args(ap):
    ap.add_argument("--batch", type=int, default=256, help="c4: registrations in the batch")
    ap.add_argument("--pair-points", type=int, default=200_000, help="c4: points per submap")
    ap.add_argument("--tile-points", type=int, default=25_000_000, help="c5: map points per rank")
    ap.add_argument("--max-dist", type=float, default=0.5, help="c3/c4/c5: max correspondence distance")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="c5: sums over NVLink peer memory inside the solve kernel, or an NCCL allreduce between kernels")


def main():
    import bench
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", choices=sorted(MODES), required=True)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--map-points", type=int, default=5_000_000)
    ap.add_argument("--scan-points", type=int, default=131072)
    ap.add_argument("--poses", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    add_mode_args(ap)
    args = ap.parse_args()
    ctx = bench.make_ctx(args)
    out = MODES[args.mode](ctx)
    if ctx.rank == 0:
        print(json.dumps({"mode": args.mode, **out}))
    if ctx.dist is not None:
        ctx.dist.destroy_process_group()


if __name__ == "__main__":
    main()
