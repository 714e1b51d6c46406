#!/usr/bin/env python
"""bench.py — BASELINE config C2: point-to-plane ICP of a synthetic 64-beam LiDAR scan (131 072 points) against a
5 000 000-point map on one B200, `param/icp.yaml` modules (KDTreeMatcher knn 1 / maxDist 0.5 / epsilon 0,
TrimmedDist 0.90, PointToPlane, Differential{1e-3, 8e-3, 3} + Counter{30}), degeneracyAwareness None.

A "step" is ONE registration (one scan -> pose). Metric: ICP registrations/s (ms per registration = ms_per_step).
  value     registrations/s with map index AND scan resident in HBM (b200_icp_run_resident), all device work of a
            registration timed with CUDA events on the handle's stream; L2 flushed between steps (outside the events)
  e2e       the same through the reference-facing call b200_icp_compute(host buffers): H2D of the scan from pinned
            host memory + registration + D2H of the result struct, CUDA events inside the library (first event before
            the H2D, last after the D2H)
  roofline  the persistent registration kernel: algorithmic bytes = iterations x N x 64 B over its
            CUDA-event duration, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline / --impl reference: the CPU oracle (restatement of the reference path; the reference itself cannot be
            built here) on the host cores, bounded sample of the same scans
N > 1: independent registrations (different scans, one map replica per GPU) sharded round-robin over ranks, no
data-path collective (SURVEY.md §8(e) "batched independent registrations") -> weak scaling.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BYTES_PER_QUERY_ITER = 64  # 16 query + 4 id + 4 dist + 16 matched point + 16 matched normal + 2x4 weight
METRIC = "icp_registrations_per_s_128k_scan_vs_5M_map"


# dram bytes per launch of icp_persistent_kernel<1> on the default workload, from the committed ncu capture
# (round-2 capture, two launches: 7.83 + 0.44 MB and 9.81 + 0.38 MB; their mean)
NCU_TRAFFIC_BYTES_PER_LAUNCH = 9.23e6
NCU_TRAFFIC_SOURCE = ("dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full, "
                      "profiles/r02_icp_kernel_ncu_summary.txt (default workload only)")



def _iter_summary(iters):
    """ICP iterations per timed step, summarised (the full list is as long as --steps)."""
    it = [int(v) for v in iters]
    if not it:
        return {}
    return {"mean": sum(it) / len(it), "min": min(it), "max": max(it), "first_8": it[:8]}

def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--map-points", type=int, default=5_000_000)
    ap.add_argument("--scan-points", type=int, default=131072)
    ap.add_argument("--poses", type=int, default=32, help="distinct synthetic scans (cycled over the steps; SURVEY 8(d): 32)")
    ap.add_argument("--modes", default="auto",
                    help="comma list of bench_modes.py modes folded into the line under \"modes\" (index_build, normals, c3, "
                         "c4, c5, map); auto: index_build,normals,c3,c4,map on one GPU, c4,c5 on several; none: headline only")
    import bench_modes
    bench_modes.add_mode_args(ap)
    ap.add_argument("--cpu-sample", type=int, default=3, help="registrations timed for the CPU baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--degeneracy", default="none", choices=["none", "xicp"],
                    help="xicp: degeneracyAwareness OptimizedEqualityConstraints with param/icp.yaml's thresholds")
    ap.add_argument("--no-flush", action="store_true", help="keep L2 warm between steps (reported in config)")
    return ap.parse_args()


def icp_kwargs():
    return dict(knn=1, maxDist=0.5, use_trimmed=1, trimmed_ratio=0.9, max_iterations=30, use_differential=1,
                min_diff_rot=1e-3, min_diff_trans=8e-3, smooth_length=3, minimizer=0)


def xicp_kwargs(args, oracle):
    """degeneracyAwareness of open3d_slam_ros/param/icp.yaml (250 / 180 / 80 deg / 45 deg) when --degeneracy xicp"""
    if getattr(args, "degeneracy", "none") != "xicp":
        return {}
    if oracle:
        return dict(degeneracy=1, xicp_enough=250.0, xicp_insufficient=180.0)   # angles: oracle defaults (80 / 45 deg)
    return dict(degeneracy=1, xicp_enough_information=250.0, xicp_insufficient_information=180.0,
                xicp_min_alignment_angle_deg=80.0, xicp_strong_alignment_angle_deg=45.0)


def workload(args, ctx=None):
    """the C2 clouds. Under torchrun the ranks share the ray casting (rank r makes scans r, r + world, ...) and hand
    the scans to each other through /tmp; every scan is bit-identical whoever makes it."""
    from open3d_slam_private_b200 import synth
    t0 = time.time()
    world = ctx.world if ctx is not None else 1
    rank = ctx.rank if ctx is not None else 0
    only = None if world == 1 else set(range(rank, args.poses, world))
    d = synth.make_c2(M=args.map_points, N=args.scan_points, n_poses=args.poses, only=only)
    if world > 1:
        import tempfile
        tag = os.path.join(tempfile.gettempdir(), "b200_bench_scan_%s_%d_%d" % (os.environ.get("MASTER_PORT", "0"),
                                                                              args.scan_points, args.poses))
        for i in sorted(only):
            s = d["scans"][i]
            np.savez(tag + "_%d.tmp.npz" % i, xyz=s["xyz"], T_true=s["T_true"], T_init=s["T_init"])
            os.replace(tag + "_%d.tmp.npz" % i, tag + "_%d.npz" % i)
        ctx.barrier()
        for i in range(args.poses):
            if d["scans"][i] is None:
                z = np.load(tag + "_%d.npz" % i)
                d["scans"][i] = dict(xyz=z["xyz"], T_true=z["T_true"], T_init=z["T_init"])
        ctx.barrier()
        for i in sorted(only):
            os.remove(tag + "_%d.npz" % i)
    ref4 = synth.feat4(d["map_xyz"])
    scans = [(synth.feat4(s["xyz"]), s["T_init"], s["T_true"]) for s in d["scans"]]
    return ref4, d["map_normals"], scans, time.time() - t0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.rows, self.proc, self.th = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def mark(self):
        return time.time()

    def stop(self, t0=None, t1=None):
        """summary over the samples taken in [t0, t1] (the timed region); all samples when none fall inside"""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [r for (t, r) in self.rows if t0 is not None and t0 - 0.05 <= t <= t1 + 0.15]
        window = "timed region"
        if not inside:
            inside, window = [r for (_, r) in self.rows], "whole run (timed region shorter than the sampling period)"
        for r in inside:
            if len(r) < 8:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "window": window,
                "reasons": sorted(reasons)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def cpu_baseline(ref4, refn, scans, n_sample, threads, args=None, device_results=None, cross_check=False):
    """The CPU oracle on the host cores: initReference (kd-tree build) timed apart from compute(), like the SLAM
    (Mapper.cpp:343 vs :373). Returns registrations/s over the bounded sample; with device_results (the device's
    result for the same scans) also the parity of the two on this very workload."""
    from oracle import oracle_py as O
    t0 = time.perf_counter()
    ses = O.IcpSession(ref4, refn)
    build_s = time.perf_counter() - t0
    p = O.default_params(counter_first=0, knn_threads=threads, **icp_kwargs(), **xicp_kwargs(args, True))
    times, iters, res = [], [], []
    for i in range(n_sample):
        r4, Ti, _ = scans[i % len(scans)]
        t0 = time.perf_counter()
        o = ses.compute(p, r4, None, Ti)
        times.append(time.perf_counter() - t0)
        iters.append(o["iterations"])
        res.append(o)
    ses.close()
    tot = float(np.sum(times))
    out = dict(value=n_sample / tot, unit="registrations/s", cores=threads, kind="port",
               sample="%d registrations of the same C2 scans (%d pts vs %d-pt map), oracle kd-tree kNN on %d thread(s); "
                      "kd-tree build %.1f s not included" % (n_sample, scans[0][0].shape[0], ref4.shape[0], threads, build_s),
               ms_per_registration=1e3 * tot / n_sample, iterations=iters, build_s=build_s)
    if device_results is not None:
        dT = [float(np.abs(o["T"] - g["T"]).max()) for o, g in zip(res, device_results)]
        out["parity"] = {"iterations_equal": all(o["iterations"] == g["iterations"] for o, g in zip(res, device_results)),
                         "max_abs_dT_vs_oracle": max(dT), "registrations_compared": len(dT),
                         "iterations_oracle": iters, "iterations_device": [int(g["iterations"]) for g in device_results]}
    if cross_check:
        out["knn_cross_check"] = knn_cross_check(ref4, scans[0], threads)
    return out


def knn_cross_check(ref4, scan, threads):
    """so that the CPU arm is provably not a strawman, the oracle's kd-tree 1-NN (all threads) beside
    scipy.spatial.cKDTree(workers=-1) on the same queries (one scan at its initial guess, maxDist 0.5)."""
    from oracle import oracle_py as O
    r4, Ti, _ = scan
    q4 = np.ascontiguousarray(r4 @ np.asarray(Ti, np.float32).T, np.float32)
    q4[:, 3] = 1.0
    t0 = time.perf_counter()
    kd = O.KdTree(ref4)
    ob = time.perf_counter() - t0
    t0 = time.perf_counter()
    oi, od = kd.knn(q4, 1, 0.5, threads=threads)
    oq = time.perf_counter() - t0
    out = {"queries": int(q4.shape[0]), "oracle_build_s": ob, "oracle_query_ms": oq * 1e3, "threads": threads}
    try:
        from scipy.spatial import cKDTree
        t0 = time.perf_counter()
        tree = cKDTree(ref4[:, :3])
        sb = time.perf_counter() - t0
        t0 = time.perf_counter()
        sd, si = tree.query(q4[:, :3], k=1, distance_upper_bound=0.5, workers=-1)
        sq = time.perf_counter() - t0
        si = np.where(np.isfinite(sd), si, -1)
        out.update(ckdtree_build_s=sb, ckdtree_query_ms=sq * 1e3,
                   ids_agree_fraction=float((si == oi.ravel()).mean()))
    except Exception as e:   # scipy missing: report, do not fail the bench
        out["ckdtree"] = "unavailable: %s" % e
    return out


def config_dict(args, extra=None):
    c = {"workload": "C2: synthetic 64-beam LiDAR scan (%d pts) point-to-plane ICP vs %d-pt map, param/icp.yaml modules "
                     "(knn 1, maxDist 0.5, eps 0, TrimmedDist 0.90, Differential{1e-3,8e-3,3}+Counter{30})"
                     % (args.scan_points, args.map_points),
         "map_points": args.map_points, "scan_points": args.scan_points, "distinct_scans": args.poses,
         "degeneracyAwareness": "OptimizedEqualityConstraints{250,180,80,45}" if getattr(args, "degeneracy", "none") == "xicp" else "None",
         "step": "one registration (one scan -> pose)",
         "multi_gpu": "independent registrations sharded round-robin over ranks, one map replica per GPU, no collective"}
    if extra:
        c.update(extra)
    return c


def run_reference(args):
    """--impl reference: the reference's CPU path for this hot path = the oracle port (the reference itself needs
    Eigen/Boost/libnabo and cannot be built on this image), all host threads, bounded sample per step."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    ref4, refn, scans, gen_s = workload(args)
    threads = os.cpu_count() or 1
    from oracle import oracle_py as O
    t0 = time.perf_counter()
    ses = O.IcpSession(ref4, refn)
    build_s = time.perf_counter() - t0
    p = O.default_params(counter_first=0, knn_threads=threads, **icp_kwargs(), **xicp_kwargs(args, True))
    for i in range(args.warmup):
        ses.compute(p, scans[i % len(scans)][0], None, scans[i % len(scans)][1])
    t0 = time.perf_counter()
    iters = []
    for i in range(args.steps):
        r4, Ti, _ = scans[i % len(scans)]
        iters.append(ses.compute(p, r4, None, Ti)["iterations"])
    dt = time.perf_counter() - t0
    v = args.steps / dt
    sample = ("every step = one full registration of the C2 workload on the CPU oracle, kd-tree kNN over %d host "
              "threads; kd-tree build (%.1f s) outside the timed region" % (threads, build_s))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "registrations/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args, {"iterations_per_step": _iter_summary(iters)}),
        "cpu_baseline": {"value": v, "unit": "registrations/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "registrations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def pinned_copy(a):
    """host buffer in pinned memory (what a real-time caller would hand over); torch is plumbing only"""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t, t.numpy()


def make_ctx(args, c2=None):
    """device + process-group plumbing shared by the headline and the modes (torch is plumbing only)"""
    import torch
    import bench_modes
    from open3d_slam_private_b200 import capi as B
    rank, world, local = dist_env()
    B.require_device()  # no CPU fallback: fail loudly
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return bench_modes.Ctx(args, torch, B, dist, rank, world, local, c2)


def mode_list(args, world):
    if args.modes == "none":
        return []
    if args.modes == "auto":
        return ["index_build", "normals", "c3", "c4", "map", "lc"] if world == 1 else ["c4", "c5"]
    return [m for m in args.modes.split(",") if m]


def run_b200(args):
    import bench_modes
    ctx = make_ctx(args)
    ref4, refn, scans, gen_s = workload(args, ctx)
    ctx._c2 = (ref4, refn, scans[:4])
    torch, B, dist, rank, world, local = ctx.torch, ctx.B, ctx.dist, ctx.rank, ctx.world, ctx.local
    barrier = ctx.barrier
    N = scans[0][0].shape[0]
    icp = B.ICP(B.default_icp_params(**icp_kwargs(), **xicp_kwargs(args, False)), device=local)
    t0 = time.perf_counter()
    icp.init_reference(ref4, refn)
    init_ref_s = time.perf_counter() - t0
    keep, host_scans = [], []
    for r4, Ti, Tt in scans:
        t, a = pinned_copy(r4)
        keep.append(t)
        host_scans.append((a, Ti, Tt))
    # registration i of this rank: every rank walks through ALL distinct scans (offset by its rank), so the ranks do
    # the same work over the run; (rank + i * world) % poses would pin one scan (and its iteration count) to a rank
    # whenever world divides the number of scans
    mine = lambda i: host_scans[(rank + i) % len(host_scans)]
    flush = (lambda: None) if args.no_flush else icp.flush_l2

    clk = ClockSampler(local) if rank == 0 else None   # started early: nvidia-smi needs a moment to come up
    for i in range(args.warmup):
        a, Ti, _ = mine(i)
        icp.compute(a, None, Ti)
    # ---- value: everything resident, CUDA events around all device work of one registration
    barrier()
    t_mark0 = time.time()
    t_wall0 = time.perf_counter()
    step_ms, kern_ms = [], []
    iters, launches, touched, perr, grid = [], 0, 0, 0.0, 0
    for i in range(args.steps):
        a, Ti, Tt = mine(i)
        icp.upload_reading(a)          # not timed: inputs resident before the timed region of the step
        flush()
        r = icp.run_resident(Ti)
        step_ms.append(r["total_ms"]); kern_ms.append(r["device_ms"])
        iters.append(r["iterations"]); launches += r["kernel_launches"]; touched += r["points_touched"]
        grid = r["grid_ctas"]
        perr = max(perr, float(np.abs(r["T"][:3, 3] - Tt[:3, 3]).max()))
    barrier()
    # ---- e2e: host buffers through the reference-facing call, H2D + D2H inside the timed events; beside it the
    # caller's own wall clock around the same call (the L2 flush is drained first)
    e2e_step, e2e_wall = [], []
    for i in range(args.steps):
        a, Ti, _ = mine(i)
        flush()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = icp.compute(a, None, Ti)
        e2e_wall.append((time.perf_counter() - t0) * 1e3)
        e2e_step.append(r["e2e_ms"])
    barrier()
    wall_s = time.perf_counter() - t_wall0
    clocks = clk.stop(t_mark0, time.time()) if clk else None
    tot_ms, e2e_ms, kern_tot, e2e_wall_ms = float(np.sum(step_ms)), float(np.sum(e2e_step)), float(np.sum(kern_ms)), float(np.sum(e2e_wall))
    if dist is not None:
        tot_ms, e2e_ms, _, e2e_wall_ms = ctx.max_over_ranks([tot_ms, e2e_ms, kern_tot, e2e_wall_ms])
        cnt = torch.tensor([launches], device="cuda", dtype=torch.int64)
        dist.all_reduce(cnt)
        launches = int(cnt.item())
    # the device's results for the scans the CPU baseline will repeat (parity on the benchmarked workload itself)
    dev_results = []
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        for i in range(args.cpu_sample):
            a, Ti, _ = host_scans[i % len(host_scans)]
            dev_results.append(icp.compute(a, None, Ti))
    icp.close()
    del icp
    # ---- the other BASELINE configs, folded into the same line (every rank takes part; rank 0 reports)
    modes = {}
    for m in mode_list(args, world):
        t0 = time.perf_counter()
        try:
            modes[m] = bench_modes.MODES[m](ctx)
        except Exception as e:   # a mode must not take the headline down with it
            modes[m] = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
        modes[m]["bench_wall_s"] = time.perf_counter() - t0
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    K = args.steps
    value = world * K / (tot_ms * 1e-3)
    e2e_v = world * K / (e2e_ms * 1e-3)
    peaks, peak_src = {}, "fallback (B200_PROFILING.md)"
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak_src = "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    alg_bytes = float(sum(iters)) * N * BYTES_PER_QUERY_ITER          # rank 0's launches
    achieved = alg_bytes / (kern_tot * 1e-3) / 1e9
    pct = lambda v, q: float(np.percentile(np.asarray(v, np.float64), q))
    out = {
        "metric": METRIC, "value": value, "unit": "registrations/s", "n_gpus": world, "steps": K,
        "warmup": args.warmup, "ms_per_step": tot_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args, {
            "l2": "kept warm" if args.no_flush else "flushed (256 MiB write) before every timed step, outside the events",
            "iterations_per_step": _iter_summary(iters), "max_abs_translation_error_vs_truth_m": perr,
            "map_index_build_s_incl_h2d": init_ref_s, "synthetic_generation_s": gen_s, "grid_ctas": grid,
            "timed_region_wall_s": wall_s}),
        "ms_per_step_stats": {"median": pct(step_ms, 50), "p95": pct(step_ms, 95), "min": float(np.min(step_ms)),
                              "max": float(np.max(step_ms)), "of": "rank 0's steps, CUDA events"},
        "e2e": {"value": e2e_v, "unit": "registrations/s", "ms_per_step": e2e_ms / K,
                "ms_per_step_median": pct(e2e_step, 50), "ms_per_step_p95": pct(e2e_step, 95),
                "timed_with": "CUDA events inside b200_icp_compute: before the H2D of the scan .. after the D2H of the result",
                "e2e_wall_ms": e2e_wall_ms / K, "e2e_wall_ms_median": pct(e2e_wall, 50), "e2e_wall_ms_p95": pct(e2e_wall, 95),
                "e2e_wall_value": world * K / (e2e_wall_ms * 1e-3),
                "e2e_wall_what": "the caller's clock (time.perf_counter) around the same b200_icp_compute call through ctypes",
                "h2d_bytes_per_step": N * 16 + 64, "d2h_bytes_per_step": 160 + 256},
        "gpu_launches": launches,
        "roofline": {"kernel": "icp_persistent_kernel<1>", "bound": "hbm", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": NCU_TRAFFIC_BYTES_PER_LAUNCH,
                     "traffic_source": NCU_TRAFFIC_SOURCE,
                     "limiter": "issue/latency and grid barriers, not HBM: the touched map cells stay in L2",
                     "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg_bytes / K, "avg_launch_ms": kern_tot / K,
                     "knn_candidate_GBps": touched * 16 / (kern_tot * 1e-3) / 1e9,
                     "candidates_per_query_iteration": touched / max(1.0, float(sum(iters)) * N)},
        "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(ref4, refn, scans, args.cpu_sample, os.cpu_count() or 1, args, dev_results, cross_check=True)
        out["parity"] = cb.pop("parity", None)
        out["cpu_baseline"] = cb
        one = cpu_baseline(ref4, refn, scans, 1, 1, args)
        out["cpu_baseline"]["single_thread_value"] = one["value"]
    if modes:
        out["modes"] = modes
    print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()

def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
