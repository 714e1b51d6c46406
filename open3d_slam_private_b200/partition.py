"""Spatially partitioned map registration (BASELINE config C5): the map is cut into one slab per
rank along its longest horizontal axis, every slab is padded by a halo of the correspondence distance, the scan is
replicated, and per ICP iteration the ranks exchange ONE allreduce of 32 doubles (JtJ, Jtr, correspondence count,
sum of squared distances). Every rank then applies the same update redundantly — no broadcast.

Exactness: a source point is answered by exactly one rank (the one whose un-padded interval [lo, hi) holds its
transformed coordinate), and every map point within the correspondence distance of it lies inside that rank's padded
slab, so the union of the ranks' correspondences is the single-device result (ids are global, ties go to the smallest
global index because slabs keep the map order).

Two loops run partitioned. SlabICP: boundary B's (Open3D RegistrationICP: 1-NN within max_correspondence_distance,
no global quantile) with one 32-double exchange per iteration. SlabPmICP: the libpointmatcher chain of caller A
(TrimmedDist's global quantile, float H/b, the reference's checkers): the ranks exchange the MATCHES of an iteration
(every query is matched by the one rank that owns it; the sum over the ranks of the zero-filled arrays is the complete
set), then every rank runs the rest of the iteration on all of them with the same persistent kernel that serves the
un-partitioned registration — so the result is that registration's, bit for bit.
"""
import numpy as np

from . import capi as B

N_SUMS = 32


def plan_slabs(map_xyz, n_slabs, axis=None):
    """-> (axis, edges): edges has n_slabs + 1 float32 entries, edges[0] = -inf, edges[-1] = +inf; slab g owns
    coordinates in [edges[g], edges[g + 1]). Cuts at equal point counts along the longest horizontal axis."""
    xyz = np.asarray(map_xyz, np.float32)[:, :3]
    if axis is None:
        ext = xyz[:, :2].max(0) - xyz[:, :2].min(0)
        axis = int(np.argmax(ext))
    c = xyz[:, axis]
    edges = np.empty(n_slabs + 1, np.float32)
    edges[0], edges[-1] = -np.inf, np.inf
    if n_slabs > 1:
        qs = np.arange(1, n_slabs) * (c.size / n_slabs)
        part = np.partition(c, qs.astype(np.int64))
        edges[1:-1] = part[qs.astype(np.int64)]
    if np.any(np.diff(edges) <= 0):
        raise ValueError("degenerate slabs: too many identical coordinates along axis %d" % axis)
    return axis, edges


def halo_width(max_dist, coord_scale):
    """halo that is safe against the float32 rounding of the distance test: a neighbour accepted with a computed
    squared distance <= max_dist^2 is never more than this far along one axis"""
    return np.float32(max_dist * (1.0 + 1e-5) + 4e-7 * coord_scale)


def slab_members(map_xyz, axis, edges, g, max_dist):
    """global indices (ascending) of the map points in slab g's padded interval"""
    c = np.asarray(map_xyz, np.float32)[:, axis]
    h = halo_width(max_dist, float(np.abs(c).max()) if c.size else 0.0)
    return np.flatnonzero((c >= edges[g] - h) & (c < edges[g + 1] + h)).astype(np.int32)


class SlabICP:
    """One rank's share: the padded slab as target of a device registration handle, plus the loop that alternates the
    device steps with an allreduce. `allreduce(tensor)` sums a CUDA double tensor of N_SUMS entries over the ranks in
    place on the current stream (torch.distributed.all_reduce with NCCL); None = single slab."""

    def __init__(self, slab4, slab_normals, gid, axis, lo, hi, kind=B.O3D_POINT_TO_PLANE, max_correspondence_distance=1.0,
                 max_iteration=30, relative_fitness=1e-6, relative_rmse=1e-6, device=0, allreduce=None, poll_every=4):
        import torch
        self.torch = torch
        self.device = device
        self.h = B.O3dICP(kind=kind, max_correspondence_distance=max_correspondence_distance,
                          max_iteration=max_iteration, relative_fitness=relative_fitness, relative_rmse=relative_rmse,
                          device=device)
        self.h.set_target(slab4, slab_normals)
        self.h.set_partition(axis, lo, hi, gid)
        self.max_iteration = max_iteration
        self.allreduce = allreduce
        self.poll_every = max(1, int(poll_every))
        self.sums = torch.zeros(N_SUMS, dtype=torch.float64, device="cuda:%d" % device)
        self.collectives = 0
        self.peers = False

    def connect_peers(self, dist):
        """Exchange the sums over NVLink / NVSwitch peer memory inside the solve kernel instead of an allreduce between
        kernels: every rank allocates a mailbox, the CUDA IPC handles travel once through dist.all_gather_object, and
        from then on a registration is enqueued whole (no collective, no host polling)."""
        rank, world = dist.get_rank(), dist.get_world_size()
        _, handle = self.h.peer_init(rank, world)
        handles = [None] * world
        dist.all_gather_object(handles, handle)
        self.h.peer_connect(handles=handles)
        dist.barrier()
        self.peers = True

    def register(self, src4, src_normals=None, init=None):
        torch = self.torch
        if self.peers:
            self.h.enqueue_peers(src4, src_normals, init)
            res = self.h.finish()
            res["collectives"] = 0
            return res
        with torch.cuda.device(self.device):
            # torch's default stream is the legacy NULL stream: name it explicitly (cudaStreamLegacy == 0x1), a NULL
            # argument would select the handle's own stream and lose the ordering with the allreduce
            stream = torch.cuda.current_stream().cuda_stream or 1
            self.h.begin(src4, src_normals, init, stream=stream)
            ptr = self.sums.data_ptr()
            self.collectives = 0
            for it in range(self.max_iteration + 1):
                self.h.step_linearize(ptr)
                if self.allreduce is not None:
                    self.allreduce(self.sums)
                    self.collectives += 1
                self.h.step_update(ptr)
                # every rank sees the same totals, so `done` flips on the same iteration everywhere
                if (it + 1) % self.poll_every == 0 and self.h.poll()[0]:
                    break
            res = self.h.finish()
        res["collectives"] = self.collectives
        return res

    def correspondences(self):
        """(ids, squared dists) of the last match; ids are global map indices, -1 no neighbour, -2 another rank's point"""
        return self.h.correspondences()

    def close(self):
        self.h.close()


def merge_correspondences(ids_list):
    """union of the per-slab id arrays (each source point is owned by exactly one slab)"""
    ids = np.stack(ids_list)
    owners = (ids != -2).sum(0)
    if not (owners == 1).all():
        raise AssertionError("ownership is not a partition: %d points with %s owners" % ((owners != 1).sum(), set(owners[owners != 1])))
    return ids.max(0)


def register_partitioned(map4, map_normals, src4, src_normals=None, init=None, dist=None, device=None, **kw):
    """Driver for one process per GPU: every rank passes the same map (or at least its own slab's points in map
    order), builds its slab and runs the loop with an NCCL allreduce (exchange="nccl") or with the sums exchanged over
    peer memory inside the solve kernel (exchange="peer"). dist = torch.distributed (initialised) or None."""
    import torch
    world = dist.get_world_size() if dist is not None else 1
    rank = dist.get_rank() if dist is not None else 0
    if device is None:
        device = rank % max(1, torch.cuda.device_count())
    axis, edges = plan_slabs(map4, world)
    md = kw.get("max_correspondence_distance", 1.0)
    gid = slab_members(map4, axis, edges, rank, md)
    ar = None
    if world > 1:
        def ar(t):
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
    exchange = kw.pop("exchange", "nccl")
    slab = SlabICP(np.ascontiguousarray(map4[gid]), None if map_normals is None else np.ascontiguousarray(map_normals[gid]),
                   gid, axis, float(edges[rank]), float(edges[rank + 1]), device=device, allreduce=ar, **kw)
    if world > 1 and exchange == "peer":
        slab.connect_peers(dist)
    res = slab.register(src4, src_normals, init)
    ids, _ = slab.correspondences()
    if world > 1:
        t = torch.from_numpy(ids).to("cuda:%d" % device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ids = t.cpu().numpy()
    res["ids"] = ids
    res["slab_points"] = int(gid.size)
    slab.close()
    return res


class SlabPmICP:
    """One rank's share of the libpointmatcher chain over a partitioned map (b200_icp_slab_*, include/b200reg.h).

    allreduce(tensor): sums a CUDA tensor over the ranks in place (torch.distributed.all_reduce with NCCL), None for a
    single slab. The exchange arrays (9 N floats + N int32 per iteration) are torch tensors on the handle's device."""

    def __init__(self, slab4, slab_normals, gid, map_mean, axis, lo, hi, params=None, device=0, allreduce=None):
        import torch
        self.torch = torch
        self.device = device
        self.icp = B.ICP(params, device=device)
        self.icp.init_reference_slab(slab4, slab_normals, gid, map_mean, axis, lo, hi)
        self.allreduce = allreduce
        self.fbuf = self.ibuf = None
        self._ext = None
        self.collectives = 0

    def _buffers(self, n):
        torch = self.torch
        if self.fbuf is None or self.fbuf.numel() != 9 * n:
            dev = "cuda:%d" % self.device
            self.fbuf = torch.zeros(9 * n, dtype=torch.float32, device=dev)
            self.ibuf = torch.zeros(n, dtype=torch.int32, device=dev)
            torch.cuda.synchronize(self.device)   # the fills ran on torch's stream, the kernels use the handle's
        return self.fbuf, self.ibuf

    def begin(self, read4, normals=None, T_init=None):
        self.icp.slab_begin(read4, normals, T_init)
        self.collectives = 0
        return self._buffers(int(np.asarray(read4).shape[0]))

    def match(self):
        self.icp.slab_match(self.fbuf.data_ptr(), self.ibuf.data_ptr())

    def iterate(self, fbuf=None, ibuf=None):
        fbuf = self.fbuf if fbuf is None else fbuf
        ibuf = self.ibuf if ibuf is None else ibuf
        return self.icp.slab_iterate(fbuf.data_ptr(), ibuf.data_ptr())

    def register(self, read4, normals=None, T_init=None, raise_on_error=True):
        """the whole loop of one rank: match -> all-reduce of the exchange arrays -> rest of the iteration"""
        torch = self.torch
        with torch.cuda.device(self.device):
            self.begin(read4, normals, T_init)
            # the exchange runs on the handle's own stream (made torch's current stream, so NCCL orders itself against
            # it): match kernel -> all-reduces -> rest of the iteration with one host synchronisation per iteration (the
            # read-back of `iterate`) instead of three
            if self._ext is None:
                self._ext = torch.cuda.ExternalStream(self.icp.slab_stream(), device=self.device)
            with torch.cuda.stream(self._ext):
                go = True
                while go:
                    self.icp.slab_match_async(self.fbuf.data_ptr(), self.ibuf.data_ptr())
                    if self.allreduce is not None:
                        self.allreduce(self.fbuf)
                        self.allreduce(self.ibuf)
                        self.collectives += 2
                    go = self.iterate()
            res = self.icp.slab_finish(raise_on_error)
        res["collectives"] = self.collectives
        return res

    def matches(self):
        return self.icp.matches()

    def close(self):
        self.icp.close()


def map_mean(map4):
    """centroid of the whole map in double (ICP.cpp:883-890 subtracts its float rounding from the reference)"""
    return np.asarray(map4, np.float32)[:, :3].astype(np.float64).mean(0)


def register_slabs_pm_one_device(map4, map_normals, read4, n_slabs, params=None, T_init=None, mean=None, device=0,
                                 read_normals=None, plan=None):
    """All slabs on ONE device, one after the other (tests, and the proof that the partition reproduces the
    un-partitioned run): the exchange is a plain sum of the slabs' arrays."""
    import torch
    axis, edges = plan_slabs(map4, n_slabs) if plan is None else plan     # plan = (axis, edges): cuts chosen by the caller
    md = float(params.maxDist) if params is not None else 1.0
    mean = map_mean(map4) if mean is None else mean
    slabs, owned = [], []
    for g in range(n_slabs):
        gid = slab_members(map4, axis, edges, g, md)
        slabs.append(SlabPmICP(np.ascontiguousarray(map4[gid]), None if map_normals is None else np.ascontiguousarray(map_normals[gid]),
                               gid, mean, axis, float(edges[g]), float(edges[g + 1]), params=params, device=device))
    with torch.cuda.device(device):
        for sl in slabs:
            sl.begin(read4, read_normals, T_init)
        go, iters = True, 0
        while go:
            for sl in slabs:
                sl.match()
            if iters == 0:
                owned = [int((sl.ibuf != 0).sum().item()) for sl in slabs]   # (id 0 of slab 0 also reads as "not owned")
            ftot = torch.stack([sl.fbuf for sl in slabs]).sum(0)
            itot = torch.stack([sl.ibuf for sl in slabs]).sum(0, dtype=torch.int32)
            torch.cuda.synchronize()
            flags = [sl.iterate(ftot, itot) for sl in slabs]
            assert len(set(flags)) == 1, "the slabs disagree on whether to go on"
            go = flags[0]
            iters += 1
        results = [sl.icp.slab_finish(False) for sl in slabs]
    out = results[0]
    out["all_T"] = [r["T"] for r in results]
    out["matches"] = slabs[0].matches()
    out["slab_points"] = [int(sl.icp._ref4.shape[0]) for sl in slabs]
    out["owned_first_iteration"] = owned
    for sl in slabs:
        sl.close()
    return out


def register_partitioned_pm(map4, map_normals, read4, params=None, T_init=None, dist=None, device=None, mean=None,
                            read_normals=None):
    """One process per GPU: every rank passes the same map (or at least its own slab's points in map order and the
    whole map's centroid), builds its slab and runs the loop with two NCCL all-reduces per iteration."""
    import torch
    world = dist.get_world_size() if dist is not None else 1
    rank = dist.get_rank() if dist is not None else 0
    if device is None:
        device = rank % max(1, torch.cuda.device_count())
    axis, edges = plan_slabs(map4, world)
    md = float(params.maxDist) if params is not None else 1.0
    gid = slab_members(map4, axis, edges, rank, md)
    ar = None
    if world > 1:
        def ar(t):
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
    slab = SlabPmICP(np.ascontiguousarray(map4[gid]), None if map_normals is None else np.ascontiguousarray(map_normals[gid]),
                     gid, map_mean(map4) if mean is None else mean, axis, float(edges[rank]), float(edges[rank + 1]),
                     params=params, device=device, allreduce=ar)
    res = slab.register(read4, read_normals, T_init, raise_on_error=False)
    res["slab_points"] = int(gid.size)
    slab.close()
    return res
