"""Synthetic LiDAR-shaped workloads (BASELINE configs C2-C5): an analytic street-canyon scene
(ground, two facing walls, axis-aligned boxes, vertical cylinders), a jittered-lattice map with analytic normals and
ray-cast 64-beam scans. Pure numpy, seeded, generated at run time (nothing of this size is committed).
"""
import numpy as np

SEED = 20261104


class Scene:
    def __init__(self, seed=SEED, length=200.0, width=30.0, wall_h=12.0, n_boxes=200, n_cyl=100):
        rng = np.random.default_rng(seed)
        self.length, self.width, self.wall_h = float(length), float(width), float(wall_h)
        hx, hy = self.length / 2, self.width / 2
        # boxes: centre (x,y), half sizes, height; sit on the ground
        e = rng.uniform(0.5, 6.0, size=(n_boxes, 3))
        cx = rng.uniform(-hx + 4, hx - 4, n_boxes)
        cy = rng.uniform(-hy + 1, hy - 1, n_boxes)
        self.box_min = np.stack([cx - e[:, 0] / 2, cy - e[:, 1] / 2, np.zeros(n_boxes)], 1)
        self.box_max = np.stack([cx + e[:, 0] / 2, cy + e[:, 1] / 2, e[:, 2]], 1)
        self.cyl_c = np.stack([rng.uniform(-hx + 2, hx - 2, n_cyl), rng.uniform(-hy + 1, hy - 1, n_cyl)], 1)
        self.cyl_r = rng.uniform(0.15, 0.5, n_cyl)
        self.cyl_h = rng.uniform(2.0, 8.0, n_cyl)

    # ---- surfaces as parametric patches ----
    def _rects(self):
        """list of (origin, u, v, normal) rectangles"""
        hx, hy, H = self.length / 2, self.width / 2, self.wall_h
        R = [(np.array([-hx, -hy, 0.0]), np.array([2 * hx, 0, 0.0]), np.array([0, 2 * hy, 0.0]), np.array([0, 0, 1.0])),
             (np.array([-hx, -hy, 0.0]), np.array([2 * hx, 0, 0.0]), np.array([0, 0, H]), np.array([0, 1.0, 0])),
             (np.array([-hx, hy, 0.0]), np.array([2 * hx, 0, 0.0]), np.array([0, 0, H]), np.array([0, -1.0, 0]))]
        for mn, mx in zip(self.box_min, self.box_max):
            d = mx - mn
            R.append((np.array([mn[0], mn[1], mx[2]]), np.array([d[0], 0, 0]), np.array([0, d[1], 0]), np.array([0, 0, 1.0])))
            R.append((mn.copy(), np.array([d[0], 0, 0]), np.array([0, 0, d[2]]), np.array([0, -1.0, 0])))
            R.append((np.array([mn[0], mx[1], mn[2]]), np.array([d[0], 0, 0]), np.array([0, 0, d[2]]), np.array([0, 1.0, 0])))
            R.append((mn.copy(), np.array([0, d[1], 0]), np.array([0, 0, d[2]]), np.array([-1.0, 0, 0])))
            R.append((np.array([mx[0], mn[1], mn[2]]), np.array([0, d[1], 0]), np.array([0, 0, d[2]]), np.array([1.0, 0, 0])))
        return R

    def area(self):
        a = sum(np.linalg.norm(u) * np.linalg.norm(v) for _, u, v, _ in self._rects())
        a += float(np.sum(2 * np.pi * self.cyl_r * self.cyl_h))
        return a

    def sample_map(self, M, seed=SEED + 1, jitter=0.01):
        """exactly M points on the surfaces (jittered lattice), float32 xyz + analytic normals, shuffled."""
        rng = np.random.default_rng(seed)
        s = np.sqrt(self.area() / (M * 1.02))
        pts, nrm = [], []
        for o, u, v, n in self._rects():
            lu, lv = np.linalg.norm(u), np.linalg.norm(v)
            nu, nv = max(1, int(np.ceil(lu / s))), max(1, int(np.ceil(lv / s)))
            a = (np.arange(nu) + 0.5) / nu
            b = (np.arange(nv) + 0.5) / nv
            A, B = np.meshgrid(a, b, indexing="ij")
            p = o[None, :] + A.reshape(-1, 1) * u[None, :] + B.reshape(-1, 1) * v[None, :]
            pts.append(p)
            nrm.append(np.broadcast_to(n, p.shape))
        for c, r, h in zip(self.cyl_c, self.cyl_r, self.cyl_h):
            na, nh = max(3, int(np.ceil(2 * np.pi * r / s))), max(1, int(np.ceil(h / s)))
            ang = (np.arange(na) + 0.5) / na * 2 * np.pi
            zz = (np.arange(nh) + 0.5) / nh * h
            A, Z = np.meshgrid(ang, zz, indexing="ij")
            d = np.stack([np.cos(A.ravel()), np.sin(A.ravel()), np.zeros(A.size)], 1)
            p = np.stack([c[0] + r * d[:, 0], c[1] + r * d[:, 1], Z.ravel()], 1)
            pts.append(p)
            nrm.append(d)
        P = np.concatenate(pts)
        Nn = np.concatenate(nrm)
        if P.shape[0] < M:  # top up with random duplicates of the lattice (never happens with the 2 % margin)
            extra = rng.integers(0, P.shape[0], M - P.shape[0])
            P = np.concatenate([P, P[extra]])
            Nn = np.concatenate([Nn, Nn[extra]])
        perm = rng.permutation(P.shape[0])[:M]
        P = P[perm] + rng.uniform(-jitter, jitter, size=(M, 3))
        return P.astype(np.float32), np.ascontiguousarray(Nn[perm]).astype(np.float32)

    # ---- ray casting ----
    def raycast(self, origin, dirs, max_range=100.0):
        """distance along each ray to the first surface (inf when none within max_range)"""
        n = dirs.shape[0]
        t = np.full(n, np.inf)
        hx, hy, H = self.length / 2, self.width / 2, self.wall_h
        with np.errstate(divide="ignore", invalid="ignore"):
            tg = -origin[2] / dirs[:, 2]
            pg = origin[None, :2] + tg[:, None] * dirs[:, :2]
            ok = (tg > 0) & (np.abs(pg[:, 0]) <= hx) & (np.abs(pg[:, 1]) <= hy)
            t = np.where(ok & (tg < t), tg, t)
            for ywall in (-hy, hy):
                tw = (ywall - origin[1]) / dirs[:, 1]
                pw_x = origin[0] + tw * dirs[:, 0]
                pw_z = origin[2] + tw * dirs[:, 2]
                ok = (tw > 0) & (np.abs(pw_x) <= hx) & (pw_z >= 0) & (pw_z <= H)
                t = np.where(ok & (tw < t), tw, t)
            # boxes and cylinders: chunked over rays, the chunks spread over host threads (numpy releases the GIL; every
            # ray sees the same arithmetic whatever the chunking, so the scan does not depend on the thread count)
            inv = 1.0 / dirs

            def chunk(lo):
                sl = slice(lo, min(n, lo + 8192))
                t1 = (self.box_min[None, :, :] - origin[None, None, :]) * inv[sl, None, :]
                t2 = (self.box_max[None, :, :] - origin[None, None, :]) * inv[sl, None, :]
                tn = np.nanmax(np.minimum(t1, t2), axis=2)
                tf = np.nanmin(np.maximum(t1, t2), axis=2)
                hit = (tf >= np.maximum(tn, 0)) & (tn > 0)
                tb = np.where(hit, tn, np.inf).min(axis=1)
                # cylinders
                dxy = dirs[sl, :2]
                a = np.sum(dxy * dxy, axis=1)[:, None]
                oc = origin[None, None, :2] - self.cyl_c[None, :, :]
                b = 2 * np.sum(oc * dxy[:, None, :], axis=2)
                c = np.sum(oc * oc, axis=2) - self.cyl_r[None, :] ** 2
                disc = b * b - 4 * a * c
                tc = (-b - np.sqrt(np.where(disc >= 0, disc, np.nan))) / (2 * a)
                z = origin[2] + tc * dirs[sl, 2][:, None]
                hitc = (disc >= 0) & (tc > 0) & (z >= 0) & (z <= self.cyl_h[None, :])
                tcm = np.where(hitc, tc, np.inf).min(axis=1)
                return sl, np.minimum(tb, tcm)

            import os
            from concurrent.futures import ThreadPoolExecutor

            def guarded(lo):   # numpy's error state is per thread
                with np.errstate(divide="ignore", invalid="ignore"):
                    return chunk(lo)

            with ThreadPoolExecutor(max_workers=min(32, os.cpu_count() or 1)) as ex:
                for sl, tb in ex.map(guarded, range(0, n, 8192)):
                    t[sl] = np.minimum(t[sl], tb)
        t[t > max_range] = np.inf
        return t

    def scan(self, pose_xyyaw, n_points=131072, rings=64, seed=0, noise=0.02, height=1.8, max_range=100.0):
        """64-beam scan from a sensor pose; returns (points in the SENSOR frame (n,3) float32, T_map_sensor 4x4)."""
        rng = np.random.default_rng(SEED + 1000 + seed)
        x, y, yaw = pose_xyyaw
        az_n = int(np.ceil(n_points / rings * 1.25))
        elev = np.deg2rad(np.linspace(-24.8, 2.0, rings))
        az = (np.arange(az_n) + rng.uniform(0, 1)) / az_n * 2 * np.pi
        E, A = np.meshgrid(elev, az, indexing="ij")
        d_s = np.stack([np.cos(E) * np.cos(A), np.cos(E) * np.sin(A), np.sin(E)], -1).reshape(-1, 3)
        c, s = np.cos(yaw), np.sin(yaw)
        Rm = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1.0]])
        origin = np.array([x, y, height])
        t = self.raycast(origin, d_s @ Rm.T, max_range)
        hit = np.isfinite(t)
        idx = np.flatnonzero(hit)
        if idx.size < n_points:
            raise RuntimeError("scene too open for this pose: %d hits < %d" % (idx.size, n_points))
        # keep an even subset of the hits (preserves the ring structure)
        idx = idx[np.linspace(0, idx.size - 1, n_points).astype(np.int64)]
        r = t[idx] + rng.normal(0, noise, n_points)
        pts_s = d_s[idx] * r[:, None]
        T = np.eye(4)
        T[:3, :3] = Rm
        T[:3, 3] = origin
        return pts_s.astype(np.float32), T


def perturbation(seed, trans=0.25, rot_deg=2.0):
    """initial-guess error: translation U(-trans, trans) per axis, rotation rot_deg about a random axis"""
    rng = np.random.default_rng(SEED + 5000 + seed)
    tr = rng.uniform(-trans, trans, 3)
    ax = rng.normal(size=3)
    ax /= np.linalg.norm(ax)
    th = np.deg2rad(rot_deg)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = tr
    return T


def feat4(xyz):
    out = np.ones((xyz.shape[0], 4), np.float32)
    out[:, :3] = xyz
    return out


def make_c2(M=5_000_000, N=131072, n_poses=32, seed=SEED, scene_kw=None, only=None):
    """BASELINE config C2: map (M points + normals), n_poses scans of N points with truth and initial guesses.
    only: indices of the scans to ray-cast (the others are None) — several processes can share the work, every scan
    is the same whoever makes it."""
    sc = Scene(seed=seed, **(scene_kw or {}))
    map_xyz, map_n = sc.sample_map(M)
    rng = np.random.default_rng(seed + 77)
    scans = []
    for i in range(n_poses):
        pose = (rng.uniform(-sc.length / 2 + 20, sc.length / 2 - 20), rng.uniform(-sc.width / 2 + 4, sc.width / 2 - 4),
                rng.uniform(-np.pi, np.pi))
        if only is not None and i not in only:
            scans.append(None)
            continue
        pts, T_true = sc.scan(pose, n_points=N, seed=i)
        T_init = T_true @ perturbation(i)
        scans.append(dict(xyz=pts, T_true=T_true, T_init=T_init.astype(np.float32)))
    return dict(scene=sc, map_xyz=map_xyz, map_normals=map_n, scans=scans)


# ------------------------------------------------------------------------------------------- C4: submap pairs
def make_c4(n_distinct=8, N=200_000, seed=SEED, map_points=3_000_000):
    """BASELINE config C4 (loop-closure candidates): n_distinct source/target submap pairs of N points each. A submap
    is a voxelised map patch (uniform density, like the reference's submaps): the N map points nearest to a pose,
    expressed in that pose's frame. Source and target come from two independent samplings of the same street, from
    poses ~1.5 m apart, with the true relative transform. A batch of 256 cycles through them."""
    sc = Scene(seed=seed + 3)
    map_a, _ = sc.sample_map(map_points, seed=seed + 11)
    map_b, _ = sc.sample_map(map_points, seed=seed + 12)
    rng = np.random.default_rng(seed + 91)

    def patch(points, pose):
        x, y, yaw = pose
        c, s_ = np.cos(yaw), np.sin(yaw)
        T = np.eye(4)
        T[:3, :3] = np.array([[c, -s_, 0], [s_, c, 0], [0, 0, 1.0]])
        T[:3, 3] = [x, y, 0.0]
        d2 = (points[:, 0] - x) ** 2 + (points[:, 1] - y) ** 2
        keep = np.argpartition(d2, N)[:N]
        keep.sort()
        local = (points[keep].astype(np.float64) - T[:3, 3]) @ T[:3, :3]
        return local.astype(np.float32), T

    pairs = []
    for i in range(n_distinct):
        x, y = rng.uniform(-sc.length / 2 + 40, sc.length / 2 - 40), rng.uniform(-sc.width / 2 + 5, sc.width / 2 - 5)
        yaw = rng.uniform(-np.pi, np.pi)
        a, Ta = patch(map_a, (x, y, yaw))
        b, Tb = patch(map_b, (x + rng.uniform(-1.5, 1.5), y + rng.uniform(-1.0, 1.0), yaw + rng.uniform(-0.05, 0.05)))
        T_true = np.linalg.inv(Ta) @ Tb          # source (b) -> target (a) frame
        pairs.append(dict(tgt_xyz=a, src_xyz=b, T_true=T_true, T_init=T_true @ perturbation(300 + i, 0.15, 1.0)))
    return pairs


# ------------------------------------------------------------------------------------------- C5: tiled city map
class TiledStreet:
    """A street made of `n_tiles` street-canyon scenes of equal length laid end to end along x. Tile g is Scene(seed
    + 100 + g) shifted to x in [g L - L/2, g L + L/2). Each rank of the partitioned mode materialises only its own tile
    and the halo strips of its two neighbours."""

    def __init__(self, n_tiles, seed=SEED, **scene_kw):
        self.n = int(n_tiles)
        self.tiles = [Scene(seed=seed + 100 + g, **scene_kw) for g in range(self.n)]
        self.L = self.tiles[0].length
        self.width = self.tiles[0].width

    def edges(self):
        e = np.empty(self.n + 1, np.float32)
        e[0], e[-1] = -np.inf, np.inf
        for g in range(1, self.n):
            e[g] = np.float32(g * self.L - self.L / 2)
        return e

    def tile_points(self, g, m_per_tile):
        xyz, nrm = self.tiles[g].sample_map(m_per_tile, seed=SEED + 1 + g)
        xyz[:, 0] += np.float32(g * self.L)
        return xyz, nrm

    def slab(self, g, m_per_tile, halo):
        """-> (xyz, normals, gid) of tile g padded by the neighbours' points within `halo` of its borders; gid =
        tile index * m_per_tile + index inside the tile, ascending"""
        e = self.edges()
        parts = []
        for t in (g - 1, g, g + 1):
            if t < 0 or t >= self.n:
                continue
            xyz, nrm = self.tile_points(t, m_per_tile)
            gid = np.arange(t * m_per_tile, (t + 1) * m_per_tile, dtype=np.int64)
            keep = (xyz[:, 0] >= e[g] - halo) & (xyz[:, 0] < e[g + 1] + halo)
            parts.append((xyz[keep], nrm[keep], gid[keep]))
        xyz = np.concatenate([p[0] for p in parts])
        nrm = np.concatenate([p[1] for p in parts])
        gid = np.concatenate([p[2] for p in parts])
        if gid.max() >= 2 ** 31:
            raise ValueError("global ids exceed int32")
        return np.ascontiguousarray(xyz), np.ascontiguousarray(nrm), gid.astype(np.int32)

    def raycast(self, origin, dirs, max_range=100.0):
        t = np.full(dirs.shape[0], np.inf)
        for g, sc in enumerate(self.tiles):
            if abs(origin[0] - g * self.L) > self.L / 2 + max_range:
                continue
            o = origin.copy()
            o[0] -= g * self.L
            t = np.minimum(t, sc.raycast(o, dirs, max_range))
        return t

    def scan(self, pose_xyyaw, **kw):
        proxy = Scene.__new__(Scene)     # reuse Scene.scan's beam pattern with this street's ray caster
        proxy.raycast = self.raycast
        return Scene.scan(proxy, pose_xyyaw, **kw)


def make_c5_scans(street, n_poses=4, N=131072):
    """scans taken astride the border between the two middle tiles (both neighbours own part of every scan)"""
    rng = np.random.default_rng(SEED + 177)
    xb = (street.n // 2) * street.L - street.L / 2 if street.n > 1 else 0.0
    out = []
    for i in range(n_poses):
        pose = (xb + rng.uniform(-3, 3), rng.uniform(-street.width / 2 + 4, street.width / 2 - 4), rng.uniform(-np.pi, np.pi))
        pts, T_true = street.scan(pose, n_points=N, seed=700 + i)
        out.append(dict(xyz=pts, T_true=T_true, T_init=T_true @ perturbation(700 + i)))
    return out
