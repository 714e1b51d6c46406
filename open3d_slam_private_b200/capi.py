"""ctypes binding of the C ABI (include/b200reg.h). Thin: numpy arrays in the reference's memory layout go straight
through as pointers. No CPU fallback — a missing library or a missing device raises.

Layouts (identical to the reference's Eigen images): features (N,4) float32 C-contiguous == 4xN column-major;
normals (N,3); ids/dists/weights (N,k) == k x N column-major; 4x4 matrices are passed column-major and returned as
ordinary (row, col) numpy matrices.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBPATH = os.path.join(_HERE, "lib", "libb200reg.so")
_lib = None

(B200_OK, B200_E_NO_MATCHES, B200_E_BAD_PARAM, B200_E_NAN, B200_E_CUDA, B200_E_UNSUPPORTED, B200_E_NOT_INITIALIZED,
 B200_E_BOUND) = range(8)
OUTLIER_NULL, OUTLIER_MAXDIST, OUTLIER_MINDIST, OUTLIER_MEDIANDIST, OUTLIER_TRIMMEDDIST = range(5)

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)


class B200Error(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("b200 status %d: %s" % (status, msg))
        self.status = status


class ConvergenceError(B200Error):
    """PointMatcher::ConvergenceError analogue (status NO_MATCHES / NAN)."""


class IndexOpts(C.Structure):
    _fields_ = [("cell_size", C.c_float), ("target_per_cell", C.c_float), ("max_cells", C.c_uint64)]


class SurfaceNormalParams(C.Structure):
    _fields_ = [("knn", C.c_int), ("maxDist", C.c_float), ("epsilon", C.c_float), ("sortEigen", C.c_int),
                ("smoothNormals", C.c_int)]


class IcpParams(C.Structure):
    _fields_ = [
        ("knn", C.c_int), ("epsilon", C.c_float), ("maxDist", C.c_float),
        ("use_trimmed", C.c_int), ("trimmed_ratio", C.c_float),
        ("use_maxdist_filter", C.c_int), ("maxdist_filter", C.c_float),
        ("use_surface_normal_filter", C.c_int), ("surface_normal_max_angle", C.c_float),
        ("minimizer", C.c_int), ("max_iterations", C.c_int),
        ("use_differential", C.c_int), ("min_diff_rot", C.c_float), ("min_diff_trans", C.c_float),
        ("smooth_length", C.c_int), ("cell_size", C.c_float),
        ("degeneracy", C.c_int), ("xicp_enough_information", C.c_float), ("xicp_insufficient_information", C.c_float),
        ("xicp_min_alignment_angle_deg", C.c_float), ("xicp_strong_alignment_angle_deg", C.c_float),
        ("use_bound", C.c_int), ("max_rotation_norm", C.c_float), ("max_translation_norm", C.c_float),
    ]


class IcpResult(C.Structure):
    _fields_ = [
        ("status", C.c_int), ("iterations", C.c_int), ("max_iterations_reached", C.c_int), ("n_matches_last", C.c_int),
        ("T", C.c_float * 16), ("T_iter", C.c_float * 16), ("T_refIn_refMean", C.c_float * 16),
        ("T_refMean_readMean", C.c_float * 16), ("weighted_point_used_ratio", C.c_float),
        ("point_used_ratio", C.c_float), ("points_touched", C.c_uint64), ("device_ms", C.c_float),
        ("h2d_ms", C.c_float), ("total_ms", C.c_float), ("e2e_ms", C.c_float), ("kernel_launches", C.c_int),
        ("grid_ctas", C.c_int),
        ("returned_prior", C.c_int), ("xicp_borderline", C.c_int), ("loc_xyz", C.c_int * 3), ("loc_rpy", C.c_int * 3),
        ("eig_t", C.c_float * 9), ("eig_r", C.c_float * 9), ("xicp_sums", C.c_double * 12),
    ]


class O3dParams(C.Structure):
    _fields_ = [("kind", C.c_int), ("max_correspondence_distance", C.c_float), ("max_iteration", C.c_int),
                ("relative_fitness", C.c_double), ("relative_rmse", C.c_double), ("gicp_epsilon", C.c_double),
                ("cell_size", C.c_float)]


class O3dResult(C.Structure):
    _fields_ = [("T", C.c_double * 16), ("fitness", C.c_double), ("inlier_rmse", C.c_double), ("iterations", C.c_int),
                ("n_correspondences", C.c_int), ("kernel_launches", C.c_int), ("device_ms", C.c_float),
                ("total_ms", C.c_float)]


class RansacParams(C.Structure):
    _fields_ = [("mutual_filter", C.c_int), ("max_correspondence_distance", C.c_double), ("ransac_n", C.c_int),
                ("edge_length_similarity", C.c_double), ("distance_threshold", C.c_double), ("max_iteration", C.c_int),
                ("confidence", C.c_double), ("seed", C.c_uint64), ("batch", C.c_int)]


class RansacResult(C.Structure):
    _fields_ = [("T", C.c_double * 16), ("fitness", C.c_double), ("inlier_rmse", C.c_double),
                ("n_correspondences", C.c_int), ("n_feature_correspondences", C.c_int), ("mutual_used", C.c_int),
                ("iterations_run", C.c_int64), ("validations", C.c_int64), ("device_ms", C.c_float),
                ("kernel_launches", C.c_int)]


def lib_path():
    return _LIBPATH


def load():
    """Load the CUDA library; build it in-tree if sources are newer. Raises if it cannot be had."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("B200REG_LIB") or _LIBPATH   # (experiments: a differently compiled build of the same sources)
    if not os.path.exists(path):
        from . import build as _b
        _b.build()
    lib = C.CDLL(path)
    lib.b200_last_error.restype = C.c_char_p
    _lib = lib
    return lib


def device_count():
    return load().b200_device_count()


def require_device():
    n = device_count()
    if n < 1:
        raise B200Error(B200_E_CUDA, "no CUDA device visible; the B200 registration core has no CPU fallback")
    return n


def _check(st):
    if st == B200_OK:
        return
    msg = load().b200_last_error().decode("utf-8", "replace")
    if st in (B200_E_NO_MATCHES, B200_E_NAN, B200_E_BOUND):
        raise ConvergenceError(st, msg)
    raise B200Error(st, msg)


def _fp(a):
    return None if a is None else a.ctypes.data_as(_f32p)


def _ip(a):
    return None if a is None else a.ctypes.data_as(_i32p)


def f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def feat4(xyz):
    xyz = np.asarray(xyz, dtype=np.float32)
    out = np.ones((xyz.shape[0], 4), dtype=np.float32)
    out[:, :3] = xyz[:, :3]
    return out


def _colmajor(T):
    return np.ascontiguousarray(np.asarray(T, np.float32).T).ravel()


def _mat(a):
    return np.array(a, np.float32).reshape(4, 4).T.copy()


class Index:
    """KDTreeMatcher replacement: init() at construction, find_closests() == findClosests()."""

    def __init__(self, ref4, normals=None, device=0, cell_size=0.0, target_per_cell=0.0, max_cells=0):
        require_device()
        self.ref4 = f32(ref4)
        self.normals = f32(normals)
        self.h = C.c_void_p()
        opts = IndexOpts(cell_size, target_per_cell, max_cells)
        _check(load().b200_index_create(C.byref(self.h), device, _fp(self.ref4), _fp(self.normals),
                                        self.ref4.shape[0], C.byref(opts)))
        self.visit_counter = 0

    def find_closests(self, q4, knn=1, epsilon=0.0, max_dist=np.inf, max_dist_per_point=None):
        q4 = f32(q4)
        n = q4.shape[0]
        ids = np.empty((n, knn), np.int32)
        d = np.empty((n, knn), np.float32)
        mdpp = f32(max_dist_per_point)
        vis = C.c_uint64()
        _check(load().b200_index_knn(self.h, _fp(q4), n, knn, C.c_float(epsilon), C.c_float(max_dist), _fp(mdpp),
                                     _ip(ids), _fp(d), C.byref(vis)))
        self.visit_counter += vis.value
        return ids, d

    def update(self, ref4, normals=None, cell_size=0.0, target_per_cell=0.0, max_cells=0):
        """re-initialise with a new cloud, reusing the device buffers"""
        self.ref4, self.normals = f32(ref4), f32(normals)
        opts = IndexOpts(cell_size, target_per_cell, max_cells)
        _check(load().b200_index_update(self.h, _fp(self.ref4), _fp(self.normals), self.ref4.shape[0], C.byref(opts)))

    def self_normals(self, knn=5, max_dist=np.inf, sort_eigen=0, keep=("normals",), smooth=0):
        """SurfaceNormalDataPointsFilter over the indexed cloud itself"""
        n = self.ref4.shape[0]
        shapes = dict(normals=(n, 3), densities=(n,), eigValues=(n, 3), eigVectors=(n, 9), matchedIds=(n, knn),
                      meanDists=(n,))
        out = {k_: (np.empty(shapes[k_], np.float32) if k_ in keep else None) for k_ in shapes}
        prm = SurfaceNormalParams(knn, max_dist, 0.0, sort_eigen, smooth)
        _check(load().b200_index_self_normals(self.h, C.byref(prm), _fp(out["normals"]), _fp(out["densities"]),
                                              _fp(out["eigValues"]), _fp(out["eigVectors"]), _fp(out["matchedIds"]),
                                              _fp(out["meanDists"])))
        return {k_: v for k_, v in out.items() if v is not None}

    def last_normals_fallback(self):
        n = C.c_int()
        _check(load().b200_index_last_normals_fallback(self.h, C.byref(n)))
        return n.value

    def last_device_ms(self):
        ms = C.c_float()
        _check(load().b200_index_last_device_ms(self.h, C.byref(ms)))
        return ms.value

    def info(self):
        cell = C.c_float()
        dims = (C.c_int32 * 3)()
        occ = C.c_int64()
        ms = C.c_float()
        _check(load().b200_index_info(self.h, C.byref(cell), dims, C.byref(occ), C.byref(ms)))
        return dict(cell_size=cell.value, dims=tuple(dims), occupied_cells=occ.value, build_ms=ms.value)

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            load().b200_index_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def surface_normals(feat4_, knn=5, max_dist=np.inf, sort_eigen=0, smooth=0, device=0, keep=("normals",)):
    require_device()
    f = f32(feat4_)
    n = f.shape[0]
    shapes = dict(normals=(n, 3), densities=(n,), eigValues=(n, 3), eigVectors=(n, 9), matchedIds=(n, knn),
                  meanDists=(n,))
    out = {k_: (np.empty(shapes[k_], np.float32) if k_ in keep else None) for k_ in shapes}
    prm = SurfaceNormalParams(knn, max_dist, 0.0, sort_eigen, smooth)
    _check(load().b200_surface_normals(device, _fp(f), n, C.byref(prm), _fp(out["normals"]), _fp(out["densities"]),
                                       _fp(out["eigValues"]), _fp(out["eigVectors"]), _fp(out["matchedIds"]),
                                       _fp(out["meanDists"])))
    return {k_: v for k_, v in out.items() if v is not None}


def dists_quantile(dists, q, device=0):
    require_device()
    d = f32(dists).ravel()
    out = C.c_float()
    _check(load().b200_dists_quantile(device, _fp(d), C.c_int64(d.size), C.c_float(q), C.byref(out)))
    return out.value


def outlier_dist(kind, param, dists, device=0):
    require_device()
    d = f32(dists)
    w = np.empty_like(d)
    _check(load().b200_outlier_dist(device, kind, C.c_float(param), _fp(d), C.c_int64(d.size), _fp(w)))
    return w


NOALLOC_MAX_N = 1 << 24


def outlier_trimmed_noalloc(d_dists_ptr, n, ratio, d_weights_ptr, d_status_limit_ptr=0, stream=0, device=0):
    """b200_outlier_trimmed_noalloc: DEVICE pointers (integers), only enqueues on `stream` (0: legacy default stream);
    allocates nothing. Returns nothing; the status / limit land in the caller's device uint32[2]."""
    _check(load().b200_outlier_trimmed_noalloc(device, C.c_void_p(d_dists_ptr), C.c_int64(n), C.c_float(ratio),
                                               C.c_void_p(d_weights_ptr), C.c_void_p(d_status_limit_ptr or None),
                                               C.c_void_p(stream or None)))


def outlier_surface_normal(nread, nref, ids, max_angle, device=0):
    require_device()
    nread, nref = f32(nread), f32(nref)
    ids = np.ascontiguousarray(ids, np.int32)
    n, k = ids.shape
    w = np.empty((n, k), np.float32)
    m = 0 if nref is None else nref.shape[0]
    _check(load().b200_outlier_surface_normal(device, _fp(nread), _fp(nref), m, _ip(ids), k, n, C.c_float(max_angle),
                                              _fp(w)))
    return w


def error_elements(read4, ref4, refn, w, ids, dists, device=0):
    require_device()
    read4, ref4, refn, w, dists = f32(read4), f32(ref4), f32(refn), f32(w), f32(dists)
    ids = np.ascontiguousarray(ids, np.int32)
    n, k = ids.shape
    cap = n * k
    readP = np.empty((cap, 4), np.float32)
    refP = np.empty((cap, 4), np.float32)
    refnP = np.empty((cap, 3), np.float32) if refn is not None else None
    wP = np.empty(cap, np.float32)
    idsP = np.empty(cap, np.int32)
    dP = np.empty(cap, np.float32)
    P = C.c_int()
    pur, wpur = C.c_float(), C.c_float()
    _check(load().b200_error_elements(device, _fp(read4), _fp(ref4), _fp(refn), ref4.shape[0], _fp(w), _ip(ids),
                                      _fp(dists), k, n, _fp(readP), _fp(refP), _fp(refnP), _fp(wP), _ip(idsP), _fp(dP),
                                      C.byref(P), C.byref(pur), C.byref(wpur)))
    p = P.value
    return dict(reading=readP[:p], reference=refP[:p], ref_normals=None if refnP is None else refnP[:p],
                weights=wP[:p], ids=idsP[:p], dists=dP[:p], point_used_ratio=pur.value,
                weighted_point_used_ratio=wpur.value)


def p2plane_hessian(readP, refP, refnP, wP, device=0):
    require_device()
    readP, refP, refnP, wP = f32(readP), f32(refP), f32(refnP), f32(wP)
    H = np.empty(36, np.float32)
    b = np.empty(6, np.float32)
    _check(load().b200_p2plane_hessian(device, _fp(readP), _fp(refP), _fp(refnP), _fp(wP), readP.shape[0], _fp(H),
                                       _fp(b)))
    return H.reshape(6, 6).T.copy(), b


def p2plane_solve(H, b, device=0):
    require_device()
    A = np.ascontiguousarray(np.asarray(H, np.float32).T).ravel()
    b = f32(b)
    x = np.empty(6, np.float32)
    T = np.empty(16, np.float32)
    info = C.c_int()
    _check(load().b200_p2plane_solve(device, _fp(A), _fp(b), _fp(x), _fp(T), C.byref(info)))
    return x, T.reshape(4, 4).T.copy(), info.value


def p2point(readP, refP, wP, device=0):
    require_device()
    readP, refP, wP = f32(readP), f32(refP), f32(wP)
    T = np.empty(16, np.float32)
    _check(load().b200_p2point(device, _fp(readP), _fp(refP), _fp(wP), readP.shape[0], _fp(T)))
    return T.reshape(4, 4).T.copy()


def rigid_apply(T, feat4_, normals=None, device=0):
    require_device()
    f = f32(feat4_).copy()
    n = None if normals is None else f32(normals).copy()
    Tc = _colmajor(T)
    _check(load().b200_rigid_apply(device, _fp(Tc), _fp(f), _fp(n), f.shape[0]))
    return f, n


def default_icp_params(**kw):
    p = IcpParams()
    load().b200_icp_default_params(C.byref(p))
    for k_, v in kw.items():
        assert hasattr(p, k_), k_
        setattr(p, k_, v)
    return p


class ICP:
    """PointMatcher<float>::ICP replacement for the supported chain (see include/b200reg.h)."""

    def __init__(self, params=None, device=0):
        require_device()
        self.params = params if params is not None else default_icp_params()
        self.h = C.c_void_p()
        _check(load().b200_icp_create(C.byref(self.h), device, C.byref(self.params)))
        self._n = 0

    def init_reference(self, ref4, normals=None):
        self._ref4, self._refn = f32(ref4), f32(normals)
        _check(load().b200_icp_init_reference(self.h, _fp(self._ref4), _fp(self._refn), self._ref4.shape[0]))

    def _result(self, res, st, raise_on_error):
        out = dict(status=st, iterations=res.iterations, max_iterations_reached=bool(res.max_iterations_reached),
                   n_matches_last=res.n_matches_last, T=_mat(res.T), T_iter=_mat(res.T_iter),
                   T_refIn_refMean=_mat(res.T_refIn_refMean), T_refMean_readMean=_mat(res.T_refMean_readMean),
                   weighted_point_used_ratio=res.weighted_point_used_ratio, point_used_ratio=res.point_used_ratio,
                   points_touched=res.points_touched, device_ms=res.device_ms, h2d_ms=res.h2d_ms,
                   total_ms=res.total_ms, e2e_ms=res.e2e_ms, kernel_launches=res.kernel_launches,
                   grid_ctas=res.grid_ctas, returned_prior=bool(res.returned_prior),
                   xicp_borderline=int(res.xicp_borderline), loc_xyz=np.array(res.loc_xyz[:], np.int32),
                   loc_rpy=np.array(res.loc_rpy[:], np.int32),
                   eig_t=np.array(res.eig_t[:], np.float32).reshape(3, 3).T.copy(),
                   eig_r=np.array(res.eig_r[:], np.float32).reshape(3, 3).T.copy(),
                   xicp_sums=np.array(res.xicp_sums[:], np.float64))
        if raise_on_error:
            _check(st)
        return out

    def compute(self, read4, normals=None, T_init=None, raise_on_error=True):
        read4, normals = f32(read4), f32(normals)
        self._n = read4.shape[0]
        Tc = None if T_init is None else _colmajor(T_init)
        res = IcpResult()
        st = load().b200_icp_compute(self.h, _fp(read4), _fp(normals), self._n, _fp(Tc), C.byref(res))
        return self._result(res, st, raise_on_error)

    def init_reference_o3d(self, pts3, normals3=None):
        """Open3D double clouds (N, 3) float64 straight in: open3dToPointmatcher's cast happens on the device"""
        dp = C.POINTER(C.c_double)
        self._p3 = np.ascontiguousarray(pts3, np.float64)
        self._n3 = None if normals3 is None else np.ascontiguousarray(normals3, np.float64)
        _check(load().b200_icp_init_reference_o3d(self.h, self._p3.ctypes.data_as(dp),
                                                  None if self._n3 is None else self._n3.ctypes.data_as(dp), self._p3.shape[0]))

    def compute_o3d(self, pts3, normals3=None, T_init=None, raise_on_error=True):
        dp = C.POINTER(C.c_double)
        p = np.ascontiguousarray(pts3, np.float64)
        n = None if normals3 is None else np.ascontiguousarray(normals3, np.float64)
        self._n = p.shape[0]
        Tc = None if T_init is None else _colmajor(T_init)
        res = IcpResult()
        st = load().b200_icp_compute_o3d(self.h, p.ctypes.data_as(dp), None if n is None else n.ctypes.data_as(dp), self._n,
                                         _fp(Tc), C.byref(res))
        return self._result(res, st, raise_on_error)

    # ---- partitioned map (b200_icp_slab_*): this handle holds one padded slab; d_* are device pointers (int)
    def init_reference_slab(self, slab4, normals, gid, map_mean, axis, lo, hi):
        self._ref4, self._refn = f32(slab4), f32(normals)
        self._gid = None if gid is None else np.ascontiguousarray(gid, np.int32)
        mean = np.ascontiguousarray(map_mean, np.float64)
        _check(load().b200_icp_init_reference_slab(self.h, _fp(self._ref4), _fp(self._refn), self._ref4.shape[0],
                                                   _ip(self._gid), mean.ctypes.data_as(C.POINTER(C.c_double)), int(axis),
                                                   C.c_float(lo), C.c_float(hi)))

    def slab_begin(self, read4, normals=None, T_init=None):
        read4, normals = f32(read4), f32(normals)
        self._n = read4.shape[0]
        Tc = None if T_init is None else _colmajor(T_init)
        _check(load().b200_icp_slab_begin(self.h, _fp(read4), _fp(normals), self._n, _fp(Tc)))

    def slab_match(self, d_match, d_ids):
        _check(load().b200_icp_slab_match(self.h, C.c_void_p(int(d_match)), C.c_void_p(int(d_ids))))

    def slab_match_async(self, d_match, d_ids):
        _check(load().b200_icp_slab_match_async(self.h, C.c_void_p(int(d_match)), C.c_void_p(int(d_ids))))

    def slab_stream(self):
        f = load().b200_icp_slab_stream
        f.restype = C.c_void_p
        return f(self.h) or 0

    def slab_iterate(self, d_match, d_ids):
        it = C.c_int()
        _check(load().b200_icp_slab_iterate(self.h, C.c_void_p(int(d_match)), C.c_void_p(int(d_ids)), C.byref(it)))
        return bool(it.value)

    def slab_finish(self, raise_on_error=True):
        res = IcpResult()
        st = load().b200_icp_slab_finish(self.h, C.byref(res))
        return self._result(res, st, raise_on_error)

    def upload_reading(self, read4, normals=None):
        read4, normals = f32(read4), f32(normals)
        self._n = read4.shape[0]
        _check(load().b200_icp_upload_reading(self.h, _fp(read4), _fp(normals), self._n))

    def run_resident(self, T_init=None, raise_on_error=True):
        Tc = None if T_init is None else _colmajor(T_init)
        res = IcpResult()
        st = load().b200_icp_run_resident(self.h, _fp(Tc), C.byref(res))
        return self._result(res, st, raise_on_error)

    def matches(self):
        k = self.params.knn
        ids = np.empty((self._n, k), np.int32)
        d = np.empty((self._n, k), np.float32)
        w = np.empty((self._n, k), np.float32)
        _check(load().b200_icp_get_matches(self.h, _ip(ids), _fp(d), _fp(w)))
        return ids, d, w

    def log(self, max_log=64):
        H = np.zeros((max_log, 36), np.float32)
        b = np.zeros((max_log, 6), np.float32)
        x = np.zeros((max_log, 6), np.float32)
        T = np.zeros((max_log, 16), np.float32)
        P = np.zeros(max_log, np.int32)
        n = C.c_int()
        _check(load().b200_icp_get_log(self.h, max_log, _fp(H), _fp(b), _fp(x), _fp(T), _ip(P), C.byref(n)))
        it = n.value
        return dict(H=H[:it].reshape(it, 6, 6).transpose(0, 2, 1).copy(), b=b[:it], x=x[:it],
                    T_iters=T[:it].reshape(it, 4, 4).transpose(0, 2, 1).copy(), P=P[:it])

    def phase_times(self, max_log=64):
        """(iterations, 6) microseconds: match, trimmed select, linearise, wait, reduce+solve, whole iteration"""
        us = np.zeros((max_log, 6), np.float32)
        n = C.c_int()
        _check(load().b200_icp_get_phase_times(self.h, max_log, _fp(us), C.byref(n)))
        return us[:n.value]

    def debug_stats(self, max_log=64):
        """(iterations, 4): rows, candidates, slowest CTA match us, mean CTA match us (needs B200_ICP_DEBUG=1)"""
        st = np.zeros((max_log, 12), np.float64)
        n = C.c_int()
        _check(load().b200_icp_get_debug_stats(self.h, max_log, st.ctypes.data_as(C.POINTER(C.c_double)), C.byref(n)))
        return st[:n.value]

    def debug_warps(self):
        slots = (self._n + 31) // 32
        out = np.zeros((slots, 4), np.uint32)
        n = C.c_int()
        _check(load().b200_icp_get_debug_warps(self.h, out.ctypes.data_as(C.POINTER(C.c_uint32)), slots, C.byref(n)))
        return out[:n.value]

    def flush_l2(self):
        _check(load().b200_icp_flush_l2(self.h))

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            load().b200_icp_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


O3D_POINT_TO_PLANE, O3D_POINT_TO_POINT, O3D_GENERALIZED = 0, 1, 2


class O3dICP:
    """Boundary B: Open3D-semantics RegistrationICP / RegistrationGeneralizedICP on the device (b200_o3d_*)."""

    def __init__(self, kind=O3D_POINT_TO_PLANE, max_correspondence_distance=1.0, max_iteration=30,
                 relative_fitness=1e-6, relative_rmse=1e-6, gicp_epsilon=1e-3, cell_size=0.0, device=0):
        require_device()
        p = O3dParams()
        load().b200_o3d_default_params(C.byref(p))
        p.kind, p.max_correspondence_distance, p.max_iteration = kind, max_correspondence_distance, max_iteration
        p.relative_fitness, p.relative_rmse, p.gicp_epsilon, p.cell_size = relative_fitness, relative_rmse, gicp_epsilon, cell_size
        self.params = p
        self._h = C.c_void_p()
        self._n = 0
        _check(load().b200_o3d_create(C.byref(self._h), device, C.byref(p)))

    def set_target(self, tgt4, normals):
        tgt4, normals = f32(tgt4), f32(normals)
        _check(load().b200_o3d_set_target(self._h, _fp(tgt4), _fp(normals), tgt4.shape[0]))

    def set_target_estimate_normals(self, tgt4, knn=10, max_dist=np.inf):
        tgt4 = f32(tgt4)
        _check(load().b200_o3d_set_target_estimate_normals(self._h, _fp(tgt4), tgt4.shape[0], int(knn), C.c_float(max_dist)))

    def estimate_normals(self, pts4, knn=20, radius=1.0):
        """estimateNormalsOrCovariancesIfNeeded for any cloud (Open3D EstimateNormals + normalise + orient): (N, 3) float64"""
        pts4 = f32(pts4)
        n = pts4.shape[0]
        out = np.empty((n, 3), np.float64)
        _check(load().b200_o3d_estimate_normals(self._h, _fp(pts4), n, int(knn), C.c_float(radius),
                                                out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def register(self, src4, normals=None, init=None):
        src4, normals = f32(src4), f32(normals)
        Ti = np.eye(4) if init is None else np.asarray(init, np.float64)
        Tc = np.ascontiguousarray(Ti.T).ravel()
        r = O3dResult()
        self._n = src4.shape[0]
        _check(load().b200_o3d_register(self._h, _fp(src4), _fp(normals), self._n,
                                        Tc.ctypes.data_as(C.POINTER(C.c_double)), C.byref(r)))
        return dict(T=np.array(r.T, np.float64).reshape(4, 4).T.copy(), fitness=r.fitness, inlier_rmse=r.inlier_rmse,
                    iterations=r.iterations, n_correspondences=r.n_correspondences, device_ms=r.device_ms,
                    total_ms=r.total_ms, kernel_launches=r.kernel_launches)

    def information_matrix(self, src4, T, max_correspondence_distance):
        """GetInformationMatrixFromPointClouds against the handle's target -> (6x6 float64, fitness, rmse, n_corr)"""
        src4 = f32(src4)
        Tc = np.ascontiguousarray(np.asarray(T, np.float64).T).ravel()
        info = np.empty(36, np.float64)
        fit, rmse, n = C.c_double(), C.c_double(), C.c_int()
        dp = C.POINTER(C.c_double)
        _check(load().b200_o3d_information_matrix(self._h, _fp(src4), src4.shape[0], Tc.ctypes.data_as(dp),
                                                  C.c_float(max_correspondence_distance), info.ctypes.data_as(dp),
                                                  C.byref(fit), C.byref(rmse), C.byref(n)))
        self._n = src4.shape[0]
        return info.reshape(6, 6).T.copy(), fit.value, rmse.value, n.value

    def register_estimate_normals(self, src4, knn=20, radius=1.0, init=None):
        """registerClouds for a source without normals: estimated on the device (Open3D semantics), no host round trip"""
        src4 = f32(src4)
        Ti = np.eye(4) if init is None else np.asarray(init, np.float64)
        Tc = np.ascontiguousarray(Ti.T).ravel()
        r = O3dResult()
        self._n = src4.shape[0]
        _check(load().b200_o3d_register_estimate_normals(self._h, _fp(src4), self._n, int(knn), C.c_float(radius),
                                                         Tc.ctypes.data_as(C.POINTER(C.c_double)), C.byref(r)))
        return dict(T=np.array(r.T, np.float64).reshape(4, 4).T.copy(), fitness=r.fitness, inlier_rmse=r.inlier_rmse,
                    iterations=r.iterations, n_correspondences=r.n_correspondences, device_ms=r.device_ms,
                    total_ms=r.total_ms, kernel_launches=r.kernel_launches)

    # ---- partitioned map: stepwise run, the caller sums the 32 doubles over the slabs between the two steps
    def set_partition(self, axis, lo, hi, gid=None):
        g = None if gid is None else np.ascontiguousarray(gid, np.int32)
        _check(load().b200_o3d_set_partition(self._h, int(axis), C.c_float(lo), C.c_float(hi), _ip(g)))

    def begin(self, src4, normals=None, init=None, stream=None):
        src4, normals = f32(src4), f32(normals)
        Ti = np.eye(4) if init is None else np.asarray(init, np.float64)
        Tc = np.ascontiguousarray(Ti.T).ravel()
        self._n = src4.shape[0]
        _check(load().b200_o3d_begin(self._h, _fp(src4), _fp(normals), self._n, Tc.ctypes.data_as(C.POINTER(C.c_double)),
                                     C.c_void_p(stream or 0)))

    # ---- peer exchange over NVLink / NVSwitch peer memory (include/b200reg.h: b200_o3d_peer_*)
    def peer_init(self, rank, world):
        """-> (device pointer of this rank's mailbox, its 64-byte CUDA IPC handle)"""
        ptr = C.c_void_p()
        handle = C.create_string_buffer(64)
        _check(load().b200_o3d_peer_init(self._h, int(rank), int(world), C.byref(ptr), handle))
        return ptr.value, handle.raw

    def peer_connect(self, pointers=None, devices=None, handles=None):
        """pointers (+ devices): mailboxes of handles living in this process; handles: IPC handles of other processes"""
        if pointers is not None:
            arr = (C.c_void_p * len(pointers))(*[C.c_void_p(p_) for p_ in pointers])
            dev = None if devices is None else (C.c_int * len(devices))(*[int(d) for d in devices])
            _check(load().b200_o3d_peer_connect(self._h, arr, dev, None))
        else:
            _check(load().b200_o3d_peer_connect(self._h, None, None, b"".join(handles)))

    def enqueue_peers(self, src4, normals=None, init=None, stream=None):
        """every iteration of one registration enqueued without host synchronisation; finish() fetches the result"""
        src4, normals = f32(src4), f32(normals)
        Ti = np.eye(4) if init is None else np.asarray(init, np.float64)
        Tc = np.ascontiguousarray(Ti.T).ravel()
        self._n = src4.shape[0]
        self._keep = (src4, normals, Tc)   # the uploads are asynchronous
        _check(load().b200_o3d_enqueue_peers(self._h, _fp(src4), _fp(normals), self._n,
                                             Tc.ctypes.data_as(C.POINTER(C.c_double)), C.c_void_p(stream or 0)))

    def step_linearize(self, sums_ptr):
        _check(load().b200_o3d_step_linearize(self._h, C.c_void_p(sums_ptr)))

    def step_update(self, sums_ptr):
        _check(load().b200_o3d_step_update(self._h, C.c_void_p(sums_ptr)))

    def poll(self):
        done, upd = C.c_int(), C.c_int()
        _check(load().b200_o3d_poll(self._h, C.byref(done), C.byref(upd)))
        return bool(done.value), upd.value

    def finish(self):
        r = O3dResult()
        _check(load().b200_o3d_finish(self._h, C.byref(r)))
        return dict(T=np.array(r.T, np.float64).reshape(4, 4).T.copy(), fitness=r.fitness, inlier_rmse=r.inlier_rmse,
                    iterations=r.iterations, n_correspondences=r.n_correspondences, device_ms=r.device_ms,
                    total_ms=r.total_ms, kernel_launches=r.kernel_launches)

    def correspondences(self):
        ids = np.empty(self._n, np.int32)
        d = np.empty(self._n, np.float32)
        _check(load().b200_o3d_get_correspondences(self._h, _ip(ids), _fp(d)))
        return ids, d

    def close(self):
        if self._h:
            load().b200_o3d_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- scan pre-processing (include/b200reg.h: b200_prep_*) -----------------------------------------------------------
CROP_NONE, CROP_CYLINDER, CROP_MIN_RADIUS, CROP_MAX_RADIUS, CROP_MINMAX_RADIUS = 0, 1, 2, 3, 4


class CropParams(C.Structure):
    _fields_ = [("kind", C.c_int), ("invert", C.c_int), ("radius_min", C.c_double), ("radius_max", C.c_double),
                ("min_z", C.c_double), ("max_z", C.c_double), ("center", C.c_double * 3)]


class LoopClosure:
    """Loop-closure front half on the device (b200_lc_*): ComputeFPFHFeature (Submap.cpp:272) and
    RegistrationRANSACBasedOnFeatureMatching (PlaceRecognition.cpp:81-85)."""

    def __init__(self, device=0):
        require_device()
        self._h = C.c_void_p()
        _check(load().b200_lc_create(C.byref(self._h), device))

    def compute_fpfh(self, pts4, normals, radius=2.5, max_nn=100):
        """-> (features (N, 33) float64, neighbourhood sizes (N,) int32)"""
        pts4 = f32(pts4)
        nrm = np.ascontiguousarray(normals, np.float64)
        n = pts4.shape[0]
        out = np.empty((n, 33), np.float64)
        cnt = np.empty(n, np.int32)
        dp = C.POINTER(C.c_double)
        _check(load().b200_lc_compute_fpfh(self._h, _fp(pts4), nrm.ctypes.data_as(dp), n, C.c_float(radius), int(max_nn),
                                           out.ctypes.data_as(dp), _ip(cnt)))
        return out, cnt

    def compute_features(self, pts4, normal_radius=1.0, normal_knn=10, fpfh_radius=2.5, fpfh_max_nn=100):
        """normals (Open3D semantics) + FPFH on one index -> (normals (N, 3) float64, features (N, 33) float64)"""
        pts4 = f32(pts4)
        n = pts4.shape[0]
        nrm, out = np.empty((n, 3), np.float64), np.empty((n, 33), np.float64)
        dp = C.POINTER(C.c_double)
        _check(load().b200_lc_compute_features(self._h, _fp(pts4), n, C.c_float(normal_radius), int(normal_knn),
                                               C.c_float(fpfh_radius), int(fpfh_max_nn), nrm.ctypes.data_as(dp),
                                               out.ctypes.data_as(dp)))
        return nrm, out

    def feature_nn(self, a, b):
        """exact nearest row of b (M, 33) for every row of a (N, 33) -> (ids int32, squared distances float64)"""
        a, b = np.ascontiguousarray(a, np.float64), np.ascontiguousarray(b, np.float64)
        ids = np.empty(a.shape[0], np.int32)
        d2 = np.empty(a.shape[0], np.float64)
        dp = C.POINTER(C.c_double)
        _check(load().b200_lc_feature_nn(self._h, a.ctypes.data_as(dp), a.shape[0], b.ctypes.data_as(dp), b.shape[0],
                                         _ip(ids), d2.ctypes.data_as(dp)))
        return ids, d2

    def ransac(self, src4, tgt4, sf, tf, **kw):
        """-> dict (T 4x4, fitness, inlier_rmse, correspondences (K, 2), ...); keywords = b200_ransac_params fields"""
        src4, tgt4 = f32(src4), f32(tgt4)
        sf, tf = np.ascontiguousarray(sf, np.float64), np.ascontiguousarray(tf, np.float64)
        prm = RansacParams()
        load().b200_lc_default_ransac_params(C.byref(prm))
        for k, v in kw.items():
            if not hasattr(prm, k):
                raise TypeError("unknown RANSAC parameter %r" % k)
            setattr(prm, k, int(v) if k in ("mutual_filter", "ransac_n", "max_iteration", "seed", "batch") else float(v))
        res = RansacResult()
        corr = np.empty((max(src4.shape[0], 1), 2), np.int32)
        dp = C.POINTER(C.c_double)
        _check(load().b200_lc_ransac_feature_matching(self._h, _fp(src4), src4.shape[0], _fp(tgt4), tgt4.shape[0],
                                                      sf.ctypes.data_as(dp), tf.ctypes.data_as(dp), C.byref(prm),
                                                      C.byref(res), _ip(corr)))
        return dict(T=np.array(res.T).reshape(4, 4).T.copy(), fitness=res.fitness, inlier_rmse=res.inlier_rmse,
                    correspondences=corr[:res.n_correspondences].copy(),
                    n_feature_correspondences=res.n_feature_correspondences, mutual_used=bool(res.mutual_used),
                    iterations_run=res.iterations_run, validations=res.validations, device_ms=res.device_ms,
                    kernel_launches=res.kernel_launches)

    def overlap_indices(self, src4, tgt4, T, voxel_size, min_points=1):
        """computeIndicesOfOverlappingPoints (helpers.cpp:320-345) -> (source indices, target indices) int32, ascending"""
        src4, tgt4 = f32(src4), f32(tgt4)
        Tc = np.ascontiguousarray(np.asarray(T, np.float64).T).ravel()
        si, ti = np.empty(max(len(src4), 1), np.int32), np.empty(max(len(tgt4), 1), np.int32)
        ns, nt = C.c_int(), C.c_int()
        _check(load().b200_lc_overlap_indices(self._h, _fp(src4), len(src4), _fp(tgt4), len(tgt4),
                                              Tc.ctypes.data_as(C.POINTER(C.c_double)), C.c_double(voxel_size), int(min_points),
                                              _ip(si), C.byref(ns), _ip(ti), C.byref(nt)))
        return si[:ns.value].copy(), ti[:nt.value].copy()

    def last_device_ms(self):
        ms = C.c_float()
        _check(load().b200_lc_last_device_ms(self._h, C.byref(ms)))
        return ms.value

    def close(self):
        if self._h:
            load().b200_lc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Prep:
    """open3d_slam's per-scan pre-processing on the device: croppers (croppers.cpp) and voxelize (helpers.cpp:108-115)."""

    def __init__(self, device=0):
        require_device()
        self._h = C.c_void_p()
        _check(load().b200_prep_create(C.byref(self._h), device))

    def crop(self, kind, pts4, normals=None, radius_min=0.0, radius_max=0.0, min_z=0.0, max_z=0.0,
             center=(0.0, 0.0, 0.0), invert=False):
        """-> (points K x 4 float32, normals K x 3 or None, kept indices K int32)"""
        pts4, normals = f32(pts4), f32(normals)
        n = pts4.shape[0]
        prm = CropParams(int(kind), int(bool(invert)), radius_min, radius_max, min_z, max_z, (C.c_double * 3)(*center))
        out = np.empty((max(n, 1), 4), np.float32)
        outn = np.empty((max(n, 1), 3), np.float32) if normals is not None else None
        idx = np.empty(max(n, 1), np.int32)
        k = C.c_int()
        _check(load().b200_prep_crop(self._h, C.byref(prm), _fp(pts4), _fp(normals), n, _fp(out), _fp(outn), _ip(idx),
                                     C.byref(k)))
        k = k.value
        return out[:k].copy(), (None if outn is None else outn[:k].copy()), idx[:k].copy()

    def random_downsample(self, ratio, pts4, normals=None, seed=0):
        """-> (points K x 4, normals K x 3 or None, kept indices K int32), K = int(ratio * N)"""
        pts4, normals = f32(pts4), f32(normals)
        n = pts4.shape[0]
        out = np.empty((max(n, 1), 4), np.float32)
        outn = np.empty((max(n, 1), 3), np.float32) if normals is not None else None
        idx = np.empty(max(n, 1), np.int32)
        k = C.c_int()
        _check(load().b200_prep_random_downsample(self._h, C.c_double(ratio), C.c_uint32(seed), _fp(pts4), _fp(normals), n,
                                                  _fp(out), _fp(outn), _ip(idx), C.byref(k)))
        k = k.value
        return out[:k].copy(), (None if outn is None else outn[:k].copy()), idx[:k].copy()

    def voxel_downsample(self, voxel_size, pts4, normals=None):
        """-> (points K x 3 float64, normals K x 3 float64 or None, index of the first member K int32)"""
        pts4, normals = f32(pts4), f32(normals)
        n = pts4.shape[0]
        out = np.empty((max(n, 1), 3), np.float64)
        outn = np.empty((max(n, 1), 3), np.float64) if normals is not None else None
        first = np.empty(max(n, 1), np.int32)
        k = C.c_int()
        dp = C.POINTER(C.c_double)
        _check(load().b200_prep_voxel_downsample(self._h, C.c_double(voxel_size), _fp(pts4), _fp(normals), n,
                                                 out.ctypes.data_as(dp), None if outn is None else outn.ctypes.data_as(dp),
                                                 _ip(first), C.byref(k)))
        k = k.value
        return out[:k].copy(), (None if outn is None else outn[:k].copy()), first[:k].copy()

    def last_ordered_voxels(self):
        n = C.c_int()
        _check(load().b200_prep_last_ordered_voxels(self._h, C.byref(n)))
        return n.value

    def last_device_ms(self):
        ms = C.c_float()
        _check(load().b200_prep_last_device_ms(self._h, C.byref(ms)))
        return ms.value

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            load().b200_prep_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def crop_params(kind=0, radius_min=0.0, radius_max=0.0, min_z=0.0, max_z=0.0, center=(0.0, 0.0, 0.0), invert=False):
    return CropParams(int(kind), int(bool(invert)), float(radius_min), float(radius_max), float(min_z), float(max_z),
                      (C.c_double * 3)(*[float(v) for v in center]))


class Map:
    """Device-resident map (b200_map_*): Submap::insertScan, cropSubmap and the hand-over of the patch to an ICP
    handle without the map crossing PCIe. Points / normals are (M, 3) float64 like Open3D's."""

    def __init__(self, device=0):
        require_device()
        self._h = C.c_void_p()
        _check(load().b200_map_create(C.byref(self._h), device))

    @staticmethod
    def _dp(a):
        return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))

    def set(self, pts3, normals3=None):
        p = np.ascontiguousarray(pts3, np.float64)
        n = None if normals3 is None else np.ascontiguousarray(normals3, np.float64)
        _check(load().b200_map_set(self._h, self._dp(p), self._dp(n), p.shape[0]))

    def size(self):
        m, hn = C.c_int(), C.c_int()
        _check(load().b200_map_size(self._h, C.byref(m), C.byref(hn)))
        return m.value, bool(hn.value)

    def download(self):
        m, hn = self.size()
        p = np.empty((m, 3), np.float64)
        n = np.empty((m, 3), np.float64) if hn else None
        _check(load().b200_map_download(self._h, self._dp(p), self._dp(n)))
        return p, n

    def insert_scan(self, scan4, scan_normals, T, voxel_size, crop):
        """-> (points passed through, new map size)"""
        scan4, scan_normals = f32(scan4), f32(scan_normals)
        Tc = np.ascontiguousarray(np.asarray(T, np.float64).T).ravel()
        no, ns = C.c_int(), C.c_int()
        _check(load().b200_map_insert_scan(self._h, _fp(scan4), _fp(scan_normals), scan4.shape[0], self._dp(Tc),
                                           C.c_double(voxel_size), C.byref(crop), C.byref(no), C.byref(ns)))
        return no.value, ns.value

    def crop(self, crop):
        m, hn = self.size()
        out = np.empty((max(m, 1), 4), np.float32)
        outn = np.empty((max(m, 1), 3), np.float32) if hn else None
        idx = np.empty(max(m, 1), np.int32)
        k = C.c_int()
        _check(load().b200_map_crop(self._h, C.byref(crop), _fp(out), _fp(outn), _ip(idx), C.byref(k)))
        k = k.value
        return out[:k].copy(), (None if outn is None else outn[:k].copy()), idx[:k].copy()

    def crop_to_icp(self, crop, icp):
        k = C.c_int()
        _check(load().b200_map_crop_to_icp(self._h, C.byref(crop), icp.h, C.byref(k)))
        return k.value

    def last_device_ms(self):
        ms = C.c_float()
        _check(load().b200_map_last_device_ms(self._h, C.byref(ms)))
        return ms.value

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            load().b200_map_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
