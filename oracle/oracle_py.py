"""ORACLE — TEST INFRASTRUCTURE ONLY. ctypes binding of oracle/liboracle.so (see oracle/oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
The product package (open3d_slam_private_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

ORC_OK, ORC_E_NO_MATCHES, ORC_E_BAD_PARAM, ORC_E_NAN = 0, 1, 2, 3
ORC_E_BOUND = 7

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)


class IcpParams(C.Structure):
    _fields_ = [
        ("knn", C.c_int), ("maxDist", C.c_float),
        ("use_trimmed", C.c_int), ("trimmed_ratio", C.c_float),
        ("use_maxdist_filter", C.c_int), ("maxdist_filter", C.c_float),
        ("use_surface_normal_filter", C.c_int), ("surface_normal_max_angle", C.c_float),
        ("minimizer", C.c_int), ("max_iterations", C.c_int),
        ("use_differential", C.c_int), ("min_diff_rot", C.c_float), ("min_diff_trans", C.c_float),
        ("smooth_length", C.c_int), ("counter_first", C.c_int), ("accum", C.c_int), ("mean_mode", C.c_int),
        ("knn_threads", C.c_int),
        ("degeneracy", C.c_int), ("xicp_enough", C.c_float), ("xicp_insufficient", C.c_float),
        ("xicp_cos_min", C.c_float), ("xicp_cos_strong", C.c_float),
        ("use_bound", C.c_int), ("max_rotation_norm", C.c_float), ("max_translation_norm", C.c_float),
    ]


class IcpResult(C.Structure):
    _fields_ = [
        ("iterations", C.c_int), ("max_iterations_reached", C.c_int), ("n_matches_last", C.c_int),
        ("T", C.c_float * 16), ("T_iter", C.c_float * 16), ("T_refIn_refMean", C.c_float * 16),
        ("T_refMean_readMean", C.c_float * 16), ("weighted_point_used_ratio", C.c_float),
        ("points_touched", C.c_long),
        ("returned_prior", C.c_int), ("loc_xyz", C.c_int * 3), ("loc_rpy", C.c_int * 3),
        ("eig_t", C.c_float * 9), ("eig_r", C.c_float * 9), ("xicp_sums", C.c_double * 12),
    ]


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("knn.cpp", "modules.cpp", "normals.cpp", "icp.cpp", "o3d.cpp", "prep.cpp", "fpfh.cpp", "oracle.h",
                                             "la_small.hpp")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_kdtree_build.restype = C.c_void_p
        _LIB.orc_kdtree_knn.restype = C.c_long
    return _LIB


def _fp(a):
    return None if a is None else a.ctypes.data_as(_f32p)


def _ip(a):
    return None if a is None else a.ctypes.data_as(_i32p)


def feat4(xyz):
    """(N,3) -> reference `features` memory image: N x {x,y,z,1} float32 (4 x N column-major)."""
    xyz = np.asarray(xyz, dtype=np.float32)
    out = np.ones((xyz.shape[0], 4), dtype=np.float32)
    out[:, :3] = xyz
    return np.ascontiguousarray(out)


def f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def knn_brute(ref4, q4, k, max_dist=np.inf):
    ref4, q4 = f32(ref4), f32(q4)
    n = q4.shape[0]
    ids = np.empty((n, k), np.int32)
    d = np.empty((n, k), np.float32)
    st = lib().orc_knn_brute(_fp(ref4), ref4.shape[0], _fp(q4), n, k, C.c_float(max_dist), _ip(ids), _fp(d))
    assert st == 0
    return ids, d


class KdTree:
    def __init__(self, ref4):
        self.ref4 = f32(ref4)
        self.h = C.c_void_p(lib().orc_kdtree_build(_fp(self.ref4), self.ref4.shape[0]))

    def knn(self, q4, k, max_dist=np.inf, threads=1):
        q4 = f32(q4)
        n = q4.shape[0]
        ids = np.empty((n, k), np.int32)
        d = np.empty((n, k), np.float32)
        self.touched = lib().orc_kdtree_knn(self.h, _fp(q4), n, k, C.c_float(max_dist), threads, _ip(ids), _fp(d))
        return ids, d

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_kdtree_free(self.h)
            self.h = None


def dists_quantile(dists, q):
    dists = f32(dists).ravel()
    out = C.c_float()
    st = lib().orc_dists_quantile(_fp(dists), C.c_long(dists.size), C.c_float(q), C.byref(out))
    return st, out.value


def _outlier(fn, dists, param):
    dists = f32(dists)
    w = np.empty_like(dists)
    st = fn(_fp(dists), C.c_long(dists.size), C.c_float(param), _fp(w))
    return st, w


def outlier_trimmed(dists, ratio):
    return _outlier(lib().orc_outlier_trimmed, dists, ratio)


def outlier_maxdist(dists, max_dist):
    return _outlier(lib().orc_outlier_maxdist, dists, max_dist)


def outlier_mindist(dists, min_dist):
    return _outlier(lib().orc_outlier_mindist, dists, min_dist)


def outlier_median(dists, factor):
    return _outlier(lib().orc_outlier_median, dists, factor)


def outlier_null(dists):
    dists = f32(dists)
    w = np.empty_like(dists)
    lib().orc_outlier_null(_fp(dists), C.c_long(dists.size), _fp(w))
    return 0, w


def outlier_surface_normal(nread, nref, ids, max_angle):
    nread, nref = f32(nread), f32(nref)
    ids = np.ascontiguousarray(ids, np.int32)
    n, k = ids.shape
    w = np.empty((n, k), np.float32)
    st = lib().orc_outlier_surface_normal(_fp(nread), _fp(nref), _ip(ids), k, n, C.c_float(max_angle), _fp(w))
    return st, w


def error_elements(read4, ref4, refn, w, ids, dists):
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
    st = lib().orc_error_elements(_fp(read4), _fp(ref4), _fp(refn), _fp(w), _ip(ids), _fp(dists), k, n, _fp(readP),
                                  _fp(refP), _fp(refnP), _fp(wP), _ip(idsP), _fp(dP), C.byref(P), C.byref(pur),
                                  C.byref(wpur))
    p = P.value
    return st, dict(reading=readP[:p], reference=refP[:p], ref_normals=None if refnP is None else refnP[:p],
                    weights=wP[:p], ids=idsP[:p], dists=dP[:p], point_used_ratio=pur.value,
                    weighted_point_used_ratio=wpur.value)


def p2plane_hb(readP, refP, refnP, wP, accum=1):
    readP, refP, refnP, wP = f32(readP), f32(refP), f32(refnP), f32(wP)
    H = np.empty(36, np.float32)
    b = np.empty(6, np.float32)
    lib().orc_p2plane_hb(_fp(readP), _fp(refP), _fp(refnP), _fp(wP), readP.shape[0], accum, _fp(H), _fp(b))
    return H.reshape(6, 6).T.copy(), b  # H returned as a normal (row, col) numpy matrix


def solve6(H, b):
    A = np.ascontiguousarray(np.asarray(H, np.float32).T).ravel()  # column-major
    b = f32(b)
    x = np.empty(6, np.float32)
    info = C.c_int()
    lib().orc_solve6(_fp(A), _fp(b), _fp(x), C.byref(info))
    return x, info.value


def solve_n(A, b):
    """orc_solve_n: the (6 + c) x (6 + c) equality-constrained system of PointToPlane.cpp:461-505."""
    A = np.asarray(A, np.float32)
    n = A.shape[0]
    Ac = np.ascontiguousarray(A.T).ravel()
    b = f32(b)
    x = np.empty(n, np.float32)
    info = C.c_int()
    st = lib().orc_solve_n(_fp(Ac), _fp(b), n, _fp(x), C.byref(info))
    assert st == 0
    return x, info.value


def p2plane_T(x):
    x = f32(x)
    T = np.empty(16, np.float32)
    lib().orc_p2plane_T(_fp(x), _fp(T))
    return T.reshape(4, 4).T.copy()


def p2point_T(readP, refP, wP, accum=0):
    readP, refP, wP = f32(readP), f32(refP), f32(wP)
    T = np.empty(16, np.float32)
    lib().orc_p2point_T2(_fp(readP), _fp(refP), _fp(wP), readP.shape[0], accum, _fp(T))
    return T.reshape(4, 4).T.copy()


def rigid_apply(T, feat4_, normals=None):
    Tc = np.ascontiguousarray(np.asarray(T, np.float32).T).ravel()
    f = f32(feat4_).copy()
    n = None if normals is None else f32(normals).copy()
    lib().orc_rigid_apply(_fp(Tc), _fp(f), _fp(n), f.shape[0])
    return f, n


def surface_normals(feat4_, knn, max_dist=np.inf, sort_eigen=0, smooth=0):
    f = f32(feat4_)
    n = f.shape[0]
    out = dict(normals=np.empty((n, 3), np.float32), densities=np.empty(n, np.float32),
               eigValues=np.empty((n, 3), np.float32), eigVectors=np.empty((n, 9), np.float32),
               matchedIds=np.empty((n, knn), np.float32), meanDists=np.empty(n, np.float32),
               ids=np.empty((n, knn), np.int32), dists=np.empty((n, knn), np.float32))
    st = lib().orc_surface_normals(_fp(f), n, knn, C.c_float(max_dist), sort_eigen, smooth, _fp(out["normals"]),
                                   _fp(out["densities"]), _fp(out["eigValues"]), _fp(out["eigVectors"]),
                                   _fp(out["matchedIds"]), _fp(out["meanDists"]), _ip(out["ids"]), _fp(out["dists"]))
    return st, out


def sampling_surface_normal(feat4_, knn=7, ratio=0.5, sampling_method=0, max_box_dim=np.inf):
    f = f32(feat4_)
    n = f.shape[0]
    fo = np.empty((n, 4), np.float32)
    no = np.empty((n, 3), np.float32)
    m = lib().orc_sampling_surface_normal(_fp(f), n, knn, C.c_float(ratio), sampling_method, C.c_float(max_box_dim),
                                          _fp(fo), _fp(no))
    return fo[:m].copy(), no[:m].copy()


def default_params(**kw):
    """PointMatcher ICPChainBase::setDefault (ICP.cpp:100-113) + module defaults."""
    p = IcpParams(knn=1, maxDist=np.inf, use_trimmed=1, trimmed_ratio=0.85, use_maxdist_filter=0, maxdist_filter=1.0,
                  use_surface_normal_filter=0, surface_normal_max_angle=1.57, minimizer=0, max_iterations=40,
                  use_differential=1, min_diff_rot=0.001, min_diff_trans=0.001, smooth_length=3, counter_first=1,
                  accum=1, mean_mode=1, knn_threads=1, degeneracy=0, xicp_enough=250.0, xicp_insufficient=180.0,
                  xicp_cos_min=float(np.cos(np.float32(np.deg2rad(np.float32(80.0))))),
                  xicp_cos_strong=float(np.cos(np.float32(np.deg2rad(np.float32(45.0))))),
                  use_bound=0, max_rotation_norm=1.0, max_translation_norm=1.0)
    for k_, v in kw.items():
        assert hasattr(p, k_), k_
        setattr(p, k_, v)
    return p


def _icp_run(call, params, N, T_init, max_log, forced_T_iters):
    """shared driver of orc_icp_forced / orc_icp_compute: allocates the logs, unpacks the result"""
    k = params.knn
    Ti = np.eye(4, dtype=np.float32) if T_init is None else np.asarray(T_init, np.float32)
    Tc = np.ascontiguousarray(Ti.T).ravel()
    res = IcpResult()
    logH = np.zeros((max_log, 36), np.float32)
    logb = np.zeros((max_log, 6), np.float32)
    logx = np.zeros((max_log, 6), np.float32)
    logT = np.zeros((max_log, 16), np.float32)
    logP = np.zeros(max_log, np.int32)
    ids = np.empty((N, k), np.int32)
    dists = np.empty((N, k), np.float32)
    w = np.empty((N, k), np.float32)
    if forced_T_iters is None:
        fT, nf = None, 0
    else:
        fT = np.ascontiguousarray(np.asarray(forced_T_iters, np.float32).transpose(0, 2, 1)).reshape(-1)
        nf = len(forced_T_iters)
    st = call(_fp(Tc), C.byref(res), max_log, _fp(logH), _fp(logb), _fp(logx), _fp(logT), _ip(logP), _ip(ids),
              _fp(dists), _fp(w), _fp(fT), nf)
    it = min(res.iterations, max_log)
    m = lambda a: np.array(a, np.float32).reshape(4, 4).T.copy()
    return dict(status=st, iterations=res.iterations, max_iterations_reached=bool(res.max_iterations_reached),
                n_matches_last=res.n_matches_last, T=m(res.T), T_iter=m(res.T_iter),
                T_refIn_refMean=m(res.T_refIn_refMean), T_refMean_readMean=m(res.T_refMean_readMean),
                weighted_point_used_ratio=res.weighted_point_used_ratio, points_touched=res.points_touched,
                returned_prior=bool(res.returned_prior), loc_xyz=np.array(res.loc_xyz[:], np.int32),
                loc_rpy=np.array(res.loc_rpy[:], np.int32),
                eig_t=np.array(res.eig_t[:], np.float32).reshape(3, 3).T.copy(),
                eig_r=np.array(res.eig_r[:], np.float32).reshape(3, 3).T.copy(),
                xicp_sums=np.array(res.xicp_sums[:], np.float64),
                H=logH[:it].reshape(it, 6, 6).transpose(0, 2, 1).copy(), b=logb[:it], x=logx[:it],
                T_iters=logT[:it].reshape(it, 4, 4).transpose(0, 2, 1).copy(), P=logP[:it], ids=ids, dists=dists,
                weights=w)


def icp(params, ref4, refn, read4, readn=None, T_init=None, max_log=64, forced_T_iters=None):
    """forced_T_iters (n,4,4): one-step mode, see orc_icp_forced — T_iters[i] of the result is the oracle's own
    update computed from forced_T_iters[i-1]; ids/dists/weights are those of the last iteration."""
    ref4, refn, read4, readn = f32(ref4), f32(refn), f32(read4), f32(readn)
    M, N = ref4.shape[0], read4.shape[0]
    call = lambda *a: lib().orc_icp_forced(C.byref(params), _fp(ref4), _fp(refn), M, _fp(read4), _fp(readn), N, *a)
    return _icp_run(call, params, N, T_init, max_log, forced_T_iters)


class IcpSession:
    """initReference once, compute many (Mapper.cpp:343,373 pattern); lets the CPU baseline time compute() apart
    from the kd-tree build."""

    def __init__(self, ref4, refn, mean_mode=1):
        self.ref4, self.refn = f32(ref4), f32(refn)
        lib().orc_icp_ref_create.restype = C.c_void_p
        self.h = C.c_void_p(lib().orc_icp_ref_create(_fp(self.ref4), _fp(self.refn), self.ref4.shape[0], mean_mode))

    def compute(self, params, read4, readn=None, T_init=None):
        read4, readn = f32(read4), f32(readn)
        Ti = np.eye(4, dtype=np.float32) if T_init is None else np.asarray(T_init, np.float32)
        Tc = np.ascontiguousarray(Ti.T).ravel()
        res = IcpResult()
        st = lib().orc_icp_compute(self.h, C.byref(params), _fp(read4), _fp(readn), read4.shape[0], _fp(Tc),
                                   C.byref(res), 0, None, None, None, None, None, None, None, None, None, 0)
        m = lambda a: np.array(a, np.float32).reshape(4, 4).T.copy()
        return dict(status=st, iterations=res.iterations, max_iterations_reached=bool(res.max_iterations_reached),
                    n_matches_last=res.n_matches_last, T=m(res.T), points_touched=res.points_touched)

    def icp(self, params, read4, readn=None, T_init=None, max_log=64, forced_T_iters=None):
        """like the module-level icp() (logs, last matches, one-step mode) on the kept reference"""
        read4, readn = f32(read4), f32(readn)
        N = read4.shape[0]
        call = lambda *a: lib().orc_icp_compute(self.h, C.byref(params), _fp(read4), _fp(readn), N, *a)
        return _icp_run(call, params, N, T_init, max_log, forced_T_iters)

    def close(self):
        if getattr(self, "h", None):
            lib().orc_icp_ref_free(self.h)
            self.h = None

    def __del__(self):
        self.close()


def o3d_covariances_from_normals(normals, eps=1e-3):
    n = f32(normals)
    cov = np.empty((n.shape[0], 9), np.float64)
    lib().orc_o3d_covariances_from_normals(_fp(n), n.shape[0], C.c_double(eps), cov.ctypes.data_as(C.POINTER(C.c_double)))
    return cov


def o3d_icp(kind, src4, tgt4, tgt_normals=None, src_normals=None, max_dist=1.0, init=None, max_iteration=30,
            rel_fitness=1e-6, rel_rmse=1e-6, eps=1e-3, threads=4):
    """Open3D RegistrationICP (kind 0 point-to-plane) / RegistrationGeneralizedICP (kind 2, covariances from normals)."""
    src4, tgt4, tn, sn = f32(src4), f32(tgt4), f32(tgt_normals), f32(src_normals)
    n = src4.shape[0]
    dp = C.POINTER(C.c_double)
    Ti = np.eye(4) if init is None else np.asarray(init, np.float64)
    Tc = np.ascontiguousarray(Ti.T).ravel()
    sc = tc = None
    if kind == 2:
        sc = o3d_covariances_from_normals(sn, eps)
        tc = o3d_covariances_from_normals(tn, eps)
    T = np.empty(16, np.float64)
    fit, rmse, it = C.c_double(), C.c_double(), C.c_int()
    ids = np.empty(n, np.int32)
    st = lib().orc_o3d_icp(kind, _fp(src4), n, _fp(tgt4), tgt4.shape[0], _fp(tn),
                           None if sc is None else sc.ctypes.data_as(dp), None if tc is None else tc.ctypes.data_as(dp),
                           C.c_float(max_dist), Tc.ctypes.data_as(dp), max_iteration, C.c_double(rel_fitness),
                           C.c_double(rel_rmse), threads, T.ctypes.data_as(dp), C.byref(fit), C.byref(rmse),
                           C.byref(it), _ip(ids))
    return dict(status=st, T=T.reshape(4, 4).T.copy(), fitness=fit.value, inlier_rmse=rmse.value,
                iterations=it.value, ids=ids)


def o3d_estimate_normals(pts4, knn=20, radius=1.0, threads=4):
    """-> (normals (N,3) float64, neighbourhood sizes (N,) int32); see orc_o3d_estimate_normals"""
    pts4 = f32(pts4)
    n = pts4.shape[0]
    out = np.empty((n, 3), np.float64)
    cnt = np.empty(n, np.int32)
    st = lib().orc_o3d_estimate_normals(_fp(pts4), n, int(knn), C.c_float(radius), int(threads),
                                        out.ctypes.data_as(C.POINTER(C.c_double)), _ip(cnt))
    assert st == ORC_OK, st
    return out, cnt


def o3d_information_matrix(src4, tgt4, T, max_dist, threads=4):
    """-> (6x6 float64, fitness, rmse, matched target ids (N,) int32)"""
    src4, tgt4 = f32(src4), f32(tgt4)
    dp = C.POINTER(C.c_double)
    Tc = np.ascontiguousarray(np.asarray(T, np.float64).T).ravel()
    info = np.empty(36, np.float64)
    fit, rmse = C.c_double(), C.c_double()
    ids = np.empty(src4.shape[0], np.int32)
    st = lib().orc_o3d_information_matrix(_fp(src4), src4.shape[0], _fp(tgt4), tgt4.shape[0], Tc.ctypes.data_as(dp),
                                          C.c_float(max_dist), int(threads), info.ctypes.data_as(dp), C.byref(fit),
                                          C.byref(rmse), _ip(ids))
    assert st == ORC_OK, st
    return info.reshape(6, 6).T.copy(), fit.value, rmse.value, ids


class RansacParams(C.Structure):
    _fields_ = [("mutual_filter", C.c_int), ("max_correspondence_distance", C.c_double), ("ransac_n", C.c_int),
                ("edge_length_similarity", C.c_double), ("distance_threshold", C.c_double), ("max_iteration", C.c_int),
                ("confidence", C.c_double), ("seed", C.c_uint64)]


class RansacResult(C.Structure):
    _fields_ = [("T", C.c_double * 16), ("fitness", C.c_double), ("inlier_rmse", C.c_double),
                ("n_correspondences", C.c_int), ("n_feature_correspondences", C.c_int), ("mutual_used", C.c_int),
                ("iterations_run", C.c_int64), ("validations", C.c_int64)]


def o3d_fpfh(pts4, normals, radius, max_nn, threads=4):
    """orc_o3d_fpfh -> (features (N,33) float64, neighbourhood sizes (N,) int32)"""
    pts4 = f32(pts4)
    nrm = np.ascontiguousarray(normals, np.float64)
    n = pts4.shape[0]
    out = np.empty((n, 33), np.float64)
    cnt = np.empty(n, np.int32)
    dp = C.POINTER(C.c_double)
    st = lib().orc_o3d_fpfh(_fp(pts4), nrm.ctypes.data_as(dp), n, C.c_float(radius), int(max_nn), int(threads),
                            out.ctypes.data_as(dp), _ip(cnt))
    assert st == ORC_OK, st
    return out, cnt


def o3d_feature_nn(a, b, threads=4):
    """exact 1-NN of every row of a (N,33) in b (M,33) -> (ids int32, squared distances float64)"""
    a, b = np.ascontiguousarray(a, np.float64), np.ascontiguousarray(b, np.float64)
    ids = np.empty(a.shape[0], np.int32)
    d2 = np.empty(a.shape[0], np.float64)
    dp = C.POINTER(C.c_double)
    st = lib().orc_o3d_feature_nn(a.ctypes.data_as(dp), a.shape[0], b.ctypes.data_as(dp), b.shape[0], int(threads),
                                  _ip(ids), d2.ctypes.data_as(dp))
    assert st == ORC_OK, st
    return ids, d2


def o3d_ransac(src4, tgt4, sf, tf, mutual_filter=True, max_correspondence_distance=0.75, ransac_n=3,
               edge_length_similarity=0.5, distance_threshold=0.75, max_iteration=1000000, confidence=0.99, seed=1,
               threads=4):
    """orc_o3d_ransac_feature_matching -> dict (T 4x4, fitness, inlier_rmse, correspondences (K,2), ...)"""
    src4, tgt4 = f32(src4), f32(tgt4)
    sf, tf = np.ascontiguousarray(sf, np.float64), np.ascontiguousarray(tf, np.float64)
    prm = RansacParams(int(bool(mutual_filter)), max_correspondence_distance, int(ransac_n), edge_length_similarity,
                       distance_threshold, int(max_iteration), confidence, int(seed))
    res = RansacResult()
    corr = np.empty((max(src4.shape[0], 1), 2), np.int32)
    dp = C.POINTER(C.c_double)
    st = lib().orc_o3d_ransac_feature_matching(_fp(src4), src4.shape[0], _fp(tgt4), tgt4.shape[0], sf.ctypes.data_as(dp),
                                               tf.ctypes.data_as(dp), C.byref(prm), int(threads), C.byref(res), _ip(corr))
    assert st == ORC_OK, st
    return dict(T=np.array(res.T).reshape(4, 4).T.copy(), fitness=res.fitness, inlier_rmse=res.inlier_rmse,
                correspondences=corr[:res.n_correspondences].copy(), n_feature_correspondences=res.n_feature_correspondences,
                mutual_used=bool(res.mutual_used), iterations_run=res.iterations_run, validations=res.validations)


def overlap_indices(src4, tgt4, T, voxel_size, min_points=1):
    """orc_overlap_indices -> (source indices, target indices) int32, ascending"""
    src4, tgt4 = f32(src4), f32(tgt4)
    Tc = np.ascontiguousarray(np.asarray(T, np.float64).T).ravel()
    si, ti = np.empty(max(len(src4), 1), np.int32), np.empty(max(len(tgt4), 1), np.int32)
    ns, nt = C.c_int(), C.c_int()
    st = lib().orc_overlap_indices(_fp(src4), len(src4), _fp(tgt4), len(tgt4), Tc.ctypes.data_as(C.POINTER(C.c_double)),
                                   C.c_double(voxel_size), int(min_points), _ip(si), C.byref(ns), _ip(ti), C.byref(nt))
    assert st == ORC_OK, st
    return si[:ns.value].copy(), ti[:nt.value].copy()


def crop(kind, pts4, radius_min=0.0, radius_max=0.0, min_z=0.0, max_z=0.0, center=(0.0, 0.0, 0.0), invert=False):
    """orc_crop -> indices of the kept points, input order (croppers.cpp:66-95)"""
    pts4 = f32(pts4)
    n = pts4.shape[0]
    idx = np.empty(max(n, 1), np.int32)
    c = (C.c_double * 3)(*[float(v) for v in center])
    k = lib().orc_crop(int(kind), C.c_double(radius_min), C.c_double(radius_max), C.c_double(min_z), C.c_double(max_z), c,
                       int(bool(invert)), _fp(pts4), n, _ip(idx))
    return idx[:k].copy()


def voxel_downsample(voxel_size, pts4, normals=None):
    """orc_voxel_downsample -> (points Kx3 float64, normals Kx3 float64 or None, first index K int32)"""
    pts4, normals = f32(pts4), f32(normals)
    n = pts4.shape[0]
    out = np.empty((max(n, 1), 3), np.float64)
    outn = np.empty((max(n, 1), 3), np.float64) if normals is not None else None
    first = np.empty(max(n, 1), np.int32)
    dp = C.POINTER(C.c_double)
    k = lib().orc_voxel_downsample(C.c_double(voxel_size), _fp(pts4), _fp(normals), n, out.ctypes.data_as(dp),
                                   None if outn is None else outn.ctypes.data_as(dp), _ip(first))
    if k < 0:
        return k, None, None
    return out[:k].copy(), (None if outn is None else outn[:k].copy()), first[:k].copy()


def map_insert_scan(map3, mapn3, scan4, scann3, T, voxel_size, kind=0, radius_min=0.0, radius_max=0.0, min_z=0.0,
                    max_z=0.0, center=(0.0, 0.0, 0.0), invert=False):
    """Submap::insertScan restated (orc_map_insert_scan) -> (points (K,3) f64, normals or None, n_outside)"""
    dp = C.POINTER(C.c_double)
    map3 = np.ascontiguousarray(map3, np.float64).reshape(-1, 3)
    mapn3 = None if mapn3 is None else np.ascontiguousarray(mapn3, np.float64).reshape(-1, 3)
    scan4, scann3 = f32(scan4), f32(scann3)
    M, N = map3.shape[0], scan4.shape[0]
    has_n = scann3 is not None
    if has_n and mapn3 is None:
        mapn3 = np.zeros((0, 3), np.float64)
    out = np.empty((M + N, 3), np.float64)
    outn = np.empty((M + N, 3), np.float64) if has_n else None
    Tc = np.ascontiguousarray(np.asarray(T, np.float64).T).ravel()
    c = (C.c_double * 3)(*[float(v) for v in center])
    no = C.c_int()
    lib().orc_map_insert_scan.restype = C.c_int
    k = lib().orc_map_insert_scan(map3.ctypes.data_as(dp), None if mapn3 is None else mapn3.ctypes.data_as(dp), M,
                                  _fp(scan4), _fp(scann3), N, Tc.ctypes.data_as(dp), C.c_double(voxel_size), int(kind),
                                  C.c_double(radius_min), C.c_double(radius_max), C.c_double(min_z), C.c_double(max_z), c,
                                  int(bool(invert)), out.ctypes.data_as(dp), None if outn is None else outn.ctypes.data_as(dp),
                                  C.byref(no))
    return out[:k].copy(), (None if outn is None else outn[:k].copy()), no.value


def crop_d(kind, pts3, radius_min=0.0, radius_max=0.0, min_z=0.0, max_z=0.0, center=(0.0, 0.0, 0.0), invert=False):
    pts3 = np.ascontiguousarray(pts3, np.float64)
    n = pts3.shape[0]
    idx = np.empty(max(n, 1), np.int32)
    c = (C.c_double * 3)(*[float(v) for v in center])
    k = lib().orc_crop_d(int(kind), C.c_double(radius_min), C.c_double(radius_max), C.c_double(min_z), C.c_double(max_z), c,
                         int(bool(invert)), pts3.ctypes.data_as(C.POINTER(C.c_double)), n, _ip(idx))
    return idx[:k].copy()


def random_downsample_indices(n, ratio, seed=0):
    """The selection rule of b200_prep_random_downsample restated in numpy (Open3D's own RandomDownSample draws from
    std::random_device and cannot be reproduced): the int(ratio * n) indices with the smallest fmix32(i ^ seed * phi)."""
    if ratio >= 1.0:
        return np.arange(n, dtype=np.int32)
    k = int(ratio * n)
    x = np.arange(n, dtype=np.uint32) ^ np.uint32((int(seed) * 0x9E3779B9) & 0xFFFFFFFF)
    x ^= x >> np.uint32(16)
    x = (x.astype(np.uint64) * np.uint64(0x85ebca6b) & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    x ^= x >> np.uint32(13)
    x = (x.astype(np.uint64) * np.uint64(0xc2b2ae35) & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    x ^= x >> np.uint32(16)
    assert np.unique(x).size == n          # a bijection: no ties
    return np.sort(np.argsort(x, kind="stable")[:k]).astype(np.int32)
