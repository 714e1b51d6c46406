// C ABI of the B200 registration core (include/b200reg.h). Owns device memory, streams and host<->device staging;
// every compute step is a CUDA kernel from grid_index.cu / modules.cu / icp_fused.cu. There is no CPU fallback:
// without a device the compute entry points return B200_E_CUDA.
#include "common.cuh"
#include "grid_index.cuh"
#include "modules.cuh"
#include "normals.cuh"
#include "icp_kernel.cuh"
#include "solve.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <new>
#include <string>
#include <vector>

namespace b200 {

static thread_local std::string g_err;

void set_error(const std::string& msg) { g_err = msg; }

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  char buf[512];
  std::snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  g_err = buf;
  cudaGetLastError();  // clear sticky-less errors
  return B200_E_CUDA;
}

namespace {

// stream-ordered temporary device buffer
struct Tmp {
  void* p = nullptr;
  cudaStream_t s = nullptr;
  int alloc(size_t bytes, cudaStream_t stream) {
    s = stream;
    B200_CUDA(cudaMallocAsync(&p, std::max<size_t>(bytes, 16), stream));
    return B200_OK;
  }
  template <typename T> T* as() const { return static_cast<T*>(p); }
  ~Tmp() {
    if (p) cudaFreeAsync(p, s);
  }
};

struct CallCtx {  // one stream per modular call
  int device;
  DeviceGuard dg;
  cudaStream_t s = nullptr;
  explicit CallCtx(int dev) : device(dev), dg(dev) {}
  int open() {
    if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
    B200_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    return B200_OK;
  }
  ~CallCtx() {
    if (s) { cudaStreamSynchronize(s); cudaStreamDestroy(s); }
  }
};

int h2d(void* dst, const void* src, size_t bytes, cudaStream_t s) {
  if (bytes) B200_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s));
  return B200_OK;
}
int d2h(void* dst, const void* src, size_t bytes, cudaStream_t s) {
  if (bytes) B200_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s));
  return B200_OK;
}

}  // namespace
}  // namespace b200

using namespace b200;

// =================================================================================================================
struct b200_index {
  MapIndex map;
  cudaStream_t stream = nullptr;
  // query staging (grown on demand)
  float4* d_q = nullptr;
  int32_t* d_ids = nullptr;
  float* d_dists = nullptr;
  float* d_mdpp = nullptr;
  unsigned long long* d_visited = nullptr;
  size_t cap_q = 0, cap_kq = 0;
  // outputs of b200_index_self_normals (grown on demand, kept between calls)
  float* d_sn = nullptr;
  size_t cap_sn = 0;
  // CUDA events around the kernels of the last b200_index_knn / b200_index_self_normals (b200_index_last_device_ms)
  cudaEvent_t ev[2] = {nullptr, nullptr};
  float last_ms = 0.f;
};

extern "C" {

const char* b200_last_error(void) { return g_err.c_str(); }
int b200_version(void) { return 102; }
int b200_struct_size(int which) {
  switch (which) {
    case 0: return (int)sizeof(b200_icp_params);
    case 1: return (int)sizeof(b200_icp_result);
    case 2: return (int)sizeof(b200_index_opts);
    case 3: return (int)sizeof(b200_surface_normal_params);
    case 4: return (int)sizeof(b200_o3d_params);
    case 5: return (int)sizeof(b200_o3d_result);
    case 6: return (int)sizeof(b200_ransac_params);
    case 7: return (int)sizeof(b200_ransac_result);
    default: return -1;
  }
}

int b200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int b200_index_create(b200_index** out, int device, const float* ref4xM, const float* normals3xM, int M,
                      const b200_index_opts* opts) {
  if (!out || M < 0 || (M > 0 && !ref4xM)) { set_error("b200_index_create: bad arguments"); return B200_E_BAD_PARAM; }
  *out = nullptr;
  DeviceGuard dg(device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  b200_index* h = new (std::nothrow) b200_index();
  if (!h) { set_error("out of host memory"); return B200_E_CUDA; }
  h->map.device = device;
  auto fail = [&](int st) { b200_index_destroy(h); return st; };
  if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess)
    return fail(cuda_fail(cudaGetLastError(), "cudaStreamCreate", __FILE__, __LINE__));
  h->map.stream = h->stream;
  int st = index_alloc_points(h->map, M, normals3xM != nullptr);
  if (st) return fail(st);
  if ((st = upload_features(h->map.d_ref, ref4xM, M, h->stream))) return fail(st);
  if (normals3xM && M > 0) {
    Tmp stage;
    if ((st = stage.alloc((size_t)M * 3 * sizeof(float), h->stream))) return fail(st);
    if ((st = upload_normals(h->map.d_nrm_orig, normals3xM, M, stage.as<float>(), h->stream))) return fail(st);
  }
  if ((st = index_build(h->map, opts))) return fail(st);
  if (cudaMalloc(&h->d_visited, sizeof(unsigned long long)) != cudaSuccess)
    return fail(cuda_fail(cudaGetLastError(), "cudaMalloc", __FILE__, __LINE__));
  for (int i = 0; i < 2; ++i)
    if (cudaEventCreate(&h->ev[i]) != cudaSuccess) return fail(cuda_fail(cudaGetLastError(), "cudaEventCreate", __FILE__, __LINE__));
  *out = h;
  return B200_OK;
}

int b200_index_update(b200_index* h, const float* ref4xM, const float* normals3xM, int M, const b200_index_opts* opts) {
  if (!h || M < 0 || (M > 0 && !ref4xM)) { set_error("b200_index_update: bad arguments"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->map.device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  B200_TRY(index_alloc_points(h->map, M, normals3xM != nullptr));
  B200_TRY(upload_features(h->map.d_ref, ref4xM, M, h->stream));
  if (normals3xM && M > 0) {
    Tmp stage;
    B200_TRY(stage.alloc((size_t)M * 3 * sizeof(float), h->stream));
    B200_TRY(upload_normals(h->map.d_nrm_orig, normals3xM, M, stage.as<float>(), h->stream));
  }
  return index_build(h->map, opts);
}

void b200_index_destroy(b200_index* h) {
  if (!h) return;
  {
    DeviceGuard dg(h->map.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->d_q); cudaFree(h->d_ids); cudaFree(h->d_dists); cudaFree(h->d_mdpp); cudaFree(h->d_visited);
    cudaFree(h->d_sn);
    for (int i = 0; i < 2; ++i)
      if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    h->map.release();
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
}

int b200_index_knn(b200_index* h, const float* query4xN, int N, int knn, float epsilon, float maxDist,
                   const float* maxDistPerPoint, int32_t* ids, float* dists, uint64_t* visit_count) {
  (void)epsilon;  // exact search satisfies every epsilon
  if (!h || N < 0 || knn < 1 || (N > 0 && (!query4xN || !ids || !dists))) {
    set_error("b200_index_knn: bad arguments");
    return B200_E_BAD_PARAM;
  }
  if (knn > 64) { set_error("knn > 64 is not supported by the B200 matcher"); return B200_E_UNSUPPORTED; }
  if (!(maxDist >= 0.f)) { set_error("maxDist must be >= 0"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->map.device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  if (visit_count) *visit_count = 0;
  if (N == 0) return B200_OK;
  cudaStream_t s = h->stream;
  if ((size_t)N > h->cap_q) {
    cudaFree(h->d_q); cudaFree(h->d_mdpp);
    h->d_q = nullptr; h->d_mdpp = nullptr; h->cap_q = 0;
    B200_CUDA(cudaMalloc(&h->d_q, (size_t)N * sizeof(float4)));
    B200_CUDA(cudaMalloc(&h->d_mdpp, (size_t)N * sizeof(float)));
    h->cap_q = (size_t)N;
  }
  if ((size_t)N * knn > h->cap_kq) {
    cudaFree(h->d_ids); cudaFree(h->d_dists);
    h->d_ids = nullptr; h->d_dists = nullptr; h->cap_kq = 0;
    B200_CUDA(cudaMalloc(&h->d_ids, (size_t)N * knn * sizeof(int32_t)));
    B200_CUDA(cudaMalloc(&h->d_dists, (size_t)N * knn * sizeof(float)));
    h->cap_kq = (size_t)N * knn;
  }
  B200_TRY(h2d(h->d_q, query4xN, (size_t)N * sizeof(float4), s));
  if (maxDistPerPoint) B200_TRY(h2d(h->d_mdpp, maxDistPerPoint, (size_t)N * sizeof(float), s));
  B200_CUDA(cudaMemsetAsync(h->d_visited, 0, sizeof(unsigned long long), s));
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  B200_TRY(knn_device(h->map.view(), h->d_q, N, knn, maxDist, maxDistPerPoint ? h->d_mdpp : nullptr, h->d_ids,
                      h->d_dists, h->d_visited, s));
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  B200_TRY(d2h(ids, h->d_ids, (size_t)N * knn * sizeof(int32_t), s));
  B200_TRY(d2h(dists, h->d_dists, (size_t)N * knn * sizeof(float), s));
  unsigned long long vis = 0;
  B200_TRY(d2h(&vis, h->d_visited, sizeof(vis), s));
  B200_CUDA(cudaStreamSynchronize(s));
  B200_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]));
  if (visit_count) *visit_count = (uint64_t)vis;
  return B200_OK;
}

int b200_index_last_normals_fallback(b200_index* h, int* n) {
  if (!h || !n) { set_error("b200_index_last_normals_fallback: bad arguments"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->map.device);
  return surface_normals_last_fallback(h->map, n, h->stream);
}

int b200_index_last_device_ms(const b200_index* h, float* ms) {
  if (!h || !ms) { set_error("b200_index_last_device_ms: bad arguments"); return B200_E_BAD_PARAM; }
  *ms = h->last_ms;
  return B200_OK;
}

int b200_index_info(const b200_index* h, float* cell_size, int32_t dims[3], int64_t* occupied_cells, float* build_ms) {
  if (!h) { set_error("null index"); return B200_E_BAD_PARAM; }
  if (cell_size) *cell_size = h->map.cell;
  if (dims) { dims[0] = h->map.nx; dims[1] = h->map.ny; dims[2] = h->map.nz; }
  if (occupied_cells) *occupied_cells = h->map.occupied;
  if (build_ms) *build_ms = h->map.build_ms;
  return B200_OK;
}

}  // extern "C"

// =================================================================================================================
// per-module entry points (host buffers in, host buffers out)
// =================================================================================================================
extern "C" {

static int check_sn_params(const b200_surface_normal_params* prm) {
  if (!prm) { set_error("b200_surface_normals: bad arguments"); return B200_E_BAD_PARAM; }
  if (prm->knn < 3) { set_error("SurfaceNormalDataPointsFilter: knn must be >= 3"); return B200_E_BAD_PARAM; }
  if (prm->knn > 64) { set_error("SurfaceNormal knn > 64 is not supported"); return B200_E_UNSUPPORTED; }
  if (prm->smoothNormals && prm->knn > 32) {   // the smoothing pass takes its neighbour ids from the k <= 32 kernels
    set_error("smoothNormals = 1 with knn > 32 is not supported by the B200 filter");
    return B200_E_UNSUPPORTED;
  }
  return B200_OK;
}

int b200_index_self_normals(b200_index* h, const b200_surface_normal_params* prm, float* normals, float* densities,
                            float* eigValues, float* eigVectors, float* matchedIds, float* meanDists) {
  if (!h) { set_error("b200_index_self_normals: bad arguments"); return B200_E_BAD_PARAM; }
  B200_TRY(check_sn_params(prm));
  const int N = h->map.M, k = prm->knn;
  if (N == 0) return B200_OK;
  DeviceGuard dg(h->map.device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  // one grown-on-demand buffer, carved per requested output
  if (prm->smoothNormals && !normals) { set_error("smoothNormals needs the normals output"); return B200_E_BAD_PARAM; }
  const size_t cnt[6] = {normals ? (size_t)3 * N : 0, densities ? (size_t)N : 0, eigValues ? (size_t)3 * N : 0,
                         eigVectors ? (size_t)9 * N : 0, matchedIds ? (size_t)k * N : 0, meanDists ? (size_t)N : 0};
  size_t total = 0;
  for (size_t c : cnt) total += c;
  if (total > h->cap_sn) {
    cudaFree(h->d_sn);
    h->d_sn = nullptr; h->cap_sn = 0;
    B200_CUDA(cudaMalloc(&h->d_sn, total * sizeof(float)));
    h->cap_sn = total;
  }
  float* dp[6];
  size_t off = 0;
  for (int i = 0; i < 6; ++i) { dp[i] = cnt[i] ? h->d_sn + off : nullptr; off += cnt[i]; }
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  B200_TRY(surface_normals_index(h->map, h->map.d_ref, false, N, k, prm->maxDist, prm->sortEigen, prm->smoothNormals,
                                 dp[0], dp[1], dp[2], dp[3], dp[4], dp[5], s));
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  float* host[6] = {normals, densities, eigValues, eigVectors, matchedIds, meanDists};
  for (int i = 0; i < 6; ++i)
    if (host[i]) B200_TRY(d2h(host[i], dp[i], cnt[i] * sizeof(float), s));
  B200_CUDA(cudaStreamSynchronize(s));
  B200_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]));
  return B200_OK;
}

int b200_surface_normals(int device, const float* feat4xN, int N, const b200_surface_normal_params* prm,
                         float* normals, float* densities, float* eigValues, float* eigVectors, float* matchedIds,
                         float* meanDists) {
  if (N < 0 || (N > 0 && !feat4xN)) { set_error("b200_surface_normals: bad arguments"); return B200_E_BAD_PARAM; }
  B200_TRY(check_sn_params(prm));
  if (N == 0) return B200_OK;
  b200_index* idx = nullptr;
  B200_TRY(b200_index_create(&idx, device, feat4xN, nullptr, N, nullptr));
  const int st = b200_index_self_normals(idx, prm, normals, densities, eigValues, eigVectors, matchedIds, meanDists);
  b200_index_destroy(idx);
  return st;
}

int b200_dists_quantile(int device, const float* dists, int64_t n, float quantile, float* out) {
  if (n < 0 || (n > 0 && !dists) || !out) { set_error("b200_dists_quantile: bad arguments"); return B200_E_BAD_PARAM; }
  CallCtx c(device);
  B200_TRY(c.open());
  Tmp td, ts;
  B200_TRY(td.alloc((size_t)n * sizeof(float), c.s));
  B200_TRY(ts.alloc(kQuantileScratchWords * sizeof(uint32_t), c.s));
  B200_TRY(h2d(td.p, dists, (size_t)n * sizeof(float), c.s));
  // emptiness is checked before the range (Matches.cpp:75-79): run with a clamped quantile, then report
  const bool bad_q = quantile < 0.f || quantile > 1.f;
  B200_TRY(quantile_device(td.as<float>(), n, bad_q ? 0.5f : quantile, ts.as<uint32_t>(), c.s));
  uint32_t res[2];
  B200_TRY(d2h(res, ts.p, sizeof(res), c.s));
  B200_CUDA(cudaStreamSynchronize(c.s));
  if (res[0] == B200_E_NO_MATCHES) { set_error("No matches available for computing distance quantiles"); return B200_E_NO_MATCHES; }
  if (bad_q) { set_error("Distance quantile of matches must lie in the range [0,1]"); return B200_E_BAD_PARAM; }
  std::memcpy(out, &res[1], 4);
  return B200_OK;
}

int b200_outlier_dist(int device, int kind, float param, const float* dists, int64_t n, float* weights) {
  if (n < 0 || (n > 0 && (!dists || !weights)) || kind < B200_OUTLIER_NULL || kind > B200_OUTLIER_TRIMMEDDIST) {
    set_error("b200_outlier_dist: bad arguments");
    return B200_E_BAD_PARAM;
  }
  CallCtx c(device);
  B200_TRY(c.open());
  Tmp td, tw, ts;
  B200_TRY(td.alloc((size_t)n * sizeof(float), c.s));
  B200_TRY(tw.alloc((size_t)n * sizeof(float), c.s));
  B200_TRY(ts.alloc(kQuantileScratchWords * sizeof(uint32_t), c.s));
  B200_TRY(h2d(td.p, dists, (size_t)n * sizeof(float), c.s));
  float p = param;
  if (kind == B200_OUTLIER_MAXDIST || kind == B200_OUTLIER_MINDIST) p = (float)std::pow(param, 2);  // OutlierFiltersImpl.cpp:69,91
  bool bad_q = false;
  if (kind == B200_OUTLIER_MEDIANDIST || kind == B200_OUTLIER_TRIMMEDDIST) {
    const float q = kind == B200_OUTLIER_MEDIANDIST ? 0.5f : param;
    bad_q = q < 0.f || q > 1.f;
    B200_TRY(quantile_device(td.as<float>(), n, bad_q ? 0.5f : q, ts.as<uint32_t>(), c.s));
  } else {
    B200_CUDA(cudaMemsetAsync(ts.p, 0, 16 * sizeof(uint32_t), c.s));
  }
  B200_TRY(outlier_dist_device(kind, p, td.as<float>(), n, ts.as<uint32_t>(), tw.as<float>(), c.s));
  uint32_t res[2];
  B200_TRY(d2h(res, ts.p, sizeof(res), c.s));
  B200_TRY(d2h(weights, tw.p, (size_t)n * sizeof(float), c.s));
  B200_CUDA(cudaStreamSynchronize(c.s));
  if (res[0] == B200_E_NO_MATCHES) { set_error("No matches available for computing distance quantiles"); return B200_E_NO_MATCHES; }
  if (bad_q) { set_error("Distance quantile of matches must lie in the range [0,1]"); return B200_E_BAD_PARAM; }
  return B200_OK;
}

int b200_outlier_surface_normal(int device, const float* nread, const float* nref, int M, const int32_t* ids, int knn,
                                int N, float maxAngle, float* weights) {
  if (N < 0 || knn < 1 || M < 0 || (N > 0 && (!ids || !weights))) { set_error("b200_outlier_surface_normal: bad arguments"); return B200_E_BAD_PARAM; }
  CallCtx c(device);
  B200_TRY(c.open());
  const size_t n = (size_t)knn * N;
  Tmp ta, tb, ti, tw;
  B200_TRY(ti.alloc(n * sizeof(int32_t), c.s));
  B200_TRY(tw.alloc(n * sizeof(float), c.s));
  B200_TRY(h2d(ti.p, ids, n * sizeof(int32_t), c.s));
  if (nread && nref) {
    B200_TRY(ta.alloc((size_t)3 * N * sizeof(float), c.s));
    B200_TRY(tb.alloc((size_t)3 * M * sizeof(float), c.s));
    B200_TRY(h2d(ta.p, nread, (size_t)3 * N * sizeof(float), c.s));
    B200_TRY(h2d(tb.p, nref, (size_t)3 * M * sizeof(float), c.s));
  }
  B200_TRY(outlier_surface_normal_device(ta.as<float>(), tb.as<float>(), ti.as<int32_t>(), knn, N, std::cos(maxAngle),
                                         tw.as<float>(), c.s));
  B200_TRY(d2h(weights, tw.p, n * sizeof(float), c.s));
  B200_CUDA(cudaStreamSynchronize(c.s));
  return B200_OK;
}

int b200_error_elements(int device, const float* read4xN, const float* ref4xM, const float* refn3xM, int M,
                        const float* w, const int32_t* ids, const float* dists, int knn, int N, float* readP,
                        float* refP, float* refnP, float* wP, int32_t* idsP, float* distsP, int* P,
                        float* pointUsedRatio, float* weightedPointUsedRatio) {
  if (N <= 0 || knn < 1 || M <= 0 || !read4xN || !ref4xM || !w || !ids || !dists || !P) {
    set_error("b200_error_elements: bad arguments");
    return B200_E_BAD_PARAM;
  }
  CallCtx c(device);
  B200_TRY(c.open());
  const size_t n = (size_t)knn * N;
  Tmp tr, tf, tn, tw, ti, td, tfl, ttile, oR, oF, oN, oW, oI, oD;
  B200_TRY(tr.alloc((size_t)N * 16, c.s)); B200_TRY(tf.alloc((size_t)M * 16, c.s));
  B200_TRY(tw.alloc(n * 4, c.s)); B200_TRY(ti.alloc(n * 4, c.s)); B200_TRY(td.alloc(n * 4, c.s));
  B200_TRY(tfl.alloc(n * 4, c.s)); B200_TRY(ttile.alloc(std::max<size_t>(1, scan_tiles_for(n)) * 4, c.s));
  B200_TRY(oR.alloc(n * 16, c.s)); B200_TRY(oF.alloc(n * 16, c.s)); B200_TRY(oW.alloc(n * 4, c.s));
  B200_TRY(oI.alloc(n * 4, c.s)); B200_TRY(oD.alloc(n * 4, c.s));
  B200_TRY(h2d(tr.p, read4xN, (size_t)N * 16, c.s)); B200_TRY(h2d(tf.p, ref4xM, (size_t)M * 16, c.s));
  B200_TRY(h2d(tw.p, w, n * 4, c.s)); B200_TRY(h2d(ti.p, ids, n * 4, c.s)); B200_TRY(h2d(td.p, dists, n * 4, c.s));
  if (refn3xM) {
    B200_TRY(tn.alloc((size_t)M * 12, c.s)); B200_TRY(oN.alloc(n * 12, c.s));
    B200_TRY(h2d(tn.p, refn3xM, (size_t)M * 12, c.s));
  }
  B200_TRY(error_elements_device(tr.as<float4>(), tf.as<float4>(), tn.as<float>(), tw.as<float>(), ti.as<int32_t>(),
                                 td.as<float>(), knn, N, tfl.as<uint32_t>(), ttile.as<uint32_t>(), oR.as<float4>(),
                                 oF.as<float4>(), oN.as<float>(), oW.as<float>(), oI.as<int32_t>(), oD.as<float>(), c.s));
  uint32_t total = 0;
  B200_TRY(d2h(&total, tfl.as<uint32_t>() + (n - 1), 4, c.s));
  B200_CUDA(cudaStreamSynchronize(c.s));
  const size_t p = total;
  *P = (int)p;
  if (readP) B200_TRY(d2h(readP, oR.p, p * 16, c.s));
  if (refP) B200_TRY(d2h(refP, oF.p, p * 16, c.s));
  if (refnP && refn3xM) B200_TRY(d2h(refnP, oN.p, p * 12, c.s));
  std::vector<float> wtmp(p);
  B200_TRY(d2h(wtmp.data(), oW.p, p * 4, c.s));
  if (idsP) B200_TRY(d2h(idsP, oI.p, p * 4, c.s));
  if (distsP) B200_TRY(d2h(distsP, oD.p, p * 4, c.s));
  B200_CUDA(cudaStreamSynchronize(c.s));
  if (wP) std::memcpy(wP, wtmp.data(), p * 4);
  // ratios as the reference accumulates them (sequential float sum, ErrorMinimizer.cpp:121,139-140)
  float wsum = 0.f;
  for (size_t i = 0; i < p; ++i) wsum += wtmp[i];
  if (pointUsedRatio) *pointUsedRatio = (float)p / (float)(knn * N);
  if (weightedPointUsedRatio) *weightedPointUsedRatio = wsum / (float)(knn * N);
  if (p == 0) { set_error("ErrorMnimizer: no point to minimize"); return B200_E_NO_MATCHES; }
  return B200_OK;
}

int b200_p2plane_hessian(int device, const float* readP, const float* refP, const float* refnP, const float* wP, int P,
                         float* H36, float* b6) {
  if (P <= 0 || !readP || !refP || !refnP || !wP || !H36 || !b6) { set_error("b200_p2plane_hessian: bad arguments"); return B200_E_BAD_PARAM; }
  CallCtx c(device);
  B200_TRY(c.open());
  Tmp tr, tf, tn, tw, tp, to;
  B200_TRY(tr.alloc((size_t)P * 16, c.s)); B200_TRY(tf.alloc((size_t)P * 16, c.s)); B200_TRY(tn.alloc((size_t)P * 12, c.s));
  B200_TRY(tw.alloc((size_t)P * 4, c.s)); B200_TRY(tp.alloc(1024 * 32 * sizeof(double), c.s)); B200_TRY(to.alloc(42 * 4, c.s));
  B200_TRY(h2d(tr.p, readP, (size_t)P * 16, c.s)); B200_TRY(h2d(tf.p, refP, (size_t)P * 16, c.s));
  B200_TRY(h2d(tn.p, refnP, (size_t)P * 12, c.s)); B200_TRY(h2d(tw.p, wP, (size_t)P * 4, c.s));
  B200_TRY(p2plane_hessian_device(tr.as<float4>(), tf.as<float4>(), tn.as<float>(), tw.as<float>(), P, tp.as<double>(),
                                  to.as<float>(), c.s));
  float out[42];
  B200_TRY(d2h(out, to.p, sizeof(out), c.s));
  B200_CUDA(cudaStreamSynchronize(c.s));
  std::memcpy(H36, out, 36 * 4);
  std::memcpy(b6, out + 36, 6 * 4);
  return B200_OK;
}

int b200_p2plane_solve(int device, const float* H36, const float* b6, float* x6, float* T16, int* info) {
  if (!H36 || !b6) { set_error("b200_p2plane_solve: bad arguments"); return B200_E_BAD_PARAM; }
  CallCtx c(device);
  B200_TRY(c.open());
  Tmp io;
  B200_TRY(io.alloc(66 * 4, c.s));
  float buf[66] = {0};   // 36 H, 6 b | 6 x, 16 T, solve info, the warp path's verdict
  std::memcpy(buf, H36, 36 * 4);
  std::memcpy(buf + 36, b6, 6 * 4);
  B200_TRY(h2d(io.p, buf, sizeof(buf), c.s));
  B200_TRY(p2plane_solve_device(io.as<float>(), c.s));
  B200_TRY(d2h(buf, io.p, sizeof(buf), c.s));
  B200_CUDA(cudaStreamSynchronize(c.s));
  if (x6) std::memcpy(x6, buf + 42, 6 * 4);
  if (T16) std::memcpy(T16, buf + 48, 16 * 4);
  if (info) std::memcpy(info, buf + 64, 4);
  return B200_OK;
}

int b200_p2point(int device, const float* readP, const float* refP, const float* wP, int P, float* T16) {
  if (P <= 0 || !readP || !refP || !wP || !T16) { set_error("b200_p2point: bad arguments"); return B200_E_BAD_PARAM; }
  CallCtx c(device);
  B200_TRY(c.open());
  Tmp tr, tf, tw, tp, to;
  B200_TRY(tr.alloc((size_t)P * 16, c.s)); B200_TRY(tf.alloc((size_t)P * 16, c.s)); B200_TRY(tw.alloc((size_t)P * 4, c.s));
  B200_TRY(tp.alloc(1024 * 32 * sizeof(double), c.s)); B200_TRY(to.alloc(16 * 4, c.s));
  B200_TRY(h2d(tr.p, readP, (size_t)P * 16, c.s)); B200_TRY(h2d(tf.p, refP, (size_t)P * 16, c.s));
  B200_TRY(h2d(tw.p, wP, (size_t)P * 4, c.s));
  B200_TRY(p2point_device(tr.as<float4>(), tf.as<float4>(), tw.as<float>(), P, tp.as<double>(), to.as<float>(), c.s));
  B200_TRY(d2h(T16, to.p, 16 * 4, c.s));
  B200_CUDA(cudaStreamSynchronize(c.s));
  return B200_OK;
}

int b200_rigid_apply(int device, const float* T16, float* feat4xN, float* normals3xN, int N) {
  if (N < 0 || !T16 || (N > 0 && !feat4xN)) { set_error("b200_rigid_apply: bad arguments"); return B200_E_BAD_PARAM; }
  if (N == 0) return B200_OK;
  CallCtx c(device);
  B200_TRY(c.open());
  Tmp tT, tf, tn;
  B200_TRY(tT.alloc(16 * 4, c.s)); B200_TRY(tf.alloc((size_t)N * 16, c.s));
  B200_TRY(h2d(tT.p, T16, 16 * 4, c.s)); B200_TRY(h2d(tf.p, feat4xN, (size_t)N * 16, c.s));
  if (normals3xN) { B200_TRY(tn.alloc((size_t)N * 12, c.s)); B200_TRY(h2d(tn.p, normals3xN, (size_t)N * 12, c.s)); }
  B200_TRY(rigid_apply_device(tT.as<float>(), tf.as<float4>(), tn.as<float>(), N, c.s));
  B200_TRY(d2h(feat4xN, tf.p, (size_t)N * 16, c.s));
  if (normals3xN) B200_TRY(d2h(normals3xN, tn.p, (size_t)N * 12, c.s));
  B200_CUDA(cudaStreamSynchronize(c.s));
  return B200_OK;
}

}  // extern "C"
