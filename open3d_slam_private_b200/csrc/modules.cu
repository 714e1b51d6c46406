// Stand-alone module kernels (see modules.cuh). HBM-bound streaming / gather kernels; grids sized as multiples of
// the SM count; per-thread partials reduced with warp shuffles and one fixed-order final pass (deterministic).
#include "modules.cuh"
#include "normals.cuh"
#include "knn.cuh"
#include "solve.cuh"
#include <algorithm>
#include <cmath>

namespace b200 {

namespace {

constexpr int kThreads = 256;
constexpr int kBins = 2048;

inline int blocks_for(size_t n, int threads) {
  const size_t b = (n + threads - 1) / threads;
  return (int)std::max<size_t>(1, std::min<size_t>(b, 148 * 16));
}

// ---------------------------------------------------------------------------------------------------------------
// kNN
// ---------------------------------------------------------------------------------------------------------------
template <int K>
__global__ void __launch_bounds__(kThreads) knn_kernel(const __grid_constant__ GridView g,
                                                       const float4* __restrict__ query, int N, int kq, float md2,
                                                       const float* __restrict__ mdpp, int32_t* __restrict__ ids,
                                                       float* __restrict__ dists,
                                                       unsigned long long* __restrict__ visited) {
  unsigned long long vis_total = 0ull;
  __shared__ Warp1Scratch s_ws[K == 1 ? kThreads / 32 : 1];
  // warp-uniform trip count: the k = 1 search is a warp collective
  const int lane = threadIdx.x & 31;
  for (int base = (blockIdx.x * blockDim.x + threadIdx.x) - lane; base < N; base += gridDim.x * blockDim.x) {
    const int i = base + lane;
    const bool in = i < N;
    const float4 q = in ? __ldg(&query[i]) : make_float4(0.f, 0.f, 0.f, 1.f);
    float m2 = md2;
    if (mdpp && in) {  // KDTreeVarDistMatcher: per-point radius
      const float r = __ldg(&mdpp[i]);
      m2 = isinf(r) ? INFINITY : __fmul_rn(r, r);
    }
    unsigned vis = 0;
    if (K == 1) {
      float bd; int bid, bsp;
      knn_query_warp1(g, in, q.x, q.y, q.z, m2, bd, bid, bsp, vis, s_ws[K == 1 ? (threadIdx.x >> 5) : 0]);
      if (in) { ids[i] = bid; dists[i] = bd; }
    } else if (in) {
      KList<K> L;
      knn_query<K>(g, q.x, q.y, q.z, m2, kq, L, vis);
#pragma unroll
      for (int j = 0; j < K; ++j)
        if (j < kq) {
          ids[(size_t)i * kq + j] = L.id[j];
          dists[(size_t)i * kq + j] = L.d[j];
        }
    }
    vis_total += vis;
  }
  for (int o = 16; o > 0; o >>= 1) vis_total += __shfl_xor_sync(0xffffffffu, vis_total, o);
  if (visited && (threadIdx.x & 31) == 0 && vis_total) atomicAdd(visited, vis_total);
}

// 2 <= kq <= 32: one warp per query, the k-best list spread over the lanes (knn_query_warpk)
__global__ void __launch_bounds__(kThreads) knn_warpk_kernel(const __grid_constant__ GridView g,
                                                             const float4* __restrict__ query, int N, int kq, float md2,
                                                             const float* __restrict__ mdpp, int32_t* __restrict__ ids,
                                                             float* __restrict__ dists,
                                                             unsigned long long* __restrict__ visited) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  unsigned long long vis_total = 0ull;
  for (int i = warp; i < N; i += nwarps) {
    const float4 q = __ldg(&query[i]);
    float m2 = md2;
    if (mdpp) {  // KDTreeVarDistMatcher: per-point radius
      const float r = __ldg(&mdpp[i]);
      m2 = isinf(r) ? INFINITY : __fmul_rn(r, r);
    }
    unsigned vis = 0;
    float d; int id, sp;
    knn_query_warpk(g, q.x, q.y, q.z, m2, kq, d, id, sp, vis);
    if (lane < kq) {
      ids[(size_t)i * kq + lane] = id;
      dists[(size_t)i * kq + lane] = d;
    }
    vis_total += vis;
  }
  if (visited && lane == 0 && vis_total) atomicAdd(visited, vis_total);
}

int launch_knn_warpk(const GridView& g, const float4* q, int N, int kq, float md2, const float* mdpp, int32_t* ids,
                     float* dists, unsigned long long* visited, cudaStream_t s) {
  const int warps_per_block = kThreads / 32;
  const int blocks = (int)std::min<long long>(((long long)N + warps_per_block - 1) / warps_per_block, 148LL * 64);
  knn_warpk_kernel<<<std::max(1, blocks), kThreads, 0, s>>>(g, q, N, kq, md2, mdpp, ids, dists, visited);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

template <int K>
int launch_knn(const GridView& g, const float4* q, int N, int kq, float md2, const float* mdpp, int32_t* ids,
               float* dists, unsigned long long* visited, cudaStream_t s) {
  knn_kernel<K><<<blocks_for(N, kThreads), kThreads, 0, s>>>(g, q, N, kq, md2, mdpp, ids, dists, visited);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// quantile: 3-pass radix select over the bits of non-negative floats (11 + 11 + 9 bits)
// state words: [0] status, [1] result bits, [2] prefix, [3] rank, [4] total; histograms at +16
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) qhist_kernel(const float* __restrict__ d, int64_t n, int pass,
                                                         uint32_t* __restrict__ state) {
  __shared__ unsigned sh[kBins];
  for (int b = threadIdx.x; b < kBins; b += blockDim.x) sh[b] = 0u;
  __syncthreads();
  if (state[0] != B200_OK) return;
  const unsigned prefix = state[2];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = __ldg(&d[i]);
    if (v == INFINITY) continue;
    const unsigned bits = __float_as_uint(v);
    if (pass == 0) atomicAdd(&sh[(bits >> 20) & 0x7ffu], 1u);
    else if (pass == 1) { if (((bits >> 20) & 0x7ffu) == prefix) atomicAdd(&sh[(bits >> 9) & 0x7ffu], 1u); }
    else { if (((bits >> 9) & 0x3fffffu) == prefix) atomicAdd(&sh[bits & 0x1ffu], 1u); }
  }
  __syncthreads();
  unsigned* gh = state + 16 + pass * kBins;
  for (int b = threadIdx.x; b < kBins; b += blockDim.x)
    if (sh[b]) atomicAdd(&gh[b], sh[b]);
}

__global__ void qpick_kernel(int pass, float quantile, uint32_t* __restrict__ state) {
  // single thread: 2048 bins, negligible
  if (state[0] != B200_OK) return;
  const unsigned* gh = state + 16 + pass * kBins;
  unsigned rank = state[3];
  if (pass == 0) {
    unsigned total = 0;
    for (int b = 0; b < kBins; ++b) total += gh[b];
    state[4] = total;
    if (total == 0u) { state[0] = B200_E_NO_MATCHES; return; }
    if (quantile == 1.0f) rank = total - 1u;
    else {
      rank = (unsigned)__fmul_rn((float)total, quantile);
      if (rank > total - 1u) rank = total - 1u;
    }
  }
  unsigned ex = 0;
  int bin = 0;
  for (int b = 0; b < kBins; ++b) {
    if (rank < ex + gh[b]) { bin = b; break; }
    ex += gh[b];
  }
  rank -= ex;
  state[3] = rank;
  if (pass == 0) state[2] = (unsigned)bin;
  else if (pass == 1) state[2] = (state[2] << 11) | (unsigned)bin;
  else state[1] = (state[2] << 9) | (unsigned)bin;
}

__global__ void outlier_dist_kernel(int kind, float param, const float* __restrict__ d, int64_t n,
                                    const uint32_t* __restrict__ state, float* __restrict__ w) {
  float lim = param;
  if (kind == B200_OUTLIER_MEDIANDIST || kind == B200_OUTLIER_TRIMMEDDIST) {
    if (state[0] != B200_OK) return;
    const float qv = __uint_as_float(state[1]);
    lim = kind == B200_OUTLIER_MEDIANDIST ? __fmul_rn(param, qv) : qv;
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = __ldg(&d[i]);
    float r;
    switch (kind) {
      case B200_OUTLIER_NULL: r = (v == INFINITY) ? 0.f : 1.f; break;
      case B200_OUTLIER_MINDIST: r = (v >= lim) ? 1.f : 0.f; break;
      default: r = (v <= lim) ? 1.f : 0.f; break;
    }
    w[i] = r;
  }
}

__global__ void outlier_sn_kernel(const float* __restrict__ nread, const float* __restrict__ nref,
                                  const int32_t* __restrict__ ids, int k, int N, float eps, float* __restrict__ w) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < (int64_t)k * N;
       e += (int64_t)gridDim.x * blockDim.x) {
    if (!nread || !nref) { w[e] = 1.f; continue; }
    const int x = (int)(e / k);
    const int id = ids[e];
    if (id == -1) { w[e] = 0.f; continue; }
    const float a[3] = {nread[3 * (size_t)x], nread[3 * (size_t)x + 1], nread[3 * (size_t)x + 2]};
    const float b[3] = {nref[3 * (size_t)id], nref[3 * (size_t)id + 1], nref[3 * (size_t)id + 2]};
    float na[3], nb[3];
    normalized3_dev(a, na);
    normalized3_dev(b, nb);
    float v = __fmul_rn(na[0], nb[0]);
    v = __fadd_rn(v, __fmul_rn(na[1], nb[1]));
    v = __fadd_rn(v, __fmul_rn(na[2], nb[2]));
    w[e] = (v < eps) ? 0.f : 1.f;
  }
}

}  // namespace

int knn_device(const GridView& g, const float4* q, int N, int kq, float maxDist, const float* mdpp, int32_t* ids,
               float* dists, unsigned long long* visited, cudaStream_t s) {
  if (N <= 0) return B200_OK;
  const float md2 = std::isinf(maxDist) ? INFINITY : maxDist * maxDist;
  if (kq <= 1) return launch_knn<1>(g, q, N, kq, md2, mdpp, ids, dists, visited, s);
  if (kq <= 32) return launch_knn_warpk(g, q, N, kq, md2, mdpp, ids, dists, visited, s);
  if (kq <= 64) return launch_knn<64>(g, q, N, kq, md2, mdpp, ids, dists, visited, s);
  set_error("knn > 64 is not supported by the B200 matcher");
  return B200_E_UNSUPPORTED;
}

int quantile_device(const float* d, int64_t n, float quantile, uint32_t* state, cudaStream_t s) {
  B200_CUDA(cudaMemsetAsync(state, 0, kQuantileScratchWords * sizeof(uint32_t), s));
  if (quantile < 0.f || quantile > 1.f) {  // ConvergenceError, Matches.cpp:78-79 (after the empty check)
    // the reference checks emptiness first; resolve on the host side of the C ABI
    set_error("Distance quantile of matches must lie in the range [0,1]");
    return B200_E_BAD_PARAM;
  }
  for (int pass = 0; pass < 3; ++pass) {
    qhist_kernel<<<blocks_for((size_t)n, kThreads), kThreads, 0, s>>>(d, n, pass, state);
    qpick_kernel<<<1, 1, 0, s>>>(pass, quantile, state);
  }
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int outlier_dist_device(int kind, float param, const float* d, int64_t n, const uint32_t* state, float* w,
                        cudaStream_t s) {
  if (n <= 0) return B200_OK;
  outlier_dist_kernel<<<blocks_for((size_t)n, kThreads), kThreads, 0, s>>>(kind, param, d, n, state, w);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int outlier_surface_normal_device(const float* nread, const float* nref, const int32_t* ids, int k, int N,
                                  float cos_max_angle, float* w, cudaStream_t s) {
  if ((int64_t)k * N <= 0) return B200_OK;
  outlier_sn_kernel<<<blocks_for((size_t)k * N, kThreads), kThreads, 0, s>>>(nread, nref, ids, k, N, cos_max_angle, w);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// =================================================================================================================
// ErrorElements, linearisation, solve, point-to-point, rigid transform
// =================================================================================================================
namespace {

// flags[e] = 1 iff pair e survives (valid distance and non-zero weight), ErrorMinimizer.cpp:103-107
__global__ void ee_flag_kernel(const float* __restrict__ w, const float* __restrict__ d, int64_t n,
                               uint32_t* __restrict__ flags) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x)
    flags[e] = (d[e] != INFINITY && w[e] != 0.0f) ? 1u : 0u;
}

// after the inclusive scan flags[e] = number of kept pairs in [0..e]; kept pair e goes to slot flags[e] - 1
__global__ void ee_scatter_kernel(const float4* __restrict__ read, const float4* __restrict__ ref,
                                  const float* __restrict__ refn, const float* __restrict__ w,
                                  const int32_t* __restrict__ ids, const float* __restrict__ d, int k, int64_t n,
                                  const uint32_t* __restrict__ incl, float4* __restrict__ readP,
                                  float4* __restrict__ refP, float* __restrict__ refnP, float* __restrict__ wP,
                                  int32_t* __restrict__ idsP, float* __restrict__ dP) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t cur = incl[e], prev = e ? incl[e - 1] : 0u;
    if (cur == prev) continue;
    const uint32_t j = cur - 1u;
    const int i = (int)(e / k);
    const int id = ids[e];
    if (readP) readP[j] = read[i];
    if (refP) refP[j] = ref[id];
    if (refnP && refn) {
      refnP[3 * (size_t)j] = refn[3 * (size_t)id];
      refnP[3 * (size_t)j + 1] = refn[3 * (size_t)id + 1];
      refnP[3 * (size_t)j + 2] = refn[3 * (size_t)id + 2];
    }
    if (wP) wP[j] = w[e];
    if (idsP) idsP[j] = id;
    if (dP) dP[j] = d[e];
  }
}

constexpr int kRedThreads = 256;
constexpr int kRedBlocksMax = 1024;

template <int NV>
__device__ __forceinline__ void block_reduce_store(double* v, double* __restrict__ partials) {
  __shared__ double sm[kRedThreads / 32][NV];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int t = 0; t < NV; ++t) {
    double s = v[t];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) sm[wid][t] = s;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    double s = 0.0;
    for (int w2 = 0; w2 < kRedThreads / 32; ++w2) s += sm[w2][threadIdx.x];
    partials[(size_t)blockIdx.x * 32 + threadIdx.x] = s;
  }
}

// H (21 unique) and b (6) per pair in float exactly like the reference's element-wise products, summed in double
__global__ void __launch_bounds__(kRedThreads) hessian_kernel(const float4* __restrict__ read,
                                                              const float4* __restrict__ ref,
                                                              const float* __restrict__ n3,
                                                              const float* __restrict__ w, int P,
                                                              double* __restrict__ partials) {
  double acc[27];
#pragma unroll
  for (int t = 0; t < 27; ++t) acc[t] = 0.0;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < P; p += gridDim.x * blockDim.x) {
    const float4 q = __ldg(&read[p]);
    const float4 r = __ldg(&ref[p]);
    const float nx = __ldg(&n3[3 * (size_t)p]), ny = __ldg(&n3[3 * (size_t)p + 1]), nz = __ldg(&n3[3 * (size_t)p + 2]);
    const float ww = __ldg(&w[p]);
    float F[6], wF[6];
    F[0] = __fsub_rn(__fmul_rn(q.y, nz), __fmul_rn(q.z, ny));
    F[1] = __fsub_rn(__fmul_rn(q.z, nx), __fmul_rn(q.x, nz));
    F[2] = __fsub_rn(__fmul_rn(q.x, ny), __fmul_rn(q.y, nx));
    F[3] = nx; F[4] = ny; F[5] = nz;
#pragma unroll
    for (int a = 0; a < 6; ++a) wF[a] = __fmul_rn(ww, F[a]);
    float dot = __fmul_rn(__fsub_rn(q.x, r.x), nx);
    dot = __fadd_rn(dot, __fmul_rn(__fsub_rn(q.y, r.y), ny));
    dot = __fadd_rn(dot, __fmul_rn(__fsub_rn(q.z, r.z), nz));
    int t = 0;
#pragma unroll
    for (int c = 0; c < 6; ++c)
#pragma unroll
      for (int a = 0; a <= c; ++a) acc[t++] += (double)__fmul_rn(wF[a], F[c]);
#pragma unroll
    for (int a = 0; a < 6; ++a) acc[21 + a] += (double)__fmul_rn(wF[a], dot);
  }
  block_reduce_store<27>(acc, partials);
}

// NOTE: the reference forms the full 36 products wF_i * F_j; (i,j) and (j,i) differ only when weights are not
// 0/1 ((w F_i) F_j vs (w F_j) F_i). The lower triangle is mirrored from the upper one here.
__global__ void hessian_final_kernel(const double* __restrict__ partials, int nblocks, float* __restrict__ out) {
  __shared__ double s[27];
  if (threadIdx.x < 27) {
    double a = 0.0;
    for (int b = 0; b < nblocks; ++b) a += partials[(size_t)b * 32 + threadIdx.x];
    s[threadIdx.x] = a;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int c = 0; c < 6; ++c)
      for (int r = 0; r <= c; ++r) {
        const float h = (float)s[t++];
        out[r + 6 * c] = h; out[c + 6 * r] = h;
      }
    for (int r = 0; r < 6; ++r) out[36 + r] = -(float)s[21 + r];
  }
}

// One warp, the same sequence the registration kernel runs (icp_fused.cu): the warp solve with its rank certificate
// first, the serial reference-literal path only when that decides nothing (0) or replays a rank-deficient QR (2).
// io[65] receives the warp path's return value (tests: certificate vs replay across the rank threshold).
__global__ void __launch_bounds__(32) solve_kernel(float* __restrict__ io) {
  __shared__ float H[36], b[6], x[6], x_spd[6];
  __shared__ double W[98];
  const int lane = threadIdx.x;
  for (int i = lane; i < 36; i += 32) H[i] = io[i];
  if (lane < 6) { b[lane] = io[36 + lane]; x[lane] = 0.f; x_spd[lane] = 0.f; }
  __syncwarp();
  const int fast = solve6_warp(H, b, x, W, x_spd);
  __syncwarp();
  if (lane == 0) {
    int info = 0;
    if (fast != 1) solve6_dev(H, b, x, &info, fast == 2 ? x_spd : nullptr);
    float T[16];
    p2plane_T_dev(x, T);
    for (int i = 0; i < 6; ++i) io[42 + i] = x[i];
    for (int i = 0; i < 16; ++i) io[48 + i] = T[i];
    io[64] = __int_as_float(info);
    io[65] = __int_as_float(fast);
  }
}

__global__ void __launch_bounds__(kRedThreads) moments_kernel(const float4* __restrict__ read,
                                                              const float4* __restrict__ ref,
                                                              const float* __restrict__ w, int P,
                                                              double* __restrict__ partials) {
  double m[16];
#pragma unroll
  for (int t = 0; t < 16; ++t) m[t] = 0.0;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < P; p += gridDim.x * blockDim.x) {
    const float4 q = __ldg(&read[p]);
    const float4 r = __ldg(&ref[p]);
    const double wd = (double)__ldg(&w[p]);
    const double rr[3] = {(double)q.x, (double)q.y, (double)q.z};
    const double ff[3] = {(double)r.x, (double)r.y, (double)r.z};
    m[0] += wd;
#pragma unroll
    for (int d = 0; d < 3; ++d) { m[1 + d] += wd * rr[d]; m[4 + d] += wd * ff[d]; }
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int a = 0; a < 3; ++a) m[7 + a + 3 * c] += wd * ff[a] * rr[c];
  }
  block_reduce_store<16>(m, partials);
}

__global__ void p2point_final_kernel(const double* __restrict__ partials, int nblocks, float* __restrict__ T) {
  __shared__ double s[16];
  if (threadIdx.x < 16) {
    double a = 0.0;
    for (int b = 0; b < nblocks; ++b) a += partials[(size_t)b * 32 + threadIdx.x];
    s[threadIdx.x] = a;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float Tl[16];
    p2point_from_moments(s, Tl);
    for (int i = 0; i < 16; ++i) T[i] = Tl[i];
  }
}

__global__ void rigid_apply_kernel(const float* __restrict__ T16, float4* __restrict__ feat,
                                   float* __restrict__ n3, int N) {
  __shared__ float Tc[16], R9[9];
  if (threadIdx.x == 0) rigid_prepare(T16, Tc, R9);
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    feat[i] = apply_T(Tc, feat[i]);
    if (n3) {
      const float3 r = apply_R(R9, make_float3(n3[3 * (size_t)i], n3[3 * (size_t)i + 1], n3[3 * (size_t)i + 2]));
      n3[3 * (size_t)i] = r.x; n3[3 * (size_t)i + 1] = r.y; n3[3 * (size_t)i + 2] = r.z;
    }
  }
}

inline int red_blocks(int P) { return std::max(1, std::min(kRedBlocksMax, (P + kRedThreads - 1) / kRedThreads)); }

}  // namespace

int error_elements_device(const float4* read, const float4* ref, const float* refn, const float* w,
                          const int32_t* ids, const float* d, int k, int N, uint32_t* flags, uint32_t* tile,
                          float4* readP, float4* refP, float* refnP, float* wP, int32_t* idsP, float* dP,
                          cudaStream_t s) {
  const int64_t n = (int64_t)k * N;
  if (n <= 0) return B200_OK;
  ee_flag_kernel<<<blocks_for((size_t)n, kThreads), kThreads, 0, s>>>(w, d, n, flags);
  B200_CUDA(cudaGetLastError());
  B200_TRY(inclusive_scan_u32(flags, (size_t)n, tile, s));
  ee_scatter_kernel<<<blocks_for((size_t)n, kThreads), kThreads, 0, s>>>(read, ref, refn, w, ids, d, k, n, flags,
                                                                        readP, refP, refnP, wP, idsP, dP);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int p2plane_hessian_device(const float4* read, const float4* ref, const float* n3, const float* w, int P,
                           double* partials, float* out, cudaStream_t s) {
  const int nb = red_blocks(P);
  hessian_kernel<<<nb, kRedThreads, 0, s>>>(read, ref, n3, w, P, partials);
  hessian_final_kernel<<<1, 32, 0, s>>>(partials, nb, out);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int p2plane_solve_device(float* io, cudaStream_t s) {
  solve_kernel<<<1, 32, 0, s>>>(io);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int p2point_device(const float4* read, const float4* ref, const float* w, int P, double* partials, float* T,
                   cudaStream_t s) {
  const int nb = red_blocks(P);
  moments_kernel<<<nb, kRedThreads, 0, s>>>(read, ref, w, P, partials);
  p2point_final_kernel<<<1, 32, 0, s>>>(partials, nb, T);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int rigid_apply_device(const float* T16, float4* feat, float* n3, int N, cudaStream_t s) {
  if (N <= 0) return B200_OK;
  rigid_apply_kernel<<<blocks_for((size_t)N, kThreads), kThreads, 0, s>>>(T16, feat, n3, N);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// =================================================================================================================
// SurfaceNormalDataPointsFilter (SurfaceNormal.cpp:82-290, utils.h:84-150): self-kNN + per-point patch statistics
// =================================================================================================================
namespace {

template <int K>
__global__ void __launch_bounds__(128) surface_normal_kernel(const __grid_constant__ GridView g,
                                                             const float4* __restrict__ feat, int N, int knn,
                                                             float md2, int sortEigen, float* __restrict__ normals,
                                                             float* __restrict__ dens, float* __restrict__ eigva,
                                                             float* __restrict__ eigve, float* __restrict__ matched,
                                                             float* __restrict__ meandist) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const float4 q = __ldg(&feat[i]);
    KList<K> L;
    unsigned vis = 0;
    knn_query<K>(g, q.x, q.y, q.z, md2, knn, L, vis);
    // mean over the valid neighbours in list order (d.rowwise().sum() / realKnn, :181)
    int realKnn = 0;
    float sx = 0.f, sy = 0.f, sz = 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j)
      if (j < knn && L.d[j] != INFINITY) {
        const float4 p = __ldg(&g.pts[L.sp[j]]);
        sx = __fadd_rn(sx, p.x); sy = __fadd_rn(sy, p.y); sz = __fadd_rn(sz, p.z);
        ++realKnn;
      }
    const float mx = sx / (float)realKnn, my = sy / (float)realKnn, mz = sz / (float)realKnn;
    float C[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) C[t] = 0.f;
    float maxSq = 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j)
      if (j < knn && L.d[j] != INFINITY) {
        const float4 p = __ldg(&g.pts[L.sp[j]]);
        const float nn[3] = {__fsub_rn(p.x, mx), __fsub_rn(p.y, my), __fsub_rn(p.z, mz)};
        float sq = __fmul_rn(nn[0], nn[0]);
        sq = __fadd_rn(sq, __fmul_rn(nn[1], nn[1]));
        sq = __fadd_rn(sq, __fmul_rn(nn[2], nn[2]));
        maxSq = fmaxf(maxSq, sq);
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int r = 0; r < 3; ++r) C[r + 3 * c] = __fadd_rn(C[r + 3 * c], __fmul_rn(nn[r], nn[c]));
      }
    sn_finish(i, q, C, realKnn, maxSq, mx, my, mz, sortEigen, normals, dens, eigva, eigve, meandist);
    if (matched) {
#pragma unroll
      for (int j = 0; j < K; ++j)
        if (j < knn) matched[(size_t)i * knn + j] = (float)L.id[j];
    }
  }
}

// knn <= 32: a warp takes 32 consecutive points. Phase 1, one point at a time: warp-cooperative k-NN (list across
// the lanes), mean and covariance accumulated in LIST ORDER with the same float operations as the per-thread kernel
// (lanes 0..8 own one covariance entry each); the 9 entries move to the lane that owns the point. Phase 2: every lane
// finishes its own point (rank test, eigen-decomposition, outputs).
#ifndef B200_SN_MINBLOCKS
#define B200_SN_MINBLOCKS 8   // 64 registers: 32 resident warps per SM (measured on the 5 M-point map, k = 20: 17.7 -> 14.6 ms)
#endif
__global__ void __launch_bounds__(128, B200_SN_MINBLOCKS) surface_normal_warp_kernel(const __grid_constant__ GridView g,
                                                                  const float4* __restrict__ feat, int N, int knn,
                                                                  float md2, int sortEigen, float* __restrict__ normals,
                                                                  float* __restrict__ dens, float* __restrict__ eigva,
                                                                  float* __restrict__ eigve, float* __restrict__ matched,
                                                                  float* __restrict__ meandist,
                                                                  const int* __restrict__ list,
                                                                  const int* __restrict__ list_count,
                                                                  int* __restrict__ ids_out, int out_by_id) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  const int er = lane % 3, ec = (lane / 3) % 3;   // covariance entry of lanes 0..8 (column-major r + 3 c)
  // list != null: only the points feat[list[0 .. *list_count)] (the ones the tile kernel could not settle)
  if (list) N = *list_count;
  const bool want_stats = normals || dens || eigva || eigve || meandist;
  for (int base = warp * 32; base < N; base += nwarps * 32) {
    float Cm[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) Cm[t] = 0.f;
    int m_real = 0;
    float m_maxSq = 0.f, m_mx = 0.f, m_my = 0.f, m_mz = 0.f;
    float4 m_q = make_float4(0.f, 0.f, 0.f, 1.f);
    const int cnt = min(32, N - base);
    const int my_i = lane < cnt ? (list ? list[base + lane] : base + lane) : 0;
    for (int t = 0; t < cnt; ++t) {
      const int i = __shfl_sync(FULL, my_i, t);
      const float4 q = __ldg(&feat[i]);
      unsigned vis = 0;
      float d; int id, sp;
      knn_query_warpk(g, q.x, q.y, q.z, md2, knn, d, id, sp, vis);
      if (!want_stats) {   // only the neighbour ids are wanted (normals_tile.cu: self_knn_ids_index)
        const int io = out_by_id ? __float_as_int(q.w) : i;
        if (matched && lane < knn) matched[(size_t)io * knn + lane] = (float)id;
        if (ids_out && lane < knn) ids_out[(size_t)io * knn + lane] = id;
        continue;
      }
      const bool valid = lane < knn && d != INFINITY;
      const int realKnn = __popc(__ballot_sync(FULL, valid));   // a prefix of the lanes (the list is sorted)
      float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
      if (valid) p = __ldg(&g.pts[sp]);
      float sx = 0.f, sy = 0.f, sz = 0.f;
      for (int j = 0; j < realKnn; ++j) {   // d.rowwise().sum() in list order (:181)
        sx = __fadd_rn(sx, __shfl_sync(FULL, p.x, j));
        sy = __fadd_rn(sy, __shfl_sync(FULL, p.y, j));
        sz = __fadd_rn(sz, __shfl_sync(FULL, p.z, j));
      }
      const float mx = sx / (float)realKnn, my = sy / (float)realKnn, mz = sz / (float)realKnn;
      const float nx = __fsub_rn(p.x, mx), ny = __fsub_rn(p.y, my), nz = __fsub_rn(p.z, mz);
      float sq = __fmul_rn(nx, nx);
      sq = __fadd_rn(sq, __fmul_rn(ny, ny));
      sq = __fadd_rn(sq, __fmul_rn(nz, nz));
      float maxSq = valid ? sq : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) maxSq = fmaxf(maxSq, __shfl_xor_sync(FULL, maxSq, o));
      float acc = 0.f;
      for (int j = 0; j < realKnn; ++j) {
        const float a = __shfl_sync(FULL, nx, j), b = __shfl_sync(FULL, ny, j), c = __shfl_sync(FULL, nz, j);
        const float vr = er == 0 ? a : (er == 1 ? b : c), vc = ec == 0 ? a : (ec == 1 ? b : c);
        acc = __fadd_rn(acc, __fmul_rn(vr, vc));
      }
#pragma unroll
      for (int e = 0; e < 9; ++e) {
        const float v = __shfl_sync(FULL, acc, e);
        if (lane == t) Cm[e] = v;
      }
      if (lane == t) { m_real = realKnn; m_maxSq = maxSq; m_mx = mx; m_my = my; m_mz = mz; m_q = q; }
      // out_by_id: feat is the cell-sorted array (w = bits of the original index) and the outputs go to that index
      const int io = out_by_id ? __float_as_int(q.w) : i;
      if (matched && lane < knn) matched[(size_t)io * knn + lane] = (float)id;
      if (ids_out && lane < knn) ids_out[(size_t)io * knn + lane] = id;   // -1 where there is no neighbour
    }
    if (lane < cnt && want_stats)
      sn_finish(out_by_id ? __float_as_int(m_q.w) : my_i, m_q, Cm, m_real, m_maxSq, m_mx, m_my, m_mz, sortEigen, normals, dens, eigva, eigve, meandist);
  }
}

template <int K>
int launch_sn(const GridView& g, const float4* feat, int N, int knn, float md2, int sortEigen, float* normals,
              float* dens, float* eigva, float* eigve, float* matched, float* meandist, cudaStream_t s) {
  surface_normal_kernel<K><<<blocks_for((size_t)N, 128), 128, 0, s>>>(g, feat, N, knn, md2, sortEigen, normals, dens,
                                                                      eigva, eigve, matched, meandist);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace

int surface_normals_device(const GridView& g, const float4* feat, int N, int knn, float maxDist, int sortEigen,
                           float* normals, float* dens, float* eigva, float* eigve, float* matched, float* meandist,
                           cudaStream_t s) {
  if (N <= 0) return B200_OK;
  const float md2 = std::isinf(maxDist) ? INFINITY : maxDist * maxDist;
  if (knn >= 2 && knn <= 32) {
    const long long chunks = ((long long)N + 31) / 32;   // one warp per 32 points, 4 warps per block
    const int blocks = (int)std::min<long long>((chunks + 3) / 4, 148LL * 32);
    surface_normal_warp_kernel<<<std::max(1, blocks), 128, 0, s>>>(g, feat, N, knn, md2, sortEigen, normals, dens, eigva,
                                                                   eigve, matched, meandist, nullptr, nullptr, nullptr, 0);
    B200_CUDA(cudaGetLastError());
    return B200_OK;
  }
#define B200_SN(KK) return launch_sn<KK>(g, feat, N, knn, md2, sortEigen, normals, dens, eigva, eigve, matched, meandist, s)
  if (knn <= 64) B200_SN(64);
#undef B200_SN
  set_error("SurfaceNormal knn > 64 is not supported by the B200 filter");
  return B200_E_UNSUPPORTED;
}

// the exact warp-cooperative path for a LIST of points (device-side count; see normals_tile.cu) or, with d_list ==
// null, for all N points of feat; out_by_id: feat is the cell-sorted array and outputs are indexed by original index
int surface_normals_list_device(const GridView& g, const float4* feat, const int* d_list, const int* d_count, int knn,
                                float maxDist, int sortEigen, float* normals, float* dens, float* eigva, float* eigve,
                                float* matched, float* meandist, int* ids_out, int out_by_id, int N, cudaStream_t s) {
  const float md2 = std::isinf(maxDist) ? INFINITY : maxDist * maxDist;
  int blocks = 148 * 8;
  if (!d_list) blocks = (int)std::max<long long>(1, std::min<long long>((((long long)N + 31) / 32 + 3) / 4, 148LL * 32));
  surface_normal_warp_kernel<<<blocks, 128, 0, s>>>(g, feat, N, knn, md2, sortEigen, normals, dens, eigva, eigve, matched,
                                                    meandist, d_list, d_count, ids_out, out_by_id);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
