// C ABI of the whole-registration chain (b200_icp_*): replaces PointMatcher<float>::ICP::{initReference, compute}
// (libpointmatcher/pointmatcher/ICP.cpp:813-1349). One call = 1 H2D of the reading + 4 kernels
// (centroid sums -> frames -> reading preparation -> persistent ICP kernel) + 1 D2H of the result struct, with no
// intermediate host synchronisation.
#include "common.cuh"
#include "grid_index.cuh"
#include "icp_kernel.cuh"
#include "knn.cuh"
#include "solve.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <new>

using namespace b200;

namespace b200 {
int icp_reference_from_device(b200_icp* h, int M, bool has_normals, const double* mean_given = nullptr);
float4* icp_reference_points(b200_icp* h, int M, bool with_normals, float4** normals_out, cudaStream_t* stream_out);
}  // namespace b200

namespace {

constexpr int kMeanBlocks = 256;
constexpr int kMeanThreads = 256;

// per-block partial sums of xyz in double (fixed order => deterministic)
__global__ void __launch_bounds__(kMeanThreads) read_sum_kernel(const float4* __restrict__ p, int N,
                                                                double* __restrict__ partials) {
  double s[3] = {0.0, 0.0, 0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const float4 v = __ldg(&p[i]);
    s[0] += (double)v.x; s[1] += (double)v.y; s[2] += (double)v.z;
  }
  __shared__ double sm[3][kMeanThreads / 32];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    for (int o = 16; o > 0; o >>= 1) s[d] += __shfl_xor_sync(0xffffffffu, s[d], o);
    if ((threadIdx.x & 31) == 0) sm[d][threadIdx.x >> 5] = s[d];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int w = 0; w < kMeanThreads / 32; ++w) t += sm[threadIdx.x][w];
    partials[3 * (size_t)blockIdx.x + threadIdx.x] = t;
  }
}

// frames[0..15] = T_refIn_refMean (set at initReference); this kernel fills
// frames[16..31] = T_refMean_readMean = T_refIn_refMean^-1 * T_init * T_readIn_readMean (ICP.cpp:983)
// frames[32..47] = T_readIn_readMean (ICP.cpp:968-970), frames[48..50] = reading centroid,
// frames[64..79] = T_refMean_dataIn = T_refIn_refMean^-1 * T_init (ICP.cpp:975), frames[80..95] = its inverse
// (transformationToOptimizationFrame.inverse() of the X-ICP detection, ICP.cpp:2236)
__global__ void frames_kernel(const double* __restrict__ partials, int nblocks, int N, const float* __restrict__ Tinit,
                              float* __restrict__ frames, int want_data_in) {
  // one warp: lane l sums blocks l, l + 32, ... (fixed order), then a fixed shuffle tree
  const int lane = threadIdx.x & 31;
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  double s[3] = {0.0, 0.0, 0.0};
  for (int b = lane; b < nblocks; b += 32)
    for (int d = 0; d < 3; ++d) s[d] += partials[3 * (size_t)b + d];
  for (int d = 0; d < 3; ++d)
    for (int o = 16; o > 0; o >>= 1) s[d] += __shfl_xor_sync(0xffffffffu, s[d], o);
  if (lane != 0) return;
  float mean[3];
  for (int d = 0; d < 3; ++d) mean[d] = (float)(s[d] / (double)N);
  float Tread[16], invRef[16], tmp[16], Tpre[16], Tref[16], Ti[16];
  eye4(Tread);
  for (int d = 0; d < 3; ++d) Tread[12 + d] = mean[d];
  for (int t = 0; t < 16; ++t) { Tref[t] = frames[t]; Ti[t] = Tinit[t]; }
  inv4_dev(Tref, invRef);
  mul4_dev(invRef, Ti, tmp);
  mul4_dev(tmp, Tread, Tpre);
  for (int t = 0; t < 16; ++t) { frames[16 + t] = Tpre[t]; frames[32 + t] = Tread[t]; }
  for (int d = 0; d < 3; ++d) frames[48 + d] = mean[d];
  if (want_data_in) {   // only the X-ICP detection needs T_refMean_dataIn and its inverse
    float tinv[16];
    if (!inv4_dev(tmp, tinv)) eye4(tinv);
    for (int t = 0; t < 16; ++t) { frames[64 + t] = tmp[t]; frames[80 + t] = tinv[t]; }
  }
}

// reading.features.topRows(3).colwise() -= mean (ICP.cpp:973), then transformations.apply(reading,
// T_refMean_readMean) (ICP.cpp:984) incl. the rotation of the "normals" descriptor
__global__ void prepare_reading_kernel(const float4* __restrict__ raw, const float4* __restrict__ raw_n, int N,
                                       const float* __restrict__ frames, float4* __restrict__ out,
                                       float4* __restrict__ out_n) {
  __shared__ float Tc[16], R9[9], mean[3];
  if (threadIdx.x == 0) {
    rigid_prepare(frames + 16, Tc, R9);
    mean[0] = frames[48]; mean[1] = frames[49]; mean[2] = frames[50];
  }
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    float4 v = __ldg(&raw[i]);
    v.x = __fsub_rn(v.x, mean[0]); v.y = __fsub_rn(v.y, mean[1]); v.z = __fsub_rn(v.z, mean[2]);
    out[i] = apply_T(Tc, v);
    if (raw_n) {
      const float4 n = __ldg(&raw_n[i]);
      const float3 r = apply_R(R9, make_float3(n.x, n.y, n.z));
      out_n[i] = make_float4(r.x, r.y, r.z, 0.f);
    }
  }
}

// open3dToPointmatcher (open3d_utils/open3d_conversions/src/open3d_conversions.cpp:57-118): features(r, i) =
// float(points_[i][r]), normals.col(i) = normals_[i].cast<float>() — done on the device, the cloud crosses PCIe once
// as it lies in Open3D's std::vector<Eigen::Vector3d>
__global__ void pack_o3d_kernel(const double* __restrict__ p3, const double* __restrict__ n3, int N,
                                float4* __restrict__ out, float4* __restrict__ outn) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    out[i] = make_float4((float)p3[3 * (size_t)i], (float)p3[3 * (size_t)i + 1], (float)p3[3 * (size_t)i + 2], 1.f);
    if (n3) outn[i] = make_float4((float)n3[3 * (size_t)i], (float)n3[3 * (size_t)i + 1], (float)n3[3 * (size_t)i + 2], 0.f);
  }
}

__global__ void flush_kernel(float4* __restrict__ buf, size_t n, float v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    buf[i] = make_float4(v, v, v, v);
}

// Partitioned map, first half of an iteration on one rank: every query the rank owns (its position under the current
// pose lies in the rank's interval [lo, hi) along `axis`) is matched against the rank's padded slab; the distance, the
// GLOBAL index, the matched point and its normal go to the exchange arrays, all other entries are zero — so that the
// sum over the ranks (one all-reduce) is the complete set of matches, bit for bit what one device holding the whole map
// would have found (a neighbour within maxDist of an owned query lies inside the padded slab; ties go to the smallest
// global index because a slab keeps the map's order). state: the resumable block state of the persistent kernel
// (its first 16 floats are T_iter), null before the first iteration (T_iter = identity).
__global__ void __launch_bounds__(256) slab_match_kernel(GridView g, const int* __restrict__ gid,
                                                         const float4* __restrict__ reading, int N, float md2,
                                                         const unsigned* __restrict__ state, int axis, float lo, float hi,
                                                         float4* __restrict__ m_pt, float4* __restrict__ m_nrm,
                                                         float* __restrict__ m_dist, int* __restrict__ m_id) {
  __shared__ float T[16], Tc[16], R9[9];
  __shared__ Warp1Scratch s_ws[256 / 32];
  if (threadIdx.x == 0) {
    if (state) for (int t = 0; t < 16; ++t) T[t] = __uint_as_float(state[t]);
    else eye4(T);
    rigid_prepare(T, Tc, R9);   // transformations.apply incl. correctParameters, as the persistent kernel does
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int nthreads = gridDim.x * blockDim.x;
  // lane l of warp slot w takes query l * nslots + w: scan neighbours (equally expensive searches) are spread over all
  // warps — on the dense C5 slabs 32 consecutive points per warp cost twice as much per pass (o3d_match_lin_kernel,
  // B200_O3D_QMAP sweep); every output is written by query index, so the map does not change a bit of the result
  const int nslots = (N + 31) >> 5;
  for (int slot = blockIdx.x * (blockDim.x >> 5) + wid; slot < nslots; slot += nthreads >> 5) {
    const int i = lane * nslots + (slot + lane * 97) % nslots;
    const bool in = i < N;
    const float4 q = in ? apply_T(Tc, __ldg(&reading[i])) : make_float4(0.f, 0.f, 0.f, 1.f);
    const float c = axis == 0 ? q.x : (axis == 1 ? q.y : q.z);
    const bool mine = in && c >= lo && c < hi;   // (a NaN coordinate belongs to nobody: no match, like the search)
    float bd; int bid, bsp;
    unsigned vis = 0;
    knn_query_warp1(g, mine, q.x, q.y, q.z, md2, bd, bid, bsp, vis, s_ws[wid]);
    if (!in) continue;
    float4 pt = make_float4(0.f, 0.f, 0.f, 0.f), nr = pt;
    float d = 0.f;
    int id = 0;
    if (mine) {
      d = bd; id = -1;
      if (bid >= 0) {
        id = gid ? __ldg(&gid[bid]) : bid;
        const float4 t = __ldg(&g.pts[bsp]);
        pt = make_float4(t.x, t.y, t.z, 0.f);
        if (g.nrm) { const float4 n = __ldg(&g.nrm[bsp]); nr = make_float4(n.x, n.y, n.z, 0.f); }
      }
    }
    m_pt[i] = pt; m_nrm[i] = nr; m_dist[i] = d; m_id[i] = id;
  }
}

}  // namespace

constexpr int kFrameFloats = 96;

struct b200_icp {
  int device = 0;
  b200_icp_params prm;
  cudaStream_t stream = nullptr;
  MapIndex map;
  bool ref_ready = false, ref_has_normals = false;
  // reading
  float4 *d_raw = nullptr, *d_raw_n = nullptr, *d_read = nullptr, *d_read_n = nullptr;
  float* d_stage3 = nullptr;
  size_t cap_read = 0;
  int N = 0;
  bool has_read_n = false;
  // work
  int *d_ids = nullptr, *d_spos = nullptr;
  float *d_dists = nullptr, *d_w = nullptr;
  size_t cap_work = 0;
  float* d_cand_d = nullptr; int* d_cand_q = nullptr; float4* d_cand_rec = nullptr;
  size_t cap_cand = 0;
  int fast_path = 1;  // env B200_ICP_FAST=0 forces the general iteration
  int warm_start = -1;  // env B200_ICP_WARM=0/1 pins the warm start of the searches; default: on for small scans (launcher)
  int qpw_log2 = -1;  // env B200_ICP_QPW=2..5 pins log2 of the queries per warp; default: chosen by the launcher from N
  int coh_log2 = -1;  // env B200_ICP_COH=0..5 pins the coherent run of query_of (icp_fused.cu); default: chosen per iteration
  unsigned* d_hist = nullptr;
  double *d_partials = nullptr, *d_partials2 = nullptr, *d_meanpart = nullptr;
  IcpDeviceResult* d_out = nullptr;
  IcpLogEntry* d_log = nullptr;
  float *d_frames = nullptr, *d_Tinit = nullptr;
  // pinned host mirrors
  IcpDeviceResult* h_out = nullptr;
  float* h_frames = nullptr;  // kFrameFloats floats
  float* h_Tinit = nullptr;
  cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  int upload_launches = 0;
  float4* d_flush = nullptr;
  size_t flush_elems = 0;
  double* d_o3d = nullptr;   // staging of Open3D double clouds (3 + 3 doubles per point)
  size_t cap_o3d = 0;
  int last_grid = 0;
  int debug_stats = 0;  // env B200_ICP_DEBUG=1
  unsigned* d_dbg_warp = nullptr;
  size_t cap_dbg_warp = 0;
  // partitioned map (b200_icp_slab_*): this handle's index holds one padded slab of the map
  bool slab = false;
  int slab_axis = 0;
  float slab_lo = 0.f, slab_hi = 0.f;      // owned interval along slab_axis, in the <refMean> frame
  int* d_gid = nullptr;                    // global index of every slab point (original order of the slab)
  size_t cap_gid = 0;
  unsigned* d_state = nullptr;             // resumable block state of the persistent kernel (kIcpStateWords)
  int slab_started = 0;                    // 0: no registration open, 1: begun, no iteration yet, 2: iterating
  int slab_iterate = 0;
  bool slab_has_gid = false;
};

namespace {

int icp_alloc_fixed(b200_icp* h) {
  B200_CUDA(cudaMalloc(&h->d_hist, ((size_t)2 * 3 * kHistBins + 2) * sizeof(unsigned)));
  B200_CUDA(cudaMalloc(&h->d_partials, icp_partials_capacity(h->device) * sizeof(double)));
  B200_CUDA(cudaMalloc(&h->d_partials2, icp_partials_capacity(h->device) * sizeof(double)));
  B200_CUDA(cudaMalloc(&h->d_meanpart, (size_t)3 * 1024 * sizeof(double)));
  B200_CUDA(cudaMalloc(&h->d_out, sizeof(IcpDeviceResult)));
  B200_CUDA(cudaMalloc(&h->d_log, (size_t)kMaxLog * sizeof(IcpLogEntry)));
  B200_CUDA(cudaMalloc(&h->d_frames, kFrameFloats * sizeof(float)));
  B200_CUDA(cudaMalloc(&h->d_Tinit, 16 * sizeof(float)));
  B200_CUDA(cudaMallocHost(&h->h_out, sizeof(IcpDeviceResult)));
  B200_CUDA(cudaMallocHost(&h->h_frames, kFrameFloats * sizeof(float)));
  B200_CUDA(cudaMallocHost(&h->h_Tinit, 16 * sizeof(float)));
  for (int i = 0; i < 5; ++i) B200_CUDA(cudaEventCreate(&h->ev[i]));
  B200_CUDA(cudaMemset(h->d_log, 0, (size_t)kMaxLog * sizeof(IcpLogEntry)));
  return B200_OK;
}

int icp_ensure_reading(b200_icp* h, int N, bool normals) {
  if ((size_t)N > h->cap_read) {
    cudaFree(h->d_raw); cudaFree(h->d_raw_n); cudaFree(h->d_read); cudaFree(h->d_read_n); cudaFree(h->d_stage3);
    h->d_raw = h->d_raw_n = h->d_read = h->d_read_n = nullptr;
    h->d_stage3 = nullptr;
    h->cap_read = 0;
    const size_t b = (size_t)N * sizeof(float4);
    B200_CUDA(cudaMalloc(&h->d_raw, b));
    B200_CUDA(cudaMalloc(&h->d_raw_n, b));
    B200_CUDA(cudaMalloc(&h->d_read, b));
    B200_CUDA(cudaMalloc(&h->d_read_n, b));
    B200_CUDA(cudaMalloc(&h->d_stage3, (size_t)N * 3 * sizeof(float)));
    h->cap_read = (size_t)N;
  }
  const size_t need = (size_t)N * (size_t)h->prm.knn;
  if (need > h->cap_work) {
    cudaFree(h->d_ids); cudaFree(h->d_spos); cudaFree(h->d_dists); cudaFree(h->d_w);
    h->d_ids = h->d_spos = nullptr;
    h->d_dists = h->d_w = nullptr;
    h->cap_work = 0;
    B200_CUDA(cudaMalloc(&h->d_ids, need * sizeof(int)));
    B200_CUDA(cudaMalloc(&h->d_spos, need * sizeof(int)));
    B200_CUDA(cudaMalloc(&h->d_dists, need * sizeof(float)));
    B200_CUDA(cudaMalloc(&h->d_w, need * sizeof(float)));
    h->cap_work = need;
  }
  const size_t cneed = icp_cand_capacity(N);
  if (cneed > h->cap_cand) {
    cudaFree(h->d_cand_d); cudaFree(h->d_cand_q); cudaFree(h->d_cand_rec);
    h->d_cand_d = nullptr; h->d_cand_q = nullptr; h->d_cand_rec = nullptr;
    h->cap_cand = 0;
    // two parities each (icp_fused.cu: double-buffered by iteration)
    B200_CUDA(cudaMalloc(&h->d_cand_d, 2 * cneed * sizeof(float)));
    B200_CUDA(cudaMalloc(&h->d_cand_q, 2 * cneed * sizeof(int)));
    B200_CUDA(cudaMalloc(&h->d_cand_rec, 2 * cneed * 2 * sizeof(float4)));
    h->cap_cand = cneed;
  }
  (void)normals;
  return B200_OK;
}

int check_params(const b200_icp_params* p) {
  if (p->knn < 1) { set_error("KDTreeMatcher: knn must be >= 1"); return B200_E_BAD_PARAM; }
  if (p->knn > 64) { set_error("ICP chain supports knn <= 64 (the limit of the matcher)"); return B200_E_UNSUPPORTED; }
  if (!(p->maxDist >= 0.f)) { set_error("KDTreeMatcher: maxDist must be >= 0"); return B200_E_BAD_PARAM; }
  if (p->use_trimmed && !(p->trimmed_ratio > 0.f && p->trimmed_ratio <= 1.f)) {
    set_error("TrimmedDistOutlierFilter: ratio must lie in (0, 1]");
    return B200_E_BAD_PARAM;
  }
  if (p->minimizer != 0 && p->minimizer != 1) { set_error("unknown error minimizer"); return B200_E_BAD_PARAM; }
  if (p->degeneracy != B200_DEGENERACY_NONE && p->degeneracy != B200_DEGENERACY_OPTIMIZED_EQUALITY) {
    set_error("degeneracyAwareness: only None and OptimizedEqualityConstraints are built");
    return B200_E_UNSUPPORTED;
  }
  if (p->degeneracy == B200_DEGENERACY_OPTIMIZED_EQUALITY && p->knn != 1) {
    set_error("degeneracyAwareness OptimizedEqualityConstraints: knn must be 1");
    return B200_E_UNSUPPORTED;
  }
  if (p->use_bound && !(p->max_rotation_norm >= 0.f && p->max_translation_norm >= 0.f)) {
    set_error("BoundTransformationChecker: maxRotationNorm / maxTranslationNorm must be >= 0");
    return B200_E_BAD_PARAM;
  }
  if (p->max_iterations <= 0 && !p->use_differential) {
    set_error("ICP chain needs at least one transformation checker");
    return B200_E_BAD_PARAM;
  }
  if (p->use_differential && (p->smooth_length < 0 || p->smooth_length > kMaxHist - 2)) {
    set_error("DifferentialTransformationChecker: smoothLength must lie in [0, 62]");
    return B200_E_UNSUPPORTED;
  }
  return B200_OK;
}

int icp_fill_params(b200_icp* h, IcpKernelParams& p);
int icp_prologue(b200_icp* h, const float* Tinit16);
int icp_result(b200_icp* h, b200_icp_result* res, bool timed_upload);

int icp_launch(b200_icp* h, const float* Tinit16, b200_icp_result* res, bool timed_upload) {
  if (h->slab) { set_error("this handle holds one slab of a partitioned map: use b200_icp_slab_*"); return B200_E_BAD_PARAM; }
  cudaStream_t s = h->stream;
  B200_TRY(icp_prologue(h, Tinit16));
  IcpKernelParams p;
  B200_TRY(icp_fill_params(h, p));
  B200_CUDA(cudaEventRecord(h->ev[2], s));
  B200_TRY(launch_icp_kernel(p, h->device, s, &h->last_grid));
  B200_CUDA(cudaEventRecord(h->ev[3], s));
  B200_CUDA(cudaMemcpyAsync(h->h_out, h->d_out, sizeof(IcpDeviceResult), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaMemcpyAsync(h->h_frames, h->d_frames, kFrameFloats * sizeof(float), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaEventRecord(h->ev[4], s));
  B200_CUDA(cudaStreamSynchronize(s));
  return icp_result(h, res, timed_upload);
}

// mean of the reading, frames, reading in the <refMean> frame; work buffers cleared (one registration = one prologue)
int icp_prologue(b200_icp* h, const float* Tinit16) {
  cudaStream_t s = h->stream;
  const int N = h->N;
  if (Tinit16) std::memcpy(h->h_Tinit, Tinit16, 16 * sizeof(float));
  else eye4(h->h_Tinit);
  B200_CUDA(cudaMemcpyAsync(h->d_Tinit, h->h_Tinit, 16 * sizeof(float), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  const int mb = std::min(kMeanBlocks, std::max(1, (N + kMeanThreads - 1) / kMeanThreads));
  read_sum_kernel<<<mb, kMeanThreads, 0, s>>>(h->d_raw, N, h->d_meanpart);
  frames_kernel<<<1, 32, 0, s>>>(h->d_meanpart, mb, N, h->d_Tinit, h->d_frames,
                                 h->prm.degeneracy == B200_DEGENERACY_OPTIMIZED_EQUALITY ? 1 : 0);
  const int pb = std::max(1, std::min(148 * 8, (N + 255) / 256));
  prepare_reading_kernel<<<pb, 256, 0, s>>>(h->d_raw, h->has_read_n ? h->d_raw_n : nullptr, N, h->d_frames, h->d_read,
                                            h->d_read_n);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaMemsetAsync(h->d_hist, 0, ((size_t)2 * 3 * kHistBins + 2) * sizeof(unsigned), s));
  B200_CUDA(cudaMemsetAsync(h->d_out, 0, sizeof(IcpDeviceResult), s));
  if (h->debug_stats) B200_CUDA(cudaMemsetAsync(h->d_log, 0, (size_t)kMaxLog * sizeof(IcpLogEntry), s));
  return B200_OK;
}

int icp_fill_params(b200_icp* h, IcpKernelParams& p) {
  const int N = h->N;
  std::memset(&p, 0, sizeof(p));
  p.g = h->map.view();
  p.reading = h->d_read;
  p.reading_n = h->has_read_n ? h->d_read_n : nullptr;
  p.N = N; p.k = h->prm.knn;
  p.md2 = std::isinf(h->prm.maxDist) ? INFINITY : h->prm.maxDist * h->prm.maxDist;
  p.use_trimmed = h->prm.use_trimmed; p.trimmed_ratio = h->prm.trimmed_ratio;
  p.use_maxdist = h->prm.use_maxdist_filter; p.maxdist_lim2 = (float)std::pow(h->prm.maxdist_filter, 2);
  p.use_sn = h->prm.use_surface_normal_filter; p.sn_cos = std::cos(h->prm.surface_normal_max_angle);
  p.minimizer = h->prm.minimizer;
  p.max_iterations = h->prm.max_iterations; p.use_differential = h->prm.use_differential;
  p.min_rot = h->prm.min_diff_rot; p.min_trans = h->prm.min_diff_trans; p.smooth = h->prm.smooth_length;
  p.use_bound = h->prm.use_bound; p.bound_rot = h->prm.max_rotation_norm; p.bound_trans = h->prm.max_translation_norm;
  p.ids = h->d_ids; p.dists = h->d_dists; p.spos = h->d_spos; p.weights = h->d_w;
  p.cand_d = h->d_cand_d; p.cand_q = h->d_cand_q; p.cand_rec = h->d_cand_rec; p.cand_seg = 0;
  p.fast_path = (h->fast_path && h->prm.knn == 1) ? 1 : 0;
  p.coh_log2 = h->coh_log2;
  p.qpw_log2 = h->qpw_log2;
  p.warm_start = h->warm_start;   // < 0: the launcher decides
  p.debug_stats = h->debug_stats;
  p.dbg_warp = nullptr;
  if (h->debug_stats) {
    const size_t need = (size_t)((N + 3) / 4) * 4;   // one entry per warp slot, at most N / 4 slots
    if (need > h->cap_dbg_warp) {
      cudaFree(h->d_dbg_warp);
      h->d_dbg_warp = nullptr;
      B200_CUDA(cudaMalloc(&h->d_dbg_warp, need * sizeof(unsigned)));
      h->cap_dbg_warp = need;
    }
    p.dbg_warp = h->d_dbg_warp;
  }
  p.xicp = h->prm.degeneracy == B200_DEGENERACY_OPTIMIZED_EQUALITY ? 1 : 0;
  p.x_enough = h->prm.xicp_enough_information; p.x_insufficient = h->prm.xicp_insufficient_information;
  // ICP.cpp:648-665: cos(convertAngleInDegreesToRadians(angle)) in float
  p.x_cos_min = std::cos(h->prm.xicp_min_alignment_angle_deg * (float)M_PI / 180.f);
  p.x_cos_strong = std::cos(h->prm.xicp_strong_alignment_angle_deg * (float)M_PI / 180.f);
  p.partials2 = h->d_partials2;
  p.list_count = h->d_hist + (size_t)2 * 3 * kHistBins;
  p.hist = h->d_hist; p.partials = h->d_partials; p.out = h->d_out; p.log = h->d_log; p.frames = h->d_frames;
  return B200_OK;
}

int icp_result(b200_icp* h, b200_icp_result* res, bool timed_upload) {
  if (res) {
    std::memset(res, 0, sizeof(*res));
    const IcpDeviceResult& o = *h->h_out;
    res->status = o.status; res->iterations = o.iterations;
    res->max_iterations_reached = o.max_iterations_reached; res->n_matches_last = o.n_matches_last;
    std::memcpy(res->T, o.T, sizeof(o.T));
    std::memcpy(res->T_iter, o.T_iter, sizeof(o.T_iter));
    std::memcpy(res->T_refIn_refMean, h->h_frames, 16 * sizeof(float));
    std::memcpy(res->T_refMean_readMean, h->h_frames + 16, 16 * sizeof(float));
    res->weighted_point_used_ratio = o.weighted_point_used_ratio;
    res->point_used_ratio = o.point_used_ratio;
    res->points_touched = o.points_touched;
    res->returned_prior = o.returned_prior; res->xicp_borderline = o.xicp_borderline;
    for (int i = 0; i < 3; ++i) { res->loc_xyz[i] = o.loc_xyz[i]; res->loc_rpy[i] = o.loc_rpy[i]; }
    std::memcpy(res->eig_t, o.eig_t, sizeof(o.eig_t));
    std::memcpy(res->eig_r, o.eig_r, sizeof(o.eig_r));
    std::memcpy(res->xicp_sums, o.xicp_sums, sizeof(o.xicp_sums));
    if (o.returned_prior) std::memcpy(res->T, h->h_Tinit, 16 * sizeof(float));   // ICP.cpp:1335-1341
    res->grid_ctas = h->last_grid;
    B200_CUDA(cudaEventElapsedTime(&res->device_ms, h->ev[2], h->ev[3]));
    float pre = 0.f;
    B200_CUDA(cudaEventElapsedTime(&pre, h->ev[1], h->ev[3]));
    if (timed_upload) {
      B200_CUDA(cudaEventElapsedTime(&res->h2d_ms, h->ev[0], h->ev[1]));
      B200_CUDA(cudaEventElapsedTime(&res->total_ms, h->ev[0], h->ev[3]));
      B200_CUDA(cudaEventElapsedTime(&res->e2e_ms, h->ev[0], h->ev[4]));
      res->kernel_launches = 4 + h->upload_launches;  // centroid sums, frames, reading preparation, registration
    } else {
      res->kernel_launches = 4;
      res->h2d_ms = 0.f;
      res->total_ms = pre;
    }
  }
  const int st = h->h_out->status;
  if (st == B200_E_NO_MATCHES) set_error("ConvergenceError: no point to minimize / no matches");
  else if (st == B200_E_NAN) set_error("ConvergenceError: abs rotation norm not a number");
  else if (st == B200_E_BOUND) set_error("ConvergenceError: limit out of bounds (BoundTransformationChecker)");
  return st;
}

int icp_upload(b200_icp* h, const float* read4xN, const float* readn3xN, int N) {
  B200_TRY(icp_ensure_reading(h, N, readn3xN != nullptr));
  B200_CUDA(cudaEventRecord(h->ev[0], h->stream));
  B200_TRY(upload_features(h->d_raw, read4xN, N, h->stream));
  if (readn3xN) B200_TRY(upload_normals(h->d_raw_n, readn3xN, N, h->d_stage3, h->stream));
  h->upload_launches = readn3xN ? 1 : 0;  // normals repack kernel
  h->N = N;
  h->has_read_n = readn3xN != nullptr;
  return B200_OK;
}

}  // namespace

extern "C" {

void b200_icp_default_params(b200_icp_params* p) {
  if (!p) return;
  std::memset(p, 0, sizeof(*p));
  p->knn = 1; p->epsilon = 0.f; p->maxDist = INFINITY;            // MatchersImpl.h:83-86
  p->use_trimmed = 1; p->trimmed_ratio = 0.85f;                    // OutlierFiltersImpl.h:137
  p->use_maxdist_filter = 0; p->maxdist_filter = 1.f;
  p->use_surface_normal_filter = 0; p->surface_normal_max_angle = 1.57f;
  p->minimizer = 0;
  p->max_iterations = 40;                                          // TransformationCheckersImpl.h:70
  p->use_differential = 1; p->min_diff_rot = 0.001f; p->min_diff_trans = 0.001f; p->smooth_length = 3;  // :90-92
  p->cell_size = 0.f;
  // degeneracyAwareness absent => kNone (ICP.cpp:586-590); the thresholds below are param/icp.yaml's
  p->degeneracy = B200_DEGENERACY_NONE;
  p->xicp_enough_information = 250.f; p->xicp_insufficient_information = 180.f;
  p->xicp_min_alignment_angle_deg = 80.f; p->xicp_strong_alignment_angle_deg = 45.f;
  p->use_bound = 0; p->max_rotation_norm = 1.f; p->max_translation_norm = 1.f;   // TransformationCheckersImpl.h:117-123
}

int b200_icp_create(b200_icp** out, int device, const b200_icp_params* prm) {
  if (!out || !prm) { set_error("b200_icp_create: bad arguments"); return B200_E_BAD_PARAM; }
  *out = nullptr;
  B200_TRY(check_params(prm));
  DeviceGuard dg(device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  int coop = 0;
  if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device) != cudaSuccess || !coop) {
    cudaGetLastError();
    set_error("device does not support cooperative launch");
    return B200_E_CUDA;
  }
  b200_icp* h = new (std::nothrow) b200_icp();
  if (!h) { set_error("out of host memory"); return B200_E_CUDA; }
  h->device = device;
  h->prm = *prm;
  { const char* e = getenv("B200_ICP_DEBUG"); h->debug_stats = (e && e[0] == '1') ? 1 : 0; }
  { const char* e = getenv("B200_ICP_FAST"); h->fast_path = (e && e[0] == '0') ? 0 : 1; }
  { const char* e = getenv("B200_ICP_COH"); if (e && e[0] >= '0' && e[0] <= '5') h->coh_log2 = e[0] - '0'; }
  { const char* e = getenv("B200_ICP_WARM"); if (e && (e[0] == '0' || e[0] == '1')) h->warm_start = e[0] - '0'; }
  { const char* e = getenv("B200_ICP_QPW"); if (e && e[0] >= '2' && e[0] <= '5') h->qpw_log2 = e[0] - '0'; }
  h->map.device = device;
  int st = B200_OK;
  if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess)
    st = cuda_fail(cudaGetLastError(), "cudaStreamCreate", __FILE__, __LINE__);
  h->map.stream = h->stream;
  if (!st) st = icp_alloc_fixed(h);
  if (st) { b200_icp_destroy(h); return st; }
  *out = h;
  return B200_OK;
}

void b200_icp_destroy(b200_icp* h) {
  if (!h) return;
  {
    DeviceGuard dg(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->d_raw); cudaFree(h->d_raw_n); cudaFree(h->d_read); cudaFree(h->d_read_n); cudaFree(h->d_stage3);
    cudaFree(h->d_ids); cudaFree(h->d_spos); cudaFree(h->d_dists); cudaFree(h->d_w);
    cudaFree(h->d_cand_d); cudaFree(h->d_cand_q); cudaFree(h->d_cand_rec);
    cudaFree(h->d_hist); cudaFree(h->d_partials); cudaFree(h->d_partials2); cudaFree(h->d_meanpart); cudaFree(h->d_out); cudaFree(h->d_log);
    cudaFree(h->d_frames); cudaFree(h->d_Tinit); cudaFree(h->d_flush); cudaFree(h->d_dbg_warp); cudaFree(h->d_o3d);
    cudaFree(h->d_gid); cudaFree(h->d_state);
    cudaFreeHost(h->h_out); cudaFreeHost(h->h_frames); cudaFreeHost(h->h_Tinit);
    for (int i = 0; i < 5; ++i)
      if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    h->map.release();
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
}

int b200_icp_init_reference(b200_icp* h, const float* ref4xM, const float* refn3xM, int M) {
  if (!h || M < 0 || (M > 0 && !ref4xM)) { set_error("b200_icp_init_reference: bad arguments"); return B200_E_BAD_PARAM; }
  h->ref_ready = false;
  if (M == 0) {  // "The reference point cloud is empty" — initReference returns false, ICP.cpp:850-855
    set_error("The reference point cloud is empty");
    return B200_E_NOT_INITIALIZED;
  }
  if (h->prm.minimizer == 0 && !refn3xM) {  // InvalidField at getDescriptorViewByName("normals"), ICP.cpp:1519
    set_error("PointToPlaneErrorMinimizer needs the 'normals' descriptor on the reference");
    return B200_E_BAD_PARAM;
  }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  B200_TRY(index_alloc_points(h->map, M, refn3xM != nullptr));
  B200_TRY(upload_features(h->map.d_ref, ref4xM, M, s));
  if (refn3xM) {
    float* stage = nullptr;
    B200_CUDA(cudaMallocAsync(&stage, (size_t)M * 3 * sizeof(float), s));
    const int st = upload_normals(h->map.d_nrm_orig, refn3xM, M, stage, s);
    cudaFreeAsync(stage, s);
    if (st) return st;
  }
  return icp_reference_from_device(h, M, refn3xM != nullptr);
}

}  // extern "C"

namespace b200 {

// buffers of the handle's index for a reference of M points that the caller fills ON THE DEVICE (map_maint.cu)
float4* icp_reference_points(b200_icp* h, int M, bool with_normals, float4** normals_out, cudaStream_t* stream_out) {
  DeviceGuard dg(h->device);
  h->ref_ready = false;
  if (index_alloc_points(h->map, M, with_normals) != B200_OK) return nullptr;
  if (normals_out) *normals_out = with_normals ? h->map.d_nrm_orig : nullptr;
  if (stream_out) *stream_out = h->stream;
  return h->map.d_ref;
}

// the rest of ICP::initReference (ICP.cpp:847-898) on a reference resident in h->map.d_ref / d_nrm_orig:
// mean-centre, build the index, store T_refIn_refMean
int icp_reference_from_device(b200_icp* h, int M, bool has_normals, const double* mean_given) {
  if (h->prm.minimizer == 0 && !has_normals) {  // InvalidField at getDescriptorViewByName("normals"), ICP.cpp:1519
    set_error("PointToPlaneErrorMinimizer needs the 'normals' descriptor on the reference");
    return B200_E_BAD_PARAM;
  }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  h->ref_ready = false;
  h->slab = false;
  h->ref_has_normals = has_normals;
  // centroid in double, rounded to float; reference expressed in <refMean> (ICP.cpp:883-890)
  double mean[3];
  if (mean_given) { mean[0] = mean_given[0]; mean[1] = mean_given[1]; mean[2] = mean_given[2]; }   // (a slab: the whole map's)
  else B200_TRY(device_mean3(h->map.d_ref, M, h->d_meanpart, mean, s));
  const float mf[3] = {(float)mean[0], (float)mean[1], (float)mean[2]};
  B200_TRY(device_subtract3(h->map.d_ref, M, mf[0], mf[1], mf[2], s));
  b200_index_opts opts;
  std::memset(&opts, 0, sizeof(opts));
  opts.cell_size = h->prm.cell_size;
  B200_TRY(index_build(h->map, &opts));
  float frames[kFrameFloats];
  std::memset(frames, 0, sizeof(frames));
  eye4(frames);
  frames[12] = mf[0]; frames[13] = mf[1]; frames[14] = mf[2];
  std::memcpy(h->h_frames, frames, sizeof(frames));
  B200_CUDA(cudaMemcpyAsync(h->d_frames, h->h_frames, sizeof(frames), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaStreamSynchronize(s));
  h->ref_ready = true;
  return B200_OK;
}

}  // namespace b200

extern "C" {

int b200_icp_compute(b200_icp* h, const float* read4xN, const float* readn3xN, int N, const float* Tinit16,
                     b200_icp_result* res) {
  if (!h || N < 0 || (N > 0 && !read4xN)) { set_error("b200_icp_compute: bad arguments"); return B200_E_BAD_PARAM; }
  if (!h->ref_ready) { set_error("You must setup a matcher before running ICP (no reference)"); return B200_E_NOT_INITIALIZED; }
  if (N == 0) { set_error("The reading point cloud is empty."); return B200_E_BAD_PARAM; }  // ICP.cpp:958-960
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  B200_TRY(icp_upload(h, read4xN, readn3xN, N));
  return icp_launch(h, Tinit16, res, true);
}

static int icp_stage_o3d(b200_icp* h, const double* p3, const double* n3, int N, float4* out, float4* outn) {
  if ((size_t)N > h->cap_o3d) {
    cudaFree(h->d_o3d);
    h->d_o3d = nullptr; h->cap_o3d = 0;
    B200_CUDA(cudaMalloc(&h->d_o3d, (size_t)N * 6 * sizeof(double)));
    h->cap_o3d = (size_t)N;
  }
  cudaStream_t s = h->stream;
  double* dn = n3 ? h->d_o3d + 3 * (size_t)N : nullptr;
  B200_CUDA(cudaMemcpyAsync(h->d_o3d, p3, (size_t)N * 3 * sizeof(double), cudaMemcpyHostToDevice, s));
  if (n3) B200_CUDA(cudaMemcpyAsync(dn, n3, (size_t)N * 3 * sizeof(double), cudaMemcpyHostToDevice, s));
  pack_o3d_kernel<<<std::max(1, std::min(148 * 8, (N + 255) / 256)), 256, 0, s>>>(h->d_o3d, dn, N, out, outn);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int b200_icp_compute_o3d(b200_icp* h, const double* pts3xN, const double* nrm3xN, int N, const float* Tinit16,
                         b200_icp_result* res) {
  if (!h || N < 0 || (N > 0 && !pts3xN)) { set_error("b200_icp_compute_o3d: bad arguments"); return B200_E_BAD_PARAM; }
  if (!h->ref_ready) { set_error("You must setup a matcher before running ICP (no reference)"); return B200_E_NOT_INITIALIZED; }
  if (N == 0) { set_error("The reading point cloud is empty."); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  B200_TRY(icp_ensure_reading(h, N, nrm3xN != nullptr));
  B200_CUDA(cudaEventRecord(h->ev[0], h->stream));
  B200_TRY(icp_stage_o3d(h, pts3xN, nrm3xN, N, h->d_raw, h->d_raw_n));
  h->upload_launches = 1;
  h->N = N;
  h->has_read_n = nrm3xN != nullptr;
  return icp_launch(h, Tinit16, res, true);
}

int b200_icp_init_reference_o3d(b200_icp* h, const double* pts3xM, const double* nrm3xM, int M) {
  if (!h || M < 0 || (M > 0 && !pts3xM)) { set_error("b200_icp_init_reference_o3d: bad arguments"); return B200_E_BAD_PARAM; }
  h->ref_ready = false;
  if (M == 0) { set_error("The reference point cloud is empty"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  B200_TRY(index_alloc_points(h->map, M, nrm3xM != nullptr));
  B200_TRY(icp_stage_o3d(h, pts3xM, nrm3xM, M, h->map.d_ref, h->map.d_nrm_orig));
  return icp_reference_from_device(h, M, nrm3xM != nullptr);
}

int b200_icp_upload_reading(b200_icp* h, const float* read4xN, const float* readn3xN, int N) {
  if (!h || N <= 0 || !read4xN) { set_error("b200_icp_upload_reading: bad arguments"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  B200_TRY(icp_upload(h, read4xN, readn3xN, N));
  B200_CUDA(cudaStreamSynchronize(h->stream));
  return B200_OK;
}

int b200_icp_run_resident(b200_icp* h, const float* Tinit16, b200_icp_result* res) {
  if (!h) { set_error("null handle"); return B200_E_BAD_PARAM; }
  if (!h->ref_ready || h->N <= 0) { set_error("reference or reading not resident"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  return icp_launch(h, Tinit16, res, false);
}

int b200_icp_get_matches(b200_icp* h, int32_t* ids, float* dists, float* weights) {
  if (!h || h->N <= 0) { set_error("no registration has run"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  const size_t n = (size_t)h->N * h->prm.knn;
  if (ids) B200_CUDA(cudaMemcpyAsync(ids, h->d_ids, n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
  if (dists) B200_CUDA(cudaMemcpyAsync(dists, h->d_dists, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
  if (weights) B200_CUDA(cudaMemcpyAsync(weights, h->d_w, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
  B200_CUDA(cudaStreamSynchronize(h->stream));
  return B200_OK;
}

int b200_icp_get_log(b200_icp* h, int max_log, float* H, float* b, float* x, float* T_iter, int32_t* P, int* n) {
  if (!h || !n) { set_error("b200_icp_get_log: bad arguments"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  IcpLogEntry* tmp = new (std::nothrow) IcpLogEntry[kMaxLog];
  if (!tmp) { set_error("out of host memory"); return B200_E_CUDA; }
  // on the handle's own stream: a blocking cudaMemcpy would go through the legacy default stream and serialise
  // against other handles' work (handles are re-entrant across threads, b200reg.h)
  if (cudaMemcpyAsync(tmp, h->d_log, (size_t)kMaxLog * sizeof(IcpLogEntry), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
      cudaStreamSynchronize(h->stream) != cudaSuccess) {
    delete[] tmp;
    return cuda_fail(cudaGetLastError(), "cudaMemcpyAsync(log)", __FILE__, __LINE__);
  }
  const int cnt = std::min(std::min(max_log, kMaxLog), h->h_out ? h->h_out->iterations : 0);
  for (int i = 0; i < cnt; ++i) {
    if (H) std::memcpy(H + 36 * (size_t)i, tmp[i].H, sizeof(tmp[i].H));
    if (b) std::memcpy(b + 6 * (size_t)i, tmp[i].b, sizeof(tmp[i].b));
    if (x) std::memcpy(x + 6 * (size_t)i, tmp[i].x, sizeof(tmp[i].x));
    if (T_iter) std::memcpy(T_iter + 16 * (size_t)i, tmp[i].T_iter, sizeof(tmp[i].T_iter));
    if (P) P[i] = tmp[i].P;
  }
  *n = cnt;
  delete[] tmp;
  return B200_OK;
}

int b200_icp_get_phase_times(b200_icp* h, int max_log, float* us, int* n) {
  if (!h || !n || !us) { set_error("b200_icp_get_phase_times: bad arguments"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  IcpLogEntry* tmp = new (std::nothrow) IcpLogEntry[kMaxLog];
  if (!tmp) { set_error("out of host memory"); return B200_E_CUDA; }
  // on the handle's own stream: a blocking cudaMemcpy would go through the legacy default stream and serialise
  // against other handles' work (handles are re-entrant across threads, b200reg.h)
  if (cudaMemcpyAsync(tmp, h->d_log, (size_t)kMaxLog * sizeof(IcpLogEntry), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
      cudaStreamSynchronize(h->stream) != cudaSuccess) {
    delete[] tmp;
    return cuda_fail(cudaGetLastError(), "cudaMemcpyAsync(log)", __FILE__, __LINE__);
  }
  const int cnt = std::min(std::min(max_log, kMaxLog), h->h_out ? h->h_out->iterations : 0);
  for (int i = 0; i < cnt; ++i) {
    for (int t = 0; t < 5; ++t) us[6 * (size_t)i + t] = (float)((double)(tmp[i].ts[t + 1] - tmp[i].ts[t]) * 1e-3);
    us[6 * (size_t)i + 5] = (i + 1 < cnt) ? (float)((double)(tmp[i + 1].ts[0] - tmp[i].ts[0]) * 1e-3)
                                          : (float)((double)(tmp[i].ts[5] - tmp[i].ts[0]) * 1e-3);
  }
  *n = cnt;
  delete[] tmp;
  return B200_OK;
}

int b200_icp_get_debug_stats(b200_icp* h, int max_log, double* st, int* n) {
  if (!h || !n || !st) { set_error("b200_icp_get_debug_stats: bad arguments"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  IcpLogEntry* tmp = new (std::nothrow) IcpLogEntry[kMaxLog];
  if (!tmp) { set_error("out of host memory"); return B200_E_CUDA; }
  // on the handle's own stream: a blocking cudaMemcpy would go through the legacy default stream and serialise
  // against other handles' work (handles are re-entrant across threads, b200reg.h)
  if (cudaMemcpyAsync(tmp, h->d_log, (size_t)kMaxLog * sizeof(IcpLogEntry), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
      cudaStreamSynchronize(h->stream) != cudaSuccess) {
    delete[] tmp;
    return cuda_fail(cudaGetLastError(), "cudaMemcpyAsync(log)", __FILE__, __LINE__);
  }
  const int cnt = std::min(std::min(max_log, kMaxLog), h->h_out ? h->h_out->iterations : 0);
  for (int i = 0; i < cnt; ++i) {
    st[12 * (size_t)i + 0] = (double)tmp[i].dbg_rows;
    st[12 * (size_t)i + 1] = (double)tmp[i].dbg_cands;
    st[12 * (size_t)i + 2] = (double)tmp[i].dbg_match_ns_max * 1e-3;
    st[12 * (size_t)i + 3] = (double)tmp[i].dbg_match_ns_sum * 1e-3 / std::max(1, h->last_grid);
    for (int t = 0; t < 4; ++t) st[12 * (size_t)i + 4 + t] = (double)tmp[i].sub_ns[t] * 1e-3;
    st[12 * (size_t)i + 8] = (double)tmp[i].path;
    st[12 * (size_t)i + 9] = (double)tmp[i].n_cand;
    st[12 * (size_t)i + 10] = (double)tmp[i].limit;
    st[12 * (size_t)i + 11] = (double)tmp[i].delta;
  }
  *n = cnt;
  delete[] tmp;
  return B200_OK;
}

int b200_icp_get_debug_warps(b200_icp* h, uint32_t* out, int max_slots, int* n) {
  if (!h || !out || !n) { set_error("b200_icp_get_debug_warps: bad arguments"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  const int slots = std::min(max_slots, (h->N + 31) / 32);
  *n = 0;
  if (!h->d_dbg_warp || slots <= 0) return B200_OK;
  B200_CUDA(cudaMemcpyAsync(out, h->d_dbg_warp, (size_t)slots * 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
  B200_CUDA(cudaStreamSynchronize(h->stream));
  *n = slots;
  return B200_OK;
}

int b200_icp_flush_l2(b200_icp* h) {
  if (!h) { set_error("null handle"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!h->d_flush) {
    h->flush_elems = ((size_t)256 << 20) / sizeof(float4);  // 256 MiB > 126 MB L2
    B200_CUDA(cudaMalloc(&h->d_flush, h->flush_elems * sizeof(float4)));
  }
  flush_kernel<<<148 * 8, 256, 0, h->stream>>>(h->d_flush, h->flush_elems, 1.0f);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// ---- partitioned map (BASELINE C5) with the libpointmatcher chain ----------------------------------------
// One handle per rank holds one padded slab of the map. An iteration is: b200_icp_slab_match on every rank (searches
// for the queries the rank owns, zeros elsewhere) -> the caller sums the exchange arrays over the ranks (ONE all-reduce
// of 9 N floats and one of N int32; NCCL or a plain add when several slabs share a device) -> b200_icp_slab_iterate on
// every rank: the persistent kernel, told to take the matches as given, runs the rest of the iteration (outlier chain
// incl. TrimmedDist's global quantile, ErrorElements, H/b, solve, pose update, checkers) and stops; its block state
// (pose, checker history, ...) stays on the device for the next launch. Every rank computes the same pose from the
// same sums: no broadcast. The result equals the un-partitioned b200_icp_compute bit for bit (tests/test_gpu_slab_pm.py).

int b200_icp_init_reference_slab(b200_icp* h, const float* ref4xM, const float* refn3xM, int M, const int32_t* gid_M,
                                 const double* map_mean3, int axis, float lo, float hi) {
  if (!h || M <= 0 || !ref4xM || !map_mean3 || axis < 0 || axis > 2 || !(lo < hi)) {
    set_error("b200_icp_init_reference_slab: bad arguments");
    return B200_E_BAD_PARAM;
  }
  if (h->prm.knn != 1 || h->prm.degeneracy != B200_DEGENERACY_NONE) {
    set_error("partitioned map: knn = 1 and degeneracyAwareness None only");
    return B200_E_UNSUPPORTED;
  }
  if (h->prm.minimizer == 0 && !refn3xM) {
    set_error("PointToPlaneErrorMinimizer needs the 'normals' descriptor on the reference");
    return B200_E_BAD_PARAM;
  }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  h->ref_ready = false;
  B200_TRY(index_alloc_points(h->map, M, refn3xM != nullptr));
  B200_TRY(upload_features(h->map.d_ref, ref4xM, M, s));
  if (refn3xM) {
    float* stage = nullptr;
    B200_CUDA(cudaMallocAsync(&stage, (size_t)M * 3 * sizeof(float), s));
    const int st = upload_normals(h->map.d_nrm_orig, refn3xM, M, stage, s);
    cudaFreeAsync(stage, s);
    if (st) return st;
  }
  if ((size_t)M > h->cap_gid) {
    cudaFree(h->d_gid);
    h->d_gid = nullptr; h->cap_gid = 0;
    B200_CUDA(cudaMalloc(&h->d_gid, (size_t)M * sizeof(int)));
    h->cap_gid = (size_t)M;
  }
  if (gid_M) B200_CUDA(cudaMemcpyAsync(h->d_gid, gid_M, (size_t)M * sizeof(int), cudaMemcpyHostToDevice, s));
  if (!h->d_state) B200_CUDA(cudaMalloc(&h->d_state, (size_t)kIcpStateWords * sizeof(unsigned)));
  B200_TRY(icp_reference_from_device(h, M, refn3xM != nullptr, map_mean3));
  // the interval in the frame the queries live in (<refMean>): the same float subtraction on every rank
  const float mf = (float)map_mean3[axis];
  h->slab = true;
  h->slab_axis = axis;
  h->slab_lo = std::isinf(lo) ? lo : lo - mf;
  h->slab_hi = std::isinf(hi) ? hi : hi - mf;
  h->slab_started = 0;
  h->slab_has_gid = gid_M != nullptr;
  return B200_OK;
}

int b200_icp_slab_begin(b200_icp* h, const float* read4xN, const float* readn3xN, int N, const float* Tinit16) {
  if (!h || N <= 0 || !read4xN) { set_error("b200_icp_slab_begin: bad arguments"); return B200_E_BAD_PARAM; }
  if (!h->ref_ready || !h->slab) { set_error("b200_icp_slab_begin: no slab reference (b200_icp_init_reference_slab)"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  B200_TRY(icp_upload(h, read4xN, readn3xN, N));
  B200_TRY(icp_prologue(h, Tinit16));
  h->slab_started = 1;
  h->slab_iterate = 1;
  return B200_OK;
}

static int slab_match_impl(b200_icp* h, float* d_match9xN, int32_t* d_idN, bool sync);
int b200_icp_slab_match(b200_icp* h, float* d_match9xN, int32_t* d_idN) { return slab_match_impl(h, d_match9xN, d_idN, true); }
int b200_icp_slab_match_async(b200_icp* h, float* d_match9xN, int32_t* d_idN) { return slab_match_impl(h, d_match9xN, d_idN, false); }
void* b200_icp_slab_stream(b200_icp* h) { return h ? (void*)h->stream : nullptr; }

static int slab_match_impl(b200_icp* h, float* d_match9xN, int32_t* d_idN, bool sync) {
  if (!h || !d_match9xN || !d_idN) { set_error("b200_icp_slab_match: bad arguments"); return B200_E_BAD_PARAM; }
  if (!h->slab || h->slab_started == 0) { set_error("b200_icp_slab_match: b200_icp_slab_begin first"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  const int N = h->N;
  float4* m_pt = reinterpret_cast<float4*>(d_match9xN);
  float4* m_nrm = m_pt + N;
  float* m_dist = d_match9xN + (size_t)8 * N;
  const float md2 = std::isinf(h->prm.maxDist) ? INFINITY : h->prm.maxDist * h->prm.maxDist;
  const int blocks = std::max(1, std::min(148 * 8, (N + 255) / 256));
  slab_match_kernel<<<blocks, 256, 0, h->stream>>>(h->map.view(), h->slab_has_gid ? h->d_gid : nullptr, h->d_read, N, md2,
                                                  h->slab_started == 2 ? h->d_state : nullptr, h->slab_axis, h->slab_lo,
                                                  h->slab_hi, m_pt, m_nrm, m_dist, d_idN);
  B200_CUDA(cudaGetLastError());
  if (sync) B200_CUDA(cudaStreamSynchronize(h->stream));   // the caller's exchange runs on its own stream
  return B200_OK;
}

int b200_icp_slab_iterate(b200_icp* h, const float* d_match9xN, const int32_t* d_idN, int* iterate) {
  if (!h || !d_match9xN || !d_idN || !iterate) { set_error("b200_icp_slab_iterate: bad arguments"); return B200_E_BAD_PARAM; }
  if (!h->slab || h->slab_started == 0) { set_error("b200_icp_slab_iterate: b200_icp_slab_begin first"); return B200_E_NOT_INITIALIZED; }
  if (!h->slab_iterate) { set_error("b200_icp_slab_iterate: the registration has ended"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  const int N = h->N;
  IcpKernelParams p;
  B200_TRY(icp_fill_params(h, p));
  p.forced = 1;
  p.forced_dist = d_match9xN + (size_t)8 * N;
  p.forced_id = d_idN;
  p.g.pts = reinterpret_cast<const float4*>(d_match9xN);
  p.g.nrm = h->ref_has_normals ? reinterpret_cast<const float4*>(d_match9xN) + N : nullptr;
  p.max_launch_iters = 1;
  p.resume = h->slab_started == 2 ? 1 : 0;
  p.state = h->d_state;
  B200_CUDA(cudaMemsetAsync(h->d_hist, 0, ((size_t)2 * 3 * kHistBins + 2) * sizeof(unsigned), s));
  B200_TRY(launch_icp_kernel(p, h->device, s, &h->last_grid));
  B200_CUDA(cudaMemcpyAsync(h->h_out, h->d_out, sizeof(IcpDeviceResult), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  h->slab_started = 2;
  h->slab_iterate = h->h_out->iterate;
  *iterate = h->slab_iterate;
  return B200_OK;
}

int b200_icp_slab_finish(b200_icp* h, b200_icp_result* res) {
  if (!h || !res) { set_error("b200_icp_slab_finish: bad arguments"); return B200_E_BAD_PARAM; }
  if (!h->slab || h->slab_started != 2) { set_error("b200_icp_slab_finish: no iteration has run"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  cudaStream_t s = h->stream;
  B200_CUDA(cudaEventRecord(h->ev[2], s));
  B200_CUDA(cudaEventRecord(h->ev[3], s));
  B200_CUDA(cudaMemcpyAsync(h->h_out, h->d_out, sizeof(IcpDeviceResult), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaMemcpyAsync(h->h_frames, h->d_frames, kFrameFloats * sizeof(float), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaEventRecord(h->ev[4], s));
  B200_CUDA(cudaStreamSynchronize(s));
  h->slab_started = 0;
  return icp_result(h, res, false);
}

}  // extern "C"
