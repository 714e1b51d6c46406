// Loop-closure front half of open3d_slam on the device:
//   Submap::computeFeatures (open3d_slam/src/Submap.cpp:255-275)
//       registration::ComputeFPFHFeature(sparseMapCloud_, KDTreeSearchParamHybrid(featureRadius_, featureKnn_))
//   PlaceRecognition::buildLoopClosureConstraints (open3d_slam/src/PlaceRecognition.cpp:56-59, 81-85)
//       RegistrationRANSACBasedOnFeatureMatching(sourceSparse, targetSparse, sourceFeature, targetFeature, true,
//           ransacMaxCorrespondenceDistance_, TransformationEstimationPointToPoint(false), ransacModelSize_,
//           {CorrespondenceCheckerBasedOnDistance, CorrespondenceCheckerBasedOnEdgeLength},
//           RANSACConvergenceCriteria(ransacNumIter_, ransacProbability_))
// Open3D v0.15.1 is not in the reference tree; semantics recalled from upstream (pipelines/registration/Feature.cpp,
// Registration.cpp, CorrespondenceChecker.cpp; parity unpinned; the test suite's CPU checker restates the same
// semantics independently), with these stated deviations: float-converted clouds, float neighbour search with the k-NN
// contract of knn.cuh, doubles afterwards; exact feature-space 1-NN (ties -> smallest index); a counter-based sample
// stream consumed in index order with the single-thread semantics of the upstream loop.
//
// Shape on the device:
//   hybrid_knn_kernel     one CTA per query: the cells of the radius ball are scanned twice — a 1024-bin histogram of
//                         the squared distances bounds the k-th one, then the candidates up to that bin are collected
//                         in shared memory, sorted by (distance, index) with a bitonic network and the first k kept.
//                         Any k (featureKnn_ is 100; the warp-list search of knn.cuh stops at 32).
//   spfh_kernel           one warp per point: the pair features of its neighbours in double, 33 shared counters; the
//                         reference adds 100/(n-1) once per neighbour, so a bin's value is that increment added
//                         count times in sequence — reproduced from the count.
//   fpfh_kernel           36 threads per point: 33 run the reference's neighbour-order recurrence of one bin each,
//                         3 run the group sums' (k, bin)-order recurrence; same additions in the same order.
//   feature_nn_kernel     brute-force 1-NN in the 33-bin space in double (tiles of the other set in shared memory,
//                         the query column in registers), chunked over the other set + a fixed-order reduction.
//   ransac_*              hypotheses in batches: one thread per hypothesis draws the samples and runs the edge-length
//                         check, survivors estimate the transformation (Umeyama over the samples) and run the distance
//                         check; one warp per surviving hypothesis counts the inliers over all correspondences; one
//                         CTA then walks the batch in iteration order exactly like the sequential loop would (best
//                         result, exit estimate), so the answer does not depend on the batch size.
#include <cstring>
#include <cstdlib>
#include <new>
#include <algorithm>
#include "grid_index.cuh"
#include "knn.cuh"
#include "normals.cuh"

namespace b200 {
int inclusive_scan_u32(uint32_t* d_data, size_t n, uint32_t* d_tile, cudaStream_t s);   // preprocess.cu
size_t scan_tiles_for(size_t n);
int o3d_normals_from_ids_device(const int* ids, int knn, const float4* pts_by_id, const float4* q, int M, double* out_d,
                                int out_by_id, cudaStream_t s);   // o3d_icp.cu
namespace {

constexpr int kHKThreads = 128;
constexpr int kHKCap = 2048;
constexpr int kHKBins = 1024;

__global__ void __launch_bounds__(kHKThreads) hybrid_knn_kernel(const GridView g, const float4* __restrict__ q_by_id,
                                                                int N, int k, float md2, int* __restrict__ ids,
                                                                int* __restrict__ cnt, int* __restrict__ overflow) {
  __shared__ unsigned long long keys[kHKCap];
  __shared__ unsigned hist[kHKBins];
  __shared__ int s_n, s_thr, s_bad;
  const int i = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float4 q4 = __ldg(&q_by_id[i]);
  QueryGeom Q;
  const bool valid = make_geom(g, q4.x, q4.y, q4.z, md2, Q);
  for (int b = tid; b < kHKBins; b += kHKThreads) hist[b] = 0u;
  if (tid == 0) { s_n = 0; s_thr = kHKBins - 1; s_bad = 0; }
  __syncthreads();
  const CellBox box = ball_box(g, Q.ux, Q.uy, Q.uz, Q.rmd + Q.slack);
  const int rows = valid ? box.rows() : 0;
  const float scale = (float)kHKBins / md2;
  for (int pass = 0; pass < 2; ++pass) {
    const int thr = s_thr;
    for (int r = warp; r < rows; r += kHKThreads / 32) {
      const int iy = box.y0 + r % box.ny(), iz = box.z0 + r / box.ny();
      const uint32_t* row = g.G + ((size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz));
      const uint32_t a = __ldg(row + box.x0), b = __ldg(row + box.x1 + 1);
      for (uint32_t s = a + lane; s < b; s += 32) {
        const float4 p = __ldg(&g.pts[s]);
        const float d = dist2_rn(Q.qx, Q.qy, Q.qz, p.x, p.y, p.z);
        if (!(d <= md2)) continue;
        const int bin = min(kHKBins - 1, (int)(d * scale));   // monotone in d
        if (pass == 0) atomicAdd(&hist[bin], 1u);
        else if (bin <= thr) {
          const int slot = atomicAdd(&s_n, 1);
          if (slot < kHKCap) keys[slot] = ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)__float_as_int(p.w);
        }
      }
    }
    __syncthreads();
    if (pass == 0 && warp == 0) {
      // first bin at which the running count reaches k: a lane sums its 32 bins (rotated start: no bank conflicts), the
      // warp scans the 32 sums, the lane the count crosses in walks its own bins (one thread walking all 1024 left
      // the other 127 at the barrier for most of the CTA's life: ncu, profiles/r02_lc_kernels_ncu_summary.txt)
      static_assert(kHKBins == 32 * 32, "one lane per 32 bins");
      unsigned own = 0;
      for (int j = 0; j < 32; ++j) own += hist[lane * 32 + ((j + lane) & 31)];
      unsigned inc = own;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
      }
      const unsigned crossed = __ballot_sync(0xffffffffu, inc >= (unsigned)k);
      if (crossed == 0u) {            // fewer than k in the ball: all of them
        if (lane == 31 && inc > (unsigned)kHKCap) s_bad = 1;
      } else if (lane == __ffs(crossed) - 1) {   // at least k: everything up to the k-th one's bin
        unsigned cum = inc - own;
        int b = lane * 32;
        for (;; ++b) {
          cum += hist[b];
          if (cum >= (unsigned)k) break;
        }
        s_thr = b;
        if (cum > (unsigned)kHKCap) s_bad = 1;
      }
    }
    __syncthreads();
    if (s_bad) {   // more equal-distance candidates than the sort buffer holds: reported, not guessed
      if (tid == 0) { atomicExch(overflow, 1); cnt[i] = 0; }
      for (int t = tid; t < k; t += kHKThreads) ids[(size_t)i * k + t] = -1;
      return;
    }
  }
  const int n = s_n;
  if (n <= 2 * kHKThreads) {
    // short lists (the usual case: ~100 in the ball): a key's place is the number of smaller keys (keys are distinct:
    // the index is part of them) — broadcast reads, no barrier; same order as the sort below
    const int m = min(n, k);
    for (int t = tid; t < n; t += kHKThreads) {
      const unsigned long long mine = keys[t];
      int rank = 0;
      for (int j = 0; j < n; ++j) rank += keys[j] < mine ? 1 : 0;
      if (rank < k) ids[(size_t)i * k + rank] = (int)(unsigned)(mine & 0xffffffffull);
    }
    for (int t = m + tid; t < k; t += kHKThreads) ids[(size_t)i * k + t] = -1;
    if (tid == 0) cnt[i] = m;
    return;
  }
  int P = 1;
  while (P < n) P <<= 1;
  for (int t = n + tid; t < P; t += kHKThreads) keys[t] = ~0ull;
  __syncthreads();
  for (int size = 2; size <= P; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = tid; t < P / 2; t += kHKThreads) {
        const int lo = 2 * t - (t & (stride - 1));
        const int hi = lo + stride;
        const bool up = (lo & size) == 0;
        const unsigned long long x = keys[lo], y = keys[hi];
        if ((x > y) == up) { keys[lo] = y; keys[hi] = x; }
      }
      __syncthreads();
    }
  const int m = min(n, k);
  for (int t = tid; t < k; t += kHKThreads) ids[(size_t)i * k + t] = t < m ? (int)(unsigned)(keys[t] & 0xffffffffull) : -1;
  if (tid == 0) cnt[i] = m;
}

#define B200_PI 3.14159265358979323846

__device__ __forceinline__ int bin11(double x) {
  const int h = (int)floor(x);
  return h < 0 ? 0 : (h >= 11 ? 10 : h);
}

// ComputePairFeatures (Feature.cpp) -> the three histogram bins (f0 = atan2 angle, f1 = v . n2, f2 = n1 . d / |d|)
__device__ void pair_bins(const double* p1, const double* n1, const double* p2, const double* n2, int* b) {
  double dp[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  double f0 = 0.0, f1 = 0.0, f2 = 0.0;
  const double dist = sqrt(dp[0] * dp[0] + dp[1] * dp[1] + dp[2] * dp[2]);
  if (dist != 0.0) {
    const double* a = n1;
    const double* c = n2;
    const double angle1 = (n1[0] * dp[0] + n1[1] * dp[1] + n1[2] * dp[2]) / dist;
    const double angle2 = (n2[0] * dp[0] + n2[1] * dp[1] + n2[2] * dp[2]) / dist;
    double g2;
    if (acos(fabs(angle1)) > acos(fabs(angle2))) {
      a = n2; c = n1;
      dp[0] = -dp[0]; dp[1] = -dp[1]; dp[2] = -dp[2];
      g2 = -angle2;
    } else {
      g2 = angle1;
    }
    double v[3] = {dp[1] * a[2] - dp[2] * a[1], dp[2] * a[0] - dp[0] * a[2], dp[0] * a[1] - dp[1] * a[0]};
    const double vn = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (vn != 0.0) {
      v[0] /= vn; v[1] /= vn; v[2] /= vn;
      const double w[3] = {a[1] * v[2] - a[2] * v[1], a[2] * v[0] - a[0] * v[2], a[0] * v[1] - a[1] * v[0]};
      f2 = g2;
      f1 = v[0] * c[0] + v[1] * c[1] + v[2] * c[2];
      f0 = atan2(w[0] * c[0] + w[1] * c[1] + w[2] * c[2], a[0] * c[0] + a[1] * c[1] + a[2] * c[2]);
    }
  }
  b[0] = bin11(11.0 * (f0 + B200_PI) / (2.0 * B200_PI));
  b[1] = 11 + bin11(11.0 * (f1 + 1.0) * 0.5);
  b[2] = 22 + bin11(11.0 * (f2 + 1.0) * 0.5);
}

constexpr int kSpfhWarps = 8;
__global__ void __launch_bounds__(kSpfhWarps * 32) spfh_kernel(const float4* __restrict__ pts, const double* __restrict__ nrm,
                                                              const int* __restrict__ ids, const int* __restrict__ cnt,
                                                              int N, int k, double* __restrict__ spfh) {
  __shared__ int s_h[kSpfhWarps][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * kSpfhWarps + warp;
  if (i >= N) return;
  for (int j = lane; j < 33; j += 32) s_h[warp][j] = 0;
  __syncwarp();
  const int n = cnt[i];
  const float4 q = __ldg(&pts[i]);
  const double p1[3] = {(double)q.x, (double)q.y, (double)q.z};
  const double n1[3] = {nrm[3 * (size_t)i], nrm[3 * (size_t)i + 1], nrm[3 * (size_t)i + 2]};
  for (int t = 1 + lane; t < n; t += 32) {
    const int id = ids[(size_t)i * k + t];
    const float4 p = __ldg(&pts[id]);
    const double p2[3] = {(double)p.x, (double)p.y, (double)p.z};
    const double n2[3] = {nrm[3 * (size_t)id], nrm[3 * (size_t)id + 1], nrm[3 * (size_t)id + 2]};
    int b[3];
    pair_bins(p1, n1, p2, n2, b);
    atomicAdd(&s_h[warp][b[0]], 1);
    atomicAdd(&s_h[warp][b[1]], 1);
    atomicAdd(&s_h[warp][b[2]], 1);
  }
  __syncwarp();
  const double incr = n > 1 ? 100.0 / (double)(n - 1) : 0.0;
  for (int j = lane; j < 33; j += 32) {
    const int c = s_h[warp][j];
    double v = 0.0;
    for (int t = 0; t < c; ++t) v += incr;   // `feature(h, i) += hist_incr` once per neighbour
    spfh[(size_t)33 * i + j] = v;
  }
}

constexpr int kFpfhPts = 7;
// Fallback for neighbourhoods too large for the shared-memory tile of fpfh_cta_kernel: 36 threads per point, every
// thread walks the neighbours itself (dependent loads and one division per step: latency-bound, 0.73 ms for a
// 6 000-point submap at knn 100 where the tile kernel takes a tenth of that).
__global__ void __launch_bounds__(36 * kFpfhPts) fpfh_kernel(const float4* __restrict__ pts, const int* __restrict__ ids,
                                                            const int* __restrict__ cnt, int N, int k,
                                                            const double* __restrict__ spfh, double* __restrict__ out) {
  __shared__ double s_sum[kFpfhPts][3];
  const int x = threadIdx.x, y = threadIdx.y;
  const int i = blockIdx.x * kFpfhPts + y;
  const bool live = i < N;
  const int n = live ? cnt[i] : 0;
  double acc = 0.0;
  if (n > 1) {
    const float4 q = __ldg(&pts[i]);
    const double qx = (double)q.x, qy = (double)q.y, qz = (double)q.z;
    for (int t = 1; t < n; ++t) {
      const int id = ids[(size_t)i * k + t];
      const float4 p = __ldg(&pts[id]);
      const double dx = (double)p.x - qx, dy = (double)p.y - qy, dz = (double)p.z - qz;
      const double dist = dx * dx + dy * dy + dz * dz;
      if (dist == 0.0) continue;
      const double* col = spfh + (size_t)33 * id;
      if (x < 33) acc += col[x] / dist;
      else {
        const int g0 = 11 * (x - 33);
        for (int j = 0; j < 11; ++j) acc += col[g0 + j] / dist;
      }
    }
  }
  if (x >= 33) s_sum[y][x - 33] = acc;
  __syncthreads();
  if (live && x < 33) {
    double v = 0.0;
    if (n > 1) {
      double s = s_sum[y][x / 11];
      if (s != 0.0) s = 100.0 / s;
      v = acc * s;
      v += spfh[(size_t)33 * i + x];
    }
    out[(size_t)33 * i + x] = v;
  }
}

// One CTA per point. Phase A: the (neighbour, bin) quotients spfh[bin, id_k] / d_k are independent of one another, so
// all 128 threads compute them with their loads in flight together and park them in a shared-memory tile [k][33]
// (a skipped neighbour, d = 0, parks zeros: adding +0.0 to these non-negative sums changes nothing). Phase B: 33 threads
// run the reference's neighbour-order recurrence of one bin each from the tile, 3 threads the group sums' (k, bin)
// recurrence — the same additions in the same order as the fallback above and as the reference, with no load or
// division left on the dependent chain.
constexpr int kFpfhCtaThreads = 128;
__global__ void __launch_bounds__(kFpfhCtaThreads) fpfh_cta_kernel(const float4* __restrict__ pts, const int* __restrict__ ids,
                                                                  const int* __restrict__ cnt, int N, int k,
                                                                  const double* __restrict__ spfh, double* __restrict__ out) {
  extern __shared__ __align__(16) double s_val[];   // k x 33 quotients, then k distances, then k ids
  __shared__ double s_sum[3];
  const int i = blockIdx.x, tid = threadIdx.x;
  const int n = cnt[i];
  if (n <= 1) {
    if (tid < 33) out[(size_t)33 * i + tid] = 0.0;
    return;
  }
  double* s_dist = s_val + (size_t)33 * k;
  int* s_id = reinterpret_cast<int*>(s_dist + k);
  const float4 q = __ldg(&pts[i]);
  const double qx = (double)q.x, qy = (double)q.y, qz = (double)q.z;
  for (int t = 1 + tid; t < n; t += kFpfhCtaThreads) {
    const int id = ids[(size_t)i * k + t];
    const float4 p = __ldg(&pts[id]);
    const double dx = (double)p.x - qx, dy = (double)p.y - qy, dz = (double)p.z - qz;
    s_dist[t] = dx * dx + dy * dy + dz * dz;
    s_id[t] = id;
  }
  __syncthreads();
  const int pairs = (n - 1) * 33;
  for (int p = tid; p < pairs; p += kFpfhCtaThreads) {
    const int t = 1 + p / 33, j = p - (t - 1) * 33;
    const double d = s_dist[t];
    s_val[(size_t)33 * t + j] = d == 0.0 ? 0.0 : spfh[(size_t)33 * s_id[t] + j] / d;
  }
  __syncthreads();
  double acc = 0.0;
  if (tid < 33) {
    for (int t = 1; t < n; ++t) acc += s_val[(size_t)33 * t + tid];
  } else if (tid < 36) {
    const int g0 = 11 * (tid - 33);
    for (int t = 1; t < n; ++t) {
      const double* v = s_val + (size_t)33 * t + g0;
#pragma unroll
      for (int j = 0; j < 11; ++j) acc += v[j];
    }
    s_sum[tid - 33] = acc;
  }
  __syncthreads();
  if (tid < 33) {
    double s = s_sum[tid / 11];
    if (s != 0.0) s = 100.0 / s;
    double v = acc * s;
    v += spfh[(size_t)33 * i + tid];
    out[(size_t)33 * i + tid] = v;
  }
}

// ---- feature-space 1-NN -------------------------------------------------------------------------------------------
constexpr int kNNThreads = 128;
constexpr int kNNTile = 32;
constexpr int kNNTileBytes = kNNTile * 33 * (int)sizeof(double);   // 8448 = 528 x 16

// 1-D TMA (cp.async.bulk) + mbarrier helpers: a tile of 32 feature columns is ONE contiguous, 16-byte aligned run of
// 8448 bytes in global memory — exactly what the bulk-copy engine moves without occupying a single load slot.
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra LAB_DONE;\n"
      "bra LAB_WAIT;\n"
      "LAB_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// TMA = true: two shared-memory stages filled by cp.async.bulk (one elected thread, completion on an mbarrier), the
// copy of tile t + 1 in flight while tile t is scanned, one __syncthreads per tile (stage release). TMA = false: every
// thread loads its share of the tile with ordinary loads, two __syncthreads per tile. Same arithmetic, same results.
// b must be readable up to the next multiple of 16 bytes behind column M - 1 (the callers allocate the slack).
template <bool TMA>
__global__ void __launch_bounds__(kNNThreads) feature_nn_kernel(const double* __restrict__ a, int N,
                                                                const double* __restrict__ b, int M, int chunk,
                                                                double* __restrict__ part_d, int* __restrict__ part_j) {
  __shared__ __align__(128) double s_b[TMA ? 2 : 1][kNNTile * 33];
  __shared__ __align__(8) unsigned long long s_bar[2];
  const int i = blockIdx.x * kNNThreads + threadIdx.x;
  const int j0 = blockIdx.y * chunk, j1 = min(M, j0 + chunk);
  double f[33];
#pragma unroll
  for (int t = 0; t < 33; ++t) f[t] = i < N ? a[(size_t)33 * i + t] : 0.0;
  double best = INFINITY;
  int bj = -1;
  const int ntiles = (j1 - j0 + kNNTile - 1) / kNNTile;
  auto tile_bytes = [&](int t) {
    const int nt = min(kNNTile, j1 - (j0 + t * kNNTile));
    return (unsigned)((nt * 33 * (int)sizeof(double) + 15) & ~15);
  };
  if (TMA) {
    if (threadIdx.x == 0) {
      mbar_init(&s_bar[0], 1);
      mbar_init(&s_bar[1], 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0)
      for (int t = 0; t < 2 && t < ntiles; ++t) {
        mbar_expect_tx(&s_bar[t], tile_bytes(t));
        bulk_g2s(s_b[t], b + (size_t)33 * (j0 + t * kNNTile), tile_bytes(t), &s_bar[t]);
      }
  }
  for (int t = 0; t < ntiles; ++t) {
    const int j = j0 + t * kNNTile;
    const int nt = min(kNNTile, j1 - j);
    const int st = TMA ? (t & 1) : 0;
    if (TMA) {
      mbar_wait(&s_bar[st], (unsigned)((t >> 1) & 1));
    } else {
      __syncthreads();
      for (int u = threadIdx.x; u < nt * 33; u += kNNThreads) s_b[0][u] = b[(size_t)33 * j + u];
      __syncthreads();
    }
    const double* tb = s_b[st];
    for (int jj = 0; jj < nt; ++jj) {
      double d = 0.0;
#pragma unroll
      for (int u = 0; u < 33; ++u) { const double e = f[u] - tb[jj * 33 + u]; d += e * e; }
      if (d < best) { best = d; bj = j + jj; }
    }
    if (TMA) {
      __syncthreads();   // every thread is done with this stage: refill it with tile t + 2
      if (threadIdx.x == 0 && t + 2 < ntiles) {
        mbar_expect_tx(&s_bar[st], tile_bytes(t + 2));
        bulk_g2s(s_b[st], b + (size_t)33 * (j0 + (t + 2) * kNNTile), tile_bytes(t + 2), &s_bar[st]);
      }
    }
  }
  if (i < N) {
    part_d[(size_t)blockIdx.y * N + i] = best;
    part_j[(size_t)blockIdx.y * N + i] = bj;
  }
}

__global__ void feature_nn_reduce_kernel(const double* __restrict__ part_d, const int* __restrict__ part_j, int N,
                                         int chunks, int* __restrict__ nn, double* __restrict__ d2) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  double best = INFINITY;
  int bj = -1;
  for (int c = 0; c < chunks; ++c) {   // ascending chunks, strict <: ties -> smallest index
    const double d = part_d[(size_t)c * N + i];
    if (d < best) { best = d; bj = part_j[(size_t)c * N + i]; }
  }
  nn[i] = bj;
  if (d2) d2[i] = best;
}

// ---- correspondences: ordered compaction by one CTA ---------------------------------------------------------------
// mode 0: every source point with its feature neighbour; mode 1: the mutual ones. cs / ct: indices, ps / pt: points.
__global__ void __launch_bounds__(1024) corr_build_kernel(const int* __restrict__ ij, const int* __restrict__ ji, int Ns,
                                                          int mode, const float4* __restrict__ src,
                                                          const float4* __restrict__ tgt, int* __restrict__ cs,
                                                          int* __restrict__ ct, double* __restrict__ ps,
                                                          double* __restrict__ pt, int* __restrict__ n_out) {
  __shared__ int s_w[32];
  __shared__ int s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int b0 = 0; b0 < Ns; b0 += 1024) {
    const int i = b0 + tid;
    int j = -1;
    bool keep = false;
    if (i < Ns) {
      j = ij[i];
      keep = j >= 0 && (mode == 0 || ji[j] == i);
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_w[warp] = __popc(bal);
    __syncthreads();
    int off = s_base;
    for (int w = 0; w < warp; ++w) off += s_w[w];
    if (keep) {
      const int o = off + __popc(bal & ((1u << lane) - 1u));
      cs[o] = i; ct[o] = j;
      const float4 a = __ldg(&src[i]), c = __ldg(&tgt[j]);
      ps[3 * (size_t)o] = (double)a.x; ps[3 * (size_t)o + 1] = (double)a.y; ps[3 * (size_t)o + 2] = (double)a.z;
      pt[3 * (size_t)o] = (double)c.x; pt[3 * (size_t)o + 1] = (double)c.y; pt[3 * (size_t)o + 2] = (double)c.z;
    }
    __syncthreads();
    if (tid == 0) {
      int t = 0;
      for (int w = 0; w < 32; ++w) t += s_w[w];
      s_base += t;
    }
    __syncthreads();
  }
  if (tid == 0) *n_out = s_base;
}

// ---- RANSAC -------------------------------------------------------------------------------------------------------
struct RansacState {
  long long exit_itr;     // iterations >= this are not consumed
  long long itr_done;     // iterations consumed so far
  long long validations;
  int best_good;
  int n_surv;             // survivors of the batch being processed
  double best_err2, best_rmse;
  double best_T[16];
};

struct RansacParamsDev {
  int nc, rn, max_iteration;
  double edge, dist_thr, max2, confidence;
  unsigned long long seed;
};

__device__ __forceinline__ unsigned long long splitmix(unsigned long long x) {
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
  x ^= x >> 27; x *= 0x94d049bb133111ebull;
  x ^= x >> 31;
  return x;
}

// Eigen::umeyama(src, dst, with_scaling = false) over n pairs (TransformationEstimationPointToPoint): sigma =
// 1/n sum (d - md)(s - ms)^T = U D V^T, R = U diag(1, 1, sign(det U det V)) V^T, t = md - R ms. One-sided Jacobi on the
// columns of sigma; the left vectors are rebuilt as u1, u2, u1 x u2 (three pairs give a rank-2 sigma; Eigen's full U
// holds +-(u1 x u2) and the sign rule makes R the same proper rotation either way).
__device__ void umeyama_rigid_dev(const double* s3xn, const double* d3xn, int n, double* T16) {
  double ms[3] = {0, 0, 0}, md[3] = {0, 0, 0};
  for (int i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) { ms[a] += s3xn[3 * i + a]; md[a] += d3xn[3 * i + a]; }
  for (int a = 0; a < 3; ++a) { ms[a] /= (double)n; md[a] /= (double)n; }
  double A[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  for (int i = 0; i < n; ++i)
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r) A[r + 3 * c] += (d3xn[3 * i + r] - md[r]) * (s3xn[3 * i + c] - ms[c]);
  for (int a = 0; a < 9; ++a) A[a] /= (double)n;
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rotated = false;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        double al = 0, be = 0, ga = 0;
        for (int k = 0; k < 3; ++k) { al += A[k + 3 * p] * A[k + 3 * p]; be += A[k + 3 * q] * A[k + 3 * q]; ga += A[k + 3 * p] * A[k + 3 * q]; }
        if (ga == 0.0 || fabs(ga) <= 1e-16 * sqrt(al * be)) continue;
        rotated = true;
        const double zeta = (be - al) / (2.0 * ga);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
        for (int k = 0; k < 3; ++k) {
          const double ap = A[k + 3 * p], aq = A[k + 3 * q];
          A[k + 3 * p] = c * ap - sn * aq; A[k + 3 * q] = sn * ap + c * aq;
          const double vp = V[k + 3 * p], vq = V[k + 3 * q];
          V[k + 3 * p] = c * vp - sn * vq; V[k + 3 * q] = sn * vp + c * vq;
        }
      }
    if (!rotated) break;
  }
  double sg[3];
  int o[3] = {0, 1, 2};
  for (int j = 0; j < 3; ++j) sg[j] = sqrt(A[3 * j] * A[3 * j] + A[1 + 3 * j] * A[1 + 3 * j] + A[2 + 3 * j] * A[2 + 3 * j]);
  for (int a = 0; a < 2; ++a)
    for (int b = a + 1; b < 3; ++b)
      if (sg[o[b]] > sg[o[a]]) { const int t = o[a]; o[a] = o[b]; o[b] = t; }
  double u1[3], u2[3], u3[3], v1[3], v2[3], v3[3];
  for (int k = 0; k < 3; ++k) {
    u1[k] = sg[o[0]] > 0 ? A[k + 3 * o[0]] / sg[o[0]] : (k == 0 ? 1.0 : 0.0);
    v1[k] = V[k + 3 * o[0]]; v2[k] = V[k + 3 * o[1]]; v3[k] = V[k + 3 * o[2]];
  }
  if (sg[o[1]] > 1e-300) {
    for (int k = 0; k < 3; ++k) u2[k] = A[k + 3 * o[1]] / sg[o[1]];
  } else {
    const int m = fabs(u1[0]) <= fabs(u1[1]) ? (fabs(u1[0]) <= fabs(u1[2]) ? 0 : 2) : (fabs(u1[1]) <= fabs(u1[2]) ? 1 : 2);
    double e[3] = {0, 0, 0};
    e[m] = 1.0;
    const double d = u1[m];
    double n2 = 0;
    for (int k = 0; k < 3; ++k) { u2[k] = e[k] - d * u1[k]; n2 += u2[k] * u2[k]; }
    for (int k = 0; k < 3; ++k) u2[k] /= sqrt(n2);
  }
  u3[0] = u1[1] * u2[2] - u1[2] * u2[1]; u3[1] = u1[2] * u2[0] - u1[0] * u2[2]; u3[2] = u1[0] * u2[1] - u1[1] * u2[0];
  const double detV = v1[0] * (v2[1] * v3[2] - v2[2] * v3[1]) - v2[0] * (v1[1] * v3[2] - v1[2] * v3[1]) +
                      v3[0] * (v1[1] * v2[2] - v1[2] * v2[1]);
  const double sgn = detV < 0 ? -1.0 : 1.0;
  for (int i = 0; i < 16; ++i) T16[i] = 0.0;
  T16[15] = 1.0;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) T16[r + 4 * c] = u1[r] * v1[c] + u2[r] * v2[c] + sgn * u3[r] * v3[c];
  for (int r = 0; r < 3; ++r) T16[r + 12] = md[r] - (T16[r] * ms[0] + T16[r + 4] * ms[1] + T16[r + 8] * ms[2]);
}

__device__ __forceinline__ void xform3(const double* T, const double* p, double* out) {
  for (int r = 0; r < 3; ++r) out[r] = T[r] * p[0] + T[r + 4] * p[1] + T[r + 8] * p[2] + T[r + 12];
}

constexpr int kMaxRn = 8;

// one thread per hypothesis of the batch [start, start + B): samples, edge-length check, transformation, distance
// check; survivors are appended (unordered) with their transformation
__global__ void __launch_bounds__(128) ransac_hypo_kernel(const RansacParamsDev p, RansacState* __restrict__ st,
                                                          long long start, int B, const double* __restrict__ ps,
                                                          const double* __restrict__ pt, int* __restrict__ good,
                                                          int* __restrict__ slot_of, int* __restrict__ surv_h,
                                                          double* __restrict__ surv_T) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= B) return;
  good[h] = -1;
  slot_of[h] = -1;
  const long long itr = start + h;
  const long long limit = min((long long)p.max_iteration, st->exit_itr);
  if (itr >= limit) return;
  int sidx[kMaxRn];
  double s[3 * kMaxRn], d[3 * kMaxRn];
  for (int j = 0; j < p.rn; ++j) {
    const unsigned long long x = splitmix(p.seed + ((unsigned long long)itr * (unsigned long long)p.rn + (unsigned long long)j) * 0x9e3779b97f4a7c15ull);
    sidx[j] = (int)(((x >> 32) * (unsigned long long)p.nc) >> 32);
    for (int a = 0; a < 3; ++a) {
      s[3 * j + a] = ps[3 * (size_t)sidx[j] + a];
      d[3 * j + a] = pt[3 * (size_t)sidx[j] + a];
    }
  }
  if (p.edge > 0.0) {
    for (int a = 0; a < p.rn; ++a)
      for (int b = a + 1; b < p.rn; ++b) {
        double ds = 0, dt = 0;
        for (int t = 0; t < 3; ++t) {
          ds += (s[3 * a + t] - s[3 * b + t]) * (s[3 * a + t] - s[3 * b + t]);
          dt += (d[3 * a + t] - d[3 * b + t]) * (d[3 * a + t] - d[3 * b + t]);
        }
        ds = sqrt(ds); dt = sqrt(dt);
        if (ds < dt * p.edge || dt < ds * p.edge) return;
      }
  }
  double T[16];
  umeyama_rigid_dev(s, d, p.rn, T);
  if (p.dist_thr > 0.0) {
    for (int j = 0; j < p.rn; ++j) {
      double q[3];
      xform3(T, s + 3 * j, q);
      double e = 0;
      for (int t = 0; t < 3; ++t) e += (d[3 * j + t] - q[t]) * (d[3 * j + t] - q[t]);
      if (sqrt(e) > p.dist_thr) return;
    }
  }
  const int slot = atomicAdd(&st->n_surv, 1);
  slot_of[h] = slot;
  surv_h[slot] = h;
  for (int t = 0; t < 16; ++t) surv_T[16 * (size_t)slot + t] = T[t];
}

// one warp per surviving hypothesis: EvaluateRANSACBasedOnCorrespondence over all correspondences
__global__ void __launch_bounds__(256) ransac_eval_kernel(const RansacParamsDev p, const RansacState* __restrict__ st,
                                                          const double* __restrict__ ps, const double* __restrict__ pt,
                                                          const int* __restrict__ surv_h,
                                                          const double* __restrict__ surv_T, int* __restrict__ good,
                                                          double* __restrict__ err2) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  const int ns = st->n_surv;
  for (int sv = warp; sv < ns; sv += nwarps) {
    double T[16];
    for (int t = 0; t < 16; ++t) T[t] = surv_T[16 * (size_t)sv + t];
    int g = 0;
    double e2 = 0.0;
    for (int c = lane; c < p.nc; c += 32) {
      double q[3];
      xform3(T, ps + 3 * (size_t)c, q);
      double e = 0;
      for (int t = 0; t < 3; ++t) { const double v = q[t] - pt[3 * (size_t)c + t]; e += v * v; }
      if (e < p.max2) { ++g; e2 += e; }
    }
    for (int o = 16; o > 0; o >>= 1) {
      g += __shfl_xor_sync(0xffffffffu, g, o);
      e2 += __shfl_xor_sync(0xffffffffu, e2, o);
    }
    if (lane == 0) { good[surv_h[sv]] = g; err2[surv_h[sv]] = e2; }
  }
}

struct Cand {
  int good;
  double rmse;
  int h;
};
__device__ __forceinline__ bool better(int g, double r, int bg, double br) {   // IsBetterRANSACThan
  return g > bg || (g == bg && g > 0 && r < br);
}

// one CTA walks the batch in iteration order: which hypotheses the sequential loop would have adopted, and where it
// would have stopped
constexpr int kWalkThreads = 1024;
__global__ void __launch_bounds__(kWalkThreads) ransac_walk_kernel(const RansacParamsDev p, RansacState* __restrict__ st,
                                                                   long long start, int B, const int* __restrict__ good,
                                                                   const double* __restrict__ err2,
                                                                   const int* __restrict__ slot_of,
                                                                   const double* __restrict__ surv_T) {
  extern __shared__ unsigned s_rec[];          // B / 32 record bits
  __shared__ int s_g[kWalkThreads];
  __shared__ double s_r[kWalkThreads];
  __shared__ long long s_stop;
  __shared__ int s_val;
  const int tid = threadIdx.x;
  const long long limit = min((long long)p.max_iteration, st->exit_itr);
  if (start >= limit) {
    if (tid == 0) st->n_surv = 0;
    return;
  }
  const int nb = (int)min((long long)B, limit - start);   // entries of this batch that may be consumed
  const int per = (B + kWalkThreads - 1) / kWalkThreads;
  const int h0 = tid * per, h1 = min(nb, h0 + per);
  for (int w = tid; w < (B + 31) / 32; w += kWalkThreads) s_rec[w] = 0u;
  if (tid == 0) s_val = 0;
  // segment bests
  int bg = 0;
  double br = 0.0;
  for (int h = h0; h < h1; ++h) {
    const int g = good[h];
    if (g <= 0) continue;
    const double r = sqrt(err2[h] / (double)g);
    if (better(g, r, bg, br)) { bg = g; br = r; }
  }
  s_g[tid] = bg; s_r[tid] = br;
  __syncthreads();
  if (tid == 0) {   // exclusive prefix of the segment bests, seeded with the best of the earlier batches
    int cg = st->best_good;
    double cr = st->best_rmse;
    for (int t = 0; t < kWalkThreads; ++t) {
      const int g = s_g[t];
      const double r = s_r[t];
      s_g[t] = cg; s_r[t] = cr;
      if (better(g, r, cg, cr)) { cg = g; cr = r; }
    }
  }
  __syncthreads();
  bg = s_g[tid]; br = s_r[tid];
  for (int h = h0; h < h1; ++h) {
    const int g = good[h];
    if (g <= 0) continue;
    const double r = sqrt(err2[h] / (double)g);
    if (better(g, r, bg, br)) { bg = g; br = r; atomicOr(&s_rec[h >> 5], 1u << (h & 31)); }
  }
  __syncthreads();
  if (tid == 0) {
    long long exit_itr = st->exit_itr;
    long long stop = start + nb;      // first iteration not consumed
    long long last = start;           // one past the last adopted hypothesis
    for (int w = 0; w < (nb + 31) / 32; ++w) {
      unsigned m = s_rec[w];
      bool out = false;
      while (m) {
        const int h = 32 * w + __ffs(m) - 1;
        m &= m - 1;
        if (start + h >= exit_itr) { out = true; break; }
        const int g = good[h];
        last = start + h + 1;
        st->best_good = g;
        st->best_err2 = err2[h];
        st->best_rmse = sqrt(err2[h] / (double)g);
        const double* T = surv_T + 16 * (size_t)slot_of[h];
        for (int t = 0; t < 16; ++t) st->best_T[t] = T[t];
        const double ratio = (double)g / (double)p.nc;
        const double pw = pow(ratio, (double)p.rn);
        if (p.confidence < 1.0 && pw > 0.0) {
          if (pw >= 1.0) exit_itr = 0;
          else {
            const double kk = log(1.0 - p.confidence) / log(1.0 - pw);
            if (kk >= 0.0 && kk < (double)exit_itr) exit_itr = (long long)ceil(kk);
          }
        }
      }
      if (out || start + 32LL * (w + 1) >= exit_itr) break;
    }
    if (exit_itr < stop) stop = max(exit_itr, last);   // the loop finishes the iteration that lowered the estimate
    st->exit_itr = exit_itr;
    st->itr_done = stop;
    st->n_surv = 0;
    s_stop = stop;
  }
  __syncthreads();
  const long long stop = s_stop;
  int v = 0;
  for (int h = h0; h < h1; ++h)
    if (start + h < stop && good[h] >= 0) ++v;
  if (v) atomicAdd(&s_val, v);
  __syncthreads();
  if (tid == 0) st->validations += s_val;
}

// inlier correspondences of the adopted hypothesis, in correspondence order (one CTA)
__global__ void __launch_bounds__(1024) ransac_final_kernel(const RansacParamsDev p, const RansacState* __restrict__ st,
                                                            const double* __restrict__ ps, const double* __restrict__ pt,
                                                            const int* __restrict__ cs, const int* __restrict__ ct,
                                                            int* __restrict__ corr2, int* __restrict__ n_out) {
  __shared__ int s_w[32];
  __shared__ int s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_base = 0;
  __syncthreads();
  if (st->best_good <= 0) {
    if (tid == 0) *n_out = 0;
    return;
  }
  double T[16];
  for (int t = 0; t < 16; ++t) T[t] = st->best_T[t];
  for (int b0 = 0; b0 < p.nc; b0 += 1024) {
    const int c = b0 + tid;
    bool keep = false;
    if (c < p.nc) {
      double q[3];
      xform3(T, ps + 3 * (size_t)c, q);
      double e = 0;
      for (int t = 0; t < 3; ++t) { const double v = q[t] - pt[3 * (size_t)c + t]; e += v * v; }
      keep = e < p.max2;
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_w[warp] = __popc(bal);
    __syncthreads();
    int off = s_base;
    for (int w = 0; w < warp; ++w) off += s_w[w];
    if (keep) {
      const int o = off + __popc(bal & ((1u << lane) - 1u));
      corr2[2 * (size_t)o] = cs[c];
      corr2[2 * (size_t)o + 1] = ct[c];
    }
    __syncthreads();
    if (tid == 0) {
      int t = 0;
      for (int w = 0; w < 32; ++w) t += s_w[w];
      s_base += t;
    }
    __syncthreads();
  }
  if (tid == 0) *n_out = s_base;
}


// ---- computeIndicesOfOverlappingPoints (open3d_slam/src/helpers.cpp:320-345) ----------------------------------------
// Both clouds go into one voxel table (key = floor(p / voxelSize) per axis, VoxelHashMap.hpp:48-51, 21 bits per axis;
// the source transformed first, helpers.cpp:328-330) with one counter per cloud; a point is kept when its voxel holds
// at least min_points of BOTH clouds.
constexpr unsigned long long kVoxEmpty = ~0ull;
__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
  return x;
}
struct Mat4 { double m[16]; };
__global__ void overlap_insert_kernel(const float4* __restrict__ src, int Ns, const float4* __restrict__ tgt, int Nt,
                                      const Mat4 T, double inv, unsigned long long mask,
                                      unsigned long long* __restrict__ keys, uint32_t* __restrict__ cnt2,
                                      uint32_t* __restrict__ slot_of, int* __restrict__ bad) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < Ns + Nt; i += gridDim.x * blockDim.x) {
    const bool is_src = i < Ns;
    const float4 q = __ldg(is_src ? &src[i] : &tgt[i - Ns]);
    double p[3] = {(double)q.x, (double)q.y, (double)q.z};
    if (is_src) { double t[3]; xform3(T.m, p, t); p[0] = t[0]; p[1] = t[1]; p[2] = t[2]; }
    const double fx = floor(p[0] * inv), fy = floor(p[1] * inv), fz = floor(p[2] * inv);
    if (!(fabs(fx) < 1048575.0 && fabs(fy) < 1048575.0 && fabs(fz) < 1048575.0)) { *bad = 1; slot_of[i] = 0xffffffffu; continue; }
    const unsigned long long key = ((unsigned long long)((long long)fx + 1048576ll) << 42) |
                                   ((unsigned long long)((long long)fy + 1048576ll) << 21) |
                                   (unsigned long long)((long long)fz + 1048576ll);
    unsigned long long s = mix64(key) & mask;
    while (true) {
      const unsigned long long prev = atomicCAS(&keys[s], kVoxEmpty, key);
      if (prev == kVoxEmpty || prev == key) break;
      s = (s + 1ull) & mask;
    }
    slot_of[i] = (uint32_t)s;
    atomicAdd(&cnt2[2 * s + (is_src ? 0 : 1)], 1u);
  }
}
__global__ void overlap_flag_kernel(const uint32_t* __restrict__ slot_of, const uint32_t* __restrict__ cnt2, int n,
                                    uint32_t min_points, uint32_t* __restrict__ flag) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t s = slot_of[i];
    flag[i] = (s != 0xffffffffu && cnt2[2 * (size_t)s] >= min_points && cnt2[2 * (size_t)s + 1] >= min_points) ? 1u : 0u;
  }
}
// scan = inclusive scan of the flags over [source | target]; kept indices of each cloud in index order
__global__ void overlap_scatter_kernel(const uint32_t* __restrict__ scan, int Ns, int Nt, int* __restrict__ out_src,
                                       int* __restrict__ out_tgt) {
  const uint32_t ns_kept = Ns ? scan[Ns - 1] : 0u;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < Ns + Nt; i += gridDim.x * blockDim.x) {
    const uint32_t a = scan[i], a0 = i ? scan[i - 1] : 0u;
    if (a == a0) continue;
    if (i < Ns) out_src[a0] = i;
    else out_tgt[a0 - ns_kept] = i - Ns;
  }
}
__global__ void fill_u64_lc_kernel(unsigned long long* __restrict__ p, size_t n, unsigned long long v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

template <class T>
int grow(T** p, size_t* cap, size_t n) {
  if (n <= *cap) return B200_OK;
  cudaFree(*p);
  *p = nullptr;
  *cap = 0;
  B200_CUDA(cudaMalloc(p, n * sizeof(T)));
  *cap = n;
  return B200_OK;
}

}  // namespace
}  // namespace b200

using namespace b200;

struct b200_lc {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[2] = {nullptr, nullptr};
  float last_ms = 0.f;
  MapIndex idx;                 // neighbour index of compute_fpfh
  // FPFH scratch
  double *d_nrm = nullptr, *d_spfh = nullptr, *d_fpfh = nullptr;
  int *d_ids = nullptr, *d_cnt = nullptr, *d_flag = nullptr;
  size_t cap_nrm = 0, cap_spfh = 0, cap_fpfh = 0, cap_ids = 0, cap_cnt = 0;
  // feature matching / RANSAC scratch
  float4 *d_src = nullptr, *d_tgt = nullptr;
  double *d_sf = nullptr, *d_tf = nullptr, *d_part_d = nullptr, *d_ps = nullptr, *d_pt = nullptr, *d_err2 = nullptr,
         *d_surv_T = nullptr;
  int *d_part_j = nullptr, *d_ij = nullptr, *d_ji = nullptr, *d_cs = nullptr, *d_ct = nullptr, *d_good = nullptr,
      *d_slot = nullptr, *d_surv_h = nullptr, *d_corr = nullptr, *d_n = nullptr;
  size_t cap_src = 0, cap_tgt = 0, cap_sf = 0, cap_tf = 0, cap_part_d = 0, cap_part_j = 0, cap_ij = 0, cap_ji = 0,
         cap_cs = 0, cap_ct = 0, cap_ps = 0, cap_pt = 0, cap_good = 0, cap_err2 = 0, cap_slot = 0, cap_surv_h = 0,
         cap_surv_T = 0, cap_corr = 0;
  RansacState* d_state = nullptr;
  RansacState* h_state = nullptr;   // pinned
  // overlap scratch
  unsigned long long* d_vkeys = nullptr;
  uint32_t *d_vcnt = nullptr, *d_vslot = nullptr, *d_vflag = nullptr, *d_vtile = nullptr;
  int *d_vsrc = nullptr, *d_vtgt = nullptr;
  size_t cap_vkeys = 0, cap_vcnt = 0, cap_vslot = 0, cap_vflag = 0, cap_vtile = 0, cap_vsrc = 0, cap_vtgt = 0;
};

static int feature_nn(b200_lc* h, const double* d_a, int N, const double* d_b, int M, int* d_nn, double* d_d2) {
  cudaStream_t s = h->stream;
  const int blocks_x = (N + kNNThreads - 1) / kNNThreads;
  int chunks = std::max(1, std::min(64, (148 * 8 + blocks_x - 1) / blocks_x));
  int chunk = (M + chunks - 1) / chunks;
  chunk = std::max(kNNTile, (chunk + kNNTile - 1) / kNNTile * kNNTile);
  chunks = (M + chunk - 1) / chunk;
  B200_TRY(grow(&h->d_part_d, &h->cap_part_d, (size_t)chunks * N));
  B200_TRY(grow(&h->d_part_j, &h->cap_part_j, (size_t)chunks * N));
  static const bool use_tma = [] { const char* e = std::getenv("B200_LC_TMA"); return !e || std::atoi(e) != 0; }();
  if (use_tma) feature_nn_kernel<true><<<dim3(blocks_x, chunks), kNNThreads, 0, s>>>(d_a, N, d_b, M, chunk, h->d_part_d, h->d_part_j);
  else feature_nn_kernel<false><<<dim3(blocks_x, chunks), kNNThreads, 0, s>>>(d_a, N, d_b, M, chunk, h->d_part_d, h->d_part_j);
  B200_CUDA(cudaGetLastError());
  feature_nn_reduce_kernel<<<(N + 255) / 256, 256, 0, s>>>(h->d_part_d, h->d_part_j, N, chunks, d_nn, d_d2);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

extern "C" {

int b200_lc_create(b200_lc** out, int device) {
  if (!out) { set_error("b200_lc_create: bad arguments"); return B200_E_BAD_PARAM; }
  *out = nullptr;
  DeviceGuard dg(device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  b200_lc* h = new (std::nothrow) b200_lc();
  if (!h) { set_error("out of host memory"); return B200_E_CUDA; }
  h->device = device;
  h->idx.device = device;
  auto fail = [&](cudaError_t e, const char* what) { const int st = cuda_fail(e, what, __FILE__, __LINE__); b200_lc_destroy(h); return st; };
  cudaError_t e;
  if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(e, "cudaStreamCreate");
  h->idx.stream = h->stream;
  for (int i = 0; i < 2; ++i)
    if ((e = cudaEventCreate(&h->ev[i])) != cudaSuccess) return fail(e, "cudaEventCreate");
  if ((e = cudaMalloc(&h->d_state, sizeof(RansacState))) != cudaSuccess) return fail(e, "cudaMalloc");
  if ((e = cudaMallocHost(&h->h_state, sizeof(RansacState))) != cudaSuccess) return fail(e, "cudaMallocHost");
  if ((e = cudaMalloc(&h->d_n, 4 * sizeof(int))) != cudaSuccess) return fail(e, "cudaMalloc");
  if ((e = cudaMalloc(&h->d_flag, sizeof(int))) != cudaSuccess) return fail(e, "cudaMalloc");
  *out = h;
  return B200_OK;
}

void b200_lc_destroy(b200_lc* h) {
  if (!h) return;
  {
    DeviceGuard dg(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->d_nrm); cudaFree(h->d_spfh); cudaFree(h->d_fpfh); cudaFree(h->d_ids); cudaFree(h->d_cnt); cudaFree(h->d_flag);
    cudaFree(h->d_src); cudaFree(h->d_tgt); cudaFree(h->d_sf); cudaFree(h->d_tf); cudaFree(h->d_part_d); cudaFree(h->d_ps);
    cudaFree(h->d_pt); cudaFree(h->d_err2); cudaFree(h->d_surv_T); cudaFree(h->d_part_j); cudaFree(h->d_ij); cudaFree(h->d_ji);
    cudaFree(h->d_cs); cudaFree(h->d_ct); cudaFree(h->d_good); cudaFree(h->d_slot); cudaFree(h->d_surv_h); cudaFree(h->d_corr);
    cudaFree(h->d_n); cudaFree(h->d_state);
    cudaFree(h->d_vkeys); cudaFree(h->d_vcnt); cudaFree(h->d_vslot); cudaFree(h->d_vflag); cudaFree(h->d_vtile);
    cudaFree(h->d_vsrc); cudaFree(h->d_vtgt);
    cudaFreeHost(h->h_state);
    for (int i = 0; i < 2; ++i)
      if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    h->idx.release();
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
}

// normals3xN == nullptr: the normals are estimated on the device first (est_knn, est_radius; Open3D semantics) on the
// same index and optionally returned through normals_out3xN
static int lc_fpfh_impl(b200_lc* h, const float* pts4xN, const double* normals3xN, int N, float radius, int max_nn,
                        double* fpfh33xN, int32_t* nn_count_N, int est_knn, float est_radius, double* normals_out3xN) {
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  B200_TRY(index_alloc_points(h->idx, N, false));
  B200_TRY(upload_features(h->idx.d_ref, pts4xN, N, s));
  B200_TRY(grow(&h->d_nrm, &h->cap_nrm, (size_t)3 * N));
  B200_TRY(grow(&h->d_spfh, &h->cap_spfh, (size_t)33 * N));
  B200_TRY(grow(&h->d_fpfh, &h->cap_fpfh, (size_t)33 * N));
  B200_TRY(grow(&h->d_ids, &h->cap_ids, (size_t)max_nn * N));
  B200_TRY(grow(&h->d_cnt, &h->cap_cnt, (size_t)N));
  if (normals3xN) B200_CUDA(cudaMemcpyAsync(h->d_nrm, normals3xN, (size_t)3 * N * sizeof(double), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaMemsetAsync(h->d_flag, 0, sizeof(int), s));
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  b200_index_opts opts;
  std::memset(&opts, 0, sizeof(opts));
  B200_TRY(index_build(h->idx, &opts));
  if (!normals3xN) {
    int* nids = nullptr;
    B200_TRY(self_knn_ids_index(h->idx, est_knn, est_radius, true, &nids, s));
    B200_TRY(o3d_normals_from_ids_device(nids, est_knn, h->idx.d_ref, h->idx.d_pts, N, h->d_nrm, 1, s));
  }
  const GridView g = h->idx.view();
  hybrid_knn_kernel<<<N, kHKThreads, 0, s>>>(g, h->idx.d_ref, N, max_nn, radius * radius, h->d_ids, h->d_cnt, h->d_flag);
  B200_CUDA(cudaGetLastError());
  spfh_kernel<<<(N + kSpfhWarps - 1) / kSpfhWarps, kSpfhWarps * 32, 0, s>>>(h->idx.d_ref, h->d_nrm, h->d_ids, h->d_cnt, N, max_nn, h->d_spfh);
  B200_CUDA(cudaGetLastError());
  const size_t tile = (size_t)max_nn * (33 * sizeof(double) + sizeof(double) + sizeof(int));
  if (tile <= 200 * 1024) {
    if (tile > 48 * 1024) B200_CUDA(cudaFuncSetAttribute(fpfh_cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile));
    fpfh_cta_kernel<<<N, kFpfhCtaThreads, tile, s>>>(h->idx.d_ref, h->d_ids, h->d_cnt, N, max_nn, h->d_spfh, h->d_fpfh);
  } else {
    fpfh_kernel<<<(N + kFpfhPts - 1) / kFpfhPts, dim3(36, kFpfhPts), 0, s>>>(h->idx.d_ref, h->d_ids, h->d_cnt, N, max_nn, h->d_spfh, h->d_fpfh);
  }
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  int flag = 0;
  B200_CUDA(cudaMemcpyAsync(&flag, h->d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaMemcpyAsync(fpfh33xN, h->d_fpfh, (size_t)33 * N * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (nn_count_N) B200_CUDA(cudaMemcpyAsync(nn_count_N, h->d_cnt, (size_t)N * sizeof(int), cudaMemcpyDeviceToHost, s));
  if (normals_out3xN) B200_CUDA(cudaMemcpyAsync(normals_out3xN, h->d_nrm, (size_t)3 * N * sizeof(double), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]);
  if (flag) {
    set_error("b200_lc_compute_fpfh: a neighbourhood holds more than 2048 points at the distance of its max_nn-th neighbour");
    return B200_E_UNSUPPORTED;
  }
  return B200_OK;
}

int b200_lc_compute_fpfh(b200_lc* h, const float* pts4xN, const double* normals3xN, int N, float radius, int max_nn,
                         double* fpfh33xN, int32_t* nn_count_N) {
  if (!h || N < 0 || (N > 0 && (!pts4xN || !normals3xN || !fpfh33xN))) { set_error("b200_lc_compute_fpfh: bad arguments"); return B200_E_BAD_PARAM; }
  if (max_nn < 1 || max_nn > kHKCap) { set_error("b200_lc_compute_fpfh: max_nn must be in 1..2048"); return B200_E_BAD_PARAM; }
  if (!(radius > 0.f) || std::isinf(radius)) { set_error("b200_lc_compute_fpfh: radius must be positive and finite"); return B200_E_BAD_PARAM; }
  if (N == 0) return B200_OK;
  return lc_fpfh_impl(h, pts4xN, normals3xN, N, radius, max_nn, fpfh33xN, nn_count_N, 0, 0.f, nullptr);
}

int b200_lc_compute_features(b200_lc* h, const float* pts4xN, int N, float normal_radius, int normal_knn, float fpfh_radius,
                             int fpfh_max_nn, double* normals3xN, double* fpfh33xN) {
  if (!h || N < 0 || (N > 0 && (!pts4xN || !fpfh33xN))) { set_error("b200_lc_compute_features: bad arguments"); return B200_E_BAD_PARAM; }
  if (normal_knn < 1 || normal_knn > 32 || !(normal_radius > 0.f)) { set_error("b200_lc_compute_features: normal_knn in 1..32, normal_radius > 0"); return B200_E_BAD_PARAM; }
  if (fpfh_max_nn < 1 || fpfh_max_nn > kHKCap) { set_error("b200_lc_compute_features: fpfh_max_nn must be in 1..2048"); return B200_E_BAD_PARAM; }
  if (!(fpfh_radius > 0.f) || std::isinf(fpfh_radius)) { set_error("b200_lc_compute_features: fpfh_radius must be positive and finite"); return B200_E_BAD_PARAM; }
  if (N == 0) return B200_OK;
  return lc_fpfh_impl(h, pts4xN, nullptr, N, fpfh_radius, fpfh_max_nn, fpfh33xN, nullptr, normal_knn, normal_radius, normals3xN);
}

int b200_lc_feature_nn(b200_lc* h, const double* a33xN, int N, const double* b33xM, int M, int32_t* nn_N, double* d2_N) {
  if (!h || N <= 0 || M <= 0 || !a33xN || !b33xM || !nn_N) { set_error("b200_lc_feature_nn: bad arguments"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  B200_TRY(grow(&h->d_sf, &h->cap_sf, (size_t)33 * N + 2));   // + 16 bytes: the bulk copies round a tile up to 16
  B200_TRY(grow(&h->d_tf, &h->cap_tf, (size_t)33 * M + 2));
  B200_TRY(grow(&h->d_ij, &h->cap_ij, (size_t)N));
  B200_TRY(grow(&h->d_err2, &h->cap_err2, (size_t)N));
  B200_CUDA(cudaMemcpyAsync(h->d_sf, a33xN, (size_t)33 * N * sizeof(double), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaMemcpyAsync(h->d_tf, b33xM, (size_t)33 * M * sizeof(double), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  B200_TRY(feature_nn(h, h->d_sf, N, h->d_tf, M, h->d_ij, h->d_err2));
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  B200_CUDA(cudaMemcpyAsync(nn_N, h->d_ij, (size_t)N * sizeof(int), cudaMemcpyDeviceToHost, s));
  if (d2_N) B200_CUDA(cudaMemcpyAsync(d2_N, h->d_err2, (size_t)N * sizeof(double), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]);
  return B200_OK;
}

void b200_lc_default_ransac_params(b200_ransac_params* p) {
  if (!p) return;
  std::memset(p, 0, sizeof(*p));
  p->mutual_filter = 1;                      // PlaceRecognition.cpp:82
  p->max_correspondence_distance = 0.75;     // ransacMaxCorrespondenceDistance_ (Parameters.hpp:130)
  p->ransac_n = 3;                           // ransacModelSize_ (:129)
  p->edge_length_similarity = 0.5;           // correspondenceCheckerEdgeLength_ (:132)
  p->distance_threshold = 0.75;              // correspondenceCheckerDistance_ (:131)
  p->max_iteration = 1000000;                // ransacNumIter_ (:127)
  p->confidence = 0.99;                      // ransacProbability_ (:128)
  p->seed = 1;
  p->batch = 65536;
}

int b200_lc_ransac_feature_matching(b200_lc* h, const float* src4xN, int Ns, const float* tgt4xM, int Nt,
                                    const double* src_fpfh33xN, const double* tgt_fpfh33xM, const b200_ransac_params* prm,
                                    b200_ransac_result* res, int32_t* corr2xK) {
  if (!h || !prm || !res || Ns <= 0 || Nt <= 0 || !src4xN || !tgt4xM || !src_fpfh33xN || !tgt_fpfh33xM) {
    set_error("b200_lc_ransac_feature_matching: bad arguments");
    return B200_E_BAD_PARAM;
  }
  std::memset(res, 0, sizeof(*res));
  for (int d = 0; d < 4; ++d) res->T[5 * d] = 1.0;
  // Open3D returns an empty RegistrationResult for these (Registration.cpp)
  if (prm->ransac_n < 3 || !(prm->max_correspondence_distance > 0.0)) return B200_OK;
  if (prm->ransac_n > kMaxRn) { set_error("b200_lc_ransac_feature_matching: ransac_n must be <= 8"); return B200_E_UNSUPPORTED; }
  if (prm->max_iteration < 0 || !(prm->confidence >= 0.0 && prm->confidence <= 1.0)) { set_error("b200_lc_ransac_feature_matching: bad criteria"); return B200_E_BAD_PARAM; }
  int B = prm->batch > 0 ? prm->batch : 65536;
  B = std::max(1024, std::min(B, 1 << 20));
  B = (B + 1023) / 1024 * 1024;
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  B200_TRY(grow(&h->d_src, &h->cap_src, (size_t)Ns));
  B200_TRY(grow(&h->d_tgt, &h->cap_tgt, (size_t)Nt));
  B200_TRY(grow(&h->d_sf, &h->cap_sf, (size_t)33 * Ns + 2));
  B200_TRY(grow(&h->d_tf, &h->cap_tf, (size_t)33 * Nt + 2));
  B200_TRY(grow(&h->d_ij, &h->cap_ij, (size_t)Ns));
  B200_TRY(grow(&h->d_ji, &h->cap_ji, (size_t)Nt));
  B200_TRY(grow(&h->d_cs, &h->cap_cs, (size_t)Ns));
  B200_TRY(grow(&h->d_ct, &h->cap_ct, (size_t)Ns));
  B200_TRY(grow(&h->d_ps, &h->cap_ps, (size_t)3 * Ns));
  B200_TRY(grow(&h->d_pt, &h->cap_pt, (size_t)3 * Ns));
  B200_TRY(grow(&h->d_corr, &h->cap_corr, (size_t)2 * Ns));
  B200_TRY(grow(&h->d_good, &h->cap_good, (size_t)B));
  B200_TRY(grow(&h->d_err2, &h->cap_err2, (size_t)std::max(B, Ns)));
  B200_TRY(grow(&h->d_slot, &h->cap_slot, (size_t)B));
  B200_TRY(grow(&h->d_surv_h, &h->cap_surv_h, (size_t)B));
  B200_TRY(grow(&h->d_surv_T, &h->cap_surv_T, (size_t)16 * B));
  B200_TRY(upload_features(h->d_src, src4xN, Ns, s));
  B200_TRY(upload_features(h->d_tgt, tgt4xM, Nt, s));
  B200_CUDA(cudaMemcpyAsync(h->d_sf, src_fpfh33xN, (size_t)33 * Ns * sizeof(double), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaMemcpyAsync(h->d_tf, tgt_fpfh33xM, (size_t)33 * Nt * sizeof(double), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  int launches = 0;
  B200_TRY(feature_nn(h, h->d_sf, Ns, h->d_tf, Nt, h->d_ij, nullptr));
  launches += 2;
  int nc = 0, mutual_used = 0;
  if (prm->mutual_filter) {
    B200_TRY(feature_nn(h, h->d_tf, Nt, h->d_sf, Ns, h->d_ji, nullptr));
    corr_build_kernel<<<1, 1024, 0, s>>>(h->d_ij, h->d_ji, Ns, 1, h->d_src, h->d_tgt, h->d_cs, h->d_ct, h->d_ps, h->d_pt, h->d_n);
    B200_CUDA(cudaGetLastError());
    launches += 3;
    B200_CUDA(cudaMemcpyAsync(&nc, h->d_n, sizeof(int), cudaMemcpyDeviceToHost, s));
    B200_CUDA(cudaStreamSynchronize(s));
    mutual_used = nc >= prm->ransac_n * 3;   // "mutual correspondence set should not be too small"
  }
  if (!mutual_used) {
    corr_build_kernel<<<1, 1024, 0, s>>>(h->d_ij, h->d_ji, Ns, 0, h->d_src, h->d_tgt, h->d_cs, h->d_ct, h->d_ps, h->d_pt, h->d_n);
    B200_CUDA(cudaGetLastError());
    ++launches;
    B200_CUDA(cudaMemcpyAsync(&nc, h->d_n, sizeof(int), cudaMemcpyDeviceToHost, s));
    B200_CUDA(cudaStreamSynchronize(s));
  }
  res->mutual_used = mutual_used;
  res->n_feature_correspondences = nc;
  if (nc < prm->ransac_n) {
    B200_CUDA(cudaEventRecord(h->ev[1], s));
    B200_CUDA(cudaStreamSynchronize(s));
    cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]);
    res->device_ms = h->last_ms;
    res->kernel_launches = launches;
    return B200_OK;
  }
  RansacParamsDev p;
  p.nc = nc; p.rn = prm->ransac_n; p.max_iteration = prm->max_iteration;
  p.edge = prm->edge_length_similarity; p.dist_thr = prm->distance_threshold;
  p.max2 = prm->max_correspondence_distance * prm->max_correspondence_distance;
  p.confidence = prm->confidence; p.seed = prm->seed;
  RansacState init;
  std::memset(&init, 0, sizeof(init));
  init.exit_itr = prm->max_iteration;
  for (int d = 0; d < 4; ++d) init.best_T[5 * d] = 1.0;
  *h->h_state = init;
  B200_CUDA(cudaMemcpyAsync(h->d_state, h->h_state, sizeof(RansacState), cudaMemcpyHostToDevice, s));
  const size_t walk_smem = (size_t)(B / 32) * sizeof(unsigned);
  if (walk_smem > 48 * 1024) B200_CUDA(cudaFuncSetAttribute(ransac_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)walk_smem));
  // The answer does not depend on how the iterations are cut into batches, so the first batches are small (a good
  // pair exits after a few dozen iterations: log(1 - confidence) / log(1 - ratio^n)) and grow fourfold up to B;
  // the state is read back after two small batches, then after every eight; batches behind the exit estimate return
  // at once.
  long long start = 0;
  int cur = std::min(B, 1024), since_read = 0;
  while (start < prm->max_iteration) {
    ransac_hypo_kernel<<<cur / 128, 128, 0, s>>>(p, h->d_state, start, cur, h->d_ps, h->d_pt, h->d_good, h->d_slot, h->d_surv_h, h->d_surv_T);
    ransac_eval_kernel<<<148 * 4, 256, 0, s>>>(p, h->d_state, h->d_ps, h->d_pt, h->d_surv_h, h->d_surv_T, h->d_good, h->d_err2);
    ransac_walk_kernel<<<1, kWalkThreads, walk_smem, s>>>(p, h->d_state, start, cur, h->d_good, h->d_err2, h->d_slot, h->d_surv_T);
    launches += 3;
    start += cur;
    ++since_read;
    const bool growing = cur < B;
    cur = (int)std::min<long long>(B, 4LL * cur);
    if (since_read >= (growing ? 2 : 8) || start >= prm->max_iteration) {
      since_read = 0;
      B200_CUDA(cudaGetLastError());
      B200_CUDA(cudaMemcpyAsync(h->h_state, h->d_state, sizeof(RansacState), cudaMemcpyDeviceToHost, s));
      B200_CUDA(cudaStreamSynchronize(s));
      if (h->h_state->exit_itr <= start) break;
    }
  }
  ransac_final_kernel<<<1, 1024, 0, s>>>(p, h->d_state, h->d_ps, h->d_pt, h->d_cs, h->d_ct, h->d_corr, h->d_n);
  B200_CUDA(cudaGetLastError());
  ++launches;
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  int k = 0;
  B200_CUDA(cudaMemcpyAsync(&k, h->d_n, sizeof(int), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaMemcpyAsync(h->h_state, h->d_state, sizeof(RansacState), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  if (corr2xK && k > 0) {
    B200_CUDA(cudaMemcpyAsync(corr2xK, h->d_corr, (size_t)2 * k * sizeof(int), cudaMemcpyDeviceToHost, s));
    B200_CUDA(cudaStreamSynchronize(s));
  }
  cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]);
  const RansacState& st = *h->h_state;
  std::memcpy(res->T, st.best_T, sizeof(res->T));
  res->fitness = (double)st.best_good / (double)nc;
  res->inlier_rmse = st.best_rmse;
  res->n_correspondences = k;
  res->iterations_run = std::min<long long>(st.itr_done, prm->max_iteration);
  res->validations = st.validations;
  res->device_ms = h->last_ms;
  res->kernel_launches = launches;
  return B200_OK;
}

int b200_lc_overlap_indices(b200_lc* h, const float* src4xN, int Ns, const float* tgt4xM, int Nt, const double* T16,
                            double voxel_size, int min_points, int32_t* src_idx, int* n_src, int32_t* tgt_idx, int* n_tgt) {
  if (!h || Ns < 0 || Nt < 0 || (Ns > 0 && !src4xN) || (Nt > 0 && !tgt4xM) || !T16 || !n_src || !n_tgt ||
      (Ns > 0 && !src_idx) || (Nt > 0 && !tgt_idx)) {
    set_error("b200_lc_overlap_indices: bad arguments");
    return B200_E_BAD_PARAM;
  }
  if (!(voxel_size > 0.0) || min_points < 1) { set_error("b200_lc_overlap_indices: voxel_size > 0 and min_points >= 1 (helpers.cpp:323)"); return B200_E_BAD_PARAM; }
  *n_src = *n_tgt = 0;
  const int n = Ns + Nt;
  if (n == 0) return B200_OK;
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  size_t slots = 1024;
  while (slots < 2 * (size_t)n) slots <<= 1;
  B200_TRY(grow(&h->d_src, &h->cap_src, (size_t)std::max(Ns, 1)));
  B200_TRY(grow(&h->d_tgt, &h->cap_tgt, (size_t)std::max(Nt, 1)));
  B200_TRY(grow(&h->d_vkeys, &h->cap_vkeys, slots));
  B200_TRY(grow(&h->d_vcnt, &h->cap_vcnt, 2 * slots));
  B200_TRY(grow(&h->d_vslot, &h->cap_vslot, (size_t)n));
  B200_TRY(grow(&h->d_vflag, &h->cap_vflag, (size_t)n));
  B200_TRY(grow(&h->d_vtile, &h->cap_vtile, scan_tiles_for((size_t)n) + 1));
  B200_TRY(grow(&h->d_vsrc, &h->cap_vsrc, (size_t)std::max(Ns, 1)));
  B200_TRY(grow(&h->d_vtgt, &h->cap_vtgt, (size_t)std::max(Nt, 1)));
  if (Ns) B200_TRY(upload_features(h->d_src, src4xN, Ns, s));
  if (Nt) B200_TRY(upload_features(h->d_tgt, tgt4xM, Nt, s));
  B200_CUDA(cudaMemsetAsync(h->d_flag, 0, sizeof(int), s));
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  fill_u64_lc_kernel<<<296, 256, 0, s>>>(h->d_vkeys, slots, kVoxEmpty);
  B200_CUDA(cudaMemsetAsync(h->d_vcnt, 0, 2 * slots * sizeof(uint32_t), s));
  Mat4 T;
  std::memcpy(T.m, T16, sizeof(T.m));
  const int blocks = std::min(148 * 8, (n + 255) / 256);
  overlap_insert_kernel<<<blocks, 256, 0, s>>>(h->d_src, Ns, h->d_tgt, Nt, T, 1.0 / voxel_size, (unsigned long long)slots - 1ull,
                                               h->d_vkeys, h->d_vcnt, h->d_vslot, h->d_flag);
  overlap_flag_kernel<<<blocks, 256, 0, s>>>(h->d_vslot, h->d_vcnt, n, (uint32_t)min_points, h->d_vflag);
  B200_CUDA(cudaGetLastError());
  B200_TRY(inclusive_scan_u32(h->d_vflag, (size_t)n, h->d_vtile, s));
  overlap_scatter_kernel<<<blocks, 256, 0, s>>>(h->d_vflag, Ns, Nt, h->d_vsrc, h->d_vtgt);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  uint32_t tot[2] = {0u, 0u};
  int bad = 0;
  if (Ns) B200_CUDA(cudaMemcpyAsync(&tot[0], h->d_vflag + (Ns - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaMemcpyAsync(&tot[1], h->d_vflag + (n - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaMemcpyAsync(&bad, h->d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]);
  if (bad) { set_error("b200_lc_overlap_indices: a voxel index lies beyond +-2^20"); return B200_E_BAD_PARAM; }
  *n_src = (int)tot[0];
  *n_tgt = (int)(tot[1] - tot[0]);
  if (*n_src) B200_CUDA(cudaMemcpyAsync(src_idx, h->d_vsrc, (size_t)*n_src * sizeof(int), cudaMemcpyDeviceToHost, s));
  if (*n_tgt) B200_CUDA(cudaMemcpyAsync(tgt_idx, h->d_vtgt, (size_t)*n_tgt * sizeof(int), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  return B200_OK;
}

int b200_lc_last_device_ms(const b200_lc* h, float* ms) {
  if (!h || !ms) { set_error("b200_lc_last_device_ms: bad arguments"); return B200_E_BAD_PARAM; }
  *ms = h->last_ms;
  return B200_OK;
}

}  // extern "C"
