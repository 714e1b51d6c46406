// Boundary B: the registration loop behind open3d_slam's CloudRegistration::registerClouds
// (open3d_slam/src/CloudRegistration.cpp:17-23 RegistrationGeneralizedICP, :62-66 / :89-94 RegistrationICP), i.e.
// Open3D v0.15.1's RegistrationICP with TransformationEstimationPointToPlane or ...ForGeneralizedICP. Open3D is not in
// the reference tree; the algorithm is recalled from upstream (parity unpinned).
//
//   result = match(T src); repeat { x = -(JtJ)^-1 Jtr; T = T(x) T; result' = match(T src);
//                                   stop when |d fitness| < rel_fitness and |d rmse| < rel_rmse }  (<= max_iteration updates)
//
// Device shape: every iteration is two kernels on one stream, enqueued eight iterations at a time without any host
// synchronisation; a device-side `done` flag turns the launches after convergence into no-ops, and the state read back
// after a chunk is the result (a registration of 4-6 updates: one chunk, one read-back).
//   o3d_match_lin_kernel  transform (double T, float result like the cloud Open3D hands to its kd-tree after our
//                         double->float conversion) -> exact 1-NN within max_correspondence_distance (warp-cooperative
//                         search of knn.cuh) -> residual rows and Jacobians in double -> 21 + 6 + 2 sums per CTA
//   o3d_solve_kernel      fixed-order reduction of the CTA partials, fitness / inlier RMSE, convergence test against
//                         the previous match, LDL^T solve of the 6x6 system, T(x) = Rz Ry Rx, T <- T(x) T
// GICP covariances are Open3D's "from normals" form C = R diag(eps, 1, 1) R^T = I - (1 - eps) n n^T (with its quirk:
// normals within ~8 degrees of -e1 are treated as e1), so M = C_t + C_s has the closed-form eigen-decomposition used
// below for W = M^-1/2 (eigenvectors (a + b), (a - b), a x b): no per-pair iterative eigen-solver.
#include "common.cuh"
#include "grid_index.cuh"
#include "knn.cuh"
#include "modules.cuh"
#include "normals.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <cstdlib>

using namespace b200;

namespace {

constexpr int kThreads = 256;
constexpr int kVals = 32;  // 21 JtJ (upper, column-major packed) + 6 Jtr + count + sum d^2 (+ 3 spare)

struct O3dState {
  double T[16];          // current transformation (column-major)
  double fitness, rmse;  // of the last match
  double prev_fitness, prev_rmse;
  int updates;           // transformation updates applied so far (Open3D's loop counter)
  int matches_done;      // match kernels executed
  int done;
  int n_corr;
  int status;
};

struct O3dParams {
  GridView g;
  const float4* src;     // N source points
  const float4* src_n;   // N source normals (GICP) or null
  int N, kind;           // kind 0 point-to-plane, 1 point-to-point, 2 generalized
  float md2;
  int max_iteration;
  double rel_fitness, rel_rmse, eps;
  int* ids; float* dists;
  double* partials;      // gridDim x kVals
  O3dState* st;
  // spatially partitioned target (one slab of a larger map per device): this device answers only for the queries whose
  // transformed coordinate along own_axis lies in [own_lo, own_hi); gid maps slab-local ids to ids in the whole map
  int own_axis;          // -1: not partitioned
  float own_lo, own_hi;
  const int* gid;
  // query -> lane map of the match kernel: 0 = 32 consecutive source points per warp, 1 = lane l of warp slot w takes
  // point l * nslots + w (spatially coherent, equally expensive points spread over all warps), 2 = 1 until the
  // previous match's RMSE is below 0.4 cells (every query settled by its own 2x2x2 block), then 0
  int qmap;
};

__device__ __forceinline__ double3 gicp_axis(float4 n) {
  // unit length restored in double (the reference's normals are unit doubles; the float boundary rounds them).
  // Open3D GetRotationFromE1ToX: c = e1 . x < -0.99 -> identity rotation, i.e. the covariance of the e1 normal;
  // a zero normal gives the identity rotation too.
  const double x = (double)n.x, y = (double)n.y, z = (double)n.z;
  const double nn = sqrt(x * x + y * y + z * z);
  if (!(nn > 0.0)) return make_double3(1.0, 0.0, 0.0);
  const double c = x / nn;
  if (c < -0.99) return make_double3(1.0, 0.0, 0.0);
  return make_double3(c, y / nn, z / nn);
}

// The match kernel is a chain of dependent L2 gathers (the 1-NN search), so what it needs is resident warps. The 29
// double sums of the linearisation used to live in registers (128 per thread: 16 warps per SM). Now a lane keeps ONE
// sum: every lane stages the values of its correspondence in a per-warp shared-memory tile (two batches of at most 16
// rows x 32 lanes, column index rotated by the row so that neither the writes nor the row reads conflict by more than
// two ways), and lane r adds row r over the 32 lanes in lane order — same additions for every launch geometry, 64
// registers, 32 warps per SM.
constexpr int kStageRows = 16;
struct StageTile { double v[kStageRows][32]; };

// one batch: rows [T0, T1) of the values `emit` enumerates (emit calls put(t, value) with compile-time t after
// unrolling); lane (T0 ? 16 : 0) + r accumulates row r
template <int T0, int T1, class Emit>
__device__ __forceinline__ void stage_batch(StageTile& tile, int lane, bool active, Emit&& emit, double& acc) {
  static_assert(T1 - T0 <= kStageRows, "batch too tall");
  emit([&](int t, double v) {
    if (t >= T0 && t < T1) tile.v[t - T0][(lane + (t - T0)) & 31] = active ? v : 0.0;
  });
  __syncwarp();
  const int r = lane - (T0 ? 16 : 0);
  if (r >= 0 && r < T1 - T0) {
    double a = acc;
#pragma unroll 8
    for (int c = 0; c < 32; ++c) a += tile.v[r][(c + r) & 31];
    acc = a;
  }
  __syncwarp();
}

template <int KIND>
__global__ void __launch_bounds__(kThreads, 4) o3d_match_lin_kernel(const __grid_constant__ O3dParams p) {
  // the search's per-warp scratch and the staging tile are never live at the same time (__syncwarp in between)
  __shared__ __align__(16) unsigned char s_raw[kThreads / 32][sizeof(StageTile)];
  static_assert(sizeof(Warp1Scratch) <= sizeof(StageTile), "the search scratch must fit into the staging tile");
  __shared__ double s_T[12];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (p.st->done) return;
  if (tid < 12) s_T[tid] = p.st->T[(tid % 3) + 4 * (tid / 3)];
  __syncthreads();
  // lane l < 15 holds value l, lane 16 + u holds value 15 + u (u < 14): 29 sums per warp
  double acc = 0.0;
  const int nthreads = gridDim.x * kThreads;
  const int nslots = (p.N + 31) >> 5;
  const bool scatter = p.qmap == 1 || (p.qmap == 2 && (p.st->matches_done == 0 || !(p.st->rmse < 0.4 * (double)p.g.cell)));
  StageTile& tile = *reinterpret_cast<StageTile*>(s_raw[wid]);
  Warp1Scratch& scratch = *reinterpret_cast<Warp1Scratch*>(s_raw[wid]);
  for (int slot = blockIdx.x * (kThreads / 32) + wid; slot < nslots; slot += nthreads >> 5) {
    const int i = scatter ? lane * nslots + (slot + lane * 97) % nslots : slot * 32 + lane;
    const bool in = i < p.N;
    float4 q = make_float4(0.f, 0.f, 0.f, 1.f);
    if (in) {
      const float4 s = __ldg(&p.src[i]);
      q.x = (float)(s_T[0] * s.x + s_T[3] * s.y + s_T[6] * s.z + s_T[9]);
      q.y = (float)(s_T[1] * s.x + s_T[4] * s.y + s_T[7] * s.z + s_T[10]);
      q.z = (float)(s_T[2] * s.x + s_T[5] * s.y + s_T[8] * s.z + s_T[11]);
    }
    bool mine = in;
    if (p.own_axis >= 0) {
      const float c = p.own_axis == 0 ? q.x : (p.own_axis == 1 ? q.y : q.z);
      mine = in && c >= p.own_lo && c < p.own_hi;
    }
    float bd; int bid, bsp;
    unsigned vis = 0;
    knn_query_warp1(p.g, mine, q.x, q.y, q.z, p.md2, bd, bid, bsp, vis, scratch);
    __syncwarp();
    if (in) {
      if (!mine) { p.ids[i] = -2; p.dists[i] = __int_as_float(0x7f800000); }   // another slab's query
      else { p.ids[i] = (bid >= 0 && p.gid) ? __ldg(&p.gid[bid]) : bid; p.dists[i] = bd; }
    }
    const bool hit = in && mine && bid >= 0;
    if (!__any_sync(0xffffffffu, hit)) continue;
    float4 t4 = make_float4(0.f, 0.f, 0.f, 0.f), n4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (hit) {
      t4 = __ldg(&p.g.pts[bsp]);
      if (KIND == 0 || KIND == 2) n4 = __ldg(&p.g.nrm[bsp]);
    }
    const double s[3] = {(double)q.x, (double)q.y, (double)q.z};
    const double d[3] = {s[0] - (double)t4.x, s[1] - (double)t4.y, s[2] - (double)t4.z};
    // values 27 (count) and 28 (sum d^2) ride in the second batch of the first (or only) row
    if (KIND == 3) {
      // GetInformationMatrixFromPointClouds (Open3D v0.15.1, as called at PlaceRecognition.cpp:148-149): G^T G over the
      // correspondences with the three rows G = [-[t]x | I] of the matched TARGET point t
      const double x = (double)t4.x, y = (double)t4.y, z = (double)t4.z;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const double G[6] = {r == 0 ? 0.0 : (r == 1 ? -z : y), r == 0 ? z : (r == 1 ? 0.0 : -x), r == 0 ? -y : (r == 1 ? x : 0.0),
                             r == 0 ? 1.0 : 0.0, r == 1 ? 1.0 : 0.0, r == 2 ? 1.0 : 0.0};
        auto emit = [&](auto&& put) {
          int t = 0;
#pragma unroll
          for (int c = 0; c < 6; ++c)
#pragma unroll
            for (int a = 0; a <= c; ++a) put(t++, G[a] * G[c]);
#pragma unroll
          for (int u = 21; u < 27; ++u) put(u, 0.0);
          put(27, r == 0 ? 1.0 : 0.0); put(28, r == 0 ? (double)bd : 0.0);   // every row of the tile is rewritten
        };
        stage_batch<0, 15>(tile, lane, hit, emit, acc);
        stage_batch<15, 29>(tile, lane, hit, emit, acc);
      }
    } else if (KIND == 1) {
      // Umeyama moments: sum s (0..2), sum t (3..5), sum t s^T (6..14, column-major)
      const double t[3] = {(double)t4.x, (double)t4.y, (double)t4.z};
      auto emit = [&](auto&& put) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { put(a, s[a]); put(3 + a, t[a]); }
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int r = 0; r < 3; ++r) put(6 + r + 3 * c, t[r] * s[c]);
#pragma unroll
        for (int u = 15; u < 27; ++u) put(u, 0.0);
        put(27, 1.0); put(28, (double)bd);
      };
      stage_batch<0, 15>(tile, lane, hit, emit, acc);
      stage_batch<15, 29>(tile, lane, hit, emit, acc);
    } else if (KIND == 0) {
      const double n[3] = {(double)n4.x, (double)n4.y, (double)n4.z};
      const double r = d[0] * n[0] + d[1] * n[1] + d[2] * n[2];
      const double J[6] = {s[1] * n[2] - s[2] * n[1], s[2] * n[0] - s[0] * n[2], s[0] * n[1] - s[1] * n[0], n[0], n[1], n[2]};
      auto emit = [&](auto&& put) {
        int t = 0;
#pragma unroll
        for (int c = 0; c < 6; ++c)
#pragma unroll
          for (int a = 0; a <= c; ++a) put(t++, J[a] * J[c]);
#pragma unroll
        for (int a = 0; a < 6; ++a) put(21 + a, J[a] * r);
        put(27, 1.0); put(28, (double)bd);
      };
      stage_batch<0, 15>(tile, lane, hit, emit, acc);
      stage_batch<15, 29>(tile, lane, hit, emit, acc);
    } else {
      // M = C_t + R C_s R^T = 2 I - (1 - eps) (a a^T + b b^T), a = target axis, b = rotated source axis
      const double3 a = gicp_axis(n4);
      const double3 b0 = gicp_axis(hit ? __ldg(&p.src_n[i]) : make_float4(1.f, 0.f, 0.f, 0.f));
      const double3 b = make_double3(s_T[0] * b0.x + s_T[3] * b0.y + s_T[6] * b0.z, s_T[1] * b0.x + s_T[4] * b0.y + s_T[7] * b0.z,
                                     s_T[2] * b0.x + s_T[5] * b0.y + s_T[8] * b0.z);
      const double k1 = 1.0 - p.eps;
      const double ab = a.x * b.x + a.y * b.y + a.z * b.z;
      // eigenpairs in span{a, b}: u+ ~ a + b (lambda 2 - k1 (1 + ab)), u- ~ a - b (lambda 2 - k1 (1 - ab)); rest: 2
      const double up[3] = {a.x + b.x, a.y + b.y, a.z + b.z}, um[3] = {a.x - b.x, a.y - b.y, a.z - b.z};
      const double np2 = up[0] * up[0] + up[1] * up[1] + up[2] * up[2], nm2 = um[0] * um[0] + um[1] * um[1] + um[2] * um[2];
      const double w0 = 0.70710678118654752440;  // 2^-1/2
      const double wp = np2 > 1e-24 ? 1.0 / sqrt(2.0 - k1 * (1.0 + ab)) - w0 : 0.0;
      const double wm = nm2 > 1e-24 ? 1.0 / sqrt(2.0 - k1 * (1.0 - ab)) - w0 : 0.0;
      const double ip = np2 > 1e-24 ? wp / np2 : 0.0, im = nm2 > 1e-24 ? wm / nm2 : 0.0;
      // J = W [ -[s]x  I ], r = W d, W = w0 I + wp u+ u+^T + wm u- u-^T (symmetric); three rows, staged one at a time
#pragma unroll
      for (int row = 0; row < 3; ++row) {
        const double wr[3] = {(row == 0 ? w0 : 0.0) + ip * up[row] * up[0] + im * um[row] * um[0],
                              (row == 1 ? w0 : 0.0) + ip * up[row] * up[1] + im * um[row] * um[1],
                              (row == 2 ? w0 : 0.0) + ip * up[row] * up[2] + im * um[row] * um[2]};
        // -[s]x has columns (0, -s2, s1), (s2, 0, -s0), (-s1, s0, 0)
        const double J[6] = {wr[2] * s[1] - wr[1] * s[2], wr[0] * s[2] - wr[2] * s[0], wr[1] * s[0] - wr[0] * s[1],
                             wr[0], wr[1], wr[2]};
        const double r = wr[0] * d[0] + wr[1] * d[1] + wr[2] * d[2];
        auto emit = [&](auto&& put) {
          int t = 0;
#pragma unroll
          for (int c = 0; c < 6; ++c)
#pragma unroll
            for (int a2 = 0; a2 <= c; ++a2) put(t++, J[a2] * J[c]);
#pragma unroll
          for (int a2 = 0; a2 < 6; ++a2) put(21 + a2, J[a2] * r);
          put(27, row == 0 ? 1.0 : 0.0); put(28, row == 0 ? (double)bd : 0.0);
        };
        stage_batch<0, 15>(tile, lane, hit, emit, acc);
        stage_batch<15, 29>(tile, lane, hit, emit, acc);
      }
    }
  }
  // CTA partial: the warps' sums in warp order (deterministic); value t sits on lane t (t < 15) or lane t + 1
  __syncthreads();
  double* part = reinterpret_cast<double*>(&s_raw[0][0]);   // reused: [warp][32]
  part[wid * 32 + lane] = acc;
  __syncthreads();
  if (tid < kVals) {
    double v = 0.0;
    if (tid < 29) {
      const int src_lane = tid < 15 ? tid : tid + 1;
      for (int w = 0; w < kThreads / 32; ++w) v += part[w * 32 + src_lane];
    }
    p.partials[(size_t)blockIdx.x * kVals + tid] = v;
  }
}

static void launch_match_lin(const O3dParams& p, int blocks, cudaStream_t s) {
  switch (p.kind) {
    case 0: o3d_match_lin_kernel<0><<<blocks, kThreads, 0, s>>>(p); break;
    case 1: o3d_match_lin_kernel<1><<<blocks, kThreads, 0, s>>>(p); break;
    case 2: o3d_match_lin_kernel<2><<<blocks, kThreads, 0, s>>>(p); break;
    default: o3d_match_lin_kernel<3><<<blocks, kThreads, 0, s>>>(p); break;
  }
}

// R = U diag(1, 1, det(U) det(V)) V^T for the SVD S = U diag(sigma) V^T, by one-sided Jacobi on the columns of S
// (sigma sorted descending; the third left vector is rebuilt as u1 x u2, which also covers rank-2 S).
__device__ void kabsch3(const double* S, double* R) {
  double A[9], V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  for (int i = 0; i < 9; ++i) A[i] = S[i];
  for (int sweep = 0; sweep < 40; ++sweep) {
    bool rotated = false;
    for (int pc = 0; pc < 2; ++pc)
      for (int qc = pc + 1; qc < 3; ++qc) {
        double al = 0, be = 0, ga = 0;
        for (int k = 0; k < 3; ++k) { al += A[k + 3 * pc] * A[k + 3 * pc]; be += A[k + 3 * qc] * A[k + 3 * qc]; ga += A[k + 3 * pc] * A[k + 3 * qc]; }
        if (ga == 0.0 || fabs(ga) <= 1e-16 * sqrt(al * be)) continue;
        rotated = true;
        const double zeta = (be - al) / (2.0 * ga);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
        for (int k = 0; k < 3; ++k) {
          const double ap = A[k + 3 * pc], aq = A[k + 3 * qc];
          A[k + 3 * pc] = c * ap - sn * aq; A[k + 3 * qc] = sn * ap + c * aq;
          const double vp = V[k + 3 * pc], vq = V[k + 3 * qc];
          V[k + 3 * pc] = c * vp - sn * vq; V[k + 3 * qc] = sn * vp + c * vq;
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
    u1[k] = sg[o[0]] > 0 ? A[k + 3 * o[0]] / sg[o[0]] : (k == 0);
    v1[k] = V[k + 3 * o[0]]; v2[k] = V[k + 3 * o[1]]; v3[k] = V[k + 3 * o[2]];
  }
  if (sg[o[1]] > 1e-300) {
    for (int k = 0; k < 3; ++k) u2[k] = A[k + 3 * o[1]] / sg[o[1]];
  } else {  // rank <= 1: any unit vector orthogonal to u1
    const int m = fabs(u1[0]) <= fabs(u1[1]) ? (fabs(u1[0]) <= fabs(u1[2]) ? 0 : 2) : (fabs(u1[1]) <= fabs(u1[2]) ? 1 : 2);
    double e[3] = {0, 0, 0};
    e[m] = 1.0;
    const double d = u1[m];
    double n2 = 0;
    for (int k = 0; k < 3; ++k) { u2[k] = e[k] - d * u1[k]; n2 += u2[k] * u2[k]; }
    for (int k = 0; k < 3; ++k) u2[k] /= sqrt(n2);
  }
  u3[0] = u1[1] * u2[2] - u1[2] * u2[1]; u3[1] = u1[2] * u2[0] - u1[0] * u2[2]; u3[2] = u1[0] * u2[1] - u1[1] * u2[0];
  const double detV = v1[0] * (v2[1] * v3[2] - v2[2] * v3[1]) - v2[0] * (v1[1] * v3[2] - v1[2] * v3[1]) + v3[0] * (v1[1] * v2[2] - v1[2] * v2[1]);
  const double sgn = detV < 0 ? -1.0 : 1.0;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) R[r + 3 * c] = u1[r] * v1[c] + u2[r] * v2[c] + sgn * u3[r] * v3[c];
}

// This device's sums from the CTA partials, fixed order (slice-major over CTAs, eight loads in flight per thread, then
// the slices in order): the reduction is a chain of L2 latencies, so it is spread over kRedThreads = 1024 threads
// (32 slices x 32 values: 19 partials per thread at 592 CTAs instead of 74 with 256 threads).
constexpr int kRedThreads = 1024;
constexpr int kRedSlices = kRedThreads / kVals;
__device__ __forceinline__ void reduce_partials_cta(const double* __restrict__ partials, int nblocks, bool live,
                                                    double (*s_sum)[kVals], double* s_tot) {
  const int tid = threadIdx.x;
  const int c = tid & (kVals - 1), slice = tid / kVals;
  double s = 0.0;
  if (live) {
    int b = slice;
    for (; b + 7 * kRedSlices < nblocks; b += 8 * kRedSlices) {
      double v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldcg(&partials[(size_t)(b + u * kRedSlices) * kVals + c]);
#pragma unroll
      for (int u = 0; u < 8; ++u) s += v[u];
    }
    for (; b < nblocks; b += kRedSlices) s += __ldcg(&partials[(size_t)b * kVals + c]);
  }
  s_sum[slice][c] = s;
  __syncthreads();
  if (tid < kVals) {
    double t = 0.0;
    for (int sl = 0; sl < kRedSlices; ++sl) t += s_sum[sl][tid];
    s_tot[tid] = t;
  }
  __syncthreads();
}

// one CTA: this device's partials -> out[kVals] (zeros once the registration is done, so a late allreduce is harmless)
__global__ void __launch_bounds__(kRedThreads) o3d_reduce_kernel(const __grid_constant__ O3dParams p, int nblocks,
                                                                 double* __restrict__ out) {
  __shared__ double s_sum[kRedSlices][kVals];
  __shared__ double s_tot[kVals];
  reduce_partials_cta(p.partials, nblocks, !p.st->done, s_sum, s_tot);
  if (threadIdx.x < kVals) out[threadIdx.x] = s_tot[threadIdx.x];
}

__device__ void o3d_solve_tail(const O3dParams& p, const double* s_tot) {
  // the state is worked on in registers / local memory and written back once: every st-> access used to be a
  // dependent round trip to L2 in this single-thread tail
  O3dState loc = *p.st;
  O3dState* st = &loc;
  const double cnt = s_tot[27];
  const double fit = cnt / (double)p.N;
  const double err = cnt > 0.0 ? sqrt(s_tot[28] / cnt) : 0.0;
  const int j = st->matches_done;  // this is match number j (0 = the match before the first update)
  st->matches_done = j + 1;
  st->prev_fitness = st->fitness; st->prev_rmse = st->rmse;
  st->fitness = fit; st->rmse = err;
  st->n_corr = (int)(cnt + 0.5);
  bool stop = false;
  if (j >= 1 && fabs(st->prev_fitness - fit) < p.rel_fitness && fabs(st->prev_rmse - err) < p.rel_rmse) stop = true;
  if (st->updates >= p.max_iteration) stop = true;
  if (cnt == 0.0) stop = true;   // nothing to estimate from (Open3D would apply identity updates until it "converges")
  if (!stop && p.kind == 1) {
    // TransformationEstimationPointToPoint: Eigen::umeyama(source, target, with_scaling = false)
    double ms[3], mt[3], S[9], U[16];
    for (int a = 0; a < 3; ++a) { ms[a] = s_tot[a] / cnt; mt[a] = s_tot[3 + a] / cnt; }
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r) S[r + 3 * c] = s_tot[6 + r + 3 * c] / cnt - mt[r] * ms[c];
    double R3[9];
    kabsch3(S, R3);
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r) U[r + 4 * c] = R3[r + 3 * c];
    for (int r = 0; r < 3; ++r) U[r + 12] = mt[r] - (R3[r] * ms[0] + R3[r + 3] * ms[1] + R3[r + 6] * ms[2]);
    U[3] = 0.0; U[7] = 0.0; U[11] = 0.0; U[15] = 1.0;
    double R[16];
    for (int c = 0; c < 4; ++c)
      for (int r = 0; r < 4; ++r) {
        double a = 0.0;
        for (int k = 0; k < 4; ++k) a += U[r + 4 * k] * st->T[k + 4 * c];
        R[r + 4 * c] = a;
      }
    for (int i = 0; i < 16; ++i) st->T[i] = R[i];
    st->updates += 1;
  } else if (!stop) {
    // x = -(JtJ)^-1 Jtr by LDL^T (Eigen ldlt(), no pivoting needed for these positive definite systems)
    double A[36], L[36], D[6], y[6], x[6];
    int t = 0;
    for (int c = 0; c < 6; ++c)
      for (int r = 0; r <= c; ++r) { A[r + 6 * c] = s_tot[t]; A[c + 6 * r] = s_tot[t]; ++t; }
    bool ok = true;
    for (int i = 0; i < 36; ++i) L[i] = 0.0;
    for (int c = 0; c < 6 && ok; ++c) {
      double d = A[c + 6 * c];
      for (int k = 0; k < c; ++k) d -= L[c + 6 * k] * L[c + 6 * k] * D[k];
      if (!(fabs(d) > 0.0)) { ok = false; break; }
      D[c] = d;
      L[c + 6 * c] = 1.0;
      for (int i = c + 1; i < 6; ++i) {
        double s = A[i + 6 * c];
        for (int k = 0; k < c; ++k) s -= L[i + 6 * k] * L[c + 6 * k] * D[k];
        L[i + 6 * c] = s / d;
      }
    }
    if (!ok) stop = true;
    else {
      for (int i = 0; i < 6; ++i) {
        double s = -s_tot[21 + i];
        for (int k = 0; k < i; ++k) s -= L[i + 6 * k] * y[k];
        y[i] = s;
      }
      for (int i = 0; i < 6; ++i) y[i] /= D[i];
      for (int i = 5; i >= 0; --i) {
        double s = y[i];
        for (int k = i + 1; k < 6; ++k) s -= L[k + 6 * i] * x[k];
        x[i] = s;
      }
      // TransformVector6dToMatrix4d: R = Rz(x2) Ry(x1) Rx(x0), t = x3..5; T <- U T
      const double cx = cos(x[0]), sx = sin(x[0]), cy = cos(x[1]), sy = sin(x[1]), cz = cos(x[2]), sz = sin(x[2]);
      double U[16];
      U[0] = cz * cy; U[4] = cz * sy * sx - sz * cx; U[8] = cz * sy * cx + sz * sx; U[12] = x[3];
      U[1] = sz * cy; U[5] = sz * sy * sx + cz * cx; U[9] = sz * sy * cx - cz * sx; U[13] = x[4];
      U[2] = -sy;     U[6] = cy * sx;                U[10] = cy * cx;               U[14] = x[5];
      U[3] = 0.0; U[7] = 0.0; U[11] = 0.0; U[15] = 1.0;
      double R[16];
      for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r) {
          double a = 0.0;
          for (int k = 0; k < 4; ++k) a += U[r + 4 * k] * st->T[k + 4 * c];
          R[r + 4 * c] = a;
        }
      for (int i = 0; i < 16; ++i) st->T[i] = R[i];
      st->updates += 1;
    }
  }
  if (stop) st->done = 1;
  *p.st = loc;
}

// one CTA: reduce, test convergence, solve, update
// sums == nullptr: reduce this device's partials and go on; otherwise `sums` already holds the totals over all slabs
// (after the allreduce) and the partials are not read
__global__ void __launch_bounds__(kRedThreads) o3d_solve_kernel(const __grid_constant__ O3dParams p, int nblocks,
                                                                const double* __restrict__ sums) {
  __shared__ double s_sum[kRedSlices][kVals];
  __shared__ double s_tot[kVals];
  O3dState* st = p.st;
  if (st->done) return;
  const int tid = threadIdx.x;
  if (sums) {
    if (tid < kVals) s_tot[tid] = sums[tid];
    __syncthreads();
  } else {
    reduce_partials_cta(p.partials, nblocks, true, s_sum, s_tot);
  }
  if (tid != 0) return;
  o3d_solve_tail(p, s_tot);
}

// ---- peer exchange of the sums over NVLink / NVSwitch peer memory (partitioned map, one slab per GPU) ---------------
// Replaces "reduce kernel -> NCCL allreduce of 32 doubles -> solve kernel" by ONE kernel per iteration: reduce this
// device's partials, store them straight into every peer's mailbox (plain 8-byte stores to peer memory mapped with
// CUDA IPC or, inside one process, peer access), wait until every peer's contribution has arrived in the own mailbox,
// add the contributions in rank order (bit-identical totals on every rank, so no broadcast) and solve. The transfer
// is 64 packets of 8 bytes per peer in NCCL's "LL" style: {32 data bits, 32-bit sequence number} written with one
// 8-byte store, so a packet is valid exactly when its sequence number is the expected one — no fence, no flag word,
// no host round trip; the sequence number (steps since the mailbox was created) doubles as the parity that keeps a
// rank that is one step ahead from overwriting what a slower rank still has to read.
constexpr int kPeerMax = 16;
constexpr int kPeerChunk = 8;         // iterations enqueued at a time when the peers are other processes
constexpr int kPackets = 2 * kVals;   // a double travels as two packets
struct PeerView {
  unsigned long long* box[kPeerMax];  // box[r]: rank r's mailbox [2 parities][world][kPackets] as seen from this device
  int rank, world;
  unsigned seq_base;                  // sequence number of the step before this registration's first
  long long timeout_cycles;
};

__device__ __forceinline__ void peer_store(unsigned long long* a, unsigned long long v) {
  asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(a), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long peer_load(const unsigned long long* a) {
  unsigned long long v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(a) : "memory");
  return v;
}

__global__ void __launch_bounds__(kRedThreads) o3d_exchange_solve_kernel(const __grid_constant__ O3dParams p, int nblocks,
                                                                         const __grid_constant__ PeerView pv) {
  __shared__ double s_sum[kRedSlices][kVals];
  __shared__ double s_tot[kVals];
  __shared__ unsigned s_half[kPeerMax][kPackets];
  __shared__ int s_fail;
  O3dState* st = p.st;
  if (st->done) return;
  const int tid = threadIdx.x;
  if (tid == 0) s_fail = 0;
  reduce_partials_cta(p.partials, nblocks, true, s_sum, s_tot);   // ends with a barrier: s_fail is visible too
  const unsigned seq = pv.seq_base + (unsigned)st->matches_done + 1u;
  const size_t par = (size_t)(seq & 1u) * (size_t)pv.world;
  // scatter: packet k of my sums into slot [parity][my rank] of every rank's mailbox (mine included)
  for (int t = tid; t < pv.world * kPackets; t += kRedThreads) {
    const int r = t / kPackets, k = t - r * kPackets;
    const unsigned long long bits = (unsigned long long)__double_as_longlong(s_tot[k >> 1]);
    const unsigned half = (k & 1) ? (unsigned)(bits >> 32) : (unsigned)bits;
    peer_store(pv.box[r] + (par + (size_t)pv.rank) * kPackets + k, ((unsigned long long)seq << 32) | half);
  }
  // gather: wait for packet k of every rank in my own mailbox
  const long long t0 = clock64();
  for (int t = tid; t < pv.world * kPackets; t += kRedThreads) {
    const int r = t / kPackets, k = t - r * kPackets;
    const unsigned long long* a = pv.box[pv.rank] + (par + (size_t)r) * kPackets + k;
    unsigned long long v = peer_load(a);
    while ((unsigned)(v >> 32) != seq) {
      if (clock64() - t0 > pv.timeout_cycles) { s_fail = 1; break; }
      __nanosleep(20);
      v = peer_load(a);
    }
    s_half[r][k] = (unsigned)v;
  }
  __syncthreads();
  if (s_fail) {  // a peer never delivered (crashed / out of step): stop instead of hanging the device
    if (tid == 0) { st->status = B200_E_CUDA; st->done = 1; }
    return;
  }
  if (tid < kVals) {  // totals in rank order: the same additions on every rank
    double t = 0.0;
    for (int r = 0; r < pv.world; ++r)
      t += __longlong_as_double((long long)(((unsigned long long)s_half[r][2 * tid + 1] << 32) | s_half[r][2 * tid]));
    s_tot[tid] = t;
  }
  __syncthreads();
  if (tid != 0) return;
  o3d_solve_tail(p, s_tot);
}


// Open3D v0.15.1 PointCloud::EstimateNormals(KDTreeSearchParamHybrid(radius, max_nn)) + NormalizeNormals +
// OrientNormalsTowardsCameraLocation(origin), the sequence of estimateNormalsOrCovariancesIfNeeded
// (CloudRegistration.cpp:27-46, 67-79). Open3D is not in the reference tree: restated from upstream —
//   EstimatePerPointCovariances: neighbours = hybrid search (within radius, at most max_nn, the point itself
//     included, nearest first); fewer than 3 -> covariance = Identity; else utility::ComputeCovariance: the nine
//     cumulants (x, y, z, xx, xy, xz, yy, yz, zz) summed in DOUBLE in neighbour order, divided by the count,
//     cov = E[p p^T] - E[p] E[p]^T;
//   ComputeNormal: eigenvector of the smallest eigenvalue (Identity -> (0, 0, 1)); zero vector -> (0, 0, 1);
//   NormalizeNormals; OrientNormalsTowardsCameraLocation(0): zero normal -> direction to the camera, else flipped when
//     normal . (camera - point) < 0.
// ids: k per point (original indices, nearest first, -1 = none) at row `j`; pts_by_id: the cloud in original order;
// q: the points the rows belong to. Output: float4 at row j (registration kernels) and / or double 3 x N at out_d row.
__global__ void __launch_bounds__(kThreads) o3d_estimate_normals_kernel(const int* __restrict__ ids, int knn,
                                                                        const float4* __restrict__ pts_by_id,
                                                                        const float4* __restrict__ q, int M,
                                                                        float4* __restrict__ out4,
                                                                        double* __restrict__ out_d, int out_by_id) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= M) return;
  double cum[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  int n = 0;
  for (int t = 0; t < knn; ++t) {
    const int id = ids[(size_t)j * knn + t];
    if (id < 0) continue;
    const float4 p = __ldg(&pts_by_id[id]);
    const double x = (double)p.x, y = (double)p.y, z = (double)p.z;
    cum[0] += x; cum[1] += y; cum[2] += z;
    cum[3] += x * x; cum[4] += x * y; cum[5] += x * z;
    cum[6] += y * y; cum[7] += y * z; cum[8] += z * z;
    ++n;
  }
  double v[3] = {0.0, 0.0, 1.0};
  if (n >= 3) {
    for (int t = 0; t < 9; ++t) cum[t] /= (double)n;
    double C[9], w[3], V[9];
    C[0] = cum[3] - cum[0] * cum[0];
    C[4] = cum[6] - cum[1] * cum[1];
    C[8] = cum[8] - cum[2] * cum[2];
    C[1] = C[3] = cum[4] - cum[0] * cum[1];
    C[2] = C[6] = cum[5] - cum[0] * cum[2];
    C[5] = C[7] = cum[7] - cum[1] * cum[2];
    la::symEigen3(C, w, V);
    int c = 0;
    if (w[1] < w[c]) c = 1;
    if (w[2] < w[c]) c = 2;
    v[0] = V[3 * c]; v[1] = V[3 * c + 1]; v[2] = V[3 * c + 2];
    if (v[0] * v[0] + v[1] * v[1] + v[2] * v[2] == 0.0) { v[0] = 0.0; v[1] = 0.0; v[2] = 1.0; }
  }
  const double nn = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (nn > 0.0) { v[0] /= nn; v[1] /= nn; v[2] /= nn; }
  const float4 p = __ldg(&q[j]);
  const double r[3] = {-(double)p.x, -(double)p.y, -(double)p.z};   // camera (origin) - point
  if (v[0] * r[0] + v[1] * r[1] + v[2] * r[2] < 0.0) { v[0] = -v[0]; v[1] = -v[1]; v[2] = -v[2]; }
  // out_by_id: q is the cell-sorted array (w = bits of the original index) and the outputs go to that index
  const size_t o = out_by_id ? (size_t)__float_as_int(p.w) : (size_t)j;
  if (out4) out4[o] = make_float4((float)v[0], (float)v[1], (float)v[2], 0.f);
  if (out_d) {
    out_d[3 * o] = v[0]; out_d[3 * o + 1] = v[1]; out_d[3 * o + 2] = v[2];
  }
}

}  // namespace

namespace b200 {
// for loop_closure.cu: Open3D normals (o3d_estimate_normals_kernel) from neighbour lists that are already on the device
int o3d_normals_from_ids_device(const int* ids, int knn, const float4* pts_by_id, const float4* q, int M, double* out_d,
                                int out_by_id, cudaStream_t s) {
  o3d_estimate_normals_kernel<<<(M + kThreads - 1) / kThreads, kThreads, 0, s>>>(ids, knn, pts_by_id, q, M, nullptr, out_d,
                                                                                 out_by_id);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}
}  // namespace b200

struct b200_o3d {
  int device = 0;
  b200_o3d_params prm;
  cudaStream_t stream = nullptr;
  MapIndex map;
  MapIndex nmap;               // scratch index of b200_o3d_estimate_normals
  double* d_nd = nullptr;      // its double normals
  size_t cap_nd = 0;
  bool target_ready = false;
  float4 *d_src = nullptr, *d_src_n = nullptr;
  float* d_stage3 = nullptr;
  int* d_ids = nullptr;
  float* d_dists = nullptr;
  size_t cap_src = 0;
  int N = 0;
  double* d_partials = nullptr;
  int max_blocks = 0;
  O3dState* d_state = nullptr;
  O3dState* h_state = nullptr;  // pinned
  cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
  // partition (one slab of a larger map)
  int qmap = 2;                // O3dParams::qmap (B200_O3D_QMAP overrides)
  int own_axis = -1;
  float own_lo = 0.f, own_hi = 0.f;
  int* d_gid = nullptr;
  // stepwise run
  O3dParams p;
  int blocks = 0;
  cudaStream_t run_stream = nullptr;
  bool running = false;
  int launches = 0;
  // peer exchange (o3d_exchange_solve_kernel)
  unsigned long long* d_mailbox = nullptr;
  PeerView pv;
  bool peers_ready = false;
  bool peers_ipc = false;     // peers live in other processes: b200_o3d_finish may enqueue the iterations in chunks
  int peer_pending = 0;       // iterations of the running registration not enqueued yet
  unsigned peer_seq = 0;
  bool peer_run = false;      // the running registration exchanges over the mailboxes
  void* opened[kPeerMax] = {};   // mailboxes mapped with cudaIpcOpenMemHandle
};

extern "C" {

void b200_o3d_default_params(b200_o3d_params* p) {
  if (!p) return;
  std::memset(p, 0, sizeof(*p));
  p->kind = 0;
  p->max_correspondence_distance = 1.0f;   // RegistrationIcp*::maxCorrespondenceDistance_ (CloudRegistration.hpp:38)
  p->max_iteration = 30;                   // open3d ICPConvergenceCriteria defaults: 1e-6, 1e-6, 30
  p->relative_fitness = 1e-6;
  p->relative_rmse = 1e-6;
  p->gicp_epsilon = 1e-3;                  // TransformationEstimationForGeneralizedICP::epsilon_
  p->cell_size = 0.f;
}

int b200_o3d_create(b200_o3d** out, int device, const b200_o3d_params* prm) {
  if (!out || !prm) { set_error("b200_o3d_create: bad arguments"); return B200_E_BAD_PARAM; }
  *out = nullptr;
  if (prm->kind < 0 || prm->kind > 2) { set_error("b200_o3d: kind must be 0 (point-to-plane), 1 (point-to-point) or 2 (generalized)"); return B200_E_BAD_PARAM; }
  if (!(prm->max_correspondence_distance > 0.f) || prm->max_iteration < 0) { set_error("b200_o3d: bad parameters"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  b200_o3d* h = new (std::nothrow) b200_o3d();
  if (!h) { set_error("out of host memory"); return B200_E_CUDA; }
  h->device = device;
  h->prm = *prm;
  if (const char* e = std::getenv("B200_O3D_QMAP")) h->qmap = std::atoi(e);
  h->map.device = device;
  h->nmap.device = device;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  h->max_blocks = sms * 8;
  auto fail = [&](cudaError_t e, const char* what) { const int st = cuda_fail(e, what, __FILE__, __LINE__); b200_o3d_destroy(h); return st; };
  cudaError_t e;
  if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(e, "cudaStreamCreate");
  h->map.stream = h->stream;
  h->nmap.stream = h->stream;
  if ((e = cudaMalloc(&h->d_partials, ((size_t)h->max_blocks + 1) * kVals * sizeof(double))) != cudaSuccess) return fail(e, "cudaMalloc");
  if ((e = cudaMalloc(&h->d_state, sizeof(O3dState))) != cudaSuccess) return fail(e, "cudaMalloc");
  if ((e = cudaMallocHost(&h->h_state, sizeof(O3dState))) != cudaSuccess) return fail(e, "cudaMallocHost");
  for (int i = 0; i < 3; ++i)
    if ((e = cudaEventCreate(&h->ev[i])) != cudaSuccess) return fail(e, "cudaEventCreate");
  *out = h;
  return B200_OK;
}

void b200_o3d_destroy(b200_o3d* h) {
  if (!h) return;
  {
    DeviceGuard dg(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->d_src); cudaFree(h->d_src_n); cudaFree(h->d_stage3); cudaFree(h->d_ids); cudaFree(h->d_dists);
    cudaFree(h->d_partials); cudaFree(h->d_state); cudaFree(h->d_gid);
    cudaFreeHost(h->h_state);
    for (int r = 0; r < kPeerMax; ++r)
      if (h->opened[r]) cudaIpcCloseMemHandle(h->opened[r]);
    cudaFree(h->d_mailbox);
    for (int i = 0; i < 3; ++i)
      if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    h->map.release();
    h->nmap.release();
    cudaFree(h->d_nd);
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
}

int b200_o3d_set_target(b200_o3d* h, const float* tgt4xM, const float* tgtn3xM, int M) {
  if (!h || M <= 0 || !tgt4xM) { set_error("b200_o3d_set_target: bad arguments"); return B200_E_BAD_PARAM; }
  if (!tgtn3xM && h->prm.kind != 1) { set_error("b200_o3d: the target cloud needs normals (point-to-plane residuals / GICP covariances)"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  h->target_ready = false;
  cudaStream_t s = h->stream;
  B200_TRY(index_alloc_points(h->map, M, tgtn3xM != nullptr));
  B200_TRY(upload_features(h->map.d_ref, tgt4xM, M, s));
  if (tgtn3xM) {
    float* stage = nullptr;
    B200_CUDA(cudaMallocAsync(&stage, (size_t)M * 3 * sizeof(float), s));
    const int st = upload_normals(h->map.d_nrm_orig, tgtn3xM, M, stage, s);
    cudaFreeAsync(stage, s);
    if (st) return st;
  }
  b200_index_opts opts;
  std::memset(&opts, 0, sizeof(opts));
  opts.cell_size = h->prm.cell_size;
  B200_TRY(index_build(h->map, &opts));   // no mean-centring: Open3D registers in the callers' frame
  B200_CUDA(cudaStreamSynchronize(s));
  cudaFree(h->d_gid);
  h->d_gid = nullptr;
  h->own_axis = -1;
  h->target_ready = true;
  return B200_OK;
}

int b200_o3d_set_target_estimate_normals(b200_o3d* h, const float* tgt4xM, int M, int knn, float maxDist) {
  if (!h || M <= 0 || !tgt4xM) { set_error("b200_o3d_set_target_estimate_normals: bad arguments"); return B200_E_BAD_PARAM; }
  if (knn < 1 || knn > 32) { set_error("b200_o3d_set_target_estimate_normals: knn must be in 1..32"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  h->target_ready = false;
  cudaStream_t s = h->stream;
  B200_TRY(index_alloc_points(h->map, M, true));
  B200_TRY(upload_features(h->map.d_ref, tgt4xM, M, s));
  B200_CUDA(cudaMemsetAsync(h->map.d_nrm_orig, 0, (size_t)M * sizeof(float4), s));
  b200_index_opts opts;
  std::memset(&opts, 0, sizeof(opts));
  opts.cell_size = h->prm.cell_size;
  B200_TRY(index_build(h->map, &opts));
  // the target's own index answers the neighbour queries; rows in cell-sorted order, so the normals land where the
  // registration kernels read them (no second index, no host round trip)
  int* ids = nullptr;
  B200_TRY(self_knn_ids_index(h->map, knn, maxDist, true, &ids, s));
  o3d_estimate_normals_kernel<<<(M + kThreads - 1) / kThreads, kThreads, 0, s>>>(ids, knn, h->map.d_ref, h->map.d_pts, M,
                                                                                 h->map.d_nrm, nullptr, 0);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaStreamSynchronize(s));
  cudaFree(h->d_gid);
  h->d_gid = nullptr;
  h->own_axis = -1;
  h->target_ready = true;
  return B200_OK;
}

int b200_o3d_estimate_normals(b200_o3d* h, const float* pts4xN, int N, int knn, float radius, double* normals3xN) {
  if (!h || N < 0 || (N > 0 && (!pts4xN || !normals3xN))) { set_error("b200_o3d_estimate_normals: bad arguments"); return B200_E_BAD_PARAM; }
  if (knn < 1 || knn > 32) { set_error("b200_o3d_estimate_normals: knn must be in 1..32"); return B200_E_BAD_PARAM; }
  if (!(radius > 0.f)) { set_error("b200_o3d_estimate_normals: radius must be > 0"); return B200_E_BAD_PARAM; }
  if (N == 0) return B200_OK;   // the reference returns without doing anything for an empty cloud
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  // a scratch index of its own: the registration target stays untouched
  B200_TRY(index_alloc_points(h->nmap, N, false));
  B200_TRY(upload_features(h->nmap.d_ref, pts4xN, N, s));
  b200_index_opts opts;
  std::memset(&opts, 0, sizeof(opts));
  B200_TRY(index_build(h->nmap, &opts));
  int* ids = nullptr;
  B200_TRY(self_knn_ids_index(h->nmap, knn, radius, true, &ids, s));
  if ((size_t)N > h->cap_nd) {
    cudaFree(h->d_nd);
    h->d_nd = nullptr; h->cap_nd = 0;
    B200_CUDA(cudaMalloc(&h->d_nd, (size_t)N * 3 * sizeof(double)));
    h->cap_nd = (size_t)N;
  }
  o3d_estimate_normals_kernel<<<(N + kThreads - 1) / kThreads, kThreads, 0, s>>>(ids, knn, h->nmap.d_ref, h->nmap.d_pts, N,
                                                                                 nullptr, h->d_nd, 1);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaMemcpyAsync(normals3xN, h->d_nd, (size_t)N * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  return B200_OK;
}

int b200_o3d_set_partition(b200_o3d* h, int axis, float lo, float hi, const int32_t* gid_M) {
  if (!h || axis < -1 || axis > 2) { set_error("b200_o3d_set_partition: bad arguments"); return B200_E_BAD_PARAM; }
  if (!h->target_ready) { set_error("b200_o3d_set_partition: set the slab as target first"); return B200_E_NOT_INITIALIZED; }
  if (axis >= 0 && !(lo < hi)) { set_error("b200_o3d_set_partition: empty ownership interval"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaFree(h->d_gid);
  h->d_gid = nullptr;
  h->own_axis = axis; h->own_lo = lo; h->own_hi = hi;
  if (axis >= 0 && gid_M) {
    const size_t bytes = (size_t)h->map.M * sizeof(int32_t);
    B200_CUDA(cudaMalloc(&h->d_gid, bytes));
    B200_CUDA(cudaMemcpy(h->d_gid, gid_M, bytes, cudaMemcpyHostToDevice));
  }
  return B200_OK;
}

}  // extern "C"

// est_knn > 0: the source normals are estimated on the device (Open3D semantics, o3d_estimate_normals_kernel) from a
// scratch index over the uploaded source instead of being handed over
static int o3d_begin_impl(b200_o3d* h, const float* src4xN, const float* srcn3xN, int N, const double* init16,
                          void* stream, int est_knn, float est_radius) {
  if (!h || N <= 0 || !src4xN) { set_error("b200_o3d_begin: bad arguments"); return B200_E_BAD_PARAM; }
  if (!h->target_ready) { set_error("b200_o3d: no target set"); return B200_E_NOT_INITIALIZED; }
  if (h->prm.kind == 2 && !srcn3xN && est_knn <= 0) { set_error("b200_o3d: generalized ICP needs source normals (covariances are derived from them)"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
  h->running = false;
  h->peer_pending = 0;
  if ((size_t)N > h->cap_src) {
    cudaFree(h->d_src); cudaFree(h->d_src_n); cudaFree(h->d_stage3); cudaFree(h->d_ids); cudaFree(h->d_dists);
    h->d_src = h->d_src_n = nullptr; h->d_stage3 = nullptr; h->d_ids = nullptr; h->d_dists = nullptr;
    h->cap_src = 0;
    B200_CUDA(cudaMalloc(&h->d_src, (size_t)N * sizeof(float4)));
    B200_CUDA(cudaMalloc(&h->d_src_n, (size_t)N * sizeof(float4)));
    B200_CUDA(cudaMalloc(&h->d_stage3, (size_t)N * 3 * sizeof(float)));
    B200_CUDA(cudaMalloc(&h->d_ids, (size_t)N * sizeof(int)));
    B200_CUDA(cudaMalloc(&h->d_dists, (size_t)N * sizeof(float)));
    h->cap_src = (size_t)N;
  }
  h->N = N;
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  B200_TRY(upload_features(h->d_src, src4xN, N, s));
  if (srcn3xN) B200_TRY(upload_normals(h->d_src_n, srcn3xN, N, h->d_stage3, s));
  int est_launches = 0;
  if (!srcn3xN && est_knn > 0) {
    h->nmap.stream = s;
    B200_TRY(index_alloc_points(h->nmap, N, false));
    B200_CUDA(cudaMemcpyAsync(h->nmap.d_ref, h->d_src, (size_t)N * sizeof(float4), cudaMemcpyDeviceToDevice, s));
    b200_index_opts opts;
    std::memset(&opts, 0, sizeof(opts));
    B200_TRY(index_build(h->nmap, &opts));
    int* nids = nullptr;
    B200_TRY(self_knn_ids_index(h->nmap, est_knn, est_radius, true, &nids, s));
    o3d_estimate_normals_kernel<<<(N + kThreads - 1) / kThreads, kThreads, 0, s>>>(nids, est_knn, h->nmap.d_ref, h->nmap.d_pts,
                                                                                   N, h->d_src_n, nullptr, 1);
    B200_CUDA(cudaGetLastError());
    h->nmap.stream = h->stream;
    est_launches = 8;   // index build (bbox, 2 histograms, scan passes, scatter) + k-NN ids + normals
  }
  const bool have_n = srcn3xN || est_knn > 0;
  std::memset(h->h_state, 0, sizeof(O3dState));
  for (int i = 0; i < 16; ++i) h->h_state->T[i] = init16 ? init16[i] : ((i % 5 == 0) ? 1.0 : 0.0);
  B200_CUDA(cudaMemcpyAsync(h->d_state, h->h_state, sizeof(O3dState), cudaMemcpyHostToDevice, s));
  O3dParams& p = h->p;
  std::memset(&p, 0, sizeof(p));
  p.g = h->map.view();
  p.src = h->d_src;
  p.src_n = have_n ? h->d_src_n : nullptr;
  p.N = N; p.kind = h->prm.kind;
  p.md2 = h->prm.max_correspondence_distance * h->prm.max_correspondence_distance;
  p.max_iteration = h->prm.max_iteration;
  p.rel_fitness = h->prm.relative_fitness; p.rel_rmse = h->prm.relative_rmse; p.eps = h->prm.gicp_epsilon;
  p.ids = h->d_ids; p.dists = h->d_dists; p.partials = h->d_partials; p.st = h->d_state;
  p.own_axis = h->own_axis; p.own_lo = h->own_lo; p.own_hi = h->own_hi; p.gid = h->d_gid;
  p.qmap = h->qmap;
  // four resident CTAs per SM (64 registers, 25 KB of shared memory each): one wave, every warp walks its slots
  h->blocks = std::max(1, std::min(h->max_blocks / 2, (N + kThreads - 1) / kThreads));
  h->run_stream = s;
  h->launches = srcn3xN ? 1 : est_launches;
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  h->running = true;
  return B200_OK;
}

extern "C" {

int b200_o3d_begin(b200_o3d* h, const float* src4xN, const float* srcn3xN, int N, const double* init16, void* stream) {
  return o3d_begin_impl(h, src4xN, srcn3xN, N, init16, stream, 0, 0.f);
}

int b200_o3d_step_linearize(b200_o3d* h, double* d_sums32) {
  if (!h || !h->running || !d_sums32) { set_error("b200_o3d_step_linearize: no registration begun"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  launch_match_lin(h->p, h->blocks, h->run_stream);
  o3d_reduce_kernel<<<1, kRedThreads, 0, h->run_stream>>>(h->p, h->blocks, d_sums32);
  h->launches += 2;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int b200_o3d_step_update(b200_o3d* h, const double* d_sums32) {
  if (!h || !h->running || !d_sums32) { set_error("b200_o3d_step_update: no registration begun"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  o3d_solve_kernel<<<1, kRedThreads, 0, h->run_stream>>>(h->p, h->blocks, d_sums32);
  h->launches += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int b200_o3d_peer_init(b200_o3d* h, int rank, int world, void** mailbox_dev, void* ipc_handle_64) {
  if (!h || world < 1 || world > kPeerMax || rank < 0 || rank >= world) {
    set_error("b200_o3d_peer_init: need 0 <= rank < world <= 16");
    return B200_E_BAD_PARAM;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the C ABI hands IPC handles around as 64 bytes");
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  h->peers_ready = false;
  for (int r = 0; r < kPeerMax; ++r)
    if (h->opened[r]) { cudaIpcCloseMemHandle(h->opened[r]); h->opened[r] = nullptr; }
  cudaFree(h->d_mailbox);
  h->d_mailbox = nullptr;
  const size_t bytes = (size_t)2 * world * kPackets * sizeof(unsigned long long);
  B200_CUDA(cudaMalloc(&h->d_mailbox, bytes));
  B200_CUDA(cudaMemset(h->d_mailbox, 0, bytes));   // sequence number 0 is never sent
  B200_CUDA(cudaDeviceSynchronize());
  std::memset(&h->pv, 0, sizeof(h->pv));
  h->pv.rank = rank; h->pv.world = world;
  h->peer_seq = 0;
  if (mailbox_dev) *mailbox_dev = h->d_mailbox;
  if (ipc_handle_64) {
    cudaIpcMemHandle_t ih;
    B200_CUDA(cudaIpcGetMemHandle(&ih, h->d_mailbox));
    std::memcpy(ipc_handle_64, &ih, 64);
  }
  return B200_OK;
}

int b200_o3d_peer_connect(b200_o3d* h, void* const* mailboxes_dev, const int* devices, const void* ipc_handles_64) {
  if (!h || !h->d_mailbox || (!mailboxes_dev && !ipc_handles_64)) {
    set_error("b200_o3d_peer_connect: call b200_o3d_peer_init first and pass pointers or IPC handles");
    return B200_E_BAD_PARAM;
  }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  for (int r = 0; r < h->pv.world; ++r) {
    if (r == h->pv.rank) { h->pv.box[r] = h->d_mailbox; continue; }
    if (mailboxes_dev) {  // same process: the pointer is valid here once peer access is on
      if (devices && devices[r] != h->device) {
        const cudaError_t e = cudaDeviceEnablePeerAccess(devices[r], 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
        cudaGetLastError();
      }
      h->pv.box[r] = (unsigned long long*)mailboxes_dev[r];
      h->peers_ipc = false;
    } else {
      h->peers_ipc = true;
      cudaIpcMemHandle_t ih;
      std::memcpy(&ih, (const char*)ipc_handles_64 + (size_t)64 * r, 64);
      void* ptr = nullptr;
      B200_CUDA(cudaIpcOpenMemHandle(&ptr, ih, cudaIpcMemLazyEnablePeerAccess));
      h->opened[r] = ptr;
      h->pv.box[r] = (unsigned long long*)ptr;
    }
  }
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, h->device);
  long long timeout_ms = 5000;   // 5 s of SM clock; B200_O3D_PEER_TIMEOUT_MS overrides (tests of the failure path)
  if (const char* e = std::getenv("B200_O3D_PEER_TIMEOUT_MS")) timeout_ms = std::max(1LL, std::atoll(e));
  h->pv.timeout_cycles = (long long)(khz > 0 ? khz : 1500000) * timeout_ms;
  h->peers_ready = true;
  return B200_OK;
}

int b200_o3d_enqueue_peers(b200_o3d* h, const float* src4xN, const float* srcn3xN, int N, const double* init16,
                           void* stream) {
  if (!h || !h->peers_ready) { set_error("b200_o3d_enqueue_peers: peers not connected"); return B200_E_NOT_INITIALIZED; }
  if (h->own_axis < 0 && h->pv.world > 1) { set_error("b200_o3d_enqueue_peers: no partition set"); return B200_E_NOT_INITIALIZED; }
  B200_TRY(b200_o3d_begin(h, src4xN, srcn3xN, N, init16, stream));
  DeviceGuard dg(h->device);
  h->pv.seq_base = h->peer_seq;
  // max_iteration updates need max_iteration + 1 matches; launches after `done` return at once (on every rank at the
  // same step: the totals are bit-identical), so nothing is sent or awaited any more. One process per rank: the first
  // eight iterations now, b200_o3d_finish adds eight more whenever the state it reads back is not done (every rank
  // reads the same `done`, so all of them do) — a registration of 4-6 updates never pays for 2 x 25 idle launches.
  // Several handles in one process: everything now (finish of one handle must not wait for work another handle's
  // finish would have to enqueue).
  const int total = h->prm.max_iteration + 1;
  const int now = h->peers_ipc ? std::min(kPeerChunk, total) : total;
  h->peer_pending = total - now;
  for (int it = 0; it < now; ++it) {
    launch_match_lin(h->p, h->blocks, h->run_stream);
    o3d_exchange_solve_kernel<<<1, kRedThreads, 0, h->run_stream>>>(h->p, h->blocks, h->pv);
  }
  h->peer_run = true;   // b200_o3d_finish advances peer_seq by the steps actually taken
  h->launches += 2 * now;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int b200_o3d_poll(b200_o3d* h, int* done, int* updates) {
  if (!h || !h->running) { set_error("b200_o3d_poll: no registration begun"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  B200_CUDA(cudaMemcpyAsync(h->h_state, h->d_state, sizeof(O3dState), cudaMemcpyDeviceToHost, h->run_stream));
  B200_CUDA(cudaStreamSynchronize(h->run_stream));
  if (done) *done = h->h_state->done;
  if (updates) *updates = h->h_state->updates;
  return B200_OK;
}

static int o3d_fill_result(b200_o3d* h, b200_o3d_result* res) {
  std::memset(res, 0, sizeof(*res));
  for (int i = 0; i < 16; ++i) res->T[i] = h->h_state->T[i];
  res->fitness = h->h_state->fitness;
  res->inlier_rmse = h->h_state->rmse;
  res->iterations = h->h_state->updates;
  res->n_correspondences = h->h_state->n_corr;
  res->kernel_launches = h->launches;
  B200_CUDA(cudaEventElapsedTime(&res->device_ms, h->ev[1], h->ev[2]));
  B200_CUDA(cudaEventElapsedTime(&res->total_ms, h->ev[0], h->ev[2]));
  if (h->h_state->status != 0) {
    set_error("b200_o3d: a peer's sums did not arrive within the time limit (5 s by default; rank missing or out of step): re-run b200_o3d_peer_init / peer_connect on all ranks");
    return h->h_state->status;
  }
  return B200_OK;
}

int b200_o3d_finish(b200_o3d* h, b200_o3d_result* res) {
  if (!h || !h->running || !res) { set_error("b200_o3d_finish: no registration begun"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  cudaStream_t s = h->run_stream;
  while (true) {
    B200_CUDA(cudaEventRecord(h->ev[2], s));
    B200_CUDA(cudaMemcpyAsync(h->h_state, h->d_state, sizeof(O3dState), cudaMemcpyDeviceToHost, s));
    B200_CUDA(cudaStreamSynchronize(s));
    if (h->h_state->done || h->peer_pending <= 0) break;
    const int n = std::min(kPeerChunk, h->peer_pending);   // peer exchange across processes, not converged yet
    for (int it = 0; it < n; ++it) {
      launch_match_lin(h->p, h->blocks, s);
      o3d_exchange_solve_kernel<<<1, kRedThreads, 0, s>>>(h->p, h->blocks, h->pv);
    }
    h->peer_pending -= n;
    h->launches += 2 * n;
    B200_CUDA(cudaGetLastError());
  }
  h->peer_pending = 0;
  if (h->peer_run) {
    // Sequence numbers stay consecutive across registrations (every rank takes the same number of steps: the totals
    // are bit-identical), so the two parity slots of the mailbox are enough at the boundary as well: a rank that has
    // started the next registration is one step ahead of a peer still reading the last step, never two.
    h->peer_seq = h->pv.seq_base + (unsigned)h->h_state->matches_done;
    h->peer_run = false;
  }
  return o3d_fill_result(h, res);
}

static int o3d_register_impl(b200_o3d* h, const float* src4xN, const float* srcn3xN, int N, const double* init16,
                             b200_o3d_result* res, int est_knn, float est_radius);

int b200_o3d_information_matrix(b200_o3d* h, const float* src4xN, int N, const double* T16,
                                float max_correspondence_distance, double* info36, double* fitness, double* inlier_rmse,
                                int* n_correspondences) {
  if (!h || N <= 0 || !src4xN || !info36) { set_error("b200_o3d_information_matrix: bad arguments"); return B200_E_BAD_PARAM; }
  if (!(max_correspondence_distance > 0.f)) { set_error("b200_o3d_information_matrix: max_correspondence_distance must be > 0"); return B200_E_BAD_PARAM; }
  if (h->own_axis >= 0) { set_error("b200_o3d_information_matrix: not for a slab of a partitioned map"); return B200_E_BAD_PARAM; }
  const int kind0 = h->prm.kind;
  h->prm.kind = 1;   // the upload path of a registration that needs no normals
  const int st = o3d_begin_impl(h, src4xN, nullptr, N, T16, nullptr, 0, 0.f);
  h->prm.kind = kind0;
  B200_TRY(st);
  DeviceGuard dg(h->device);
  h->running = false;
  cudaStream_t s = h->run_stream;
  O3dParams p = h->p;
  p.kind = 3;
  p.md2 = max_correspondence_distance * max_correspondence_distance;
  launch_match_lin(p, h->blocks, s);
  o3d_reduce_kernel<<<1, kRedThreads, 0, s>>>(p, h->blocks, h->d_partials + (size_t)h->max_blocks * kVals);
  B200_CUDA(cudaGetLastError());
  double sums[kVals];
  B200_CUDA(cudaMemcpyAsync(sums, h->d_partials + (size_t)h->max_blocks * kVals, sizeof(sums), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  int t = 0;
  for (int c = 0; c < 6; ++c)
    for (int a = 0; a <= c; ++a) { info36[a + 6 * c] = sums[t]; info36[c + 6 * a] = sums[t]; ++t; }
  const double cnt = sums[27];
  if (fitness) *fitness = cnt / (double)N;
  if (inlier_rmse) *inlier_rmse = cnt > 0.0 ? std::sqrt(sums[28] / cnt) : 0.0;
  if (n_correspondences) *n_correspondences = (int)(cnt + 0.5);
  return B200_OK;
}

int b200_o3d_register(b200_o3d* h, const float* src4xN, const float* srcn3xN, int N, const double* init16,
                      b200_o3d_result* res) {
  return o3d_register_impl(h, src4xN, srcn3xN, N, init16, res, 0, 0.f);
}

int b200_o3d_register_estimate_normals(b200_o3d* h, const float* src4xN, int N, int knn, float radius,
                                       const double* init16, b200_o3d_result* res) {
  if (knn < 1 || knn > 32 || !(radius > 0.f)) { set_error("b200_o3d_register_estimate_normals: 1 <= knn <= 32, radius > 0"); return B200_E_BAD_PARAM; }
  return o3d_register_impl(h, src4xN, nullptr, N, init16, res, knn, radius);
}

static int o3d_register_impl(b200_o3d* h, const float* src4xN, const float* srcn3xN, int N, const double* init16,
                             b200_o3d_result* res, int est_knn, float est_radius) {
  if (!res) { set_error("b200_o3d_register: bad arguments"); return B200_E_BAD_PARAM; }
  if (h && h->own_axis >= 0) { set_error("b200_o3d_register: the target is one slab of a partitioned map; drive it with begin/step/finish"); return B200_E_BAD_PARAM; }
  B200_TRY(o3d_begin_impl(h, src4xN, srcn3xN, N, init16, nullptr, est_knn, est_radius));
  DeviceGuard dg(h->device);
  // max_iteration updates need max_iteration + 1 matches. The iterations are enqueued kChunk at a time without any
  // host synchronisation inside a chunk; launches after `done` return at once, and the read-back of the state that
  // ends a chunk is the one the result needs anyway — a registration that converges within the first chunk (the usual
  // 4-6 updates) pays neither for 2 x 25 idle launches (~3 us each) nor for an extra round trip.
  constexpr int kChunk = 8;
  cudaStream_t s = h->run_stream;
  for (int it = 0; it <= h->prm.max_iteration;) {
    const int n = std::min(kChunk, h->prm.max_iteration + 1 - it);
    for (int j = 0; j < n; ++j) {
      launch_match_lin(h->p, h->blocks, s);
      o3d_solve_kernel<<<1, kRedThreads, 0, s>>>(h->p, h->blocks, nullptr);
    }
    it += n;
    h->launches += 2 * n;
    B200_CUDA(cudaGetLastError());
    B200_CUDA(cudaEventRecord(h->ev[2], s));
    B200_CUDA(cudaMemcpyAsync(h->h_state, h->d_state, sizeof(O3dState), cudaMemcpyDeviceToHost, s));
    B200_CUDA(cudaStreamSynchronize(s));
    if (h->h_state->done) break;
  }
  return o3d_fill_result(h, res);
}

int b200_o3d_get_correspondences(b200_o3d* h, int32_t* ids_N, float* dists_N) {
  if (!h || h->N <= 0) { set_error("no registration has run"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  if (ids_N) B200_CUDA(cudaMemcpyAsync(ids_N, h->d_ids, (size_t)h->N * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
  if (dists_N) B200_CUDA(cudaMemcpyAsync(dists_N, h->d_dists, (size_t)h->N * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
  B200_CUDA(cudaStreamSynchronize(h->stream));
  return B200_OK;
}

}  // extern "C"
