// SurfaceNormalDataPointsFilter over the indexed cloud itself (SurfaceNormal.cpp:82-290), cell-block tiles.
//
// The queries ARE the indexed points, so queries that are neighbours in space share their candidate sets: a CTA takes
// a core block of kCore^3 cells, stages every point of the block and of a halo of kHalo cells around it in shared
// memory ONCE (x-runs of cells are contiguous in the cell-sorted array: 64 coalesced row copies), and every thread
// then answers one query of the core by scanning the staged candidates (broadcast shared-memory reads, no divergence)
// into a register-resident sorted k-best list ordered by (dist2, original index) — the reference's order, ties ->
// smallest index. A query is settled when the radius it still needs, min(sqrt(k-th best), maxDist), provably lies
// inside the staged region (its distance in cell units to the nearest face of the region that is not also the face of
// the grid); otherwise, or when a region holds more points than fit in shared memory (crowded cells of a raw scan
// near the sensor), its index goes to a list that the exact warp-cooperative kernel of modules.cu answers afterwards.
// Mean, covariance, rank test, eigen-decomposition and the descriptors are the SAME float operations in the same
// (list) order as the other two kernels, so the outputs are bit-identical whichever path answered a point.
//
// smoothNormals (SurfaceNormal.cpp:265-288) smooths IN PLACE in index order: point i averages the normals of its
// neighbours as they stand when i is reached — already smoothed for lower indices, raw for higher ones. That is a
// dependency chain, reproduced exactly by smooth_normals_kernel: warps take points in index order (a ticket counter)
// and wait until every lower-index neighbour has published its smoothed normal.
#include "normals.cuh"
#include "knn.cuh"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

namespace b200 {
namespace {

constexpr int kCore = 4;                      // core block: kCore^3 cells
constexpr int kHalo = 2;                      // staged around it
constexpr int kSpan = kCore + 2 * kHalo;      // 8 cells per axis
constexpr int kRows = kSpan * kSpan;          // 64 x-runs
constexpr int kCoreRows = kCore * kCore;      // 16 x-runs hold the queries
constexpr int kTileThreads = 128;
constexpr int kTileCap = 2304;                // staged points (36 KB of float4)
constexpr int kBins = 32;                     // distance bins of the threshold pass (uint16 counters, 8 KB)

struct SnOut {
  float *normals, *dens, *eigva, *eigve, *matched, *meandist;
  int* ids;   // k x N neighbour ids (-1: none) for the smoothing pass, or null
};

template <int K>
__global__ void __launch_bounds__(kTileThreads) sn_tile_kernel(const __grid_constant__ GridView g, int nbx, int nby,
                                                               int knn, float md2, int sortEigen, int out_by_id,
                                                               SnOut o, int* __restrict__ list,
                                                               int* __restrict__ list_count) {
  __shared__ float4 s_pts[kTileCap];
  __shared__ uint32_t s_row_a[kRows], s_row_off[kRows + 1];
  __shared__ uint32_t s_core_a[kCoreRows], s_core_off[kCoreRows + 1];
  __shared__ int s_base;
  __shared__ unsigned short s_cnt[kBins][kTileThreads];   // per-thread histogram of dist2 (column = thread: no conflicts)
  const int tid = threadIdx.x;
  const int bx = blockIdx.x % nbx, by = (blockIdx.x / nbx) % nby, bz = blockIdx.x / (nbx * nby);
  const int x0 = bx * kCore, y0 = by * kCore, z0 = bz * kCore;
  // ---- the 16 core rows (queries)
  if (tid < kCoreRows) {
    const int iy = y0 + (tid & (kCore - 1)), iz = z0 + (tid / kCore);
    uint32_t a = 0, b = 0;
    if (iy < g.ny && iz < g.nz) {
      const uint32_t* row = g.G + ((size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz));
      a = __ldg(row + x0);
      b = __ldg(row + min(x0 + kCore, g.nx));
    }
    s_core_a[tid] = a;
    s_core_off[tid + 1] = b - a;
  }
  __syncthreads();
  if (tid == 0) {
    uint32_t acc = 0;
    s_core_off[0] = 0;
    for (int r = 0; r < kCoreRows; ++r) { acc += s_core_off[r + 1]; s_core_off[r + 1] = acc; }
  }
  __syncthreads();
  const int Q = (int)s_core_off[kCoreRows];
  if (Q == 0) return;
  // ---- the 64 rows of the staged region
  const int rx0 = max(x0 - kHalo, 0), rx1 = min(x0 + kCore + kHalo, g.nx);   // cells [rx0, rx1)
  if (tid < kRows) {
    const int iy = y0 - kHalo + (tid & (kSpan - 1)), iz = z0 - kHalo + (tid / kSpan);
    uint32_t a = 0, b = 0;
    if (iy >= 0 && iy < g.ny && iz >= 0 && iz < g.nz) {
      const uint32_t* row = g.G + ((size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz));
      a = __ldg(row + rx0);
      b = __ldg(row + rx1);
    }
    s_row_a[tid] = a;
    s_row_off[tid + 1] = b - a;
  }
  __syncthreads();
  if (tid == 0) {
    uint32_t acc = 0;
    s_row_off[0] = 0;
    for (int r = 0; r < kRows; ++r) { acc += s_row_off[r + 1]; s_row_off[r + 1] = acc; }
  }
  __syncthreads();
  const int C = (int)s_row_off[kRows];
  auto core_sp = [&](int q) {   // sorted position of the q-th query of this block
    int r = 0;
#pragma unroll
    for (int t = 1; t < kCoreRows; ++t) r += (s_core_off[t] <= (uint32_t)q) ? 1 : 0;
    return (int)(s_core_a[r] + ((uint32_t)q - s_core_off[r]));
  };
  if (C > kTileCap) {   // crowded region: every query of the block goes to the exact warp path
    if (tid == 0) s_base = atomicAdd(list_count, Q);
    __syncthreads();
    for (int q = tid; q < Q; q += kTileThreads) {
      const int sp = core_sp(q);
      list[s_base + q] = out_by_id ? __float_as_int(__ldg(&g.pts[sp]).w) : sp;
    }
    return;
  }
  for (int c = tid; c < C; c += kTileThreads) {
    int lo = 0, hi = kRows - 1;   // last row whose offset is <= c
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (s_row_off[mid] <= (uint32_t)c) lo = mid; else hi = mid - 1;
    }
    s_pts[c] = __ldg(&g.pts[s_row_a[lo] + ((uint32_t)c - s_row_off[lo])]);
  }
  __syncthreads();
  // faces of the region that are not faces of the grid (beyond the grid there are no points)
  const float INF = INFINITY;
  const float fx0 = rx0 > 0 ? (float)rx0 : -INF, fx1 = rx1 < g.nx ? (float)rx1 : INF;
  const float fy0 = y0 - kHalo > 0 ? (float)(y0 - kHalo) : -INF, fy1 = y0 + kCore + kHalo < g.ny ? (float)(y0 + kCore + kHalo) : INF;
  const float fz0 = z0 - kHalo > 0 ? (float)(z0 - kHalo) : -INF, fz1 = z0 + kCore + kHalo < g.nz ? (float)(z0 + kCore + kHalo) : INF;
  for (int q = tid; q < Q; q += kTileThreads) {
    const int sp = core_sp(q);
    const float4 qp = __ldg(&g.pts[sp]);
    const int qid = __float_as_int(qp.w);
    const int i = out_by_id ? qid : sp;
    QueryGeom qg;
    const bool ok_geom = make_geom(g, qp.x, qp.y, qp.z, md2, qg);
    const float margin = fminf(fminf(fminf(qg.ux - fx0, fx1 - qg.ux), fminf(qg.uy - fy0, fy1 - qg.uy)),
                               fminf(qg.uz - fz0, fz1 - qg.uz));
    // Pass 1: histogram of dist2 over [0, R2), R = the radius that is provably inside the region (the whole cloud's
    // diagonal when the region is the whole grid). The staged points arrive in a spatial sweep, which is the worst
    // order for an insertion list (half of them would be accepted); the histogram tells the bin that holds the k-th
    // nearest, and pass 2 only offers the candidates up to that bin (a few more than k).
    float R2 = md2;
    if (margin != INFINITY) { const float r = fmaxf(margin - qg.slack, 0.f) * g.cell; R2 = fminf(R2, r * r); }
    if (R2 == INFINITY) {
      const float ex = (float)g.nx * g.cell, ey = (float)g.ny * g.cell, ez = (float)g.nz * g.cell;
      R2 = ex * ex + ey * ey + ez * ez;
    }
    const float bscale = R2 > 0.f ? (float)kBins / R2 : 0.f;
#pragma unroll
    for (int b = 0; b < kBins; ++b) s_cnt[b][tid] = 0;
    for (int c = 0; c < C; ++c) {
      const float4 p = s_pts[c];
      const float d = dist2_rn(qp.x, qp.y, qp.z, p.x, p.y, p.z);
      const float fb = d * bscale;
      if (d <= md2 && fb < (float)kBins) s_cnt[(int)fb][tid] += 1;
    }
    int tb = kBins, cum = 0;   // first bin at which the cumulative count reaches knn (kBins: never)
#pragma unroll
    for (int b = 0; b < kBins; ++b) {
      cum += s_cnt[b][tid];
      if (cum >= knn && tb == kBins) tb = b;
    }
    // fewer than knn points within the radius that can be vouched for: only settled when a finite maxDist caps the
    // search inside the region (then every point within maxDist is in the bins)
    const bool capped = md2 != INFINITY && sqrtf(md2) * g.inv_cell * 1.00001f + qg.slack < margin;
    if (!ok_geom || (tb == kBins && !capped && margin != INFINITY)) {
      list[atomicAdd(list_count, 1)] = i;
      continue;
    }
    // no threshold bin (fewer than knn points in the bins: whole grid staged, or maxDist caps the search): offer all
    const bool use_bins = tb < kBins;
    const float flim = (float)(tb + 1);
    KList<K> L;
    L.init();
    for (int c = 0; c < C; ++c) {
      const float4 p = s_pts[c];
      const float d = dist2_rn(qp.x, qp.y, qp.z, p.x, p.y, p.z);
      if (d <= md2 && (!use_bins || d * bscale < flim)) L.offer(d, __float_as_int(p.w), c);   // sp field = position in the tile
    }
    float kth = L.d[K - 1];
#pragma unroll
    for (int j = 0; j < K - 1; ++j)
      if (j == knn - 1) kth = L.d[j];
    // is everything within the radius still needed inside the staged region?
    const float need = need_radius(g, qg, kth);
    if (!(need < margin)) {
      list[atomicAdd(list_count, 1)] = i;
      continue;
    }
    // mean over the valid neighbours in list order (d.rowwise().sum() / realKnn, :181)
    int realKnn = 0;
    float sx = 0.f, sy = 0.f, sz = 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j)
      if (j < knn && L.d[j] != INFINITY) {
        const float4 p = s_pts[L.sp[j]];
        sx = __fadd_rn(sx, p.x); sy = __fadd_rn(sy, p.y); sz = __fadd_rn(sz, p.z);
        ++realKnn;
      }
    const float mx = sx / (float)realKnn, my = sy / (float)realKnn, mz = sz / (float)realKnn;
    float Cm[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) Cm[t] = 0.f;
    float maxSq = 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j)
      if (j < knn && L.d[j] != INFINITY) {
        const float4 p = s_pts[L.sp[j]];
        const float nn[3] = {__fsub_rn(p.x, mx), __fsub_rn(p.y, my), __fsub_rn(p.z, mz)};
        float sq = __fmul_rn(nn[0], nn[0]);
        sq = __fadd_rn(sq, __fmul_rn(nn[1], nn[1]));
        sq = __fadd_rn(sq, __fmul_rn(nn[2], nn[2]));
        maxSq = fmaxf(maxSq, sq);
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int r = 0; r < 3; ++r) Cm[r + 3 * c] = __fadd_rn(Cm[r + 3 * c], __fmul_rn(nn[r], nn[c]));
      }
    sn_finish(i, make_float4(qp.x, qp.y, qp.z, 1.f), Cm, realKnn, maxSq, mx, my, mz, sortEigen, o.normals, o.dens, o.eigva,
              o.eigve, o.meandist);
#pragma unroll
    for (int j = 0; j < K; ++j)
      if (j < knn) {
        if (o.matched) o.matched[(size_t)i * knn + j] = (float)L.id[j];
        if (o.ids) o.ids[(size_t)i * knn + j] = L.id[j];
      }
  }
}

// ---- smoothNormals, SurfaceNormal.cpp:265-288 ------------------------------------------------------------------------
// One warp per point, points handed out in index order by a ticket counter (a warp that holds ticket i only ever waits
// for points < i, whose warps hold earlier tickets and are therefore already running: no deadlock). flags[i] = 1 once
// normals[i] is the smoothed one. The neighbour sum runs sequentially in list order (mean += / -= normal), like the
// reference; the dot product uses the raw normal of i (currentNormal) and the neighbour's normal as it stands.
__global__ void __launch_bounds__(128) smooth_normals_kernel(float* __restrict__ normals, const float* __restrict__ raw,
                                                             const int* __restrict__ ids, int N, int knn,
                                                             int* __restrict__ ticket, volatile int* __restrict__ flags) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  while (true) {
    int i = 0;
    if (lane == 0) i = atomicAdd(ticket, 1);
    i = __shfl_sync(FULL, i, 0);
    if (i >= N) return;
    const float c0 = raw[3 * (size_t)i], c1 = raw[3 * (size_t)i + 1], c2 = raw[3 * (size_t)i + 2];
    float m0 = 0.f, m1 = 0.f, m2 = 0.f;
    int n = 0;
    for (int jb = 0; jb < knn; jb += 32) {
      const int j = jb + lane;
      const int id = j < knn ? ids[(size_t)i * knn + j] : -1;
      // wait for the lower-index neighbours (already smoothed when the reference reaches i); i itself and higher
      // indices are read raw
      if (id >= 0 && id < i)
        while (flags[id] == 0) __nanosleep(20);
      __threadfence();
      float v0 = 0.f, v1 = 0.f, v2 = 0.f;
      if (id >= 0) {
        // lower indices: smoothed (published); i itself and higher indices: RAW, whatever their warps did meanwhile
        const float* src = id < i ? normals : raw;
        v0 = __ldcg(&src[3 * (size_t)id]); v1 = __ldcg(&src[3 * (size_t)id + 1]); v2 = __ldcg(&src[3 * (size_t)id + 2]);
      }
      const int cnt = min(32, knn - jb);
      for (int t = 0; t < cnt; ++t) {
        const int tid_ = __shfl_sync(FULL, id, t);
        const float a0 = __shfl_sync(FULL, v0, t), a1 = __shfl_sync(FULL, v1, t), a2 = __shfl_sync(FULL, v2, t);
        if (tid_ < 0) continue;   // matches.dists(j, i) == InvalidDist
        float dot = __fmul_rn(c0, a0);
        dot = __fadd_rn(dot, __fmul_rn(c1, a1));
        dot = __fadd_rn(dot, __fmul_rn(c2, a2));
        if (dot > 0.f) { m0 = __fadd_rn(m0, a0); m1 = __fadd_rn(m1, a1); m2 = __fadd_rn(m2, a2); }
        else { m0 = __fsub_rn(m0, a0); m1 = __fsub_rn(m1, a1); m2 = __fsub_rn(m2, a2); }
        ++n;
      }
    }
    if (lane == 0) {
      const float fn = (float)n;
      __stcg(&normals[3 * (size_t)i], m0 / fn);
      __stcg(&normals[3 * (size_t)i + 1], m1 / fn);
      __stcg(&normals[3 * (size_t)i + 2], m2 / fn);
      __threadfence();
      flags[i] = 1;
    }
    __syncwarp();
  }
}

template <int K>
int launch_tile(const GridView& g, int knn, float md2, int sortEigen, int out_by_id, const SnOut& o, int* list,
                int* count, cudaStream_t s) {
  const int nbx = (g.nx + kCore - 1) / kCore, nby = (g.ny + kCore - 1) / kCore, nbz = (g.nz + kCore - 1) / kCore;
  const long long nb = (long long)nbx * nby * nbz;
  if (nb > 0x7fffffffLL) { set_error("surface normals: too many cell blocks for one launch"); return B200_E_UNSUPPORTED; }
  sn_tile_kernel<K><<<(unsigned)nb, kTileThreads, 0, s>>>(g, nbx, nby, knn, md2, sortEigen, out_by_id, o, list, count);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace

int surface_normals_device(const GridView& g, const float4* feat, int N, int knn, float maxDist, int sortEigen,
                           float* normals, float* dens, float* eigva, float* eigve, float* matched, float* meandist,
                           cudaStream_t s);   // modules.cu

int surface_normals_index(MapIndex& idx, const float4* feat, bool feat_sorted, int N, int knn, float maxDist,
                          int sortEigen, int smooth, float* normals, float* dens, float* eigva, float* eigve,
                          float* matched, float* meandist, cudaStream_t s) {
  if (N <= 0) return B200_OK;
  if (N != idx.M) { set_error("surface_normals_index: the queries must be the indexed cloud"); return B200_E_BAD_PARAM; }
  if (smooth && !normals) { set_error("smoothNormals needs the normals output"); return B200_E_BAD_PARAM; }
  const GridView g = idx.view();
  const float md2 = std::isinf(maxDist) ? INFINITY : maxDist * maxDist;
  // B200_SN_TILE: 1 = cell-block tiles for uniformly sampled clouds, 2 = for every cloud. Default OFF: measured on the
  // 5 M-point map at k = 20 the tiles take 48 ms against 20.6 ms of the warp-cooperative kernel: a tile
  // evaluates every staged point of an 8^3-cell region per query (5-10x the candidates the box-growing search
  // touches) and its register-resident list caps the occupancy at 8 warps per SM.
  static const int tile_mode = [] { const char* e = getenv("B200_SN_TILE"); return (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 0; }();
  const bool uniform = idx.occupied > 0 && (double)idx.M <= 8.0 * (double)idx.occupied;
  const bool tile = knn >= 2 && knn <= 32 && (tile_mode == 2 || (tile_mode == 1 && uniform));
  if (knn > 32) {
    if (smooth) {   // the smoothing pass takes its integer neighbour ids from the k <= 32 kernels
      set_error("smoothNormals with knn > 32 is not supported by the B200 filter");
      return B200_E_UNSUPPORTED;
    }
    return surface_normals_device(g, feat, N, knn, maxDist, sortEigen, normals, dens, eigva, eigve, matched, meandist, s);
  }
  // scratch: fallback list + counter + ticket; neighbour ids, done flags and raw normals for the smoothing pass
  if ((size_t)N + 2 > idx.cap_sn_list) {
    cudaFree(idx.d_sn_list);
    idx.d_sn_list = nullptr; idx.cap_sn_list = 0;
    B200_CUDA(cudaMalloc(&idx.d_sn_list, ((size_t)N + 2) * sizeof(int)));
    idx.cap_sn_list = (size_t)N + 2;
  }
  int* ids = nullptr;
  if (smooth) {
    const size_t need = (size_t)N * knn + (size_t)N + (size_t)3 * N;   // ids, done flags, raw normals
    if (need > idx.cap_sn_ids) {
      cudaFree(idx.d_sn_ids);
      idx.d_sn_ids = nullptr; idx.cap_sn_ids = 0;
      B200_CUDA(cudaMalloc(&idx.d_sn_ids, need * sizeof(int)));
      idx.cap_sn_ids = need;
    }
    ids = idx.d_sn_ids;
  }
  int* list = idx.d_sn_list + 2;
  int* count = idx.d_sn_list;       // [0] fallback count, [1] smoothing ticket
  B200_CUDA(cudaMemsetAsync(count, 0, 2 * sizeof(int), s));
  // Queries are always walked in CELL-SORTED order (consecutive queries are neighbours in space: their cell rows and
  // candidates are still in L1 / L2); outputs land at the original index unless the caller wants them sorted too.
  const int ob = feat_sorted ? 0 : 1;
  (void)feat;
  if (tile) {
    SnOut o{normals, dens, eigva, eigve, matched, meandist, ids};
    if (knn <= 8) B200_TRY(launch_tile<8>(g, knn, md2, sortEigen, ob, o, list, count, s));
    else if (knn <= 16) B200_TRY(launch_tile<16>(g, knn, md2, sortEigen, ob, o, list, count, s));
    else if (knn <= 24) B200_TRY(launch_tile<24>(g, knn, md2, sortEigen, ob, o, list, count, s));
    else B200_TRY(launch_tile<32>(g, knn, md2, sortEigen, ob, o, list, count, s));
    // the tile kernel's list holds OUTPUT indices; with outputs by original index the exact path reads idx.d_ref
    B200_TRY(surface_normals_list_device(g, feat_sorted ? g.pts : idx.d_ref, list, count, knn, maxDist, sortEigen, normals,
                                         dens, eigva, eigve, matched, meandist, ids, 0, 0, s));
  } else {
    B200_TRY(surface_normals_list_device(g, g.pts, nullptr, nullptr, knn, maxDist, sortEigen, normals, dens, eigva, eigve,
                                         matched, meandist, ids, ob, N, s));
  }
  if (smooth) {
    int* flags = ids + (size_t)N * knn;
    float* raw = reinterpret_cast<float*>(flags + N);
    B200_CUDA(cudaMemsetAsync(flags, 0, (size_t)N * sizeof(int), s));
    B200_CUDA(cudaMemcpyAsync(raw, normals, (size_t)3 * N * sizeof(float), cudaMemcpyDeviceToDevice, s));
    smooth_normals_kernel<<<148 * 4, 128, 0, s>>>(normals, raw, ids, N, knn, count + 1, flags);
    B200_CUDA(cudaGetLastError());
  }
  return B200_OK;
}

// k nearest neighbours of every indexed point among the indexed points (the point itself included), within maxDist:
// ids (original indices, -1 = none) k per point, indexed by sorted position (out_sorted) or by original index.
// *ids_out points into idx's scratch and stays valid until the next normals / ids call on idx.
int self_knn_ids_index(MapIndex& idx, int knn, float maxDist, bool out_sorted, int** ids_out, cudaStream_t s) {
  const int N = idx.M;
  if (knn < 1 || knn > 32) { set_error("self_knn_ids_index: 1 <= knn <= 32"); return B200_E_UNSUPPORTED; }
  const size_t need = (size_t)N * knn;
  if (need > idx.cap_sn_ids) {
    cudaFree(idx.d_sn_ids);
    idx.d_sn_ids = nullptr; idx.cap_sn_ids = 0;
    B200_CUDA(cudaMalloc(&idx.d_sn_ids, need * sizeof(int)));
    idx.cap_sn_ids = need;
  }
  const GridView g = idx.view();
  *ids_out = idx.d_sn_ids;
  if (N == 0) return B200_OK;
  return surface_normals_list_device(g, g.pts, nullptr, nullptr, knn, maxDist, 0, nullptr, nullptr, nullptr, nullptr,
                                     nullptr, nullptr, idx.d_sn_ids, out_sorted ? 0 : 1, N, s);
}

// points the tile kernel handed to the exact path in the last call (diagnostics; synchronises the stream)
int surface_normals_last_fallback(const MapIndex& idx, int* n, cudaStream_t s) {
  *n = 0;
  if (!idx.d_sn_list) return B200_OK;
  B200_CUDA(cudaMemcpyAsync(n, idx.d_sn_list, sizeof(int), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  return B200_OK;
}

}  // namespace b200
