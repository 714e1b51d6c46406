// Exact k-nearest-neighbour query over the device-resident uniform grid (replaces libnabo's NNS::knn as called by
// KDTreeMatcher::findClosests, libpointmatcher/pointmatcher/MatchersImpl.cpp:85-102).
//
// Semantics (pinned bit-exact by the parity tests): squared L2 in float accumulated x->y->z with separate multiply and
// add (the reference is compiled without FMA, so __fmul_rn/__fadd_rn are used to stop nvcc contracting), candidates
// accepted iff dist2 <= maxDist^2, result ascending by (dist2, original index), ties -> smallest index, unfilled
// slots id = -1 / dist = +inf.
//
// Search. Everything is done in "cell units" u = (x - origin) * inv_cell, computed with the SAME float expression
// the index build uses, so the point -> cell map is monotone. A query scans the cells of a box
// [floor(u - h), floor(u + h)]^3, which contains every map point whose cell coordinates are within h of the query's
// on every axis, hence every point within distance (h - slack) * cell. After a box is scanned the radius still
// needed is r = min(sqrt(k-th best), maxDist); the query is finished as soon as h covers r, otherwise the box grows
// to r (or doubles while no bound exists) and only the new shell is scanned. Rows (runs of cells along x, ONE
// contiguous candidate range in the cell-sorted point array) farther than r from the query in (y, z) are skipped and
// the x-extent of a row is clipped to the ball. The answer is exact for ANY cell size and ANY maxDist (incl. inf).
//
// Two drivers share the row scanner:
//   knn_query<K>      one thread per query (k > 1; register-resident k-best list)
//   knn_query_warp1   k = 1, warp-cooperative: the (query, row) work items of the 32 queries of a warp are
//                     flattened and dealt round-robin to the 32 lanes, so a query with a far neighbour (or none
//                     within maxDist) no longer stalls the other 31 lanes — the tail that dominated the first version
//                     (ncu: 80 % of warp samples waiting at barriers).
#pragma once
#include "common.cuh"

namespace b200 {

template <int K>
struct KList {
  float d[K];
  int id[K];  // original index
  int sp[K];  // sorted position (for gathering normals)
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int j = 0; j < K; ++j) { d[j] = INFINITY; id[j] = -1; sp[j] = -1; }
  }
  __device__ __forceinline__ void offer(float dd, int iid, int ssp) {
    if (!(dd < d[K - 1] || (dd == d[K - 1] && iid < id[K - 1]))) return;
    d[K - 1] = dd; id[K - 1] = iid; sp[K - 1] = ssp;
#pragma unroll
    for (int j = K - 1; j > 0; --j) {
      const bool sw = d[j] < d[j - 1] || (d[j] == d[j - 1] && id[j] < id[j - 1]);
      if (sw) {
        const float td = d[j]; d[j] = d[j - 1]; d[j - 1] = td;
        const int ti = id[j]; id[j] = id[j - 1]; id[j - 1] = ti;
        const int ts = sp[j]; sp[j] = sp[j - 1]; sp[j - 1] = ts;
      }
    }
  }
};

__device__ __forceinline__ float dist2_rn(float qx, float qy, float qz, float px, float py, float pz) {
  const float dx = __fsub_rn(qx, px), dy = __fsub_rn(qy, py), dz = __fsub_rn(qz, pz);
  float d = __fmul_rn(dx, dx);
  d = __fadd_rn(d, __fmul_rn(dy, dy));
  d = __fadd_rn(d, __fmul_rn(dz, dz));
  return d;
}

// cell coordinate of a position along one axis — the SAME expression is used when the index is built, which makes
// the point -> cell map monotone and lets the query bound rounding with a relative slack.
__device__ __forceinline__ float cell_coord(float x, float o, float inv_cell) {
  return __fmul_rn(__fsub_rn(x, o), inv_cell);
}
__device__ __forceinline__ int cell_of(float u, int n) {
  int c = (int)floorf(u);
  return c < 0 ? 0 : (c >= n ? n - 1 : c);
}

// inclusive cell box, clamped to the grid; empty when lo > hi on any axis
struct CellBox {
  int x0, x1, y0, y1, z0, z1;
  __device__ __forceinline__ bool empty() const { return x0 > x1 || y0 > y1 || z0 > z1; }
  __device__ __forceinline__ int ny() const { return y1 - y0 + 1; }
  __device__ __forceinline__ int rows() const { return empty() ? 0 : (y1 - y0 + 1) * (z1 - z0 + 1); }
};

__device__ __forceinline__ void axis_range(float u, float h, int n, int& lo, int& hi) {
  // floor(u -+ h) clamped in float first (u, h may be huge), then to the grid
  const float a = fminf(fmaxf(floorf(u - h), -1.0f), (float)n);
  const float b = fminf(fmaxf(floorf(u + h), -1.0f), (float)n);
  lo = max((int)a, 0);
  hi = min((int)b, n - 1);
}

__device__ __forceinline__ CellBox ball_box(const GridView& g, float ux, float uy, float uz, float h) {
  CellBox b;
  axis_range(ux, h, g.nx, b.x0, b.x1);
  axis_range(uy, h, g.ny, b.y0, b.y1);
  axis_range(uz, h, g.nz, b.z0, b.z1);
  return b;
}

__device__ __forceinline__ bool covers_grid(const GridView& g, const CellBox& b) {
  return b.x0 <= 0 && b.y0 <= 0 && b.z0 <= 0 && b.x1 >= g.nx - 1 && b.y1 >= g.ny - 1 && b.z1 >= g.nz - 1;
}

// per-query constants of the search
struct QueryGeom {
  float qx, qy, qz;   // query position
  float ux, uy, uz;   // in cell units
  float slack;        // bound (cell units) on the rounding of any cell-unit difference involving this query
  float md2;          // maxDist^2 (inf allowed)
  float rmd;          // maxDist in cell units incl. slack (inf allowed)
};

__device__ __forceinline__ bool make_geom(const GridView& g, float qx, float qy, float qz, float md2, QueryGeom& q) {
  q.qx = qx; q.qy = qy; q.qz = qz; q.md2 = md2;
  q.ux = cell_coord(qx, g.ox, g.inv_cell);
  q.uy = cell_coord(qy, g.oy, g.inv_cell);
  q.uz = cell_coord(qz, g.oz, g.inv_cell);
  const float nmax = (float)max(g.nx, max(g.ny, g.nz));
  q.slack = fmaxf(fmaxf(fabsf(q.ux), fabsf(q.uy)), fmaxf(fabsf(q.uz), nmax)) * 4.8e-7f;  // ~2^-21 relative
  q.rmd = (md2 == INFINITY) ? INFINITY : sqrtf(md2) * g.inv_cell * 1.00001f + q.slack;
  return q.ux == q.ux && q.uy == q.uy && q.uz == q.uz;  // false for a NaN query: it has no matches
}

// radius (cell units, conservative) that still has to be covered given the current k-th best distance
__device__ __forceinline__ float need_radius(const GridView& g, const QueryGeom& q, float kth) {
  const float rk = (kth == INFINITY) ? INFINITY : sqrtf(kth) * g.inv_cell * 1.00001f + q.slack;
  return fminf(rk, q.rmd);
}

// first box half-width: the 2x2x2 block around the query (h = 0.5), smaller when maxDist is smaller; when the
// query lies outside the grid, far enough to reach it
__device__ __forceinline__ float first_halfwidth(const GridView& g, const QueryGeom& q) {
  float h = 0.5f;
  const float ox = fmaxf(fmaxf(-q.ux, q.ux - (float)g.nx), 0.f);
  const float oy = fmaxf(fmaxf(-q.uy, q.uy - (float)g.ny), 0.f);
  const float oz = fmaxf(fmaxf(-q.uz, q.uz - (float)g.nz), 0.f);
  const float out = fmaxf(ox, fmaxf(oy, oz));
  if (out > 0.f) h = out + 1.0f + q.slack;
  return fminf(h, q.rmd);
}

// ---- candidate range / row scanners ------------------------------------------------------------------------------
// visit(d, original index, sorted position) is called for every candidate with d <= maxDist^2
template <class Visit>
__device__ __forceinline__ void scan_range(const GridView& g, const QueryGeom& q, uint32_t a, uint32_t b,
                                           unsigned& visited, Visit&& visit) {
  visited += (b - a);
  uint32_t j = a;
  for (; j + 4u <= b; j += 4u) {  // four independent 16-byte loads in flight
    const float4 p0 = __ldg(&g.pts[j]), p1 = __ldg(&g.pts[j + 1]), p2 = __ldg(&g.pts[j + 2]), p3 = __ldg(&g.pts[j + 3]);
    const float d0 = dist2_rn(q.qx, q.qy, q.qz, p0.x, p0.y, p0.z);
    const float d1 = dist2_rn(q.qx, q.qy, q.qz, p1.x, p1.y, p1.z);
    const float d2 = dist2_rn(q.qx, q.qy, q.qz, p2.x, p2.y, p2.z);
    const float d3 = dist2_rn(q.qx, q.qy, q.qz, p3.x, p3.y, p3.z);
    if (d0 <= q.md2) visit(d0, __float_as_int(p0.w), (int)j);
    if (d1 <= q.md2) visit(d1, __float_as_int(p1.w), (int)j + 1);
    if (d2 <= q.md2) visit(d2, __float_as_int(p2.w), (int)j + 2);
    if (d3 <= q.md2) visit(d3, __float_as_int(p3.w), (int)j + 3);
  }
  if (j < b) {  // 1..3 left: one batch of clamped loads (a serial loop would pay the memory latency per candidate)
    const uint32_t j1 = min(j + 1u, b - 1u), j2 = min(j + 2u, b - 1u);
    const float4 p0 = __ldg(&g.pts[j]), p1 = __ldg(&g.pts[j1]), p2 = __ldg(&g.pts[j2]);
    const float d0 = dist2_rn(q.qx, q.qy, q.qz, p0.x, p0.y, p0.z);
    const float d1 = dist2_rn(q.qx, q.qy, q.qz, p1.x, p1.y, p1.z);
    const float d2 = dist2_rn(q.qx, q.qy, q.qz, p2.x, p2.y, p2.z);
    if (d0 <= q.md2) visit(d0, __float_as_int(p0.w), (int)j);
    if (j + 1u < b && d1 <= q.md2) visit(d1, __float_as_int(p1.w), (int)j1);
    if (j + 2u < b && d2 <= q.md2) visit(d2, __float_as_int(p2.w), (int)j2);
  }
}

// one row (iy, iz) of box B; cells of the previously scanned box P are skipped; rc = radius (cell units) beyond
// which nothing is needed any more (inf: no pruning)
template <class Visit>
__device__ __forceinline__ void scan_row(const GridView& g, const QueryGeom& q, int bx0, int bx1, int px0, int px1,
                                         int py0, int py1, int pz0, int pz1, float rc, int iy, int iz,
                                         unsigned& visited, Visit&& visit) {
  int xa = bx0, xb = bx1;
  if (rc != INFINITY) {
    const float dy = fmaxf(fmaxf((float)iy - q.uy, q.uy - (float)(iy + 1)) - q.slack, 0.f);
    const float dz = fmaxf(fmaxf((float)iz - q.uz, q.uz - (float)(iz + 1)) - q.slack, 0.f);
    const float rem = rc * rc - dy * dy - dz * dz;
    if (rem < 0.f) return;
    const float xh = sqrtf(rem) * 1.00001f + q.slack;
    int lo, hi;
    axis_range(q.ux, xh, g.nx, lo, hi);
    xa = max(xa, lo); xb = min(xb, hi);
    if (xa > xb) return;
  }
  const uint32_t* row = g.G + ((size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz));
  if (px0 <= px1 && iy >= py0 && iy <= py1 && iz >= pz0 && iz <= pz1) {
    const int e1 = min(xb, px0 - 1), s2 = max(xa, px1 + 1);
    const bool left = xa <= e1, right = s2 <= xb;
    uint32_t a0 = 0, b0 = 0, a1 = 0, b1 = 0;
    if (left) { a0 = __ldg(row + xa); b0 = __ldg(row + e1 + 1); }
    if (right) { a1 = __ldg(row + s2); b1 = __ldg(row + xb + 1); }
    if (left) scan_range(g, q, a0, b0, visited, visit);
    if (right) scan_range(g, q, a1, b1, visited, visit);
  } else {
    const uint32_t a = __ldg(row + xa), b = __ldg(row + xb + 1);
    scan_range(g, q, a, b, visited, visit);
  }
}

// ---- one thread per query, k-best list in registers --------------------------------------------------------------
// kq = number of neighbours actually requested (<= K)
template <int K>
__device__ void knn_query(const GridView& g, float qx, float qy, float qz, float md2, int kq, KList<K>& L,
                          unsigned& visited) {
  L.init();
  QueryGeom q;
  if (!make_geom(g, qx, qy, qz, md2, q)) return;  // NaN query: no matches
  auto visit = [&](float d, int id, int sp) { L.offer(d, id, sp); };
  if (fabsf(q.ux) > 1.0e9f || fabsf(q.uy) > 1.0e9f || fabsf(q.uz) > 1.0e9f) {
    // absurdly far query (cell arithmetic no longer resolves cells): scan everything, still exact
    scan_range(g, q, 0u, (uint32_t)g.M, visited, visit);
    return;
  }
  auto kth_of = [&]() {
    float kth = L.d[K - 1];
#pragma unroll
    for (int j = 0; j < K - 1; ++j)
      if (j == kq - 1) kth = L.d[j];  // register-resident select (no dynamic indexing)
    return kth;
  };
  float h = first_halfwidth(g, q);
  CellBox P;
  P.x0 = P.y0 = P.z0 = 1; P.x1 = P.y1 = P.z1 = 0;
  while (true) {
    const CellBox B = ball_box(g, q.ux, q.uy, q.uz, h + q.slack);
    if (!B.empty()) {
      for (int iz = B.z0; iz <= B.z1; ++iz)
        for (int iy = B.y0; iy <= B.y1; ++iy) {
          const float rc = need_radius(g, q, kth_of());  // shrinks while the scan finds closer points
          scan_row(g, q, B.x0, B.x1, P.x0, P.x1, P.y0, P.y1, P.z0, P.z1, rc, iy, iz, visited, visit);
        }
      P = B;
    }
    const float need = need_radius(g, q, kth_of());
    if (h >= need || covers_grid(g, B)) break;
    h = (need <= fmaxf(4.f * h, 3.f)) ? need : fmaxf(2.f * h, 1.f);
  }
}

// ---- k = 1, warp-cooperative -------------------------------------------------------------------------------------
// Must be called by all 32 lanes of a converged warp; lanes with active == false only help. Each pass, every lane
// that still needs cells owns nrows(lane) rows of its current box; the rows of all lanes are numbered consecutively
// (warp prefix sum) and dealt 32 x kItems at a time. Phase A resolves kItems rows per lane to candidate ranges (all
// cell-table loads of a round are in flight together); phase B scans them. A lane scanning a row for another lane's
// query keeps a local best and folds it into the owner's slot of the per-warp scratch with one shared-memory
// atomicMin (Warp1Scratch). (d, id) ordering with ties -> smallest id keeps the result independent of which lane
// scanned what. Cells already scanned by an earlier pass are scanned again
// (harmless for k = 1: the same (d, id) never replaces itself), which keeps an item one contiguous range.
__device__ __forceinline__ bool better1(float d, int id, float bd, int bid) {
  return d < bd || (d == bd && id < bid);
}

// next half-width once a box of half-width h did not settle the query: straight to the needed radius when that is
// known and not much larger; doubling otherwise (unknown radius, or a radius so large that a nearer shell probably
// already holds the neighbour)
__device__ __forceinline__ float next_halfwidth(float h, float need) {
  return (need <= fmaxf(4.f * h, 3.f)) ? need : fmaxf(2.f * h, 1.f);
}

constexpr int kItems = 4;

// Per-warp shared-memory scratch of the cooperative passes: every lane publishes the geometry of its query once per
// pass; the lanes that scan a row for it read what they need with three 16-byte loads (instead of 15 shuffles), and
// fold their findings with ONE shared-memory atomicMin on the key (bits(d) << 32 | id) — d >= 0, so the unsigned order
// of the key is the (d, id) order of the reference, ties -> smallest index — instead of a 20-shuffle segmented minimum.
struct __align__(16) Warp1Scratch {
  float4 geo[32];               // ux, uy, uz, slack
  float4 pos[32];               // qx, qy, qz, radius still needed (cell units)
  int4 box[32];                 // x0, x1, y0, z0 of the box
  float2 misc[32];              // maxDist^2, bits(number of rows along y)
  unsigned long long best[32];  // key of the best candidate so far
  int bsp[32];                  // its position in the cell-sorted array
  int incl[32];                 // inclusive prefix of the row counts
};

__device__ __forceinline__ unsigned long long key1(float d, int id) {
  return ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)id;
}

// Rows longer than this are scanned by the whole warp (32 consecutive candidates per load, one 512-byte request)
// instead of by the lane that owns them: on a non-uniform target (a raw LiDAR scan: cells near the sensor hold tens
// of points, far ones a single point) one lane walking a crowded row kept the other 31 waiting.
constexpr uint32_t kLongRow = 16u;

// Which rows to share: the warp scans the shared rows one after the other (~1.5 load round trips each), while a lane
// on its own needs len / 4 of them. Sharing the rows longer than theta pays while their number stays below theta / 6
// (one crowded cell among sparse ones: yes; a coherent warp that meets a dense patch as a whole: no). `len` = this
// lane's longest row (0: none). Returns the threshold (warp-uniform), 0xffffffff = share nothing.
__device__ __forceinline__ uint32_t long_row_threshold(uint32_t len) {
  uint32_t theta = kLongRow;
  while (theta <= 512u) {
    const int c = __popc(__ballot_sync(0xffffffffu, len > theta));
    if (c == 0) return 0xffffffffu;
    if ((uint32_t)c * 6u <= theta) return theta;
    theta <<= 1;
  }
  return 0xffffffffu;
}

// All 32 lanes scan candidates [a, b) for ONE query (warp-uniform arguments); every lane returns the warp's best
// (d, id, sorted position), id = -1 when no candidate lies within maxDist.
__device__ __forceinline__ void warp_scan_row1(const GridView& g, float qx, float qy, float qz, float md2, uint32_t a,
                                               uint32_t b, float& d, int& id, int& sp) {
  constexpr unsigned FULL = 0xffffffffu;
  const uint32_t lane = threadIdx.x & 31u;
  float ld = INFINITY;
  int lid = -1, lsp = -1;
  for (uint32_t j0 = a; j0 < b; j0 += 64u) {   // two independent loads per lane in flight
    const uint32_t j = j0 + lane, j2 = j + 32u;
    const float4 p0 = __ldg(&g.pts[min(j, b - 1u)]), p1 = __ldg(&g.pts[min(j2, b - 1u)]);
    const float d0 = dist2_rn(qx, qy, qz, p0.x, p0.y, p0.z), d1 = dist2_rn(qx, qy, qz, p1.x, p1.y, p1.z);
    if (j < b && d0 <= md2 && better1(d0, __float_as_int(p0.w), ld, lid)) { ld = d0; lid = __float_as_int(p0.w); lsp = (int)j; }
    if (j2 < b && d1 <= md2 && better1(d1, __float_as_int(p1.w), ld, lid)) { ld = d1; lid = __float_as_int(p1.w); lsp = (int)j2; }
  }
  const unsigned long long mine = key1(ld, lid);
  unsigned long long k = mine;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = __shfl_xor_sync(FULL, k, o);
    k = t < k ? t : k;
  }
  const unsigned who = __ballot_sync(FULL, mine == k);   // ids are unique: one lane unless nothing was found
  d = __uint_as_float((unsigned)(k >> 32));
  id = (int)(unsigned)(k & 0xffffffffull);
  sp = __shfl_sync(FULL, lsp, __ffs(who) - 1);
}

__device__ inline void knn_query_warp1(const GridView& g, bool active, float qx, float qy, float qz, float md2,
                                       float& best_d, int& best_id, int& best_sp, unsigned& visited, Warp1Scratch& ws,
                                       unsigned* rows_out = nullptr, unsigned* passes_out = nullptr,
                                       bool coop_first = false, int hint_sp = -1) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  best_d = INFINITY; best_id = -1; best_sp = -1;
  QueryGeom q;
  bool need = make_geom(g, qx, qy, qz, md2, q) && active;
  if (need && (fabsf(q.ux) > 1.0e9f || fabsf(q.uy) > 1.0e9f || fabsf(q.uz) > 1.0e9f)) {
    // absurdly far query: scan everything (exact); rare enough not to be shared
    auto visit = [&](float d, int id, int sp) { if (better1(d, id, best_d, best_id)) { best_d = d; best_id = id; best_sp = sp; } };
    scan_range(g, q, 0u, (uint32_t)g.M, visited, visit);
    need = false;
  }
  float h = need ? first_halfwidth(g, q) : 0.f;
  // Warm start: hint_sp is the sorted position of a point the caller believes to be close (the previous iteration's
  // match). Its distance bounds the neighbour's, so the first box shrinks from the 2x2x2 block to the cells the ball of
  // that radius touches — still every point at that distance or closer, so the answer (ties included) is unchanged.
  if (need && hint_sp >= 0) {
    const float4 hp = __ldg(&g.pts[hint_sp]);
    const float hd = dist2_rn(q.qx, q.qy, q.qz, hp.x, hp.y, hp.z);
    if (hd <= q.md2) {
      best_d = hd; best_id = __float_as_int(hp.w); best_sp = hint_sp;
      h = fminf(h, need_radius(g, q, hd));
    }
  }
  // ---- pass 1, per lane: the 2x2x2 block around the query. Every lane has the same <= 4 rows of work, so there is
  // nothing to balance and no shuffle traffic; near convergence this pass settles every query. All cell-table
  // loads are issued before the first candidate is touched.
  {
    CellBox B;
    B.x0 = B.y0 = B.z0 = 1; B.x1 = B.y1 = B.z1 = 0;
    // coop_first (warp-uniform): few lanes carry a query, so even the first box is dealt row by row to all 32 lanes
    bool p1 = need && h <= 0.5f && !coop_first;
    if (p1) {
      B = ball_box(g, q.ux, q.uy, q.uz, h + q.slack);
      p1 = B.y1 - B.y0 <= 1 && B.z1 - B.z0 <= 1;
    }
    uint32_t ra[4] = {0u, 0u, 0u, 0u}, rb[4] = {0u, 0u, 0u, 0u};
    if (p1) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int iy = B.y0 + (r & 1), iz = B.z0 + (r >> 1);
        if (iy <= B.y1 && iz <= B.z1 && B.x0 <= B.x1) {
          const uint32_t* row = g.G + ((size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz));
          ra[r] = __ldg(row + B.x0); rb[r] = __ldg(row + B.x1 + 1);
          if (rows_out) ++*rows_out;
        }
      }
    }
    // crowded rows: the whole warp scans them, one (lane, row) after the other
    uint32_t mylen = 0u;
#pragma unroll
    for (int r = 0; r < 4; ++r) mylen = max(mylen, rb[r] - ra[r]);
    if (__any_sync(FULL, mylen > kLongRow)) {
    const uint32_t theta = long_row_threshold(mylen);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const bool longr = rb[r] - ra[r] > theta;
      unsigned m = __ballot_sync(FULL, longr);
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const uint32_t a = __shfl_sync(FULL, ra[r], src), b = __shfl_sync(FULL, rb[r], src);
        const float sx = __shfl_sync(FULL, q.qx, src), sy = __shfl_sync(FULL, q.qy, src), sz = __shfl_sync(FULL, q.qz, src);
        const float smd = __shfl_sync(FULL, q.md2, src);
        float d; int id, sp;
        warp_scan_row1(g, sx, sy, sz, smd, a, b, d, id, sp);
        if (lane == src) {
          visited += b - a;
          if (id >= 0 && better1(d, id, best_d, best_id)) { best_d = d; best_id = id; best_sp = sp; }
        }
      }
      if (longr) rb[r] = ra[r];   // done
    }
    }
    if (p1) {
      auto visit = [&](float d, int id, int sp) { if (better1(d, id, best_d, best_id)) { best_d = d; best_id = id; best_sp = sp; } };
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (rb[r] > ra[r]) scan_range(g, q, ra[r], rb[r], visited, visit);
      const float nd = need_radius(g, q, best_d);
      if (h >= nd || covers_grid(g, B)) need = false;
      else h = next_halfwidth(h, nd);
    }
  }
  while (__any_sync(FULL, need)) {
    if (passes_out) ++*passes_out;
    CellBox B;
    B.x0 = B.y0 = B.z0 = 1; B.x1 = B.y1 = B.z1 = 0;
    float rc = INFINITY;
    if (need) {
      B = ball_box(g, q.ux, q.uy, q.uz, h + q.slack);
      rc = need_radius(g, q, best_d);
    }
    const int nrows = need ? B.rows() : 0;
    int incl = nrows;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(FULL, incl, o);
      if (lane >= o) incl += t;
    }
    const int excl = incl - nrows;
    const int total = __shfl_sync(FULL, incl, 31);
    __syncwarp();   // the previous pass (or call) is done with the scratch
    ws.incl[lane] = incl;
    if (need) {
      ws.geo[lane] = make_float4(q.ux, q.uy, q.uz, q.slack);
      ws.pos[lane] = make_float4(q.qx, q.qy, q.qz, rc);
      ws.box[lane] = make_int4(B.x0, B.x1, B.y0, B.z0);
      ws.misc[lane] = make_float2(q.md2, __int_as_float(B.ny()));
      // an unmatched query carries id -1 = 0xffffffff in the key: above every real id at the same (infinite) distance
      ws.best[lane] = key1(best_d, best_id);
      ws.bsp[lane] = best_sp;
    }
    __syncwarp();
    for (int base = 0; base < total; base += 32 * kItems) {
      // ---- phase A: item -> owner -> row -> candidate range (all cell-table loads of the round in flight together)
      uint32_t ra[kItems], rb[kItems];
      int own[kItems];
#pragma unroll
      for (int s = 0; s < kItems; ++s) {
        const int t = base + s * 32 + lane;
        int o = 0;  // owner of item t: the first lane whose inclusive prefix exceeds t
#pragma unroll
        for (int step = 16; step > 0; step >>= 1)
          if (ws.incl[(o + step - 1) & 31] <= t) o += step;
        const bool work = t < total;
        o &= 31;
        ra[s] = 0u; rb[s] = 0u;
        own[s] = work ? o : -1;
        if (work && rows_out) ++*rows_out;
        if (work) {
          const float4 og = ws.geo[o];   // ux, uy, uz, slack
          const int4 ob = ws.box[o];     // x0, x1, y0, z0
          const float orc = ws.pos[o].w;
          const int ony = __float_as_int(ws.misc[o].y);
          const int r = t - (o ? ws.incl[o - 1] : 0);
          const int rz = (int)(((float)r + 0.5f) * __frcp_rn((float)ony));  // exact: r, ony < 2^20
          const int iz = ob.w + rz, iy = ob.z + (r - rz * ony);
          int xa = ob.x, xb = ob.y;
          bool skip = false;
          if (orc != INFINITY) {  // clip the row to the ball of the radius still needed
            const float dy = fmaxf(fmaxf((float)iy - og.y, og.y - (float)(iy + 1)) - og.w, 0.f);
            const float dz = fmaxf(fmaxf((float)iz - og.z, og.z - (float)(iz + 1)) - og.w, 0.f);
            const float rem = orc * orc - dy * dy - dz * dz;
            if (rem < 0.f) skip = true;
            else {
              const float xh = sqrtf(rem) * 1.00001f + og.w;
              int lo, hi;
              axis_range(og.x, xh, g.nx, lo, hi);
              xa = max(xa, lo); xb = min(xb, hi);
              skip = xa > xb;
            }
          }
          if (!skip) {
            const uint32_t* row = g.G + ((size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz));
            ra[s] = __ldg(row + xa); rb[s] = __ldg(row + xb + 1);
          }
        }
      }
      // ---- phase B: scan, fold into the owner's key
      uint32_t mylen = 0u;
#pragma unroll
      for (int s = 0; s < kItems; ++s) mylen = max(mylen, rb[s] - ra[s]);
      const uint32_t theta = __any_sync(FULL, mylen > kLongRow) ? long_row_threshold(mylen) : 0xffffffffu;
#pragma unroll
      for (int s = 0; s < kItems; ++s) {
        if (base + s * 32 >= total) break;  // warp-uniform
        const int o = own[s] & 31;
        float ld = INFINITY;
        int lid = -1, lsp = -1;
        const bool has = own[s] >= 0 && rb[s] > ra[s];
        const bool longr = has && rb[s] - ra[s] > theta;
        unsigned lm = theta != 0xffffffffu ? __ballot_sync(FULL, longr) : 0u;
        while (lm) {   // crowded rows: the whole warp scans them, one item after the other
          const int src = __ffs(lm) - 1;
          lm &= lm - 1u;
          const uint32_t a = __shfl_sync(FULL, ra[s], src), b = __shfl_sync(FULL, rb[s], src);
          const int so = __shfl_sync(FULL, o, src);
          const float4 op = ws.pos[so];
          float d; int id, sp;
          warp_scan_row1(g, op.x, op.y, op.z, ws.misc[so].x, a, b, d, id, sp);
          if (lane == src) { visited += b - a; ld = d; lid = id; lsp = sp; }
        }
        if (has && !longr) {
          const float4 op = ws.pos[o];
          QueryGeom oq;
          oq.qx = op.x; oq.qy = op.y; oq.qz = op.z; oq.md2 = ws.misc[o].x;
          auto visit = [&](float d, int id, int sp) { if (better1(d, id, ld, lid)) { ld = d; lid = id; lsp = sp; } };
          scan_range(g, oq, ra[s], rb[s], visited, visit);
        }
        const unsigned long long k = key1(ld, lid);
        if (lid >= 0) atomicMin(&ws.best[o], k);
        __syncwarp();
        if (lid >= 0 && ws.best[o] == k) ws.bsp[o] = lsp;   // the one lane that holds the minimum (ids are unique)
      }
    }
    __syncwarp();
    if (need) {
      const unsigned long long k = ws.best[lane];
      best_d = __uint_as_float((unsigned)(k >> 32));
      best_id = (int)(unsigned)(k & 0xffffffffull);
      best_sp = ws.bsp[lane];
      const float nd = need_radius(g, q, best_d);
      if (h >= nd || covers_grid(g, B)) need = false;
      else h = next_halfwidth(h, nd);
    }
  }
}

// ---- 2 <= k <= 32, one warp per query ---------------------------------------------------------------------------
// The k-best list lives across the warp: lane j holds the j-th best (d, id, sorted position), ascending in (d, id)
// — the reference's order (ties -> smallest index). Candidates arrive 32 at a time, one per lane: when only a few of
// them beat the current k-th best they are inserted one by one (ballot + shift), otherwise the batch is sorted with a
// bitonic network and merged (min(L[i], C[31 - i]) + bitonic merge). Empty slots carry (inf, INT_MAX, -1).
constexpr int kNoId = 0x7fffffff;

__device__ __forceinline__ bool key_less(float d, int id, float od, int oid) { return d < od || (d == od && id < oid); }

__device__ __forceinline__ void warp_cmpx(float& d, int& id, int& sp, int j, bool keep_min) {
  const float od = __shfl_xor_sync(0xffffffffu, d, j);
  const int oid = __shfl_xor_sync(0xffffffffu, id, j);
  const int osp = __shfl_xor_sync(0xffffffffu, sp, j);
  const bool oless = key_less(od, oid, d, id);
  if (keep_min == oless) { d = od; id = oid; sp = osp; }   // equal keys only occur between empty slots
}

struct WarpKList {
  float d; int id; int sp;        // this lane's element
  float kth_d; int kth_id;        // element kq - 1 (warp-uniform)
  int kq;
  __device__ __forceinline__ void init(int kq_) { d = INFINITY; id = kNoId; sp = -1; kth_d = INFINITY; kth_id = kNoId; kq = kq_; }
  __device__ __forceinline__ void refresh() {
    kth_d = __shfl_sync(0xffffffffu, d, kq - 1);
    kth_id = __shfl_sync(0xffffffffu, id, kq - 1);
  }
  // all 32 lanes call; has = this lane carries a candidate
  __device__ __forceinline__ void offer(bool has, float cd, int cid, int csp) {
    const int lane = threadIdx.x & 31;
    has = has && cd != INFINITY && key_less(cd, cid, kth_d, kth_id);
    unsigned m = __ballot_sync(0xffffffffu, has);
    if (m == 0u) return;
    if (__popc(m) <= 6) {
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1u;
        const float bd = __shfl_sync(0xffffffffu, cd, src);
        const int bid = __shfl_sync(0xffffffffu, cid, src);
        const int bsp = __shfl_sync(0xffffffffu, csp, src);
        if (!key_less(bd, bid, kth_d, kth_id)) continue;   // the k-th best moved meanwhile (warp-uniform test)
        const int pos = __popc(__ballot_sync(0xffffffffu, key_less(d, id, bd, bid)));   // list sorted: a prefix
        const float ud = __shfl_up_sync(0xffffffffu, d, 1);
        const int uid = __shfl_up_sync(0xffffffffu, id, 1);
        const int usp = __shfl_up_sync(0xffffffffu, sp, 1);
        if (lane > pos) { d = ud; id = uid; sp = usp; }
        else if (lane == pos) { d = bd; id = bid; sp = bsp; }
        refresh();
      }
      return;
    }
    if (!has) { cd = INFINITY; cid = kNoId; csp = -1; }
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1)
#pragma unroll
      for (int j = k >> 1; j > 0; j >>= 1) warp_cmpx(cd, cid, csp, j, ((lane & j) == 0) == ((lane & k) == 0));
    const float rd = __shfl_sync(0xffffffffu, cd, 31 - lane);
    const int rid = __shfl_sync(0xffffffffu, cid, 31 - lane);
    const int rsp = __shfl_sync(0xffffffffu, csp, 31 - lane);
    if (key_less(rd, rid, d, id)) { d = rd; id = rid; sp = rsp; }
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) warp_cmpx(d, id, sp, j, (lane & j) == 0);
    refresh();
  }
};

// candidate ranges of one row (iy, iz) of box B for a query: like scan_row, but only resolves the (at most two)
// ranges of sorted positions; n0 / n1 = their lengths
__device__ __forceinline__ void row_ranges(const GridView& g, const QueryGeom& q, const CellBox& B, const CellBox& P,
                                           float rc, int iy, int iz, uint32_t& a0, uint32_t& n0, uint32_t& a1,
                                           uint32_t& n1) {
  a0 = a1 = n0 = n1 = 0u;
  int xa = B.x0, xb = B.x1;
  if (rc != INFINITY) {
    const float dy = fmaxf(fmaxf((float)iy - q.uy, q.uy - (float)(iy + 1)) - q.slack, 0.f);
    const float dz = fmaxf(fmaxf((float)iz - q.uz, q.uz - (float)(iz + 1)) - q.slack, 0.f);
    const float rem = rc * rc - dy * dy - dz * dz;
    if (rem < 0.f) return;
    const float xh = sqrtf(rem) * 1.00001f + q.slack;
    int lo, hi;
    axis_range(q.ux, xh, g.nx, lo, hi);
    xa = max(xa, lo); xb = min(xb, hi);
    if (xa > xb) return;
  }
  const uint32_t* row = g.G + ((size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz));
  if (P.x0 <= P.x1 && iy >= P.y0 && iy <= P.y1 && iz >= P.z0 && iz <= P.z1) {
    const int e1 = min(xb, P.x0 - 1), s2 = max(xa, P.x1 + 1);
    if (xa <= e1) { a0 = __ldg(row + xa); n0 = __ldg(row + e1 + 1) - a0; }
    if (s2 <= xb) { a1 = __ldg(row + s2); n1 = __ldg(row + xb + 1) - a1; }
  } else {
    a0 = __ldg(row + xa);
    n0 = __ldg(row + xb + 1) - a0;
  }
}

// All 32 lanes call with the SAME query. On return lane j holds the j-th nearest neighbour (d = inf, id = -1 when
// there is none); lanes >= kq hold leftovers that must be ignored.
__device__ inline void knn_query_warpk(const GridView& g, float qx, float qy, float qz, float md2, int kq,
                                       float& out_d, int& out_id, int& out_sp, unsigned& visited) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  WarpKList L;
  L.init(kq);
  QueryGeom q;
  const bool valid = make_geom(g, qx, qy, qz, md2, q);
  if (valid && (fabsf(q.ux) > 1.0e9f || fabsf(q.uy) > 1.0e9f || fabsf(q.uz) > 1.0e9f)) {
    // absurdly far query (cell arithmetic no longer resolves cells): scan everything, still exact
    for (uint32_t base = 0; base < (uint32_t)g.M; base += 32u) {
      const uint32_t j = base + lane;
      bool has = j < (uint32_t)g.M;
      float d = INFINITY; int id = kNoId;
      if (has) {
        const float4 p = __ldg(&g.pts[j]);
        d = dist2_rn(qx, qy, qz, p.x, p.y, p.z);
        id = __float_as_int(p.w);
        has = d <= md2;
      }
      L.offer(has, d, id, (int)j);
    }
    if (lane == 0) visited += (unsigned)g.M;
  } else if (valid) {
    float h = first_halfwidth(g, q);
    // First box for a long list: the 2x2x2 block holds ~24 points at the index's 3 points per cell, too few to bound the
    // kq-th neighbour of a kq >= 8 query, so a second and a third pass follow. Where the query's own cell is sparse
    // (<= kq / 3 points) the search starts one cell wide instead (27 cells, fewer passes: 5 M-point map, k = 20:
    // 14.6 -> 11.8 ms); where it is crowded (a raw scan near the sensor) the narrow box stays — a wide one there
    // costs 20 % (measured with the width fixed at 1.0 and 1.5 cells). Any width is exact.
    if (kq >= 8 && q.ux >= 0.f && q.uy >= 0.f && q.uz >= 0.f && q.ux < (float)g.nx && q.uy < (float)g.ny && q.uz < (float)g.nz) {
      const size_t c = (size_t)cell_of(q.ux, g.nx) + (size_t)g.nx * ((size_t)cell_of(q.uy, g.ny) + (size_t)g.ny * (size_t)cell_of(q.uz, g.nz));
      const uint32_t own = __ldg(g.G + c + 1) - __ldg(g.G + c);
      if (3u * own <= (uint32_t)kq) h = fminf(fmaxf(h, 1.0f), q.rmd);
    }
    CellBox P;
    P.x0 = P.y0 = P.z0 = 1; P.x1 = P.y1 = P.z1 = 0;
    while (true) {
      const CellBox B = ball_box(g, q.ux, q.uy, q.uz, h + q.slack);
      const int nrows = B.rows(), ny = B.ny();
      for (int rb = 0; rb < nrows; rb += 32) {
        const float rc = need_radius(g, q, L.kth_d);   // shrinks while the scan finds closer points
        uint32_t a0 = 0, n0 = 0, a1 = 0, n1 = 0;
        const int r = rb + lane;
        if (r < nrows) row_ranges(g, q, B, P, rc, B.y0 + r % ny, B.z0 + r / ny, a0, n0, a1, n1);
        const uint32_t tot = n0 + n1;
        uint32_t incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t v = __shfl_up_sync(FULL, incl, o);
          if (lane >= o) incl += v;
        }
        const uint32_t total = __shfl_sync(FULL, incl, 31);
        if (lane == 0) visited += total;
        for (uint32_t base = 0; base < total; base += 32u) {
          const uint32_t t = base + lane;
          // owner row of candidate t: the first lane whose inclusive prefix exceeds t
          int lo = 0;
#pragma unroll
          for (int step = 16; step > 0; step >>= 1) {
            const uint32_t v = __shfl_sync(FULL, incl, lo + step - 1);
            if (v <= t) lo += step;
          }
          lo = min(lo, 31);
          const uint32_t o_incl = __shfl_sync(FULL, incl, lo), o_tot = __shfl_sync(FULL, tot, lo);
          const uint32_t o_a0 = __shfl_sync(FULL, a0, lo), o_n0 = __shfl_sync(FULL, n0, lo), o_a1 = __shfl_sync(FULL, a1, lo);
          bool has = t < total;
          float d = INFINITY; int id = kNoId; int sp = -1;
          if (has) {
            const uint32_t k = t - (o_incl - o_tot);
            const uint32_t j = k < o_n0 ? o_a0 + k : o_a1 + (k - o_n0);
            const float4 p = __ldg(&g.pts[j]);
            d = dist2_rn(qx, qy, qz, p.x, p.y, p.z);
            id = __float_as_int(p.w);
            sp = (int)j;
            has = d <= md2;
          }
          L.offer(has, d, id, sp);
        }
      }
      if (!B.empty()) P = B;
      const float need = need_radius(g, q, L.kth_d);
      if (h >= need || covers_grid(g, B)) break;
      h = next_halfwidth(h, need);
    }
  }
  out_d = L.d;
  out_id = L.d == INFINITY ? -1 : L.id;
  out_sp = L.d == INFINITY ? -1 : L.sp;
}

}  // namespace b200
