// Exact k-nearest-neighbour query over the device-resident uniform grid (replaces libnabo's NNS::knn as called by
// KDTreeMatcher::findClosests, libpointmatcher/pointmatcher/MatchersImpl.cpp:85-102).
//
// Semantics (pinned bit-exact by the parity tests): squared L2 in float accumulated x->y->z with separate multiply and
// add (the reference is compiled without FMA, so __fmul_rn/__fadd_rn are used to stop nvcc contracting), candidates
// accepted iff dist2 <= maxDist^2, result ascending by (dist2, original index), ties -> smallest index, unfilled
// slots id = -1 / dist = +inf.
//
// Search: cells are visited in growing cubic blocks around the query; a block of half-width s guarantees that every
// unvisited point is farther than a conservative radius r(s). The query stops as soon as its k-th best is within
// r(s) or r(s) exceeds maxDist, so the answer is exact for ANY cell size and ANY maxDist (including inf).
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

template <int K>
__device__ __forceinline__ void scan_range(const GridView& g, uint32_t a, uint32_t b, float qx, float qy, float qz,
                                           float md2, KList<K>& L, unsigned& visited) {
  for (uint32_t j = a; j < b; ++j) {
    const float4 p = __ldg(&g.pts[j]);
    const float d = dist2_rn(qx, qy, qz, p.x, p.y, p.z);
    if (d <= md2) L.offer(d, __float_as_int(p.w), (int)j);
  }
  visited += (b - a);
}

// kq = number of neighbours actually requested (<= K)
template <int K>
__device__ void knn_query(const GridView& g, float qx, float qy, float qz, float md2, int kq, KList<K>& L,
                          unsigned& visited) {
  L.init();
  const float ux = cell_coord(qx, g.ox, g.inv_cell);
  const float uy = cell_coord(qy, g.oy, g.inv_cell);
  const float uz = cell_coord(qz, g.oz, g.inv_cell);
  if (!(ux == ux && uy == uy && uz == uz)) return;  // NaN query: no matches
  const float lim = 1.0e9f;
  if (fabsf(ux) > lim || fabsf(uy) > lim || fabsf(uz) > lim) {
    // absurdly far query: every map point is a candidate only if maxDist allows it; scan everything (exact)
    scan_range<K>(g, 0u, (uint32_t)g.M, qx, qy, qz, md2, L, visited);
    return;
  }
  const float nmax = (float)max(g.nx, max(g.ny, g.nz));
  const float slack = fmaxf(fmaxf(fabsf(ux), fabsf(uy)), fmaxf(fabsf(uz), nmax)) * 4.8e-7f;  // ~2^-21 relative
  const int bx = (int)floorf(ux - 0.5f), by = (int)floorf(uy - 0.5f), bz = (int)floorf(uz - 0.5f);
  // first stage whose block touches the grid on every axis
  int s = 0;
  s = max(s, bx - (g.nx - 1)); s = max(s, -(bx + 1));
  s = max(s, by - (g.ny - 1)); s = max(s, -(by + 1));
  s = max(s, bz - (g.nz - 1)); s = max(s, -(bz + 1));
  // previously scanned block (clamped, inclusive); empty at start
  int px0 = 1, px1 = 0, py0 = 1, py1 = 0, pz0 = 1, pz1 = 0;
  while (true) {
    const int lx = bx - s, hx = bx + 1 + s, ly = by - s, hy = by + 1 + s, lz = bz - s, hz = bz + 1 + s;
    const int cx0 = max(lx, 0), cx1 = min(hx, g.nx - 1);
    const int cy0 = max(ly, 0), cy1 = min(hy, g.ny - 1);
    const int cz0 = max(lz, 0), cz1 = min(hz, g.nz - 1);
    for (int iz = cz0; iz <= cz1; ++iz) {
      const bool zin = (iz >= pz0 && iz <= pz1);
      for (int iy = cy0; iy <= cy1; ++iy) {
        const uint32_t* row = g.G + ((size_t)g.nx * ((size_t)iy + (size_t)g.ny * (size_t)iz));
        if (zin && iy >= py0 && iy <= py1) {
          // row crosses the already scanned block: only the two new end segments
          if (cx0 < px0) scan_range<K>(g, __ldg(row + cx0), __ldg(row + px0), qx, qy, qz, md2, L, visited);
          if (cx1 > px1) scan_range<K>(g, __ldg(row + px1 + 1), __ldg(row + cx1 + 1), qx, qy, qz, md2, L, visited);
        } else {
          scan_range<K>(g, __ldg(row + cx0), __ldg(row + cx1 + 1), qx, qy, qz, md2, L, visited);
        }
      }
    }
    px0 = cx0; px1 = cx1; py0 = cy0; py1 = cy1; pz0 = cz0; pz1 = cz1;
    // conservative distance (in cells) from the query to the nearest face beyond which unvisited points may lie
    float gmin = INFINITY;
    if (lx > 0) gmin = fminf(gmin, ux - (float)lx);
    if (hx < g.nx - 1) gmin = fminf(gmin, (float)(hx + 1) - ux);
    if (ly > 0) gmin = fminf(gmin, uy - (float)ly);
    if (hy < g.ny - 1) gmin = fminf(gmin, (float)(hy + 1) - uy);
    if (lz > 0) gmin = fminf(gmin, uz - (float)lz);
    if (hz < g.nz - 1) gmin = fminf(gmin, (float)(hz + 1) - uz);
    if (gmin == INFINITY) break;  // the block covers the whole grid
    const float gc = (gmin - slack) * g.cell * 0.999998f;
    const float r2 = gc > 0.f ? gc * gc : 0.f;
    float kth = L.d[K - 1];
#pragma unroll
    for (int j = 0; j < K - 1; ++j)
      if (j == kq - 1) kth = L.d[j];  // register-resident select (no dynamic indexing)
    if (kth <= r2) break;  // k-th best is closer than anything unvisited
    if (r2 > md2) break;           // nothing unvisited can be accepted
    s += 1 + (s >> 1);
  }
}

}  // namespace b200
