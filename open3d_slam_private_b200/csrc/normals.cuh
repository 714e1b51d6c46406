// SurfaceNormalDataPointsFilter on the device (SurfaceNormal.cpp:82-290, utils.h:84-150): pieces shared by the
// kernels of modules.cu (exact warp-cooperative / per-thread paths) and normals_tile.cu (cell-block tiles).
#pragma once
#include <cfloat>
#include "grid_index.cuh"
#include "small_la.cuh"

namespace b200 {

// SurfaceNormalDataPointsFilter::inPlaceFilter over the cloud the index was built on. feat = the queries: idx.d_ref
// (original order; outputs indexed by original index) or idx.d_pts (cell-sorted order; outputs indexed by sorted
// position) — feat_sorted says which. 2 <= knn <= 32 runs the cell-block tile kernel with the exact warp path as
// fallback for the points it cannot settle; other knn the per-thread path. smooth: SurfaceNormal.cpp:259-290 (needs
// normals != null). Grows idx's scratch on demand.
int surface_normals_index(MapIndex& idx, const float4* feat, bool feat_sorted, int N, int knn, float maxDist,
                          int sortEigen, int smooth, float* normals, float* dens, float* eigva, float* eigve,
                          float* matched, float* meandist, cudaStream_t s);

int surface_normals_last_fallback(const MapIndex& idx, int* n, cudaStream_t s);
int self_knn_ids_index(MapIndex& idx, int knn, float maxDist, bool out_sorted, int** ids_out, cudaStream_t s);

// modules.cu
int surface_normals_list_device(const GridView& g, const float4* feat, const int* d_list, const int* d_count, int knn,
                                float maxDist, int sortEigen, float* normals, float* dens, float* eigva, float* eigve,
                                float* matched, float* meandist, int* ids_out, int out_by_id, int N, cudaStream_t s);

// eigen-analysis of one neighbourhood covariance and the descriptor outputs (one thread per point)
__device__ __forceinline__ void sn_finish(int i, float4 q, const float* C, int realKnn, float maxSq, float mx, float my,
                                          float mz, int sortEigen, float* __restrict__ normals,
                                          float* __restrict__ dens, float* __restrict__ eigva,
                                          float* __restrict__ eigve, float* __restrict__ meandist) {
  la::PivQR<float, 3> qr;
  qr.compute(C);
  const bool degenerate = !(qr.rank() + 1 >= 3);   // :191
  float va[3] = {0.f, 0.f, 0.f}, ve[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) ve[t] = 0.f;
  if (!degenerate) {
    double Cd[9], wd[3], Vd[9];
    for (int t = 0; t < 9; ++t) Cd[t] = (double)C[t];
    la::symEigen3(Cd, wd, Vd);
    float v0[3], e0[9];
    for (int t = 0; t < 3; ++t) v0[t] = (float)wd[t];
    for (int t = 0; t < 9; ++t) e0[t] = (float)Vd[t];
    int idx[3] = {0, 1, 2};
    if (sortEigen) {  // stable ascending (sortIndexes, :199-210)
      if (v0[idx[1]] < v0[idx[0]]) { const int t = idx[0]; idx[0] = idx[1]; idx[1] = t; }
      if (v0[idx[2]] < v0[idx[1]]) { const int t = idx[1]; idx[1] = idx[2]; idx[2] = t; }
      if (v0[idx[1]] < v0[idx[0]]) { const int t = idx[0]; idx[0] = idx[1]; idx[1] = t; }
    }
    for (int c = 0; c < 3; ++c) {
      va[c] = v0[idx[c]];
      for (int r = 0; r < 3; ++r) ve[r + 3 * c] = e0[r + 3 * idx[c]];
    }
  }
  if (normals) {
    int col = 0;
    if (!sortEigen) {  // computeNormal, utils.h:131-150
      float best = FLT_MAX;
      for (int j = 0; j < 3; ++j)
        if (va[j] < best) { col = j; best = va[j]; }
    }
    for (int r = 0; r < 3; ++r) normals[3 * (size_t)i + r] = fminf(1.f, fmaxf(-1.f, ve[r + 3 * col]));
  }
  if (dens) {
    float dv = 0.f;
    if (!degenerate) {  // computeDensity, utils.h:105-129
      const float volume = __fmul_rn(__fmul_rn((float)(4. / 3.), (float)M_PI), __fmul_rn(maxSq, sqrtf(maxSq)));
      dv = volume > 0.f ? (float)realKnn / volume : 0.f;
    }
    dens[i] = dv;
  }
  if (eigva) for (int r = 0; r < 3; ++r) eigva[3 * (size_t)i + r] = va[r];
  if (eigve)
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) eigve[9 * (size_t)i + 3 * r + c] = ve[r + 3 * c];
  if (meandist) {
    float md;
    if (degenerate) md = 18446744073709551615.0f;  // float(numeric_limits<size_t>::max()), :245
    else {
      const float dx = __fsub_rn(q.x, mx), dy = __fsub_rn(q.y, my), dz = __fsub_rn(q.z, mz);
      md = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
    }
    meandist[i] = md;
  }
}


}  // namespace b200
