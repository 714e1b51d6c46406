// Stand-alone (one call = one module) device entry points behind the per-module C ABI. All pointers are DEVICE
// pointers; the C ABI layer (c_api.cu) owns staging. The fused whole-registration path lives in icp_fused.cu.
#pragma once
#include "common.cuh"

namespace b200 {

// exposed by grid_index.cu
int inclusive_scan_u32(uint32_t* d_data, size_t n, uint32_t* d_tile /* >= ceil(n/4096) */, cudaStream_t s);
size_t scan_tiles_for(size_t n);

// KDTreeMatcher::findClosests (MatchersImpl.cpp:85-102); kq <= 64
int knn_device(const GridView& g, const float4* d_query, int N, int kq, float maxDist, const float* d_maxDistPerPoint,
               int32_t* d_ids, float* d_dists, unsigned long long* d_visited, cudaStream_t s);

// Matches::getDistsQuantile (Matches.cpp:60-87). d_state: >= 3*2048 + 16 uint32 scratch. Result: d_state[0] =
// status (b200_status), d_state[1] = bits of the quantile value.
constexpr size_t kQuantileScratchWords = 3 * 2048 + 16;
int quantile_device(const float* d_dists, int64_t n, float quantile, uint32_t* d_state, cudaStream_t s);

// weights from a distance predicate. kind: b200_outlier_kind. For MEDIAN/TRIMMED the limit is read from
// d_state[1] (and multiplied by `param` for MEDIAN); nothing is written when d_state[0] != OK.
int outlier_dist_device(int kind, float param, const float* d_dists, int64_t n, const uint32_t* d_state, float* d_w,
                        cudaStream_t s);
int outlier_surface_normal_device(const float* d_nread3, const float* d_nref3, const int32_t* d_ids, int k, int N,
                                  float cos_max_angle, float* d_w, cudaStream_t s);

// ErrorElements (ErrorMinimizer.cpp:59-193): order-preserving compaction. d_flags/d_tile scratch: k*N + tiles words.
// d_sums: 2 doubles (P as double, sum of kept weights).
int error_elements_device(const float4* d_read, const float4* d_ref, const float* d_refn3, const float* d_w,
                          const int32_t* d_ids, const float* d_dists, int k, int N, uint32_t* d_flags,
                          uint32_t* d_tile, float4* d_readP, float4* d_refP, float* d_refnP, float* d_wP,
                          int32_t* d_idsP, float* d_distsP, cudaStream_t s);

// calculateOptimizationHessian (ICP.cpp:1513-1566) on compacted pairs; d_partials >= 1024*32 doubles;
// d_out: 36 floats H (column-major) + 6 floats b
int p2plane_hessian_device(const float4* d_read, const float4* d_ref, const float* d_refn3, const float* d_w, int P,
                           double* d_partials, float* d_out, cudaStream_t s);
// solve + SE(3): d_io layout: in H[36] b[6]; out x[6] T[16] info(int bits)
int p2plane_solve_device(float* d_io, cudaStream_t s);
// PointToPoint (PointToPoint.cpp:62-103); d_T: 16 floats
int p2point_device(const float4* d_read, const float4* d_ref, const float* d_w, int P, double* d_partials, float* d_T,
                   cudaStream_t s);
// RigidTransformation::inPlaceCompute (TransformationsImpl.cpp:61-102); d_T16 device pointer to the 4x4
int rigid_apply_device(const float* d_T16, float4* d_feat, float* d_normals3, int N, cudaStream_t s);

// SurfaceNormalDataPointsFilter (SurfaceNormal.cpp:82-290) over an index built on the same cloud.
// Any output may be null. knn <= 64.
int surface_normals_device(const GridView& g, const float4* d_feat, int N, int knn, float maxDist, int sortEigen,
                           float* d_normals3, float* d_dens, float* d_eigva3, float* d_eigve9, float* d_matched,
                           float* d_meandist, cudaStream_t s);

}  // namespace b200
