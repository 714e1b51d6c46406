// ORACLE — TEST INFRASTRUCTURE ONLY. Scan pre-processing that runs before the registration hot path
// (SURVEY 8(f) rank 2): open3d_slam's croppers and Open3D's VoxelDownSample as called from
// open3d_slam/src/ScanToMapRegistration.cpp:36-54, Odometry.cpp:22-26 and helpers.cpp:108-115.
//
// Croppers are in the reference tree (open3d_slam/src/croppers.cpp) and restated literally. VoxelDownSample lives in
// Open3D v0.15.1, which the reference fetches at build time and is not in the tree: restated as recalled from upstream
// (geometry/PointCloud.cpp, VoxelDownSample): voxel_min_bound = min_bound - voxel_size / 2, voxel index =
// floor((p - voxel_min_bound) / voxel_size) per axis, points (and normals) of a voxel averaged, accumulated in the
// order of the input; the output order of Open3D is the iteration order of an std::unordered_map (unspecified) —
// here: ascending index of the first point of the voxel. Parity unpinned to the reference binary for this function.
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <tuple>
#include <vector>

extern "C" {

// CroppingVolume::isWithinVolume (croppers.cpp:53-55) for the volumes of croppingVolumeFactory (:25-47).
// kind 0 base class (everything), 1 Cylinder (:162-164), 2 MinRadius (:149-151), 3 MaxRadius (:131-133),
// 4 MinMaxRadius (:115-118). Points are floats widened to double (the clouds reach Open3D as doubles made from the
// sensor's floats). Returns the number of kept points; kept_idx (capacity N) in input order (crop, :66-95).
int orc_crop(int kind, double radius_min, double radius_max, double min_z, double max_z, const double* center3,
             int invert, const float* pts4xN, int N, int32_t* kept_idx) {
  int n = 0;
  for (int i = 0; i < N; ++i) {
    const double px = pts4xN[4 * (size_t)i], py = pts4xN[4 * (size_t)i + 1], pz = pts4xN[4 * (size_t)i + 2];
    const double dx = px - center3[0], dy = py - center3[1], dz = pz - center3[2];
    bool in = true;
    switch (kind) {
      case 1: in = pz >= min_z && pz <= max_z && std::sqrt(dx * dx + dy * dy) <= radius_max; break;
      case 2: in = std::sqrt(dx * dx + dy * dy + dz * dz) >= radius_min; break;
      case 3: in = std::sqrt(dx * dx + dy * dy + dz * dz) <= radius_max; break;
      case 4: { const double d = std::sqrt(dx * dx + dy * dy + dz * dz); in = d <= radius_max && d >= radius_min; } break;
      default: in = true;
    }
    if (invert) in = !in;
    if (in) kept_idx[n++] = i;
  }
  return n;
}

// PointCloud::VoxelDownSample(voxel_size). out3xK / outn3xK (nullable) / first_idx have capacity N. Returns K, or -1
// for voxel_size <= 0 ("voxel_size <= 0") and -2 when the voxel size is too small for the extent (Open3D's check
// voxel_size * INT_MAX < max extent).
int orc_voxel_downsample(double voxel_size, const float* pts4xN, const float* nrm3xN, int N, double* out3xK,
                         double* outn3xK, int32_t* first_idx) {
  if (!(voxel_size > 0.0)) return -1;
  if (N <= 0) return 0;
  double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = 0; i < N; ++i)
    for (int d = 0; d < 3; ++d) {
      const double v = pts4xN[4 * (size_t)i + d];
      mn[d] = std::min(mn[d], v); mx[d] = std::max(mx[d], v);
    }
  double ext = 0.0;
  for (int d = 0; d < 3; ++d) ext = std::max(ext, mx[d] - mn[d]);
  if (voxel_size * 2147483647.0 < ext) return -2;
  const double mb[3] = {mn[0] - voxel_size * 0.5, mn[1] - voxel_size * 0.5, mn[2] - voxel_size * 0.5};
  struct Acc { double p[3] = {0, 0, 0}, n[3] = {0, 0, 0}; int cnt = 0; int first = 0; };
  std::map<std::tuple<int, int, int>, Acc> vox;
  for (int i = 0; i < N; ++i) {
    int ix[3];
    for (int d = 0; d < 3; ++d) ix[d] = (int)std::floor(((double)pts4xN[4 * (size_t)i + d] - mb[d]) / voxel_size);
    auto key = std::make_tuple(ix[0], ix[1], ix[2]);
    auto it = vox.find(key);
    if (it == vox.end()) { it = vox.emplace(key, Acc()).first; it->second.first = i; }
    Acc& a = it->second;
    for (int d = 0; d < 3; ++d) a.p[d] += (double)pts4xN[4 * (size_t)i + d];
    if (nrm3xN)
      for (int d = 0; d < 3; ++d) a.n[d] += (double)nrm3xN[3 * (size_t)i + d];
    a.cnt += 1;
  }
  std::vector<const Acc*> order;
  order.reserve(vox.size());
  for (auto& kv : vox) order.push_back(&kv.second);
  std::sort(order.begin(), order.end(), [](const Acc* a, const Acc* b) { return a->first < b->first; });
  int k = 0;
  for (const Acc* a : order) {
    for (int d = 0; d < 3; ++d) out3xK[3 * (size_t)k + d] = a->p[d] / (double)a->cnt;
    if (nrm3xN && outn3xK)
      for (int d = 0; d < 3; ++d) outn3xK[3 * (size_t)k + d] = a->n[d] / (double)a->cnt;
    if (first_idx) first_idx[k] = a->first;
    ++k;
  }
  return k;
}

// ---- map maintenance: Submap::insertScan (open3d_slam/src/Submap.cpp:39-96) ----------------------
//   transformedCloud = o3d_slam::transform(mapToRangeSensor.matrix(), preProcessedScan)   (:52; Open3D
//     PointCloud::Transform: p' = (T [p; 1]).head<3>() / w in double, n' = T.block<3,3>(0,0) n)
//   mapCloud_ += *transformedCloud; mapBuilderCropper_->setPose(mapToRangeSensor)           (:85-86)
//   voxelizeInsideCroppingVolume -> voxelizeWithinCroppingVolume(mapVoxelSize, cropper, map) (:162, helpers.cpp:117-200):
//     points OUTSIDE the cropper are copied through in input order; points inside are accumulated per voxel,
//     voxel index = floor(p * (1 / voxelSize)) per axis (VoxelHashMap.hpp:48-51), averaged (AccumulatedPoint: sums in
//     double in input order), normals averaged and then normalised; the voxels follow the outside points in the
//     iteration order of an std::unordered_map (unspecified) — here: ascending index of the first member.
// The map is doubles (Open3D's points_ / normals_); the scan arrives as the float images of the C ABI.
// out arrays have capacity M + N points. Returns the new map size; *n_outside = points copied through.
int orc_map_insert_scan(const double* map3xM, const double* mapn3xM, int M, const float* scan4xN, const float* scann3xN,
                        int N, const double* T16, double voxel_size, int crop_kind, double radius_min, double radius_max,
                        double min_z, double max_z, const double* center3, int invert, double* out3xK, double* outn3xK,
                        int* n_outside) {
  const int tot = M + N;
  const bool hasN = mapn3xM != nullptr && (scann3xN != nullptr || N == 0);
  std::vector<double> p(3 * (size_t)tot), nn(hasN ? 3 * (size_t)tot : 0);
  std::memcpy(p.data(), map3xM, 3 * (size_t)M * sizeof(double));
  if (hasN) std::memcpy(nn.data(), mapn3xM, 3 * (size_t)M * sizeof(double));
  for (int i = 0; i < N; ++i) {
    const double x = scan4xN[4 * (size_t)i], y = scan4xN[4 * (size_t)i + 1], z = scan4xN[4 * (size_t)i + 2];
    double q[4];
    for (int r = 0; r < 4; ++r) q[r] = T16[r] * x + T16[r + 4] * y + T16[r + 8] * z + T16[r + 12] * 1.0;
    for (int r = 0; r < 3; ++r) p[3 * (size_t)(M + i) + r] = q[r] / q[3];
    if (hasN) {
      const double a = scann3xN[3 * (size_t)i], b = scann3xN[3 * (size_t)i + 1], c = scann3xN[3 * (size_t)i + 2];
      for (int r = 0; r < 3; ++r) nn[3 * (size_t)(M + i) + r] = T16[r] * a + T16[r + 4] * b + T16[r + 8] * c;
    }
  }
  if (!(voxel_size > 0.0)) {   // "*output = cloud" (helpers.cpp:121-124)
    std::memcpy(out3xK, p.data(), p.size() * sizeof(double));
    if (hasN && outn3xK) std::memcpy(outn3xK, nn.data(), nn.size() * sizeof(double));
    if (n_outside) *n_outside = tot;
    return tot;
  }
  const double inv = 1.0 / voxel_size;
  struct Acc { double p[3] = {0, 0, 0}, n[3] = {0, 0, 0}; int cnt = 0; int first = 0; };
  std::map<std::tuple<int, int, int>, Acc> vox;
  int k = 0;
  for (int i = 0; i < tot; ++i) {
    const double px = p[3 * (size_t)i], py = p[3 * (size_t)i + 1], pz = p[3 * (size_t)i + 2];
    const double dx = px - center3[0], dy = py - center3[1], dz = pz - center3[2];
    bool in = true;
    switch (crop_kind) {
      case 1: in = pz >= min_z && pz <= max_z && std::sqrt(dx * dx + dy * dy) <= radius_max; break;
      case 2: in = std::sqrt(dx * dx + dy * dy + dz * dz) >= radius_min; break;
      case 3: in = std::sqrt(dx * dx + dy * dy + dz * dz) <= radius_max; break;
      case 4: { const double d = std::sqrt(dx * dx + dy * dy + dz * dz); in = d <= radius_max && d >= radius_min; } break;
      default: in = true;
    }
    if (invert) in = !in;
    if (!in) {
      for (int d = 0; d < 3; ++d) out3xK[3 * (size_t)k + d] = p[3 * (size_t)i + d];
      if (hasN && outn3xK)
        for (int d = 0; d < 3; ++d) outn3xK[3 * (size_t)k + d] = nn[3 * (size_t)i + d];
      ++k;
      continue;
    }
    auto key = std::make_tuple((int)std::floor(px * inv), (int)std::floor(py * inv), (int)std::floor(pz * inv));
    auto it = vox.find(key);
    if (it == vox.end()) { it = vox.emplace(key, Acc()).first; it->second.first = i; }
    Acc& a = it->second;
    for (int d = 0; d < 3; ++d) a.p[d] += p[3 * (size_t)i + d];
    if (hasN)
      for (int d = 0; d < 3; ++d) a.n[d] += nn[3 * (size_t)i + d];
    a.cnt += 1;
  }
  if (n_outside) *n_outside = k;
  std::vector<const Acc*> order;
  order.reserve(vox.size());
  for (auto& kv : vox) order.push_back(&kv.second);
  std::sort(order.begin(), order.end(), [](const Acc* a, const Acc* b) { return a->first < b->first; });
  for (const Acc* a : order) {
    for (int d = 0; d < 3; ++d) out3xK[3 * (size_t)k + d] = a->p[d] / (double)a->cnt;
    if (hasN && outn3xK) {
      double v[3] = {a->n[0] / (double)a->cnt, a->n[1] / (double)a->cnt, a->n[2] / (double)a->cnt};
      const double z2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
      if (z2 > 0.0) { const double s = std::sqrt(z2); v[0] /= s; v[1] /= s; v[2] /= s; }   // Eigen normalized()
      for (int d = 0; d < 3; ++d) outn3xK[3 * (size_t)k + d] = v[d];
    }
    ++k;
  }
  return k;
}

// CroppingVolume::crop on a double cloud (the map; croppers.cpp:66-95): indices kept, input order
int orc_crop_d(int kind, double radius_min, double radius_max, double min_z, double max_z, const double* center3,
               int invert, const double* pts3xN, int N, int32_t* kept_idx) {
  int n = 0;
  for (int i = 0; i < N; ++i) {
    const double px = pts3xN[3 * (size_t)i], py = pts3xN[3 * (size_t)i + 1], pz = pts3xN[3 * (size_t)i + 2];
    const double dx = px - center3[0], dy = py - center3[1], dz = pz - center3[2];
    bool in = true;
    switch (kind) {
      case 1: in = pz >= min_z && pz <= max_z && std::sqrt(dx * dx + dy * dy) <= radius_max; break;
      case 2: in = std::sqrt(dx * dx + dy * dy + dz * dz) >= radius_min; break;
      case 3: in = std::sqrt(dx * dx + dy * dy + dz * dz) <= radius_max; break;
      case 4: { const double d = std::sqrt(dx * dx + dy * dy + dz * dz); in = d <= radius_max && d >= radius_min; } break;
      default: in = true;
    }
    if (invert) in = !in;
    if (in) kept_idx[n++] = i;
  }
  return n;
}

}  // extern "C"
