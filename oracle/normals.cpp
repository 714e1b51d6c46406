// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
//
// SurfaceNormalDataPointsFilter::inPlaceFilter — libpointmatcher/pointmatcher/DataPointsFilters/SurfaceNormal.cpp:82-290
// helpers                                       — libpointmatcher/pointmatcher/DataPointsFilters/utils/utils.h:84-150
// SamplingSurfaceNormalDataPointsFilter         — .../DataPointsFilters/SamplingSurfaceNormal.cpp:76-344
//
// The reference runs the *non-symmetric* Eigen::EigenSolver on the symmetric 3x3 scatter matrix; eigenvector
// sign and the order of (near-)equal eigenvalues are implementation artefacts. The oracle uses a symmetric Jacobi
// solve in double on the float matrix; comparisons are modulo normal sign ("parity unpinned").
#include "oracle.h"
#include "la_small.hpp"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

namespace {

const float kInf = std::numeric_limits<float>::infinity();

// neighbourhood statistics shared by both filters. d: 3 x n (column-major), n >= 1
struct Patch {
  float mean[3];
  float C[9];
  float eigVa[3];
  float eigVe[9];  // columns
  bool degenerate;
  float maxNNSq;   // max squared norm of a centred neighbour (computeDensity)
};

void patchStats(const float* d, int n, Patch& p) {
  for (int r = 0; r < 3; ++r) {
    float s = 0.f;
    for (int j = 0; j < n; ++j) s += d[r + 3 * j];
    p.mean[r] = s / (float)n;     // d.rowwise().sum() / T(realKnn), SurfaceNormal.cpp:181
  }
  for (int i = 0; i < 9; ++i) p.C[i] = 0.f;
  p.maxNNSq = 0.f;
  for (int j = 0; j < n; ++j) {
    float nn[3];
    for (int r = 0; r < 3; ++r) nn[r] = d[r + 3 * j] - p.mean[r];   // :182
    float sq = nn[0] * nn[0];
    sq = sq + nn[1] * nn[1];
    sq = sq + nn[2] * nn[2];
    p.maxNNSq = std::max(p.maxNNSq, sq);
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r) p.C[r + 3 * c] += nn[r] * nn[c];   // C = NN * NN^T, :184
  }
  orc::FullPivQR<float, 3> qr;
  qr.compute(p.C);
  p.degenerate = !(qr.rank() + 1 >= 3);                              // :191
  for (int i = 0; i < 3; ++i) p.eigVa[i] = 0.f;
  for (int i = 0; i < 9; ++i) p.eigVe[i] = 0.f;
  if (!p.degenerate) {
    double Cd[9], w[3], V[9];
    for (int i = 0; i < 9; ++i) Cd[i] = p.C[i];
    orc::symEigen<double, 3>(Cd, w, V);
    for (int i = 0; i < 3; ++i) p.eigVa[i] = (float)w[i];
    for (int i = 0; i < 9; ++i) p.eigVe[i] = (float)V[i];
  }
}

int smallestEig(const float* va) {  // computeNormal, utils.h:131-150
  int id = 0;
  float best = std::numeric_limits<float>::max();
  for (int j = 0; j < 3; ++j)
    if (va[j] < best) { id = j; best = va[j]; }
  return id;
}

float density(float maxNNSq, int n) {  // computeDensity, utils.h:105-129
  const float fourThirds = (float)(4. / 3.);
  const float pi = (float)M_PI;
  const float volume = fourThirds * pi * (maxNNSq * std::sqrt(maxNNSq));
  return volume > 0.f ? (float)n / volume : 0.f;
}

}  // namespace

extern "C" int orc_surface_normals(const float* feat, int N, int knn, float maxDist, int sortEigen,
                                   int smoothNormals, float* normals, float* densities, float* eigValues,
                                   float* eigVectors, float* matchedIds, float* meanDists, int32_t* ids_out,
                                   float* dists_out) {
  if (knn < 3) return ORC_E_BAD_PARAM;  // parameter minimum, SurfaceNormal.h:70
  std::vector<int32_t> ids((size_t)knn * N);
  std::vector<float> dists((size_t)knn * N);
  void* tree = orc_kdtree_build(feat, N);                                  // matcher.init(cloud), :159
  orc_kdtree_knn(tree, feat, N, knn, maxDist, 1, ids.data(), dists.data()); // :162
  orc_kdtree_free(tree);
  std::vector<float> nrm;
  if (smoothNormals && !normals) { nrm.resize(3 * (size_t)N); normals = nrm.data(); }
  std::vector<float> d(3 * (size_t)knn);
  for (int i = 0; i < N; ++i) {
    int realKnn = 0;
    for (int j = 0; j < knn; ++j)
      if (dists[(size_t)i * knn + j] != kInf) {
        const int r = ids[(size_t)i * knn + j];
        for (int c = 0; c < 3; ++c) d[c + 3 * realKnn] = feat[4 * (size_t)r + c];
        ++realKnn;
      }
    Patch p;
    patchStats(d.data(), realKnn, p);
    float va[3], ve[9];
    std::memcpy(va, p.eigVa, sizeof(va));
    std::memcpy(ve, p.eigVe, sizeof(ve));
    if (sortEigen && !p.degenerate) {   // sortIndexes ascending, :199-210
      int idx[3] = {0, 1, 2};
      std::stable_sort(idx, idx + 3, [&](int a, int b) { return p.eigVa[a] < p.eigVa[b]; });
      for (int c = 0; c < 3; ++c) {
        va[c] = p.eigVa[idx[c]];
        for (int r = 0; r < 3; ++r) ve[r + 3 * c] = p.eigVe[r + 3 * idx[c]];
      }
    }
    if (normals) {
      const int col = sortEigen ? 0 : smallestEig(va);
      for (int r = 0; r < 3; ++r) {
        float v = ve[r + 3 * col];
        v = std::min(1.f, std::max(-1.f, v));   // cwiseMax(-1).cwiseMin(1), :227
        normals[3 * (size_t)i + r] = v;
      }
    }
    if (densities) densities[i] = p.degenerate ? 0.f : density(p.maxNNSq, realKnn);
    if (eigValues) std::memcpy(eigValues + 3 * (size_t)i, va, sizeof(va));
    if (eigVectors)  // row-major serialisation
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) eigVectors[9 * (size_t)i + 3 * r + c] = ve[r + 3 * c];
    if (meanDists) {
      if (p.degenerate) meanDists[i] = (float)std::numeric_limits<std::size_t>::max();
      else {
        const float dx = feat[4 * (size_t)i] - p.mean[0], dy = feat[4 * (size_t)i + 1] - p.mean[1],
                    dz = feat[4 * (size_t)i + 2] - p.mean[2];
        meanDists[i] = std::sqrt(dx * dx + dy * dy + dz * dz);
      }
    }
  }
  if (matchedIds)
    for (size_t e = 0; e < (size_t)knn * N; ++e) matchedIds[e] = (float)ids[e];   // :260-263
  if (smoothNormals) {   // :265-288
    // the reference smooths in place, reading already-smoothed normals of lower indices
    for (int i = 0; i < N; ++i) {
      const float cur[3] = {normals[3 * (size_t)i], normals[3 * (size_t)i + 1], normals[3 * (size_t)i + 2]};
      float mean[3] = {0, 0, 0};
      int n = 0;
      for (int j = 0; j < knn; ++j)
        if (dists[(size_t)i * knn + j] != kInf) {
          const float* nb = normals + 3 * (size_t)ids[(size_t)i * knn + j];
          const float dot = cur[0] * nb[0] + cur[1] * nb[1] + cur[2] * nb[2];
          for (int r = 0; r < 3; ++r) mean[r] = dot > 0.f ? mean[r] + nb[r] : mean[r] - nb[r];
          ++n;
        }
      for (int r = 0; r < 3; ++r) normals[3 * (size_t)i + r] = mean[r] / (float)n;
    }
  }
  if (ids_out) std::memcpy(ids_out, ids.data(), ids.size() * sizeof(int32_t));
  if (dists_out) std::memcpy(dists_out, dists.data(), dists.size() * sizeof(float));
  return ORC_OK;
}

// ---------------------------------------------------------------------------------------------------------------
namespace {
struct Build {
  std::vector<float> feat;      // 4 x N working copy (samplingMethod 1 overwrites kept points with the box mean)
  std::vector<float> normals;   // 3 x N
  std::vector<int> indices;
  std::vector<int> keep;
  int knn;
  float ratio;
  int samplingMethod;
  float maxBoxDim;
};

void fuseRange(Build& b, int first, int last) {  // SamplingSurfaceNormal.cpp:234-344
  const int n = last - first;
  std::vector<float> d(3 * (size_t)n);
  float mn[3] = {kInf, kInf, kInf}, mx[3] = {-kInf, -kInf, -kInf};
  for (int i = 0; i < n; ++i)
    for (int r = 0; r < 3; ++r) {
      const float v = b.feat[4 * (size_t)b.indices[first + i] + r];
      d[r + 3 * i] = v;
      mn[r] = std::min(mn[r], v); mx[r] = std::max(mx[r], v);
    }
  const float boxDim = std::max(mx[0] - mn[0], std::max(mx[1] - mn[1], mx[2] - mn[2]));
  if (boxDim > b.maxBoxDim) return;   // :248-253
  Patch p;
  patchStats(d.data(), n, p);
  if (p.degenerate) return;           // :272-276
  const int col = smallestEig(p.eigVa);
  const float normal[3] = {p.eigVe[3 * col], p.eigVe[3 * col + 1], p.eigVe[3 * col + 2]};
  if (b.samplingMethod == 0) {        // :292-313
    for (int i = 0; i < n; ++i) {
      const float r = (float)std::rand() / (float)RAND_MAX;
      if (r < b.ratio) {
        const int kk = b.indices[first + i];
        b.keep.push_back(kk);
        std::memcpy(&b.normals[3 * (size_t)kk], normal, sizeof(normal));
      }
    }
  } else {                            // :314-343
    const int kk = b.indices[first];
    b.keep.push_back(kk);
    for (int r = 0; r < 3; ++r) b.feat[4 * (size_t)kk + r] = p.mean[r];
    b.feat[4 * (size_t)kk + 3] = 1.f;
    std::memcpy(&b.normals[3 * (size_t)kk], normal, sizeof(normal));
  }
}

void buildNew(Build& b, int first, int last, const float* minV, const float* maxV) {  // :184-232
  const int count = last - first;
  if (count <= b.knn) { fuseRange(b, first, last); return; }
  int cutDim = 0;  // argMax over all 4 rows of (max - min); the homogeneous row is 0
  for (int r = 1; r < 4; ++r)
    if (maxV[r] - minV[r] > maxV[cutDim] - minV[cutDim]) cutDim = r;
  const int rightCount = count / 2, leftCount = count - rightCount;
  std::nth_element(b.indices.begin() + first, b.indices.begin() + first + leftCount, b.indices.begin() + last,
                   [&](int a, int c) { return b.feat[4 * (size_t)a + cutDim] < b.feat[4 * (size_t)c + cutDim]; });
  const int cutIndex = b.indices[first + leftCount];
  const float cutVal = b.feat[4 * (size_t)cutIndex + cutDim];
  float leftMax[4], rightMin[4];
  std::memcpy(leftMax, maxV, sizeof(leftMax));
  std::memcpy(rightMin, minV, sizeof(rightMin));
  leftMax[cutDim] = cutVal;
  rightMin[cutDim] = cutVal;
  buildNew(b, first, first + leftCount, minV, leftMax);
  buildNew(b, first + leftCount, last, rightMin, maxV);
}
}  // namespace

extern "C" int orc_sampling_surface_normal(const float* feat, int N, int knn, float ratio, int samplingMethod,
                                           float maxBoxDim, float* featOut, float* normalsOut) {
  Build b;
  b.feat.assign(feat, feat + 4 * (size_t)N);
  b.normals.assign(3 * (size_t)N, 0.f);
  b.indices.resize(N);
  for (int i = 0; i < N; ++i) b.indices[i] = i;
  b.knn = knn; b.ratio = ratio; b.samplingMethod = samplingMethod; b.maxBoxDim = maxBoxDim;
  float mn[4] = {kInf, kInf, kInf, kInf}, mx[4] = {-kInf, -kInf, -kInf, -kInf};
  for (int i = 0; i < N; ++i)
    for (int r = 0; r < 4; ++r) {
      mn[r] = std::min(mn[r], feat[4 * (size_t)i + r]);
      mx[r] = std::max(mx[r], feat[4 * (size_t)i + r]);
    }
  if (N > 0) buildNew(b, 0, N, mn, mx);
  std::sort(b.keep.begin(), b.keep.end());   // :160
  const int out = (int)b.keep.size();
  for (int i = 0; i < out; ++i) {
    std::memcpy(featOut + 4 * (size_t)i, &b.feat[4 * (size_t)b.keep[i]], 16);
    std::memcpy(normalsOut + 3 * (size_t)i, &b.normals[3 * (size_t)b.keep[i]], 12);
  }
  return out;
}
