// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
//
// Loop-closure front half of open3d_slam:
//   Submap::computeFeatures                     open3d_slam/src/Submap.cpp:255-275
//       registration::ComputeFPFHFeature(sparseMapCloud_, KDTreeSearchParamHybrid(featureRadius_, featureKnn_))
//   PlaceRecognition::buildLoopClosureConstraints   open3d_slam/src/PlaceRecognition.cpp:56-59, 81-85
//       RegistrationRANSACBasedOnFeatureMatching(sourceSparse, targetSparse, sourceFeature, targetFeature,
//           mutual_filter = true, ransacMaxCorrespondenceDistance_, TransformationEstimationPointToPoint(false),
//           ransacModelSize_, {CorrespondenceCheckerBasedOnDistance, CorrespondenceCheckerBasedOnEdgeLength},
//           RANSACConvergenceCriteria(ransacNumIter_, ransacProbability_))
// Open3D v0.15.1 is fetched by the reference at build time and is NOT in /root/reference: both functions are RESTATED
// AS RECALLED FROM UPSTREAM (pipelines/registration/Feature.cpp, Registration.cpp, CorrespondenceChecker.cpp) —
// "parity unpinned" like o3d.cpp; the reference holds no test or golden for them.
//
// Stated deviations (shared with the device path, which is compared against this file):
//   * clouds are the float-converted ones (boundary B's convention, o3d.cpp); the FPFH neighbour search runs on them
//     with the k-NN contract of knn.cpp (float squared L2, ties -> smallest index; Open3D searches doubles with FLANN),
//     everything after the search is double;
//   * the feature-space 1-NN is exact with squared L2 summed over the 33 bins in index order, ties -> smallest index
//     (FLANN's exact kd-tree search returns the same neighbour up to ties and rounding of its 4-wide accumulation);
//   * Open3D draws the RANSAC samples from a std::mt19937 seeded by std::random_device inside an OpenMP loop: it is
//     not reproducible by design. Here sample j of iteration itr is a counter-based hash of (seed, itr, j)
//     (splitmix64 finaliser, multiply-shift onto the correspondence count), and the iterations are consumed in
//     index order with the single-thread semantics of the upstream loop (one best result, one exit estimate).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <thread>
#include <tuple>
#include <vector>
#include "oracle.h"

namespace {

// ComputePairFeatures (Feature.cpp): the Darboux-frame angles of a point pair; (f0, f1, f2, distance)
void pair_features(const double* p1, const double* n1, const double* p2, const double* n2, double* out) {
  double dp[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  out[0] = out[1] = out[2] = 0.0;
  out[3] = std::sqrt(dp[0] * dp[0] + dp[1] * dp[1] + dp[2] * dp[2]);
  if (out[3] == 0.0) { out[3] = 0.0; return; }
  const double* a = n1;
  const double* b = n2;
  const double angle1 = (n1[0] * dp[0] + n1[1] * dp[1] + n1[2] * dp[2]) / out[3];
  const double angle2 = (n2[0] * dp[0] + n2[1] * dp[1] + n2[2] * dp[2]) / out[3];
  double f2;
  if (std::acos(std::fabs(angle1)) > std::acos(std::fabs(angle2))) {
    a = n2; b = n1;
    dp[0] = -dp[0]; dp[1] = -dp[1]; dp[2] = -dp[2];
    f2 = -angle2;
  } else {
    f2 = angle1;
  }
  double v[3] = {dp[1] * a[2] - dp[2] * a[1], dp[2] * a[0] - dp[0] * a[2], dp[0] * a[1] - dp[1] * a[0]};
  const double vn = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (vn == 0.0) { out[3] = 0.0; return; }   // Vector4d::Zero()
  v[0] /= vn; v[1] /= vn; v[2] /= vn;
  const double w[3] = {a[1] * v[2] - a[2] * v[1], a[2] * v[0] - a[0] * v[2], a[0] * v[1] - a[1] * v[0]};
  out[2] = f2;
  out[1] = v[0] * b[0] + v[1] * b[1] + v[2] * b[2];
  out[0] = std::atan2(w[0] * b[0] + w[1] * b[1] + w[2] * b[2], a[0] * b[0] + a[1] * b[1] + a[2] * b[2]);
}

inline int bin11(double x) {
  int h = (int)std::floor(x);
  return h < 0 ? 0 : (h >= 11 ? 10 : h);
}

inline uint64_t splitmix(uint64_t x) {
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
  x ^= x >> 27; x *= 0x94d049bb133111ebull;
  x ^= x >> 31;
  return x;
}

// Eigen::umeyama(src, dst, with_scaling = false) over n point pairs (TransformationEstimationPointToPoint):
// sigma = 1/n sum (d - md)(s - ms)^T = U D V^T, R = U diag(1, 1, sign(det U det V)) V^T, t = md - R ms.
// One-sided Jacobi on the columns of sigma; the left vectors are rebuilt as u1, u2, u1 x u2 (three point pairs give a
// rank-2 sigma; Eigen's full U holds +-(u1 x u2) and the sign rule makes R the same proper rotation either way).
void umeyama_rigid(const double* s3xn, const double* d3xn, int n, double* T16) {
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
        if (ga == 0.0 || std::fabs(ga) <= 1e-16 * std::sqrt(al * be)) continue;
        rotated = true;
        const double zeta = (be - al) / (2.0 * ga);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / std::sqrt(1.0 + t * t), sn = c * t;
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
  for (int j = 0; j < 3; ++j) sg[j] = std::sqrt(A[3 * j] * A[3 * j] + A[1 + 3 * j] * A[1 + 3 * j] + A[2 + 3 * j] * A[2 + 3 * j]);
  for (int a = 0; a < 2; ++a)
    for (int b = a + 1; b < 3; ++b)
      if (sg[o[b]] > sg[o[a]]) std::swap(o[a], o[b]);
  double u1[3], u2[3], u3[3], v1[3], v2[3], v3[3];
  for (int k = 0; k < 3; ++k) {
    u1[k] = sg[o[0]] > 0 ? A[k + 3 * o[0]] / sg[o[0]] : (k == 0 ? 1.0 : 0.0);
    v1[k] = V[k + 3 * o[0]]; v2[k] = V[k + 3 * o[1]]; v3[k] = V[k + 3 * o[2]];
  }
  if (sg[o[1]] > 1e-300) {
    for (int k = 0; k < 3; ++k) u2[k] = A[k + 3 * o[1]] / sg[o[1]];
  } else {   // rank <= 1 (collinear or coincident samples): any unit vector orthogonal to u1
    const int m = std::fabs(u1[0]) <= std::fabs(u1[1]) ? (std::fabs(u1[0]) <= std::fabs(u1[2]) ? 0 : 2)
                                                       : (std::fabs(u1[1]) <= std::fabs(u1[2]) ? 1 : 2);
    double e[3] = {0, 0, 0};
    e[m] = 1.0;
    const double d = u1[m];
    double n2 = 0;
    for (int k = 0; k < 3; ++k) { u2[k] = e[k] - d * u1[k]; n2 += u2[k] * u2[k]; }
    for (int k = 0; k < 3; ++k) u2[k] /= std::sqrt(n2);
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

inline void xform(const double* T, const float* p4, double* out) {
  const double x = p4[0], y = p4[1], z = p4[2];
  for (int r = 0; r < 3; ++r) out[r] = T[r] * x + T[r + 4] * y + T[r + 8] * z + T[r + 12];
}

template <class F>
void parallel_for(int n, int threads, F f) {
  threads = std::max(1, std::min(threads, n));
  if (threads == 1) { f(0, n); return; }
  std::vector<std::thread> th;
  for (int t = 0; t < threads; ++t) {
    const int a = (int)((int64_t)n * t / threads), b = (int)((int64_t)n * (t + 1) / threads);
    th.emplace_back([=] { f(a, b); });
  }
  for (auto& t : th) t.join();
}

}  // namespace

// ComputeFPFHFeature (Feature.cpp): SPFH of every point over its hybrid neighbourhood (within radius, at most max_nn,
// nearest first, the first entry — the point itself — skipped), then the distance-weighted sum of the neighbours'
// SPFHs, each 11-bin group normalised to 100, plus the point's own SPFH. nn_count (nullable): neighbourhood sizes.
extern "C" int orc_o3d_fpfh(const float* pts4xN, const double* nrm3xN, int N, float radius, int max_nn, int threads,
                            double* fpfh33xN, int32_t* nn_count) {
  if (N < 0 || max_nn < 1 || !(radius > 0.f) || (N > 0 && (!pts4xN || !nrm3xN || !fpfh33xN))) return ORC_E_BAD_PARAM;
  if (N == 0) return ORC_OK;
  std::vector<int32_t> ids((size_t)max_nn * N);
  std::vector<float> dists((size_t)max_nn * N);
  void* tree = orc_kdtree_build(pts4xN, N);
  orc_kdtree_knn(tree, pts4xN, N, max_nn, radius, threads > 0 ? threads : 1, ids.data(), dists.data());
  orc_kdtree_free(tree);
  std::vector<double> spfh((size_t)33 * N, 0.0);
  std::vector<int> cnt((size_t)N);
  auto P = [&](int i, double* p) { p[0] = pts4xN[4 * (size_t)i]; p[1] = pts4xN[4 * (size_t)i + 1]; p[2] = pts4xN[4 * (size_t)i + 2]; };
  parallel_for(N, threads, [&](int a, int b) {
    for (int i = a; i < b; ++i) {
      int n = 0;
      while (n < max_nn && ids[(size_t)i * max_nn + n] >= 0) ++n;
      cnt[i] = n;
      if (nn_count) nn_count[i] = n;
      if (n <= 1) continue;
      const double incr = 100.0 / (double)(n - 1);
      double p1[3], p2[3], pf[4];
      P(i, p1);
      double* h = &spfh[(size_t)33 * i];
      for (int k = 1; k < n; ++k) {
        const int j = ids[(size_t)i * max_nn + k];
        P(j, p2);
        pair_features(p1, nrm3xN + 3 * (size_t)i, p2, nrm3xN + 3 * (size_t)j, pf);
        if (pf[3] == 0.0) pf[0] = pf[1] = pf[2] = 0.0;
        h[bin11(11.0 * (pf[0] + M_PI) / (2.0 * M_PI))] += incr;
        h[11 + bin11(11.0 * (pf[1] + 1.0) * 0.5)] += incr;
        h[22 + bin11(11.0 * (pf[2] + 1.0) * 0.5)] += incr;
      }
    }
  });
  parallel_for(N, threads, [&](int a, int b) {
    for (int i = a; i < b; ++i) {
      double* f = fpfh33xN + (size_t)33 * i;
      for (int j = 0; j < 33; ++j) f[j] = 0.0;
      const int n = cnt[i];
      if (n <= 1) continue;
      double p1[3], p2[3];
      P(i, p1);
      double sum[3] = {0.0, 0.0, 0.0};
      for (int k = 1; k < n; ++k) {
        const int id = ids[(size_t)i * max_nn + k];
        P(id, p2);
        const double dx = p2[0] - p1[0], dy = p2[1] - p1[1], dz = p2[2] - p1[2];
        const double dist = dx * dx + dy * dy + dz * dz;   // distance2[k]
        if (dist == 0.0) continue;
        for (int j = 0; j < 33; ++j) {
          const double val = spfh[(size_t)33 * id + j] / dist;
          sum[j / 11] += val;
          f[j] += val;
        }
      }
      for (int g = 0; g < 3; ++g)
        if (sum[g] != 0.0) sum[g] = 100.0 / sum[g];
      for (int j = 0; j < 33; ++j) {
        f[j] *= sum[j / 11];
        f[j] += spfh[(size_t)33 * i + j];
      }
    }
  });
  return ORC_OK;
}

// exact 1-NN of every column of a in b (33-bin histograms): squared L2 in double summed in bin order, ties -> smallest
// index (stands in for KDTreeFlann(feature).SearchKNN(col, 1, ...))
extern "C" int orc_o3d_feature_nn(const double* a33xN, int N, const double* b33xM, int M, int threads, int32_t* nn_N,
                                  double* d2_N) {
  if (N < 0 || M <= 0) return ORC_E_BAD_PARAM;
  parallel_for(N, threads, [&](int lo, int hi) {
    for (int i = lo; i < hi; ++i) {
      const double* a = a33xN + (size_t)33 * i;
      double best = INFINITY;
      int bj = -1;
      for (int j = 0; j < M; ++j) {
        const double* b = b33xM + (size_t)33 * j;
        double d = 0.0;
        for (int t = 0; t < 33; ++t) { const double e = a[t] - b[t]; d += e * e; }
        if (d < best) { best = d; bj = j; }
      }
      nn_N[i] = bj;
      if (d2_N) d2_N[i] = best;
    }
  });
  return ORC_OK;
}

// RegistrationRANSACBasedOnFeatureMatching (Registration.cpp), see the header of this file. corr2xK (capacity 2 * Ns):
// the inlier correspondences (source, target) of the best hypothesis in correspondence order.
extern "C" int orc_o3d_ransac_feature_matching(const float* src4xN, int Ns, const float* tgt4xM, int Nt,
                                               const double* sf33xN, const double* tf33xM, const orc_ransac_params* prm,
                                               int threads, orc_ransac_result* res, int32_t* corr2xK) {
  if (!prm || !res || Ns <= 0 || Nt <= 0) return ORC_E_BAD_PARAM;
  std::memset(res, 0, sizeof(*res));
  for (int d = 0; d < 4; ++d) res->T[5 * d] = 1.0;   // RegistrationResult(): identity, fitness 0, rmse 0
  const int rn = prm->ransac_n;
  if (rn < 3 || rn > 8 || !(prm->max_correspondence_distance > 0.0)) return ORC_OK;
  std::vector<int32_t> ij((size_t)Ns), ji;
  orc_o3d_feature_nn(sf33xN, Ns, tf33xM, Nt, threads, ij.data(), nullptr);
  std::vector<int32_t> cs, ct;
  if (prm->mutual_filter) {
    ji.resize((size_t)Nt);
    orc_o3d_feature_nn(tf33xM, Nt, sf33xN, Ns, threads, ji.data(), nullptr);
    for (int i = 0; i < Ns; ++i)
      if (ji[ij[i]] == i) { cs.push_back(i); ct.push_back(ij[i]); }
    if ((int)cs.size() >= rn * 3) res->mutual_used = 1;
    else { cs.clear(); ct.clear(); }
  }
  if (!res->mutual_used) {
    for (int i = 0; i < Ns; ++i) { cs.push_back(i); ct.push_back(ij[i]); }
  }
  const int nc = (int)cs.size();
  res->n_feature_correspondences = nc;
  if (nc < rn) return ORC_OK;
  const double max2 = prm->max_correspondence_distance * prm->max_correspondence_distance;
  int best_good = 0;          // RegistrationResult(): fitness 0, rmse 0 — IsBetterRANSACThan needs fitness > 0
  double best_rmse = 0.0, best_T[16];
  for (int i = 0; i < 16; ++i) best_T[i] = res->T[i];
  int64_t exit_itr = prm->max_iteration, itr = 0, validations = 0;
  for (; itr < prm->max_iteration && itr < exit_itr; ++itr) {
    int sidx[8];
    double s[24], d[24];
    for (int j = 0; j < rn; ++j) {
      const uint64_t h = splitmix(prm->seed + ((uint64_t)itr * (uint64_t)rn + (uint64_t)j) * 0x9e3779b97f4a7c15ull);
      sidx[j] = (int)(((h >> 32) * (uint64_t)nc) >> 32);
      for (int a = 0; a < 3; ++a) {
        s[3 * j + a] = src4xN[4 * (size_t)cs[sidx[j]] + a];
        d[3 * j + a] = tgt4xM[4 * (size_t)ct[sidx[j]] + a];
      }
    }
    bool ok = true;
    if (prm->edge_length_similarity > 0.0) {   // CorrespondenceCheckerBasedOnEdgeLength::Check
      for (int a = 0; a < rn && ok; ++a)
        for (int b = a + 1; b < rn; ++b) {
          double ds = 0, dt = 0;
          for (int t = 0; t < 3; ++t) {
            ds += (s[3 * a + t] - s[3 * b + t]) * (s[3 * a + t] - s[3 * b + t]);
            dt += (d[3 * a + t] - d[3 * b + t]) * (d[3 * a + t] - d[3 * b + t]);
          }
          ds = std::sqrt(ds); dt = std::sqrt(dt);
          if (ds < dt * prm->edge_length_similarity || dt < ds * prm->edge_length_similarity) { ok = false; break; }
        }
    }
    if (!ok) continue;
    double T[16];
    umeyama_rigid(s, d, rn, T);
    if (prm->distance_threshold > 0.0) {       // CorrespondenceCheckerBasedOnDistance::Check
      for (int j = 0; j < rn; ++j) {
        double q[3];
        xform(T, src4xN + 4 * (size_t)cs[sidx[j]], q);
        double e = 0;
        for (int t = 0; t < 3; ++t) e += (d[3 * j + t] - q[t]) * (d[3 * j + t] - q[t]);
        if (std::sqrt(e) > prm->distance_threshold) { ok = false; break; }
      }
    }
    if (!ok) continue;
    ++validations;
    // EvaluateRANSACBasedOnCorrespondence
    int good = 0;
    double err2 = 0.0;
    for (int c = 0; c < nc; ++c) {
      double q[3];
      xform(T, src4xN + 4 * (size_t)cs[c], q);
      double e = 0;
      for (int t = 0; t < 3; ++t) { const double v = q[t] - (double)tgt4xM[4 * (size_t)ct[c] + t]; e += v * v; }
      if (e < max2) { ++good; err2 += e; }
    }
    const double rmse = good ? std::sqrt(err2 / (double)good) : 0.0;
    if (good > best_good || (good == best_good && good > 0 && rmse < best_rmse)) {   // IsBetterRANSACThan
      best_good = good; best_rmse = rmse;
      std::memcpy(best_T, T, sizeof(T));
      const double ratio = (double)good / (double)nc;
      const double pw = std::pow(ratio, (double)rn);
      if (prm->confidence < 1.0 && pw > 0.0) {
        if (pw >= 1.0) exit_itr = 0;
        else {
          const double k = std::log(1.0 - prm->confidence) / std::log(1.0 - pw);
          if (k >= 0.0 && k < (double)exit_itr) exit_itr = (int64_t)std::ceil(k);
        }
      }
    }
  }
  res->iterations_run = itr;
  res->validations = validations;
  std::memcpy(res->T, best_T, sizeof(best_T));
  res->fitness = (double)best_good / (double)nc;
  res->inlier_rmse = best_rmse;
  int k = 0;
  if (best_good > 0)
    for (int c = 0; c < nc; ++c) {
      double q[3];
      xform(best_T, src4xN + 4 * (size_t)cs[c], q);
      double e = 0;
      for (int t = 0; t < 3; ++t) { const double v = q[t] - (double)tgt4xM[4 * (size_t)ct[c] + t]; e += v * v; }
      if (e < max2) {
        if (corr2xK) { corr2xK[2 * k] = cs[c]; corr2xK[2 * k + 1] = ct[c]; }
        ++k;
      }
    }
  res->n_correspondences = k;
  return ORC_OK;
}

// computeIndicesOfOverlappingPoints (open3d_slam/src/helpers.cpp:320-345, IN the reference tree): VoxelMap with the
// layers "target" and "source" (the source transformed by sourceToTarget first), key = getVoxelIdx(p, inverse voxel
// size) = floor(p * (1 / voxelSize)) per axis (VoxelHashMap.hpp:43-51); the indices of every voxel that holds at least
// minNumPointsPerVoxel points of both layers. The reference appends them in std::unordered_map order; here ascending.
extern "C" int orc_overlap_indices(const float* src4xN, int Ns, const float* tgt4xM, int Nt, const double* T16,
                                   double voxel_size, int min_points, int32_t* src_idx, int* n_src, int32_t* tgt_idx,
                                   int* n_tgt) {
  if (!(voxel_size > 0.0) || min_points < 1) return ORC_E_BAD_PARAM;
  const double inv = 1.0 / voxel_size;
  typedef std::tuple<int, int, int> Key;
  std::map<Key, std::pair<int, int>> vox;
  std::vector<Key> ks((size_t)Ns), kt((size_t)Nt);
  for (int i = 0; i < Ns; ++i) {
    double q[3];
    xform(T16, src4xN + 4 * (size_t)i, q);
    ks[i] = Key((int)std::floor(q[0] * inv), (int)std::floor(q[1] * inv), (int)std::floor(q[2] * inv));
    ++vox[ks[i]].first;
  }
  for (int i = 0; i < Nt; ++i) {
    const float* p = tgt4xM + 4 * (size_t)i;
    kt[i] = Key((int)std::floor((double)p[0] * inv), (int)std::floor((double)p[1] * inv), (int)std::floor((double)p[2] * inv));
    ++vox[kt[i]].second;
  }
  int a = 0, b = 0;
  for (int i = 0; i < Ns; ++i) {
    const auto& c = vox[ks[i]];
    if (c.first >= min_points && c.second >= min_points) src_idx[a++] = i;
  }
  for (int i = 0; i < Nt; ++i) {
    const auto& c = vox[kt[i]];
    if (c.first >= min_points && c.second >= min_points) tgt_idx[b++] = i;
  }
  *n_src = a; *n_tgt = b;
  return ORC_OK;
}
