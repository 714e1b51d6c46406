// ORACLE — TEST INFRASTRUCTURE ONLY.
// Restatement of the Open3D v0.15.1 registration pipeline that open3d_slam's CloudRegistration calls
// (open3d_slam/src/CloudRegistration.cpp:17-23 RegistrationGeneralizedICP, :62-66 / :89-94 RegistrationICP;
// constraint_builders.cpp:65-68). Open3D is NOT in the reference tree (fetched at build time, tag v0.15.1,
// open3d_catkin/CMakeLists.txt:103-104) and not installed here: the formulae below are RECALLED from upstream
// (pipelines/registration/Registration.cpp, TransformationEstimation.cpp, GeneralizedICP.cpp, utility/Eigen.cpp) and
// marked as such in DESIGN.md — parity for this boundary is "unpinned" (the reference holds no test for it either).
//
//   RegistrationICP: T = init; result = match(T src); loop i < max_iteration { update = estimate(corr);
//     T = update T; result' = match(T src); stop when |fitness' - fitness| < rel_fitness and |rmse' - rmse| < rel_rmse }
//   match: 1-NN within max_correspondence_distance; fitness = #corr / #src; inlier_rmse = sqrt(sum d^2 / #corr)
//   point-to-plane: r = (s - t).n_t, J = [s x n_t ; n_t], x = -(JtJ)^-1 Jtr (LDLT), T(x) = Rz(x2) Ry(x1) Rx(x0), t = x3..5
//   point-to-point: Eigen::umeyama without scaling over the correspondences
//   GICP: C = R diag(eps,1,1) R^T from the normal (GetRotationFromE1ToX), M = C_t + C_s, W = M^-1/2, 3 residual rows
//     r = W (s - t), J = W [-[s]x  I]; source covariances rotate with the cloud (PointCloud::Transform)
//
// One deliberate difference, stated: clouds are the float-converted ones the B200 boundary receives
// (open3dToPointmatcher, open3d_conversions.cpp:57-118, converts double -> float the same way), correspondences are
// searched in float with the k-NN contract of oracle/knn.cpp; everything after the match is double like Open3D
// (normals are re-normalised in double first, see orc_o3d_covariances_from_normals).
#include "oracle.h"
#include "la_small.hpp"

#include <cmath>
#include <cstring>
#include <vector>

namespace {

void mat4_mul(const double* A, const double* B, double* C) {  // column-major
  double R[16];
  for (int j = 0; j < 4; ++j)
    for (int i = 0; i < 4; ++i) {
      double a = 0;
      for (int k = 0; k < 4; ++k) a += A[i + 4 * k] * B[k + 4 * j];
      R[i + 4 * j] = a;
    }
  std::memcpy(C, R, sizeof(R));
}

// TransformVector6dToMatrix4d (utility/Eigen.cpp): R = Rz(x2) * Ry(x1) * Rx(x0)
void vec6_to_T(const double* x, double* T) {
  const double cx = std::cos(x[0]), sx = std::sin(x[0]), cy = std::cos(x[1]), sy = std::sin(x[1]), cz = std::cos(x[2]),
               sz = std::sin(x[2]);
  const double Rx[9] = {1, 0, 0, 0, cx, sx, 0, -sx, cx};     // column-major
  const double Ry[9] = {cy, 0, -sy, 0, 1, 0, sy, 0, cy};
  const double Rz[9] = {cz, sz, 0, -sz, cz, 0, 0, 0, 1};
  double A[9], R[9];
  auto mul3 = [](const double* P, const double* Q, double* O) {
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < 3; ++i) {
        double a = 0;
        for (int k = 0; k < 3; ++k) a += P[i + 3 * k] * Q[k + 3 * j];
        O[i + 3 * j] = a;
      }
  };
  mul3(Rz, Ry, A);
  mul3(A, Rx, R);
  for (int i = 0; i < 16; ++i) T[i] = 0;
  T[15] = 1;
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) T[i + 4 * j] = R[i + 3 * j];
  T[12] = x[3]; T[13] = x[4]; T[14] = x[5];
}

// A.ldlt().solve(b) for a symmetric 6x6 (no pivoting needed for the positive definite systems met here)
bool ldlt_solve6(const double* A, const double* b, double* x) {
  double L[36] = {0}, D[6];
  for (int j = 0; j < 6; ++j) {
    double d = A[j + 6 * j];
    for (int k = 0; k < j; ++k) d -= L[j + 6 * k] * L[j + 6 * k] * D[k];
    if (!(std::fabs(d) > 0)) return false;
    D[j] = d;
    L[j + 6 * j] = 1;
    for (int i = j + 1; i < 6; ++i) {
      double s = A[i + 6 * j];
      for (int k = 0; k < j; ++k) s -= L[i + 6 * k] * L[j + 6 * k] * D[k];
      L[i + 6 * j] = s / d;
    }
  }
  double y[6];
  for (int i = 0; i < 6; ++i) {
    double s = b[i];
    for (int k = 0; k < i; ++k) s -= L[i + 6 * k] * y[k];
    y[i] = s;
  }
  for (int i = 0; i < 6; ++i) y[i] /= D[i];
  for (int i = 5; i >= 0; --i) {
    double s = y[i];
    for (int k = i + 1; k < 6; ++k) s -= L[k + 6 * i] * x[k];
    x[i] = s;
  }
  return true;
}

// W = M^-1/2 for a symmetric positive definite 3x3 (M.inverse().sqrt() in GeneralizedICP.cpp)
void inv_sqrt3(const double* M, double* W) {
  double w[3], V[9];
  orc::symEigen<double, 3>(M, w, V);
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) {
      double a = 0;
      for (int k = 0; k < 3; ++k) a += V[i + 3 * k] * (1.0 / std::sqrt(w[k])) * V[j + 3 * k];
      W[i + 3 * j] = a;
    }
}

}  // namespace

// GetRotationFromE1ToX + C = R diag(eps,1,1) R^T (GeneralizedICP.cpp, InitializePointCloudForGeneralizedICP)
extern "C" void orc_o3d_covariances_from_normals(const float* n3xN, int N, double eps, double* cov9xN) {
  for (int p = 0; p < N; ++p) {
    double x[3] = {n3xN[3 * (size_t)p], n3xN[3 * (size_t)p + 1], n3xN[3 * (size_t)p + 2]};
    // the reference's normals are unit-length doubles (NormalizeNormals); the float boundary rounds them, so the
    // unit length is restored in double before the covariance is formed
    const double nn = std::sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
    if (nn > 0) { x[0] /= nn; x[1] /= nn; x[2] /= nn; }
    double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    const double c = x[0];                       // e1 . x
    if (!(c < -0.99)) {
      const double v[3] = {0.0, -x[2], x[1]};    // e1 x x
      const double sv[9] = {0, v[2], -v[1], -v[2], 0, v[0], v[1], -v[0], 0};  // skew, column-major
      const double f = 1.0 / (1.0 + c);
      for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i) {
          double s2 = 0;
          for (int k = 0; k < 3; ++k) s2 += sv[i + 3 * k] * sv[k + 3 * j];
          R[i + 3 * j] = (i == j ? 1.0 : 0.0) + sv[i + 3 * j] + s2 * f;
        }
    }
    const double Cd[3] = {eps, 1.0, 1.0};
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < 3; ++i) {
        double a = 0;
        for (int k = 0; k < 3; ++k) a += R[i + 3 * k] * Cd[k] * R[j + 3 * k];
        cov9xN[9 * (size_t)p + i + 3 * j] = a;
      }
  }
}

// kind: 0 point-to-plane (tgt normals required), 1 point-to-point, 2 generalized (both covariances required)
extern "C" int orc_o3d_icp(int kind, const float* src4xN, int N, const float* tgt4xM, int M, const float* tgtn3xM,
                           const double* srccov9xN, const double* tgtcov9xM, float max_dist, const double* init16,
                           int max_iteration, double rel_fitness, double rel_rmse, int knn_threads, double* T16,
                           double* fitness, double* rmse, int* iterations, int32_t* corr_ids) {
  if (N <= 0 || M <= 0) return ORC_E_BAD_PARAM;
  if (kind == 0 && !tgtn3xM) return ORC_E_BAD_PARAM;
  if (kind == 2 && (!srccov9xN || !tgtcov9xM)) return ORC_E_BAD_PARAM;
  void* tree = orc_kdtree_build(tgt4xM, M);
  double T[16];
  std::memcpy(T, init16, sizeof(T));
  std::vector<float> cur(4 * (size_t)N), dists(N);
  std::vector<int32_t> ids(N);
  double fit = 0, err = 0;
  auto match = [&]() {
    // the cloud handed to the matcher is float: T (double) applied to the float source, rounded to float
    for (int i = 0; i < N; ++i) {
      const double x = src4xN[4 * (size_t)i], y = src4xN[4 * (size_t)i + 1], z = src4xN[4 * (size_t)i + 2];
      for (int d = 0; d < 3; ++d) cur[4 * (size_t)i + d] = (float)(T[d] * x + T[d + 4] * y + T[d + 8] * z + T[d + 12]);
      cur[4 * (size_t)i + 3] = 1.f;
    }
    orc_kdtree_knn(tree, cur.data(), N, 1, max_dist, knn_threads, ids.data(), dists.data());
    long c = 0;
    double e2 = 0;
    for (int i = 0; i < N; ++i)
      if (ids[i] >= 0) { ++c; e2 += (double)dists[i]; }
    fit = (double)c / (double)N;
    err = c > 0 ? std::sqrt(e2 / (double)c) : 0.0;
    return c;
  };
  long ncorr = match();
  int it = 0;
  for (; it < max_iteration; ++it) {
    if (ncorr == 0) break;
    double JTJ[36] = {0}, JTr[6] = {0};
    const double R[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
    for (int i = 0; i < N; ++i) {
      if (ids[i] < 0) continue;
      const double s[3] = {cur[4 * (size_t)i], cur[4 * (size_t)i + 1], cur[4 * (size_t)i + 2]};
      const float* tp = tgt4xM + 4 * (size_t)ids[i];
      const double t[3] = {tp[0], tp[1], tp[2]};
      const double d[3] = {s[0] - t[0], s[1] - t[1], s[2] - t[2]};
      if (kind == 1) continue;   // point-to-point is estimated from the correspondences below (Umeyama)
      if (kind == 0) {
        const float* np = tgtn3xM + 3 * (size_t)ids[i];
        const double n[3] = {np[0], np[1], np[2]};
        const double r = d[0] * n[0] + d[1] * n[1] + d[2] * n[2];
        const double J[6] = {s[1] * n[2] - s[2] * n[1], s[2] * n[0] - s[0] * n[2], s[0] * n[1] - s[1] * n[0], n[0], n[1], n[2]};
        for (int a = 0; a < 6; ++a) {
          for (int b = 0; b < 6; ++b) JTJ[a + 6 * b] += J[a] * J[b];
          JTr[a] += J[a] * r;
        }
      } else {
        // source covariance rotated with the cloud: Cs' = R Cs R^T
        const double* Cs0 = srccov9xN + 9 * (size_t)i;
        const double* Ct = tgtcov9xM + 9 * (size_t)ids[i];
        double RC[9], Mm[9], W[9];
        for (int j = 0; j < 3; ++j)
          for (int a = 0; a < 3; ++a) {
            double v = 0;
            for (int k = 0; k < 3; ++k) v += R[a + 3 * k] * Cs0[k + 3 * j];
            RC[a + 3 * j] = v;
          }
        for (int j = 0; j < 3; ++j)
          for (int a = 0; a < 3; ++a) {
            double v = 0;
            for (int k = 0; k < 3; ++k) v += RC[a + 3 * k] * R[j + 3 * k];
            Mm[a + 3 * j] = v + Ct[a + 3 * j];
          }
        inv_sqrt3(Mm, W);
        // J = W [ -[s]x  I ]
        const double nsk[9] = {0, -s[2], s[1], s[2], 0, -s[0], -s[1], s[0], 0};  // -[s]x, column-major
        for (int row = 0; row < 3; ++row) {
          double J[6];
          for (int c = 0; c < 3; ++c) {
            double v = 0;
            for (int k = 0; k < 3; ++k) v += W[row + 3 * k] * nsk[k + 3 * c];
            J[c] = v;
            J[3 + c] = W[row + 3 * c];
          }
          const double r = W[row] * d[0] + W[row + 3] * d[1] + W[row + 6] * d[2];
          for (int a = 0; a < 6; ++a) {
            for (int b = 0; b < 6; ++b) JTJ[a + 6 * b] += J[a] * J[b];
            JTr[a] += J[a] * r;
          }
        }
      }
    }
    double nb[6], x[6], U[16];
    if (kind == 1) {
      // TransformationEstimationPointToPoint::ComputeTransformation: Eigen::umeyama(src, dst, with_scaling = false):
      // sigma = 1/n sum (t - mt)(s - ms)^T = U D V^T, S = diag(1, 1, sign(det U det V)), R = U S V^T, t = mt - R ms
      double ms[3] = {0, 0, 0}, mt[3] = {0, 0, 0}, sig[9] = {0};
      for (int i = 0; i < N; ++i) {
        if (ids[i] < 0) continue;
        for (int a = 0; a < 3; ++a) { ms[a] += cur[4 * (size_t)i + a]; mt[a] += tgt4xM[4 * (size_t)ids[i] + a]; }
      }
      for (int a = 0; a < 3; ++a) { ms[a] /= (double)ncorr; mt[a] /= (double)ncorr; }
      for (int i = 0; i < N; ++i) {
        if (ids[i] < 0) continue;
        for (int c = 0; c < 3; ++c)
          for (int r = 0; r < 3; ++r)
            sig[r + 3 * c] += ((double)tgt4xM[4 * (size_t)ids[i] + r] - mt[r]) * ((double)cur[4 * (size_t)i + c] - ms[c]);
      }
      for (int a = 0; a < 9; ++a) sig[a] /= (double)ncorr;
      double Us[9], Ds[3], Vs[9];
      orc::jacobiSVD<double, 3>(sig, Us, Ds, Vs);
      const double sg = orc::det3(Us) * orc::det3(Vs) < 0 ? -1.0 : 1.0;
      for (int i = 0; i < 16; ++i) U[i] = 0;
      U[15] = 1;
      for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r)
          U[r + 4 * c] = Us[r] * Vs[c] + Us[r + 3] * Vs[c + 3] + sg * Us[r + 6] * Vs[c + 6];
      for (int r = 0; r < 3; ++r) U[r + 12] = mt[r] - (U[r] * ms[0] + U[r + 4] * ms[1] + U[r + 8] * ms[2]);
    } else {
    for (int a = 0; a < 6; ++a) nb[a] = -JTr[a];
    if (!ldlt_solve6(JTJ, nb, x)) break;   // solution_exist == false: identity update, loop would stall — stop
    vec6_to_T(x, U);
    }
    mat4_mul(U, T, T);
    const double fit0 = fit, err0 = err;
    ncorr = match();
    if (std::fabs(fit0 - fit) < rel_fitness && std::fabs(err0 - err) < rel_rmse) { ++it; break; }
  }
  orc_kdtree_free(tree);
  std::memcpy(T16, T, sizeof(T));
  if (fitness) *fitness = fit;
  if (rmse) *rmse = err;
  if (iterations) *iterations = it;
  if (corr_ids) std::memcpy(corr_ids, ids.data(), sizeof(int32_t) * (size_t)N);
  return ORC_OK;
}

// ---- estimateNormalsOrCovariancesIfNeeded (CloudRegistration.cpp:27-46, 67-79) ---------------------------------------
// Open3D v0.15.1 PointCloud::EstimateNormals(KDTreeSearchParamHybrid(radius, max_nn), fast = true),
// NormalizeNormals, OrientNormalsTowardsCameraLocation(Zero) — RECALLED from upstream (geometry/EstimateNormals.cpp,
// utility/Eigen.cpp ComputeCovariance, geometry/PointCloud.cpp): "parity unpinned" like the rest of this file.
//   EstimatePerPointCovariances: indices = hybrid search around the point (itself included, nearest first, at most
//     max_nn, within radius); size < 3 -> Identity; else cumulants (x y z xx xy xz yy yz zz) summed in double over
//     the indices, divided by the count: cov = E[p p^T] - E[p] E[p]^T
//   ComputeNormal: eigenvector of the smallest eigenvalue (FastEigen3x3; Identity -> (0,0,1)); zero -> (0,0,1)
//   NormalizeNormals; orientation: zero normal -> direction to the camera ((0,0,1) when that is zero too), else
//     flipped when normal . (camera - point) < 0
// Stated deviations: the neighbour search runs on the float-converted points with the k-NN contract of knn.cpp
// (Open3D searches the double points with FLANN); the eigenvector comes from a cyclic Jacobi iteration instead of
// FastEigen3x3's closed form — both return the smallest eigenvector of the same matrix, they differ by rounding
// and, for (near-)equal smallest eigenvalues, by the choice within the eigenspace (the tests exclude those points).
extern "C" int orc_o3d_estimate_normals(const float* pts4xN, int N, int max_nn, float radius, int threads,
                                        double* normals3xN, int32_t* n_neighbours) {
  if (N < 0 || max_nn < 1 || !(radius > 0.f)) return ORC_E_BAD_PARAM;
  if (N == 0) return ORC_OK;
  std::vector<int32_t> ids((size_t)max_nn * N);
  std::vector<float> dists((size_t)max_nn * N);
  void* tree = orc_kdtree_build(pts4xN, N);
  orc_kdtree_knn(tree, pts4xN, N, max_nn, radius, threads > 0 ? threads : 1, ids.data(), dists.data());
  orc_kdtree_free(tree);
  for (int i = 0; i < N; ++i) {
    double cum[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    int n = 0;
    for (int j = 0; j < max_nn; ++j) {
      const int id = ids[(size_t)i * max_nn + j];
      if (id < 0) continue;
      const double x = pts4xN[4 * (size_t)id], y = pts4xN[4 * (size_t)id + 1], z = pts4xN[4 * (size_t)id + 2];
      cum[0] += x; cum[1] += y; cum[2] += z;
      cum[3] += x * x; cum[4] += x * y; cum[5] += x * z;
      cum[6] += y * y; cum[7] += y * z; cum[8] += z * z;
      ++n;
    }
    if (n_neighbours) n_neighbours[i] = n;
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
      orc::symEigen<double, 3>(C, w, V);
      int c = 0;
      if (w[1] < w[c]) c = 1;
      if (w[2] < w[c]) c = 2;
      v[0] = V[3 * c]; v[1] = V[3 * c + 1]; v[2] = V[3 * c + 2];
      if (v[0] * v[0] + v[1] * v[1] + v[2] * v[2] == 0.0) { v[0] = 0.0; v[1] = 0.0; v[2] = 1.0; }
    }
    const double nn = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (nn > 0.0) { v[0] /= nn; v[1] /= nn; v[2] /= nn; }
    const double r[3] = {-(double)pts4xN[4 * (size_t)i], -(double)pts4xN[4 * (size_t)i + 1], -(double)pts4xN[4 * (size_t)i + 2]};
    if (v[0] * r[0] + v[1] * r[1] + v[2] * r[2] < 0.0) { v[0] = -v[0]; v[1] = -v[1]; v[2] = -v[2]; }
    normals3xN[3 * (size_t)i] = v[0]; normals3xN[3 * (size_t)i + 1] = v[1]; normals3xN[3 * (size_t)i + 2] = v[2];
  }
  return ORC_OK;
}

// ---- GetInformationMatrixFromPointClouds (Open3D v0.15.1 pipelines/registration/Registration.cpp, recalled; called at
// open3d_slam/src/PlaceRecognition.cpp:148-149) --------------------------------------------------------------------
//   pcd = source transformed by T; correspondences = GetRegistrationResultAndCorrespondences(pcd, target, kdtree,
//   max_correspondence_distance): 1-NN within the distance; GTG = sum over correspondences of G_r G_r^T for the three
//   rows (0, z, -y, 1, 0, 0), (-z, 0, x, 0, 1, 0), (y, -x, 0, 0, 0, 1) of the matched TARGET point (x, y, z).
// Same float-cloud convention as orc_o3d_icp: the transformed source is rounded to float for the match.
extern "C" int orc_o3d_information_matrix(const float* src4xN, int N, const float* tgt4xM, int M, const double* T16,
                                          float max_dist, int threads, double* info36, double* fitness, double* rmse,
                                          int32_t* corr_ids) {
  if (N <= 0 || M <= 0) return ORC_E_BAD_PARAM;
  std::vector<float> q(4 * (size_t)N);
  for (int i = 0; i < N; ++i) {
    const double x = src4xN[4 * (size_t)i], y = src4xN[4 * (size_t)i + 1], z = src4xN[4 * (size_t)i + 2];
    for (int r = 0; r < 3; ++r) q[4 * (size_t)i + r] = (float)(T16[r] * x + T16[r + 4] * y + T16[r + 8] * z + T16[r + 12]);
    q[4 * (size_t)i + 3] = 1.f;
  }
  std::vector<int32_t> ids((size_t)N);
  std::vector<float> d((size_t)N);
  void* tree = orc_kdtree_build(tgt4xM, M);
  orc_kdtree_knn(tree, q.data(), N, 1, max_dist, threads > 0 ? threads : 1, ids.data(), d.data());
  orc_kdtree_free(tree);
  double GTG[36];
  for (double& v : GTG) v = 0.0;
  double cnt = 0.0, s2 = 0.0;
  for (int i = 0; i < N; ++i) {
    if (corr_ids) corr_ids[i] = ids[i];
    if (ids[i] < 0) continue;
    const double x = tgt4xM[4 * (size_t)ids[i]], y = tgt4xM[4 * (size_t)ids[i] + 1], z = tgt4xM[4 * (size_t)ids[i] + 2];
    const double G[3][6] = {{0.0, z, -y, 1.0, 0.0, 0.0}, {-z, 0.0, x, 0.0, 1.0, 0.0}, {y, -x, 0.0, 0.0, 0.0, 1.0}};
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 6; ++c)
        for (int a = 0; a < 6; ++a) GTG[a + 6 * c] += G[r][a] * G[r][c];
    cnt += 1.0;
    s2 += (double)d[i];
  }
  std::memcpy(info36, GTG, sizeof(GTG));
  if (fitness) *fitness = cnt / (double)N;
  if (rmse) *rmse = cnt > 0.0 ? std::sqrt(s2 / cnt) : 0.0;
  return ORC_OK;
}
