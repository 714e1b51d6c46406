// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
//
// Restatement of the per-iteration ICP modules of the reference (float32, column-major). Each function names the
// reference lines it follows. Compiled with -ffp-contract=off: the reference is built -O3 without -march, so its
// float arithmetic has no fused multiply-add.
#include "oracle.h"
#include "la_small.hpp"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

static const float kInf = std::numeric_limits<float>::infinity();

extern "C" {

// Matches::getDistsQuantile — libpointmatcher/pointmatcher/Matches.cpp:60-87
int orc_dists_quantile(const float* dists, long n, float quantile, float* out) {
  std::vector<float> values;
  values.reserve(n);
  for (long i = 0; i < n; ++i)
    if (dists[i] != kInf) values.push_back(dists[i]);
  if (values.empty()) return ORC_E_NO_MATCHES;          // ConvergenceError, Matches.cpp:75-76
  if (quantile < 0.0 || quantile > 1.0) return ORC_E_BAD_PARAM;  // :78-79
  if (quantile == 1.0) {
    *out = *std::max_element(values.begin(), values.end());      // :82-83
    return ORC_OK;
  }
  // `values.size() * quantile` is size_t * float -> float, then truncated (T = float), :84-85
  const float pos = (float)values.size() * quantile;
  const size_t idx = (size_t)pos;
  std::nth_element(values.begin(), values.begin() + idx, values.end());
  *out = values[idx];
  return ORC_OK;
}

// MaxDistOutlierFilter — OutlierFiltersImpl.cpp:66-81 (threshold squared with pow(x,2) in float)
int orc_outlier_maxdist(const float* dists, long n, float maxDist, float* w) {
  const float lim = (float)std::pow(maxDist, 2);
  for (long i = 0; i < n; ++i) w[i] = dists[i] <= lim ? 1.f : 0.f;
  return ORC_OK;
}
// MinDistOutlierFilter — OutlierFiltersImpl.cpp:87-101
int orc_outlier_mindist(const float* dists, long n, float minDist, float* w) {
  const float lim = (float)std::pow(minDist, 2);
  for (long i = 0; i < n; ++i) w[i] = dists[i] >= lim ? 1.f : 0.f;
  return ORC_OK;
}
// MedianDistOutlierFilter — OutlierFiltersImpl.cpp:117-125
int orc_outlier_median(const float* dists, long n, float factor, float* w) {
  float median;
  const int st = orc_dists_quantile(dists, n, 0.5f, &median);
  if (st) return st;
  const float lim = factor * median;
  for (long i = 0; i < n; ++i) w[i] = dists[i] <= lim ? 1.f : 0.f;
  return ORC_OK;
}
// TrimmedDistOutlierFilter — OutlierFiltersImpl.cpp:140-147
int orc_outlier_trimmed(const float* dists, long n, float ratio, float* w) {
  float lim;
  const int st = orc_dists_quantile(dists, n, ratio, &lim);
  if (st) return st;
  for (long i = 0; i < n; ++i) w[i] = dists[i] <= lim ? 1.f : 0.f;
  return ORC_OK;
}
// OutlierFilters::compute, empty chain — OutlierFilter.cpp:66-83
int orc_outlier_null(const float* dists, long n, float* w) {
  for (long i = 0; i < n; ++i) w[i] = dists[i] == kInf ? 0.f : 1.f;
  return ORC_OK;
}

static inline void normalized3(const float* v, float* o) {
  // Eigen normalized(): z = squaredNorm(); z > 0 ? v / sqrt(z) : v
  float z = v[0] * v[0];
  z = z + v[1] * v[1];
  z = z + v[2] * v[2];
  if (z > 0.f) {
    const float s = std::sqrt(z);
    o[0] = v[0] / s; o[1] = v[1] / s; o[2] = v[2] / s;
  } else {
    o[0] = v[0]; o[1] = v[1]; o[2] = v[2];
  }
}

// SurfaceNormalOutlierFilter — OutlierFiltersImpl.cpp:236-288 (signed cosine, "new version")
int orc_outlier_surface_normal(const float* nread, const float* nref, const int32_t* ids, int k, int N,
                               float maxAngle, float* w) {
  const float eps = std::cos(maxAngle);  // :229
  if (!nread || !nref) {                 // :273-283
    for (long i = 0; i < (long)k * N; ++i) w[i] = 1.f;
    return ORC_OK;
  }
  for (int x = 0; x < N; ++x) {
    float nr[3];
    normalized3(nread + 3 * (size_t)x, nr);
    for (int y = 0; y < k; ++y) {
      const int id = ids[(size_t)x * k + y];
      if (id == -1) { w[(size_t)x * k + y] = 0.f; continue; }
      float nf[3];
      normalized3(nref + 3 * (size_t)id, nf);
      float v = nr[0] * nf[0];
      v = v + nr[1] * nf[1];
      v = v + nr[2] * nf[2];
      w[(size_t)x * k + y] = (v < eps) ? 0.f : 1.f;
    }
  }
  return ORC_OK;
}

// ErrorElements ctor — ErrorMinimizer.cpp:59-193
int orc_error_elements(const float* read, const float* ref, const float* refn, const float* w, const int32_t* ids,
                       const float* dists, int k, int N, float* readP, float* refP, float* refnP, float* wP,
                       int32_t* idsP, float* distsP, int* P_out, float* pointUsedRatio,
                       float* weightedPointUsedRatio) {
  int j = 0;
  float wsum = 0.f;
  for (int i = 0; i < N; ++i)
    for (int kk = 0; kk < k; ++kk) {
      const size_t e = (size_t)i * k + kk;
      if (dists[e] == kInf) continue;       // :103-105
      if (w[e] != 0.0f) {                   // :107
        if (readP) std::memcpy(readP + 4 * (size_t)j, read + 4 * (size_t)i, 16);
        const int id = ids[e];
        if (refP) std::memcpy(refP + 4 * (size_t)j, ref + 4 * (size_t)id, 16);
        if (refnP && refn) std::memcpy(refnP + 3 * (size_t)j, refn + 3 * (size_t)id, 12);
        if (wP) wP[j] = w[e];
        if (idsP) idsP[j] = id;
        if (distsP) distsP[j] = dists[e];
        ++j;
        wsum += w[e];                       // :121
      }
    }
  *P_out = j;
  if (pointUsedRatio) *pointUsedRatio = (float)j / (float)(k * N);          // :139
  if (weightedPointUsedRatio) *weightedPointUsedRatio = wsum / (float)(k * N);  // :140
  if (j == 0) return ORC_E_NO_MATCHES;      // :75-77
  return ORC_OK;
}

// ICP::calculateOptimizationHessian — ICP.cpp:1513-1566, crossProduct — ErrorMinimizer.cpp:355-388
int orc_p2plane_hb(const float* read, const float* ref, const float* n, const float* w, int P, int accum, float* H,
                   float* b) {
  double Hd[36] = {0}, bd[6] = {0};
  float Hf[36] = {0}, bf[6] = {0};
  for (int p = 0; p < P; ++p) {
    const float* a = read + 4 * (size_t)p;
    const float* q = ref + 4 * (size_t)p;
    const float* nn = n + 3 * (size_t)p;
    float F[6], wF[6];
    F[0] = a[1] * nn[2] - a[2] * nn[1];   // ErrorMinimizer.cpp:376-378
    F[1] = a[2] * nn[0] - a[0] * nn[2];
    F[2] = a[0] * nn[1] - a[1] * nn[0];
    F[3] = nn[0]; F[4] = nn[1]; F[5] = nn[2];
    for (int i = 0; i < 6; ++i) wF[i] = w[p] * F[i];  // ICP.cpp:1534-1543
    // deltas and dot product, ICP.cpp:1555-1563 (dotProd starts at zero and adds x, y, z terms)
    float dot = 0.f;
    for (int i = 0; i < 3; ++i) dot = dot + (a[i] - q[i]) * nn[i];
    for (int j = 0; j < 6; ++j)
      for (int i = 0; i < 6; ++i) {
        const float prod = wF[i] * F[j];   // hessian = wF * F^T, ICP.cpp:1546
        Hf[i + 6 * j] += prod;
        Hd[i + 6 * j] += (double)prod;
      }
    for (int i = 0; i < 6; ++i) {
      const float prod = wF[i] * dot;      // constraints = -(wF * dot^T), ICP.cpp:1565
      bf[i] += prod;
      bd[i] += (double)prod;
    }
  }
  for (int i = 0; i < 36; ++i) H[i] = accum ? (float)Hd[i] : Hf[i];
  for (int i = 0; i < 6; ++i) b[i] = accum ? -(float)bd[i] : -bf[i];
  return ORC_OK;
}

// solvePossiblyUnderdeterminedLinearSystem — PointToPlane.cpp:199-256. N = 6 for the plain system, 6 + c for the
// equality-constrained (KKT) system of solvePossiblyUnderdeterminedLinearSystemWithEqualityConstraints (:461-505);
// isConstrained only changes what the reference prints.
extern "C++" {
namespace {
template <int N>
int solveN(const float* A, const float* b, float* x, int* info) {
  orc::FullPivQR<float, N> qr;
  qr.compute(A);
  auto svd_double = [&]() {
    double Ad[N * N], bd[N], xd[N];
    for (int i = 0; i < N * N; ++i) Ad[i] = A[i];
    for (int i = 0; i < N; ++i) bd[i] = b[i];
    orc::svdSolve<double, N>(Ad, bd, xd);
    for (int i = 0; i < N; ++i) x[i] = (float)xd[i];
  };
  if (qr.isInvertible()) {   // :203 / :249-255
    svd_double();
    if (info) *info = 0;
    return ORC_OK;
  }
  // rank-deficient: R1 y = Q1^T b, minimal norm — :205-229
  const int rank = qr.rank();
  if (info) *info = 1;
  if (rank == 0) {
    for (int i = 0; i < N; ++i) x[i] = 0.f;  // llt of a 0x0 system: x = 0 (then NaN guard downstream is moot)
  } else {
    float Q[N * N];
    qr.matrixQ(Q);
    // AP = A * colsPermutation
    float AP[N * N];
    for (int j = 0; j < N; ++j)
      for (int i = 0; i < N; ++i) AP[i + N * j] = A[i + N * qr.colPerm[j]];
    // R1 = (Q1^T * A * P).block(0,0,rank,6), Q1t = Q^T rows 0..rank-1
    std::vector<float> R1(rank * N), qtb(rank);
    for (int r = 0; r < rank; ++r) {
      for (int j = 0; j < N; ++j) {
        float acc = 0.f;
        for (int i = 0; i < N; ++i) acc += Q[i + N * r] * AP[i + N * j];
        R1[r + rank * j] = acc;
      }
      float acc = 0.f;
      for (int i = 0; i < N; ++i) acc += Q[i + N * r] * b[i];
      qtb[r] = acc;
    }
    // G = R1 * R1^T (rank x rank), Cholesky solve G z = Q1^T b  (llt().solve, :222)
    std::vector<float> G(rank * rank), z(rank);
    for (int a = 0; a < rank; ++a)
      for (int c = 0; c < rank; ++c) {
        float acc = 0.f;
        for (int j = 0; j < N; ++j) acc += R1[a + rank * j] * R1[c + rank * j];
        G[a + rank * c] = acc;
      }
    std::vector<float> L(rank * rank, 0.f);
    for (int j = 0; j < rank; ++j) {
      float d = G[j + rank * j];
      for (int kk = 0; kk < j; ++kk) d -= L[j + rank * kk] * L[j + rank * kk];
      d = std::sqrt(d);
      L[j + rank * j] = d;
      for (int i = j + 1; i < rank; ++i) {
        float s = G[i + rank * j];
        for (int kk = 0; kk < j; ++kk) s -= L[i + rank * kk] * L[j + rank * kk];
        L[i + rank * j] = s / d;
      }
    }
    for (int i = 0; i < rank; ++i) {
      float s = qtb[i];
      for (int kk = 0; kk < i; ++kk) s -= L[i + rank * kk] * z[kk];
      z[i] = s / L[i + rank * i];
    }
    for (int i = rank - 1; i >= 0; --i) {
      float s = z[i];
      for (int kk = i + 1; kk < rank; ++kk) s -= L[kk + rank * i] * z[kk];
      z[i] = s / L[i + rank * i];
    }
    // y = R1.triangularView<Upper>()^T * z : only the upper-triangular part (j >= r) of R1 is used, :222
    float y[N];
    for (int j = 0; j < N; ++j) {
      float acc = 0.f;
      for (int r = 0; r < rank && r <= j; ++r) acc += R1[r + rank * j] * z[r];
      y[j] = acc;
    }
    // x = colsPermutation * y : x[colPerm[j]] = y[j], :231
    for (int j = 0; j < N; ++j) x[qr.colPerm[j]] = y[j];
  }
  // accuracy check b.isApprox(A x, 1e-5) — :233-234 (||b - ax||^2 <= prec^2 * min(||b||^2, ||ax||^2))
  float ax[N];
  for (int i = 0; i < N; ++i) {
    float acc = 0.f;
    for (int j = 0; j < N; ++j) acc += A[i + N * j] * x[j];
    ax[i] = acc;
  }
  float e2 = 0.f, b2 = 0.f, a2 = 0.f;
  for (int i = 0; i < N; ++i) { e2 += (b[i] - ax[i]) * (b[i] - ax[i]); b2 += b[i] * b[i]; a2 += ax[i] * ax[i]; }
  if (!(e2 <= 1e-5f * 1e-5f * std::min(b2, a2))) {
    svd_double();  // :236-238
    if (info) *info = 2;
  }
  return ORC_OK;
}
}  // namespace
}  // extern "C++"

int orc_solve6(const float* A, const float* b, float* x, int* info) { return solveN<6>(A, b, x, info); }

int orc_solve_n(const float* A, const float* b, int n, float* x, int* info) {
  switch (n) {
    case 6: return solveN<6>(A, b, x, info);
    case 7: return solveN<7>(A, b, x, info);
    case 8: return solveN<8>(A, b, x, info);
    case 9: return solveN<9>(A, b, x, info);
    case 10: return solveN<10>(A, b, x, info);
    case 11: return solveN<11>(A, b, x, info);
    case 12: return solveN<12>(A, b, x, info);
    default: return ORC_E_BAD_PARAM;
  }
}

// x -> T : PointToPlane.cpp:337-381 (6-DOF branch; force2D/force4DOF unsupported)
int orc_p2plane_T(const float* x, float* T) {
  for (int i = 0; i < 16; ++i) T[i] = 0.f;
  T[0] = T[5] = T[10] = T[15] = 1.f;
  float n2 = x[0] * x[0];
  n2 = n2 + x[1] * x[1];
  n2 = n2 + x[2] * x[2];
  const float nrm = std::sqrt(n2);
  const float angle = std::atan(nrm);            // :341
  // stableNormalized(): w = max|x_i|; z = (x / w).squaredNorm(); z > 0 ? (x / w) / sqrt(z) : x
  float ax[3] = {x[0], x[1], x[2]};
  const float wmax = std::max(std::fabs(x[0]), std::max(std::fabs(x[1]), std::fabs(x[2])));
  {
    const float s0 = x[0] / wmax, s1 = x[1] / wmax, s2 = x[2] / wmax;
    float z = s0 * s0;
    z = z + s1 * s1;
    z = z + s2 * s2;
    if (z > 0.f) {
      const float sq = std::sqrt(z);
      ax[0] = s0 / sq; ax[1] = s1 / sq; ax[2] = s2 / sq;
    }
  }
  // AngleAxis::toRotationMatrix
  const float s = std::sin(angle), c = std::cos(angle);
  const float sx = s * ax[0], sy = s * ax[1], sz = s * ax[2];
  const float cx = (1.f - c) * ax[0], cy = (1.f - c) * ax[1], cz = (1.f - c) * ax[2];
  float R[9];
  float tmp;
  tmp = cx * ax[1]; R[0 + 3 * 1] = tmp - sz; R[1 + 3 * 0] = tmp + sz;
  tmp = cx * ax[2]; R[0 + 3 * 2] = tmp + sy; R[2 + 3 * 0] = tmp - sy;
  tmp = cy * ax[2]; R[1 + 3 * 2] = tmp - sx; R[2 + 3 * 1] = tmp + sx;
  R[0] = cx * ax[0] + c; R[4] = cy * ax[1] + c; R[8] = cz * ax[2] + c;
  bool nan = false;
  for (int i = 0; i < 9; ++i) nan |= (R[i] != R[i]);
  for (int i = 3; i < 6; ++i) nan |= (x[i] != x[i]);
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) T[i + 4 * j] = R[i + 3 * j];
  T[12] = x[3]; T[13] = x[4]; T[14] = x[5];     // :357-359
  if (nan) {                                     // :374-380: rotation block reset to identity
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < 3; ++i) T[i + 4 * j] = (i == j) ? 1.f : 0.f;
  }
  return ORC_OK;
}

// PointToPointErrorMinimizer::compute_in_place — PointToPoint.cpp:62-103
int orc_p2point_T(const float* read, const float* ref, const float* w, int P, float* T) {
  return orc_p2point_T2(read, ref, w, P, 0, T);
}

// accum 0: reference-literal float sums in index order (the order is an Eigen artefact); accum 1: the same float
// inputs summed in double ("truth"), rounded to float before the SVD — what the device path is compared with.
int orc_p2point_T2(const float* read, const float* ref, const float* w, int P, int accum, float* T) {
  float mr[3] = {0, 0, 0}, mf[3] = {0, 0, 0};
  float m[9] = {0};  // m = ref_c * diag(w) * read_c^T, column-major m(i,j)
  if (!accum) {
    float wsum = 0.f;
    for (int p = 0; p < P; ++p) wsum += w[p];
    const float winv = 1.f / wsum;
    for (int p = 0; p < P; ++p)
      for (int d = 0; d < 3; ++d) {
        mr[d] += read[4 * (size_t)p + d] * w[p];
        mf[d] += ref[4 * (size_t)p + d] * w[p];
      }
    for (int d = 0; d < 3; ++d) { mr[d] *= winv; mf[d] *= winv; }
    for (int p = 0; p < P; ++p) {
      float rc[3], fc[3];
      for (int d = 0; d < 3; ++d) { rc[d] = read[4 * (size_t)p + d] - mr[d]; fc[d] = ref[4 * (size_t)p + d] - mf[d]; }
      for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i) m[i + 3 * j] += (fc[i] * w[p]) * rc[j];
    }
  } else {
    double wsum = 0.0, dr[3] = {0, 0, 0}, df[3] = {0, 0, 0}, dm[9] = {0};
    for (int p = 0; p < P; ++p) {
      const double wd = (double)w[p];
      wsum += wd;
      for (int d = 0; d < 3; ++d) { dr[d] += wd * (double)read[4 * (size_t)p + d]; df[d] += wd * (double)ref[4 * (size_t)p + d]; }
    }
    for (int d = 0; d < 3; ++d) { dr[d] /= wsum; df[d] /= wsum; }
    for (int p = 0; p < P; ++p) {
      const double wd = (double)w[p];
      for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i)
          dm[i + 3 * j] += ((double)ref[4 * (size_t)p + i] - df[i]) * wd * ((double)read[4 * (size_t)p + j] - dr[j]);
    }
    for (int d = 0; d < 3; ++d) { mr[d] = (float)dr[d]; mf[d] = (float)df[d]; }
    for (int i = 0; i < 9; ++i) m[i] = (float)dm[i];
  }
  float U[9], S[3], V[9];
  orc::jacobiSVD<float, 3>(m, U, S, V);
  float R[9];
  auto mulUVt = [&](const float* Vt) {  // R = U * Vt, Vt given column-major
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < 3; ++i) {
        float acc = 0.f;
        for (int kk = 0; kk < 3; ++kk) acc += U[i + 3 * kk] * Vt[kk + 3 * j];
        R[i + 3 * j] = acc;
      }
  };
  float Vt[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Vt[i + 3 * j] = V[j + 3 * i];
  mulUVt(Vt);
  if (orc::det3(R) < 0.f) {  // :92-97: negate last row of V^T
    for (int j = 0; j < 3; ++j) Vt[2 + 3 * j] = -Vt[2 + 3 * j];
    mulUVt(Vt);
  }
  for (int i = 0; i < 16; ++i) T[i] = 0.f;
  T[15] = 1.f;
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) T[i + 4 * j] = R[i + 3 * j];
  for (int i = 0; i < 3; ++i) {
    float acc = 0.f;
    for (int j = 0; j < 3; ++j) acc += R[i + 3 * j] * mr[j];
    T[12 + i] = mf[i] - acc;
  }
  return ORC_OK;
}

// RigidTransformation::inPlaceCompute — TransformationsImpl.cpp:61-102, checkParameters :105-119,
// correctParameters :122-170. Product order: acc = 0; acc += T(i,0)*x; += T(i,1)*y; += T(i,2)*z; += T(i,3)*w
// (depth-sequential GEMM accumulation, no FMA).
int orc_rigid_apply(const float* Tin, float* feat, float* normals, int N) {
  float T[16];
  std::memcpy(T, Tin, sizeof(T));
  float R[9];
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) R[i + 3 * j] = T[i + 4 * j];
  if (std::fabs(1.f - orc::det3(R)) > 0.001f) {
    float c1[3], c2[3], n0[3], n1[3];
    normalized3(&T[4], c1);
    normalized3(&T[8], c2);
    n0[0] = c1[1] * c2[2] - c1[2] * c2[1]; n0[1] = c1[2] * c2[0] - c1[0] * c2[2]; n0[2] = c1[0] * c2[1] - c1[1] * c2[0];
    n1[0] = c2[1] * n0[2] - c2[2] * n0[1]; n1[1] = c2[2] * n0[0] - c2[0] * n0[2]; n1[2] = c2[0] * n0[1] - c2[1] * n0[0];
    for (int i = 0; i < 3; ++i) { T[i] = n0[i]; T[4 + i] = n1[i]; T[8 + i] = c2[i]; }
  }
  for (int p = 0; p < N; ++p) {
    float* f = feat + 4 * (size_t)p;
    const float x = f[0], y = f[1], z = f[2], w = f[3];
    for (int i = 0; i < 4; ++i) {
      float acc = T[i] * x;
      acc = acc + T[i + 4] * y;
      acc = acc + T[i + 8] * z;
      acc = acc + T[i + 12] * w;
      f[i] = acc;
    }
  }
  if (normals) {
    // descriptors rotate with R of the *uncorrected* parameters (TransformationsImpl.cpp:72, 98)
    for (int p = 0; p < N; ++p) {
      float* f = normals + 3 * (size_t)p;
      const float x = f[0], y = f[1], z = f[2];
      for (int i = 0; i < 3; ++i) {
        float acc = R[i] * x;
        acc = acc + R[i + 3] * y;
        acc = acc + R[i + 6] * z;
        f[i] = acc;
      }
    }
  }
  return ORC_OK;
}

}  // extern "C"
