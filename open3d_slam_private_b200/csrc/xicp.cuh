// X-ICP localizability on the device: degeneracyAwareness = OptimizedEqualityConstraints of the reference's ICP
// (param/icp.yaml of open3d_slam_ros; ICP.cpp:629-670 parameters, :1112-1141 call site).
//
//   ICP::eigenAnalysis                                   ICP.cpp:1580-1591      -> xicp_eigen_analysis
//   ICP::detectLocalizabilityWithOptimizedMethod          ICP.cpp:2187-2444      -> alignment pass in icp_fused.cu +
//                                                                                   xicp_decide
//   ICP::detectLocalizability                             ICP.cpp:2128-2155      -> xicp_decide (see below)
//   solvePossiblyUnderdeterminedLinearSystemWithEqualityConstraints, readLocalizabilityFlags,
//   generateConstrainedOptimizationProblem                PointToPlane.cpp:461-505, 570-626, 757-778 -> xicp_solve
//
// detectLocalizability is a sequential float loop whose exit test sits in the loop condition. Every term is
// non-negative, so both running sums are monotone and "some prefix reached a threshold" is the same statement as
// "the full sum reached it". The device therefore forms the two full sums per direction in parallel (double) and
// certifies the decision against the worst-case rounding of the reference's float recurrence (xicp_band); a direction
// whose sums fall inside that band is decided by replaying the reference's own sequential recurrence (icp_fused.cu),
// so the flags are the reference's by construction.
#pragma once
#include "small_la.cuh"
#include "solve.cuh"

namespace b200 {

// Relative half-width of the band around a threshold inside which the parallel double sums do not certify the
// reference's decision: the float recurrence of n <= P non-negative terms that stops at the threshold is off by at most
// n * 2^-24 relative (half an ulp per addition of a running sum <= threshold); 20 % on top for the terms' own rounding
// (the centre of calculatePointCloudCenter is summed in double here, sequentially in float there), floor 0.5 %.
__device__ __forceinline__ double xicp_band(int P) { return fmax(5e-3, 1.2 * (double)P * 5.9604644775390625e-8); }
constexpr int kKktCap = 12;          // 6 unknowns + up to 6 equality constraints

// state of LocalizabilityParametersForErrorMinimization (PointMatcher.h:585-660) for one registration
struct XicpState {
  int repeat_eigen;     // debugging_.whetherToRepeatEigenAnalysis_ (:592), starts true
  int return_prior;     // debugging_.whetherToReturnPrior_ (:595)
  int borderline;       // decisions that fell inside the rounding band and were replayed exactly
  int border_mask;      // bit e: translation eigenvector e, bit 3 + e: rotation eigenvector e
  int loc_xyz[3], loc_rpy[3];   // 1 = kLocalizable (:638-639)
  float cT[3], cR[3];   // translation/rotationConstraintValues_ (:625-626)
  float Ut[9], Ur[9];   // translation/rotationEigenvectors_ (:634-635), columns, optimisation (refMean) frame
  float UtB[9], UrB[9]; // the same in the frame the reading came in (first iteration only)
  float center[3];      // calculatePointCloudCenter (ICP.cpp:1594-1618) of the matched reading points, that frame
  float R[9], Rinv[9];  // rotation of T_refMean_dataIn and its inverse
  double sums[12];
};

__device__ inline void xicp_state_init(XicpState& X) {
  X.repeat_eigen = 1; X.return_prior = 0; X.borderline = 0; X.border_mask = 0;
  for (int i = 0; i < 3; ++i) { X.loc_xyz[i] = 1; X.loc_rpy[i] = 1; X.cT[i] = 0.f; X.cR[i] = 0.f; X.center[i] = 0.f; }
  for (int i = 0; i < 9; ++i) { X.Ut[i] = 0.f; X.Ur[i] = 0.f; X.UtB[i] = 0.f; X.UrB[i] = 0.f; }
  for (int i = 0; i < 12; ++i) X.sums[i] = 0.0;
}

// dynamic-size 3x3 .inverse() (PartialPivLU): Gauss-Jordan with partial pivoting
__device__ inline bool inv3_dev(const float* A, float* Ainv) {
  float a[3][6];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { a[i][j] = A[i + 3 * j]; a[i][3 + j] = (i == j) ? 1.f : 0.f; }
  for (int c = 0; c < 3; ++c) {
    int piv = c;
    for (int r = c + 1; r < 3; ++r)
      if (fabsf(a[r][c]) > fabsf(a[piv][c])) piv = r;
    if (a[piv][c] == 0.f) return false;
    if (piv != c)
      for (int j = 0; j < 6; ++j) { const float t = a[piv][j]; a[piv][j] = a[c][j]; a[c][j] = t; }
    const float d = a[c][c];
    for (int j = 0; j < 6; ++j) a[c][j] /= d;
    for (int r = 0; r < 3; ++r)
      if (r != c) {
        const float f = a[r][c];
        for (int j = 0; j < 6; ++j) a[r][j] -= f * a[c][j];
      }
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Ainv[i + 3 * j] = a[i][3 + j];
  return true;
}

__device__ inline void mul3v_dev(const float* R, const float* v, float* o) {
  for (int i = 0; i < 3; ++i) o[i] = R[i] * v[0] + R[i + 3] * v[1] + R[i + 6] * v[2];
}

// Eigen's U is a full orthonormal basis even when the smallest singular value is exactly zero
__device__ inline void complete_u3(float* U, const float* S) {
  if (S[2] == 0.f && S[1] > 0.f) {
    U[6] = U[1] * U[5] - U[2] * U[4];
    U[7] = U[2] * U[3] - U[0] * U[5];
    U[8] = U[0] * U[4] - U[1] * U[3];
  }
}

// ICP::eigenAnalysis(hessian, translationEigenvectors33, rotationEigenvectors33), ICP.cpp:1580-1591
__device__ inline void xicp_eigen_analysis(const float* H, float* Ut, float* Ur) {
  float B[9], S[3], V[9];
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) B[i + 3 * j] = H[(3 + i) + 6 * (3 + j)];
  svd3f(B, Ut, S, V);
  complete_u3(Ut, S);
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) B[i + 3 * j] = H[i + 6 * j];
  svd3f(B, Ur, S, V);
  complete_u3(Ur, S);
}

// outcome of ICP::detectLocalizability from the two full sums: 1 enough information, 0 not; *borderline set when the
// sums are within the rounding band of a threshold the decision depends on
__device__ inline int xicp_decide(double combined, double high, float enough, float insufficient, int P,
                                  int* borderline) {
  const double e = (double)enough, u = (double)insufficient, band = xicp_band(P);
  const bool yes = (combined >= e) || (high >= u);
  const bool sure_yes = (combined >= e * (1.0 + band)) || (high >= u * (1.0 + band));
  const bool sure_no = (combined < e * (1.0 - band)) && (high < u * (1.0 - band));
  if (!(sure_yes || sure_no)) *borderline = 1;
  return yes ? 1 : 0;
}

// ---- solvePossiblyUnderdeterminedLinearSystem for the (6 + c) x (6 + c) equality-constrained system, one thread,
// work arrays in caller-provided scratch (shared memory). Same operations in the same order as la::PivQR /
// la::svdSolve / solve6_dev with N = n.
struct KktScratch {
  float A[kKktCap * kKktCap], qr[kKktCap * kKktCap], Q[kKktCap * kKktCap], AP[kKktCap * kKktCap];
  float R1[kKktCap * kKktCap], G[kKktCap * kKktCap], L[kKktCap * kKktCap];
  double W[kKktCap * kKktCap], V[kKktCap * kKktCap];
  float b[kKktCap], x[kKktCap], tau[kKktCap], qtb[kKktCap], z[kKktCap], y[kKktCap];
  double S[kKktCap], xd[kKktCap];
  int rowT[kKktCap], colT[kKktCap], perm[kKktCap];
};

// one-sided Jacobi SVD solve in double (la::svdSolve with runtime n), A column-major n x n with stride n
__device__ inline void svd_solve_rt(const float* A, const float* b, int n, KktScratch& k) {
  double* W = k.W; double* V = k.V;
  for (int i = 0; i < n * n; ++i) { W[i] = (double)A[i]; V[i] = 0.0; }
  for (int i = 0; i < n; ++i) V[i + i * n] = 1.0;
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rot = false;
    for (int p = 0; p < n - 1; ++p)
      for (int q = p + 1; q < n; ++q) {
        double al = 0, be = 0, ga = 0;
        for (int i = 0; i < n; ++i) {
          al += W[i + p * n] * W[i + p * n];
          be += W[i + q * n] * W[i + q * n];
          ga += W[i + p * n] * W[i + q * n];
        }
        if (ga == 0.0 || fabs(ga) <= DBL_EPSILON * sqrt(al * be)) continue;
        rot = true;
        const double zeta = (be - al) / (2.0 * ga);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
        for (int i = 0; i < n; ++i) {
          const double wp = W[i + p * n], wq = W[i + q * n];
          W[i + p * n] = c * wp - s * wq; W[i + q * n] = s * wp + c * wq;
          const double vp = V[i + p * n], vq = V[i + q * n];
          V[i + p * n] = c * vp - s * vq; V[i + q * n] = s * vp + c * vq;
        }
      }
    if (!rot) break;
  }
  double smax = 0.0;
  for (int j = 0; j < n; ++j) {
    double n2 = 0;
    for (int i = 0; i < n; ++i) n2 += W[i + j * n] * W[i + j * n];
    k.S[j] = sqrt(n2);
    if (k.S[j] > smax) smax = k.S[j];
  }
  double thr = smax * (double)n * DBL_EPSILON;
  if (thr < DBL_MIN) thr = DBL_MIN;
  for (int i = 0; i < n; ++i) k.xd[i] = 0.0;
  for (int j = 0; j < n; ++j) {
    if (!(k.S[j] > thr)) continue;
    double d = 0;
    for (int i = 0; i < n; ++i) d += W[i + j * n] * (double)b[i];
    d = d / (k.S[j] * k.S[j]);
    for (int i = 0; i < n; ++i) k.xd[i] += V[i + j * n] * d;
  }
  for (int i = 0; i < n; ++i) k.x[i] = (float)k.xd[i];
}

// info: 0 full rank, 1 rank-deficient minimal-norm, 2 double-SVD fallback (as solve6_dev). Input k.A / k.b (stride n),
// output k.x.
__device__ inline void solve_rt_dev(int n, KktScratch& k, int* info) {
  const float* A = k.A; const float* b = k.b;
  float* qr = k.qr;
  // ---- la::PivQR<float, n>::compute
  for (int i = 0; i < n * n; ++i) qr[i] = A[i];
  const float prec = FLT_EPSILON * (float)n;
  int nonzero = n;
  float maxpivot = 0.f, biggest = 0.f;
  for (int kk = 0; kk < n; ++kk) {
    int rb = kk, cb = kk;
    float big = -1.f;
    for (int j = kk; j < n; ++j)
      for (int i = kk; i < n; ++i) {
        const float v = fabsf(qr[i + j * n]);
        if (v > big) { big = v; rb = i; cb = j; }
      }
    if (kk == 0) biggest = big;
    if (fabsf(big) <= fabsf(biggest) * prec) {
      nonzero = kk;
      for (int i = kk; i < n; ++i) { k.rowT[i] = i; k.colT[i] = i; k.tau[i] = 0.f; }
      break;
    }
    k.rowT[kk] = rb; k.colT[kk] = cb;
    if (rb != kk)
      for (int j = kk; j < n; ++j) { const float t = qr[kk + j * n]; qr[kk + j * n] = qr[rb + j * n]; qr[rb + j * n] = t; }
    if (cb != kk)
      for (int i = 0; i < n; ++i) { const float t = qr[i + kk * n]; qr[i + kk * n] = qr[i + cb * n]; qr[i + cb * n] = t; }
    float tail = 0.f;
    for (int i = kk + 1; i < n; ++i) tail += qr[i + kk * n] * qr[i + kk * n];
    const float c0 = qr[kk + kk * n];
    float beta, tk;
    if (tail <= FLT_MIN) {
      tk = 0.f; beta = c0;
      for (int i = kk + 1; i < n; ++i) qr[i + kk * n] = 0.f;
    } else {
      beta = sqrtf(c0 * c0 + tail);
      if (c0 >= 0.f) beta = -beta;
      for (int i = kk + 1; i < n; ++i) qr[i + kk * n] = qr[i + kk * n] / (c0 - beta);
      tk = (beta - c0) / beta;
    }
    qr[kk + kk * n] = beta;
    k.tau[kk] = tk;
    if (fabsf(beta) > maxpivot) maxpivot = fabsf(beta);
    if (n - kk > 1 && tk != 0.f) {
      for (int j = kk + 1; j < n; ++j) {
        float tmp = 0.f;
        for (int i = kk + 1; i < n; ++i) tmp += qr[i + kk * n] * qr[i + j * n];
        tmp += qr[kk + j * n];
        qr[kk + j * n] -= tk * tmp;
        for (int i = kk + 1; i < n; ++i) qr[i + j * n] -= tk * qr[i + kk * n] * tmp;
      }
    }
  }
  for (int i = 0; i < n; ++i) k.perm[i] = i;
  for (int kk = 0; kk < n; ++kk) { const int t = k.perm[kk]; k.perm[kk] = k.perm[k.colT[kk]]; k.perm[k.colT[kk]] = t; }
  int rank = 0;
  {
    const float pre = fabsf(maxpivot) * (FLT_EPSILON * (float)n);
    for (int i = 0; i < nonzero; ++i) rank += (fabsf(qr[i + i * n]) > pre) ? 1 : 0;
  }
  if (rank == n) {  // isInvertible(): double-precision SVD solve, PointToPlane.cpp:249-255
    svd_solve_rt(A, b, n, k);
    *info = 0;
    return;
  }
  *info = 1;
  if (rank == 0) {
    for (int i = 0; i < n; ++i) k.x[i] = 0.f;
  } else {
    // ---- matrixQ
    float* Q = k.Q;
    for (int i = 0; i < n * n; ++i) Q[i] = 0.f;
    for (int i = 0; i < n; ++i) Q[i + i * n] = 1.f;
    for (int kk = n - 1; kk >= 0; --kk) {
      const float tk = k.tau[kk];
      if (tk != 0.f) {
        if (n - kk > 1) {
          for (int j = kk; j < n; ++j) {
            float tmp = Q[kk + j * n];
            for (int i = kk + 1; i < n; ++i) tmp += qr[i + kk * n] * Q[i + j * n];
            Q[kk + j * n] -= tk * tmp;
            for (int i = kk + 1; i < n; ++i) Q[i + j * n] -= tk * qr[i + kk * n] * tmp;
          }
        } else {
          for (int j = kk; j < n; ++j) Q[kk + j * n] *= (1.f - tk);
        }
      }
      const int r = k.rowT[kk];
      if (r != kk)
        for (int j = 0; j < n; ++j) { const float t = Q[kk + j * n]; Q[kk + j * n] = Q[r + j * n]; Q[r + j * n] = t; }
    }
    // ---- minimal-norm solution of R1 y = Q1^T b, PointToPlane.cpp:205-231
    float *AP = k.AP, *R1 = k.R1, *G = k.G, *L = k.L;
    for (int j = 0; j < n; ++j)
      for (int i = 0; i < n; ++i) AP[i + n * j] = A[i + n * k.perm[j]];
    for (int r = 0; r < rank; ++r) {
      for (int j = 0; j < n; ++j) {
        float acc = 0.f;
        for (int i = 0; i < n; ++i) acc += Q[i + n * r] * AP[i + n * j];
        R1[r + n * j] = acc;
      }
      float acc = 0.f;
      for (int i = 0; i < n; ++i) acc += Q[i + n * r] * b[i];
      k.qtb[r] = acc;
    }
    for (int a = 0; a < rank; ++a)
      for (int c = 0; c < rank; ++c) {
        float acc = 0.f;
        for (int j = 0; j < n; ++j) acc += R1[a + n * j] * R1[c + n * j];
        G[a + n * c] = acc;
      }
    for (int i = 0; i < n * n; ++i) L[i] = 0.f;
    for (int j = 0; j < rank; ++j) {
      float d = G[j + n * j];
      for (int kk = 0; kk < j; ++kk) d -= L[j + n * kk] * L[j + n * kk];
      d = sqrtf(d);
      L[j + n * j] = d;
      for (int i = j + 1; i < rank; ++i) {
        float s = G[i + n * j];
        for (int kk = 0; kk < j; ++kk) s -= L[i + n * kk] * L[j + n * kk];
        L[i + n * j] = s / d;
      }
    }
    for (int i = 0; i < rank; ++i) {
      float s = k.qtb[i];
      for (int kk = 0; kk < i; ++kk) s -= L[i + n * kk] * k.z[kk];
      k.z[i] = s / L[i + n * i];
    }
    for (int i = rank - 1; i >= 0; --i) {
      float s = k.z[i];
      for (int kk = i + 1; kk < rank; ++kk) s -= L[kk + n * i] * k.z[kk];
      k.z[i] = s / L[i + n * i];
    }
    for (int j = 0; j < n; ++j) {
      float acc = 0.f;
      for (int r = 0; r < rank && r <= j; ++r) acc += R1[r + n * j] * k.z[r];
      k.y[j] = acc;
    }
    for (int j = 0; j < n; ++j) k.x[k.perm[j]] = k.y[j];
  }
  float e2 = 0.f, b2 = 0.f, a2 = 0.f;
  for (int i = 0; i < n; ++i) {
    float ax = 0.f;
    for (int j = 0; j < n; ++j) ax += A[i + n * j] * k.x[j];
    e2 += (b[i] - ax) * (b[i] - ax); b2 += b[i] * b[i]; a2 += ax * ax;
  }
  if (!(e2 <= 1e-5f * 1e-5f * fminf(b2, a2))) {
    svd_solve_rt(A, b, n, k);
    *info = 2;
  }
}

// number of equality constraints = non-localizable directions (PointToPlane.cpp:468-471)
__device__ inline int xicp_num_constraints(const XicpState& X) {
  int c = 0;
  for (int i = 0; i < 3; ++i) c += (X.loc_rpy[i] == 0) + (X.loc_xyz[i] == 0);
  return c;
}

// builds and solves the augmented system; x6 receives the first six unknowns (rotation, translation)
static __device__ __noinline__ void xicp_solve(const float* H, const float* b, const XicpState& X, float* x6, int* info,
                                               KktScratch& k) {
  int idx[6], c = 0;
  for (int i = 0; i < 3; ++i) if (X.loc_rpy[i] == 0) idx[c++] = i;          // readLocalizabilityFlags: rotation first
  for (int i = 0; i < 3; ++i) if (X.loc_xyz[i] == 0) idx[c++] = i + 3;
  const int n = 6 + c;
  for (int i = 0; i < n * n; ++i) k.A[i] = 0.f;
  for (int i = 0; i < n; ++i) k.b[i] = 0.f;
  for (int j = 0; j < 6; ++j)
    for (int i = 0; i < 6; ++i) k.A[i + n * j] = H[i + 6 * j];
  for (int i = 0; i < 6; ++i) k.b[i] = b[i];
  for (int cc = 0; cc < c; ++cc) {
    const int index = idx[cc];
    if (index < 3) {
      for (int i = 0; i < 3; ++i) { k.A[i + n * (6 + cc)] = X.Ur[i + 3 * index]; k.A[(6 + cc) + n * i] = X.Ur[i + 3 * index]; }
      k.b[6 + cc] = X.cR[index];
    } else {
      for (int i = 0; i < 3; ++i) {
        k.A[(i + 3) + n * (6 + cc)] = X.Ut[i + 3 * (index - 3)];
        k.A[(6 + cc) + n * (i + 3)] = X.Ut[i + 3 * (index - 3)];
      }
      k.b[6 + cc] = X.cT[index - 3];
    }
  }
  solve_rt_dev(n, k, info);
  for (int i = 0; i < 6; ++i) x6[i] = k.x[i];
}

}  // namespace b200
