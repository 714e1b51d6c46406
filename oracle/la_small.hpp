// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product; only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load anything under oracle/.
//
// Small dense linear algebra restating the Eigen routines the reference path calls (Eigen itself is absent
// from this container):
//   * FullPivHouseholderQR rank / isInvertible / Q / permutation  — used by
//       libpointmatcher/pointmatcher/ErrorMinimizers/PointToPlane.cpp:206-247 (6x6, float) and
//       libpointmatcher/pointmatcher/DataPointsFilters/SurfaceNormal.cpp:191 (3x3, float)
//   * JacobiSVD<double>::solve                                    — PointToPlane.cpp:236,255
//   * symmetric eigen-decomposition (stands in for the non-symmetric EigenSolver on a symmetric 3x3,
//       SurfaceNormal.cpp:193-195; eigenvector sign/order are artefacts, see DESIGN.md "parity unpinned")
//   * JacobiSVD<float> of a 3x3                                   — ErrorMinimizers/PointToPoint.cpp:86
// Matrices are column-major like Eigen: a(i,j) = a[i + j*n].
#pragma once
#include <cmath>
#include <cfloat>
#include <algorithm>
#include <limits>

namespace orc {

// ---------------------------------------------------------------------------------------------
// Full-pivoting Householder QR of an n x n matrix (Eigen::FullPivHouseholderQR::computeInPlace).
// ---------------------------------------------------------------------------------------------
template <typename T, int N>
struct FullPivQR {
  T qr[N * N];      // Householder vectors below the diagonal, R on and above
  T hcoef[N];       // Householder coefficients (tau)
  int rowTransp[N]; // row transpositions, step k swaps rows k <-> rowTransp[k]
  int colTransp[N];
  int colPerm[N];   // colsPermutation(): A * P = Q * R, column j of (A*P) is column colPerm[j] of A
  int nonzeroPivots;
  T maxPivot;

  T& at(int i, int j) { return qr[i + j * N]; }
  T at(int i, int j) const { return qr[i + j * N]; }

  void compute(const T* A) {
    for (int i = 0; i < N * N; ++i) qr[i] = A[i];
    const T precision = std::numeric_limits<T>::epsilon() * T(N);
    nonzeroPivots = N;
    maxPivot = T(0);
    T biggest = T(0);
    int numTransp = 0;
    for (int k = 0; k < N; ++k) {
      // biggest |coefficient| in the bottom-right corner; Eigen's maxCoeff visits column by column and
      // keeps the first maximum
      int rb = k, cb = k;
      T big = T(-1);
      for (int j = k; j < N; ++j)
        for (int i = k; i < N; ++i) {
          const T v = std::fabs(at(i, j));
          if (v > big) { big = v; rb = i; cb = j; }
        }
      if (k == 0) biggest = big;
      if (std::fabs(big) <= std::fabs(biggest) * precision) {  // isMuchSmallerThan
        nonzeroPivots = k;
        for (int i = k; i < N; ++i) { rowTransp[i] = i; colTransp[i] = i; hcoef[i] = T(0); }
        break;
      }
      rowTransp[k] = rb;
      colTransp[k] = cb;
      if (k != rb) { for (int j = k; j < N; ++j) std::swap(at(k, j), at(rb, j)); ++numTransp; }  // tail only, as Eigen
      if (k != cb) { for (int i = 0; i < N; ++i) std::swap(at(i, k), at(i, cb)); ++numTransp; }
      // makeHouseholderInPlace on column k, rows k..N-1
      T tailSq = T(0);
      for (int i = k + 1; i < N; ++i) tailSq += at(i, k) * at(i, k);
      const T c0 = at(k, k);
      T tau, beta;
      if (tailSq <= std::numeric_limits<T>::min()) {
        tau = T(0);
        beta = c0;
        for (int i = k + 1; i < N; ++i) at(i, k) = T(0);
      } else {
        beta = std::sqrt(c0 * c0 + tailSq);
        if (c0 >= T(0)) beta = -beta;
        for (int i = k + 1; i < N; ++i) at(i, k) = at(i, k) / (c0 - beta);
        tau = (beta - c0) / beta;
      }
      at(k, k) = beta;
      hcoef[k] = tau;
      if (std::fabs(beta) > maxPivot) maxPivot = std::fabs(beta);
      // apply H = I - tau v v^T (v = [1; essential]) to the trailing columns
      if (N - k == 1) {
        // single row: nothing to the right
      } else if (tau != T(0)) {
        for (int j = k + 1; j < N; ++j) {
          T tmp = T(0);
          for (int i = k + 1; i < N; ++i) tmp += at(i, k) * at(i, j);
          tmp += at(k, j);
          at(k, j) -= tau * tmp;
          for (int i = k + 1; i < N; ++i) at(i, j) -= tau * at(i, k) * tmp;
        }
      }
    }
    // colsPermutation: identity, then applyTranspositionOnTheRight(k, colTransp[k]) for k = 0..N-1
    for (int i = 0; i < N; ++i) colPerm[i] = i;
    for (int k = 0; k < N; ++k) std::swap(colPerm[k], colPerm[colTransp[k]]);
    (void)numTransp;
  }

  T threshold() const { return std::numeric_limits<T>::epsilon() * T(N); }

  int rank() const {
    const T pre = std::fabs(maxPivot) * threshold();
    int r = 0;
    for (int i = 0; i < nonzeroPivots; ++i) r += (std::fabs(at(i, i)) > pre) ? 1 : 0;
    return r;
  }
  bool isInvertible() const { return rank() == N; }

  // Q (N x N, column-major) such that A * P = Q * R; includes the row transpositions exactly like
  // Eigen's matrixQ(): start from identity, for k = N-1..0 apply H_k on the left to the trailing block,
  // then swap rows k <-> rowTransp[k].
  void matrixQ(T* Q) const {
    for (int i = 0; i < N * N; ++i) Q[i] = T(0);
    for (int i = 0; i < N; ++i) Q[i + i * N] = T(1);
    for (int k = N - 1; k >= 0; --k) {
      const T tau = hcoef[k];
      if (tau != T(0) && N - k > 1) {
        for (int j = k; j < N; ++j) {
          T tmp = Q[k + j * N];
          for (int i = k + 1; i < N; ++i) tmp += at(i, k) * Q[i + j * N];
          Q[k + j * N] -= tau * tmp;
          for (int i = k + 1; i < N; ++i) Q[i + j * N] -= tau * at(i, k) * tmp;
        }
      } else if (tau != T(0)) {
        for (int j = k; j < N; ++j) Q[k + j * N] *= (T(1) - tau);
      }
      const int r = rowTransp[k];
      if (r != k)
        for (int j = 0; j < N; ++j) std::swap(Q[k + j * N], Q[r + j * N]);
    }
  }
};

// ---------------------------------------------------------------------------------------------
// One-sided Jacobi (Hestenes) SVD of an n x n matrix, A = U diag(s) V^T. Plays JacobiSVD.
// ---------------------------------------------------------------------------------------------
template <typename T, int N>
void jacobiSVD(const T* A, T* U, T* S, T* V) {
  T W[N * N];
  for (int i = 0; i < N * N; ++i) W[i] = A[i];
  for (int i = 0; i < N * N; ++i) V[i] = T(0);
  for (int i = 0; i < N; ++i) V[i + i * N] = T(1);
  const T eps = std::numeric_limits<T>::epsilon();
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rotated = false;
    for (int p = 0; p < N - 1; ++p)
      for (int q = p + 1; q < N; ++q) {
        T alpha = 0, beta = 0, gamma = 0;
        for (int i = 0; i < N; ++i) {
          alpha += W[i + p * N] * W[i + p * N];
          beta += W[i + q * N] * W[i + q * N];
          gamma += W[i + p * N] * W[i + q * N];
        }
        if (gamma == T(0) || std::fabs(gamma) <= eps * std::sqrt(alpha * beta)) continue;
        rotated = true;
        const T zeta = (beta - alpha) / (T(2) * gamma);
        const T t = (zeta >= 0 ? T(1) : T(-1)) / (std::fabs(zeta) + std::sqrt(T(1) + zeta * zeta));
        const T c = T(1) / std::sqrt(T(1) + t * t), s = c * t;
        for (int i = 0; i < N; ++i) {
          const T wp = W[i + p * N], wq = W[i + q * N];
          W[i + p * N] = c * wp - s * wq;
          W[i + q * N] = s * wp + c * wq;
          const T vp = V[i + p * N], vq = V[i + q * N];
          V[i + p * N] = c * vp - s * vq;
          V[i + q * N] = s * vp + c * vq;
        }
      }
    if (!rotated) break;
  }
  // singular values = column norms; sort descending like Eigen
  int order[N];
  T sv[N];
  for (int j = 0; j < N; ++j) {
    T n2 = 0;
    for (int i = 0; i < N; ++i) n2 += W[i + j * N] * W[i + j * N];
    sv[j] = std::sqrt(n2);
    order[j] = j;
  }
  std::stable_sort(order, order + N, [&](int a, int b) { return sv[a] > sv[b]; });
  T Vs[N * N];
  for (int jj = 0; jj < N; ++jj) {
    const int j = order[jj];
    S[jj] = sv[j];
    for (int i = 0; i < N; ++i) {
      U[i + jj * N] = sv[j] > T(0) ? W[i + j * N] / sv[j] : T(0);
      Vs[i + jj * N] = V[i + j * N];
    }
  }
  for (int i = 0; i < N * N; ++i) V[i] = Vs[i];
}

// JacobiSVD::solve — pseudo-inverse with Eigen's rank threshold (diagSize * epsilon, relative to the
// largest singular value).
template <typename T, int N>
void svdSolve(const T* A, const T* b, T* x) {
  T U[N * N], S[N], V[N * N];
  jacobiSVD<T, N>(A, U, S, V);
  const T thr = std::max(S[0] * T(N) * std::numeric_limits<T>::epsilon(), std::numeric_limits<T>::min());
  T y[N];
  for (int j = 0; j < N; ++j) {
    T d = 0;
    for (int i = 0; i < N; ++i) d += U[i + j * N] * b[i];
    y[j] = (S[j] > thr) ? d / S[j] : T(0);
  }
  for (int i = 0; i < N; ++i) {
    T acc = 0;
    for (int j = 0; j < N; ++j) acc += V[i + j * N] * y[j];
    x[i] = acc;
  }
}

// ---------------------------------------------------------------------------------------------
// Cyclic Jacobi eigen-decomposition of a symmetric n x n matrix (uses the upper triangle averaged with
// the lower). Eigenvalues unsorted in `w`, eigenvectors in the columns of `Vec`.
// ---------------------------------------------------------------------------------------------
template <typename T, int N>
void symEigen(const T* Ain, T* w, T* Vec) {
  T A[N * N];
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) A[i + j * N] = T(0.5) * (Ain[i + j * N] + Ain[j + i * N]);
  for (int i = 0; i < N * N; ++i) Vec[i] = T(0);
  for (int i = 0; i < N; ++i) Vec[i + i * N] = T(1);
  for (int sweep = 0; sweep < 64; ++sweep) {
    T off = 0, diag = 0;
    for (int i = 0; i < N; ++i)
      for (int j = 0; j < N; ++j) (i == j ? diag : off) += A[i + j * N] * A[i + j * N];
    if (off <= std::numeric_limits<T>::epsilon() * std::numeric_limits<T>::epsilon() * diag || off == T(0)) break;
    for (int p = 0; p < N - 1; ++p)
      for (int q = p + 1; q < N; ++q) {
        const T apq = A[p + q * N];
        if (apq == T(0)) continue;
        const T theta = (A[q + q * N] - A[p + p * N]) / (T(2) * apq);
        const T t = (theta >= 0 ? T(1) : T(-1)) / (std::fabs(theta) + std::sqrt(theta * theta + T(1)));
        const T c = T(1) / std::sqrt(t * t + T(1)), s = t * c;
        for (int k = 0; k < N; ++k) {  // A <- A * J
          const T akp = A[k + p * N], akq = A[k + q * N];
          A[k + p * N] = c * akp - s * akq;
          A[k + q * N] = s * akp + c * akq;
        }
        for (int k = 0; k < N; ++k) {  // A <- J^T * A
          const T apk = A[p + k * N], aqk = A[q + k * N];
          A[p + k * N] = c * apk - s * aqk;
          A[q + k * N] = s * apk + c * aqk;
        }
        for (int k = 0; k < N; ++k) {
          const T vkp = Vec[k + p * N], vkq = Vec[k + q * N];
          Vec[k + p * N] = c * vkp - s * vkq;
          Vec[k + q * N] = s * vkp + c * vkq;
        }
      }
  }
  for (int i = 0; i < N; ++i) w[i] = A[i + i * N];
}

template <typename T>
inline T det3(const T* m) {  // column-major 3x3
  return m[0] * (m[4] * m[8] - m[7] * m[5]) - m[3] * (m[1] * m[8] - m[7] * m[2]) + m[6] * (m[1] * m[5] - m[4] * m[2]);
}

}  // namespace orc
