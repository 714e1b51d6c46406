// Small fixed-size dense linear algebra for the on-device solve stage (one thread per problem).
// Column-major like Eigen: a(i,j) = a[i + j*N].
//
// Reference routines being replaced:
//   Eigen::FullPivHouseholderQR rank / isInvertible   — PointToPlane.cpp:202-203 (6x6), SurfaceNormal.cpp:191 (3x3)
//   JacobiSVD<double>::solve                           — PointToPlane.cpp:236, 255
//   minimal-norm solve of the rank-deficient system    — PointToPlane.cpp:205-231
//   EigenSolver on the symmetric 3x3 scatter matrix    — SurfaceNormal.cpp:193-195
//   JacobiSVD<float> of the 3x3 cross-covariance       — PointToPoint.cpp:86
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <float.h>

namespace b200 {
namespace la {

template <typename T> struct Eps;
template <> struct Eps<float> { __host__ __device__ static float eps() { return FLT_EPSILON; } __host__ __device__ static float tiny() { return FLT_MIN; } };
template <> struct Eps<double> { __host__ __device__ static double eps() { return DBL_EPSILON; } __host__ __device__ static double tiny() { return DBL_MIN; } };

// Full-pivoting Householder QR, following Eigen's computeInPlace step by step so that the rank decision on
// borderline matrices matches the reference's.
template <typename T, int N>
struct PivQR {
  T qr[N * N];
  T tau[N];
  int rowT[N], colT[N], perm[N];
  int nonzero;
  T maxpivot;

  __host__ __device__ T& at(int i, int j) { return qr[i + j * N]; }
  __host__ __device__ T at(int i, int j) const { return qr[i + j * N]; }

  __host__ __device__ void compute(const T* A) {
    for (int i = 0; i < N * N; ++i) qr[i] = A[i];
    const T prec = Eps<T>::eps() * T(N);
    nonzero = N;
    maxpivot = T(0);
    T biggest = T(0);
    for (int k = 0; k < N; ++k) {
      int rb = k, cb = k;
      T big = T(-1);
      for (int j = k; j < N; ++j)
        for (int i = k; i < N; ++i) {
          const T v = fabs(at(i, j));
          if (v > big) { big = v; rb = i; cb = j; }
        }
      if (k == 0) biggest = big;
      if (fabs(big) <= fabs(biggest) * prec) {
        nonzero = k;
        for (int i = k; i < N; ++i) { rowT[i] = i; colT[i] = i; tau[i] = T(0); }
        break;
      }
      rowT[k] = rb; colT[k] = cb;
      if (rb != k)
        for (int j = k; j < N; ++j) { const T t = at(k, j); at(k, j) = at(rb, j); at(rb, j) = t; }  // tail only
      if (cb != k)
        for (int i = 0; i < N; ++i) { const T t = at(i, k); at(i, k) = at(i, cb); at(i, cb) = t; }
      T tail = T(0);
      for (int i = k + 1; i < N; ++i) tail += at(i, k) * at(i, k);
      const T c0 = at(k, k);
      T beta, tk;
      if (tail <= Eps<T>::tiny()) {
        tk = T(0); beta = c0;
        for (int i = k + 1; i < N; ++i) at(i, k) = T(0);
      } else {
        beta = sqrt(c0 * c0 + tail);
        if (c0 >= T(0)) beta = -beta;
        for (int i = k + 1; i < N; ++i) at(i, k) = at(i, k) / (c0 - beta);
        tk = (beta - c0) / beta;
      }
      at(k, k) = beta;
      tau[k] = tk;
      if (fabs(beta) > maxpivot) maxpivot = fabs(beta);
      if (N - k > 1 && tk != T(0)) {
        for (int j = k + 1; j < N; ++j) {
          T tmp = T(0);
          for (int i = k + 1; i < N; ++i) tmp += at(i, k) * at(i, j);
          tmp += at(k, j);
          at(k, j) -= tk * tmp;
          for (int i = k + 1; i < N; ++i) at(i, j) -= tk * at(i, k) * tmp;
        }
      }
    }
    for (int i = 0; i < N; ++i) perm[i] = i;
    for (int k = 0; k < N; ++k) { const int t = perm[k]; perm[k] = perm[colT[k]]; perm[colT[k]] = t; }
  }

  __host__ __device__ int rank() const {
    const T pre = fabs(maxpivot) * (Eps<T>::eps() * T(N));
    int r = 0;
    for (int i = 0; i < nonzero; ++i) r += (fabs(at(i, i)) > pre) ? 1 : 0;
    return r;
  }

  // Q with A * P = Q * R, row transpositions folded in (Eigen matrixQ())
  __host__ __device__ void matrixQ(T* Q) const {
    for (int i = 0; i < N * N; ++i) Q[i] = T(0);
    for (int i = 0; i < N; ++i) Q[i + i * N] = T(1);
    for (int k = N - 1; k >= 0; --k) {
      const T tk = tau[k];
      if (tk != T(0)) {
        if (N - k > 1) {
          for (int j = k; j < N; ++j) {
            T tmp = Q[k + j * N];
            for (int i = k + 1; i < N; ++i) tmp += at(i, k) * Q[i + j * N];
            Q[k + j * N] -= tk * tmp;
            for (int i = k + 1; i < N; ++i) Q[i + j * N] -= tk * at(i, k) * tmp;
          }
        } else {
          for (int j = k; j < N; ++j) Q[k + j * N] *= (T(1) - tk);
        }
      }
      const int r = rowT[k];
      if (r != k)
        for (int j = 0; j < N; ++j) { const T t = Q[k + j * N]; Q[k + j * N] = Q[r + j * N]; Q[r + j * N] = t; }
    }
  }
};

// Rank decision of PivQR<T, N> (compute() + rank()) with every loop unrolled and every pivot swap written as a
// predicated swap over compile-time indices, so the N x N working matrix lives in registers (PivQR's dynamic
// row/column indices push it to local memory, ~10x slower for one thread). Same operations in the same order =>
// bit-identical pivots and the same rank as PivQR.
template <typename T, int N>
__device__ __forceinline__ int pivqr_rank(const T* A) {
  T q[N * N];
#pragma unroll
  for (int i = 0; i < N * N; ++i) q[i] = A[i];
  const T prec = Eps<T>::eps() * T(N);
  int nonzero = N;
  bool stopped = false;
  T maxpivot = T(0), biggest = T(0);
  T diag[N];
#pragma unroll
  for (int k = 0; k < N; ++k) {
    diag[k] = T(0);
    if (!stopped) {
      int rb = k, cb = k;
      T big = T(-1);
#pragma unroll
      for (int j = k; j < N; ++j)
#pragma unroll
        for (int i = k; i < N; ++i) {
          const T v = fabs(q[i + j * N]);
          if (v > big) { big = v; rb = i; cb = j; }
        }
      if (k == 0) biggest = big;
      if (fabs(big) <= fabs(biggest) * prec) {
        nonzero = k;
        stopped = true;
      } else {
#pragma unroll
        for (int r = k + 1; r < N; ++r)
          if (r == rb) {
#pragma unroll
            for (int j = k; j < N; ++j) { const T t = q[k + j * N]; q[k + j * N] = q[r + j * N]; q[r + j * N] = t; }
          }
#pragma unroll
        for (int c = k + 1; c < N; ++c)
          if (c == cb) {
#pragma unroll
            for (int i = 0; i < N; ++i) { const T t = q[i + k * N]; q[i + k * N] = q[i + c * N]; q[i + c * N] = t; }
          }
        T tail = T(0);
#pragma unroll
        for (int i = k + 1; i < N; ++i) tail += q[i + k * N] * q[i + k * N];
        const T c0 = q[k + k * N];
        T beta, tk;
        if (tail <= Eps<T>::tiny()) {
          tk = T(0); beta = c0;
#pragma unroll
          for (int i = k + 1; i < N; ++i) q[i + k * N] = T(0);
        } else {
          beta = sqrt(c0 * c0 + tail);
          if (c0 >= T(0)) beta = -beta;
#pragma unroll
          for (int i = k + 1; i < N; ++i) q[i + k * N] = q[i + k * N] / (c0 - beta);
          tk = (beta - c0) / beta;
        }
        q[k + k * N] = beta;
        diag[k] = beta;
        if (fabs(beta) > maxpivot) maxpivot = fabs(beta);
        if (N - k > 1 && tk != T(0)) {
#pragma unroll
          for (int j = k + 1; j < N; ++j) {
            T tmp = T(0);
#pragma unroll
            for (int i = k + 1; i < N; ++i) tmp += q[i + k * N] * q[i + j * N];
            tmp += q[k + j * N];
            q[k + j * N] -= tk * tmp;
#pragma unroll
            for (int i = k + 1; i < N; ++i) q[i + j * N] -= tk * q[i + k * N] * tmp;
          }
        }
      }
    }
  }
  const T pre = fabs(maxpivot) * (Eps<T>::eps() * T(N));
  int r = 0;
#pragma unroll
  for (int i = 0; i < N; ++i) r += (i < nonzero && fabs(diag[i]) > pre) ? 1 : 0;
  return r;
}

// Cholesky solve of a symmetric positive definite system in double. Returns false on a non-positive pivot.
template <int N>
__host__ __device__ bool cholSolve(const double* A, const double* b, double* x) {
  double L[N * N];
#pragma unroll
  for (int j = 0; j < N; ++j) {
    double d = A[j + j * N];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= L[j + k * N] * L[j + k * N];
    if (!(d > 0.0)) return false;
    d = sqrt(d);
    L[j + j * N] = d;
    const double inv = 1.0 / d;
#pragma unroll
    for (int i = j + 1; i < N; ++i) {
      double s = A[i + j * N];
#pragma unroll
      for (int k = 0; k < j; ++k) s -= L[i + k * N] * L[j + k * N];
      L[i + j * N] = s * inv;
    }
  }
  double y[N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double s = b[i];
#pragma unroll
    for (int k = 0; k < i; ++k) s -= L[i + k * N] * y[k];
    y[i] = s / L[i + i * N];
  }
#pragma unroll
  for (int i = N - 1; i >= 0; --i) {
    double s = y[i];
#pragma unroll
    for (int k = i + 1; k < N; ++k) s -= L[k + i * N] * x[k];
    x[i] = s / L[i + i * N];
  }
  return true;
}

// One-sided Jacobi SVD, A = U diag(S) V^T (unsorted), and the pseudo-inverse solve with Eigen's default threshold.
template <typename T, int N>
__host__ __device__ void svdSolve(const T* A, const T* b, T* x) {
  T W[N * N], V[N * N];
  for (int i = 0; i < N * N; ++i) { W[i] = A[i]; V[i] = T(0); }
  for (int i = 0; i < N; ++i) V[i + i * N] = T(1);
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rot = false;
    for (int p = 0; p < N - 1; ++p)
      for (int q = p + 1; q < N; ++q) {
        T al = 0, be = 0, ga = 0;
        for (int i = 0; i < N; ++i) {
          al += W[i + p * N] * W[i + p * N];
          be += W[i + q * N] * W[i + q * N];
          ga += W[i + p * N] * W[i + q * N];
        }
        if (ga == T(0) || fabs(ga) <= Eps<T>::eps() * sqrt(al * be)) continue;
        rot = true;
        const T zeta = (be - al) / (T(2) * ga);
        const T t = (zeta >= 0 ? T(1) : T(-1)) / (fabs(zeta) + sqrt(T(1) + zeta * zeta));
        const T c = T(1) / sqrt(T(1) + t * t), s = c * t;
        for (int i = 0; i < N; ++i) {
          const T wp = W[i + p * N], wq = W[i + q * N];
          W[i + p * N] = c * wp - s * wq; W[i + q * N] = s * wp + c * wq;
          const T vp = V[i + p * N], vq = V[i + q * N];
          V[i + p * N] = c * vp - s * vq; V[i + q * N] = s * vp + c * vq;
        }
      }
    if (!rot) break;
  }
  T S[N], smax = T(0);
  for (int j = 0; j < N; ++j) {
    T n2 = 0;
    for (int i = 0; i < N; ++i) n2 += W[i + j * N] * W[i + j * N];
    S[j] = sqrt(n2);
    if (S[j] > smax) smax = S[j];
  }
  T thr = smax * T(N) * Eps<T>::eps();
  if (thr < Eps<T>::tiny()) thr = Eps<T>::tiny();
  for (int i = 0; i < N; ++i) x[i] = T(0);
  for (int j = 0; j < N; ++j) {
    if (!(S[j] > thr)) continue;
    T d = 0;  // u_j . b with u_j = W_j / S_j
    for (int i = 0; i < N; ++i) d += W[i + j * N] * b[i];
    d = d / (S[j] * S[j]);
    for (int i = 0; i < N; ++i) x[i] += V[i + j * N] * d;
  }
}

// Cyclic Jacobi eigen-decomposition of a symmetric 3x3 in double; w unsorted, eigenvectors in columns of V.
__host__ __device__ inline void symEigen3(const double* Ain, double* w, double* V) {
  double A[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) A[i + 3 * j] = 0.5 * (Ain[i + 3 * j] + Ain[j + 3 * i]);
  for (int i = 0; i < 9; ++i) V[i] = 0.0;
  V[0] = V[4] = V[8] = 1.0;
  for (int sweep = 0; sweep < 64; ++sweep) {
    double off = 0, diag = 0;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        const double v = A[i + 3 * j] * A[i + 3 * j];
        if (i == j) diag += v; else off += v;
      }
    if (off == 0.0 || off <= DBL_EPSILON * DBL_EPSILON * diag) break;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        const double apq = A[p + 3 * q];
        if (apq == 0.0) continue;
        const double theta = (A[q + 3 * q] - A[p + 3 * p]) / (2.0 * apq);
        const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < 3; ++k) {
          const double akp = A[k + 3 * p], akq = A[k + 3 * q];
          A[k + 3 * p] = c * akp - s * akq; A[k + 3 * q] = s * akp + c * akq;
        }
        for (int k = 0; k < 3; ++k) {
          const double apk = A[p + 3 * k], aqk = A[q + 3 * k];
          A[p + 3 * k] = c * apk - s * aqk; A[q + 3 * k] = s * apk + c * aqk;
        }
        for (int k = 0; k < 3; ++k) {
          const double vkp = V[k + 3 * p], vkq = V[k + 3 * q];
          V[k + 3 * p] = c * vkp - s * vkq; V[k + 3 * q] = s * vkp + c * vkq;
        }
      }
  }
  w[0] = A[0]; w[1] = A[4]; w[2] = A[8];
}

template <typename T>
__host__ __device__ inline T det3(const T* m) {
  return m[0] * (m[4] * m[8] - m[7] * m[5]) - m[3] * (m[1] * m[8] - m[7] * m[2]) + m[6] * (m[1] * m[5] - m[4] * m[2]);
}

}  // namespace la
}  // namespace b200
