// On-device solve stage of one ICP iteration: 6x6 solve, SE(3) construction, pose composition, convergence checks.
// Runs in one thread (negligible work); shared by the modular kernels and the persistent ICP kernel.
#pragma once
#include "small_la.cuh"

namespace b200 {

// ---- 4x4 helpers (column-major, depth-sequential accumulation without FMA) --------------------
__host__ __device__ inline void eye4(float* T) {
  for (int i = 0; i < 16; ++i) T[i] = 0.f;
  T[0] = T[5] = T[10] = T[15] = 1.f;
}
__device__ inline void mul4_dev(const float* A, const float* B, float* C) {
  float R[16];
  for (int j = 0; j < 4; ++j)
    for (int i = 0; i < 4; ++i) {
      float acc = __fmul_rn(A[i], B[4 * j]);
      acc = __fadd_rn(acc, __fmul_rn(A[i + 4], B[1 + 4 * j]));
      acc = __fadd_rn(acc, __fmul_rn(A[i + 8], B[2 + 4 * j]));
      acc = __fadd_rn(acc, __fmul_rn(A[i + 12], B[3 + 4 * j]));
      R[i + 4 * j] = acc;
    }
  for (int i = 0; i < 16; ++i) C[i] = R[i];
}

__device__ inline void normalized3_dev(const float* v, float* o) {
  float z = __fmul_rn(v[0], v[0]);
  z = __fadd_rn(z, __fmul_rn(v[1], v[1]));
  z = __fadd_rn(z, __fmul_rn(v[2], v[2]));
  if (z > 0.f) {
    const float s = sqrtf(z);
    o[0] = v[0] / s; o[1] = v[1] / s; o[2] = v[2] / s;
  } else {
    o[0] = v[0]; o[1] = v[1]; o[2] = v[2];
  }
}

// RigidTransformation::checkParameters / correctParameters — TransformationsImpl.cpp:105-170.
// Tc = matrix applied to the features; R9 = rotation applied to normals (the UNcorrected block, :72,98).
__device__ inline void rigid_prepare(const float* T, float* Tc, float* R9) {
  for (int i = 0; i < 16; ++i) Tc[i] = T[i];
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) R9[i + 3 * j] = T[i + 4 * j];
  float det = R9[0] * (R9[4] * R9[8] - R9[7] * R9[5]) - R9[3] * (R9[1] * R9[8] - R9[7] * R9[2]) +
              R9[6] * (R9[1] * R9[5] - R9[4] * R9[2]);
  if (fabsf(1.f - det) > 0.001f) {
    float c1[3], c2[3], n0[3], n1[3];
    normalized3_dev(&T[4], c1);
    normalized3_dev(&T[8], c2);
    n0[0] = c1[1] * c2[2] - c1[2] * c2[1]; n0[1] = c1[2] * c2[0] - c1[0] * c2[2]; n0[2] = c1[0] * c2[1] - c1[1] * c2[0];
    n1[0] = c2[1] * n0[2] - c2[2] * n0[1]; n1[1] = c2[2] * n0[0] - c2[0] * n0[2]; n1[2] = c2[0] * n0[1] - c2[1] * n0[0];
    for (int i = 0; i < 3; ++i) { Tc[i] = n0[i]; Tc[4 + i] = n1[i]; Tc[8 + i] = c2[i]; }
  }
}

// features <- T * features for one point: acc = T(i,0)x; += T(i,1)y; += T(i,2)z; += T(i,3)w (no FMA)
__device__ __forceinline__ float4 apply_T(const float* T, float4 p) {
  float4 r;
  float a;
  a = __fmul_rn(T[0], p.x); a = __fadd_rn(a, __fmul_rn(T[4], p.y)); a = __fadd_rn(a, __fmul_rn(T[8], p.z)); r.x = __fadd_rn(a, __fmul_rn(T[12], p.w));
  a = __fmul_rn(T[1], p.x); a = __fadd_rn(a, __fmul_rn(T[5], p.y)); a = __fadd_rn(a, __fmul_rn(T[9], p.z)); r.y = __fadd_rn(a, __fmul_rn(T[13], p.w));
  a = __fmul_rn(T[2], p.x); a = __fadd_rn(a, __fmul_rn(T[6], p.y)); a = __fadd_rn(a, __fmul_rn(T[10], p.z)); r.z = __fadd_rn(a, __fmul_rn(T[14], p.w));
  a = __fmul_rn(T[3], p.x); a = __fadd_rn(a, __fmul_rn(T[7], p.y)); a = __fadd_rn(a, __fmul_rn(T[11], p.z)); r.w = __fadd_rn(a, __fmul_rn(T[15], p.w));
  return r;
}
__device__ __forceinline__ float3 apply_R(const float* R9, float3 n) {
  float3 r;
  float a;
  a = __fmul_rn(R9[0], n.x); a = __fadd_rn(a, __fmul_rn(R9[3], n.y)); r.x = __fadd_rn(a, __fmul_rn(R9[6], n.z));
  a = __fmul_rn(R9[1], n.x); a = __fadd_rn(a, __fmul_rn(R9[4], n.y)); r.y = __fadd_rn(a, __fmul_rn(R9[7], n.z));
  a = __fmul_rn(R9[2], n.x); a = __fadd_rn(a, __fmul_rn(R9[5], n.y)); r.z = __fadd_rn(a, __fmul_rn(R9[8], n.z));
  return r;
}

// solvePossiblyUnderdeterminedLinearSystem(A, b, x, false) — PointToPlane.cpp:199-256.
// info: 0 full rank, 1 rank-deficient minimal-norm, 2 double-SVD fallback.
// x_spd (nullable): A^-1 b from a double LDL^T of the symmetrised A whose pivots are well conditioned (solve6_warp
// returned 2). It stands in for the reference's double-precision JacobiSVD fallback: for a symmetric positive definite
// matrix with cond <= 1e9 both are backward-stable solves of the same system and agree to cond * eps_double <= 2e-7
// relative, below the rounding of the cast to float that follows. (A scan far from the map's centroid makes the
// rotation block of H dwarf the translation block, the reference's FLOAT full-pivot QR then calls the matrix rank
// deficient, its minimal-norm solution fails the isApprox test and every iteration ends in that SVD: measured 110 us
// per iteration as a serial double Jacobi SVD, on 5 of the 32 benchmark poses.)
static __device__ __noinline__ void solve6_dev(const float* A, const float* b, float* x, int* info,
                                               const float* x_spd = nullptr) {
  // x_spd given: the warp replayed the float QR and found rank < 6, so go straight to the full factorisation (whose
  // rank() is the same decision) instead of running the register-resident rank pass first
  la::PivQR<float, 6> qr;
  int rank;
  if (x_spd) { qr.compute(A); rank = qr.rank(); }
  else rank = la::pivqr_rank<float, 6>(A);
  double Ad[36], bd[6], xd[6];
  for (int i = 0; i < 36; ++i) Ad[i] = (double)A[i];
  for (int i = 0; i < 6; ++i) bd[i] = (double)b[i];
  if (rank == 6) {
    // the reference solves with a double-precision SVD (least squares); for a numerically full-rank matrix this
    // is A^-1 b. Cholesky needs exact symmetry: symmetrise, and fall back to the SVD if a pivot fails.
    double As[36];
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 6; ++j) As[i + 6 * j] = 0.5 * (Ad[i + 6 * j] + Ad[j + 6 * i]);
    bool sym = true;
    for (int i = 0; i < 6 && sym; ++i)
      for (int j = 0; j < i; ++j)
        if (fabs(Ad[i + 6 * j] - Ad[j + 6 * i]) > 1e-6 * (fabs(Ad[i + 6 * i]) + fabs(Ad[j + 6 * j]))) { sym = false; break; }
    if (!(sym && la::cholSolve<6>(As, bd, xd))) la::svdSolve<double, 6>(Ad, bd, xd);
    for (int i = 0; i < 6; ++i) x[i] = (float)xd[i];
    *info = 0;
    return;
  }
  *info = 1;
  if (!x_spd) qr.compute(A);
  if (rank == 0) {
    for (int i = 0; i < 6; ++i) x[i] = 0.f;
  } else {
    float Q[36], AP[36], R1[36], qtb[6], G[36], L[36], z[6], y[6];
    qr.matrixQ(Q);
    for (int j = 0; j < 6; ++j)
      for (int i = 0; i < 6; ++i) AP[i + 6 * j] = A[i + 6 * qr.perm[j]];
    for (int r = 0; r < rank; ++r) {
      for (int j = 0; j < 6; ++j) {
        float acc = 0.f;
        for (int i = 0; i < 6; ++i) acc += Q[i + 6 * r] * AP[i + 6 * j];
        R1[r + 6 * j] = acc;
      }
      float acc = 0.f;
      for (int i = 0; i < 6; ++i) acc += Q[i + 6 * r] * b[i];
      qtb[r] = acc;
    }
    for (int a = 0; a < rank; ++a)
      for (int c = 0; c < rank; ++c) {
        float acc = 0.f;
        for (int j = 0; j < 6; ++j) acc += R1[a + 6 * j] * R1[c + 6 * j];
        G[a + 6 * c] = acc;
      }
    for (int i = 0; i < 36; ++i) L[i] = 0.f;
    for (int j = 0; j < rank; ++j) {
      float d = G[j + 6 * j];
      for (int k = 0; k < j; ++k) d -= L[j + 6 * k] * L[j + 6 * k];
      d = sqrtf(d);
      L[j + 6 * j] = d;
      for (int i = j + 1; i < rank; ++i) {
        float s = G[i + 6 * j];
        for (int k = 0; k < j; ++k) s -= L[i + 6 * k] * L[j + 6 * k];
        L[i + 6 * j] = s / d;
      }
    }
    for (int i = 0; i < rank; ++i) {
      float s = qtb[i];
      for (int k = 0; k < i; ++k) s -= L[i + 6 * k] * z[k];
      z[i] = s / L[i + 6 * i];
    }
    for (int i = rank - 1; i >= 0; --i) {
      float s = z[i];
      for (int k = i + 1; k < rank; ++k) s -= L[k + 6 * i] * z[k];
      z[i] = s / L[i + 6 * i];
    }
    for (int j = 0; j < 6; ++j) {
      float acc = 0.f;
      for (int r = 0; r < rank && r <= j; ++r) acc += R1[r + 6 * j] * z[r];
      y[j] = acc;
    }
    for (int j = 0; j < 6; ++j) x[qr.perm[j]] = y[j];
  }
  float e2 = 0.f, b2 = 0.f, a2 = 0.f;
  for (int i = 0; i < 6; ++i) {
    float ax = 0.f;
    for (int j = 0; j < 6; ++j) ax += A[i + 6 * j] * x[j];
    e2 += (b[i] - ax) * (b[i] - ax); b2 += b[i] * b[i]; a2 += ax * ax;
  }
  if (!(e2 <= 1e-5f * 1e-5f * fminf(b2, a2))) {
    if (x_spd) {
      for (int i = 0; i < 6; ++i) x[i] = x_spd[i];
    } else {
      la::svdSolve<double, 6>(Ad, bd, xd);
      for (int i = 0; i < 6; ++i) x[i] = (float)xd[i];
    }
    *info = 2;
  }
}

// Rank decision of la::pivqr_rank<float, 6> replayed by one warp on a shared-memory work matrix (36 floats): lane j
// owns column j, rolled loops, same operations per element in the same order => the same pivots and the same rank as
// Eigen's FullPivHouseholderQR restatement, in compact code. Returns true iff the rank is 6.
__device__ inline bool pivqr_full_rank6_warp(const float* A, float* Q) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  if (lane < 6)
    for (int i = 0; i < 6; ++i) Q[i + 6 * lane] = A[i + 6 * lane];
  __syncwarp();
  const float prec = FLT_EPSILON * 6.f;
  float maxpivot = 0.f, minpivot = INFINITY, biggest = 0.f;
  for (int k = 0; k < 6; ++k) {
    float big = -1.f;
    int rb = k, cb = lane;
    if (lane >= k && lane < 6)
      for (int i = k; i < 6; ++i) {
        const float v = fabsf(Q[i + 6 * lane]);
        if (v > big) { big = v; rb = i; }
      }
    for (int o = 4; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(FULL, big, o);
      const int orb = __shfl_xor_sync(FULL, rb, o), ocb = __shfl_xor_sync(FULL, cb, o);
      if (ob > big || (ob == big && ocb < cb)) { big = ob; rb = orb; cb = ocb; }
    }
    big = __shfl_sync(FULL, big, 0); rb = __shfl_sync(FULL, rb, 0); cb = __shfl_sync(FULL, cb, 0);
    if (k == 0) biggest = big;
    if (fabsf(big) <= fabsf(biggest) * prec) return false;
    if (lane >= k && lane < 6 && rb != k) { const float t = Q[k + 6 * lane]; Q[k + 6 * lane] = Q[rb + 6 * lane]; Q[rb + 6 * lane] = t; }
    __syncwarp();
    if (cb != k && lane < 6) { const float t = Q[lane + 6 * k]; Q[lane + 6 * k] = Q[lane + 6 * cb]; Q[lane + 6 * cb] = t; }
    __syncwarp();
    float tail = 0.f;
    for (int i = k + 1; i < 6; ++i) tail += Q[i + 6 * k] * Q[i + 6 * k];
    const float c0 = Q[k + 6 * k];
    float beta, tk;
    const bool tiny = tail <= FLT_MIN;
    if (tiny) { tk = 0.f; beta = c0; }
    else {
      beta = sqrtf(c0 * c0 + tail);
      if (c0 >= 0.f) beta = -beta;
      tk = (beta - c0) / beta;
    }
    __syncwarp();
    if (lane > k && lane < 6) Q[lane + 6 * k] = tiny ? 0.f : Q[lane + 6 * k] / (c0 - beta);  // essential part, one entry per lane
    if (lane == k) Q[k + 6 * k] = beta;
    __syncwarp();
    maxpivot = fmaxf(maxpivot, fabsf(beta));
    minpivot = fminf(minpivot, fabsf(beta));
    if (k < 5 && tk != 0.f && lane > k && lane < 6) {
      float tmp = 0.f;
      for (int i = k + 1; i < 6; ++i) tmp += Q[i + 6 * k] * Q[i + 6 * lane];
      tmp += Q[k + 6 * lane];
      Q[k + 6 * lane] -= tk * tmp;
      for (int i = k + 1; i < 6; ++i) Q[i + 6 * lane] -= tk * Q[i + 6 * k] * tmp;
    }
    __syncwarp();
  }
  return minpivot > fabsf(maxpivot) * prec;
}

// Warp-parallel fast path of solve6_dev for the common case (numerically full rank, symmetric positive definite).
// Deliberately COMPACT code: rolled loops over a shared-memory work matrix. One warp executes this once per
// iteration, so its cost is dominated by instruction fetch of straight-line code, not by arithmetic: the fully
// unrolled register version (and the serial reference path) took 7-9 us per iteration, this one well under 2 us.
//
// Lanes 0..5 factor A = L D L^T in double (column j on lane j); lanes 8..13 run the SAME instructions on A - s I,
// s = 2.8 * (6 eps_f) * cmax with cmax = the largest column norm of A, which is the rank certificate: every diagonal
// entry of any QR factor satisfies |R_kk| >= sigma_min(A), and the reference's rank threshold is 6 eps_f * max|R_kk|
// <= 6 eps_f * cmax (Householder reflections preserve column norms); if A - s I is still positive definite then
// lambda_min(A) > s and FullPivHouseholderQR (PointToPlane.cpp:202-203) finds full rank. The factor 2.8 = 1 (the
// threshold itself) + 1.8 of allowance for the ~eps_f |A| rounding of a FLOAT QR: an empirical margin, not a proven
// bound — which is why a failed certificate does not decide anything: the QR rank decision is then replayed exactly
// (pivqr_full_rank6_warp), and tests sweep lambda_min across 1x..4x the threshold against the oracle's rank decision
// (tests/test_gpu_parity.py::test_solve_rank_threshold_band). 1/d_k comes from a float seed
// and three Newton steps in double (IEEE double division/sqrt are long dependent sequences here); against
// cond(A) ~ 1e4..1e5 the ~1 ulp quotient error is ~1e-11 relative in x, invisible after the cast to float.
// Returns 1: x is final (full rank by the reference's criterion). 0: nothing decided, x untouched (asymmetric input or
// a non-positive pivot): the serial reference routine decides. 2: the reference's float QR finds A rank deficient
// (replayed exactly), but the double LDL^T went through with well-conditioned pivots (min d / max d > 1e-9): x_spd
// receives A^-1 b for solve6_dev to use where the reference falls back to its double SVD; x untouched.
// W: 98 doubles of shared memory. Must be called by all 32 lanes of one warp.
__device__ inline int solve6_warp(const float* A, const float* b, float* x, double* W, float* x_spd = nullptr) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int blk = (lane >> 3) & 1, j = lane & 7;   // block 0: A, block 1: A - s I
  const bool mine = lane < 16 && j < 6;
  double* M = W + 36 * blk;                          // column-major 6x6, lower triangle used
  // shift s = 2.8 * (6 eps_f) * cmax, cmax = largest column norm >= max|R_kk| (Householder reflections preserve
  // column norms): 1x is the reference's threshold, the rest covers the ~eps_f * |A| rounding of a float QR
  float cn = 0.f;
  for (int i = 0; i < 6; ++i) { const float t = A[i + 6 * (j < 6 ? j : 0)]; cn += t * t; }
  for (int o = 4; o > 0; o >>= 1) cn = fmaxf(cn, __shfl_xor_sync(FULL, cn, o));
  const float cmax = sqrtf(__shfl_sync(FULL, cn, 0));
  const double sft = (double)(2.8f * 6.f * FLT_EPSILON * cmax);
  bool good = cmax > 0.f;
  if (mine) {
    for (int i = 0; i < 6; ++i) {
      const double u = (double)A[i + 6 * j], w = (double)A[j + 6 * i];
      if (fabs(u - w) > 1e-6 * (fabs((double)A[i + 6 * i]) + fabs((double)A[j + 6 * j]))) good = false;
      M[i + 6 * j] = 0.5 * (u + w) - ((blk == 1 && i == j) ? sft : 0.0);
    }
  }
  __syncwarp();
  double* rinv = W + 72;  // 6 reciprocals of block 0
  for (int k = 0; k < 6; ++k) {
    const double d = M[k + 6 * k];
    if (!(d > 0.0)) good = false;
    double rk = (double)(1.0f / (float)d);
    rk = rk * (2.0 - d * rk);
    rk = rk * (2.0 - d * rk);
    rk = rk * (2.0 - d * rk);
    if (mine && j > k) {  // a(i, j) -= L(i, k) * d_k * L(j, k), i >= j, with column k still holding L * d
      const double cjk = M[j + 6 * k];
      for (int i = j; i < 6; ++i) M[i + 6 * j] -= (M[i + 6 * k] * rk) * cjk;
    }
    __syncwarp();
    if (mine && j > k) M[j + 6 * k] *= rk;  // column k becomes L(:, k); one entry per lane
    if (lane == 0) rinv[k] = rk;
    __syncwarp();
  }
  // block 0 must have factored; block 1 is only the certificate
  const bool good0 = __all_sync(FULL, good || blk == 1);
  if (!good0) return 0;
  bool full = true;
  if (!__all_sync(FULL, good)) {
    // certificate failed: replay the reference's rank decision exactly (work matrix after the two double blocks)
    full = pivqr_full_rank6_warp(A, reinterpret_cast<float*>(W + 80));
    if (!full) {
      if (!x_spd) return 0;
      double rmin = rinv[0], rmax = rinv[0];   // reciprocals of the pivots
      for (int k = 1; k < 6; ++k) { rmin = fmin(rmin, rinv[k]); rmax = fmax(rmax, rinv[k]); }
      if (!(rmin > 1e-9 * rmax)) return 0;
      x = x_spd;
    }
  }
  // substitutions on block 0: lane i owns component i
  double v = (lane < 6) ? (double)b[lane] : 0.0;
  for (int k = 0; k < 6; ++k) {  // forward, unit L
    const double yk = __shfl_sync(FULL, v, k);
    if (lane < 6 && lane > k) v -= W[lane + 6 * k] * yk;
  }
  if (lane < 6) v *= rinv[lane];
  for (int k = 5; k >= 0; --k) {  // backward, L^T
    const double xk = __shfl_sync(FULL, v, k);
    if (lane < k) v -= W[k + 6 * lane] * xk;
  }
  if (lane < 6) x[lane] = (float)v;
  __syncwarp();
  return full ? 1 : 2;
}

// x -> 4x4 — PointToPlane.cpp:337-381 (6-DOF branch)
__device__ inline void p2plane_T_dev(const float* x, float* T) {
  eye4(T);
  float n2 = __fmul_rn(x[0], x[0]);
  n2 = __fadd_rn(n2, __fmul_rn(x[1], x[1]));
  n2 = __fadd_rn(n2, __fmul_rn(x[2], x[2]));
  const float angle = atanf(sqrtf(n2));
  float ax[3] = {x[0], x[1], x[2]};
  const float wmax = fmaxf(fabsf(x[0]), fmaxf(fabsf(x[1]), fabsf(x[2])));
  {
    const float s0 = x[0] / wmax, s1 = x[1] / wmax, s2 = x[2] / wmax;
    float z = __fmul_rn(s0, s0);
    z = __fadd_rn(z, __fmul_rn(s1, s1));
    z = __fadd_rn(z, __fmul_rn(s2, s2));
    if (z > 0.f) {
      const float sq = sqrtf(z);
      ax[0] = s0 / sq; ax[1] = s1 / sq; ax[2] = s2 / sq;
    }
  }
  const float s = sinf(angle), c = cosf(angle);
  const float sx = __fmul_rn(s, ax[0]), sy = __fmul_rn(s, ax[1]), sz = __fmul_rn(s, ax[2]);
  const float omc = __fsub_rn(1.f, c);
  const float cx = __fmul_rn(omc, ax[0]), cy = __fmul_rn(omc, ax[1]), cz = __fmul_rn(omc, ax[2]);
  float R[9], tmp;
  tmp = __fmul_rn(cx, ax[1]); R[3] = __fsub_rn(tmp, sz); R[1] = __fadd_rn(tmp, sz);
  tmp = __fmul_rn(cx, ax[2]); R[6] = __fadd_rn(tmp, sy); R[2] = __fsub_rn(tmp, sy);
  tmp = __fmul_rn(cy, ax[2]); R[7] = __fsub_rn(tmp, sx); R[5] = __fadd_rn(tmp, sx);
  R[0] = __fadd_rn(__fmul_rn(cx, ax[0]), c);
  R[4] = __fadd_rn(__fmul_rn(cy, ax[1]), c);
  R[8] = __fadd_rn(__fmul_rn(cz, ax[2]), c);
  bool nan = false;
  for (int i = 0; i < 9; ++i) nan |= (R[i] != R[i]);
  for (int i = 3; i < 6; ++i) nan |= (x[i] != x[i]);
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) T[i + 4 * j] = nan ? (i == j ? 1.f : 0.f) : R[i + 3 * j];
  T[12] = x[3]; T[13] = x[4]; T[14] = x[5];
}

// ---- transformation checkers — TransformationCheckersImpl.cpp:46-151 --------------------------------------------
struct Quat { float w, x, y, z; };

__device__ inline Quat quat_from_T(const float* T) {
  Quat q;
  float t = T[0] + T[5] + T[10];
  if (t > 0.f) {
    t = sqrtf(t + 1.f);
    q.w = 0.5f * t;
    t = 0.5f / t;
    q.x = (T[2 + 4 * 1] - T[1 + 4 * 2]) * t;
    q.y = (T[0 + 4 * 2] - T[2 + 4 * 0]) * t;
    q.z = (T[1 + 4 * 0] - T[0 + 4 * 1]) * t;
  } else {
    int i = 0;
    if (T[5] > T[0]) i = 1;
    if (T[10] > T[i + 4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = sqrtf(T[i + 4 * i] - T[j + 4 * j] - T[k + 4 * k] + 1.f);
    float v[3];
    v[i] = 0.5f * t;
    t = 0.5f / t;
    q.w = (T[k + 4 * j] - T[j + 4 * k]) * t;
    v[j] = (T[j + 4 * i] + T[i + 4 * j]) * t;
    v[k] = (T[k + 4 * i] + T[i + 4 * k]) * t;
    q.x = v[0]; q.y = v[1]; q.z = v[2];
  }
  return q;
}
__device__ inline float angular_distance(const Quat& a, const Quat& b0) {
  const Quat b{b0.w, -b0.x, -b0.y, -b0.z};
  const float w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  const float x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  const float y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
  const float z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
  return 2.f * atan2f(sqrtf(x * x + y * y + z * z), fabsf(w));
}

// one-sided Jacobi SVD of a 3x3 in FLOAT with singular values sorted descending (stable): plays
// Eigen::JacobiSVD<Matrix3f> of PointToPoint.cpp:86 in the reference's own precision. A = U diag(S) V^T.
__device__ inline void svd3f(const float* A, float* U, float* S, float* V) {
  float W[9], Vw[9];
  for (int i = 0; i < 9; ++i) { W[i] = A[i]; Vw[i] = 0.f; }
  Vw[0] = Vw[4] = Vw[8] = 1.f;
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rot = false;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        float al = 0.f, be = 0.f, ga = 0.f;
        for (int i = 0; i < 3; ++i) {
          al += W[i + 3 * p] * W[i + 3 * p]; be += W[i + 3 * q] * W[i + 3 * q]; ga += W[i + 3 * p] * W[i + 3 * q];
        }
        if (ga == 0.f || fabsf(ga) <= FLT_EPSILON * sqrtf(al * be)) continue;
        rot = true;
        const float zeta = (be - al) / (2.f * ga);
        const float t = (zeta >= 0.f ? 1.f : -1.f) / (fabsf(zeta) + sqrtf(1.f + zeta * zeta));
        const float c = 1.f / sqrtf(1.f + t * t), sn = c * t;
        for (int i = 0; i < 3; ++i) {
          const float wp = W[i + 3 * p], wq = W[i + 3 * q];
          W[i + 3 * p] = c * wp - sn * wq; W[i + 3 * q] = sn * wp + c * wq;
          const float vp = Vw[i + 3 * p], vq = Vw[i + 3 * q];
          Vw[i + 3 * p] = c * vp - sn * vq; Vw[i + 3 * q] = sn * vp + c * vq;
        }
      }
    if (!rot) break;
  }
  float sv[3];
  int order[3] = {0, 1, 2};
  for (int j = 0; j < 3; ++j) {
    float n2 = 0.f;
    for (int i = 0; i < 3; ++i) n2 += W[i + 3 * j] * W[i + 3 * j];
    sv[j] = sqrtf(n2);
  }
  // stable insertion sort, descending
  for (int a = 1; a < 3; ++a)
    for (int b = a; b > 0 && sv[order[b]] > sv[order[b - 1]]; --b) { const int t = order[b]; order[b] = order[b - 1]; order[b - 1] = t; }
  for (int jj = 0; jj < 3; ++jj) {
    const int j = order[jj];
    S[jj] = sv[j];
    for (int i = 0; i < 3; ++i) {
      U[i + 3 * jj] = sv[j] > 0.f ? W[i + 3 * j] / sv[j] : 0.f;
      V[i + 3 * jj] = Vw[i + 3 * j];
    }
  }
}

// PointToPointErrorMinimizer::compute_in_place (PointToPoint.cpp:62-103) from raw weighted moments summed in double:
// m[0] = sum w, m[1..3] = sum w*r, m[4..6] = sum w*f, m[7..15] = sum w * f_i * r_j (column-major i + 3j).
// Centroids and the centred cross-covariance are rounded to float, and SVD / R = U V^T / reflection fix / t are done
// in float exactly like the reference (Matrix3f JacobiSVD).
__device__ inline void p2point_from_moments(const double* m, float* T) {
  const double winv = 1.0 / m[0];
  double mrd[3], mfd[3];
  float mr[3], mf[3], C[9];
  for (int d = 0; d < 3; ++d) { mrd[d] = m[1 + d] * winv; mfd[d] = m[4 + d] * winv; mr[d] = (float)mrd[d]; mf[d] = (float)mfd[d]; }
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) C[i + 3 * j] = (float)(m[7 + i + 3 * j] - m[0] * mfd[i] * mrd[j]);
  float U[9], S[3], V[9], Vt[9], R[9];
  svd3f(C, U, S, V);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Vt[i + 3 * j] = V[j + 3 * i];
  auto compose = [&]() {
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < 3; ++i) {
        float a = 0.f;
        for (int k = 0; k < 3; ++k) a += U[i + 3 * k] * Vt[k + 3 * j];
        R[i + 3 * j] = a;
      }
  };
  compose();
  if (la::det3(R) < 0.f) {  // PointToPoint.cpp:92-97: negate the last row of V^T
    for (int j = 0; j < 3; ++j) Vt[2 + 3 * j] = -Vt[2 + 3 * j];
    compose();
  }
  eye4(T);
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) T[i + 4 * j] = R[i + 3 * j];
  for (int i = 0; i < 3; ++i) {
    float a = 0.f;
    for (int j = 0; j < 3; ++j) a += R[i + 3 * j] * mr[j];
    T[12 + i] = mf[i] - a;
  }
}


// general 4x4 inverse, Gauss-Jordan with partial pivoting (plays Eigen's PartialPivLU inverse on the frame
// matrices of ICP.cpp:983,1345; exact for pure translations)
__device__ inline bool inv4_dev(const float* A, float* Ainv) {
  float a[4][8];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) { a[i][j] = A[i + 4 * j]; a[i][4 + j] = (i == j) ? 1.f : 0.f; }
  for (int c = 0; c < 4; ++c) {
    int piv = c;
    for (int r = c + 1; r < 4; ++r)
      if (fabsf(a[r][c]) > fabsf(a[piv][c])) piv = r;
    if (a[piv][c] == 0.f) return false;
    if (piv != c)
      for (int j = 0; j < 8; ++j) { const float t = a[piv][j]; a[piv][j] = a[c][j]; a[c][j] = t; }
    const float d = a[c][c];
    for (int j = 0; j < 8; ++j) a[c][j] /= d;
    for (int r = 0; r < 4; ++r)
      if (r != c) {
        const float f = a[r][c];
        if (f != 0.f)
          for (int j = 0; j < 8; ++j) a[r][j] = __fsub_rn(a[r][j], __fmul_rn(f, a[c][j]));
      }
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) Ainv[i + 4 * j] = a[i][4 + j];
  return true;
}

}  // namespace b200
