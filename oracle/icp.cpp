// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
//
// Restatement of PointMatcher<float>::ICP — libpointmatcher/pointmatcher/ICP.cpp:813-1349 (compute,
// initReference, computeWithTransformedReference, degeneracyAwareness = kNone) and of the transformation
// checkers — TransformationCheckersImpl.cpp:46-151, TransformationChecker.cpp:109-121.
#include "oracle.h"
#include "la_small.hpp"
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

namespace {

typedef float M4[16];  // column-major

void eye4(float* T) {
  for (int i = 0; i < 16; ++i) T[i] = 0.f;
  T[0] = T[5] = T[10] = T[15] = 1.f;
}
void mul4(const float* A, const float* B, float* C) {  // C = A * B, depth-sequential accumulation
  float R[16];
  for (int j = 0; j < 4; ++j)
    for (int i = 0; i < 4; ++i) {
      float acc = A[i] * B[4 * j];
      acc = acc + A[i + 4] * B[1 + 4 * j];
      acc = acc + A[i + 8] * B[2 + 4 * j];
      acc = acc + A[i + 12] * B[3 + 4 * j];
      R[i + 4 * j] = acc;
    }
  std::memcpy(C, R, sizeof(R));
}
bool inv4(const float* A, float* Ainv) {  // Gauss-Jordan with partial pivoting (plays PartialPivLU::inverse)
  float a[4][8];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) { a[i][j] = A[i + 4 * j]; a[i][4 + j] = (i == j) ? 1.f : 0.f; }
  for (int c = 0; c < 4; ++c) {
    int piv = c;
    for (int r = c + 1; r < 4; ++r)
      if (std::fabs(a[r][c]) > std::fabs(a[piv][c])) piv = r;
    if (a[piv][c] == 0.f) return false;
    if (piv != c)
      for (int j = 0; j < 8; ++j) std::swap(a[piv][j], a[c][j]);
    const float d = a[c][c];
    for (int j = 0; j < 8; ++j) a[c][j] /= d;
    for (int r = 0; r < 4; ++r)
      if (r != c) {
        const float f = a[r][c];
        if (f != 0.f)
          for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j];
      }
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) Ainv[i + 4 * j] = a[i][4 + j];
  return true;
}

struct Quat { float w, x, y, z; };

// Eigen::Quaternion(Matrix3) — Shoemake's algorithm as in Eigen/src/Geometry/Quaternion.h
Quat quatFromT(const float* T) {
  auto m = [&](int i, int j) { return T[i + 4 * j]; };
  Quat q;
  float t = m(0, 0) + m(1, 1) + m(2, 2);
  if (t > 0.f) {
    t = std::sqrt(t + 1.f);
    q.w = 0.5f * t;
    t = 0.5f / t;
    q.x = (m(2, 1) - m(1, 2)) * t;
    q.y = (m(0, 2) - m(2, 0)) * t;
    q.z = (m(1, 0) - m(0, 1)) * t;
  } else {
    int i = 0;
    if (m(1, 1) > m(0, 0)) i = 1;
    if (m(2, 2) > m(i, i)) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(m(i, i) - m(j, j) - m(k, k) + 1.f);
    float v[3];
    v[i] = 0.5f * t;
    t = 0.5f / t;
    q.w = (m(k, j) - m(j, k)) * t;
    v[j] = (m(j, i) + m(i, j)) * t;
    v[k] = (m(k, i) + m(i, k)) * t;
    q.x = v[0]; q.y = v[1]; q.z = v[2];
  }
  return q;
}
// Quaternion::angularDistance (Eigen >= 3.3): d = a * conj(b); 2 * atan2(|d.vec|, |d.w|)
float angularDistance(const Quat& a, const Quat& b0) {
  const Quat b{b0.w, -b0.x, -b0.y, -b0.z};
  const float w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  const float x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  const float y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
  const float z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
  const float vn = std::sqrt(x * x + y * y + z * z);
  return 2.f * std::atan2(vn, std::fabs(w));
}

struct Checkers {
  const orc_icp_params* p;
  float counter = 0.f;
  std::vector<Quat> rot;
  Quat rot0;              // BoundTransformationChecker::initialRotation3D
  std::vector<float> tr;  // 3 per entry
  void init(const float* T) {
    counter = 0.f;
    rot.clear(); tr.clear();
    rot.push_back(quatFromT(T));
    rot0 = rot[0];
    tr.push_back(T[12]); tr.push_back(T[13]); tr.push_back(T[14]);
  }
  // returns status; sets iterate / maxReached
  int check(const float* T, bool& iterate, bool& maxReached) {
    bool counterHit = false;
    if (p->max_iterations > 0) {
      counter += 1.f;
      if (counter >= (float)p->max_iterations) counterHit = true;  // throws MaxNumIterationsReached
    }
    if (p->use_differential) {
      rot.push_back(quatFromT(T));
      tr.push_back(T[12]); tr.push_back(T[13]); tr.push_back(T[14]);
      float cv0 = 0.f, cv1 = 0.f;
      const size_t L = (size_t)p->smooth_length;
      if (rot.size() > L) {
        for (size_t i = rot.size() - 1; i >= rot.size() - L; --i) {
          cv0 += std::fabs(angularDistance(rot[i], rot[i - 1]));
          const float dx = tr[3 * i] - tr[3 * (i - 1)], dy = tr[3 * i + 1] - tr[3 * (i - 1) + 1],
                      dz = tr[3 * i + 2] - tr[3 * (i - 1) + 2];
          cv1 += std::fabs(std::sqrt(dx * dx + dy * dy + dz * dz));
        }
        cv0 /= (float)L;
        cv1 /= (float)L;
        if (cv0 < p->min_diff_rot && cv1 < p->min_diff_trans) iterate = false;
      }
      if (cv0 != cv0 || cv1 != cv1) return ORC_E_NAN;  // ConvergenceError, TransformationCheckersImpl.cpp:147-150
    }
    if (p->use_bound) {
      // BoundTransformationChecker::check, TransformationCheckersImpl.cpp:198-222; init saw T_iter = identity
      // (ICP.cpp:997): initialRotation3D = Quaternion(I), initialTranslation = 0
      const float cvr = angularDistance(quatFromT(T), rot0);
      const float cvt = std::sqrt(T[12] * T[12] + T[13] * T[13] + T[14] * T[14]);
      if (cvr > p->max_rotation_norm || cvt > p->max_translation_norm) return ORC_E_BOUND;  // ConvergenceError, :214-221
    }
    if (counterHit) { iterate = false; maxReached = true; }
    return ORC_OK;
  }
};

void meanOf(const float* feat, int N, int mode, float* mean3) {
  if (mode == 0) {
    float s[3] = {0, 0, 0};
    for (int i = 0; i < N; ++i)
      for (int d = 0; d < 3; ++d) s[d] += feat[4 * (size_t)i + d];
    for (int d = 0; d < 3; ++d) mean3[d] = s[d] / (float)N;
  } else {
    double s[3] = {0, 0, 0};
    for (int i = 0; i < N; ++i)
      for (int d = 0; d < 3; ++d) s[d] += (double)feat[4 * (size_t)i + d];
    for (int d = 0; d < 3; ++d) mean3[d] = (float)(s[d] / (double)N);
  }
}


// ---------------------------------------------------------------------------------------------------------------
// X-ICP localizability, degeneracyAwareness = OptimizedEqualityConstraints (param/icp.yaml of open3d_slam_ros;
// loadAdditionalYAMLContent ICP.cpp:629-670). State = LocalizabilityParametersForErrorMinimization
// (PointMatcher.h:585-660), local to one computeWithTransformedReference call (ICP.cpp:921).
// ---------------------------------------------------------------------------------------------------------------
struct XicpState {
  bool repeatEigen = true;      // debugging_.whetherToRepeatEigenAnalysis_ (PointMatcher.h:592)
  bool returnPrior = false;     // debugging_.whetherToReturnPrior_ (:595)
  float Ut[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // translationEigenvectors_ (:634), columns
  float Ur[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // rotationEigenvectors_ (:635)
  int locXyz[3] = {1, 1, 1};    // localizabilityXyz_ (:638), 1 = kLocalizable
  int locRpy[3] = {1, 1, 1};
  float cT[3] = {0, 0, 0}, cR[3] = {0, 0, 0};  // translation/rotationConstraintValues_ (:625-626)
  double sums[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // diagnostics: full (no early exit) sums, see oracle.h
};

bool inv3(const float* A, float* Ainv) {  // Gauss-Jordan with partial pivoting (dynamic 3x3 .inverse() = PartialPivLU)
  float M[3][6];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { M[i][j] = A[i + 3 * j]; M[i][3 + j] = (i == j) ? 1.f : 0.f; }
  for (int c = 0; c < 3; ++c) {
    int piv = c;
    for (int r = c + 1; r < 3; ++r)
      if (std::fabs(M[r][c]) > std::fabs(M[piv][c])) piv = r;
    if (M[piv][c] == 0.f) return false;
    if (piv != c)
      for (int j = 0; j < 6; ++j) std::swap(M[piv][j], M[c][j]);
    const float d = M[c][c];
    for (int j = 0; j < 6; ++j) M[c][j] /= d;
    for (int r = 0; r < 3; ++r)
      if (r != c) {
        const float f = M[r][c];
        for (int j = 0; j < 6; ++j) M[r][j] -= f * M[c][j];
      }
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) Ainv[i + 3 * j] = M[i][3 + j];
  return true;
}

void mul3v(const float* R, float* v) {  // v <- R v (column-major 3x3)
  float o[3];
  for (int i = 0; i < 3; ++i) o[i] = R[i] * v[0] + R[i + 3] * v[1] + R[i + 6] * v[2];
  v[0] = o[0]; v[1] = o[1]; v[2] = o[2];
}

// ICP::eigenAnalysis, ICP.cpp:1580-1591: U of the SVD of the translation (bottom-right) and rotation (top-left)
// 3x3 blocks of the 6x6 Hessian, singular values descending.
// Eigen's U is a full orthonormal basis even when the smallest singular value is exactly zero (an exactly planar
// set of normals); the one-sided Jacobi of la_small.hpp leaves that column zero, so it is completed here.
void completeU3(float* U, const float* S) {
  if (S[2] == 0.f && S[1] > 0.f) {
    U[6] = U[1] * U[5] - U[2] * U[4];
    U[7] = U[2] * U[3] - U[0] * U[5];
    U[8] = U[0] * U[4] - U[1] * U[3];
  }
}
void xicpEigenAnalysis(const float* H, float* Ut, float* Ur) {
  float B[9], S[3], V[9];
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) B[i + 3 * j] = H[(3 + i) + 6 * (3 + j)];
  orc::jacobiSVD<float, 3>(B, Ut, S, V);
  completeU3(Ut, S);
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) B[i + 3 * j] = H[i + 6 * j];
  orc::jacobiSVD<float, 3>(B, Ur, S, V);
  completeU3(Ur, S);
}

// ICP::detectLocalizability, ICP.cpp:2128-2155: sequential float sums with the early exit in the loop condition.
// full2 (diagnostics): the two sums over ALL points, in double.
bool xicpDetect(const float* align3xP, int P, const float* v, float cosMin, float cosStrong, float enough,
                float insufficient, double* full2) {
  bool ok = false;
  float combined = 0.f, high = 0.f;
  for (int i = 0; i < P && !ok; ++i) {
    const float* a = align3xP + 3 * (size_t)i;
    const float al = std::fabs(a[0] * v[0] + a[1] * v[1] + a[2] * v[2]);
    if (al > cosMin) combined += al;
    if (al > cosStrong) high += al;
    ok = (combined >= enough) || (high >= insufficient);
  }
  if (full2) {
    double c = 0, h = 0;
    for (int i = 0; i < P; ++i) {
      const float* a = align3xP + 3 * (size_t)i;
      const float al = std::fabs(a[0] * v[0] + a[1] * v[1] + a[2] * v[2]);
      if (al > cosMin) c += (double)al;
      if (al > cosStrong) h += (double)al;
    }
    full2[0] = c; full2[1] = h;
  }
  return ok;
}

// ICP::detectLocalizabilityWithOptimizedMethod, ICP.cpp:2187-2444. readP (4 x P) / refnP (3 x P) are the matched
// reading points (refMean frame) and matched reference normals of ErrorElements; T_opt = T_refMean_dataIn
// (transformationToOptimizationFrame, ICP.cpp:1067); T_convert_from_base_to_lidar_mat is the identity (:1012-1024).
bool xicpDetectOptimized(const orc_icp_params* prm, const float* H, const float* readP, const float* refnP, int P,
                         const float* T_opt, bool first, XicpState& st) {
  if (P == 0) return false;                                      // :2191
  bool allZero = true;
  for (int i = 0; i < 36; ++i) allZero = allZero && (H[i] == 0.f);
  if (allZero) return false;                                     // :2197
  if (st.repeatEigen) xicpEigenAnalysis(H, st.Ut, st.Ur);        // :2205-2210
  bool zt = true, zr = true;
  for (int i = 0; i < 9; ++i) { zt = zt && (st.Ut[i] == 0.f); zr = zr && (st.Ur[i] == 0.f); }
  if (zt || zr) return false;                                    // :2212-2219
  if (!first) return true;                                       // :2221-2226

  // :2236-2240: matched points and matched normals back into the frame the reading came in
  float Tinv[16];
  inv4(T_opt, Tinv);
  std::vector<float> pts(readP, readP + 4 * (size_t)P), nrm(refnP, refnP + 3 * (size_t)P);
  orc_rigid_apply(Tinv, pts.data(), nrm.data(), P);
  // :2248-2267: eigenvectors into the same frame
  float R[9], Rinv[9];
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) R[i + 3 * j] = T_opt[i + 4 * j];
  if (!inv3(R, Rinv)) return false;
  for (int c = 0; c < 3; ++c) { mul3v(Rinv, st.Ur + 3 * c); mul3v(Rinv, st.Ut + 3 * c); }
  // :2270-2271 calculateCrossProductElements (:1669-1697): centre (sequential float sums, :1604-1617), cross(p - c, n)
  float center[3] = {0.f, 0.f, 0.f};
  for (int i = 0; i < P; ++i)
    for (int f = 0; f < 3; ++f) center[f] += pts[4 * (size_t)i + f];
  for (int f = 0; f < 3; ++f) center[f] /= (float)P;
  std::vector<float> rotAlign(3 * (size_t)P);
  for (int i = 0; i < P; ++i) {
    const float ax = pts[4 * (size_t)i] - center[0], ay = pts[4 * (size_t)i + 1] - center[1],
                az = pts[4 * (size_t)i + 2] - center[2];
    const float* n = nrm.data() + 3 * (size_t)i;
    float c[3] = {ay * n[2] - az * n[1], az * n[0] - ax * n[2], ax * n[1] - ay * n[0]};  // ErrorMinimizer.cpp:376-378
    const float nn = std::sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    if (!(nn < 1.0f)) { c[0] /= nn; c[1] /= nn; c[2] /= nn; }  // :2283-2291
    rotAlign[3 * (size_t)i] = c[0]; rotAlign[3 * (size_t)i + 1] = c[1]; rotAlign[3 * (size_t)i + 2] = c[2];
  }
  // :2294-2385
  for (int e = 0; e < 3; ++e) {
    const bool okT = xicpDetect(nrm.data(), P, st.Ut + 3 * e, prm->xicp_cos_min, prm->xicp_cos_strong, prm->xicp_enough,
                                prm->xicp_insufficient, st.sums + 2 * e);
    st.locXyz[e] = okT ? 1 : 0;
    st.cT[e] = okT ? 1.f : 0.f;
    const bool okR = xicpDetect(rotAlign.data(), P, st.Ur + 3 * e, prm->xicp_cos_min, prm->xicp_cos_strong,
                                prm->xicp_enough, prm->xicp_insufficient, st.sums + 6 + 2 * e);
    st.locRpy[e] = okR ? 1 : 0;
    st.cR[e] = okR ? 1.f : 0.f;
  }
  // :2389-2411 eigenvectors back to the optimisation (refMean) frame
  for (int c = 0; c < 3; ++c) { mul3v(R, st.Ur + 3 * c); mul3v(R, st.Ut + 3 * c); }
  // :2415-2429
  st.repeatEigen = false;
  for (int i = 0; i < 3; ++i)
    if (st.locXyz[i] == 0 || st.locRpy[i] == 0) { st.repeatEigen = true; break; }
  return true;
}

// solvePossiblyUnderdeterminedLinearSystemWithEqualityConstraints, PointToPlane.cpp:461-505, with
// readLocalizabilityFlags (:757-778) and generateConstrainedOptimizationProblem (:570-626)
int xicpSolve(const float* H, const float* b, const XicpState& st, float* x6, int* info, int* nConstraints) {
  int idx[6], c = 0;
  for (int i = 0; i < 3; ++i) if (st.locRpy[i] == 0) idx[c++] = i;
  for (int i = 0; i < 3; ++i) if (st.locXyz[i] == 0) idx[c++] = i + 3;
  if (nConstraints) *nConstraints = c;
  if (c == 0) return orc_solve6(H, b, x6, info);
  const int n = 6 + c;
  std::vector<float> A((size_t)n * n, 0.f), rhs(n, 0.f), x(n, 0.f);
  for (int j = 0; j < 6; ++j)
    for (int i = 0; i < 6; ++i) A[i + n * j] = H[i + 6 * j];
  for (int i = 0; i < 6; ++i) rhs[i] = b[i];
  for (int cc = 0; cc < c; ++cc) {
    const int index = idx[cc];
    if (index < 3) {
      for (int i = 0; i < 3; ++i) { A[i + n * (6 + cc)] = st.Ur[i + 3 * index]; A[(6 + cc) + n * i] = st.Ur[i + 3 * index]; }
      rhs[6 + cc] = st.cR[index];
    } else {
      for (int i = 0; i < 3; ++i) {
        A[(i + 3) + n * (6 + cc)] = st.Ut[i + 3 * (index - 3)];
        A[(6 + cc) + n * (i + 3)] = st.Ut[i + 3 * (index - 3)];
      }
      rhs[6 + cc] = st.cT[index - 3];
    }
  }
  const int r = orc_solve_n(A.data(), rhs.data(), n, x.data(), info);
  for (int i = 0; i < 6; ++i) x6[i] = x[i];
  return r;
}

}  // namespace

extern "C" int orc_icp(const orc_icp_params* prm, const float* ref4xM, const float* refn3xM, int M,
                       const float* read4xN, const float* readn3xN, int N, const float* Tinit, orc_icp_result* res,
                       int max_log, float* logH, float* logb, float* logx, float* logT, int* logP,
                       int32_t* last_ids, float* last_dists, float* last_w) {
  return orc_icp_forced(prm, ref4xM, refn3xM, M, read4xN, readn3xN, N, Tinit, res, max_log, logH, logb, logx, logT,
                        logP, last_ids, last_dists, last_w, nullptr, 0);
}

// One-step ("teacher-forced") variant used by the parity tests: with forcedT (n_forced column-major 4x4), iteration
// i logs the oracle's own T_iter computed from the current state and then CONTINUES from forcedT[i] (the pose another
// implementation reached after its iteration i). Runs exactly n_forced iterations; the checkers see the forced poses.
// ---- initReference, ICP.cpp:847-898: mean-centre the reference and build the matcher's tree. Kept as a handle so
// that one reference serves many compute() calls (Mapper.cpp:343 vs :373) and the baseline can time them apart.
struct OrcRef {
  std::vector<float> ref;
  std::vector<float> refn;
  bool has_n = false;
  int M = 0;
  float T_refIn_refMean[16];
  void* tree = nullptr;
};

extern "C" void* orc_icp_ref_create(const float* ref4xM, const float* refn3xM, int M, int mean_mode) {
  if (M <= 0) return nullptr;
  OrcRef* r = new OrcRef();
  r->M = M;
  r->ref.assign(ref4xM, ref4xM + 4 * (size_t)M);
  if (refn3xM) { r->refn.assign(refn3xM, refn3xM + 3 * (size_t)M); r->has_n = true; }
  float meanRef[3];
  meanOf(r->ref.data(), M, mean_mode, meanRef);
  eye4(r->T_refIn_refMean);
  for (int d = 0; d < 3; ++d) r->T_refIn_refMean[12 + d] = meanRef[d];
  for (int i = 0; i < M; ++i)
    for (int d = 0; d < 3; ++d) r->ref[4 * (size_t)i + d] -= meanRef[d];
  r->tree = orc_kdtree_build(r->ref.data(), M);
  return r;
}

extern "C" void orc_icp_ref_free(void* h) {
  OrcRef* r = (OrcRef*)h;
  if (!r) return;
  orc_kdtree_free(r->tree);
  delete r;
}

extern "C" int orc_icp_forced(const orc_icp_params* prm, const float* ref4xM, const float* refn3xM, int M,
                              const float* read4xN, const float* readn3xN, int N, const float* Tinit,
                              orc_icp_result* res, int max_log, float* logH, float* logb, float* logx, float* logT,
                              int* logP, int32_t* last_ids, float* last_dists, float* last_w, const float* forcedT,
                              int n_forced) {
  if (M <= 0 || N <= 0) return ORC_E_BAD_PARAM;
  void* h = orc_icp_ref_create(ref4xM, refn3xM, M, prm->mean_mode);
  const int st = orc_icp_compute(h, prm, read4xN, readn3xN, N, Tinit, res, max_log, logH, logb, logx, logT, logP,
                                 last_ids, last_dists, last_w, forcedT, n_forced);
  orc_icp_ref_free(h);
  return st;
}

// ICP::compute on a prepared reference (ICP.cpp:813-844 with initializeMatcherWithInputReference = false)
extern "C" int orc_icp_compute(void* href, const orc_icp_params* prm, const float* read4xN, const float* readn3xN,
                               int N, const float* Tinit, orc_icp_result* res, int max_log, float* logH, float* logb,
                               float* logx, float* logT, int* logP, int32_t* last_ids, float* last_dists,
                               float* last_w, const float* forcedT, int n_forced) {
  OrcRef* R = (OrcRef*)href;
  if (!R || N <= 0) return ORC_E_BAD_PARAM;
  const int k = prm->knn;
  const std::vector<float>& ref = R->ref;
  const float* refn3xM = R->has_n ? R->refn.data() : nullptr;
  const float* T_refIn_refMean = R->T_refIn_refMean;
  void* tree = R->tree;

  // ---- computeWithTransformedReference, ICP.cpp:902-1000
  std::vector<float> reading(read4xN, read4xN + 4 * (size_t)N);
  std::vector<float> readingN;
  if (readn3xN) readingN.assign(readn3xN, readn3xN + 3 * (size_t)N);
  float meanRead[3];
  meanOf(reading.data(), N, prm->mean_mode, meanRead);
  float T_readIn_readMean[16];
  eye4(T_readIn_readMean);
  for (int d = 0; d < 3; ++d) T_readIn_readMean[12 + d] = meanRead[d];
  for (int i = 0; i < N; ++i)
    for (int d = 0; d < 3; ++d) reading[4 * (size_t)i + d] -= meanRead[d];
  float invRef[16], T_refMean_readMean[16], tmp[16];
  inv4(T_refIn_refMean, invRef);
  mul4(invRef, Tinit, tmp);
  mul4(tmp, T_readIn_readMean, T_refMean_readMean);       // ICP.cpp:983
  orc_rigid_apply(T_refMean_readMean, reading.data(), readingN.empty() ? nullptr : readingN.data(), N);  // :984

  float T_refMean_dataIn[16];                              // :975
  mul4(invRef, Tinit, T_refMean_dataIn);
  XicpState xs;
  float T_iter[16];
  eye4(T_iter);
  bool iterate = true, maxReached = false;
  Checkers chk;
  chk.p = prm;
  chk.init(T_iter);                                        // :997

  std::vector<float> step(4 * (size_t)N), stepN;
  std::vector<int32_t> ids((size_t)k * N);
  std::vector<float> dists((size_t)k * N), w((size_t)k * N), w2((size_t)k * N);
  std::vector<float> readP(4 * (size_t)k * N), refP(4 * (size_t)k * N), refnP(3 * (size_t)k * N), wP((size_t)k * N);
  int iterationCount = 0, status = ORC_OK, P = 0;
  long touched = 0;
  float wpur = 0.f;
  while (iterate) {                                        // :1027
    step = reading;                                        // :1034
    if (!readingN.empty()) stepN = readingN;
    orc_rigid_apply(T_iter, step.data(), stepN.empty() ? nullptr : stepN.data(), N);  // :1047
    touched += orc_kdtree_knn(tree, step.data(), N, k, prm->maxDist, prm->knn_threads, ids.data(), dists.data());
    // outlier filter chain: element-wise product, OutlierFilter.cpp:84-99
    const long n = (long)k * N;
    bool first = true;
    auto chain = [&](int st) {
      if (st) { status = st; return; }
      if (first) { w = w2; first = false; }
      else for (long i = 0; i < n; ++i) w[i] = w[i] * w2[i];
    };
    if (prm->use_trimmed) chain(orc_outlier_trimmed(dists.data(), n, prm->trimmed_ratio, w2.data()));
    if (!status && prm->use_maxdist_filter) chain(orc_outlier_maxdist(dists.data(), n, prm->maxdist_filter, w2.data()));
    if (!status && prm->use_surface_normal_filter)
      chain(orc_outlier_surface_normal(stepN.empty() ? nullptr : stepN.data(), refn3xM, ids.data(), k, N,
                                       prm->surface_normal_max_angle, w2.data()));
    if (!status && first) orc_outlier_null(dists.data(), n, w.data());
    if (status) break;
    float pur;
    status = orc_error_elements(step.data(), ref.data(), refn3xM, w.data(), ids.data(), dists.data(), k, N,
                                readP.data(), refP.data(), refn3xM ? refnP.data() : nullptr, wP.data(), nullptr,
                                nullptr, &P, &pur, &wpur);  // :1061
    if (status) break;
    float H[36] = {0}, b[6] = {0}, x[6] = {0}, real[16];
    if (prm->minimizer == 0) {
      if (!refn3xM) { status = ORC_E_BAD_PARAM; break; }   // InvalidField "normals", ICP.cpp:1519
      orc_p2plane_hb(readP.data(), refP.data(), refnP.data(), wP.data(), P, prm->accum, H, b);  // :1071
      int info;
      if (prm->degeneracy == 1) {                          // :1112-1141
        if (!xicpDetectOptimized(prm, H, readP.data(), refnP.data(), P, T_refMean_dataIn, iterationCount == 0, xs))
          xs.returnPrior = true;
        if (xs.returnPrior) break;                         // :1175-1179
        xicpSolve(H, b, xs, x, &info, nullptr);            // :1205 -> PointToPlane.cpp:311-316
      } else {
        orc_solve6(H, b, x, &info);                        // :1205 -> PointToPlane.cpp:299
      }
      orc_p2plane_T(x, real);
    } else {
      orc_p2point_T2(readP.data(), refP.data(), wP.data(), P, prm->accum, real);
    }
    mul4(real, T_iter, T_iter);                            // :1213
    if (iterationCount < max_log) {
      if (logH) std::memcpy(logH + 36 * (size_t)iterationCount, H, sizeof(H));
      if (logb) std::memcpy(logb + 6 * (size_t)iterationCount, b, sizeof(b));
      if (logx) std::memcpy(logx + 6 * (size_t)iterationCount, x, sizeof(x));
      if (logT) std::memcpy(logT + 16 * (size_t)iterationCount, T_iter, sizeof(T_iter));
      if (logP) logP[iterationCount] = P;
    }
    if (forcedT) std::memcpy(T_iter, forcedT + 16 * (size_t)iterationCount, sizeof(T_iter));
    status = chk.check(T_iter, iterate, maxReached);       // :1292-1298
    ++iterationCount;
    if (status) break;
    if (forcedT) iterate = iterationCount < n_forced;
  }
  // ---- result, ICP.cpp:1345
  float invRead[16], r1[16], r2[16];
  inv4(T_readIn_readMean, invRead);
  mul4(T_refIn_refMean, T_iter, r1);
  mul4(r1, T_refMean_readMean, r2);
  mul4(r2, invRead, res->T);
  if (xs.returnPrior) std::memcpy(res->T, Tinit, 16 * sizeof(float));  // :1335-1341
  res->returned_prior = xs.returnPrior ? 1 : 0;
  for (int i = 0; i < 3; ++i) { res->loc_xyz[i] = xs.locXyz[i]; res->loc_rpy[i] = xs.locRpy[i]; }
  std::memcpy(res->eig_t, xs.Ut, sizeof(xs.Ut));
  std::memcpy(res->eig_r, xs.Ur, sizeof(xs.Ur));
  std::memcpy(res->xicp_sums, xs.sums, sizeof(xs.sums));
  res->iterations = iterationCount;
  res->max_iterations_reached = maxReached ? 1 : 0;
  res->n_matches_last = P;
  res->weighted_point_used_ratio = wpur;
  res->points_touched = touched;
  std::memcpy(res->T_iter, T_iter, sizeof(T_iter));
  std::memcpy(res->T_refIn_refMean, T_refIn_refMean, 16 * sizeof(float));
  std::memcpy(res->T_refMean_readMean, T_refMean_readMean, sizeof(T_refMean_readMean));
  if (last_ids) std::memcpy(last_ids, ids.data(), ids.size() * sizeof(int32_t));
  if (last_dists) std::memcpy(last_dists, dists.data(), dists.size() * sizeof(float));
  if (last_w) std::memcpy(last_w, w.data(), w.size() * sizeof(float));
  return status;
}
