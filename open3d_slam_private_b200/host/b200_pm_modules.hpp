// b200_pm_modules.hpp — libpointmatcher modules backed by the B200 registration core (include/b200reg.h).
//
// Header-only C++ shims, templated over the PointMatcher type `PM` (the reference instantiates them with
// PointMatcher<float>; this repo's self-test with pmmirror::PM). Each class derives from the reference's interface
// struct, keeps the reference module's parameter names, and forwards to the C ABI with the Eigen memory images
// passed straight through (features.data() is N x {x,y,z,1}; ids/dists/weights are k x N column-major). Status codes
// are mapped back to the exception types the reference throws. See INTEGRATION.md for the registrar lines.
//
//   B200KDTreeMatcher                  <- MatchersImpl<T>::KDTreeMatcher            (MatchersImpl.h:70-103, .cpp:60-102)
//   B200SurfaceNormalDataPointsFilter  <- SurfaceNormalDataPointsFilter<T>          (DataPointsFilters/SurfaceNormal.h/.cpp)
//   B200{MaxDist,MinDist,MedianDist,TrimmedDist,SurfaceNormal}OutlierFilter  <- OutlierFiltersImpl<T>::* (OutlierFiltersImpl.cpp)
//   B200PointToPlaneErrorMinimizer / B200PointToPointErrorMinimizer  <- ErrorMinimizers/PointToPlane.cpp, PointToPoint.cpp
//   B200ICP                            <- PointMatcher<T>::ICP::{initReference, compute, operator()} (ICP.cpp:793-1349)
#pragma once
#include <cmath>
#include <cstdlib>
#include <limits>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/b200reg.h"

namespace b200pm {

// ---- status -> exception (SURVEY.md §8(b) error convention) -----------------------------------------------------
template <class PM>
inline void check(int st) {
  if (st == B200_OK) return;
  const std::string msg = b200_last_error();
  switch (st) {
    case B200_E_NO_MATCHES:
    case B200_E_NAN: throw typename PM::ConvergenceError(msg);
    case B200_E_BAD_PARAM:
    case B200_E_UNSUPPORTED: throw typename PM::InvalidParameter(msg);
    default: throw std::runtime_error(msg);
  }
}

inline float param_f(const std::string& s) {
  if (s == "inf") return std::numeric_limits<float>::infinity();
  if (s == "-inf") return -std::numeric_limits<float>::infinity();
  return std::strtof(s.c_str(), nullptr);
}

// contiguous 3 x N copy of a 3-row descriptor block (descriptor rows of different labels are interleaved per column)
template <class PM>
inline bool pack_descriptor3(const typename PM::DataPoints& c, const std::string& name, std::vector<float>& out) {
  if (!c.descriptorExists(name)) return false;
  size_t span = 0;
  const size_t row = c.descriptorStartingRow(name, &span);
  if (span != 3) return false;
  const size_t n = (size_t)c.features.cols();
  out.resize(3 * n);
  for (size_t j = 0; j < n; ++j)
    for (size_t i = 0; i < 3; ++i) out[3 * j + i] = c.descriptors(row + i, j);
  return true;
}

template <class PM>
inline void require_3d(const typename PM::DataPoints& c) {
  if (c.features.rows() != 4) throw typename PM::InvalidParameter("B200 modules support 3D clouds (4 x N homogeneous features) only");
}

// ---- Matcher -----------------------------------------------------------------------------------------------------
template <class PM>
struct B200KDTreeMatcher : public PM::Matcher {
  static std::string description() { return "Exact k-nearest-neighbour matcher on a B200 (device-resident uniform-grid index); drop-in for KDTreeMatcher."; }
  const int knn;
  const float epsilon;     // accepted; the search is exact, which satisfies every epsilon
  const int searchType;    // accepted for YAML compatibility, ignored
  const float maxDist;
  const int device;

  explicit B200KDTreeMatcher(const typename PM::Parameters& params = typename PM::Parameters())
      : PM::Matcher(params),
        knn(std::atoi(this->get("knn", "1").c_str())), epsilon(param_f(this->get("epsilon", "0"))),
        searchType(std::atoi(this->get("searchType", "1").c_str())), maxDist(param_f(this->get("maxDist", "inf"))),
        device(std::atoi(this->get("device", "0").c_str())) {
    if (knn < 1) throw typename PM::InvalidParameter("B200KDTreeMatcher: knn must be >= 1");
  }
  ~B200KDTreeMatcher() override { b200_index_destroy(idx_); }
  B200KDTreeMatcher(const B200KDTreeMatcher&) = delete;
  B200KDTreeMatcher& operator=(const B200KDTreeMatcher&) = delete;

  void init(const typename PM::DataPoints& filteredReference) override {
    require_3d<PM>(filteredReference);
    b200_index_destroy(idx_);
    idx_ = nullptr;
    check<PM>(b200_index_create(&idx_, device, filteredReference.features.data(), nullptr,
                                (int)filteredReference.features.cols(), nullptr));
  }

  typename PM::Matches findClosests(const typename PM::DataPoints& filteredReading) override {
    if (!idx_) throw std::runtime_error("B200KDTreeMatcher: init() was not called");
    require_3d<PM>(filteredReading);
    const int n = (int)filteredReading.features.cols();
    typename PM::Matches m(knn, n);
    uint64_t visited = 0;
    check<PM>(b200_index_knn(idx_, filteredReading.features.data(), n, knn, epsilon, maxDist, nullptr, m.ids.data(),
                             m.dists.data(), &visited));
    this->visitCounter += (unsigned long)visited;
    return m;
  }

 private:
  b200_index* idx_ = nullptr;
};

// ---- DataPointsFilter ----------------------------------------------------------------------------------------------
template <class PM>
struct B200SurfaceNormalDataPointsFilter : public PM::DataPointsFilter {
  static std::string description() { return "Normals / densities / eigen-decomposition from k nearest neighbours on a B200; drop-in for SurfaceNormalDataPointsFilter."; }
  const int knn;
  const float maxDist, epsilon;
  const bool keepNormals, keepDensities, keepEigenValues, keepEigenVectors, keepMatchedIds, keepMeanDist, sortEigen,
      smoothNormals;
  const int device;

  explicit B200SurfaceNormalDataPointsFilter(const typename PM::Parameters& params = typename PM::Parameters())
      : PM::DataPointsFilter(params),
        knn(std::atoi(this->get("knn", "5").c_str())), maxDist(param_f(this->get("maxDist", "inf"))),
        epsilon(param_f(this->get("epsilon", "0"))), keepNormals(this->get("keepNormals", "1") != "0"),
        keepDensities(this->get("keepDensities", "0") != "0"), keepEigenValues(this->get("keepEigenValues", "0") != "0"),
        keepEigenVectors(this->get("keepEigenVectors", "0") != "0"), keepMatchedIds(this->get("keepMatchedIds", "0") != "0"),
        keepMeanDist(this->get("keepMeanDist", "0") != "0"), sortEigen(this->get("sortEigen", "0") != "0"),
        smoothNormals(this->get("smoothNormals", "0") != "0"), device(std::atoi(this->get("device", "0").c_str())) {}

  typename PM::DataPoints filter(const typename PM::DataPoints& input) override {
    typename PM::DataPoints out(input);
    inPlaceFilter(out);
    return out;
  }

  void inPlaceFilter(typename PM::DataPoints& cloud) override {
    require_3d<PM>(cloud);
    typedef typename PM::Matrix Matrix;
    const int n = (int)cloud.features.cols();
    Matrix normals, dens, eigva, eigve, matched, meand;
    if (keepNormals) normals.resize(3, n);
    if (keepDensities) dens.resize(1, n);
    if (keepEigenValues) eigva.resize(3, n);
    if (keepEigenVectors) eigve.resize(9, n);
    if (keepMatchedIds) matched.resize(knn, n);
    if (keepMeanDist) meand.resize(1, n);
    b200_surface_normal_params prm;
    prm.knn = knn; prm.maxDist = maxDist; prm.epsilon = epsilon; prm.sortEigen = sortEigen; prm.smoothNormals = smoothNormals;
    // the filter keeps its device index between scans: buffers are reused, nothing is allocated per call
    if (!idx_) check<PM>(b200_index_create(&idx_, device, cloud.features.data(), nullptr, n, nullptr));
    else check<PM>(b200_index_update(idx_, cloud.features.data(), nullptr, n, nullptr));
    check<PM>(b200_index_self_normals(idx_, &prm, keepNormals ? normals.data() : nullptr,
                                      keepDensities ? dens.data() : nullptr, keepEigenValues ? eigva.data() : nullptr,
                                      keepEigenVectors ? eigve.data() : nullptr, keepMatchedIds ? matched.data() : nullptr,
                                      keepMeanDist ? meand.data() : nullptr));
    // same descriptor names as SurfaceNormal.cpp:118-150
    if (keepNormals) cloud.addDescriptor("normals", normals);
    if (keepDensities) cloud.addDescriptor("densities", dens);
    if (keepEigenValues) cloud.addDescriptor("eigValues", eigva);
    if (keepEigenVectors) cloud.addDescriptor("eigVectors", eigve);
    if (keepMatchedIds) cloud.addDescriptor("matchedIds", matched);
    if (keepMeanDist) cloud.addDescriptor("meanDists", meand);
  }
  ~B200SurfaceNormalDataPointsFilter() override { b200_index_destroy(idx_); }
  B200SurfaceNormalDataPointsFilter(const B200SurfaceNormalDataPointsFilter&) = delete;
  B200SurfaceNormalDataPointsFilter& operator=(const B200SurfaceNormalDataPointsFilter&) = delete;

 private:
  b200_index* idx_ = nullptr;
};

// ---- OutlierFilters ------------------------------------------------------------------------------------------------
template <class PM>
struct B200DistOutlierFilter : public PM::OutlierFilter {
  const int kind;      // b200_outlier_kind
  const float param;
  const int device;
  B200DistOutlierFilter(int kind_, const char* pname, const char* def, const typename PM::Parameters& params)
      : PM::OutlierFilter(params), kind(kind_), param(param_f(this->get(pname, def))),
        device(std::atoi(this->get("device", "0").c_str())) {}
  typename PM::OutlierWeights compute(const typename PM::DataPoints&, const typename PM::DataPoints&,
                                      const typename PM::Matches& input) override {
    typename PM::OutlierWeights w(input.dists.rows(), input.dists.cols());
    check<PM>(b200_outlier_dist(device, kind, param, input.dists.data(),
                                (int64_t)input.dists.rows() * (int64_t)input.dists.cols(), w.data()));
    return w;
  }
};
// parameter names and defaults of OutlierFiltersImpl.h
template <class PM> struct B200MaxDistOutlierFilter : B200DistOutlierFilter<PM> {
  explicit B200MaxDistOutlierFilter(const typename PM::Parameters& p = typename PM::Parameters())
      : B200DistOutlierFilter<PM>(B200_OUTLIER_MAXDIST, "maxDist", "1", p) {}
};
template <class PM> struct B200MinDistOutlierFilter : B200DistOutlierFilter<PM> {
  explicit B200MinDistOutlierFilter(const typename PM::Parameters& p = typename PM::Parameters())
      : B200DistOutlierFilter<PM>(B200_OUTLIER_MINDIST, "minDist", "1", p) {}
};
template <class PM> struct B200MedianDistOutlierFilter : B200DistOutlierFilter<PM> {
  explicit B200MedianDistOutlierFilter(const typename PM::Parameters& p = typename PM::Parameters())
      : B200DistOutlierFilter<PM>(B200_OUTLIER_MEDIANDIST, "factor", "3", p) {}
};
template <class PM> struct B200TrimmedDistOutlierFilter : B200DistOutlierFilter<PM> {
  explicit B200TrimmedDistOutlierFilter(const typename PM::Parameters& p = typename PM::Parameters())
      : B200DistOutlierFilter<PM>(B200_OUTLIER_TRIMMEDDIST, "ratio", "0.85", p) {}
};

template <class PM>
struct B200SurfaceNormalOutlierFilter : public PM::OutlierFilter {
  const float maxAngle;
  const int device;
  explicit B200SurfaceNormalOutlierFilter(const typename PM::Parameters& params = typename PM::Parameters())
      : PM::OutlierFilter(params), maxAngle(param_f(this->get("maxAngle", "1.57"))),
        device(std::atoi(this->get("device", "0").c_str())) {}
  typename PM::OutlierWeights compute(const typename PM::DataPoints& reading, const typename PM::DataPoints& reference,
                                      const typename PM::Matches& input) override {
    std::vector<float> nread, nref;
    const bool has = pack_descriptor3<PM>(reading, "normals", nread) && pack_descriptor3<PM>(reference, "normals", nref);
    typename PM::OutlierWeights w(input.ids.rows(), input.ids.cols());
    check<PM>(b200_outlier_surface_normal(device, has ? nread.data() : nullptr, has ? nref.data() : nullptr,
                                          (int)reference.features.cols(), input.ids.data(), (int)input.ids.rows(),
                                          (int)input.ids.cols(), maxAngle, w.data()));
    return w;
  }
};

// ---- ErrorMinimizers -----------------------------------------------------------------------------------------------
// In this fork ICP::calculateOptimizationHessian (ICP.cpp:1513-1566) forms A/b and the minimiser solves them
// (degeneracyAwareness None); the shim does both on the device from the matched pairs it is handed.
template <class PM>
struct B200PointToPlaneErrorMinimizer : public PM::ErrorMinimizer {
  const int device;
  float H[36], b[6], x[6];
  int solveInfo = 0;
  explicit B200PointToPlaneErrorMinimizer(const typename PM::Parameters& params = typename PM::Parameters())
      : PM::ErrorMinimizer(params), device(std::atoi(this->get("device", "0").c_str())) {
    if (this->get("force2D", "0") != "0" || this->get("force4DOF", "0") != "0")
      throw typename PM::InvalidParameter("B200PointToPlaneErrorMinimizer supports force2D = 0 and force4DOF = 0 only");
  }
  typename PM::TransformationParameters compute(const typename PM::ErrorMinimizer::ErrorElements& m) override {
    require_3d<PM>(m.reading);
    std::vector<float> n;
    if (!pack_descriptor3<PM>(m.reference, "normals", n)) throw typename PM::InvalidField("Cannot find descriptor normals");
    const int P = (int)m.reading.features.cols();
    check<PM>(b200_p2plane_hessian(device, m.reading.features.data(), m.reference.features.data(), n.data(),
                                   m.weights.data(), P, H, b));
    typename PM::TransformationParameters T(4, 4);
    check<PM>(b200_p2plane_solve(device, H, b, x, T.data(), &solveInfo));
    return T;
  }
};

template <class PM>
struct B200PointToPointErrorMinimizer : public PM::ErrorMinimizer {
  const int device;
  explicit B200PointToPointErrorMinimizer(const typename PM::Parameters& params = typename PM::Parameters())
      : PM::ErrorMinimizer(params), device(std::atoi(this->get("device", "0").c_str())) {}
  typename PM::TransformationParameters compute(const typename PM::ErrorMinimizer::ErrorElements& m) override {
    require_3d<PM>(m.reading);
    typename PM::TransformationParameters T(4, 4);
    check<PM>(b200_p2point(device, m.reading.features.data(), m.reference.features.data(), m.weights.data(),
                           (int)m.reading.features.cols(), T.data()));
    return T;
  }
};

// ---- whole chain ---------------------------------------------------------------------------------------------------
// Same call shape as PM::ICP (PointMatcher.h:1027-1042): initReference once (Mapper.cpp:343), compute per scan
// (Mapper.cpp:373) or operator() for a one-shot registration. All iterations run in one device kernel; matches and
// outlier weights of the last iteration are fetched on demand.
template <class PM>
struct B200ICP {
  b200_icp_params prm;
  b200_icp_result last;
  explicit B200ICP(int device = 0) : device_(device) { b200_icp_default_params(&prm); }
  ~B200ICP() { b200_icp_destroy(h_); }
  B200ICP(const B200ICP&) = delete;
  B200ICP& operator=(const B200ICP&) = delete;

  // "degeneracyAwareness" node of the ICP YAML (ICP::loadAdditionalYAMLContent, ICP.cpp:575-760): method name and its
  // string parameters, e.g. ("OptimizedEqualityConstraints", {{"enoughInformationThreshold", "250"}, ...}) as in
  // open3d_slam_ros/param/icp.yaml. Returns false exactly where the reference's loader does (a parameter missing,
  // ICP.cpp:636-669); methods that are not built throw InvalidParameter. Call before initReference.
  bool setDegeneracyAwareness(const std::string& method, const std::map<std::string, std::string>& params) {
    if (method == "None") { prm.degeneracy = B200_DEGENERACY_NONE; return true; }
    if (method == "OptimizedEqualityConstraints") {
      const char* keys[4] = {"enoughInformationThreshold", "insufficientInformationThreshold",
                             "point2NormalMinimalAlignmentAngleThreshold", "point2NormalStrongAlignmentAngleThreshold"};
      float v[4];
      for (int i = 0; i < 4; ++i) {
        const auto it = params.find(keys[i]);
        if (it == params.end()) return false;
        v[i] = std::strtof(it->second.c_str(), nullptr);
      }
      prm.degeneracy = B200_DEGENERACY_OPTIMIZED_EQUALITY;
      prm.xicp_enough_information = v[0]; prm.xicp_insufficient_information = v[1];
      prm.xicp_min_alignment_angle_deg = v[2]; prm.xicp_strong_alignment_angle_deg = v[3];
      return true;
    }
    throw typename PM::InvalidParameter("degeneracyAwareness: " + method + " is not built (None, OptimizedEqualityConstraints)");
  }
  // LocalizabilityAnalysisResults::localizabilityXyz_/Rpy_ of the last compute (PointMatcher.h:638-639), 1 = localizable
  void localizability(int xyz[3], int rpy[3]) const {
    for (int i = 0; i < 3; ++i) { xyz[i] = last.loc_xyz[i]; rpy[i] = last.loc_rpy[i]; }
  }

  bool initReference(const typename PM::DataPoints& referenceIn) {
    require_3d<PM>(referenceIn);
    ensure();
    std::vector<float> n;
    const bool hasN = pack_descriptor3<PM>(referenceIn, "normals", n);
    const int st = b200_icp_init_reference(h_, referenceIn.features.data(), hasN ? n.data() : nullptr,
                                           (int)referenceIn.features.cols());
    if (st == B200_E_NOT_INITIALIZED) return false;  // "The reference point cloud is empty", ICP.cpp:850-855
    check<PM>(st);
    return true;
  }
  typename PM::TransformationParameters compute(const typename PM::DataPoints& readingIn,
                                                const typename PM::TransformationParameters& T_refIn_dataIn) {
    require_3d<PM>(readingIn);
    ensure();
    std::vector<float> n;
    const bool hasN = pack_descriptor3<PM>(readingIn, "normals", n);
    nRead_ = (int)readingIn.features.cols();
    check<PM>(b200_icp_compute(h_, readingIn.features.data(), hasN ? n.data() : nullptr, nRead_, T_refIn_dataIn.data(), &last));
    typename PM::TransformationParameters T(4, 4);
    for (int i = 0; i < 16; ++i) T.data()[i] = last.T[i];
    return T;
  }
  typename PM::TransformationParameters operator()(const typename PM::DataPoints& readingIn,
                                                   const typename PM::DataPoints& referenceIn) {
    if (!initReference(referenceIn)) throw std::runtime_error("The reference point cloud is empty");
    return compute(readingIn, PM::TransformationParameters::Identity(4, 4));
  }
  typename PM::Matches matches(typename PM::OutlierWeights* weights = nullptr) {
    typename PM::Matches m(prm.knn, nRead_);
    if (weights) weights->resize(prm.knn, nRead_);
    check<PM>(b200_icp_get_matches(h_, m.ids.data(), m.dists.data(), weights ? weights->data() : nullptr));
    return m;
  }

 private:
  void ensure() {
    if (!h_) check<PM>(b200_icp_create(&h_, device_, &prm));
  }
  b200_icp* h_ = nullptr;
  int device_ = 0, nRead_ = 0;
};

}  // namespace b200pm
