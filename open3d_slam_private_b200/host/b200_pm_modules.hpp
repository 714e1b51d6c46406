// b200_pm_modules.hpp — libpointmatcher modules backed by the B200 registration core (include/b200reg.h).
//
// Header-only C++ shims written against libpointmatcher's OWN interface: include pointmatcher/PointMatcher.h before
// this header in the reference tree (host/pm_mirror.h here, a literal mirror of the same names — there is no #if
// between the two). Each class derives from the reference's interface struct with the reference's constructor
// convention `C(const Parameters& params)` -> `Interface(className, C::availableParameters(), params)`, supplies the
// static description() / availableParameters() that Registrar::GenericClassDescriptor<C> needs (Registrar.h:95-123),
// reads every parameter through Parametrizable::get<S>() so that parametersUsed is filled (a parameter that is set
// but never read makes the registrar throw, Registrar.h:103-111), and forwards to the C ABI with the Eigen memory
// images passed straight through (features.data() is N x {x,y,z,1}; ids/dists/weights are k x N column-major).
// Status codes are mapped back to the exception types the reference throws. Registrar lines: INTEGRATION.md.
//
//   B200KDTreeMatcher                  <- MatchersImpl<T>::KDTreeMatcher            (MatchersImpl.h:70-103, .cpp:60-102)
//   B200SurfaceNormalDataPointsFilter  <- SurfaceNormalDataPointsFilter<T>          (DataPointsFilters/SurfaceNormal.h/.cpp)
//   B200{MaxDist,MinDist,MedianDist,TrimmedDist,SurfaceNormal}OutlierFilter  <- OutlierFiltersImpl<T>::* (OutlierFiltersImpl.h/.cpp)
//   B200PointToPlaneErrorMinimizer / B200PointToPointErrorMinimizer  <- ErrorMinimizers/PointToPlane.{h,cpp}, PointToPoint.{h,cpp}
//   B200ICP                            <- PointMatcher<T>::ICP::{initReference, compute, operator()} (PointMatcher.h:1027-1042, ICP.cpp:793-1349)
// Every module takes one parameter the reference's does not have: "device" (CUDA device ordinal, default 0).
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

typedef PointMatcherSupport::Parametrizable P;

// ---- status -> exception (SURVEY.md §8(b) error convention) -----------------------------------------------------
template <class PM>
inline void check(int st) {
  if (st == B200_OK) return;
  const std::string msg = b200_last_error();
  switch (st) {
    case B200_E_NO_MATCHES:
    case B200_E_NAN:
    case B200_E_BOUND: throw typename PM::ConvergenceError(msg);
    case B200_E_BAD_PARAM:
    case B200_E_UNSUPPORTED: throw P::InvalidParameter(msg);
    default: throw std::runtime_error(msg);
  }
}

inline P::ParameterDoc device_doc() {
  return {"device", "CUDA device ordinal the module runs on (B200 modules only)", "0", "0", "2147483647", &P::Comp<unsigned>};
}

// contiguous 3 x N copy of a 3-row descriptor block (descriptor rows of different labels are interleaved per column)
template <class PM>
inline bool pack_descriptor3(const typename PM::DataPoints& c, const std::string& name, std::vector<float>& out) {
  if (!c.descriptorExists(name) || c.getDescriptorDimension(name) != 3) return false;
  const size_t row = c.getDescriptorStartingRow(name);
  const size_t n = (size_t)c.features.cols();
  out.resize(3 * n);
  for (size_t j = 0; j < n; ++j)
    for (size_t i = 0; i < 3; ++i) out[3 * j + i] = c.descriptors(row + i, j);
  return true;
}

template <class PM>
inline void require_3d(const typename PM::DataPoints& c) {
  if (c.features.rows() != 4) throw P::InvalidParameter("B200 modules support 3D clouds (4 x N homogeneous features) only");
}

// ---- Matcher -----------------------------------------------------------------------------------------------------
template <class PM>
struct B200KDTreeMatcher : public PM::Matcher {
  typedef typename PM::Parameters Parameters;
  typedef typename PM::ParametersDoc ParametersDoc;
  inline static const std::string description() {
    return "Exact k-nearest-neighbour matcher on a B200 (device-resident uniform-grid index); drop-in for KDTreeMatcher.";
  }
  // names, defaults and bounds of KDTreeMatcher (MatchersImpl.h:81-88)
  inline static const ParametersDoc availableParameters() {
    return {
        {"knn", "number of nearest neighbors to consider it the reference", "1", "1", "2147483647", &P::Comp<unsigned>},
        {"epsilon", "approximation to use for the nearest-neighbor search (accepted; the search is exact, which satisfies every epsilon)", "0", "0", "inf", &P::Comp<float>},
        {"searchType", "Nabo search type (accepted for YAML compatibility, ignored)", "1", "0", "2", &P::Comp<unsigned>},
        {"maxDist", "maximum distance to consider for neighbors", "inf", "0", "inf", &P::Comp<float>},
        device_doc()};
  }
  const int knn;
  const float epsilon;
  const int searchType;
  const float maxDist;
  const int device;

  B200KDTreeMatcher(const Parameters& params = Parameters())
      : PM::Matcher("B200KDTreeMatcher", B200KDTreeMatcher::availableParameters(), params),
        knn(P::get<int>("knn")), epsilon(P::get<float>("epsilon")), searchType(P::get<int>("searchType")),
        maxDist(P::get<float>("maxDist")), device(P::get<int>("device")) {}
  virtual ~B200KDTreeMatcher() { b200_index_destroy(idx_); }
  B200KDTreeMatcher(const B200KDTreeMatcher&) = delete;
  B200KDTreeMatcher& operator=(const B200KDTreeMatcher&) = delete;

  virtual void init(const typename PM::DataPoints& filteredReference) {
    require_3d<PM>(filteredReference);
    b200_index_destroy(idx_);
    idx_ = nullptr;
    check<PM>(b200_index_create(&idx_, device, filteredReference.features.data(), nullptr,
                                (int)filteredReference.features.cols(), nullptr));
  }

  virtual typename PM::Matches findClosests(const typename PM::DataPoints& filteredReading) {
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
  typedef typename PM::Parameters Parameters;
  typedef typename PM::ParametersDoc ParametersDoc;
  inline static const std::string description() {
    return "Normals / densities / eigen-decomposition from k nearest neighbours on a B200; drop-in for SurfaceNormalDataPointsFilter.";
  }
  // the 11 parameters of SurfaceNormalDataPointsFilter (SurfaceNormal.h:65-81)
  inline static const ParametersDoc availableParameters() {
    return {
        {"knn", "number of nearest neighbors to consider, including the point itself", "5", "3", "2147483647", &P::Comp<unsigned>},
        {"maxDist", "maximum distance to consider for neighbors", "inf", "0", "inf", &P::Comp<float>},
        {"epsilon", "approximation to use for the nearest-neighbor search", "0", "0", "inf", &P::Comp<float>},
        {"keepNormals", "whether the normals should be added as descriptors to the resulting cloud", "1"},
        {"keepDensities", "whether the point densities should be added as descriptors to the resulting cloud", "0"},
        {"keepEigenValues", "whether the eigen values should be added as descriptors to the resulting cloud", "0"},
        {"keepEigenVectors", "whether the eigen vectors should be added as descriptors to the resulting cloud", "0"},
        {"keepMatchedIds", "whether the identifiers of matches points should be added as descriptors to the resulting cloud", "0"},
        {"keepMeanDist", "whether the distance to the nearest neighbor mean should be added as descriptors to the resulting cloud", "0"},
        {"sortEigen", "whether the eigenvalues and eigenvectors should be sorted (ascending) based on the eigenvalues", "0"},
        {"smoothNormals", "whether the normal vector should be average with the nearest neighbors", "0"},
        device_doc()};
  }
  const unsigned knn;
  const float maxDist, epsilon;
  const bool keepNormals, keepDensities, keepEigenValues, keepEigenVectors, keepMatchedIds, keepMeanDist, sortEigen,
      smoothNormals;
  const int device;

  B200SurfaceNormalDataPointsFilter(const Parameters& params = Parameters())
      : PM::DataPointsFilter("B200SurfaceNormalDataPointsFilter", B200SurfaceNormalDataPointsFilter::availableParameters(), params),
        knn(P::get<unsigned>("knn")), maxDist(P::get<float>("maxDist")), epsilon(P::get<float>("epsilon")),
        keepNormals(P::get<bool>("keepNormals")), keepDensities(P::get<bool>("keepDensities")),
        keepEigenValues(P::get<bool>("keepEigenValues")), keepEigenVectors(P::get<bool>("keepEigenVectors")),
        keepMatchedIds(P::get<bool>("keepMatchedIds")), keepMeanDist(P::get<bool>("keepMeanDist")),
        sortEigen(P::get<bool>("sortEigen")), smoothNormals(P::get<bool>("smoothNormals")), device(P::get<int>("device")) {}
  virtual ~B200SurfaceNormalDataPointsFilter() { b200_index_destroy(idx_); }
  B200SurfaceNormalDataPointsFilter(const B200SurfaceNormalDataPointsFilter&) = delete;
  B200SurfaceNormalDataPointsFilter& operator=(const B200SurfaceNormalDataPointsFilter&) = delete;

  virtual typename PM::DataPoints filter(const typename PM::DataPoints& input) {
    typename PM::DataPoints out(input);
    inPlaceFilter(out);
    return out;
  }

  virtual void inPlaceFilter(typename PM::DataPoints& cloud) {
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
    prm.knn = (int)knn; prm.maxDist = maxDist; prm.epsilon = epsilon; prm.sortEigen = sortEigen; prm.smoothNormals = smoothNormals;
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

 private:
  b200_index* idx_ = nullptr;
};

// ---- OutlierFilters ------------------------------------------------------------------------------------------------
// one helper for the four distance filters; the concrete classes carry the reference's names, parameter names,
// defaults and bounds (OutlierFiltersImpl.h:76-81, 96-101, 115-120, 134-139)
template <class PM>
struct B200DistOutlierFilter : public PM::OutlierFilter {
  const int kind;      // b200_outlier_kind
  const float param;
  const int device;
  B200DistOutlierFilter(const std::string& className, const typename PM::ParametersDoc doc, int kind_, const char* pname,
                        const typename PM::Parameters& params)
      : PM::OutlierFilter(className, doc, params), kind(kind_), param(P::get<float>(pname)), device(P::get<int>("device")) {}
  virtual typename PM::OutlierWeights compute(const typename PM::DataPoints&, const typename PM::DataPoints&,
                                              const typename PM::Matches& input) {
    typename PM::OutlierWeights w(input.dists.rows(), input.dists.cols());
    check<PM>(b200_outlier_dist(device, kind, param, input.dists.data(),
                                (int64_t)input.dists.rows() * (int64_t)input.dists.cols(), w.data()));
    return w;
  }
};
#define B200_DIST_FILTER(Class, Kind, pname, pdoc, pdef, pmin, pmax, text)                                              \
  template <class PM> struct Class : B200DistOutlierFilter<PM> {                                                        \
    inline static const std::string description() { return text; }                                                     \
    inline static const typename PM::ParametersDoc availableParameters() {                                             \
      return {{pname, pdoc, pdef, pmin, pmax, &P::Comp<float>}, device_doc()};                                         \
    }                                                                                                                   \
    Class(const typename PM::Parameters& params = typename PM::Parameters())                                           \
        : B200DistOutlierFilter<PM>(#Class, Class::availableParameters(), Kind, pname, params) {}                      \
  };
B200_DIST_FILTER(B200MaxDistOutlierFilter, B200_OUTLIER_MAXDIST, "maxDist", "threshold distance (Euclidean norm)", "1",
                 "0.0000001", "inf", "MaxDistOutlierFilter on a B200: links longer than maxDist get weight 0.")
B200_DIST_FILTER(B200MinDistOutlierFilter, B200_OUTLIER_MINDIST, "minDist", "threshold distance (Euclidean norm)", "1",
                 "0.0000001", "inf", "MinDistOutlierFilter on a B200: links shorter than minDist get weight 0.")
B200_DIST_FILTER(B200MedianDistOutlierFilter, B200_OUTLIER_MEDIANDIST, "factor",
                 "points farther away factor * median will be considered outliers.", "3", "0.0000001", "inf",
                 "MedianDistOutlierFilter on a B200.")
B200_DIST_FILTER(B200TrimmedDistOutlierFilter, B200_OUTLIER_TRIMMEDDIST, "ratio", "percentage to keep", "0.85",
                 "0.0000001", "1.0", "TrimmedDistOutlierFilter on a B200 (exact quantile by radix select).")
#undef B200_DIST_FILTER

template <class PM>
struct B200SurfaceNormalOutlierFilter : public PM::OutlierFilter {
  inline static const std::string description() {
    return "SurfaceNormalOutlierFilter on a B200: weight 0 when the angle between the two normals exceeds maxAngle.";
  }
  inline static const typename PM::ParametersDoc availableParameters() {   // OutlierFiltersImpl.h:182-187
    return {{"maxAngle", "Maximum authorised angle between the 2 surface normals (in radian)", "1.57", "0.0", "3.1416", &P::Comp<float>},
            device_doc()};
  }
  const float maxAngle;
  const int device;
  B200SurfaceNormalOutlierFilter(const typename PM::Parameters& params = typename PM::Parameters())
      : PM::OutlierFilter("B200SurfaceNormalOutlierFilter", B200SurfaceNormalOutlierFilter::availableParameters(), params),
        maxAngle(P::get<float>("maxAngle")), device(P::get<int>("device")) {}
  virtual typename PM::OutlierWeights compute(const typename PM::DataPoints& reading, const typename PM::DataPoints& reference,
                                              const typename PM::Matches& input) {
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
// This fork's pure virtual is compute(const ErrorElements&, LocalizabilityParametersForErrorMinimization&)
// (PointMatcher.h:790): ICP::calculateOptimizationHessian (ICP.cpp:1513-1566) has filled A_ / b_ before the call and
// PointToPlaneErrorMinimizer::compute_in_place solves THEM (PointToPlane.cpp:290-300). The shim does the same on the
// device (b200_p2plane_solve = solvePossiblyUnderdeterminedLinearSystem + the SE(3) construction of :337-381).
// accumulate() is the device stand-in for calculateOptimizationHessian for callers that want A_ / b_ formed there too.
template <class PM>
struct B200PointToPlaneErrorMinimizer : public PM::ErrorMinimizer {
  typedef typename PM::ErrorMinimizer::ErrorElements ErrorElements;
  typedef typename PM::LocalizabilityParametersForErrorMinimization LocalizabilityParametersForErrorMinimization;
  inline static const std::string description() {
    return "Point-to-plane error (6x6 solve and SE(3) construction on a B200); drop-in for PointToPlaneErrorMinimizer, force2D = force4DOF = 0.";
  }
  inline static const typename PM::ParametersDoc availableParameters() {   // ErrorMinimizers/PointToPlane.h:70-77
    return {{"force2D", "If set to true(1), the minimization will be forced to give a solution in 2D (i.e., on the XY-plane) even with 3D inputs.", "0", "0", "1", &P::Comp<bool>},
            {"force4DOF", "If set to true(1), the minimization will optimize only yaw and translation, pitch and roll will follow the prior.", "0", "0", "1", &P::Comp<bool>},
            device_doc()};
  }
  const bool force2D, force4DOF;
  const int device;
  float x[6];
  int solveInfo = 0;   // 0 full rank, 1 rank-deficient minimal norm, 2 double-SVD fallback (PointToPlane.cpp:199-256)
  B200PointToPlaneErrorMinimizer(const typename PM::Parameters& params = typename PM::Parameters())
      : PM::ErrorMinimizer("B200PointToPlaneErrorMinimizer", B200PointToPlaneErrorMinimizer::availableParameters(), params),
        force2D(P::get<bool>("force2D")), force4DOF(P::get<bool>("force4DOF")), device(P::get<int>("device")) {
    if (force2D || force4DOF)
      throw P::InvalidParameter("B200PointToPlaneErrorMinimizer supports force2D = 0 and force4DOF = 0 only");
  }
  // A_ = sum w F F^T, b_ = -sum w F r over the matched pairs (ICP.cpp:1527-1565), formed on the device
  void accumulate(const ErrorElements& m, LocalizabilityParametersForErrorMinimization& lp) {
    require_3d<PM>(m.reading);
    std::vector<float> n;
    if (!pack_descriptor3<PM>(m.reference, "normals", n)) throw typename PM::DataPoints::InvalidField("Cannot find descriptor normals");
    lp.A_.resize(6, 6);
    lp.b_.resize(6, 1);
    check<PM>(b200_p2plane_hessian(device, m.reading.features.data(), m.reference.features.data(), n.data(),
                                   m.weights.data(), (int)m.reading.features.cols(), lp.A_.data(), lp.b_.data()));
  }
  virtual typename PM::TransformationParameters compute(const ErrorElements& mPts,
                                                        LocalizabilityParametersForErrorMinimization& lp) {
    require_3d<PM>(mPts.reading);
    if (lp.degeneracyAwarenessMethod != PM::DegeneracyAwarenessMethod::kNone)
      throw P::InvalidParameter("B200PointToPlaneErrorMinimizer: constrained solves run inside B200ICP (setDegeneracyAwareness)");
    if (lp.A_.rows() != 6 || lp.A_.cols() != 6 || lp.b_.rows() != 6)
      throw P::InvalidParameter("B200PointToPlaneErrorMinimizer: A_ (6x6) / b_ (6x1) were not set");
    typename PM::TransformationParameters T(4, 4);
    check<PM>(b200_p2plane_solve(device, lp.A_.data(), lp.b_.data(), x, T.data(), &solveInfo));
    return T;
  }
};

template <class PM>
struct B200PointToPointErrorMinimizer : public PM::ErrorMinimizer {
  typedef typename PM::ErrorMinimizer::ErrorElements ErrorElements;
  typedef typename PM::LocalizabilityParametersForErrorMinimization LocalizabilityParametersForErrorMinimization;
  inline static const std::string description() { return "Point-to-point error (weighted Kabsch / SVD on a B200); drop-in for PointToPointErrorMinimizer."; }
  inline static const typename PM::ParametersDoc availableParameters() { return {device_doc()}; }
  const int device;
  B200PointToPointErrorMinimizer(const typename PM::Parameters& params = typename PM::Parameters())
      : PM::ErrorMinimizer("B200PointToPointErrorMinimizer", B200PointToPointErrorMinimizer::availableParameters(), params),
        device(P::get<int>("device")) {}
  virtual typename PM::TransformationParameters compute(const ErrorElements& m, LocalizabilityParametersForErrorMinimization&) {
    require_3d<PM>(m.reading);
    typename PM::TransformationParameters T(4, 4);
    check<PM>(b200_p2point(device, m.reading.features.data(), m.reference.features.data(), m.weights.data(),
                           (int)m.reading.features.cols(), T.data()));
    return T;
  }
};

// ---- whole chain ---------------------------------------------------------------------------------------------------
// Same call shape as PM::ICP (PointMatcher.h:1027-1042): initReference once (Mapper.cpp:343), compute per scan —
// `icp_.compute(cloud, <unused>, init, false)` is the call Mapper.cpp:372-373 makes — or operator() for a one-shot
// registration. All iterations run in one device kernel; matches and outlier weights of the last iteration are
// fetched on demand.
template <class PM>
struct B200ICP {
  typedef typename PM::DataPoints DataPoints;
  typedef typename PM::TransformationParameters TransformationParameters;
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
        v[i] = PointMatcherSupport::lexical_cast<float>(it->second);
      }
      prm.degeneracy = B200_DEGENERACY_OPTIMIZED_EQUALITY;
      prm.xicp_enough_information = v[0]; prm.xicp_insufficient_information = v[1];
      prm.xicp_min_alignment_angle_deg = v[2]; prm.xicp_strong_alignment_angle_deg = v[3];
      return true;
    }
    throw P::InvalidParameter("degeneracyAwareness: " + method + " is not built (None, OptimizedEqualityConstraints)");
  }
  // LocalizabilityAnalysisResults::localizabilityXyz_/Rpy_ of the last compute (PointMatcher.h:638-639), 1 = localizable
  void localizability(int xyz[3], int rpy[3]) const {
    for (int i = 0; i < 3; ++i) { xyz[i] = last.loc_xyz[i]; rpy[i] = last.loc_rpy[i]; }
  }

  bool initReference(const DataPoints& referenceIn) {                                   // ICP.cpp:847-898
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
  // PM::ICP::compute(readingIn, referenceIn, T_refIn_dataIn, initializeMatcherWithInputReference = true)
  // (PointMatcher.h:1037-1041, ICP.cpp:813-844): with the flag set the reference is (re)initialised from referenceIn
  // first, otherwise referenceIn is unused and the reference of the last initReference stands.
  TransformationParameters compute(const DataPoints& readingIn, const DataPoints& referenceIn,
                                   const TransformationParameters& T_refIn_dataIn,
                                   const bool initializeMatcherWithInputReference = true) {
    if (initializeMatcherWithInputReference && !initReference(referenceIn))
      throw std::runtime_error("The reference point cloud is empty");
    require_3d<PM>(readingIn);
    ensure();
    std::vector<float> n;
    const bool hasN = pack_descriptor3<PM>(readingIn, "normals", n);
    nRead_ = (int)readingIn.features.cols();
    check<PM>(b200_icp_compute(h_, readingIn.features.data(), hasN ? n.data() : nullptr, nRead_, T_refIn_dataIn.data(), &last));
    TransformationParameters T(4, 4);
    for (int i = 0; i < 16; ++i) T.data()[i] = last.T[i];
    return T;
  }
  TransformationParameters operator()(const DataPoints& readingIn, const DataPoints& referenceIn) {   // ICP.cpp:793-801
    return compute(readingIn, referenceIn, TransformationParameters::Identity(4, 4), true);
  }
  TransformationParameters operator()(const DataPoints& readingIn, const DataPoints& referenceIn,
                                      const TransformationParameters& initialTransformationParameters) {   // :803-811
    return compute(readingIn, referenceIn, initialTransformationParameters, true);
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
