// b200_cloud_registration.hpp — boundary B: a registration backend for open3d_slam's CloudRegistration interface
// (open3d_slam/include/open3d_slam/CloudRegistration.hpp:19-28; factory src/CloudRegistration.cpp:105-121) backed by
// the B200 registration core.
//
// Header-only and templated so that it compiles without Open3D: in the reference tree
//   Cloud  = open3d::geometry::PointCloud   (points_, normals_ : std::vector<Eigen::Vector3d>)
//   Result = open3d::pipelines::registration::RegistrationResult (transformation_, correspondence_set_, fitness_,
//            inlier_rmse_)
//   Mat4   = Eigen::Matrix4d (Transform::matrix())
// and the class below is wrapped by a 10-line subclass of o3d_slam::CloudRegistration. Clouds are
// converted double -> float exactly like open3dToPointmatcher (open3d_conversions.cpp:57-118).
//
// Semantics: Open3D's own loop (b200_o3d_* in include/b200reg.h): 1-NN within maxCorrespondenceDistance_, the
// TransformationEstimation{PointToPlane, PointToPoint, ForGeneralizedICP} update, T <- update * T in double, and the
// ICPConvergenceCriteria stopping rule (relative fitness / relative RMSE / max_iteration). fitness_, inlier_rmse_ and
// correspondence_set_ are those of the last match, as in RegistrationResult.
#pragma once
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/b200reg.h"

namespace b200o3d {

enum class Kind { PointToPlane = 0, PointToPoint = 1, Generalized = 2 };

template <class Cloud, class Mat4, class Result>
class RegistrationIcpB200 {
 public:
  double maxCorrespondenceDistance_ = 1.0;   // IcpParameters::maxCorrespondenceDistance_ (Parameters.hpp:66-72)
  int maxNumIter_ = 50;                      // IcpParameters::maxNumIter_ -> ICPConvergenceCriteria::max_iteration_
  double relativeFitness_ = 1e-6;            // ICPConvergenceCriteria::relative_fitness_ (Open3D default)
  double relativeRmse_ = 1e-6;               // ICPConvergenceCriteria::relative_rmse_ (Open3D default)
  double gicpEpsilon_ = 1e-3;                // TransformationEstimationForGeneralizedICP::epsilon_
  int knnNormalEstimation_ = 10;             // IcpParameters::knn_
  double maxRadiusNormalEstimation_ = 2.0;   // IcpParameters::maxDistanceKnn_
  Kind kind_ = Kind::PointToPlane;
  int device_ = 0;

  RegistrationIcpB200() = default;
  ~RegistrationIcpB200() { b200_o3d_destroy(h_); }
  RegistrationIcpB200(const RegistrationIcpB200&) = delete;
  RegistrationIcpB200& operator=(const RegistrationIcpB200&) = delete;

  // CloudRegistration::registerClouds(source, target, init)
  Result registerClouds(const Cloud& source, const Cloud& target, const Mat4& init) {
    ensure();
    std::vector<float> tgt4, tgtn, src4, srcn;
    to_feat4(target.points_, tgt4);
    const bool tgtHasN = target.normals_.size() == target.points_.size() && !target.normals_.empty();
    const bool srcHasN = source.normals_.size() == source.points_.size() && !source.normals_.empty();
    if (kind_ != Kind::PointToPoint && !tgtHasN) throw std::runtime_error("RegistrationIcpB200: target cloud has no normals");
    if (kind_ == Kind::Generalized && !srcHasN) throw std::runtime_error("RegistrationIcpB200: source cloud has no normals");
    if (tgtHasN) to_xyz3(target.normals_, tgtn);
    if (srcHasN && kind_ == Kind::Generalized) to_xyz3(source.normals_, srcn);
    to_feat4(source.points_, src4);
    check(b200_o3d_set_target(h_, tgt4.data(), tgtHasN ? tgtn.data() : nullptr, (int)target.points_.size()));
    double Ti[16];
    for (int j = 0; j < 4; ++j)
      for (int i = 0; i < 4; ++i) Ti[i + 4 * j] = (double)init(i, j);
    b200_o3d_result r;
    const int n = (int)source.points_.size();
    check(b200_o3d_register(h_, src4.data(), srcn.empty() ? nullptr : srcn.data(), n, Ti, &r));
    Result out;
    for (int j = 0; j < 4; ++j)
      for (int i = 0; i < 4; ++i) out.transformation_(i, j) = r.T[i + 4 * j];
    std::vector<int32_t> ids(n);
    check(b200_o3d_get_correspondences(h_, ids.data(), nullptr));
    out.correspondence_set_.clear();
    for (int i = 0; i < n; ++i)
      if (ids[i] >= 0) out.correspondence_set_.emplace_back(i, ids[i]);
    out.fitness_ = r.fitness;
    out.inlier_rmse_ = r.inlier_rmse;
    lastIterations_ = r.iterations;
    return out;
  }
  int lastIterations_ = 0;

  // CloudRegistration::estimateNormalsOrCovariancesIfNeeded (CloudRegistration.cpp:27-46, 67-79): nothing when the
  // cloud has normals or is empty; else Open3D's EstimateNormals(KDTreeSearchParamHybrid(maxRadiusNormalEstimation_,
  // knnNormalEstimation_)) + NormalizeNormals + OrientNormalsTowardsCameraLocation on the device
  // (b200_o3d_estimate_normals: double covariance normalised by the count, < 3 neighbours -> (0, 0, 1)).
  void estimateNormalsOrCovariancesIfNeeded(Cloud* cloud) {
    if (!cloud || cloud->points_.empty()) return;
    if (cloud->normals_.size() == cloud->points_.size()) return;   // HasNormals()
    ensure();
    std::vector<float> f4;
    to_feat4(cloud->points_, f4);
    const int n = (int)cloud->points_.size();
    std::vector<double> nrm(3 * (size_t)n);
    check(b200_o3d_estimate_normals(h_, f4.data(), n, knnNormalEstimation_, (float)maxRadiusNormalEstimation_, nrm.data()));
    cloud->normals_.resize(n);
    for (int i = 0; i < n; ++i)
      for (int d = 0; d < 3; ++d) cloud->normals_[i][d] = nrm[3 * (size_t)i + d];
  }

 private:
  static void check(int st) {
    if (st != B200_OK) throw std::runtime_error(std::string("b200: ") + b200_last_error());
  }
  template <class V>
  static void to_feat4(const V& pts, std::vector<float>& out) {
    out.resize(4 * pts.size());
    for (size_t i = 0; i < pts.size(); ++i) {
      out[4 * i] = (float)pts[i][0]; out[4 * i + 1] = (float)pts[i][1]; out[4 * i + 2] = (float)pts[i][2]; out[4 * i + 3] = 1.f;
    }
  }
  template <class V>
  static void to_xyz3(const V& v, std::vector<float>& out) {
    out.resize(3 * v.size());
    for (size_t i = 0; i < v.size(); ++i) { out[3 * i] = (float)v[i][0]; out[3 * i + 1] = (float)v[i][1]; out[3 * i + 2] = (float)v[i][2]; }
  }
  void ensure() {
    if (h_) return;
    b200_o3d_params p;
    b200_o3d_default_params(&p);
    p.kind = (int)kind_;
    p.max_correspondence_distance = (float)maxCorrespondenceDistance_;
    p.max_iteration = maxNumIter_;
    p.relative_fitness = relativeFitness_; p.relative_rmse = relativeRmse_; p.gicp_epsilon = gicpEpsilon_;
    check(b200_o3d_create(&h_, device_, &p));
  }
  b200_o3d* h_ = nullptr;
};

}  // namespace b200o3d
