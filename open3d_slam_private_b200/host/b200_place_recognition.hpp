// b200_place_recognition.hpp — the loop-closure front half of open3d_slam backed by the B200 registration core
// (b200_lc_* in include/b200reg.h). Drop-ins for the three calls of the place-recognition path:
//   computeFpfhFeature        registration::ComputeFPFHFeature(cloud, KDTreeSearchParamHybrid(radius, knn))
//                             — Submap::computeFeatures, open3d_slam/src/Submap.cpp:272
//   ransacFeatureMatching     RegistrationRANSACBasedOnFeatureMatching(source, target, sourceFeature, targetFeature, true,
//                             maxCorrespondenceDistance, TransformationEstimationPointToPoint(false), ransacModelSize,
//                             {distanceChecker, edgeLengthChecker}, RANSACConvergenceCriteria(numIter, probability))
//                             — PlaceRecognition::buildLoopClosureConstraints, src/PlaceRecognition.cpp:81-85
//   overlappingIndices        computeIndicesOfOverlappingPoints (src/helpers.cpp:320-345; PlaceRecognition.cpp:103,
//                             constraint_builders.cpp:54)
// Header-only and templated so that it compiles without Open3D: in the reference tree
//   Cloud   = open3d::geometry::PointCloud                      (points_, normals_ : std::vector<Eigen::Vector3d>)
//   Feature = open3d::pipelines::registration::Feature          (data_ : Eigen::MatrixXd 33 x N; Resize(dim, n))
//   Result  = open3d::pipelines::registration::RegistrationResult, Mat4 = Eigen::Matrix4d
// Clouds are converted double -> float exactly like open3dToPointmatcher (open3d_conversions.cpp:57-118).
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/b200reg.h"

namespace b200o3d {

class PlaceRecognitionB200 {
 public:
  // PlaceRecognitionParameters (open3d_slam/include/open3d_slam/Parameters.hpp:122-135)
  double featureRadius_ = 2.5;
  int featureKnn_ = 100;
  int ransacNumIter_ = 1000000;
  double ransacProbability_ = 0.99;
  int ransacModelSize_ = 3;
  double ransacMaxCorrespondenceDistance_ = 0.75;
  double correspondenceCheckerDistance_ = 0.75;
  double correspondenceCheckerEdgeLength_ = 0.5;
  uint64_t seed_ = 1;          // sample stream (Open3D's own is seeded from std::random_device)
  int device_ = 0;

  PlaceRecognitionB200() = default;
  ~PlaceRecognitionB200() { b200_lc_destroy(h_); }
  PlaceRecognitionB200(const PlaceRecognitionB200&) = delete;
  PlaceRecognitionB200& operator=(const PlaceRecognitionB200&) = delete;

  template <class Cloud, class Feature>
  void computeFpfhFeature(const Cloud& cloud, Feature* feature) {
    if (cloud.normals_.size() != cloud.points_.size())
      throw std::runtime_error("[ComputeFPFHFeature] Failed because input point cloud has no normal.");   // Open3D's own error
    ensure();
    const int n = (int)cloud.points_.size();
    feature->Resize(33, n);
    if (n == 0) return;
    std::vector<float> f4;
    to_feat4(cloud.points_, f4);
    std::vector<double> nrm(3 * (size_t)n), out(33 * (size_t)n);
    for (int i = 0; i < n; ++i)
      for (int d = 0; d < 3; ++d) nrm[3 * (size_t)i + d] = cloud.normals_[i][d];
    check(b200_lc_compute_fpfh(h_, f4.data(), nrm.data(), n, (float)featureRadius_, featureKnn_, out.data(), nullptr));
    for (int i = 0; i < n; ++i)
      for (int r = 0; r < 33; ++r) feature->data_(r, i) = out[33 * (size_t)i + r];
  }

  // Submap::computeFeatures after the voxel down-sample (Submap.cpp:269-272) in one call: EstimateNormals(Hybrid(
  // normalEstimationRadius_, normalKnn_)) + NormalizeNormals + OrientNormalsTowardsCameraLocation(0) + ComputeFPFHFeature
  // on one device index; fills cloud->normals_ like the reference's sequence does
  double normalEstimationRadius_ = 1.0;   // Parameters.hpp:122
  int normalKnn_ = 10;                    // Parameters.hpp:126
  template <class Cloud, class Feature>
  void estimateNormalsAndComputeFpfh(Cloud* cloud, Feature* feature) {
    ensure();
    const int n = (int)cloud->points_.size();
    feature->Resize(33, n);
    cloud->normals_.resize(n);
    if (n == 0) return;
    std::vector<float> f4;
    to_feat4(cloud->points_, f4);
    std::vector<double> nrm(3 * (size_t)n), out(33 * (size_t)n);
    check(b200_lc_compute_features(h_, f4.data(), n, (float)normalEstimationRadius_, normalKnn_, (float)featureRadius_, featureKnn_,
                                   nrm.data(), out.data()));
    for (int i = 0; i < n; ++i) {
      for (int d = 0; d < 3; ++d) cloud->normals_[i][d] = nrm[3 * (size_t)i + d];
      for (int r = 0; r < 33; ++r) feature->data_(r, i) = out[33 * (size_t)i + r];
    }
  }

  template <class Cloud, class Feature, class Result>
  Result ransacFeatureMatching(const Cloud& source, const Cloud& target, const Feature& sourceFeature,
                               const Feature& targetFeature) {
    ensure();
    const int ns = (int)source.points_.size(), nt = (int)target.points_.size();
    std::vector<float> s4, t4;
    to_feat4(source.points_, s4);
    to_feat4(target.points_, t4);
    std::vector<double> sf(33 * (size_t)ns), tf(33 * (size_t)nt);
    for (int i = 0; i < ns; ++i)
      for (int r = 0; r < 33; ++r) sf[33 * (size_t)i + r] = sourceFeature.data_(r, i);
    for (int i = 0; i < nt; ++i)
      for (int r = 0; r < 33; ++r) tf[33 * (size_t)i + r] = targetFeature.data_(r, i);
    b200_ransac_params p;
    b200_lc_default_ransac_params(&p);
    p.mutual_filter = 1;
    p.max_correspondence_distance = ransacMaxCorrespondenceDistance_;
    p.ransac_n = ransacModelSize_;
    p.edge_length_similarity = correspondenceCheckerEdgeLength_;
    p.distance_threshold = correspondenceCheckerDistance_;
    p.max_iteration = ransacNumIter_;
    p.confidence = ransacProbability_;
    p.seed = seed_++;
    b200_ransac_result r;
    std::vector<int32_t> corr(2 * (size_t)(ns > 0 ? ns : 1));
    check(b200_lc_ransac_feature_matching(h_, s4.data(), ns, t4.data(), nt, sf.data(), tf.data(), &p, &r, corr.data()));
    Result out;
    for (int j = 0; j < 4; ++j)
      for (int i = 0; i < 4; ++i) out.transformation_(i, j) = r.T[i + 4 * j];
    out.fitness_ = r.fitness;
    out.inlier_rmse_ = r.inlier_rmse;
    out.correspondence_set_.clear();
    for (int k = 0; k < r.n_correspondences; ++k) out.correspondence_set_.emplace_back(corr[2 * k], corr[2 * k + 1]);
    lastIterations_ = r.iterations_run;
    return out;
  }
  int64_t lastIterations_ = 0;

  template <class Cloud, class Mat4>
  void overlappingIndices(const Cloud& source, const Cloud& target, const Mat4& sourceToTarget, double voxelSize,
                          size_t minNumPointsPerVoxel, std::vector<size_t>* idxsSource, std::vector<size_t>* idxsTarget) {
    ensure();
    const int ns = (int)source.points_.size(), nt = (int)target.points_.size();
    std::vector<float> s4, t4;
    to_feat4(source.points_, s4);
    to_feat4(target.points_, t4);
    double T[16];
    for (int j = 0; j < 4; ++j)
      for (int i = 0; i < 4; ++i) T[i + 4 * j] = (double)sourceToTarget(i, j);
    std::vector<int32_t> si((size_t)(ns > 0 ? ns : 1)), ti((size_t)(nt > 0 ? nt : 1));
    int a = 0, b = 0;
    check(b200_lc_overlap_indices(h_, s4.data(), ns, t4.data(), nt, T, voxelSize, (int)minNumPointsPerVoxel, si.data(), &a,
                                  ti.data(), &b));
    idxsSource->assign(si.begin(), si.begin() + a);
    idxsTarget->assign(ti.begin(), ti.begin() + b);
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
  void ensure() {
    if (!h_) check(b200_lc_create(&h_, device_));
  }
  b200_lc* h_ = nullptr;
};

}  // namespace b200o3d
