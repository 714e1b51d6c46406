// Self-test of the host-side C++ shims against pm_mirror.h (a literal mirror of the reference's PointMatcher.h names). Without a CUDA device it checks that every module
// fails loudly with the mapped exception (no CPU fallback); with a device it runs the modules on a small synthetic
// cloud and checks them against a brute-force computation done right here.
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <utility>

#include "pm_mirror.h"   // in the reference tree: #include "pointmatcher/PointMatcher.h"
#include "b200_cloud_registration.hpp"
#include "b200_place_recognition.hpp"
#include "b200_pm_modules.hpp"

using PM = PointMatcher<float>;
using PointMatcherSupport::Parametrizable;

static int fails = 0;
#define EXPECT(c) do { if (!(c)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c); ++fails; } } while (0)

static PM::DataPoints plane_cloud(int n, unsigned seed, float dz, bool normals) {
  std::mt19937 rng(seed);
  std::uniform_real_distribution<float> U(-5.f, 5.f), J(-0.01f, 0.01f);
  PM::DataPoints c;
  c.features.resize(4, n);
  for (int i = 0; i < n; ++i) {
    const float x = U(rng), y = U(rng);
    c.features(0, i) = x; c.features(1, i) = y;
    c.features(2, i) = 0.05f * std::sin(x) + 0.03f * y * y * 0.1f + dz + J(rng);
    c.features(3, i) = 1.f;
  }
  if (normals) {
    PM::Matrix nn(3, n);
    for (int i = 0; i < n; ++i) {
      const float x = c.features(0, i), y = c.features(1, i);
      float v[3] = {-0.05f * std::cos(x), -0.006f * y, 1.f};
      const float s = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
      for (int d = 0; d < 3; ++d) nn(d, i) = v[d] / s;
    }
    c.addDescriptor("normals", nn);
  }
  return c;
}

struct V3 { double v[3]; double& operator[](int i) { return v[i]; } const double& operator[](int i) const { return v[i]; } };
struct Cloud { std::vector<V3> points_, normals_; };
struct M4 { double m[16]; double& operator()(int i, int j) { return m[i + 4 * j]; } const double& operator()(int i, int j) const { return m[i + 4 * j]; } };
// open3d::pipelines::registration::Feature as far as the shim touches it: data_(row, col), Resize(dim, n)
struct FeatureData { std::vector<double> v; int rows = 0; double& operator()(int r, int c) { return v[r + (size_t)rows * c]; } const double& operator()(int r, int c) const { return v[r + (size_t)rows * c]; } };
struct Feature { FeatureData data_; void Resize(int dim, int n) { data_.rows = dim; data_.v.assign((size_t)dim * n, 0.0); } };
struct Result { M4 transformation_; std::vector<std::pair<int, int>> correspondence_set_; double fitness_ = 0, inlier_rmse_ = 0; };

// The shims go through the reference's own plugin mechanism: ADD_TO_REGISTRAR lines (what INTEGRATION.md adds to
// PointMatcher<T>::PointMatcher(), Registry.cpp:61-128), creation BY NAME with the parameter strings of
// open3d_slam_ros/param/icp.yaml, the "was set but is not used" exception (Registrar.h:103-111) and the documented
// bounds (Parametrizable.cpp:179-184). No device is needed for any of this.
template <class T>
struct Registered : PointMatcher<T> {
  typedef PointMatcher<T> PM_;
  typedef typename PM_::Matcher Matcher;
  typedef typename PM_::DataPointsFilter DataPointsFilter;
  typedef typename PM_::OutlierFilter OutlierFilter;
  typedef typename PM_::ErrorMinimizer ErrorMinimizer;
  using PM_::MatcherRegistrar; using PM_::DataPointsFilterRegistrar; using PM_::OutlierFilterRegistrar; using PM_::ErrorMinimizerRegistrar;
  Registered() {
    ADD_TO_REGISTRAR(Matcher, B200KDTreeMatcher, typename b200pm::B200KDTreeMatcher<PM_>)
    ADD_TO_REGISTRAR(DataPointsFilter, B200SurfaceNormalDataPointsFilter, typename b200pm::B200SurfaceNormalDataPointsFilter<PM_>)
    ADD_TO_REGISTRAR(OutlierFilter, B200MaxDistOutlierFilter, typename b200pm::B200MaxDistOutlierFilter<PM_>)
    ADD_TO_REGISTRAR(OutlierFilter, B200MinDistOutlierFilter, typename b200pm::B200MinDistOutlierFilter<PM_>)
    ADD_TO_REGISTRAR(OutlierFilter, B200MedianDistOutlierFilter, typename b200pm::B200MedianDistOutlierFilter<PM_>)
    ADD_TO_REGISTRAR(OutlierFilter, B200TrimmedDistOutlierFilter, typename b200pm::B200TrimmedDistOutlierFilter<PM_>)
    ADD_TO_REGISTRAR(OutlierFilter, B200SurfaceNormalOutlierFilter, typename b200pm::B200SurfaceNormalOutlierFilter<PM_>)
    ADD_TO_REGISTRAR(ErrorMinimizer, B200PointToPlaneErrorMinimizer, typename b200pm::B200PointToPlaneErrorMinimizer<PM_>)
    ADD_TO_REGISTRAR(ErrorMinimizer, B200PointToPointErrorMinimizer, typename b200pm::B200PointToPointErrorMinimizer<PM_>)
  }
};

static const Registered<float>& registry() {
  static Registered<float> r;
  return r;
}

static int registrar_checks() {
  int before = fails;
  const Registered<float>& pm = registry();
  // param/icp.yaml: matcher: KDTreeMatcher {knn: 1, maxDist: 0.5, epsilon: 0.01} with the B200 name
  std::shared_ptr<PM::Matcher> m = pm.MatcherRegistrar.create("B200KDTreeMatcher", {{"knn", "1"}, {"maxDist", "0.5"}, {"epsilon", "0.01"}});
  auto* km = dynamic_cast<b200pm::B200KDTreeMatcher<PM>*>(m.get());
  EXPECT(km && km->knn == 1 && km->maxDist == 0.5f && km->epsilon == 0.01f && km->searchType == 1 && km->className == "B200KDTreeMatcher");
  EXPECT(m->parametersUsed.count("knn") && m->parametersUsed.count("searchType") && m->parameters.at("maxDist") == "0.5");
  // outlierFilters: TrimmedDistOutlierFilter {ratio: 0.90}, SurfaceNormalOutlierFilter {maxAngle: 1.57}
  std::shared_ptr<PM::OutlierFilter> tf = pm.OutlierFilterRegistrar.create("B200TrimmedDistOutlierFilter", {{"ratio", "0.90"}});
  EXPECT(dynamic_cast<b200pm::B200TrimmedDistOutlierFilter<PM>*>(tf.get()) && dynamic_cast<b200pm::B200DistOutlierFilter<PM>*>(tf.get())->param == 0.90f);
  std::shared_ptr<PM::OutlierFilter> sf = pm.OutlierFilterRegistrar.create("B200SurfaceNormalOutlierFilter", {{"maxAngle", "1.57"}});
  EXPECT(dynamic_cast<b200pm::B200SurfaceNormalOutlierFilter<PM>*>(sf.get())->maxAngle == 1.57f);
  // errorMinimizer: PointToPlaneErrorMinimizer (no parameters); the Mapper's normal filter (Mapper.cpp:44-52)
  std::shared_ptr<PM::ErrorMinimizer> em = pm.ErrorMinimizerRegistrar.create("B200PointToPlaneErrorMinimizer");
  EXPECT(em && em->className == "B200PointToPlaneErrorMinimizer");
  std::shared_ptr<PM::DataPointsFilter> nf = pm.DataPointsFilterRegistrar.create(
      "B200SurfaceNormalDataPointsFilter", {{"knn", "10"}, {"epsilon", "0.03"}, {"keepNormals", "1"}, {"keepDensities", "1"}});
  auto* nfp = dynamic_cast<b200pm::B200SurfaceNormalDataPointsFilter<PM>*>(nf.get());
  EXPECT(nfp && nfp->knn == 10u && nfp->keepDensities && !nfp->keepEigenValues && nfp->epsilon == 0.03f);
  EXPECT(pm.MatcherRegistrar.getDescription("B200KDTreeMatcher").find("B200") != std::string::npos);
  EXPECT(pm.MatcherRegistrar.getDescriptor("B200KDTreeMatcher")->availableParameters().size() == 5);
  // a parameter that no constructor reads: "was set but is not used" (Registrar.h:103-111)
  bool threw = false;
  try { pm.MatcherRegistrar.create("B200KDTreeMatcher", {{"knn", "1"}, {"bogusKey", "3"}}); }
  catch (const Parametrizable::InvalidParameter& e) { threw = std::string(e.what()).find("was set but is not used") != std::string::npos; }
  EXPECT(threw);
  // documented bounds (Parametrizable.cpp:179-184): ratio above 1.0, knn below 3 for the normal filter
  threw = false;
  try { pm.OutlierFilterRegistrar.create("B200TrimmedDistOutlierFilter", {{"ratio", "1.5"}}); }
  catch (const Parametrizable::InvalidParameter& e) { threw = std::string(e.what()).find("larger than maximum") != std::string::npos; }
  EXPECT(threw);
  threw = false;
  try { pm.DataPointsFilterRegistrar.create("B200SurfaceNormalDataPointsFilter", {{"knn", "2"}}); }
  catch (const Parametrizable::InvalidParameter& e) { threw = std::string(e.what()).find("smaller than minimum") != std::string::npos; }
  EXPECT(threw);
  threw = false;
  try { pm.MatcherRegistrar.create("NoSuchMatcher"); } catch (const PointMatcherSupport::InvalidElement&) { threw = true; }
  EXPECT(threw);
  threw = false;   // accepted by the reference, refused loudly here
  try { pm.ErrorMinimizerRegistrar.create("B200PointToPlaneErrorMinimizer", {{"force2D", "1"}}); }
  catch (const Parametrizable::InvalidParameter&) { threw = true; }
  EXPECT(threw);
  return fails - before;
}

int main() {
  const int ndev = b200_device_count();
  std::printf("devices: %d\n", ndev);
  PM::DataPoints ref = plane_cloud(4000, 1, 0.f, true), read = plane_cloud(1500, 2, 0.08f, false);
  if (ndev < 1) {
    // no CPU fallback: every module must throw
    bool threw = false;
    try { b200pm::B200KDTreeMatcher<PM> m; m.init(ref); } catch (const std::runtime_error&) { threw = true; }
    EXPECT(threw);
    threw = false;
    try { b200pm::B200ICP<PM> icp; icp.initReference(ref); } catch (const std::runtime_error&) { threw = true; }
    EXPECT(threw);
    threw = false;
    try { b200pm::B200KDTreeMatcher<PM> m(PM::Parameters{{"knn", "0"}}); } catch (const PM::InvalidParameter&) { threw = true; }   // below the documented minimum
    EXPECT(threw);
    fails += registrar_checks();
    std::printf(fails ? "selftest FAILED\n" : "selftest ok (no device: loud failure verified)\n");
    return fails ? 1 : 0;
  }
  registrar_checks();
  // ---- matcher vs brute force (bit-exact: same float arithmetic, ties -> smallest index)
  b200pm::B200KDTreeMatcher<PM> matcher(PM::Parameters{{"knn", "3"}, {"maxDist", "0.5"}});
  matcher.init(ref);
  PM::Matches m = matcher.findClosests(read);
  EXPECT(m.ids.rows() == 3 && m.ids.cols() == 1500 && matcher.visitCounter > 0);
  int bad = 0;
  for (int i = 0; i < 1500; i += 7) {
    float bd[3] = {INFINITY, INFINITY, INFINITY};
    int bi[3] = {-1, -1, -1};
    for (int j = 0; j < 4000; ++j) {
      const float dx = read.features(0, i) - ref.features(0, j), dy = read.features(1, i) - ref.features(1, j),
                  dz = read.features(2, i) - ref.features(2, j);
      volatile float a = dx * dx, b = dy * dy, c = dz * dz;  // volatile: no FMA contraction
      volatile float s1 = a + b;
      const float d = s1 + c;
      if (d > 0.25f) continue;
      for (int k = 0; k < 3; ++k)
        if (d < bd[k] || (d == bd[k] && j < bi[k])) {
          for (int t = 2; t > k; --t) { bd[t] = bd[t - 1]; bi[t] = bi[t - 1]; }
          bd[k] = d; bi[k] = j;
          break;
        }
    }
    for (int k = 0; k < 3; ++k) bad += (m.ids(k, i) != bi[k]) || (m.dists(k, i) != bd[k]);
  }
  EXPECT(bad == 0);
  // ---- outlier filters
  b200pm::B200TrimmedDistOutlierFilter<PM> trimmed(PM::Parameters{{"ratio", "0.8"}});
  PM::OutlierWeights w = trimmed.compute(read, ref, m);
  long kept = 0, valid = 0;
  for (int i = 0; i < 1500; ++i)
    for (int k = 0; k < 3; ++k) { kept += w(k, i) != 0.f; valid += std::isfinite(m.dists(k, i)); }
  EXPECT(kept > 0 && kept <= valid && kept >= (long)(0.78 * valid));
  b200pm::B200MaxDistOutlierFilter<PM> maxd(PM::Parameters{{"maxDist", "0.2"}});
  PM::OutlierWeights w2 = maxd.compute(read, ref, m);
  for (int i = 0; i < 1500; i += 11) EXPECT((w2(0, i) != 0.f) == (m.dists(0, i) <= 0.04f));
  // ---- surface normals
  b200pm::B200SurfaceNormalDataPointsFilter<PM> sn(PM::Parameters{{"knn", "10"}, {"keepDensities", "1"}});
  PM::DataPoints withN = sn.filter(read);
  EXPECT(withN.descriptorExists("normals") && withN.descriptorExists("densities") && withN.descriptors.rows() == 4);
  int upright = 0;
  for (int i = 0; i < 1500; ++i) upright += std::fabs(withN.descriptors(2, i)) > 0.9f;
  EXPECT(upright > 1400);
  // ---- whole chain: recovers the 0.08 m offset along z
  b200pm::B200ICP<PM> icp;
  icp.prm.maxDist = 0.5f; icp.prm.trimmed_ratio = 0.9f;
  PM::TransformationParameters T = icp(read, ref);
  EXPECT(std::fabs(T(2, 3) + 0.08f) < 0.01f && icp.last.iterations > 1);
  // the call Mapper.cpp:372-373 makes: compute(cloud, <unused>, init, false) on the reference set before
  PM::TransformationParameters T2 = icp.compute(read, PM::DataPoints(), PM::TransformationParameters::Identity(4, 4), false);
  int sameT = 0;
  for (int i = 0; i < 16; ++i) sameT += T2.data()[i] == T.data()[i];
  EXPECT(sameT == 16);
  // ---- error minimiser with this fork's signature: A_ / b_ formed (here on the device), then solved by compute()
  {
    std::shared_ptr<PM::ErrorMinimizer> em = registry().ErrorMinimizerRegistrar.create("B200PointToPlaneErrorMinimizer");
    auto* p2p = dynamic_cast<b200pm::B200PointToPlaneErrorMinimizer<PM>*>(em.get());
    PM::ErrorMinimizer::ErrorElements ee;
    PM::OutlierWeights wl0;
    PM::Matches ml0 = icp.matches(&wl0);
    int P = 0;
    for (int i = 0; i < 1500; ++i) P += (wl0(0, i) != 0.f && ml0.ids(0, i) >= 0);
    ee.reading.features.resize(4, P); ee.reference.features.resize(4, P); ee.weights.resize(1, P);
    PM::Matrix nn(3, P);
    for (int i = 0, c = 0; i < 1500; ++i) {
      if (!(wl0(0, i) != 0.f && ml0.ids(0, i) >= 0)) continue;
      const int j = ml0.ids(0, i);
      for (int d = 0; d < 4; ++d) { ee.reading.features(d, c) = read.features(d, i); ee.reference.features(d, c) = ref.features(d, j); }
      for (int d = 0; d < 3; ++d) nn(d, c) = ref.descriptors(d, j);
      ee.weights(0, c) = 1.f;
      ++c;
    }
    ee.reference.addDescriptor("normals", nn);
    PM::LocalizabilityParametersForErrorMinimization lp;
    p2p->accumulate(ee, lp);
    PM::TransformationParameters Te = em->compute(ee, lp);
    std::printf("minimiser shim: A_(5,5) %.1f tz %.5f solveInfo %d (a near-planar scene may take the rank-deficient branch)\n",
                lp.A_(5, 5), Te(2, 3), p2p->solveInfo);
    EXPECT(lp.A_(5, 5) > 0.f && std::fabs(Te(2, 3) + 0.08f) < 0.02f && p2p->solveInfo >= 0 && p2p->solveInfo <= 2);
    lp.degeneracyAwarenessMethod = PM::DegeneracyAwarenessMethod::kOptimizedEqualityConstraints;
    bool threwE = false;
    try { em->compute(ee, lp); } catch (const Parametrizable::InvalidParameter&) { threwE = true; }
    EXPECT(threwE);
  }
  PM::OutlierWeights wl;
  PM::Matches ml = icp.matches(&wl);
  EXPECT(ml.ids.cols() == 1500 && wl.cols() == 1500);
  // ---- X-ICP (degeneracyAwareness: OptimizedEqualityConstraints, param/icp.yaml): a near-planar scene observes only
  // the translation along its normal and the two tilts; the other three directions are held by equality constraints
  {
    b200pm::B200ICP<PM> xicp;
    xicp.prm.maxDist = 0.5f; xicp.prm.trimmed_ratio = 0.9f;
    EXPECT(!xicp.setDegeneracyAwareness("OptimizedEqualityConstraints", {{"enoughInformationThreshold", "250"}}));
    EXPECT(xicp.setDegeneracyAwareness("OptimizedEqualityConstraints",
                                       {{"enoughInformationThreshold", "250"},
                                        {"insufficientInformationThreshold", "180"},
                                        {"point2NormalMinimalAlignmentAngleThreshold", "80"},
                                        {"point2NormalStrongAlignmentAngleThreshold", "45"}}));
    PM::TransformationParameters Tx = xicp(read, ref);
    int lx[3], lr[3];
    xicp.localizability(lx, lr);
    std::printf("X-ICP: xyz %d%d%d rpy %d%d%d tz %.5f tx %.2e ty %.2e\n", lx[0], lx[1], lx[2], lr[0], lr[1], lr[2], Tx(2, 3),
                Tx(0, 3), Tx(1, 3));
    EXPECT(lx[0] == 1 && lx[1] == 0 && lx[2] == 0 && lr[0] == 1 && lr[1] == 1 && lr[2] == 0);
    EXPECT(std::fabs(Tx(2, 3) + 0.08f) < 0.01f && std::fabs(Tx(0, 3)) < 2e-3f && std::fabs(Tx(1, 3)) < 2e-3f);
    bool threwX = false;
    try { xicp.setDegeneracyAwareness("SolutionRemapping", {}); } catch (const PM::InvalidParameter&) { threwX = true; }
    EXPECT(threwX);
  }
  // ---- boundary B
  Cloud src, tgt;
  for (int i = 0; i < 1500; ++i) src.points_.push_back(V3{{read.features(0, i), read.features(1, i), read.features(2, i)}});
  for (int i = 0; i < 4000; ++i) {
    tgt.points_.push_back(V3{{ref.features(0, i), ref.features(1, i), ref.features(2, i)}});
    tgt.normals_.push_back(V3{{ref.descriptors(0, i), ref.descriptors(1, i), ref.descriptors(2, i)}});
  }
  b200o3d::RegistrationIcpB200<Cloud, M4, Result> reg;
  reg.maxCorrespondenceDistance_ = 0.5;
  M4 I{};
  for (int i = 0; i < 4; ++i) I(i, i) = 1.0;
  Result rr = reg.registerClouds(src, tgt, I);
  std::printf("boundary B: tz %.5f fitness %.4f rmse %.5f corr %zu\n", rr.transformation_(2, 3), rr.fitness_, rr.inlier_rmse_,
              rr.correspondence_set_.size());
  EXPECT(std::fabs(rr.transformation_(2, 3) + 0.08) < 0.01 && rr.fitness_ > 0.9 && rr.inlier_rmse_ < 0.2);
  Cloud bare;
  bare.points_ = tgt.points_;
  reg.estimateNormalsOrCovariancesIfNeeded(&bare);
  EXPECT(bare.normals_.size() == bare.points_.size());
  // generalized and point-to-point estimators through the same class
  Cloud srcN = src;
  reg.estimateNormalsOrCovariancesIfNeeded(&srcN);
  for (b200o3d::Kind k : {b200o3d::Kind::Generalized, b200o3d::Kind::PointToPoint}) {
    b200o3d::RegistrationIcpB200<Cloud, M4, Result> rk;
    rk.kind_ = k;
    rk.maxCorrespondenceDistance_ = 0.5;
    Result r2 = rk.registerClouds(srcN, tgt, I);
    std::printf("boundary B kind %d: tz %.5f fitness %.4f iterations %d\n", (int)k, r2.transformation_(2, 3), r2.fitness_, rk.lastIterations_);
    EXPECT(std::fabs(r2.transformation_(2, 3) + 0.08) < 0.02 && r2.fitness_ > 0.9);
  }
  // ---- loop-closure front half: FPFH of the target and of a moved copy of it, RANSAC brings the copy back
  {
    Cloud a = bare, b;              // `bare` got its normals from estimateNormalsOrCovariancesIfNeeded above
    const double yaw = 0.2, c = std::cos(yaw), sn = std::sin(yaw), t[3] = {0.7, -0.4, 0.1};
    for (size_t i = 0; i < a.points_.size(); ++i) {   // b = T^-1 a, so that T maps b (source) onto a (target)
      const V3& p = a.points_[i];
      const V3& n = a.normals_[i];
      const double q[3] = {p[0] - t[0], p[1] - t[1], p[2] - t[2]};
      b.points_.push_back(V3{{c * q[0] + sn * q[1], -sn * q[0] + c * q[1], q[2]}});
      b.normals_.push_back(V3{{c * n[0] + sn * n[1], -sn * n[0] + c * n[1], n[2]}});
    }
    b200o3d::PlaceRecognitionB200 pr;
    pr.featureRadius_ = 1.5; pr.featureKnn_ = 50;
    pr.ransacMaxCorrespondenceDistance_ = 0.05; pr.correspondenceCheckerDistance_ = 0.05; pr.correspondenceCheckerEdgeLength_ = 0.9;
    Feature fa, fb;
    pr.computeFpfhFeature(a, &fa);
    pr.computeFpfhFeature(b, &fb);
    {   // the one-call variant on a cloud without normals gives the same features as normals-then-FPFH on `a`
      Cloud a2;
      a2.points_ = a.points_;
      Feature fa2;
      pr.normalEstimationRadius_ = reg.maxRadiusNormalEstimation_; pr.normalKnn_ = reg.knnNormalEstimation_;
      pr.estimateNormalsAndComputeFpfh(&a2, &fa2);
      bool same = a2.normals_.size() == a.normals_.size();
      for (size_t i = 0; same && i < a.normals_.size(); ++i)
        for (int d = 0; d < 3; ++d) same = same && a2.normals_[i][d] == a.normals_[i][d];
      for (size_t i = 0; same && i < fa.data_.v.size(); ++i) same = same && fa2.data_.v[i] == fa.data_.v[i];
      EXPECT(same);
    }
    double colsum = 0.0;
    for (int r = 0; r < 33; ++r) colsum += fa.data_(r, 0);
    Result lc = pr.ransacFeatureMatching<Cloud, Feature, Result>(b, a, fb, fa);
    std::printf("loop closure: fpfh column sum %.1f, ransac fitness %.3f corr %zu iterations %lld t (%.3f %.3f %.3f)\n", colsum,
                lc.fitness_, lc.correspondence_set_.size(), (long long)pr.lastIterations_, lc.transformation_(0, 3),
                lc.transformation_(1, 3), lc.transformation_(2, 3));
    EXPECT(std::fabs(colsum - 600.0) < 1e-6 || colsum == 0.0);
    EXPECT(lc.fitness_ > 0.5 && lc.correspondence_set_.size() >= 25);
    EXPECT(std::fabs(lc.transformation_(0, 3) - t[0]) < 0.02 && std::fabs(lc.transformation_(1, 3) - t[1]) < 0.02 &&
           std::fabs(lc.transformation_(2, 3) - t[2]) < 0.02);
    std::vector<size_t> is, it;
    pr.overlappingIndices(b, a, lc.transformation_, 0.3, 1, &is, &it);
    EXPECT(is.size() > b.points_.size() * 9 / 10 && it.size() > a.points_.size() * 9 / 10);
    bool threwF = false;
    Cloud noN;
    noN.points_ = a.points_;
    try { pr.computeFpfhFeature(noN, &fa); } catch (const std::runtime_error&) { threwF = true; }
    EXPECT(threwF);
  }
  std::printf(fails ? "selftest FAILED\n" : "selftest ok (device path)\n");
  return fails ? 1 : 0;
}
