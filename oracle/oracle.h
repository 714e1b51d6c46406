/* ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's registration hot path (libpointmatcher fork inside
 * leggedrobotics/open3d_slam_private). It is the checker for the CUDA path, never the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * Pinning status:
 *   - whole-ICP results are pinned against the reference's own goldens: validT3d (utest.cpp:351-360),
 *     icpSingular / icpIdentity known answers (utest.cpp:163-221) and the icp_data ref_trans files.
 *   - kNN ids/dists, normals, weights, H, b, iteration counts: the reference has no test that pins them and
 *     libnabo / Eigen are not in /root/reference => "parity unpinned" for those quantities; the oracle fixes
 *     them by the published algorithm (exact kNN at epsilon = 0, ties -> smallest index).
 *
 * All matrices are column-major exactly like the reference's Eigen matrices:
 *   features 4 x N  (x,y,z,1 per point),  normals 3 x N,  dists/ids/weights k x N.
 */
#ifndef ORACLE_H_
#define ORACLE_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_OK 0
#define ORC_E_NO_MATCHES 1   /* PointMatcher::ConvergenceError */
#define ORC_E_BAD_PARAM 2
#define ORC_E_NAN 3
#define ORC_E_BOUND 7   /* ConvergenceError("limit out of bounds"), TransformationCheckersImpl.cpp:214-221 */

/* ---- kNN (stands in for libnabo NNS::knn as called at MatchersImpl.cpp:98) ---- */
int orc_knn_brute(const float* ref4xM, int M, const float* query4xN, int N, int k, float maxDist,
                  int32_t* ids_kxN, float* dists_kxN);
void* orc_kdtree_build(const float* ref4xM, int M);
void orc_kdtree_free(void* tree);
/* returns number of points touched (libnabo's visit statistic, our tree's own count) */
long orc_kdtree_knn(const void* tree, const float* query4xN, int N, int k, float maxDist, int nthreads,
                    int32_t* ids_kxN, float* dists_kxN);

/* ---- Matches::getDistsQuantile (Matches.cpp:60-87) ---- */
int orc_dists_quantile(const float* dists, long n, float quantile, float* out);

/* ---- outlier filters (OutlierFiltersImpl.cpp:75-81, 95-101, 117-125, 140-147, 236-288) ---- */
int orc_outlier_maxdist(const float* dists, long n, float maxDist, float* w);
int orc_outlier_mindist(const float* dists, long n, float minDist, float* w);
int orc_outlier_median(const float* dists, long n, float factor, float* w);
int orc_outlier_trimmed(const float* dists, long n, float ratio, float* w);
int orc_outlier_surface_normal(const float* nread3xN, const float* nref3xM, const int32_t* ids_kxN, int k, int N,
                               float maxAngle, float* w);
/* OutlierFilters::compute with an empty chain (OutlierFilter.cpp:66-83) */
int orc_outlier_null(const float* dists, long n, float* w);

/* ---- ErrorElements ctor (ErrorMinimizer.cpp:59-193): order-preserving compaction.
 * outputs sized for k*N pairs; returns P through *P_out. */
int orc_error_elements(const float* read4xN, const float* ref4xM, const float* refn3xM /*nullable*/,
                       const float* w_kxN, const int32_t* ids_kxN, const float* dists_kxN, int k, int N,
                       float* read4xP, float* ref4xP, float* refn3xP, float* w_P, int32_t* ids_P, float* dists_P,
                       int* P_out, float* pointUsedRatio, float* weightedPointUsedRatio);

/* ---- ICP::calculateOptimizationHessian (ICP.cpp:1513-1566). accum: 0 = float sequential (reference-like
 * float GEMM, order is an Eigen artefact), 1 = float products accumulated in double ("truth"). */
int orc_p2plane_hb(const float* read4xP, const float* ref4xP, const float* refn3xP, const float* w_P, int P,
                   int accum, float* H36, float* b6);

/* ---- solvePossiblyUnderdeterminedLinearSystem (PointToPlane.cpp:113-265, unconstrained branch).
 * info: 0 = full-rank double-SVD path, 1 = rank-deficient minimal-norm path, 2 = SVD fallback */
int orc_solve6(const float* A36, const float* b6, float* x6, int* info);
/* the same routine for the (6 + c) x (6 + c) equality-constrained system (PointToPlane.cpp:461-505), 6 <= n <= 12 */
int orc_solve_n(const float* Ann, const float* bn, int n, float* xn, int* info);
/* x -> 4x4 (PointToPlane.cpp:337-381) */
int orc_p2plane_T(const float* x6, float* T16);

/* ---- PointToPointErrorMinimizer::compute_in_place (PointToPoint.cpp:62-103) ---- */
int orc_p2point_T(const float* read4xP, const float* ref4xP, const float* w_P, int P, float* T16);
/* accum as in orc_p2plane_hb: 0 = float sums in index order, 1 = float inputs summed in double */
int orc_p2point_T2(const float* read4xP, const float* ref4xP, const float* w_P, int P, int accum, float* T16);

/* ---- RigidTransformation::inPlaceCompute (TransformationsImpl.cpp:61-102) incl. check/correct ---- */
int orc_rigid_apply(const float* T16, float* feat4xN, float* normals3xN /*nullable*/, int N);

/* ---- SurfaceNormalDataPointsFilter::inPlaceFilter (SurfaceNormal.cpp:82-290).
 * Any output may be NULL. eigVectors row-major serialised (utils.h:89-103). */
int orc_surface_normals(const float* feat4xN, int N, int knn, float maxDist, int sortEigen, int smoothNormals,
                        float* normals3xN, float* densities1xN, float* eigValues3xN, float* eigVectors9xN,
                        float* matchedIds_kxN, float* meanDists1xN, int32_t* ids_kxN, float* dists_kxN);

/* ---- SamplingSurfaceNormalDataPointsFilter (SamplingSurfaceNormal.cpp:76-344); reference pre-filter of the
 * golden configs. Returns number of output points; feat/normals are written compacted (capacity N).
 * samplingMethod 0 keeps every point whose rand()/RAND_MAX < ratio (std::rand, unseeded like the reference). */
int orc_sampling_surface_normal(const float* feat4xN, int N, int knn, float ratio, int samplingMethod,
                                float maxBoxDim, float* featOut4xN, float* normalsOut3xN);

/* ---- whole ICP (ICP.cpp:813-1349) ---- */
typedef struct {
  int knn;              /* KDTreeMatcher knn */
  float maxDist;        /* KDTreeMatcher maxDist (inf allowed) */
  int use_trimmed;      /* TrimmedDistOutlierFilter */
  float trimmed_ratio;
  int use_maxdist_filter; /* MaxDistOutlierFilter */
  float maxdist_filter;
  int use_surface_normal_filter; /* SurfaceNormalOutlierFilter */
  float surface_normal_max_angle;
  int minimizer;        /* 0 = PointToPlane, 1 = PointToPoint */
  int max_iterations;   /* CounterTransformationChecker; <=0: absent */
  int use_differential; /* DifferentialTransformationChecker */
  float min_diff_rot, min_diff_trans;
  int smooth_length;
  int counter_first;    /* 1: Counter is checked before Differential (setDefault order), 0: after (icp.yaml) */
  int accum;            /* see orc_p2plane_hb */
  int mean_mode;        /* 0 = reference-literal sequential float mean, 1 = double sum rounded to float */
  int knn_threads;      /* kd-tree query threads (1 = like the single-threaded reference) */
  /* degeneracyAwareness (ICP.cpp:575-760): 0 = None, 1 = OptimizedEqualityConstraints (X-ICP, :629-670) */
  int degeneracy;
  float xicp_enough;        /* enoughInformationThreshold */
  float xicp_insufficient;  /* insufficientInformationThreshold */
  float xicp_cos_min;       /* cos(point2NormalMinimalAlignmentAngleThreshold) */
  float xicp_cos_strong;    /* cos(point2NormalStrongAlignmentAngleThreshold) */
  /* BoundTransformationChecker (TransformationCheckersImpl.cpp:165-222), evaluated after Differential */
  int use_bound;
  float max_rotation_norm, max_translation_norm;
} orc_icp_params;

typedef struct {
  int iterations;
  int max_iterations_reached;
  int n_matches_last;   /* P of the last iteration */
  float T[16];          /* result, column-major 4x4 */
  float T_iter[16];
  float T_refIn_refMean[16];
  float T_refMean_readMean[16];
  float weighted_point_used_ratio;
  long points_touched;
  /* X-ICP (degeneracy = 1): flags of the first iteration's detection (1 = localizable), eigenvectors as they stand
   * after the last iteration (columns, refMean frame), whether the prior was returned (ICP.cpp:1335-1341), and as
   * diagnostics the alignment sums over ALL matched pairs in double (the reference stops early):
   * xicp_sums[2e], [2e+1] = combined/high for translation eigenvector e; [6+2e], [7+2e] for rotation eigenvector e */
  int returned_prior;
  int loc_xyz[3], loc_rpy[3];
  float eig_t[9], eig_r[9];
  double xicp_sums[12];
} orc_icp_result;

/* log_* (nullable): per-iteration dumps, capacity max_log iterations: H 36, b 6, x 6, T_iter 16 floats, P int */
int orc_icp(const orc_icp_params* prm, const float* ref4xM, const float* refn3xM, int M, const float* read4xN,
            const float* readn3xN /*nullable*/, int N, const float* Tinit16, orc_icp_result* res, int max_log,
            float* logH, float* logb, float* logx, float* logT, int* logP, int32_t* last_ids, float* last_dists,
            float* last_w);
/* teacher-forced variant (see icp.cpp): after logging its own result of iteration i the oracle continues from
 * forcedT[i]; runs exactly n_forced iterations */
int orc_icp_forced(const orc_icp_params* prm, const float* ref4xM, const float* refn3xM, int M, const float* read4xN,
                   const float* readn3xN, int N, const float* Tinit16, orc_icp_result* res, int max_log, float* logH,
                   float* logb, float* logx, float* logT, int* logP, int32_t* last_ids, float* last_dists,
                   float* last_w, const float* forcedT, int n_forced);

/* ---- boundary B: Open3D v0.15.1 RegistrationICP / RegistrationGeneralizedICP as called by
 * open3d_slam/src/CloudRegistration.cpp:17-23, 62-66 (recalled from upstream, see o3d.cpp). kind 0 point-to-plane,
 * 2 generalized. Matrices column-major double. corr_ids (nullable): matched target index per source point or -1. */
void orc_o3d_covariances_from_normals(const float* n3xN, int N, double eps, double* cov9xN);
int orc_o3d_icp(int kind, const float* src4xN, int N, const float* tgt4xM, int M, const float* tgtn3xM,
                const double* srccov9xN, const double* tgtcov9xM, float max_dist, const double* init16,
                int max_iteration, double rel_fitness, double rel_rmse, int knn_threads, double* T16, double* fitness,
                double* rmse, int* iterations, int32_t* corr_ids);

/* estimateNormalsOrCovariancesIfNeeded (CloudRegistration.cpp:27-46, 67-79): Open3D EstimateNormals with
 * KDTreeSearchParamHybrid(radius, max_nn) + NormalizeNormals + OrientNormalsTowardsCameraLocation(origin), see o3d.cpp.
 * normals3xN: doubles; n_neighbours (nullable): neighbourhood size per point */
int orc_o3d_estimate_normals(const float* pts4xN, int N, int max_nn, float radius, int threads, double* normals3xN,
                             int32_t* n_neighbours);

/* GetInformationMatrixFromPointClouds as called at PlaceRecognition.cpp:148-149 (see o3d.cpp); corr_ids nullable */
int orc_o3d_information_matrix(const float* src4xN, int N, const float* tgt4xM, int M, const double* T16, float max_dist,
                               int threads, double* info36, double* fitness, double* rmse, int32_t* corr_ids);

/* ---- loop-closure front half (see fpfh.cpp): Open3D v0.15.1 ComputeFPFHFeature as called at
 * Submap.cpp:272 and RegistrationRANSACBasedOnFeatureMatching as called at PlaceRecognition.cpp:81-85 (recalled from
 * upstream: parity unpinned). Features are 33 x N doubles, column-major like open3d Feature::data_. ---- */
int orc_o3d_fpfh(const float* pts4xN, const double* nrm3xN, int N, float radius, int max_nn, int threads,
                 double* fpfh33xN, int32_t* nn_count);
int orc_o3d_feature_nn(const double* a33xN, int N, const double* b33xM, int M, int threads, int32_t* nn_N, double* d2_N);
typedef struct {
  int mutual_filter;
  double max_correspondence_distance;
  int ransac_n;                    /* 3..8 */
  double edge_length_similarity;   /* CorrespondenceCheckerBasedOnEdgeLength, <= 0: absent */
  double distance_threshold;       /* CorrespondenceCheckerBasedOnDistance, <= 0: absent */
  int max_iteration;               /* RANSACConvergenceCriteria */
  double confidence;
  uint64_t seed;                   /* counter-based sample stream, see fpfh.cpp */
} orc_ransac_params;
typedef struct {
  double T[16];                    /* column-major */
  double fitness, inlier_rmse;
  int n_correspondences;           /* correspondence_set_.size() */
  int n_feature_correspondences;   /* size of the feature correspondence set the hypotheses were drawn from */
  int mutual_used;
  int64_t iterations_run, validations;
} orc_ransac_result;
int orc_o3d_ransac_feature_matching(const float* src4xN, int Ns, const float* tgt4xM, int Nt, const double* sf33xN,
                                    const double* tf33xM, const orc_ransac_params* prm, int threads,
                                    orc_ransac_result* res, int32_t* corr2xK);

/* computeIndicesOfOverlappingPoints (open3d_slam/src/helpers.cpp:320-345); outputs ascending, capacity Ns / Nt */
int orc_overlap_indices(const float* src4xN, int Ns, const float* tgt4xM, int Nt, const double* T16, double voxel_size,
                        int min_points, int32_t* src_idx, int* n_src, int32_t* tgt_idx, int* n_tgt);

/* ---- scan pre-processing (see prep.cpp): open3d_slam croppers (croppers.cpp:25-164) and Open3D
 * VoxelDownSample (recalled from upstream v0.15.1) ---- */
int orc_crop(int kind, double radius_min, double radius_max, double min_z, double max_z, const double* center3,
             int invert, const float* pts4xN, int N, int32_t* kept_idx);
int orc_voxel_downsample(double voxel_size, const float* pts4xN, const float* nrm3xN, int N, double* out3xK,
                         double* outn3xK, int32_t* first_idx);

/* map maintenance, Submap::insertScan (Submap.cpp:39-96) and the cropper on a double cloud; see prep.cpp */
int orc_map_insert_scan(const double* map3xM, const double* mapn3xM, int M, const float* scan4xN, const float* scann3xN,
                        int N, const double* T16, double voxel_size, int crop_kind, double radius_min, double radius_max,
                        double min_z, double max_z, const double* center3, int invert, double* out3xK, double* outn3xK,
                        int* n_outside);
int orc_crop_d(int kind, double radius_min, double radius_max, double min_z, double max_z, const double* center3,
               int invert, const double* pts3xN, int N, int32_t* kept_idx);

/* the same split as the reference's initReference / compute (one reference, many readings) */
void* orc_icp_ref_create(const float* ref4xM, const float* refn3xM, int M, int mean_mode);
void orc_icp_ref_free(void* h);
int orc_icp_compute(void* href, const orc_icp_params* prm, const float* read4xN, const float* readn3xN, int N,
                    const float* Tinit16, orc_icp_result* res, int max_log, float* logH, float* logb, float* logx,
                    float* logT, int* logP, int32_t* last_ids, float* last_dists, float* last_w, const float* forcedT,
                    int n_forced);

#ifdef __cplusplus
}
#endif
#endif
