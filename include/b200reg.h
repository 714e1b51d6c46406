/* b200reg.h — C ABI of the B200-native point-cloud registration core.
 *
 * Drop-in boundary for ONE hot path of leggedrobotics/open3d_slam_private: the modules that run inside each ICP
 * iteration of libpointmatcher's PointMatcher<float>::ICP. Every entry point names the reference interface it
 * replaces (paths relative to the reference root). Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions (identical to the reference's Eigen memory images, so shims pass `.data()` straight through):
 *   features   4 x N float, column-major  => N x {x,y,z,1}           (PointMatcher.h:222-403, DataPoints::features)
 *   normals    3 x N float, column-major  => N x {nx,ny,nz}          (descriptor block "normals")
 *   dists/ids/weights  k x N, column-major => per query k contiguous  (PointMatcher.h:416-442, Matches/OutlierWeights)
 *   ids are int32, invalid = -1; dists are SQUARED L2 in float, invalid = +inf (MatchersImpl.cpp:96-97)
 *   4x4 transforms are column-major float[16] (Eigen default)
 * All host arrays are caller-owned and caller-sized; the library never frees caller memory. Device memory is owned
 * by the handles. Handles are single-owner (not internally locked); calls on different handles are re-entrant.
 * Every function returns a b200_status; b200_last_error() gives the message for the calling thread.
 * There is NO CPU fallback: without a CUDA device every compute entry point returns B200_E_CUDA.
 */
#ifndef B200REG_H_
#define B200REG_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  B200_OK = 0,
  B200_E_NO_MATCHES = 1,   /* PointMatcher::ConvergenceError ("no point to minimize", Matches.cpp:75, ErrorMinimizer.cpp:76) */
  B200_E_BAD_PARAM = 2,    /* PointMatcherSupport::InvalidParameter / InvalidField */
  B200_E_NAN = 3,          /* ConvergenceError("abs rotation norm not a number"), TransformationCheckersImpl.cpp:147 */
  B200_E_CUDA = 4,         /* CUDA runtime failure or no device */
  B200_E_UNSUPPORTED = 5,  /* accepted by the reference, not by this core (e.g. knn > 64, force2D) */
  B200_E_NOT_INITIALIZED = 6, /* std::runtime_error("You must setup a matcher before running ICP") analogue */
  B200_E_BOUND = 7         /* ConvergenceError("limit out of bounds: ..."), BoundTransformationChecker,
                              TransformationCheckersImpl.cpp:214-221 */
} b200_status;

const char* b200_last_error(void);
int b200_version(void);
/* sizeof of the ABI structs as compiled (0 b200_icp_params, 1 b200_icp_result, 2 b200_index_opts,
 * 3 b200_surface_normal_params, 4 b200_o3d_params, 5 b200_o3d_result, 6 b200_ransac_params, 7 b200_ransac_result), so a
 * binding can verify its mirror */
int b200_struct_size(int which);
/* number of visible CUDA devices (0 when none / driver missing) */
int b200_device_count(void);

/* ------------------------------------------------------------------------------------------------------------
 * Matcher — replaces MatchersImpl<float>::KDTreeMatcher (MatchersImpl.h:70-101, MatchersImpl.cpp:60-102) and
 * KDTreeVarDistMatcher (MatchersImpl.cpp:108-150). The libnabo kd-tree becomes a device-resident uniform-grid index.
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct b200_index b200_index;

typedef struct {
  float cell_size;       /* <= 0: choose automatically from the point density */
  float target_per_cell; /* <= 0: default (points per occupied cell aimed at by the automatic choice) */
  uint64_t max_cells;    /* 0: default cap on the dense cell table */
} b200_index_opts;

/* KDTreeMatcher::init(filteredReference) — MatchersImpl.cpp:78-83. normals3xM may be NULL. opts may be NULL. */
int b200_index_create(b200_index** out, int device, const float* ref4xM, const float* normals3xM, int M,
                      const b200_index_opts* opts);
void b200_index_destroy(b200_index* idx);
/* Re-initialise an existing index with a new cloud (KDTreeMatcher::init on the next reference, Mapper.cpp:329-347's
 * periodic rebuild; also the per-scan cloud whose normals are wanted): device buffers are reused when large enough. */
int b200_index_update(b200_index* idx, const float* ref4xM, const float* normals3xM, int M, const b200_index_opts* opts);
/* KDTreeMatcher::findClosests(filteredReading) — MatchersImpl.cpp:85-102. epsilon is accepted for interface
 * compatibility; the search is always exact (the exact answer satisfies every epsilon). maxDistPerPoint (1 x N,
 * nullable) reproduces KDTreeVarDistMatcher (MatchersImpl.cpp:133-150). visit_count (nullable) receives the number
 * of candidate points examined (Matcher::visitCounter). */
int b200_index_knn(b200_index* idx, const float* query4xN, int N, int knn, float epsilon, float maxDist,
                   const float* maxDistPerPoint, int32_t* ids_kxN, float* dists_kxN, uint64_t* visit_count);
/* measurement aid: CUDA-event time (ms) of the kernels of the last b200_index_knn / b200_index_self_normals call on
 * this handle (host<->device copies excluded) */
int b200_index_last_device_ms(const b200_index* idx, float* ms);
/* measurement aid: how many points of the last b200_index_self_normals call the cell-block tile kernel handed to the
 * exact warp-cooperative kernel (k-th neighbour not provably inside the staged region, or a crowded region) */
int b200_index_last_normals_fallback(b200_index* idx, int* n);
/* build statistics: cell size, grid dims, occupied cells, build time (ms, device) */
int b200_index_info(const b200_index* idx, float* cell_size, int32_t dims[3], int64_t* occupied_cells,
                    float* build_ms);

/* ------------------------------------------------------------------------------------------------------------
 * DataPointsFilter — replaces SurfaceNormalDataPointsFilter::inPlaceFilter (SurfaceNormal.cpp:82-290).
 * Any output pointer may be NULL (the corresponding keep* flag of the reference is false).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct {
  int knn;            /* default 5 (SurfaceNormal.h:70) */
  float maxDist;      /* default inf */
  float epsilon;      /* accepted, search is exact */
  int sortEigen;      /* 0/1 */
  int smoothNormals;  /* 0/1 (SurfaceNormal.cpp:265-288: in place, in index order; knn <= 32) */
} b200_surface_normal_params;

int b200_surface_normals(int device, const float* feat4xN, int N, const b200_surface_normal_params* prm,
                         float* normals3xN, float* densities1xN, float* eigValues3xN, float* eigVectors9xN,
                         float* matchedIds_kxN, float* meanDists1xN);
/* SurfaceNormalDataPointsFilter::inPlaceFilter (SurfaceNormal.cpp:112-262) of the indexed cloud itself, outputs as
 * b200_surface_normals above; with a kept index (b200_index_update per scan) nothing is allocated per call. */
int b200_index_self_normals(b200_index* idx, const b200_surface_normal_params* prm, float* normals, float* densities,
                            float* eigValues, float* eigVectors, float* matchedIds, float* meanDists);

/* ------------------------------------------------------------------------------------------------------------
 * OutlierFilter — replaces OutlierFiltersImpl<float>::{MaxDist,MinDist,MedianDist,TrimmedDist,SurfaceNormal}
 * OutlierFilter::compute (OutlierFiltersImpl.cpp:75-81, 95-101, 117-125, 140-147, 236-288) and the empty chain of
 * OutlierFilters::compute (OutlierFilter.cpp:66-83). n = k*N entries.
 * ---------------------------------------------------------------------------------------------------------- */
typedef enum {
  B200_OUTLIER_NULL = 0,
  B200_OUTLIER_MAXDIST = 1,
  B200_OUTLIER_MINDIST = 2,
  B200_OUTLIER_MEDIANDIST = 3,
  B200_OUTLIER_TRIMMEDDIST = 4
} b200_outlier_kind;
int b200_outlier_dist(int device, int kind, float param, const float* dists, int64_t n, float* weights);
/* Matches::getDistsQuantile (Matches.cpp:60-87) */
int b200_dists_quantile(int device, const float* dists, int64_t n, float quantile, float* out);
/* SurfaceNormalOutlierFilter::compute; either normals pointer NULL => all ones (OutlierFiltersImpl.cpp:273-283) */
int b200_outlier_surface_normal(int device, const float* nread3xN, const float* nref3xM, int M,
                                const int32_t* ids_kxN, int knn, int N, float maxAngle, float* weights);

/* b200_outlier_trimmed_noalloc — NO COUNTERPART IN THE REFERENCE; added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). The same
 * result as b200_outlier_dist(B200_OUTLIER_TRIMMEDDIST, ratio, ...) with these differences:
 *   - d_dists (n floats, squared distances, +inf = invalid) and d_weights (n floats) are DEVICE pointers owned by the
 *     caller; d_status_limit (nullable) is a DEVICE uint32[2] that receives {b200_status decided on the device
 *     (B200_OK or B200_E_NO_MATCHES), bits of the float limit = Matches::getDistsQuantile}; when the status is
 *     B200_E_NO_MATCHES no weight is written;
 *   - the work is only ENQUEUED on cuda_stream (cudaStream_t; NULL = the legacy default stream): no synchronisation,
 *     no host read-back, and no cudaMalloc / cudaFree / allocating routine of any kind. The radix-select scratch
 *     (24 640 bytes) is a file-scope __device__ array every GPU that loads the library holds until the process exits;
 *   - n must not exceed B200_NOALLOC_MAX_N (larger inputs: B200_E_UNSUPPORTED); ratio outside [0, 1] is
 *     B200_E_BAD_PARAM (checked on the host, before the emptiness the device detects);
 *   - every call shares that one array per device (each call clears it first): calls whose GPU work overlaps — other
 *     streams, other host threads on the same device — RACE; serialise them per device. The first call on a device
 *     resolves the array's device address (cudaGetSymbolAddress); make that call once before a graph capture. */
#define B200_NOALLOC_MAX_N (1 << 24)
int b200_outlier_trimmed_noalloc(int device, const float* d_dists, int64_t n, float ratio, float* d_weights,
                                 uint32_t* d_status_limit, void* cuda_stream);

/* ------------------------------------------------------------------------------------------------------------
 * ErrorMinimizer — replaces ErrorMinimizer::ErrorElements (ErrorMinimizer.cpp:59-193),
 * ICP::calculateOptimizationHessian (ICP.cpp:1513-1566), PointToPlaneErrorMinimizer::compute_in_place with
 * degeneracyAwareness = kNone (PointToPlane.cpp:113-265, 275-400) and PointToPointErrorMinimizer::compute_in_place
 * (PointToPoint.cpp:62-103).
 * ---------------------------------------------------------------------------------------------------------- */
/* order-preserving compaction; outputs have capacity k*N columns; any output may be NULL */
int b200_error_elements(int device, const float* read4xN, const float* ref4xM, const float* refn3xM, int M,
                        const float* w_kxN, const int32_t* ids_kxN, const float* dists_kxN, int knn, int N,
                        float* read4xP, float* ref4xP, float* refn3xP, float* w_P, int32_t* ids_P, float* dists_P,
                        int* P, float* pointUsedRatio, float* weightedPointUsedRatio);
/* H (6x6 column-major) and b from already matched pairs */
int b200_p2plane_hessian(int device, const float* read4xP, const float* ref4xP, const float* refn3xP,
                         const float* w_P, int P, float* H36, float* b6);
/* 6x6 solve + SE(3) construction on device. info: 0 full rank, 1 rank-deficient minimal norm, 2 SVD fallback */
int b200_p2plane_solve(int device, const float* H36, const float* b6, float* x6, float* T16, int* info);
int b200_p2point(int device, const float* read4xP, const float* ref4xP, const float* w_P, int P, float* T16);

/* ------------------------------------------------------------------------------------------------------------
 * Transformation — replaces TransformationsImpl<float>::RigidTransformation::inPlaceCompute incl.
 * checkParameters / correctParameters (TransformationsImpl.cpp:61-170). In place; normals may be NULL.
 * ---------------------------------------------------------------------------------------------------------- */
int b200_rigid_apply(int device, const float* T16, float* feat4xN, float* normals3xN, int N);

/* ------------------------------------------------------------------------------------------------------------
 * ICP chain — replaces PointMatcher<float>::ICP::{initReference, compute, operator()} (ICP.cpp:793-1349) for the
 * chain {RigidTransformation, KDTreeMatcher, [TrimmedDist][MaxDist][SurfaceNormal]OutlierFilter,
 * PointToPlane|PointToPoint ErrorMinimizer, [Counter][Differential][Bound]TransformationChecker}, degeneracyAwareness
 * None. All iterations run in ONE persistent cooperative kernel; the host reads back one result struct.
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct b200_icp b200_icp;

#define B200_DEGENERACY_NONE 0
#define B200_DEGENERACY_OPTIMIZED_EQUALITY 1

typedef struct {
  int knn;                       /* KDTreeMatcher.knn (1..64) */
  float epsilon;                 /* KDTreeMatcher.epsilon (accepted; exact search) */
  float maxDist;                 /* KDTreeMatcher.maxDist (inf allowed) */
  int use_trimmed;               /* TrimmedDistOutlierFilter */
  float trimmed_ratio;
  int use_maxdist_filter;        /* MaxDistOutlierFilter */
  float maxdist_filter;
  int use_surface_normal_filter; /* SurfaceNormalOutlierFilter */
  float surface_normal_max_angle;
  int minimizer;                 /* 0 PointToPlaneErrorMinimizer, 1 PointToPointErrorMinimizer */
  int max_iterations;            /* CounterTransformationChecker.maxIterationCount; <= 0: absent */
  int use_differential;          /* DifferentialTransformationChecker */
  float min_diff_rot, min_diff_trans;
  int smooth_length;
  float cell_size;               /* index option, <= 0 automatic */
  /* degeneracyAwareness (ICP::loadAdditionalYAMLContent, ICP.cpp:575-760): B200_DEGENERACY_NONE or
   * B200_DEGENERACY_OPTIMIZED_EQUALITY = "OptimizedEqualityConstraints" (X-ICP; ICP.cpp:629-670, the setting of
   * open3d_slam_ros/param/icp.yaml). Needs knn = 1 and the PointToPlaneErrorMinimizer (ICP.cpp:1114). */
  int degeneracy;
  float xicp_enough_information;        /* enoughInformationThreshold */
  float xicp_insufficient_information;  /* insufficientInformationThreshold */
  float xicp_min_alignment_angle_deg;   /* point2NormalMinimalAlignmentAngleThreshold (degrees) */
  float xicp_strong_alignment_angle_deg;/* point2NormalStrongAlignmentAngleThreshold (degrees) */
  /* BoundTransformationChecker (TransformationCheckersImpl.cpp:165-222): init stores rotation and translation of the
   * first T_iter (identity, ICP.cpp:997); check raises ConvergenceError -> B200_E_BOUND as soon as
   * angularDistance(q(T_iter), q0) > maxRotationNorm or |t - t0| > maxTranslationNorm. Evaluated after the
   * Differential checker and before the Counter's stop takes effect (a chain listing Counter first would report
   * MaxNumIterationsReached instead when both trigger in the same iteration). */
  int use_bound;
  float max_rotation_norm, max_translation_norm;   /* defaults 1, 1 (TransformationCheckersImpl.h:117-123) */
} b200_icp_params;

typedef struct {
  int status;                    /* b200_status decided on device */
  int iterations;                /* Inspector stat "IterationsCount" */
  int max_iterations_reached;    /* ICP::maxNumIterationsReached */
  int n_matches_last;            /* ErrorElements size of the last iteration */
  float T[16];                   /* T_refIn_readIn result (ICP.cpp:1345) */
  float T_iter[16];
  float T_refIn_refMean[16];
  float T_refMean_readMean[16];
  float weighted_point_used_ratio; /* "OverlapRatio" */
  float point_used_ratio;
  uint64_t points_touched;       /* "PointCountTouched" */
  float device_ms;               /* CUDA-event time of the registration kernel */
  float h2d_ms, total_ms;        /* upload time; first event .. end of the registration kernel */
  float e2e_ms;                  /* first event .. after the device->host copy of this result (0 for run_resident) */
  int kernel_launches;           /* kernels this call launched on the device (uploads/memsets not counted) */
  int grid_ctas;                 /* CTAs of the persistent registration kernel */
  /* X-ICP (degeneracy = B200_DEGENERACY_OPTIMIZED_EQUALITY): localizabilityXyz_/Rpy_ of the first iteration's
   * detection (1 = kLocalizable, PointMatcher.h:602-607, 638-639) per eigenvector index; translation/rotation
   * eigenvectors (columns, column-major 3x3, optimisation frame) after the last iteration; whether the prior was
   * returned (ICP.cpp:1335-1341: T = T_refIn_readIn); the alignment sums over all matched pairs (double):
   * xicp_sums[2e], [2e+1] = combined/high contribution for translation eigenvector e, [6+2e], [7+2e] for rotation
   * eigenvector e; xicp_borderline = number of flags whose sums lay within the rounding band (0.5 - 1 %) of a threshold and were therefore
   * decided by replaying the reference's sequential float recurrence (ICP.cpp:2128-2155) on the device. */
  int returned_prior, xicp_borderline;
  int loc_xyz[3], loc_rpy[3];
  float eig_t[9], eig_r[9];
  double xicp_sums[12];
} b200_icp_result;

void b200_icp_default_params(b200_icp_params* prm); /* ICPChainBase::setDefault, ICP.cpp:100-113 */
int b200_icp_create(b200_icp** out, int device, const b200_icp_params* prm);
void b200_icp_destroy(b200_icp* icp);
/* ICP::initReference(referenceIn) — ICP.cpp:847-898: mean-centre, build the index. refn3xM may be NULL for
 * PointToPoint. */
int b200_icp_init_reference(b200_icp* icp, const float* ref4xM, const float* refn3xM, int M);
/* ICP::compute(readingIn, <unused>, T_refIn_readIn, initializeMatcherWithInputReference=false) — ICP.cpp:813-844,
 * the call Mapper.cpp:372-373 makes. readn3xN may be NULL. Tinit16 NULL = identity (ICP::operator(), ICP.cpp:793). */
int b200_icp_compute(b200_icp* icp, const float* read4xN, const float* readn3xN, int N, const float* Tinit16,
                     b200_icp_result* res);
/* The same two calls taking Open3D clouds as they lie in memory (std::vector<Eigen::Vector3d>: 3 x N column-major
 * DOUBLES; normals nullable): open3dToPointmatcher (open3d_utils/open3d_conversions/src/open3d_conversions.cpp:57-118),
 * the per-point double -> float cast the Mapper runs on the host before every initReference / compute
 * (Mapper.cpp:337-343 "This is time consuming"), happens on the device — the cloud crosses PCIe once, unconverted. */
int b200_icp_init_reference_o3d(b200_icp* icp, const double* pts3xM, const double* nrm3xM, int M);
int b200_icp_compute_o3d(b200_icp* icp, const double* pts3xN, const double* nrm3xN, int N, const float* Tinit16,
                         b200_icp_result* res);
/* Same registration with the reading already resident on the device (for kernel-only timing):
 * upload once, then run any number of times. */
int b200_icp_upload_reading(b200_icp* icp, const float* read4xN, const float* readn3xN, int N);
int b200_icp_run_resident(b200_icp* icp, const float* Tinit16, b200_icp_result* res);
/* The same chain over a SPATIALLY PARTITIONED map (BASELINE config C5): one handle per rank holds one
 * slab of the map padded by maxDist; the scan is replicated. Per iteration of ICP::computeWithTransformedReference
 * (ICP.cpp:1027-1311):
 *   b200_icp_slab_match    matcher->findClosests (:1051) for the queries the rank owns under the current pose; writes
 *                          the exchange arrays on the device: d_match9xN = N float4 matched points | N float4 matched
 *                          normals | N float squared distances, d_idN = N global indices; zeros for queries of other
 *                          ranks. The caller sums both arrays over the ranks (one all-reduce each; nothing to do for
 *                          a single slab);
 *   b200_icp_slab_iterate  outlierFilters.compute incl. TrimmedDist's quantile over ALL pairs (:1055,
 *                          OutlierFiltersImpl.cpp:140-147, Matches.cpp:60-87), ErrorElements, errorMinimizer->compute,
 *                          T_iter update, transformationCheckers.check (:1292) on the summed arrays; *iterate = 0 when
 *                          a checker stopped the loop (or it failed: see the result's status).
 * Every rank computes the same pose from the same sums (no broadcast). Result, iteration count, matches and weights
 * are bit-identical to b200_icp_compute on one device holding the whole map. knn = 1, degeneracyAwareness None.
 * gid_M: global index of every slab point (null: identity); map_mean3: centroid of the WHOLE map (the reference frame
 * of ICP.cpp:883-890 is the whole map's); [lo, hi): the rank's own interval along axis (0..2), map frame, +-inf at
 * the ends. d_* are DEVICE pointers of the caller (e.g. torch tensors) on the handle's device. */
int b200_icp_init_reference_slab(b200_icp* icp, const float* ref4xM, const float* refn3xM, int M, const int32_t* gid_M,
                                 const double* map_mean3, int axis, float lo, float hi);
int b200_icp_slab_begin(b200_icp* icp, const float* read4xN, const float* readn3xN, int N, const float* Tinit16);
int b200_icp_slab_match(b200_icp* icp, float* d_match9xN, int32_t* d_idN);
/* the same, only enqueued on the handle's stream, and that stream (a cudaStream_t): a caller that runs its exchange on
 * this stream (NCCL through torch: make it the current stream) needs no host synchronisation between the match, the
 * all-reduces and b200_icp_slab_iterate, which is enqueued on it as well */
int b200_icp_slab_match_async(b200_icp* icp, float* d_match9xN, int32_t* d_idN);
void* b200_icp_slab_stream(b200_icp* icp);
int b200_icp_slab_iterate(b200_icp* icp, const float* d_match9xN, const int32_t* d_idN, int* iterate);
int b200_icp_slab_finish(b200_icp* icp, b200_icp_result* res);
/* ICP::matches / ICP::outlierWeights of the last iteration (PointMatcher.h ICPChainBase members) */
int b200_icp_get_matches(b200_icp* icp, int32_t* ids_kxN, float* dists_kxN, float* weights_kxN);
/* per-iteration trace (first max_log iterations): H 36, b 6, x 6, T_iter 16 floats, P int. Returns count in *n. */
int b200_icp_get_log(b200_icp* icp, int max_log, float* H, float* b, float* x, float* T_iter, int32_t* P, int* n);
/* profiling aid: per-iteration phase durations in microseconds as seen by CTA 0 of the registration kernel
 * (%globaltimer): us[6*i + {0 match, 1 trimmed select, 2 linearise, 3 wait for all partials, 4 reduce+solve+checks,
 * 5 whole iteration}] */
int b200_icp_get_phase_times(b200_icp* icp, int max_log, float* us, int* n);
/* profiling aid, filled only when the handle was created with env B200_ICP_DEBUG=1 (adds global atomics to the
 * kernel): st[12*i + {0 cell rows resolved, 1 candidates examined, 2 slowest CTA's matching phase (us), 3 mean CTA
 * matching phase (us), 4 reduction of partials, 5 6x6 solve, 6 pose update, 7 checkers (us, CTA 0), 8 iteration shape (1 single-barrier, 0 general, 2 single-barrier attempt fell back),
 * 9 bracket candidates, 10 TrimmedDist limit, 11 bracket half-width used by the previous attempt}]. Columns 4-9 are filled without the env switch as well. */
int b200_icp_get_debug_stats(b200_icp* icp, int max_log, double* st, int* n);
/* profiling aid (B200_ICP_DEBUG=1): per warp slot of the first iteration {rows, candidates, passes, ns} */
int b200_icp_get_debug_warps(b200_icp* icp, uint32_t* out4xS, int max_slots, int* n);
/* evict L2 (writes a buffer larger than L2) on the handle's stream; for benchmarking hygiene */
int b200_icp_flush_l2(b200_icp* icp);

/* ------------------------------------------------------------------------------------------------------------
 * Boundary B — replaces the Open3D v0.15.1 pipeline behind o3d_slam::CloudRegistration::registerClouds
 * (open3d_slam/src/CloudRegistration.cpp:17-23 RegistrationGeneralizedICP, :62-66 and :89-94 RegistrationICP;
 * constraint_builders.cpp:65-68): 1-NN within max_correspondence_distance, point-to-plane or generalized-ICP
 * linearisation (covariances C = R diag(eps,1,1) R^T derived from the normals, as Open3D does when the clouds carry
 * normals), x = -(JtJ)^-1 Jtr, T(x) = Rz Ry Rx, stop on |d fitness| and |d inlier_rmse| below the criteria or after
 * max_iteration updates. Transformations are column-major DOUBLE 4x4 (Eigen::Matrix4d).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct b200_o3d b200_o3d;
typedef struct {
  int kind;                          /* 0 TransformationEstimationPointToPlane, 1 ...PointToPoint, 2 ...ForGeneralizedICP */
  float max_correspondence_distance; /* IcpParameters::maxCorrespondenceDistance_ (Parameters.hpp:66-72) */
  int max_iteration;                 /* ICPConvergenceCriteria::max_iteration_ <- IcpParameters::maxNumIter_ */
  double relative_fitness, relative_rmse; /* ICPConvergenceCriteria, Open3D defaults 1e-6 */
  double gicp_epsilon;               /* TransformationEstimationForGeneralizedICP::epsilon_, 1e-3 */
  float cell_size;                   /* index option, <= 0 automatic */
} b200_o3d_params;
typedef struct {
  double T[16];            /* RegistrationResult::transformation_ */
  double fitness;          /* #correspondences / #source points */
  double inlier_rmse;      /* sqrt(mean squared correspondence distance) */
  int iterations;          /* transformation updates applied */
  int n_correspondences;   /* RegistrationResult::correspondence_set_.size() */
  int kernel_launches;
  float device_ms, total_ms;
} b200_o3d_result;
void b200_o3d_default_params(b200_o3d_params* prm);
int b200_o3d_create(b200_o3d** out, int device, const b200_o3d_params* prm);
void b200_o3d_destroy(b200_o3d* h);
/* target cloud with normals (estimateNormalsOrCovariancesIfNeeded has run on it, CloudRegistration.cpp:27-46);
 * normals may be NULL for kind 1 */
int b200_o3d_set_target(b200_o3d* h, const float* tgt4xM, const float* tgtn3xM, int M);
/* target cloud WITHOUT normals: estimateNormalsOrCovariancesIfNeeded (CloudRegistration.cpp:67-79) runs on the device
 * against the target's own index -- Open3D semantics as in b200_o3d_estimate_normals below (at most knn neighbours
 * within maxDist, IcpParameters::knn_ / maxDistanceKnn_; 1 <= knn <= 32) -- and the normals never leave the device */
int b200_o3d_set_target_estimate_normals(b200_o3d* h, const float* tgt4xM, int M, int knn, float maxDist);
/* estimateNormalsOrCovariancesIfNeeded (CloudRegistration.cpp:27-46, 67-79) for a cloud that is not the target (the
 * SOURCE of a generalized-ICP registration, or any cloud handed to the virtual): Open3D EstimateNormals with
 * KDTreeSearchParamHybrid(radius, knn) — neighbourhood = at most knn nearest points within radius, the point itself
 * included; covariance in double, normalised by the count; fewer than 3 neighbours -> (0, 0, 1) — then
 * NormalizeNormals and OrientNormalsTowardsCameraLocation(origin). normals3xN: column-major doubles like Open3D's
 * normals_. N = 0 returns B200_OK without touching anything (the reference does nothing for an empty cloud). */
int b200_o3d_estimate_normals(b200_o3d* h, const float* pts4xN, int N, int knn, float radius, double* normals3xN);
/* registerClouds(source, target, init); srcn3xN is required for kind 2 only */
int b200_o3d_register(b200_o3d* h, const float* src4xN, const float* srcn3xN, int N, const double* init16,
                      b200_o3d_result* res);
/* registerClouds for a source WITHOUT normals (generalized ICP: estimateNormalsOrCovariancesIfNeeded would have run on
 * it first, CloudRegistration.cpp:27-46): the source is indexed on the device, its normals are estimated there with
 * the semantics of b200_o3d_estimate_normals and go straight into the registration — one H2D of the points, no host
 * round trip of the normals. total_ms / device time of the result include the estimation. */
int b200_o3d_register_estimate_normals(b200_o3d* h, const float* src4xN, int N, int knn, float radius,
                                       const double* init16, b200_o3d_result* res);
/* Loop-closure constraint (first slice): open3d GetInformationMatrixFromPointClouds(source, target,
 * max_correspondence_distance, T) as called at PlaceRecognition.cpp:148-149 (Open3D v0.15.1, not in the reference
 * tree): correspondences = 1-NN of the transformed source in the handle's target within the distance; information =
 * sum over them of G^T G with the rows G = [-[t]x | I] of the matched TARGET point t. info36: column-major 6x6 doubles.
 * Also the fitness / inlier RMSE of that match (minRefinementFitness_ test, PlaceRecognition.cpp:118-128); nullable. */
int b200_o3d_information_matrix(b200_o3d* h, const float* src4xN, int N, const double* T16,
                                float max_correspondence_distance, double* info36, double* fitness, double* inlier_rmse,
                                int* n_correspondences);
/* Spatially partitioned map: each device holds one slab of the map (padded by a halo of
 * max_correspondence_distance) as its target and answers only for the source points whose transformed coordinate along
 * `axis` lies in [lo, hi) -- exactly one owner per point, so the union of the slabs' correspondences is the
 * single-device result. gid_M (nullable) maps slab-local target indices to indices in the whole map. The host drives
 * the loop and sums the 32 doubles of every slab between the two steps (NCCL allreduce on `stream`):
 *   begin; repeat max_iteration + 1 times { step_linearize(sums); allreduce(sums); step_update(sums) }; finish.
 * sums layout: 21 JtJ upper (column-major packed) + 6 Jtr (kind 1: Umeyama moments in 0..14), [27] correspondence
 * count, [28] sum of squared distances, [29..31] zero. Steps after convergence write zeros / do nothing. */
int b200_o3d_set_partition(b200_o3d* h, int axis, float lo, float hi, const int32_t* gid_M);
int b200_o3d_begin(b200_o3d* h, const float* src4xN, const float* srcn3xN, int N, const double* init16,
                   void* cuda_stream /* cudaStream_t; NULL: the handle's own; cudaStreamLegacy (0x1) names the default stream */);
int b200_o3d_step_linearize(b200_o3d* h, double* d_sums32 /* DEVICE pointer */);
int b200_o3d_step_update(b200_o3d* h, const double* d_sums32 /* DEVICE pointer, totals over all slabs */);
int b200_o3d_poll(b200_o3d* h, int* done, int* updates); /* synchronises the stream */
/* The same loop with the sums exchanged over NVLink / NVSwitch PEER MEMORY inside the solve kernel instead of an
 * allreduce between kernels (one process per GPU, or several handles in one process). Every rank:
 *   b200_o3d_peer_init    allocates its mailbox; returns its device pointer and/or its 64-byte CUDA IPC handle;
 *   (exchange the pointers / handles: torch.distributed.all_gather_object, MPI, ...)
 *   b200_o3d_peer_connect maps the peers' mailboxes: mailboxes_dev[world] (+ devices[world], same process) or
 *                         ipc_handles (world x 64 bytes, other processes); entry `rank` is ignored;
 *   b200_o3d_enqueue_peers per registration: enqueues every iteration (match+linearise, exchange+solve) on `stream`
 *                         (NULL: the handle's own) without any host synchronisation; b200_o3d_finish fetches the result.
 * All ranks must run the same registrations in the same order with the same max_iteration (the packets carry a
 * sequence number, consecutive across registrations); a peer that does not deliver within 5 s (B200_O3D_PEER_TIMEOUT_MS)
 * makes b200_o3d_finish return B200_E_CUDA, after which the ranks are out of step: b200_o3d_peer_init +
 * b200_o3d_peer_connect must be re-run on ALL ranks, collectively. The same holds for calling b200_o3d_peer_init again
 * for any other reason: it frees this rank's mailbox, which the peers still have mapped — every rank must re-connect
 * before the next registration. */
int b200_o3d_peer_init(b200_o3d* h, int rank, int world, void** mailbox_dev /* nullable */, void* ipc_handle_64 /* nullable */);
int b200_o3d_peer_connect(b200_o3d* h, void* const* mailboxes_dev /* nullable */, const int* devices /* nullable */,
                          const void* ipc_handles_64 /* nullable */);
int b200_o3d_enqueue_peers(b200_o3d* h, const float* src4xN, const float* srcn3xN, int N, const double* init16,
                           void* stream);
int b200_o3d_finish(b200_o3d* h, b200_o3d_result* res);
/* correspondence_set_: matched target index per source point (-2: owned by another slab) (-1: none) and squared distances; either may be NULL */
int b200_o3d_get_correspondences(b200_o3d* h, int32_t* ids_N, float* dists_N);

/* ------------------------------------------------------------------------------------------------------------
 * Scan pre-processing — what open3d_slam runs on every scan before it registers it
 * (ScanToMapRegistration.cpp:36-54, Odometry.cpp:22-26) and on the map before a patch is cut out (:90-104).
 * Clouds are the float memory images used everywhere else in this header (4 x N features, 3 x N normals).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct b200_prep b200_prep;
int b200_prep_create(b200_prep** out, int device);
void b200_prep_destroy(b200_prep* h);

/* croppingVolumeFactory's volumes (open3d_slam/src/croppers.cpp:25-47) */
#define B200_CROP_NONE 0          /* CroppingVolume base class: everything is inside (:49-51) */
#define B200_CROP_CYLINDER 1      /* CylinderCroppingVolume (:160-164): radius_max, min_z, max_z */
#define B200_CROP_MIN_RADIUS 2    /* MinRadiusCroppingVolume (:147-151): radius_min */
#define B200_CROP_MAX_RADIUS 3    /* MaxRadiusCroppingVolume (:129-133): radius_max */
#define B200_CROP_MINMAX_RADIUS 4 /* MinMaxRadiusCroppingVolume (:113-118) */
typedef struct {
  int kind;
  int invert;            /* CroppingVolume::setIsInvertVolume (:56-58) */
  double radius_min, radius_max, min_z, max_z;
  double center[3];      /* pose_.translation() (CroppingVolume::setPose, :59-61) */
} b200_crop_params;
/* CroppingVolume::crop(cloud) (croppers.cpp:66-95): the points inside the volume, in input order, with their normals
 * (nrm3xN / outn3xN nullable) and — CroppingVolume::getIndicesWithinVolume (:62-71) — their indices (nullable).
 * Outputs have capacity N; *n_out receives the number kept. */
int b200_prep_crop(b200_prep* h, const b200_crop_params* prm, const float* pts4xN, const float* nrm3xN, int N,
                   float* out4xN, float* outn3xN, int32_t* kept_idx, int* n_out);
/* o3d_slam::randomDownSample (helpers.cpp:100-106; ScanToMapRegistration.cpp:44,54, Odometry.cpp:26) -> open3d
 * PointCloud::RandomDownSample(ratio): size_t(ratio * N) points drawn without replacement (Open3D shuffles with an
 * mt19937 seeded from std::random_device, so the reference's own choice is not reproducible); ratio >= 1 leaves the
 * cloud as it is (helpers.cpp:101-103), ratio <= 0 is Open3D's "sampling_ratio must lie in (0, 1]" error. Here the
 * sample is the size_t(ratio * N) points with the smallest keys of a bijective 32-bit hash of (index, seed): uniform,
 * reproducible for a given seed; output in input order. */
int b200_prep_random_downsample(b200_prep* h, double ratio, uint32_t seed, const float* pts4xN, const float* nrm3xN, int N,
                                float* out4xN, float* outn3xN, int32_t* kept_idx, int* n_out);
/* o3d_slam::voxelize (helpers.cpp:108-115) -> open3d PointCloud::VoxelDownSample(voxel_size) (Open3D v0.15.1, not in
 * the reference tree): voxel index floor((p - (min_bound - voxel_size / 2)) / voxel_size), points and normals of a
 * voxel averaged in double in input order. Output (capacity N, column-major 3 x K doubles like Open3D's points_) is
 * ordered by the index of the first point of each voxel (Open3D's order is unspecified); first_idx nullable. */
int b200_prep_voxel_downsample(b200_prep* h, double voxel_size, const float* pts4xN, const float* nrm3xN, int N,
                               double* out3xK, double* outn3xK, int32_t* first_idx, int* n_out);
/* how many voxels of the last b200_prep_voxel_downsample had to be summed in input order (their values span so many
 * binades that a double sum could round; every other voxel's sum is exact in any order and is formed with atomics) */
int b200_prep_last_ordered_voxels(const b200_prep* h, int* n);
/* device time of the last call's kernels (CUDA events, uploads and read-backs excluded) */
int b200_prep_last_device_ms(const b200_prep* h, float* ms);

/* ------------------------------------------------------------------------------------------------------------
 * Map maintenance on the device — what open3d_slam does to the active submap between two
 * registrations, with the map resident in HBM (stored as doubles like Open3D's points_ / normals_):
 *   b200_map_insert_scan  = Submap::insertScan (open3d_slam/src/Submap.cpp:39-96): o3d_slam::transform of the scan by
 *       mapToRangeSensor (helpers.cpp:283-318), mapCloud_ += scan, voxelizeWithinCroppingVolume (helpers.cpp:117-200):
 *       points outside `crop` (the map-builder cropper, its centre set to the sensor position by the caller,
 *       Submap.cpp:86) pass through in input order; the points inside are averaged per voxel floor(p / voxel_size)
 *       (VoxelHashMap.hpp:48-51), normals averaged then normalised. New map = pass-through points, then the voxels
 *       ordered by the index of their first member (the reference's order is that of an std::unordered_map).
 *       voxel_size <= 0: the map is just the concatenation (helpers.cpp:121-124). An empty scan changes nothing.
 *   b200_map_crop         = CroppingVolume::crop of the map (ScanToMapIcp::cropSubmap, ScanToMapRegistration.cpp:90-96)
 *       as the float images of this header (open3dToPointmatcher's double -> float, open3d_conversions.cpp:57-118)
 *   b200_map_crop_to_icp  = the same patch handed to ICP::initReference (Mapper.cpp:329-347) WITHOUT leaving the
 *       device: cut straight into the registration handle's reference buffers, mean-centred and indexed there.
 * Voxel membership, counts and the pass-through set are exact; voxel means are double atomics (unspecified order):
 * equal to the reference's sequential sums to a few ulps.
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct b200_map b200_map;
int b200_map_create(b200_map** out, int device);
void b200_map_destroy(b200_map* m);
/* replace the map (mapCloud_ = cloud; e.g. the initial map, Submap.cpp:47-51); nrm3xM nullable; column-major doubles */
int b200_map_set(b200_map* m, const double* pts3xM, const double* nrm3xM, int M);
int b200_map_size(const b200_map* m, int* M, int* has_normals);
int b200_map_download(b200_map* m, double* pts3xM, double* nrm3xM);
/* scan: float images (4 x N features, 3 x N normals or NULL); T16: mapToRangeSensor, column-major double 4x4 */
int b200_map_insert_scan(b200_map* m, const float* scan4xN, const float* scann3xN, int N, const double* T16,
                         double voxel_size, const b200_crop_params* crop, int* n_outside, int* new_size);
/* outputs have capacity b200_map_size points; any of them may be NULL; kept_idx = indices into the map */
int b200_map_crop(b200_map* m, const b200_crop_params* crop, float* out4xK, float* outn3xK, int32_t* kept_idx, int* n_out);
int b200_map_crop_to_icp(b200_map* m, const b200_crop_params* crop, b200_icp* icp, int* n_patch);
/* device time of the kernels of the last insert_scan / crop_to_icp (uploads and read-backs excluded) */
int b200_map_last_device_ms(const b200_map* m, float* ms);

/* ------------------------------------------------------------------------------------------------------------
 * Loop-closure front half — what open3d_slam runs once per finished submap and once per
 * loop-closure candidate before the ICP refinement (b200_o3d_register) and the information matrix
 * (b200_o3d_information_matrix). Open3D v0.15.1 is not in the reference tree: semantics recalled from upstream
 * (pipelines/registration/Feature.cpp, Registration.cpp, CorrespondenceChecker.cpp), parity unpinned.
 * Clouds are the float memory images of this header (4 x N); features are 33 x N doubles, column-major like
 * open3d::pipelines::registration::Feature::data_.
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct b200_lc b200_lc;
int b200_lc_create(b200_lc** out, int device);
void b200_lc_destroy(b200_lc* h);
/* Submap::computeFeatures (open3d_slam/src/Submap.cpp:255-275): registration::ComputeFPFHFeature(cloud,
 * KDTreeSearchParamHybrid(radius, max_nn)) — featureRadius_ 2.5, featureKnn_ 100 (Parameters.hpp:124-125). The
 * neighbourhood of a point = at most max_nn nearest points within radius, nearest first, the first one (the point
 * itself) skipped; SPFH = 3 x 11-bin histograms of the pair features, 100 / (n - 1) per neighbour; FPFH = sum of the
 * neighbours' SPFHs weighted by 1 / squared distance, each 11-bin group normalised to 100, plus the point's own SPFH.
 * normals3xN: column-major doubles like Open3D's normals_ (b200_o3d_estimate_normals returns them that way).
 * max_nn <= 2048; nn_count_N (nullable) receives the neighbourhood sizes. N = 0 returns B200_OK. */
int b200_lc_compute_fpfh(b200_lc* h, const float* pts4xN, const double* normals3xN, int N, float radius, int max_nn,
                         double* fpfh33xN, int32_t* nn_count_N);
/* The whole feature step of Submap::computeFeatures after the voxel down-sample (Submap.cpp:269-272) in one call and on
 * one index: EstimateNormals(KDTreeSearchParamHybrid(normalEstimationRadius_, normalKnn_)) + NormalizeNormals +
 * OrientNormalsTowardsCameraLocation(origin) (the semantics of b200_o3d_estimate_normals; normal_knn <= 32), then
 * ComputeFPFHFeature as above on those normals, which never leave the device (normals3xN nullable: returned when the
 * caller wants sparseMapCloud_.normals_ filled). Equal to b200_o3d_estimate_normals followed by b200_lc_compute_fpfh. */
int b200_lc_compute_features(b200_lc* h, const float* pts4xN, int N, float normal_radius, int normal_knn, float fpfh_radius,
                             int fpfh_max_nn, double* normals3xN, double* fpfh33xN);
/* the feature-space search of RegistrationRANSACBasedOnFeatureMatching (KDTreeFlann(feature).SearchKNN(col, 1)):
 * exact nearest column of b for every column of a, squared L2 in double summed in bin order, ties -> smallest
 * index. d2_N nullable. */
int b200_lc_feature_nn(b200_lc* h, const double* a33xN, int N, const double* b33xM, int M, int32_t* nn_N, double* d2_N);
typedef struct {
  int mutual_filter;                  /* PlaceRecognition.cpp:82 passes true */
  double max_correspondence_distance; /* ransacMaxCorrespondenceDistance_ */
  int ransac_n;                       /* ransacModelSize_; 3..8 here (< 3: Open3D returns an empty result) */
  double edge_length_similarity;      /* CorrespondenceCheckerBasedOnEdgeLength(correspondenceCheckerEdgeLength_); <= 0: absent */
  double distance_threshold;          /* CorrespondenceCheckerBasedOnDistance(correspondenceCheckerDistance_); <= 0: absent */
  int max_iteration;                  /* RANSACConvergenceCriteria(ransacNumIter_, ransacProbability_) */
  double confidence;
  uint64_t seed;                      /* sample stream: see below */
  int batch;                          /* hypotheses per launch (0: 65536); does not change the result */
} b200_ransac_params;
typedef struct {
  double T[16];                       /* RegistrationResult::transformation_, column-major */
  double fitness, inlier_rmse;        /* over the feature correspondences (EvaluateRANSACBasedOnCorrespondence) */
  int n_correspondences;              /* correspondence_set_.size(): what PlaceRecognition.cpp:87 tests against
                                         ransacMinCorrespondenceSetSize_ */
  int n_feature_correspondences;      /* size of the set the hypotheses were drawn from */
  int mutual_used;                    /* 0: the mutual set had fewer than 3 * ransac_n pairs, all pairs were used */
  int64_t iterations_run, validations;
  float device_ms;
  int kernel_launches;
} b200_ransac_result;
void b200_lc_default_ransac_params(b200_ransac_params* prm);
/* RegistrationRANSACBasedOnFeatureMatching as called at PlaceRecognition.cpp:81-85 (TransformationEstimationPointToPoint
 * without scaling; both checkers). Open3D draws its samples from an mt19937 seeded by std::random_device inside an
 * OpenMP loop and is not reproducible by design; here sample j of iteration i is a counter-based hash of (seed, i, j)
 * and the iterations are consumed in index order with the semantics of the upstream loop run by one thread (one best
 * result by (fitness, rmse), exit estimate log(1 - confidence) / log(1 - inlier_ratio^ransac_n)): reproducible for a
 * given seed and independent of `batch`. corr2xK (nullable, capacity 2 x Ns): the inlier pairs (source, target). */
int b200_lc_ransac_feature_matching(b200_lc* h, const float* src4xN, int Ns, const float* tgt4xM, int Nt,
                                    const double* src_fpfh33xN, const double* tgt_fpfh33xM, const b200_ransac_params* prm,
                                    b200_ransac_result* res, int32_t* corr2xK);
/* computeIndicesOfOverlappingPoints(source, target, sourceToTarget, voxelSize, minNumPointsPerVoxel, &idxsSource,
 * &idxsTarget) (open3d_slam/src/helpers.cpp:320-345; called between RANSAC and the refinement, PlaceRecognition.cpp:
 * 98-108, and by the odometry constraints, constraint_builders.cpp:54): both clouds into one VoxelMap (key =
 * floor(p / voxelSize) per axis, VoxelHashMap.hpp:48-51; the source transformed by T16 first), kept = the points whose
 * voxel holds at least min_points of BOTH clouds. Outputs (capacity Ns / Nt) in ascending index order (the reference's
 * order is that of an std::unordered_map walk). */
int b200_lc_overlap_indices(b200_lc* h, const float* src4xN, int Ns, const float* tgt4xM, int Nt, const double* T16,
                            double voxel_size, int min_points, int32_t* src_idx, int* n_src, int32_t* tgt_idx, int* n_tgt);
/* device time of the last call's kernels (CUDA events; uploads and read-backs excluded) */
int b200_lc_last_device_ms(const b200_lc* h, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* B200REG_H_ */
