// Parameter block and result layout of the persistent whole-registration kernel (icp_fused.cu).
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int kIcpThreads = 512;       // threads per CTA of the persistent kernel
constexpr int kHistBins = 2048;        // radix-select bins per pass (11 + 11 + 9 bits of a non-negative float)
constexpr int kAccum = 32;             // doubles per CTA partial (29 used)
constexpr int kMaxHist = 64;           // ring of past poses kept for DifferentialTransformationChecker
constexpr int kMaxLog = 64;            // per-iteration trace capacity
constexpr int kFastVals = 32;          // values reduced per pair on the single-barrier path (27 H/b + 5 counters)
constexpr int kIcpStateWords = 4096;    // capacity (unsigned) of the resumable block state
constexpr int kListCap = 16384;        // distances of the selected first-level bin resolved in shared memory (general shape)
constexpr int kCandCap = 2048;         // bracket candidates the single-barrier path resolves in shared memory
// 64 KiB staging tile + candidate distances and positions
constexpr size_t kIcpDynSmem = (size_t)kFastVals * kIcpThreads * sizeof(float) + (size_t)kCandCap * 8;

struct IcpLogEntry {
  float H[36];
  float b[6];
  float x[6];
  float T_iter[16];
  int P;
  int solve_info;
  int path;        // 1: single-barrier iteration, 0: general iteration, 2: single-barrier attempt fell back
  int n_cand;      // bracket candidates of the attempt
  // %globaltimer (ns) seen by thread 0 of CTA 0: iteration start, own matching done, trimmed limit known,
  // own linearisation done, all partials visible, solve + checks done
  unsigned long long ts[6];
  // debug statistics (only with IcpKernelParams::debug_stats): cell rows resolved, candidates examined, and the
  // slowest / summed per-CTA duration of the matching phase (ns)
  unsigned long long dbg_rows, dbg_cands, dbg_match_ns_max, dbg_match_ns_sum;
  float limit, delta;  // TrimmedDist limit of this iteration, bracket half-width prepared for the next
  unsigned sub_ns[4];  // CTA 0 inside "reduce+solve": reduction of partials, 6x6 solve, pose update, checkers
};

struct IcpDeviceResult {
  int status;
  int iterations;
  int max_iterations_reached;
  int n_matches_last;
  float T_iter[16];
  float T[16];                 // T_refIn_refMean * T_iter * T_refMean_readMean * T_readIn_readMean^-1 (ICP.cpp:1345)
  float weighted_point_used_ratio;
  float point_used_ratio;
  unsigned long long points_touched;
  int iterate;                 // 1: the launch stopped at its iteration budget, the registration goes on (resume)
  // X-ICP (IcpKernelParams::xicp): detection flags of the first iteration (1 = localizable), eigenvectors as they
  // stand after the last iteration, full alignment sums (double), prior returned (ICP.cpp:1335-1341), and whether a
  // decision fell inside the rounding band of the reference's sequential float sums
  int returned_prior, xicp_borderline;
  int loc_xyz[3], loc_rpy[3];
  float eig_t[9], eig_r[9];
  double xicp_sums[12];
};

struct IcpKernelParams {
  GridView g;
  const float4* reading;    // N points already expressed in the refMean frame (ICP.cpp:984)
  const float4* reading_n;  // N normals rotated alike, or null
  int N, k;
  float md2;                // KDTreeMatcher maxDist^2 (inf allowed)
  int use_trimmed; float trimmed_ratio;
  int use_maxdist; float maxdist_lim2;
  int use_sn; float sn_cos;
  int minimizer;            // 0 point-to-plane, 1 point-to-point
  int max_iterations; int use_differential; float min_rot, min_trans; int smooth;
  int use_bound; float bound_rot, bound_trans;   // BoundTransformationChecker (TransformationCheckersImpl.cpp:165-222)
  // work buffers (k x N)
  int* ids; float* dists; int* spos; float* weights;
  unsigned* hist;           // [2][3][kHistBins]
  unsigned* list_count;     // [2]: entries of the per-parity list of the selected bin's distances (cand_d, kListCap each)
  double* partials;         // [2 parities][gridDim.x][kAccum]
  IcpDeviceResult* out;
  IcpLogEntry* log;         // kMaxLog entries
  // single-barrier iteration (see icp_fused.cu): per-CTA segments of TrimmedDist bracket candidates
  float* cand_d; int* cand_q; float4* cand_rec;   // [2 parities][cand_cap] (cand_rec: 2 float4 per candidate)
  int cand_seg;             // entries per CTA segment (set by the launcher)
  size_t cand_cap;          // entries per parity, >= gridDim * cand_seg (set by the launcher)
  int fast_path;            // 0: always run the general 4-barrier iteration
  int coh_log2;             // log2 of the run of consecutive scan points a warp keeps together (0..5), see query_of;
                            // < 0: chosen per iteration by the kernel
  int qpw_log2;             // log2 of the queries a warp holds (2..5); < 0 on input: chosen by the launcher
  int warm_start;           // seed every search after the first iteration with the query's previous match
  unsigned* dbg_warp;       // debug: per warp slot of iteration 0: {rows, candidates, passes, ns} (nullable)
  int debug_stats;          // collect IcpLogEntry::dbg_* (adds global atomics; profiling aid only)
  const float* frames;      // 16-float blocks: T_refIn_refMean, T_refMean_readMean, T_readIn_readMean, (centroid),
                            // T_refMean_dataIn (ICP.cpp:975), its inverse
  // Partitioned map: the matches of an iteration come from outside — every rank searches its own slab
  // for the queries it owns (slab_match_kernel), the ranks sum their contributions, and this kernel runs the rest of
  // the iteration on the complete set. forced = 1: no search; query i is matched to (forced_id[i], forced_dist[i]),
  // and g.pts / g.nrm point at N-entry arrays holding the matched point and normal of query i (so every gather of
  // "the match at sorted position sp" works unchanged with sp = i). k = 1 only.
  int forced;
  const float* forced_dist;
  const int* forced_id;
  // Launch budget and resumable state: at most max_launch_iters iterations per launch (0: no limit); with resume = 1
  // the block state (pose, checker history, TrimmedDist bracket, ...) is restored from `state` instead of initialised,
  // and it is saved there whenever `state` is given. kIcpStateWords unsigned.
  int max_launch_iters, resume;
  unsigned* state;
  // X-ICP, degeneracyAwareness = OptimizedEqualityConstraints (ICP.cpp:629-670); k = 1, point-to-plane only
  int xicp;
  float x_enough, x_insufficient, x_cos_min, x_cos_strong;
  double* partials2;        // [gridDim.x][16]: alignment sums of the detection pass
};

// launches the persistent kernel for capacity K in {1, 4, 16} (K >= p.k); cooperative launch, grid chosen from
// the occupancy calculator. Returns the grid size used through *grid_out.
int launch_icp_kernel(const IcpKernelParams& p, int device, cudaStream_t s, int* grid_out);
// entries of the cand_* arrays the launch may use for N queries (upper bound over all grids)
size_t icp_cand_capacity(int N);
// how many doubles of `partials` the launch needs
size_t icp_partials_capacity(int device);

}  // namespace b200
