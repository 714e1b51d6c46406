// Whole multi-iteration ICP registration as ONE persistent cooperative kernel (no host round trips).
//
// Replaces the hot loop of PointMatcher<float>::ICP::computeWithTransformedReference
// (libpointmatcher/pointmatcher/ICP.cpp:1027-1311) for degeneracyAwareness = None:
//   transformations.apply      (:1047)  -> apply_T on the fly, never materialised
//   matcher->findClosests      (:1051)  -> knn_query over the grid index
//   outlierFilters.compute     (:1055)  -> radix select (TrimmedDist) + per-pair predicates
//   ErrorElements              (:1061)  -> implicit (pairs with valid dist and non-zero weight)
//   calculateOptimizationHessian (:1071) -> 27-term accumulation, warp shuffles + one CTA partial
//   errorMinimizer->compute    (:1205)  -> 6x6 solve + SE(3) in-kernel
//   T_iter = real * T_iter     (:1213), transformationCheckers.check (:1292)
// Every CTA redundantly reduces the CTA partials in a fixed order and solves, so all CTAs hold bit-identical
// state and only grid-wide barriers (no broadcasts) separate the phases.
//
// Two iteration shapes. GENERAL: match | barrier | first histogram level of the TrimmedDist quantile, the selected
// bin's distances to a list | barrier | radix select of the list in shared memory, weights + H/b partials from the
// registers that still hold the match | barrier | reduce + solve: 3 grid barriers (k > 1, more than one query per
// thread or a bin above kListCap: three histogram levels, 4 barriers). SINGLE-BARRIER (k = 1,
// point-to-plane): the quantile moves little between iterations, so the limit is bracketed by
// [lim_prev (1 - delta), lim_prev (1 + delta)]; during matching every pair is classified as certainly inside (d <= lo,
// accumulated at once), certainly outside (d > hi) or a candidate (few hundred; d and the pair's linearisation are
// parked in a per-CTA segment). After ONE barrier every CTA sums the partials and the class counts, checks that the
// exact rank of Matches::getDistsQuantile falls among the candidates, selects it by counting in shared memory, adds
// the accepted candidates in a fixed order and solves. If the bracket missed (or no bracket exists yet) the
// iteration falls through to the general shape, so the result is exact either way.
#include "icp_kernel.cuh"
#include "knn.cuh"
#include "solve.cuh"
#include "xicp.cuh"
#include <cooperative_groups.h>
#include <algorithm>

namespace cg = cooperative_groups;

namespace b200 {
int icp_cand_segment(int nslots, int grid);
size_t icp_cand_capacity(int N);
namespace {

struct BlockState {
  float T_iter[16];
  float Tc[16];
  float R9[9];
  int iterate, status, iteration, maxReached;
  float counter;
  int nrot;
  Quat rot[kMaxHist];
  Quat rot0;
  float tr[kMaxHist][3];
  unsigned sel_bin, sel_rank, sel_total, sel_cnt, list_base;
  int P, solve_info;
  float wsum;
  // TrimmedDist bracket predicted from the previous iteration's limit (single-barrier path)
  int have_bracket, try_fast, path;
  int coh;   // log2 of the run of consecutive scan points per warp in this iteration (query_of)
  float lim_prev, delta, lo, hi, dens, step_prev;
  unsigned limit_bits;
  float H[36], b[6], x[6], x_spd[6];
  // X-ICP: pose that takes the matched points back to the frame the reading came in (ICP.cpp:2236)
  float TcInv[16], R9Inv[9];
  XicpState X;
};

constexpr int kWarps = kIcpThreads / 32;
constexpr float kBracketCands = 300.f;   // candidates the TrimmedDist bracket of the single-barrier shape aims at
static_assert(sizeof(Warp1Scratch) * kWarps <= (size_t)kFastVals * kIcpThreads * sizeof(float),
              "the search scratch must fit into the staging tile it shares");

__device__ __forceinline__ unsigned block_excl_scan(unsigned v, unsigned* s, unsigned& total) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  unsigned inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    const unsigned w = lane < kWarps ? s[lane] : 0u;
    unsigned winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    s[lane] = winc - w;
    if (lane == 31) s[32] = winc;
  }
  __syncthreads();
  const unsigned res = s[wid] + inc - v;
  total = s[32];
  __syncthreads();
  return res;
}

// every CTA finds the bin that holds element `rank` of the global histogram (kHistBins bins). With `first` the rank
// is derived from the histogram total exactly like Matches::getDistsQuantile (Matches.cpp:82-86):
// index = (size_t)(float(n_valid) * ratio), the maximum when ratio == 1.
__device__ __forceinline__ void select_bin(const unsigned* __restrict__ ghist, unsigned rank, bool first, float ratio,
                                           unsigned* s_scan, BlockState& S) {
  constexpr int per = kHistBins / kIcpThreads;
  unsigned c[per], sum = 0;
#pragma unroll
  for (int j = 0; j < per; ++j) {
    c[j] = __ldcg(&ghist[threadIdx.x * per + j]);
    sum += c[j];
  }
  unsigned total;
  unsigned ex = block_excl_scan(sum, s_scan, total);
  if (threadIdx.x == 0) S.sel_total = total;
  if (first && total > 0u) {
    if (ratio == 1.0f) rank = total - 1u;
    else {
      rank = (unsigned)__fmul_rn((float)total, ratio);
      if (rank > total - 1u) rank = total - 1u;
    }
  }
  if (total > 0u && rank >= ex && rank < ex + sum) {
#pragma unroll
    for (int j = 0; j < per; ++j) {
      if (rank >= ex && rank < ex + c[j]) { S.sel_bin = threadIdx.x * per + j; S.sel_rank = rank - ex; S.sel_cnt = c[j]; }
      ex += c[j];
    }
  }
  __syncthreads();
}

// rank-th smallest (0-based) key among vals[0..n), key = bits(val) - base_bits < 2^nbits: radix select in shared memory,
// 9 bits per pass (bins: 512 counters, one per thread). Called by the whole CTA; every thread returns the key.
__device__ __forceinline__ unsigned smem_radix_select(const float* vals, unsigned n, unsigned base_bits, int nbits,
                                                      unsigned rank, unsigned* bins, unsigned* s_scan, BlockState& S) {
  const int tid = threadIdx.x;
  unsigned prefix = 0u, mask = 0u;
  int shift = nbits > 9 ? nbits - 9 : 0;
  while (true) {
    bins[tid] = 0u;
    __syncthreads();
    for (unsigned c = tid; c < n; c += kIcpThreads) {
      const unsigned key = __float_as_uint(vals[c]) - base_bits;
      if ((key & mask) == prefix) atomicAdd(&bins[(key >> shift) & 511u], 1u);
    }
    __syncthreads();
    const unsigned cnt = bins[tid];
    unsigned tot2;
    const unsigned ex = block_excl_scan(cnt, s_scan, tot2);
    if (rank >= ex && rank < ex + cnt) { S.sel_bin = (unsigned)tid; S.sel_rank = rank - ex; }
    __syncthreads();
    prefix |= S.sel_bin << shift;
    mask |= 511u << shift;
    rank = S.sel_rank;
    __syncthreads();
    if (shift == 0) break;
    shift = shift > 9 ? shift - 9 : 0;
  }
  return prefix;
}

__device__ __forceinline__ void flush_hist(unsigned* s_hist, unsigned* __restrict__ ghist) {
  __syncthreads();
  for (int b = threadIdx.x; b < kHistBins; b += kIcpThreads) {
    const unsigned v = s_hist[b];
    if (v) atomicAdd(&ghist[b], v);
  }
}
__device__ __forceinline__ void zero_hist(unsigned* s_hist) {
  for (int b = threadIdx.x; b < kHistBins; b += kIcpThreads) s_hist[b] = 0u;
  __syncthreads();
}

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Staging tile s_v[value][thread] (kFastVals x kIcpThreads floats). Thread (value = tid / 16, segment = tid % 16)
// sums its value over the 32 threads of its segment in double, in a fixed (rotated, bank-conflict-free) order.
__device__ __forceinline__ double role_sum(const float* s_v, int tid) {
  const int vi = tid >> 4, seg = tid & 15;
  const float* base = s_v + vi * kIcpThreads + seg * 32;
  double a = 0.0;
#pragma unroll 8
  for (int j = 0; j < 32; ++j) a += (double)base[(j + seg) & 31];
  return a;
}

// Query -> (warp slot, lane) map. A warp holds 32 >> c groups of G = 1 << c CONSECUTIVE scan points; group g of slot w
// is chunk (w + 97 g) mod nslots of part g of the scan. Consecutive scan points are neighbours in space, so the lanes
// of a group read the same cell rows and the same 128-byte lines of the cell-sorted points (one L1 wavefront instead
// of G); the groups of a warp come from 32 / G distant parts of the scan, so equally expensive queries (no neighbour
// within maxDist, far neighbours) are still spread over all warps. c = 0 is one query per part (no coherence).
// With c > 0 the slots are scattered over the chunks with a multiplicative hash (A coprime to nslots): the 16 warps
// of a CTA hold chunks from all over the scan, so no CTA is handed a whole expensive region (measured without the
// hash: the slowest CTA of the first iteration took 3x the mean).
__device__ __forceinline__ unsigned slot_stride(unsigned nslots) {
  const unsigned cand[4] = {2654435761u % 1000003u, 7919u, 104729u, 1299709u};
  for (int t = 0; t < 4; ++t) {
    unsigned a = cand[t] % nslots, b = nslots;
    if (a == 0u) continue;
    const unsigned a0 = a;
    while (b) { const unsigned r = a % b; a = b; b = r; }
    if (a == 1u) return a0;
  }
  return 1u;
}
// A warp may hold fewer than 32 queries (qpw = 4, 8, 16: small scans spread over all SMs; the idle lanes help in the
// cooperative parts of the search): lanes >= qpw get no query, and there are qpw >> c groups per slot.
__device__ __forceinline__ int query_of(int slot, int lane, int nslots, int c, unsigned stride, int qpw) {
  if (lane >= qpw) return 0x7fffffff;
  const unsigned g = (unsigned)lane >> c, j = (unsigned)lane & ((1u << c) - 1u);
  const unsigned s2 = c ? (unsigned)(((unsigned long long)(unsigned)slot * stride) % (unsigned)nslots) : (unsigned)slot;
  return (int)(((g * (unsigned)nslots + (s2 + g * 97u) % (unsigned)nslots) << c) + j);
}

template <int K>
__global__ void __launch_bounds__(kIcpThreads, 2) icp_persistent_kernel(const __grid_constant__ IcpKernelParams p) {
  cg::grid_group grid = cg::this_grid();
  __shared__ BlockState S;
  __shared__ unsigned s_hist[kHistBins];
  __shared__ double s_red[kWarps][kAccum];
  __shared__ unsigned s_scan[40];
  __shared__ double s_tot[kAccum];   // grid totals of the 27 H/b terms, pair count, weight sum (+ class counts)
  extern __shared__ __align__(16) float s_dyn[];   // kFastVals x kIcpThreads staging tile / candidate staging
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int nthreads = gridDim.x * kIcpThreads;
  const int gtid = blockIdx.x * kIcpThreads + tid;
  const int k = p.k;

  static_assert(sizeof(BlockState) <= (size_t)kIcpStateWords * sizeof(unsigned), "resumable state does not fit");
  if (p.resume) {
    for (int t = tid; t < (int)(sizeof(BlockState) / sizeof(unsigned)); t += kIcpThreads)
      reinterpret_cast<unsigned*>(&S)[t] = __ldcg(&p.state[t]);
  } else if (tid == 0) {
    eye4(S.T_iter);
    S.iterate = 1; S.status = B200_OK; S.iteration = 0; S.maxReached = 0;
    S.counter = 0.f;
    S.nrot = 1;
    S.rot[0] = quat_from_T(S.T_iter);   // transformationCheckers.init, ICP.cpp:997
    S.rot0 = S.rot[0];                  // BoundTransformationChecker::init (initialTranslation = 0)
    S.tr[0][0] = S.tr[0][1] = S.tr[0][2] = 0.f;
    S.P = 0; S.wsum = 0.f; S.solve_info = 0;
    S.have_bracket = 0; S.try_fast = 0; S.path = 0; S.lim_prev = 0.f; S.delta = 0.25f; S.lo = 0.f; S.hi = 0.f; S.dens = 0.f; S.step_prev = 0.f;
    S.limit_bits = 0u;
    xicp_state_init(S.X);
    if (K == 1 && p.xicp) {
      rigid_prepare(p.frames + 80, S.TcInv, S.R9Inv);
      for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i) S.X.R[i + 3 * j] = p.frames[64 + i + 4 * j];
      if (!inv3_dev(S.X.R, S.X.Rinv)) S.X.return_prior = 1;
    }
  }
  __syncthreads();
  const bool xicp = (K == 1) && p.xicp && p.minimizer == 0;
  const int qpw = 1 << p.qpw_log2;   // queries per warp (launcher: 32 unless the scan is too small to fill the SMs)
  const unsigned qstride = slot_stride((unsigned)((p.N + qpw - 1) >> p.qpw_log2));
  unsigned long long visited_total = 0ull;

  const bool stamp = (blockIdx.x == 0 && tid == 0 && p.log != nullptr);
  unsigned long long ts[6] = {0ull, 0ull, 0ull, 0ull, 0ull, 0ull};
  int launch_iters = 0;
  while (true) {
    if (stamp) ts[0] = gtimer();
    if (tid == 0) {
      rigid_prepare(S.T_iter, S.Tc, S.R9);
      // p.coh_log2 < 0: automatic. The previous TrimmedDist limit (a high quantile of the match distances) tells how far
      // the queries still are from their neighbours, in cells; without it, from the third iteration on.
      if (p.coh_log2 >= 0) S.coh = p.coh_log2;
      else if (S.iteration == 0) S.coh = 0;
      else if (p.use_trimmed) S.coh = (S.have_bracket && sqrtf(S.lim_prev) * p.g.inv_cell < 0.4f) ? 5 : 0;
      else S.coh = S.iteration >= 2 ? 5 : 0;
      if (S.coh > p.qpw_log2) S.coh = p.qpw_log2;
    }
    zero_hist(s_hist);  // includes __syncthreads
    const int par = S.iteration & 1;
    unsigned* gh = p.hist + (size_t)par * 3 * kHistBins;
    // Everything one CTA writes and ANOTHER CTA reads within an iteration is double-buffered by the iteration's parity:
    // a single-barrier iteration has no closing barrier, so a fast CTA may already be writing iteration n + 1 while a
    // slow one still reads iteration n; it cannot reach iteration n + 2 before every CTA has passed barrier n + 1.
    double* const partials = p.partials + (size_t)par * gridDim.x * kAccum;
    float* const cand_d = p.cand_d + (size_t)par * p.cand_cap;
    int* const cand_q = p.cand_q + (size_t)par * p.cand_cap;
    float4* const cand_rec = p.cand_rec + (size_t)par * 2 * p.cand_cap;
    // zero the other parity's histograms for the next iteration: their last readers finished before the barrier
    // that ended the previous iteration, their next writers start after this iteration's barrier
    if (p.use_trimmed) {
      unsigned* gz = p.hist + (size_t)(par ^ 1) * 3 * kHistBins;
      for (int b = gtid; b < 3 * kHistBins; b += nthreads) gz[b] = 0u;
      if (gtid == 0) p.list_count[par ^ 1] = 0u;
    }

    // ---------------- phase 1: transform + match ----------------
    unsigned dbg_rows = 0, dbg_cands = 0;
    const unsigned long long dbg_t0 = p.debug_stats ? gtimer() : 0ull;
    // Query -> thread map: lane l of warp slot w takes query l * nslots + w. Consecutive scan points (spatially
    // coherent, so equally cheap or equally expensive) are spread over different warps; a warp holding 32 expensive
    // queries was the tail every other warp waited for at the grid barrier. Warp-uniform trip count: the k = 1 search
    // is a warp collective.
    const int nslots = (p.N + qpw - 1) >> p.qpw_log2;
    const int rounds = (nslots + (nthreads >> 5) - 1) / (nthreads >> 5);  // grid-uniform trip count
    // Coherent warps (32 consecutive scan points: their cell rows and candidates share 128-byte lines, 1 L1 wavefront
    // instead of up to 32) once nearly every query is settled by its own 2x2x2 block, i.e. when all queries cost the
    // same; scattered lanes while some queries still need the cooperative passes (a coherent warp full of expensive
    // queries is a tail the whole grid waits for: measured 3x on the first iteration).
    const int coh = S.coh;
    // single-barrier shape for this iteration? (needs the previous limit when TrimmedDist is in the chain)
    const bool fast = (K == 1) && p.fast_path && p.minimizer == 0 && (!p.use_trimmed || (S.have_bracket && S.try_fast)) &&
                      !(xicp && S.iteration == 0);
    const float br_lo = p.use_trimmed ? S.lo : INFINITY, br_hi = p.use_trimmed ? S.hi : INFINITY;
    double role_acc = 0.0;    // thread (value v = tid / 16, segment = tid % 16) sums value v over 32 threads
    unsigned cand_base = 0u;  // candidates this CTA has parked so far
    // General shape, one query per thread (rounds == 1), point-to-plane: the match, the transformed query and the matched
    // point and normal stay in registers across the quantile barriers, so that the linearisation phase neither re-reads
    // its inputs nor waits for the gather (the loads are issued here, before the barriers).
    const bool carry = (K == 1) && rounds == 1 && !fast && p.minimizer == 0;
    bool st_in = false;
    int st_i = 0, st_sp = -1;
    float st_d = INFINITY;
    float4 st_q = make_float4(0.f, 0.f, 0.f, 1.f), st_r4 = make_float4(0.f, 0.f, 0.f, 0.f), st_n4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int round = 0; round < rounds; ++round) {
      const int slot = (gtid >> 5) + round * (nthreads >> 5);
      const int i = slot < nslots ? query_of(slot, lane, nslots, coh, qstride, qpw) : p.N;
      const bool in = i < p.N;
      const float4 q = in ? apply_T(S.Tc, __ldg(&p.reading[i])) : make_float4(0.f, 0.f, 0.f, 1.f);
      unsigned vis = 0;
      if (K == 1) {
        float bd; int bid, bsp;
        unsigned w_rows = 0, w_pass = 0;
        const unsigned long long w_t0 = p.debug_stats ? gtimer() : 0ull;
        // the search's per-warp scratch shares the staging tile: the tile is written only after the barrier below
        // warm start from the previous iteration's match of the same query (its own thread wrote it or, after a change
        // of the query map, another CTA did before the barrier that closed that iteration)
        const int hint = (in && S.iteration > 0 && p.warm_start) ? __ldcg(&p.spos[i]) : -1;
        if (p.forced) {   // the match was found elsewhere (partitioned map); its point and normal sit at position i
          bd = INFINITY; bid = -1; bsp = -1;
          if (in) {
            bd = __ldcg(&p.forced_dist[i]); bid = __ldcg(&p.forced_id[i]);
            if (bid >= 0) bsp = i; else bd = INFINITY;
          }
        } else
        knn_query_warp1(p.g, in, q.x, q.y, q.z, p.md2, bd, bid, bsp, vis, reinterpret_cast<Warp1Scratch*>(s_dyn)[wid],
                        p.debug_stats ? &w_rows : nullptr, p.debug_stats ? &w_pass : nullptr, qpw < 32, hint);
        if (p.debug_stats) {
          dbg_rows += w_rows;
          if (p.dbg_warp && S.iteration == 0) {
            const unsigned long long w_dt = gtimer() - w_t0;
            unsigned r = w_rows, c = vis;
            for (int o = 16; o > 0; o >>= 1) { r += __shfl_xor_sync(0xffffffffu, r, o); c += __shfl_xor_sync(0xffffffffu, c, o); }
            if (lane == 0) {
              p.dbg_warp[4 * slot + 0] = r; p.dbg_warp[4 * slot + 1] = c; p.dbg_warp[4 * slot + 2] = w_pass;
              p.dbg_warp[4 * slot + 3] = (unsigned)w_dt;
            }
          }
        }
        if (in) {
          p.ids[i] = bid; p.dists[i] = bd; p.spos[i] = bsp;
          if (p.use_trimmed && bd != INFINITY) atomicAdd(&s_hist[(__float_as_uint(bd) >> 20) & 0x7ffu], 1u);
        }
        if (carry) {
          st_in = in; st_i = i; st_sp = bsp; st_d = bd; st_q = q;
          if (in && bsp >= 0) {
            st_r4 = __ldg(&p.g.pts[bsp]);
            if (p.g.nrm) st_n4 = __ldg(&p.g.nrm[bsp]);
          }
        }
        if (fast) {
          // weight factors that do not depend on the quantile, pair linearisation, class w.r.t. the bracket
          float F[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, dot = 0.f, w_other = 0.f, w = 0.f;
          unsigned is_cand = 0u;
          float c_valid = 0.f, c_below = 0.f;
          if (in) {
            const bool valid = bd != INFINITY;
            const bool any = p.use_trimmed || p.use_maxdist || p.use_sn;
            w_other = any ? 1.f : (valid ? 1.f : 0.f);           // OutlierFilter.cpp:66-99
            if (p.use_maxdist) w_other *= (bd <= p.maxdist_lim2) ? 1.f : 0.f;
            float4 r4 = make_float4(0.f, 0.f, 0.f, 0.f), n4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (bsp >= 0) {
              r4 = __ldg(&p.g.pts[bsp]);
              if (p.g.nrm) n4 = __ldg(&p.g.nrm[bsp]);
            }
            if (p.use_sn && p.reading_n && p.g.nrm) {
              if (bsp < 0) w_other = 0.f;
              else {
                const float4 rn = __ldg(&p.reading_n[i]);
                const float3 rr = apply_R(S.R9, make_float3(rn.x, rn.y, rn.z));
                float a3[3] = {rr.x, rr.y, rr.z}, qn[3], nv[3] = {n4.x, n4.y, n4.z}, no[3];
                normalized3_dev(a3, qn);
                normalized3_dev(nv, no);
                float dotn = __fmul_rn(qn[0], no[0]);
                dotn = __fadd_rn(dotn, __fmul_rn(qn[1], no[1]));
                dotn = __fadd_rn(dotn, __fmul_rn(qn[2], no[2]));
                w_other *= (dotn < p.sn_cos) ? 0.f : 1.f;
              }
            }
            // crossProduct (ErrorMinimizer.cpp:376-378) and calculateOptimizationHessian (ICP.cpp:1527-1565)
            F[0] = __fsub_rn(__fmul_rn(q.y, n4.z), __fmul_rn(q.z, n4.y));
            F[1] = __fsub_rn(__fmul_rn(q.z, n4.x), __fmul_rn(q.x, n4.z));
            F[2] = __fsub_rn(__fmul_rn(q.x, n4.y), __fmul_rn(q.y, n4.x));
            F[3] = n4.x; F[4] = n4.y; F[5] = n4.z;
            dot = __fmul_rn(__fsub_rn(q.x, r4.x), n4.x);
            dot = __fadd_rn(dot, __fmul_rn(__fsub_rn(q.y, r4.y), n4.y));
            dot = __fadd_rn(dot, __fmul_rn(__fsub_rn(q.z, r4.z), n4.z));
            float trim = 1.f;
            if (p.use_trimmed) {
              if (!valid) trim = 0.f;
              else {
                c_valid = 1.f;
                if (bd <= br_lo) c_below = 1.f;
                else if (bd <= br_hi) { is_cand = 1u; trim = 0.f; }   // decided after the barrier
                else trim = 0.f;
              }
            }
            w = w_other * trim;
            if (!is_cand) p.weights[i] = w;
            if (!valid) w = 0.f;   // ErrorElements keeps pairs with a valid dist and a non-zero weight
          }
          // park the candidates of this round in the CTA's segment, in thread order (deterministic)
          if (!p.use_trimmed) __syncthreads();   // (the scan below has its own barriers) every warp has left the search
          if (p.use_trimmed) {
            unsigned tot;
            const unsigned pos = block_excl_scan(is_cand, s_scan, tot);
            if (is_cand) {
              const size_t ci = (size_t)blockIdx.x * p.cand_seg + cand_base + pos;
              cand_d[ci] = bd; cand_q[ci] = i;
              cand_rec[2 * ci] = make_float4(F[0], F[1], F[2], F[3]);
              cand_rec[2 * ci + 1] = make_float4(F[4], F[5], dot, w_other);
            }
            cand_base += tot;
          }
          // stage the 32 values of this thread, then sum them per role in double (fixed order)
          {
            float* col = s_dyn + tid;
            const bool kept = w != 0.f;
            int t = 0;
#pragma unroll
            for (int c = 0; c < 6; ++c)
#pragma unroll
              for (int a = 0; a <= c; ++a) { col[t * kIcpThreads] = kept ? __fmul_rn(__fmul_rn(w, F[a]), F[c]) : 0.f; ++t; }
#pragma unroll
            for (int a = 0; a < 6; ++a) col[(21 + a) * kIcpThreads] = kept ? __fmul_rn(__fmul_rn(w, F[a]), dot) : 0.f;
            col[27 * kIcpThreads] = kept ? 1.f : 0.f;
            col[28 * kIcpThreads] = kept ? w : 0.f;
            col[29 * kIcpThreads] = is_cand ? 1.f : 0.f;
            col[30 * kIcpThreads] = c_valid;
            col[31 * kIcpThreads] = c_below;
          }
          __syncthreads();
          role_acc += role_sum(s_dyn, tid);
          __syncthreads();
        }
      } else if (in) {
        KList<K> L;
        knn_query<K>(p.g, q.x, q.y, q.z, p.md2, k, L, vis);
#pragma unroll
        for (int j = 0; j < K; ++j)
          if (j < k) {
            const size_t e = (size_t)i * k + j;
            p.ids[e] = L.id[j]; p.dists[e] = L.d[j]; p.spos[e] = L.sp[j];
            if (p.use_trimmed && L.d[j] != INFINITY) atomicAdd(&s_hist[(__float_as_uint(L.d[j]) >> 20) & 0x7ffu], 1u);
          }
      }
      visited_total += vis;
      dbg_cands += vis;
    }
    if (p.debug_stats && p.log && S.iteration < kMaxLog) {
      IcpLogEntry& e = p.log[S.iteration];
      for (int o = 16; o > 0; o >>= 1) {
        dbg_rows += __shfl_xor_sync(0xffffffffu, dbg_rows, o);
        dbg_cands += __shfl_xor_sync(0xffffffffu, dbg_cands, o);
      }
      if (lane == 0) { atomicAdd(&e.dbg_rows, (unsigned long long)dbg_rows); atomicAdd(&e.dbg_cands, (unsigned long long)dbg_cands); }
      __syncthreads();
      if (tid == 0) {
        const unsigned long long dt = gtimer() - dbg_t0;
        atomicMax(&e.dbg_match_ns_max, dt);
        atomicAdd(&e.dbg_match_ns_sum, dt);
      }
    }

    float limit = INFINITY;
    bool have_totals = false;
    int n_cand_seen = 0;
    // ---------------- single-barrier shape: partials + class counts -> barrier -> resolve the bracket ----------------
    if (fast) {
      double* red = &s_red[0][0];  // kFastVals x 16 segment sums, index = value * 16 + segment = tid
      red[tid] = role_acc;
      __syncthreads();
      if (tid < kFastVals) {
        double t = 0.0;
        for (int sg = 0; sg < 16; ++sg) t += red[tid * 16 + sg];
        __stcg(&partials[(size_t)blockIdx.x * kAccum + tid], t);
      }
      if (stamp) ts[1] = gtimer();
      grid.sync();
      if (stamp) ts[2] = gtimer();
      {  // fixed-order sum of the CTA partials (same order in every CTA)
        const int c = tid & (kAccum - 1), slice = tid / kAccum;
        constexpr int kSlices = kIcpThreads / kAccum;
        double sacc = 0.0;
        for (unsigned b = slice; b < gridDim.x; b += kSlices) sacc += __ldcg(&partials[(size_t)b * kAccum + c]);
        red[slice * kAccum + c] = sacc;
        __syncthreads();
        if (tid < kAccum) {
          double t = 0.0;
          for (int sl = 0; sl < kSlices; ++sl) t += red[sl * kAccum + tid];
          s_tot[tid] = t;
        }
        __syncthreads();
      }
      bool ok = true;
      unsigned r_in = 0u;
      const unsigned n_cand = (unsigned)(s_tot[29] + 0.5), n_valid = (unsigned)(s_tot[30] + 0.5),
                     n_below = (unsigned)(s_tot[31] + 0.5);
      n_cand_seen = (int)n_cand;
      if (p.use_trimmed) {
        if (n_valid == 0u) {  // ConvergenceError("No matches available..."), Matches.cpp:75-76
          if (tid == 0) { S.status = B200_E_NO_MATCHES; S.iterate = 0; }
          __syncthreads();
          break;
        }
        // Matches::getDistsQuantile, Matches.cpp:82-86: index = (size_t)(float(n_valid) * ratio), max if ratio == 1
        unsigned rank = n_valid - 1u;
        if (p.trimmed_ratio != 1.0f) {
          rank = (unsigned)__fmul_rn((float)n_valid, p.trimmed_ratio);
          if (rank > n_valid - 1u) rank = n_valid - 1u;
        }
        ok = rank >= n_below && (rank - n_below) < n_cand && n_cand <= (unsigned)kCandCap;
        r_in = rank - n_below;
      }
      if (ok) {
        if (p.use_trimmed) {
          float* s_cd = s_dyn + kFastVals * kIcpThreads;              // candidate distances (kCandCap)
          int* s_ci = reinterpret_cast<int*>(s_cd + kCandCap);         // their positions in the global segments
          // compact numbering of the parked candidates: offsets of the CTA segments by a block scan, then every thread
          // resolves its candidates' segment by bisection (all loads of a round are independent)
          unsigned* s_off = reinterpret_cast<unsigned*>(s_hist);   // the level-0 histogram is not needed any more
          const unsigned nb = tid < (int)gridDim.x ? (unsigned)(__ldcg(&partials[(size_t)tid * kAccum + 29]) + 0.5) : 0u;
          unsigned tot;
          const unsigned off = block_excl_scan(nb, s_scan, tot);
          s_off[tid] = off;
          if (tid == 0) s_off[kIcpThreads] = tot;
          __syncthreads();
          for (unsigned c = tid; c < n_cand; c += kIcpThreads) {
            int lo_b = 0, hi_b = (int)gridDim.x - 1;   // last segment whose offset is <= c
            while (lo_b < hi_b) {
              const int mid = (lo_b + hi_b + 1) >> 1;
              if (s_off[mid] <= c) lo_b = mid; else hi_b = mid - 1;
            }
            const int ci = lo_b * p.cand_seg + (int)(c - s_off[lo_b]);
            s_cd[c] = __ldcg(&cand_d[ci]);
            s_ci[c] = ci;
          }
          __syncthreads();
          // the r_in-th smallest candidate: radix select, 9 bits per pass, on key = bits(d) - bits(lo) (monotone for
          // positive floats; the bracket is narrow, so 2-3 passes)
          {
            const unsigned base_bits = __float_as_uint(br_lo);
            const unsigned range = __float_as_uint(br_hi) - base_bits;
            const int nbits = 32 - __clz((int)(range | 1u));
            const unsigned key = smem_radix_select(s_cd, n_cand, base_bits, nbits, r_in, reinterpret_cast<unsigned*>(red), s_scan, S);
            if (tid == 0) S.limit_bits = base_bits + key;
            __syncthreads();
          }
          limit = __uint_as_float(S.limit_bits);
          // add the accepted candidates (fixed order: candidate c is staged by thread c mod 512)
          double acc2 = 0.0;
          for (unsigned c0 = 0; c0 < n_cand; c0 += kIcpThreads) {
            const unsigned c = c0 + tid;
            float F[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, dot = 0.f, w = 0.f;
            if (c < n_cand) {
              const int ci = s_ci[c];
              const float4 ra = __ldcg(&cand_rec[2 * (size_t)ci]), rb = __ldcg(&cand_rec[2 * (size_t)ci + 1]);
              F[0] = ra.x; F[1] = ra.y; F[2] = ra.z; F[3] = ra.w; F[4] = rb.x; F[5] = rb.y; dot = rb.z;
              w = (s_cd[c] <= limit) ? rb.w : 0.f;
            }
            float* col = s_dyn + tid;
            const bool kept = w != 0.f;
            int t = 0;
#pragma unroll
            for (int cc = 0; cc < 6; ++cc)
#pragma unroll
              for (int a = 0; a <= cc; ++a) { col[t * kIcpThreads] = kept ? __fmul_rn(__fmul_rn(w, F[a]), F[cc]) : 0.f; ++t; }
#pragma unroll
            for (int a = 0; a < 6; ++a) col[(21 + a) * kIcpThreads] = kept ? __fmul_rn(__fmul_rn(w, F[a]), dot) : 0.f;
            col[27 * kIcpThreads] = kept ? 1.f : 0.f;
            col[28 * kIcpThreads] = kept ? w : 0.f;
            col[29 * kIcpThreads] = 0.f; col[30 * kIcpThreads] = 0.f; col[31 * kIcpThreads] = 0.f;
            __syncthreads();
            acc2 += role_sum(s_dyn, tid);
            __syncthreads();
          }
          // the weights of the parked pairs are written by the CTA that owns the queries (the next iteration's weight
          // of a query is written by the same CTA, so program order within the CTA is all that is needed)
          for (unsigned c = tid; c < cand_base; c += kIcpThreads) {
            const size_t ci = (size_t)blockIdx.x * p.cand_seg + c;
            p.weights[cand_q[ci]] = (cand_d[ci] <= limit) ? cand_rec[2 * ci + 1].w : 0.f;
          }
          red[tid] = acc2;
          __syncthreads();
          if (tid < 29) {
            double t = 0.0;
            for (int sg = 0; sg < 16; ++sg) t += red[tid * 16 + sg];
            s_tot[tid] += t;
          }
          __syncthreads();
        }
        have_totals = true;
        if (stamp) { ts[3] = gtimer(); ts[4] = ts[3]; }
      }
      if (tid == 0) S.path = ok ? 1 : 2;
    } else if (tid == 0) S.path = 0;

    if (!have_totals) {
    // ---------------- phase 2: TrimmedDist quantile by 3-pass radix select ----------------
    if (stamp && !fast) ts[1] = gtimer();
    if (p.use_trimmed) {
      flush_hist(s_hist, gh);
      grid.sync();
      // Matches::getDistsQuantile, Matches.cpp:60-87: index = (size_t)(float(n_valid) * ratio), max if ratio == 1
      select_bin(gh, 0u, true, p.trimmed_ratio, s_scan, S);
      if (S.sel_total == 0u) {  // ConvergenceError("No matches available..."), Matches.cpp:75-76
        __syncthreads();
        if (tid == 0) { S.status = B200_E_NO_MATCHES; S.iterate = 0; }
        __syncthreads();
        break;
      }
      const unsigned bin0 = S.sel_bin, rank0 = S.sel_rank, cnt0 = S.sel_cnt;
      if (carry && cnt0 <= (unsigned)kListCap) {
        // Second and third level in one step: the few distances of the selected bin (every thread still holds its own
        // in a register) are appended to a list, and after ONE more barrier every CTA picks element rank0 of the list by
        // a radix select in shared memory — instead of two more histogram passes with a barrier each.
        const unsigned mine = (st_in && st_d != INFINITY && ((__float_as_uint(st_d) >> 20) & 0x7ffu) == bin0) ? 1u : 0u;
        unsigned tot;
        const unsigned pos = block_excl_scan(mine, s_scan, tot);
        if (tid == 0 && tot) S.list_base = atomicAdd(&p.list_count[par], tot);
        __syncthreads();
        float* list = cand_d;   // this parity's candidate segment is free in a general iteration
        if (mine) __stcg(&list[S.list_base + pos], st_d);
        grid.sync();
        float* s_list = s_dyn;
        for (unsigned c = tid; c < cnt0; c += kIcpThreads) s_list[c] = __ldcg(&list[c]);
        __syncthreads();
        const unsigned base_bits = bin0 << 20;
        const unsigned key = smem_radix_select(s_list, cnt0, base_bits, 20, rank0, reinterpret_cast<unsigned*>(&s_red[0][0]), s_scan, S);
        limit = __uint_as_float(base_bits + key);
      } else {
      zero_hist(s_hist);
      for (int i = gtid; i < p.N; i += nthreads)
        for (int j = 0; j < k; ++j) {
          const float d = p.dists[(size_t)i * k + j];
          const unsigned bits = __float_as_uint(d);
          if (d != INFINITY && ((bits >> 20) & 0x7ffu) == bin0) atomicAdd(&s_hist[(bits >> 9) & 0x7ffu], 1u);
        }
      flush_hist(s_hist, gh + kHistBins);
      grid.sync();
      select_bin(gh + kHistBins, rank0, false, 0.f, s_scan, S);
      const unsigned bin1 = S.sel_bin, rank1 = S.sel_rank;
      const unsigned pre = (bin0 << 11) | bin1;
      zero_hist(s_hist);
      for (int i = gtid; i < p.N; i += nthreads)
        for (int j = 0; j < k; ++j) {
          const float d = p.dists[(size_t)i * k + j];
          const unsigned bits = __float_as_uint(d);
          if (d != INFINITY && ((bits >> 9) & 0x3fffffu) == pre) atomicAdd(&s_hist[bits & 0x1ffu], 1u);
        }
      flush_hist(s_hist, gh + 2 * kHistBins);
      grid.sync();
      select_bin(gh + 2 * kHistBins, rank1, false, 0.f, s_scan, S);
      limit = __uint_as_float((pre << 9) | S.sel_bin);
      }
    }
    if (stamp) ts[2] = gtimer();

    // ---------------- phase 3: weights + linearisation, reduced to one partial per CTA ----------------
    if (carry) {
      float F[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, dot = 0.f, w = 0.f;
      bool kept = false;
      if (st_in) {
        const float d = st_d;
        const float4 q = st_q, r4 = st_r4, n4 = st_n4;
        // outlier filter chain = product of the filters' weights (OutlierFilter.cpp:84-99); empty chain: (dist != inf)
        const bool any = p.use_trimmed || p.use_maxdist || p.use_sn;
        w = any ? 1.f : (d == INFINITY ? 0.f : 1.f);
        if (p.use_trimmed) w *= (d <= limit) ? 1.f : 0.f;
        if (p.use_maxdist) w *= (d <= p.maxdist_lim2) ? 1.f : 0.f;
        if (p.use_sn && p.reading_n && p.g.nrm) {
          if (st_sp < 0) w = 0.f;
          else {
            const float4 rn = __ldg(&p.reading_n[st_i]);
            const float3 rr = apply_R(S.R9, make_float3(rn.x, rn.y, rn.z));
            float a3[3] = {rr.x, rr.y, rr.z}, qn[3], nv[3] = {n4.x, n4.y, n4.z}, no[3];
            normalized3_dev(a3, qn);
            normalized3_dev(nv, no);
            float dotn = __fmul_rn(qn[0], no[0]);
            dotn = __fadd_rn(dotn, __fmul_rn(qn[1], no[1]));
            dotn = __fadd_rn(dotn, __fmul_rn(qn[2], no[2]));
            w *= (dotn < p.sn_cos) ? 0.f : 1.f;
          }
        }
        p.weights[st_i] = w;
        kept = (d != INFINITY) && (w != 0.f);  // ErrorElements, ErrorMinimizer.cpp:103-107
        // crossProduct (ErrorMinimizer.cpp:376-378) and calculateOptimizationHessian (ICP.cpp:1527-1565)
        F[0] = __fsub_rn(__fmul_rn(q.y, n4.z), __fmul_rn(q.z, n4.y));
        F[1] = __fsub_rn(__fmul_rn(q.z, n4.x), __fmul_rn(q.x, n4.z));
        F[2] = __fsub_rn(__fmul_rn(q.x, n4.y), __fmul_rn(q.y, n4.x));
        F[3] = n4.x; F[4] = n4.y; F[5] = n4.z;
        dot = __fmul_rn(__fsub_rn(q.x, r4.x), n4.x);
        dot = __fadd_rn(dot, __fmul_rn(__fsub_rn(q.y, r4.y), n4.y));
        dot = __fadd_rn(dot, __fmul_rn(__fsub_rn(q.z, r4.z), n4.z));
      }
      float4 pb = make_float4(0.f, 0.f, 0.f, 0.f);
      if (xicp && S.iteration == 0 && kept) pb = apply_T(S.TcInv, st_q);   // calculatePointCloudCenter, ICP.cpp:1594-1618
      // The staging tile doubles as the search's per-warp scratch: without TrimmedDist no barrier lies between the
      // search and this point, and a warp that is done must not overwrite the scratch of one still searching.
      if (!p.use_trimmed) __syncthreads();
      {
        float* col = s_dyn + tid;
        int t = 0;
#pragma unroll
        for (int c = 0; c < 6; ++c)
#pragma unroll
          for (int a = 0; a <= c; ++a) { col[t * kIcpThreads] = kept ? __fmul_rn(__fmul_rn(w, F[a]), F[c]) : 0.f; ++t; }
#pragma unroll
        for (int a = 0; a < 6; ++a) col[(21 + a) * kIcpThreads] = kept ? __fmul_rn(__fmul_rn(w, F[a]), dot) : 0.f;
        col[27 * kIcpThreads] = kept ? 1.f : 0.f;
        col[28 * kIcpThreads] = kept ? w : 0.f;
        col[29 * kIcpThreads] = pb.x; col[30 * kIcpThreads] = pb.y; col[31 * kIcpThreads] = pb.z;
      }
      __syncthreads();
      const double role = role_sum(s_dyn, tid);
      double* red = &s_red[0][0];   // kFastVals x 16 segment sums, index = value * 16 + segment = tid
      red[tid] = role;
      __syncthreads();
      if (tid < kAccum) {
        double t = 0.0;
        for (int sg = 0; sg < 16; ++sg) t += red[tid * 16 + sg];
        __stcg(&partials[(size_t)blockIdx.x * kAccum + tid], t);
      }
    } else {
    for (int t = tid; t < kWarps * kAccum; t += kIcpThreads) (&s_red[0][0])[t] = 0.0;
    __syncthreads();
    for (int slot = gtid >> 5; slot < nslots; slot += nthreads >> 5) {  // same map as phase 1 (no barrier needed in between)
      const int i = query_of(slot, lane, nslots, coh, qstride, qpw);
      const bool in = i < p.N;
      float4 q = make_float4(0.f, 0.f, 0.f, 1.f);
      float3 nq = make_float3(0.f, 0.f, 0.f);
      if (in) {
        q = apply_T(S.Tc, __ldg(&p.reading[i]));
        if (p.use_sn && p.reading_n) {
          const float4 n4 = __ldg(&p.reading_n[i]);
          const float3 r = apply_R(S.R9, make_float3(n4.x, n4.y, n4.z));
          float v[3] = {r.x, r.y, r.z}, o[3];
          normalized3_dev(v, o);
          nq = make_float3(o[0], o[1], o[2]);
        }
      }
      for (int j = 0; j < k; ++j) {
        float v[29];
#pragma unroll
        for (int t = 0; t < 29; ++t) v[t] = 0.f;
        double mom[16];
        bool kept = false;
        float w = 0.f;
        float4 r4 = make_float4(0.f, 0.f, 0.f, 0.f), n4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (in) {
          const size_t e = (size_t)i * k + j;
          const float d = p.dists[e];
          const int sp = p.spos[e];
          // outlier filter chain = product of the filters' weights (OutlierFilter.cpp:84-99); empty chain:
          // weight = (dist != inf) (:66-83)
          const bool any = p.use_trimmed || p.use_maxdist || p.use_sn;
          w = any ? 1.f : (d == INFINITY ? 0.f : 1.f);
          if (p.use_trimmed) w *= (d <= limit) ? 1.f : 0.f;
          if (p.use_maxdist) w *= (d <= p.maxdist_lim2) ? 1.f : 0.f;
          if (sp >= 0) {
            r4 = __ldg(&p.g.pts[sp]);
            if (p.g.nrm) n4 = __ldg(&p.g.nrm[sp]);
          }
          if (p.use_sn) {
            if (p.reading_n && p.g.nrm) {
              if (sp < 0) w = 0.f;
              else {
                float nv[3] = {n4.x, n4.y, n4.z}, no[3];
                normalized3_dev(nv, no);
                float dotn = __fmul_rn(nq.x, no[0]);
                dotn = __fadd_rn(dotn, __fmul_rn(nq.y, no[1]));
                dotn = __fadd_rn(dotn, __fmul_rn(nq.z, no[2]));
                w *= (dotn < p.sn_cos) ? 0.f : 1.f;
              }
            }
          }
          p.weights[e] = w;
          kept = (d != INFINITY) && (w != 0.f);  // ErrorElements, ErrorMinimizer.cpp:103-107
        }
        if (kept) {
          if (p.minimizer == 0) {
            // crossProduct (ErrorMinimizer.cpp:376-378) and calculateOptimizationHessian (ICP.cpp:1527-1565)
            float F[6], wF[6];
            F[0] = __fsub_rn(__fmul_rn(q.y, n4.z), __fmul_rn(q.z, n4.y));
            F[1] = __fsub_rn(__fmul_rn(q.z, n4.x), __fmul_rn(q.x, n4.z));
            F[2] = __fsub_rn(__fmul_rn(q.x, n4.y), __fmul_rn(q.y, n4.x));
            F[3] = n4.x; F[4] = n4.y; F[5] = n4.z;
#pragma unroll
            for (int a = 0; a < 6; ++a) wF[a] = __fmul_rn(w, F[a]);
            float dot = __fmul_rn(__fsub_rn(q.x, r4.x), n4.x);
            dot = __fadd_rn(dot, __fmul_rn(__fsub_rn(q.y, r4.y), n4.y));
            dot = __fadd_rn(dot, __fmul_rn(__fsub_rn(q.z, r4.z), n4.z));
            int t = 0;
#pragma unroll
            for (int c = 0; c < 6; ++c)
#pragma unroll
              for (int a = 0; a <= c; ++a) v[t++] = __fmul_rn(wF[a], F[c]);
#pragma unroll
            for (int a = 0; a < 6; ++a) v[21 + a] = __fmul_rn(wF[a], dot);
          }
          v[27] = 1.f;
          v[28] = w;
        }
        if (p.minimizer == 0) {
#pragma unroll
          for (int t = 0; t < 29; ++t) {
            const double s = warp_sum((double)v[t]);
            if (lane == 0) s_red[wid][t] += s;
          }
        } else {
#pragma unroll
          for (int t = 0; t < 16; ++t) mom[t] = 0.0;
          if (kept) {
            const double wd = (double)w;
            const double rr[3] = {(double)q.x, (double)q.y, (double)q.z};
            const double ff[3] = {(double)r4.x, (double)r4.y, (double)r4.z};
            mom[0] = wd;
            for (int d = 0; d < 3; ++d) { mom[1 + d] = wd * rr[d]; mom[4 + d] = wd * ff[d]; }
            for (int c = 0; c < 3; ++c)
              for (int a = 0; a < 3; ++a) mom[7 + a + 3 * c] = wd * ff[a] * rr[c];
          }
#pragma unroll
          for (int t = 0; t < 16; ++t) {
            const double s = warp_sum(mom[t]);
            if (lane == 0) s_red[wid][t] += s;
          }
          const double s27 = warp_sum((double)v[27]), s28 = warp_sum((double)v[28]);
          if (lane == 0) { s_red[wid][27] += s27; s_red[wid][28] += s28; }
        }
        if (xicp && S.iteration == 0) {
          // calculatePointCloudCenter (ICP.cpp:1594-1618) of the matched reading points in the frame the reading came
          // in: sums in double here, sequential float in the reference
          float4 pb = make_float4(0.f, 0.f, 0.f, 0.f);
          if (kept) pb = apply_T(S.TcInv, q);
          const double cx = warp_sum((double)pb.x), cy = warp_sum((double)pb.y), cz = warp_sum((double)pb.z);
          if (lane == 0) { s_red[wid][29] += cx; s_red[wid][30] += cy; s_red[wid][31] += cz; }
        }
      }
    }
    __syncthreads();
    if (tid < kAccum) {
      double s = 0.0;
      for (int w2 = 0; w2 < kWarps; ++w2) s += s_red[w2][tid];
      __stcg(&partials[(size_t)blockIdx.x * kAccum + tid], s);
    }
    }  // !carry
    if (stamp) ts[3] = gtimer();
    grid.sync();
    if (stamp) ts[4] = gtimer();

    // ---------------- phase 4: every CTA reduces all partials in block order, solves, updates, checks ----------------
    {  // fixed order (slice-major, then slices in order) => every CTA gets bit-identical sums
      const int c = tid & (kAccum - 1), slice = tid / kAccum;
      constexpr int kSlices = kIcpThreads / kAccum;
      double s = 0.0;
      for (unsigned b = slice; b < gridDim.x; b += kSlices) s += __ldcg(&partials[(size_t)b * kAccum + c]);
      s_red[slice][c] = s;
      __syncthreads();
      if (tid < kAccum) {
        double t = 0.0;
        for (int sl = 0; sl < kSlices; ++sl) t += s_red[sl][tid];
        s_tot[tid] = t;
      }
    }
    }  // !have_totals
    __syncthreads();
    // ---------------- X-ICP: eigen-analysis, first-iteration detection (one more pass + barrier) ----------------
    if (xicp) {
      const int P0 = (int)(s_tot[27] + 0.5);
      if (tid == 0 && P0 != 0) {
        XicpState& X = S.X;
        S.P = P0;
        int t = 0;
        for (int c = 0; c < 6; ++c)
          for (int r = 0; r <= c; ++r) {
            const float h = (float)s_tot[t++];
            S.H[r + 6 * c] = h; S.H[c + 6 * r] = h;
          }
        bool allZero = true;
        for (int i = 0; i < 36; ++i) allZero = allZero && (S.H[i] == 0.f);
        if (allZero) X.return_prior = 1;                                   // ICP.cpp:2197-2201
        else {
          if (X.repeat_eigen) xicp_eigen_analysis(S.H, X.Ut, X.Ur);        // :2205-2210
          bool zt = true, zr = true;
          for (int i = 0; i < 9; ++i) { zt = zt && (X.Ut[i] == 0.f); zr = zr && (X.Ur[i] == 0.f); }
          if (zt || zr) X.return_prior = 1;                                // :2212-2219
        }
        if (!X.return_prior && S.iteration == 0) {
          for (int c = 0; c < 3; ++c) {                                    // :2248-2267
            mul3v_dev(X.Rinv, X.Ur + 3 * c, X.UrB + 3 * c);
            mul3v_dev(X.Rinv, X.Ut + 3 * c, X.UtB + 3 * c);
          }
          for (int d = 0; d < 3; ++d) X.center[d] = (float)(s_tot[29 + d] / (double)P0);
        }
      }
      __syncthreads();
      if (P0 != 0 && S.X.return_prior) {   // ICP.cpp:1117-1121, 1175-1179: leave the loop before the minimiser
        if (tid == 0) S.iterate = 0;
        __syncthreads();
        break;
      }
      if (P0 != 0 && S.iteration == 0) {
        // alignment of every matched pair with the six eigenvectors (ICP.cpp:2270-2385)
        for (int t = tid; t < kWarps * kAccum; t += kIcpThreads) (&s_red[0][0])[t] = 0.0;
        __syncthreads();
        const float cmin = p.x_cos_min, cstr = p.x_cos_strong;
        for (int slot = gtid >> 5; slot < nslots; slot += nthreads >> 5) {
          const int i = query_of(slot, lane, nslots, coh, qstride, qpw);
          double acc[12];
#pragma unroll
          for (int t = 0; t < 12; ++t) acc[t] = 0.0;
          if (i < p.N) {
            const float d = p.dists[i], w = p.weights[i];
            const int sp = p.spos[i];
            if (d != INFINITY && w != 0.f && sp >= 0) {
              const float4 q = apply_T(S.Tc, __ldg(&p.reading[i]));
              const float4 n4 = __ldg(&p.g.nrm[sp]);
              const float4 pb = apply_T(S.TcInv, q);
              const float3 nb = apply_R(S.R9Inv, make_float3(n4.x, n4.y, n4.z));
              const float ax = pb.x - S.X.center[0], ay = pb.y - S.X.center[1], az = pb.z - S.X.center[2];
              float c0 = ay * nb.z - az * nb.y, c1 = az * nb.x - ax * nb.z, c2 = ax * nb.y - ay * nb.x;
              const float nn = sqrtf(c0 * c0 + c1 * c1 + c2 * c2);
              if (!(nn < 1.0f)) { c0 /= nn; c1 /= nn; c2 /= nn; }          // :2283-2291
#pragma unroll
              for (int e = 0; e < 3; ++e) {
                const float* vt = S.X.UtB + 3 * e;
                const float* vr = S.X.UrB + 3 * e;
                const float at = fabsf(nb.x * vt[0] + nb.y * vt[1] + nb.z * vt[2]);
                const float ar = fabsf(c0 * vr[0] + c1 * vr[1] + c2 * vr[2]);
                if (at > cmin) acc[2 * e] = (double)at;
                if (at > cstr) acc[2 * e + 1] = (double)at;
                if (ar > cmin) acc[6 + 2 * e] = (double)ar;
                if (ar > cstr) acc[7 + 2 * e] = (double)ar;
              }
            }
          }
#pragma unroll
          for (int t = 0; t < 12; ++t) {
            const double sv = warp_sum(acc[t]);
            if (lane == 0) s_red[wid][t] += sv;
          }
        }
        __syncthreads();
        if (tid < 16) {
          double sv = 0.0;
          for (int w2 = 0; w2 < kWarps; ++w2) sv += s_red[w2][tid];
          __stcg(&p.partials2[(size_t)blockIdx.x * 16 + tid], sv);
        }
        grid.sync();
        {  // fixed order (slice-major over the CTAs, then the slices in order): the same sums in every CTA
          const int c = tid & 15, slice = tid >> 4;
          constexpr int kSl = kIcpThreads / 16;
          double sv = 0.0;
          for (unsigned b = slice; b < gridDim.x; b += kSl) sv += __ldcg(&p.partials2[(size_t)b * 16 + c]);
          double* red2 = &s_red[0][0];   // kSl x 16 doubles
          red2[slice * 16 + c] = sv;
          __syncthreads();
          if (tid < 12) {
            double t = 0.0;
            for (int sl = 0; sl < kSl; ++sl) t += red2[sl * 16 + tid];
            S.X.sums[tid] = t;
          }
        }
        __syncthreads();
        if (tid == 0) {
          XicpState& X = S.X;
          X.border_mask = 0;
          for (int e = 0; e < 3; ++e) {
            int bl = 0;
            X.loc_xyz[e] = xicp_decide(X.sums[2 * e], X.sums[2 * e + 1], p.x_enough, p.x_insufficient, P0, &bl);
            if (bl) X.border_mask |= 1 << e;
            bl = 0;
            X.loc_rpy[e] = xicp_decide(X.sums[6 + 2 * e], X.sums[7 + 2 * e], p.x_enough, p.x_insufficient, P0, &bl);
            if (bl) X.border_mask |= 8 << e;
          }
        }
        __syncthreads();
        if (S.X.border_mask != 0 && wid == 0) {
          // A sum lies within the rounding band of a threshold: replay the reference's own recurrence for that
          // direction (ICP.cpp:2128-2155: sequential float sums over the matched pairs in ErrorElements order, exit
          // test in the loop condition). One warp per CTA, every CTA alike (no further barrier): the lanes evaluate
          // 32 consecutive pairs, then fold them in index order. Rare (band = 0.5 % of a threshold) and slow (~1 ms
          // per direction), exact by construction.
          const int mask = S.X.border_mask;
          float cen[3] = {S.X.center[0], S.X.center[1], S.X.center[2]};
          for (int pass = (mask & 0x38) ? -1 : 0; pass < 6; ++pass) {
            if (pass >= 0 && !((mask >> pass) & 1)) continue;
            const float* v = pass < 0 ? nullptr : (pass < 3 ? S.X.UtB + 3 * pass : S.X.UrB + 3 * (pass - 3));
            float combined = 0.f, high = 0.f, c3[3] = {0.f, 0.f, 0.f};
            bool done = false;
            for (int base = 0; base < p.N && !done; base += 32) {
              const int i = base + lane;
              bool kept = false;
              float al = 0.f, px = 0.f, py = 0.f, pz = 0.f;
              if (i < p.N) {
                const float d = __ldcg(&p.dists[i]), w = __ldcg(&p.weights[i]);
                const int sp = __ldcg(&p.spos[i]);
                if (d != INFINITY && w != 0.f && sp >= 0) {
                  kept = true;
                  const float4 q = apply_T(S.Tc, __ldg(&p.reading[i]));
                  const float4 pb = apply_T(S.TcInv, q);
                  px = pb.x; py = pb.y; pz = pb.z;
                  if (pass >= 0) {
                    const float4 n4 = __ldg(&p.g.nrm[sp]);
                    const float3 nb = apply_R(S.R9Inv, make_float3(n4.x, n4.y, n4.z));
                    if (pass < 3) al = fabsf(nb.x * v[0] + nb.y * v[1] + nb.z * v[2]);
                    else {
                      const float ax = pb.x - cen[0], ay = pb.y - cen[1], az = pb.z - cen[2];
                      float c0 = ay * nb.z - az * nb.y, c1 = az * nb.x - ax * nb.z, c2 = ax * nb.y - ay * nb.x;
                      const float nn = sqrtf(c0 * c0 + c1 * c1 + c2 * c2);
                      if (!(nn < 1.0f)) { c0 /= nn; c1 /= nn; c2 /= nn; }
                      al = fabsf(c0 * v[0] + c1 * v[1] + c2 * v[2]);
                    }
                  }
                }
              }
              unsigned km = __ballot_sync(0xffffffffu, kept);
              while (km && !done) {
                const int l = __ffs(km) - 1;
                km &= km - 1u;
                if (pass < 0) {   // calculatePointCloudCenter, ICP.cpp:1604-1611
                  c3[0] += __shfl_sync(0xffffffffu, px, l);
                  c3[1] += __shfl_sync(0xffffffffu, py, l);
                  c3[2] += __shfl_sync(0xffffffffu, pz, l);
                } else {
                  const float a = __shfl_sync(0xffffffffu, al, l);
                  if (a > p.x_cos_min) combined += a;
                  if (a > p.x_cos_strong) high += a;
                  done = (combined >= p.x_enough) || (high >= p.x_insufficient);
                }
              }
            }
            if (pass < 0) {
              const float Pf = (float)P0;
              cen[0] = c3[0] / Pf; cen[1] = c3[1] / Pf; cen[2] = c3[2] / Pf;
            } else if (lane == 0) {
              if (pass < 3) S.X.loc_xyz[pass] = done ? 1 : 0;
              else S.X.loc_rpy[pass - 3] = done ? 1 : 0;
              S.X.borderline += 1;
            }
          }
        }
        // the replay reads every CTA's matches: nobody may start the next iteration (which overwrites them) before
        // all are done
        if (S.X.border_mask != 0) grid.sync();
        else __syncthreads();
        if (tid == 0) {
          XicpState& X = S.X;
          for (int e = 0; e < 3; ++e) { X.cT[e] = X.loc_xyz[e] ? 1.f : 0.f; X.cR[e] = X.loc_rpy[e] ? 1.f : 0.f; }
          for (int c = 0; c < 3; ++c) {                                    // :2389-2411 back to the optimisation frame
            mul3v_dev(X.R, X.UrB + 3 * c, X.Ur + 3 * c);
            mul3v_dev(X.R, X.UtB + 3 * c, X.Ut + 3 * c);
          }
          X.repeat_eigen = 0;                                              // :2415-2429
          for (int i = 0; i < 3; ++i)
            if (X.loc_xyz[i] == 0 || X.loc_rpy[i] == 0) X.repeat_eigen = 1;
        }
        __syncthreads();
      }
    }
    const int n_constr = xicp ? xicp_num_constraints(S.X) : 0;
    if (wid == 0) {
      // lane 0 forms H and b; the whole warp solves (solve6_warp); lane 0 updates the pose and runs the checkers
      if (lane == 0) {
        const double* a = s_tot;
        S.P = (int)(a[27] + 0.5);
        S.wsum = (float)a[28];
        for (int t = 0; t < 36; ++t) S.H[t] = 0.f;
        for (int t = 0; t < 6; ++t) { S.b[t] = 0.f; S.x[t] = 0.f; }
        if (S.P != 0 && p.minimizer == 0) {
          int t = 0;
          for (int c = 0; c < 6; ++c)
            for (int r = 0; r <= c; ++r) {
              const float h = (float)a[t++];
              S.H[r + 6 * c] = h; S.H[c + 6 * r] = h;
            }
          for (int r = 0; r < 6; ++r) S.b[r] = -(float)a[21 + r];
        }
      }
      __syncwarp();
      unsigned long long tsub[5] = {0ull, 0ull, 0ull, 0ull, 0ull};
      if (stamp) tsub[0] = gtimer();
      int fast = 0;
      if (S.P != 0 && p.minimizer == 0 && n_constr == 0) fast = solve6_warp(S.H, S.b, S.x, &s_red[1][0], S.x_spd);
      __syncwarp();
     if (lane == 0) {
      const double* a = s_tot;
      const int P = S.P;
      float real[16];
      eye4(real);
      if (P == 0) {
        S.status = B200_E_NO_MATCHES;  // ErrorElements: "no point to minimize", ErrorMinimizer.cpp:75-77
        S.iterate = 0;
      } else {
        if (p.minimizer == 0) {
          if (fast == 1) S.solve_info = 0;
          else if (n_constr > 0)   // equality-constrained (KKT) system, PointToPlane.cpp:461-505
            xicp_solve(S.H, S.b, S.X, S.x, &S.solve_info, *reinterpret_cast<KktScratch*>(s_dyn));
          else solve6_dev(S.H, S.b, S.x, &S.solve_info, fast == 2 ? S.x_spd : nullptr);  // rank-deficient / borderline: serial reference path
          if (stamp) tsub[1] = gtimer();
          p2plane_T_dev(S.x, real);
        } else {
          p2point_from_moments(a, real);
        }
        mul4_dev(real, S.T_iter, S.T_iter);  // ICP.cpp:1213
        if (stamp) tsub[2] = gtimer();
        // transformationCheckers.check, ICP.cpp:1292-1298
        bool counterHit = false;
        if (p.max_iterations > 0) {
          S.counter += 1.f;
          if (S.counter >= (float)p.max_iterations) counterHit = true;
        }
        if (p.use_differential) {
          const int n = S.nrot;
          S.rot[n % kMaxHist] = quat_from_T(S.T_iter);
          S.tr[n % kMaxHist][0] = S.T_iter[12]; S.tr[n % kMaxHist][1] = S.T_iter[13]; S.tr[n % kMaxHist][2] = S.T_iter[14];
          S.nrot = n + 1;
          float cv0 = 0.f, cv1 = 0.f;
          const int Ls = p.smooth;
          if (S.nrot > Ls) {
            for (int i = S.nrot - 1; i >= S.nrot - Ls; --i) {
              const int c = i % kMaxHist, pr = (i - 1) % kMaxHist;
              cv0 += fabsf(angular_distance(S.rot[c], S.rot[pr]));
              const float dx = S.tr[c][0] - S.tr[pr][0], dy = S.tr[c][1] - S.tr[pr][1], dz = S.tr[c][2] - S.tr[pr][2];
              cv1 += fabsf(sqrtf(dx * dx + dy * dy + dz * dz));
            }
            cv0 /= (float)Ls;
            cv1 /= (float)Ls;
            if (cv0 < p.min_rot && cv1 < p.min_trans) S.iterate = 0;
          }
          if (cv0 != cv0 || cv1 != cv1) { S.status = B200_E_NAN; S.iterate = 0; }
        }
        if (p.use_bound && S.status == B200_OK) {
          // BoundTransformationChecker::check (TransformationCheckersImpl.cpp:198-222) against the pose stored by init
          // (T_iter = identity, ICP.cpp:997): throws ConvergenceError when a bound is exceeded
          const float cvr = angular_distance(quat_from_T(S.T_iter), S.rot0);
          const float cvt = sqrtf(S.T_iter[12] * S.T_iter[12] + S.T_iter[13] * S.T_iter[13] + S.T_iter[14] * S.T_iter[14]);
          if (cvr > p.bound_rot || cvt > p.bound_trans) { S.status = B200_E_BOUND; S.iterate = 0; }
        }
        if (counterHit && S.status == B200_OK) { S.iterate = 0; S.maxReached = 1; }
        if (stamp) tsub[3] = gtimer();
      }
      if (blockIdx.x == 0 && S.iteration < kMaxLog && p.log) {
        IcpLogEntry& e = p.log[S.iteration];
        for (int t = 0; t < 36; ++t) e.H[t] = S.H[t];
        for (int t = 0; t < 6; ++t) { e.b[t] = S.b[t]; e.x[t] = S.x[t]; }
        for (int t = 0; t < 16; ++t) e.T_iter[t] = S.T_iter[t];
        e.P = P; e.solve_info = S.solve_info; e.path = S.path; e.n_cand = n_cand_seen; e.limit = limit; e.delta = S.delta;
        ts[5] = gtimer();
        for (int t = 0; t < 6; ++t) e.ts[t] = ts[t];
        e.sub_ns[0] = (unsigned)(tsub[0] - ts[4]);
        e.sub_ns[1] = (unsigned)(tsub[1] - tsub[0]);
        e.sub_ns[2] = (unsigned)(tsub[2] - tsub[1]);
        e.sub_ns[3] = (unsigned)(tsub[3] - tsub[2]);
      }
      if (p.use_trimmed) {
        // bracket for the next iteration: wide enough for the movement just seen (4x), narrow enough to keep the
        // candidate set at a few hundred; if both cannot be had the next iteration takes the general shape at once
        const float rel = S.have_bracket ? fabsf(limit - S.lim_prev) / S.lim_prev : 1.f;
        if (S.path != 0 && n_cand_seen > 0) S.dens = (float)n_cand_seen / S.delta;   // candidates per unit of delta
        else if (S.dens == 0.f) S.dens = 0.5f * (float)p.N * (float)p.k;
        const float dmax = fminf(kBracketCands / S.dens, 0.25f);
        // predicted movement of the limit: what was just seen, scaled by how fast the pose updates are shrinking
        float step = 0.f;
        for (int t = 0; t < 6; ++t) step += S.x[t] * S.x[t];
        step = sqrtf(step);
        const float shrink = (S.step_prev > 0.f) ? fminf(step / S.step_prev, 1.f) : 1.f;
        S.step_prev = step;
        S.try_fast = (2.f * rel * shrink <= dmax) ? 1 : 0;
        S.delta = dmax;
        S.lim_prev = limit;
        S.have_bracket = (limit > 0.f && limit < INFINITY) ? 1 : 0;
        S.lo = limit * (1.f - dmax);
        S.hi = limit * (1.f + dmax);
      }
      S.iteration += 1;
     }
    }
    __syncthreads();
    if (!S.iterate) break;
    if (p.max_launch_iters > 0 && ++launch_iters >= p.max_launch_iters) break;
  }
  if (p.state && blockIdx.x == 0) {   // every CTA holds the same state; CTA 0 saves it for the next launch
    __syncthreads();
    for (int t = tid; t < (int)(sizeof(BlockState) / sizeof(unsigned)); t += kIcpThreads)
      __stcg(&p.state[t], reinterpret_cast<const unsigned*>(&S)[t]);
  }

  // points touched: one atomic per warp
  {
    unsigned long long v = visited_total;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0 && v) atomicAdd(&p.out->points_touched, v);
  }
  if (blockIdx.x == 0 && tid == 0) {
    IcpDeviceResult* o = p.out;
    o->status = S.status;
    o->iterations = S.iteration;
    o->max_iterations_reached = S.maxReached;
    o->iterate = S.iterate;
    o->n_matches_last = S.P;
    for (int t = 0; t < 16; ++t) o->T_iter[t] = S.T_iter[t];
    // move the result back to the callers' frames, ICP.cpp:1345
    float invRead[16], r1[16], r2[16], fr[48];
    for (int t = 0; t < 48; ++t) fr[t] = p.frames[t];
    inv4_dev(fr + 32, invRead);
    mul4_dev(fr, S.T_iter, r1);
    mul4_dev(r1, fr + 16, r2);
    mul4_dev(r2, invRead, r1);
    for (int t = 0; t < 16; ++t) o->T[t] = r1[t];
    const float denom = (float)((long long)p.k * (long long)p.N);
    o->weighted_point_used_ratio = S.wsum / denom;   // ErrorMinimizer.cpp:139-140
    o->point_used_ratio = (float)S.P / denom;
    o->returned_prior = S.X.return_prior; o->xicp_borderline = S.X.borderline;
    for (int i = 0; i < 3; ++i) { o->loc_xyz[i] = S.X.loc_xyz[i]; o->loc_rpy[i] = S.X.loc_rpy[i]; }
    for (int i = 0; i < 9; ++i) { o->eig_t[i] = S.X.Ut[i]; o->eig_r[i] = S.X.Ur[i]; }
    for (int i = 0; i < 12; ++i) o->xicp_sums[i] = S.X.sums[i];
  }
}

template <int K>
int launch_k(const IcpKernelParams& p, int device, cudaStream_t s, int* grid_out) {
  int per_sm = 0, sms = 0;
  B200_CUDA(cudaFuncSetAttribute(icp_persistent_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kIcpDynSmem));
  B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, icp_persistent_kernel<K>, kIcpThreads, kIcpDynSmem));
  B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  if (per_sm < 1) { set_error("persistent ICP kernel does not fit on an SM"); return B200_E_CUDA; }
  int grid = per_sm * sms;
  IcpKernelParams pc = p;
  // A small scan does not fill the machine with 32 queries per warp (25 k points = 49 CTAs on 148 SMs, and the
  // matching phase lasts as long as the slowest warp): give a warp fewer queries, down to 4, while the warps still
  // fit in one wave. k = 1 only (the k > 1 search is per thread).
  pc.qpw_log2 = 5;
  if (K == 1) {
    const int warps_cap = grid * (kIcpThreads / 32);
    while (pc.qpw_log2 > 2 && ((p.N + (1 << (pc.qpw_log2 - 1)) - 1) >> (pc.qpw_log2 - 1)) <= warps_cap) --pc.qpw_log2;
    if (p.qpw_log2 >= 2 && p.qpw_log2 <= 5) pc.qpw_log2 = p.qpw_log2;   // pinned by the caller (B200_ICP_QPW)
  }
  // Warm start of the searches (knn.cuh): one more dependent load per query and iteration against fewer cells to
  // visit. Measured: a loss on the dense, uniform map (C2 0.46 -> 0.50 ms: the matching phase is a chain of memory
  // latencies there, not a volume of candidates), a gain on a raw scan as target (car_cloud 1.91 -> 1.67 ms:
  // crowded cells). Small scans are registered against such targets, so: on iff the warps were thinned out.
  if (pc.warm_start < 0) pc.warm_start = pc.qpw_log2 < 5 ? 1 : 0;
  // do not launch more CTAs than there is work for (fewer CTAs = cheaper grid barriers)
  const int nslots = (p.N + (1 << pc.qpw_log2) - 1) >> pc.qpw_log2;
  const int need = std::max(1, (nslots + kIcpThreads / 32 - 1) / (kIcpThreads / 32));
  grid = std::min(grid, need);
  pc.cand_seg = icp_cand_segment(nslots, grid);
  pc.cand_cap = icp_cand_capacity(p.N);   // per parity
  if (grid > kIcpThreads || (size_t)grid * (size_t)pc.cand_seg > icp_cand_capacity(p.N)) pc.fast_path = 0;
  void* args[] = {(void*)&pc};
  B200_CUDA(cudaLaunchCooperativeKernel((void*)icp_persistent_kernel<K>, dim3(grid), dim3(kIcpThreads), args,
                                        kIcpDynSmem, s));
  if (grid_out) *grid_out = grid;
  return B200_OK;
}

}  // namespace

// The kernel is instantiated for five list sizes; the two large ones (k-best lists of 32 and 64) are compiled as a
// translation unit of their own (icp_fused_big.cu includes this file with B200_ICP_FUSED_PART = 1) so that the two
// halves build in parallel.
#ifndef B200_ICP_FUSED_PART
#define B200_ICP_FUSED_PART 0
#endif

#if B200_ICP_FUSED_PART == 1
int launch_icp_kernel_big(const IcpKernelParams& p, int device, cudaStream_t s, int* grid_out) {
  if (p.k <= 32) return launch_k<32>(p, device, s, grid_out);
  return launch_k<64>(p, device, s, grid_out);   // the k-best list lives in local memory from here on
}
#else
int launch_icp_kernel_big(const IcpKernelParams& p, int device, cudaStream_t s, int* grid_out);

// candidates one CTA can park: every query it owns (rounds x threads)
int icp_cand_segment(int nslots, int grid) {
  const int warps = grid * (kIcpThreads / 32);
  const int rounds = (nslots + warps - 1) / warps;
  return rounds * kIcpThreads;
}
size_t icp_cand_capacity(int N) { return (size_t)N + (size_t)kIcpThreads * 1024; }

size_t icp_partials_capacity(int device) {
  int sms = 148;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) { cudaGetLastError(); sms = 148; }
  return (size_t)sms * 8 * kAccum * 2;  // up to 8 CTAs per SM, two parities
}

int launch_icp_kernel(const IcpKernelParams& p, int device, cudaStream_t s, int* grid_out) {
  if (p.k <= 1) return launch_k<1>(p, device, s, grid_out);
  if (p.k <= 4) return launch_k<4>(p, device, s, grid_out);
  if (p.k <= 16) return launch_k<16>(p, device, s, grid_out);
  if (p.k <= 64) return launch_icp_kernel_big(p, device, s, grid_out);
  set_error("ICP chain supports knn <= 64 (the limit of the matcher, b200_index_knn)");
  return B200_E_UNSUPPORTED;
}
#endif

}  // namespace b200
