// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
//
// Exact k-nearest-neighbour search standing in for libnabo (un-vendored dependency, >= 1.0.7, version floor
// only: libpointmatcher/pointmatcher/MatchersImpl.h:41-44). Call site restated:
//   KDTreeMatcher::findClosests, libpointmatcher/pointmatcher/MatchersImpl.cpp:85-102
//     featureNNS->knn(query, ids, dists, knn, epsilon, NNS::ALLOW_SELF_MATCH, maxDist)
// Contract at epsilon = 0 (libnabo's published algorithm; parity unpinned, no reference test pins ids):
//   * squared L2 in float, accumulated x -> y -> z, separate multiply and add (no FMA contraction; this file is
//     compiled with -ffp-contract=off)
//   * a candidate is accepted iff dist2 <= maxDist^2 and it beats the current k-th best
//   * results ascending by (dist2, index); ties -> smallest index (north_star rule)
//   * unfilled slots: id = -1 (Matches::InvalidId), dist = +inf (Matches::InvalidDist), MatchersImpl.cpp:96-97
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <limits>
#include <thread>
#include <vector>

namespace {

struct KBest {
  int k;
  int32_t* ids;
  float* dists;
  void init(int k_, int32_t* i, float* d) {
    k = k_; ids = i; dists = d;
    for (int j = 0; j < k; ++j) { ids[j] = -1; dists[j] = std::numeric_limits<float>::infinity(); }
  }
  inline bool better(float d, int32_t id, int slot) const {
    return d < dists[slot] || (d == dists[slot] && id < ids[slot]);
  }
  inline void offer(float d, int32_t id) {
    if (!better(d, id, k - 1)) return;
    int j = k - 1;
    while (j > 0 && better(d, id, j - 1)) { dists[j] = dists[j - 1]; ids[j] = ids[j - 1]; --j; }
    dists[j] = d; ids[j] = id;
  }
  inline float worst() const { return dists[k - 1]; }
};

inline float dist2(const float* q, const float* p) {
  const float dx = q[0] - p[0], dy = q[1] - p[1], dz = q[2] - p[2];
  float d = dx * dx;
  d = d + dy * dy;
  d = d + dz * dz;
  return d;
}

inline float maxDist2(float maxDist) {
  return std::isinf(maxDist) ? std::numeric_limits<float>::infinity() : maxDist * maxDist;
}

struct Node {
  int32_t left, right;  // children (internal) or [begin, end) into pts (leaf, dim == -1)
  int32_t dim;
  float split;
};

struct Tree {
  std::vector<Node> nodes;
  std::vector<float> pts;     // reordered xyz + original index bits, 4 floats per point
  std::vector<int32_t> orig;  // original index of reordered point
  int M = 0;
};

constexpr int kBucket = 8;

int buildRec(Tree& t, const float* ref, std::vector<int32_t>& idx, int first, int last) {
  const int me = (int)t.nodes.size();
  t.nodes.push_back(Node{});
  if (last - first <= kBucket) {
    t.nodes[me] = Node{first, last, -1, 0.f};
    return me;
  }
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = first; i < last; ++i)
    for (int d = 0; d < 3; ++d) {
      const float v = ref[4 * (size_t)idx[i] + d];
      mn[d] = std::min(mn[d], v); mx[d] = std::max(mx[d], v);
    }
  int dim = 0;
  if (mx[1] - mn[1] > mx[dim] - mn[dim]) dim = 1;
  if (mx[2] - mn[2] > mx[dim] - mn[dim]) dim = 2;
  const int mid = first + (last - first) / 2;
  std::nth_element(idx.begin() + first, idx.begin() + mid, idx.begin() + last, [&](int32_t a, int32_t b) {
    const float va = ref[4 * (size_t)a + dim], vb = ref[4 * (size_t)b + dim];
    return va < vb || (va == vb && a < b);
  });
  const float split = ref[4 * (size_t)idx[mid] + dim];
  const int l = buildRec(t, ref, idx, first, mid);
  const int r = buildRec(t, ref, idx, mid, last);
  t.nodes[me] = Node{l, r, dim, split};
  return me;
}

// off[d] = distance from the query to the current cell along axis d (0 if inside). The bound is re-summed in the
// same x,y,z order as dist2, so bound <= any computed dist2 of a point in the cell (monotone rounding).
void searchRec(const Tree& t, int n, const float* q, float* off, KBest& best, float md2, long& touched) {
  const Node& nd = t.nodes[n];
  if (nd.dim < 0) {
    for (int i = nd.left; i < nd.right; ++i) {
      const float d = dist2(q, &t.pts[4 * (size_t)i]);
      ++touched;
      if (d <= md2) best.offer(d, t.orig[i]);
    }
    return;
  }
  const float diff = q[nd.dim] - nd.split;
  const int nearC = diff < 0.f ? nd.left : nd.right;
  const int farC = diff < 0.f ? nd.right : nd.left;
  searchRec(t, nearC, q, off, best, md2, touched);
  const float old = off[nd.dim];
  // points in the far child satisfy |q - p| >= |diff| along dim (left: p <= split, right: p >= split)
  const float ad = std::fabs(diff);
  if (ad > old) off[nd.dim] = ad;
  float bound = off[0] * off[0];
  bound = bound + off[1] * off[1];
  bound = bound + off[2] * off[2];
  if (bound <= md2 && bound <= best.worst()) searchRec(t, farC, q, off, best, md2, touched);
  off[nd.dim] = old;
}

}  // namespace

extern "C" {

int orc_knn_brute(const float* ref, int M, const float* query, int N, int k, float maxDist, int32_t* ids,
                  float* dists) {
  if (k <= 0 || M < 0 || N < 0) return ORC_E_BAD_PARAM;
  const float md2 = maxDist2(maxDist);
  for (int i = 0; i < N; ++i) {
    KBest best;
    best.init(k, ids + (size_t)i * k, dists + (size_t)i * k);
    const float* q = query + 4 * (size_t)i;
    for (int j = 0; j < M; ++j) {
      const float d = dist2(q, ref + 4 * (size_t)j);
      if (d <= md2) best.offer(d, j);
    }
  }
  return ORC_OK;
}

void* orc_kdtree_build(const float* ref, int M) {
  Tree* t = new Tree();
  t->M = M;
  std::vector<int32_t> idx(M);
  for (int i = 0; i < M; ++i) idx[i] = i;
  t->nodes.reserve((size_t)M / kBucket * 4 + 16);
  if (M > 0) buildRec(*t, ref, idx, 0, M);
  t->pts.resize(4 * (size_t)M);
  t->orig = idx;
  for (int i = 0; i < M; ++i)
    for (int d = 0; d < 4; ++d) t->pts[4 * (size_t)i + d] = ref[4 * (size_t)idx[i] + d];
  return t;
}

void orc_kdtree_free(void* tree) { delete static_cast<Tree*>(tree); }

long orc_kdtree_knn(const void* tree, const float* query, int N, int k, float maxDist, int nthreads, int32_t* ids,
                    float* dists) {
  const Tree& t = *static_cast<const Tree*>(tree);
  const float md2 = maxDist2(maxDist);
  if (nthreads < 1) nthreads = 1;
  std::vector<long> touched(nthreads, 0);
  auto work = [&](int tid) {
    const int lo = (int)((long)N * tid / nthreads), hi = (int)((long)N * (tid + 1) / nthreads);
    long tc = 0;
    for (int i = lo; i < hi; ++i) {
      KBest best;
      best.init(k, ids + (size_t)i * k, dists + (size_t)i * k);
      if (t.M == 0) continue;
      float off[3] = {0.f, 0.f, 0.f};
      searchRec(t, 0, query + 4 * (size_t)i, off, best, md2, tc);
    }
    touched[tid] = tc;
  };
  if (nthreads == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int i = 0; i < nthreads; ++i) th.emplace_back(work, i);
    for (auto& x : th) x.join();
  }
  long total = 0;
  for (long v : touched) total += v;
  return total;
}

}  // extern "C"
