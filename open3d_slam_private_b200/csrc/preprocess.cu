// Scan pre-processing on the device: what open3d_slam runs on every scan before the registration
// (ScanToMapRegistration.cpp:36-54, Odometry.cpp:22-26) and on the map before a patch is registered against
// (ScanToMapRegistration.cpp:90-104):
//   CroppingVolume::crop / getIndicesWithinVolume   open3d_slam/src/croppers.cpp:53-95, volumes :115-164
//   o3d_slam::voxelize -> PointCloud::VoxelDownSample   helpers.cpp:108-115 (Open3D v0.15.1, recalled from upstream)
// Both are HBM-bound streaming passes. Crop: predicate in double on the float coordinates, order-preserving
// compaction (flag -> scan -> scatter). Voxel down-sample: voxel key per point, open-addressing hash table built with
// 64-bit atomicCAS, voxels numbered by the index of their first point (deterministic; Open3D's own order is that of an
// std::unordered_map). The reference sums a voxel's points in double while it walks the cloud; a double sum of floats is
// exact in EVERY order when the values' binades and their count leave room in the 53-bit mantissa (sum_is_exact), so
// those voxels — practically all — are summed with warp-folded atomics; for the rest one CTA per voxel collects the
// members in index order and one thread sums them in input order. Either way the means are bit-identical to the oracle's.
#include "common.cuh"
#include "grid_index.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>

using namespace b200;

namespace b200 {
size_t scan_tiles_for(size_t n);
int inclusive_scan_u32(uint32_t* d_data, size_t n, uint32_t* d_tile, cudaStream_t s);
}  // namespace b200

namespace {

constexpr int kThreads = 256;
constexpr unsigned long long kEmpty = 0xffffffffffffffffull;
constexpr uint32_t kSmallVoxel = 64;   // ordered voxels up to this size are sorted and summed by one thread

inline int blocks_for(size_t n) { return (int)std::max<size_t>(1, std::min<size_t>((n + kThreads - 1) / kThreads, 148 * 16)); }

struct CropP {
  int kind, invert;
  double rmin, rmax, zmin, zmax, cx, cy, cz;
};

__global__ void crop_flag_kernel(const float4* __restrict__ p, int N, CropP c, uint32_t* __restrict__ flag) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const float4 v = __ldg(&p[i]);
    const double px = (double)v.x, py = (double)v.y, pz = (double)v.z;
    const double dx = px - c.cx, dy = py - c.cy, dz = pz - c.cz;
    bool in = true;
    if (c.kind == 1) in = pz >= c.zmin && pz <= c.zmax && sqrt(dx * dx + dy * dy) <= c.rmax;        // croppers.cpp:162-164
    else if (c.kind == 2) in = sqrt(dx * dx + dy * dy + dz * dz) >= c.rmin;                          // :149-151
    else if (c.kind == 3) in = sqrt(dx * dx + dy * dy + dz * dz) <= c.rmax;                          // :131-133
    else if (c.kind == 4) { const double d = sqrt(dx * dx + dy * dy + dz * dz); in = d <= c.rmax && d >= c.rmin; }  // :115-118
    if (c.invert) in = !in;                                                                           // :53-55
    flag[i] = in ? 1u : 0u;
  }
}

// scan = inclusive scan of the flags: point i is kept iff scan[i] != scan[i - 1]; its slot is scan[i] - 1
__global__ void crop_scatter_kernel(const float4* __restrict__ p, const float* __restrict__ n3, int N,
                                    const uint32_t* __restrict__ scan, float4* __restrict__ out,
                                    float* __restrict__ outn3, int* __restrict__ idx) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const uint32_t s = scan[i], s0 = i ? scan[i - 1] : 0u;
    if (s == s0) continue;
    const uint32_t o = s - 1u;
    out[o] = __ldg(&p[i]);
    idx[o] = i;
    if (n3) {
      outn3[3 * (size_t)o] = n3[3 * (size_t)i];
      outn3[3 * (size_t)o + 1] = n3[3 * (size_t)i + 1];
      outn3[3 * (size_t)o + 2] = n3[3 * (size_t)i + 2];
    }
  }
}

__device__ __forceinline__ uint32_t f2ord(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// min and max per axis: 6 ordered-uint atomics per block
__global__ void minmax_kernel(const float4* __restrict__ p, int N, uint32_t* __restrict__ out6) {
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const float4 v = __ldg(&p[i]);
    mn[0] = fminf(mn[0], v.x); mn[1] = fminf(mn[1], v.y); mn[2] = fminf(mn[2], v.z);
    mx[0] = fmaxf(mx[0], v.x); mx[1] = fmaxf(mx[1], v.y); mx[2] = fmaxf(mx[2], v.z);
  }
  __shared__ float s[6][kThreads / 32];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    for (int o = 16; o > 0; o >>= 1) {
      mn[d] = fminf(mn[d], __shfl_xor_sync(0xffffffffu, mn[d], o));
      mx[d] = fmaxf(mx[d], __shfl_xor_sync(0xffffffffu, mx[d], o));
    }
    if ((threadIdx.x & 31) == 0) { s[d][threadIdx.x >> 5] = mn[d]; s[3 + d][threadIdx.x >> 5] = mx[d]; }
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float v = s[threadIdx.x][0];
    for (int w = 1; w < kThreads / 32; ++w) v = threadIdx.x < 3 ? fminf(v, s[threadIdx.x][w]) : fmaxf(v, s[threadIdx.x][w]);
    if (threadIdx.x < 3) atomicMin(&out6[threadIdx.x], f2ord(v));
    else atomicMax(&out6[threadIdx.x], f2ord(v));
  }
}

struct VoxP {
  double mbx, mby, mbz, vs;
  unsigned long long mask;   // table size - 1 (power of two)
};

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
  return x;
}

// biased exponent range of the non-zero values (denormals count as exponent 1: same quantum 2^-149)
__device__ __forceinline__ void exp_range(float v, int& lo, int& hi) {
  const unsigned b = __float_as_uint(v) & 0x7fffffffu;
  if (b == 0u) return;
  const int e = max((int)(b >> 23), 1);
  lo = min(lo, e); hi = max(hi, e);
}

// A sum of n floats in double is EXACT in every order (all partial sums representable) when
// (emax - emin) + 24 + ceil(log2 n) <= 53; then the order-free atomic sums below equal the reference's sequential sum
// bit for bit. Voxels that fail the test (values spanning many binades) take the ordered path.
__device__ __forceinline__ bool sum_is_exact(int emin, int emax, unsigned n) {
  if (emax == 0) return true;   // all zeros
  const int clog = n <= 1u ? 0 : 32 - __clz(n - 1u);
  return (emax - emin) + 24 + clog <= 53;
}

// voxel index = floor((p - voxel_min_bound) / voxel_size) per axis (double, like Open3D); key = 21 bits per axis.
// Inserts the key (linear probing) and remembers the slot of every point. The lanes of a warp that fell into the same
// voxel (consecutive scan points usually do) are folded first, in lane = index order, and ONE lane per voxel and warp
// issues the atomics: member count, smallest member index, sums of the points (and normals) in double, exponent range.
__global__ void voxel_insert_kernel(const float4* __restrict__ p, const float* __restrict__ n3, int N, VoxP v,
                                    unsigned long long* __restrict__ keys, uint32_t* __restrict__ cnt,
                                    int* __restrict__ first, uint32_t* __restrict__ slot_of, double* __restrict__ sums,
                                    int* __restrict__ exps, int* __restrict__ err) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  for (int base = (blockIdx.x * blockDim.x + threadIdx.x) - lane; base < N; base += gridDim.x * blockDim.x) {
    const int i = base + lane;
    uint32_t s = 0xffffffffu;
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    float nx = 0.f, ny = 0.f, nz = 0.f;
    if (i < N) {
      q = __ldg(&p[i]);
      if (n3) { nx = n3[3 * (size_t)i]; ny = n3[3 * (size_t)i + 1]; nz = n3[3 * (size_t)i + 2]; }
      const double fx = floor(((double)q.x - v.mbx) / v.vs), fy = floor(((double)q.y - v.mby) / v.vs),
                   fz = floor(((double)q.z - v.mbz) / v.vs);
      if (!(fx >= 0.0 && fx < 2097152.0 && fy >= 0.0 && fy < 2097152.0 && fz >= 0.0 && fz < 2097152.0)) {
        *err = 1;   // NaN coordinate or more than 2^21 voxels along an axis
      } else {
        const unsigned long long key = (unsigned long long)fx | ((unsigned long long)fy << 21) | ((unsigned long long)fz << 42);
        unsigned long long t = mix64(key) & v.mask;
        while (true) {
          const unsigned long long old = atomicCAS(&keys[t], kEmpty, key);
          if (old == kEmpty || old == key) break;
          t = (t + 1ull) & v.mask;
        }
        s = (uint32_t)t;
      }
      slot_of[i] = s;
    }
    // fold the lanes of the same voxel, ascending lane order
    double sx = 0.0, sy = 0.0, sz = 0.0, tx = 0.0, ty = 0.0, tz = 0.0;
    int plo = 255, phi = 0, nlo = 255, nhi = 0, leader = -1;
    unsigned members = 0;
    for (int l = 0; l < 32; ++l) {
      const uint32_t sl = __shfl_sync(FULL, s, l);
      const float ax = __shfl_sync(FULL, q.x, l), ay = __shfl_sync(FULL, q.y, l), az = __shfl_sync(FULL, q.z, l);
      float bx = 0.f, by = 0.f, bz = 0.f;
      if (n3) { bx = __shfl_sync(FULL, nx, l); by = __shfl_sync(FULL, ny, l); bz = __shfl_sync(FULL, nz, l); }
      if (sl == s && s != 0xffffffffu) {
        if (leader < 0) leader = l;
        ++members;
        sx += (double)ax; sy += (double)ay; sz += (double)az;
        exp_range(ax, plo, phi); exp_range(ay, plo, phi); exp_range(az, plo, phi);
        if (n3) {
          tx += (double)bx; ty += (double)by; tz += (double)bz;
          exp_range(bx, nlo, nhi); exp_range(by, nlo, nhi); exp_range(bz, nlo, nhi);
        }
      }
    }
    if (leader == lane) {
      atomicAdd(&cnt[s], members);
      atomicMin(&first[s], i);
      double* a = sums + 6 * (size_t)s;
      atomicAdd(a, sx); atomicAdd(a + 1, sy); atomicAdd(a + 2, sz);
      int* e = exps + 4 * (size_t)s;
      if (phi) { atomicMin(e, plo); atomicMax(e + 1, phi); }
      if (n3) {
        atomicAdd(a + 3, tx); atomicAdd(a + 4, ty); atomicAdd(a + 5, tz);
        if (nhi) { atomicMin(e + 2, nlo); atomicMax(e + 3, nhi); }
      }
    }
  }
}

__global__ void exps_init_kernel(int* __restrict__ exps, size_t T) {
  for (size_t s = blockIdx.x * (size_t)blockDim.x + threadIdx.x; s < T; s += (size_t)gridDim.x * blockDim.x) {
    exps[4 * s] = 255; exps[4 * s + 1] = 0; exps[4 * s + 2] = 255; exps[4 * s + 3] = 0;
  }
}

// flag[first point of every voxel] = 1
__global__ void voxel_mark_kernel(const unsigned long long* __restrict__ keys, const int* __restrict__ first,
                                  size_t T, uint32_t* __restrict__ flag) {
  for (size_t s = blockIdx.x * (size_t)blockDim.x + threadIdx.x; s < T; s += (size_t)gridDim.x * blockDim.x)
    if (keys[s] != kEmpty) flag[first[s]] = 1u;
}

// voxel number = (inclusive scan of the flags)[first] - 1. Voxels whose sums are not provably exact (sum_is_exact) keep
// their member count in offs[v + 1] (for the scan that lays out the member lists of the ordered path); the others get 0.
__global__ void voxel_number_kernel(const unsigned long long* __restrict__ keys, const int* __restrict__ first,
                                    const uint32_t* __restrict__ cnt, const int* __restrict__ exps, int with_normals,
                                    size_t T, const uint32_t* __restrict__ scan, uint32_t* __restrict__ vox_of_slot,
                                    uint32_t* __restrict__ slot_of_vox, uint32_t* __restrict__ offs,
                                    int* __restrict__ firstK, int* __restrict__ n_ordered,
                                    uint32_t* __restrict__ ordered_list) {
  for (size_t s = blockIdx.x * (size_t)blockDim.x + threadIdx.x; s < T; s += (size_t)gridDim.x * blockDim.x)
    if (keys[s] != kEmpty) {
      const uint32_t v = scan[first[s]] - 1u;
      vox_of_slot[s] = v;
      slot_of_vox[v] = (uint32_t)s;
      firstK[v] = first[s];
      const int* e = exps + 4 * s;
      const bool exact = sum_is_exact(e[0], e[1], cnt[s]) && (!with_normals || sum_is_exact(e[2], e[3], cnt[s]));
      offs[v + 1] = exact ? 0u : cnt[s];
      if (!exact) {
        atomicAdd(n_ordered, 1);
        if (cnt[s] > kSmallVoxel) ordered_list[atomicAdd(n_ordered + 1, 1)] = v;   // large: one CTA each
      }
    }
}

// member lists of the SMALL ordered voxels (any order; voxel_mean_kernel sorts them)
__global__ void voxel_members_kernel(const uint32_t* __restrict__ slot_of, int N, const uint32_t* __restrict__ vox_of_slot,
                                     const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ offs,
                                     uint32_t* __restrict__ cursor, int* __restrict__ members) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const uint32_t s = slot_of[i];
    if (s == 0xffffffffu) continue;
    const uint32_t v = vox_of_slot[s];
    if (offs[v + 1] == offs[v] || cnt[s] > kSmallVoxel) continue;
    members[offs[v] + atomicAdd(&cursor[v], 1u)] = i;
  }
}

// Ordered path for LARGE voxels, one CTA per voxel that failed sum_is_exact (offs = inclusive scan of their member counts => offs[v] is
// the start of voxel v's member list). The warps walk disjoint, ascending chunks of the cloud, pick the points of this
// voxel (count, then write: ballot prefix keeps the index order inside a warp, the chunk order keeps it across warps),
// and thread 0 then runs the reference's recurrence s <- fl(s + x_i) over the list. The search is the expensive part and
// is parallel; the additions are few and sequential by nature.
__global__ void __launch_bounds__(kThreads) voxel_ordered_kernel(const float4* __restrict__ p, const float* __restrict__ n3,
                                                                 int N, const uint32_t* __restrict__ ordered_list,
                                                                 const uint32_t* __restrict__ slot_of_vox,
                                                                 const uint32_t* __restrict__ slot_of,
                                                                 const uint32_t* __restrict__ offs, int* __restrict__ members,
                                                                 double* __restrict__ out3, double* __restrict__ outn3) {
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int kW = kThreads / 32;
  __shared__ uint32_t s_cnt[kW];
  const uint32_t v = ordered_list[blockIdx.x];
  const uint32_t s = slot_of_vox[v];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int chunk = (((N + kW - 1) / kW) + 31) & ~31;
  const int lo = w * chunk, hi = min(N, lo + chunk);
  uint32_t c = 0;
  for (int base = lo; base < hi; base += 32) {
    const int i = base + lane;
    c += __popc(__ballot_sync(FULL, i < hi && slot_of[i] == s));
  }
  if (lane == 0) s_cnt[w] = c;
  __syncthreads();
  uint32_t o = offs[v];
  for (int k = 0; k < w; ++k) o += s_cnt[k];
  for (int base = lo; base < hi; base += 32) {
    const int i = base + lane;
    const bool m = i < hi && slot_of[i] == s;
    const unsigned bal = __ballot_sync(FULL, m);
    if (m) members[o + __popc(bal & ((1u << lane) - 1u))] = i;
    o += __popc(bal);
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const uint32_t a = offs[v], b = offs[v + 1];
  double sx = 0.0, sy = 0.0, sz = 0.0, nx = 0.0, ny = 0.0, nz = 0.0;
  for (uint32_t i = a; i < b; ++i) {
    const int m = members[i];
    const float4 q = __ldg(&p[m]);
    sx += (double)q.x; sy += (double)q.y; sz += (double)q.z;
    if (n3) { nx += (double)n3[3 * (size_t)m]; ny += (double)n3[3 * (size_t)m + 1]; nz += (double)n3[3 * (size_t)m + 2]; }
  }
  const double cn = (double)(b - a);
  out3[3 * (size_t)v] = sx / cn; out3[3 * (size_t)v + 1] = sy / cn; out3[3 * (size_t)v + 2] = sz / cn;
  if (n3) { outn3[3 * (size_t)v] = nx / cn; outn3[3 * (size_t)v + 1] = ny / cn; outn3[3 * (size_t)v + 2] = nz / cn; }
}

// one thread per voxel. Passed sum_is_exact: average of the order-free sums. Small ordered voxel: members into ascending
// index order (insertion sort), sums in double in that order. Large ordered voxels: voxel_ordered_kernel.
__global__ void voxel_mean_kernel(const float4* __restrict__ p, const float* __restrict__ n3, int K,
                                  const uint32_t* __restrict__ offs, int* __restrict__ members,
                                  const uint32_t* __restrict__ slot_of_vox, const uint32_t* __restrict__ cnt,
                                  const double* __restrict__ sums, double* __restrict__ out3, double* __restrict__ outn3) {
  for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < K; v += gridDim.x * blockDim.x) {
    const uint32_t a = offs[v], b = offs[v + 1];
    const uint32_t s = slot_of_vox[v];
    if (a != b && cnt[s] > kSmallVoxel) continue;
    double sx, sy, sz, nx = 0.0, ny = 0.0, nz = 0.0;
    if (a == b) {
      const double* q = sums + 6 * (size_t)s;
      sx = q[0]; sy = q[1]; sz = q[2];
      if (n3) { nx = q[3]; ny = q[4]; nz = q[5]; }
    } else {
      for (uint32_t i = a + 1; i < b; ++i) {
        const int m = members[i];
        uint32_t j = i;
        while (j > a && members[j - 1] > m) { members[j] = members[j - 1]; --j; }
        members[j] = m;
      }
      sx = sy = sz = 0.0;
      for (uint32_t i = a; i < b; ++i) {
        const int m = members[i];
        const float4 q = __ldg(&p[m]);
        sx += (double)q.x; sy += (double)q.y; sz += (double)q.z;
        if (n3) { nx += (double)n3[3 * (size_t)m]; ny += (double)n3[3 * (size_t)m + 1]; nz += (double)n3[3 * (size_t)m + 2]; }
      }
    }
    const double c = (double)cnt[s];
    out3[3 * (size_t)v] = sx / c; out3[3 * (size_t)v + 1] = sy / c; out3[3 * (size_t)v + 2] = sz / c;
    if (n3) { outn3[3 * (size_t)v] = nx / c; outn3[3 * (size_t)v + 1] = ny / c; outn3[3 * (size_t)v + 2] = nz / c; }
  }
}

// ---- RandomDownSample ------------------------------------------------------------------------------------------------
// key(i) = fmix32(i ^ seed * 0x9E3779B9): fmix32 (murmur3's finaliser) is a bijection of the 32-bit integers, so the
// keys of distinct points are distinct and "the K smallest keys" is a uniformly random K-subset without replacement.
__device__ __forceinline__ uint32_t rds_key(uint32_t i, uint32_t seed) {
  uint32_t x = i ^ (seed * 0x9E3779B9u);
  x ^= x >> 16; x *= 0x85ebca6bu; x ^= x >> 13; x *= 0xc2b2ae35u; x ^= x >> 16;
  return x;
}
// st: [0] wanted rank (K - 1, then the rank inside the selected bin), [1] selected high half, [2] K-th smallest key
__global__ void rds_hist_kernel(int N, uint32_t seed, int pass, const uint32_t* __restrict__ st, uint32_t* __restrict__ hist) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const uint32_t k = rds_key((uint32_t)i, seed);
    if (pass == 0) atomicAdd(&hist[k >> 16], 1u);
    else if ((k >> 16) == st[1]) atomicAdd(&hist[65536u + (k & 0xffffu)], 1u);
  }
}
__global__ void __launch_bounds__(1024) rds_pick_kernel(int pass, uint32_t* __restrict__ st, const uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_sum[1024];
  const int tid = threadIdx.x;
  const uint32_t* hh = hist + (pass ? 65536 : 0);
  uint32_t c[64], sum = 0;
  for (int j = 0; j < 64; ++j) { c[j] = hh[tid * 64 + j]; sum += c[j]; }
  s_sum[tid] = sum;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {   // inclusive scan (Hillis-Steele; 10 steps)
    const uint32_t v = tid >= o ? s_sum[tid - o] : 0u;
    __syncthreads();
    s_sum[tid] += v;
    __syncthreads();
  }
  const uint32_t rank = st[0];
  uint32_t ex = s_sum[tid] - sum;
  __syncthreads();
  if (rank >= ex && rank < ex + sum) {
    for (int j = 0; j < 64; ++j) {
      if (rank < ex + c[j]) {
        const uint32_t bin = (uint32_t)(tid * 64 + j);
        if (pass == 0) { st[1] = bin; st[0] = rank - ex; }
        else st[2] = (st[1] << 16) | bin;
        break;
      }
      ex += c[j];
    }
  }
}
__global__ void rds_flag_kernel(int N, uint32_t seed, const uint32_t* __restrict__ st, uint32_t* __restrict__ flag) {
  const uint32_t kth = st[2];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x)
    flag[i] = rds_key((uint32_t)i, seed) <= kth ? 1u : 0u;
}

}  // namespace

// =================================================================================================================
// host side
// =================================================================================================================
struct b200_prep {
  int device = 0;
  cudaStream_t stream = nullptr;
  struct Buf {
    void* p = nullptr;
    size_t cap = 0;
  };
  enum { PTS, N3, FLAG, TILE, OUT4, OUTN3, IDX, KEYS, CNT, FIRST, SLOT, VOX, OFFS, CUR, MEMB, OUT3D, OUTN3D, FIRSTK, SCR, SUMS, EXPS, SLOTV, ORDL, NBUF };
  int last_ordered = 0;   // voxels of the last down-sample that took the ordered path
  Buf buf[NBUF];
  float last_ms = 0.f;
  cudaEvent_t ev[2] = {nullptr, nullptr};
};

namespace {

int ensure(b200_prep* h, int which, size_t bytes) {
  b200_prep::Buf& b = h->buf[which];
  if (bytes <= b.cap) return B200_OK;
  cudaFree(b.p);
  b.p = nullptr; b.cap = 0;
  const size_t want = bytes + bytes / 4 + 256;
  B200_CUDA(cudaMalloc(&b.p, want));
  b.cap = want;
  return B200_OK;
}
template <class T> T* ptr(b200_prep* h, int which) { return reinterpret_cast<T*>(h->buf[which].p); }

int upload_cloud(b200_prep* h, const float* pts4xN, const float* nrm3xN, int N) {
  B200_TRY(ensure(h, b200_prep::PTS, (size_t)N * sizeof(float4)));
  B200_CUDA(cudaMemcpyAsync(h->buf[b200_prep::PTS].p, pts4xN, (size_t)N * sizeof(float4), cudaMemcpyHostToDevice, h->stream));
  if (nrm3xN) {
    B200_TRY(ensure(h, b200_prep::N3, (size_t)N * 3 * sizeof(float)));
    B200_CUDA(cudaMemcpyAsync(h->buf[b200_prep::N3].p, nrm3xN, (size_t)N * 3 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
  }
  return B200_OK;
}

}  // namespace

extern "C" {

int b200_prep_create(b200_prep** out, int device) {
  if (!out) { set_error("b200_prep_create: bad arguments"); return B200_E_BAD_PARAM; }
  *out = nullptr;
  DeviceGuard dg(device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  b200_prep* h = new (std::nothrow) b200_prep();
  if (!h) { set_error("out of host memory"); return B200_E_CUDA; }
  h->device = device;
  cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreate(&h->ev[0]);
  if (e == cudaSuccess) e = cudaEventCreate(&h->ev[1]);
  if (e != cudaSuccess) { const int st = cuda_fail(e, "b200_prep_create", __FILE__, __LINE__); b200_prep_destroy(h); return st; }
  *out = h;
  return B200_OK;
}

void b200_prep_destroy(b200_prep* h) {
  if (!h) return;
  {
    DeviceGuard dg(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (int i = 0; i < b200_prep::NBUF; ++i) cudaFree(h->buf[i].p);
    for (int i = 0; i < 2; ++i)
      if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
}

int b200_prep_crop(b200_prep* h, const b200_crop_params* prm, const float* pts4xN, const float* nrm3xN, int N,
                   float* out4xN, float* outn3xN, int32_t* kept_idx, int* n_out) {
  if (!h || !prm || N < 0 || (N > 0 && !pts4xN) || !n_out || prm->kind < B200_CROP_NONE || prm->kind > B200_CROP_MINMAX_RADIUS) {
    set_error("b200_prep_crop: bad arguments");
    return B200_E_BAD_PARAM;
  }
  *n_out = 0;
  if (N == 0) return B200_OK;
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  B200_TRY(upload_cloud(h, pts4xN, nrm3xN, N));
  B200_TRY(ensure(h, b200_prep::FLAG, ((size_t)N + 2) * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::TILE, std::max<size_t>(1, scan_tiles_for((size_t)N)) * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::OUT4, (size_t)N * sizeof(float4)));
  B200_TRY(ensure(h, b200_prep::IDX, (size_t)N * sizeof(int)));
  if (nrm3xN) B200_TRY(ensure(h, b200_prep::OUTN3, (size_t)N * 3 * sizeof(float)));
  CropP c{prm->kind, prm->invert ? 1 : 0, prm->radius_min, prm->radius_max, prm->min_z, prm->max_z,
          prm->center[0], prm->center[1], prm->center[2]};
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  uint32_t* flag = ptr<uint32_t>(h, b200_prep::FLAG);
  crop_flag_kernel<<<blocks_for(N), kThreads, 0, s>>>(ptr<float4>(h, b200_prep::PTS), N, c, flag);
  B200_TRY(inclusive_scan_u32(flag, (size_t)N, ptr<uint32_t>(h, b200_prep::TILE), s));
  crop_scatter_kernel<<<blocks_for(N), kThreads, 0, s>>>(ptr<float4>(h, b200_prep::PTS),
                                                         nrm3xN ? ptr<float>(h, b200_prep::N3) : nullptr, N, flag,
                                                         ptr<float4>(h, b200_prep::OUT4),
                                                         nrm3xN ? ptr<float>(h, b200_prep::OUTN3) : nullptr,
                                                         ptr<int>(h, b200_prep::IDX));
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  uint32_t K = 0;
  B200_CUDA(cudaMemcpyAsync(&K, flag + (N - 1), sizeof(K), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  if (K > 0) {
    if (out4xN) B200_CUDA(cudaMemcpyAsync(out4xN, h->buf[b200_prep::OUT4].p, (size_t)K * sizeof(float4), cudaMemcpyDeviceToHost, s));
    if (outn3xN && nrm3xN) B200_CUDA(cudaMemcpyAsync(outn3xN, h->buf[b200_prep::OUTN3].p, (size_t)K * 3 * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (kept_idx) B200_CUDA(cudaMemcpyAsync(kept_idx, h->buf[b200_prep::IDX].p, (size_t)K * sizeof(int), cudaMemcpyDeviceToHost, s));
    B200_CUDA(cudaStreamSynchronize(s));
  }
  B200_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]));
  *n_out = (int)K;
  return B200_OK;
}

// o3d_slam::randomDownSample (helpers.cpp:100-106) -> open3d PointCloud::RandomDownSample(ratio) (v0.15.1, not in the
// reference tree: sample_size = size_t(ratio * n) indices drawn without replacement by shuffling with an mt19937 seeded
// from std::random_device — the reference's own output is not reproducible). Here: the sample_size points with the
// smallest keys of a bijective hash of (index, seed), in input order.
int b200_prep_random_downsample(b200_prep* h, double ratio, uint32_t seed, const float* pts4xN, const float* nrm3xN, int N,
                                float* out4xN, float* outn3xN, int32_t* kept_idx, int* n_out) {
  if (!h || N < 0 || (N > 0 && !pts4xN) || !n_out) { set_error("b200_prep_random_downsample: bad arguments"); return B200_E_BAD_PARAM; }
  if (!(ratio > 0.0)) { set_error("RandomDownSample: sampling_ratio must lie in (0, 1]"); return B200_E_BAD_PARAM; }
  *n_out = 0;
  if (N == 0) return B200_OK;
  if (ratio >= 1.0) {   // "if (downSamplingRatio >= 1.0) return" (helpers.cpp:101-103): the cloud is left as it is
    if (out4xN) std::memcpy(out4xN, pts4xN, (size_t)N * 4 * sizeof(float));
    if (outn3xN && nrm3xN) std::memcpy(outn3xN, nrm3xN, (size_t)N * 3 * sizeof(float));
    if (kept_idx) for (int i = 0; i < N; ++i) kept_idx[i] = i;
    *n_out = N;
    h->last_ms = 0.f;
    return B200_OK;
  }
  const size_t K = (size_t)(ratio * (double)N);
  if (K == 0) return B200_OK;
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  B200_TRY(upload_cloud(h, pts4xN, nrm3xN, N));
  B200_TRY(ensure(h, b200_prep::FLAG, ((size_t)N + 2) * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::TILE, std::max<size_t>(1, scan_tiles_for((size_t)N)) * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::OUT4, (size_t)N * sizeof(float4)));
  B200_TRY(ensure(h, b200_prep::IDX, (size_t)N * sizeof(int)));
  B200_TRY(ensure(h, b200_prep::SCR, (2 * 65536 + 4) * sizeof(uint32_t)));
  if (nrm3xN) B200_TRY(ensure(h, b200_prep::OUTN3, (size_t)N * 3 * sizeof(float)));
  uint32_t* hist = ptr<uint32_t>(h, b200_prep::SCR);
  uint32_t* st = hist + 2 * 65536;
  const uint32_t st0[4] = {(uint32_t)(K - 1), 0u, 0u, 0u};
  B200_CUDA(cudaMemsetAsync(hist, 0, 2 * 65536 * sizeof(uint32_t), s));
  B200_CUDA(cudaMemcpyAsync(st, st0, sizeof(st0), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  uint32_t* flag = ptr<uint32_t>(h, b200_prep::FLAG);
  for (int pass = 0; pass < 2; ++pass) {
    rds_hist_kernel<<<blocks_for(N), kThreads, 0, s>>>(N, seed, pass, st, hist);
    rds_pick_kernel<<<1, 1024, 0, s>>>(pass, st, hist);
  }
  rds_flag_kernel<<<blocks_for(N), kThreads, 0, s>>>(N, seed, st, flag);
  B200_TRY(inclusive_scan_u32(flag, (size_t)N, ptr<uint32_t>(h, b200_prep::TILE), s));
  crop_scatter_kernel<<<blocks_for(N), kThreads, 0, s>>>(ptr<float4>(h, b200_prep::PTS),
                                                         nrm3xN ? ptr<float>(h, b200_prep::N3) : nullptr, N, flag,
                                                         ptr<float4>(h, b200_prep::OUT4),
                                                         nrm3xN ? ptr<float>(h, b200_prep::OUTN3) : nullptr,
                                                         ptr<int>(h, b200_prep::IDX));
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  if (out4xN) B200_CUDA(cudaMemcpyAsync(out4xN, h->buf[b200_prep::OUT4].p, K * sizeof(float4), cudaMemcpyDeviceToHost, s));
  if (outn3xN && nrm3xN) B200_CUDA(cudaMemcpyAsync(outn3xN, h->buf[b200_prep::OUTN3].p, K * 3 * sizeof(float), cudaMemcpyDeviceToHost, s));
  if (kept_idx) B200_CUDA(cudaMemcpyAsync(kept_idx, h->buf[b200_prep::IDX].p, K * sizeof(int), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  B200_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]));
  *n_out = (int)K;
  return B200_OK;
}

int b200_prep_voxel_downsample(b200_prep* h, double voxel_size, const float* pts4xN, const float* nrm3xN, int N,
                               double* out3xK, double* outn3xK, int32_t* first_idx, int* n_out) {
  if (!h || N < 0 || (N > 0 && (!pts4xN || !out3xK)) || !n_out) { set_error("b200_prep_voxel_downsample: bad arguments"); return B200_E_BAD_PARAM; }
  *n_out = 0;
  if (!(voxel_size > 0.0)) { set_error("VoxelDownSample: voxel_size <= 0"); return B200_E_BAD_PARAM; }
  if (N == 0) return B200_OK;
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  B200_TRY(upload_cloud(h, pts4xN, nrm3xN, N));
  size_t T = 1024;
  while (T < 2 * (size_t)N) T <<= 1;
  B200_TRY(ensure(h, b200_prep::SCR, 64));
  B200_TRY(ensure(h, b200_prep::FLAG, ((size_t)N + 2) * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::TILE, std::max<size_t>(1, scan_tiles_for((size_t)N + 2)) * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::KEYS, T * sizeof(unsigned long long)));
  B200_TRY(ensure(h, b200_prep::CNT, T * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::FIRST, T * sizeof(int)));
  B200_TRY(ensure(h, b200_prep::VOX, T * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::SUMS, T * 6 * sizeof(double)));
  B200_TRY(ensure(h, b200_prep::EXPS, T * 4 * sizeof(int)));
  B200_TRY(ensure(h, b200_prep::SLOTV, (size_t)N * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::ORDL, (size_t)N * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::SLOT, (size_t)N * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::OFFS, ((size_t)N + 2) * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::CUR, (size_t)N * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_prep::MEMB, (size_t)N * sizeof(int)));
  B200_TRY(ensure(h, b200_prep::FIRSTK, (size_t)N * sizeof(int)));
  B200_TRY(ensure(h, b200_prep::OUT3D, (size_t)N * 3 * sizeof(double)));
  if (nrm3xN) B200_TRY(ensure(h, b200_prep::OUTN3D, (size_t)N * 3 * sizeof(double)));
  const float4* d_pts = ptr<float4>(h, b200_prep::PTS);
  const float* d_n3 = nrm3xN ? ptr<float>(h, b200_prep::N3) : nullptr;
  uint32_t* scr = ptr<uint32_t>(h, b200_prep::SCR);
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  // ---- bounds: voxel_min_bound = min_bound - voxel_size / 2
  uint32_t init[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
  B200_CUDA(cudaMemcpyAsync(scr, init, sizeof(init), cudaMemcpyHostToDevice, s));
  minmax_kernel<<<blocks_for(N), kThreads, 0, s>>>(d_pts, N, scr);
  B200_CUDA(cudaGetLastError());
  uint32_t mm[6];
  B200_CUDA(cudaMemcpyAsync(mm, scr, sizeof(mm), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  double mn[3], mx[3], ext = 0.0;
  for (int d = 0; d < 3; ++d) {
    mn[d] = (double)ordered_uint_as_float(mm[d]); mx[d] = (double)ordered_uint_as_float(mm[3 + d]);
    ext = std::max(ext, mx[d] - mn[d]);
  }
  if (!(ext == ext) || voxel_size * 2147483647.0 < ext) { set_error("VoxelDownSample: voxel_size is too small"); return B200_E_BAD_PARAM; }
  VoxP v{mn[0] - voxel_size * 0.5, mn[1] - voxel_size * 0.5, mn[2] - voxel_size * 0.5, voxel_size, (unsigned long long)T - 1ull};
  // ---- hash insert, voxel numbering by first member
  B200_CUDA(cudaMemsetAsync(h->buf[b200_prep::KEYS].p, 0xff, T * sizeof(unsigned long long), s));
  B200_CUDA(cudaMemsetAsync(h->buf[b200_prep::CNT].p, 0, T * sizeof(uint32_t), s));
  B200_CUDA(cudaMemsetAsync(h->buf[b200_prep::FIRST].p, 0x7f, T * sizeof(int), s));   // 0x7f7f7f7f > any index
  B200_CUDA(cudaMemsetAsync(h->buf[b200_prep::FLAG].p, 0, ((size_t)N + 2) * sizeof(uint32_t), s));
  B200_CUDA(cudaMemsetAsync(scr + 8, 0, 3 * sizeof(int), s));
  B200_CUDA(cudaMemsetAsync(h->buf[b200_prep::SUMS].p, 0, T * 6 * sizeof(double), s));
  exps_init_kernel<<<blocks_for(T), kThreads, 0, s>>>(ptr<int>(h, b200_prep::EXPS), T);
  unsigned long long* keys = ptr<unsigned long long>(h, b200_prep::KEYS);
  uint32_t* cnt = ptr<uint32_t>(h, b200_prep::CNT);
  int* first = ptr<int>(h, b200_prep::FIRST);
  uint32_t* flag = ptr<uint32_t>(h, b200_prep::FLAG);
  uint32_t* offs = ptr<uint32_t>(h, b200_prep::OFFS);
  voxel_insert_kernel<<<blocks_for(N), kThreads, 0, s>>>(d_pts, d_n3, N, v, keys, cnt, first, ptr<uint32_t>(h, b200_prep::SLOT),
                                                         ptr<double>(h, b200_prep::SUMS), ptr<int>(h, b200_prep::EXPS),
                                                         reinterpret_cast<int*>(scr + 8));
  voxel_mark_kernel<<<blocks_for(T), kThreads, 0, s>>>(keys, first, T, flag);
  B200_TRY(inclusive_scan_u32(flag, (size_t)N, ptr<uint32_t>(h, b200_prep::TILE), s));
  uint32_t K = 0;
  int err = 0;
  B200_CUDA(cudaMemcpyAsync(&K, flag + (N - 1), sizeof(K), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaMemcpyAsync(&err, scr + 8, sizeof(err), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  if (err) { set_error("VoxelDownSample: a coordinate is not finite or the cloud spans more than 2^21 voxels along an axis"); return B200_E_BAD_PARAM; }
  B200_CUDA(cudaMemsetAsync(offs, 0, ((size_t)K + 1) * sizeof(uint32_t), s));
  voxel_number_kernel<<<blocks_for(T), kThreads, 0, s>>>(keys, first, cnt, ptr<int>(h, b200_prep::EXPS), nrm3xN ? 1 : 0, T, flag,
                                                         ptr<uint32_t>(h, b200_prep::VOX), ptr<uint32_t>(h, b200_prep::SLOTV),
                                                         offs, ptr<int>(h, b200_prep::FIRSTK), reinterpret_cast<int*>(scr + 9),
                                                         ptr<uint32_t>(h, b200_prep::ORDL));
  B200_TRY(inclusive_scan_u32(offs, (size_t)K + 1, ptr<uint32_t>(h, b200_prep::TILE), s));
  int n_ord[2] = {0, 0};   // voxels on the ordered path, and how many of them are large
  B200_CUDA(cudaMemcpyAsync(n_ord, scr + 9, sizeof(n_ord), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  if (n_ord[0] > n_ord[1]) {
    B200_CUDA(cudaMemsetAsync(h->buf[b200_prep::CUR].p, 0, (size_t)K * sizeof(uint32_t), s));
    voxel_members_kernel<<<blocks_for(N), kThreads, 0, s>>>(ptr<uint32_t>(h, b200_prep::SLOT), N, ptr<uint32_t>(h, b200_prep::VOX),
                                                            cnt, offs, ptr<uint32_t>(h, b200_prep::CUR), ptr<int>(h, b200_prep::MEMB));
  }
  voxel_mean_kernel<<<blocks_for(K), kThreads, 0, s>>>(d_pts, d_n3, (int)K, offs, ptr<int>(h, b200_prep::MEMB),
                                                       ptr<uint32_t>(h, b200_prep::SLOTV), cnt, ptr<double>(h, b200_prep::SUMS),
                                                       ptr<double>(h, b200_prep::OUT3D),
                                                       nrm3xN ? ptr<double>(h, b200_prep::OUTN3D) : nullptr);
  if (n_ord[1] > 0)
    voxel_ordered_kernel<<<n_ord[1], kThreads, 0, s>>>(d_pts, d_n3, N, ptr<uint32_t>(h, b200_prep::ORDL),
                                                       ptr<uint32_t>(h, b200_prep::SLOTV), ptr<uint32_t>(h, b200_prep::SLOT), offs,
                                                       ptr<int>(h, b200_prep::MEMB), ptr<double>(h, b200_prep::OUT3D),
                                                       nrm3xN ? ptr<double>(h, b200_prep::OUTN3D) : nullptr);
  h->last_ordered = n_ord[0];
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  B200_CUDA(cudaMemcpyAsync(out3xK, h->buf[b200_prep::OUT3D].p, (size_t)K * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (outn3xK && nrm3xN) B200_CUDA(cudaMemcpyAsync(outn3xK, h->buf[b200_prep::OUTN3D].p, (size_t)K * 3 * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (first_idx) B200_CUDA(cudaMemcpyAsync(first_idx, h->buf[b200_prep::FIRSTK].p, (size_t)K * sizeof(int), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  B200_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]));
  *n_out = (int)K;
  return B200_OK;
}

int b200_prep_last_ordered_voxels(const b200_prep* h, int* n) {
  if (!h || !n) { set_error("b200_prep_last_ordered_voxels: bad arguments"); return B200_E_BAD_PARAM; }
  *n = h->last_ordered;
  return B200_OK;
}

int b200_prep_last_device_ms(const b200_prep* h, float* ms) {
  if (!h || !ms) { set_error("b200_prep_last_device_ms: bad arguments"); return B200_E_BAD_PARAM; }
  *ms = h->last_ms;
  return B200_OK;
}

}  // extern "C"
