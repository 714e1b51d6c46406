// Device-side construction of the uniform-grid map index (counting sort by cell: histogram -> scan -> scatter).
// Replaces KDTreeMatcher::init / Nabo::NNS::create (libpointmatcher/pointmatcher/MatchersImpl.cpp:78-83).
// All passes are HBM-bound streaming kernels; no library sort/scan is used.
#include "grid_index.cuh"
#include "modules.cuh"
#include "knn.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace b200 {

namespace {

constexpr int kThreads = 256;
constexpr int kScanThreads = 1024;
constexpr int kScanItems = 4;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ uint32_t float_as_ordered_uint(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

inline int grid_for(size_t n, int threads, int device_sms = 148) {
  size_t b = (n + threads - 1) / threads;
  const size_t cap = (size_t)device_sms * 16;
  return (int)std::max<size_t>(1, std::min(b, cap));
}

// ---- host 4xN -> float4 is a straight copy; normals 3xN -> float4 needs a repack -------------------------------
__global__ void pack_normals_kernel(const float* __restrict__ n3, float4* __restrict__ out, int N) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x)
    out[i] = make_float4(n3[3 * (size_t)i], n3[3 * (size_t)i + 1], n3[3 * (size_t)i + 2], 0.f);
}

// ---- bounding box over finite coordinates: 6 ordered-uint atomics per BLOCK (one per warp made the six addresses the
// bottleneck: 87 us for 5 M points, 20 times the time of reading them)
__global__ void bbox_kernel(const float4* __restrict__ p, int N, unsigned long long* __restrict__ scratch) {
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const float4 v = __ldg(&p[i]);
    const float c[3] = {v.x, v.y, v.z};
#pragma unroll
    for (int d = 0; d < 3; ++d)
      if (isfinite(c[d])) { mn[d] = fminf(mn[d], c[d]); mx[d] = fmaxf(mx[d], c[d]); }
  }
  __shared__ float sm[6][kThreads / 32];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    for (int o = 16; o > 0; o >>= 1) {
      mn[d] = fminf(mn[d], __shfl_xor_sync(0xffffffffu, mn[d], o));
      mx[d] = fmaxf(mx[d], __shfl_xor_sync(0xffffffffu, mx[d], o));
    }
    if ((threadIdx.x & 31) == 0) { sm[d][threadIdx.x >> 5] = mn[d]; sm[3 + d][threadIdx.x >> 5] = mx[d]; }
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    const int d = threadIdx.x;
    float v = sm[d][0];
    for (int w = 1; w < kThreads / 32; ++w) v = d < 3 ? fminf(v, sm[d][w]) : fmaxf(v, sm[d][w]);
    uint32_t* s = reinterpret_cast<uint32_t*>(scratch);
    if (d < 3) atomicMin(&s[d], float_as_ordered_uint(v));
    else atomicMax(&s[d], float_as_ordered_uint(v));
  }
}

struct GridDims {
  float ox, oy, oz, inv_cell;
  int nx, ny, nz;
};

__device__ __forceinline__ size_t cell_index(const GridDims& g, float x, float y, float z) {
  const int cx = cell_of(cell_coord(x, g.ox, g.inv_cell), g.nx);
  const int cy = cell_of(cell_coord(y, g.oy, g.inv_cell), g.ny);
  const int cz = cell_of(cell_coord(z, g.oz, g.inv_cell), g.nz);
  return (size_t)cx + (size_t)g.nx * ((size_t)cy + (size_t)g.ny * (size_t)cz);
}

// A[c + 1] += 1 per point; counts cells that become occupied
__global__ void hist_kernel(const float4* __restrict__ p, int N, GridDims g, uint32_t* __restrict__ A,
                            unsigned long long* __restrict__ occupied) {
  unsigned occ = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const float4 v = __ldg(&p[i]);
    const size_t c = cell_index(g, v.x, v.y, v.z);
    if (atomicAdd(&A[c + 1], 1u) == 0u) ++occ;
  }
  for (int o = 16; o > 0; o >>= 1) occ += __shfl_xor_sync(0xffffffffu, occ, o);
  if ((threadIdx.x & 31) == 0 && occ) atomicAdd(occupied, (unsigned long long)occ);
}

// ---- inclusive scan of A[0..n) in place: tile sums -> scan of sums -> apply ---------------------------------
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* smem, uint32_t& total) {
  // 1024 threads: warp scan + scan of warp totals
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) smem[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    uint32_t w = (lane < (int)(blockDim.x >> 5)) ? smem[lane] : 0u;
    uint32_t winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    smem[lane] = winc - w;          // exclusive warp offsets
    if (lane == 31) smem[32] = winc;  // block total
  }
  __syncthreads();
  const uint32_t res = smem[wid] + inc - v;
  total = smem[32];
  __syncthreads();
  return res;
}

__global__ void __launch_bounds__(kScanThreads) scan_tile_sums_kernel(const uint32_t* __restrict__ A, size_t n,
                                                                      uint32_t* __restrict__ tile_sums) {
  __shared__ uint32_t smem[33];
  const size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
  uint32_t s = 0;
#pragma unroll
  for (int j = 0; j < kScanItems; ++j)
    if (base + j < n) s += A[base + j];
  uint32_t total;
  block_exclusive_scan(s, smem, total);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kScanThreads) scan_sums_kernel(uint32_t* __restrict__ tile_sums, size_t ntiles) {
  __shared__ uint32_t smem[33];
  uint32_t carry = 0;
  for (size_t base = 0; base < ntiles; base += kScanThreads) {
    const size_t i = base + threadIdx.x;
    const uint32_t v = i < ntiles ? tile_sums[i] : 0u;
    uint32_t total;
    const uint32_t ex = block_exclusive_scan(v, smem, total);
    if (i < ntiles) tile_sums[i] = carry + ex;  // exclusive offset of tile i
    carry += total;
  }
}

__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(uint32_t* __restrict__ A, size_t n,
                                                                  const uint32_t* __restrict__ tile_offsets) {
  __shared__ uint32_t smem[33];
  const size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
  uint32_t v[kScanItems];
  uint32_t s = 0;
#pragma unroll
  for (int j = 0; j < kScanItems; ++j) {
    v[j] = (base + j < n) ? A[base + j] : 0u;
    s += v[j];
  }
  uint32_t total;
  uint32_t run = block_exclusive_scan(s, smem, total) + tile_offsets[blockIdx.x];
#pragma unroll
  for (int j = 0; j < kScanItems; ++j) {
    run += v[j];
    if (base + j < n) A[base + j] = run;  // inclusive
  }
}

// after the inclusive scan A[c + 1] = end(c); each point takes the slot atomicSub(...) - 1, so A[c + 1] ends as
// start(c) and G = A + 1 satisfies G[c] = start(c), G[c + 1] = end(c).
__global__ void scatter_kernel(const float4* __restrict__ p, const float4* __restrict__ nrm, int N, GridDims g,
                               uint32_t* __restrict__ A, float4* __restrict__ out_p, float4* __restrict__ out_n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const float4 v = __ldg(&p[i]);
    const size_t c = cell_index(g, v.x, v.y, v.z);
    const uint32_t pos = atomicSub(&A[c + 1], 1u) - 1u;
    out_p[pos] = make_float4(v.x, v.y, v.z, __int_as_float(i));
    if (nrm) out_n[pos] = __ldg(&nrm[i]);
  }
}

__global__ void set_u32_kernel(uint32_t* p, uint32_t v) { *p = v; }

// ---- mean / subtract -------------------------------------------------------------------------------------------
__global__ void sum3_kernel(const float4* __restrict__ p, int N, double* __restrict__ partials) {
  double s[3] = {0.0, 0.0, 0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const float4 v = __ldg(&p[i]);
    s[0] += (double)v.x; s[1] += (double)v.y; s[2] += (double)v.z;
  }
  __shared__ double sm[3][kThreads / 32];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    for (int o = 16; o > 0; o >>= 1) s[d] += __shfl_xor_sync(0xffffffffu, s[d], o);
    if ((threadIdx.x & 31) == 0) sm[d][threadIdx.x >> 5] = s[d];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int w = 0; w < kThreads / 32; ++w) t += sm[threadIdx.x][w];
    partials[3 * (size_t)blockIdx.x + threadIdx.x] = t;
  }
}

__global__ void subtract3_kernel(float4* __restrict__ p, int N, float mx, float my, float mz) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    float4 v = p[i];
    v.x = __fsub_rn(v.x, mx); v.y = __fsub_rn(v.y, my); v.z = __fsub_rn(v.z, mz);
    p[i] = v;
  }
}

}  // namespace

// =================================================================================================================
// host side
// =================================================================================================================

void MapIndex::release() {
  DeviceGuard dg(device);
  cudaFree(d_ref); cudaFree(d_nrm_orig); cudaFree(d_pts); cudaFree(d_nrm); cudaFree(d_A); cudaFree(d_tile);
  cudaFree(d_scratch); cudaFree(d_sn_list); cudaFree(d_sn_ids);
  d_sn_list = d_sn_ids = nullptr;
  cap_sn_list = cap_sn_ids = 0;
  d_ref = d_nrm_orig = d_pts = d_nrm = nullptr;
  d_A = d_tile = nullptr;
  d_scratch = nullptr;
  cap_cells = cap_pts = cap_tiles = 0;
  M = 0;
}

int upload_features(float4* d_dst, const float* h_feat4xN, int N, cudaStream_t s) {
  if (N <= 0) return B200_OK;
  B200_CUDA(cudaMemcpyAsync(d_dst, h_feat4xN, (size_t)N * sizeof(float4), cudaMemcpyHostToDevice, s));
  return B200_OK;
}

int upload_normals(float4* d_dst, const float* h_nrm3xN, int N, float* d_stage, cudaStream_t s) {
  if (N <= 0) return B200_OK;
  B200_CUDA(cudaMemcpyAsync(d_stage, h_nrm3xN, (size_t)N * 3 * sizeof(float), cudaMemcpyHostToDevice, s));
  pack_normals_kernel<<<grid_for(N, kThreads), kThreads, 0, s>>>(d_stage, d_dst, N);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int index_alloc_points(MapIndex& idx, int M, bool with_normals) {
  if ((size_t)M > idx.cap_pts) {
    // A handle that is re-initialised with clouds of varying size (the map patch cut around the robot at every
    // reference update, a scan indexed per registration) grows with a quarter of headroom, so that a patch a few
    // points larger than the largest so far does not cost four cudaFree + four cudaMalloc (milliseconds, and a device
    // synchronisation) on the per-scan path; the first allocation is exact (a 5 M-point map is loaded once).
    const size_t want = idx.cap_pts ? (size_t)M + (size_t)M / 4 : (size_t)M;
    cudaFree(idx.d_ref); cudaFree(idx.d_pts); cudaFree(idx.d_nrm_orig); cudaFree(idx.d_nrm);
    idx.d_ref = idx.d_pts = idx.d_nrm_orig = idx.d_nrm = nullptr;
    idx.cap_pts = 0;
    const size_t bytes = std::max<size_t>(1, want) * sizeof(float4);
    B200_CUDA(cudaMalloc(&idx.d_ref, bytes));
    B200_CUDA(cudaMalloc(&idx.d_pts, bytes));
    idx.cap_pts = want;
  }
  if (with_normals && !idx.d_nrm_orig) {
    const size_t bytes = std::max<size_t>(1, idx.cap_pts) * sizeof(float4);
    B200_CUDA(cudaMalloc(&idx.d_nrm_orig, bytes));
    B200_CUDA(cudaMalloc(&idx.d_nrm, bytes));
  }
  if (!idx.d_scratch) B200_CUDA(cudaMalloc(&idx.d_scratch, 8 * sizeof(unsigned long long)));
  idx.M = M;
  return B200_OK;
}

namespace {

struct Extents {
  float mn[3], mx[3];
  bool valid;
};

int compute_bbox(MapIndex& idx, Extents& e) {
  uint32_t init[16];
  std::memset(init, 0, sizeof(init));
  init[0] = init[1] = init[2] = 0xffffffffu;
  B200_CUDA(cudaMemcpyAsync(idx.d_scratch, init, sizeof(init), cudaMemcpyHostToDevice, idx.stream));
  bbox_kernel<<<grid_for(idx.M, kThreads), kThreads, 0, idx.stream>>>(idx.d_ref, idx.M, idx.d_scratch);
  B200_CUDA(cudaGetLastError());
  uint32_t out[6];
  B200_CUDA(cudaMemcpyAsync(out, idx.d_scratch, sizeof(out), cudaMemcpyDeviceToHost, idx.stream));
  B200_CUDA(cudaStreamSynchronize(idx.stream));
  e.valid = true;
  for (int d = 0; d < 3; ++d) {
    e.mn[d] = ordered_uint_as_float(out[d]);
    e.mx[d] = ordered_uint_as_float(out[3 + d]);
    if (!(std::isfinite(e.mn[d]) && std::isfinite(e.mx[d]) && e.mn[d] <= e.mx[d])) e.valid = false;
  }
  return B200_OK;
}

// choose dims for a cell size; returns false when the table would exceed max_cells
bool dims_for(const Extents& e, float cell, uint64_t max_cells, MapIndex& idx) {
  const float inv = 1.0f / cell;
  // one cell of margin on the low side, so every point has cell coordinate >= 1 - tiny and < n - 1
  double n[3];
  for (int d = 0; d < 3; ++d) n[d] = std::floor(((double)e.mx[d] - (double)e.mn[d]) * (double)inv) + 3.0;
  if (n[0] * n[1] * n[2] > (double)max_cells || n[0] > 2.0e9 || n[1] > 2.0e9 || n[2] > 2.0e9) return false;
  idx.cell = cell;
  idx.inv_cell = inv;
  idx.ox = e.mn[0] - cell; idx.oy = e.mn[1] - cell; idx.oz = e.mn[2] - cell;
  idx.nx = (int)n[0]; idx.ny = (int)n[1]; idx.nz = (int)n[2];
  return true;
}

int ensure_table(MapIndex& idx) {
  const size_t need = idx.ncells() + 2;
  if (need > idx.cap_cells) {
    const size_t want = idx.cap_cells ? need + need / 4 : need;   // same headroom rule as the point buffers
    cudaFree(idx.d_A); cudaFree(idx.d_tile);
    idx.d_A = idx.d_tile = nullptr;
    idx.cap_cells = 0;
    B200_CUDA(cudaMalloc(&idx.d_A, want * sizeof(uint32_t)));
    const size_t tiles = (want + kScanTile - 1) / kScanTile;
    B200_CUDA(cudaMalloc(&idx.d_tile, std::max<size_t>(1, tiles) * sizeof(uint32_t)));
    idx.cap_cells = want;
    idx.cap_tiles = tiles;
  }
  return B200_OK;
}

int histogram_pass(MapIndex& idx, int64_t& occupied) {
  B200_TRY(ensure_table(idx));
  const size_t n = idx.ncells() + 2;
  B200_CUDA(cudaMemsetAsync(idx.d_A, 0, n * sizeof(uint32_t), idx.stream));
  B200_CUDA(cudaMemsetAsync(idx.d_scratch + 4, 0, sizeof(unsigned long long), idx.stream));
  GridDims g{idx.ox, idx.oy, idx.oz, idx.inv_cell, idx.nx, idx.ny, idx.nz};
  hist_kernel<<<grid_for(idx.M, kThreads), kThreads, 0, idx.stream>>>(idx.d_ref, idx.M, g, idx.d_A,
                                                                      idx.d_scratch + 4);
  B200_CUDA(cudaGetLastError());
  unsigned long long occ = 0;
  B200_CUDA(cudaMemcpyAsync(&occ, idx.d_scratch + 4, sizeof(occ), cudaMemcpyDeviceToHost, idx.stream));
  B200_CUDA(cudaStreamSynchronize(idx.stream));
  occupied = (int64_t)occ;
  return B200_OK;
}

}  // namespace

int index_build(MapIndex& idx, const b200_index_opts* opts) {
  DeviceGuard dg(idx.device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  const float target = (opts && opts->target_per_cell > 0.f) ? opts->target_per_cell : 3.0f;
  // default cap on the dense table: 2^29 cells, and no more than 64 cells per point (>= 2^22) — a small cloud in a
  // large bounding box (a raw scan) must not pay for clearing and scanning a table a thousand times its own size
  const uint64_t auto_cap = std::min<uint64_t>((uint64_t)1 << 29, std::max<uint64_t>((uint64_t)1 << 22, 64ull * (uint64_t)std::max(idx.M, 1)));
  const uint64_t max_cells = (opts && opts->max_cells > 0) ? opts->max_cells : auto_cap;
  cudaEvent_t e0, e1;
  B200_CUDA(cudaEventCreate(&e0));
  B200_CUDA(cudaEventCreate(&e1));
  B200_CUDA(cudaEventRecord(e0, idx.stream));
  idx.occupied = 0;
  if (idx.M == 0) {
    idx.nx = idx.ny = idx.nz = 1;
    idx.cell = idx.inv_cell = 1.f;
    idx.ox = idx.oy = idx.oz = 0.f;
    B200_TRY(ensure_table(idx));
    B200_CUDA(cudaMemsetAsync(idx.d_A, 0, 3 * sizeof(uint32_t), idx.stream));
  } else {
    Extents e;
    B200_TRY(compute_bbox(idx, e));
    if (!e.valid) {  // no finite point at all: a single cell holds everything (they can never match)
      for (int d = 0; d < 3; ++d) { e.mn[d] = 0.f; e.mx[d] = 0.f; }
    }
    const float ext = std::max(e.mx[0] - e.mn[0], std::max(e.mx[1] - e.mn[1], e.mx[2] - e.mn[2]));
    float amax = 0.f;
    for (int d = 0; d < 3; ++d) amax = std::max(amax, std::max(std::fabs(e.mn[d]), std::fabs(e.mx[d])));
    const float min_cell = std::max(amax * 1.6e-5f, 1e-30f);  // keep >= 2^-16 of the coordinate magnitude
    float cell;
    int64_t occ = 0;
    if (opts && opts->cell_size > 0.f) {
      cell = std::max(opts->cell_size, min_cell);
      while (!dims_for(e, cell, max_cells, idx)) cell *= 1.26f;
      B200_TRY(histogram_pass(idx, occ));
    } else {
      // surfaces: points per occupied cell ~ M c^2 / area; start from the largest extent, refine from the measured
      // occupancy (2-D scaling law), at most 4 histogram passes
      cell = ext > 0.f ? ext * std::sqrt(target / (float)idx.M) : 1.f;
      cell = std::max(cell, min_cell);
      bool capped = false;   // the cell is the finest one the table cap allows: nothing finer to try
      for (int pass = 0; pass < 4; ++pass) {
        while (!dims_for(e, cell, max_cells, idx)) cell *= 1.26f;
        cell = idx.cell;
        B200_TRY(histogram_pass(idx, occ));
        if (capped) break;
        const float per = (float)idx.M / (float)std::max<int64_t>(1, occ);
        if (per > target / 1.5f && per < target * 1.5f) break;
        float next = cell * std::sqrt(target / per);
        next = std::max(next, min_cell);
        if (std::fabs(next - cell) < 0.05f * cell || pass == 3) break;
        // Would the finer table still fit? If not, go straight to the finest one that does and stop there (a raw scan:
        // a few dense square metres around the sensor inside a bounding box of hectares — keeping the coarse cell left
        // 28 points per occupied cell and 614 candidates per k = 20 query where the finest feasible table has ~150).
        MapIndex probe = idx;
        if (next < cell && !dims_for(e, next, max_cells, probe)) {
          float lo = next, hi = cell;   // lo does not fit, hi does
          for (int it = 0; it < 16; ++it) {
            const float mid = std::sqrt(lo * hi);
            if (dims_for(e, mid, max_cells, probe)) hi = mid; else lo = mid;
          }
          if (!(hi < 0.95f * cell)) break;
          next = hi;
          capped = true;
        }
        cell = next;
      }
    }
    idx.occupied = occ;
    // inclusive scan of A[0 .. ncells] (A[c+1] holds count(c))
    const size_t n = idx.ncells() + 1;
    const size_t tiles = (n + kScanTile - 1) / kScanTile;
    scan_tile_sums_kernel<<<(unsigned)tiles, kScanThreads, 0, idx.stream>>>(idx.d_A, n, idx.d_tile);
    scan_sums_kernel<<<1, kScanThreads, 0, idx.stream>>>(idx.d_tile, tiles);
    scan_apply_kernel<<<(unsigned)tiles, kScanThreads, 0, idx.stream>>>(idx.d_A, n, idx.d_tile);
    B200_CUDA(cudaGetLastError());
    GridDims g{idx.ox, idx.oy, idx.oz, idx.inv_cell, idx.nx, idx.ny, idx.nz};
    scatter_kernel<<<grid_for(idx.M, kThreads), kThreads, 0, idx.stream>>>(idx.d_ref, idx.d_nrm_orig, idx.M, g,
                                                                         idx.d_A, idx.d_pts, idx.d_nrm);
    set_u32_kernel<<<1, 1, 0, idx.stream>>>(idx.d_A + idx.ncells() + 1, (uint32_t)idx.M);
    B200_CUDA(cudaGetLastError());
  }
  B200_CUDA(cudaEventRecord(e1, idx.stream));
  B200_CUDA(cudaStreamSynchronize(idx.stream));
  B200_CUDA(cudaEventElapsedTime(&idx.build_ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return B200_OK;
}

int device_mean3(const float4* d_pts, int N, double* d_partials, double mean_out[3], cudaStream_t s) {
  const int blocks = std::min(1024, std::max(1, (N + kThreads - 1) / kThreads));
  sum3_kernel<<<blocks, kThreads, 0, s>>>(d_pts, N, d_partials);
  B200_CUDA(cudaGetLastError());
  std::vector<double> h(3 * (size_t)blocks);
  B200_CUDA(cudaMemcpyAsync(h.data(), d_partials, h.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  double acc[3] = {0, 0, 0};
  for (int b = 0; b < blocks; ++b)
    for (int d = 0; d < 3; ++d) acc[d] += h[3 * (size_t)b + d];
  for (int d = 0; d < 3; ++d) mean_out[d] = acc[d] / (double)N;
  return B200_OK;
}

int device_subtract3(float4* d_pts, int N, float mx, float my, float mz, cudaStream_t s) {
  if (N <= 0) return B200_OK;
  subtract3_kernel<<<grid_for(N, kThreads), kThreads, 0, s>>>(d_pts, N, mx, my, mz);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

size_t scan_tiles_for(size_t n) { return (n + kScanTile - 1) / kScanTile; }

int inclusive_scan_u32(uint32_t* d_data, size_t n, uint32_t* d_tile, cudaStream_t s) {
  if (n == 0) return B200_OK;
  const size_t tiles = scan_tiles_for(n);
  scan_tile_sums_kernel<<<(unsigned)tiles, kScanThreads, 0, s>>>(d_data, n, d_tile);
  scan_sums_kernel<<<1, kScanThreads, 0, s>>>(d_tile, tiles);
  scan_apply_kernel<<<(unsigned)tiles, kScanThreads, 0, s>>>(d_data, n, d_tile);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
