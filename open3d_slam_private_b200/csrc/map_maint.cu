// Device-resident map maintenance: what open3d_slam does to the active submap between two
// registrations, without the map ever crossing PCIe.
//
//   b200_map_insert_scan   Submap::insertScan (open3d_slam/src/Submap.cpp:39-96): o3d_slam::transform of the scan
//                          (helpers.cpp:283-318), mapCloud_ += scan, voxelizeWithinCroppingVolume (helpers.cpp:117-200):
//                          points outside the cropper pass through in order, points inside are averaged per voxel
//                          (index floor(p / voxelSize), VoxelHashMap.hpp:48-51), normals averaged then normalised
//   b200_map_crop_to_icp   ScanToMapIcp::cropSubmap (ScanToMapRegistration.cpp:90-96) + open3dToPointmatcher
//                          (open3d_conversions.cpp:57-118: double -> float) + ICP::initReference (Mapper.cpp:329-347):
//                          the patch is cut on the device straight into the registration handle's reference buffers
//                          and indexed there — today's path re-uploads 160 MB of map and rebuilds (22 ms at 5 M points)
//
// The map is stored as DOUBLES like Open3D's points_ / normals_ (3 + 3 per point), so repeated averaging behaves as in
// the reference; floats are made only when a patch goes to the registration (as open3dToPointmatcher does).
// Voxel membership, counts and the pass-through set are index work and are exact; the voxel sums are accumulated with
// double atomics, i.e. in an unspecified order: the means equal the reference's sequential sums to a few ulps (1e-15
// relative), not bit for bit. Output order: pass-through points in input order, then the voxels by the index of their
// first member (the reference's order is that of an std::unordered_map).
#include "common.cuh"
#include "grid_index.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>

namespace b200 {
int inclusive_scan_u32(uint32_t* d_data, size_t n, uint32_t* d_tile, cudaStream_t s);
size_t scan_tiles_for(size_t n);
// icp_host.cu: finishes ICP::initReference on a reference already resident in the handle's index buffers
int icp_reference_from_device(b200_icp* h, int M, bool has_normals, const double* mean_given = nullptr);
float4* icp_reference_points(b200_icp* h, int M, bool with_normals, float4** normals_out, cudaStream_t* stream_out);
}  // namespace b200

using namespace b200;

namespace {

constexpr int kThreads = 256;
constexpr unsigned long long kEmpty = 0xffffffffffffffffull;

inline int blocks_for(size_t n) { return (int)std::max<size_t>(1, std::min<size_t>((n + kThreads - 1) / kThreads, 148 * 16)); }

struct CropD {
  int kind, invert;
  double rmin, rmax, zmin, zmax, cx, cy, cz;
};

__device__ __forceinline__ bool inside_volume(const CropD& c, double px, double py, double pz) {
  const double dx = px - c.cx, dy = py - c.cy, dz = pz - c.cz;
  bool in = true;
  if (c.kind == 1) in = pz >= c.zmin && pz <= c.zmax && sqrt(dx * dx + dy * dy) <= c.rmax;        // croppers.cpp:162-164
  else if (c.kind == 2) in = sqrt(dx * dx + dy * dy + dz * dz) >= c.rmin;                          // :149-151
  else if (c.kind == 3) in = sqrt(dx * dx + dy * dy + dz * dz) <= c.rmax;                          // :131-133
  else if (c.kind == 4) { const double d = sqrt(dx * dx + dy * dy + dz * dz); in = d <= c.rmax && d >= c.rmin; }  // :115-118
  return c.invert ? !in : in;                                                                      // :53-55
}

// o3d_slam::transform (helpers.cpp:299-310): p' = (T [p; 1]).head<3>() / w, n' = (T [n; 0]).head<3>(), in double;
// the scan (float images of the C ABI) is appended behind the M map points
__global__ void append_scan_kernel(const float4* __restrict__ scan, const float* __restrict__ scan_n3, int N,
                                   const double* __restrict__ T, double* __restrict__ xyz, double* __restrict__ nrm,
                                   int M) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
    const float4 v = __ldg(&scan[i]);
    const double x = (double)v.x, y = (double)v.y, z = (double)v.z;
    double q[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) q[r] = T[r] * x + T[r + 4] * y + T[r + 8] * z + T[r + 12] * 1.0;
    double* o = xyz + 3 * (size_t)(M + i);
    o[0] = q[0] / q[3]; o[1] = q[1] / q[3]; o[2] = q[2] / q[3];
    if (nrm) {
      const double a = (double)scan_n3[3 * (size_t)i], b = (double)scan_n3[3 * (size_t)i + 1], c = (double)scan_n3[3 * (size_t)i + 2];
      double* on = nrm + 3 * (size_t)(M + i);
#pragma unroll
      for (int r = 0; r < 3; ++r) on[r] = T[r] * a + T[r + 4] * b + T[r + 8] * c;
    }
  }
}

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
  return x;
}

// outside[i] = 1 for the points that pass through; the points inside the cropper are inserted into the voxel table
// (key = 21 bits per axis of floor(p * inv) + 2^20; linear probing) and accumulated: count, smallest member index,
// sums of the point and the normal. bad: a voxel index beyond +-2^20.
__global__ void classify_insert_kernel(const double* __restrict__ xyz, const double* __restrict__ nrm, int n, CropD c,
                                       double inv, unsigned long long mask, uint32_t* __restrict__ outside,
                                       unsigned long long* __restrict__ keys, uint32_t* __restrict__ cnt,
                                       int* __restrict__ first, double* __restrict__ sums, uint32_t* __restrict__ slot_of,
                                       int* __restrict__ bad) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double px = xyz[3 * (size_t)i], py = xyz[3 * (size_t)i + 1], pz = xyz[3 * (size_t)i + 2];
    if (!inside_volume(c, px, py, pz)) { outside[i] = 1u; slot_of[i] = 0xffffffffu; continue; }
    outside[i] = 0u;
    const double fx = floor(px * inv), fy = floor(py * inv), fz = floor(pz * inv);   // VoxelHashMap.hpp:48-51
    if (!(fabs(fx) < 1048575.0 && fabs(fy) < 1048575.0 && fabs(fz) < 1048575.0)) { *bad = 1; slot_of[i] = 0xffffffffu; continue; }
    const unsigned long long key = ((unsigned long long)((long long)fx + 1048576ll) << 42) |
                                   ((unsigned long long)((long long)fy + 1048576ll) << 21) |
                                   (unsigned long long)((long long)fz + 1048576ll);
    unsigned long long s = mix64(key) & mask;
    while (true) {
      const unsigned long long prev = atomicCAS(&keys[s], kEmpty, key);
      if (prev == kEmpty || prev == key) break;
      s = (s + 1ull) & mask;
    }
    slot_of[i] = (uint32_t)s;
    atomicAdd(&cnt[s], 1u);
    atomicMin(&first[s], i);
    double* a = sums + 6 * s;
    atomicAdd(&a[0], px); atomicAdd(&a[1], py); atomicAdd(&a[2], pz);
    if (nrm) { atomicAdd(&a[3], nrm[3 * (size_t)i]); atomicAdd(&a[4], nrm[3 * (size_t)i + 1]); atomicAdd(&a[5], nrm[3 * (size_t)i + 2]); }
  }
}

// flag[i] = 1 for the first member of every voxel
__global__ void mark_first_kernel(const uint32_t* __restrict__ slot_of, const int* __restrict__ first, int n,
                                  uint32_t* __restrict__ flag) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t s = slot_of[i];
    flag[i] = (s != 0xffffffffu && first[s] == i) ? 1u : 0u;
  }
}

// scan_out / scan_vox: inclusive scans of the pass-through flags and of the first-member flags
__global__ void write_map_kernel(const double* __restrict__ xyz, const double* __restrict__ nrm, int n,
                                 const uint32_t* __restrict__ scan_out, const uint32_t* __restrict__ scan_vox,
                                 const uint32_t* __restrict__ slot_of, const uint32_t* __restrict__ cnt,
                                 const double* __restrict__ sums, uint32_t n_outside, double* __restrict__ oxyz,
                                 double* __restrict__ onrm) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t so = scan_out[i], so0 = i ? scan_out[i - 1] : 0u;
    if (so != so0) {   // passes through, in input order
      const size_t o = so - 1u;
      for (int d = 0; d < 3; ++d) oxyz[3 * o + d] = xyz[3 * (size_t)i + d];
      if (nrm) for (int d = 0; d < 3; ++d) onrm[3 * o + d] = nrm[3 * (size_t)i + d];
      continue;
    }
    const uint32_t sv = scan_vox[i], sv0 = i ? scan_vox[i - 1] : 0u;
    if (sv == sv0) continue;   // a later member of some voxel
    const size_t o = (size_t)n_outside + sv - 1u;
    const uint32_t s = slot_of[i];
    const double m = (double)cnt[s];
    const double* a = sums + 6 * (size_t)s;
    oxyz[3 * o] = a[0] / m; oxyz[3 * o + 1] = a[1] / m; oxyz[3 * o + 2] = a[2] / m;   // GetAveragePoint
    if (nrm) {                                                                         // GetAverageNormal().normalized()
      double v[3] = {a[3] / m, a[4] / m, a[5] / m};
      const double z2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
      if (z2 > 0.0) { const double r = sqrt(z2); v[0] /= r; v[1] /= r; v[2] /= r; }
      onrm[3 * o] = v[0]; onrm[3 * o + 1] = v[1]; onrm[3 * o + 2] = v[2];
    }
  }
}

__global__ void crop_flag_d_kernel(const double* __restrict__ xyz, int n, CropD c, uint32_t* __restrict__ flag) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    flag[i] = inside_volume(c, xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2]) ? 1u : 0u;
}

// the kept points as the float images the registration core takes (open3dToPointmatcher: double -> float)
__global__ void crop_scatter_f4_kernel(const double* __restrict__ xyz, const double* __restrict__ nrm, int n,
                                       const uint32_t* __restrict__ scan, float4* __restrict__ out,
                                       float4* __restrict__ outn, int* __restrict__ idx) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t s = scan[i], s0 = i ? scan[i - 1] : 0u;
    if (s == s0) continue;
    const size_t o = s - 1u;
    out[o] = make_float4((float)xyz[3 * (size_t)i], (float)xyz[3 * (size_t)i + 1], (float)xyz[3 * (size_t)i + 2], 1.f);
    if (outn && nrm) outn[o] = make_float4((float)nrm[3 * (size_t)i], (float)nrm[3 * (size_t)i + 1], (float)nrm[3 * (size_t)i + 2], 0.f);
    if (idx) idx[o] = i;
  }
}

__global__ void fill_u64_kernel(unsigned long long* __restrict__ p, size_t n, unsigned long long v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void fill_i32_kernel(int* __restrict__ p, size_t n, int v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

}  // namespace

// =================================================================================================================
struct b200_map {
  int device = 0;
  cudaStream_t stream = nullptr;
  double *xyz[2] = {nullptr, nullptr}, *nrm[2] = {nullptr, nullptr};   // ping-pong: cur and the next map
  size_t cap = 0;        // points each of the four arrays can hold
  int cur = 0, M = 0;
  bool has_normals = false;
  // scratch, grown on demand
  struct Buf { void* p = nullptr; size_t cap = 0; };
  enum { OUTF, VOXF, TILE, SLOT, KEYS, CNT, FIRST, SUMS, SCAN4, SCANN, IDX, NBUF };
  Buf buf[NBUF];
  double* d_T = nullptr;
  int* d_bad = nullptr;
  cudaEvent_t ev[2] = {nullptr, nullptr};
  float last_ms = 0.f;
};

namespace {

int ensure(b200_map* h, int which, size_t bytes) {
  b200_map::Buf& b = h->buf[which];
  if (bytes <= b.cap) return B200_OK;
  cudaFree(b.p);
  b.p = nullptr; b.cap = 0;
  const size_t want = bytes + bytes / 4 + 256;
  B200_CUDA(cudaMalloc(&b.p, want));
  b.cap = want;
  return B200_OK;
}
template <class T> T* ptr(b200_map* h, int which) { return reinterpret_cast<T*>(h->buf[which].p); }

// room for n points in all four map arrays; the current map survives a growth
int reserve_points(b200_map* h, size_t n) {
  if (n <= h->cap) return B200_OK;
  const size_t want = n + n / 2 + 1024;
  for (int k = 0; k < 2; ++k) {
    double *nx = nullptr, *nn = nullptr;
    B200_CUDA(cudaMalloc(&nx, want * 3 * sizeof(double)));
    B200_CUDA(cudaMalloc(&nn, want * 3 * sizeof(double)));
    if (k == h->cur && h->M > 0) {
      B200_CUDA(cudaMemcpyAsync(nx, h->xyz[k], (size_t)h->M * 3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
      B200_CUDA(cudaMemcpyAsync(nn, h->nrm[k], (size_t)h->M * 3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
      B200_CUDA(cudaStreamSynchronize(h->stream));
    }
    cudaFree(h->xyz[k]); cudaFree(h->nrm[k]);
    h->xyz[k] = nx; h->nrm[k] = nn;
  }
  h->cap = want;
  return B200_OK;
}

CropD to_crop(const b200_crop_params* p) {
  CropD c;
  c.kind = p->kind; c.invert = p->invert;
  c.rmin = p->radius_min; c.rmax = p->radius_max; c.zmin = p->min_z; c.zmax = p->max_z;
  c.cx = p->center[0]; c.cy = p->center[1]; c.cz = p->center[2];
  return c;
}

int check_crop(const b200_crop_params* p, const char* who) {
  if (!p || p->kind < B200_CROP_NONE || p->kind > B200_CROP_MINMAX_RADIUS) {
    set_error(std::string(who) + ": unknown cropping volume");
    return B200_E_BAD_PARAM;
  }
  return B200_OK;
}

}  // namespace

extern "C" {

int b200_map_create(b200_map** out, int device) {
  if (!out) { set_error("b200_map_create: bad arguments"); return B200_E_BAD_PARAM; }
  *out = nullptr;
  DeviceGuard dg(device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  b200_map* h = new (std::nothrow) b200_map();
  if (!h) { set_error("out of host memory"); return B200_E_CUDA; }
  h->device = device;
  auto fail = [&](cudaError_t e, const char* what) { b200_map_destroy(h); return cuda_fail(e, what, __FILE__, __LINE__); };
  cudaError_t e;
  if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(e, "cudaStreamCreate");
  if ((e = cudaMalloc(&h->d_T, 16 * sizeof(double))) != cudaSuccess) return fail(e, "cudaMalloc");
  if ((e = cudaMalloc(&h->d_bad, sizeof(int))) != cudaSuccess) return fail(e, "cudaMalloc");
  for (int i = 0; i < 2; ++i)
    if ((e = cudaEventCreate(&h->ev[i])) != cudaSuccess) return fail(e, "cudaEventCreate");
  *out = h;
  return B200_OK;
}

void b200_map_destroy(b200_map* h) {
  if (!h) return;
  {
    DeviceGuard dg(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (int k = 0; k < 2; ++k) { cudaFree(h->xyz[k]); cudaFree(h->nrm[k]); }
    for (auto& b : h->buf) cudaFree(b.p);
    cudaFree(h->d_T); cudaFree(h->d_bad);
    for (int i = 0; i < 2; ++i)
      if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
}

int b200_map_set(b200_map* h, const double* pts3xM, const double* nrm3xM, int M) {
  if (!h || M < 0 || (M > 0 && !pts3xM)) { set_error("b200_map_set: bad arguments"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  h->M = 0;
  B200_TRY(reserve_points(h, (size_t)M));
  if (M > 0) {
    B200_CUDA(cudaMemcpyAsync(h->xyz[h->cur], pts3xM, (size_t)M * 3 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (nrm3xM) B200_CUDA(cudaMemcpyAsync(h->nrm[h->cur], nrm3xM, (size_t)M * 3 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    B200_CUDA(cudaStreamSynchronize(h->stream));
  }
  h->M = M;
  h->has_normals = nrm3xM != nullptr;
  return B200_OK;
}

int b200_map_size(const b200_map* h, int* M, int* has_normals) {
  if (!h) { set_error("null map"); return B200_E_BAD_PARAM; }
  if (M) *M = h->M;
  if (has_normals) *has_normals = h->has_normals ? 1 : 0;
  return B200_OK;
}

int b200_map_download(b200_map* h, double* pts3xM, double* nrm3xM) {
  if (!h) { set_error("null map"); return B200_E_BAD_PARAM; }
  DeviceGuard dg(h->device);
  if (h->M > 0) {
    if (pts3xM) B200_CUDA(cudaMemcpyAsync(pts3xM, h->xyz[h->cur], (size_t)h->M * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (nrm3xM && h->has_normals) B200_CUDA(cudaMemcpyAsync(nrm3xM, h->nrm[h->cur], (size_t)h->M * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    B200_CUDA(cudaStreamSynchronize(h->stream));
  }
  return B200_OK;
}

int b200_map_last_device_ms(const b200_map* h, float* ms) {
  if (!h || !ms) { set_error("b200_map_last_device_ms: bad arguments"); return B200_E_BAD_PARAM; }
  *ms = h->last_ms;
  return B200_OK;
}

int b200_map_insert_scan(b200_map* h, const float* scan4xN, const float* scann3xN, int N, const double* T16,
                         double voxel_size, const b200_crop_params* crop, int* n_outside, int* new_size) {
  if (!h || N < 0 || (N > 0 && !scan4xN) || !T16) { set_error("b200_map_insert_scan: bad arguments"); return B200_E_BAD_PARAM; }
  B200_TRY(check_crop(crop, "b200_map_insert_scan"));
  if (N == 0) {   // "if (preProcessedScan.IsEmpty()) return true" (Submap.cpp:41-43)
    if (n_outside) *n_outside = 0;
    if (new_size) *new_size = h->M;
    return B200_OK;
  }
  if (h->M > 0 && h->has_normals != (scann3xN != nullptr)) {
    set_error("b200_map_insert_scan: the map and the scan must both carry normals or both carry none");
    return B200_E_BAD_PARAM;
  }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  const bool hasN = scann3xN != nullptr;
  const int M = h->M, n = M + N;
  B200_TRY(reserve_points(h, (size_t)n));
  B200_TRY(ensure(h, b200_map::SCAN4, (size_t)N * sizeof(float4)));
  if (hasN) B200_TRY(ensure(h, b200_map::SCANN, (size_t)N * 3 * sizeof(float)));
  B200_CUDA(cudaMemcpyAsync(ptr<float4>(h, b200_map::SCAN4), scan4xN, (size_t)N * sizeof(float4), cudaMemcpyHostToDevice, s));
  if (hasN) B200_CUDA(cudaMemcpyAsync(ptr<float>(h, b200_map::SCANN), scann3xN, (size_t)N * 3 * sizeof(float), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaMemcpyAsync(h->d_T, T16, 16 * sizeof(double), cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  double* xyz = h->xyz[h->cur];
  double* nrm = hasN ? h->nrm[h->cur] : nullptr;
  append_scan_kernel<<<blocks_for((size_t)N), kThreads, 0, s>>>(ptr<float4>(h, b200_map::SCAN4),
                                                               hasN ? ptr<float>(h, b200_map::SCANN) : nullptr, N, h->d_T,
                                                               xyz, nrm, M);
  h->has_normals = hasN;
  if (!(voxel_size > 0.0)) {   // "*output = cloud" (helpers.cpp:121-124): the map is just the concatenation
    B200_CUDA(cudaGetLastError());
    B200_CUDA(cudaEventRecord(h->ev[1], s));
    B200_CUDA(cudaStreamSynchronize(s));
    B200_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]));
    h->M = n;
    if (n_outside) *n_outside = n;
    if (new_size) *new_size = n;
    return B200_OK;
  }
  size_t T = 1;
  while (T < 2 * (size_t)n) T <<= 1;
  const size_t tiles = scan_tiles_for((size_t)n);
  B200_TRY(ensure(h, b200_map::OUTF, (size_t)n * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_map::VOXF, (size_t)n * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_map::TILE, (tiles + 1) * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_map::SLOT, (size_t)n * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_map::KEYS, T * sizeof(unsigned long long)));
  B200_TRY(ensure(h, b200_map::CNT, T * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_map::FIRST, T * sizeof(int)));
  B200_TRY(ensure(h, b200_map::SUMS, T * 6 * sizeof(double)));
  fill_u64_kernel<<<blocks_for(T), kThreads, 0, s>>>(ptr<unsigned long long>(h, b200_map::KEYS), T, kEmpty);
  fill_i32_kernel<<<blocks_for(T), kThreads, 0, s>>>(ptr<int>(h, b200_map::FIRST), T, 0x7fffffff);
  B200_CUDA(cudaMemsetAsync(ptr<uint32_t>(h, b200_map::CNT), 0, T * sizeof(uint32_t), s));
  B200_CUDA(cudaMemsetAsync(ptr<double>(h, b200_map::SUMS), 0, T * 6 * sizeof(double), s));
  B200_CUDA(cudaMemsetAsync(h->d_bad, 0, sizeof(int), s));
  uint32_t* outf = ptr<uint32_t>(h, b200_map::OUTF);
  uint32_t* voxf = ptr<uint32_t>(h, b200_map::VOXF);
  uint32_t* slot = ptr<uint32_t>(h, b200_map::SLOT);
  classify_insert_kernel<<<blocks_for((size_t)n), kThreads, 0, s>>>(xyz, nrm, n, to_crop(crop), 1.0 / voxel_size, T - 1, outf,
                                                                    ptr<unsigned long long>(h, b200_map::KEYS),
                                                                    ptr<uint32_t>(h, b200_map::CNT), ptr<int>(h, b200_map::FIRST),
                                                                    ptr<double>(h, b200_map::SUMS), slot, h->d_bad);
  mark_first_kernel<<<blocks_for((size_t)n), kThreads, 0, s>>>(slot, ptr<int>(h, b200_map::FIRST), n, voxf);
  B200_CUDA(cudaGetLastError());
  B200_TRY(inclusive_scan_u32(outf, (size_t)n, ptr<uint32_t>(h, b200_map::TILE), s));
  B200_TRY(inclusive_scan_u32(voxf, (size_t)n, ptr<uint32_t>(h, b200_map::TILE), s));
  uint32_t tot[2];
  int bad = 0;
  B200_CUDA(cudaMemcpyAsync(&tot[0], outf + (n - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaMemcpyAsync(&tot[1], voxf + (n - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaMemcpyAsync(&bad, h->d_bad, sizeof(int), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  if (bad) { h->M = M; set_error("b200_map_insert_scan: voxel_size is too small for the extent of the map (voxel index beyond 2^20)"); return B200_E_BAD_PARAM; }
  const int nxt = h->cur ^ 1;
  write_map_kernel<<<blocks_for((size_t)n), kThreads, 0, s>>>(xyz, nrm, n, outf, voxf, slot, ptr<uint32_t>(h, b200_map::CNT),
                                                              ptr<double>(h, b200_map::SUMS), tot[0], h->xyz[nxt],
                                                              hasN ? h->nrm[nxt] : nullptr);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  B200_CUDA(cudaStreamSynchronize(s));
  B200_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]));
  h->cur = nxt;
  h->M = (int)(tot[0] + tot[1]);
  if (n_outside) *n_outside = (int)tot[0];
  if (new_size) *new_size = h->M;
  return B200_OK;
}

int b200_map_crop(b200_map* h, const b200_crop_params* crop, float* out4xK, float* outn3xK, int32_t* kept_idx, int* n_out) {
  if (!h || !n_out) { set_error("b200_map_crop: bad arguments"); return B200_E_BAD_PARAM; }
  B200_TRY(check_crop(crop, "b200_map_crop"));
  *n_out = 0;
  const int n = h->M;
  if (n == 0) return B200_OK;
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  B200_TRY(ensure(h, b200_map::OUTF, (size_t)n * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_map::TILE, (scan_tiles_for((size_t)n) + 1) * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_map::SCAN4, (size_t)n * sizeof(float4)));
  B200_TRY(ensure(h, b200_map::VOXF, (size_t)n * sizeof(float4)));   // normals as float4
  B200_TRY(ensure(h, b200_map::IDX, (size_t)n * sizeof(int)));
  uint32_t* flag = ptr<uint32_t>(h, b200_map::OUTF);
  crop_flag_d_kernel<<<blocks_for((size_t)n), kThreads, 0, s>>>(h->xyz[h->cur], n, to_crop(crop), flag);
  B200_CUDA(cudaGetLastError());
  B200_TRY(inclusive_scan_u32(flag, (size_t)n, ptr<uint32_t>(h, b200_map::TILE), s));
  crop_scatter_f4_kernel<<<blocks_for((size_t)n), kThreads, 0, s>>>(h->xyz[h->cur], h->has_normals ? h->nrm[h->cur] : nullptr, n,
                                                                    flag, ptr<float4>(h, b200_map::SCAN4),
                                                                    ptr<float4>(h, b200_map::VOXF), ptr<int>(h, b200_map::IDX));
  uint32_t K = 0;
  B200_CUDA(cudaMemcpyAsync(&K, flag + (n - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  *n_out = (int)K;
  if (K == 0) return B200_OK;
  if (out4xK) B200_CUDA(cudaMemcpyAsync(out4xK, ptr<float4>(h, b200_map::SCAN4), (size_t)K * sizeof(float4), cudaMemcpyDeviceToHost, s));
  if (kept_idx) B200_CUDA(cudaMemcpyAsync(kept_idx, ptr<int>(h, b200_map::IDX), (size_t)K * sizeof(int), cudaMemcpyDeviceToHost, s));
  if (outn3xK && h->has_normals) {
    // float4 -> packed 3 floats on the host side of the copy (2-D copy: 12 of every 16 bytes)
    B200_CUDA(cudaMemcpy2DAsync(outn3xK, 12, ptr<float4>(h, b200_map::VOXF), 16, 12, K, cudaMemcpyDeviceToHost, s));
  }
  B200_CUDA(cudaStreamSynchronize(s));
  return B200_OK;
}

int b200_map_crop_to_icp(b200_map* h, const b200_crop_params* crop, b200_icp* icp, int* n_patch) {
  if (!h || !icp) { set_error("b200_map_crop_to_icp: bad arguments"); return B200_E_BAD_PARAM; }
  B200_TRY(check_crop(crop, "b200_map_crop_to_icp"));
  const int n = h->M;
  if (n_patch) *n_patch = 0;
  if (n == 0) { set_error("The reference point cloud is empty"); return B200_E_NOT_INITIALIZED; }
  DeviceGuard dg(h->device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  cudaStream_t s = h->stream;
  B200_TRY(ensure(h, b200_map::OUTF, (size_t)n * sizeof(uint32_t)));
  B200_TRY(ensure(h, b200_map::TILE, (scan_tiles_for((size_t)n) + 1) * sizeof(uint32_t)));
  uint32_t* flag = ptr<uint32_t>(h, b200_map::OUTF);
  B200_CUDA(cudaEventRecord(h->ev[0], s));
  crop_flag_d_kernel<<<blocks_for((size_t)n), kThreads, 0, s>>>(h->xyz[h->cur], n, to_crop(crop), flag);
  B200_CUDA(cudaGetLastError());
  B200_TRY(inclusive_scan_u32(flag, (size_t)n, ptr<uint32_t>(h, b200_map::TILE), s));
  uint32_t K = 0;
  B200_CUDA(cudaMemcpyAsync(&K, flag + (n - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  if (K == 0) { set_error("The reference point cloud is empty (no map point inside the cropping volume)"); return B200_E_NOT_INITIALIZED; }
  float4* nrm4 = nullptr;
  cudaStream_t icp_stream = nullptr;
  float4* ref4 = icp_reference_points(icp, (int)K, h->has_normals, &nrm4, &icp_stream);
  if (!ref4) return B200_E_CUDA;
  crop_scatter_f4_kernel<<<blocks_for((size_t)n), kThreads, 0, s>>>(h->xyz[h->cur], h->has_normals ? h->nrm[h->cur] : nullptr, n,
                                                                    flag, ref4, nrm4, nullptr);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaEventRecord(h->ev[1], s));
  B200_CUDA(cudaStreamSynchronize(s));   // the registration handle continues on its own stream
  B200_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev[0], h->ev[1]));
  if (n_patch) *n_patch = (int)K;
  return icp_reference_from_device(icp, (int)K, h->has_normals);
}

}  // extern "C"
