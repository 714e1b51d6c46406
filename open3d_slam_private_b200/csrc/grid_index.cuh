// Host-side owner of the device-resident map index (replaces the libnabo kd-tree built by KDTreeMatcher::init,
// libpointmatcher/pointmatcher/MatchersImpl.cpp:78-83).
#pragma once
#include "common.cuh"

namespace b200 {

struct MapIndex {
  int device = 0;
  int M = 0;
  cudaStream_t stream = nullptr;  // not owned
  // original order (ids returned to the caller are positions in these arrays)
  float4* d_ref = nullptr;        // M x {x,y,z,1}
  float4* d_nrm_orig = nullptr;   // M x {nx,ny,nz,0} (null when the cloud has no normals)
  // sorted by cell
  float4* d_pts = nullptr;        // M x {x,y,z,bits(original index)}
  float4* d_nrm = nullptr;        // M x {nx,ny,nz,0}
  uint32_t* d_A = nullptr;        // ncells + 2 counters; G = d_A + 1
  uint32_t* d_tile = nullptr;     // scan scratch
  unsigned long long* d_scratch = nullptr;  // 8 x u64: bbox (6 ordered uints), occupied count
  size_t cap_cells = 0, cap_pts = 0, cap_tiles = 0;
  // scratch of surface_normals_index (normals_tile.cu), grown on demand: fallback list (+ counter), neighbour ids and
  // done flags of the smoothing pass
  int* d_sn_list = nullptr;
  int* d_sn_ids = nullptr;
  size_t cap_sn_list = 0, cap_sn_ids = 0;
  // grid parameters
  float ox = 0, oy = 0, oz = 0, cell = 1.f, inv_cell = 1.f;
  int nx = 1, ny = 1, nz = 1;
  int64_t occupied = 0;
  float build_ms = 0.f;

  GridView view() const {
    GridView g;
    g.pts = d_pts; g.nrm = d_nrm; g.G = d_A + 1;
    g.ox = ox; g.oy = oy; g.oz = oz; g.inv_cell = inv_cell; g.cell = cell;
    g.nx = nx; g.ny = ny; g.nz = nz; g.M = M;
    return g;
  }
  size_t ncells() const { return (size_t)nx * ny * nz; }
  void release();
};

// Upload helpers: host 4xN features -> device float4; host 3xN normals -> device float4 (w = 0).
int upload_features(float4* d_dst, const float* h_feat4xN, int N, cudaStream_t s);
int upload_normals(float4* d_dst, const float* h_nrm3xN, int N, float* d_stage, cudaStream_t s);

// Allocate (or grow) the point buffers of `idx` for M points on idx.device.
int index_alloc_points(MapIndex& idx, int M, bool with_normals);
// Build the grid from idx.d_ref / idx.d_nrm_orig (already on the device, final coordinates). Synchronises the
// stream a few times (bounding box and cell-size refinement need host decisions); build is not on the
// per-iteration path.
int index_build(MapIndex& idx, const b200_index_opts* opts);

// sum of xyz over N points in double (deterministic two-stage tree); result on host
int device_mean3(const float4* d_pts, int N, double* d_partials /* >= 3*1024 doubles */, double mean_out[3],
                 cudaStream_t s);
// p.xyz -= m (float subtraction, like `features.topRows(3).colwise() -= mean`, ICP.cpp:890)
int device_subtract3(float4* d_pts, int N, float mx, float my, float mz, cudaStream_t s);

}  // namespace b200
