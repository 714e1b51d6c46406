// Common definitions of the B200 registration core (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <string>
#include "../../include/b200reg.h"

namespace b200 {

// ---- error plumbing (C ABI returns codes; message kept per thread) ------------------------------------------
void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define B200_CUDA(call)                                                        \
  do {                                                                         \
    cudaError_t _e = (call);                                                   \
    if (_e != cudaSuccess) return ::b200::cuda_fail(_e, #call, __FILE__, __LINE__); \
  } while (0)

#define B200_TRY(expr)              \
  do {                              \
    int _s = (expr);                \
    if (_s != B200_OK) return _s;   \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); }
    ok = (cudaSetDevice(dev) == cudaSuccess);
    if (!ok) cudaGetLastError();
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// ---- device-resident uniform-grid index over the (mean-centred) reference cloud --------------------------------
// Dense cell table, x fastest: cell(ix,iy,iz) = ix + nx * (iy + ny * iz).
// G[c] = first sorted position of cell c, G[c+1] = one past its last (G has ncells+1 entries, G[ncells] = M), so an
// x-row of cells [c0..c1] is ONE contiguous candidate range [G[c0], G[c1+1]).
// pts[s] = {x, y, z, bits(original index)} sorted by cell: a candidate costs one 16-byte load.
struct GridView {
  const float4* pts;       // M, sorted by cell
  const float4* nrm;       // M, sorted alike (may be null)
  const uint32_t* G;       // ncells + 1
  float ox, oy, oz;        // grid origin (min corner)
  float inv_cell, cell;
  int nx, ny, nz;
  int M;
};

__host__ __device__ inline float ordered_uint_as_float(uint32_t u) {
  // inverse of float_as_ordered_uint
  uint32_t b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f;
  memcpy(&f, &b, 4);
  return f;
#endif
}

}  // namespace b200
