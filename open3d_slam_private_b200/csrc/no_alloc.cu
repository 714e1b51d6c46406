// Allocation-free TrimmedDist outlier filter: b200_outlier_trimmed_noalloc (include/b200reg.h).
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same arithmetic as
// b200_outlier_dist(B200_OUTLIER_TRIMMEDDIST) — Matches::getDistsQuantile (Matches.cpp:60-87) by a 3-pass radix select
// over the bits of the non-negative floats (11 + 11 + 9), then weight = (dist <= limit)
// (OutlierFiltersImpl.cpp:140-147) — but inputs and outputs are DEVICE pointers owned by the caller, the work is only
// enqueued on the caller's stream (no synchronisation, no host read-back), and the radix-select scratch is the
// file-scope __device__ array below instead of a cudaMalloc'd block.
//
// The array exists once per GPU that loads the library and is held until the process exits. Every call uses it and it
// keeps the previous call's contents, so each call clears it first (a memset node on the same stream), and two calls
// whose GPU work OVERLAPS (different streams, or different host threads on the same device) race on it: callers
// serialise their calls per device.
#include "common.cuh"
#include <mutex>

namespace b200 {
namespace {

constexpr int kNaThreads = 256;
constexpr int kNaBins = 2048;
// state words: [0] status, [1] result bits, [2] prefix, [3] rank, [4] total; three histograms at +16
constexpr size_t kNaScratchWords = 3 * (size_t)kNaBins + 16;   // 24 640 bytes

static __device__ __align__(256) uint32_t g_noalloc_scratch[kNaScratchWords];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
uint32_t* g_scratch_addr[kMaxDevices];   // DEVICE address of g_noalloc_scratch on each device (null: not set up yet)

// Once per device, with that device current: in host code the array's name is only a stand-in, the kernels need the
// address cudaGetSymbolAddress returns for the CURRENT device.
int scratch_for_current_device(int device, uint32_t** out) {
  if (device < 0 || device >= kMaxDevices) { set_error("b200_outlier_trimmed_noalloc: device index out of range"); return B200_E_BAD_PARAM; }
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (!g_scratch_addr[device]) {
    uint32_t* ptr = nullptr;
    B200_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
    if (!ptr) { set_error("cudaGetSymbolAddress returned a null address for the no-alloc scratch"); return B200_E_CUDA; }
    g_scratch_addr[device] = ptr;
  }
  *out = g_scratch_addr[device];
  return B200_OK;
}

__global__ void __launch_bounds__(kNaThreads) na_hist_kernel(const float* __restrict__ d, int64_t n, int pass,
                                                             uint32_t* __restrict__ state) {
  __shared__ unsigned sh[kNaBins];
  for (int b = threadIdx.x; b < kNaBins; b += blockDim.x) sh[b] = 0u;
  __syncthreads();
  if (state[0] != B200_OK) return;
  const unsigned prefix = state[2];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = __ldg(&d[i]);
    if (v == INFINITY) continue;
    const unsigned bits = __float_as_uint(v);
    if (pass == 0) atomicAdd(&sh[(bits >> 20) & 0x7ffu], 1u);
    else if (pass == 1) { if (((bits >> 20) & 0x7ffu) == prefix) atomicAdd(&sh[(bits >> 9) & 0x7ffu], 1u); }
    else { if (((bits >> 9) & 0x3fffffu) == prefix) atomicAdd(&sh[bits & 0x1ffu], 1u); }
  }
  __syncthreads();
  unsigned* gh = state + 16 + pass * kNaBins;
  for (int b = threadIdx.x; b < kNaBins; b += blockDim.x)
    if (sh[b]) atomicAdd(&gh[b], sh[b]);
}

// one warp: bin of the wanted rank by a chunked prefix sum (64 bins per lane)
__global__ void na_pick_kernel(int pass, float quantile, uint32_t* __restrict__ state) {
  if (state[0] != B200_OK) return;
  const int lane = threadIdx.x;
  const unsigned* gh = state + 16 + pass * kNaBins;
  constexpr int per = kNaBins / 32;
  unsigned sum = 0;
  for (int j = 0; j < per; ++j) sum += gh[lane * per + j];
  unsigned incl = sum;
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const unsigned total = __shfl_sync(0xffffffffu, incl, 31);
  unsigned rank = state[3];
  __syncwarp();
  if (pass == 0) {
    if (total == 0u) {   // ConvergenceError("No matches available..."), Matches.cpp:75-76
      if (lane == 0) { state[4] = 0u; state[0] = B200_E_NO_MATCHES; }
      return;
    }
    // Matches.cpp:82-86: index = (size_t)(float(n_valid) * quantile), the maximum when quantile == 1
    if (quantile == 1.0f) rank = total - 1u;
    else {
      rank = (unsigned)__fmul_rn((float)total, quantile);
      if (rank > total - 1u) rank = total - 1u;
    }
  }
  unsigned ex = incl - sum;
  if (rank >= ex && rank < ex + sum) {   // exactly one lane
    int bin = lane * per;
    for (int j = 0; j < per; ++j) {
      const unsigned c = gh[lane * per + j];
      if (rank < ex + c) { bin = lane * per + j; break; }
      ex += c;
    }
    state[3] = rank - ex;
    if (pass == 0) { state[4] = total; state[2] = (unsigned)bin; }
    else if (pass == 1) state[2] = (state[2] << 11) | (unsigned)bin;
    else state[1] = (state[2] << 9) | (unsigned)bin;
  }
}

__global__ void __launch_bounds__(kNaThreads) na_weights_kernel(const float* __restrict__ d, int64_t n,
                                                                const uint32_t* __restrict__ state,
                                                                float* __restrict__ w, uint32_t* __restrict__ out2) {
  const uint32_t st = state[0], lim_bits = state[1];
  if (blockIdx.x == 0 && threadIdx.x == 0 && out2) { out2[0] = st; out2[1] = st == B200_OK ? lim_bits : 0u; }
  if (st != B200_OK) return;   // like the allocating entry point: no weights without a quantile
  const float lim = __uint_as_float(lim_bits);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    w[i] = (__ldg(&d[i]) <= lim) ? 1.f : 0.f;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" int b200_outlier_trimmed_noalloc(int device, const float* d_dists, int64_t n, float ratio, float* d_weights,
                                            uint32_t* d_status_limit, void* cuda_stream) {
  if (n < 0 || (n > 0 && (!d_dists || !d_weights))) {
    set_error("b200_outlier_trimmed_noalloc: bad arguments (null device pointer or negative size)");
    return B200_E_BAD_PARAM;
  }
  if (n > (int64_t)B200_NOALLOC_MAX_N) {
    set_error("b200_outlier_trimmed_noalloc: n exceeds B200_NOALLOC_MAX_N (16 777 216 entries); use b200_outlier_dist");
    return B200_E_UNSUPPORTED;
  }
  if (!(ratio >= 0.f && ratio <= 1.f)) {   // Matches.cpp:78-79
    set_error("Distance quantile of matches must lie in the range [0,1]");
    return B200_E_BAD_PARAM;
  }
  DeviceGuard dg(device);
  if (!dg.ok) return cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice", __FILE__, __LINE__);
  uint32_t* scratch = nullptr;
  B200_TRY(scratch_for_current_device(device, &scratch));
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : cudaStreamLegacy;
  // the array keeps the previous call's contents: status, prefix, rank and the three histograms must start at zero
  B200_CUDA(cudaMemsetAsync(scratch, 0, kNaScratchWords * sizeof(uint32_t), s));
  const int64_t want = (n + kNaThreads - 1) / kNaThreads;
  const int blocks = (int)(want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want));
  for (int pass = 0; pass < 3; ++pass) {
    na_hist_kernel<<<blocks, kNaThreads, 0, s>>>(d_dists, n, pass, scratch);
    na_pick_kernel<<<1, 32, 0, s>>>(pass, ratio, scratch);
  }
  na_weights_kernel<<<blocks, kNaThreads, 0, s>>>(d_dists, n, scratch, d_weights, d_status_limit);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}
