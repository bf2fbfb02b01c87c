// vk_logsumexp_static: logsumexp (backend/lite/utils.py:70-91) of a DEVICE array with NO allocation anywhere on the
// path -- for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
// It has no counterpart in the reference.  vk_logsumexp_array is
// the same operation with cudaMalloc'd scratch.
//
// Scratch: the file-scope array `vk_static_scratch` below (VK_STATIC_SCRATCH_BYTES = 256 + 32 * 4096 bytes = 128.25
// KiB; every GPU that loads the library holds it until exit).  Layout: a 256-byte header whose first word is the
// arrival counter of the grid, then one 32-byte record (max, sum exp(x - max), min, NaN flag) per tile of
// VK_STATIC_TILE = 4096 elements.  The record count bounds the input: n <= VK_STATIC_MAX_N = 4096 * 4096 = 2^24.
// The last block to arrive combines the records IN TILE ORDER (the result does not depend on block scheduling).
//
// In host code `vk_static_scratch` is a host-side stand-in, not a device address: its device address is looked up
// once per device with cudaGetSymbolAddress (scratch_of) and that pointer is what the kernel gets.
// Every call shares the array and finds the previous call's contents in it: the header is cleared on the caller's
// stream before every launch, the records are overwritten before they are read.  Calls that overlap on the GPU (two
// streams, two threads, two processes sharing a context) RACE on it: serialise them.
#include <mutex>
#include "vk_common.cuh"

namespace vk {

static const int VK_STATIC_TILE = 4096;
static const int VK_STATIC_THREADS = 256;
static const int VK_STATIC_MAX_TILES = 4096;
static const int64_t VK_STATIC_MAX_N = (int64_t)VK_STATIC_TILE * VK_STATIC_MAX_TILES;
static const size_t VK_STATIC_HEADER_BYTES = 256;
static const size_t VK_STATIC_SCRATCH_BYTES = VK_STATIC_HEADER_BYTES + 32 * (size_t)VK_STATIC_MAX_TILES;

static __device__ __align__(256) unsigned char vk_static_scratch[VK_STATIC_SCRATCH_BYTES];

// one tile per block; scratch = the DEVICE address of vk_static_scratch
__global__ void __launch_bounds__(VK_STATIC_THREADS)
lse_static_kernel(const double *__restrict__ x, int64_t n, unsigned char *__restrict__ scratch, double *__restrict__ out) {
  unsigned int *counter = (unsigned int *)scratch;
  double *rec = (double *)(scratch + VK_STATIC_HEADER_BYTES);
  __shared__ double sh[4][VK_STATIC_THREADS / 32];
  __shared__ double m_sh;
  __shared__ bool last;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = VK_STATIC_THREADS / 32;
  const int64_t base = (int64_t)blockIdx.x * VK_STATIC_TILE;
  double v[VK_STATIC_TILE / VK_STATIC_THREADS];
  double mx = -INFINITY, mn = INFINITY, nanf = 0.0;
#pragma unroll
  for (int j = 0; j < VK_STATIC_TILE / VK_STATIC_THREADS; ++j) {
    const int64_t i = base + (int64_t)j * VK_STATIC_THREADS + threadIdx.x;
    v[j] = i < n ? x[i] : NAN;                       // NaN marks "no element" below; real NaNs raise the flag here
    if (i < n) { if (v[j] != v[j]) nanf = 1.0; else { mx = fmax(mx, v[j]); mn = fmin(mn, v[j]); } }
  }
  mx = warp_max(mx); mn = -warp_max(-mn); nanf = warp_max(nanf);
  if (lane == 0) { sh[0][w] = mx; sh[2][w] = mn; sh[3][w] = nanf; }
  __syncthreads();
  if (w == 0) {
    mx = lane < nw ? sh[0][lane] : -INFINITY; mn = lane < nw ? sh[2][lane] : INFINITY; nanf = lane < nw ? sh[3][lane] : 0.0;
    mx = warp_max(mx); mn = -warp_max(-mn); nanf = warp_max(nanf);
    if (lane == 0) { m_sh = mx; sh[2][0] = mn; sh[3][0] = nanf; }
  }
  __syncthreads();
  const double m = m_sh;
  double s = 0.0;
  if (m > -INFINITY && m < INFINITY) {
#pragma unroll
    for (int j = 0; j < VK_STATIC_TILE / VK_STATIC_THREADS; ++j)
      if (v[j] == v[j]) s += exp(v[j] - m);
  }
  s = warp_sum(s);
  if (lane == 0) sh[1][w] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < nw; ++k) t += sh[1][k];      // fixed order
    rec[4 * blockIdx.x + 0] = m; rec[4 * blockIdx.x + 1] = t;
    rec[4 * blockIdx.x + 2] = sh[2][0]; rec[4 * blockIdx.x + 3] = sh[3][0];
    __threadfence();
    last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  // the last block: global (max, min, NaN flag), then the sum of the tiles' sums rescaled to the global max
  const int n_tiles = (int)gridDim.x;
  mx = -INFINITY; mn = INFINITY; nanf = 0.0;
  for (int b = threadIdx.x; b < n_tiles; b += VK_STATIC_THREADS) {
    mx = fmax(mx, rec[4 * b]); mn = fmin(mn, rec[4 * b + 2]); nanf = fmax(nanf, rec[4 * b + 3]);
  }
  mx = warp_max(mx); mn = -warp_max(-mn); nanf = warp_max(nanf);
  __syncthreads();
  if (lane == 0) { sh[0][w] = mx; sh[2][w] = mn; sh[3][w] = nanf; }
  __syncthreads();
  mx = -INFINITY; mn = INFINITY; nanf = 0.0;
  for (int k = 0; k < nw; ++k) { mx = fmax(mx, sh[0][k]); mn = fmin(mn, sh[2][k]); nanf = fmax(nanf, sh[3][k]); }
  // thread t sums tiles t, t + 256, ...; the partial sums are then added in thread order: a fixed order overall
  double part = 0.0;
  if (mx > -INFINITY && mx < INFINITY)
    for (int b = threadIdx.x; b < n_tiles; b += VK_STATIC_THREADS) {
      const double mb = rec[4 * b];
      if (mb > -INFINITY) part += rec[4 * b + 1] * exp(mb - mx);
    }
  __shared__ double part_sh[VK_STATIC_THREADS];
  part_sh[threadIdx.x] = part;
  __syncthreads();
  if (threadIdx.x == 0) {
    double se = 0.0;
    for (int k = 0; k < VK_STATIC_THREADS; ++k) se += part_sh[k];
    double lse;                                       // utils.py:77-91
    if (nanf != 0.0) lse = NAN;
    else if (isinf(mx) && mn != -mx) lse = mx;
    else if (isinf(mx)) lse = NAN;
    else lse = mx + log(se);
    *out = lse;
    *counter = 0;
  }
}

__global__ void lse_static_empty_kernel(double *out) { *out = -INFINITY; }

// the device address of vk_static_scratch on the CURRENT device, looked up once per device
static unsigned char *scratch_of(int device) {
  static std::mutex mu;
  static unsigned char *ptr[64] = {nullptr};
  VK_REQUIRE(device >= 0 && device < 64, VK_EINVAL, "vk_logsumexp_static: device index out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[device]) {
    unsigned char *p = nullptr;
    VK_CUDA(cudaGetSymbolAddress((void **)&p, vk_static_scratch));
    VK_REQUIRE(p != nullptr, VK_ECUDA, "vk_logsumexp_static: cudaGetSymbolAddress returned a null address");
    ptr[device] = p;
  }
  return ptr[device];
}

}  // namespace vk

using namespace vk;

extern "C" int vk_logsumexp_static(const double *in_dev, int64_t n, double *out_dev, void *cuda_stream) {
  try {
    VK_REQUIRE(out_dev != nullptr, VK_EINVAL, "vk_logsumexp_static: out_dev is NULL");
    VK_REQUIRE(n >= 0 && (in_dev != nullptr || n == 0), VK_EINVAL, "vk_logsumexp_static: bad input array");
    if (n > VK_STATIC_MAX_N) {
      char b[256];
      snprintf(b, sizeof b, "vk_logsumexp_static: n = %lld exceeds the static scratch's maximum of %lld elements "
                            "(use vk_logsumexp_array or a particle handle for larger arrays)",
               (long long)n, (long long)VK_STATIC_MAX_N);
      throw Error(VK_EINVAL, b);
    }
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) throw Error(VK_ECUDA, "vk_logsumexp_static: no CUDA device available (this library has no CPU fallback)");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (n == 0) {
      lse_static_empty_kernel<<<1, 1, 0, st>>>(out_dev);
      VK_CUDA(cudaGetLastError());
      return VK_OK;
    }
    unsigned char *scratch = scratch_of(device);
    VK_CUDA(cudaMemsetAsync(scratch, 0, VK_STATIC_HEADER_BYTES, st));      // the arrival counter starts at zero
    const int n_tiles = (int)((n + VK_STATIC_TILE - 1) / VK_STATIC_TILE);
    lse_static_kernel<<<n_tiles, VK_STATIC_THREADS, 0, st>>>(in_dev, n, scratch, out_dev);
    VK_CUDA(cudaGetLastError());
  } catch (const vk::Error &e) {
    vk::set_last_error(e.what());
    return e.code;
  } catch (const std::exception &e) {
    vk::set_last_error(e.what());
    return VK_ECUDA;
  }
  return VK_OK;
}
