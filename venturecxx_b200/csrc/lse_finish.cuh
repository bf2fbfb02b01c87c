// Block / grid reduction of (max, min, NaN flag) over the log weights into red[0..2] -- pass 1 of
// logsumexp (backend/lite/utils.py:70-91 needs the max for the shift and the min / NaN flag for its
// edge cases).  Shared by lse_pass1_kernel (k_weights.cu) and by kernels that PRODUCE the weights
// (hmm_step_kernel): a producer that reduces what it writes saves the following resample a full read
// pass over the weights.
// red layout (doubles): [0]=max [1]=min [2]=nan flag [3]=sum exp [4]=lse; [8..] per-block partials (3 per block)
#pragma once
#include "vk_common.cuh"
#include "p2p_common.cuh"

namespace vk {

static const int RED_PART = 8;

// every thread of the block calls this with its partial (max, min, nan); block size <= 1024.
// xchg.tab != nullptr (particle set sharded over peer GPUs, p2p_common.cuh): the last block then SENDS exchange 0
// in its tail; red[0..2] keep this rank's values until the consumer (p2p_tile_sums_kernel) has collected the
// peers' and stores the global ones.
__device__ __forceinline__ void lse_minmax_finish(double mx, double mn, double nanf, double *red,
                                                  unsigned int *counter, const P2PXchg &xchg) {
  __shared__ double sh[3][32];
  __shared__ bool last;
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  mx = warp_max(mx); mn = -warp_max(-mn); nanf = warp_max(nanf);
  if (lane == 0) { sh[0][w] = mx; sh[1][w] = mn; sh[2][w] = nanf; }
  __syncthreads();
  if (w == 0) {
    int nw = blockDim.x >> 5;
    mx = lane < nw ? sh[0][lane] : -INFINITY;
    mn = lane < nw ? sh[1][lane] : INFINITY;
    nanf = lane < nw ? sh[2][lane] : 0.0;
    mx = warp_max(mx); mn = -warp_max(-mn); nanf = warp_max(nanf);
    if (lane == 0) {
      red[RED_PART + 3 * blockIdx.x + 0] = mx;
      red[RED_PART + 3 * blockIdx.x + 1] = mn;
      red[RED_PART + 3 * blockIdx.x + 2] = nanf;
    }
  }
  __threadfence();
  if (threadIdx.x == 0) last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (last) {
    // the whole block reads the per-block partials (a single warp walking them was a chain of
    // gridDim.x/32 dependent L2 round trips at the tail of every producer kernel)
    __threadfence();
    mx = -INFINITY; mn = INFINITY; nanf = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
      mx = fmax(mx, red[RED_PART + 3 * b + 0]);
      mn = fmin(mn, red[RED_PART + 3 * b + 1]);
      nanf = fmax(nanf, red[RED_PART + 3 * b + 2]);
    }
    mx = warp_max(mx); mn = -warp_max(-mn); nanf = warp_max(nanf);
    __syncthreads();                      // sh[] was last read before the counter increment
    if (lane == 0) { sh[0][w] = mx; sh[1][w] = mn; sh[2][w] = nanf; }
    __syncthreads();
    if (w == 0) {
      int nw = blockDim.x >> 5;
      mx = lane < nw ? sh[0][lane] : -INFINITY;
      mn = lane < nw ? sh[1][lane] : INFINITY;
      nanf = lane < nw ? sh[2][lane] : 0.0;
      mx = warp_max(mx); mn = -warp_max(-mn); nanf = warp_max(nanf);
      if (lane == 0) { red[0] = mx; red[1] = mn; red[2] = nanf; *counter = 0; }
    }
    if (xchg.tab) {
      __syncthreads();
      if (w == 0) p2p_send_warp(xchg, 0, red[0], red[1], red[2] != 0.0 ? 1ull : 0ull);
    }
  }
}
__device__ __forceinline__ void lse_minmax_finish(double mx, double mn, double nanf, double *red,
                                                  unsigned int *counter) {
  P2PXchg none;
  none.tab = nullptr;
  lse_minmax_finish(mx, mn, nanf, red, counter, none);
}

}  // namespace vk
