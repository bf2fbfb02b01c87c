// Weight kernels: logsumexp, normalised probabilities, fixed-point prefix
// scan, ancestor search (systematic / multinomial) and ancestor gather.
//
// Reference semantics restated (paths under /root/reference):
//   backend/lite/utils.py:70-91    logsumexp (max-shift; +-inf / NaN policy)
//   backend/lite/utils.py:93-99    normalizeList (non-positive sum -> uniform)
//   backend/lite/utils.py:152-157  log_domain_even_out
//   backend/lite/utils.py:159-180  logWeightsToNormalizedDirect, sampleLogCategorical
//   python/lib/engine/trace_set.py:202-235  resample / _resample_traces
//
// B200 mapping: all kernels are HBM streaming passes over P doubles with
// grid-stride loops sized to a multiple of the SM count; block partials are
// combined by the last block to finish (no second launch, deterministic
// order).  The CDF is a 64-bit FIXED-POINT prefix sum so that ancestor
// indices do not depend on the association order of a parallel scan, on the
// GPU count, or on the platform's rounding: exact integers everywhere.
#include <string.h>
#include "vk_common.cuh"
#include "weights.cuh"
#include "models.cuh"
#include "lse_finish.cuh"

namespace vk {

// red layout (doubles): [0]=max [1]=min [2]=nan flag [3]=sum exp [4]=lse
// [8 ..] per-block partials (3 per block)
__device__ __forceinline__ void lse_pass1_body(const double *__restrict__ a, int64_t n, double &mx, double &mn,
                                               double &nanf) {
  mx = -INFINITY; mn = INFINITY; nanf = 0.0;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
  if ((((uintptr_t)a) & 15) == 0) {      // 16-byte loads, four in flight per thread
    const double2 *a2 = reinterpret_cast<const double2 *>(a);
    const int64_t n2 = n >> 1;
#pragma unroll 4
    for (int64_t i = tid; i < n2; i += stride) {
      double2 v = a2[i];
      if (v.x != v.x) nanf = 1.0; else { mx = fmax(mx, v.x); mn = fmin(mn, v.x); }
      if (v.y != v.y) nanf = 1.0; else { mx = fmax(mx, v.y); mn = fmin(mn, v.y); }
    }
    if ((n & 1) && tid == 0) {
      double v = a[n - 1];
      if (v != v) nanf = 1.0; else { mx = fmax(mx, v); mn = fmin(mn, v); }
    }
  } else {
    for (int64_t i = tid; i < n; i += stride) {
      double v = a[i];
      if (v != v) nanf = 1.0;
      else { mx = fmax(mx, v); mn = fmin(mn, v); }
    }
  }
}
__global__ void __launch_bounds__(256)
lse_pass1_kernel(const double *__restrict__ a, int64_t n, double *red, unsigned int *counter) {
  double mx, mn, nanf;
  lse_pass1_body(a, n, mx, mn, nanf);
  lse_minmax_finish(mx, mn, nanf, red, counter);
}
// the same pass for a particle set sharded over peer GPUs: the last block runs exchange 0 (p2p_common.cuh)
__global__ void __launch_bounds__(256)
lse_pass1_xchg_kernel(const double *__restrict__ a, int64_t n, double *red, unsigned int *counter, P2PXchg xchg) {
  double mx, mn, nanf;
  lse_pass1_body(a, n, mx, mn, nanf);
  lse_minmax_finish(mx, mn, nanf, red, counter, xchg);
}

// sum exp(a - m) with per-thread accumulation, block tree, last-block finish.
__global__ void __launch_bounds__(256)
lse_pass2_kernel(const double *__restrict__ a, int64_t n, double *red, unsigned int *counter) {
  __shared__ double sh[33];
  const double m = red[0];
  double s = 0.0;
  // if m is +-inf the sum is not used (edge-case branch below) -> skip work
  if (!isinf(m)) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
      s += exp(a[i] - m);
  }
  s = block_sum(s, sh);
  if (threadIdx.x == 0) red[RED_PART + blockIdx.x] = s;
  __shared__ bool last;
  __threadfence();
  if (threadIdx.x == 0) last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (last) {
    __threadfence();
    double t = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) t += red[RED_PART + b];
    t = block_sum(t, sh);
    if (threadIdx.x == 0) {
      double mn = red[1], nanf = red[2], lse;
      // utils.py:77-91
      if (n == 0) lse = -INFINITY;
      else if (nanf != 0.0) lse = NAN;
      else if (isinf(m) && mn != -m) lse = m;
      else if (isinf(m)) lse = NAN;          // +inf and -inf both present
      else lse = m + log(t);
      red[3] = t; red[4] = lse; *counter = 0;
    }
  }
}

void device_logsumexp(const double *a, int64_t n, double *red, unsigned int *counter,
                      int sm_count, cudaStream_t st) {
  int grid = grid_for(n, 256, sm_count, 4);
  lse_pass1_kernel<<<grid, 256, 0, st>>>(a, n, red, counter);
  lse_pass2_kernel<<<grid, 256, 0, st>>>(a, n, red, counter + 1);
  VK_CUDA(cudaGetLastError());
}

// utils.py:159-168: all -inf -> zeros; else exp(lg - max) normalised
// (normalizeList: sum <= 0 -> uniform, cannot happen when max is finite).
__global__ void normalized_probs_kernel(const double *__restrict__ a, int64_t n,
                                        const double *__restrict__ red, double *out) {
  const double m = red[0], s = red[3];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    double v;
    if (!(m > -INFINITY)) v = 0.0;
    else v = exp(a[i] - m) / s;
    out[i] = v;
  }
}

// ---- fixed-point weights ---------------------------------------------------
// q_i = rint(exp(lw_i - m) * 2^S); NaN or -inf -> 0; all -inf (m = -inf) -> 1
// for everybody (normalizeList's "equally impossible"); m = +inf -> 1 for the +inf entries, 0 for the rest
// (the mass is never 0 for a non-empty particle set, so every child gets an ancestor).
__device__ __forceinline__ uint64_t fixed_weight(double lw, double m, double scale) {
  if (!(m > -INFINITY)) return 1ull;
  if (m == INFINITY) return lw == INFINITY ? 1ull : 0ull;
  double d = lw - m;
  if (!(d == d)) return 0ull;
  return __double2ull_rn(exp(d) * scale);
}

static const int SCAN_THREADS = 256;
static const int SCAN_ITEMS = 8;
static const int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint64_t warp_scan_incl(uint64_t v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint64_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// pass A: per-tile totals
__global__ void __launch_bounds__(SCAN_THREADS)
fixed_tile_sums_kernel(const double *__restrict__ lw, int64_t n, double m_in,
                       const double *__restrict__ red, double scale, uint64_t *tile_sums) {
  __shared__ uint64_t sh[32];
  const double m = red ? red[0] : m_in;
  int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
  uint64_t s = 0;
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; ++j) {
    int64_t i = base + (int64_t)j * SCAN_THREADS + threadIdx.x;
    if (i < n) s += fixed_weight(lw[i], m, scale);
  }
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) sh[w] = s;
  __syncthreads();
  if (w == 0) {
    s = lane < (SCAN_THREADS >> 5) ? sh[lane] : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) tile_sums[blockIdx.x] = s;
  }
}

// pass B: exclusive scan of the tile totals by one block; total -> tile_sums[n_tiles]
__global__ void __launch_bounds__(1024)
scan_tile_sums_kernel(uint64_t *tile_sums, int n_tiles) {
  __shared__ uint64_t sh[32];
  __shared__ uint64_t carry_sh;
  if (threadIdx.x == 0) carry_sh = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int b0 = 0; b0 < n_tiles; b0 += 1024) {
    int b = b0 + threadIdx.x;
    uint64_t v = b < n_tiles ? tile_sums[b] : 0ull;
    uint64_t inc = warp_scan_incl(v, lane);
    if (lane == 31) sh[w] = inc;
    __syncthreads();
    if (w == 0) {
      uint64_t t = sh[lane];
      t = warp_scan_incl(t, lane);
      sh[lane] = t;
    }
    __syncthreads();
    uint64_t woff = w > 0 ? sh[w - 1] : 0ull;
    uint64_t carry = carry_sh;
    uint64_t excl = carry + woff + inc - v;
    if (b < n_tiles) tile_sums[b] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry_sh = carry + woff + inc;
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_sums[n_tiles] = carry_sh;
}

// pass C: inclusive CDF. Element order inside a tile is blocked per thread
// (thread t owns items t*8 .. t*8+7) so the prefix is over natural index order.
__global__ void __launch_bounds__(SCAN_THREADS)
fixed_cdf_kernel(const double *__restrict__ lw, int64_t n, double m_in,
                 const double *__restrict__ red, double scale,
                 const uint64_t *__restrict__ tile_sums, uint64_t *__restrict__ cdf) {
  __shared__ uint64_t sh[32];
  const double m = red ? red[0] : m_in;
  int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  uint64_t q[SCAN_ITEMS];
  uint64_t s = 0;
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; ++j) {
    int64_t i = base + j;
    q[j] = i < n ? fixed_weight(lw[i], m, scale) : 0ull;
    s += q[j];
  }
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint64_t inc = warp_scan_incl(s, lane);
  if (lane == 31) sh[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint64_t t = lane < (SCAN_THREADS >> 5) ? sh[lane] : 0ull;
    t = warp_scan_incl(t, lane);
    sh[lane] = t;
  }
  __syncthreads();
  uint64_t run = tile_sums[blockIdx.x] + (w > 0 ? sh[w - 1] : 0ull) + inc - s;
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; ++j) {
    int64_t i = base + j;
    run += q[j];
    if (i < n) cdf[i] = run;
  }
}

// smallest i in [0,n) with cdf[i] > thr  (thr < cdf[n-1] guaranteed)
__device__ __forceinline__ int64_t upper_bound_u64(const uint64_t *__restrict__ cdf, int64_t n,
                                                   uint64_t thr) {
  int64_t lo = 0, hi = n - 1;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (cdf[mid] > thr) hi = mid; else lo = mid + 1;
  }
  return lo;
}

// Ancestor search. `total` = W = mass of ALL shards, `offset` = mass of the
// shards before this one (0 on one GPU).  Child k of n_new (global index
// k_base + local thread) gets threshold
//   systematic : floor(((k<<32) + U0) * W / (n_new<<32)),  U0 = floor(u*2^32)
//   multinomial: floor(U_k * W / 2^53),                    U_k = floor(u_k*2^53)
// and ancestor = first local i with offset + cdf[i] > thr.  Children whose
// threshold lies outside this shard's [offset, offset + cdf[n-1]) get -1.
__global__ void ancestor_search_kernel(const uint64_t *__restrict__ cdf, int64_t n,
                                       uint64_t offset, uint64_t total,
                                       int64_t n_new, int64_t k_begin, int64_t k_count,
                                       int mode, uint64_t seed, uint32_t step,
                                       const double *__restrict__ uniforms,
                                       int32_t *__restrict__ anc_out) {
  uint64_t shard_mass = cdf[n - 1];
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < k_count;
       t += (int64_t)gridDim.x * blockDim.x) {
    int64_t k = k_begin + t;
    uint64_t thr;
    if (mode == VK_RESAMPLE_SYSTEMATIC) {
      double u;
      if (uniforms) u = uniforms[0];
      else { Stream r(seed, GLOBAL_PARTICLE, step); u = r.uniform(); }
      uint64_t U0 = (uint64_t)(u * 4294967296.0);
      if (U0 > 0xFFFFFFFFull) U0 = 0xFFFFFFFFull;
      unsigned __int128 num = (((unsigned __int128)(uint64_t)k << 32) | U0) * (unsigned __int128)total;
      unsigned __int128 den = (unsigned __int128)(uint64_t)n_new << 32;
      thr = (uint64_t)(num / den);
    } else {
      double u;
      if (uniforms) u = uniforms[k];
      else { Stream r(seed, (uint64_t)k, step); u = r.uniform(); }
      uint64_t U = (uint64_t)(u * 9007199254740992.0);
      if (U > 9007199254740991ull) U = 9007199254740991ull;
      thr = (uint64_t)(((unsigned __int128)U * (unsigned __int128)total) >> 53);
    }
    int32_t a = -1;
    if (thr >= offset && thr - offset < shard_mass)
      a = (int32_t)upper_bound_u64(cdf, n, thr - offset);
    anc_out[t] = a;
  }
}

// ---- ancestor gather --------------------------------------------------------
// particle-major rows of row_bytes (multiple of 16): 16-byte units
__global__ void gather_rows16_kernel(const uint4 *__restrict__ in, uint4 *__restrict__ out,
                                     const int32_t *__restrict__ anc, int64_t n_new,
                                     int64_t units_per_row) {
  int64_t total = n_new * units_per_row;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    int64_t k = t / units_per_row, c = t - k * units_per_row;
    out[t] = in[(int64_t)anc[k] * units_per_row + c];
  }
}
// generic byte rows (row_bytes not a multiple of 16)
__global__ void gather_rows1_kernel(const unsigned char *__restrict__ in,
                                    unsigned char *__restrict__ out,
                                    const int32_t *__restrict__ anc, int64_t n_new,
                                    int64_t row_bytes) {
  int64_t total = n_new * row_bytes;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    int64_t k = t / row_bytes, c = t - k * row_bytes;
    out[t] = in[(int64_t)anc[k] * row_bytes + c];
  }
}
// particle-minor SoA: n_rows rows of P elements
template <typename T>
__global__ void gather_soa_kernel(const T *__restrict__ in, T *__restrict__ out,
                                  const int32_t *__restrict__ anc, int64_t P_old,
                                  int64_t n_new, int64_t n_rows) {
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n_new;
       k += (int64_t)gridDim.x * blockDim.x) {
    int64_t a = anc[k];
    for (int64_t r = 0; r < n_rows; ++r) out[r * n_new + k] = in[r * P_old + a];
  }
}

__global__ void fill_kernel(double *out, int64_t n, const double *red, double log_n_new) {
  double v = red[4] - log_n_new;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    out[i] = v;
}

int frac_bits_for(int64_t P_total) {
  int lg = 0;
  while (((int64_t)1 << lg) < P_total) ++lg;
  int S = 62 - lg;
  return S > 52 ? 52 : S;
}

void build_fixed_cdf(vk_particles *h, const double *red_or_null, double m, int frac_bits) {
  cudaStream_t st = h->stream;
  int64_t P = h->P;
  int n_tiles = (int)((P + SCAN_TILE - 1) / SCAN_TILE);
  if (h->red_u.n < (size_t)n_tiles + 1) h->red_u.alloc((size_t)n_tiles + 1);
  double scale = ldexp(1.0, frac_bits);
  fixed_tile_sums_kernel<<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, m, red_or_null, scale, h->red_u.p);
  scan_tile_sums_kernel<<<1, 1024, 0, st>>>(h->red_u.p, n_tiles);
  fixed_cdf_kernel<<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, m, red_or_null, scale, h->red_u.p, h->cdf.p);
  VK_CUDA(cudaGetLastError());
}

int gather_state(vk_particles *h, const int32_t *anc_dev, int64_t n_new, double *lw_out, double log_n_new,
                 std::vector<StateArray> *explicit_arrs) {
  cudaStream_t st = h->stream;
  if (!explicit_arrs)
    if (int done = h->model->gather(anc_dev, n_new, lw_out, h->red.p, log_n_new)) return done;
  std::vector<StateArray> arrs = explicit_arrs ? *explicit_arrs : h->model->state_arrays();
  for (auto &a : arrs) {
    if (a.elem_bytes == 0) {
      if (a.row_bytes % 16 == 0) {
        int64_t units = (int64_t)(a.row_bytes / 16);
        int grid = grid_for(n_new * units, 256, h->sm_count, 8);
        gather_rows16_kernel<<<grid, 256, 0, st>>>((const uint4 *)*a.cur, (uint4 *)*a.alt, anc_dev, n_new, units);
      } else {
        int grid = grid_for(n_new * (int64_t)a.row_bytes, 256, h->sm_count, 8);
        gather_rows1_kernel<<<grid, 256, 0, st>>>((const unsigned char *)*a.cur, (unsigned char *)*a.alt, anc_dev, n_new, (int64_t)a.row_bytes);
      }
    } else {
      int grid = grid_for(n_new, 256, h->sm_count, 8);
      if (a.elem_bytes == 8)
        gather_soa_kernel<uint64_t><<<grid, 256, 0, st>>>((const uint64_t *)*a.cur, (uint64_t *)*a.alt, anc_dev, h->P, n_new, (int64_t)a.n_rows);
      else if (a.elem_bytes == 4)
        gather_soa_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t *)*a.cur, (uint32_t *)*a.alt, anc_dev, h->P, n_new, (int64_t)a.n_rows);
      else if (a.elem_bytes == 1)
        gather_soa_kernel<uint8_t><<<grid, 256, 0, st>>>((const uint8_t *)*a.cur, (uint8_t *)*a.alt, anc_dev, h->P, n_new, (int64_t)a.n_rows);
      else
        throw Error(VK_EINVAL, "unsupported SoA element size");
    }
    void *t = *a.cur; *a.cur = *a.alt; *a.alt = t;
  }
  VK_CUDA(cudaGetLastError());
  return 1;
}

// ---- systematic resampling without a CDF array or a search ------------------------
// Thresholds thr_k = floor(((k<<32)+U0) W / (n_new<<32)) increase with k, so the children
// of particle i are the contiguous range [kf(cdf[i-1]), kf(cdf[i])) with
//   kf(c) = smallest k with thr_k >= c = ceil((ceil(c*den/W) - U0) / 2^32),  den = n_new<<32
// (exact 128-bit integers; identical to searching the CDF for every child).
typedef unsigned __int128 u128_t;
// Evaluated without a 128-bit division (hundreds of instructions in software): thr_k >= c  <=>
// ((k<<32)+U0) * W >= c * (n_new<<32), two 64x64->128 products; k starts from a double-precision estimate
// that is off by at most one and is corrected with that exact test.  Same integer as the closed form.
__device__ __forceinline__ bool child_reaches(uint64_t k, uint64_t c, uint64_t W, int64_t n_new, uint64_t U0) {
  u128_t lhs = (u128_t)((k << 32) + U0) * (u128_t)W;
  u128_t rhs = ((u128_t)c * (u128_t)(uint64_t)n_new) << 32;
  return lhs >= rhs;
}
// `ratio` = n_new / W and `u0f` = U0 / 2^32 in double (loop invariants of the caller).
// thr_k >= c  <=>  k + U0/2^32 >= c n_new / W, so the answer is ceil(x), x = c n_new / W - U0/2^32.  The
// double evaluation of x is off by a few ulp (<< delta = (|x|+1) 2^-45): when x is further than delta
// from an integer, ceil of the double IS the exact answer; otherwise (probability ~1e-7 per call) the
// exact 128-bit test decides.  Same integer either way.
__device__ __forceinline__ int64_t first_child_at_least(uint64_t c, uint64_t W, int64_t n_new, uint64_t U0,
                                                        double ratio, double u0f) {
  if (c == 0) return 0;
  const double est = fma((double)c, ratio, -u0f);
  const double fl = floor(est), frac = est - fl, delta = (fabs(est) + 1.0) * 2.842170943040401e-14;
  if (frac > delta && frac < 1.0 - delta) {
    if (fl < 0.0) return 0;
    if (fl >= (double)n_new) return n_new;
    return (int64_t)fl + 1;
  }
  int64_t k = est <= 0.0 ? 0 : (est >= (double)n_new ? n_new : (int64_t)est);
  while (k < n_new && !child_reaches((uint64_t)k, c, W, n_new, U0)) ++k;
  while (k > 0 && child_reaches((uint64_t)(k - 1), c, W, n_new, U0)) --k;
  return k;
}
__device__ __forceinline__ int64_t first_child_at_least(uint64_t c, uint64_t W, int64_t n_new, uint64_t U0) {
  return first_child_at_least(c, W, n_new, U0, (double)n_new / (double)W, (double)U0 * (1.0 / 4294967296.0));
}
__device__ __forceinline__ uint64_t systematic_U0(const double *uniforms, uint64_t seed, uint32_t step) {
  double u;
  if (uniforms) u = uniforms[0];
  else { Stream r(seed, GLOBAL_PARTICLE, step); u = r.uniform(); }
  uint64_t U0 = (uint64_t)(u * 4294967296.0);
  return U0 > 0xFFFFFFFFull ? 0xFFFFFFFFull : U0;
}

// pass A': per-tile totals of the fixed-point weights AND of exp(lw - m) (for the LSE)
// ITEMS particles per thread: 8 for large sets; the peer-memory path uses 2 for small per-GPU sets (four times the
// CTAs, a quarter of the serial work per thread -- integer sums, so the CDF does not depend on the tiling)
template <int ITEMS = SCAN_ITEMS>
__device__ __forceinline__ void
fixed_tile_sums_exp_body(const double *__restrict__ lw, int64_t n, const double m,
                         double scale, uint64_t *tile_sums, double *tile_exp) {
  __shared__ uint64_t sh[32];
  __shared__ double shd[33];
  int64_t base = (int64_t)blockIdx.x * (SCAN_THREADS * ITEMS);
  uint64_t s = 0;
  double se = 0.0;
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    int64_t i = base + (int64_t)j * SCAN_THREADS + threadIdx.x;
    if (i < n) {
      double v = lw[i];
      if (!(m > -INFINITY)) s += 1ull;
      else if (m == INFINITY) s += v == INFINITY ? 1ull : 0ull;
      else {
        double d = v - m;
        if (d == d) { double e = exp(d); s += __double2ull_rn(e * scale); se += e; }
      }
    }
  }
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) sh[w] = s;
  se = block_sum(se, shd);
  if (w == 0) {
    s = lane < (SCAN_THREADS >> 5) ? sh[lane] : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) { tile_sums[blockIdx.x] = s; tile_exp[blockIdx.x] = se; }
  }
}
__global__ void __launch_bounds__(SCAN_THREADS)
fixed_tile_sums_exp_kernel(const double *__restrict__ lw, int64_t n, const double *__restrict__ red,
                           double scale, uint64_t *tile_sums, double *tile_exp) {
  fixed_tile_sums_exp_body(lw, n, red[0], scale, tile_sums, tile_exp);
}

// pass B': scan of tile totals (as scan_tile_sums_kernel) + LSE from the tile exp sums
// (fixed order) with the edge cases of utils.py:77-91 -> red[3], red[4]
__device__ __forceinline__ void scan_tile_sums_lse_body(uint64_t *tile_sums, const double *tile_exp, int n_tiles,
                                                        int64_t n, double *red) {
  __shared__ uint64_t sh[32];
  __shared__ double shd[33];
  __shared__ uint64_t carry_sh;
  if (threadIdx.x == 0) carry_sh = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int nthr = (int)blockDim.x, nw = nthr >> 5;      // any block size that is a multiple of 32, <= 1024
  double se = 0.0;
  for (int b0 = 0; b0 < n_tiles; b0 += nthr) {
    int b = b0 + threadIdx.x;
    uint64_t v = b < n_tiles ? tile_sums[b] : 0ull;
    if (b < n_tiles) se += tile_exp[b];
    uint64_t inc = warp_scan_incl(v, lane);
    if (lane == 31) sh[w] = inc;
    __syncthreads();
    if (w == 0) { uint64_t t = lane < nw ? sh[lane] : 0ull; t = warp_scan_incl(t, lane); sh[lane] = t; }
    __syncthreads();
    uint64_t woff = w > 0 ? sh[w - 1] : 0ull;
    uint64_t carry = carry_sh;
    if (b < n_tiles) tile_sums[b] = carry + woff + inc - v;
    __syncthreads();
    if (threadIdx.x == nthr - 1) carry_sh = carry + woff + inc;
    __syncthreads();
  }
  se = block_sum(se, shd);
  if (threadIdx.x == 0) {
    tile_sums[n_tiles] = carry_sh;
    double m = red[0], mn = red[1], nanf = red[2], lse;
    if (n == 0) lse = -INFINITY;
    else if (nanf != 0.0) lse = NAN;
    else if (isinf(m) && mn != -m) lse = m;
    else if (isinf(m)) lse = NAN;
    else lse = m + log(se);
    red[3] = se; red[4] = lse;
  }
}
__global__ void __launch_bounds__(1024)
scan_tile_sums_lse_kernel(uint64_t *tile_sums, const double *tile_exp, int n_tiles, int64_t n, double *red) {
  scan_tile_sums_lse_body(tile_sums, tile_exp, n_tiles, n, red);
}

// pass C': running CDF in registers -> children ranges -> anc[k] = i  (no CDF array, no search)
// RAW_TILES: tile_sums holds the per-tile totals as fixed_tile_sums_exp_kernel wrote them; every CTA sums
// the tiles before its own (and all of them, for W) itself -- integer sums, so any order gives the same
// CDF -- and CTA 0 finishes the logsumexp from tile_exp (what scan_tile_sums_lse_kernel does), which
// removes that single-CTA kernel from the chain when there are few tiles.
template <bool RAW_TILES>
__global__ void __launch_bounds__(SCAN_THREADS)
systematic_offspring_kernel(const double *__restrict__ lw, int64_t n, double *__restrict__ red,
                            double scale, uint64_t *__restrict__ tile_sums, const double *__restrict__ tile_exp,
                            int n_tiles, int64_t n_new, uint64_t seed, uint32_t step,
                            const double *__restrict__ uniforms, int32_t *__restrict__ anc) {
  __shared__ uint64_t sh[32];
  __shared__ uint64_t sh2[2][SCAN_THREADS / 32];
  __shared__ double shd[33];
  const double m = red[0];
  uint64_t W, tile_base;
  if (RAW_TILES) {
    uint64_t pre = 0, tot = 0;
    double se = 0.0;
    for (int b = threadIdx.x; b < n_tiles; b += SCAN_THREADS) {
      const uint64_t v = tile_sums[b];
      tot += v;
      if (b < (int)blockIdx.x) pre += v;
      if (blockIdx.x == 0) se += tile_exp[b];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { pre += __shfl_xor_sync(0xffffffffu, pre, o); tot += __shfl_xor_sync(0xffffffffu, tot, o); }
    if ((threadIdx.x & 31) == 0) { sh2[0][threadIdx.x >> 5] = pre; sh2[1][threadIdx.x >> 5] = tot; }
    __syncthreads();
    pre = 0; tot = 0;
#pragma unroll
    for (int k = 0; k < SCAN_THREADS / 32; ++k) { pre += sh2[0][k]; tot += sh2[1][k]; }
    W = tot; tile_base = pre;
    if (blockIdx.x == 0) {     // uniform branch: the whole CTA
      se = block_sum(se, shd);
      if (threadIdx.x == 0) {
        const double mn = red[1], nanf = red[2];
        double lse;
        if (n == 0) lse = -INFINITY;
        else if (nanf != 0.0) lse = NAN;
        else if (isinf(m) && mn != -m) lse = m;
        else if (isinf(m)) lse = NAN;
        else lse = m + log(se);
        red[3] = se; red[4] = lse;
      }
    }
  } else {
    W = tile_sums[n_tiles]; tile_base = tile_sums[blockIdx.x];
  }
  const uint64_t U0 = systematic_U0(uniforms, seed, step);
  int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  uint64_t q[SCAN_ITEMS];
  uint64_t s = 0;
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; ++j) {
    int64_t i = base + j;
    q[j] = i < n ? fixed_weight(lw[i], m, scale) : 0ull;
    s += q[j];
  }
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint64_t inc = warp_scan_incl(s, lane);
  if (lane == 31) sh[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint64_t t = lane < (SCAN_THREADS >> 5) ? sh[lane] : 0ull;
    t = warp_scan_incl(t, lane);
    sh[lane] = t;
  }
  __syncthreads();
  if (s == 0) return;   // no mass in this thread's items: no children
  uint64_t run = tile_base + (w > 0 ? sh[w - 1] : 0ull) + inc - s;
  const double ratio = (double)n_new / (double)W, u0f = (double)U0 * (1.0 / 4294967296.0);
  int64_t k_lo = first_child_at_least(run, W, n_new, U0, ratio, u0f);
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; ++j) {
    if (q[j] == 0) continue;
    run += q[j];
    int64_t k_hi = first_child_at_least(run, W, n_new, U0, ratio, u0f);
    for (int64_t k = k_lo; k < k_hi; ++k) anc[k] = (int32_t)(base + j);
    k_lo = k_hi;
  }
}

// ---- small particle sets: the whole resample in ONE single-CTA kernel ----------------
// (C1 at P = 1000 is launch-latency bound: 1 launch instead of ~10.)  lw, the fixed-point
// CDF and the ancestors live in shared memory; state arrays are gathered cur -> alt.
static const int SMALL_P_MAX = 2048;
static const int SMALL_THREADS = 512;
static const int SMALL_MAX_ARRAYS = 4;
struct SmallArrays {
  int n;
  const unsigned char *cur[SMALL_MAX_ARRAYS];
  unsigned char *alt[SMALL_MAX_ARRAYS];
  long long row_bytes[SMALL_MAX_ARRAYS];   // particle-major rows (elem_bytes == 0)
  long long n_rows[SMALL_MAX_ARRAYS];      // SoA rows
  int elem_bytes[SMALL_MAX_ARRAYS];
};
__global__ void __launch_bounds__(SMALL_THREADS)
resample_small_kernel(double *__restrict__ lw, int P, int n_new, int mode, int frac_bits, uint64_t seed,
                      uint32_t step, const double *__restrict__ uniforms, int32_t *__restrict__ anc_out,
                      double *__restrict__ red, SmallArrays arrs) {
  __shared__ uint64_t cdf[SMALL_P_MAX];
  __shared__ int32_t anc[SMALL_P_MAX];
  __shared__ double shd[3][33];
  __shared__ uint64_t shu[33];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  constexpr int ITEMS = SMALL_P_MAX / SMALL_THREADS;   // consecutive items per thread
  double v[ITEMS];
  double mx = -INFINITY, mn = INFINITY, nanf = 0.0;
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    int i = tid * ITEMS + j;
    v[j] = i < P ? lw[i] : NAN;
    if (i < P) { if (v[j] != v[j]) nanf = 1.0; else { mx = fmax(mx, v[j]); mn = fmin(mn, v[j]); } }
  }
  mx = warp_max(mx); mn = -warp_max(-mn); nanf = warp_max(nanf);
  if (lane == 0) { shd[0][w] = mx; shd[1][w] = mn; shd[2][w] = nanf; }
  __syncthreads();
  { const int nw = SMALL_THREADS / 32;
    mx = lane < nw ? shd[0][lane] : -INFINITY; mn = lane < nw ? shd[1][lane] : INFINITY; nanf = lane < nw ? shd[2][lane] : 0.0; }
  mx = warp_max(mx); mn = -warp_max(-mn); nanf = warp_max(nanf);
  const double m = mx;
  const double scale = ldexp(1.0, frac_bits);
  uint64_t q[ITEMS], s = 0;
  double se = 0.0;
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    int i = tid * ITEMS + j;
    q[j] = 0;
    if (i < P) {
      if (!(m > -INFINITY)) q[j] = 1ull;
      else if (m == INFINITY) q[j] = v[j] == INFINITY ? 1ull : 0ull;
      else { double d = v[j] - m; if (d == d) { double e = exp(d); q[j] = __double2ull_rn(e * scale); se += e; } }
    }
    s += q[j];
  }
  __syncthreads();   // shd reuse
  se = warp_sum(se);
  uint64_t inc = warp_scan_incl(s, lane);
  if (lane == 31) shu[w] = inc;
  if (lane == 0) shd[0][w] = se;
  __syncthreads();
  if (w == 0) { uint64_t t = lane < SMALL_THREADS / 32 ? shu[lane] : 0ull; t = warp_scan_incl(t, lane); shu[lane] = t; }
  se = warp_sum(lane < SMALL_THREADS / 32 ? shd[0][lane] : 0.0);
  __syncthreads();
  const uint64_t W = shu[31];
  uint64_t run = (w > 0 ? shu[w - 1] : 0ull) + inc - s;
  double lse;
  if (nanf != 0.0) lse = NAN;
  else if (isinf(m) && mn != -m) lse = m;
  else if (isinf(m)) lse = NAN;
  else lse = m + log(se);
  if (mode == VK_RESAMPLE_SYSTEMATIC) {
    const uint64_t U0 = systematic_U0(uniforms, seed, step);
    if (s != 0) {
      int64_t k_lo = first_child_at_least(run, W, n_new, U0);
#pragma unroll
      for (int j = 0; j < ITEMS; ++j) {
        if (q[j] == 0) continue;
        run += q[j];
        int64_t k_hi = first_child_at_least(run, W, n_new, U0);
        for (int64_t k = k_lo; k < k_hi; ++k) anc[k] = tid * ITEMS + j;
        k_lo = k_hi;
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) { run += q[j]; int i = tid * ITEMS + j; if (i < P) cdf[i] = run; }
    __syncthreads();
    for (int k = tid; k < n_new; k += SMALL_THREADS) {
      double u;
      if (uniforms) u = uniforms[k];
      else { Stream r(seed, (uint64_t)k, step); u = r.uniform(); }
      uint64_t U = (uint64_t)(u * 9007199254740992.0);
      if (U > 9007199254740991ull) U = 9007199254740991ull;
      uint64_t thr = (uint64_t)(((u128_t)U * (u128_t)W) >> 53);
      int lo = 0, hi = P - 1;
      while (lo < hi) { int mid = (lo + hi) >> 1; if (cdf[mid] > thr) hi = mid; else lo = mid + 1; }
      anc[k] = lo;
    }
  }
  __syncthreads();
  const double even = lse - log((double)n_new);
  for (int k = tid; k < n_new; k += SMALL_THREADS) { anc_out[k] = anc[k]; lw[k] = even; }
  if (tid == 0) { red[0] = m; red[1] = mn; red[2] = nanf; red[3] = se; red[4] = lse; }
  for (int a = 0; a < arrs.n; ++a) {
    if (arrs.elem_bytes[a] == 0) {
      long long rb = arrs.row_bytes[a];
      if ((rb & 15) == 0) {
        long long units = rb >> 4, total = (long long)n_new * units;
        const uint4 *in = (const uint4 *)arrs.cur[a]; uint4 *out = (uint4 *)arrs.alt[a];
        for (long long t = tid; t < total; t += SMALL_THREADS) { long long k = t / units, c = t - k * units; out[t] = in[(long long)anc[k] * units + c]; }
      } else {
        long long total = (long long)n_new * rb;
        for (long long t = tid; t < total; t += SMALL_THREADS) { long long k = t / rb, c = t - k * rb; arrs.alt[a][t] = arrs.cur[a][(long long)anc[k] * rb + c]; }
      }
    } else {
      int eb = arrs.elem_bytes[a];
      long long total = (long long)n_new * arrs.n_rows[a];
      for (long long t = tid; t < total; t += SMALL_THREADS) {
        long long r = t / n_new, k = t - r * n_new;
        const unsigned char *src = arrs.cur[a] + (r * P + anc[k]) * eb;
        unsigned char *dst = arrs.alt[a] + (r * n_new + k) * eb;
        if (eb == 8) *(uint64_t *)dst = *(const uint64_t *)src;
        else if (eb == 4) *(uint32_t *)dst = *(const uint32_t *)src;
        else for (int b = 0; b < eb; ++b) dst[b] = src[b];
      }
    }
  }
}

static bool resample_small(vk_particles *h, int64_t n_new, int mode, const double *uniforms_dev, uint32_t step,
                           bool resize) {
  if (h->P > SMALL_P_MAX || n_new > SMALL_P_MAX) return false;
  // resize: state_arrays(false) reports the fresh n_new-sized buffers the caller installed as back buffers
  std::vector<StateArray> arrs = resize ? h->model->state_arrays(false) : h->model->state_arrays();
  if (arrs.size() > (size_t)SMALL_MAX_ARRAYS) return false;
  SmallArrays sa; sa.n = (int)arrs.size();
  for (int a = 0; a < sa.n; ++a) {
    sa.cur[a] = (const unsigned char *)*arrs[a].cur; sa.alt[a] = (unsigned char *)*arrs[a].alt;
    sa.row_bytes[a] = (long long)arrs[a].row_bytes; sa.n_rows[a] = (long long)arrs[a].n_rows;
    sa.elem_bytes[a] = (int)arrs[a].elem_bytes;
    if (sa.elem_bytes[a] == 0 && (long long)n_new * sa.row_bytes[a] > (1ll << 26)) return false;  // big rows: use the wide path
  }
  resample_small_kernel<<<1, SMALL_THREADS, 0, h->stream>>>(h->lw.p, (int)h->P, (int)n_new, mode, frac_bits_for(h->P), h->seed,
                                                   step, uniforms_dev, h->anc.p, h->red.p, sa);
  VK_CUDA(cudaGetLastError());
  h->lw_minmax_fresh = false;
  for (auto &a : arrs) { void *t = *a.cur; *a.cur = *a.alt; *a.alt = t; }
  return true;
}

// exact-size replacement of a per-particle buffer, keeping the first min(old, new) elements
template <typename T>
static void resize_keep(DevBuf<T> &b, size_t n_new, cudaStream_t st) {
  if (b.n == n_new) return;
  DevBuf<T> fresh;
  fresh.alloc(n_new);
  const size_t keep = b.n < n_new ? b.n : n_new;
  if (keep && b.p) VK_CUDA(cudaMemcpyAsync(fresh.p, b.p, keep * sizeof(T), cudaMemcpyDeviceToDevice, st));
  VK_CUDA(cudaStreamSynchronize(st));
  b.swap(fresh);
}
static size_t state_array_bytes(const StateArray &a, int64_t n) {
  return a.elem_bytes ? (size_t)n * a.n_rows * a.elem_bytes : (size_t)n * a.row_bytes;
}

// Whole single-GPU resample: max -> fixed-point tile sums (+ LSE) -> ancestors -> gather -> even weights.
// n_new != P (trace_set.py:202-219: `(resample 1000)` on a single-particle set is how the reference's
// programs grow the particle set, benchmarks/Makefile:6-9): the children are gathered into fresh buffers
// of n_new particles which the model then adopts (Model::resized); every per-particle array of the handle
// is re-sized and h->P becomes n_new.
void resample(vk_particles *h, int64_t n_new, int mode, const double *uniforms_dev) {
  VK_REQUIRE(n_new >= 1 && n_new <= ((int64_t)1 << 30), VK_EINVAL, "resample: n_new must be in [1, 2^30]");
  VK_REQUIRE(mode == VK_RESAMPLE_MULTINOMIAL || mode == VK_RESAMPLE_SYSTEMATIC, VK_EINVAL, "resample: bad mode");
  cudaStream_t st = h->stream;
  const int64_t P = h->P;
  const bool resize = n_new != P;
  std::vector<void *> fresh_cur, fresh_alt;
  std::vector<StateArray> rs_arrs;
  if (resize) {
    VK_REQUIRE(!h->p2p || !p2p_is_open(h), VK_EUNSUPPORTED, "resample: a peer-sharded particle set cannot change its size");
    rs_arrs = h->model->state_arrays(false);
    const int64_t cap = n_new > P ? n_new : P;
    if ((int64_t)h->lw.n < cap) resize_keep(h->lw, (size_t)cap, st);
    if ((int64_t)h->anc.n < n_new) h->anc.alloc((size_t)n_new);
    try {
      for (auto &a : rs_arrs) {                 // children land in fresh buffers of n_new particles
        void *c = nullptr, *b = nullptr;
        VK_CUDA(cudaMalloc(&c, state_array_bytes(a, n_new))); fresh_cur.push_back(c);
        VK_CUDA(cudaMalloc(&b, state_array_bytes(a, n_new))); fresh_alt.push_back(b);
      }
    } catch (...) {
      for (void *q : fresh_cur) cudaFree(q);
      for (void *q : fresh_alt) cudaFree(q);
      throw;
    }
    for (size_t i = 0; i < rs_arrs.size(); ++i) *rs_arrs[i].alt = fresh_cur[i];   // the old back buffer stays owned by the model
  }
  uint32_t step = h->next_step();
  bool done = resample_small(h, n_new, mode, uniforms_dev, step, resize);
  if (!done) {
    int S = frac_bits_for(P);
    double scale = ldexp(1.0, S);
    int n_tiles = (int)((P + SCAN_TILE - 1) / SCAN_TILE);
    if (h->red_u.n < (size_t)n_tiles + 1) h->red_u.alloc((size_t)n_tiles + 1);
    if (h->tile_exp.n < (size_t)n_tiles) h->tile_exp.alloc((size_t)n_tiles);
    // pass 1 (max / min / NaN flag) is skipped when the kernel that wrote the weights already reduced them
    if (!h->lw_minmax_fresh)
      lse_pass1_kernel<<<grid_for(P, 256, h->sm_count, 4), 256, 0, st>>>(h->lw.p, P, h->red.p, h->counters.p);
    h->lw_minmax_fresh = false;
    h->prof.mark(st, "(gap)");
    fixed_tile_sums_exp_kernel<<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, h->red.p, scale, h->red_u.p, h->tile_exp.p);
    h->prof.mark(st, "tile_sums");
    const bool raw_tiles = mode == VK_RESAMPLE_SYSTEMATIC && n_tiles <= 1024;   // each CTA reads all tile totals
    if (!raw_tiles) scan_tile_sums_lse_kernel<<<1, 1024, 0, st>>>(h->red_u.p, h->tile_exp.p, n_tiles, P, h->red.p);
    if (raw_tiles) {
      systematic_offspring_kernel<true><<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, h->red.p, scale, h->red_u.p, h->tile_exp.p,
                                                                          n_tiles, n_new, h->seed, step, uniforms_dev, h->anc.p);
    } else if (mode == VK_RESAMPLE_SYSTEMATIC) {
      systematic_offspring_kernel<false><<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, h->red.p, scale, h->red_u.p, h->tile_exp.p,
                                                                           n_tiles, n_new, h->seed, step, uniforms_dev, h->anc.p);
    } else {
      fixed_cdf_kernel<<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, 0.0, h->red.p, scale, h->red_u.p, h->cdf.p);
      launch_ancestor_search_single(h, n_new, mode, step, uniforms_dev, grid_for(n_new, 256, h->sm_count, 8));
    }
    VK_CUDA(cudaGetLastError());
    h->prof.mark(st, "offspring / cdf + search");
    const double log_n_new = log((double)n_new);
    if (!(gather_state(h, h->anc.p, n_new, h->lw.p, log_n_new, resize ? &rs_arrs : nullptr) & 2))
      fill_kernel<<<grid_for(n_new, 256, h->sm_count, 4), 256, 0, st>>>(h->lw.p, n_new, h->red.p, log_n_new);
    VK_CUDA(cudaGetLastError());
    h->prof.mark(st, "gather");
  }
  if (resize) {
    VK_CUDA(cudaStreamSynchronize(st));          // the gathers read the old buffers, which the model frees now
    h->model->resized(n_new, fresh_cur, fresh_alt);
    h->P = n_new;
    resize_keep(h->lw, (size_t)n_new, st);
    h->scratch_d.alloc((size_t)n_new);
    h->cdf.alloc((size_t)n_new);
    if ((int64_t)h->anc.n != n_new) resize_keep(h->anc, (size_t)n_new, st);
    if (h->acc_d.p) h->acc_d.alloc((size_t)n_new);
  }
  h->model->after_resample();
}

// single-shard variant: total read from cdf[n-1] on the device
__global__ void ancestor_search_single_kernel(const uint64_t *__restrict__ cdf, int64_t n,
                                              int64_t n_new, int mode, uint64_t seed,
                                              uint32_t step, const double *__restrict__ uniforms,
                                              int32_t *__restrict__ anc_out) {
  const uint64_t total = cdf[n - 1];
  double u_sys = 0.0;
  if (mode == VK_RESAMPLE_SYSTEMATIC) {
    if (uniforms) u_sys = uniforms[0];
    else { Stream r(seed, GLOBAL_PARTICLE, step); u_sys = r.uniform(); }
  }
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n_new;
       k += (int64_t)gridDim.x * blockDim.x) {
    uint64_t thr;
    if (mode == VK_RESAMPLE_SYSTEMATIC) {
      uint64_t U0 = (uint64_t)(u_sys * 4294967296.0);
      if (U0 > 0xFFFFFFFFull) U0 = 0xFFFFFFFFull;
      unsigned __int128 num = (((unsigned __int128)(uint64_t)k << 32) | U0) * (unsigned __int128)total;
      unsigned __int128 den = (unsigned __int128)(uint64_t)n_new << 32;
      thr = (uint64_t)(num / den);
    } else {
      double u;
      if (uniforms) u = uniforms[k];
      else { Stream r(seed, (uint64_t)k, step); u = r.uniform(); }
      uint64_t U = (uint64_t)(u * 9007199254740992.0);
      if (U > 9007199254740991ull) U = 9007199254740991ull;
      thr = (uint64_t)(((unsigned __int128)U * (unsigned __int128)total) >> 53);
    }
    anc_out[k] = (int32_t)upper_bound_u64(cdf, n, thr);
  }
}

void launch_ancestor_search_single(vk_particles *h, int64_t n_new, int mode, uint32_t step,
                                   const double *uniforms_dev, int grid) {
  ancestor_search_single_kernel<<<grid, 256, 0, h->stream>>>(h->cdf.p, h->P, n_new, mode, h->seed,
                                                             step, uniforms_dev, h->anc.p);
  VK_CUDA(cudaGetLastError());
}

__global__ void fill_value_kernel(double *out, int64_t n, double v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    out[i] = v;
}
void launch_fill_even_weights(vk_particles *h, int64_t n, double lse, double log_n_new) {
  fill_value_kernel<<<grid_for(n, 256, h->sm_count, 4), 256, 0, h->stream>>>(h->lw.p, n, lse - log_n_new);
  VK_CUDA(cudaGetLastError());
}

__global__ void axpy_kernel(double *y, const double *x, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    y[i] += x[i];
}
void launch_axpy(vk_particles *h, double *y, const double *x, int64_t n) {
  axpy_kernel<<<grid_for(n, 256, h->sm_count, 4), 256, 0, h->stream>>>(y, x, n);
  VK_CUDA(cudaGetLastError());
}

void launch_normalized_probs(vk_particles *h, double *out_dev) {
  device_logsumexp(h->lw.p, h->P, h->red.p, h->counters.p, h->sm_count, h->stream);
  normalized_probs_kernel<<<grid_for(h->P, 256, h->sm_count, 4), 256, 0, h->stream>>>(h->lw.p, h->P, h->red.p, out_dev);
  VK_CUDA(cudaGetLastError());
}

void launch_ancestor_search_shard(vk_particles *h, uint64_t offset, uint64_t total, int64_t n_new,
                                  int64_t k_begin, int64_t k_count, int mode, uint32_t step,
                                  const double *uniforms_dev, int32_t *anc_out_dev) {
  if (k_count <= 0) return;
  ancestor_search_kernel<<<grid_for(k_count, 256, h->sm_count, 8), 256, 0, h->stream>>>(
      h->cdf.p, h->P, offset, total, n_new, k_begin, k_count, mode, h->seed, step, uniforms_dev, anc_out_dev);
  VK_CUDA(cudaGetLastError());
}


// ---- multi-GPU resampling helpers ---------------------------
// Packed row of one particle = the concatenation of its slice of every state
// array, each segment padded to 16 bytes (what crosses NVLink in all_to_all).
static size_t seg_bytes(const StateArray &a) {
  size_t b = a.elem_bytes == 0 ? a.row_bytes : a.n_rows * a.elem_bytes;
  return (b + 15) & ~(size_t)15;
}
size_t packed_row_bytes(vk_particles *h) {
  size_t tot = 0;
  for (auto &a : h->model->state_arrays()) tot += seg_bytes(a);
  return tot;
}
// pack: out[k] <- particle anc[k];  unpack: particle slots[k] <- in[k]
__global__ void pack_rows_kernel(const unsigned char *__restrict__ arr, unsigned char *__restrict__ packed,
                                 const int32_t *__restrict__ idx, int64_t n, int64_t row_bytes,
                                 int64_t packed_bytes, int64_t off, int unpack) {
  if ((row_bytes & 15) == 0) {
    int64_t units = row_bytes >> 4, total = n * units;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
      int64_t k = t / units, c = t - k * units;
      uint4 *pk = (uint4 *)(packed + k * packed_bytes + off) + c;
      uint4 *ar = (uint4 *)(const_cast<unsigned char *>(arr) + (int64_t)idx[k] * row_bytes) + c;
      if (unpack) *ar = *pk; else *pk = *ar;
    }
  } else {
    int64_t total = n * row_bytes;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
      int64_t k = t / row_bytes, c = t - k * row_bytes;
      unsigned char *pk = packed + k * packed_bytes + off + c;
      unsigned char *ar = const_cast<unsigned char *>(arr) + (int64_t)idx[k] * row_bytes + c;
      if (unpack) *ar = *pk; else *pk = *ar;
    }
  }
}
__global__ void pack_soa_kernel(const unsigned char *__restrict__ arr, unsigned char *__restrict__ packed,
                                const int32_t *__restrict__ idx, int64_t n, int64_t P, int64_t n_rows,
                                int64_t elem, int64_t packed_bytes, int64_t off, int unpack) {
  int64_t total = n * n_rows;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = t / n, k = t - r * n;   // k fastest: coalesced on the SoA side
    unsigned char *pk = packed + k * packed_bytes + off + r * elem;
    unsigned char *ar = const_cast<unsigned char *>(arr) + (r * P + (int64_t)idx[k]) * elem;
    for (int64_t b = 0; b < elem; ++b) { if (unpack) ar[b] = pk[b]; else pk[b] = ar[b]; }
  }
}
void pack_state(vk_particles *h, const int32_t *idx_dev, int64_t n, void *packed_dev, bool unpack) {
  if (n <= 0) return;
  size_t pb = packed_row_bytes(h), off = 0;
  for (auto &a : h->model->state_arrays()) {
    if (a.elem_bytes == 0) {
      int64_t work = (a.row_bytes & 15) == 0 ? n * (int64_t)(a.row_bytes >> 4) : n * (int64_t)a.row_bytes;
      pack_rows_kernel<<<grid_for(work, 256, h->sm_count, 8), 256, 0, h->stream>>>(
          (const unsigned char *)*a.cur, (unsigned char *)packed_dev, idx_dev, n, (int64_t)a.row_bytes, (int64_t)pb,
          (int64_t)off, unpack ? 1 : 0);
    } else {
      pack_soa_kernel<<<grid_for(n * (int64_t)a.n_rows, 256, h->sm_count, 8), 256, 0, h->stream>>>(
          (const unsigned char *)*a.cur, (unsigned char *)packed_dev, idx_dev, n, h->P, (int64_t)a.n_rows,
          (int64_t)a.elem_bytes, (int64_t)pb, (int64_t)off, unpack ? 1 : 0);
    }
    off += seg_bytes(a);
  }
  VK_CUDA(cudaGetLastError());
}
// local max of the log weights (NaN ignored) -> red[0]
void shard_weight_max(vk_particles *h) {
  int grid = grid_for(h->P, 256, h->sm_count, 4);
  lse_pass1_kernel<<<grid, 256, 0, h->stream>>>(h->lw.p, h->P, h->red.p, h->counters.p);
  VK_CUDA(cudaGetLastError());
}
// fixed-point CDF relative to the GLOBAL max; red[0] := global max, red[3] := sum exp(lw - max)
__global__ void set_red0_kernel(double *red, double m) { red[0] = m; red[1] = m; red[2] = 0.0; }
void shard_build_cdf(vk_particles *h, double global_max, int frac_bits) {
  set_red0_kernel<<<1, 1, 0, h->stream>>>(h->red.p, global_max);
  int grid = grid_for(h->P, 256, h->sm_count, 4);
  lse_pass2_kernel<<<grid, 256, 0, h->stream>>>(h->lw.p, h->P, h->red.p, h->counters.p + 1);
  build_fixed_cdf(h, nullptr, global_max, frac_bits);
  VK_CUDA(cudaGetLastError());
}

#include "p2p_resample.cuh"

}  // namespace vk
