// Global resampling across the GPUs of one NVSwitch node WITHOUT collectives or host round trips:
// the ranks (one process per GPU) exchange three scalars through mailboxes in each other's memory and
// move ancestor state with direct peer loads over NVLink.  Included at the end of k_weights.cu (it reuses
// that file's tile-sum / scan / offspring machinery); namespace vk.  Mailbox protocol, failure policy and the
// exchange itself: p2p_common.cuh.
//
// What it replaces: python/lib/engine/trace_set.py:202-235 (resample -> _resample_traces -> copy_trace)
// when the particle set is sharded.  The NCCL formulation of the same step
// (distributed.py sharded_resample: all_reduce(max), all_gather(mass), all_to_all(rows), three host
// synchronisations) stays as the portable path and as this one's parity reference.
//
// Resample number q on every rank is the SAME FOUR LAUNCHES as on one GPU; the exchanges ride in kernel tails:
//   A  the kernel that wrote the weights        local max / min / NaN flag; its last block runs EXCHANGE 0
//      (hmm_step_kernel) or lse_pass1           -> global max, min, NaN flag (every rank reduces the same
//                                                  'world' values in rank order: identical numbers everywhere)
//   B  p2p_tile_sums_kernel                     local fixed-point mass W_r and sum exp(lw - gmax) against the
//                                               GLOBAL max; its last block scans the tile totals and runs
//                                               EXCHANGE 1 -> offset_r = sum_{r'<r} W_r', W, LSE, every rank's
//                                               live-buffer index
//   C  offspring (systematic) or                each rank finds the children whose integer threshold falls in
//      cdf + search (multinomial)               ITS mass interval and stores the GLOBAL ancestor index straight
//                                               into the ancestor array of the rank that owns the child slot
//                                               (4-byte peer stores over NVLink)
//   D  gather                                   slot s waits until its ancestor has ARRIVED (the array holds -1
//                                               until a peer stores into it -- no barrier between C and D), then
//                                               pulls the ancestor's rows from the owning rank's LIVE buffers
//                                               (peer loads) into the local back buffers; buffers swap;
//                                               lw := LSE - log P_total
// Buffer safety: a rank's live buffers are read by peers only during D of resample q; they are next written
// by D of resample q+1, which that rank reaches only after exchange 1 of q+1, i.e. after every peer has
// finished D of q (stream order on the peer).  The ancestor array is double-buffered on the parity of q: D of
// q resets the OTHER half to -1 (its readers finished in q-1; its next writers, C of q+1, start only after
// this rank has posted exchange 1 of q+1).  Mailbox slots are double-buffered on the parity of the phase's
// own exchange counter.
//
// Integer thresholds on a fixed-point CDF make the ancestors independent of the number of GPUs: the result
// is bit-identical to vk_resample on one GPU holding all particles (tools/check_multi_gpu.py,
// tests/test_gpu_peer_memory.py).
#pragma once
#include <stdlib.h>
#include "p2p_common.cuh"

struct P2PArrays {                     // by-value kernel argument of the gather
  int n;
  unsigned char *alt[P2P_MAX_ARRAYS];
  long long row_bytes[P2P_MAX_ARRAYS];
  long long n_rows[P2P_MAX_ARRAYS];
  int elem_bytes[P2P_MAX_ARRAYS];
  int live_units[P2P_MAX_ARRAYS];      // row-major arrays: leading 16-byte units of a row that hold data (StateArray::live_bytes)
  const unsigned char *own_live[P2P_MAX_ARRAYS];   // this rank's live buffers (most ancestors are this rank's own)
  int patch_array, patch_src, patch_byte;   // Model::state_arrays_patched: byte patch_byte of every row of
                                            // array patch_array := low byte of the ancestor's element of
                                            // the 4-byte particle-minor array patch_src (patch_array < 0: none)
  double *lw; const double *red; double log_p_total;   // evened-out weights written by the same kernel
  int32_t *anc_other;                  // the ancestor half of the previous resample: reset to -1 here
  int world, rank;
  long long timeout_ns;
  volatile int *err_host;
  unsigned long long *prof;            // VK_P2P_PROFILE=1: phase times of CTA 0 of the fused kernel (ns, accumulated)
};
struct P2P {
  int world = 0, rank = 0, n_arrays = 0;
  bool opened = false, replicated = false;
  unsigned long long n_resamples = 0;
  unsigned long long phase_seq[P2P_PHASES] = {0, 0};
  unsigned long long ss_seq = 0;                 // exchanges of the subsampled-MH mailbox so far
  long long ss_stride = 0;                       // doubles per (parity, sender) slot of the ss mailbox
  long long ss_pcap = 1;                         // sequence words per (parity, sender): one per particle
  long long timeout_ns = 20000000000ll;
  DevBuf<P2PMail> mail;
  DevBuf<int32_t> anc_g;                         // [2][P]
  DevBuf<P2PTables> tab;
  DevBuf<P2PParams> par;
  DevBuf<unsigned long long> prof;               // VK_P2P_PROFILE=1
  DevBuf<unsigned long long> bar;                // arrival counter of the fused resample kernel's grid barrier (monotone)
  unsigned long long bar_count = 0;
  DevBuf<double> ss_mail;                        // [2][P2P_MAX_WORLD][ss_stride]
  DevBuf<unsigned long long> ss_seq_words;       // [2][P2P_MAX_WORLD]
  int *err_host = nullptr, *err_host_dev = nullptr;   // mapped pinned word and its device alias
  void *own_buf[P2P_MAX_ARRAYS][2];
  std::vector<void *> imported;
};
static const int P2P_FIXED_HANDLES = 4;          // mailbox, ancestor array, ss mailbox, ss sequence words

// Tile totals against the GLOBAL max (fixed_tile_sums_exp_kernel's work).  Prologue of every CTA: collect exchange 0
// (the peers' max / min / NaN flag, sent from the tail of the kernel that wrote the weights).  The last block then
// stores the global values to red[0..2], scans the tile totals, forms this rank's sum exp (scan_tile_sums_lse_body)
// and SENDS exchange 1 -- one launch, no single-CTA kernel of its own, no wait in the tail.
template <int ITEMS>
__global__ void __launch_bounds__(SCAN_THREADS)
p2p_tile_sums_kernel(const double *__restrict__ lw, int64_t n, double *__restrict__ red, double scale,
                     uint64_t *__restrict__ tile_sums, double *__restrict__ tile_exp, int n_tiles,
                     unsigned int *__restrict__ counter, const P2PXchg x0, const P2PXchg xchg) {
  __shared__ double mm_sh[3];
  __shared__ int ok_sh;
  if (threadIdx.x < 32) {
    double ra, rb, mx, mn, nanf;
    unsigned long long rc;
    const bool ok = p2p_collect_warp(x0, 0, ra, rb, rc);
    p2p_reduce_minmax(x0, ra, rb, rc, mx, mn, nanf);
    if (threadIdx.x == 0) { mm_sh[0] = mx; mm_sh[1] = mn; mm_sh[2] = nanf; ok_sh = ok ? 1 : 0; }
  }
  __syncthreads();
  if (!ok_sh) return;
  fixed_tile_sums_exp_body<ITEMS>(lw, n, mm_sh[0], scale, tile_sums, tile_exp);
  __shared__ bool last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!last) return;
  __threadfence();
  if (threadIdx.x == 0) { red[0] = mm_sh[0]; red[1] = mm_sh[1]; red[2] = mm_sh[2]; }   // every CTA has finished reading
  __syncthreads();
  scan_tile_sums_lse_body(tile_sums, tile_exp, n_tiles, n, red);
  __syncthreads();
  if (threadIdx.x == 0) *counter = 0;
  if (threadIdx.x < 32) p2p_send_warp(xchg, 1, red[3], (double)xchg.my_cur_idx, tile_sums[n_tiles]);
}

// Prologue of the kernels that consume exchange 1: every CTA's first warp collects the ranks' (sum exp, live buffer,
// mass) and forms this rank's mass offset and the global mass; CTA 0 also stores them (par, red[3..4]) for the gather
// and for vk_p2p_status.  Returns false (whole CTA) after a timeout.
__device__ __forceinline__ bool p2p_collect_mass(const P2PXchg &x, double *red, unsigned long long &offset,
                                                 unsigned long long &total) {
  __shared__ unsigned long long xm_sh[2];
  __shared__ int xok_sh;
  if (threadIdx.x < 32) {
    double ra, rb, se;
    unsigned long long rc, off, tot;
    const bool ok = p2p_collect_warp(x, 1, ra, rb, rc);
    p2p_reduce_mass(x, ra, rc, off, tot, se);
    if (threadIdx.x == 0) { xm_sh[0] = off; xm_sh[1] = tot; xok_sh = ok ? 1 : 0; }
    if (ok && blockIdx.x == 0) {
      if (threadIdx.x < x.world) x.par->cur_idx[threadIdx.x] = (int)rb;
      if (threadIdx.x == 0) {
        x.par->offset = off; x.par->total = tot;
        red[3] = se; red[4] = p2p_lse(x.n_total, red[0], red[1], red[2], se);
      }
    }
  }
  __syncthreads();
  offset = xm_sh[0]; total = xm_sh[1];
  return xok_sh != 0;
}

// Step C, systematic: as systematic_offspring_kernel, against the GLOBAL mass: this rank's running sum starts at
// par->offset, thresholds are those of n_total children over par->total, and child k's ancestor goes to the rank
// that owns slot k (k / P_local) as a GLOBAL particle index.
template <int ITEMS>
__global__ void __launch_bounds__(SCAN_THREADS)
p2p_offspring_kernel(const double *__restrict__ lw, int64_t n, const double *__restrict__ red, double scale,
                     const uint64_t *__restrict__ tile_sums, const P2PParams *__restrict__ par,
                     P2PTables *__restrict__ tab, int anc_half, int64_t n_total, int64_t p_base, uint64_t seed,
                     uint32_t step, const P2PXchg x1) {
  __shared__ uint64_t sh[32];
  if (((const volatile P2PParams *)par)->err) return;
  unsigned long long x_offset, W;
  if (!p2p_collect_mass(x1, const_cast<double *>(red), x_offset, W)) return;
  const double m = red[0];
  const uint64_t U0 = systematic_U0(nullptr, seed, step);
  int64_t base = (int64_t)blockIdx.x * (SCAN_THREADS * ITEMS) + (int64_t)threadIdx.x * ITEMS;
  uint64_t q[ITEMS];
  uint64_t s = 0;
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
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
  if (s == 0) return;
  uint64_t run = x_offset + tile_sums[blockIdx.x] + (w > 0 ? sh[w - 1] : 0ull) + inc - s;
  const double ratio = (double)n_total / (double)W, u0f = (double)U0 * (1.0 / 4294967296.0);
  int64_t k_lo = first_child_at_least(run, W, n_total, U0, ratio, u0f);
  const int64_t half_off = (int64_t)anc_half * n;
  // the thread's children are one contiguous range of slots: owner rank and offset advance by counting, the owner's
  // array pointer is fetched when the range crosses into the next rank (no division, no table load per child)
  int64_t d = k_lo / n, rem = k_lo - d * n;
  int32_t *dst = k_lo < n_total ? tab->anc[d] + half_off : nullptr;
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    if (q[j] == 0) continue;
    run += q[j];
    int64_t k_hi = first_child_at_least(run, W, n_total, U0, ratio, u0f);
    const int32_t g = (int32_t)(p_base + base + j);
    for (int64_t k = k_lo; k < k_hi; ++k) {
      dst[rem] = g;
      if (++rem == n) { rem = 0; ++d; dst = k + 1 < n_total ? tab->anc[d] + half_off : nullptr; }
    }
    k_lo = k_hi;
  }
}

// Step C, multinomial (the reference's law, trace_set.py:209-219): child k's threshold floor(U_k W / 2^53) with
// U_k from Philox(seed, k, step); every rank scans ALL children and keeps those landing in its mass interval
// (a local binary search each -- no remote CDF reads), storing the global ancestor at the owner of slot k.
__global__ void __launch_bounds__(256)
p2p_multinomial_kernel(const uint64_t *__restrict__ cdf, int64_t n, const P2PParams *__restrict__ par,
                       P2PTables *__restrict__ tab, int anc_half, int64_t n_total, int64_t p_base, uint64_t seed,
                       uint32_t step, double *__restrict__ red, const P2PXchg x1) {
  if (((const volatile P2PParams *)par)->err) return;
  unsigned long long offset, total;
  if (!p2p_collect_mass(x1, red, offset, total)) return;
  const uint64_t mass = cdf[n - 1];
  const int64_t half_off = (int64_t)anc_half * n;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n_total;
       k += (int64_t)gridDim.x * blockDim.x) {
    Stream r(seed, (uint64_t)k, step);
    double u = r.uniform();
    uint64_t U = (uint64_t)(u * 9007199254740992.0);
    if (U > 9007199254740991ull) U = 9007199254740991ull;
    uint64_t thr = (uint64_t)(((unsigned __int128)U * (unsigned __int128)total) >> 53);
    if (thr < offset || thr - offset >= mass) continue;
    int64_t i = upper_bound_u64(cdf, n, thr - offset);
    int64_t d = k / n;
    tab->anc[d][half_off + k - d * n] = (int32_t)(p_base + i);
  }
}

// the ancestor of child slot s, once a peer (or this rank) has stored it; -1 after a timeout
__device__ __forceinline__ int64_t p2p_wait_ancestor(const int32_t *anc_g, int64_t s, P2PParams *par,
                                                     long long timeout_ns, volatile int *err_host) {
  const volatile int32_t *slot = anc_g + s;
  int32_t g = *slot;
  if (g >= 0) return g;
  const unsigned long long t0 = p2p_now_ns();
  int spins = 0;
  while ((g = *slot) < 0) {
    if ((++spins & 63) == 0) {
      if (((volatile P2PParams *)par)->err) return -1;
      if ((long long)(p2p_now_ns() - t0) > timeout_ns) { par->err = 1; *err_host = 1; __threadfence_system(); return -1; }
    }
    __nanosleep(20);
  }
  return g;
}

// Slot-major gather of small state.  A thread owns GS child slots of a CTA tile: it loads their ancestor indices
// first, then every row unit and element of all GS slots, then stores -- GS independent chains of dependent loads
// per thread instead of one.  MAXU: the most 16-byte units any row-major array has live.  Returns false after a
// timed-out wait.
template <int GS, int MAXU>
__device__ __forceinline__ bool p2p_gather_slots(const int32_t *__restrict__ anc_g, int64_t n, P2PParams *__restrict__ par,
                                                 const P2PArrays &arrs, const P2PTables *__restrict__ tab,
                                                 const int *cur_idx, double even) {
  // the live buffer of array a on rank d: this rank's is a kernel argument; a peer's (children at a shard boundary)
  // costs two dependent loads from the tables.  cur_idx: every rank's live-buffer index (exchange 1)
  auto live = [&](int a, int d) -> const unsigned char * {
    return d == arrs.rank ? arrs.own_live[a] : tab->buf[a][d][cur_idx[d]];
  };
  const int64_t tile = (int64_t)blockDim.x * GS;
  const volatile int32_t *anc_v = anc_g;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < n; base += (int64_t)gridDim.x * tile) {
    int32_t i[GS];                       // global ancestor index, then the index within its owner's shard (P <= 2^30)
    int d[GS];
#pragma unroll
    for (int j = 0; j < GS; ++j) {
      const int64_t s = base + (int64_t)j * blockDim.x + threadIdx.x;
      i[j] = s < n ? anc_v[s] : 0;
    }
    const int32_t n32 = (int32_t)n, mine_lo = (int32_t)(arrs.rank * n);
#pragma unroll
    for (int j = 0; j < GS; ++j) {
      const int64_t s = base + (int64_t)j * blockDim.x + threadIdx.x;
      if (s < n && i[j] < 0) {
        i[j] = (int32_t)p2p_wait_ancestor(anc_g, s, par, arrs.timeout_ns, arrs.err_host);
        if (i[j] < 0) return false;
      }
      // owner rank of a global particle index: most ancestors are this rank's own
      const int32_t g = i[j];
      d[j] = arrs.rank;
      i[j] = g - mine_lo;
      if (i[j] < 0 || i[j] >= n32) { d[j] = g / n32; i[j] = g - d[j] * n32; }
    }
    for (int a = 0; a < arrs.n; ++a) {
      if (arrs.elem_bytes[a] == 0) {
        const long long rb = arrs.row_bytes[a];
        const int units = arrs.live_units[a];
        uint4 v[GS][MAXU];
        uint32_t sv[GS];
        const bool patched = a == arrs.patch_array;
#pragma unroll
        for (int j = 0; j < GS; ++j) {
          const int64_t s = base + (int64_t)j * blockDim.x + threadIdx.x;
          if (s >= n) continue;
          const uint4 *src = (const uint4 *)(live(a, d[j]) + (long long)i[j] * rb);
#pragma unroll
          for (int u = 0; u < MAXU; ++u) if (u < units) v[j][u] = src[u];
          if (patched) sv[j] = ((const uint32_t *)live(arrs.patch_src, d[j]))[i[j]] & 0xFFu;
        }
        const int pu = arrs.patch_byte >> 4, pw = (arrs.patch_byte >> 2) & 3, psh = (arrs.patch_byte & 3) * 8;
        const uint32_t keep = ~(0xFFu << psh);
#pragma unroll
        for (int j = 0; j < GS; ++j) {
          const int64_t s = base + (int64_t)j * blockDim.x + threadIdx.x;
          if (s >= n) continue;
          uint4 *dst = (uint4 *)(arrs.alt[a] + s * rb);
#pragma unroll
          for (int u = 0; u < MAXU; ++u) {
            if (u >= units) continue;
            if (patched && u == pu) {
              const uint32_t ins = sv[j] << psh;
              if (pw == 0) v[j][u].x = (v[j][u].x & keep) | ins;
              else if (pw == 1) v[j][u].y = (v[j][u].y & keep) | ins;
              else if (pw == 2) v[j][u].z = (v[j][u].z & keep) | ins;
              else v[j][u].w = (v[j][u].w & keep) | ins;
            }
            dst[u] = v[j][u];
          }
        }
      } else {
        const long long nr = arrs.n_rows[a];
        const int eb = arrs.elem_bytes[a];
        for (long long r = 0; r < nr; ++r) {
          uint64_t e8[GS];
#pragma unroll
          for (int j = 0; j < GS; ++j) {
            const int64_t s = base + (int64_t)j * blockDim.x + threadIdx.x;
            if (s >= n) continue;
            const unsigned char *src = live(a, d[j]) + (r * n + i[j]) * eb;
            if (eb == 8) e8[j] = *(const uint64_t *)src;
            else if (eb == 4) e8[j] = *(const uint32_t *)src;
            else { unsigned char *dst = arrs.alt[a] + (r * n + s) * eb; for (int e = 0; e < eb; ++e) dst[e] = src[e]; }
          }
#pragma unroll
          for (int j = 0; j < GS; ++j) {
            const int64_t s = base + (int64_t)j * blockDim.x + threadIdx.x;
            if (s >= n) continue;
            unsigned char *dst = arrs.alt[a] + (r * n + s) * eb;
            if (eb == 8) *(uint64_t *)dst = e8[j];
            else if (eb == 4) *(uint32_t *)dst = (uint32_t)e8[j];
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < GS; ++j) {
      const int64_t s = base + (int64_t)j * blockDim.x + threadIdx.x;
      if (s < n) { arrs.anc_other[s] = -1; arrs.lw[s] = even; }
    }
  }
  return true;
}

// Step D for small state (every array particle-minor, or rows of at most 64 live bytes in 16-byte units -- the HMM, the
// normal model, compiled programs): ONE THREAD PER CHILD SLOT copies everything that belongs to the slot; the ancestor
// is waited for and decoded once per slot, and the slot's weight and the other ancestor half are reset by the same
// thread.  A kernel of its own so that its register count (not the generic path's) sets the occupancy.
template <int GS, int MAXU>
__global__ void __launch_bounds__(256, 3)
p2p_gather_slots_kernel(const int32_t *__restrict__ anc_g, int64_t n, const P2PTables *__restrict__ tab,
                        P2PParams *__restrict__ par, P2PArrays arrs) {
  if (!((volatile P2PParams *)par)->err) {
    const double even = arrs.red[4] - arrs.log_p_total;      // lw := LSE - log P_total (utils.py:152-157)
    if (p2p_gather_slots<GS, MAXU>(anc_g, n, par, arrs, tab, par->cur_idx, even) && !((volatile P2PParams *)par)->err) return;
  }
  // a failed exchange poisons the weights instead
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
    arrs.anc_other[t] = -1; arrs.lw[t] = NAN;
  }
}

// Grid barrier of a cooperative launch: a monotone 64-bit counter (never reset; `target` = arrivals expected once this
// barrier is complete, tracked by the host).  Gives up when the sticky error word is set or the wall-clock bound trips
// (a CTA that timed out in a collect has left the kernel).
__device__ __forceinline__ bool p2p_grid_barrier(unsigned long long *bar, unsigned long long target, P2PParams *par,
                                                 long long timeout_ns, volatile int *err_host) {
  __shared__ int bar_ok_sh;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1ull);
    const unsigned long long t0 = p2p_now_ns();
    int ok = 1, spins = 0;
    while (*(volatile unsigned long long *)bar < target) {
      if ((++spins & 63) == 0) {
        if (((volatile P2PParams *)par)->err) { ok = 0; break; }
        if ((long long)(p2p_now_ns() - t0) > timeout_ns) { par->err = 1; *err_host = 1; __threadfence_system(); ok = 0; break; }
      }
    }
    __threadfence();
    bar_ok_sh = ok;
  }
  __syncthreads();
  return bar_ok_sh != 0;
}

// Steps B + C + D of a SYSTEMATIC sharded resample of small state in ONE cooperative launch (one tile per CTA, every CTA
// resident): collect exchange 0; fixed-point weights of the tile ONCE, kept in registers (the separate kernels compute
// exp(lw - m) twice and read the weights twice); tile totals -> grid barrier -> every CTA sums the tile totals itself
// (integers: any order gives the same CDF), CTA 0 sends exchange 1, every CTA collects it; offspring from the registers;
// a second grid barrier (this rank's ancestors are all stored); then the CTA turns to the gather of ITS child slots,
// waiting per slot only for ancestors that other RANKS are still storing.  Two kernel boundaries and one pass over the
// weights less than p2p_tile_sums_kernel + p2p_offspring_kernel + p2p_gather_slots_kernel.
template <int ITEMS, int GS, int MAXU>
__global__ void __launch_bounds__(SCAN_THREADS, 4)
p2p_resample_fused_kernel(const double *__restrict__ lw_in, int64_t n, double *__restrict__ red, double scale,
                          uint64_t *__restrict__ tile_sums, double *__restrict__ tile_exp,
                          unsigned long long *__restrict__ bar, unsigned long long bar_target, const P2PXchg x0,
                          const P2PXchg x1, P2PParams *__restrict__ par, P2PTables *__restrict__ tab, int anc_half,
                          int64_t n_total, int64_t p_base, uint64_t seed, uint32_t step,
                          const int32_t *__restrict__ anc_mine, P2PArrays arrs) {
  __shared__ double mm_sh[3];
  __shared__ int ok_sh;
  __shared__ uint64_t sh[32];
  __shared__ double shd[33];
  __shared__ uint64_t tot_sh[2];
  __shared__ double se_sh;
  __shared__ int cur_sh[P2P_MAX_WORLD];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  // a timed-out exchange or barrier: the particle set is invalid from here on; the weights say so (the sticky error
  // word makes the next host-synchronising call fail)
  auto poison = [&]() {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) arrs.lw[t] = NAN;
  };
  const bool prof = arrs.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0;
  unsigned long long tp = prof ? p2p_now_ns() : 0ull;
#define P2P_PROF(slot) do { if (prof) { const unsigned long long now_ = p2p_now_ns(); arrs.prof[slot] += now_ - tp; tp = now_; } } while (0)
  // ---- exchange 0 (sent from the tail of the kernel that wrote the weights)
  if (threadIdx.x < 32) {
    double ra, rb, mx, mn, nanf;
    unsigned long long rc;
    const bool ok = p2p_collect_warp(x0, 0, ra, rb, rc);
    p2p_reduce_minmax(x0, ra, rb, rc, mx, mn, nanf);
    if (threadIdx.x == 0) { mm_sh[0] = mx; mm_sh[1] = mn; mm_sh[2] = nanf; ok_sh = ok ? 1 : 0; }
  }
  __syncthreads();
  if (!ok_sh) { poison(); return; }
  P2P_PROF(0);
  const double m = mm_sh[0];
  // ---- fixed-point weights of this thread's particles (thread-contiguous, as p2p_offspring_kernel)
  const int64_t base = (int64_t)blockIdx.x * (SCAN_THREADS * ITEMS) + (int64_t)threadIdx.x * ITEMS;
  uint64_t q[ITEMS];
  uint64_t s = 0;
  double se = 0.0;
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const int64_t i = base + j;
    q[j] = 0ull;
    if (i < n) {
      const double v = lw_in[i];
      if (!(m > -INFINITY)) q[j] = 1ull;
      else if (m == INFINITY) q[j] = v == INFINITY ? 1ull : 0ull;
      else {
        const double d = v - m;
        if (d == d) { const double e = exp(d); q[j] = __double2ull_rn(e * scale); se += e; }
      }
    }
    s += q[j];
  }
  const uint64_t inc = warp_scan_incl(s, lane);
  if (lane == 31) sh[w] = inc;
  se = block_sum(se, shd);                       // (two __syncthreads inside: sh[] is complete after it)
  if (w == 0) {
    uint64_t t = lane < (SCAN_THREADS >> 5) ? sh[lane] : 0ull;
    t = warp_scan_incl(t, lane);
    sh[lane] = t;
    if (lane == (SCAN_THREADS >> 5) - 1) tile_sums[blockIdx.x] = t;
    if (lane == 0) tile_exp[blockIdx.x] = se;
  }
  // ---- every tile total is in memory
  P2P_PROF(1);
  if (!p2p_grid_barrier(bar, bar_target - gridDim.x, par, arrs.timeout_ns, arrs.err_host)) { poison(); return; }
  P2P_PROF(2);
  const int n_tiles = (int)gridDim.x;
  {
    uint64_t pre = 0, tot = 0;
    double te = 0.0;
    for (int b = threadIdx.x; b < n_tiles; b += SCAN_THREADS) {
      const uint64_t v = ((const volatile uint64_t *)tile_sums)[b];
      tot += v;
      if (b < (int)blockIdx.x) pre += v;
      te += ((const volatile double *)tile_exp)[b];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { pre += __shfl_xor_sync(0xffffffffu, pre, o); tot += __shfl_xor_sync(0xffffffffu, tot, o); }
    __shared__ uint64_t pt_sh[2][SCAN_THREADS / 32];
    if (lane == 0) { pt_sh[0][w] = pre; pt_sh[1][w] = tot; }
    te = block_sum(te, shd);                     // the same order in every CTA
    if (threadIdx.x == 0) {
      pre = 0; tot = 0;
      for (int k = 0; k < SCAN_THREADS / 32; ++k) { pre += pt_sh[0][k]; tot += pt_sh[1][k]; }
      tot_sh[0] = pre; tot_sh[1] = tot; se_sh = te;
    }
    __syncthreads();
  }
  P2P_PROF(3);
  // ---- exchange 1: CTA 0 sends (sum exp, live buffer, mass); every CTA collects
  if (blockIdx.x == 0 && threadIdx.x < 32) p2p_send_warp(x1, 1, se_sh, (double)x1.my_cur_idx, tot_sh[1]);
  if (threadIdx.x < 32) {
    double ra, rb, seg;
    unsigned long long rc, off, tot;
    const bool ok = p2p_collect_warp(x1, 1, ra, rb, rc);
    p2p_reduce_mass(x1, ra, rc, off, tot, seg);
    if (threadIdx.x < x1.world) cur_sh[threadIdx.x] = (int)rb;
    if (threadIdx.x == 0) { tot_sh[0] += off; tot_sh[1] = tot; se_sh = seg; ok_sh = ok ? 1 : 0; }
    if (ok && blockIdx.x == 0) {                 // for vk_p2p_status and the unfused kernels of a later call
      if (threadIdx.x < x1.world) par->cur_idx[threadIdx.x] = (int)rb;
      if (threadIdx.x == 0) {
        par->offset = off; par->total = tot;
        red[0] = mm_sh[0]; red[1] = mm_sh[1]; red[2] = mm_sh[2];
        red[3] = seg; red[4] = p2p_lse(x1.n_total, mm_sh[0], mm_sh[1], mm_sh[2], seg);
      }
    }
  }
  __syncthreads();
  if (!ok_sh) { poison(); return; }
  P2P_PROF(4);
  const uint64_t W = tot_sh[1];
  const double even = p2p_lse(x1.n_total, mm_sh[0], mm_sh[1], mm_sh[2], se_sh) - arrs.log_p_total;
  // ---- offspring (p2p_offspring_kernel's loop on the registers)
  if (s != 0) {
    const uint64_t U0 = systematic_U0(nullptr, seed, step);
    uint64_t run = tot_sh[0] + (w > 0 ? sh[w - 1] : 0ull) + inc - s;
    const double ratio = (double)n_total / (double)W, u0f = (double)U0 * (1.0 / 4294967296.0);
    int64_t k_lo = first_child_at_least(run, W, n_total, U0, ratio, u0f);
    const int64_t half_off = (int64_t)anc_half * n;
    int64_t d = k_lo / n, rem = k_lo - d * n;
    int32_t *dst = k_lo < n_total ? tab->anc[d] + half_off : nullptr;
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
      if (q[j] == 0) continue;
      run += q[j];
      const int64_t k_hi = first_child_at_least(run, W, n_total, U0, ratio, u0f);
      const int32_t g = (int32_t)(p_base + base + j);
      for (int64_t k = k_lo; k < k_hi; ++k) {
        dst[rem] = g;
        if (++rem == n) { rem = 0; ++d; dst = k + 1 < n_total ? tab->anc[d] + half_off : nullptr; }
      }
      k_lo = k_hi;
    }
  }
  P2P_PROF(5);
  // ---- every ancestor this rank produces is stored: the gather below polls only for what peers still owe (spinning
  // on local ancestors while neighbouring CTAs are still in their offspring loop costs more than this barrier:
  // 62.7 -> 59.7 us per step at a world of one)
  if (!p2p_grid_barrier(bar, bar_target, par, arrs.timeout_ns, arrs.err_host)) { poison(); return; }
  P2P_PROF(6);
  // ---- gather of this CTA's share of the child slots
  if (!p2p_gather_slots<GS, MAXU>(anc_mine, n, par, arrs, tab, cur_sh, even)) {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) arrs.lw[t] = NAN;
  }
  __syncthreads();
  P2P_PROF(7);
  if (prof) arrs.prof[8] += 1;
#undef P2P_PROF
}

// Step D: slot s copies its ancestor's state from the owning rank's live buffers (NVLink peer loads; rows that
// are multiples of 16 bytes move as uint4) into the local back buffers.
__global__ void __launch_bounds__(256)
p2p_gather_kernel(const int32_t *__restrict__ anc_g, int64_t n, const P2PTables *__restrict__ tab,
                  P2PParams *__restrict__ par, P2PArrays arrs) {
  const int64_t tid0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
  // every rank's LIVE buffer of every array, once per CTA (the tables live in global memory: read per element they
  // put two dependent loads in front of every peer load)
  __shared__ const unsigned char *live_sh[P2P_MAX_ARRAYS][P2P_MAX_WORLD];
  for (int e = threadIdx.x; e < P2P_MAX_ARRAYS * P2P_MAX_WORLD; e += blockDim.x) {
    const int a = e / P2P_MAX_WORLD, d = e - a * P2P_MAX_WORLD;
    live_sh[a][d] = (a < arrs.n && d < arrs.world) ? tab->buf[a][d][par->cur_idx[d]] : nullptr;
  }
  __syncthreads();
  for (int64_t t = tid0; t < n; t += stride) arrs.anc_other[t] = -1;
  bool failed = ((volatile P2PParams *)par)->err != 0;
  // owner rank of a global particle index: most ancestors are this rank's own
  auto owner = [&](int64_t g, int64_t &i) -> int {
    int d = arrs.rank;
    i = g - (int64_t)d * n;
    if (i < 0 || i >= n) { d = (int)(g / n); i = g - (int64_t)d * n; }
    return d;
  };
  for (int a = 0; a < arrs.n && !failed; ++a) {
    if (arrs.elem_bytes[a] == 0) {
      const long long rb = arrs.row_bytes[a];
      if ((rb & 15) == 0) {
        const long long units = rb >> 4;
        for (int64_t t = tid0; t < n * units; t += stride) {
          int64_t s = t / units, u = t - s * units;
          int64_t g = p2p_wait_ancestor(anc_g, s, par, arrs.timeout_ns, arrs.err_host);
          if (g < 0) { failed = true; break; }
          int64_t i;
          const int d = owner(g, i);
          const uint4 *src = (const uint4 *)(live_sh[a][d] + i * rb);
          uint4 v = src[u];
          if (a == arrs.patch_array && u == (arrs.patch_byte >> 4)) {
            const uint32_t sv = ((const uint32_t *)live_sh[arrs.patch_src][d])[i] & 0xFFu;
            const int pw = (arrs.patch_byte >> 2) & 3, psh = (arrs.patch_byte & 3) * 8;
            const uint32_t keep = ~(0xFFu << psh), ins = sv << psh;
            if (pw == 0) v.x = (v.x & keep) | ins;
            else if (pw == 1) v.y = (v.y & keep) | ins;
            else if (pw == 2) v.z = (v.z & keep) | ins;
            else v.w = (v.w & keep) | ins;
          }
          ((uint4 *)(arrs.alt[a] + s * rb))[u] = v;
        }
      } else {
        for (int64_t t = tid0; t < n * rb; t += stride) {
          int64_t s = t / rb, u = t - s * rb;
          int64_t g = p2p_wait_ancestor(anc_g, s, par, arrs.timeout_ns, arrs.err_host);
          if (g < 0) { failed = true; break; }
          int64_t i;
          const int d = owner(g, i);
          arrs.alt[a][s * rb + u] = live_sh[a][d][i * rb + u];
        }
      }
    } else {
      const long long nr = arrs.n_rows[a];
      const int eb = arrs.elem_bytes[a];
      for (int64_t t = tid0; t < n * nr; t += stride) {
        int64_t r = t / n, s = t - r * n;
        int64_t g = p2p_wait_ancestor(anc_g, s, par, arrs.timeout_ns, arrs.err_host);
        if (g < 0) { failed = true; break; }
        int64_t i;
        const int d = owner(g, i);
        const unsigned char *src = live_sh[a][d] + (r * n + i) * eb;
        unsigned char *dst = arrs.alt[a] + (r * n + s) * eb;
        if (eb == 8) *(uint64_t *)dst = *(const uint64_t *)src;
        else if (eb == 4) *(uint32_t *)dst = *(const uint32_t *)src;
        else for (int e = 0; e < eb; ++e) dst[e] = src[e];
      }
    }
  }
  // lw := LSE - log P_total (utils.py:152-157); a failed exchange poisons the weights instead
  const double even = (failed || ((volatile P2PParams *)par)->err) ? NAN : arrs.red[4] - arrs.log_p_total;
  for (int64_t t = tid0; t < n; t += stride) arrs.lw[t] = even;
}

// ---- host side -------------------------------------------------------------------------------------
static P2P *p2p_of(vk_particles *h) { return (P2P *)h->p2p; }

// raised by host-synchronising entry points once the stream has drained
void p2p_check_error(vk_particles *h) {
  P2P *s = p2p_of(h);
  if (s && s->err_host && *(volatile int *)s->err_host)
    throw Error(VK_ECUDA, "peer-memory exchange timed out: a peer GPU never arrived (rank " + std::to_string(s->rank) +
                              " of " + std::to_string(s->world) + "); the particle set of this handle is invalid");
}

// handles_out: (P2P_FIXED_HANDLES + 2 * n_arrays) IPC handles of 64 bytes: mailbox, ancestor array, ss mailbox,
// ss sequence words, then both physical buffers of every state array in state_arrays() order.  Returns the count.
int p2p_export(vk_particles *h, unsigned char *handles_out, int max_handles) {
  VK_REQUIRE(sizeof(cudaIpcMemHandle_t) == 64, VK_ECUDA, "unexpected cudaIpcMemHandle_t size");
  P2P *s = p2p_of(h);
  if (!s) { s = new P2P(); h->p2p = s; }
  VK_REQUIRE(!s->opened, VK_ESTATE, "p2p: already opened");
  std::vector<StateArray> arrs = h->model->state_arrays();
  VK_REQUIRE((int)arrs.size() <= P2P_MAX_ARRAYS, VK_EUNSUPPORTED, "p2p: too many state arrays");
  s->n_arrays = (int)arrs.size();
  int need = P2P_FIXED_HANDLES + 2 * s->n_arrays;
  VK_REQUIRE(max_handles >= need, VK_EINVAL, "p2p_export: handle buffer too small");
  VK_CUDA(cudaStreamSynchronize(h->stream));
  if (!s->mail.p) {
    s->mail.alloc((size_t)2 * P2P_PHASES * P2P_MAX_WORLD);
    VK_CUDA(cudaMemset(s->mail.p, 0, s->mail.bytes()));
    s->anc_g.alloc((size_t)2 * h->P);
    VK_CUDA(cudaMemset(s->anc_g.p, 0xFF, s->anc_g.bytes()));       // -1: nothing stored yet
    s->tab.alloc(1);
    s->par.alloc(1);
    VK_CUDA(cudaMemset(s->par.p, 0, sizeof(P2PParams)));
    // subsampled MH exchanges (sum dllh, sum dllh^2) of P2P_SS_MAX_ROUNDS minibatches per particle
    const bool ss = h->kind == VK_MODEL_BLR && h->P <= 4096;
    s->ss_stride = ss ? (long long)P2P_SS_MAX_ROUNDS * h->P * 2 : 2;
    s->ss_pcap = ss ? (long long)h->P : 1;
    s->ss_mail.alloc((size_t)2 * P2P_MAX_WORLD * s->ss_stride);
    s->ss_seq_words.alloc((size_t)2 * P2P_MAX_WORLD * s->ss_pcap);
    VK_CUDA(cudaMemset(s->ss_seq_words.p, 0, s->ss_seq_words.bytes()));
    VK_CUDA(cudaHostAlloc((void **)&s->err_host, sizeof(int), cudaHostAllocMapped));
    *s->err_host = 0;
    VK_CUDA(cudaHostGetDevicePointer((void **)&s->err_host_dev, s->err_host, 0));
    if (const char *e = getenv("VK_P2P_TIMEOUT_S")) {
      double sec = atof(e);
      if (sec > 0.0) s->timeout_ns = (long long)(sec * 1e9);
    }
  }
  cudaIpcMemHandle_t *out = (cudaIpcMemHandle_t *)handles_out;
  VK_CUDA(cudaIpcGetMemHandle(&out[0], s->mail.p));
  VK_CUDA(cudaIpcGetMemHandle(&out[1], s->anc_g.p));
  VK_CUDA(cudaIpcGetMemHandle(&out[2], s->ss_mail.p));
  VK_CUDA(cudaIpcGetMemHandle(&out[3], s->ss_seq_words.p));
  for (int a = 0; a < s->n_arrays; ++a) {
    s->own_buf[a][0] = *arrs[a].cur; s->own_buf[a][1] = *arrs[a].alt;
    VK_CUDA(cudaIpcGetMemHandle(&out[P2P_FIXED_HANDLES + 2 * a], s->own_buf[a][0]));
    VK_CUDA(cudaIpcGetMemHandle(&out[P2P_FIXED_HANDLES + 1 + 2 * a], s->own_buf[a][1]));
  }
  return need;
}

// all_handles: world blocks of n_handles handles (rank order), as gathered from p2p_export on every rank.
void p2p_open(vk_particles *h, int world, int rank, const unsigned char *all_handles, int n_handles) {
  P2P *s = p2p_of(h);
  VK_REQUIRE(s && s->mail.p, VK_ESTATE, "p2p_open: call vk_p2p_export first");
  VK_REQUIRE(world >= 1 && world <= P2P_MAX_WORLD && rank >= 0 && rank < world, VK_EINVAL, "p2p_open: bad world / rank");
  VK_REQUIRE(n_handles == P2P_FIXED_HANDLES + 2 * s->n_arrays, VK_EINVAL, "p2p_open: handle count differs from this model's");
  VK_REQUIRE((int64_t)world * h->P < ((int64_t)1 << 31), VK_EUNSUPPORTED, "p2p: global particle index must fit int32");
  // particle base rank * P: the particle set is block-partitioned (resample, E2); base 0 on every rank: the
  // particles are replicated and only the subsampled-MH mailbox is used (observations are sharded, E3)
  VK_REQUIRE(h->p_base == (int64_t)rank * h->P || h->p_base == 0, VK_ESTATE,
             "p2p_open: particle base must be rank * P (sharded) or 0 (replicated) (vk_set_particle_base)");
  s->replicated = h->p_base != (int64_t)rank * h->P;
  P2PTables t;
  memset(&t, 0, sizeof t);
  const cudaIpcMemHandle_t *hs = (const cudaIpcMemHandle_t *)all_handles;
  for (int r = 0; r < world; ++r) {
    void *ptrs[P2P_FIXED_HANDLES + 2 * P2P_MAX_ARRAYS];
    for (int k = 0; k < n_handles; ++k) {
      if (r == rank) {
        ptrs[k] = k == 0 ? (void *)s->mail.p : k == 1 ? (void *)s->anc_g.p : k == 2 ? (void *)s->ss_mail.p
                  : k == 3 ? (void *)s->ss_seq_words.p
                           : s->own_buf[(k - P2P_FIXED_HANDLES) / 2][(k - P2P_FIXED_HANDLES) & 1];
      } else {
        VK_CUDA(cudaIpcOpenMemHandle(&ptrs[k], hs[(size_t)r * n_handles + k], cudaIpcMemLazyEnablePeerAccess));
        s->imported.push_back(ptrs[k]);
      }
    }
    t.mail[r] = (P2PMail *)ptrs[0];
    t.anc[r] = (int32_t *)ptrs[1];
    t.ss_mail[r] = (double *)ptrs[2];
    t.ss_seq[r] = (unsigned long long *)ptrs[3];
    for (int a = 0; a < s->n_arrays; ++a) {
      t.buf[a][r][0] = (unsigned char *)ptrs[P2P_FIXED_HANDLES + 2 * a];
      t.buf[a][r][1] = (unsigned char *)ptrs[P2P_FIXED_HANDLES + 1 + 2 * a];
    }
  }
  VK_CUDA(cudaMemcpy(s->tab.p, &t, sizeof t, cudaMemcpyHostToDevice));
  s->world = world; s->rank = rank; s->opened = true;
  s->n_resamples = 0; s->phase_seq[0] = s->phase_seq[1] = 0; s->ss_seq = 0;
}

// the by-value description of the NEXT exchange of `phase` (advances that phase's counter); false when this
// handle's particle set is not sharded over peers
bool p2p_begin_exchange(vk_particles *h, int phase, P2PXchg *x) {
  P2P *s = p2p_of(h);
  x->tab = nullptr;
  if (!s || !s->opened) return false;
  x->tab = s->tab.p; x->par = s->par.p; x->err_host = s->err_host_dev;
  x->world = s->world; x->rank = s->rank; x->my_cur_idx = 0;
  x->seq = ++s->phase_seq[phase];
  x->n_total = (long long)s->world * h->P;
  x->timeout_ns = s->timeout_ns;
  return true;
}
bool p2p_is_open(vk_particles *h) { P2P *s = p2p_of(h); return s && s->opened; }
bool p2p_ss_view(vk_particles *h, P2PSSView *v) {
  P2P *s = p2p_of(h);
  if (!s || !s->opened || s->ss_pcap < h->P) return false;
  v->tab = s->tab.p; v->par = s->par.p; v->err_host_dev = s->err_host_dev;
  v->world = s->world; v->rank = s->rank;
  v->ss_stride = s->ss_stride; v->ss_pcap = s->ss_pcap; v->timeout_ns = s->timeout_ns;
  return true;
}

void p2p_resample(vk_particles *h, int mode) {
  P2P *s = p2p_of(h);
  VK_REQUIRE(s && s->opened, VK_ESTATE, "p2p_resample: peers not opened (vk_p2p_export / vk_p2p_open)");
  VK_REQUIRE(mode == VK_RESAMPLE_MULTINOMIAL || mode == VK_RESAMPLE_SYSTEMATIC, VK_EINVAL, "resample: bad mode");
  VK_REQUIRE(!s->replicated, VK_ESTATE, "p2p_resample: this handle's particles are replicated, not sharded");
  // validate everything that can throw BEFORE the step / exchange counters move or the model's pending
  // history column is handed over: a rank that throws here is still in lock step with its peers
  {
    std::vector<StateArray> chk = h->model->state_arrays_peek();
    VK_REQUIRE((int)chk.size() == s->n_arrays, VK_ESTATE, "p2p: state arrays changed since export");
    int ci = *chk[0].cur == s->own_buf[0][0] ? 0 : 1;
    for (int a = 0; a < s->n_arrays; ++a)
      VK_REQUIRE(*chk[a].cur == s->own_buf[a][ci] && *chk[a].alt == s->own_buf[a][1 - ci], VK_ESTATE,
                 "p2p: state buffers were re-allocated since export");
  }
  cudaStream_t st = h->stream;
  const int64_t P = h->P, P_total = (int64_t)s->world * P;
  int patch[3];
  std::vector<StateArray> arrs = h->model->state_arrays_patched(patch);
  if (patch[0] >= 0)
    VK_REQUIRE(arrs[patch[0]].elem_bytes == 0 && (arrs[patch[0]].row_bytes & 15) == 0 && arrs[patch[1]].elem_bytes == 4 &&
                   arrs[patch[1]].n_rows == 1, VK_EUNSUPPORTED, "p2p: pending-column patch needs 16-byte rows and a 4-byte source");
  const int cur_idx = *arrs[0].cur == s->own_buf[0][0] ? 0 : 1;
  const uint32_t step = h->next_step();
  const int half = (int)(++s->n_resamples & 1ull);
  const int S = frac_bits_for(P_total);
  const double scale = ldexp(1.0, S);
  // small per-GPU sets: 2 particles per thread instead of 8 (each kernel of the step is a single wave whose latency
  // is the serial work of one thread); the multinomial path shares fixed_cdf_kernel's tiling and keeps 8
  static const int items_env = getenv("VK_P2P_ITEMS") ? atoi(getenv("VK_P2P_ITEMS")) : 0;
  const bool small_tiles = mode == VK_RESAMPLE_SYSTEMATIC && (items_env ? items_env == 2 : P <= ((int64_t)1 << 18));
  const int tile = SCAN_THREADS * (small_tiles ? 2 : SCAN_ITEMS);
  const int n_tiles = (int)((P + tile - 1) / tile);
  if (h->red_u.n < (size_t)n_tiles + 1) h->red_u.alloc((size_t)n_tiles + 1);
  if (h->tile_exp.n < (size_t)n_tiles) h->tile_exp.alloc((size_t)n_tiles);
  P2PXchg x;
  h->prof.mark(st, "(gap)");
  if (!(h->lw_minmax_fresh && h->lw_x0_done)) {     // the producer of the weights did not run exchange 0 itself
    p2p_begin_exchange(h, 0, &x);
    lse_pass1_xchg_kernel<<<grid_for(P, 256, h->sm_count, 4), 256, 0, st>>>(h->lw.p, P, h->red.p, h->counters.p, x);
    h->prof.mark(st, "lse_pass1+send0");
  }
  h->lw_minmax_fresh = false; h->lw_x0_done = false;
  p2p_begin_exchange(h, 1, &x);
  x.my_cur_idx = cur_idx;
  P2PXchg x0 = x;                        // the exchange 0 just sent (by the weights' producer or by lse_pass1_xchg_kernel)
  x0.seq = s->phase_seq[0];
  // what the gather needs
  P2PArrays pa;
  memset(&pa, 0, sizeof pa);
  pa.n = s->n_arrays;
  pa.patch_array = patch[0]; pa.patch_src = patch[1]; pa.patch_byte = patch[2];
  pa.lw = h->lw.p; pa.red = h->red.p; pa.log_p_total = log((double)P_total);
  pa.anc_other = s->anc_g.p + (int64_t)(1 - half) * P;
  pa.timeout_ns = s->timeout_ns; pa.err_host = s->err_host_dev;
  pa.world = s->world; pa.rank = s->rank;
  int64_t work = P;
  bool slot_major = true;
  int max_units = 0;
  for (int a = 0; a < s->n_arrays; ++a) {
    pa.alt[a] = (unsigned char *)*arrs[a].alt;
    pa.own_live[a] = (const unsigned char *)*arrs[a].cur;
    pa.row_bytes[a] = (long long)arrs[a].row_bytes; pa.n_rows[a] = (long long)arrs[a].n_rows;
    pa.elem_bytes[a] = (int)arrs[a].elem_bytes;
    int64_t wa = arrs[a].elem_bytes ? P * (int64_t)arrs[a].n_rows
                                    : ((arrs[a].row_bytes & 15) == 0 ? P * (int64_t)(arrs[a].row_bytes >> 4) : P * (int64_t)arrs[a].row_bytes);
    work = wa > work ? wa : work;
    const size_t live = arrs[a].live_bytes ? arrs[a].live_bytes : arrs[a].row_bytes;
    pa.live_units[a] = arrs[a].elem_bytes ? 0 : (int)((live + 15) >> 4);
    if (arrs[a].elem_bytes == 0) {
      if ((arrs[a].row_bytes & 15) != 0 || pa.live_units[a] > 4) slot_major = false;
      max_units = pa.live_units[a] > max_units ? pa.live_units[a] : max_units;
    }
  }
  const int64_t anc_off = (int64_t)half * P;
  const int32_t *anc_mine = s->anc_g.p + anc_off;
  // small state, systematic: steps B + C + D as ONE cooperative launch when a CTA per tile fits on the GPU
  static const bool unfused_env = getenv("VK_P2P_UNFUSED") != nullptr;
  const void *fk = nullptr;
  int fused_tiles = n_tiles;
  if (mode == VK_RESAMPLE_SYSTEMATIC && slot_major && !unfused_env) {
    // 2 slots per thread and batch in the gather: 4 and 8 were measured slower (68 / 88 us per step against 60 at a
    // world of one: register spills, and the phases are bound by per-thread serial work, not by memory latency)
    if (small_tiles) fk = max_units <= 2 ? (const void *)p2p_resample_fused_kernel<2, 2, 2> : (const void *)p2p_resample_fused_kernel<2, 2, 4>;
    else fk = max_units <= 2 ? (const void *)p2p_resample_fused_kernel<SCAN_ITEMS, 2, 2> : (const void *)p2p_resample_fused_kernel<SCAN_ITEMS, 2, 4>;
    int coop = 0, per_sm = 0;
    VK_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device));
    VK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fk, SCAN_THREADS, 0));
    if (!coop || (int64_t)per_sm * h->sm_count < fused_tiles) fk = nullptr;
  }
  if (fk) {
    static const bool prof_env = getenv("VK_P2P_PROFILE") != nullptr;
    if (prof_env && !s->prof.p) { s->prof.alloc(16); VK_CUDA(cudaMemsetAsync(s->prof.p, 0, s->prof.bytes(), st)); }
    pa.prof = s->prof.p;
    if (!s->bar.p) { s->bar.alloc(2); VK_CUDA(cudaMemsetAsync(s->bar.p, 0, s->bar.bytes(), st)); s->bar_count = 0; }
    s->bar_count += 2ull * (unsigned long long)fused_tiles;      // two grid barriers per launch
    const double *lw_in = h->lw.p; double *red_p = h->red.p; uint64_t *ts_p = h->red_u.p; double *te_p = h->tile_exp.p;
    unsigned long long *bar_p = s->bar.p, bar_target = s->bar_count;
    P2PParams *par_p = s->par.p; P2PTables *tab_p = s->tab.p;
    int64_t n_arg = P, n_total_arg = P_total, p_base_arg = h->p_base;
    uint64_t seed_arg = h->seed; uint32_t step_arg = step; int half_arg = half; double scale_arg = scale;
    void *args[] = {&lw_in, &n_arg, &red_p, &scale_arg, &ts_p, &te_p, &bar_p, &bar_target, &x0, &x, &par_p, &tab_p, &half_arg,
                    &n_total_arg, &p_base_arg, &seed_arg, &step_arg, &anc_mine, &pa};
    VK_CUDA(cudaLaunchCooperativeKernel(fk, dim3(fused_tiles), dim3(SCAN_THREADS), args, 0, st));
    h->prof.mark(st, "p2p_resample_fused");
  } else {
  if (small_tiles)
    p2p_tile_sums_kernel<2><<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, h->red.p, scale, h->red_u.p, h->tile_exp.p, n_tiles,
                                                              h->counters.p + 2, x0, x);
  else
    p2p_tile_sums_kernel<SCAN_ITEMS><<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, h->red.p, scale, h->red_u.p, h->tile_exp.p,
                                                                       n_tiles, h->counters.p + 2, x0, x);
  h->prof.mark(st, "collect0+p2p_tile_sums+send1");
  if (mode == VK_RESAMPLE_SYSTEMATIC) {
    if (small_tiles)
      p2p_offspring_kernel<2><<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, h->red.p, scale, h->red_u.p, s->par.p, s->tab.p,
                                                                half, P_total, h->p_base, h->seed, step, x);
    else
      p2p_offspring_kernel<SCAN_ITEMS><<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, h->red.p, scale, h->red_u.p, s->par.p,
                                                                         s->tab.p, half, P_total, h->p_base, h->seed, step, x);
  } else {
    fixed_cdf_kernel<<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, 0.0, h->red.p, scale, h->red_u.p, h->cdf.p);
    p2p_multinomial_kernel<<<grid_for(P_total, 256, h->sm_count, 8), 256, 0, st>>>(h->cdf.p, P, s->par.p, s->tab.p, half,
                                                                                  P_total, h->p_base, h->seed, step, h->red.p, x);
  }
  h->prof.mark(st, mode == VK_RESAMPLE_SYSTEMATIC ? "collect1+p2p_offspring" : "cdf+collect1+p2p_multinomial");
  if (slot_major && max_units <= 2)
    p2p_gather_slots_kernel<4, 2><<<grid_for((P + 3) / 4, 256, h->sm_count, 3), 256, 0, st>>>(anc_mine, P, s->tab.p, s->par.p, pa);
  else if (slot_major)
    p2p_gather_slots_kernel<2, 4><<<grid_for((P + 1) / 2, 256, h->sm_count, 3), 256, 0, st>>>(anc_mine, P, s->tab.p, s->par.p, pa);
  else
    p2p_gather_kernel<<<grid_for(work, 256, h->sm_count, 8), 256, 0, st>>>(anc_mine, P, s->tab.p, s->par.p, pa);
  h->prof.mark(st, "p2p_gather");
  }
  VK_CUDA(cudaGetLastError());
  for (auto &a : arrs) { void *t = *a.cur; *a.cur = *a.alt; *a.alt = t; }
  h->model->after_resample();
}

// synchronises the stream; err = 1 if any exchange of this handle ever timed out; lse = global LSE of the last resample
void p2p_status(vk_particles *h, int *err_out, double *lse_out) {
  P2P *s = p2p_of(h);
  VK_REQUIRE(s && s->par.p, VK_ESTATE, "p2p_status: p2p not initialised");
  double red[5];
  VK_CUDA(cudaMemcpyAsync(red, h->red.p, sizeof red, cudaMemcpyDeviceToHost, h->stream));
  VK_CUDA(cudaStreamSynchronize(h->stream));
  if (err_out) *err_out = *(volatile int *)s->err_host;
  if (lse_out) *lse_out = red[4];
}

void p2p_release(vk_particles *h) {
  if (P2P *sp = p2p_of(h)) if (sp->prof.p) {     // phase times of CTA 0 of the fused resample kernel
    unsigned long long pr[16];
    if (cudaMemcpy(pr, sp->prof.p, sizeof pr, cudaMemcpyDeviceToHost) == cudaSuccess && pr[8])
      fprintf(stderr, "[vk p2p profile rank %d] fused launches %llu | us per launch: collect0 %.2f weights+tile %.2f barrier1 %.2f tile totals %.2f "
                      "send1+collect1 %.2f offspring %.2f barrier2 %.2f gather %.2f\n", sp->rank, pr[8], pr[0] * 1e-3 / pr[8], pr[1] * 1e-3 / pr[8],
              pr[2] * 1e-3 / pr[8], pr[3] * 1e-3 / pr[8], pr[4] * 1e-3 / pr[8], pr[5] * 1e-3 / pr[8], pr[6] * 1e-3 / pr[8], pr[7] * 1e-3 / pr[8]);
  }
  P2P *s = p2p_of(h);
  if (!s) return;
  for (void *p : s->imported) cudaIpcCloseMemHandle(p);
  if (s->err_host) cudaFreeHost(s->err_host);
  delete s;
  h->p2p = nullptr;
}
