// Global resampling across the GPUs of one NVSwitch node WITHOUT collectives or host round trips:
// the ranks (one process per GPU) exchange three scalars through mailboxes in each other's memory and
// move ancestor state with direct peer loads over NVLink.  Included at the end of k_weights.cu (it reuses
// that file's tile-sum / scan / offspring machinery); namespace vk.
//
// What it replaces: python/lib/engine/trace_set.py:202-235 (resample -> _resample_traces -> copy_trace)
// when the particle set is sharded.  The NCCL formulation of the same step
// (distributed.py sharded_resample: all_reduce(max), all_gather(mass), all_to_all(rows), three host
// synchronisations) stays as the portable path and as this one's parity reference.
//
// Protocol of resample number q (every rank runs the same stream-ordered kernel sequence):
//   A  lse_pass1                      local max / min / NaN flag
//   X0 exchange(phase 0)              -> global max, min, NaN flag (every rank reduces the same 'world' values
//                                        in rank order, so all ranks hold identical numbers)
//   B  fixed_tile_sums_exp, scan      local fixed-point mass W_r and sum exp(lw - gmax), against the GLOBAL max
//   X1 exchange(phase 1)              -> offset_r = sum_{r'<r} W_r', W = sum W_r, LSE, every rank's live-buffer index
//   C  offspring (systematic) or      each rank finds the children whose integer threshold falls in ITS mass
//      cdf + search (multinomial)     interval and stores the GLOBAL ancestor index straight into the anc array of
//                                     the rank that owns the child slot (4-byte peer stores over NVLink)
//   X2 exchange(phase 2)              barrier: every child slot everywhere has its ancestor
//   D  gather                         slot s pulls its ancestor's rows from the owning rank's LIVE buffers (peer
//                                     loads) into the local back buffers; buffers swap; lw := LSE - log P_total
// Buffer safety: a rank's live buffers are read by peers only during D of resample q; they are next written
// by D of resample q+1, which that rank reaches only after X0..X2 of q+1, i.e. after every peer has finished D
// of q (stream order on the peer).  Mailbox slots are double-buffered on the parity of q on top of that.
//
// Integer thresholds on a fixed-point CDF make the ancestors independent of the number of GPUs: the result
// is bit-identical to vk_resample on one GPU holding all particles (tools/check_multi_gpu.py).
#pragma once

static const int P2P_MAX_WORLD = 16;
static const int P2P_MAX_ARRAYS = 4;
static const int P2P_PHASES = 3;

struct P2PMail {                       // 32 bytes, written by ONE sender into its slot of a peer's mailbox
  unsigned long long seq;              // written last (after a system fence): 1 + resample number
  double a, b;
  unsigned long long c;
};
struct P2PParams {                     // device-resident results of the exchanges
  unsigned long long offset, total;    // this rank's mass offset, global mass
  int cur_idx[P2P_MAX_WORLD];          // which of its two physical buffers is each rank's live one
  int err;                             // 1 = a wait timed out (peer never arrived)
};
struct P2PTables {                     // device-resident peer pointers
  P2PMail *mail[P2P_MAX_WORLD];        // [parity][phase][sender] inside
  int32_t *anc[P2P_MAX_WORLD];
  unsigned char *buf[P2P_MAX_ARRAYS][P2P_MAX_WORLD][2];
};
struct P2PArrays {                     // by-value kernel argument of the gather
  int n;
  unsigned char *alt[P2P_MAX_ARRAYS];
  long long row_bytes[P2P_MAX_ARRAYS];
  long long n_rows[P2P_MAX_ARRAYS];
  int elem_bytes[P2P_MAX_ARRAYS];
  int patch_array, patch_src, patch_byte;   // Model::state_arrays_patched: byte patch_byte of every row of
                                            // array patch_array := low byte of the ancestor's element of
                                            // the 4-byte particle-minor array patch_src (patch_array < 0: none)
  double *lw; const double *red; double log_p_total;   // evened-out weights written by the same kernel
};
struct P2P {
  int world = 0, rank = 0, n_arrays = 0;
  bool opened = false;
  unsigned long long seq = 0;
  DevBuf<P2PMail> mail;
  DevBuf<int32_t> anc_g;
  DevBuf<P2PTables> tab;
  DevBuf<P2PParams> par;
  void *own_buf[P2P_MAX_ARRAYS][2];
  std::vector<void *> imported;
};

__device__ __forceinline__ int p2p_mail_index(int parity, int phase, int sender) {
  return (parity * P2P_PHASES + phase) * P2P_MAX_WORLD + sender;
}

// One warp; lane r talks to rank r.  Phase 0 sends (max, min, nan) from red[0..2]; phase 1 sends
// (sum exp, live-buffer index, mass); phase 2 is a pure barrier.  Results: phase 0 -> red[0..2]; phase 1 ->
// par (offset, total, cur_idx) and red[3] (global sum exp), red[4] (LSE, utils.py:77-91 edge cases).
__device__ __forceinline__ void
p2p_exchange_warp(P2PTables *tab, P2PParams *par, double *red, const uint64_t *tile_sums, int n_tiles,
                  int world, int rank, int phase, unsigned long long seq, int my_cur_idx, long long n_total,
                  long long spin_limit) {
  const int lane = threadIdx.x;
  const int parity = (int)(seq & 1ull);
  double a = 0.0, b = 0.0;
  unsigned long long c = 0ull;
  if (phase == 0) { a = red[0]; b = red[1]; c = red[2] != 0.0 ? 1ull : 0ull; }
  else if (phase == 1) { a = red[3]; b = (double)my_cur_idx; c = tile_sums[n_tiles]; }
  if (lane < world) {
    volatile P2PMail *dst = tab->mail[lane] + p2p_mail_index(parity, phase, rank);
    dst->a = a; dst->b = b; dst->c = c;
    __threadfence_system();
    dst->seq = seq;
  }
  double ra = 0.0, rb = 0.0;
  unsigned long long rc = 0ull;
  int bad = 0;
  if (lane < world) {
    volatile P2PMail *src = tab->mail[rank] + p2p_mail_index(parity, phase, lane);
    long long t0 = clock64();
    while (src->seq != seq) {
      if (clock64() - t0 > spin_limit) { bad = 1; break; }
      __nanosleep(64);
    }
    __threadfence_system();
    ra = src->a; rb = src->b; rc = src->c;
  }
  bad = __any_sync(0xffffffffu, bad);
  if (bad && lane == 0) par->err = 1;
  if (phase == 0) {
    double mx = lane < world ? ra : -INFINITY, mn = lane < world ? rb : INFINITY;
    unsigned long long nf = lane < world ? rc : 0ull;
    for (int o = 16; o > 0; o >>= 1) {
      mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      nf |= __shfl_xor_sync(0xffffffffu, nf, o);
    }
    if (lane == 0) { red[0] = mx; red[1] = mn; red[2] = nf ? 1.0 : 0.0; }
  } else if (phase == 1) {
    // rank-ordered sums (lane 0 walks the ranks) so that every GPU computes identical values
    unsigned long long off = 0ull, tot = 0ull;
    double se = 0.0;
    for (int r = 0; r < world; ++r) {
      unsigned long long m_r = __shfl_sync(0xffffffffu, rc, r);
      double se_r = __shfl_sync(0xffffffffu, ra, r);
      if (r < rank) off += m_r;
      tot += m_r;
      se += se_r;
    }
    if (lane < world) par->cur_idx[lane] = (int)rb;
    if (lane == 0) {
      par->offset = off; par->total = tot;
      double m = red[0], mn = red[1], nanf = red[2], lse;
      if (n_total == 0) lse = -INFINITY;
      else if (nanf != 0.0) lse = NAN;
      else if (isinf(m) && mn != -m) lse = m;
      else if (isinf(m)) lse = NAN;
      else lse = m + log(se);
      red[3] = se; red[4] = lse;
    }
  }
}

__global__ void __launch_bounds__(32)
p2p_exchange_kernel(P2PTables *tab, P2PParams *par, double *red, const uint64_t *tile_sums, int n_tiles,
                    int world, int rank, int phase, unsigned long long seq, int my_cur_idx, long long n_total,
                    long long spin_limit) {
  p2p_exchange_warp(tab, par, red, tile_sums, n_tiles, world, rank, phase, seq, my_cur_idx, n_total, spin_limit);
}
// scan of this rank's tile totals + local sum exp (scan_tile_sums_lse_kernel's work), then exchange 1 by
// warp 0 of the same single CTA: one launch instead of two on the latency-bound path
__global__ void __launch_bounds__(1024)
p2p_scan_exchange1_kernel(P2PTables *tab, P2PParams *par, double *red, uint64_t *tile_sums, const double *tile_exp,
                          int n_tiles, int64_t n_local, int world, int rank, unsigned long long seq, int my_cur_idx,
                          long long n_total, long long spin_limit) {
  scan_tile_sums_lse_body(tile_sums, tile_exp, n_tiles, n_local, red);
  __syncthreads();
  if (threadIdx.x < 32)
    p2p_exchange_warp(tab, par, red, tile_sums, n_tiles, world, rank, 1, seq, my_cur_idx, n_total, spin_limit);
}

// Step C, systematic: as systematic_offspring_kernel, against the GLOBAL mass: this rank's running sum starts at
// par->offset, thresholds are those of n_total children over par->total, and child k's ancestor goes to the rank
// that owns slot k (k / P_local) as a GLOBAL particle index.
__global__ void __launch_bounds__(SCAN_THREADS)
p2p_offspring_kernel(const double *__restrict__ lw, int64_t n, const double *__restrict__ red, double scale,
                     const uint64_t *__restrict__ tile_sums, const P2PParams *__restrict__ par,
                     P2PTables *__restrict__ tab, int64_t n_total, int64_t p_base, uint64_t seed, uint32_t step) {
  __shared__ uint64_t sh[32];
  const double m = red[0];
  const uint64_t W = par->total;
  const uint64_t U0 = systematic_U0(nullptr, seed, step);
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
  if (s == 0) return;
  uint64_t run = par->offset + tile_sums[blockIdx.x] + (w > 0 ? sh[w - 1] : 0ull) + inc - s;
  int64_t k_lo = first_child_at_least(run, W, n_total, U0);
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; ++j) {
    if (q[j] == 0) continue;
    run += q[j];
    int64_t k_hi = first_child_at_least(run, W, n_total, U0);
    const int32_t g = (int32_t)(p_base + base + j);
    for (int64_t k = k_lo; k < k_hi; ++k) {
      int64_t d = k / n;
      tab->anc[d][k - d * n] = g;
    }
    k_lo = k_hi;
  }
}

// Step C, multinomial (the reference's law, trace_set.py:209-219): child k's threshold floor(U_k W / 2^53) with
// U_k from Philox(seed, k, step); every rank scans ALL children and keeps those landing in its mass interval
// (a local binary search each -- no remote CDF reads), storing the global ancestor at the owner of slot k.
__global__ void __launch_bounds__(256)
p2p_multinomial_kernel(const uint64_t *__restrict__ cdf, int64_t n, const P2PParams *__restrict__ par,
                       P2PTables *__restrict__ tab, int64_t n_total, int64_t p_base, uint64_t seed, uint32_t step) {
  const uint64_t offset = par->offset, total = par->total;
  const uint64_t mass = cdf[n - 1];
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
    tab->anc[d][k - d * n] = (int32_t)(p_base + i);
  }
}

// Step D: slot s copies its ancestor's state from the owning rank's live buffers (NVLink peer loads; rows that
// are multiples of 16 bytes move as uint4) into the local back buffers.
__global__ void __launch_bounds__(256)
p2p_gather_kernel(const int32_t *__restrict__ anc_g, int64_t n, const P2PTables *__restrict__ tab,
                  const P2PParams *__restrict__ par, P2PArrays arrs) {
  for (int a = 0; a < arrs.n; ++a) {
    if (arrs.elem_bytes[a] == 0) {
      const long long rb = arrs.row_bytes[a];
      if ((rb & 15) == 0) {
        const long long units = rb >> 4;
        for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n * units;
             t += (int64_t)gridDim.x * blockDim.x) {
          int64_t s = t / units, u = t - s * units;
          int64_t g = anc_g[s], d = g / n, i = g - d * n;
          const uint4 *src = (const uint4 *)(tab->buf[a][d][par->cur_idx[d]] + i * rb);
          uint4 v = src[u];
          if (a == arrs.patch_array && u == (arrs.patch_byte >> 4)) {
            const uint32_t sv = ((const uint32_t *)tab->buf[arrs.patch_src][d][par->cur_idx[d]])[i] & 0xFFu;
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
        for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n * rb;
             t += (int64_t)gridDim.x * blockDim.x) {
          int64_t s = t / rb, u = t - s * rb;
          int64_t g = anc_g[s], d = g / n, i = g - d * n;
          arrs.alt[a][s * rb + u] = tab->buf[a][d][par->cur_idx[d]][i * rb + u];
        }
      }
    } else {
      const long long nr = arrs.n_rows[a];
      const int eb = arrs.elem_bytes[a];
      for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n * nr;
           t += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = t / n, s = t - r * n;
        int64_t g = anc_g[s], d = g / n, i = g - d * n;
        const unsigned char *src = tab->buf[a][d][par->cur_idx[d]] + (r * n + i) * eb;
        unsigned char *dst = arrs.alt[a] + (r * n + s) * eb;
        if (eb == 8) *(uint64_t *)dst = *(const uint64_t *)src;
        else if (eb == 4) *(uint32_t *)dst = *(const uint32_t *)src;
        else for (int e = 0; e < eb; ++e) dst[e] = src[e];
      }
    }
  }
  const double even = arrs.red[4] - arrs.log_p_total;   // lw := LSE - log P_total (utils.py:152-157)
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x)
    arrs.lw[t] = even;
}

// ---- host side -------------------------------------------------------------------------------------
static P2P *p2p_of(vk_particles *h) { return (P2P *)h->p2p; }

// handles_out: (2 + 2 * n_arrays) IPC handles of 64 bytes: mailbox, anc, then both physical buffers of every
// state array in state_arrays() order.  Returns the count.
int p2p_export(vk_particles *h, unsigned char *handles_out, int max_handles) {
  VK_REQUIRE(sizeof(cudaIpcMemHandle_t) == 64, VK_ECUDA, "unexpected cudaIpcMemHandle_t size");
  P2P *s = p2p_of(h);
  if (!s) { s = new P2P(); h->p2p = s; }
  VK_REQUIRE(!s->opened, VK_ESTATE, "p2p: already opened");
  std::vector<StateArray> arrs = h->model->state_arrays();
  VK_REQUIRE((int)arrs.size() <= P2P_MAX_ARRAYS, VK_EUNSUPPORTED, "p2p: too many state arrays");
  s->n_arrays = (int)arrs.size();
  int need = 2 + 2 * s->n_arrays;
  VK_REQUIRE(max_handles >= need, VK_EINVAL, "p2p_export: handle buffer too small");
  VK_CUDA(cudaStreamSynchronize(h->stream));
  if (!s->mail.p) {
    s->mail.alloc((size_t)2 * P2P_PHASES * P2P_MAX_WORLD);
    VK_CUDA(cudaMemset(s->mail.p, 0, s->mail.bytes()));
    s->anc_g.alloc((size_t)h->P);
    s->tab.alloc(1);
    s->par.alloc(1);
    VK_CUDA(cudaMemset(s->par.p, 0, sizeof(P2PParams)));
  }
  cudaIpcMemHandle_t *out = (cudaIpcMemHandle_t *)handles_out;
  VK_CUDA(cudaIpcGetMemHandle(&out[0], s->mail.p));
  VK_CUDA(cudaIpcGetMemHandle(&out[1], s->anc_g.p));
  for (int a = 0; a < s->n_arrays; ++a) {
    s->own_buf[a][0] = *arrs[a].cur; s->own_buf[a][1] = *arrs[a].alt;
    VK_CUDA(cudaIpcGetMemHandle(&out[2 + 2 * a], s->own_buf[a][0]));
    VK_CUDA(cudaIpcGetMemHandle(&out[3 + 2 * a], s->own_buf[a][1]));
  }
  return need;
}

// all_handles: world blocks of n_handles handles (rank order), as gathered from p2p_export on every rank.
void p2p_open(vk_particles *h, int world, int rank, const unsigned char *all_handles, int n_handles) {
  P2P *s = p2p_of(h);
  VK_REQUIRE(s && s->mail.p, VK_ESTATE, "p2p_open: call vk_p2p_export first");
  VK_REQUIRE(world >= 1 && world <= P2P_MAX_WORLD && rank >= 0 && rank < world, VK_EINVAL, "p2p_open: bad world / rank");
  VK_REQUIRE(n_handles == 2 + 2 * s->n_arrays, VK_EINVAL, "p2p_open: handle count differs from this model's");
  VK_REQUIRE((int64_t)world * h->P < ((int64_t)1 << 31), VK_EUNSUPPORTED, "p2p: global particle index must fit int32");
  VK_REQUIRE(h->p_base == (int64_t)rank * h->P, VK_ESTATE, "p2p_open: particle base must be rank * P (vk_set_particle_base)");
  P2PTables t;
  memset(&t, 0, sizeof t);
  const cudaIpcMemHandle_t *hs = (const cudaIpcMemHandle_t *)all_handles;
  for (int r = 0; r < world; ++r) {
    void *ptrs[2 + 2 * P2P_MAX_ARRAYS];
    for (int k = 0; k < n_handles; ++k) {
      if (r == rank) {
        ptrs[k] = k == 0 ? (void *)s->mail.p : k == 1 ? (void *)s->anc_g.p : s->own_buf[(k - 2) / 2][(k - 2) & 1];
      } else {
        VK_CUDA(cudaIpcOpenMemHandle(&ptrs[k], hs[(size_t)r * n_handles + k], cudaIpcMemLazyEnablePeerAccess));
        s->imported.push_back(ptrs[k]);
      }
    }
    t.mail[r] = (P2PMail *)ptrs[0];
    t.anc[r] = (int32_t *)ptrs[1];
    for (int a = 0; a < s->n_arrays; ++a) {
      t.buf[a][r][0] = (unsigned char *)ptrs[2 + 2 * a];
      t.buf[a][r][1] = (unsigned char *)ptrs[3 + 2 * a];
    }
  }
  VK_CUDA(cudaMemcpy(s->tab.p, &t, sizeof t, cudaMemcpyHostToDevice));
  s->world = world; s->rank = rank; s->opened = true; s->seq = 0;
}

void p2p_resample(vk_particles *h, int mode) {
  P2P *s = p2p_of(h);
  VK_REQUIRE(s && s->opened, VK_ESTATE, "p2p_resample: peers not opened (vk_p2p_export / vk_p2p_open)");
  VK_REQUIRE(mode == VK_RESAMPLE_MULTINOMIAL || mode == VK_RESAMPLE_SYSTEMATIC, VK_EINVAL, "resample: bad mode");
  cudaStream_t st = h->stream;
  const int64_t P = h->P, P_total = (int64_t)s->world * P;
  const uint32_t step = h->next_step();
  const unsigned long long seq = ++s->seq;
  const long long spin = 4000000000ll;   // ~2 s of SM clocks: a peer that never arrives sets par->err
  int patch[3];
  std::vector<StateArray> arrs = h->model->state_arrays_patched(patch);
  VK_REQUIRE((int)arrs.size() == s->n_arrays, VK_ESTATE, "p2p: state arrays changed since export");
  if (patch[0] >= 0)
    VK_REQUIRE(arrs[patch[0]].elem_bytes == 0 && (arrs[patch[0]].row_bytes & 15) == 0 && arrs[patch[1]].elem_bytes == 4 &&
                   arrs[patch[1]].n_rows == 1, VK_EUNSUPPORTED, "p2p: pending-column patch needs 16-byte rows and a 4-byte source");
  int cur_idx = *arrs[0].cur == s->own_buf[0][0] ? 0 : 1;
  for (int a = 0; a < s->n_arrays; ++a)
    VK_REQUIRE(*arrs[a].cur == s->own_buf[a][cur_idx] && *arrs[a].alt == s->own_buf[a][1 - cur_idx], VK_ESTATE,
               "p2p: state buffers were re-allocated since export");
  const int S = frac_bits_for(P_total);
  const double scale = ldexp(1.0, S);
  const int n_tiles = (int)((P + SCAN_TILE - 1) / SCAN_TILE);
  if (h->red_u.n < (size_t)n_tiles + 1) h->red_u.alloc((size_t)n_tiles + 1);
  if (h->tile_exp.n < (size_t)n_tiles) h->tile_exp.alloc((size_t)n_tiles);
  if (!h->lw_minmax_fresh)
    lse_pass1_kernel<<<grid_for(P, 256, h->sm_count, 4), 256, 0, st>>>(h->lw.p, P, h->red.p, h->counters.p);
  h->lw_minmax_fresh = false;
  p2p_exchange_kernel<<<1, 32, 0, st>>>(s->tab.p, s->par.p, h->red.p, h->red_u.p, n_tiles, s->world, s->rank, 0, seq,
                                        cur_idx, P_total, spin);
  fixed_tile_sums_exp_kernel<<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, h->red.p, scale, h->red_u.p, h->tile_exp.p);
  p2p_scan_exchange1_kernel<<<1, 1024, 0, st>>>(s->tab.p, s->par.p, h->red.p, h->red_u.p, h->tile_exp.p, n_tiles, P,
                                                s->world, s->rank, seq, cur_idx, P_total, spin);
  if (mode == VK_RESAMPLE_SYSTEMATIC) {
    p2p_offspring_kernel<<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, h->red.p, scale, h->red_u.p, s->par.p, s->tab.p,
                                                           P_total, h->p_base, h->seed, step);
  } else {
    fixed_cdf_kernel<<<n_tiles, SCAN_THREADS, 0, st>>>(h->lw.p, P, 0.0, h->red.p, scale, h->red_u.p, h->cdf.p);
    p2p_multinomial_kernel<<<grid_for(P_total, 256, h->sm_count, 8), 256, 0, st>>>(h->cdf.p, P, s->par.p, s->tab.p, P_total,
                                                                                  h->p_base, h->seed, step);
  }
  p2p_exchange_kernel<<<1, 32, 0, st>>>(s->tab.p, s->par.p, h->red.p, h->red_u.p, n_tiles, s->world, s->rank, 2, seq,
                                        cur_idx, P_total, spin);
  P2PArrays pa;
  memset(&pa, 0, sizeof pa);
  pa.n = s->n_arrays;
  pa.patch_array = patch[0]; pa.patch_src = patch[1]; pa.patch_byte = patch[2];
  pa.lw = h->lw.p; pa.red = h->red.p; pa.log_p_total = log((double)P_total);
  int64_t work = P;
  for (int a = 0; a < s->n_arrays; ++a) {
    pa.alt[a] = (unsigned char *)*arrs[a].alt;
    pa.row_bytes[a] = (long long)arrs[a].row_bytes; pa.n_rows[a] = (long long)arrs[a].n_rows;
    pa.elem_bytes[a] = (int)arrs[a].elem_bytes;
    int64_t wa = arrs[a].elem_bytes ? P * (int64_t)arrs[a].n_rows
                                    : ((arrs[a].row_bytes & 15) == 0 ? P * (int64_t)(arrs[a].row_bytes >> 4) : P * (int64_t)arrs[a].row_bytes);
    work = wa > work ? wa : work;
  }
  p2p_gather_kernel<<<grid_for(work, 256, h->sm_count, 8), 256, 0, st>>>(s->anc_g.p, P, s->tab.p, s->par.p, pa);
  VK_CUDA(cudaGetLastError());
  for (auto &a : arrs) { void *t = *a.cur; *a.cur = *a.alt; *a.alt = t; }
  h->model->after_resample();
}

// synchronises the stream; err = 1 if any exchange of this handle ever timed out; lse = global LSE of the last resample
void p2p_status(vk_particles *h, int *err_out, double *lse_out) {
  P2P *s = p2p_of(h);
  VK_REQUIRE(s && s->par.p, VK_ESTATE, "p2p_status: p2p not initialised");
  P2PParams p;
  VK_CUDA(cudaMemcpyAsync(&p, s->par.p, sizeof p, cudaMemcpyDeviceToHost, h->stream));
  double red[5];
  VK_CUDA(cudaMemcpyAsync(red, h->red.p, sizeof red, cudaMemcpyDeviceToHost, h->stream));
  VK_CUDA(cudaStreamSynchronize(h->stream));
  if (err_out) *err_out = p.err;
  if (lse_out) *lse_out = red[4];
}

void p2p_release(vk_particles *h) {
  P2P *s = p2p_of(h);
  if (!s) return;
  for (void *p : s->imported) cudaIpcCloseMemHandle(p);
  delete s;
  h->p2p = nullptr;
}
