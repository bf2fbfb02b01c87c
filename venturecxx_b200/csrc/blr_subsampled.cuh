// Config C5: subsampled (austerity) MH for the logistic-regression model.
// Reference: backend/lite/infer/subsampled_mh.py
//   subsampledMixMH :76-111   global proposal, mu_0 = (log u - alpha)/N, permutation
//   sequentialTest  :118-171  minibatched t-test with finite-population correction
//   evalOneLocalSection :377-409  dllh_i = xi_i - rho_i of one local section
//   findGlobalBorder :326-368 global border = the `w` array node; local sections =
//                             (lookup w -> vector_dot -> log_odds_flip), one per observation
// evalOneLocalSection first refreshes the section to the OLD global value
// (updateValuesAtScaffold) and then regenerates it under the NEW one, so
//   dllh_i = log_logistic(s_i x_i.w') - log_logistic(s_i x_i.w)
// depends only on (w, w', x_i): no per-particle logit cache is read or written
// and "stale" sections need no bookkeeping -- the cache is simply marked dirty.
//
// Mapping.  Observations (local sections) shard across GPUs; the P particles
// are replicated.  LANES ARE PARTICLES: a warp streams minibatch observations,
// every lane holds one particle's weight vector in registers, the 8D-byte
// feature row of an observation is one broadcast load shared by the 32
// particles of the warp, and each lane accumulates its own (sum dllh,
// sum dllh^2) in registers -- no shuffles, no atomics.  Algorithmic traffic
// (8D+1)/P_tile bytes per (particle, observation) evaluation.
// The minibatch order is a keyed Feistel bijection on [0, N) (sampling without
// replacement, O(1) per index) shared by the particle batch; the reference
// draws np_rng.permutation per trace (:103) -- each particle's marginal chain
// law is the same.
#pragma once
#include "vk_common.cuh"
#include "sp_math.cuh"
#include "fast_math.cuh"
#include "subsample_common.cuh"

namespace vk {

// per-particle state of the transition in flight (all arrays of length P)
struct SSState {
  int *site; double *w_new; double *delta; double *mu0;
  double *cum, *cum2, *n, *k, *tstat;
  int *status;          // 0 active, 1 accept, 2 reject
  double *acc_total, *n_total;   // statistics over the transitions of a call
};

// global proposal (:79-88,102): site, w', log u, mu_0
// delta_mode 0: prior resimulation (the reference's behaviour for `normal` even with
//   useDeltaKernels, since NormalOutputPSP.hasDeltaKernel() is False, continuous.py:352-353);
// delta_mode 2: NormalDriftKernel (continuous.py:68-84), weight 0.
__global__ void ss_propose_kernel(const double *__restrict__ w, SSState S, int D, int64_t P, int64_t p_base,
                                  double mu0_prior, double s0_prior, double N_total, int block_mode,
                                  int t_index, int delta_mode, double eps, uint64_t seed, uint32_t step) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    Stream r(seed, (uint64_t)(p_base + p), step);
    double u_sel = r.uniform();
    double log_u = log(r.uniform());
    int j;
    if (block_mode == VK_BLOCK_ONE) { j = (int)(u_sel * (double)D); if (j >= D) j = D - 1; }
    else if (block_mode == VK_BLOCK_EACH) j = t_index % D;
    else j = block_mode;
    double w_old = w[(int64_t)j * P + p];
    double w_new;
    if (delta_mode == 2) {
      double nu = normal_sim(r, 0.0, s0_prior);
      w_new = mu0_prior + sqrt(1.0 - eps * eps) * (w_old - mu0_prior) + eps * nu;
    } else {
      w_new = normal_sim(r, mu0_prior, s0_prior);
    }
    S.site[p] = j; S.w_new[p] = w_new; S.delta[p] = w_new - w_old;
    S.mu0[p] = (log_u - 0.0) / N_total;   // alpha of the global section is 0 (no absorbing node in it)
    S.cum[p] = 0.0; S.cum2[p] = 0.0; S.n[p] = 0.0; S.k[p] = 0.0; S.tstat[p] = 0.0; S.status[p] = 0;
  }
}

// one minibatch round: positions [q0, q1) of the permutation; this shard owns global
// observation indices [obs_lo, obs_lo + N_local).  partial: [gridDim.x][P][2].
//
// A warp owns 32 particles (one per lane, weights in registers) and a strided stream of
// minibatch positions.  It works in chunks of 32 positions: every lane runs the Feistel
// permutation for ONE position (lane-parallel index generation), the 32 feature rows are
// fetched together with coalesced loads into the warp's shared-memory tile (32 rows in
// flight instead of one dependent DRAM round trip per observation), and the tile is then
// consumed row by row with broadcast reads.
template <int DT>
__global__ void __launch_bounds__(256)
ss_round_kernel(const double *__restrict__ Xr /*[N_local][D] signed rows*/, const double *__restrict__ w,
                SSState S, int D, int64_t P, int64_t q0, int64_t q1, int64_t obs_lo, int64_t N_local,
                FeistelKey key, double *__restrict__ partial, const int *__restrict__ gate) {
  if (*gate == 0) return;               // every sequential test of this transition has decided: a no-op round
  extern __shared__ double sm_dyn[];    // [warps][32 rows][D] tiles, then [warps][32][2] partial sums
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
  const int n_groups = (int)((P + 31) / 32);
  // warps are split between particle groups; streams of observations are interleaved over
  // (block, warp-within-group)
  const int g = warp % n_groups;
  const int streams_per_block = n_warps / n_groups;      // >= 1 (host guarantees n_groups <= n_warps)
  const int s_in_block = warp / n_groups;
  const int64_t p = (int64_t)g * 32 + lane;
  const bool live = (s_in_block < streams_per_block) && p < P && S.status[p] == 0;
  double wl[DT];
  double delta = 0.0; int site = 0;
#pragma unroll
  for (int j = 0; j < DT; ++j) wl[j] = (live && j < D) ? w[(int64_t)j * P + p] : 0.0;
  if (live) { delta = S.delta[p]; site = S.site[p]; }
  double s1 = 0.0, s2 = 0.0;
  const bool any_live = __any_sync(0xffffffffu, live);
  double *tile = sm_dyn + (size_t)warp * 32 * D;
  if (any_live && s_in_block < streams_per_block) {
    const int64_t stream = (int64_t)blockIdx.x * streams_per_block + s_in_block;
    const int64_t n_streams = (int64_t)gridDim.x * streams_per_block;
    for (int64_t qc = q0 + stream; qc < q1; qc += 32 * n_streams) {
      const int64_t q = qc + (int64_t)lane * n_streams;          // this lane's position of the chunk
      int64_t i = -1;
      if (q < q1) {
        i = (int64_t)feistel_perm((uint64_t)q, key) - obs_lo;
        if (i < 0 || i >= N_local) i = -1;                       // another shard's observation
      }
      const unsigned valid = __ballot_sync(0xffffffffu, i >= 0);
      if (valid == 0u) continue;
      __syncwarp();
      for (int idx = lane; idx < 32 * D; idx += 32) {            // D iterations, 32/D rows each
        int r = idx / D, j = idx - r * D;
        int64_t ir = __shfl_sync(0xffffffffu, i, r);
        tile[idx] = ir >= 0 ? __ldg(&Xr[ir * D + j]) : 0.0;
      }
      __syncwarp();
      for (int r = 0; r < 32; ++r) {
        if (!((valid >> r) & 1u)) continue;                      // warp-uniform
        const double *x = tile + r * D;
        double eta = 0.0;
#pragma unroll
        for (int j = 0; j < DT; ++j)
          if (j < D) eta = fma(wl[j], x[j], eta);                // broadcast reads
        if (live) {
          double d = fast_log_logistic(fma(delta, x[site], eta)) - fast_log_logistic(eta);
          s1 += d; s2 += d * d;
        }
      }
    }
  }
  // block combine in fixed (warp) order -> deterministic
  __syncthreads();
  double *sm_part = sm_dyn;
  double *slot = sm_part + ((size_t)warp * 32 + lane) * 2;
  slot[0] = s1; slot[1] = s2;
  __syncthreads();
  for (int64_t pp = threadIdx.x; pp < P; pp += blockDim.x) {
    int gg = (int)(pp / 32), ll = (int)(pp % 32);
    double a = 0.0, b = 0.0;
    for (int s = 0; s < streams_per_block; ++s) {
      int wv = s * n_groups + gg;
      a += sm_part[((size_t)wv * 32 + ll) * 2];
      b += sm_part[((size_t)wv * 32 + ll) * 2 + 1];
    }
    partial[((size_t)blockIdx.x * P + pp) * 2] = a;
    partial[((size_t)blockIdx.x * P + pp) * 2 + 1] = b;
  }
}

// sums[P][2] = sum over blocks of partial; one warp per output, lanes stride the blocks and
// combine with a fixed shuffle tree (deterministic)
__global__ void ss_reduce_kernel(const double *__restrict__ partial, int n_blocks, int64_t P,
                                 double *__restrict__ sums, const int *__restrict__ gate) {
  if (*gate == 0) return;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t t = warp; t < P * 2; t += n_warps) {
    double a = 0.0;
    for (int b = lane; b < n_blocks; b += 32) a += partial[(size_t)b * P * 2 + t];
    a = warp_sum(a);
    if (lane == 0) sums[t] = a;
  }
}

// sequentialTest bookkeeping of one particle for one finished minibatch (:132-168); returns the new status
__device__ __forceinline__ int ss_decide_particle(SSState &S, int64_t p, double sum1, double sum2, double n_end,
                                                  double N, int k0, double epsilon) {
  double cum = S.cum[p] + sum1, cum2 = S.cum2[p] + sum2;
  double k = S.k[p] + 1.0, n = n_end;
  S.cum[p] = cum; S.cum2[p] = cum2; S.k[p] = k; S.n[p] = n;
  double mx = cum / n, mx2 = cum2 / n, mu0 = S.mu0[p];
  int st = 0;
  if (k < (double)k0 && n < N) { /* no test during the first k0 minibatches */ }
  else if (n == N) st = (mx >= mu0) ? 1 : 2;
  else {
    double sx = sqrt((1.0 - (n - 1.0) / (N - 1.0)) * (mx2 - mx * mx) / (n - 1.0));
    double q;
    if (sx == 0.0) q = (mx >= mu0) ? 1.0 : 0.0;
    else { double ts = (mx - mu0) / sx; S.tstat[p] = ts; q = dev_student_t_cdf(ts, n - 1.0); }
    if (q < epsilon) st = 2;
    else if (q > 1.0 - epsilon) st = 1;
  }
  S.status[p] = st;
  return st;
}
__global__ void ss_decide_kernel(SSState S, const double *__restrict__ sums, int64_t P, double n_end,
                                 double N, int k0, double epsilon, int *__restrict__ n_active) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    if (S.status[p] != 0) continue;
    if (ss_decide_particle(S, p, sums[2 * p], sums[2 * p + 1], n_end, N, k0, epsilon) == 0) atomicAdd(n_active, 1);
  }
}

// Straggler rounds.  The loop over minibatches runs until the slowest sequential test has decided, and
// in ss_round_kernel a round with two undecided particles costs what a round with 64 costs (lanes are
// particles).  Once at most SS_COMPACT_MAX particles are undecided the lanes become OBSERVATIONS: every
// lane loads the feature row of its own minibatch position and evaluates the (few) undecided particles one
// after the other, their weights staged in shared memory.  Same per-minibatch sums (to rounding: the
// summation order differs), same sequence of decisions.
static const int SS_COMPACT_MAX = 8;
template <int DT>
__global__ void __launch_bounds__(256)
ss_round_compact_kernel(const double *__restrict__ Xr, const double *__restrict__ w, SSState S, int D, int64_t P,
                        int64_t q0, int64_t q1, int64_t obs_lo, int64_t N_local, FeistelKey key,
                        double *__restrict__ partial, int *__restrict__ gate,
                        int fused, unsigned int *__restrict__ counter, double N_total, int k0, double epsilon) {
  if (*gate == 0) return;
  __shared__ double w_sh[SS_COMPACT_MAX][DT];
  __shared__ double delta_sh[SS_COMPACT_MAX];
  __shared__ int site_sh[SS_COMPACT_MAX], pid_sh[SS_COMPACT_MAX], n_act_sh;
  __shared__ double red_sh[8][SS_COMPACT_MAX][2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
  {                                             // undecided particles in index order (P <= 256 = blockDim.x)
    __shared__ int cnt_sh[8];
    const bool und = threadIdx.x < P && S.status[threadIdx.x] == 0;
    const unsigned b = __ballot_sync(0xffffffffu, und);
    if (lane == 0) cnt_sh[warp] = __popc(b);
    __syncthreads();
    int before = 0, total = 0;
    for (int wv = 0; wv < n_warps; ++wv) { if (wv < warp) before += cnt_sh[wv]; total += cnt_sh[wv]; }
    const int slot = before + __popc(b & ((1u << lane) - 1u));
    if (und && slot < SS_COMPACT_MAX) {
      pid_sh[slot] = (int)threadIdx.x; delta_sh[slot] = S.delta[threadIdx.x]; site_sh[slot] = S.site[threadIdx.x];
    }
    if (threadIdx.x == 0) n_act_sh = total < SS_COMPACT_MAX ? total : SS_COMPACT_MAX;
  }
  __syncthreads();
  const int n_act = n_act_sh;
  for (int t = threadIdx.x; t < n_act * DT; t += blockDim.x) {
    const int a = t / DT, j = t - a * DT;
    w_sh[a][j] = j < D ? w[(int64_t)j * P + pid_sh[a]] : 0.0;
  }
  __syncthreads();
  double s1[SS_COMPACT_MAX], s2[SS_COMPACT_MAX];
#pragma unroll
  for (int a = 0; a < SS_COMPACT_MAX; ++a) { s1[a] = 0.0; s2[a] = 0.0; }
  const int64_t n_thr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t q = q0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < q1; q += n_thr) {
    const int64_t i = (int64_t)feistel_perm((uint64_t)q, key) - obs_lo;
    if (i < 0 || i >= N_local) continue;        // another shard's observation
    double x[DT];
    const double *row = Xr + i * D;
#pragma unroll
    for (int j = 0; j < DT; ++j) x[j] = j < D ? __ldg(&row[j]) : 0.0;
#pragma unroll
    for (int a = 0; a < SS_COMPACT_MAX; ++a) {
      if (a >= n_act) break;
      double eta = 0.0;
#pragma unroll
      for (int j = 0; j < DT; ++j) eta = fma(w_sh[a][j], x[j], eta);
      const double xs = __ldg(&row[site_sh[a]]);
      const double d = fast_log_logistic(fma(delta_sh[a], xs, eta)) - fast_log_logistic(eta);
      s1[a] += d; s2[a] += d * d;
    }
  }
#pragma unroll
  for (int a = 0; a < SS_COMPACT_MAX; ++a) {
    double u = warp_sum(s1[a]), v = warp_sum(s2[a]);
    if (lane == 0) { red_sh[warp][a][0] = u; red_sh[warp][a][1] = v; }
  }
  __syncthreads();
  // partial[block][p][2]: zero for decided particles, warp-ordered sums for the undecided ones
  for (int64_t pp = threadIdx.x; pp < P; pp += blockDim.x) {
    partial[((size_t)blockIdx.x * P + pp) * 2] = 0.0;
    partial[((size_t)blockIdx.x * P + pp) * 2 + 1] = 0.0;
  }
  __syncthreads();
  if (threadIdx.x < n_act) {
    double a_ = 0.0, b_ = 0.0;
    for (int wv = 0; wv < n_warps; ++wv) { a_ += red_sh[wv][threadIdx.x][0]; b_ += red_sh[wv][threadIdx.x][1]; }
    partial[((size_t)blockIdx.x * P + pid_sh[threadIdx.x]) * 2] = a_;
    partial[((size_t)blockIdx.x * P + pid_sh[threadIdx.x]) * 2 + 1] = b_;
  }
  if (!fused) return;
  // fused tail (single-GPU loop): the LAST block sums the blocks' partials in ss_reduce_kernel's order
  // (lanes stride the blocks, shuffle tree) and runs the sequential-test bookkeeping of the <= 8
  // undecided particles, so a straggler round is one launch instead of round + reduce + memset + decide
  __shared__ bool last_sh;
  __shared__ int still_sh[SS_COMPACT_MAX];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last_sh = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!last_sh) return;
  __threadfence();
  if (warp < n_act) {
    const int64_t p = pid_sh[warp];
    double a_ = 0.0, b_ = 0.0;
    for (int bb = lane; bb < (int)gridDim.x; bb += 32) {
      a_ += partial[((size_t)bb * P + p) * 2];
      b_ += partial[((size_t)bb * P + p) * 2 + 1];
    }
    a_ = warp_sum(a_); b_ = warp_sum(b_);
    if (lane == 0) still_sh[warp] = ss_decide_particle(S, p, a_, b_, (double)q1, N_total, k0, epsilon) == 0 ? 1 : 0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int c = 0;
    for (int a = 0; a < n_act; ++a) c += still_sh[a];
    *gate = c;            // = the n_active the host reads back
    *counter = 0u;
  }
}

// accept: install w' (operator.accept, :108-109); reject: nothing to restore (the
// proposal never left the transition state).  Accumulate per-call statistics.
__global__ void ss_commit_kernel(double *__restrict__ w, SSState S, int64_t P) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    if (S.status[p] == 1) { w[(int64_t)S.site[p] * P + p] = S.w_new[p]; S.acc_total[p] += 1.0; }
    S.n_total[p] += S.n[p];
  }
}

__global__ void ss_stats_kernel(SSState S, int64_t P, int K, double *__restrict__ out) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    out[3 * p] = S.acc_total[p];
    out[3 * p + 1] = K > 0 ? S.n_total[p] / (double)K : 0.0;
    double ts = S.tstat[p];
    // :170  tstat if tstat else (inf if accept else -inf)
    out[3 * p + 2] = (ts != 0.0) ? ts : (S.status[p] == 1 ? INFINITY : -INFINITY);
  }
}

// Xr[i][j] = Xs[j][i]  (row-major copy of the signed features for row-contiguous access)
__global__ void ss_transpose_kernel(const double *__restrict__ Xs, double *__restrict__ Xr, int D,
                                    int64_t Ncap, int64_t N) {
  int64_t total = N * D;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    int64_t i = t / D; int j = (int)(t - i * D);
    Xr[t] = Xs[(int64_t)j * Ncap + i];
  }
}

}  // namespace vk
