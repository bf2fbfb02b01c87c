// Config C3: the HMM of examples/hmm.vnt:31-60 as a particle filter.
//   state(1) = init;  state(t) ~ categorical(T[state(t-1)-1], [1..S])
//   observe categorical(E[state(t)-1], [1..O]) = d_t
// Reference flow per time step (hmm.vnt:66-92, SURVEY App. A "C3"):
//   observe   -> evaluating state(t) SIMULATES the transition from the prior
//                (regen.py eval of a new mem family), the emission node waits
//                as an unpropagated observation            lite/trace.py:415-420
//   incorporate -> lw += log E[s_t][d_t]                    lite/trace.py:422-448,
//                                                          utils.py:106-118
//   resample  -> k_weights.cu
// Here the transition draw and the emission weight are ONE kernel (the draw
// is deferred from observe to incorporate; nothing can read the state in
// between).  Layout: cur[P] int32 (state at the newest time), hist[P][Tpad]
// uint8 rows (the whole mem table of `state`, which the reference copies with
// the trace on resample).  One thread per particle; HBM-bound:
// 4+4 B cur, 1 B hist, 8+8 B lw per particle-step.
#include "vk_common.cuh"
#include <utility>
#include "sp_math.cuh"
#include "lse_finish.cuh"
#include "weights.cuh"

namespace vk {

static const int HMM_MAX_S = 32;

struct HMMTables {
  int S, O, n_rows, init;
  const double *trans;  // n_rows x S
  const double *emit;   // S x O
};

// Step tables, built once per model by hmm_step_tables_kernel with the same device functions a
// per-particle evaluation would call (so the filter is bit-identical to the per-particle form):
//   [cdf n_rows*S: running sums categorical_sim compares against][denom n_rows][last_pos n_rows (as double)]
//   [lw_emit S*O: categorical_logpmf(o | state s), NaN -> -inf (lite/trace.py:438-448)]
__host__ __device__ inline int hmm_step_table_len(int n_rows, int S, int O) { return n_rows * S + 2 * n_rows + S * O; }
__global__ void hmm_step_tables_kernel(HMMTables tb, double *__restrict__ out) {
  double *cdf = out, *denom = cdf + tb.n_rows * tb.S, *last_pos = denom + tb.n_rows, *lw_emit = last_pos + tb.n_rows;
  for (int r_ = blockIdx.x * blockDim.x + threadIdx.x; r_ < tb.n_rows; r_ += gridDim.x * blockDim.x) {
    const double *ps = tb.trans + r_ * tb.S;
    double d = 0.0;
    for (int i = 0; i < tb.S; ++i) d += ps[i];
    double acc = 0.0; int last = 0;
    for (int i = 0; i < tb.S; ++i) { if (ps[i] > 0.0) last = i; acc += ps[i]; cdf[r_ * tb.S + i] = acc; }
    denom[r_] = d; last_pos[r_] = (double)last;
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < tb.S * tb.O; i += gridDim.x * blockDim.x) {
    double w = categorical_logpmf(i % tb.O, tb.emit + (i / tb.O) * tb.O, tb.O, 1);
    lw_emit[i] = (w != w) ? -INFINITY : w;
  }
}

// propagate one time step and weight by the emission.  Each thread owns HMM_STEP_ITEMS particles of a
// CTA tile and issues all of their loads before any arithmetic (one DRAM latency per thread instead of
// two per particle); the particle loop is Philox + S comparisons + two shared-memory table reads.
static const int HMM_STEP_ITEMS = 4;
__global__ void __launch_bounds__(256)
hmm_step_kernel(int32_t *__restrict__ cur, uint8_t *__restrict__ hist, int Tpad, int t /*0-based*/,
                double *__restrict__ lw, double *__restrict__ incr, HMMTables tb, const double *__restrict__ tables,
                int symbol /*1-based*/, int64_t P, int64_t p_base, uint64_t seed, uint32_t step,
                int accumulate_incr, double *__restrict__ red, unsigned int *__restrict__ counter, int write_hist,
                const P2PXchg xchg) {
  extern __shared__ double sh[];
  const int n_tr = tb.n_rows * tb.S + 2 * tb.n_rows;
  double *cdf = sh, *denom = cdf + tb.n_rows * tb.S, *last_pos = denom + tb.n_rows, *lw_emit = last_pos + tb.n_rows;
  for (int i = threadIdx.x; i < n_tr; i += blockDim.x) sh[i] = tables[i];
  for (int i = threadIdx.x; i < tb.S; i += blockDim.x) lw_emit[i] = tables[n_tr + i * tb.O + (symbol - 1)];
  __syncthreads();
  double mx = -INFINITY, mn = INFINITY, nanf = 0.0;
  const int64_t tile = (int64_t)blockDim.x * HMM_STEP_ITEMS;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < P; base += (int64_t)gridDim.x * tile) {
    int prev[HMM_STEP_ITEMS];
    double lwv[HMM_STEP_ITEMS], inc[HMM_STEP_ITEMS];
#pragma unroll
    for (int j = 0; j < HMM_STEP_ITEMS; ++j) {
      const int64_t p = base + (int64_t)j * blockDim.x + threadIdx.x;
      prev[j] = 1; lwv[j] = 0.0; inc[j] = 0.0;
      if (p < P) {
        if (t > 0) prev[j] = cur[p];
        lwv[j] = lw[p];
        if (incr && accumulate_incr) inc[j] = incr[p];
      }
    }
#pragma unroll
    for (int j = 0; j < HMM_STEP_ITEMS; ++j) {
      const int64_t p = base + (int64_t)j * blockDim.x + threadIdx.x;
      if (p >= P) continue;
      int s;
      if (t == 0) {
        s = tb.init;  // hmm.vnt:57-59: (if (<= t 1) 1 ...)
      } else {
        Stream r(seed, (uint64_t)(p_base + p), step);
        const int row = prev[j] - 1;
        const double u = r.uniform(), d = denom[row];
        if (!(d > 0.0)) {                       // categorical_sim's all-zero row: uniform
          int k = (int)(u * (double)tb.S);
          s = 1 + (k < tb.S ? k : tb.S - 1);
        } else {
          const double thr = u * d;
          const double *c = cdf + row * tb.S;
          int k = (int)last_pos[row];
          for (int i = tb.S - 1; i >= 0; --i) if (thr < c[i]) k = i;   // first i with thr < running sum
          s = 1 + k;
        }
      }
      cur[p] = s;
      // one byte into a particle-major row is a partial-sector write (a 32-byte read-modify-write per
      // particle); the newest column is left "pending" in cur[] when the caller can fold it into the
      // resampling gather that follows (hmm_gather_kernel), see HMMModel::hist_pending
      if (write_hist) hist[p * Tpad + t] = (uint8_t)s;
      const double w = lw_emit[s - 1];
      const double v = lwv[j] + w;
      lw[p] = v;
      if (v != v) nanf = 1.0; else { mx = fmax(mx, v); mn = fmin(mn, v); }
      if (incr) incr[p] = inc[j] + w;
    }
  }
  // the filter resamples right after this step: reduce (max, min, NaN flag) of the weights just written so
  // that vk_resample / vk_p2p_resample can skip their own read pass (lse_finish.cuh); when the particle set
  // is sharded over peer GPUs the last block also runs exchange 0 of that resample (xchg.tab != nullptr)
  lse_minmax_finish(mx, mn, nanf, red, counter, xchg);
}

// likelihood_weight: resimulate states 2..t from the prior, lw := sum_t log E
__global__ void __launch_bounds__(256)
hmm_lw_kernel(int32_t *__restrict__ cur, uint8_t *__restrict__ hist, int Tpad, int T,
              double *__restrict__ lw, HMMTables tb, const int32_t *__restrict__ symbols, int64_t P,
              int64_t p_base, uint64_t seed, uint32_t step, int write_lw) {
  extern __shared__ double sh[];
  double *tr = sh;
  double *em = sh + tb.n_rows * tb.S;
  for (int i = threadIdx.x; i < tb.n_rows * tb.S; i += blockDim.x) tr[i] = tb.trans[i];
  for (int i = threadIdx.x; i < tb.S * tb.O; i += blockDim.x) em[i] = tb.emit[i];
  __syncthreads();
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    Stream r(seed, (uint64_t)(p_base + p), step);
    int s = tb.init;
    double w = 0.0;
    for (int t = 0; t < T; ++t) {
      if (t > 0) s = 1 + categorical_sim(r, tr + (s - 1) * tb.S, tb.S, 1);
      hist[p * Tpad + t] = (uint8_t)s;
      w += categorical_logpmf(symbols[t] - 1, em + (s - 1) * tb.O, tb.O, 1);
    }
    if (T > 0) cur[p] = s;
    if (write_lw) lw[p] = w;
  }
}

#define HMM_NODES_INTERIOR 13
#define HMM_NODES_LAST 9
// single-site resimulation MH over the state sequence (scope default: one
// block per random choice = state(2..T)).  scaffold(state(t)): DRG {state(t)},
// absorbing {state(t+1) (its categorical parameter row changes), obs_t}.
__global__ void __launch_bounds__(256)
hmm_mh_kernel(int32_t *__restrict__ cur, uint8_t *__restrict__ hist, int Tpad, int T, HMMTables tb,
              const int32_t *__restrict__ symbols, int64_t P, int64_t p_base, uint64_t seed,
              uint32_t step0, int n_steps, int block_mode, long long *__restrict__ acc_out,
              double *__restrict__ nodes_out) {
  extern __shared__ double sh[];
  double *tr = sh;
  double *em = sh + tb.n_rows * tb.S;
  for (int i = threadIdx.x; i < tb.n_rows * tb.S; i += blockDim.x) tr[i] = tb.trans[i];
  for (int i = threadIdx.x; i < tb.S * tb.O; i += blockDim.x) em[i] = tb.emit[i];
  __syncthreads();
  const int nb = T - 1;  // blocks: state(2..T)
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    uint8_t *h = hist + p * Tpad;
    long long acc = 0;
    int n_last = 0;
    for (int k = 0; k < n_steps; ++k) {
      Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)k);
      double u_sel = r.uniform();
      double logu = log(r.uniform());
      int b;
      if (block_mode == VK_BLOCK_ONE) { b = (int)(u_sel * (double)nb); if (b >= nb) b = nb - 1; }
      else if (block_mode == VK_BLOCK_EACH) b = k % nb;
      else b = block_mode;
      n_last += (b == nb - 1);
      int t = b + 1;  // 0-based time index of the site
      int prev = h[t - 1], old = h[t];
      int nxt = (t + 1 < T) ? h[t + 1] : 0;
      double rho = categorical_logpmf(symbols[t] - 1, em + (old - 1) * tb.O, tb.O, 1);
      if (nxt) rho += categorical_logpmf(nxt - 1, tr + (old - 1) * tb.S, tb.S, 1);
      int s = 1 + categorical_sim(r, tr + (prev - 1) * tb.S, tb.S, 1);
      double xi = categorical_logpmf(symbols[t] - 1, em + (s - 1) * tb.O, tb.O, 1);
      if (nxt) xi += categorical_logpmf(nxt - 1, tr + (s - 1) * tb.S, tb.S, 1);
      double alpha;
      if (rho == -INFINITY && xi == -INFINITY) alpha = INFINITY;
      else if (rho == INFINITY && xi == INFINITY) alpha = -INFINITY;
      else alpha = xi - rho;
      if (logu < alpha) { h[t] = (uint8_t)s; ++acc; }
    }
    if (T > 0) cur[p] = h[T - 1];
    if (acc_out) acc_out[p] = acc;
    // transloop's return value (dispatch.py:109-128) = mean |regenCounts| (scaffold.py:67-68).  Measured
    // on the reference's Lite backend for examples/hmm.vnt (tests/golden/lite_chains.json, c3): the
    // scaffold of an interior state(t) has HMM_NODES_INTERIOR nodes, that of the newest state
    // HMM_NODES_LAST (no state(t+1) family hanging off it).  `each` floor-divides a sweep's sum by #blocks.
    if (nodes_out) {
      if (block_mode == VK_BLOCK_EACH)
        nodes_out[p] = (double)((HMM_NODES_INTERIOR * (nb - 1) + HMM_NODES_LAST) / nb);
      else
        nodes_out[p] = (double)(HMM_NODES_INTERIOR * (n_steps - n_last) + HMM_NODES_LAST * n_last) / (double)n_steps;
    }
  }
}

// Enumerative Gibbs on state(t) (infer/egibbs.py:42-107 under mixMH): candidates s = 1..S, weight
// exp(xi_s) = P(s | state(t-1)) P(obs_t | s) P(state(t+1) | s), new value drawn proportionally, always
// accepted (logAlpha = 0).  Draw layout as in MH: 0 block choice, 1 mixMH's accept uniform (unused), 2 the
// selection uniform.  Requires S <= 32.
__global__ void __launch_bounds__(256)
hmm_gibbs_kernel(int32_t *__restrict__ cur, uint8_t *__restrict__ hist, int Tpad, int T, HMMTables tb,
                 const int32_t *__restrict__ symbols, int64_t P, int64_t p_base, uint64_t seed,
                 uint32_t step0, int n_steps, int block_mode, long long *__restrict__ changed_out,
                 double *__restrict__ nodes_out) {
  extern __shared__ double sh[];
  double *tr = sh;
  double *em = sh + tb.n_rows * tb.S;
  for (int i = threadIdx.x; i < tb.n_rows * tb.S; i += blockDim.x) tr[i] = tb.trans[i];
  for (int i = threadIdx.x; i < tb.S * tb.O; i += blockDim.x) em[i] = tb.emit[i];
  __syncthreads();
  const int nb = T - 1, S = tb.S;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    uint8_t *h = hist + p * Tpad;
    long long ch = 0;
    int n_last = 0;
    double pr[32];
    for (int k = 0; k < n_steps; ++k) {
      Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)k);
      double u_sel = r.uniform();
      (void)r.uniform();
      int b;
      if (block_mode == VK_BLOCK_ONE) { b = (int)(u_sel * (double)nb); if (b >= nb) b = nb - 1; }
      else if (block_mode == VK_BLOCK_EACH) b = k % nb;
      else b = block_mode;
      n_last += (b == nb - 1);
      int t = b + 1;
      int prev = h[t - 1], old = h[t];
      int nxt = (t + 1 < T) ? h[t + 1] : 0;
      for (int s = 0; s < S; ++s) {
        double v = categorical_prob(s, tr + (prev - 1) * S, S) * categorical_prob(symbols[t] - 1, em + s * tb.O, tb.O);
        if (nxt) v *= categorical_prob(nxt - 1, tr + s * S, S);
        pr[s] = v;
      }
      int sn = 1 + gibbs_pick_small(pr, S, r.uniform());
      ch += (sn != old);
      h[t] = (uint8_t)sn;
    }
    if (T > 0) cur[p] = h[T - 1];
    if (changed_out) changed_out[p] = ch;
    if (nodes_out) {   // the scaffold is the MH one: same node counts
      if (block_mode == VK_BLOCK_EACH)
        nodes_out[p] = (double)((HMM_NODES_INTERIOR * (nb - 1) + HMM_NODES_LAST) / nb);
      else
        nodes_out[p] = (double)(HMM_NODES_INTERIOR * (n_steps - n_last) + HMM_NODES_LAST * n_last) / (double)n_steps;
    }
  }
}

// `(pgibbs seq ordered Pi K)`: PGibbsOperator under mixMH (infer/pgibbs.py:49-126, dispatch.py:197-212) -- conditional
// SMC over the time steps inside EVERY outer particle: Pi new inner particles next to the retained path.  Border segment
// t holds state(t+1) (0-based time t); segment 0 is the deterministic state(1), so nothing is drawn there and every
// inner particle starts with the weight of the first emission.  At every later segment each inner particle draws an
// ancestor from sampleLogCategorical over the Pi current weights + the retained path's weight of the previous segment
// (utils.py:170-180: exp(lw - max), normalizeList, categorical), inherits that path, draws its state from the
// transition row and is weighted by its emission (weights are NOT accumulated: they restart after the resampling).
// The final index is drawn from the Pi new weights, alpha = logsumexp(xi) - logsumexp(xi without the chosen one + rho)
// (:109-121) and mixMH accepts iff log u < alpha.  Draw layout: draw 0 unused (no block choice), draw 1 accept uniform,
// then the draws above in call order.  One thread per outer particle; paths live in local memory.
static const int HMM_PG_MAX = 16;    // inner particles
__device__ __forceinline__ int hmm_sample_log_categorical(Stream &r, const double *lw, int n) {
  double m = -INFINITY, ps[HMM_PG_MAX + 1], tot = 0.0;
  for (int i = 0; i < n; ++i) m = fmax(m, lw[i]);
  for (int i = 0; i < n; ++i) { ps[i] = m > -INFINITY ? exp(lw[i] - m) : 0.0; tot += ps[i]; }
  for (int i = 0; i < n; ++i) ps[i] = tot > 0.0 ? ps[i] / tot : 1.0 / (double)n;     // normalizeList, utils.py:93-99
  return categorical_sim(r, ps, n, 1);
}
__device__ __forceinline__ double hmm_lse(const double *v, int n) {                  // utils.py:70-91
  double m = -INFINITY;
  for (int i = 0; i < n; ++i) m = fmax(m, v[i]);
  if (isinf(m)) return m;
  double s = 0.0;
  for (int i = 0; i < n; ++i) s += exp(v[i] - m);
  return m + log(s);
}
__global__ void __launch_bounds__(128)
hmm_pgibbs_kernel(int32_t *__restrict__ cur, uint8_t *__restrict__ hist, int Tpad, int T, HMMTables tb,
                  const int32_t *__restrict__ symbols, int64_t P, int64_t p_base, uint64_t seed, uint32_t step0,
                  int n_steps, int Pi, long long *__restrict__ acc_out, double *__restrict__ nodes_out) {
  extern __shared__ double sh[];
  double *tr = sh;
  double *em = sh + tb.n_rows * tb.S;
  for (int i = threadIdx.x; i < tb.n_rows * tb.S; i += blockDim.x) tr[i] = tb.trans[i];
  for (int i = threadIdx.x; i < tb.S * tb.O; i += blockDim.x) em[i] = tb.emit[i];
  __syncthreads();
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    uint8_t *h = hist + p * Tpad;
    uint8_t path[2][HMM_PG_MAX][32];          // [buffer][inner particle][time]
    double xi[HMM_PG_MAX + 1], nw[HMM_PG_MAX];
    long long acc = 0;
    for (int k = 0; k < n_steps; ++k) {
      Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)k);
      (void)r.uniform();
      const double logu = log(r.uniform());
      int cb = 0;
      const double w0 = categorical_logpmf(symbols[0] - 1, em + (h[0] - 1) * tb.O, tb.O, 1);
      for (int i = 0; i < Pi; ++i) { xi[i] = w0; path[0][i][0] = h[0]; }
      double rho_prev = w0;
      for (int t = 1; t < T; ++t) {
        xi[Pi] = rho_prev;
        for (int i = 0; i < Pi; ++i) {
          const int a = hmm_sample_log_categorical(r, xi, Pi + 1);
          const uint8_t *src = a == Pi ? h : path[cb][a];
          for (int u = 0; u < t; ++u) path[cb ^ 1][i][u] = src[u];
          const int s = 1 + categorical_sim(r, tr + (src[t - 1] - 1) * tb.S, tb.S, 1);
          path[cb ^ 1][i][t] = (uint8_t)s;
          nw[i] = categorical_logpmf(symbols[t] - 1, em + (s - 1) * tb.O, tb.O, 1);
        }
        for (int i = 0; i < Pi; ++i) xi[i] = nw[i];
        cb ^= 1;
        rho_prev = categorical_logpmf(symbols[t] - 1, em + (h[t] - 1) * tb.O, tb.O, 1);
      }
      const int f = hmm_sample_log_categorical(r, xi, Pi);
      double others[HMM_PG_MAX + 1];
      int n_o = 0;
      for (int i = 0; i < Pi; ++i) if (i != f) others[n_o++] = xi[i];
      others[n_o++] = rho_prev;
      const double alpha = hmm_lse(xi, Pi) - hmm_lse(others, n_o);
      if (logu < alpha) { for (int u = 0; u < T; ++u) h[u] = path[cb][f][u]; ++acc; }
    }
    if (T > 0) cur[p] = h[T - 1];
    if (acc_out) acc_out[p] = acc;
    // one scaffold over every state site: the sites' node counts add up (Lite: 13 per interior state, 9 for the newest)
    if (nodes_out) nodes_out[p] = (n_steps > 0 && T > 1) ? (double)(HMM_NODES_INTERIOR * (T - 2) + HMM_NODES_LAST) : 0.0;
  }
}

__global__ void hmm_loglik_kernel(const uint8_t *__restrict__ hist, int Tpad, int T, HMMTables tb,
                                  const int32_t *__restrict__ symbols, int64_t P, int joint, int t_first,
                                  double *__restrict__ out) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    const uint8_t *h = hist + p * Tpad;
    double w = 0.0;
    for (int t = t_first; t < T; ++t) {
      w += categorical_logpmf(symbols[t] - 1, tb.emit + (h[t] - 1) * tb.O, tb.O, 1);
      if (joint && t > 0) w += categorical_logpmf(h[t] - 1, tb.trans + (h[t - 1] - 1) * tb.S, tb.S, 1);
    }
    out[p] = w;
  }
}

__global__ void hmm_read_kernel(const uint8_t *__restrict__ hist, int Tpad, int t, int64_t P, double *out) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x)
    out[p] = (double)hist[p * Tpad + t];
}
__global__ void hmm_write_kernel(uint8_t *__restrict__ hist, int32_t *__restrict__ cur, int Tpad, int t,
                                 int is_last, int64_t P, const double *in) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    int s = (int)in[p];
    hist[p * Tpad + t] = (uint8_t)s;
    if (is_last) cur[p] = s;
  }
}
__global__ void hmm_fill_kernel(double *out, int64_t n, double v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    out[i] = v;
}

// hist[p][t] := cur[p]: materialises a pending newest column
__global__ void hmm_flush_kernel(const int32_t *__restrict__ cur, uint8_t *__restrict__ hist, int Tpad, int t, int64_t P) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x)
    hist[p * Tpad + t] = (uint8_t)cur[p];
}
// Ancestor gather of the whole HMM state in one pass: child k takes parent a = anc[k]'s newest state
// (cur) and the LIVE 16-byte units of its history row (columns < T; `units` = ceil(T/16) of the
// `row_units` a row has -- columns >= T are never read by anybody); a pending newest column t_pend (>= 0)
// is inserted into the row on the way, so the row is written once, as whole units.  When lw_out is given,
// the evened-out weight lw[k] = red[4] - log_n_new (utils.py:152-157) is written by the same pass.
// One thread per live unit of a child row.
__global__ void __launch_bounds__(256)
hmm_gather_kernel(const int32_t *__restrict__ cur_in, int32_t *__restrict__ cur_out,
                  const uint4 *__restrict__ hist_in, uint4 *__restrict__ hist_out,
                  const int32_t *__restrict__ anc, int64_t n_new, int units, int row_units, int t_pend,
                  double *__restrict__ lw_out, const double *__restrict__ red, double log_n_new) {
  const int64_t total = n_new * units;
  const int pu = t_pend >= 0 ? t_pend >> 4 : -1, pw = (t_pend >> 2) & 3, psh = (t_pend & 3) * 8;
  const double even = lw_out ? red[4] - log_n_new : 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = i / units;
    const int c = (int)(i - k * units);
    const int64_t a = anc[k];
    uint4 v = hist_in[a * row_units + c];
    if (c == 0 || c == pu) {
      const uint32_t s = (uint32_t)cur_in[a];
      if (c == 0) { cur_out[k] = (int32_t)s; if (lw_out) lw_out[k] = even; }
      if (c == pu) {
        const uint32_t keep = ~(0xFFu << psh), ins = (s & 0xFFu) << psh;
        if (pw == 0) v.x = (v.x & keep) | ins;
        else if (pw == 1) v.y = (v.y & keep) | ins;
        else if (pw == 2) v.z = (v.z & keep) | ins;
        else v.w = (v.w & keep) | ins;
      }
    }
    hist_out[k * row_units + c] = v;
  }
}

struct HMMModel : Model {
  HMMTables tb;
  int T_max, Tpad, T = 0 /*propagated time steps*/;
  std::vector<int32_t> symbols_host;   // all observed symbols (propagated first, then pending)
  DevBuf<double> trans, emit, step_tables;
  DevBuf<int32_t> symbols, cura, curb;
  DevBuf<uint8_t> hista, histb;
  int32_t *cur = nullptr, *cur_alt = nullptr;
  uint8_t *hist = nullptr, *hist_alt = nullptr;
  size_t smem, smem_step;
  // column T-1 of hist is still only in cur[] (written by the last filter step); the resampling gather
  // inserts it while it copies the rows, every other reader of hist flushes it first
  bool hist_pending = false;
  void flush_hist() {
    if (!hist_pending) return;
    hmm_flush_kernel<<<grid(), 256, 0, h->stream>>>(cur, hist, Tpad, T - 1, h->P);
    VK_CUDA(cudaGetLastError());
    hist_pending = false;
  }
  int gather(const int32_t *anc_dev, int64_t n_new, double *lw_out, const double *red, double log_n_new) override {
    const int row_units = Tpad / 16;
    int units = (T + 15) / 16;
    if (units < 1) units = 1;
    hmm_gather_kernel<<<grid_for(n_new * units, 256, h->sm_count, 8), 256, 0, h->stream>>>(
        cur, cur_alt, (const uint4 *)hist, (uint4 *)hist_alt, anc_dev, n_new, units, row_units,
        hist_pending ? T - 1 : -1, lw_out, red, log_n_new);
    VK_CUDA(cudaGetLastError());
    std::swap(cur, cur_alt); std::swap(hist, hist_alt);
    hist_pending = false;
    return lw_out ? 3 : 1;
  }

  HMMModel(vk_particles *h_, int S, int O, int n_rows, const double *trans_host, const double *emit_host,
           int init, int T_max_) : Model(h_), T_max(T_max_) {
    VK_REQUIRE(S >= 1 && S <= HMM_MAX_S && O >= 1 && O <= 255, VK_EINVAL, "hmm: S must be in [1,32], O in [1,255]");
    VK_REQUIRE(n_rows >= S, VK_EINVAL, "hmm: transition matrix needs at least S rows");
    VK_REQUIRE(init >= 1 && init <= S, VK_EINVAL, "hmm: init_state out of range");
    VK_REQUIRE(T_max >= 1 && T_max <= 4096, VK_EINVAL, "hmm: T_max must be in [1,4096]");
    tb.S = S; tb.O = O; tb.n_rows = n_rows; tb.init = init;
    trans.alloc((size_t)n_rows * S); emit.alloc((size_t)S * O);
    VK_CUDA(cudaMemcpyAsync(trans.p, trans_host, trans.bytes(), cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaMemcpyAsync(emit.p, emit_host, emit.bytes(), cudaMemcpyHostToDevice, h->stream));
    tb.trans = trans.p; tb.emit = emit.p;
    Tpad = (T_max + 15) & ~15;
    symbols.alloc((size_t)T_max);
    cura.alloc((size_t)h->P); curb.alloc((size_t)h->P);
    hista.alloc((size_t)h->P * Tpad); histb.alloc((size_t)h->P * Tpad);
    cur = cura.p; cur_alt = curb.p; hist = hista.p; hist_alt = histb.p;
    VK_CUDA(cudaMemsetAsync(cur, 0, cura.bytes(), h->stream));
    VK_CUDA(cudaMemsetAsync(hist, 0, hista.bytes(), h->stream));
    VK_CUDA(cudaMemsetAsync(hist_alt, 0, histb.bytes(), h->stream));
    smem = (size_t)(n_rows * S + S * O) * sizeof(double);
    smem_step = (size_t)(n_rows * S + 2 * n_rows + S) * sizeof(double);
    step_tables.alloc((size_t)hmm_step_table_len(n_rows, S, O));
    hmm_step_tables_kernel<<<1, 256, 0, h->stream>>>(tb, step_tables.p);
    VK_CUDA(cudaGetLastError());
  }
  int grid() const { return grid_for(h->P, 256, h->sm_count, 8); }
  int step_grid() const { return grid_for((h->P + HMM_STEP_ITEMS - 1) / HMM_STEP_ITEMS, 256, h->sm_count, 8); }
  void observe(int symbol) {
    VK_REQUIRE(symbol >= 1 && symbol <= tb.O, VK_EINVAL, "hmm: observed symbol out of range");
    VK_REQUIRE((int)symbols_host.size() < T_max, VK_EINVAL, "hmm: history capacity (T_max) exceeded");
    symbols_host.push_back(symbol);
  }
  void upload_symbols() {
    if (!symbols_host.empty())
      VK_CUDA(cudaMemcpyAsync(symbols.p, symbols_host.data(), symbols_host.size() * sizeof(int32_t),
                              cudaMemcpyHostToDevice, h->stream));
  }
  void make_consistent(double *incr_dev) override {
    int n = (int)symbols_host.size();
    if (T == n) {
      if (incr_dev) VK_CUDA(cudaMemsetAsync(incr_dev, 0, (size_t)h->P * sizeof(double), h->stream));
      return;
    }
    flush_hist();
    bool first = true, x0 = false;
    for (; T < n; ++T) {
      uint32_t step = h->next_step();
      const bool last = T == n - 1;
      P2PXchg xchg;
      xchg.tab = nullptr;
      if (last) x0 = p2p_begin_exchange(h, 0, &xchg);      // sharded particle set: exchange 0 rides in this kernel's tail
      h->prof.mark(h->stream, "(gap)");
      hmm_step_kernel<<<step_grid(), 256, smem_step, h->stream>>>(cur, hist, Tpad, T, h->lw.p, incr_dev, tb, step_tables.p,
                                                             symbols_host[T], h->P, h->p_base, h->seed, step,
                                                             first ? 0 : 1, h->red.p, h->counters.p, last ? 0 : 1, xchg);
      h->prof.mark(h->stream, x0 ? "hmm_step+send0" : "hmm_step");
      first = false;
    }
    hist_pending = true;
    VK_CUDA(cudaGetLastError());
    h->lw_minmax_fresh = true;     // red[0..2] describe the weights as they are now
    h->lw_x0_done = x0;            // ... those of ALL ranks
  }
  void run_prior(bool write_lw) {
    T = (int)symbols_host.size();
    upload_symbols();
    uint32_t step = h->next_step();
    hmm_lw_kernel<<<grid(), 256, smem, h->stream>>>(cur, hist, Tpad, T, h->lw.p, tb, symbols.p, h->P, h->p_base,
                                                    h->seed, step, write_lw ? 1 : 0);
    VK_CUDA(cudaGetLastError());
    VK_CUDA(cudaStreamSynchronize(h->stream));  // symbols_host stays valid for the async copy
    hist_pending = false;                       // every column was just rewritten
  }
  void likelihood_weight() override { run_prior(true); }
  void reset_to_prior() override { run_prior(false); }
  void mh(int scope, int block, int K, double *mean_nodes_dev, long long *acc_dev) override {
    run_sites(0, scope, block, K, mean_nodes_dev, acc_dev);
  }
  void gibbs(int scope, int block, int K, double *mean_nodes_dev, long long *changed_dev) override {
    VK_REQUIRE(tb.S <= 32, VK_EUNSUPPORTED, "hmm: gibbs is compiled for at most 32 states");
    run_sites(1, scope, block, K, mean_nodes_dev, changed_dev);
  }
  void pgibbs(int scope, int block, int particles, int K, double *mean_nodes_dev, long long *acc_dev) override {
    VK_REQUIRE(scope == 0, VK_EINVAL, "hmm: only the default scope (= the ordered scope `seq` of hmm.vnt) exists");
    VK_REQUIRE(block == VK_BLOCK_ORDERED, VK_EINVAL, "pgibbs: block must be `ordered`");
    VK_REQUIRE(particles >= 1 && particles <= HMM_PG_MAX, VK_EUNSUPPORTED, "pgibbs: 1..16 inner particles are compiled");
    VK_REQUIRE(T <= 32, VK_EUNSUPPORTED, "pgibbs: at most 32 time steps are compiled");
    VK_REQUIRE(T == (int)symbols_host.size(), VK_ESTATE, "pgibbs: unpropagated observations; call vk_make_consistent first");
    flush_hist();
    if (K > 0 && T > 1) {
      uint32_t step0 = h->next_step((uint32_t)K);
      upload_symbols();
      hmm_pgibbs_kernel<<<grid_for(h->P, 128, h->sm_count, 8), 128, smem, h->stream>>>(
          cur, hist, Tpad, T, tb, symbols.p, h->P, h->p_base, h->seed, step0, K, particles, acc_dev, mean_nodes_dev);
      VK_CUDA(cudaGetLastError());
      VK_CUDA(cudaStreamSynchronize(h->stream));
    } else {
      if (acc_dev) VK_CUDA(cudaMemsetAsync(acc_dev, 0, (size_t)h->P * sizeof(long long), h->stream));
      if (mean_nodes_dev) hmm_fill_kernel<<<grid(), 256, 0, h->stream>>>(mean_nodes_dev, h->P, 0.0);
    }
  }
  void run_sites(int op, int scope, int block, int K, double *mean_nodes_dev, long long *acc_dev) {
    VK_REQUIRE(scope == 0, VK_EINVAL, "hmm: only the default scope exists");
    VK_REQUIRE(T == (int)symbols_host.size(), VK_ESTATE, "mh: unpropagated observations; call vk_make_consistent first");
    int nb = T - 1;
    VK_REQUIRE(block == VK_BLOCK_ONE || block == VK_BLOCK_EACH || (block >= 0 && block < nb), VK_EINVAL, "hmm: bad block");
    flush_hist();
    if (K > 0 && nb > 0) {  // zero-entropy scope performs no transitions (dispatch.py:55-56)
      int n_steps = block == VK_BLOCK_EACH ? K * nb : K;
      uint32_t step0 = h->next_step((uint32_t)n_steps);
      upload_symbols();
      if (op == 0)
        hmm_mh_kernel<<<grid(), 256, smem, h->stream>>>(cur, hist, Tpad, T, tb, symbols.p, h->P, h->p_base, h->seed,
                                                        step0, n_steps, block, acc_dev, mean_nodes_dev);
      else
        hmm_gibbs_kernel<<<grid(), 256, smem, h->stream>>>(cur, hist, Tpad, T, tb, symbols.p, h->P, h->p_base, h->seed,
                                                           step0, n_steps, block, acc_dev, mean_nodes_dev);
      VK_CUDA(cudaGetLastError());
      VK_CUDA(cudaStreamSynchronize(h->stream));
    } else {
      if (acc_dev) VK_CUDA(cudaMemsetAsync(acc_dev, 0, (size_t)h->P * sizeof(long long), h->stream));
      if (mean_nodes_dev) hmm_fill_kernel<<<grid(), 256, 0, h->stream>>>(mean_nodes_dev, h->P, 0.0);
    }
  }
  void log_likelihood_at(int scope, int block, bool joint, double *out_dev) override {
    VK_REQUIRE(scope == 0, VK_EINVAL, "hmm: only the default scope exists");
    VK_REQUIRE(block == VK_BLOCK_ALL, VK_EUNSUPPORTED, "hmm: log_likelihood_at is compiled for block all");
    // The scaffold's principal nodes are the random choices state(2..T); state(1) is deterministic, so
    // its emission is not on the border and does not count (Lite agrees: lite_chains.json c3 ll_at).
    loglik_from(1, joint, out_dev);
  }
  void loglik_from(int t_first, bool joint, double *out_dev) {
    flush_hist();
    upload_symbols();
    hmm_loglik_kernel<<<grid(), 256, 0, h->stream>>>(hist, Tpad, T, tb, symbols.p, h->P, joint ? 1 : 0, t_first, out_dev);
    VK_CUDA(cudaGetLastError());
    VK_CUDA(cudaStreamSynchronize(h->stream));
  }
  void forget_observations(double *w_dev) override {
    if (w_dev) loglik_from(0, false, w_dev);   // unobserve reports every observation's density
    symbols_host.clear(); T = 0; hist_pending = false;
  }
  int64_t num_sites() const override { return T; }
  int64_t num_blocks(int) const override { return T > 1 ? T - 1 : 0; }
  void read_site(int64_t site, double *out_dev) override {
    VK_REQUIRE(site >= 0 && site < T, VK_EINVAL, "hmm: site (0-based time index) out of range");
    flush_hist();
    hmm_read_kernel<<<grid(), 256, 0, h->stream>>>(hist, Tpad, (int)site, h->P, out_dev);
  }
  void write_site(int64_t site, const double *in_dev) override {
    VK_REQUIRE(site >= 0 && site < T, VK_EINVAL, "hmm: site (0-based time index) out of range");
    flush_hist();
    hmm_write_kernel<<<grid(), 256, 0, h->stream>>>(hist, cur, Tpad, (int)site, site == T - 1, h->P, in_dev);
  }
  std::vector<StateArray> state_arrays(bool) override {
    flush_hist();   // generic consumers (pack/unpack, peer-memory and small-P gathers) see complete rows
    return state_arrays_peek();
  }
  std::vector<StateArray> state_arrays_peek() override {
    StateArray a{(void **)&cur, (void **)&cur_alt, 0, 1, 4};
    StateArray b{(void **)&hist, (void **)&hist_alt, (size_t)Tpad, 0, 0};
    return {a, b};
  }
  void resized(int64_t n_new, const std::vector<void *> &nc, const std::vector<void *> &na) override {
    cura.adopt(nc[0], (size_t)n_new); curb.adopt(na[0], (size_t)n_new);
    hista.adopt(nc[1], (size_t)n_new * Tpad); histb.adopt(na[1], (size_t)n_new * Tpad);
    cur = cura.p; cur_alt = curb.p; hist = hista.p; hist_alt = histb.p;
    VK_CUDA(cudaMemsetAsync(hist_alt, 0, histb.bytes(), h->stream));
  }
  std::vector<StateArray> state_arrays_patched(int patch[3]) override {
    patch[0] = -1;
    if (hist_pending) { patch[0] = 1; patch[1] = 0; patch[2] = T - 1; hist_pending = false; }   // the caller's gather inserts it
    StateArray a{(void **)&cur, (void **)&cur_alt, 0, 1, 4};
    StateArray b{(void **)&hist, (void **)&hist_alt, (size_t)Tpad, 0, 0};
    b.live_bytes = (size_t)(T > 0 ? T : 1);      // columns T.. of both buffers are still zero (as hmm_gather_kernel's `units`)
    return {a, b};
  }
  int device_buffer(int which, void **p, int64_t *bytes) override {
    if (which == 1) { *p = cur; *bytes = (int64_t)cura.bytes(); return VK_OK; }
    if (which == 2) { flush_hist(); *p = hist; *bytes = (int64_t)hista.bytes(); return VK_OK; }
    return VK_EINVAL;
  }
};

Model *make_hmm(vk_particles *h, int S, int O, int n_trans_rows, const double *trans_host,
                const double *emit_host, int init_state, int T_max) {
  return new HMMModel(h, S, O, n_trans_rows, trans_host, emit_host, init_state, T_max);
}
void hmm_observe(Model *m, int symbol) { static_cast<HMMModel *>(m)->observe(symbol); }

}  // namespace vk
