// Configs C2/C5: Bayesian logistic regression
//   assume w_j = normal(mu0, s0)  (j < D);  assume w = array(w_0 .. w_{D-1})
//   observe log_odds_flip(vector_dot(w, x_i)) = y_i          (i < N)
// Site program derived from the reference's scaffold rules:
//   scaffold(w_j): DRG {w_j, its lookup, the array node, N lookups of w,
//                  N vector_dot nodes}; absorbing {N log_odds_flip nodes}.
//   rho = sum_i log_logistic(+-eta_i)         (detach.py:72-83, discrete.py:148-150)
//   xi  = sum_i log_logistic(+-eta'_i), eta'_i = eta_i + (w'_j - w_j) x_ij
//                                             (regen.py:120-127, utils.py:257-268)
//   accept iff log u < xi - rho               (infer/mh.py:84-89, 187-208)
//
// HBM layout
//   Xs  [D][Ncap]  signed features  s_i * x_ij  (s_i = +1 iff y_i) -- shared by
//                  all particles, 1.3 MB at C2 => L2 resident
//   w   [D][P]     particle-minor weights
//   eta [P][Ncap]  particle-major signed logits  s_i * (w . x_i)  (the cache
//                  that replaces the N vector_dot nodes of every trace)
//   ll  [P]        current sum_i log_logistic(eta_i) (= rho of the next proposal)
//
// Kernel mapping: ONE CTA PER PARTICLE, persistent over particles.  The CTA
// stages its 8N-byte logit row in shared memory ONCE, runs all K transitions
// of the launch on chip (each transition streams one 8N-byte Xs column from
// L2), and writes the row back once.  Algorithmic HBM traffic per launch is
// therefore 2*8N B per particle instead of 8N(1+a) per transition; the
// kernel is fp64-ALU bound (exp + log1p per observation), see DESIGN.md.
#include "vk_common.cuh"
#include "sp_math.cuh"
#include "blr_subsampled.cuh"
#include "fast_math.cuh"
#include "blr_subsampled_persistent.cuh"
#include "weights.cuh"
#include <stdlib.h>

namespace vk {

struct BLRParams { double mu0, s0; int D; int64_t N, Ncap, Xld; };   // Xld: column stride of Xs (zero padded)

__device__ __forceinline__ double blr_log_alpha(double rho, double xi) {
  if (rho == -INFINITY && xi == -INFINITY) return INFINITY;   // mh.py:190-197
  if (rho == INFINITY && xi == INFINITY) return -INFINITY;    // mh.py:198-206
  return xi - rho;
}

// mode 0: reset_to_prior (w ~ prior, eta/ll over propagated obs, weights untouched)
// mode 1: likelihood_weight (as 0 and lw := ll)            lite/trace.py:522-530
// mode 2: refresh (keep w; recompute eta and ll)
// mode 3: make_consistent over [n_prop, N): eta for new obs, incr, lw += incr, ll += incr
//
// eta_i = sum_j w_j Xs[j][i] and sum_i log_logistic(eta_i), BLR_ALL_PP particles per CTA pass: the D
// feature values of an observation are read from L2 ONCE and used for all PP particles (one particle per
// pass made the kernel L2-bound: 8 D bytes of L2 traffic per evaluation, 84 GB per call at C2).
static const int BLR_ALL_PP = 4;
// F32 (vk_set_precision(VK_FP32), SURVEY 8 preamble "fp32 (throughput) mode"; north_star: log weights within 1e-5):
// logits and the log-logistic in single precision (MUFU ex2 / lg2), per-thread sums in single precision, the
// block sum and everything stored (weights, increments, the logit cache as float-rounded doubles) in double.
template <bool F32>
__global__ void __launch_bounds__(256)
blr_all_kernel(double *__restrict__ w, double *__restrict__ eta, double *__restrict__ ll,
               double *__restrict__ lw, double *__restrict__ incr,
               const double *__restrict__ Xs, BLRParams prm, int64_t n_prop, int64_t P,
               int64_t p_base, uint64_t seed, uint32_t step, int mode) {
  extern __shared__ double smem[];
  const int D = prm.D;
  double *w_sh = smem;                       // PP x D
  double *red = smem + BLR_ALL_PP * D;       // 33
  const int64_t i0 = (mode == 3) ? n_prop : 0;
  const int64_t i1 = (mode == 3) ? prm.N : n_prop;
  for (int64_t p0 = (int64_t)blockIdx.x * BLR_ALL_PP; p0 < P; p0 += (int64_t)gridDim.x * BLR_ALL_PP) {
    const int np = (int)(P - p0 < BLR_ALL_PP ? P - p0 : BLR_ALL_PP);
    __syncthreads();
    for (int t = threadIdx.x; t < BLR_ALL_PP * D; t += blockDim.x) {
      const int q = t / D, j = t - q * D;
      double v = 0.0;
      if (q < np) {
        const int64_t p = p0 + q;
        if (mode <= 1) {
          Stream r(seed, (uint64_t)(p_base + p), step);
          r.seek(2u * (uint32_t)j);
          v = normal_sim(r, prm.mu0, prm.s0);
          w[(int64_t)j * P + p] = v;
        } else {
          v = w[(int64_t)j * P + p];
        }
      }
      w_sh[t] = v;
    }
    __syncthreads();
    double acc[BLR_ALL_PP];
#pragma unroll
    for (int q = 0; q < BLR_ALL_PP; ++q) acc[q] = 0.0;
    if (F32) {
      float *wf_sh = reinterpret_cast<float *>(red + 34);      // PP x D single-precision copies of the weights
      for (int t = threadIdx.x; t < BLR_ALL_PP * D; t += blockDim.x) wf_sh[t] = (float)w_sh[t];
      __syncthreads();
      float accf[BLR_ALL_PP];
#pragma unroll
      for (int q = 0; q < BLR_ALL_PP; ++q) accf[q] = 0.0f;
      int n_in_acc = 0;
      for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
        float e[BLR_ALL_PP];
#pragma unroll
        for (int q = 0; q < BLR_ALL_PP; ++q) e[q] = 0.0f;
        for (int j = 0; j < D; ++j) {
          const float x = (float)__ldg(&Xs[(int64_t)j * prm.Xld + i]);
#pragma unroll
          for (int q = 0; q < BLR_ALL_PP; ++q) e[q] = fmaf(wf_sh[q * D + j], x, e[q]);
        }
#pragma unroll
        for (int q = 0; q < BLR_ALL_PP; ++q) {
          if (q < np) {
            eta[(p0 + q) * prm.Ncap + i] = (double)e[q];
            // log_logistic(v) = min(v, 0) - log(1 + exp(-|v|))   (utils.py:257-268), MUFU ex2 / lg2
            const float a = fabsf(e[q]);
            const float t = __expf(-a);
            accf[q] += fminf(e[q], 0.0f) - __logf(1.0f + t);
          }
        }
        if (++n_in_acc == 64) {                 // bound the length of a single-precision running sum
#pragma unroll
          for (int q = 0; q < BLR_ALL_PP; ++q) { acc[q] += (double)accf[q]; accf[q] = 0.0f; }
          n_in_acc = 0;
        }
      }
#pragma unroll
      for (int q = 0; q < BLR_ALL_PP; ++q) acc[q] += (double)accf[q];
    } else
    for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
      double e[BLR_ALL_PP];
#pragma unroll
      for (int q = 0; q < BLR_ALL_PP; ++q) e[q] = 0.0;
      for (int j = 0; j < D; ++j) {
        const double x = __ldg(&Xs[(int64_t)j * prm.Xld + i]);
#pragma unroll
        for (int q = 0; q < BLR_ALL_PP; ++q) e[q] = fma(w_sh[q * D + j], x, e[q]);   // same order as one particle per pass
      }
#pragma unroll
      for (int q = 0; q < BLR_ALL_PP; ++q) {
        if (q < np) { eta[(p0 + q) * prm.Ncap + i] = e[q]; acc[q] += fast_log_logistic(e[q]); }
      }
    }
#pragma unroll
    for (int q = 0; q < BLR_ALL_PP; ++q) {
      if (q >= np) break;                    // uniform across the CTA
      const double tot = block_sum(acc[q], red);
      if (threadIdx.x == 0) {
        const int64_t p = p0 + q;
        if (mode == 3) {
          double inc = tot;
          if (inc != inc) inc = -INFINITY;  // lite/trace.py:438-448
          if (incr) incr[p] = inc;
          lw[p] += inc;
          ll[p] += tot;
        } else {
          ll[p] = tot;
          if (mode == 1) lw[p] = tot;
        }
      }
    }
  }
}

// K resimulation-MH transitions per particle, eta row staged in shared memory
// (SMEM=true) or streamed from global memory every transition (SMEM=false,
// rows too large for shared memory).
// block_mode: VK_BLOCK_ONE (uniform site per transition), VK_BLOCK_EACH
// (site = t mod D), or a fixed site index >= 0.
// Proposals come from the prior and do not depend on the chain state, so the
// (site, w', log u) triples of a whole chunk of transitions are drawn up front,
// one transition per thread, and the transition loop itself is pure fp64 math:
// per observation one FMA (eta + delta*x) and one fast_log_logistic.
static const int BLR_MH_CHUNK = 256;   // transitions whose proposals are staged at once
static const int BLR_MH_TILE = 1024;   // observations per CTA-wide loop iteration (THREADS x UNROLL)
static const int BLR_MH_DEFAULT_VARIANT = 4;   // blr_mh2_kernel (blr_mh2.cuh); 0-3: blr_mh_kernel shapes
#define BLR_PAD_ETA 800.0              // logit of a padding slot: log_logistic(800) = -exp(-800) = 0

// shared-memory carve-up shared by the kernel and the host-side size computation:
// [log table][exp table][w D'][colsum D'][red 34][slots 64][prop_w C][prop_logu C][prop_site C][eta row Npad]
__host__ __device__ inline size_t blr_mh_smem_base(int D) {
  return (size_t)FM_SMEM_LOG_BYTES + FM_SMEM_EXP_BYTES +
         (size_t)(2 * ((D + 1) & ~1) + 34 + 64 + 2 * BLR_MH_CHUNK) * sizeof(double) +
         (size_t)BLR_MH_CHUNK * sizeof(int);
}

// K resimulation-MH transitions per particle (see the header comment).
// xi = sum_i log_logistic(v_i), v_i = eta_i + delta x_ij, is evaluated as
//   sum_i min(v_i,0) - sum_i log1p(exp(-|v_i|)),
//   sum_i min(v_i,0) = (sum_i v_i - sum_i |v_i|)/2,   sum_i v_i = S_eta + delta * colsum_j
// so the per-observation fp64 work is 1 FMA (v), 1 add (|v|) and the table-driven
// softplus (fast_math.cuh: fm_softplus_accum, 17 fp64 instructions).
// The observation axis is padded to Npad = a multiple of BLR_MH_TILE: padding slots
// have x = 0 (Xs columns are zero padded) and eta = BLR_PAD_ETA, so they contribute
// exactly 0 to xi and every thread runs the same number of full iterations.  The
// Xs values of the NEXT iteration are loaded before the current one is evaluated
// (L2 latency hidden behind ~40 fp64 instructions per chain).
// Block-wide sum with ONE barrier: per-warp sums go to a double-buffered slot
// array (buffer = parity of the call), every thread then adds the NW values in
// the same fixed order, so the result is bit-identical in all threads.
template <int NW>
__device__ __forceinline__ double block_sum_1sync(double v, double *slots /* 2*NW */, int parity) {
  v = warp_sum(v);
  double *b = slots + parity * NW;
  if ((threadIdx.x & 31) == 0) b[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = b[0];
#pragma unroll
  for (int k = 1; k < NW; ++k) t += b[k];
  return t;
}

template <bool SMEM, int THREADS, int UNROLL, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
blr_mh_kernel(double *__restrict__ w, double *__restrict__ eta, double *__restrict__ ll,
              const double *__restrict__ Xs, BLRParams prm, int64_t P, int64_t p_base,
              uint64_t seed, uint32_t step0, int n_prop_steps, int block_mode,
              long long *__restrict__ acc_out, const FmCoef cf) {
  static_assert(THREADS * UNROLL == BLR_MH_TILE, "tile");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned char *log_tbl = smem_raw;
  unsigned char *exp_tbl = smem_raw + FM_SMEM_LOG_BYTES;
  const int Dp = (prm.D + 1) & ~1;
  double *w_sh = reinterpret_cast<double *>(exp_tbl + FM_SMEM_EXP_BYTES);   // D
  double *colsum = w_sh + Dp;                // D: sum_i Xs[j][i]
  double *red = colsum + Dp;                 // 34
  double *slots = red + 34;                  // 2 x 32 per-warp partial sums (block_sum_1sync)
  double *prop_w = slots + 64;               // BLR_MH_CHUNK proposed values
  double *prop_logu = prop_w + BLR_MH_CHUNK; // BLR_MH_CHUNK log-uniforms
  int *prop_site = (int *)(prop_logu + BLR_MH_CHUNK);   // BLR_MH_CHUNK sites
  double *eta_sh = (double *)(prop_site + BLR_MH_CHUNK);  // Npad (SMEM only)
  const int N = (int)prm.N;
  const int Npad = (N + BLR_MH_TILE - 1) / BLR_MH_TILE * BLR_MH_TILE;
  const int iters = Npad / BLR_MH_TILE;
  const int tid = threadIdx.x, lane = threadIdx.x & 31;
  fm_fill_smem_tables(log_tbl, exp_tbl);
  const uint32_t exp_base = (uint32_t)__cvta_generic_to_shared(exp_tbl) + (uint32_t)(lane & 15) * 8u;
  const uint32_t log_base = (uint32_t)__cvta_generic_to_shared(log_tbl) + (uint32_t)(lane & 7) * 16u;
  for (int j = 0; j < prm.D; ++j) {          // column sums, once per CTA
    const double *xc = Xs + (int64_t)j * prm.Xld;
    double s_ = 0.0;
    for (int i = tid; i < N; i += THREADS) s_ += __ldg(&xc[i]);
    s_ = block_sum(s_, red);
    if (tid == 0) colsum[j] = s_;
  }
  for (int64_t p = blockIdx.x; p < P; p += gridDim.x) {
    __syncthreads();
    double *eta_g = eta + p * prm.Ncap;
    if (tid < prm.D) w_sh[tid] = w[(int64_t)tid * P + p];
    double s_part = 0.0;
    if (SMEM) {
      // Ncap is even and rows are 16-byte aligned: 16-byte vector copies
      const double2 *src = reinterpret_cast<const double2 *>(eta_g);
      double2 *dst = reinterpret_cast<double2 *>(eta_sh);
      for (int i = tid; i < Npad / 2; i += THREADS) {
        double2 v = make_double2(BLR_PAD_ETA, BLR_PAD_ETA);
        if (2 * i < N) { v = src[i]; if (2 * i + 1 >= N) v.y = BLR_PAD_ETA; }
        dst[i] = v; s_part += v.x + v.y;
      }
    } else {
      for (int i = tid; i < Npad; i += THREADS) s_part += i < N ? eta_g[i] : BLR_PAD_ETA;
    }
    double S_eta = block_sum(s_part, red);   // sum_i eta_i of the current state (padding included)
    double rho = ll[p];
    long long acc = 0;
    bool dirty = false;
    for (int c0 = 0; c0 < n_prop_steps; c0 += BLR_MH_CHUNK) {
      int cn = n_prop_steps - c0 < BLR_MH_CHUNK ? n_prop_steps - c0 : BLR_MH_CHUNK;
      __syncthreads();
      for (int t = tid; t < cn; t += THREADS) {
        Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)(c0 + t));
        double u_sel = r.uniform();
        double u_acc = r.uniform();
        int j;
        if (block_mode == VK_BLOCK_ONE) { j = (int)(u_sel * (double)prm.D); if (j >= prm.D) j = prm.D - 1; }
        else if (block_mode == VK_BLOCK_EACH) j = (c0 + t) % prm.D;
        else j = block_mode;
        prop_site[t] = j;
        prop_logu[t] = log(u_acc);
        prop_w[t] = normal_sim(r, prm.mu0, prm.s0);
      }
      __syncthreads();
      double xn[UNROLL];               // Xs values of the tile about to be evaluated
      {
        const double *x0 = Xs + (int64_t)prop_site[0] * prm.Xld + tid;
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) xn[u] = __ldg(x0 + u * THREADS);
      }
      for (int t = 0; t < cn; ++t) {
        const int j = prop_site[t];
        const double w_new = prop_w[t];
        const double delta = w_new - w_sh[j];
        const double *__restrict__ xcol = Xs + (int64_t)j * prm.Xld;
        // first tile of the NEXT transition's column (its site does not depend on this accept)
        const double *xnext = Xs + (int64_t)prop_site[t + 1 < cn ? t + 1 : t] * prm.Xld + tid;
        double am[UNROLL], af[UNROLL];   // sum |v|, sum r2 Q(r2)
        long long aL[UNROLL];            // sum L[k], 2^-52 fixed point
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) { am[u] = 0.0; af[u] = 0.0; aL[u] = 0; }
        int an = 0, n_logs = 0;          // product logarithm: sum of biased exponents, number of logs
        double qp[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) qp[u] = 1.0;
        const double *xp = xcol + tid + BLR_MH_TILE;
        int i0 = tid;
        for (int it = 0; it < iters; ++it, xp += BLR_MH_TILE, i0 += BLR_MH_TILE) {
          double v[UNROLL];
#pragma unroll
          for (int u = 0; u < UNROLL; ++u) {
            double ev;
            if (SMEM) ev = eta_sh[i0 + u * THREADS];
            else ev = (i0 + u * THREADS < N) ? eta_g[i0 + u * THREADS] : BLR_PAD_ETA;
            v[u] = fma(delta, xn[u], ev);
          }
          {
            const double *src = it + 1 < iters ? xp : xnext;   // next tile, or the next transition's first tile
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) xn[u] = __ldg(src + u * THREADS);
          }
          // the factors 1 + exp(-|v|) go into UNROLL running products; ONE logarithm per transition
          // (fast_math.cuh: fm_log_product), flushed every 64 iterations so a product stays < 2^64
#pragma unroll
          for (int u = 0; u < UNROLL; ++u) {
            qp[u] *= 1.0 + fm_exp_neg_abs(v[u], cf, exp_base);
            am[u] += fabs(v[u]);
          }
          if ((it & 63) == 63) {
            double q = qp[0];
#pragma unroll
            for (int u = 1; u < UNROLL; ++u) q *= qp[u];
            fm_log_product(q, cf, log_base, af[0], aL[0], an);
            ++n_logs;
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) qp[u] = 1.0;
          }
        }
        if ((iters & 63) != 0) {
          double q = qp[0];
#pragma unroll
          for (int u = 1; u < UNROLL; ++u) q *= qp[u];
          fm_log_product(q, cf, log_base, af[0], aL[0], an);
          ++n_logs;
        }
#pragma unroll
        for (int u = 1; u < UNROLL; ++u) am[0] += am[u];
        double part = fma(-0.5, am[0], -af[0]) - (double)aL[0] * 2.220446049250313e-16;
        part -= (double)(an - 1023 * n_logs) * 0.6931471805599453;   // the products' exponents
        if (tid == 0) part += 0.5 * fma(delta, colsum[j], S_eta);
        double xi = block_sum_1sync<THREADS / 32>(part, slots, t & 1);
        double alpha = blr_log_alpha(rho, xi);
        if (prop_logu[t] < alpha) {   // uniform across the CTA
          double sp = 0.0;
          if (SMEM) {
            for (int k = tid; k < Npad; k += THREADS) { double e_ = fma(delta, __ldg(&xcol[k]), eta_sh[k]); eta_sh[k] = e_; sp += e_; }
          } else {
            for (int k = tid; k < Npad; k += THREADS) {
              double e_ = BLR_PAD_ETA;
              if (k < N) { e_ = fma(delta, __ldg(&xcol[k]), eta_g[k]); eta_g[k] = e_; }
              sp += e_;
            }
          }
          if (tid == 0) w_sh[j] = w_new;
          rho = xi; ++acc; dirty = true;
          S_eta = block_sum(sp, red);   // recomputed, not updated: no drift
        }
      }
    }
    if (dirty) {
      if (SMEM) {
        const double2 *src = reinterpret_cast<const double2 *>(eta_sh);
        double2 *dst = reinterpret_cast<double2 *>(eta_g);
        for (int i = tid; i < (N + 1) / 2; i += THREADS) dst[i] = src[i];
      }
      if (tid < prm.D) w[(int64_t)tid * P + p] = w_sh[tid];
      if (tid == 0) ll[p] = rho;
    }
    if (acc_out && tid == 0) acc_out[p] = acc;
  }
}

}  // namespace vk
#include "blr_mh_f32.cuh"
#include "blr_mh2.cuh"
namespace vk {

// block = all: joint proposal of every weight from the prior (one scaffold
// with D principal nodes).  eta' is recomputed into a scratch row.
__global__ void __launch_bounds__(256)
blr_mh_all_kernel(double *__restrict__ w, double *__restrict__ eta, double *__restrict__ ll,
                  const double *__restrict__ Xs, BLRParams prm, int64_t P, int64_t p_base,
                  uint64_t seed, uint32_t step0, int K, long long *__restrict__ acc_out) {
  extern __shared__ double smem[];
  double *w_sh = smem;                       // D
  double *red = w_sh + ((prm.D + 1) & ~1);   // 34
  const int64_t N = prm.N;
  for (int64_t p = blockIdx.x; p < P; p += gridDim.x) {
    double rho = ll[p];
    long long acc = 0;
    for (int t = 0; t < K; ++t) {
      __syncthreads();
      Stream r0(seed, (uint64_t)(p_base + p), step0 + (uint32_t)t);
      (void)r0.uniform();
      double u_acc = r0.uniform();
      if ((int)threadIdx.x < prm.D) {
        Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)t);
        r.seek(2u + 2u * threadIdx.x);
        w_sh[threadIdx.x] = normal_sim(r, prm.mu0, prm.s0);
      }
      __syncthreads();
      double part = 0.0;
      for (int64_t i = threadIdx.x; i < N; i += blockDim.x) {
        double e = 0.0;
        for (int j = 0; j < prm.D; ++j) e = fma(w_sh[j], __ldg(&Xs[(int64_t)j * prm.Xld + i]), e);
        part += fast_log_logistic(e);
      }
      double xi = block_sum(part, red);
      double alpha = blr_log_alpha(rho, xi);
      if (log(u_acc) < alpha) {
        double *eta_g = eta + p * prm.Ncap;
        for (int64_t i = threadIdx.x; i < N; i += blockDim.x) {
          double e = 0.0;
          for (int j = 0; j < prm.D; ++j) e = fma(w_sh[j], __ldg(&Xs[(int64_t)j * prm.Xld + i]), e);
          eta_g[i] = e;
        }
        if ((int)threadIdx.x < prm.D) w[(int64_t)threadIdx.x * P + p] = w_sh[threadIdx.x];
        rho = xi; ++acc;
      }
    }
    if (threadIdx.x == 0) { ll[p] = rho; if (acc_out) acc_out[p] = acc; }
  }
}

// log_likelihood_at / log_joint_at: recompute from the logit cache (non-mutating)
__global__ void __launch_bounds__(256)
blr_loglik_at_kernel(const double *__restrict__ w, const double *__restrict__ eta,
                     BLRParams prm, int64_t P, int joint_block, double *__restrict__ out) {
  __shared__ double red[34];
  for (int64_t p = blockIdx.x; p < P; p += gridDim.x) {
    const double *e = eta + p * prm.Ncap;
    double part = 0.0;
    for (int64_t i = threadIdx.x; i < prm.N; i += blockDim.x) part += fast_log_logistic(e[i]);
    double tot = block_sum(part, red);
    if (threadIdx.x == 0) {
      // joint: + prior density of the principal node(s) (dispatch.py:345-361)
      if (joint_block == VK_BLOCK_ALL) {
        for (int j = 0; j < prm.D; ++j) tot += normal_logpdf(w[(int64_t)j * P + p], prm.mu0, prm.s0);
      } else if (joint_block >= 0) {
        tot += normal_logpdf(w[(int64_t)joint_block * P + p], prm.mu0, prm.s0);
      }
      out[p] = tot;
    }
  }
}

__global__ void blr_sign_features_kernel(const double *__restrict__ Xrows, const uint8_t *__restrict__ y,
                                         double *__restrict__ Xs, int D, int64_t Ncap, int64_t i_base,
                                         int64_t n) {
  int64_t total = n * D;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    int64_t i = t / D; int j = (int)(t - i * D);
    double s = y[i] ? 1.0 : -1.0;
    Xs[(int64_t)j * Ncap + i_base + i] = s * Xrows[t];
  }
}

__global__ void blr_fill_const_kernel(double *out, int64_t n, double v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    out[i] = v;
}

struct BLRModel : Model {
  BLRParams prm;
  int64_t n_prop = 0;
  DevBuf<double> Xs, wa, wb, etaa, etab, lla, llb, stage;
  DevBuf<uint8_t> y;
  double *w = nullptr, *w_alt = nullptr, *eta = nullptr, *eta_alt = nullptr, *ll = nullptr, *ll_alt = nullptr;
  int smem_limit = 0, coop_launch = 0;
  FmCoef fm_coef = fm_coef_host();
  // subsampled MH (C5): observation shard [obs_lo, obs_lo + N) of N_total, transition state
  int64_t obs_lo = 0, N_total = -1;
  bool eta_dirty = false;
  // fp32 throughput mode (blr_mh_f32.cuh): float copy of Xs, and "the logit cache holds float-rounded
  // values and ll an fp32-accurate sum" -- refreshed from w in fp64 before any fp64 consumer
  int precision = VK_FP64;
  bool eta_lowprec = false;
  DevBuf<float> Xs32;
  int64_t Xs32_rows = -1;
  DevBuf<double> Xr, ss_d, ss_partial, ss_sums;
  DevBuf<int> ss_i;
  int64_t Xr_rows = 0;
  int *ss_active_host = nullptr;
  // persistent subsampled-MH kernel (blr_subsampled_persistent.cuh)
  DevBuf<double> ssp_partial, ssp_tcrit;
  DevBuf<unsigned long long> ssp_bar;
  DevBuf<unsigned long long> ssp_seq, ssp_prof;
  int *ssp_abort_host = nullptr, *ssp_abort_dev = nullptr;     // mapped pinned: a timed-out barrier / exchange
  std::vector<double> tcrit_host;
  int64_t tcrit_Nbatch = -1, tcrit_Nt = -1; double tcrit_eps = -1.0; int tcrit_k0 = -1;

  ~BLRModel() override {
    if (ss_active_host) cudaFreeHost(ss_active_host);
    if (ssp_abort_host) cudaFreeHost(ssp_abort_host);
  }
  // per-particle subsampled-MH state is sized by the particle count (resample with n_new != P)
  void ss_release() { ss_d.release(); ss_i.release(); ss_sums.release(); ss_partial.release(); ssp_partial.release(); }

  BLRModel(vk_particles *h_, int D, int64_t Ncap, double mu0, double s0) : Model(h_) {
    VK_REQUIRE(D >= 1 && D <= 256, VK_EINVAL, "blr: D must be in [1,256]");
    VK_REQUIRE(Ncap >= 1, VK_EINVAL, "blr: observation capacity must be positive");
    prm.mu0 = mu0; prm.s0 = s0; prm.D = D; prm.N = 0; prm.Ncap = (Ncap + 1) & ~(int64_t)1;
    // Xs columns are zero padded to a multiple of the MH kernel's tile plus one tile of prefetch slack
    prm.Xld = ((prm.Ncap + BLR_MH_TILE - 1) / BLR_MH_TILE + 1) * BLR_MH_TILE;
    Xs.alloc((size_t)D * prm.Xld);
    VK_CUDA(cudaMemsetAsync(Xs.p, 0, Xs.bytes(), h->stream));
    y.alloc((size_t)prm.Ncap);
    wa.alloc((size_t)D * h->P);
    etaa.alloc((size_t)h->P * prm.Ncap);
    lla.alloc((size_t)h->P);
    w = wa.p; eta = etaa.p; ll = lla.p;
    VK_CUDA(cudaMemsetAsync(w, 0, wa.bytes(), h->stream));
    VK_CUDA(cudaMemsetAsync(ll, 0, lla.bytes(), h->stream));
    VK_CUDA(cudaDeviceGetAttribute(&smem_limit, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
    VK_CUDA(cudaDeviceGetAttribute(&coop_launch, cudaDevAttrCooperativeLaunch, h->device));
  }
  void observe(const double *X_rows_host, const uint8_t *y_host, int64_t n) {
    VK_REQUIRE(prm.N + n <= prm.Ncap, VK_EINVAL, "blr: observation capacity exceeded");
    // grow-only staging buffer: no cudaMalloc / cudaFree (an implicit device synchronisation) per call
    if (stage.n < (size_t)n * prm.D) { VK_CUDA(cudaStreamSynchronize(h->stream)); stage.alloc((size_t)n * prm.D); }
    VK_CUDA(cudaMemcpyAsync(stage.p, X_rows_host, (size_t)n * prm.D * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaMemcpyAsync(y.p + prm.N, y_host, (size_t)n, cudaMemcpyHostToDevice, h->stream));
    blr_sign_features_kernel<<<grid_for(n * prm.D, 256, h->sm_count, 8), 256, 0, h->stream>>>(
        stage.p, y.p + prm.N, Xs.p, prm.D, prm.Xld, prm.N, n);
    VK_CUDA(cudaGetLastError());
    VK_CUDA(cudaStreamSynchronize(h->stream));
    prm.N += n;
    Xr_rows = -1; Xs32_rows = -1;
  }
  void set_precision(int prec) override {
    VK_REQUIRE(prec == VK_FP64 || prec == VK_FP32, VK_EINVAL, "precision must be VK_FP64 or VK_FP32");
    precision = prec;
  }
  int get_precision() const override { return precision; }
  int grid_particles(int ctas_per_sm) const {
    int64_t cap = (int64_t)h->sm_count * ctas_per_sm;
    return (int)(h->P < cap ? h->P : cap);
  }
  void launch_all(int mode, double *incr_dev, uint32_t step, bool allow_f32 = true) {
    const bool f32 = allow_f32 && precision == VK_FP32;
    size_t sm = (size_t)(BLR_ALL_PP * prm.D + 34) * sizeof(double) + (f32 ? (size_t)BLR_ALL_PP * prm.D * sizeof(float) : 0);
    int64_t groups = (h->P + BLR_ALL_PP - 1) / BLR_ALL_PP, cap = (int64_t)h->sm_count * 8;
    const int grid = (int)(groups < cap ? groups : cap);
    if (f32) {
      blr_all_kernel<true><<<grid, 256, sm, h->stream>>>(w, eta, ll, h->lw.p, incr_dev, Xs.p, prm, n_prop, h->P, h->p_base, h->seed, step, mode);
      eta_lowprec = true;          // the cache holds float-rounded logits: an fp64 consumer recomputes it first
    } else {
      blr_all_kernel<false><<<grid, 256, sm, h->stream>>>(w, eta, ll, h->lw.p, incr_dev, Xs.p, prm, n_prop, h->P, h->p_base, h->seed, step, mode);
    }
    VK_CUDA(cudaGetLastError());
  }
  void reset_to_prior() override { launch_all(0, nullptr, h->next_step()); }
  void likelihood_weight() override {
    n_prop = prm.N;
    launch_all(1, nullptr, h->next_step());
  }
  // recompute the logit cache and ll from the weights -- in DOUBLE unless the caller is the fp32 mode itself
  void refresh() override { refresh_in(false); }
  void refresh_in(bool f32) { eta_lowprec = false; launch_all(2, nullptr, 0, f32); eta_dirty = false; }
  // subsampled transitions leave the logit cache stale (subsampled_mh.py:59-73);
  // this is SubsampledInPlaceOperator.makeConsistent (:427-455) done lazily.
  // fp64 consumers need an fp64 cache; in fp32 mode a float-rounded cache is what the mode promises (1e-5)
  void ensure_fresh() {
    if (precision == VK_FP32) { if (eta_dirty) refresh_in(true); }
    else if (eta_dirty || eta_lowprec) refresh_in(false);
  }
  void set_observation_shard(int64_t lo, int64_t total) {
    VK_REQUIRE(lo >= 0 && total >= lo, VK_EINVAL, "blr: bad observation shard");
    obs_lo = lo; N_total = total;
  }
  void ss_prepare() {
    if (Xr_rows != prm.N) {
      Xr.alloc((size_t)prm.N * prm.D);
      ss_transpose_kernel<<<grid_for(prm.N * prm.D, 256, h->sm_count, 8), 256, 0, h->stream>>>(Xs.p, Xr.p, prm.D, prm.Xld, prm.N);
      VK_CUDA(cudaGetLastError());
      Xr_rows = prm.N;
    }
    if (!ss_d.p) {
      // site, status, n_active (= gate), block counter, second status array of the persistent kernel
      ss_d.alloc((size_t)h->P * 11); ss_i.alloc((size_t)h->P * 3 + 2);
      VK_CUDA(cudaMemsetAsync(ss_i.p, 0, ss_i.bytes(), h->stream));
      ss_sums.alloc((size_t)h->P * 2);
      VK_CUDA(cudaMallocHost((void **)&ss_active_host, sizeof(int)));
    }
  }
  SSState ss_state() {
    SSState S; int64_t P = h->P; double *d = ss_d.p;
    S.w_new = d; S.delta = d + P; S.mu0 = d + 2 * P; S.cum = d + 3 * P; S.cum2 = d + 4 * P; S.n = d + 5 * P;
    S.k = d + 6 * P; S.tstat = d + 7 * P; S.acc_total = d + 8 * P; S.n_total = d + 9 * P;
    S.site = ss_i.p; S.status = ss_i.p + P;
    return S;
  }
  int ss_grid() const { return h->sm_count * 2; }
  // device count of undecided particles after the last decision (set nonzero by ss_propose): the round
  // kernels return at once when it is zero, so rounds enqueued past the last decision cost a launch only
  int *ss_gate() { return ss_i.p + 2 * h->P; }
  // --- round-level primitives (also driven from the host layer when observations are sharded)
  void ss_begin_call() {
    ss_prepare();
    VK_CUDA(cudaMemsetAsync(ss_d.p + 8 * h->P, 0, (size_t)h->P * 2 * sizeof(double), h->stream));
  }
  void ss_propose(int block, int t_index, int delta_mode, double eps) {
    int64_t Nt = N_total >= 0 ? N_total : prm.N;
    VK_REQUIRE(Nt > 1, VK_ESTATE, "subsampled_mh needs more than one local section (subsampled_mh.py:100)");
    uint32_t step = h->next_step();
    ss_key = make_feistel_key(h->seed, step, (uint64_t)Nt);
    ss_last_active = -1;     // a new transition: every particle is undecided again
    VK_CUDA(cudaMemsetAsync(ss_gate(), 0x01, sizeof(int), h->stream));
    ss_propose_kernel<<<grid_for(h->P, 128, h->sm_count, 4), 128, 0, h->stream>>>(
        w, ss_state(), prm.D, h->P, h->p_base, prm.mu0, prm.s0, (double)Nt, block, t_index, delta_mode, eps, h->seed, step);
    VK_CUDA(cudaGetLastError());
  }
  FeistelKey ss_key;
  int ss_last_active = -1;   // undecided particles after the last decision that was read back (-1: unknown)
  // local partial sums of minibatch positions [q0,q1) -> ss_sums (device, P x 2)
  // fuse_decide: the caller is the single-GPU loop and wants the decision of this minibatch made too;
  // returns true when the round kernel did it (compact rounds), false when ss_decide_launch is still due
  bool ss_round(int64_t q0, int64_t q1, bool fuse_decide = false, int k0 = 0, double eps = 0.0) {
    int n_groups = (int)((h->P + 31) / 32);
    VK_REQUIRE(n_groups <= 8, VK_EUNSUPPORTED, "subsampled_mh: at most 256 particles per GPU in this version");
    int grid = ss_grid();
    if (ss_partial.n < (size_t)grid * h->P * 2) ss_partial.alloc((size_t)grid * h->P * 2);
    size_t sm = (size_t)8 * 32 * (prm.D > 2 ? prm.D : 2) * sizeof(double);   // per-warp row tiles (reused for the partial sums)
    SSState S = ss_state();
    // few undecided particles left (count of the last decision read back): lanes = observations
    const bool compact = ss_last_active >= 0 && ss_last_active <= SS_COMPACT_MAX;
    const int fused = (compact && fuse_decide) ? 1 : 0;
    const double Nt_d = (double)(N_total >= 0 ? N_total : prm.N);
    unsigned int *ss_counter = (unsigned int *)(ss_i.p + 2 * h->P + 1);
#define VK_SS_LAUNCH(DT) do { if (compact) ss_round_compact_kernel<DT><<<grid, 256, 0, h->stream>>>(Xr.p, w, S, prm.D, h->P, q0, q1, obs_lo, prm.N, ss_key, ss_partial.p, ss_gate(), fused, ss_counter, Nt_d, k0, eps); else { VK_CUDA(cudaFuncSetAttribute(ss_round_kernel<DT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); ss_round_kernel<DT><<<grid, 256, sm, h->stream>>>(Xr.p, w, S, prm.D, h->P, q0, q1, obs_lo, prm.N, ss_key, ss_partial.p, ss_gate()); } } while (0)
    if (prm.D <= 4) VK_SS_LAUNCH(4);
    else if (prm.D <= 8) VK_SS_LAUNCH(8);
    else if (prm.D <= 16) VK_SS_LAUNCH(16);
    else if (prm.D <= 32) VK_SS_LAUNCH(32);
    else if (prm.D <= 64) VK_SS_LAUNCH(64);
    else throw Error(VK_EUNSUPPORTED, "subsampled_mh: D <= 64 in this version");
#undef VK_SS_LAUNCH
    if (fused) { VK_CUDA(cudaGetLastError()); return true; }
    ss_reduce_kernel<<<grid_for(h->P * 2 * 32, 128, h->sm_count, 4), 128, 0, h->stream>>>(ss_partial.p, grid, h->P, ss_sums.p, ss_gate());
    VK_CUDA(cudaGetLastError());
    return false;
  }
  // sequential-test decision from (possibly all-reduced) sums; ss_decide returns #particles still
  // undecided (host round trip); ss_decide_launch only enqueues, ss_decide_wait reads the count of
  // the LAST enqueued decision
  void ss_decide_launch(const double *sums_dev, int64_t n_end, int k0, double eps) {
    int64_t Nt = N_total >= 0 ? N_total : prm.N;
    int *n_active = ss_i.p + 2 * h->P;
    VK_CUDA(cudaMemsetAsync(n_active, 0, sizeof(int), h->stream));
    ss_decide_kernel<<<grid_for(h->P, 32, h->sm_count, 8), 32, 0, h->stream>>>(ss_state(), sums_dev, h->P, (double)n_end,
                                                                               (double)Nt, k0, eps, n_active);
    VK_CUDA(cudaGetLastError());
  }
  int ss_decide_wait() {
    int *n_active = ss_i.p + 2 * h->P;
    VK_CUDA(cudaMemcpyAsync(ss_active_host, n_active, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    VK_CUDA(cudaStreamSynchronize(h->stream));
    ss_last_active = *ss_active_host;
    return *ss_active_host;
  }
  int ss_decide(const double *sums_dev, int64_t n_end, int k0, double eps) {
    ss_decide_launch(sums_dev, n_end, k0, eps);
    return ss_decide_wait();
  }
  void ss_commit() {
    ss_commit_kernel<<<grid_for(h->P, 128, h->sm_count, 4), 128, 0, h->stream>>>(w, ss_state(), h->P);
    VK_CUDA(cudaGetLastError());
    eta_dirty = true;
  }
  void ss_stats(int K, double *out_dev) {
    ss_stats_kernel<<<grid_for(h->P, 128, h->sm_count, 4), 128, 0, h->stream>>>(ss_state(), h->P, K, out_dev);
    VK_CUDA(cudaGetLastError());
  }
  // t.ppf(1 - eps, n_k - 1) per minibatch count k (host, cached per configuration); NaN where no t test can
  // happen at that count (k < k0, n_k == N) and beyond the table (the kernel then evaluates the cdf itself)
  void ss_tcrit_table(int64_t Nbatch, int64_t Nt, int k0, double eps) {
    if (tcrit_Nbatch == Nbatch && tcrit_Nt == Nt && tcrit_k0 == k0 && tcrit_eps == eps && ssp_tcrit.p) return;
    const int64_t rounds = (Nt + Nbatch - 1) / Nbatch;
    const int n = (int)(rounds < 4096 ? rounds : 4096);
    tcrit_host.assign((size_t)(n > 0 ? n : 1), NAN);
    for (int ki = 0; ki < n; ++ki) {
      const int64_t k = ki + 1;
      const int64_t nk = k * Nbatch < Nt ? k * Nbatch : Nt;
      if ((k < k0 && nk < Nt) || nk == Nt || nk < 2) continue;
      tcrit_host[(size_t)ki] = eps > 0.0 ? ssp_host_t_ppf(1.0 - eps, (double)(nk - 1)) : INFINITY;
    }
    VK_CUDA(cudaStreamSynchronize(h->stream));      // a previous launch may still read the old table
    ssp_tcrit.alloc(tcrit_host.size());
    VK_CUDA(cudaMemcpyAsync(ssp_tcrit.p, tcrit_host.data(), ssp_tcrit.bytes(), cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaStreamSynchronize(h->stream));
    tcrit_Nbatch = Nbatch; tcrit_Nt = Nt; tcrit_k0 = k0; tcrit_eps = eps;
  }
  void ss_check_abort() {
    if (ssp_abort_host && *(volatile int *)ssp_abort_host)
      throw Error(VK_ECUDA, "subsampled_mh: the persistent kernel gave up waiting (a grid barrier or a peer exchange timed out)");
  }
  template <int DT>
  void ss_launch_persistent(SSPersist &a) {
    const size_t sm = ssp_smem_bytes(DT, h->P);
    VK_REQUIRE(sm <= (size_t)smem_limit, VK_EUNSUPPORTED, "subsampled_mh: particle count too large for the kernel's shared memory");
    VK_CUDA(cudaFuncSetAttribute(ss_persistent_kernel<DT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    int per_sm = 0;
    VK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ss_persistent_kernel<DT>, SSP_THREADS, sm));
    VK_REQUIRE(per_sm >= 1, VK_ECUDA, "subsampled_mh: the persistent kernel does not fit an SM");
    if (per_sm > 2) per_sm = 2;
    static const int cap = getenv("VK_SS_GRID") ? atoi(getenv("VK_SS_GRID")) : 0;
    int grid = h->sm_count * per_sm;
    if (cap > 0 && cap < grid) grid = cap;
    const size_t need = (size_t)grid * SSP_GMAX * (size_t)h->P * 2;
    if (ssp_partial.n < need) { VK_CUDA(cudaStreamSynchronize(h->stream)); ssp_partial.alloc(need); }
    a.partial = ssp_partial.p;
    void *args[] = {(void *)&a};
    VK_CUDA(cudaLaunchCooperativeKernel((const void *)ss_persistent_kernel<DT>, dim3(grid), dim3(SSP_THREADS), args, sm, h->stream));
  }
  // every transition of the call in one cooperative launch (world > 1: with the peer exchange inside)
  void ss_persistent(int block, int64_t Nbatch, int k0, double eps, int delta_mode, double delta_arg, int n_steps) {
    const int64_t Nt = N_total >= 0 ? N_total : prm.N;
    VK_REQUIRE(Nt > 1, VK_ESTATE, "subsampled_mh needs more than one local section (subsampled_mh.py:100)");
    ss_begin_call();
    if (!ssp_bar.p) {
      ssp_bar.alloc(4);
      ssp_seq.alloc(1);
      VK_CUDA(cudaMemsetAsync(ssp_seq.p, 0, sizeof(unsigned long long), h->stream));
      VK_CUDA(cudaHostAlloc((void **)&ssp_abort_host, sizeof(int), cudaHostAllocMapped));
      *ssp_abort_host = 0;
      VK_CUDA(cudaHostGetDevicePointer((void **)&ssp_abort_dev, ssp_abort_host, 0));
    }
    ss_check_abort();
    ss_tcrit_table(Nbatch, Nt, k0, eps);
    SSPersist a;
    memset(&a, 0, sizeof a);
    a.Xr = Xr.p; a.w = w; a.S = ss_state(); a.D = prm.D; a.P = h->P; a.p_base = h->p_base;
    a.obs_lo = obs_lo; a.N_local = prm.N; a.Nt = Nt; a.Nbatch = Nbatch; a.k0 = k0; a.eps = eps;
    a.block_mode = block; a.n_transitions = n_steps; a.delta_mode = delta_mode; a.delta_eps = delta_arg;
    a.mu0_prior = prm.mu0; a.s0_prior = prm.s0; a.seed = h->seed; a.step0 = h->next_step((uint32_t)n_steps);
    a.tcrit = ssp_tcrit.p; a.n_tcrit = (int)tcrit_host.size();
    a.bar = ssp_bar.p; a.abort_host = ssp_abort_dev; a.seq_dev = ssp_seq.p; a.cf = fm_coef;
    a.status_alt = ss_i.p + 2 * h->P + 2;
    P2PSSView v;
    const bool peers = N_total >= 0 && p2p_ss_view(h, &v) && v.world > 1;
    a.world = 1; a.rank = 0; a.timeout_ns = 10000000000ll;
    if (peers) {
      a.tab = v.tab; a.par = v.par; a.err_host = v.err_host_dev; a.world = v.world; a.rank = v.rank;
      a.ss_stride = v.ss_stride; a.ss_pcap = v.ss_pcap; a.timeout_ns = v.timeout_ns;
    }
    // rounds evaluated per batch (speculation past the deciding round is discarded): the fewer particles are
    // still undecided, the more rounds a batch holds -- the work per batch stays near G_work particle-rounds
    static const int g_min_env = getenv("VK_SS_ROUNDS_PER_BATCH") ? atoi(getenv("VK_SS_ROUNDS_PER_BATCH")) : 0;
    static const int g_work_env = getenv("VK_SS_BATCH_WORK") ? atoi(getenv("VK_SS_BATCH_WORK")) : 0;
    a.G_first = k0 < 1 ? 1 : (k0 > SSP_GMAX ? SSP_GMAX : k0);
    a.G_min = g_min_env > 0 ? (g_min_env > SSP_GMAX ? SSP_GMAX : g_min_env) : (peers ? 4 : 2);
    a.G_work = g_work_env > 0 ? g_work_env : 128;
    VK_CUDA(cudaMemsetAsync(ssp_bar.p, 0, ssp_bar.bytes(), h->stream));
    static const bool profile = getenv("VK_SS_PROFILE") != nullptr;       // phase times of CTA 0, printed per call
    if (profile) {
      if (!ssp_prof.p) ssp_prof.alloc(16 + 2 * 1024);
      VK_CUDA(cudaMemsetAsync(ssp_prof.p, 0, ssp_prof.bytes(), h->stream));
      a.prof = ssp_prof.p;
    }
    if (prm.D <= 4) ss_launch_persistent<4>(a);
    else if (prm.D <= 8) ss_launch_persistent<8>(a);
    else if (prm.D <= 16) ss_launch_persistent<16>(a);
    else ss_launch_persistent<32>(a);
    VK_CUDA(cudaGetLastError());
    eta_dirty = true;
    if (profile) {
      static unsigned long long pr[16 + 2 * 1024];
      VK_CUDA(cudaMemcpyAsync(pr, ssp_prof.p, sizeof pr, cudaMemcpyDeviceToHost, h->stream));
      VK_CUDA(cudaStreamSynchronize(h->stream));
      for (int kind = 0; kind < 2; ++kind) {          // spread of the CTAs' evaluation times (the barrier waits for the slowest)
        unsigned long long mn = ~0ull, mx = 0, sum = 0; int imx = 0, n = 0;
        for (int b = 0; b < 1024; ++b) {
          const unsigned long long v = pr[16 + 2 * b + kind];
          if (!v) continue;
          ++n; sum += v; if (v < mn) mn = v; if (v > mx) { mx = v; imx = b; }
        }
        if (n) fprintf(stderr, "[vk ss profile] eval %s per CTA (us over the call): min %.1f mean %.1f max %.1f (CTA %d), CTA 0 %.1f, %d CTAs\n",
                       kind ? "direct" : "staged", mn * 1e-3, sum * 1e-3 / n, mx * 1e-3, imx, pr[16 + kind] * 1e-3, n);
      }
      fprintf(stderr, "[vk ss profile] transitions %d | us: propose+bar %.1f list %.1f eval_staged %.1f eval_direct %.1f bar_eval %.1f decide %.1f "
                      "bar_decide %.1f | batches staged %llu (rounds %llu) direct %llu (rounds %llu) mean undecided %.2f\n", n_steps,
              pr[6] * 1e-3, pr[0] * 1e-3, pr[1] * 1e-3, pr[2] * 1e-3, pr[3] * 1e-3, pr[4] * 1e-3, pr[5] * 1e-3, pr[8], pr[10], pr[9], pr[11],
              (pr[8] + pr[9]) ? (double)pr[12] / (double)(pr[8] + pr[9]) : 0.0);
    }
  }
  bool ss_persistent_ok() const {
    static const int off = getenv("VK_SS_HOST_LOOP") ? atoi(getenv("VK_SS_HOST_LOOP")) : 0;
    return !off && prm.D <= 32 && h->P <= 4096 && coop_launch;
  }
  void subsampled_mh(int scope, int block, int64_t Nbatch, int k0, double eps, int use_delta, double delta_arg,
                     int update_values, int K, double *mean_nodes_dev, double *stats_dev) override {
    VK_REQUIRE(scope == 0, VK_EINVAL, "blr: only the default scope exists");
    // the scaffold cut at the global border: the weight, its lookup in `(array w_0 ...)`, the array (3 nodes)
    if (mean_nodes_dev) blr_fill_const_kernel<<<grid_for(h->P, 256, h->sm_count, 4), 256, 0, h->stream>>>(mean_nodes_dev, h->P, K > 0 ? 3.0 : 0.0);
    VK_REQUIRE(block == VK_BLOCK_ONE || block == VK_BLOCK_EACH || (block >= 0 && block < prm.D), VK_EINVAL, "blr: bad block");
    VK_REQUIRE(n_prop == prm.N, VK_ESTATE, "subsampled_mh: unpropagated observations; call vk_make_consistent first");
    VK_REQUIRE(eps >= 0.0 && eps < 0.5, VK_EINVAL, "subsampled_mh: epsilon must be in [0, 0.5)");
    // updateValues (scaffold.py:119-146) refreshes a local section to the OLD global value before it is
    // regenerated; here every section is evaluated from the current global value anyway (no per-section
    // cache is read), so both settings give the same dllh
    (void)update_values;
    int64_t Nt = N_total >= 0 ? N_total : prm.N;
    int n_prop_steps = block == VK_BLOCK_EACH ? K * prm.D : K;
    if (ss_persistent_ok()) {
      if (n_prop_steps > 0) ss_persistent(block, Nbatch, k0, eps, use_delta == 2 ? 2 : 0, delta_arg, n_prop_steps);
      else ss_begin_call();
      if (stats_dev) ss_stats(n_prop_steps, stats_dev);
      return;
    }
    VK_REQUIRE(N_total < 0 || N_total == prm.N, VK_EUNSUPPORTED,
               "subsampled_mh: sharded observations need the persistent kernel (D <= 32, P <= 4096) or the round-level vk_ssmh_* calls");
    ss_begin_call();
    for (int t = 0; t < n_prop_steps; ++t) {
      ss_propose(block, t, use_delta == 2 ? 2 : 0, delta_arg);
      // Rounds are enqueued in groups and the "anybody still undecided?" count is read back once
      // per group: a round after the last decision is a no-op (decided particles are skipped by
      // both kernels), so the result is identical to checking after every round.
      int GROUP = 4;           // grows: the stragglers' rounds are cheap (compact kernel, gated no-ops)
      int pending = 0;
      for (int64_t q0 = 0; q0 < Nt; q0 += Nbatch) {
        int64_t q1 = q0 + Nbatch < Nt ? q0 + Nbatch : Nt;
        if (!ss_round(q0, q1, true, k0, eps)) ss_decide_launch(ss_sums.p, q1, k0, eps);
        if (++pending == GROUP || q1 == Nt) { pending = 0; if (ss_decide_wait() == 0) break; if (GROUP < 16) GROUP *= 2; }
      }
      ss_commit();
    }
    if (stats_dev) ss_stats(n_prop_steps, stats_dev);
  }
  void make_consistent(double *incr_dev) override {
    if (n_prop == prm.N) {
      if (incr_dev) VK_CUDA(cudaMemsetAsync(incr_dev, 0, (size_t)h->P * sizeof(double), h->stream));
      return;
    }
    launch_all(3, incr_dev, 0);
    n_prop = prm.N;
  }
  void mh(int scope, int block, int K, double *mean_nodes_dev, long long *acc_dev) override {
    VK_REQUIRE(scope == 0, VK_EINVAL, "blr: only the default scope exists");
    VK_REQUIRE(block == VK_BLOCK_ONE || block == VK_BLOCK_EACH || block == VK_BLOCK_ALL ||
                   (block >= 0 && block < prm.D), VK_EINVAL, "blr: bad block");
    VK_REQUIRE(n_prop == prm.N, VK_ESTATE, "mh: unpropagated observations; call vk_make_consistent first");
    const bool f32 = precision == VK_FP32 && block != VK_BLOCK_ALL && K > 0 && f32_fits();
    // the fp32 loop converts the cache itself; every other kernel of this entry point is double and wants a double cache
    if (f32) { if (eta_dirty) refresh(); } else if (eta_dirty || eta_lowprec) refresh_in(false);
    double nodes = 0.0;
    if (K > 0) {
      if (block == VK_BLOCK_ALL) {
        uint32_t step0 = h->next_step((uint32_t)K);
        size_t sm = (size_t)(((prm.D + 1) & ~1) + 34) * sizeof(double);
        blr_mh_all_kernel<<<grid_particles(8), 256, sm, h->stream>>>(w, eta, ll, Xs.p, prm, h->P, h->p_base,
                                                                     h->seed, step0, K, acc_dev);
        nodes = (double)(prm.D * 2 + 1 + 2 * prm.N);
      } else {
        int n_steps = (block == VK_BLOCK_EACH) ? K * prm.D : K;
        uint32_t step0 = h->next_step((uint32_t)n_steps);
        VK_REQUIRE(prm.N <= ((int64_t)1 << 19), VK_EUNSUPPORTED, "blr mh: at most 2^19 observations per particle in this version");
        if (f32) {
          launch_mh_f32(step0, n_steps, block, acc_dev);
          nodes = (double)(3 + 2 * prm.N) * (block == VK_BLOCK_EACH ? prm.D : 1);
        } else {
        size_t base = blr_mh_smem_base(prm.D);
        int64_t Npad = (prm.N + BLR_MH_TILE - 1) / BLR_MH_TILE * BLR_MH_TILE;
        size_t need = base + (size_t)Npad * sizeof(double);
        // tuning knob for experiments: VK_BLR_MH_VARIANT = 0 (512 threads x 2 chains, 2 CTAs/SM, 64 regs),
        // 1 (256 x 4, 2 CTAs/SM, 128 regs), 2 (1024 x 1, 1 CTA/SM, 64 regs), 3 (512 x 2, 1 CTA/SM, 128 regs)
        static int variant = getenv("VK_BLR_MH_VARIANT") ? atoi(getenv("VK_BLR_MH_VARIANT")) : BLR_MH_DEFAULT_VARIANT;
        bool in_smem = need <= (size_t)smem_limit;
        size_t sm_bytes = in_smem ? need : base;
        int per_sm = (int)((size_t)(227 * 1024) / (sm_bytes + 1024));
        if (per_sm < 1) per_sm = 1;
#define VK_MH_LAUNCH(S, T, U, B)                                                                                \
  do {                                                                                                         \
    int cap = B < per_sm ? B : per_sm;                                                                         \
    VK_CUDA(cudaFuncSetAttribute(blr_mh_kernel<S, T, U, B>, cudaFuncAttributeMaxDynamicSharedMemorySize,          \
                                 (int)smem_limit));                                                            \
    blr_mh_kernel<S, T, U, B><<<grid_particles(cap), T, sm_bytes, h->stream>>>(                                   \
        w, eta, ll, Xs.p, prm, h->P, h->p_base, h->seed, step0, n_steps, block, acc_dev, fm_coef);             \
  } while (0)
#define VK_MH_VARIANT(T, U, B) do { if (in_smem) VK_MH_LAUNCH(true, T, U, B); else VK_MH_LAUNCH(false, T, U, B); } while (0)
        const size_t need2 = blr_mh2_smem_base(prm.D) + (size_t)Npad * sizeof(double);
        if (variant >= 4 && need2 <= (size_t)smem_limit) {
          // two proposals per pass over the shared-memory row (blr_mh2.cuh)
          int per2 = (int)((size_t)(227 * 1024) / (need2 + 1024));
          if (per2 < 1) per2 = 1;
          VK_CUDA(cudaFuncSetAttribute(blr_mh2_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_limit));
          blr_mh2_kernel<2><<<grid_particles(per2 < 2 ? per2 : 2), MH2_THREADS, need2, h->stream>>>(
              w, eta, ll, Xs.p, prm, h->P, h->p_base, h->seed, step0, n_steps, block, acc_dev, fm_coef);
        } else
        switch (variant) {
          case 1: case 4: VK_MH_VARIANT(256, 4, 2); break;
          case 2: VK_MH_VARIANT(1024, 1, 1); break;
          case 3: VK_MH_VARIANT(512, 2, 1); break;
          default: VK_MH_VARIANT(512, 2, 2); break;
        }
#undef VK_MH_VARIANT
#undef VK_MH_LAUNCH
        // |DRG| of scaffold(w_j) = w_j + lookup + array + N lookups + N dots;
        // 'each' sums D proposals per transition (dispatch.py:122-128)
        nodes = (double)(3 + 2 * prm.N) * (block == VK_BLOCK_EACH ? prm.D : 1);
        }
      }
      VK_CUDA(cudaGetLastError());
    }
    if (mean_nodes_dev) {
      blr_fill_const_kernel<<<grid_for(h->P, 256, h->sm_count, 4), 256, 0, h->stream>>>(mean_nodes_dev, h->P, nodes);
      VK_CUDA(cudaGetLastError());
    }
  }
  // fp32 mode needs the float logit row of one particle in shared memory; longer rows run the fp64 kernel
  size_t f32_smem() const {
    int64_t Npad = (prm.N + BLR_F32_TILE - 1) / BLR_F32_TILE * BLR_F32_TILE;
    return blr_mh_f32_smem_base(prm.D) + (size_t)Npad * sizeof(float);
  }
  bool f32_fits() const { return f32_smem() <= (size_t)smem_limit; }
  void launch_mh_f32(uint32_t step0, int n_steps, int block, long long *acc_dev) {
    if (Xs32_rows != prm.N) {
      if (!Xs32.p) Xs32.alloc(Xs.n);
      blr_to_float_kernel<<<grid_for((int64_t)Xs.n, 256, h->sm_count, 8), 256, 0, h->stream>>>(Xs.p, Xs32.p, (int64_t)Xs.n);
      VK_CUDA(cudaGetLastError());
      Xs32_rows = prm.N;
    }
    size_t sm = f32_smem();
    int per_sm = (int)((size_t)(227 * 1024) / (sm + 1024));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 4) per_sm = 4;
    VK_CUDA(cudaFuncSetAttribute(blr_mh_f32_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_limit));
    blr_mh_f32_kernel<4><<<grid_particles(per_sm), BLR_F32_THREADS, sm, h->stream>>>(
        w, eta, ll, Xs32.p, prm, h->P, h->p_base, h->seed, step0, n_steps, block, acc_dev);
    eta_lowprec = true;
  }
  void log_likelihood_at(int scope, int block, bool joint, double *out_dev) override {
    VK_REQUIRE(scope == 0, VK_EINVAL, "blr: only the default scope exists");
    VK_REQUIRE(block == VK_BLOCK_ALL || (block >= 0 && block < prm.D), VK_EINVAL,
               "log_likelihood_at: block must be all or a site index");
    ensure_fresh();
    BLRParams q = prm; q.N = n_prop;
    blr_loglik_at_kernel<<<grid_particles(8), 256, 0, h->stream>>>(w, eta, q, h->P, joint ? block : -100, out_dev);
    VK_CUDA(cudaGetLastError());
  }
  void forget_observations(double *w_dev) override {
    if (w_dev) log_likelihood_at(0, VK_BLOCK_ALL, false, w_dev);
    prm.N = 0; n_prop = 0; Xr_rows = -1; Xs32_rows = -1; eta_dirty = false; eta_lowprec = false;
    VK_CUDA(cudaMemsetAsync(Xs.p, 0, Xs.bytes(), h->stream));   // keep the zero padding invariant
    VK_CUDA(cudaMemsetAsync(ll, 0, (size_t)h->P * sizeof(double), h->stream));
  }
  int64_t num_sites() const override { return prm.D; }
  int64_t num_blocks(int) const override { return prm.D; }
  void read_site(int64_t site, double *out_dev) override {
    VK_REQUIRE(site >= 0 && site < prm.D, VK_EINVAL, "blr: site out of range");
    VK_CUDA(cudaMemcpyAsync(out_dev, w + site * h->P, (size_t)h->P * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  }
  void write_site(int64_t site, const double *in_dev) override {
    VK_REQUIRE(site >= 0 && site < prm.D, VK_EINVAL, "blr: site out of range");
    VK_CUDA(cudaMemcpyAsync(w + site * h->P, in_dev, (size_t)h->P * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  }
  std::vector<StateArray> state_arrays(bool alloc_alt) override {
    ensure_fresh();
    // back buffers are allocated on first resample (eta doubles the footprint)
    if (alloc_alt && !wb.p) { wb.alloc(wa.n); etab.alloc(etaa.n); llb.alloc(lla.n); w_alt = wb.p; eta_alt = etab.p; ll_alt = llb.p; }
    StateArray a{(void **)&w, (void **)&w_alt, 0, (size_t)prm.D, 8};
    StateArray b{(void **)&eta, (void **)&eta_alt, (size_t)prm.Ncap * 8, 0, 0};
    StateArray c{(void **)&ll, (void **)&ll_alt, 0, 1, 8};
    return {a, b, c};
  }
  void resized(int64_t n_new, const std::vector<void *> &nc, const std::vector<void *> &na) override {
    const size_t n = (size_t)n_new;
    wa.adopt(nc[0], n * prm.D); wb.adopt(na[0], n * prm.D);
    etaa.adopt(nc[1], n * prm.Ncap); etab.adopt(na[1], n * prm.Ncap);
    lla.adopt(nc[2], n); llb.adopt(na[2], n);
    w = wa.p; w_alt = wb.p; eta = etaa.p; eta_alt = etab.p; ll = lla.p; ll_alt = llb.p;
    ss_release();            // per-particle subsampled-MH state is sized by the particle count
  }
  int device_buffer(int which, void **p, int64_t *bytes) override {
    if (which == 1) { *p = w; *bytes = (int64_t)wa.bytes(); return VK_OK; }
    if (which == 2) { *p = eta; *bytes = (int64_t)etaa.bytes(); return VK_OK; }
    if (which == 3) { *p = ll; *bytes = (int64_t)lla.bytes(); return VK_OK; }
    if (which == 4) { *p = Xs.p; *bytes = (int64_t)Xs.bytes(); return VK_OK; }
    return VK_EINVAL;
  }
};

Model *make_blr(vk_particles *h, int D, int64_t Ncap, double mu0, double s0) {
  return new BLRModel(h, D, Ncap, mu0, s0);
}
void blr_observe(Model *m, const double *X_rows_host, const uint8_t *y_host, int64_t n) {
  static_cast<BLRModel *>(m)->observe(X_rows_host, y_host, n);
}
void blr_set_observation_shard(Model *m, int64_t lo, int64_t total) {
  static_cast<BLRModel *>(m)->set_observation_shard(lo, total);
}
// round-level subsampled-MH primitives for the sharded host driver
void blr_ss_begin(Model *m) { static_cast<BLRModel *>(m)->ss_begin_call(); }
void blr_ss_propose(Model *m, int block, int t_index, int delta_mode, double eps) {
  static_cast<BLRModel *>(m)->ss_propose(block, t_index, delta_mode, eps);
}
double *blr_ss_round(Model *m, int64_t q0, int64_t q1) {
  BLRModel *b = static_cast<BLRModel *>(m);
  b->ss_round(q0, q1);
  return b->ss_sums.p;
}
int blr_ss_decide(Model *m, const double *sums_dev, int64_t n_end, int k0, double eps) {
  return static_cast<BLRModel *>(m)->ss_decide(sums_dev, n_end, k0, eps);
}
void blr_ss_commit(Model *m) { static_cast<BLRModel *>(m)->ss_commit(); }
void blr_ss_stats(Model *m, int K, double *out_dev) { static_cast<BLRModel *>(m)->ss_stats(K, out_dev); }

}  // namespace vk
