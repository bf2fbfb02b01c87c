// Subsampled (austerity) MH as ONE persistent kernel per vk_subsampled_mh call: every transition of the call,
// every minibatch round of every sequential test, the decisions and -- when the observations are sharded over
// the GPUs of a node -- the exchange of the per-particle partial sums through peer memory all
// happen on the device; the host enqueues one cooperative launch and reads nothing back until the call ends.
//
// Reference: backend/lite/infer/subsampled_mh.py
//   subsampledMixMH :76-111 (proposal, mu_0, permutation)   sequentialTest :118-171
//   evalOneLocalSection :377-409 (dllh of one local section)
//
// Structure (grid = co-resident CTAs, cudaLaunchCooperativeKernel; the grid barrier is a monotone counter):
//   for each transition:
//     propose            one thread per particle (site, w', log u, mu_0)                          | barrier
//     while some particle is undecided:
//       evaluate a BATCH of G minibatch rounds: CTA b owns a slice of every round's positions; the Feistel
//       bijection turns positions into observations, the rows this rank owns are compacted IN ORDER and staged
//       in shared memory; LANES ARE OBSERVATIONS (row in registers), WARPS OWN PARTICLES (weights broadcast
//       from shared memory), so the work is proportional to the number of UNDECIDED particles; per (round,
//       particle) the CTA writes (sum dllh, sum dllh^2) of its slice                                | barrier
//       decide: particle p's handler (one warp) sums the CTAs' partials in a fixed order, [posts them to every
//       peer's mailbox, waits for the peers', adds them in RANK order -- identical numbers on every rank], and
//       runs the sequential test for the G rounds one after the other (the first round that decides wins;
//       later rounds of the batch were speculation and are ignored).  accept -> w' is installed    | barrier
//   The first batch of a transition holds k0 rounds (no test is due before), later ones G_next.  The sequence
//   of per-minibatch decisions -- what parity with sequentialTest pins -- does not depend on G.
//
// The Student-t test q < eps / q > 1 - eps (scipy.stats.t.cdf, :161) is evaluated as ts < -t_crit / ts > t_crit
// with t_crit = t.ppf(1 - eps, n - 1) tabulated on the host per minibatch count (the cdf is strictly monotone
// and symmetric; df = n - 1 depends on the round only, not on the particle): no incomplete-beta continued
// fraction on the decision path.  Rounds beyond the table fall back to the exact cdf.
#pragma once
#include "blr_subsampled.cuh"
#include "p2p_common.cuh"

namespace vk {

static const int SSP_THREADS = 256;
static const int SSP_WARPS = SSP_THREADS / 32;
static const int SSP_ROWS = 256;                 // observation rows staged per pass (= one position per thread)
static const int SSP_PER_WARP = 8;               // particles a warp accumulates in registers per pass
static const int SSP_TILE = SSP_WARPS * SSP_PER_WARP;   // particles per pass of a CTA
static const int SSP_GMAX = P2P_SS_MAX_ROUNDS;   // rounds per batch

struct SSPersist {                 // by-value kernel argument
  const double *Xr; double *w; SSState S;
  int D; int64_t P, p_base, obs_lo, N_local, Nt, Nbatch;
  int k0; double eps;
  int block_mode, n_transitions, delta_mode; double delta_eps, mu0_prior, s0_prior;
  uint64_t seed; uint32_t step0;
  const double *tcrit; int n_tcrit;             // t.ppf(1 - eps, n_k - 1) for minibatch count k = index + 1
  int G_first, G_next;
  double *partial;                              // [grid][SSP_GMAX][P][2]
  unsigned int *bar;                            // grid barrier counter, zeroed before the launch
  int *abort_flag;                              // set by a barrier / exchange that timed out
  unsigned long long *seq_dev;                  // exchanges completed so far (persists across launches)
  P2PTables *tab; P2PParams *par; volatile int *err_host;
  int world, rank; long long ss_stride, ss_pcap, timeout_ns;
  FmCoef cf;
};

__host__ __device__ inline size_t ssp_smem_bytes(int DT, int64_t P) {
  return (size_t)FM_SMEM_LOG_BYTES + FM_SMEM_EXP_BYTES + (size_t)SSP_ROWS * (DT + 1) * sizeof(double) +
         (size_t)SSP_TILE * DT * sizeof(double) + (size_t)SSP_TILE * sizeof(double) +
         (size_t)(SSP_ROWS + 2 * SSP_TILE + 16) * sizeof(int) + (size_t)((P + 3) & ~(int64_t)3) * sizeof(int);
}

// all CTAs arrive, then all leave.  `epoch` counts this CTA's barriers.  false = timed out (abort flag set).
__device__ __forceinline__ bool ssp_grid_barrier(const SSPersist &a, unsigned int &epoch) {
  __shared__ int ok_sh;
  __syncthreads();
  if (threadIdx.x == 0) {
    ++epoch;
    const unsigned int target = epoch * gridDim.x;
    __threadfence();
    atomicAdd(a.bar, 1u);
    int ok = 1;
    const unsigned long long t0 = p2p_now_ns();
    int spins = 0;
    while (*(volatile unsigned int *)a.bar < target) {
      if ((++spins & 255) == 0) {
        if (*(volatile int *)a.abort_flag) { ok = 0; break; }
        if ((long long)(p2p_now_ns() - t0) > a.timeout_ns) { *(volatile int *)a.abort_flag = 1; ok = 0; break; }
      }
    }
    if (*(volatile int *)a.abort_flag) ok = 0;
    __threadfence();
    ok_sh = ok;
  } else {
    ++epoch;
  }
  __syncthreads();
  return ok_sh != 0;
}

// dllh of one local section: log_logistic(v_new) - log_logistic(v_old) with the table-driven softplus of
// fast_math.cuh (term-by-term form: full relative accuracy of each term)
__device__ __forceinline__ double ssp_dllh(double v_new, double v_old, const FmCoef &cf, uint32_t exp_base,
                                           uint32_t log_base) {
  double af_n = 0.0, af_o = 0.0;
  long long aL_n = 0, aL_o = 0;
  fm_softplus_accum(v_new, cf, exp_base, log_base, af_n, aL_n);
  fm_softplus_accum(v_old, cf, exp_base, log_base, af_o, aL_o);
  const double dsp = (af_n - af_o) + (double)(aL_n - aL_o) * 2.220446049250313e-16;
  // min(v,0) = (v - |v|)/2
  const double dmin = 0.5 * ((v_new - v_old) - (fabs(v_new) - fabs(v_old)));
  return dmin - dsp;
}

template <int DT>
__global__ void __launch_bounds__(SSP_THREADS, 2) ss_persistent_kernel(const SSPersist a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned char *log_tbl = smem_raw;
  unsigned char *exp_tbl = smem_raw + FM_SMEM_LOG_BYTES;
  double *rows_sh = reinterpret_cast<double *>(exp_tbl + FM_SMEM_EXP_BYTES);      // SSP_ROWS x (DT+1)
  double *w_sh = rows_sh + SSP_ROWS * (DT + 1);                                   // SSP_TILE x DT
  double *delta_sh = w_sh + SSP_TILE * DT;                                        // SSP_TILE
  int *idx_sh = reinterpret_cast<int *>(delta_sh + SSP_TILE);                     // SSP_ROWS local observation indices
  int *site_sh = idx_sh + SSP_ROWS;                                               // SSP_TILE
  int *pid_sh = site_sh + SSP_TILE;                                               // SSP_TILE
  int *cnt_sh = pid_sh + SSP_TILE;                                                // 16: per-warp counts, totals
  int *act_sh = cnt_sh + 16;                                                      // P: undecided particles, in order
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int D = a.D;
  const int64_t P = a.P;
  fm_fill_smem_tables(log_tbl, exp_tbl);
  const uint32_t exp_base = (uint32_t)__cvta_generic_to_shared(exp_tbl) + (uint32_t)(lane & 15) * 8u;
  const uint32_t log_base = (uint32_t)__cvta_generic_to_shared(log_tbl) + (uint32_t)(lane & 7) * 16u;
  unsigned int epoch = 0;
  unsigned long long seq = *(volatile unsigned long long *)a.seq_dev;
  const double Nd = (double)a.Nt;
  __syncthreads();

  for (int t = 0; t < a.n_transitions; ++t) {
    const uint32_t step = a.step0 + (uint32_t)t;
    // ---- global proposal (:79-88,102), one thread per particle
    for (int64_t p = (int64_t)blockIdx.x * SSP_THREADS + tid; p < P; p += (int64_t)gridDim.x * SSP_THREADS) {
      Stream r(a.seed, (uint64_t)(a.p_base + p), step);
      const double u_sel = r.uniform();
      const double log_u = log(r.uniform());
      int j;
      if (a.block_mode == VK_BLOCK_ONE) { j = (int)(u_sel * (double)D); if (j >= D) j = D - 1; }
      else if (a.block_mode == VK_BLOCK_EACH) j = t % D;
      else j = a.block_mode;
      const double w_old = a.w[(int64_t)j * P + p];
      double w_new;
      if (a.delta_mode == 2) {
        const double nu = normal_sim(r, 0.0, a.s0_prior);
        w_new = a.mu0_prior + sqrt(1.0 - a.delta_eps * a.delta_eps) * (w_old - a.mu0_prior) + a.delta_eps * nu;
      } else {
        w_new = normal_sim(r, a.mu0_prior, a.s0_prior);
      }
      a.S.site[p] = j; a.S.w_new[p] = w_new; a.S.delta[p] = w_new - w_old;
      a.S.mu0[p] = (log_u - 0.0) / Nd;
      a.S.cum[p] = 0.0; a.S.cum2[p] = 0.0; a.S.n[p] = 0.0; a.S.k[p] = 0.0; a.S.tstat[p] = 0.0; a.S.status[p] = 0;
    }
    if (!ssp_grid_barrier(a, epoch)) return;
    const FeistelKey key = make_feistel_key(a.seed, step, (uint64_t)a.Nt);
    int64_t q_done = 0;       // permutation positions covered by the rounds evaluated so far
    int k_done = 0;           // minibatches evaluated so far
    while (true) {
      // ---- undecided particles, in index order (every CTA builds the same list)
      int nA = 0;
      for (int64_t p0 = 0; p0 < P; p0 += SSP_THREADS) {
        const int64_t p = p0 + tid;
        const bool und = p < P && ((volatile int *)a.S.status)[p] == 0;
        const unsigned bal = __ballot_sync(0xffffffffu, und);
        if (lane == 0) cnt_sh[warp] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int wv = 0; wv < SSP_WARPS; ++wv) { const int c = cnt_sh[wv]; if (wv < warp) before += c; total += c; }
        if (und) act_sh[nA + before + __popc(bal & ((1u << lane) - 1u))] = (int)p;
        nA += total;
        __syncthreads();
      }
      if (nA == 0) break;
      const int64_t rounds_left = (a.Nt - q_done + a.Nbatch - 1) / a.Nbatch;
      int G = k_done == 0 ? a.G_first : a.G_next;
      if ((int64_t)G > rounds_left) G = (int)rounds_left;
      // ---- evaluate G rounds
      for (int g = 0; g < G; ++g) {
        const int64_t q0 = q_done + (int64_t)g * a.Nbatch;
        const int64_t q1 = q0 + a.Nbatch < a.Nt ? q0 + a.Nbatch : a.Nt;
        const int64_t C = (q1 - q0 + gridDim.x - 1) / gridDim.x;
        const int64_t qa = q0 + (int64_t)blockIdx.x * C;
        const int64_t qb = qa + C < q1 ? qa + C : q1;
        double *part_g = a.partial + (((size_t)blockIdx.x * SSP_GMAX + g) * (size_t)P) * 2;
        bool first_pass = true;
        for (int64_t qs = qa; first_pass || qs < qb; qs += SSP_ROWS) {
          // rows of this rank among positions [qs, qs + SSP_ROWS), compacted in position order
          int64_t i = -1;
          if (qs + tid < qb) {
            i = (int64_t)feistel_perm((uint64_t)(qs + tid), key) - a.obs_lo;
            if (i < 0 || i >= a.N_local) i = -1;
          }
          const unsigned bal = __ballot_sync(0xffffffffu, i >= 0);
          if (lane == 0) cnt_sh[warp] = __popc(bal);
          __syncthreads();
          int before = 0, n_rows = 0;
#pragma unroll
          for (int wv = 0; wv < SSP_WARPS; ++wv) { const int c = cnt_sh[wv]; if (wv < warp) before += c; n_rows += c; }
          if (i >= 0) idx_sh[before + __popc(bal & ((1u << lane) - 1u))] = (int)i;
          __syncthreads();
          for (int e = tid; e < n_rows * D; e += SSP_THREADS) {
            const int k = e / D, j = e - k * D;
            rows_sh[k * (DT + 1) + j] = __ldg(&a.Xr[(int64_t)idx_sh[k] * D + j]);
          }
          for (int tile0 = 0; tile0 < nA; tile0 += SSP_TILE) {
            __syncthreads();     // rows staged; previous tile's weights no longer read
            for (int e = tid; e < SSP_TILE * DT; e += SSP_THREADS) {
              const int s = e / DT, j = e - s * DT;
              const int ai = tile0 + s;
              w_sh[e] = (ai < nA && j < D) ? a.w[(int64_t)j * P + act_sh[ai]] : 0.0;
            }
            if (tid < SSP_TILE) {
              const int ai = tile0 + tid;
              const int p = ai < nA ? act_sh[ai] : 0;
              pid_sh[tid] = p;
              delta_sh[tid] = ai < nA ? a.S.delta[p] : 0.0;
              site_sh[tid] = ai < nA ? a.S.site[p] : 0;
            }
            __syncthreads();
            double acc1[SSP_PER_WARP], acc2[SSP_PER_WARP];
#pragma unroll
            for (int s = 0; s < SSP_PER_WARP; ++s) { acc1[s] = 0.0; acc2[s] = 0.0; }
            for (int c0 = 0; c0 < n_rows; c0 += 32) {
              const int r = c0 + lane;
              const bool valid = r < n_rows;
              const double *row = rows_sh + (valid ? r : 0) * (DT + 1);
              double x[DT];
#pragma unroll
              for (int j = 0; j < DT; ++j) x[j] = (valid && j < D) ? row[j] : 0.0;
#pragma unroll
              for (int s = 0; s < SSP_PER_WARP; ++s) {
                const int slot = warp + SSP_WARPS * s;            // interleaved: few particles spread over the warps
                if (tile0 + slot < nA) {                          // warp-uniform
                  const double *wv = w_sh + slot * DT;
                  double eta = 0.0;
#pragma unroll
                  for (int j = 0; j < DT; ++j) eta = fma(wv[j], x[j], eta);
                  const double xs = valid ? row[site_sh[slot]] : 0.0;
                  const double v_new = fma(delta_sh[slot], xs, eta);
                  double d = ssp_dllh(v_new, eta, a.cf, exp_base, log_base);
                  if (!valid) d = 0.0;
                  acc1[s] += d; acc2[s] = fma(d, d, acc2[s]);
                }
              }
            }
#pragma unroll
            for (int s = 0; s < SSP_PER_WARP; ++s) {
              const int slot = warp + SSP_WARPS * s;
              if (tile0 + slot < nA) {
                const double u1 = warp_sum(acc1[s]), u2 = warp_sum(acc2[s]);
                if (lane == 0) {
                  double *dst = part_g + (size_t)pid_sh[slot] * 2;
                  if (first_pass) { dst[0] = u1; dst[1] = u2; } else { dst[0] += u1; dst[1] += u2; }
                }
              }
            }
          }
          __syncthreads();       // before idx_sh / rows_sh / w_sh are overwritten
          first_pass = false;
        }
      }
      if (!ssp_grid_barrier(a, epoch)) return;
      // ---- decide: the handler of an undecided particle is warp 0 of CTA (list index mod grid)
      if (warp == 0) {
        for (int ai = blockIdx.x; ai < nA; ai += gridDim.x) {
          const int64_t p = act_sh[ai];
          double my1 = 0.0, my2 = 0.0;
          for (int g = 0; g < G; ++g) {
            double s1 = 0.0, s2 = 0.0;
            for (int bb = lane; bb < (int)gridDim.x; bb += 32) {
              const double *src = a.partial + (((size_t)bb * SSP_GMAX + g) * (size_t)P + (size_t)p) * 2;
              s1 += __ldcg(src); s2 += __ldcg(src + 1);
            }
            s1 = warp_sum(s1); s2 = warp_sum(s2);
            if (lane == g) { my1 = s1; my2 = s2; }
          }
          if (a.world > 1) {
            const int parity = (int)(seq & 1ull);
            if (lane < G) {
              for (int r = 0; r < a.world; ++r) {
                volatile double *dst = a.tab->ss_mail[r] + ((size_t)(parity * P2P_MAX_WORLD + a.rank) * a.ss_stride) +
                                       ((size_t)lane * P + p) * 2;
                dst[0] = my1; dst[1] = my2;
              }
            }
            __threadfence_system();
            __syncwarp();
            int bad = 0;
            if (lane < a.world) {
              volatile unsigned long long *sq = a.tab->ss_seq[lane] + (size_t)(parity * P2P_MAX_WORLD + a.rank) * a.ss_pcap + p;
              *sq = seq + 1ull;
              volatile unsigned long long *mine = a.tab->ss_seq[a.rank] + (size_t)(parity * P2P_MAX_WORLD + lane) * a.ss_pcap + p;
              const unsigned long long t0 = p2p_now_ns();
              int spins = 0;
              while (*mine != seq + 1ull) {
                if ((++spins & 63) == 0 &&
                    ((long long)(p2p_now_ns() - t0) > a.timeout_ns || *(volatile int *)a.abort_flag)) { bad = 1; break; }
              }
            }
            bad = __any_sync(0xffffffffu, bad);
            if (bad) {
              if (lane == 0) { a.par->err = 1; *a.err_host = 1; *(volatile int *)a.abort_flag = 1; __threadfence_system(); }
              break;
            }
            __threadfence_system();
            if (lane < G) {
              double t1 = 0.0, t2 = 0.0;
              for (int r = 0; r < a.world; ++r) {       // rank order: identical sums on every rank
                const volatile double *src = a.tab->ss_mail[a.rank] + ((size_t)(parity * P2P_MAX_WORLD + r) * a.ss_stride) +
                                             ((size_t)lane * P + p) * 2;
                t1 += src[0]; t2 += src[1];
              }
              my1 = t1; my2 = t2;
            }
          }
          // running sums in round order (the reference adds one minibatch at a time, :132-139)
          double cum = a.S.cum[p], cum2 = a.S.cum2[p];
          double cg = 0.0, cg2 = 0.0;
          for (int g = 0; g < G; ++g) {
            cum += __shfl_sync(0xffffffffu, my1, g);
            cum2 += __shfl_sync(0xffffffffu, my2, g);
            if (lane == g) { cg = cum; cg2 = cum2; }
          }
          int st = 0;
          bool ts_valid = false;
          double ts = 0.0, n_g = 0.0, k_g = 0.0;
          if (lane < G) {
            const int64_t q_end = q_done + (int64_t)(lane + 1) * a.Nbatch;
            n_g = (double)(q_end < a.Nt ? q_end : a.Nt);
            k_g = a.S.k[p] + (double)(lane + 1);
            const double mx = cg / n_g, mx2 = cg2 / n_g, mu0 = a.S.mu0[p];
            if (k_g < (double)a.k0 && n_g < Nd) { /* no test during the first k0 minibatches (:141) */ }
            else if (n_g == Nd) st = (mx >= mu0) ? 1 : 2;
            else {
              const double sx = sqrt((1.0 - (n_g - 1.0) / (Nd - 1.0)) * (mx2 - mx * mx) / (n_g - 1.0));
              if (sx == 0.0) {
                const double q = (mx >= mu0) ? 1.0 : 0.0;
                if (q < a.eps) st = 2; else if (q > 1.0 - a.eps) st = 1;
              } else {
                ts = (mx - mu0) / sx; ts_valid = true;
                const int ki = k_done + lane;                         // minibatch count k = ki + 1
                const double tc = ki < a.n_tcrit ? a.tcrit[ki] : NAN;
                if (tc == tc) { if (ts < -tc) st = 2; else if (ts > tc) st = 1; }
                else { const double q = dev_student_t_cdf(ts, n_g - 1.0); if (q < a.eps) st = 2; else if (q > 1.0 - a.eps) st = 1; }
              }
            }
          }
          const unsigned dec = __ballot_sync(0xffffffffu, lane < G && st != 0);
          const int gs = dec ? __ffs(dec) - 1 : G - 1;                // the round whose state the particle keeps
          const unsigned tsm = __ballot_sync(0xffffffffu, ts_valid && lane <= gs);
          const int st_f = __shfl_sync(0xffffffffu, st, gs);
          const double cum_f = __shfl_sync(0xffffffffu, cg, gs), cum2_f = __shfl_sync(0xffffffffu, cg2, gs);
          const double n_f = __shfl_sync(0xffffffffu, n_g, gs), k_f = __shfl_sync(0xffffffffu, k_g, gs);
          const double ts_f = __shfl_sync(0xffffffffu, ts, tsm ? 31 - __clz(tsm) : 0);
          if (lane == 0) {
            a.S.cum[p] = cum_f; a.S.cum2[p] = cum2_f; a.S.n[p] = n_f; a.S.k[p] = k_f;
            if (tsm) a.S.tstat[p] = ts_f;
            if (st_f != 0) {
              if (st_f == 1) { a.w[(int64_t)a.S.site[p] * P + p] = a.S.w_new[p]; a.S.acc_total[p] += 1.0; }
              a.S.n_total[p] += n_f;
            }
            a.S.status[p] = st_f;
          }
        }
      }
      q_done += (int64_t)G * a.Nbatch;
      if (q_done > a.Nt) q_done = a.Nt;
      k_done += G;
      if (a.world > 1) ++seq;
      if (!ssp_grid_barrier(a, epoch)) return;
    }
  }
  if (blockIdx.x == 0 && tid == 0) *a.seq_dev = seq;
}

// ---- host: t.ppf(1 - eps, df) by Newton on the exact cdf (same incomplete-beta code as the device) --------
inline double ssp_host_normal_ppf(double p) {
  double lo = -40.0, hi = 40.0;
  for (int i = 0; i < 200; ++i) { double mid = 0.5 * (lo + hi); if (0.5 * erfc(-mid * 0.7071067811865476) < p) lo = mid; else hi = mid; }
  return 0.5 * (lo + hi);
}
inline double ssp_host_t_ppf(double p, double df) {
  if (!(p < 1.0)) return INFINITY;
  // bracket, then bisection refined by secant steps on the monotone cdf: robust for every df > 0
  double lo = 0.0, hi = ssp_host_normal_ppf(p);
  if (hi < 1.0) hi = 1.0;
  int guard = 0;
  while (host_student_t_cdf(hi, df) < p && guard++ < 2000) { lo = hi; hi *= 2.0; }
  double flo = host_student_t_cdf(lo, df) - p, fhi = host_student_t_cdf(hi, df) - p;
  for (int it = 0; it < 200; ++it) {
    double mid = (it & 1) && fhi != flo ? hi - fhi * (hi - lo) / (fhi - flo) : 0.5 * (lo + hi);
    if (!(mid > lo && mid < hi)) mid = 0.5 * (lo + hi);
    const double fm = host_student_t_cdf(mid, df) - p;
    if (fm < 0.0) { lo = mid; flo = fm; } else { hi = mid; fhi = fm; }
    if (hi - lo <= 2.3e-16 * hi) break;
  }
  return hi;
}

}  // namespace vk
