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
static const int SSP_ROWS = 128;                 // observation rows staged per pass
static const int SSP_CHUNKS = SSP_ROWS / 32;     // 32-row chunks of a pass (one warp-wide slice each)
static const int SSP_STAGES = 3;                 // passes in flight (cp.async ring)
static const int SSP_TILE = 64;                  // undecided particles handled per sweep over the passes
static const int SSP_GMAX = P2P_SS_MAX_ROUNDS;   // rounds per batch
static const int SSP_LIST = 256;                 // direct path: local rows a warp lists before it evaluates them
static const int SSP_DIRECT = 4;                 // at most this many undecided particles: one thread per position, no staging
static const int SSP_PART = SSP_GMAX * SSP_DIRECT * SSP_WARPS * 2 > SSP_CHUNKS * SSP_TILE * 2
                                ? SSP_GMAX * SSP_DIRECT * SSP_WARPS * 2 : SSP_CHUNKS * SSP_TILE * 2;   // doubles of scratch

struct SSPersist {                 // by-value kernel argument
  const double *Xr; double *w; SSState S;
  int *status_alt;                              // second status array: transitions alternate between the two
  int D; int64_t P, p_base, obs_lo, N_local, Nt, Nbatch;
  int k0; double eps;
  int block_mode, n_transitions, delta_mode; double delta_eps, mu0_prior, s0_prior;
  uint64_t seed; uint32_t step0;
  const double *tcrit; int n_tcrit;             // t.ppf(1 - eps, n_k - 1) for minibatch count k = index + 1
  int G_first, G_min, G_work;                   // rounds per batch: k0 first, then clamp(G_work / #undecided, G_min, SSP_GMAX)
  double *partial;                              // [grid][SSP_GMAX][P][2]
  unsigned long long *bar;                      // [0] grid barrier counter, [1] abort flag; zeroed before the launch
  int *abort_host;                              // host-mapped mirror of the abort flag (written only on a timeout)
  unsigned long long *seq_dev;                  // exchanges completed so far (persists across launches)
  P2PTables *tab; P2PParams *par; volatile int *err_host;
  int world, rank; long long ss_stride, ss_pcap, timeout_ns;
  unsigned long long *prof;                     // nullable: [16] phase times (ns) and counts seen by CTA 0 (VK_SS_PROFILE)
  FmCoef cf;
};
// prof slots: 0 list, 1 eval staged, 2 eval direct, 3 barrier after eval, 4 decide, 5 barrier after decide, 6 propose + its
// barrier, 8 batches staged, 9 batches direct, 10 rounds staged, 11 rounds direct, 12 sum of #undecided over batches

__host__ __device__ inline size_t ssp_smem_bytes(int DT, int64_t P) {
  return (size_t)FM_SMEM_LOG_BYTES + FM_SMEM_EXP_BYTES +
         (size_t)SSP_STAGES * SSP_ROWS * (DT + 1) * sizeof(double) +      // staged rows
         (size_t)SSP_TILE * DT * sizeof(double) +                         // weights of the tile's particles
         (size_t)SSP_TILE * sizeof(double) +                              // their proposal deltas
         (size_t)SSP_PART * sizeof(double) +                              // per-chunk sums of a pass / per-warp sums of a batch
         (size_t)SSP_TILE * 2 * sizeof(double) +                          // the current round's running sums
         (size_t)(SSP_STAGES * SSP_ROWS + SSP_STAGES * (SSP_WARPS + 1) + 2 * SSP_TILE) * sizeof(int) +
         (size_t)((P + 3) & ~(int64_t)3) * sizeof(int);
}

__device__ __forceinline__ void ssp_set_abort(const SSPersist &a) {
  *(volatile unsigned long long *)(a.bar + 1) = 1ull;
  *(volatile int *)a.abort_host = 1;
  __threadfence_system();
}
__device__ __forceinline__ bool ssp_aborted(const SSPersist &a) { return *(volatile unsigned long long *)(a.bar + 1) != 0ull; }

// all CTAs arrive, then all leave.  `epoch` counts this CTA's barriers.  false = timed out (abort flag set).
__device__ __forceinline__ bool ssp_grid_barrier(const SSPersist &a, unsigned long long &epoch) {
  __shared__ int ok_sh;
  __syncthreads();
  ++epoch;
  if (threadIdx.x == 0) {
    const unsigned long long target = epoch * (unsigned long long)gridDim.x;
    __threadfence();
    atomicAdd(a.bar, 1ull);
    int ok = 1;
    unsigned long long t0 = 0;
    int spins = 0;
    while (*(volatile unsigned long long *)a.bar < target) {
      if ((++spins & 1023) == 0) {
        if (ssp_aborted(a)) { ok = 0; break; }
        const unsigned long long now = p2p_now_ns();
        if (t0 == 0) t0 = now;
        else if ((long long)(now - t0) > a.timeout_ns) { ssp_set_abort(a); ok = 0; break; }
      }
    }
    if (ssp_aborted(a)) ok = 0;
    __threadfence();
    ok_sh = ok;
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

__device__ __forceinline__ void ssp_cp_async8(void *dst_shared, const void *src_global) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_shared)), "l"(src_global));
}
__device__ __forceinline__ void ssp_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_PENDING> __device__ __forceinline__ void ssp_cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_PENDING) : "memory"); }

template <int DT>
__global__ void __launch_bounds__(SSP_THREADS, 2) ss_persistent_kernel(const SSPersist a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned char *log_tbl = smem_raw;
  unsigned char *exp_tbl = smem_raw + FM_SMEM_LOG_BYTES;
  double *rows_sh = reinterpret_cast<double *>(exp_tbl + FM_SMEM_EXP_BYTES);      // SSP_STAGES x SSP_ROWS x (DT+1)
  double *w_sh = rows_sh + SSP_STAGES * SSP_ROWS * (DT + 1);                      // SSP_TILE x DT
  double *delta_sh = w_sh + SSP_TILE * DT;                                        // SSP_TILE
  double *part_sh = delta_sh + SSP_TILE;                                          // SSP_PART
  double *acc_sh = part_sh + SSP_PART;                                            // SSP_TILE x 2
  int *idx_sh = reinterpret_cast<int *>(acc_sh + SSP_TILE * 2);                   // SSP_STAGES x SSP_ROWS local observation indices
  int *cnt_sh = idx_sh + SSP_STAGES * SSP_ROWS;                                   // SSP_STAGES x (SSP_WARPS + 1): per-warp counts, rows of the pass
  int *site_sh = cnt_sh + SSP_STAGES * (SSP_WARPS + 1);                           // SSP_TILE
  int *pid_sh = site_sh + SSP_TILE;                                               // SSP_TILE
  int *act_sh = pid_sh + SSP_TILE;                                                // P: undecided particles, in order
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int D = a.D;
  const int64_t P = a.P;
  fm_fill_smem_tables(log_tbl, exp_tbl);
  const uint32_t exp_base = (uint32_t)__cvta_generic_to_shared(exp_tbl) + (uint32_t)(lane & 15) * 8u;
  const uint32_t log_base = (uint32_t)__cvta_generic_to_shared(log_tbl) + (uint32_t)(lane & 7) * 16u;
  unsigned long long epoch = 0;
  unsigned long long seq = *(volatile unsigned long long *)a.seq_dev;
  const double Nd = (double)a.Nt;
  // positions of a full minibatch per CTA, and the passes that takes (the same count for a short last round)
  const int64_t C_full = (a.Nbatch + gridDim.x - 1) / gridDim.x;
  const int ppr = (int)((C_full + SSP_ROWS - 1) / SSP_ROWS) > 0 ? (int)((C_full + SSP_ROWS - 1) / SSP_ROWS) : 1;
  __syncthreads();

  const bool prof = a.prof != nullptr && blockIdx.x == 0 && tid == 0;
  unsigned long long tp = prof ? p2p_now_ns() : 0ull;
#define SSP_PROF(slot) do { if (prof) { const unsigned long long now_ = p2p_now_ns(); a.prof[slot] += now_ - tp; tp = now_; } } while (0)
  for (int t = 0; t < a.n_transitions; ++t) {
    const uint32_t step = a.step0 + (uint32_t)t;
    // the status array alternates between transitions: a CTA that is already proposing transition t + 1 must not
    // disturb a slower CTA that is still reading the statuses of transition t; the LAST transition uses S.status
    int *status = ((a.n_transitions - 1 - t) & 1) ? a.status_alt : a.S.status;
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
      a.S.cum[p] = 0.0; a.S.cum2[p] = 0.0; a.S.n[p] = 0.0; a.S.k[p] = 0.0; a.S.tstat[p] = 0.0; status[p] = 0;
    }
    if (!ssp_grid_barrier(a, epoch)) return;
    SSP_PROF(6);
    const FeistelKey key = make_feistel_key(a.seed, step, (uint64_t)a.Nt);
    int64_t q_done = 0;       // permutation positions covered by the rounds evaluated so far
    int k_done = 0;           // minibatches evaluated so far
    while (true) {
      // ---- undecided particles, in index order (every CTA builds the same list)
      int nA = 0;
      for (int64_t p0 = 0; p0 < P; p0 += SSP_THREADS) {
        const int64_t p = p0 + tid;
        const bool und = p < P && ((volatile int *)status)[p] == 0;
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
      int G;
      if (k_done == 0) G = a.G_first;
      else { G = a.G_work / nA; if (G < a.G_min) G = a.G_min; if (G > SSP_GMAX) G = SSP_GMAX; }
      if ((int64_t)G > rounds_left) G = (int)rounds_left;
      const int n_pass = G * ppr;
      const bool direct = DT <= 16 && nA <= SSP_DIRECT;
      SSP_PROF(0);
      const unsigned long long te0 = (a.prof != nullptr && tid == 0) ? p2p_now_ns() : 0ull;
      if (prof) { a.prof[direct ? 9 : 8] += 1; a.prof[direct ? 11 : 10] += (unsigned long long)G; a.prof[12] += (unsigned long long)nA; }
      if (direct) {
        // ---- few stragglers: ONE THREAD PER POSITION.  Rows go from L2/HBM straight into registers (the next
        // position's row is in flight while the current one is evaluated); per round, the warps' sums are combined
        // in warp order.  Same partial layout as the staged path below.
        const int nT = nA;
        for (int e = tid; e < SSP_DIRECT * DT; e += SSP_THREADS) {
          const int sl = e / DT, j = e - sl * DT;
          w_sh[e] = (sl < nT && j < D) ? a.w[(int64_t)j * P + act_sh[sl]] : 0.0;
        }
        if (tid < SSP_DIRECT) {
          const int p = tid < nT ? act_sh[tid] : 0;
          pid_sh[tid] = p;
          delta_sh[tid] = tid < nT ? a.S.delta[p] : 0.0;
          site_sh[tid] = tid < nT ? a.S.site[p] : 0;
        }
        __syncthreads();
        // A ROUND BELONGS TO A WARP (wpr warps share one when the batch has fewer rounds than the CTA has warps).  The
        // warp turns its positions into local rows first -- Feistel image, this rank's shard only, compacted in
        // position order into a list -- and then evaluates the list 32 rows at a time.  With the observations sharded
        // over R ranks only 1/R of the positions are local: evaluating position by position left 1 - 1/R of the
        // lanes idle through the whole softplus arithmetic (8 GPUs: 2.9 us per round against 5.9 on one).
        const int wpr = G >= SSP_WARPS ? 1 : SSP_WARPS / G;
        const int g_stride = SSP_WARPS / wpr, sub = warp % wpr;
        int *list = reinterpret_cast<int *>(rows_sh) + warp * SSP_LIST;       // the staged ring is idle on this path
        for (int g = warp / wpr; g < G; g += g_stride) {
          const int64_t q0 = q_done + (int64_t)g * a.Nbatch;
          const int64_t q1 = q0 + a.Nbatch < a.Nt ? q0 + a.Nbatch : a.Nt;
          const int64_t Cg = (q1 - q0 + gridDim.x - 1) / gridDim.x;
          const int64_t qa_cta = q0 + (int64_t)blockIdx.x * Cg;
          const int64_t qb_cta = qa_cta + Cg < q1 ? qa_cta + Cg : q1;
          const int64_t Cw = (Cg + wpr - 1) / wpr;
          const int64_t qa = qa_cta + (int64_t)sub * Cw;
          const int64_t qb = qa + Cw < qb_cta ? qa + Cw : qb_cta;
          double a1[SSP_DIRECT], a2[SSP_DIRECT];
#pragma unroll
          for (int sl = 0; sl < SSP_DIRECT; ++sl) { a1[sl] = 0.0; a2[sl] = 0.0; }
          for (int64_t qs = qa; qs < qb; qs += SSP_LIST) {                     // (warp-uniform loops throughout)
            const int64_t qe = qs + SSP_LIST < qb ? qs + SSP_LIST : qb;
            int n_rows = 0;
            for (int64_t qq = qs; qq < qe; qq += 32) {
              const int64_t q = qq + lane;
              int64_t i_loc = -1;
              if (q < qe) {
                i_loc = (int64_t)feistel_perm((uint64_t)q, key) - a.obs_lo;
                if (i_loc < 0 || i_loc >= a.N_local) i_loc = -1;
              }
              const unsigned bal = __ballot_sync(0xffffffffu, i_loc >= 0);
              if (i_loc >= 0) list[n_rows + __popc(bal & ((1u << lane) - 1u))] = (int)i_loc;
              n_rows += __popc(bal);
            }
            __syncwarp();
            const int n_chunks = (n_rows + 31) >> 5;
            auto fetch = [&](int c, double (&x)[DT]) -> bool {
              const int r = c * 32 + lane;
              const bool ok = c < n_chunks && r < n_rows;
              if (ok) {
                const double *row = a.Xr + (int64_t)list[r] * D;
                if ((D & 1) == 0) {          // rows start on 16-byte boundaries
#pragma unroll
                  for (int j = 0; j < DT; j += 2) {
                    double2 v = make_double2(0.0, 0.0);
                    if (j < D) v = __ldg(reinterpret_cast<const double2 *>(row + j));
                    x[j] = v.x; x[j + 1] = v.y;
                  }
                } else {
#pragma unroll
                  for (int j = 0; j < DT; ++j) x[j] = j < D ? __ldg(row + j) : 0.0;
                }
              }
              return ok;
            };
            auto consume = [&](bool ok, const double (&x)[DT]) {
              if (!ok) return;
#pragma unroll
              for (int sl = 0; sl < SSP_DIRECT; ++sl) {
                if (sl < nT) {
                  const double *wv = w_sh + sl * DT;
                  const int site = site_sh[sl];
                  double eta = 0.0, xs = 0.0;
#pragma unroll
                  for (int j = 0; j < DT; ++j) { eta = fma(wv[j], x[j], eta); xs = (j == site) ? x[j] : xs; }
                  const double v_new = fma(delta_sh[sl], xs, eta);
                  const double d = ssp_dllh(v_new, eta, a.cf, exp_base, log_base);
                  a1[sl] += d; a2[sl] = fma(d, d, a2[sl]);
                }
              }
            };
            double xa[DT], xb[DT];
            bool oka = fetch(0, xa), okb = false;
            for (int c = 0; c < n_chunks; c += 2) {                            // the next chunk's rows are in flight
              okb = fetch(c + 1, xb);
              consume(oka, xa);
              if (c + 1 < n_chunks) {
                oka = fetch(c + 2, xa);
                consume(okb, xb);
              }
            }
            __syncwarp();                                                      // before the list is rewritten
          }
#pragma unroll
          for (int sl = 0; sl < SSP_DIRECT; ++sl) {
            if (sl < nT) {
              const double u1 = warp_sum(a1[sl]), u2 = warp_sum(a2[sl]);
              if (lane == 0) {
                double *dst = part_sh + (((size_t)g * SSP_DIRECT + sl) * SSP_WARPS + sub) * 2;
                dst[0] = u1; dst[1] = u2;
              }
            }
          }
        }
        __syncthreads();
        for (int e = tid; e < G * nT * 2; e += SSP_THREADS) {
          const int comp = e & 1, sl = (e >> 1) % nT, g = (e >> 1) / nT;
          const double *src = part_sh + (((size_t)g * SSP_DIRECT + sl) * SSP_WARPS) * 2 + comp;
          double sum = 0.0;
          for (int wv = 0; wv < wpr; ++wv) sum += src[wv * 2];                 // the round's warps, in position order
          a.partial[(((size_t)blockIdx.x * SSP_GMAX + g) * (size_t)P + (size_t)pid_sh[sl]) * 2 + comp] = sum;
        }
        __syncthreads();
      } else
      // ---- evaluate G rounds: a sweep over this CTA's passes per tile of undecided particles
      for (int tile0 = 0; tile0 < nA; tile0 += SSP_TILE) {
        const int nT = nA - tile0 < SSP_TILE ? nA - tile0 : SSP_TILE;
        for (int e = tid; e < SSP_TILE * DT; e += SSP_THREADS) {
          const int s = e / DT, j = e - s * DT;
          w_sh[e] = (s < nT && j < D) ? a.w[(int64_t)j * P + act_sh[tile0 + s]] : 0.0;
        }
        if (tid < SSP_TILE) {
          const int p = tid < nT ? act_sh[tile0 + tid] : 0;
          pid_sh[tid] = p;
          delta_sh[tid] = tid < nT ? a.S.delta[p] : 0.0;
          site_sh[tid] = tid < nT ? a.S.site[p] : 0;
        }
        if (tid < SSP_TILE * 2) acc_sh[tid] = 0.0;
        // stage pass i into ring slot i % SSP_STAGES: Feistel positions -> this rank's rows, compacted in position
        // order -> asynchronous copies of the rows.  Every thread commits exactly one group per call.
        auto stage = [&](int i) {
          const int buf = i % SSP_STAGES;
          const int g = i / ppr, kpass = i - g * ppr;
          const int64_t q0 = q_done + (int64_t)g * a.Nbatch;
          const int64_t q1 = q0 + a.Nbatch < a.Nt ? q0 + a.Nbatch : a.Nt;
          const int64_t Cg = (q1 - q0 + gridDim.x - 1) / gridDim.x;
          const int64_t qa = q0 + (int64_t)blockIdx.x * Cg;
          const int64_t qb = qa + Cg < q1 ? qa + Cg : q1;
          const int64_t q = qa + (int64_t)kpass * SSP_ROWS + tid;
          int64_t i_loc = -1;
          if (tid < SSP_ROWS && q < qb) {
            i_loc = (int64_t)feistel_perm((uint64_t)q, key) - a.obs_lo;
            if (i_loc < 0 || i_loc >= a.N_local) i_loc = -1;
          }
          int *cnt = cnt_sh + buf * (SSP_WARPS + 1);
          int *idx = idx_sh + buf * SSP_ROWS;
          const unsigned bal = __ballot_sync(0xffffffffu, i_loc >= 0);
          if (lane == 0) cnt[warp] = __popc(bal);
          __syncthreads();
          int before = 0, n_rows = 0;
#pragma unroll
          for (int wv = 0; wv < SSP_WARPS; ++wv) { const int c = cnt[wv]; if (wv < warp) before += c; n_rows += c; }
          if (i_loc >= 0) idx[before + __popc(bal & ((1u << lane) - 1u))] = (int)i_loc;
          if (tid == 0) cnt[SSP_WARPS] = n_rows;
          __syncthreads();
          double *rows = rows_sh + (size_t)buf * SSP_ROWS * (DT + 1);
          for (int e = tid; e < n_rows * D; e += SSP_THREADS) {
            const int k = e / D, j = e - k * D;
            ssp_cp_async8(&rows[k * (DT + 1) + j], &a.Xr[(int64_t)idx[k] * D + j]);
          }
          ssp_cp_commit();
        };
        __syncthreads();
        for (int i = 0; i < SSP_STAGES - 1; ++i) { if (i < n_pass) stage(i); else ssp_cp_commit(); }
        for (int i = 0; i < n_pass; ++i) {
          if (i + SSP_STAGES - 1 < n_pass) stage(i + SSP_STAGES - 1); else ssp_cp_commit();
          ssp_cp_wait<SSP_STAGES - 1>();       // pass i's copies have landed (this thread's); the barrier makes it everyone's
          __syncthreads();
          const int buf = i % SSP_STAGES;
          const int g = i / ppr;
          const int n_rows = cnt_sh[buf * (SSP_WARPS + 1) + SSP_WARPS];
          const int n_chunks = (n_rows + 31) >> 5;
          if (n_rows > 0 && n_rows <= 16) {
            // a thin pass (sharded observations: 1/R of a pass's positions are local): the two half-warps take the
            // same <= 16 rows and DIFFERENT particles.  The 16-lane shuffle tree gives bit for bit what the 32-lane
            // tree gives with its upper half zero, so the sums do not depend on which mapping a pass used.
            const int r = lane & 15, hsel = lane >> 4;
            const bool valid = r < n_rows;
            const double *row = rows_sh + (size_t)buf * SSP_ROWS * (DT + 1) + (valid ? r : 0) * (DT + 1);
            double x[DT];
#pragma unroll
            for (int j = 0; j < DT; ++j) x[j] = (valid && j < D) ? row[j] : 0.0;
            for (int slot0 = warp * 2; slot0 < nT; slot0 += SSP_WARPS * 2) {
              const int slot = slot0 + hsel;
              const bool live = valid && slot < nT;
              const double *wv = w_sh + (slot < nT ? slot : 0) * DT;
              double eta = 0.0;
#pragma unroll
              for (int j = 0; j < DT; ++j) eta = fma(wv[j], x[j], eta);
              const double xs = live ? row[site_sh[slot]] : 0.0;
              const double v_new = fma(live ? delta_sh[slot] : 0.0, xs, eta);
              double d = ssp_dllh(v_new, eta, a.cf, exp_base, log_base);
              if (!live) d = 0.0;
              double u1 = d, u2 = d * d;
#pragma unroll
              for (int o = 8; o > 0; o >>= 1) { u1 += __shfl_xor_sync(0xffffffffu, u1, o); u2 += __shfl_xor_sync(0xffffffffu, u2, o); }
              if (r == 0 && slot < nT) { part_sh[slot * 2] = u1; part_sh[slot * 2 + 1] = u2; }
            }
          } else if (n_chunks > 0) {
            // warps_per_chunk warps share a 32-row chunk and split the tile's particles between them
            const int wpc = SSP_WARPS / n_chunks;
            const int c = warp / wpc, sub = warp - c * wpc;
            if (c < n_chunks) {
              const int r = c * 32 + lane;
              const bool valid = r < n_rows;
              const double *row = rows_sh + (size_t)buf * SSP_ROWS * (DT + 1) + (valid ? r : 0) * (DT + 1);
              double x[DT];
#pragma unroll
              for (int j = 0; j < DT; ++j) x[j] = (valid && j < D) ? row[j] : 0.0;
              for (int slot = sub; slot < nT; slot += wpc) {
                const double *wv = w_sh + slot * DT;
                double eta = 0.0;
#pragma unroll
                for (int j = 0; j < DT; ++j) eta = fma(wv[j], x[j], eta);
                const double xs = valid ? row[site_sh[slot]] : 0.0;
                const double v_new = fma(delta_sh[slot], xs, eta);
                double d = ssp_dllh(v_new, eta, a.cf, exp_base, log_base);
                if (!valid) d = 0.0;
                const double u1 = warp_sum(d), u2 = warp_sum(d * d);
                if (lane == 0) { part_sh[(c * SSP_TILE + slot) * 2] = u1; part_sh[(c * SSP_TILE + slot) * 2 + 1] = u2; }
              }
            }
          }
          __syncthreads();
          // chunk sums -> the round's running sums, in chunk order; the last pass of a round publishes them
          if (tid < nT * 2) {
            double s = acc_sh[tid];
            for (int c = 0; c < n_chunks; ++c) s += part_sh[c * SSP_TILE * 2 + tid];
            if (i - g * ppr == ppr - 1) {
              a.partial[(((size_t)blockIdx.x * SSP_GMAX + g) * (size_t)P + (size_t)pid_sh[tid >> 1]) * 2 + (tid & 1)] = s;
              s = 0.0;
            }
            acc_sh[tid] = s;
          }
          // part_sh is rewritten only after the next pass's barriers
        }
        __syncthreads();       // before w_sh / pid_sh of the next tile are written
      }
      if (a.prof != nullptr && tid == 0) a.prof[16 + 2 * blockIdx.x + (direct ? 1 : 0)] += p2p_now_ns() - te0;
      SSP_PROF(direct ? 2 : 1);
      if (!ssp_grid_barrier(a, epoch)) return;
      SSP_PROF(3);
      // ---- decide: the handler of an undecided particle is warp 0 of CTA (list index mod grid)
      for (int ai = blockIdx.x; ai < nA; ai += gridDim.x) {          // CTA-uniform loop
        const int64_t p = act_sh[ai];
        // every warp sums the CTAs' partials of some rounds (fixed order: lane-strided, then the shuffle tree)
        // (a lane's loads of a round are issued together, 12 at a time: one L2 round trip instead of a chain of them)
        for (int g = warp; g < G; g += SSP_WARPS) {
          double s1 = 0.0, s2 = 0.0;
          for (int b0 = 0; b0 < (int)gridDim.x; b0 += 32 * 12) {
            double2 v[12];
#pragma unroll
            for (int k = 0; k < 12; ++k) {
              const int bb = b0 + k * 32 + lane;
              v[k] = make_double2(0.0, 0.0);
              if (bb < (int)gridDim.x)
                v[k] = __ldcg(reinterpret_cast<const double2 *>(a.partial + (((size_t)bb * SSP_GMAX + g) * (size_t)P + (size_t)p) * 2));
            }
#pragma unroll
            for (int k = 0; k < 12; ++k) { s1 += v[k].x; s2 += v[k].y; }      // the same order as a lane-strided loop
          }
          s1 = warp_sum(s1); s2 = warp_sum(s2);
          if (lane == 0) { acc_sh[2 * g] = s1; acc_sh[2 * g + 1] = s2; }
        }
        __syncthreads();
        if (warp == 0) do {
          double my1 = 0.0, my2 = 0.0;
          if (lane < G) { my1 = acc_sh[2 * lane]; my2 = acc_sh[2 * lane + 1]; }
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
                    ((long long)(p2p_now_ns() - t0) > a.timeout_ns || ssp_aborted(a))) { bad = 1; break; }
              }
            }
            bad = __any_sync(0xffffffffu, bad);
            if (bad) {
              if (lane == 0) { a.par->err = 1; *a.err_host = 1; ssp_set_abort(a); }
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
            status[p] = st_f;
          }
        } while (false);
        __syncthreads();             // acc_sh is reused by the next particle of this handler
      }
      q_done += (int64_t)G * a.Nbatch;
      if (q_done > a.Nt) q_done = a.Nt;
      k_done += G;
      if (a.world > 1) ++seq;
      SSP_PROF(4);
      if (!ssp_grid_barrier(a, epoch)) return;
      SSP_PROF(5);
    }
  }
  if (blockIdx.x == 0 && tid == 0) *a.seq_dev = seq;
#undef SSP_PROF
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
