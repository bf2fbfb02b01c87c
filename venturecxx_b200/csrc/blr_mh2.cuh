// blr_mh2_kernel: the logistic-regression single-site MH loop (same site program, Philox streams and accept
// rule as blr_mh_kernel in k_blr.cu; infer/mh.py:84-89,187-208, regen.py:120-127, detach.py:72-83), rearranged
// around what ncu showed binds blr_mh_kernel (profiles/r1_blr_mh_v3_ncu.md): 34.7 issued instructions per
// evaluation for 13 fp64 ones, i.e. ISSUE slots, not the fp64 pipe (50.8 %).
//
//   * TWO PROPOSALS PER PASS over the shared-memory logit row.  Proposals come from the prior, so (site, w',
//     log u) of transition t+1 is known before t is decided, and xi_{t+1} computed against the CURRENT state is
//     exact whenever t is rejected (94-97 % of the time at C2).  A pass evaluates t and t+1 together: one read
//     of the logit pair, one loop, one block-wide reduction and one barrier for two transitions, and twice the
//     independent dependency chains per thread.  If t is accepted the speculative xi_{t+1} is discarded and the
//     next pass starts at t+1; otherwise t+1 is decided from it.  The chain is the sequential one, exactly.
//   * PAIRED LOADS: a thread owns observations (2i, 2i+1), so the logit pair is one 16-byte shared-memory read
//     and each proposal's feature pair one 16-byte L2 read (half the LSU instructions per evaluation).
//   * exp(-|v|): the |v| <= 700 clamp (3 integer instructions per evaluation) is gone -- the exponent index m
//     is clamped instead (1 instruction; a huge |v| then yields e ~ 1e-270, i.e. 0 next to 1), the exp table
//     holds 2^(j/128) (128 rows x 8 replicas) with j << 13 pre-subtracted from its high word so that
//     2^(m>>7) T[j] is ONE integer multiply-add, and 1 + e = fma(2^n T[j], P(r), 1) with P = e^r of degree 4 on
//     |r| <= ln2/256 (11 fp64 instructions per evaluation instead of 13; fit error 1.5e-16).  |v| beyond 2^31 ln2/64 (the index would wrap) or NaN is caught by a test on
//     the thread's sum of |v| and sends the PASS to the branch-free full-range path (fast_log_logistic).
// Everything else (neutral padding slots, running products with one logarithm per pass, sum min(v,0) from
// sum v and sum |v|, recomputed S_eta after an accept, row written back once per launch) is blr_mh_kernel's.
#pragma once

namespace vk {

static const int MH2_THREADS = 256;
static const int MH2_TILE = 4 * MH2_THREADS;   // observations per iteration: two 16-byte pairs per thread
static const int MH2_M_MIN = -128 * 1010;      // smallest exponent index: 2^-1010 is still a normal double

// exp table for this kernel: entry j = 2^(j/128) with (j << 13) subtracted from the high word, 128 rows of 8 replicas
// (the same 8 KB as the 64 x 16 table of fast_math.cuh; |r| <= ln2/256 lets e^r = 1 + r E(r) stop at degree 4)
__device__ __forceinline__ void mh2_fill_exp_table(unsigned char *exp_tbl) {
  for (int i = threadIdx.x; i < 128 * 8; i += blockDim.x) {
    const unsigned long long j = (unsigned long long)(i >> 3);
    reinterpret_cast<unsigned long long *>(exp_tbl)[i] = FM_EXP128_TABLE_DEV[i >> 3] - (j << 45);
  }
}

// 1 + exp(-|v|); the caller guarantees |v| < 2^31 ln2/64 (otherwise the pass is redone on the safe path)
__device__ __forceinline__ double mh2_one_plus_exp(double v, const FmCoef &cf, uint32_t exp_base) {
  const double MAGIC = 6755399441055744.0;                 // 1.5 * 2^52: ulp 1
  const double a = fabs(v);
  const double t = fma(a, cf.neg_scale128, MAGIC);         // low word = rint(-a 128/ln2)
  const double nf = t - MAGIC;
  const double r = fma(nf, -cf.c_ln2_128, -a);
  const int m = max(__double2loint(t), MH2_M_MIN);
  uint32_t ea, th;
  double T;
  asm("mad.lo.u32 %0, %1, 64, %2;" : "=r"(ea) : "r"((uint32_t)(m & 127)), "r"(exp_base));
  asm("ld.shared.f64 %0, [%1];" : "=d"(T) : "r"(ea));
  asm("mad.lo.u32 %0, %1, 8192, %2;" : "=r"(th) : "r"((uint32_t)m), "r"((uint32_t)__double2hiint(T)));
  const double Ts = __hiloint2double((int)th, __double2loint(T));   // 2^(m>>7) 2^((m&127)/128)
  double P = cf.E3[3];                                     // e^r = 1 + r E(r), E of degree 3
  P = fma(P, r, cf.E3[2]);
  P = fma(P, r, cf.E3[1]);
  P = fma(P, r, cf.E3[0]);
  P = fma(P, r, 1.0);
  return fma(Ts, P, 1.0);
}

// shared-memory carve-up: [log table][exp table][w D'][colsum D'][red 34][slots 2 x 3 x NW][prop_w C][prop_logu C][prop_site C][eta row Npad]
__host__ __device__ inline size_t blr_mh2_smem_base(int D) {
  return (size_t)FM_SMEM_LOG_BYTES + FM_SMEM_EXP_BYTES +
         (size_t)(2 * ((D + 1) & ~1) + 34 + 2 * 3 * (MH2_THREADS / 32) + 2 * BLR_MH_CHUNK) * sizeof(double) +
         (size_t)BLR_MH_CHUNK * sizeof(int);
}

template <int MINB>
__global__ void __launch_bounds__(MH2_THREADS, MINB)
blr_mh2_kernel(double *__restrict__ w, double *__restrict__ eta, double *__restrict__ ll,
               const double *__restrict__ Xs, BLRParams prm, int64_t P, int64_t p_base,
               uint64_t seed, uint32_t step0, int n_prop_steps, int block_mode,
               long long *__restrict__ acc_out, const FmCoef cf) {
  constexpr int NW = MH2_THREADS / 32;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned char *log_tbl = smem_raw;
  unsigned char *exp_tbl = smem_raw + FM_SMEM_LOG_BYTES;
  const int Dp = (prm.D + 1) & ~1;
  double *w_sh = reinterpret_cast<double *>(exp_tbl + FM_SMEM_EXP_BYTES);   // D
  double *colsum = w_sh + Dp;                // D: sum_i Xs[j][i]
  double *red = colsum + Dp;                 // 34
  double *slots = red + 34;                  // [parity][3][NW] per-warp partial sums of a pass
  double *prop_w = slots + 2 * 3 * NW;       // BLR_MH_CHUNK proposed values
  double *prop_logu = prop_w + BLR_MH_CHUNK; // BLR_MH_CHUNK log-uniforms
  int *prop_site = (int *)(prop_logu + BLR_MH_CHUNK);   // BLR_MH_CHUNK sites
  double *eta_sh = (double *)(prop_site + BLR_MH_CHUNK);  // Npad
  const int N = (int)prm.N;
  const int Npad = (N + MH2_TILE - 1) / MH2_TILE * MH2_TILE;
  const int iters = Npad / MH2_TILE;
  const int tid = threadIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  fm_fill_smem_tables(log_tbl, exp_tbl);
  __syncthreads();
  mh2_fill_exp_table(exp_tbl);
  const uint32_t exp_base = (uint32_t)__cvta_generic_to_shared(exp_tbl) + (uint32_t)(lane & 7) * 8u;
  const uint32_t log_base = (uint32_t)__cvta_generic_to_shared(log_tbl) + (uint32_t)(lane & 7) * 16u;
  for (int j = 0; j < prm.D; ++j) {          // column sums, once per CTA
    const double *xc = Xs + (int64_t)j * prm.Xld;
    double s_ = 0.0;
    for (int i = tid; i < N; i += MH2_THREADS) s_ += __ldg(&xc[i]);
    s_ = block_sum(s_, red);
    if (tid == 0) colsum[j] = s_;
  }
  int pass_parity = 0;
  for (int64_t p = blockIdx.x; p < P; p += gridDim.x) {
    __syncthreads();
    double *eta_g = eta + p * prm.Ncap;
    if (tid < prm.D) w_sh[tid] = w[(int64_t)tid * P + p];
    double s_part = 0.0;
    {
      // Ncap is even and rows are 16-byte aligned: 16-byte vector copies
      const double2 *src = reinterpret_cast<const double2 *>(eta_g);
      double2 *dst = reinterpret_cast<double2 *>(eta_sh);
      for (int i = tid; i < Npad / 2; i += MH2_THREADS) {
        double2 v = make_double2(BLR_PAD_ETA, BLR_PAD_ETA);
        if (2 * i < N) { v = src[i]; if (2 * i + 1 >= N) v.y = BLR_PAD_ETA; }
        dst[i] = v; s_part += v.x + v.y;
      }
    }
    double S_eta = block_sum(s_part, red);   // sum_i eta_i of the current state (padding included)
    double rho = ll[p];
    long long acc = 0;
    bool dirty = false;
    for (int c0 = 0; c0 < n_prop_steps; c0 += BLR_MH_CHUNK) {
      const int cn = n_prop_steps - c0 < BLR_MH_CHUNK ? n_prop_steps - c0 : BLR_MH_CHUNK;
      __syncthreads();
      for (int t = tid; t < cn; t += MH2_THREADS) {
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
      int t = 0;
      // feature pairs of the first tile of the pass about to run: [proposal][pair]
      double2 xn[2][2];
      bool xn_valid = false;
      while (t < cn) {
        const bool two = t + 1 < cn;
        const int j0 = prop_site[t], j1 = prop_site[two ? t + 1 : t];
        const double wn0 = prop_w[t], wn1 = prop_w[two ? t + 1 : t];
        const double d0 = wn0 - w_sh[j0], d1 = two ? wn1 - w_sh[j1] : 0.0;
        const double2 *xc0 = reinterpret_cast<const double2 *>(Xs + (int64_t)j0 * prm.Xld) + tid;
        const double2 *xc1 = reinterpret_cast<const double2 *>(Xs + (int64_t)j1 * prm.Xld) + tid;
        if (!xn_valid) {
          xn[0][0] = __ldg(xc0); xn[0][1] = __ldg(xc0 + MH2_THREADS);
          xn[1][0] = __ldg(xc1); xn[1][1] = __ldg(xc1 + MH2_THREADS);
        }
        // first tile of the pass that follows when t is REJECTED (starts at t+2): prefetched in the last iteration
        const int tn0 = t + 2 < cn ? t + 2 : t, tn1 = t + 3 < cn ? t + 3 : tn0;
        const double2 *xnx0 = reinterpret_cast<const double2 *>(Xs + (int64_t)prop_site[tn0] * prm.Xld) + tid;
        const double2 *xnx1 = reinterpret_cast<const double2 *>(Xs + (int64_t)prop_site[tn1] * prm.Xld) + tid;
        double qp[2][4], am[2][2];
#pragma unroll
        for (int s = 0; s < 2; ++s) {
#pragma unroll
          for (int c = 0; c < 4; ++c) qp[s][c] = 1.0;
          am[s][0] = 0.0; am[s][1] = 0.0;
        }
        double af[2] = {0.0, 0.0};
        long long aL[2] = {0, 0};
        int an[2] = {0, 0}, n_logs = 0;
        const double2 *ep = reinterpret_cast<const double2 *>(eta_sh) + tid;
        for (int it = 0; it < iters; ++it, ep += 2 * MH2_THREADS) {
          const double2 e0 = ep[0], e1 = ep[MH2_THREADS];
          double v[2][4];
          v[0][0] = fma(d0, xn[0][0].x, e0.x); v[0][1] = fma(d0, xn[0][0].y, e0.y);
          v[0][2] = fma(d0, xn[0][1].x, e1.x); v[0][3] = fma(d0, xn[0][1].y, e1.y);
          v[1][0] = fma(d1, xn[1][0].x, e0.x); v[1][1] = fma(d1, xn[1][0].y, e0.y);
          v[1][2] = fma(d1, xn[1][1].x, e1.x); v[1][3] = fma(d1, xn[1][1].y, e1.y);
          {
            const bool more = it + 1 < iters;
            const double2 *s0 = more ? xc0 + (it + 1) * 2 * MH2_THREADS : xnx0;
            const double2 *s1 = more ? xc1 + (it + 1) * 2 * MH2_THREADS : xnx1;
            xn[0][0] = __ldg(s0); xn[0][1] = __ldg(s0 + MH2_THREADS);
            xn[1][0] = __ldg(s1); xn[1][1] = __ldg(s1 + MH2_THREADS);
          }
#pragma unroll
          for (int s = 0; s < 2; ++s) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              qp[s][c] *= mh2_one_plus_exp(v[s][c], cf, exp_base);
              am[s][c & 1] += fabs(v[s][c]);
            }
          }
          if ((it & 63) == 63) {     // a product of 64 factors in [1,2] stays below 2^64
#pragma unroll
            for (int s = 0; s < 2; ++s) {
              fm_log_product((qp[s][0] * qp[s][1]) * (qp[s][2] * qp[s][3]), cf, log_base, af[s], aL[s], an[s]);
#pragma unroll
              for (int c = 0; c < 4; ++c) qp[s][c] = 1.0;
            }
            ++n_logs;
          }
        }
        if ((iters & 63) != 0) {
#pragma unroll
          for (int s = 0; s < 2; ++s)
            fm_log_product((qp[s][0] * qp[s][1]) * (qp[s][2] * qp[s][3]), cf, log_base, af[s], aL[s], an[s]);
          ++n_logs;
        }
        double part[2];
        bool unsafe = false;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const double a_ = am[s][0] + am[s][1];
          unsafe = unsafe || !(a_ < 1.0e6);      // also NaN: |v| may have wrapped the exponent index
          part[s] = fma(-0.5, a_, -af[s]) - (double)aL[s] * 2.220446049250313e-16;
          part[s] -= (double)(an[s] - 1023 * n_logs) * 0.6931471805599453;
        }
        if (tid == 0) { part[0] += 0.5 * fma(d0, colsum[j0], S_eta); part[1] += 0.5 * fma(d1, colsum[j1], S_eta); }
        // block-wide sums of both proposals and of the "unsafe" flag with ONE barrier: per-warp values go to
        // the parity-selected slot array, every thread adds the NW values in the same fixed order
        double xi[2], flag;
        {
          const double p0 = warp_sum(part[0]), p1 = warp_sum(part[1]);
          const double f_ = __any_sync(0xffffffffu, unsafe) ? 1.0 : 0.0;
          double *b = slots + pass_parity * 3 * NW;
          if (lane == 0) { b[warp] = p0; b[NW + warp] = p1; b[2 * NW + warp] = f_; }
          __syncthreads();
          xi[0] = b[0]; xi[1] = b[NW]; flag = b[2 * NW];
#pragma unroll
          for (int k = 1; k < NW; ++k) { xi[0] += b[k]; xi[1] += b[NW + k]; flag += b[2 * NW + k]; }
          pass_parity ^= 1;
        }
        if (flag != 0.0) {
          // full-range path for this pass (rare: |logit| >= 1e6 or NaN somewhere in the row)
          double q0 = 0.0, q1 = 0.0;
          const double *x0 = Xs + (int64_t)j0 * prm.Xld, *x1 = Xs + (int64_t)j1 * prm.Xld;
          for (int k = tid; k < N; k += MH2_THREADS) {
            q0 += fast_log_logistic(fma(d0, __ldg(&x0[k]), eta_sh[k]));
            q1 += fast_log_logistic(fma(d1, __ldg(&x1[k]), eta_sh[k]));
          }
          xi[0] = block_sum(q0, red);
          xi[1] = block_sum(q1, red);
        }
        // decide t; if it is rejected, decide t+1 from the speculative xi (exact: the state did not change)
        int which = -1;
        if (prop_logu[t] < blr_log_alpha(rho, xi[0])) which = 0;
        else if (two && prop_logu[t + 1] < blr_log_alpha(rho, xi[1])) which = 1;
        xn_valid = which != 0;        // the prefetched tile belongs to the pass starting at t+2
        if (which >= 0) {             // uniform across the CTA
          const int j = which ? j1 : j0;
          const double dl = which ? d1 : d0;
          const double *__restrict__ xcol = Xs + (int64_t)j * prm.Xld;
          double sp = 0.0;
          for (int k = tid; k < Npad; k += MH2_THREADS) { double e_ = fma(dl, __ldg(&xcol[k]), eta_sh[k]); eta_sh[k] = e_; sp += e_; }
          if (tid == 0) w_sh[j] = which ? wn1 : wn0;
          rho = xi[which]; ++acc; dirty = true;
          S_eta = block_sum(sp, red);   // recomputed, not updated: no drift
        }
        t += which == 0 ? 1 : (two ? 2 : 1);
      }
    }
    if (dirty) {
      const double2 *src = reinterpret_cast<const double2 *>(eta_sh);
      double2 *dst = reinterpret_cast<double2 *>(eta_g);
      for (int i = tid; i < (N + 1) / 2; i += MH2_THREADS) dst[i] = src[i];
      if (tid < prm.D) w[(int64_t)tid * P + p] = w_sh[tid];
      if (tid == 0) ll[p] = rho;
    }
    if (acc_out && tid == 0) acc_out[p] = acc;
  }
}

}  // namespace vk
