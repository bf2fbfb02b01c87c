// Config C4: LDA with collapsed symmetric-Dirichlet/categorical makers
// (examples/lda.vnt:18-41):
//   alpha_document_topic ~ gamma(a,b); alpha_topic_word ~ gamma(a,b)
//   doc(d)   = make_sym_dir_cat(alpha_document_topic, K)
//   topic(k) = make_sym_dir_cat(alpha_topic_word, V)
//   word(d,pos) = ((topic ((doc d))));   observe word(d,pos) = w
// Site programs (SURVEY App. A "C4"), default scope, blocks in creation order:
//   block 0: alpha_document_topic, block 1: alpha_topic_word  -- AAA sites:
//     rho/xi = sum over made SPs of CDirCatOutputPSP.logDensityOfData
//     (lkernel.py:158-171, dirichlet.py:192-201); proposal from the gamma prior
//   block 2+i: z_i = ((doc d_i)) of token i:
//     detach: unconstrain the token, rho = log((n_{z,w}^- + aw)/(n_z^- + V aw))
//             (detach.py:52-65, dirichlet.py:168-173), unincorporate token and z
//     regen : z' ~ (n_{d,k}^- + ad) (dirichlet.py:164-166), re-constrain the token,
//             xi = log((n_{z',w}^- + aw)/(n_{z'}^- + V aw))   (regen.py:68-88)
//
// State: ONE ROW PER PARTICLE in HBM (particle-major, 16-byte aligned)
//   [alpha_d, alpha_w : 2 x f64][n_k : K x u32][n_dk : n_docs*K x u16]
//   [n_kw : K*V x u16][z : max_tokens x u8]
// Kernel mapping: one warp (= one CTA) per particle.  The whole row is staged
// in shared memory (38 KB at C4: 5-6 resident particles per SM), the launch's
// transitions run on chip -- the chain of one particle is inherently
// sequential, lanes parallelise the K-way topic scan and the lgamma sums of
// the alpha sites -- and the row is written back once.  Latency-bound, not
// HBM-bound; see DESIGN.md.
#include "vk_common.cuh"
#include "sp_math.cuh"
#include <stdlib.h>

namespace vk {

// n_kw is stored WORD-major: cell (topic k, word w) at nkw[w * Kp + k], Kp = K rounded up to 16, so the K
// candidate cells a z site needs for its word are one contiguous run (a single 32-byte sector at K = 16)
// instead of K sectors a row apart; padding cells stay zero.
struct LDALayout {
  int K, Kp, V, n_docs, max_tokens;
  size_t off_nk, off_ndk, off_nkw, off_z, row_bytes;
  double hyper_a, hyper_b;
};

struct LDARow {
  double *alpha; uint32_t *nk; uint16_t *ndk; uint16_t *nkw; uint8_t *z;
  __device__ LDARow(unsigned char *base, const LDALayout &L)
      : alpha((double *)base), nk((uint32_t *)(base + L.off_nk)), ndk((uint16_t *)(base + L.off_ndk)),
        nkw((uint16_t *)(base + L.off_nkw)), z(base + L.off_z) {}
  // small sections (alpha, n_k, n_dk) and z staged at `small` / `zs`, the K x V table left in global memory
  __device__ LDARow(unsigned char *small, uint16_t *nkw_global, unsigned char *zs, const LDALayout &L)
      : alpha((double *)small), nk((uint32_t *)(small + L.off_nk)), ndk((uint16_t *)(small + L.off_ndk)),
        nkw(nkw_global), z(zs) {}
};

__device__ __forceinline__ int warp_scan_incl_i(int v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// z ~ (counts_k + alpha): first k with u*(N + K alpha) < C_k + (k+1) alpha, C = exact
// integer prefix counts (dirichlet.py:164-166; order-independent arithmetic).
__device__ __forceinline__ int lda_sample_topic(const uint16_t *row, int K, double alpha, double u, int lane) {
  int tot = 0;
  for (int b = 0; b < K; b += 32) { int k = b + lane; tot += (k < K) ? (int)row[k] : 0; }
  tot = warp_sum_i(tot);
  double t = u * ((double)tot + (double)K * alpha);
  int carry = 0;
  for (int b = 0; b < K; b += 32) {
    int k = b + lane;
    int c = (k < K) ? (int)row[k] : 0;
    int incl = warp_scan_incl_i(c, lane) + carry;
    double cum = (double)incl + (double)(k + 1) * alpha;
    unsigned m = __ballot_sync(0xffffffffu, (k < K) && (t < cum));
    if (m) return b + __ffs(m) - 1;
    carry = __shfl_sync(0xffffffffu, incl, 31);
  }
  return K - 1;
}

// same draw with this lane's count c_k already in a register (K <= 32): identical arithmetic
__device__ __forceinline__ int lda_sample_topic_reg(int c, int K, double alpha, double u, int lane) {
  int tot = warp_sum_i(c);
  double t = u * ((double)tot + (double)K * alpha);
  int incl = warp_scan_incl_i(c, lane);
  double cum = (double)incl + (double)(lane + 1) * alpha;
  unsigned m = __ballot_sync(0xffffffffu, (lane < K) && (t < cum));
  return m ? __ffs(m) - 1 : K - 1;
}

// sum over tables of logDensityOfData(counts; alpha): n_tab tables of n_cell cells
// (dirichlet.py:192-201).  totals[t] = N of table t (nullptr: derive from `fixed_tot`).
// `cells` = number of stored cells to scan (>= n_tab * n_cell when the table is padded; padding is zero).
__device__ double lda_aaa_weight(const uint16_t *counts, int n_tab, int n_cell, const uint32_t *totals,
                                 const int32_t *fixed_tot, double alpha, int lane, int64_t cells) {
  double A = alpha * (double)n_cell;
  double lgA = lgamma(A), lga = lgamma(alpha);
  double acc = 0.0;
  for (int t = lane; t < n_tab; t += 32) {
    double N = totals ? (double)totals[t] : (double)fixed_tot[t];
    acc += lgA - lgamma(N + A);
  }
  for (int64_t i = lane; i < cells; i += 32) {
    int c = counts[i];
    if (c) acc += lgamma(alpha + (double)c) - lga;
  }
  return warp_sum(acc);
}

// mode 0: make_consistent over tokens [t0,t1): z ~ prior predictive, weight by the word, incorporate
// mode 1: likelihood_weight: alphas ~ prior, counts cleared, tokens [0,t1) as mode 0, lw := sum
// mode 2: reset_to_prior: as 1 but weights untouched
// mode 3: refresh: recount from z (after write_site)
__global__ void __launch_bounds__(32)
lda_seq_kernel(unsigned char *__restrict__ rows, LDALayout L, const int32_t *__restrict__ doc,
               const int32_t *__restrict__ word, int t0, int t1, double *__restrict__ lw,
               double *__restrict__ incr, int64_t P, int64_t p_base, uint64_t seed, uint32_t step, int mode) {
  const int lane = threadIdx.x;
  for (int64_t p = blockIdx.x; p < P; p += gridDim.x) {
    LDARow R(rows + p * L.row_bytes, L);
    Stream r(seed, (uint64_t)(p_base + p), step);
    if (mode == 1 || mode == 2) {
      double ad = gamma_sim(r, L.hyper_a, L.hyper_b);
      double aw = gamma_sim(r, L.hyper_a, L.hyper_b);
      if (lane == 0) { R.alpha[0] = ad; R.alpha[1] = aw; }
    }
    if (mode >= 1) {
      for (int i = lane; i < L.K; i += 32) R.nk[i] = 0;
      for (int i = lane; i < L.n_docs * L.K; i += 32) R.ndk[i] = 0;
      for (int64_t i = lane; i < (int64_t)L.Kp * L.V; i += 32) R.nkw[i] = 0;
    }
    __syncwarp();
    double ad = R.alpha[0], aw = R.alpha[1];
    double wsum = 0.0;
    for (int i = t0; i < t1; ++i) {
      int d = doc[i], w = word[i];
      int zi;
      if (mode == 3) zi = R.z[i];
      else zi = lda_sample_topic(R.ndk + d * L.K, L.K, ad, r.uniform(), lane);
      if (lane == 0) {
        wsum += dircat_logpmf((double)R.nkw[(int64_t)w * L.Kp + zi], aw, (double)R.nk[zi], aw * (double)L.V);
        R.z[i] = (uint8_t)zi;
        R.ndk[d * L.K + zi] += 1; R.nkw[(int64_t)w * L.Kp + zi] += 1; R.nk[zi] += 1;
      }
      __syncwarp();
    }
    if (lane == 0) {
      if (mode == 0) {
        if (wsum != wsum) wsum = -INFINITY;
        if (incr) incr[p] = wsum;
        lw[p] += wsum;
      } else if (mode == 1) lw[p] = wsum;
    }
  }
}

// K resimulation-MH transitions per particle on the staged row.
// stage: 0 = whole row in global memory, 1 = whole row in shared memory (one warp per CTA),
// 2 = alpha / n_k / n_dk in shared memory, the K x V topic-word table and z in global memory
// (their accesses of one z transition are issued together): ~2 KB per particle instead of
// 38 KB at C4, so all particles are resident at once and the serial per-particle chains
// overlap each other's latency.
static const int LDA_WARPS = 4;
// Enumerative-Gibbs selection across a warp: lane k < K holds pr_k; Kogge-Stone inclusive scan (the same
// additions in the same order as sp_math.cuh gibbs_pick_small / oracle.c gibbs_pick), then the first lane
// whose running sum exceeds u * total.
__device__ __forceinline__ int lda_gibbs_pick_warp(double pr, int K, double u, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    double t = __shfl_up_sync(0xffffffffu, pr, o);
    if (lane >= o) pr += t;
  }
  const double total = __shfl_sync(0xffffffffu, pr, K - 1);
  const double thr = u * total;
  const unsigned m = __ballot_sync(0xffffffffu, lane < K && thr < pr);
  return m ? __ffs(m) - 1 : K - 1;
}

// 7 CTAs of 4 warps per SM (<= 73 registers): C4's 4096 particles = 1024 CTAs fit the 148 SMs in ONE wave
// (7 * 148 = 1036); the chain of a particle is serial, so resident warps are the only latency hiding there is.
// The rarely taken hyper-parameter branch (lgamma sums) absorbs the spills.
__global__ void __launch_bounds__(32 * LDA_WARPS, 7)
lda_mh_kernel(unsigned char *__restrict__ rows, LDALayout L, const int32_t *__restrict__ doc,
              const int32_t *__restrict__ word, const int32_t *__restrict__ doclen, int n_tok,
              int64_t P, int64_t p_base, uint64_t seed, uint32_t step0, int n_steps, int block_mode,
              int stage, int tok_shared, long long *__restrict__ acc_out, int n_docs_live,
              double *__restrict__ nodes_out, int op, int nb_sel, int b_off) {
  extern __shared__ uint4 smem_all[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, n_w = blockDim.x >> 5;
  // op 0: resimulation MH over the default scope (nb_sel = 2 + n_tok blocks, b_off = 0).
  // op 1: enumerative Gibbs (infer/egibbs.py:42-107) over the z sites (nb_sel = n_tok, b_off = 2): every
  //       topic k is weighted exp(xi_k) = P(z = k | doc counts) P(word | topic k counts), the token's own
  //       counts removed; the new topic is drawn proportionally and always accepted (logAlpha = 0).
  const int nb = nb_sel;
  const size_t warp_bytes = stage == 1 ? L.row_bytes : L.off_nkw;   // stage 2: alpha, n_k, n_dk only
  uint4 *smem_u4 = smem_all + (size_t)wid * (warp_bytes / 16);
  // CTA-shared packed (doc << 16 | word) copy of the corpus, after the per-warp sections
  uint32_t *tok_sh = nullptr;
  if (tok_shared) {
    tok_sh = reinterpret_cast<uint32_t *>(smem_all + (size_t)n_w * (warp_bytes / 16));
    for (int i = threadIdx.x; i < n_tok; i += blockDim.x) tok_sh[i] = ((uint32_t)doc[i] << 16) | (uint32_t)word[i];
    __syncthreads();
  }
  for (int64_t p = (int64_t)blockIdx.x * n_w + wid; p < P; p += (int64_t)gridDim.x * n_w) {
    unsigned char *grow = rows + p * L.row_bytes;
    unsigned char *base = grow;
    if (stage == 1) {
      const uint4 *src = (const uint4 *)grow;
      for (size_t i = lane; i < L.row_bytes / 16; i += 32) smem_u4[i] = src[i];
      base = (unsigned char *)smem_u4;
      __syncwarp();
    } else if (stage == 2) {
      const uint4 *src = (const uint4 *)grow;
      for (size_t i = lane; i < L.off_nkw / 16; i += 32) smem_u4[i] = src[i];
      __syncwarp();
    }
    LDARow R = stage == 2 ? LDARow((unsigned char *)smem_u4, (uint16_t *)(grow + L.off_nkw), grow + L.off_z, L)
                          : LDARow(base, L);
    long long acc = 0;
    // transloop's return value (dispatch.py:109-128): |regenCounts| per transition, as measured on the
    // reference's Lite backend (tests/golden/lite_chains.json c4): z site 7; alpha_document_topic
    // 1 + 2 * (#doc makers alive); alpha_topic_word 1 + 2 * (#topic makers alive = topics in use).
    // z sites all count 7, so only the (rare) hyper-parameter transitions are tallied as they happen; the z
    // transitions are added arithmetically at the end -- no per-transition work, no division in the loop.
    long long alpha_nodes = 0, sweep_alpha = 0, sweep_tot = 0;
    int n_alpha = 0, sweep_n_alpha = 0;
    int kmod = 0;                       // k mod nb, kept by increment-and-wrap
    // The three uniforms every transition starts with (block choice, accept, topic selection) are the same
    // for all 32 lanes of the particle's warp; computing Philox redundantly on every lane was most of the
    // kernel's instructions.  Instead lane l draws them for transition k + l once per 32 transitions and each
    // transition broadcasts its own with shuffles.  (Hyper-parameter sites draw their gamma proposal from
    // the same stream, from draw 2, on the rare transitions that need it.)
    double us_l = 0.0, ua_l = 0.0, uz_l = 0.0;
    for (int k = 0; k < n_steps; ++k) {
      const int jb = k & 31;
      if (jb == 0 && k + lane < n_steps) {
        Stream rl(seed, (uint64_t)(p_base + p), step0 + (uint32_t)(k + lane));
        us_l = rl.uniform(); ua_l = rl.uniform(); uz_l = rl.uniform();
      }
      const double u_sel = __shfl_sync(0xffffffffu, us_l, jb);
      const double u_acc = __shfl_sync(0xffffffffu, ua_l, jb);
      const double u_z = __shfl_sync(0xffffffffu, uz_l, jb);
      const int km = kmod;
      kmod = (km + 1 == nb) ? 0 : km + 1;
      int b;
      if (block_mode == VK_BLOCK_ONE) { b = (int)(u_sel * (double)nb); if (b >= nb) b = nb - 1; b += b_off; }
      else if (block_mode == VK_BLOCK_EACH) b = km + b_off;
      else b = block_mode;
      if (nodes_out) {
        if (b < 2) {
          int cnt;
          if (b == 0) cnt = 1 + 2 * n_docs_live;
          else {
            int live = 0;
            for (int t = lane; t < L.K; t += 32) live += R.nk[t] > 0;
            for (int o = 16; o; o >>= 1) live += __shfl_xor_sync(0xffffffffu, live, o);
            cnt = 1 + 2 * live;
          }
          alpha_nodes += cnt; ++n_alpha; sweep_alpha += cnt; ++sweep_n_alpha;
        }
        if (block_mode == VK_BLOCK_EACH && km == nb - 1) {     // `each`: floor(sum of the sweep / #blocks)
          sweep_tot += (sweep_alpha + 7ll * (nb - sweep_n_alpha)) / nb;
          sweep_alpha = 0; sweep_n_alpha = 0;
        }
      }
      if (b < 2) {
        // AAA site: alpha' from the gamma prior (no proposal weight), rho/xi = marginal
        // likelihood of every made SP's counts under the old/new concentration
        double a_old = R.alpha[b];
        Stream r(seed, (uint64_t)(p_base + p), step0 + (uint32_t)k);
        r.seek(2);
        double a_new = gamma_sim(r, L.hyper_a, L.hyper_b);
        double rho, xi;
        if (b == 0) {
          rho = lda_aaa_weight(R.ndk, L.n_docs, L.K, nullptr, doclen, a_old, lane, (int64_t)L.n_docs * L.K);
          xi = lda_aaa_weight(R.ndk, L.n_docs, L.K, nullptr, doclen, a_new, lane, (int64_t)L.n_docs * L.K);
        } else {
          rho = lda_aaa_weight(R.nkw, L.K, L.V, R.nk, nullptr, a_old, lane, (int64_t)L.Kp * L.V);
          xi = lda_aaa_weight(R.nkw, L.K, L.V, R.nk, nullptr, a_new, lane, (int64_t)L.Kp * L.V);
        }
        double alpha = (rho == -INFINITY && xi == -INFINITY) ? INFINITY
                       : (rho == INFINITY && xi == INFINITY) ? -INFINITY : xi - rho;
        if (log(u_acc) < alpha) { if (lane == 0) R.alpha[b] = a_new; ++acc; }
        __syncwarp();
      } else if (tok_sh != nullptr && L.K <= 32) {
        // z site, fast path: (doc, word) come from the CTA's shared copy, and every lane k < K loads
        // ITS candidate cell n_kw[k][w] up front -- one overlapped global round trip instead of two
        // dependent ones (old cell, then proposed cell).  Nothing is written unless the proposal is
        // accepted AND changes the topic: "remove the token, then add it back" is done on register
        // copies of the counts (detach.py:52-65 / regen.py:68-88 leave the tables unchanged on reject).
        const int i = b - 2;
        const uint32_t dw = tok_sh[i];
        const int d = (int)(dw >> 16), w = (int)(dw & 0xffffu);
        const int zo = R.z[i];
        const int kk = lane < L.K ? lane : 0;
        int cand = (int)R.nkw[(int64_t)w * L.Kp + kk];                  // global; one cell per lane
        int nk_k = (int)R.nk[kk];
        int cdk = lane < L.K ? (int)R.ndk[d * L.K + lane] : 0;
        if (lane == zo) { cand -= 1; nk_k -= 1; cdk -= 1; }          // counts with this token removed
        const double ad = R.alpha[0], aw = R.alpha[1];
        const double Aw = aw * (double)L.V;
        if (op == 1) {
          // lane k: exp(xi_k) up to the z term's common denominator (dirichlet.py:168-173, twice)
          double pr = lane < L.K ? ((double)cdk + ad) * ((double)cand + aw) / ((double)nk_k + Aw) : 0.0;
          const int zg = lda_gibbs_pick_warp(pr, L.K, u_z, lane);
          if (zg != zo) {
            ++acc;
            const int c_o = __shfl_sync(0xffffffffu, cand, zo), c_g = __shfl_sync(0xffffffffu, cand, zg);
            if (lane == 0) {
              R.z[i] = (uint8_t)zg;
              R.nkw[(int64_t)w * L.Kp + zo] = (uint16_t)c_o;
              R.nkw[(int64_t)w * L.Kp + zg] = (uint16_t)(c_g + 1);
              R.nk[zo] -= 1; R.nk[zg] += 1;
              R.ndk[d * L.K + zo] -= 1; R.ndk[d * L.K + zg] += 1;
            }
            __syncwarp();
          }
          continue;
        }
        const int zn = lda_sample_topic_reg(cdk, L.K, ad, u_z, lane);
        const int c_o = __shfl_sync(0xffffffffu, cand, zo), n_o = __shfl_sync(0xffffffffu, nk_k, zo);
        const int c_n = __shfl_sync(0xffffffffu, cand, zn), n_n = __shfl_sync(0xffffffffu, nk_k, zn);
        const double a_o = (double)c_o + aw, b_o = (double)n_o + Aw;   // rho = log(a_o / b_o)
        const double a_n = (double)c_n + aw, b_n = (double)n_n + Aw;   // xi  = log(a_n / b_n)
        const bool accept = u_acc * b_n * a_o < a_n * b_o;             // log u < xi - rho, see below
        if (accept) ++acc;
        if (accept && zn != zo) {
          if (lane == 0) {
            R.z[i] = (uint8_t)zn;
            R.nkw[(int64_t)w * L.Kp + zo] = (uint16_t)c_o;              // c_o is already "minus this token"
            R.nkw[(int64_t)w * L.Kp + zn] = (uint16_t)(c_n + 1);
            R.nk[zo] -= 1; R.nk[zn] += 1;
            R.ndk[d * L.K + zo] -= 1; R.ndk[d * L.K + zn] += 1;
          }
          __syncwarp();
        }
      } else {
        int i = b - 2;
        int d = doc[i], w = word[i];
        int zo = R.z[i];
        double ad = R.alpha[0], aw = R.alpha[1];
        double Aw = aw * (double)L.V;
        int64_t co = (int64_t)w * L.Kp + zo;
        // counts with this token (and its z) removed: rho = log(a_o / b_o)   (dirichlet.py:168-173)
        double a_o = (double)((int)R.nkw[co] - 1) + aw, b_o = (double)((int)R.nk[zo] - 1) + Aw;
        __syncwarp();
        if (lane == 0) { R.nkw[co] -= 1; R.nk[zo] -= 1; R.ndk[d * L.K + zo] -= 1; }
        __syncwarp();
        if (op == 1) {
          int zg;
          const double u_g = u_z;
          if (L.K <= 32) {
            const int kk = lane < L.K ? lane : 0;
            double pr = lane < L.K ? ((double)R.ndk[d * L.K + kk] + ad) * ((double)R.nkw[(int64_t)w * L.Kp + kk] + aw) /
                                         ((double)R.nk[kk] + Aw) : 0.0;
            zg = lda_gibbs_pick_warp(pr, L.K, u_g, lane);
          } else {
            // sequential inclusive sums (oracle.c gibbs_pick, n > 32): total first, then the first crossing
            double total = 0.0;
            for (int t = 0; t < L.K; ++t)
              total += ((double)R.ndk[d * L.K + t] + ad) * ((double)R.nkw[(int64_t)w * L.Kp + t] + aw) / ((double)R.nk[t] + Aw);
            const double thr = u_g * total;
            double cum = 0.0;
            zg = L.K - 1;
            for (int t = 0; t < L.K; ++t) {
              cum += ((double)R.ndk[d * L.K + t] + ad) * ((double)R.nkw[(int64_t)w * L.Kp + t] + aw) / ((double)R.nk[t] + Aw);
              if (thr < cum) { zg = t; break; }
            }
          }
          if (zg != zo) ++acc;
          __syncwarp();
          if (lane == 0) {
            R.z[i] = (uint8_t)zg;
            R.nkw[(int64_t)w * L.Kp + zg] += 1; R.nk[zg] += 1; R.ndk[d * L.K + zg] += 1;
          }
          __syncwarp();
          continue;
        }
        int zn = lda_sample_topic(R.ndk + d * L.K, L.K, ad, u_z, lane);
        int64_t cn = (int64_t)w * L.Kp + zn;
        double a_n = (double)R.nkw[cn] + aw, b_n = (double)R.nk[zn] + Aw;   // xi = log(a_n / b_n)
        // accept iff log u < xi - rho (mh.py:84-89), evaluated without logarithms or divisions as
        // u * b_n * a_o < a_n * b_o: all four factors are finite and positive here, so the +-inf
        // policy of mh.py:190-208 cannot trigger; every lane of the warp redundantly executes this
        // chain, and three fp64 logs + two divisions per transition saturated the fp64 pipe.
        bool accept = u_acc * b_n * a_o < a_n * b_o;
        int zf = accept ? zn : zo;
        if (accept) ++acc;
        __syncwarp();
        if (lane == 0) {
          R.z[i] = (uint8_t)zf;
          R.nkw[(int64_t)w * L.Kp + zf] += 1; R.nk[zf] += 1; R.ndk[d * L.K + zf] += 1;
        }
        __syncwarp();
      }
    }
    if (stage == 1) {
      uint4 *dst = (uint4 *)grow;
      for (size_t i = lane; i < L.row_bytes / 16; i += 32) dst[i] = smem_u4[i];
      __syncwarp();
    } else if (stage == 2) {
      uint4 *dst = (uint4 *)grow;
      for (size_t i = lane; i < L.off_nkw / 16; i += 32) dst[i] = smem_u4[i];
      __syncwarp();
    }
    if (acc_out && lane == 0) acc_out[p] = acc;
    if (nodes_out && lane == 0)
      nodes_out[p] = block_mode == VK_BLOCK_EACH ? (double)sweep_tot / (double)(n_steps / nb)
                                                 : (double)(alpha_nodes + 7ll * (n_steps - n_alpha)) / (double)n_steps;
  }
}

// log_likelihood_at(default, all): sum over tokens of the sequential predictive
// is NOT what the reference computes; it sums each token's logDensity given all
// other counts removed one at a time.  Compiled here for block = a z site or an
// alpha site (the scaffold's border weights).
__global__ void __launch_bounds__(32)
lda_loglik_site_kernel(const unsigned char *__restrict__ rows, LDALayout L, const int32_t *__restrict__ doc,
                       const int32_t *__restrict__ word, const int32_t *__restrict__ doclen, int block,
                       int joint, int64_t P, double *__restrict__ out) {
  const int lane = threadIdx.x;
  for (int64_t p = blockIdx.x; p < P; p += gridDim.x) {
    LDARow R(const_cast<unsigned char *>(rows) + p * L.row_bytes, L);
    double v;
    if (block == 0) {
      v = lda_aaa_weight(R.ndk, L.n_docs, L.K, nullptr, doclen, R.alpha[0], lane, (int64_t)L.n_docs * L.K);
      if (joint) v += gamma_logpdf(R.alpha[0], L.hyper_a, L.hyper_b);
    } else if (block == 1) {
      v = lda_aaa_weight(R.nkw, L.K, L.V, R.nk, nullptr, R.alpha[1], lane, (int64_t)L.Kp * L.V);
      if (joint) v += gamma_logpdf(R.alpha[1], L.hyper_a, L.hyper_b);
    } else {
      int i = block - 2, d = doc[i], w = word[i], zo = R.z[i];
      double aw = R.alpha[1];
      v = dircat_logpmf((double)((int)R.nkw[(int64_t)w * L.Kp + zo] - 1), aw, (double)((int)R.nk[zo] - 1),
                        aw * (double)L.V);
      if (joint) {
        int nd = 0;
        for (int k = 0; k < L.K; ++k) nd += R.ndk[d * L.K + k];
        v += dircat_logpmf((double)((int)R.ndk[d * L.K + zo] - 1), R.alpha[0], (double)(nd - 1),
                           R.alpha[0] * (double)L.K);
      }
    }
    if (lane == 0) out[p] = v;
  }
}

__global__ void lda_read_kernel(const unsigned char *__restrict__ rows, LDALayout L, int64_t site, int64_t P,
                                double *out) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    const unsigned char *b = rows + p * L.row_bytes;
    out[p] = site < 2 ? ((const double *)b)[site] : (double)b[L.off_z + (site - 2)];
  }
}
__global__ void lda_write_kernel(unsigned char *__restrict__ rows, LDALayout L, int64_t site, int64_t P,
                                 const double *in) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (int64_t)gridDim.x * blockDim.x) {
    unsigned char *b = rows + p * L.row_bytes;
    if (site < 2) ((double *)b)[site] = in[p];
    else b[L.off_z + (site - 2)] = (uint8_t)in[p];
  }
}
__global__ void lda_fill_kernel(double *out, int64_t n, double v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    out[i] = v;
}

struct LDAModel : Model {
  LDALayout L;
  int n_tok = 0, n_prop = 0;
  std::vector<int32_t> doc_host, word_host, doclen_host;
  DevBuf<int32_t> doc, word, doclen;
  DevBuf<unsigned char> rowsa, rowsb;
  unsigned char *rows = nullptr, *rows_alt = nullptr;
  int smem_limit = 0;

  LDAModel(vk_particles *h_, int K, int V, int n_docs, int max_tokens, double a, double b) : Model(h_) {
    VK_REQUIRE(K >= 1 && K <= 255, VK_EINVAL, "lda: n_topics must be in [1,255]");
    VK_REQUIRE(V >= 1 && n_docs >= 1, VK_EINVAL, "lda: vocab_size and n_docs must be positive");
    VK_REQUIRE(max_tokens >= 1 && max_tokens <= 65535, VK_EINVAL, "lda: max_tokens must be in [1,65535] (u16 counts)");
    VK_REQUIRE(a > 0 && b > 0, VK_EINVAL, "gamma: shape and rate must be positive");
    L.K = K; L.Kp = (K + 15) & ~15; L.V = V; L.n_docs = n_docs; L.max_tokens = max_tokens; L.hyper_a = a; L.hyper_b = b;
    auto up16 = [](size_t v) { return (v + 15) & ~(size_t)15; };   // every section starts 16-byte aligned
    size_t o = 16;
    L.off_nk = o; o = up16(o + (size_t)K * 4);
    L.off_ndk = o; o = up16(o + (size_t)n_docs * K * 2);
    L.off_nkw = o; o = up16(o + (size_t)L.Kp * V * 2);
    L.off_z = o; o = up16(o + (size_t)max_tokens);
    L.row_bytes = o;
    doc.alloc((size_t)max_tokens); word.alloc((size_t)max_tokens); doclen.alloc((size_t)n_docs);
    doclen_host.assign((size_t)n_docs, 0);
    rowsa.alloc((size_t)h->P * L.row_bytes);
    rows = rowsa.p;
    VK_CUDA(cudaMemsetAsync(rows, 0, rowsa.bytes(), h->stream));
    VK_CUDA(cudaDeviceGetAttribute(&smem_limit, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
    VK_CUDA(cudaFuncSetAttribute(lda_mh_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_limit));
  }
  int grid_p(int per_sm) const {
    int64_t cap = (int64_t)h->sm_count * per_sm;
    return (int)(h->P < cap ? h->P : cap);
  }
  void observe(const int32_t *d, const int32_t *w, int64_t n) {
    VK_REQUIRE(n_tok + n <= L.max_tokens, VK_EINVAL, "lda: token capacity exceeded");
    for (int64_t i = 0; i < n; ++i) {
      VK_REQUIRE(d[i] >= 0 && d[i] < L.n_docs && w[i] >= 0 && w[i] < L.V, VK_EINVAL, "lda: doc/word id out of range");
    }
    for (int64_t i = 0; i < n; ++i) { doc_host.push_back(d[i]); word_host.push_back(w[i]); doclen_host[d[i]]++; }
    VK_CUDA(cudaMemcpyAsync(doc.p + n_tok, d, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaMemcpyAsync(word.p + n_tok, w, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaMemcpyAsync(doclen.p, doclen_host.data(), doclen_host.size() * 4, cudaMemcpyHostToDevice, h->stream));
    VK_CUDA(cudaStreamSynchronize(h->stream));
    n_tok += (int)n;
  }
  void seq(int mode, int t0, int t1, double *incr_dev, uint32_t step) {
    lda_seq_kernel<<<grid_p(16), 32, 0, h->stream>>>(rows, L, doc.p, word.p, t0, t1, h->lw.p, incr_dev, h->P,
                                                     h->p_base, h->seed, step, mode);
    VK_CUDA(cudaGetLastError());
  }
  void make_consistent(double *incr_dev) override {
    if (n_prop == n_tok) {
      if (incr_dev) VK_CUDA(cudaMemsetAsync(incr_dev, 0, (size_t)h->P * sizeof(double), h->stream));
      return;
    }
    seq(0, n_prop, n_tok, incr_dev, h->next_step());
    n_prop = n_tok;
  }
  void likelihood_weight() override { n_prop = n_tok; seq(1, 0, n_tok, nullptr, h->next_step()); }
  void reset_to_prior() override { seq(2, 0, n_prop, nullptr, h->next_step()); }
  void refresh() override { seq(3, 0, n_prop, nullptr, 0); }
  void mh(int scope, int block, int K, double *mean_nodes_dev, long long *acc_dev) override {
    VK_REQUIRE(scope == 0, VK_EINVAL, "lda: resimulation_mh is compiled for the default scope");
    int nb = 2 + n_tok;
    VK_REQUIRE(block == VK_BLOCK_ONE || block == VK_BLOCK_EACH || (block >= 0 && block < nb), VK_EINVAL, "lda: bad block");
    run_sites(0, block, nb, 0, K, mean_nodes_dev, acc_dev);
  }
  // Enumerative Gibbs (infer/egibbs.py).  Scope 0 = default: its blocks include the two gamma sites, which
  // PSP.enumerateValues cannot enumerate (psp.py:291-296 raises) -- so `one` / `each` / blocks 0,1 are
  // errors there exactly as in Lite, and a z site is addressed by its default-scope block index 2 + i.
  // Scope 1 = "z": the z sites alone, block i = token i -- lda.vnt with ((doc doc_id)) wrapped in
  // (tag (quote z) <token id> ...); this is the form pinned against Lite (lite_chains.json c4_gibbs).
  void gibbs(int scope, int block, int K, double *mean_nodes_dev, long long *changed_dev) override {
    VK_REQUIRE(scope == 0 || scope == 1, VK_EINVAL, "lda: scopes are 0 (default) and 1 (z)");
    if (scope == 0) {
      VK_REQUIRE(block >= 2 && block < 2 + n_tok, VK_EUNSUPPORTED,
                 "gibbs: the default scope holds the gamma hyper-parameter sites, which cannot be enumerated; "
                 "name a z site (block 2 + token) or use scope z");
      run_sites(1, block, n_tok, 2, K, mean_nodes_dev, changed_dev);
    } else {
      VK_REQUIRE(block == VK_BLOCK_ONE || block == VK_BLOCK_EACH || (block >= 0 && block < n_tok), VK_EINVAL, "lda: bad block");
      run_sites(1, block >= 0 ? block + 2 : block, n_tok, 2, K, mean_nodes_dev, changed_dev);
    }
  }
  void run_sites(int op, int block, int nb, int b_off, int K, double *mean_nodes_dev, long long *acc_dev) {
    VK_REQUIRE(n_prop == n_tok, VK_ESTATE, "mh: unpropagated observations; call vk_make_consistent first");
    if (K > 0 && nb > 0) {
      int n_steps = block == VK_BLOCK_EACH ? K * nb : K;
      uint32_t step0 = h->next_step((uint32_t)n_steps);
      int n_docs_live = 0;   // a doc's maker family exists once a token refers to it (mem, msp.py:28-49)
      for (int32_t c : doclen_host) n_docs_live += c > 0;
      // small sections + z per warp in shared memory, K x V table in global memory (stage 2);
      // if even that does not fit, everything stays in global memory (stage 0)
      size_t small = L.off_nkw;
      static int forced = getenv("VK_LDA_STAGE") ? atoi(getenv("VK_LDA_STAGE")) : -1;
      int stage = (small * LDA_WARPS <= (size_t)smem_limit) ? 2 : 0;
      int warps = LDA_WARPS;
      if (forced == 1 && L.row_bytes <= (size_t)smem_limit) { stage = 1; warps = 1; }
      if (forced == 0) stage = 0;
      size_t sm = stage == 2 ? small * warps : (stage == 1 ? L.row_bytes : 0);
      // the fast z path keeps a packed (doc, word) copy of the corpus per CTA
      int tok_shared = (stage == 2 && L.K <= 32 && L.n_docs <= 65535 && L.V <= 65536 &&
                        sm + (size_t)n_tok * 4 + 16 <= (size_t)smem_limit / 2) ? 1 : 0;
      if (tok_shared) sm += ((size_t)n_tok * 4 + 15) & ~(size_t)15;
      int per_sm = sm ? (int)((size_t)(227 * 1024) / (sm + 1024)) : 16;
      if (per_sm < 1) per_sm = 1;
      if (per_sm * warps > 64) per_sm = 64 / warps;
      int64_t need_ctas = (h->P + warps - 1) / warps;
      int64_t cap = (int64_t)h->sm_count * per_sm;
      int grid = (int)(need_ctas < cap ? need_ctas : cap);
      lda_mh_kernel<<<grid, 32 * warps, sm, h->stream>>>(rows, L, doc.p, word.p, doclen.p, n_tok, h->P, h->p_base,
                                                         h->seed, step0, n_steps, block, stage, tok_shared, acc_dev,
                                                         n_docs_live, mean_nodes_dev, op, nb, b_off);
      VK_CUDA(cudaGetLastError());
    } else {
      if (acc_dev) VK_CUDA(cudaMemsetAsync(acc_dev, 0, (size_t)h->P * sizeof(long long), h->stream));
      if (mean_nodes_dev) {
        lda_fill_kernel<<<grid_for(h->P, 256, h->sm_count, 4), 256, 0, h->stream>>>(mean_nodes_dev, h->P, 0.0);
        VK_CUDA(cudaGetLastError());
      }
    }
  }
  void log_likelihood_at(int scope, int block, bool joint, double *out_dev) override {
    VK_REQUIRE(scope == 0, VK_EINVAL, "lda: only the default scope exists");
    VK_REQUIRE(block >= 0 && block < 2 + n_prop, VK_EUNSUPPORTED, "lda: log_likelihood_at needs a block index");
    lda_loglik_site_kernel<<<grid_p(16), 32, 0, h->stream>>>(rows, L, doc.p, word.p, doclen.p, block, joint ? 1 : 0,
                                                             h->P, out_dev);
    VK_CUDA(cudaGetLastError());
  }
  int64_t num_sites() const override { return 2 + n_prop; }
  int64_t num_blocks(int) const override { return 2 + n_prop; }
  void read_site(int64_t site, double *out_dev) override {
    VK_REQUIRE(site >= 0 && site < 2 + n_prop, VK_EINVAL, "lda: site out of range");
    lda_read_kernel<<<grid_for(h->P, 256, h->sm_count, 4), 256, 0, h->stream>>>(rows, L, site, h->P, out_dev);
  }
  void write_site(int64_t site, const double *in_dev) override {
    VK_REQUIRE(site >= 0 && site < 2 + n_prop, VK_EINVAL, "lda: site out of range");
    lda_write_kernel<<<grid_for(h->P, 256, h->sm_count, 4), 256, 0, h->stream>>>(rows, L, site, h->P, in_dev);
  }
  std::vector<StateArray> state_arrays(bool alloc_alt) override {
    if (alloc_alt && !rowsb.p) { rowsb.alloc(rowsa.n); rows_alt = rowsb.p; }
    StateArray a{(void **)&rows, (void **)&rows_alt, L.row_bytes, 0, 0};
    return {a};
  }
  void resized(int64_t n_new, const std::vector<void *> &nc, const std::vector<void *> &na) override {
    rowsa.adopt(nc[0], (size_t)n_new * L.row_bytes); rowsb.adopt(na[0], (size_t)n_new * L.row_bytes);
    rows = rowsa.p; rows_alt = rowsb.p;
  }
  int device_buffer(int which, void **p, int64_t *bytes) override {
    if (which == 1) { *p = rows; *bytes = (int64_t)rowsa.bytes(); return VK_OK; }
    return VK_EINVAL;
  }
};

Model *make_lda(vk_particles *h, int K, int V, int n_docs, int max_tokens, double a, double b) {
  return new LDAModel(h, K, V, n_docs, max_tokens, a, b);
}
void lda_observe(Model *m, const int32_t *d, const int32_t *w, int64_t n) {
  static_cast<LDAModel *>(m)->observe(d, w, n);
}

}  // namespace vk
