// Pieces of subsampled (austerity) MH shared by the logistic-regression kernels (blr_subsampled*.cuh) and the
// compiled-program kernel (k_program.cu): the keyed permutation that orders the local sections
// (infer/subsampled_mh.py:103 draws np_rng.permutation) and Student's t CDF of the sequential test (:161).
#pragma once
#include "vk_common.cuh"

namespace vk {

struct FeistelKey { uint32_t k[4]; uint32_t half_bits; uint32_t mask; uint64_t N; };

VK_HD uint32_t feistel_mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
// bijection on [0, N): 4-round Feistel on 2*half_bits bits with cycle walking
VK_HD uint64_t feistel_perm(uint64_t q, const FeistelKey &K) {
  uint64_t v = q;
  do {
    uint32_t L = (uint32_t)(v >> K.half_bits) & K.mask, R = (uint32_t)v & K.mask;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      uint32_t t = L ^ (feistel_mix(R ^ K.k[r]) & K.mask);
      L = R; R = t;
    }
    v = ((uint64_t)L << K.half_bits) | (uint64_t)R;
  } while (v >= K.N);
  return v;
}
VK_HD FeistelKey make_feistel_key(uint64_t seed, uint32_t step, uint64_t N) {
  FeistelKey K;
  Philox4 b = philox4x32_10(0xFFFFFFFFu, 0xFFFFFFFFu, step, 0x5EED0001u, (uint32_t)seed, (uint32_t)(seed >> 32));
  for (int i = 0; i < 4; ++i) K.k[i] = b.v[i];
  uint32_t bits = 1;
  while (((uint64_t)1 << bits) < N) ++bits;
  K.half_bits = (bits + 1) / 2;
  if (K.half_bits < 1) K.half_bits = 1;
  K.mask = (K.half_bits >= 32) ? 0xFFFFFFFFu : ((1u << K.half_bits) - 1u);
  K.N = N;
  return K;
}

// regularised incomplete beta / Student-t CDF (replaces scipy.stats.t.cdf, :161)
// The recurrences are evaluated WITHOUT fused multiply-add (__dmul_rn / __dadd_rn are never
// contracted): with FMA the rounding errors of the c and d recurrences stop cancelling in
// del = c*d and |del - 1| hovers around 1e-14 for thousands of iterations at large df
// (measured: 7800 iterations at df = 1e7, t = 1.76 vs 57 without contraction).
// (host + device: the host tabulates t.ppf with the same code, blr_subsampled_persistent.cuh; on the host
// plain operators are not contracted either -- x86-64 without -mfma)
#ifdef __CUDA_ARCH__
#define VK_MUL(a, b) __dmul_rn((a), (b))
#define VK_ADD(a, b) __dadd_rn((a), (b))
#define VK_DIV(a, b) __ddiv_rn((a), (b))
#else
#define VK_MUL(a, b) ((a) * (b))
#define VK_ADD(a, b) ((a) + (b))
#define VK_DIV(a, b) ((a) / (b))
#endif
__host__ __device__ inline double dev_betacf(double a, double b, double x) {
  const double TINY = 1e-300, EPS = 4.5e-16;
  double qab = a + b, qap = a + 1.0, qam = a - 1.0, c = 1.0;
  double d = VK_ADD(1.0, -VK_DIV(VK_MUL(qab, x), qap));
  if (fabs(d) < TINY) d = TINY;
  d = VK_DIV(1.0, d);
  double h = d;
  for (int m = 1; m <= 10000; ++m) {
    double dm = (double)m, m2 = (double)(2 * m);
    double aa = VK_DIV(VK_MUL(VK_MUL(dm, b - dm), x), VK_MUL(qam + m2, a + m2));
    d = VK_ADD(1.0, VK_MUL(aa, d)); if (fabs(d) < TINY) d = TINY;
    c = VK_ADD(1.0, VK_DIV(aa, c)); if (fabs(c) < TINY) c = TINY;
    d = VK_DIV(1.0, d); h = VK_MUL(h, VK_MUL(d, c));
    aa = VK_DIV(VK_MUL(VK_MUL(-(a + dm), qab + dm), x), VK_MUL(a + m2, qap + m2));
    d = VK_ADD(1.0, VK_MUL(aa, d)); if (fabs(d) < TINY) d = TINY;
    c = VK_ADD(1.0, VK_DIV(aa, c)); if (fabs(c) < TINY) c = TINY;
    d = VK_DIV(1.0, d);
    double del = VK_MUL(d, c);
    h = VK_MUL(h, del);
    if (fabs(VK_ADD(del, -1.0)) < EPS) break;
  }
  return h;
}
__host__ __device__ inline double dev_betainc(double a, double b, double x) {
  if (x <= 0.0) return 0.0;
  if (x >= 1.0) return 1.0;
  double bt = exp(lgamma(a + b) - lgamma(a) - lgamma(b) + a * log(x) + b * log1p(-x));
  if (x < (a + 1.0) / (a + b + 2.0)) return bt * dev_betacf(a, b, x) / a;
  return 1.0 - bt * dev_betacf(b, a, 1.0 - x) / b;
}
__host__ __device__ inline double dev_student_t_cdf(double t, double df) {
  if (t != t) return NAN;
  if (isinf(t)) return t > 0 ? 1.0 : 0.0;
  double x = df / (df + t * t);
  double tail = 0.5 * dev_betainc(0.5 * df, 0.5, x);
  return t > 0 ? 1.0 - tail : tail;
}
inline double host_student_t_cdf(double t, double df) { return dev_student_t_cdf(t, df); }

}  // namespace vk
