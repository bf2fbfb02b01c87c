// Hand-scheduled fp64 log_logistic for the logistic-regression hot loop.
//
//   log_logistic(x) = min(x,0) - log1p(exp(-|x|))           (lite/utils.py:257-268;
//   for x <= -37 the reference returns x, which this expression rounds to as well)
//
// CUDA's libm route (exp + log1p) costs ~220 issued instructions per evaluation,
// most of them integer/branch overhead of the two general-purpose routines
// (profiles/r1_blr_mh_ncu.md).  Here the argument ranges are known:
//   e = exp(-a), a in [0,700]   : n = rint(-a*log2 e), r = -a - n ln2 (|r| <= ln2/2),
//                                 e^r = 1 + r + r^2 q(r), deg q = 9, scale by 2^n through
//                                 the exponent field (n in [-1010,0]: never subnormal)
//   log1p(e), e in (0,1]        : 2 atanh(s), s = e/(2+e) in (0,1/3] -- no cancellation for
//                                 small e, so the result keeps full RELATIVE accuracy;
//                                 atanh(s)/s = g(s^2), deg g = 10 on [0,1/9];
//                                 1/(2+e) from MUFU.RCP64H + two Newton steps.
// ~45 fp64 instructions, no branches, no table.  Max relative error vs a 50-digit
// reference over [-745,745]: < 4e-16 (tests/test_fast_math.py runs this same
// header on the host).
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifndef VK_HD
#define VK_HD __host__ __device__ __forceinline__
#endif

namespace vk {

VK_HD double fm_rcp_seed(double d) {
#ifdef __CUDA_ARCH__
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  return r;
#else
  return (double)(1.0f / (float)d);
#endif
}
VK_HD double fm_scale2n(double m, int n) {   // m * 2^n through the exponent field
#ifdef __CUDA_ARCH__
  return __hiloint2double(__double2hiint(m) + (n << 20), __double2loint(m));
#else
  uint64_t b; memcpy(&b, &m, 8); b += ((uint64_t)(int64_t)n) << 52; memcpy(&m, &b, 8); return m;
#endif
}
VK_HD int fm_lo_int(double t) {
#ifdef __CUDA_ARCH__
  return __double2loint(t);
#else
  uint64_t b; memcpy(&b, &t, 8); return (int)(uint32_t)b;
#endif
}

// coefficient table in constant memory on the device (fp64 instructions take a
// constant-bank operand directly; literal doubles would be re-materialised with
// two UMOVs each inside the hot loop)
#define VK_FM_COEFFS                                                                                          \
  /* q(r): e^r = 1 + r + r^2 q(r), highest degree first */                                                    \
  2.5100385495510319078e-8, 2.7620088445409748161e-7, 2.7557268459997064772e-6, 0.000024801521295954375131,   \
  0.00019841269863053616878, 0.0013888888917213716901, 0.0083333333333300618325, 0.041666666666624127873,     \
  0.16666666666666667453, 0.50000000000000010221,                                                             \
  /* g(z): 2 atanh(s)/s with z = s^2, highest degree first */                                                 \
  0.17045199264557801332, 0.083239606433710133237, 0.12110290643250378952, 0.13300359007217546084,            \
  0.15386615157236227967, 0.18181740781515607832, 0.22222224082160496478, 0.28571428545431753451,             \
  0.40000000000185926041, 0.66666666666666149045, 2.0,                                                        \
  /* 21: log2(e), 22: ln2 hi, 23: ln2 lo */                                                                   \
  1.4426950408889634074, 6.93147180369123816490e-01, 1.90821492927058770002e-10
#ifdef __CUDACC__
__constant__ double FM_C_DEV[24] = {VK_FM_COEFFS};
#endif
static const double FM_C_HOST[24] = {VK_FM_COEFFS};

VK_HD double fast_log_logistic(double x) {
  const double MAGIC = 6755399441055744.0;           // 1.5 * 2^52
#ifdef __CUDA_ARCH__
  const double *FM_C = FM_C_DEV;
#else
  const double *FM_C = FM_C_HOST;
#endif
  double a = fabs(x);
  a = a < 700.0 ? a : 700.0;                           // exp(-700) is normal; beyond, the term is < 1e-304
  double t = fma(-a, FM_C[21], MAGIC);
  int n = fm_lo_int(t);
  double nf = t - MAGIC;
  double r = fma(nf, -FM_C[22], -a);
  r = fma(nf, -FM_C[23], r);
  double q = FM_C[0];
#pragma unroll
  for (int k = 1; k < 10; ++k) q = fma(q, r, FM_C[k]);
  double p = fma(r * r, q, r);                         // e^r - 1
  double e = fm_scale2n(p + 1.0, n);                   // exp(-a)
  double d = 2.0 + e;
  double rc = fm_rcp_seed(d);
  rc = fma(rc, fma(-d, rc, 1.0), rc);                  // one Newton step: ~2^-40
  double s = e * rc;
  s = fma(fma(-d, s, e), rc, s);                       // residual correction: s = e/(2+e) to ~1 ulp
  double z = s * s;
  double g = FM_C[10];
#pragma unroll
  for (int k = 11; k < 21; ++k) g = fma(g, z, FM_C[k]);
  // min(x,0) = x/2 - |x/2| exactly (halving first: no overflow near DBL_MAX)
  double hx = 0.5 * x;
  return fma(-s, g, hx - fabs(hx));
}


// ---------------------------------------------------------------------------
// Table-driven variant for the logistic-regression MH hot loop: ~20 fp64
// instructions per evaluation instead of ~42 (profiles/r1_blr_mh_ncu.md: the
// kernel is bound by the fp64 pipe, so fp64 work is moved to lookups that run
// on the LSU / integer pipes).
//
//   softplus_neg(x) = log1p(exp(-|x|)),   log_logistic(x) = min(x,0) - softplus_neg(x)
//
//   a = min(|x|, 700)                         (integer clamp of the high word)
//   m = rint(-a 64/ln2), n = m>>6, j = m&63, r = -a - m ln2/64, |r| <= ln2/128
//   e = exp(-a) = Ts + (Ts r) E(r),  Ts = 2^n T[j],  E(r) = (e^r-1)/r  (degree 4)
//   k = rint(128 e), r2 = e inv[k] + (inv[k]-1) = (1+e) inv[k] - 1 with the TRUE e
//                    (inv[k], inv[k]-1 exact 20-bit values; |r2| <= 2^-8 + 2^-20;
//                     k = 0 gives r2 = e exactly, so tiny e keeps full RELATIVE accuracy)
//   log1p(e) = L[k] + r2 Q(r2),  L[k] = -log(inv[k]),  Q(r) = log1p(r)/r (degree 5)
//
// One constant is used for ln2/64, so the relative error of e grows like
// 1.1e-16*|x|; measured against a 40-digit reference (tests/test_fast_math.py):
// relative error of log_logistic <= 4.2e-16 on [-5,5], 1.6e-15 on [-40,40],
// 2.4e-14 on [-700,700]; absolute error <= 2.1e-16*max(1,|x|) everywhere.
//
// Table layout is chosen by the caller: entry j of the exp table is read at
// expT[j*REXP + (lane & (REXP-1))], entry k of the log table at
// logT[k*RLOG + (lane & (RLOG-1))].  In shared memory REXP = 16 (8-byte entries:
// the 16 lanes of a half-warp hit 16 distinct bank pairs) and RLOG = 8 (16-byte
// entries: the 8 lanes of a quarter-warp hit 8 distinct bank quads), so the
// data-dependent lookups are bank-conflict free.  On the host REXP = RLOG = 1.
#include "fast_math_tables.h"

struct FmLogEntry { uint64_t packed; double L; };   // 16 bytes

static const uint64_t FM_E_BITS_HOST[5] = {VK_FM_E_COEFFS_BITS};
static const uint64_t FM_Q_BITS_HOST[6] = {VK_FM_Q_COEFFS_BITS};
static const uint64_t FM_EXP_TABLE_HOST[64] = {VK_FM_EXP_TABLE_BITS};
static const uint64_t FM_LOG_TABLE_HOST[2 * 129] = {VK_FM_LOG_TABLE_BITS};
static const uint64_t FM_E128_BITS_HOST[4] = {VK_FM_E128_COEFFS_BITS};
#ifdef __CUDACC__
__device__ const uint64_t FM_EXP128_TABLE_DEV[128] = {VK_FM_EXP128_TABLE_BITS};
__constant__ uint64_t FM_EQ_BITS_DEV[11] = {VK_FM_E_COEFFS_BITS, VK_FM_Q_COEFFS_BITS};
__device__ const uint64_t FM_EXP_TABLE_DEV[64] = {VK_FM_EXP_TABLE_BITS};
__device__ const uint64_t FM_LOG_TABLE_DEV[2 * 129] = {VK_FM_LOG_TABLE_BITS};
#endif

VK_HD double fm_from_bits(uint64_t b) {
#ifdef __CUDA_ARCH__
  return __longlong_as_double((long long)b);
#else
  double d; memcpy(&d, &b, 8); return d;
#endif
}
VK_HD double fm_from_hi_lo(int hi, int lo) {
#ifdef __CUDA_ARCH__
  return __hiloint2double(hi, lo);
#else
  return fm_from_bits(((uint64_t)(uint32_t)hi << 32) | (uint32_t)lo);
#endif
}
VK_HD int fm_hi_int(double t) {
#ifdef __CUDA_ARCH__
  return __double2hiint(t);
#else
  uint64_t b; memcpy(&b, &t, 8); return (int)(uint32_t)(b >> 32);
#endif
}

template <int REXP, int RLOG>
VK_HD double tbl_softplus_neg(double x, const double *__restrict__ expT,
                              const FmLogEntry *__restrict__ logT, int lane) {
  const double MAGIC = 6755399441055744.0;      // 1.5 * 2^52: ulp 1
  const double MAGIC2 = 52776558133248.0;       // 1.5 * 2^45: ulp 2^-7
#ifdef __CUDA_ARCH__
  const uint64_t *CB = FM_EQ_BITS_DEV;
#define VK_FM_CE(i) __longlong_as_double((long long)CB[i])
#define VK_FM_CQ(i) __longlong_as_double((long long)CB[5 + (i)])
#else
#define VK_FM_CE(i) fm_from_bits(FM_E_BITS_HOST[i])
#define VK_FM_CQ(i) fm_from_bits(FM_Q_BITS_HOST[i])
#endif
  int hi = fm_hi_int(x) & 0x7fffffff;
  hi = hi < 0x4085E000 ? hi : 0x4085E000;       // |x| clamped to ~700 (also inf / NaN)
  const double a = fm_from_hi_lo(hi, fm_lo_int(x));
  const double t = fma(a, -92.332482616893656, MAGIC);          // -64/ln2
  const int m = fm_lo_int(t);
  const double nf = t - MAGIC;
  const double r = fma(nf, -0.010830424696249145, -a);          // ln2/64
  const double T = expT[(m & 63) * REXP + (lane & (REXP - 1))];
  const double Ts = fm_scale2n(T, m >> 6);
  double E = VK_FM_CE(4);
  E = fma(E, r, VK_FM_CE(3));
  E = fma(E, r, VK_FM_CE(2));
  E = fma(E, r, VK_FM_CE(1));
  E = fma(E, r, VK_FM_CE(0));
  const double e = fma(Ts * r, E, Ts);
  const double tt = e + MAGIC2;
  const int k = fm_lo_int(tt);
  const FmLogEntry ent = logT[k * RLOG + (lane & (RLOG - 1))];
  const double inv = fm_from_hi_lo((int)(uint32_t)ent.packed, 0);
  const double c = fm_from_hi_lo((int)(uint32_t)(ent.packed >> 32), 0);
  const double r2 = fma(e, inv, c);
  double Q = VK_FM_CQ(5);
  Q = fma(Q, r2, VK_FM_CQ(4));
  Q = fma(Q, r2, VK_FM_CQ(3));
  Q = fma(Q, r2, VK_FM_CQ(2));
  Q = fma(Q, r2, VK_FM_CQ(1));
  Q = fma(Q, r2, VK_FM_CQ(0));
#undef VK_FM_CE
#undef VK_FM_CQ
  return fma(r2, Q, ent.L);
}

#ifdef __CUDACC__
// ---- device specialisation used by blr_mh_kernel ------------------------------
// Same arithmetic as tbl_softplus_neg, arranged for the issue-slot budget of the
// hot loop: table rows are 128 bytes (16 x 8-byte replicas of T[j], or 8 x 16-byte
// replicas of the log entry) and are addressed with 32-bit shared-window
// addresses (mask + shift-add); polynomial coefficients travel
// as a kernel parameter so fp64 instructions read them straight from the
// constant bank; L[k] is accumulated as a 2^-52 fixed-point integer on the
// integer pipe.  Returns r2*Q(r2) accumulated into `af` and L[k] into `aL`.
struct FmCoef {
  double E[5], Q[6];
  double neg_scale;   // -64/ln2
  double c_ln2;       // ln2/64
  double k128;        // 128
  // blr_mh2.cuh: 128-entry exp table, (e^r - 1)/r of degree 3 on |r| <= ln2/256
  double E3[4];
  double neg_scale128;   // -128/ln2
  double c_ln2_128;      // ln2/128
};
inline FmCoef fm_coef_host() {
  FmCoef c;
  for (int i = 0; i < 5; ++i) c.E[i] = fm_from_bits(FM_E_BITS_HOST[i]);
  for (int i = 0; i < 6; ++i) c.Q[i] = fm_from_bits(FM_Q_BITS_HOST[i]);
  c.neg_scale = -92.332482616893656;
  c.c_ln2 = 0.010830424696249145;
  c.k128 = 128.0;
  for (int i = 0; i < 4; ++i) c.E3[i] = fm_from_bits(FM_E128_BITS_HOST[i]);
  c.neg_scale128 = -184.66496523378731;
  c.c_ln2_128 = 0.0054152123481245727;
  return c;
}
static const int FM_SMEM_LOG_BYTES = 129 * 128;   // 129 rows x 8 replicas x 16 B
static const int FM_SMEM_EXP_BYTES = 64 * 128;    // 64 rows x 16 replicas x 8 B

// fill the replicated shared-memory tables (all threads of the CTA; caller syncs)
__device__ __forceinline__ void fm_fill_smem_tables(unsigned char *log_tbl, unsigned char *exp_tbl) {
  for (int i = threadIdx.x; i < 129 * 8; i += blockDim.x) {
    int k = i >> 3;
    double L = __longlong_as_double((long long)FM_LOG_TABLE_DEV[2 * k + 1]);
    ulonglong2 en;
    en.x = FM_LOG_TABLE_DEV[2 * k];
    en.y = (unsigned long long)__double2ll_rn(L * 4503599627370496.0);   // 2^52 fixed point
    reinterpret_cast<ulonglong2 *>(log_tbl)[i] = en;
  }
  for (int i = threadIdx.x; i < 64 * 16; i += blockDim.x)
    reinterpret_cast<unsigned long long *>(exp_tbl)[i] = FM_EXP_TABLE_DEV[i >> 4];
}

// exp_base / log_base: shared-window byte address of the table + this lane's replica offset
__device__ __forceinline__ void fm_softplus_accum(double v, const FmCoef &cf, uint32_t exp_base,
                                                  uint32_t log_base, double &af, long long &aL) {
  const double MAGIC = 6755399441055744.0;       // 1.5 * 2^52: ulp 1
  int hi = __double2hiint(v) & 0x7fffffff;
  hi = min(hi, 0x4085E000);
  const double a = __hiloint2double(hi, __double2loint(v));
  const double t = fma(a, cf.neg_scale, MAGIC);            // low word = m = rint(-a 64/ln2)
  const int m = __double2loint(t);
  const double nf = t - MAGIC;
  const double r = fma(nf, -cf.c_ln2, -a);
  double T;
  uint32_t ea, th;
  // mask + multiply-add (one LOP3 + one IMAD/LEA each) instead of shift / mask / add
  asm("mad.lo.u32 %0, %1, 128, %2;" : "=r"(ea) : "r"((uint32_t)(m & 63)), "r"(exp_base));
  asm("ld.shared.f64 %0, [%1];" : "=d"(T) : "r"(ea));
  asm("mad.lo.u32 %0, %1, 16384, %2;" : "=r"(th) : "r"((uint32_t)(m & ~63)), "r"((uint32_t)__double2hiint(T)));
  const double Ts = __hiloint2double((int)th, __double2loint(T));   // T * 2^(m>>6)
  double E = cf.E[4];
  E = fma(E, r, cf.E[3]);
  E = fma(E, r, cf.E[2]);
  E = fma(E, r, cf.E[1]);
  E = fma(E, r, cf.E[0]);
  const double e = fma(Ts * r, E, Ts);
  const double tt = fma(e, cf.k128, MAGIC);                // low word = k = rint(128 e) in [0,128]
  uint32_t inv_hi, c_hi, L_lo, L_hi, la;
  asm("mad.lo.u32 %0, %1, 128, %2;" : "=r"(la) : "r"((uint32_t)__double2loint(tt)), "r"(log_base));
  asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
      : "=r"(inv_hi), "=r"(c_hi), "=r"(L_lo), "=r"(L_hi)
      : "r"(la));
  const double r2 = fma(e, __hiloint2double((int)inv_hi, 0), __hiloint2double((int)c_hi, 0));
  double Q = cf.Q[5];
  Q = fma(Q, r2, cf.Q[4]);
  Q = fma(Q, r2, cf.Q[3]);
  Q = fma(Q, r2, cf.Q[2]);
  Q = fma(Q, r2, cf.Q[1]);
  Q = fma(Q, r2, cf.Q[0]);
  af = fma(r2, Q, af);
  aL += (long long)(((unsigned long long)L_hi << 32) | L_lo);
}
// ---- product variant: ONE logarithm for many observations --------------------------------
//   sum_k log1p(e_k) = log prod_k (1 + e_k)
// The exponential half is fm_softplus_accum's.  Every factor is in [1,2], so a thread multiplies the
// factors of its observations into running products (one DADD + one DMUL per observation) and takes ONE
// logarithm per transition (per <= 256 observations, long before 2^1023):
//   q = 2^n m, m in [1,2): n and the table index k = round(128 (m - 1)) come from the high word with
//   integer instructions; r2 = m inv[k] - 1 (one FMA, inv[k] is an exact 20-bit value);
//   log q = n ln2 + L[k] + r2 Q(r2).
// fp64 instructions per observation: 9 (exp) + 2 instead of 17, and no log-table read per observation.
// What is given up is RELATIVE accuracy of a single tiny term (1 + 1e-20 rounds to 1); the ABSOLUTE error
// is one rounding (1.1e-16 relative to q, i.e. 1.1e-16 absolute in log q) per factor, which is what a sum
// of N ~ 1e4 terms of magnitude ~1 sees anyway: chains and log likelihoods stay equal to the oracle's and
// to Lite's at 1e-12 (tests/test_gpu_blr.py, tests/test_gpu_lite_chains.py).
__device__ __forceinline__ double fm_exp_neg_abs(double v, const FmCoef &cf, uint32_t exp_base) {
  const double MAGIC = 6755399441055744.0;       // 1.5 * 2^52: ulp 1
  int hi = __double2hiint(v) & 0x7fffffff;
  hi = min(hi, 0x4085E000);
  const double a = __hiloint2double(hi, __double2loint(v));
  const double t = fma(a, cf.neg_scale, MAGIC);            // low word = m = rint(-a 64/ln2)
  const int m = __double2loint(t);
  const double nf = t - MAGIC;
  const double r = fma(nf, -cf.c_ln2, -a);
  double T;
  uint32_t ea, th;
  asm("mad.lo.u32 %0, %1, 128, %2;" : "=r"(ea) : "r"((uint32_t)(m & 63)), "r"(exp_base));
  asm("ld.shared.f64 %0, [%1];" : "=d"(T) : "r"(ea));
  asm("mad.lo.u32 %0, %1, 16384, %2;" : "=r"(th) : "r"((uint32_t)(m & ~63)), "r"((uint32_t)__double2hiint(T)));
  const double Ts = __hiloint2double((int)th, __double2loint(T));   // T * 2^(m>>6)
  double E = cf.E[4];
  E = fma(E, r, cf.E[3]);
  E = fma(E, r, cf.E[2]);
  E = fma(E, r, cf.E[1]);
  E = fma(E, r, cf.E[0]);
  return fma(Ts * r, E, Ts);
}
// log of a product q >= 1 of (1 + e) terms: af += r2 Q(r2), aL += L[k] (2^-52 fixed point), an += biased
// exponent of q (the caller subtracts 1023 per call and multiplies by ln2 at the end)
__device__ __forceinline__ void fm_log_product(double q, const FmCoef &cf, uint32_t log_base, double &af,
                                               long long &aL, int &an) {
  const uint32_t hi = (uint32_t)__double2hiint(q);
  an += (int)(hi >> 20);
  const uint32_t mant = hi & 0xFFFFFu;
  const double m = __hiloint2double((int)(mant | 0x3FF00000u), __double2loint(q));
  const uint32_t k = (mant + 0x1000u) >> 13;               // round(128 (m - 1)) in [0,128]
  uint32_t inv_hi, L_lo, L_hi, la;
  [[maybe_unused]] uint32_t c_hi;                          // inv[k] - 1: the term-by-term form's addend, not needed here
  asm("mad.lo.u32 %0, %1, 128, %2;" : "=r"(la) : "r"(k), "r"(log_base));
  asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
      : "=r"(inv_hi), "=r"(c_hi), "=r"(L_lo), "=r"(L_hi)
      : "r"(la));
  const double r2 = fma(m, __hiloint2double((int)inv_hi, 0), -1.0);
  double Q = cf.Q[5];
  Q = fma(Q, r2, cf.Q[4]);
  Q = fma(Q, r2, cf.Q[3]);
  Q = fma(Q, r2, cf.Q[2]);
  Q = fma(Q, r2, cf.Q[1]);
  Q = fma(Q, r2, cf.Q[0]);
  af = fma(r2, Q, af);
  aL += (long long)(((unsigned long long)L_hi << 32) | L_lo);
}
#endif  // __CUDACC__

// host-side convenience (tests): unreplicated tables
inline double tbl_log_logistic_host(double x) {
  double hx = 0.5 * x;
  return (hx - fabs(hx)) - tbl_softplus_neg<1, 1>(x, (const double *)FM_EXP_TABLE_HOST,
                                                  (const FmLogEntry *)FM_LOG_TABLE_HOST, 0);
}

}  // namespace vk
