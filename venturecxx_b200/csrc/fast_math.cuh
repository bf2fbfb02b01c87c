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

}  // namespace vk
