// SP primitives fused into the site kernels: simulate / logDensity of the
// stochastic procedures named by the north star.  fp64, one value per thread.
// Each function cites the Lite PSP it restates (paths relative to
// /root/reference/backend/lite/).
#pragma once
#include <math.h>
#include "philox.cuh"

namespace vk {

#define VK_HALF_LOG2PI 0.91893853320467274178
#define VK_NEG_INF (-INFINITY)

// utils.py:60-61  log(0) = -inf, log1p(-1) = -inf
VK_HD double vlog(double x) { return x == 0.0 ? VK_NEG_INF : log(x); }
VK_HD double vlog1p(double x) { return x == -1.0 ? VK_NEG_INF : log1p(x); }

// utils.py:257-268  log_logistic(x) = x <= -37 ? x : -log1p(exp(-x))
VK_HD double log_logistic(double x) {
  return x <= -37.0 ? x : -log1p(exp(-x));
}
// utils.py:189-203
VK_HD double logistic(double x) {
  return x <= -37.0 ? exp(x) : 1.0 / (1.0 + exp(-x));
}
// utils.py logit(p) = log(p/(1-p))
VK_HD double logit(double p) { return log(p / (1.0 - p)); }

// ---- normal: continuous.py:313-323 -------------------------------------
VK_HD double normal_logpdf(double x, double mu, double sigma) {
  double dev = x - mu;
  return -log(sigma) - VK_HALF_LOG2PI - (0.5 * dev * dev / (sigma * sigma));
}
// Box-Muller on two stream uniforms (cosine branch).
VK_HD double std_normal(Stream &r) {
  double u1 = r.uniform();
  double u2 = r.uniform();
  return sqrt(-2.0 * log(u1)) * cos(6.283185307179586476925 * u2);
}
VK_HD double normal_sim(Stream &r, double mu, double sigma) {
  return mu + sigma * std_normal(r);
}

// ---- gamma: continuous.py:1161-1167 -------------------------------------
// scipy.stats.gamma.logpdf(x, a, scale=1/b) = xlogy(a-1, y) - y - gammaln(a)
// - log(scale) with y = x/scale.
VK_HD double gamma_logpdf(double x, double a, double b) {
  if (x < 0.0) return VK_NEG_INF;
  double y = x * b;
  double t;
  if (a == 1.0) t = 0.0;                 // xlogy(0, y) = 0 even at y = 0
  else if (y == 0.0) t = (a > 1.0) ? VK_NEG_INF : INFINITY;
  else t = (a - 1.0) * log(y);
  return t - y - lgamma(a) + log(b);
}
// Marsaglia-Tsang (2000) with the shape<1 boost; shape a, rate 1.
VK_HD double std_gamma(Stream &r, double a) {
  double boost = 1.0;
  if (a < 1.0) {
    double u = r.uniform();
    boost = pow(u, 1.0 / a);
    a += 1.0;
  }
  double d = a - 1.0 / 3.0;
  double c = 1.0 / sqrt(9.0 * d);
  for (;;) {
    double x = std_normal(r);
    double v = 1.0 + c * x;
    if (v <= 0.0) continue;
    v = v * v * v;
    double u = r.uniform();
    if (log(u) < 0.5 * x * x + d - d * v + d * log(v)) return d * v * boost;
  }
}
VK_HD double gamma_sim(Stream &r, double a, double b) { return std_gamma(r, a) / b; }

// ---- beta: continuous.py:729-859 ----------------------------------------
// scipy.stats.beta.logpdf = xlog1py(b-1,-x) + xlogy(a-1,x) - betaln(a,b)
VK_HD double betaln(double a, double b) { return lgamma(a) + lgamma(b) - lgamma(a + b); }
VK_HD double xlogy(double x, double y) { return x == 0.0 ? 0.0 : x * log(y); }
VK_HD double xlog1py(double x, double y) { return x == 0.0 ? 0.0 : x * log1p(y); }
VK_HD double beta_logpdf(double x, double a, double b) {
  if (x < 0.0 || x > 1.0) return VK_NEG_INF;
  return xlog1py(b - 1.0, -x) + xlogy(a - 1.0, x) - betaln(a, b);
}
// continuous.py:731-856: degenerate spikes, two-gamma, Johnk (+ log-space).
VK_HD double beta_sim(Stream &r, double a, double b) {
  if (a == 0.0 && b == 0.0) return r.uniform() < 0.5 ? 0.0 : 1.0;
  if (a == 0.0) return 1.0;   // sic, continuous.py:739-740
  if (b == 0.0) return 0.0;
  if (fmin(a, b) < 1e-300) return r.uniform() < a / (a + b) ? 0.0 : 1.0;
  if (1.0 < a || 1.0 < b) {
    double G = std_gamma(r, a);
    double H = std_gamma(r, b);
    return G / (G + H);
  }
  for (;;) {
    double u = r.uniform();
    double v = r.uniform();
    double x = pow(u, 1.0 / a);
    double y = pow(v, 1.0 / b);
    if (1.0 < x + y) continue;
    if (0.0 < x + y) return x / (x + y);
    double log_x = log(u) / a, log_y = log(v) / b;
    double log_m = fmax(log_x, log_y);
    log_x -= log_m; log_y -= log_m;
    return exp(log_x - log(exp(log_x) + exp(log_y)));
  }
}

// ---- flip family: discrete.py:55-67, 105-113, 143-150 --------------------
// ---- uniform_continuous (continuous.py:633-640): scipy.stats.uniform.logpdf(x, low, high - low) --------
VK_HD double uniform_logpdf(double x, double lo, double hi) {
  const double scale = hi - lo, y = (x - lo) / scale;
  return (y >= 0.0 && y <= 1.0) ? -log(scale) : VK_NEG_INF;
}
VK_HD double uniform_sim(Stream &r, double lo, double hi) { return lo + (hi - lo) * r.uniform(); }
// ---- expon (continuous.py:1130-1137): scipy.stats.expon.logpdf(x, scale = 1/theta) --------------------
VK_HD double expon_logpdf(double x, double theta) {
  const double scale = 1.0 / theta, y = x / scale;
  return y >= 0.0 ? -y - log(scale) : VK_NEG_INF;
}
VK_HD double expon_sim(Stream &r, double theta) { return -log1p(-r.uniform()) * (1.0 / theta); }

VK_HD double flip_logpmf(bool val, double p) { return val ? vlog(p) : vlog1p(-p); }
VK_HD bool flip_sim(Stream &r, double p) { return r.uniform() < p; }
VK_HD double log_flip_logpmf(bool val, double logp) { return val ? logp : vlog1p(-exp(logp)); }
VK_HD bool log_flip_sim(Stream &r, double logp) { return log(r.uniform()) < logp; }
VK_HD double log_odds_flip_logpmf(bool val, double x) { return log_logistic(val ? x : -x); }
VK_HD bool log_odds_flip_sim(Stream &r, double x) { return logit(r.uniform()) < x; }

// ---- categorical: discrete.py:208-227, utils.py:93-118 -------------------
// ps unnormalised weights (stride between entries), objects default 0..n-1.
// normalizeList: sum <= 0 -> uniform.
VK_HD double categorical_logpmf(int val, const double *ps, int n, int stride) {
  double denom = 0.0;
  for (int i = 0; i < n; ++i) denom += ps[(size_t)i * stride];
  if (val < 0 || val >= n) return VK_NEG_INF;
  double p = denom > 0.0 ? ps[(size_t)val * stride] / denom : 1.0 / (double)n;
  return p == 0.0 ? VK_NEG_INF : log(p);
}
// inverse-CDF on one uniform (np_rng.multinomial(1,ps).argmax(), utils.py:104)
VK_HD int categorical_sim(Stream &r, const double *ps, int n, int stride) {
  double denom = 0.0;
  for (int i = 0; i < n; ++i) denom += ps[(size_t)i * stride];
  double u = r.uniform();
  if (!(denom > 0.0)) {
    int k = (int)(u * (double)n);
    return k < n ? k : n - 1;
  }
  double t = u * denom, acc = 0.0;
  int last = 0;
  for (int i = 0; i < n; ++i) {
    double p = ps[(size_t)i * stride];
    if (p > 0.0) last = i;
    acc += p;
    if (t < acc) return i;
  }
  return last;
}
// normalised probability of one outcome (what exp(logDensityCategorical) is), for enumerative Gibbs
VK_HD double categorical_prob(int val, const double *ps, int n) {
  double denom = 0.0;
  for (int i = 0; i < n; ++i) denom += ps[i];
  return denom > 0.0 ? ps[val] / denom : 1.0 / (double)n;
}
// Enumerative Gibbs selection (infer/egibbs.py:86-88 -> utils.py:170-180): index drawn proportionally to
// pr[0..n), n <= 32, from one uniform.  The inclusive scan is Kogge-Stone shaped -- the same additions, in
// the same order, as the warp-shuffle scan of the LDA kernel and as oracle.c's gibbs_pick -- so the index is
// bit-identical everywhere.  pr is overwritten with the scan.
VK_HD int gibbs_pick_small(double *pr, int n, double u) {
  for (int o = 1; o < 32; o <<= 1) {
    double t[32];
    for (int i = 0; i < n; ++i) t[i] = i >= o ? pr[i - o] : 0.0;
    for (int i = 0; i < n; ++i) if (i >= o) pr[i] += t[i];
  }
  double total = pr[n - 1];
  if (!(total > 0.0)) { int k = (int)(u * (double)n); return k < n ? k : n - 1; }
  double t = u * total;
  for (int i = 0; i < n; ++i) if (t < pr[i]) return i;
  return n - 1;
}
// utils.py:170-184 (log_categorical): max-shift then categorical.
VK_HD double log_categorical_logpmf(int val, const double *lps, int n, int stride) {
  double m = VK_NEG_INF;
  for (int i = 0; i < n; ++i) m = fmax(m, lps[(size_t)i * stride]);
  if (val < 0 || val >= n) return VK_NEG_INF;
  if (isinf(m)) return lps[(size_t)val * stride] - m;  // utils.logsumexp returns m for an infinite max
  double s = 0.0;
  for (int i = 0; i < n; ++i) s += exp(lps[(size_t)i * stride] - m);
  // logsumexp([lp_val]) - logsumexp(lps)
  return lps[(size_t)val * stride] - (m + log(s));
}

// ---- dirichlet: dirichlet.py:41-77, utils.py:135-141 ---------------------
VK_HD double dirichlet_logpdf(const double *theta, const double *alpha, int n) {
  double sa = 0.0, sl = 0.0, dot = 0.0;
  for (int i = 0; i < n; ++i) {
    sa += alpha[i];
    sl += lgamma(alpha[i]);
    dot += (alpha[i] - 1.0) * log(theta[i]);
  }
  return lgamma(sa) - sl + dot;
}
VK_HD void dirichlet_sim(Stream &r, const double *alpha, int n, double *theta_out) {
  double s = 0.0;
  for (int i = 0; i < n; ++i) { theta_out[i] = std_gamma(r, alpha[i]); s += theta_out[i]; }
  for (int i = 0; i < n; ++i) theta_out[i] /= s;
}

// ---- collapsed (sym_)dir_cat with counts: dirichlet.py:158-201 -----------
VK_HD double dircat_logpmf(double count_i, double alpha_i, double count_total, double alpha_total) {
  return log((count_i + alpha_i) / (count_total + alpha_total));
}
// term of logDensityOfData for one cell: gammaln(alpha+c) - gammaln(alpha)
VK_HD double dircat_cell_term(double alpha, double c) {
  return c == 0.0 ? 0.0 : lgamma(alpha + c) - lgamma(alpha);
}
// ---- collapsed beta-bernoulli: discrete.py:419-459 -----------------------
VK_HD double betabern_logpmf(bool val, double a, double b, double yes, double no) {
  double w = (a + yes) / (a + yes + b + no);
  return val ? log(w) : log1p(-w);
}
VK_HD double betabern_log_density_of_data(double a, double b, double yes, double no) {
  return betaln(yes + a, no + b) - betaln(a, b);
}

// ---- logDensityOfData of the made SPs (dirichlet.py:192-201, discrete.py:453-459, 537-541) -------------
VK_HD double dircat_log_density_of_data(const double *alpha, int alpha_stride, const double *counts, int n) {
  double N = 0.0, A = 0.0, term2 = 0.0;
  for (int i = 0; i < n; ++i) { N += counts[i]; A += alpha[i * alpha_stride]; }
  for (int i = 0; i < n; ++i) term2 += lgamma(alpha[i * alpha_stride] + counts[i]) - lgamma(alpha[i * alpha_stride]);
  return lgamma(A) - lgamma(N + A) + term2;
}
VK_HD double suff_bernoulli_log_density_of_data(double w, double yes, double no) {
  return (yes > 0.0 ? yes * log(w) : 0.0) + (no > 0.0 ? no * log1p(-w) : 0.0);
}

}  // namespace vk
