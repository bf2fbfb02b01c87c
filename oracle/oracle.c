/*
 * oracle.c -- CPU restatement (plain C, fp64) of the reference's particle
 * inference path.  TEST INFRASTRUCTURE ONLY: imported by tests/, by
 * __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference
 * legs as the checker and the timed CPU baseline; never by the product path.
 *
 * Each function cites the reference code it follows (paths relative to
 * /root/reference).  The reference draws from numpy RandomState / Python
 * random (backend/lite/trace.py:89-92), which cannot be reproduced on a GPU;
 * the oracle therefore consumes the SAME counter-based Philox4x32-10 streams
 * as the CUDA path (DESIGN.md "RNG contract": key = seed, counter =
 * (particle, step, draw>>1)), so that whole chains can be compared value by
 * value.  Everything else -- which draws happen, in which order, the weight
 * accounting of detach/regen, the accept rule, the +-inf policies -- follows
 * Lite.  Where Lite delegates to scipy (gamma/beta logpdf, gammaln, betaln,
 * t.cdf) the closed forms are restated here and pinned against scipy and
 * against the reference's own functions in tests/ (see tests/golden/make_golden.py
 * and the tests/golden JSON fixtures).
 *
 * Pinning status: PINNED against the reference run in this container.
 *  (1) Whole chains of the five site programs (likelihood_weight, makeConsistent, resimulation MH incl. the
 *      AAA sites, enumerative Gibbs, subsampled MH, log_likelihood_at / log_joint_at) and the resampling
 *      parents equal what the reference's Lite backend computes when its RNG objects are fed this oracle's
 *      draws (orc_draw): tests/golden/make_lite_chains.py -> tests/golden/lite_chains.json ->
 *      tests/test_oracle_lite_chains.py (1e-12 relative; discrete state and node counts exact).
 *  (2) logsumexp, logistic family, sequentialTest, every PSP logDensity on the path are pinned against the
 *      reference's own functions executed from its files and against the reference tests' golden tables
 *      (tests/golden/make_golden.py, tests/test_oracle_golden.py).  gamma/beta logDensity go through scipy
 *      in the reference (no pinned version; scipy 1.18.1 here).
 * Distributional only (by necessity): the samplers' maps from uniforms to values (MT19937 vs Philox).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define HALF_LOG2PI 0.91893853320467274178
typedef unsigned __int128 u128;

/* ------------------------------------------------------------------ RNG */
typedef struct { uint32_t k0, k1, p0, p1, step, d; uint32_t b[4]; } orc_stream;

static void philox(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    uint64_t m0 = (uint64_t)0xD2511F53u * c[0];
    uint64_t m1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(m1 >> 32) ^ c[1] ^ k0;
    uint32_t n2 = (uint32_t)(m0 >> 32) ^ c[3] ^ k1;
    c[0] = n0; c[1] = (uint32_t)m1; c[2] = n2; c[3] = (uint32_t)m0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}
static void st_init(orc_stream *s, uint64_t seed, uint64_t particle, uint32_t step) {
  s->k0 = (uint32_t)seed; s->k1 = (uint32_t)(seed >> 32);
  s->p0 = (uint32_t)particle; s->p1 = (uint32_t)(particle >> 32);
  s->step = step; s->d = 0;
}
static void st_fill(orc_stream *s) {
  s->b[0] = s->p0; s->b[1] = s->p1; s->b[2] = s->step; s->b[3] = s->d >> 1;
  philox(s->b, s->k0, s->k1);
}
static double st_uniform(orc_stream *s) {
  if ((s->d & 1u) == 0u) st_fill(s);
  uint32_t w = (s->d & 1u) * 2u;
  uint64_t x = (((uint64_t)s->b[w] << 32) | s->b[w + 1]) >> 11;
  s->d++;
  /* (2^53 - 1) + 0.5 rounds to 2^53: keep the result strictly below 1 (same clamp as csrc/philox.cuh) */
  { double u = ((double)x + 0.5) * (1.0 / 9007199254740992.0); return u < 1.0 ? u : 0.99999999999999988898; }
}
static void st_seek(orc_stream *s, uint32_t d) { s->d = d; if (d & 1u) st_fill(s); }

void orc_uniforms(uint64_t seed, uint64_t particle, uint32_t step, int n, double *out) {
  orc_stream s; st_init(&s, seed, particle, step);
  for (int i = 0; i < n; ++i) out[i] = st_uniform(&s);
}

/* -------------------------------------------------- lite/utils.py maths */
/* utils.py:60-61 */
static double vlog(double x) { return x == 0 ? -INFINITY : log(x); }
static double vlog1p(double x) { return x == -1 ? -INFINITY : log1p(x); }
/* utils.py:257-268 */
double orc_log_logistic(double x) { return x <= -37 ? x : -log1p(exp(-x)); }
/* utils.py:189-203 */
double orc_logistic(double x) { return x <= -37 ? exp(x) : 1 / (1 + exp(-x)); }
static double logit(double p) { return log(p / (1 - p)); }

/* utils.py:70-91 */
double orc_logsumexp(const double *a, int64_t n) {
  if (n == 0) return -INFINITY;
  double m = -INFINITY, mn = INFINITY; int has_nan = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (a[i] != a[i]) has_nan = 1;
    else { if (a[i] > m) m = a[i]; if (a[i] < mn) mn = a[i]; }
  }
  if (has_nan) return NAN;
  if (isinf(m) && mn != -m) return m;
  if (isinf(m)) return NAN;
  double s = 0;
  for (int64_t i = 0; i < n; ++i) s += exp(a[i] - m);
  return m + log(s);
}
/* utils.py:159-168 */
void orc_normalized_probs(const double *a, int64_t n, double *out) {
  double m = -INFINITY;
  for (int64_t i = 0; i < n; ++i) if (a[i] > m) m = a[i];
  if (!(m > -INFINITY)) { for (int64_t i = 0; i < n; ++i) out[i] = 0; return; }
  double s = 0;
  for (int64_t i = 0; i < n; ++i) { out[i] = exp(a[i] - m); s += out[i]; }
  for (int64_t i = 0; i < n; ++i) out[i] /= s;
}

/* ------------------------------------------------------- SP primitives */
/* continuous.py:319-323 */
static double normal_logpdf(double x, double mu, double sigma) {
  double dev = x - mu;
  return -log(sigma) - HALF_LOG2PI - (0.5 * dev * dev / (sigma * sigma));
}
static double std_normal(orc_stream *r) {
  double u1 = st_uniform(r), u2 = st_uniform(r);
  return sqrt(-2.0 * log(u1)) * cos(6.283185307179586476925 * u2);
}
static double normal_sim(orc_stream *r, double mu, double sigma) { return mu + sigma * std_normal(r); }

/* continuous.py:1166-1167 -> scipy.stats.gamma.logpdf(x, a, scale=1/b) */
static double gamma_logpdf(double x, double a, double b) {
  if (x < 0) return -INFINITY;
  double y = x * b, t;
  if (a == 1) t = 0;
  else if (y == 0) t = a > 1 ? -INFINITY : INFINITY;
  else t = (a - 1) * log(y);
  return t - y - lgamma(a) + log(b);
}
/* continuous.py:1163-1164 np_rng.gamma(shape, scale=1/b): Marsaglia-Tsang */
static double std_gamma(orc_stream *r, double a) {
  double boost = 1;
  if (a < 1) { double u = st_uniform(r); boost = pow(u, 1 / a); a += 1; }
  double d = a - 1.0 / 3.0, c = 1 / sqrt(9 * d);
  for (;;) {
    double x = std_normal(r), v = 1 + c * x;
    if (v <= 0) continue;
    v = v * v * v;
    double u = st_uniform(r);
    if (log(u) < 0.5 * x * x + d - d * v + d * log(v)) return d * v * boost;
  }
}
static double gamma_sim(orc_stream *r, double a, double b) { return std_gamma(r, a) / b; }

static double betaln(double a, double b) { return lgamma(a) + lgamma(b) - lgamma(a + b); }
/* continuous.py:858-859 -> scipy.stats.beta.logpdf */
static double beta_logpdf(double x, double a, double b) {
  if (x < 0 || x > 1) return -INFINITY;
  double t1 = (b - 1 == 0) ? 0 : (b - 1) * log1p(-x);
  double t2 = (a - 1 == 0) ? 0 : (a - 1) * log(x);
  return t1 + t2 - betaln(a, b);
}
/* continuous.py:731-856 */
static double beta_sim(orc_stream *r, double a, double b) {
  if (a == 0 && b == 0) return st_uniform(r) < 0.5 ? 0 : 1;
  if (a == 0) return 1;
  if (b == 0) return 0;
  if (fmin(a, b) < 1e-300) return st_uniform(r) < a / (a + b) ? 0 : 1;
  if (1 < a || 1 < b) { double G = std_gamma(r, a), H = std_gamma(r, b); return G / (G + H); }
  for (;;) {
    double u = st_uniform(r), v = st_uniform(r);
    double x = pow(u, 1 / a), y = pow(v, 1 / b);
    if (1 < x + y) continue;
    if (0 < x + y) return x / (x + y);
    double lx = log(u) / a, ly = log(v) / b, lm = fmax(lx, ly);
    lx -= lm; ly -= lm;
    return exp(lx - log(exp(lx) + exp(ly)));
  }
}
/* utils.py:93-118 */
static double categorical_logpmf(int val, const double *ps, int n) {
  double denom = 0;
  for (int i = 0; i < n; ++i) denom += ps[i];
  if (val < 0 || val >= n) return -INFINITY;
  double p = denom > 0 ? ps[val] / denom : 1.0 / n;
  return p == 0 ? -INFINITY : log(p);
}
static int categorical_sim(orc_stream *r, const double *ps, int n) {
  double denom = 0;
  for (int i = 0; i < n; ++i) denom += ps[i];
  double u = st_uniform(r);
  if (!(denom > 0)) { int k = (int)(u * n); return k < n ? k : n - 1; }
  double t = u * denom, acc = 0; int last = 0;
  for (int i = 0; i < n; ++i) { if (ps[i] > 0) last = i; acc += ps[i]; if (t < acc) return i; }
  return last;
}
/* utils.py:182-184 */
static double log_categorical_logpmf(int val, const double *lps, int n) {
  if (val < 0 || val >= n) return -INFINITY;
  return orc_logsumexp(&lps[val], 1) - orc_logsumexp(lps, n);
}
/* utils.py:137-141 */
static double dirichlet_logpdf(const double *theta, const double *alpha, int n) {
  double sa = 0, sl = 0, dot = 0;
  for (int i = 0; i < n; ++i) { sa += alpha[i]; sl += lgamma(alpha[i]); dot += (alpha[i] - 1) * log(theta[i]); }
  return lgamma(sa) - sl + dot;
}
static void dirichlet_sim(orc_stream *r, const double *alpha, int n, double *out) {
  double s = 0;
  for (int i = 0; i < n; ++i) { out[i] = std_gamma(r, alpha[i]); s += out[i]; }
  for (int i = 0; i < n; ++i) out[i] /= s;
}

/* Test hook for tests/golden/make_lite_chains.py: ONE draw by this oracle's own samplers from stream
 * (seed, particle, step) starting at draw index *d; *d is advanced past what the draw consumed.  The
 * fixture generator installs these draws behind the reference Lite trace's np_rng / py_rng so that Lite's
 * scaffold / detach / regen / accept code runs on the very uniforms this oracle (and the kernels) use.
 * kind: 0 uniform, 1 normal(mu, sigma), 2 gamma(shape, rate), 3 categorical(ps[n]) -> index,
 *       4 sym-dir-cat over params = {alpha, c_0..c_{n-2}} -> index (dirichlet.py:164-166), 5 beta(a, b). */
static int lda_sample_topic(const int32_t *row, int K, double alpha, double u);
double orc_draw(int kind, const double *params, int n, uint64_t seed, uint64_t particle, uint32_t step, uint32_t *d) {
  orc_stream r; st_init(&r, seed, particle, step); st_seek(&r, *d);
  double v = 0;
  switch (kind) {
    case 0: v = st_uniform(&r); break;
    case 1: v = normal_sim(&r, params[0], params[1]); break;
    case 2: v = gamma_sim(&r, params[0], params[1]); break;
    case 3: v = (double)categorical_sim(&r, params, n); break;
    case 4: {
      int32_t row[4096]; int K = n - 1;
      for (int k = 0; k < K; ++k) row[k] = (int32_t)params[1 + k];
      v = (double)lda_sample_topic(row, K, params[0], st_uniform(&r));
      break;
    }
    case 5: v = beta_sim(&r, params[0], params[1]); break;
    default: v = NAN;
  }
  *d = r.d;
  return v;
}

enum { SP_NORMAL = 1, SP_GAMMA, SP_BETA, SP_FLIP, SP_LOG_FLIP, SP_LOG_ODDS_FLIP, SP_CATEGORICAL,
       SP_LOG_CATEGORICAL, SP_DIRICHLET, SP_SYM_DIRICHLET, SP_SYM_DIR_CAT, SP_BETA_BERNOULLI,
       SP_DIR_CAT, SP_UC_DIR_CAT, SP_SUFF_BERNOULLI };

/* batched PSP.simulate: value i drawn from Philox(seed, particle=i, step) */
int orc_sp_simulate(int sp, const double *params, int n_params, int64_t n, int dim, uint64_t seed,
                    uint32_t step, double *out) {
  for (int64_t i = 0; i < n; ++i) {
    const double *q = params + i * n_params;
    orc_stream r; st_init(&r, seed, (uint64_t)i, step);
    switch (sp) {
      case SP_NORMAL: out[i] = normal_sim(&r, q[0], q[1]); break;
      case SP_GAMMA: out[i] = gamma_sim(&r, q[0], q[1]); break;
      case SP_BETA: out[i] = beta_sim(&r, q[0], q[1]); break;
      case SP_FLIP: out[i] = st_uniform(&r) < q[0]; break;                 /* discrete.py:56-59 */
      case SP_LOG_FLIP: out[i] = log(st_uniform(&r)) < q[0]; break;        /* discrete.py:106-108 */
      case SP_LOG_ODDS_FLIP: out[i] = logit(st_uniform(&r)) < q[0]; break; /* discrete.py:144-146 */
      case SP_CATEGORICAL: out[i] = categorical_sim(&r, q, n_params); break;
      case SP_LOG_CATEGORICAL: {                                           /* utils.py:170-180 */
        double m = -INFINITY; double tmp[1024];
        if (n_params > 1024) return 1;
        for (int k = 0; k < n_params; ++k) if (q[k] > m) m = q[k];
        for (int k = 0; k < n_params; ++k) tmp[k] = m > -INFINITY ? exp(q[k] - m) : 0;
        out[i] = categorical_sim(&r, tmp, n_params); break; }
      case SP_DIRICHLET: dirichlet_sim(&r, q, dim, out + i * dim); break;
      case SP_SYM_DIRICHLET: {
        double al[1024]; if (dim > 1024) return 1;
        for (int k = 0; k < dim; ++k) al[k] = q[0];
        dirichlet_sim(&r, al, dim, out + i * dim); break; }
      case SP_SYM_DIR_CAT: {  /* dirichlet.py:164-166: index ~ (alpha + counts); params = alpha, counts[dim] */
        double wts[1024]; if (dim > 1024) return 1;
        for (int k = 0; k < dim; ++k) wts[k] = q[0] + q[1 + k];
        out[i] = categorical_sim(&r, wts, dim); break; }
      case SP_BETA_BERNOULLI: {  /* discrete.py:440-443; params = a, b, yes, no */
        double w = (q[0] + q[2]) / (q[0] + q[2] + q[1] + q[3]);
        out[i] = st_uniform(&r) < w; break; }
      case SP_DIR_CAT: {  /* dirichlet.py:164-166, asymmetric alpha; params = alpha[dim], counts[dim] */
        double wts[1024]; if (dim > 1024) return 1;
        for (int k = 0; k < dim; ++k) wts[k] = q[k] + q[dim + k];
        out[i] = categorical_sim(&r, wts, dim); break; }
      case SP_UC_DIR_CAT: out[i] = categorical_sim(&r, q, dim); break;        /* dirichlet.py:271-273; params = theta[dim] */
      case SP_SUFF_BERNOULLI: out[i] = st_uniform(&r) < q[0]; break;         /* discrete.py:531-532; params = weight */
      default: return 1;
    }
  }
  return 0;
}

int orc_sp_log_density(int sp, const double *values, const double *params, int n_params, int64_t n,
                       int dim, double *out) {
  for (int64_t i = 0; i < n; ++i) {
    const double *q = params + i * n_params;
    double v = values[dim > 1 && (sp == SP_DIRICHLET || sp == SP_SYM_DIRICHLET) ? i * dim : i];
    switch (sp) {
      case SP_NORMAL: out[i] = normal_logpdf(v, q[0], q[1]); break;
      case SP_GAMMA: out[i] = gamma_logpdf(v, q[0], q[1]); break;
      case SP_BETA: out[i] = beta_logpdf(v, q[0], q[1]); break;
      case SP_FLIP: out[i] = v != 0 ? vlog(q[0]) : vlog1p(-q[0]); break;               /* discrete.py:61-67 */
      case SP_LOG_FLIP: out[i] = v != 0 ? q[0] : vlog1p(-exp(q[0])); break;           /* discrete.py:110-113 */
      case SP_LOG_ODDS_FLIP: out[i] = orc_log_logistic(v != 0 ? q[0] : -q[0]); break; /* discrete.py:148-150 */
      case SP_CATEGORICAL: out[i] = categorical_logpmf((int)v, q, n_params); break;
      case SP_LOG_CATEGORICAL: out[i] = log_categorical_logpmf((int)v, q, n_params); break;
      case SP_DIRICHLET: out[i] = dirichlet_logpdf(values + i * dim, q, dim); break;
      case SP_SYM_DIRICHLET: {
        double al[1024]; if (dim > 1024) return 1;
        for (int k = 0; k < dim; ++k) al[k] = q[0];
        out[i] = dirichlet_logpdf(values + i * dim, al, dim); break; }
      case SP_SYM_DIR_CAT: {  /* dirichlet.py:168-173 */
        double tot = 0; int idx = (int)v;
        for (int k = 0; k < dim; ++k) tot += q[1 + k];
        out[i] = log((q[1 + idx] + q[0]) / (tot + q[0] * dim)); break; }
      case SP_BETA_BERNOULLI: {  /* discrete.py:445-451 */
        double w = (q[0] + q[2]) / (q[0] + q[2] + q[1] + q[3]);
        out[i] = v != 0 ? log(w) : log1p(-w); break; }
      case SP_DIR_CAT: {  /* dirichlet.py:168-173 */
        double tot = 0, A = 0; int idx = (int)v;
        for (int k = 0; k < dim; ++k) { tot += q[dim + k]; A += q[k]; }
        out[i] = log((q[dim + idx] + q[idx]) / (tot + A)); break; }
      case SP_UC_DIR_CAT: out[i] = log(q[(int)v]); break;                    /* dirichlet.py:275-277 */
      case SP_SUFF_BERNOULLI: out[i] = v != 0 ? log(q[0]) : log1p(-q[0]); break;   /* discrete.py:534-538 */
      default: return 1;
    }
  }
  return 0;
}

/* logDensityOfData(aux) of the made SPs: dirichlet.py:192-201 (CDirCatOutputPSP; the symmetric maker builds the
 * same PSP with a constant alpha, :303-351), discrete.py:453-459 (CBetaBernoulliOutputPSP), :537-541
 * (SuffBernoulliOutputPSP).  params per row as in include/venture_b200.h vk_sp_log_density_of_data. */
static double data_term(int sp, const double *q, int dim) {
  if (sp == SP_SYM_DIR_CAT || sp == SP_DIR_CAT) {
    double N = 0, A = 0, t2 = 0;
    for (int k = 0; k < dim; ++k) {
      double al = sp == SP_SYM_DIR_CAT ? q[0] : q[k], c = sp == SP_SYM_DIR_CAT ? q[1 + k] : q[dim + k];
      N += c; A += al; t2 += lgamma(al + c) - lgamma(al);
    }
    return lgamma(A) - lgamma(N + A) + t2;
  }
  if (sp == SP_BETA_BERNOULLI) return betaln(q[2] + q[0], q[3] + q[1]) - betaln(q[0], q[1]);
  if (sp == SP_SUFF_BERNOULLI) return (q[1] > 0 ? q[1] * log(q[0]) : 0) + (q[2] > 0 ? q[2] * log1p(-q[0]) : 0);
  return NAN;
}
int orc_sp_log_density_of_data(int sp, const double *params, int n_params, int64_t n, int dim, double *out) {
  for (int64_t i = 0; i < n; ++i) out[i] = data_term(sp, params + i * n_params, dim);
  return 0;
}
/* the uncollapsed makers' AAA LKernels: UDirCatAAALKernel dirichlet.py:250-263, USymDirCatAAALKernel :395-407,
 * UBetaBernoulliAAALKernel discrete.py:502-510; weight = marginalLogDensityOfData (lkernel.py:137-156) */
int orc_sp_aaa_simulate(int sp, const double *params, int n_params, int64_t n, int dim, uint64_t seed, uint32_t step,
                        double *out, double *wout) {
  for (int64_t i = 0; i < n; ++i) {
    const double *q = params + i * n_params;
    orc_stream r; st_init(&r, seed, (uint64_t)i, step);
    double post[1024]; if (dim > 1024) return 1;
    if (sp == SP_UC_DIR_CAT) {
      for (int k = 0; k < dim; ++k) post[k] = q[dim + k] + q[k];
      dirichlet_sim(&r, post, dim, out + i * dim);
      if (wout) wout[i] = data_term(SP_DIR_CAT, q, dim);
    } else if (sp == SP_SYM_DIRICHLET) {
      for (int k = 0; k < dim; ++k) post[k] = q[1 + k] + q[0];
      dirichlet_sim(&r, post, dim, out + i * dim);
      if (wout) wout[i] = data_term(SP_SYM_DIR_CAT, q, dim);
    } else if (sp == SP_SUFF_BERNOULLI) {
      out[i] = beta_sim(&r, q[0] + q[2], q[1] + q[3]);
      if (wout) wout[i] = data_term(SP_BETA_BERNOULLI, q, dim);
    } else return 1;
  }
  return 0;
}

/* dirichlet.py:192-201 CDirCatOutputPSP.logDensityOfData, symmetric alpha */
double orc_sym_dircat_log_density_of_data(const double *counts, int n, double alpha) {
  double N = 0;
  for (int i = 0; i < n; ++i) N += counts[i];
  double A = alpha * n;
  double term1 = lgamma(A) - lgamma(N + A), term2 = 0;
  for (int i = 0; i < n; ++i) term2 += lgamma(alpha + counts[i]) - lgamma(alpha);
  return term1 + term2;
}
/* discrete.py:453-459 */
double orc_beta_bernoulli_log_density_of_data(double a, double b, double yes, double no) {
  return betaln(yes + a, no + b) - betaln(a, b);
}

/* ----------------------------------------------------------- resampling */
/* Fixed-point weights: q_i = rint(exp(lw_i - max) * 2^S).  All -inf -> all 1
 * (utils.py:93-99,176-180: "equally impossible").  NaN -> 0. */
int orc_frac_bits(int64_t P_total) {
  int lg = 0;
  while (((int64_t)1 << lg) < P_total) ++lg;
  int S = 62 - lg;
  return S > 52 ? 52 : S;
}
void orc_fixed_cdf(const double *lw, int64_t n, double m, int frac_bits, uint64_t *cdf) {
  double scale = ldexp(1.0, frac_bits);
  uint64_t run = 0;
  for (int64_t i = 0; i < n; ++i) {
    uint64_t q;
    if (!(m > -INFINITY)) q = 1;
    /* max = +inf: the reference's sampleLogCategorical (utils.py:170-180) forms exp(inf - inf) = NaN and numpy
     * raises; defined here as "all the mass on the +inf entries" (the limit), so the total is never 0 */
    else if (m == INFINITY) q = lw[i] == INFINITY ? 1 : 0;
    else { double d = lw[i] - m; q = (d != d) ? 0 : (uint64_t)nearbyint(exp(d) * scale); }
    run += q; cdf[i] = run;
  }
}
static int64_t upper_bound(const uint64_t *cdf, int64_t n, uint64_t thr) {
  int64_t lo = 0, hi = n - 1;
  while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (cdf[mid] > thr) hi = mid; else lo = mid + 1; }
  return lo;
}
/* trace_set.py:209-219 (_resample_traces): mode 0 multinomial (one uniform
 * per child, stream (seed, particle=k, step)); mode 1 systematic (one uniform,
 * stream (seed, GLOBAL, step)). uniforms may override the stream draws. */
void orc_ancestors(const double *lw, int64_t n, int64_t n_new, int mode, uint64_t seed, uint32_t step,
                   const double *uniforms, int32_t *anc) {
  double m = -INFINITY;
  for (int64_t i = 0; i < n; ++i) if (lw[i] == lw[i] && lw[i] > m) m = lw[i];
  uint64_t *cdf = (uint64_t *)malloc((size_t)n * sizeof(uint64_t));
  orc_fixed_cdf(lw, n, m, orc_frac_bits(n), cdf);
  uint64_t total = cdf[n - 1];
  double u_sys = 0;
  if (mode == 1) {
    if (uniforms) u_sys = uniforms[0];
    else { orc_stream r; st_init(&r, seed, 0xFFFFFFFFFFFFFFFFull, step); u_sys = st_uniform(&r); }
  }
  for (int64_t k = 0; k < n_new; ++k) {
    uint64_t thr;
    if (mode == 1) {
      uint64_t U0 = (uint64_t)(u_sys * 4294967296.0);
      if (U0 > 0xFFFFFFFFull) U0 = 0xFFFFFFFFull;
      u128 num = ((((u128)(uint64_t)k) << 32) | U0) * (u128)total;
      u128 den = ((u128)(uint64_t)n_new) << 32;
      thr = (uint64_t)(num / den);
    } else {
      double u;
      if (uniforms) u = uniforms[k];
      else { orc_stream r; st_init(&r, seed, (uint64_t)k, step); u = st_uniform(&r); }
      uint64_t U = (uint64_t)(u * 9007199254740992.0);
      if (U > 9007199254740991ull) U = 9007199254740991ull;
      thr = (uint64_t)(((u128)U * (u128)total) >> 53);
    }
    anc[k] = (int32_t)upper_bound(cdf, n, thr);
  }
  free(cdf);
}

/* mh.py:187-208 */
static double mh_log_alpha(double rho, double xi) {
  if (rho == -INFINITY && xi == -INFINITY) return INFINITY;
  if (rho == INFINITY && xi == INFINITY) return -INFINITY;
  return xi - rho;
}

/* ------------------------------------------------- C1: normal-normal */
static double nn_absorb(const double *y, int64_t k0, int64_t k1, double x, double s) {
  double w = 0;  /* regen.py:120-127 absorb: weight += logDensity(obs | x) */
  for (int64_t k = k0; k < k1; ++k) w += normal_logpdf(y[k], x, s);
  return w;
}
/* lite/trace.py:522-530 */
void orc_nn_likelihood_weight(double *x, double *lw, const double *y, int64_t n_obs, int64_t P,
                              int64_t p_base, double mu0, double s0, double s, uint64_t seed, uint32_t step) {
#pragma omp parallel for schedule(static)
  for (int64_t p = 0; p < P; ++p) {
    orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step);
    x[p] = normal_sim(&r, mu0, s0);
    if (lw) lw[p] = nn_absorb(y, 0, n_obs, x[p], s);
  }
}
/* lite/trace.py:422-448 */
void orc_nn_make_consistent(const double *x, double *lw, double *incr, const double *y, int64_t k0,
                            int64_t k1, int64_t P, double s) {
#pragma omp parallel for schedule(static)
  for (int64_t p = 0; p < P; ++p) {
    double w = nn_absorb(y, k0, k1, x[p], s);
    if (w != w) w = -INFINITY;
    if (incr) incr[p] = w;
    lw[p] += w;
  }
}
/* infer/mh.py:59-113 with MHOperator (:187-208); K transitions */
void orc_nn_mh(double *x, const double *y, int64_t n_obs, int64_t P, int64_t p_base, double mu0, double s0,
               double s, uint64_t seed, uint32_t step0, int K, int64_t *acc) {
#pragma omp parallel for schedule(static)
  for (int64_t p = 0; p < P; ++p) {
    int64_t a = 0;
    for (int t = 0; t < K; ++t) {
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step0 + (uint32_t)t);
      (void)st_uniform(&r);                       /* sampleBlock: a single block */
      double logu = log(st_uniform(&r));          /* mh.py:85 */
      double rho = nn_absorb(y, 0, n_obs, x[p], s);      /* detachAndExtract */
      double xn = normal_sim(&r, mu0, s0);               /* regen: simulate from prior */
      double xi = nn_absorb(y, 0, n_obs, xn, s);
      double alpha = mh_log_alpha(rho, xi);
      if (logu < alpha) { x[p] = xn; ++a; }
    }
    if (acc) acc[p] = a;
  }
}
void orc_nn_log_likelihood_at(const double *x, const double *y, int64_t n_obs, int64_t P, double mu0,
                              double s0, double s, int joint, double *out) {
  for (int64_t p = 0; p < P; ++p) {
    double v = nn_absorb(y, 0, n_obs, x[p], s);
    if (joint) v += normal_logpdf(x[p], mu0, s0);
    out[p] = v;
  }
}

/* ------------------------------------------- C2: logistic regression */
/* w[D][P] particle-minor; X[N][D] row-major; y[N] bytes.
 * eta_i = vector_dot(w, x_i) (vectors.py:159-177); obs density
 * log_logistic(y ? eta : -eta) (discrete.py:148-150). */
static double blr_eta(const double *wp, const double *xi, int D) {
  double e = 0;
  for (int j = 0; j < D; ++j) e += wp[j] * xi[j];
  return e;
}
static double blr_loglik_range(const double *wp, const double *X, const uint8_t *y, int D, int64_t i0, int64_t i1) {
  double acc = 0;
  for (int64_t i = i0; i < i1; ++i) {
    double e = blr_eta(wp, X + i * D, D);
    acc += orc_log_logistic(y[i] ? e : -e);
  }
  return acc;
}
void orc_blr_log_likelihood(const double *w, const double *X, const uint8_t *y, int D, int64_t i0, int64_t i1,
                            int64_t P, double *out) {
#pragma omp parallel for schedule(static)
  for (int64_t p = 0; p < P; ++p) {
    double wp[256];
    for (int j = 0; j < D; ++j) wp[j] = w[(int64_t)j * P + p];
    out[p] = blr_loglik_range(wp, X, y, D, i0, i1);
  }
}
/* likelihood_weight / reset_to_prior: site j uses draws 2j, 2j+1 */
void orc_blr_prior(double *w, int D, int64_t P, int64_t p_base, double mu0, double s0, uint64_t seed, uint32_t step) {
#pragma omp parallel for schedule(static)
  for (int64_t p = 0; p < P; ++p)
    for (int j = 0; j < D; ++j) {
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step); st_seek(&r, 2u * (uint32_t)j);
      w[(int64_t)j * P + p] = normal_sim(&r, mu0, s0);
    }
}
/* block_mode: -1 one, -3 each, -2 all, >=0 fixed site. n_steps proposals. */
void orc_blr_mh(double *w, const double *X, const uint8_t *y, int D, int64_t N, int64_t P, int64_t p_base,
                double mu0, double s0, uint64_t seed, uint32_t step0, int n_steps, int block_mode,
                int64_t *acc, double *ll_out) {
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t p = 0; p < P; ++p) {
    double wp[256], wn[256];
    for (int j = 0; j < D; ++j) wp[j] = w[(int64_t)j * P + p];
    double *eta = (double *)malloc((size_t)N * sizeof(double));
    for (int64_t i = 0; i < N; ++i) eta[i] = blr_eta(wp, X + i * D, D);
    int64_t a = 0;
    for (int t = 0; t < n_steps; ++t) {
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step0 + (uint32_t)t);
      double u_sel = st_uniform(&r);
      double logu = log(st_uniform(&r));
      /* detach: rho = sum of the absorbing children's logDensity at the cached dots */
      double rho = 0;
      for (int64_t i = 0; i < N; ++i) rho += orc_log_logistic(y[i] ? eta[i] : -eta[i]);
      memcpy(wn, wp, sizeof(double) * (size_t)D);
      if (block_mode == -2) {
        for (int j = 0; j < D; ++j) { st_seek(&r, 2u + 2u * (uint32_t)j); wn[j] = normal_sim(&r, mu0, s0); }
      } else {
        int j;
        if (block_mode == -1) { j = (int)(u_sel * D); if (j >= D) j = D - 1; }  /* trace.py:320 sampleBlock */
        else if (block_mode == -3) j = t % D;
        else j = block_mode;
        wn[j] = normal_sim(&r, mu0, s0);
      }
      /* regen: every vector_dot node recomputed, every flip absorbs */
      double xi = 0;
      for (int64_t i = 0; i < N; ++i) { double e = blr_eta(wn, X + i * D, D); xi += orc_log_logistic(y[i] ? e : -e); }
      if (logu < mh_log_alpha(rho, xi)) {
        memcpy(wp, wn, sizeof(double) * (size_t)D);
        for (int64_t i = 0; i < N; ++i) eta[i] = blr_eta(wp, X + i * D, D);
        ++a;
      }
    }
    for (int j = 0; j < D; ++j) w[(int64_t)j * P + p] = wp[j];
    if (acc) acc[p] = a;
    if (ll_out) { double v = 0; for (int64_t i = 0; i < N; ++i) v += orc_log_logistic(y[i] ? eta[i] : -eta[i]); ll_out[p] = v; }
    free(eta);
  }
}

/* ------------------------------------------------ subsampled MH pieces */
/* regularised incomplete beta I_x(a,b) (continued fraction, modified Lentz);
 * stands in for scipy.stats.t.cdf (infer/subsampled_mh.py:161). */
static double betacf(double a, double b, double x) {
  const double TINY = 1e-300, EPS = 4.5e-16;
  double qab = a + b, qap = a + 1, qam = a - 1, c = 1, d = 1 - qab * x / qap;
  if (fabs(d) < TINY) d = TINY;
  d = 1 / d;
  double h = d;
  for (int m = 1; m <= 10000; ++m) {
    int m2 = 2 * m;
    double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
    d = 1 + aa * d; if (fabs(d) < TINY) d = TINY;
    c = 1 + aa / c; if (fabs(c) < TINY) c = TINY;
    d = 1 / d; h *= d * c;
    aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
    d = 1 + aa * d; if (fabs(d) < TINY) d = TINY;
    c = 1 + aa / c; if (fabs(c) < TINY) c = TINY;
    d = 1 / d;
    double del = d * c;
    h *= del;
    if (fabs(del - 1) < EPS) break;
  }
  return h;
}
static double betainc(double a, double b, double x) {
  if (x <= 0) return 0;
  if (x >= 1) return 1;
  double bt = exp(lgamma(a + b) - lgamma(a) - lgamma(b) + a * log(x) + b * log1p(-x));
  if (x < (a + 1) / (a + b + 2)) return bt * betacf(a, b, x) / a;
  return 1 - bt * betacf(b, a, 1 - x) / b;
}
double orc_student_t_cdf(double t, double df) {
  if (t != t) return NAN;
  if (isinf(t)) return t > 0 ? 1 : 0;
  double x = df / (df + t * t);
  double tail = 0.5 * betainc(0.5 * df, 0.5, x);
  return t > 0 ? 1 - tail : tail;
}

/* infer/subsampled_mh.py:118-171 sequentialTest, with fun_dllh(i) = dllh[i].
 * Returns accept; *n_out = sections consumed; *tstat_out as in :170. */
int orc_sequential_test(double mu_0, int k0, double Nbatch, double N, double epsilon, const double *dllh,
                        double *n_out, double *tstat_out) {
  double mx = 0, mx2 = 0, k = 0, n = 0, cum = 0, cum2 = 0, tstat = 0;
  int have_t = 0, accept = -1;
  while (n < N) {
    double n_start = n, n_end = fmin(n + Nbatch, N);
    for (int64_t i = (int64_t)n_start; i < (int64_t)n_end; ++i) { cum += dllh[i]; cum2 += dllh[i] * dllh[i]; }
    k += 1; n = n_end; mx = cum / n_end; mx2 = cum2 / n_end;
    if (k < k0 && n < N) continue;
    if (n == N) { accept = mx >= mu_0; break; }
    double sx = sqrt((1 - (n - 1) / (N - 1)) * (mx2 - mx * mx) / (n - 1));
    double q;
    if (sx == 0) q = mx >= mu_0 ? 1 : 0;
    else { tstat = (mx - mu_0) / sx; have_t = 1; q = orc_student_t_cdf(tstat, n - 1); }
    if (q < epsilon) { accept = 0; break; }
    else if (q > 1 - epsilon) { accept = 1; break; }
  }
  *n_out = n;
  /* :170  tstat if tstat else (+inf if accept else -inf); python truthiness: 0.0 counts as "no" */
  *tstat_out = (have_t && tstat != 0) ? tstat : (accept ? INFINITY : -INFINITY);
  return accept;
}

/* ---------------------------------------------------------- C3: HMM */
/* examples/hmm.vnt:31-60.  hist[P][Tpad] bytes, 1-based states; trans n_rows x S,
 * emit S x O.  One filter step = transition draw (regen eval of the new mem
 * family, prior proposal) + emission weight (trace.py:422-448, utils.py:106-118). */
void orc_hmm_step(uint8_t *hist, int Tpad, int t, double *lw, double *incr, const double *trans, const double *emit,
                  int S, int O, int init, int symbol, int64_t P, int64_t p_base, uint64_t seed, uint32_t step) {
#pragma omp parallel for schedule(static)
  for (int64_t p = 0; p < P; ++p) {
    int s;
    if (t == 0) s = init;
    else {
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step);
      s = 1 + categorical_sim(&r, trans + (hist[p * Tpad + t - 1] - 1) * S, S);
    }
    hist[p * Tpad + t] = (uint8_t)s;
    double w = categorical_logpmf(symbol - 1, emit + (s - 1) * O, O);
    if (w != w) w = -INFINITY;
    lw[p] += w;
    if (incr) incr[p] = w;
  }
}
void orc_hmm_prior(uint8_t *hist, int Tpad, int T, double *lw, const double *trans, const double *emit, int S, int O,
                   int init, const int32_t *symbols, int64_t P, int64_t p_base, uint64_t seed, uint32_t step) {
  for (int64_t p = 0; p < P; ++p) {
    orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step);
    int s = init; double w = 0;
    for (int t = 0; t < T; ++t) {
      if (t > 0) s = 1 + categorical_sim(&r, trans + (s - 1) * S, S);
      hist[p * Tpad + t] = (uint8_t)s;
      w += categorical_logpmf(symbols[t] - 1, emit + (s - 1) * O, O);
    }
    if (lw) lw[p] = w;
  }
}
/* single-site MH on state(t), t = block+1 (0-based time), blocks = state(2..T) */
void orc_hmm_mh(uint8_t *hist, int Tpad, int T, const double *trans, const double *emit, int S, int O,
                const int32_t *symbols, int64_t P, int64_t p_base, uint64_t seed, uint32_t step0, int n_steps,
                int block_mode, int64_t *acc) {
  int nb = T - 1;
  for (int64_t p = 0; p < P; ++p) {
    uint8_t *h = hist + p * Tpad; int64_t a = 0;
    for (int k = 0; k < n_steps; ++k) {
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step0 + (uint32_t)k);
      double u_sel = st_uniform(&r), logu = log(st_uniform(&r));
      int b;
      if (block_mode == -1) { b = (int)(u_sel * nb); if (b >= nb) b = nb - 1; }
      else if (block_mode == -3) b = k % nb; else b = block_mode;
      int t = b + 1, prev = h[t - 1], old = h[t], nxt = (t + 1 < T) ? h[t + 1] : 0;
      /* detach: absorbing children = obs_t and state(t+1) */
      double rho = categorical_logpmf(symbols[t] - 1, emit + (old - 1) * O, O);
      if (nxt) rho += categorical_logpmf(nxt - 1, trans + (old - 1) * S, S);
      int s = 1 + categorical_sim(&r, trans + (prev - 1) * S, S);
      double xi = categorical_logpmf(symbols[t] - 1, emit + (s - 1) * O, O);
      if (nxt) xi += categorical_logpmf(nxt - 1, trans + (s - 1) * S, S);
      if (logu < mh_log_alpha(rho, xi)) { h[t] = (uint8_t)s; ++a; }
    }
    if (acc) acc[p] = a;
  }
}

/* ---------------------------------------------------------- C4: LDA */
/* examples/lda.vnt:18-41, collapsed makers.  Per particle p (particle-major):
 * alpha[p][2], nk[p][K], ndk[p][n_docs*K], nkw[p][K*V] (int32 here), z[p][max_tok]. */
typedef struct { int K, V, n_docs, max_tok; double a, b; } orc_lda;

/* dirichlet.py:164-166 with the order-independent arithmetic of the product:
 * first k with u*(N + K alpha) < C_k + (k+1) alpha */
static int lda_sample_topic(const int32_t *row, int K, double alpha, double u) {
  int tot = 0;
  for (int k = 0; k < K; ++k) tot += row[k];
  double t = u * ((double)tot + (double)K * alpha);
  int c = 0;
  for (int k = 0; k < K; ++k) { c += row[k]; if (t < (double)c + (double)(k + 1) * alpha) return k; }
  return K - 1;
}
/* dirichlet.py:168-173 */
static double dircat_logpmf(double c, double alpha, double N, double A) { return log((c + alpha) / (N + A)); }
/* dirichlet.py:192-201 summed over tables */
static double lda_aaa(const int32_t *counts, int n_tab, int n_cell, double alpha) {
  double A = alpha * n_cell, acc = 0;
  for (int t = 0; t < n_tab; ++t) {
    double N = 0;
    for (int i = 0; i < n_cell; ++i) N += counts[t * n_cell + i];
    acc += lgamma(A) - lgamma(N + A);
    for (int i = 0; i < n_cell; ++i) { int c = counts[t * n_cell + i]; if (c) acc += lgamma(alpha + c) - lgamma(alpha); }
  }
  return acc;
}
/* mode 0 make_consistent [t0,t1); 1 likelihood_weight; 2 reset_to_prior */
void orc_lda_seq(double *alpha, int32_t *nk, int32_t *ndk, int32_t *nkw, uint8_t *z, const orc_lda *L,
                 const int32_t *doc, const int32_t *word, int t0, int t1, double *lw, double *incr, int64_t P,
                 int64_t p_base, uint64_t seed, uint32_t step, int mode) {
  int K = L->K, V = L->V;
  for (int64_t p = 0; p < P; ++p) {
    double *al = alpha + p * 2; int32_t *NK = nk + p * K, *NDK = ndk + p * (int64_t)L->n_docs * K, *NKW = nkw + p * (int64_t)K * V;
    uint8_t *Z = z + p * L->max_tok;
    orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step);
    if (mode >= 1) {
      al[0] = gamma_sim(&r, L->a, L->b); al[1] = gamma_sim(&r, L->a, L->b);
      memset(NK, 0, sizeof(int32_t) * K); memset(NDK, 0, sizeof(int32_t) * (size_t)L->n_docs * K);
      memset(NKW, 0, sizeof(int32_t) * (size_t)K * V);
    }
    double wsum = 0;
    for (int i = t0; i < t1; ++i) {
      int d = doc[i], w = word[i];
      int zi = lda_sample_topic(NDK + d * K, K, al[0], st_uniform(&r));
      wsum += dircat_logpmf(NKW[(int64_t)zi * V + w], al[1], NK[zi], al[1] * V);
      Z[i] = (uint8_t)zi; NDK[d * K + zi]++; NKW[(int64_t)zi * V + w]++; NK[zi]++;
    }
    if (mode == 0) { if (wsum != wsum) wsum = -INFINITY; if (incr) incr[p] = wsum; lw[p] += wsum; }
    else if (mode == 1) lw[p] = wsum;
  }
}
void orc_lda_mh(double *alpha, int32_t *nk, int32_t *ndk, int32_t *nkw, uint8_t *z, const orc_lda *L,
                const int32_t *doc, const int32_t *word, int n_tok, int64_t P, int64_t p_base, uint64_t seed,
                uint32_t step0, int n_steps, int block_mode, int64_t *acc) {
  int K = L->K, V = L->V, nb = 2 + n_tok;
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t p = 0; p < P; ++p) {
    double *al = alpha + p * 2; int32_t *NK = nk + p * K, *NDK = ndk + p * (int64_t)L->n_docs * K, *NKW = nkw + p * (int64_t)K * V;
    uint8_t *Z = z + p * L->max_tok; int64_t a = 0;
    for (int k = 0; k < n_steps; ++k) {
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step0 + (uint32_t)k);
      double u_sel = st_uniform(&r), logu = log(st_uniform(&r));
      int b;
      if (block_mode == -1) { b = (int)(u_sel * nb); if (b >= nb) b = nb - 1; }
      else if (block_mode == -3) b = k % nb; else b = block_mode;
      if (b < 2) {
        /* AAA: DeterministicMakerAAALKernel weight = sum logDensityOfData (lkernel.py:158-171) */
        double a_new = gamma_sim(&r, L->a, L->b), rho, xi;
        if (b == 0) { rho = lda_aaa(NDK, L->n_docs, K, al[0]); xi = lda_aaa(NDK, L->n_docs, K, a_new); }
        else { rho = lda_aaa(NKW, K, V, al[1]); xi = lda_aaa(NKW, K, V, a_new); }
        if (logu < mh_log_alpha(rho, xi)) { al[b] = a_new; ++a; }
      } else {
        int i = b - 2, d = doc[i], w = word[i], zo = Z[i];
        /* detach: unconstrain + unincorporate token and z (detach.py:52-65,165-201) */
        NKW[(int64_t)zo * V + w]--; NK[zo]--; NDK[d * K + zo]--;
        double rho = dircat_logpmf(NKW[(int64_t)zo * V + w], al[1], NK[zo], al[1] * V);
        int zn = lda_sample_topic(NDK + d * K, K, al[0], st_uniform(&r));
        double xi = dircat_logpmf(NKW[(int64_t)zn * V + w], al[1], NK[zn], al[1] * V);
        int zf = zo;
        if (logu < mh_log_alpha(rho, xi)) { zf = zn; ++a; }
        Z[i] = (uint8_t)zf; NKW[(int64_t)zf * V + w]++; NK[zf]++; NDK[d * K + zf]++;
      }
    }
    if (acc) acc[p] = a;
  }
}

/* ------------------------------------------------ enumerative Gibbs (N1) */
/* infer/egibbs.py:42-107 under mixMH (mh.py:59-113): for a discrete principal node every candidate value v
 * (PSP.enumerateValues) is installed with a DeterministicLKernel and the scaffold regenerated, giving
 * xi_v = logDensity(v | parents) + sum of the absorbing children's logDensity; the new value is drawn
 * proportionally to exp(xi_v) (utils.py:170-180 sampleLogCategorical) and always accepted (logAlpha = 0, rhoMix =
 * xiMix), although mixMH still draws its accept uniform (draw 1).  exp(xi_v - max) is evaluated here as the
 * product of the probabilities themselves (no log / exp round trip); the selection uniform is draw 2.
 * The inclusive scan has the shape of the warp scan the kernels use (Kogge-Stone over <= 32 lanes, sequential
 * beyond), so that the selected index is bit-identical on both sides. */
static int gibbs_pick(double *pr, int n, double u) {
  if (n <= 32) {
    for (int o = 1; o < 32; o <<= 1) {
      double t[32];
      for (int i = 0; i < n; ++i) t[i] = i >= o ? pr[i - o] : 0.0;
      for (int i = 0; i < n; ++i) if (i >= o) pr[i] += t[i];
    }
  } else {
    for (int i = 1; i < n; ++i) pr[i] += pr[i - 1];
  }
  double total = pr[n - 1];
  if (!(total > 0)) { int k = (int)(u * n); return k < n ? k : n - 1; }   /* utils.py:95-99: all impossible => uniform */
  double t = u * total;
  for (int i = 0; i < n; ++i) if (t < pr[i]) return i;
  return n - 1;
}
static double cat_prob(int val, const double *ps, int n) {
  double denom = 0;
  for (int i = 0; i < n; ++i) denom += ps[i];
  return denom > 0 ? ps[val] / denom : 1.0 / n;
}
/* z sites of the LDA model; blocks are the tokens (block_mode -1 one, -3 each, >= 0 token index) */
void orc_lda_gibbs(double *alpha, int32_t *nk, int32_t *ndk, int32_t *nkw, uint8_t *z, const orc_lda *L,
                   const int32_t *doc, const int32_t *word, int n_tok, int64_t P, int64_t p_base, uint64_t seed,
                   uint32_t step0, int n_steps, int block_mode, int64_t *changed) {
  int K = L->K, V = L->V, nb = n_tok;
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t p = 0; p < P; ++p) {
    double *al = alpha + p * 2; int32_t *NK = nk + p * K, *NDK = ndk + p * (int64_t)L->n_docs * K, *NKW = nkw + p * (int64_t)K * V;
    uint8_t *Z = z + p * L->max_tok; int64_t ch = 0;
    double *pr = (double *)malloc(sizeof(double) * (size_t)K);
    for (int k = 0; k < n_steps; ++k) {
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step0 + (uint32_t)k);
      double u_sel = st_uniform(&r); (void)st_uniform(&r);       /* mh.py:85: drawn, never decisive */
      int i;
      if (block_mode == -1) { i = (int)(u_sel * nb); if (i >= nb) i = nb - 1; }
      else if (block_mode == -3) i = k % nb; else i = block_mode;
      int d = doc[i], w = word[i], zo = Z[i];
      NKW[(int64_t)zo * V + w]--; NK[zo]--; NDK[d * K + zo]--;   /* detach: unconstrain + unincorporate */
      double ad = al[0], aw = al[1], Aw = aw * (double)V;
      for (int t = 0; t < K; ++t)      /* dirichlet.py:168-173 twice; the z term's denominator is common */
        pr[t] = ((double)NDK[d * K + t] + ad) * ((double)NKW[(int64_t)t * V + w] + aw) / ((double)NK[t] + Aw);
      int zn = gibbs_pick(pr, K, st_uniform(&r));
      if (zn != zo) ++ch;
      Z[i] = (uint8_t)zn; NKW[(int64_t)zn * V + w]++; NK[zn]++; NDK[d * K + zn]++;
    }
    free(pr);
    if (changed) changed[p] = ch;
  }
}
/* state sites of the HMM (blocks = state(2..T) as in orc_hmm_mh) */
void orc_hmm_gibbs(uint8_t *hist, int Tpad, int T, const double *trans, const double *emit, int S, int O,
                   const int32_t *symbols, int64_t P, int64_t p_base, uint64_t seed, uint32_t step0, int n_steps,
                   int block_mode, int64_t *changed) {
  int nb = T - 1;
  for (int64_t p = 0; p < P; ++p) {
    uint8_t *h = hist + p * Tpad; int64_t ch = 0;
    double pr[256];
    for (int k = 0; k < n_steps; ++k) {
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step0 + (uint32_t)k);
      double u_sel = st_uniform(&r); (void)st_uniform(&r);
      int b;
      if (block_mode == -1) { b = (int)(u_sel * nb); if (b >= nb) b = nb - 1; }
      else if (block_mode == -3) b = k % nb; else b = block_mode;
      int t = b + 1, prev = h[t - 1], old = h[t], nxt = (t + 1 < T) ? h[t + 1] : 0;
      for (int s = 0; s < S; ++s) {
        double v = cat_prob(s, trans + (prev - 1) * S, S) * cat_prob(symbols[t] - 1, emit + s * O, O);
        if (nxt) v *= cat_prob(nxt - 1, trans + s * S, S);
        pr[s] = v;
      }
      int sn = 1 + gibbs_pick(pr, S, st_uniform(&r));
      if (sn != old) ++ch;
      h[t] = (uint8_t)sn;
    }
    if (changed) changed[p] = ch;
  }
}

/* `(pgibbs seq ordered Pi K)` on the HMM: PGibbsOperator under mixMH, infer/pgibbs.py:49-126 (propose: the
 * conditional SMC; _compute_alpha :109-121), dispatch.py:197-212, with sampleLogCategorical utils.py:170-180 and
 * normalizeList :93-99.  Border segment 0 is the deterministic state(1): nothing is drawn, every inner particle
 * starts at the weight of the first emission; every later segment: ancestor, then state, per inner particle;
 * weights restart after each resampling.  Draws: 0 unused, 1 accept uniform, then in call order. */
static int sample_log_categorical(orc_stream *r, const double *lw, int n) {
  double m = -INFINITY, ps[64], tot = 0;
  for (int i = 0; i < n; ++i) if (lw[i] > m) m = lw[i];
  for (int i = 0; i < n; ++i) { ps[i] = m > -INFINITY ? exp(lw[i] - m) : 0; tot += ps[i]; }
  for (int i = 0; i < n; ++i) ps[i] = tot > 0 ? ps[i] / tot : 1.0 / n;
  return categorical_sim(r, ps, n);
}
static double lse_small(const double *v, int n) {
  double m = -INFINITY, s = 0;
  for (int i = 0; i < n; ++i) if (v[i] > m) m = v[i];
  if (isinf(m)) return m;
  for (int i = 0; i < n; ++i) s += exp(v[i] - m);
  return m + log(s);
}
void orc_hmm_pgibbs(uint8_t *hist, int Tpad, int T, const double *trans, const double *emit, int S, int O,
                    const int32_t *symbols, int64_t P, int64_t p_base, uint64_t seed, uint32_t step0, int n_steps,
                    int Pi, int64_t *accepted) {
  for (int64_t p = 0; p < P; ++p) {
    uint8_t *h = hist + p * Tpad;
    static uint8_t path[2][64][256];
    double xi[65], nw[64], others[65];
    int64_t acc = 0;
    for (int k = 0; k < n_steps; ++k) {
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step0 + (uint32_t)k);
      (void)st_uniform(&r);
      double logu = log(st_uniform(&r));
      int cb = 0;
      double w0 = categorical_logpmf(symbols[0] - 1, emit + (h[0] - 1) * O, O);
      for (int i = 0; i < Pi; ++i) { xi[i] = w0; path[0][i][0] = h[0]; }
      double rho_prev = w0;
      for (int t = 1; t < T; ++t) {
        xi[Pi] = rho_prev;
        for (int i = 0; i < Pi; ++i) {
          int a = sample_log_categorical(&r, xi, Pi + 1);
          const uint8_t *src = a == Pi ? h : path[cb][a];
          for (int u = 0; u < t; ++u) path[cb ^ 1][i][u] = src[u];
          int s = 1 + categorical_sim(&r, trans + (src[t - 1] - 1) * S, S);
          path[cb ^ 1][i][t] = (uint8_t)s;
          nw[i] = categorical_logpmf(symbols[t] - 1, emit + (s - 1) * O, O);
        }
        for (int i = 0; i < Pi; ++i) xi[i] = nw[i];
        cb ^= 1;
        rho_prev = categorical_logpmf(symbols[t] - 1, emit + (h[t] - 1) * O, O);
      }
      int f = sample_log_categorical(&r, xi, Pi), n_o = 0;
      for (int i = 0; i < Pi; ++i) if (i != f) others[n_o++] = xi[i];
      others[n_o++] = rho_prev;
      double alpha = lse_small(xi, Pi) - lse_small(others, n_o);
      if (logu < alpha) { for (int u = 0; u < T; ++u) h[u] = path[cb][f][u]; ++acc; }
    }
    if (accepted) accepted[p] = acc;
  }
}

/* ------------------------------------------- C5: subsampled MH (BLR) */
/* Keyed Feistel bijection on [0,N) shared by the particle batch: the product's
 * stand-in for np_rng.permutation(local_roots) (subsampled_mh.py:103). */
typedef struct { uint32_t k[4]; uint32_t half_bits, mask; uint64_t N; } orc_feistel;
static uint32_t feistel_mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x;
}
static orc_feistel feistel_key(uint64_t seed, uint32_t step, uint64_t N) {
  orc_feistel K; uint32_t c[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, step, 0x5EED0001u};
  philox(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  for (int i = 0; i < 4; ++i) K.k[i] = c[i];
  uint32_t bits = 1; while (((uint64_t)1 << bits) < N) ++bits;
  K.half_bits = (bits + 1) / 2; if (K.half_bits < 1) K.half_bits = 1;
  K.mask = K.half_bits >= 32 ? 0xFFFFFFFFu : ((1u << K.half_bits) - 1u); K.N = N;
  return K;
}
static uint64_t feistel_perm(uint64_t q, const orc_feistel *K) {
  uint64_t v = q;
  do {
    uint32_t L = (uint32_t)(v >> K->half_bits) & K->mask, R = (uint32_t)v & K->mask;
    for (int r = 0; r < 4; ++r) { uint32_t t = L ^ (feistel_mix(R ^ K->k[r]) & K->mask); L = R; R = t; }
    v = ((uint64_t)L << K->half_bits) | R;
  } while (v >= K->N);
  return v;
}
void orc_feistel_perm(uint64_t seed, uint32_t step, uint64_t N, uint64_t *out) {
  orc_feistel K = feistel_key(seed, step, N);
  for (uint64_t q = 0; q < N; ++q) out[q] = feistel_perm(q, &K);
}
/* infer/subsampled_mh.py:76-111 on the logistic-regression model: global section =
 * {w_j .. array node} (no absorbing node, logGlobalAlpha = 0), local section i =
 * (lookup, vector_dot, log_odds_flip) with dllh_i per evalOneLocalSection (:377-409).
 * stats[p] = (accepted, mean n per transition, last tstat). */
void orc_blr_subsampled_mh(double *w, const double *X, const uint8_t *y, int D, int64_t N, int64_t P,
                           int64_t p_base, double mu0, double s0, uint64_t seed, uint32_t step0, int n_steps,
                           int block_mode, double Nbatch, int k0, double epsilon, int delta_mode, double delta_eps,
                           double *stats) {
  uint64_t *perm = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)N);
  double *acc = (double *)calloc((size_t)P, sizeof(double)), *nsum = (double *)calloc((size_t)P, sizeof(double));
  double *last_t = (double *)calloc((size_t)P, sizeof(double));
  for (int t = 0; t < n_steps; ++t) {
    uint32_t step = step0 + (uint32_t)t;
    orc_feistel_perm(seed, step, (uint64_t)N, perm);
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t p = 0; p < P; ++p) {
      double wp[256], wn[256];
      for (int j = 0; j < D; ++j) wp[j] = wn[j] = w[(int64_t)j * P + p];
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step);
      double u_sel = st_uniform(&r), log_u = log(st_uniform(&r));
      int j;
      if (block_mode == -1) { j = (int)(u_sel * D); if (j >= D) j = D - 1; }
      else if (block_mode == -3) j = t % D; else j = block_mode;
      if (delta_mode == 2) {  /* NormalDriftKernel, continuous.py:68-84 */
        double nu = normal_sim(&r, 0.0, s0);
        wn[j] = mu0 + sqrt(1 - delta_eps * delta_eps) * (wp[j] - mu0) + delta_eps * nu;
      } else wn[j] = normal_sim(&r, mu0, s0);
      double mu_0 = (log_u - 0.0) / (double)N;   /* :102 with alpha = 0 */
      double *dllh = (double *)malloc(sizeof(double) * (size_t)N);
      for (int64_t q = 0; q < N; ++q) {
        int64_t i = (int64_t)perm[q];
        double eo = blr_eta(wp, X + i * D, D), en = blr_eta(wn, X + i * D, D);
        dllh[q] = orc_log_logistic(y[i] ? en : -en) - orc_log_logistic(y[i] ? eo : -eo);
      }
      double n_used, tstat;
      int accept = orc_sequential_test(mu_0, k0, Nbatch, (double)N, epsilon, dllh, &n_used, &tstat);
      free(dllh);
      if (accept) { w[(int64_t)j * P + p] = wn[j]; acc[p] += 1; }
      nsum[p] += n_used; last_t[p] = tstat;
    }
  }
  if (stats) for (int64_t p = 0; p < P; ++p) { stats[3 * p] = acc[p]; stats[3 * p + 1] = n_steps ? nsum[p] / n_steps : 0; stats[3 * p + 2] = last_t[p]; }
  free(perm); free(acc); free(nsum); free(last_t);
}

/* The same transition with the local sections evaluated lazily, minibatch by minibatch (what Lite does and what
 * a CPU implementation would time); identical results (tests/test_oracle_golden.py). */
void orc_blr_subsampled_mh_lazy(double *w, const double *X, const uint8_t *y, int D, int64_t N, int64_t P,
                           int64_t p_base, double mu0, double s0, uint64_t seed, uint32_t step0, int n_steps,
                           int block_mode, double Nbatch, int k0, double epsilon, int delta_mode, double delta_eps,
                           double *stats) {
  uint64_t *perm = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)N);
  double *acc = (double *)calloc((size_t)P, sizeof(double)), *nsum = (double *)calloc((size_t)P, sizeof(double));
  double *last_t = (double *)calloc((size_t)P, sizeof(double));
  for (int t = 0; t < n_steps; ++t) {
    uint32_t step = step0 + (uint32_t)t;
    orc_feistel_perm(seed, step, (uint64_t)N, perm);
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t p = 0; p < P; ++p) {
      double wp[256], wn[256];
      for (int j = 0; j < D; ++j) wp[j] = wn[j] = w[(int64_t)j * P + p];
      orc_stream r; st_init(&r, seed, (uint64_t)(p_base + p), step);
      double u_sel = st_uniform(&r), log_u = log(st_uniform(&r));
      int j;
      if (block_mode == -1) { j = (int)(u_sel * D); if (j >= D) j = D - 1; }
      else if (block_mode == -3) j = t % D; else j = block_mode;
      if (delta_mode == 2) {  /* NormalDriftKernel, continuous.py:68-84 */
        double nu = normal_sim(&r, 0.0, s0);
        wn[j] = mu0 + sqrt(1 - delta_eps * delta_eps) * (wp[j] - mu0) + delta_eps * nu;
      } else wn[j] = normal_sim(&r, mu0, s0);
      double mu_0 = (log_u - 0.0) / (double)N;   /* :102 with alpha = 0 */
      /* sequentialTest (:118-171) with fun_dllh evaluated ONLY for the sections the test consumes, as
       * evalOneLocalSection is in Lite (:377-409) */
      double n_used = 0, tstat = 0, cum = 0, cum2 = 0, kk = 0, Nd = (double)N;
      int accept = -1, have_t = 0;
      while (n_used < Nd) {
        double n_end = fmin(n_used + Nbatch, Nd);
        for (int64_t q = (int64_t)n_used; q < (int64_t)n_end; ++q) {
          int64_t i = (int64_t)perm[q];
          double eo = blr_eta(wp, X + i * D, D), en = blr_eta(wn, X + i * D, D);
          double dl = orc_log_logistic(y[i] ? en : -en) - orc_log_logistic(y[i] ? eo : -eo);
          cum += dl; cum2 += dl * dl;
        }
        kk += 1; n_used = n_end;
        double mx = cum / n_end, mx2 = cum2 / n_end;
        if (kk < k0 && n_used < Nd) continue;
        if (n_used == Nd) { accept = mx >= mu_0; break; }
        double sx = sqrt((1 - (n_used - 1) / (Nd - 1)) * (mx2 - mx * mx) / (n_used - 1)), qv;
        if (sx == 0) qv = mx >= mu_0 ? 1 : 0;
        else { tstat = (mx - mu_0) / sx; have_t = 1; qv = orc_student_t_cdf(tstat, n_used - 1); }
        if (qv < epsilon) { accept = 0; break; }
        else if (qv > 1 - epsilon) { accept = 1; break; }
      }
      tstat = (have_t && tstat != 0) ? tstat : (accept ? INFINITY : -INFINITY);
      if (accept) { w[(int64_t)j * P + p] = wn[j]; acc[p] += 1; }
      nsum[p] += n_used; last_t[p] = tstat;
    }
  }
  if (stats) for (int64_t p = 0; p < P; ++p) { stats[3 * p] = acc[p]; stats[3 * p + 1] = n_steps ? nsum[p] / n_steps : 0; stats[3 * p + 2] = last_t[p]; }
  free(perm); free(acc); free(nsum); free(last_t);
}

/* Ancestor search restricted to one shard's mass interval (multi-GPU resampling):
 * child k gets the same threshold as in orc_ancestors; ancestors outside
 * [offset, offset + cdf[n-1]) are reported as -1. */
void orc_shard_ancestors(const uint64_t *cdf, int64_t n, uint64_t offset, uint64_t total, int64_t n_new,
                         int64_t k_begin, int64_t k_count, int mode, uint64_t seed, uint32_t step, int32_t *anc) {
  uint64_t mass = cdf[n - 1];
  double u_sys = 0;
  if (mode == 1) { orc_stream r; st_init(&r, seed, 0xFFFFFFFFFFFFFFFFull, step); u_sys = st_uniform(&r); }
  for (int64_t t = 0; t < k_count; ++t) {
    int64_t k = k_begin + t; uint64_t thr;
    if (mode == 1) {
      uint64_t U0 = (uint64_t)(u_sys * 4294967296.0); if (U0 > 0xFFFFFFFFull) U0 = 0xFFFFFFFFull;
      thr = (uint64_t)((((((u128)(uint64_t)k) << 32) | U0) * (u128)total) / (((u128)(uint64_t)n_new) << 32));
    } else {
      orc_stream r; st_init(&r, seed, (uint64_t)k, step); double u = st_uniform(&r);
      uint64_t U = (uint64_t)(u * 9007199254740992.0); if (U > 9007199254740991ull) U = 9007199254740991ull;
      thr = (uint64_t)(((u128)U * (u128)total) >> 53);
    }
    anc[t] = (thr >= offset && thr - offset < mass) ? (int32_t)upper_bound(cdf, n, thr - offset) : -1;
  }
}
