/* Minimal stand-in for <gsl/gsl_randist.h> (see gsl_rng.h in this directory). */
#ifndef VK_GSL_SHIM_RANDIST_H
#define VK_GSL_SHIM_RANDIST_H
#include <stddef.h>
#include <random>
#include "gsl_rng.h"
static inline double gsl_ran_gaussian(gsl_rng *r, double sigma) {
  std::normal_distribution<double> d(0.0, sigma);
  return d(r->eng);
}
static inline double gsl_ran_flat(gsl_rng *r, double a, double b) {
  std::uniform_real_distribution<double> d(a, b);
  return d(r->eng);
}
/* N trials over K categories with probabilities p (need not be normalised), counts into n */
static inline void gsl_ran_multinomial(gsl_rng *r, size_t K, unsigned int N, const double p[], unsigned int n[]) {
  double norm = 0.0;
  for (size_t k = 0; k < K; ++k) { norm += p[k]; n[k] = 0; }
  std::uniform_real_distribution<double> d(0.0, 1.0);
  for (unsigned int t = 0; t < N; ++t) {
    double u = d(r->eng) * norm, acc = 0.0;
    size_t k = 0;
    for (; k + 1 < K; ++k) { acc += p[k]; if (u < acc) break; }
    n[k] += 1;
  }
}
#endif
