/* Minimal stand-in for <gsl/gsl_rng.h>, test infrastructure only: lets the reference's own
 * benchmarks/normal-normal-chain.cxx compile UNMODIFIED from where it lies (GSL is not installed in this
 * image).  SURVEY 8(c) C3: "gsl_rng_mt19937 / gsl_ran_gaussian / gsl_ran_flat swapped for std::mt19937 /
 * std::normal_distribution / std::uniform_real_distribution".  Only the calls that file makes exist. */
#ifndef VK_GSL_SHIM_RNG_H
#define VK_GSL_SHIM_RNG_H
#include <random>
struct gsl_rng { std::mt19937 eng; };
typedef int gsl_rng_type;
static const gsl_rng_type *gsl_rng_mt19937 = 0;
static inline gsl_rng *gsl_rng_alloc(const gsl_rng_type *) { return new gsl_rng(); }
static inline void gsl_rng_set(gsl_rng *r, unsigned long seed) { r->eng.seed((std::mt19937::result_type)seed); }
static inline void gsl_rng_free(gsl_rng *r) { delete r; }
#endif
