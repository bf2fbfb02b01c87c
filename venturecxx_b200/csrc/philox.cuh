// Counter-based Philox4x32-10 streams (DESIGN.md "RNG contract").
//
// The reference gives every trace a numpy RandomState and a Python
// random.Random seeded from one seed (backend/lite/trace.py:89-92) and draws
// sequentially from them.  Here every draw is addressed by
//   (seed, particle, step, draw index)
// so a particle's stream does not depend on scheduling, on the number of
// GPUs, or on which rank owns it.
#pragma once
#include <stdint.h>

#ifndef VK_HD
#define VK_HD __host__ __device__ __forceinline__
#endif

namespace vk {

struct Philox4 { uint32_t v[4]; };

VK_HD uint32_t mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

VK_HD Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                            uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = mulhi32(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = mulhi32(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  Philox4 out;
  out.v[0] = c0; out.v[1] = c1; out.v[2] = c2; out.v[3] = c3;
  return out;
}

// 53 random bits -> a double strictly inside (0,1).  x = 2^53 - 1 would round to exactly 1.0 (x + 0.5 is a
// tie that rounds to even); that one value is clamped to 1 - 2^-53.
VK_HD double u01_from_bits(uint32_t hi, uint32_t lo) {
  uint64_t x = (((uint64_t)hi << 32) | (uint64_t)lo) >> 11;
  const double u = ((double)x + 0.5) * (1.0 / 9007199254740992.0);
  return u < 1.0 ? u : 0.99999999999999988898;
}

// Sequential view over the draws of one (seed, particle, step).
// Draw d lives in Philox block d>>1, words (2(d&1), 2(d&1)+1).
struct Stream {
  uint32_t k0, k1, p0, p1, step, d;
  Philox4 blk;
  VK_HD Stream(uint64_t seed, uint64_t particle, uint32_t step_)
      : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)),
        p0((uint32_t)particle), p1((uint32_t)(particle >> 32)), step(step_),
        d(0) {}
  VK_HD double uniform() {
    if ((d & 1u) == 0u) blk = philox4x32_10(p0, p1, step, d >> 1, k0, k1);
    uint32_t w = (d & 1u) * 2u;
    double u = u01_from_bits(blk.v[w], blk.v[w + 1]);
    ++d;
    return u;
  }
  // skip to an even draw index (start of a fresh block)
  VK_HD void seek(uint32_t draw) { d = draw; if (d & 1u) blk = philox4x32_10(p0, p1, step, d >> 1, k0, k1); }
};

// particle id reserved for draws shared by the whole particle set
// (e.g. the single uniform of systematic resampling).
static const uint64_t GLOBAL_PARTICLE = 0xFFFFFFFFFFFFFFFFull;

}  // namespace vk
