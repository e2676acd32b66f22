// csrc/rng.cuh -- counter-based random numbers for the stochastic stages (Philox4x32-10, Salmon et al. SC'11).
// Stands in for the reference's thread-local mt19937 (RNThreadableRandomScalar, RNScalar.cpp:99-131), which is
// seeded from random_device and therefore not reproducible.  Host-compilable for the CPU known-answer test.
#pragma once
#include <stdint.h>
#include "gi_math.cuh"
#ifndef GI_DEV
#if defined(__CUDACC__)
#define GI_DEV __device__ __forceinline__
#else
#define GI_DEV inline
#endif
#endif

namespace gi {

// ------------------------------------------------------------------------------------------------ RNG
struct Rng {
  uint32_t k0, k1, c0, c1, c2, c3;
  uint32_t out[4];
  int left;
};

GI_DEV void philox_round(uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3, uint32_t k0, uint32_t k1) {
  const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
  const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
  c0 = n0; c1 = n1; c2 = n2; c3 = n3;
}

GI_DEV void rng_init(Rng &r, uint64_t seed, uint64_t index, uint32_t stage, uint32_t sample) {
  r.k0 = (uint32_t)seed; r.k1 = (uint32_t)(seed >> 32);
  r.c0 = (uint32_t)index; r.c1 = (uint32_t)(index >> 32) ^ (stage << 12);      // index < 2^44, stage < 2^20
  r.c2 = sample; r.c3 = 0;
  r.left = 0;
}

GI_DEV uint32_t rng_u32(Rng &r) {
  if (r.left == 0) {
    uint32_t a = r.c0, b = r.c1, c = r.c2, d = r.c3, k0 = r.k0, k1 = r.k1;
    for (int i = 0; i < 10; i++) { philox_round(a, b, c, d, k0, k1); k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
    r.out[0] = a; r.out[1] = b; r.out[2] = c; r.out[3] = d;
    r.c3++;
    r.left = 4;
  }
  return r.out[--r.left];
}

// uniform in [0, 1) with 53 random bits (stands in for RNThreadableRandomScalar, RNScalar.cpp:125-131)
GI_DEV double rng_double(Rng &r) {
  const uint64_t hi = rng_u32(r), lo = rng_u32(r);
  return (double)(((hi << 32) | lo) >> 11) * (1.0 / 9007199254740992.0);
}

}  // namespace gi
