// rng.cuh — counter-based Philox4x32 (Salmon et al., SC'11), keyed on the render seed and
// counted by (pixel, sample, bounce, draw block).  The kernel runs 7 rounds, the fewest the paper
// reports as Crush-resistant for Philox4x32 (10 is its default safety margin): 3 rounds fewer are
// worth 3 % on C2/C5 and 6 % on C4 (profiles/README.md).  rrt_philox_kat checks the same round
// function at 10 rounds against Random123's known answers.  Replaces the reference's thread-local
// SmallRng::from_entropy() (the_rest_of_your_life/rtweekend.rs:9-25): only the distribution of
// random_double() — uniform in [0,1) — is specified there, so any partition of the frame over
// blocks, launches or GPUs sees the same numbers for the same (pixel, sample, bounce).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rrt {

struct Rand4 { float x, y, z, w, v; };  // v: a fifth uniform built from the low bytes u01 does not use

template <int ROUNDS>
__host__ __device__ inline void philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                          uint32_t k0, uint32_t k1, uint32_t out[4]) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int r = 0; r < ROUNDS; ++r) {
#if defined(__CUDA_ARCH__)
    uint32_t hi0 = __umulhi(M0, c0), hi1 = __umulhi(M1, c2);
#else
    uint32_t hi0 = (uint32_t)(((uint64_t)M0 * c0) >> 32), hi1 = (uint32_t)(((uint64_t)M1 * c2) >> 32);
#endif
    uint32_t lo0 = M0 * c0, lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__host__ __device__ inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                             uint32_t k0, uint32_t k1, uint32_t out[4]) {
  philox4x32<10>(c0, c1, c2, c3, k0, k1, out);
}

#ifndef RRT_PHILOX_ROUNDS
#define RRT_PHILOX_ROUNDS 7
#endif

// 24-bit mantissa fill: uniform in [0,1), never 1.0
__host__ __device__ inline float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// The same function with the key schedule precomputed: rk[2r], rk[2r+1] = (k0 + r W0, k1 + r W1).  The
// render kernel reads them straight from its parameter block (constant-bank operands of the LOP3s, no
// per-round key additions), and the two products of a round are one 32x32 -> 64-bit multiply each
// (mul.wide.u32 -> IMAD.WIDE.U32) instead of a low and a high half.
#if defined(__CUDACC__)
template <int ROUNDS>
__device__ __forceinline__ void philox4x32_rk(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              const uint32_t* __restrict__ rk, uint32_t out[4]) {
#pragma unroll
  for (int r = 0; r < ROUNDS; ++r) {
    unsigned long long p0, p1;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(p0) : "r"(c0), "r"(0xD2511F53u));
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(p1) : "r"(c2), "r"(0xCD9E8D57u));
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ rk[2 * r];
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ rk[2 * r + 1];
    c0 = n0; c1 = (uint32_t)p1; c2 = n2; c3 = (uint32_t)p0;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
#endif
// the key schedule (api.cu fills DevWork::rkeys with it)
__host__ __device__ inline void philox_round_keys(uint32_t k0, uint32_t k1, uint32_t* rk, int rounds) {
  for (int r = 0; r < rounds; ++r) { rk[2 * r] = k0 + (uint32_t)r * 0x9E3779B9u; rk[2 * r + 1] = k1 + (uint32_t)r * 0xBB67AE85u; }
}

#if defined(__CUDACC__)
struct PathRng {
  uint32_t pixel, sample;
  const uint32_t* rk;   // 2 x RRT_PHILOX_ROUNDS round keys (DevWork::rkeys, or a local array)
  __device__ inline Rand4 draw(uint32_t bounce, uint32_t block) const {
    uint32_t o[4];
    philox4x32_rk<RRT_PHILOX_ROUNDS>(pixel, sample, bounce, block, rk, o);
    Rand4 r; r.x = u01(o[0]); r.y = u01(o[1]); r.z = u01(o[2]); r.w = u01(o[3]);
    r.v = (float)(((o[0] & 255u) << 16) | ((o[1] & 255u) << 8) | (o[2] & 255u)) * (1.0f / 16777216.0f);
    return r;
  }
};
#endif

}  // namespace rrt
