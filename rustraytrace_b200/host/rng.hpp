// rng.hpp — the scene-construction RNG of the host mirror.
//
// The reference draws every scene random from rand 0.8.5's SmallRng (Cargo.lock: rand 0.8.5,
// rand_core 0.6.4), which on 64-bit targets is Xoshiro256++:
//   * CPU books: thread-local SmallRng::from_entropy()  (the_rest_of_your_life/rtweekend.rs:9-11)
//     -> not reproducible, only the distribution matters;
//   * GPU path:  SmallRng::seed_from_u64(0x5EED_1234)    (src/gpu/mod.rs:199) -> deterministic.
// We always seed explicitly.  The algorithm is restated from the published xoshiro256++
// (Blackman & Vigna) and rand 0.8.5's documented float mappings; no reference test pins it
// ("parity unpinned") — the known-answer vector for state {1,2,3,4} from the
// xoshiro reference implementation is checked in tests/test_host_rng.py.
#pragma once
#include <cstdint>

namespace rrt_host {

struct SmallRng {
  uint64_t s[4];

  static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }

  // rand_core::SeedableRng::seed_from_u64 as overridden for Xoshiro256PlusPlus: SplitMix64 stream.
  static SmallRng seed_from_u64(uint64_t state) {
    SmallRng r;
    for (int i = 0; i < 4; ++i) {
      state += 0x9e3779b97f4a7c15ULL;
      uint64_t z = state;
      z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
      z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
      z = z ^ (z >> 31);
      r.s[i] = z;
    }
    return r;
  }
  static SmallRng from_state(uint64_t a, uint64_t b, uint64_t c, uint64_t d) {
    SmallRng r; r.s[0] = a; r.s[1] = b; r.s[2] = c; r.s[3] = d; return r;
  }

  uint64_t next_u64() {
    const uint64_t result = rotl(s[0] + s[3], 23) + s[0];
    const uint64_t t = s[1] << 17;
    s[2] ^= s[0];
    s[3] ^= s[1];
    s[1] ^= s[2];
    s[0] ^= s[3];
    s[2] ^= t;
    s[3] = rotl(s[3], 45);
    return result;
  }
  // rand 0.8.5 xoshiro256plusplus.rs: the upper bits, "the lowest bits have linear dependencies".
  uint32_t next_u32() { return (uint32_t)(next_u64() >> 32); }

  // Standard distribution: rng.gen::<f64>() / gen::<f32>()  (src/gpu/mod.rs:212-238)
  double gen_f64() { return (double)(next_u64() >> 11) * (1.0 / 9007199254740992.0); }
  float gen_f32() { return (float)(next_u32() >> 8) * (1.0f / 16777216.0f); }

  // rng.gen_range(low..high) for f64 (UniformFloat::sample_single): a [1,2) mantissa fill,
  // minus one, scaled.  rtweekend.rs:17-25.
  double gen_range_f64(double low, double high) {
    double scale = high - low;
    for (;;) {
      union { uint64_t u; double d; } v;
      v.u = (next_u64() >> 12) | 0x3ff0000000000000ULL;
      double res = (v.d - 1.0) * scale + low;
      if (res < high) return res;
    }
  }
  float gen_range_f32(float low, float high) {
    float scale = high - low;
    for (;;) {
      union { uint32_t u; float f; } v;
      v.u = (next_u32() >> 9) | 0x3f800000u;
      float res = (v.f - 1.0f) * scale + low;
      if (res < high) return res;
    }
  }

  // rtweekend.rs:17-30
  double random_double() { return gen_range_f64(0.0, 1.0); }
  double random_double_range(double mn, double mx) { return gen_range_f64(mn, mx); }
  int random_int(int mn, int mx) { return (int)random_double_range((double)mn, (double)(mx + 1)); }
};

}  // namespace rrt_host
