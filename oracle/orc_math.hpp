// orc_math.hpp — ORACLE (test infrastructure; never linked into or imported by the product).
//
// CPU f64 restatement of the reference's math/RNG layer (L0 of SURVEY.md §1).  PARITY UNPINNED:
// the reference has no tests, golden vectors or runnable binary here (no Rust toolchain); this
// restatement is pinned only by known-answer cases derived by hand from the cited lines
// (tests/test_oracle_kat.py).
//
//   Vec3      src/books/the_rest_of_your_life/vec3.rs:5-222
//   Ray       src/books/the_rest_of_your_life/ray.rs:4-34
//   Interval  src/books/the_rest_of_your_life/interval.rs:6-62
//   RNG       src/books/the_rest_of_your_life/rtweekend.rs:9-30 (rand 0.8.5 SmallRng = Xoshiro256++,
//             seeded explicitly here: documented departure #1)
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>

namespace orc {

const double INF = std::numeric_limits<double>::infinity();
const double PI = 3.14159265358979323846264338327950288;

inline double degrees_to_radians(double d) { return d * PI / 180.0; }  // rtweekend.rs:13-15

// ---- RNG: xoshiro256++ (Blackman & Vigna), seeded through SplitMix64 like rand's seed_from_u64
struct Rng {
  uint64_t s[4];
  static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
  explicit Rng(uint64_t seed = 0) {
    for (int i = 0; i < 4; ++i) {
      seed += 0x9e3779b97f4a7c15ULL;
      uint64_t z = seed;
      z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
      z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
      s[i] = z ^ (z >> 31);
    }
  }
  uint64_t next() {
    uint64_t r = rotl(s[0] + s[3], 23) + s[0];
    uint64_t t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
    s[2] ^= t; s[3] = rotl(s[3], 45);
    return r;
  }
  // gen_range(min..max) for f64: 52 random mantissa bits in [1,2), minus 1, scaled (rand 0.8.5
  // UniformFloat::sample_single)
  double range(double mn, double mx) {
    for (;;) {
      union { uint64_t u; double d; } v;
      v.u = (next() >> 12) | 0x3ff0000000000000ULL;
      double r = (v.d - 1.0) * (mx - mn) + mn;
      if (r < mx) return r;
    }
  }
};

// Per-thread sampling stream + operation counters (departure #4).
struct Counters {
  double aabb = 0, sph_test = 0, sph_hit = 0, quad_test = 0, quad_full = 0, inst = 0, medium = 0;
  double segments = 0, paths = 0;
  double lambertian = 0, metal = 0, dielectric = 0, isotropic = 0, emitted = 0;
  double tex_checker = 0, tex_image = 0, tex_noise = 0, pdf_mix = 0, rng_draws = 0;
  void add(const Counters& o) {
    aabb += o.aabb; sph_test += o.sph_test; sph_hit += o.sph_hit; quad_test += o.quad_test;
    quad_full += o.quad_full; inst += o.inst; medium += o.medium; segments += o.segments; paths += o.paths;
    lambertian += o.lambertian; metal += o.metal; dielectric += o.dielectric; isotropic += o.isotropic;
    emitted += o.emitted; tex_checker += o.tex_checker; tex_image += o.tex_image; tex_noise += o.tex_noise;
    pdf_mix += o.pdf_mix; rng_draws += o.rng_draws;
  }
};

struct Ctx {  // what the reference keeps in thread_local RNG (rtweekend.rs:9-11)
  Rng rng;
  Counters n;
  explicit Ctx(uint64_t seed) : rng(seed) {}
  double random_double() { n.rng_draws += 1; return rng.range(0.0, 1.0); }                    // :17-20
  double random_double_range(double a, double b) { n.rng_draws += 1; return rng.range(a, b); }  // :22-25
  int random_int(int a, int b) { return (int)random_double_range((double)a, (double)(b + 1)); } // :27-30
};

// ---- Vec3
struct Vec3 {
  double e[3];
  Vec3() : e{0, 0, 0} {}
  Vec3(double a, double b, double c) : e{a, b, c} {}
  double x() const { return e[0]; }
  double y() const { return e[1]; }
  double z() const { return e[2]; }
  double operator[](int i) const { return e[i]; }
  double& operator[](int i) { return e[i]; }
  double length_squared() const { return e[0] * e[0] + e[1] * e[1] + e[2] * e[2]; }
  double length() const { return std::sqrt(length_squared()); }
  bool near_zero() const {  // vec3.rs:38-41
    double s = 1e-8;
    return std::fabs(e[0]) < s && std::fabs(e[1]) < s && std::fabs(e[2]) < s;
  }
};
typedef Vec3 Point3;
typedef Vec3 Color;
inline Vec3 operator-(Vec3 v) { return Vec3(-v.e[0], -v.e[1], -v.e[2]); }
inline Vec3 operator+(Vec3 u, Vec3 v) { return Vec3(u.e[0] + v.e[0], u.e[1] + v.e[1], u.e[2] + v.e[2]); }
inline Vec3 operator-(Vec3 u, Vec3 v) { return Vec3(u.e[0] - v.e[0], u.e[1] - v.e[1], u.e[2] - v.e[2]); }
inline Vec3 operator*(Vec3 u, Vec3 v) { return Vec3(u.e[0] * v.e[0], u.e[1] * v.e[1], u.e[2] * v.e[2]); }
inline Vec3 operator*(double t, Vec3 v) { return Vec3(t * v.e[0], t * v.e[1], t * v.e[2]); }
inline Vec3 operator*(Vec3 v, double t) { return t * v; }
inline Vec3 operator/(Vec3 v, double t) { return (1.0 / t) * v; }  // vec3.rs: Div = self * (1.0 / rhs)
inline double dot(Vec3 u, Vec3 v) { return u.e[0] * v.e[0] + u.e[1] * v.e[1] + u.e[2] * v.e[2]; }
inline Vec3 cross(Vec3 u, Vec3 v) {
  return Vec3(u.e[1] * v.e[2] - u.e[2] * v.e[1], u.e[2] * v.e[0] - u.e[0] * v.e[2], u.e[0] * v.e[1] - u.e[1] * v.e[0]);
}
inline Vec3 unit_vector(Vec3 v) { return v / v.length(); }

inline Vec3 random_vec_range(Ctx& c, double mn, double mx) {  // vec3.rs:47-53
  double a = c.random_double_range(mn, mx), b = c.random_double_range(mn, mx), d = c.random_double_range(mn, mx);
  return Vec3(a, b, d);
}
inline Vec3 random_in_unit_disk(Ctx& c) {  // vec3.rs:172-179
  for (;;) {
    double a = c.random_double_range(-1.0, 1.0), b = c.random_double_range(-1.0, 1.0);
    Vec3 p(a, b, 0.0);
    if (p.length_squared() < 1.0) return p;
  }
}
inline Vec3 random_unit_vector(Ctx& c) {  // vec3.rs:181-189
  for (;;) {
    Vec3 p = random_vec_range(c, -1.0, 1.0);
    double lensq = p.length_squared();
    if (1e-160 < lensq && lensq <= 1.0) return p / std::sqrt(lensq);
  }
}
inline Vec3 reflect(Vec3 v, Vec3 n) { return v - 2.0 * dot(v, n) * n; }  // :201-203
inline Vec3 refract(Vec3 uv, Vec3 n, double etai_over_etat) {            // :205-210
  double cos_theta = std::fmin(-dot(uv, n), 1.0);
  Vec3 r_out_perp = etai_over_etat * (uv + cos_theta * n);
  Vec3 r_out_parallel = -std::sqrt(std::fabs(1.0 - r_out_perp.length_squared())) * n;
  return r_out_perp + r_out_parallel;
}
inline Vec3 random_cosine_direction(Ctx& c) {  // :212-222
  double r1 = c.random_double(), r2 = c.random_double();
  double phi = 2.0 * PI * r1;
  double x = std::cos(phi) * std::sqrt(r2);
  double y = std::sin(phi) * std::sqrt(r2);
  double z = std::sqrt(1.0 - r2);
  return Vec3(x, y, z);
}

// ---- Ray
struct Ray {
  Point3 orig;
  Vec3 dir;
  double tm;
  Ray() : tm(0) {}
  Ray(Point3 o, Vec3 d) : orig(o), dir(d), tm(0.0) {}
  Ray(Point3 o, Vec3 d, double t) : orig(o), dir(d), tm(t) {}
  Point3 at(double t) const { return orig + t * dir; }
};

// ---- Interval
struct Interval {
  double min, max;
  Interval() : min(INF), max(-INF) {}  // EMPTY
  Interval(double a, double b) : min(a), max(b) {}
  static Interval from_intervals(Interval a, Interval b) {  // :19-24
    return Interval(a.min <= b.min ? a.min : b.min, a.max >= b.max ? a.max : b.max);
  }
  double size() const { return max - min; }
  bool contains(double x) const { return min <= x && x <= max; }   // :31-33 closed
  bool surrounds(double x) const { return min < x && x < max; }    // :35-37 open
  double clamp(double x) const { return x < min ? min : (x > max ? max : x); }
  Interval expand(double delta) const { double p = delta / 2.0; return Interval(min - p, max + p); }
  Interval shifted(double d) const { return Interval(min + d, max + d); }
};
const Interval UNIVERSE(-INF, INF);

}  // namespace orc
